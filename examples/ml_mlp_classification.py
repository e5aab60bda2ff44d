"""Spark ML pipeline, classification (reference: examples/ml_mlp_classification.py)."""
from _data import mnist_like
from elephas.ml.adapter import to_data_frame
from elephas.ml_model import ElephasEstimator
from elephas_b200.models import optimizers, zoo
from elephas_b200.spark import DoubleType, MulticlassMetrics, Pipeline, SparkConf, SparkContext, functions

sc = SparkContext(conf=SparkConf().setAppName('Mnist_Spark_MLP').setMaster('local[8]'))
x_train, y_train, x_test, y_test = mnist_like(3000, 500)
df = to_data_frame(sc, x_train, y_train, categorical=True)
test_df = to_data_frame(sc, x_test, y_test, categorical=True)
sgd_conf = optimizers.serialize(optimizers.SGD(learning_rate=0.01, decay=1e-6, momentum=0.9, nesterov=True))

estimator = ElephasEstimator()
estimator.set_keras_model_config(zoo.mnist_mlp().to_json())
estimator.set_optimizer_config(sgd_conf)
estimator.set_mode("synchronous")
estimator.set_loss("categorical_crossentropy")
estimator.set_metrics(['acc'])
estimator.set_epochs(3)
estimator.set_batch_size(64)
estimator.set_validation_split(0.1)
estimator.set_categorical_labels(True)
estimator.set_nb_classes(10)
estimator.set_num_workers(2)

fitted_pipeline = Pipeline(stages=[estimator]).fit(df)
prediction = fitted_pipeline.transform(test_df)
pnl = prediction.select("label", functions.array_argmax("prediction").astype(DoubleType()).alias("prediction"))
pnl.show(10)
metrics = MulticlassMetrics(pnl.rdd.map(lambda row: (row.prediction, row.label)))
print('precision', metrics.precision(), 'recall', metrics.recall())
