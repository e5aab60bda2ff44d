"""Spark ML pipeline, regression (reference: examples/ml_mlp_regression.py)."""
from _data import boston_like
from elephas.ml.adapter import to_data_frame
from elephas.ml_model import ElephasEstimator
from elephas_b200.models import optimizers, zoo
from elephas_b200.spark import Pipeline, RegressionMetrics, SparkConf, SparkContext

sc = SparkContext(conf=SparkConf().setAppName('Boston_Spark_MLP').setMaster('local[8]'))
x_train, y_train, x_test, y_test = boston_like()
df = to_data_frame(sc, x_train, y_train, False)
test_df = to_data_frame(sc, x_test, y_test, False)
estimator = ElephasEstimator()
estimator.set_keras_model_config(zoo.boston_mlp().to_json())
estimator.set_optimizer_config(optimizers.serialize(optimizers.SGD(learning_rate=0.00001)))
estimator.set_mode("synchronous")
estimator.set_loss("mae")
estimator.set_metrics(['mae'])
estimator.set_epochs(10)
estimator.set_batch_size(64)
estimator.set_validation_split(0.01)
estimator.set_categorical_labels(False)
estimator.set_num_workers(2)
fitted_pipeline = Pipeline(stages=[estimator]).fit(df)
pnl = fitted_pipeline.transform(test_df).select("label", "prediction")
pnl.show(10)
print('r2', RegressionMetrics(pnl.rdd.map(lambda row: (row.prediction, row.label))).r2)
