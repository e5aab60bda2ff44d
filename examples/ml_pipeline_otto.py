"""Otto-shaped pipeline: StringIndexer -> StandardScaler -> ElephasEstimator, hogwild mode
(reference: examples/ml_pipeline_otto.py and examples/Spark_ML_Pipeline.ipynb; BASELINE.json config 3)."""
from _data import otto_like
from elephas.ml_model import ElephasEstimator
from elephas_b200.models import optimizers, zoo
from elephas_b200.spark import (DoubleType, MulticlassMetrics, Pipeline, SparkSession, StandardScaler, StringIndexer,
                                Vectors, functions)

spark_session = SparkSession.builder.appName('Otto_Spark_ML_Pipeline').getOrCreate()
feats, cats = otto_like()
train_df = spark_session.createDataFrame([(Vectors.dense(f), c) for f, c in zip(feats, cats)], ['features', 'category'])

string_indexer = StringIndexer(inputCol="category", outputCol="index_category")
scaler = StandardScaler(inputCol="features", outputCol="scaled_features", withStd=True, withMean=True)
nb_classes = train_df.select("category").distinct().count()
model = zoo.otto_mlp(input_dim=93, classes=nb_classes)

estimator = ElephasEstimator(featuresCol="scaled_features", labelCol="index_category")
estimator.set_keras_model_config(model.to_json())
estimator.set_optimizer_config(optimizers.serialize(optimizers.Adam(learning_rate=0.005)))
estimator.set_mode("hogwild")
estimator.set_loss("categorical_crossentropy")
estimator.set_metrics(['acc'])
estimator.set_epochs(10)
estimator.set_batch_size(128)
estimator.set_num_workers(2)
estimator.set_verbosity(0)
estimator.set_validation_split(0.15)
estimator.set_categorical_labels(True)
estimator.set_nb_classes(nb_classes)

fitted_pipeline = Pipeline(stages=[string_indexer, scaler, estimator]).fit(train_df)
prediction = fitted_pipeline.transform(train_df)
pnl = prediction.select("index_category", functions.array_argmax("prediction").astype(DoubleType()).alias("p"))
pnl.show(10)
print('train precision:', MulticlassMetrics(pnl.rdd.map(lambda r: (r.p, r.index_category))).precision())
