"""Spark ML stages (reference: tests/test_ml_model.py)."""
import numpy as np
import pytest

from elephas.ml.adapter import to_data_frame
from elephas.ml_model import ElephasEstimator, ElephasTransformer, load_ml_estimator, load_ml_transformer
from elephas.utils.model_utils import ModelType
from elephas_b200.models import Dense, Sequential, activations, optimizers
from elephas_b200.spark import DoubleType, MulticlassMetrics, Pipeline, PipelineModel, RegressionMetrics, functions


def _estimator(model, sgd_conf, loss="categorical_crossentropy", metrics=('acc',), epochs=1, batch_size=64,
               validation_split=0.1, categorical=True, nb_classes=10, **ctor):
    estimator = ElephasEstimator(**ctor)
    estimator.set_keras_model_config(model.to_json())
    estimator.set_optimizer_config(sgd_conf)
    estimator.set_mode("synchronous")
    estimator.set_loss(loss)
    estimator.set_metrics(list(metrics))
    estimator.set_epochs(epochs)
    estimator.set_batch_size(batch_size)
    estimator.set_validation_split(validation_split)
    estimator.set_categorical_labels(categorical)
    estimator.set_nb_classes(nb_classes)
    estimator.set_num_workers(2)
    return estimator


def test_serialization_transformer(tmp_path, classification_model):
    transformer = ElephasTransformer()
    transformer.set_keras_model_config(classification_model.to_json())
    transformer.save(str(tmp_path / "test.h5"))
    loaded_model = load_ml_transformer(str(tmp_path / "test.h5"))
    assert loaded_model.get_model().to_json() == classification_model.to_json()


def test_serialization_estimator(tmp_path, classification_model):
    estimator = ElephasEstimator()
    estimator.set_keras_model_config(classification_model.to_json())
    estimator.set_loss("categorical_crossentropy")
    estimator.save(str(tmp_path / "test.h5"))
    loaded_model = load_ml_estimator(str(tmp_path / "test.h5"))
    assert loaded_model.get_model().to_json() == classification_model.to_json()
    assert loaded_model.get_loss() == "categorical_crossentropy"


def test_serialization_transformer_and_predict(tmp_path, spark_context, classification_model, mnist_data):
    _, _, x_test, y_test = mnist_data
    df = to_data_frame(spark_context, x_test, y_test, categorical=True)
    transformer = ElephasTransformer(weights=classification_model.get_weights(), model_type=ModelType.CLASSIFICATION)
    transformer.set_keras_model_config(classification_model.to_json())
    transformer.save(str(tmp_path / "test.h5"))
    loaded_transformer = load_ml_transformer(str(tmp_path / "test.h5"))
    assert loaded_transformer.model_type == ModelType.CLASSIFICATION
    out = loaded_transformer.transform(df)
    expected = classification_model.predict(x_test[:1])[0]
    assert np.allclose(out.first()[loaded_transformer.getOutputCol()], expected, atol=1e-6)


def test_spark_ml_model_classification(spark_context, classification_model, mnist_data):
    x_train, y_train, x_test, y_test = mnist_data
    df = to_data_frame(spark_context, x_train, y_train, categorical=True)
    test_df = to_data_frame(spark_context, x_test, y_test, categorical=True)
    sgd = optimizers.SGD(learning_rate=0.01, decay=1e-6, momentum=0.9, nesterov=True)
    estimator = _estimator(classification_model, optimizers.serialize(sgd), epochs=3)
    fitted_pipeline = Pipeline(stages=[estimator]).fit(df)
    assert isinstance(fitted_pipeline, PipelineModel)
    prediction = fitted_pipeline.transform(test_df)
    pnl = prediction.select("label", functions.array_argmax("prediction").astype(DoubleType()).alias("prediction"))
    metrics = MulticlassMetrics(pnl.rdd.map(lambda row: (row.label, row.prediction)))
    assert metrics.accuracy > 0.5
    transformer = fitted_pipeline.stages[0]
    assert len(transformer.history) == 2 and 'loss' in transformer.history[0]


def test_functional_model(spark_context, classification_model_functional, mnist_data):
    x_train, y_train, x_test, y_test = mnist_data
    df = to_data_frame(spark_context, x_train, y_train, categorical=True)
    test_df = to_data_frame(spark_context, x_test, y_test, categorical=True)
    estimator = _estimator(classification_model_functional, optimizers.serialize(optimizers.SGD()))
    prediction = Pipeline(stages=[estimator]).fit(df).transform(test_df)
    assert len(prediction.first().prediction) == 10


def test_regression_model(spark_context, regression_model, boston_housing_dataset):
    x_train, y_train, x_test, y_test = boston_housing_dataset
    df = to_data_frame(spark_context, x_train, y_train)
    test_df = to_data_frame(spark_context, x_test, y_test)
    sgd_conf = optimizers.serialize(optimizers.SGD(learning_rate=0.00001))
    estimator = _estimator(regression_model, sgd_conf, loss="mae", metrics=['mae'], epochs=3,
                           validation_split=0.01, categorical=False)
    prediction = Pipeline(stages=[estimator]).fit(df).transform(test_df)
    pnl = prediction.select("label", "prediction")
    assert isinstance(pnl.first().prediction, float)
    metrics = RegressionMetrics(pnl.rdd.map(lambda row: (row.prediction, row.label)))
    assert np.isfinite(metrics.r2)


def test_set_cols_deprecated(spark_context, regression_model, boston_housing_dataset):
    x_train, y_train, x_test, y_test = boston_housing_dataset
    df = to_data_frame(spark_context, x_train, y_train)
    df = df.withColumnRenamed('features', 'scaled_features').withColumnRenamed('label', 'ground_truth')
    test_df = to_data_frame(spark_context, x_test, y_test)
    test_df = test_df.withColumnRenamed('features', 'scaled_features').withColumnRenamed('label', 'ground_truth')
    sgd_conf = optimizers.serialize(optimizers.SGD(learning_rate=0.00001))
    estimator = _estimator(regression_model, sgd_conf, loss="mae", metrics=['mae'], validation_split=0.01,
                           categorical=False)
    with pytest.deprecated_call():
        estimator.setFeaturesCol('scaled_features')
    with pytest.deprecated_call():
        estimator.setOutputCol('output')
    with pytest.deprecated_call():
        estimator.setLabelCol('ground_truth')
    prediction = Pipeline(stages=[estimator]).fit(df).transform(test_df)
    assert prediction.select("ground_truth", "output").count() == len(x_test)


def test_set_cols(spark_context, regression_model, boston_housing_dataset):
    x_train, y_train, x_test, y_test = boston_housing_dataset
    df = to_data_frame(spark_context, x_train, y_train)
    df = df.withColumnRenamed('features', 'scaled_features').withColumnRenamed('label', 'ground_truth')
    test_df = to_data_frame(spark_context, x_test, y_test)
    test_df = test_df.withColumnRenamed('features', 'scaled_features').withColumnRenamed('label', 'ground_truth')
    sgd_conf = optimizers.serialize(optimizers.SGD(learning_rate=0.00001))
    estimator = _estimator(regression_model, sgd_conf, loss="mae", metrics=['mae'], validation_split=0.01,
                           categorical=False, labelCol='ground_truth', outputCol='output',
                           featuresCol='scaled_features')
    prediction = Pipeline(stages=[estimator]).fit(df).transform(test_df)
    assert prediction.columns == ['scaled_features', 'ground_truth', 'output']


def test_custom_objects(spark_context, boston_housing_dataset):
    def custom_activation(x):
        return 2 * activations.relu(x)

    model = Sequential()
    model.add(Dense(64, input_shape=(13,)))
    model.add(Dense(64, activation=custom_activation))
    model.add(Dense(1, activation='linear'))
    x_train, y_train, x_test, y_test = boston_housing_dataset
    df = to_data_frame(spark_context, x_train, y_train)
    test_df = to_data_frame(spark_context, x_test, y_test)
    sgd_conf = optimizers.serialize(optimizers.SGD(learning_rate=0.00001))
    estimator = _estimator(model, sgd_conf, loss="mae", metrics=['mae'], batch_size=32, validation_split=0.01,
                           categorical=False)
    estimator.set_custom_objects({'custom_activation': custom_activation})
    prediction = Pipeline(stages=[estimator]).fit(df).transform(test_df)
    assert prediction.count() == len(x_test)


def test_predict_classes_probability(spark_context, classification_model, mnist_data):
    x_train, y_train, x_test, y_test = mnist_data
    df = to_data_frame(spark_context, x_train, y_train, categorical=True)
    test_df = to_data_frame(spark_context, x_test, y_test, categorical=True)
    sgd = optimizers.SGD(learning_rate=0.01, decay=1e-6, momentum=0.9, nesterov=True)
    estimator = _estimator(classification_model, optimizers.serialize(sgd))
    results = Pipeline(stages=[estimator]).fit(df).transform(test_df)
    row = results.take(1)[0]
    assert len(row.prediction) == 10 and abs(sum(row.prediction) - 1.0) < 1e-4


def test_batch_predict_classes_probability(spark_context, classification_model, mnist_data):
    x_train, y_train, x_test, y_test = mnist_data
    df = to_data_frame(spark_context, x_train, y_train, categorical=True)
    test_df = to_data_frame(spark_context, x_test, y_test, categorical=True)
    sgd = optimizers.SGD(learning_rate=0.01, decay=1e-6, momentum=0.9, nesterov=True)
    fitted = _estimator(classification_model, optimizers.serialize(sgd)).fit(df)
    results = fitted.transform(test_df)
    fitted.set_params(inference_batch_size=int(len(y_test) / 10))
    fitted.set_params(outputCol="prediction_via_batch_inference")
    both = fitted.transform(results)
    for row in both.take(25):
        assert len(row.prediction) == 10 and len(row.prediction_via_batch_inference) == 10
        assert np.array_equal(row.prediction, row.prediction_via_batch_inference)  # bit-identical


def test_save_pipeline(tmp_path, classification_model):
    sgd = optimizers.SGD(learning_rate=0.01, decay=1e-6, momentum=0.9, nesterov=True)
    estimator = _estimator(classification_model, optimizers.serialize(sgd), epochs=10, batch_size=10)
    pipeline = Pipeline(stages=[estimator])
    pipeline.save(str(tmp_path / 'tmp'))
    loaded = Pipeline.load(str(tmp_path / 'tmp'))
    stage = loaded.getStages()[0]
    assert isinstance(stage, ElephasEstimator)
    assert stage.get_keras_model_config() == classification_model.to_json()
    assert stage.get_epochs() == 10 and stage.get_optimizer_config()['class_name'] == 'SGD'


def test_otto_shaped_pipeline(spark_context, spark_session):
    """examples/ml_pipeline_otto.py: StringIndexer -> StandardScaler -> ElephasEstimator."""
    from elephas_b200.models import zoo
    from elephas_b200.spark import StandardScaler, StringIndexer, Vectors
    rng = np.random.default_rng(0)
    centers = rng.normal(size=(9, 93)) * 3
    labels = rng.integers(0, 9, 600)
    feats = centers[labels] + rng.normal(size=(600, 93))
    rows = [(Vectors.dense(f), 'Class_%d' % (l + 1)) for f, l in zip(feats, labels)]
    df = spark_session.createDataFrame(rows, ['features', 'category'])
    indexer = StringIndexer(inputCol="category", outputCol="index_category")
    scaler = StandardScaler(inputCol="features", outputCol="scaled_features", withStd=True, withMean=True)
    model = zoo.otto_mlp(hidden=64)
    est = _estimator(model, optimizers.serialize(optimizers.Adam(learning_rate=0.01)), epochs=5, batch_size=64,
                     validation_split=0.15, nb_classes=9)
    est.set_mode('hogwild')
    with pytest.deprecated_call():
        est.setFeaturesCol("scaled_features")
    with pytest.deprecated_call():
        est.setLabelCol("index_category")
    fitted = Pipeline(stages=[indexer, scaler, est]).fit(df)
    pred = fitted.transform(df)
    pnl = pred.select("index_category", functions.array_argmax("prediction").astype(DoubleType()).alias("p"))
    acc = MulticlassMetrics(pnl.rdd.map(lambda r: (r.p, r.index_category))).accuracy
    assert acc > 0.8
