"""Spark ML pipeline stages: ``ElephasEstimator`` / ``ElephasTransformer`` and their loaders.

Reference: ``elephas/ml_model.py:25-269``. Same params (the fifteen mixins of ``ml/params.py`` plus
the Spark column mixins), same ``@keyword_only`` constructors, ``get_config`` / ``save`` /
``load_ml_estimator`` / ``load_ml_transformer``, deprecated ``setFeaturesCol`` / ``setLabelCol`` /
``setOutputCol`` (``:114-127``). ``_fit`` builds and compiles the model from its JSON config, trains
a :class:`~elephas_b200.spark_model.SparkModel` and returns an ``ElephasTransformer``
(``:81-112``); ``_transform`` runs distributed batch prediction and appends the output column -
``DoubleType`` for regression, ``Array<Double>`` of class probabilities for classification
(``:243-249``).

B200-side differences: prediction runs through the sharded device predict path of ``SparkModel``
instead of per-partition Keras replicas; ``inference_batch_size`` only bounds host staging - results
are bit-identical with or without it (``tests/test_ml_model.py:312-354``) because every kernel on
the predict path is row-independent. ``get_config`` additionally carries ``optimizer_config``
(SURVEY.md §7.4 D7).
"""
import copy
import json
import warnings

import numpy as np

from .io import hdf5
from .ml.adapter import df_to_simple_rdd
from .ml.params import *  # noqa: F401,F403
from .ml.params import (HasBatchSize, HasCategoricalLabels, HasCustomObjects, HasEpochs, HasFrequency,
                        HasInferenceBatchSize, HasKerasModelConfig, HasKerasOptimizerConfig, HasLoss, HasMetrics,
                        HasMode, HasNumberOfClasses, HasNumberOfWorkers, HasValidationSplit, HasVerbosity)
from .mllib.adapter import from_vector
from .models import model_from_json
from .models import optimizers as _optimizers
from .spark import (ArrayType, DefaultParamsReadable, DefaultParamsWritable, DoubleType, Estimator, HasFeaturesCol,
                    HasLabelCol, HasOutputCol, Model, SparkSession, StructField, keyword_only)
from .spark_model import SparkModel
from .utils.model_utils import LossModelTypeMapper, ModelType, ModelTypeEncoder, as_enum


def _write_config_file(file_name, payload):
    """An HDF5 file whose only content is the ``distributed_config`` root attribute
    (``ml_model.py:61-70,172-186``)."""
    root = hdf5.Group()
    root.attrs['distributed_config'] = payload.encode('utf8')
    hdf5.save_tree(file_name, root)


def _read_config_file(file_name, object_hook=None):
    raw = hdf5.load_tree(file_name).attrs['distributed_config']
    return json.loads(raw.decode('utf8') if isinstance(raw, bytes) else raw, object_hook=object_hook)


class ElephasEstimator(Estimator, HasCategoricalLabels, HasValidationSplit, HasKerasModelConfig, HasFeaturesCol,
                       HasLabelCol, HasMode, HasEpochs, HasBatchSize, HasFrequency, HasVerbosity,
                       HasNumberOfClasses, HasNumberOfWorkers, HasOutputCol, HasLoss, HasMetrics,
                       HasKerasOptimizerConfig, HasCustomObjects, DefaultParamsReadable, DefaultParamsWritable):
    """Spark ML ``Estimator``: carries everything needed to compile and train the model; ``fit(df)``
    returns an :class:`ElephasTransformer`."""

    @keyword_only
    def __init__(self, **kwargs):
        super(ElephasEstimator, self).__init__()
        self._defaultParamMap[self.outputCol] = 'prediction'
        self.set_params(**kwargs)

    def get_config(self):
        return {'keras_model_config': self.get_keras_model_config(),
                'mode': self.get_mode(),
                'frequency': self.get_frequency(),
                'num_workers': self.get_num_workers(),
                'categorical': self.get_categorical_labels(),
                'loss': self.get_loss() if self.isDefined(self.loss) else None,
                'metrics': self.get_metrics(),
                'validation_split': self.get_validation_split(),
                'featuresCol': self.getFeaturesCol(),
                'labelCol': self.getLabelCol(),
                'epochs': self.get_epochs(),
                'batch_size': self.get_batch_size(),
                'verbose': self.get_verbosity(),
                'nb_classes': self.get_nb_classes(),
                'outputCol': self.getOutputCol(),
                'optimizer_config': self.get_optimizer_config()}

    def save(self, file_name: str):
        _write_config_file(file_name, json.dumps({'class_name': self.__class__.__name__,
                                                  'config': self.get_config()}))

    @keyword_only
    def set_params(self, **kwargs):
        """Set the given params; everything else keeps its default."""
        return self._set(**kwargs)

    def get_model(self):
        return model_from_json(self.get_keras_model_config(), self.get_custom_objects())

    def _fit(self, df):
        simple_rdd = df_to_simple_rdd(df, categorical=self.get_categorical_labels(),
                                      nb_classes=self.get_nb_classes(), features_col=self.getFeaturesCol(),
                                      label_col=self.getLabelCol())
        simple_rdd = simple_rdd.repartition(self.get_num_workers())
        keras_model = model_from_json(self.get_keras_model_config(), self.get_custom_objects())
        keras_model.custom_objects = dict(self.get_custom_objects() or {})
        loss = self.get_loss()
        optimizer_config = self.get_optimizer_config()
        optimizer = _optimizers.get(copy.deepcopy(optimizer_config)) if optimizer_config is not None \
            else _optimizers.get('rmsprop')
        keras_model.compile(loss=loss, optimizer=optimizer, metrics=self.get_metrics())

        spark_model = SparkModel(model=keras_model, mode=self.get_mode(), frequency=self.get_frequency(),
                                 num_workers=self.get_num_workers(), custom_objects=self.get_custom_objects())
        spark_model.fit(simple_rdd, epochs=self.get_epochs(), batch_size=self.get_batch_size(),
                        verbose=self.get_verbosity(), validation_split=self.get_validation_split())

        return ElephasTransformer(labelCol=self.getLabelCol(), outputCol=self.getOutputCol(),
                                  featuresCol=self.getFeaturesCol(),
                                  keras_model_config=spark_model.master_network.to_json(),
                                  weights=spark_model.master_network.get_weights(),
                                  custom_objects=self.get_custom_objects(),
                                  model_type=LossModelTypeMapper().get_model_type(loss),
                                  history=spark_model.training_histories)

    def _deprecated_col(self, which, value):
        warnings.warn("set%s is deprecated in Spark 3.0.x+ - please supply %s in the constructor i.e; "
                      "ElephasEstimator(%s='foo')" % (which[0].upper() + which[1:], which, which),
                      DeprecationWarning, stacklevel=3)
        return self._set(**{which: value})

    def setFeaturesCol(self, value):
        return self._deprecated_col('featuresCol', value)

    def setLabelCol(self, value):
        return self._deprecated_col('labelCol', value)

    def setOutputCol(self, value):
        return self._deprecated_col('outputCol', value)


def load_ml_estimator(file_name: str) -> ElephasEstimator:
    config = _read_config_file(file_name).get('config')
    return ElephasEstimator(**{k: v for k, v in config.items() if v is not None})


class ElephasTransformer(Model, HasKerasModelConfig, HasLabelCol, HasOutputCol, HasFeaturesCol, HasCustomObjects,
                         HasInferenceBatchSize, DefaultParamsReadable, DefaultParamsWritable):
    """Spark ML ``Model`` holding trained weights; ``transform(df)`` appends predictions."""

    @keyword_only
    def __init__(self, **kwargs):
        super(ElephasTransformer, self).__init__()
        if "weights" in kwargs:
            self.weights = kwargs.pop('weights')
        if "model_type" in kwargs:
            self.model_type = kwargs.pop('model_type')
        self._history = kwargs.pop("history", [])
        self._predict_model = None
        self.set_params(**kwargs)

    @property
    def history(self):
        return self._history

    @keyword_only
    def set_params(self, **kwargs):
        """Set the given params; everything else keeps its default."""
        return self._set(**kwargs)

    def get_config(self):
        return {'keras_model_config': self.get_keras_model_config(),
                'labelCol': self.getLabelCol(),
                'featuresCol': self.getFeaturesCol(),
                'outputCol': self.getOutputCol(),
                'weights': [np.asarray(weight).tolist() for weight in getattr(self, 'weights', [])],
                'model_type': getattr(self, 'model_type', None)}

    def save(self, file_name: str):
        """Write config, weights (as nested JSON lists) and model type to an HDF5 file."""
        _write_config_file(file_name, json.dumps({'class_name': self.__class__.__name__,
                                                  'config': self.get_config()}, cls=ModelTypeEncoder))

    def get_model(self):
        return model_from_json(self.get_keras_model_config(), self.get_custom_objects())

    def _model_for_predict(self):
        from .parallel.runtime import get_runtime
        if self._predict_model is None:
            model = self.get_model()
            if getattr(self, 'weights', None):
                model.set_weights(self.weights)
            model.to(get_runtime().device)
            self._predict_model = model
        return self._predict_model

    def _transform(self, df):
        output_col = self.getOutputCol()
        features_col = self.getFeaturesCol()
        new_schema = copy.deepcopy(df.schema)
        model = self._model_for_predict()
        inference_batch_size = self.get_inference_batch_size()

        def predict_partition(rows):
            rows = list(rows)
            if not rows:
                return []
            feats = np.asarray([from_vector(r[features_col]) for r in rows], dtype=np.float32)
            if inference_batch_size is not None and inference_batch_size > 0:
                chunks = [model.predict(feats[i:i + inference_batch_size])
                          for i in range(0, len(feats), inference_batch_size)]
                return list(np.vstack(chunks))
            return list(model.predict(feats))

        regression = getattr(self, 'model_type', None) == ModelType.REGRESSION

        if hasattr(df, 'column_block') and output_col not in df.columns:
            # Local DataFrames: columns travel between pipeline stages as dense blocks. The features of ALL rows are one
            # matrix (free when the previous stage - e.g. StandardScaler - produced the column as a block), the model
            # predicts on it in one go (or in inference_batch_size pieces: bit-identical, tests/test_ml_model.py), and
            # the predictions are appended as a block again; Row objects are only built if a later consumer asks for
            # rows. The reference converts every row to a NumPy array and back (elephas/ml_model.py:222-250).
            feats = np.ascontiguousarray(df.column_block(features_col), dtype=np.float32)
            if feats.ndim == 2 and feats.shape[0] > 0:
                if inference_batch_size is not None and inference_batch_size > 0:
                    preds = np.vstack([model.predict(feats[i:i + inference_batch_size])
                                       for i in range(0, len(feats), inference_batch_size)])
                else:
                    preds = model.predict(feats)
                block = np.asarray(preds, dtype=np.float64).reshape(len(feats), -1)
                return df.with_block_column(output_col, block[:, 0] if regression else block,
                                            DoubleType() if regression else ArrayType(DoubleType()))

        def predict_rows(rows):
            # one predict per partition, then the output column is attached row by row: ndarray.tolist() converts the
            # whole block to Python floats at C speed (reference: one np.asarray + float() per row, ml_model.py:222-227)
            rows = list(rows)
            preds = predict_partition(rows)
            if not rows:
                return []
            block = np.asarray(preds, dtype=np.float64).reshape(len(rows), -1)
            values = block[:, 0].tolist() if regression else block.tolist()
            return [tuple(r) + (v,) for r, v in zip(rows, values)]

        results_rdd = df.rdd.mapPartitions(predict_rows)  # (rows are only materialised on this path)
        output_field = StructField(output_col, DoubleType() if regression else ArrayType(DoubleType()), True)
        new_schema.add(output_field)
        return SparkSession.builder.getOrCreate().createDataFrame(results_rdd, new_schema)


def load_ml_transformer(file_name: str) -> ElephasTransformer:
    config = _read_config_file(file_name, object_hook=as_enum).get('config')
    config['weights'] = [np.array(weight) for weight in config['weights']]
    return ElephasTransformer(**{k: v for k, v in config.items() if v is not None})
