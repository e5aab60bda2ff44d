"""Spark ML ``Params`` mixins of the Elephas estimator / transformer.

Reference: ``elephas/ml/params.py:4-259`` - fifteen hand-written mixins, each declaring one
``Param`` with a ``set_<x>`` / ``get_<x>`` pair and (mostly) a default. The mixins are generated
here from one declarative table, which keeps names, docs, accessors and defaults in one place:

=======================  ======================  ====================  =================
mixin                    param                   accessors             default
=======================  ======================  ====================  =================
HasKerasModelConfig      keras_model_config      *_keras_model_config  (none)
HasMode                  mode                    *_mode                'asynchronous'
HasFrequency             frequency               *_frequency           'epoch'
HasNumberOfClasses       nb_classes              *_nb_classes          10
HasCategoricalLabels     categorical             *_categorical_labels  True
HasEpochs                epochs                  *_epochs              10
HasBatchSize             batch_size              *_batch_size          32
HasVerbosity             verbose                 *_verbosity           0
HasValidationSplit       validation_split        *_validation_split    0.1
HasNumberOfWorkers       num_workers             *_num_workers         8
HasKerasOptimizerConfig  optimizer_config        *_optimizer_config    None
HasMetrics               metrics                 *_metrics             ['acc']
HasLoss                  loss                    *_loss                (none)
HasCustomObjects         custom_objects          *_custom_objects      {}
HasInferenceBatchSize    inference_batch_size    *_inference_batch_size None
=======================  ======================  ====================  =================
"""
from ..spark import Param, Params

_UNSET = object()


def _make_mixin(class_name, param_name, doc, accessor=None, default=_UNSET, arg_name=None):
    accessor = accessor or param_name
    arg_name = arg_name or param_name  # keyword the reference's setter takes (set_mode(mode=...), params.py:33)

    def __init__(self):
        super(cls, self).__init__()
        setattr(self, param_name, Param(self, param_name, doc))
        if default is not _UNSET:
            # mutable defaults are copied so instances never share them
            value = default.copy() if isinstance(default, (list, dict)) else default
            self._setDefault(**{param_name: value})

    def _set(self, value):
        self._paramMap[getattr(self, param_name)] = value
        return self

    # the setter's one parameter carries the reference's name, so keyword calls written against Elephas keep working
    scope = {'_set': _set}
    exec("def setter(self, %s):\n    return _set(self, %s)\n" % (arg_name, arg_name), scope)
    setter = scope['setter']

    def getter(self):
        return self.getOrDefault(getattr(self, param_name))

    setter.__name__ = 'set_' + accessor
    getter.__name__ = 'get_' + accessor
    setter.__doc__ = "Set `%s` (%s); returns self." % (param_name, doc)
    getter.__doc__ = "Current value of `%s` (its default when unset)." % param_name
    cls = type(class_name, (Params,), {'__init__': __init__, setter.__name__: setter, getter.__name__: getter,
                                       '__doc__': "Parameter mixin: %s" % doc, '__module__': __name__})
    return cls


HasKerasModelConfig = _make_mixin('HasKerasModelConfig', 'keras_model_config', "Serialized Keras model as json string")
HasMode = _make_mixin('HasMode', 'mode', "Elephas mode", default='asynchronous')
HasFrequency = _make_mixin('HasFrequency', 'frequency', "Elephas frequency", default='epoch')
HasNumberOfClasses = _make_mixin('HasNumberOfClasses', 'nb_classes', "number of classes", default=10)
HasCategoricalLabels = _make_mixin('HasCategoricalLabels', 'categorical',
                                   "Boolean to indicate if labels are categorical", 'categorical_labels', True)
HasEpochs = _make_mixin('HasEpochs', 'epochs', "Number of epochs to train", default=10)
HasBatchSize = _make_mixin('HasBatchSize', 'batch_size', "Batch size", default=32)
HasVerbosity = _make_mixin('HasVerbosity', 'verbose', "Stdout verbosity", 'verbosity', 0)
HasValidationSplit = _make_mixin('HasValidationSplit', 'validation_split', "validation split percentage",
                                 default=0.1)
HasNumberOfWorkers = _make_mixin('HasNumberOfWorkers', 'num_workers', "number of workers", default=8)
HasKerasOptimizerConfig = _make_mixin('HasKerasOptimizerConfig', 'optimizer_config',
                                      "Serialized Keras optimizer properties", default=None)
HasMetrics = _make_mixin('HasMetrics', 'metrics', "Keras metrics", default=['acc'])
HasLoss = _make_mixin('HasLoss', 'loss', "Keras loss")
HasCustomObjects = _make_mixin('HasCustomObjects', 'custom_objects',
                               "Custom objects (i.e; custom activation functions, custom layers) in a model",
                               default={})
HasInferenceBatchSize = _make_mixin('HasInferenceBatchSize', 'inference_batch_size',
                                    "Batch inference could help limit memory consumption in executors",
                                    default=None, arg_name='batch_size')

__all__ = ['HasKerasModelConfig', 'HasMode', 'HasFrequency', 'HasNumberOfClasses', 'HasCategoricalLabels',
           'HasEpochs', 'HasBatchSize', 'HasVerbosity', 'HasValidationSplit', 'HasNumberOfWorkers',
           'HasKerasOptimizerConfig', 'HasMetrics', 'HasLoss', 'HasCustomObjects', 'HasInferenceBatchSize']
