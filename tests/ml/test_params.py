from elephas.ml.params import *  # noqa: F401,F403


def test_has_keras_model_config():
    p = HasKerasModelConfig()
    p.set_keras_model_config("{}")
    assert p.get_keras_model_config() == "{}"


def test_defaults_and_setters():
    cases = [(HasMode, 'mode', 'asynchronous', 'hogwild'), (HasFrequency, 'frequency', 'epoch', 'batch'),
             (HasNumberOfClasses, 'nb_classes', 10, 3), (HasCategoricalLabels, 'categorical_labels', True, False),
             (HasEpochs, 'epochs', 10, 2), (HasBatchSize, 'batch_size', 32, 64), (HasVerbosity, 'verbosity', 0, 2),
             (HasValidationSplit, 'validation_split', 0.1, 0.3), (HasNumberOfWorkers, 'num_workers', 8, 2),
             (HasKerasOptimizerConfig, 'optimizer_config', None, {'class_name': 'SGD', 'config': {}}),
             (HasMetrics, 'metrics', ['acc'], ['mae']), (HasCustomObjects, 'custom_objects', {}, {'f': len}),
             (HasInferenceBatchSize, 'inference_batch_size', None, 128)]
    for cls, accessor, default, new in cases:
        p = cls()
        assert getattr(p, 'get_' + accessor)() == default, cls.__name__
        assert getattr(p, 'set_' + accessor)(new) is p
        assert getattr(p, 'get_' + accessor)() == new


def test_loss_has_no_default():
    p = HasLoss()
    p.set_loss('mse')
    assert p.get_loss() == 'mse'


def test_mutable_defaults_are_not_shared():
    a, b = HasMetrics(), HasMetrics()
    a.get_metrics().append('x')
    assert b.get_metrics() == ['acc']


def test_setters_accept_the_reference_keyword_names():
    """Calls written against Elephas with keywords keep working (``set_mode(mode=...)``, ``elephas/ml/params.py``)."""
    from elephas_b200.ml import params as P

    class Holder(P.HasMode, P.HasVerbosity, P.HasCategoricalLabels, P.HasInferenceBatchSize, P.HasNumberOfClasses):
        pass

    h = Holder()
    assert h.set_mode(mode='hogwild') is h and h.get_mode() == 'hogwild'
    h.set_verbosity(verbose=2)
    h.set_categorical_labels(categorical=False)
    h.set_inference_batch_size(batch_size=64)
    h.set_nb_classes(nb_classes=3)
    assert (h.get_verbosity(), h.get_categorical_labels(), h.get_inference_batch_size(), h.get_nb_classes()) == \
        (2, False, 64, 3)
    h.set_mode('synchronous')
    assert h.get_mode() == 'synchronous'
