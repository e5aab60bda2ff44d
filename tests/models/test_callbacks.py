"""Callbacks of ``Model.fit`` (``elephas_b200/models/callbacks.py``): protocol, early stopping, checkpoints, learning-rate
schedules. The reference hands ``callbacks=`` through to ``keras.Model.fit`` (``elephas/worker.py:39-42``)."""
import math

import numpy as np
import pytest

from elephas_b200.models import Dense, Sequential, load_model
from elephas_b200.models import callbacks as C
from elephas_b200.models.optimizers import SGD


def _data(n=256, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, 8)).astype(np.float32)
    w = rng.standard_normal((8, 3)).astype(np.float32)
    y = np.eye(3, dtype=np.float32)[(x @ w).argmax(1)]
    return x, y


def _model(lr=0.1):
    m = Sequential([Dense(16, activation='relu', input_shape=(8,)), Dense(3, activation='softmax')])
    m.compile(SGD(learning_rate=lr), 'categorical_crossentropy', ['acc'])
    return m


def test_hooks_are_called_in_order_and_see_the_logs():
    x, y = _data()
    calls = []

    class Spy(C.Callback):
        def on_train_begin(self, logs=None):
            calls.append(('train_begin', self.params['epochs'], self.model is not None))

        def on_epoch_begin(self, epoch, logs=None):
            calls.append(('begin', epoch))

        def on_epoch_end(self, epoch, logs=None):
            calls.append(('end', epoch, sorted(logs)))

        def on_train_end(self, logs=None):
            calls.append(('train_end',))

    h = _model().fit(x, y, batch_size=32, epochs=2, validation_split=0.25, callbacks=[Spy()])
    assert calls == [('train_begin', 2, True), ('begin', 0), ('end', 0, ['acc', 'loss', 'val_acc', 'val_loss']),
                     ('begin', 1), ('end', 1, ['acc', 'loss', 'val_acc', 'val_loss']), ('train_end',)]
    assert len(h.history['loss']) == 2


def test_plain_object_with_only_on_epoch_end_still_works():
    x, y = _data()

    class Bare:
        seen = []

        def on_epoch_end(self, epoch, logs):
            self.seen.append(epoch)

    _model().fit(x, y, epochs=3, callbacks=[Bare()])
    assert Bare.seen == [0, 1, 2]


def test_early_stopping_stops_and_restores_the_best_weights():
    x, y = _data()
    m = _model()
    losses = iter([1.0, 0.5, 0.6, 0.7, 0.8, 0.9])
    snapshots = []

    class Script(C.Callback):  # runs before EarlyStopping: overrides the monitored value with a scripted curve
        def on_epoch_end(self, epoch, logs=None):
            logs['val_loss'] = next(losses)
            snapshots.append([w.copy() for w in self.model.get_weights()])

    es = C.EarlyStopping(monitor='val_loss', patience=2, restore_best_weights=True)
    h = m.fit(x, y, epochs=6, callbacks=[Script(), es])
    assert es.stopped_epoch == 3 and len(h.history['loss']) == 4 and m.stop_training
    for a, b in zip(m.get_weights(), snapshots[1]):
        np.testing.assert_array_equal(a, b)
    # a new fit starts with the flag cleared
    assert len(m.fit(x, y, epochs=1).history['loss']) == 1


def test_early_stopping_patience_zero_and_max_mode():
    es = C.EarlyStopping(monitor='acc', patience=0)
    assert es.mode == 'max'

    class M:
        stop_training = False
    es.set_model(M())
    es.on_train_begin()
    es.on_epoch_end(0, {'acc': 0.5})
    assert not es.model.stop_training
    es.on_epoch_end(1, {'acc': 0.4})
    assert es.model.stop_training
    with pytest.raises(ValueError):
        C.EarlyStopping(mode='sideways')


def test_model_checkpoint_writes_loadable_files(tmp_path):
    x, y = _data()
    m = _model()
    pattern = str(tmp_path / 'ck-{epoch:02d}.h5')
    m.fit(x, y, epochs=2, callbacks=[C.ModelCheckpoint(pattern)])
    back = load_model(str(tmp_path / 'ck-02.h5'))
    np.testing.assert_allclose(back.predict(x[:16]), m.predict(x[:16]), rtol=1e-5, atol=1e-6)

    best = str(tmp_path / 'best.h5')
    values = iter([0.9, 0.5, 0.7])
    written = []

    class Script(C.Callback):
        def on_epoch_end(self, epoch, logs=None):
            logs['val_loss'] = next(values)

    class Watch(C.ModelCheckpoint):
        def on_epoch_end(self, epoch, logs=None):
            before = self.best
            super().on_epoch_end(epoch, logs)
            if self.best != before:
                written.append(epoch)

    m.fit(x, y, epochs=3, callbacks=[Script(), Watch(best, save_best_only=True, save_weights_only=True)])
    assert written == [0, 1]
    with pytest.raises(ValueError):
        C.ModelCheckpoint(best, save_freq=5)


def test_learning_rate_scheduler_changes_the_steps_taken():
    x, y = _data()
    m = _model(lr=0.1)
    h = m.fit(x, y, epochs=3, shuffle=False, callbacks=[C.LearningRateScheduler(lambda epoch, lr: lr * 0.5)])
    assert h.history['lr'] == pytest.approx([0.05, 0.025, 0.0125])
    assert m.optimizer.get_config()['learning_rate'] == pytest.approx(0.0125)
    # a zero learning rate really reaches the update: the weights do not move
    frozen = _model()
    before = [w.copy() for w in frozen.get_weights()]
    frozen.fit(x, y, epochs=1, callbacks=[C.LearningRateScheduler(lambda epoch: 0.0)])
    for a, b in zip(frozen.get_weights(), before):
        np.testing.assert_array_equal(a, b)


def test_reduce_lr_on_plateau_and_terminate_on_nan():
    x, y = _data()
    m = _model(lr=0.2)
    values = iter([1.0, 1.0, 1.0, 1.0, 1.0])

    class Script(C.Callback):
        def on_epoch_end(self, epoch, logs=None):
            logs['val_loss'] = next(values)

    m.fit(x, y, epochs=5, callbacks=[Script(), C.ReduceLROnPlateau(patience=2, factor=0.5, min_lr=0.06)])
    assert m.optimizer.learning_rate == pytest.approx(0.06)   # 0.2 -> 0.1 -> floor at 0.06
    with pytest.raises(ValueError):
        C.ReduceLROnPlateau(factor=1.0)

    class Poison(C.Callback):
        def on_epoch_end(self, epoch, logs=None):
            logs['loss'] = math.nan

    h = _model().fit(x, y, epochs=4, callbacks=[Poison(), C.TerminateOnNaN()])
    assert len(h.history['loss']) == 1


def test_lambda_callback_and_optimizer_lr_alias():
    x, y = _data()
    seen = []
    _model().fit(x, y, epochs=2, callbacks=[C.LambdaCallback(on_epoch_end=lambda e, logs: seen.append(e))])
    assert seen == [0, 1]
    opt = SGD(lr=0.3)
    assert opt.lr == opt.learning_rate == 0.3
    v = opt._hyper_version
    opt.lr = 0.3
    assert opt._hyper_version == v
    opt.lr = 0.1
    assert opt._hyper_version == v + 1 and opt.get_config()['learning_rate'] == 0.1


@pytest.mark.parametrize('mode,ps', [('synchronous', None), ('asynchronous', 'http'), ('hogwild', 'nvlink')])
def test_callbacks_travel_through_spark_model_fit(mode, ps):
    """``SparkModel.fit(rdd, callbacks=[...])``: the keyword reaches every worker's ``Model.fit`` in all three modes."""
    from elephas_b200.spark_model import SparkModel
    from elephas_b200.spark import SparkContext
    from elephas_b200.utils.rdd_utils import to_simple_rdd
    x, y = _data(512)
    sc = SparkContext.getOrCreate()
    rdd = to_simple_rdd(sc, x, y).repartition(2)
    seen = []
    kwargs = {'parameter_server_mode': ps} if ps else {}
    sm = SparkModel(_model(), mode=mode, frequency='epoch', num_workers=2, **kwargs)
    sm.fit(rdd, epochs=2, batch_size=32, verbose=0,
           callbacks=[C.LambdaCallback(on_epoch_end=lambda e, logs: seen.append((e, logs['loss'])))])
    assert sorted(e for e, _ in seen) == [0, 0, 1, 1]
    assert all(np.isfinite(v) for _, v in seen)


def test_csv_logger_writes_one_line_per_epoch(tmp_path):
    x, y = _data()
    path = tmp_path / 'log.csv'
    m = _model()
    m.fit(x, y, epochs=2, validation_split=0.25, callbacks=[C.CSVLogger(path)])
    lines = path.read_text().strip().split('\n')
    assert lines[0] == 'epoch,acc,loss,val_acc,val_loss' and [l.split(',')[0] for l in lines[1:]] == ['0', '1']
    assert all(len(l.split(',')) == 5 for l in lines[1:])
    m.fit(x, y, epochs=3, initial_epoch=2, validation_split=0.25, callbacks=[C.CSVLogger(path, append=True)])
    lines = path.read_text().strip().split('\n')
    assert len(lines) == 4 and lines[-1].split(',')[0] == '2'


def test_early_stopping_keeps_its_patience_across_the_epochs_of_a_synchronous_fit():
    """With callbacks every worker runs one fit over all epochs (the reference's scheme): patience accumulates."""
    from elephas_b200.spark_model import SparkModel
    from elephas_b200.spark import SparkContext
    from elephas_b200.utils.rdd_utils import to_simple_rdd
    x, y = _data(512)
    rdd = to_simple_rdd(SparkContext.getOrCreate(), x, y).repartition(2)
    seen = []

    class Flat(C.Callback):   # the monitored value never improves after the first epoch
        def on_epoch_end(self, epoch, logs=None):
            logs['val_loss'] = 1.0
            seen.append(epoch)

    sm = SparkModel(_model(), mode='synchronous', num_workers=2)
    sm.fit(rdd, epochs=10, batch_size=32, verbose=0, callbacks=[Flat(), C.EarlyStopping(patience=2)])
    assert sorted(seen) == [0, 0, 1, 1, 2, 2]     # both workers stop after epoch 2 (epochs 1 and 2 did not improve)
