"""Keras-style callbacks for ``Model.fit``.

The reference forwards whatever keyword arguments ``SparkModel.fit`` receives to ``keras.Model.fit`` on the workers
(``elephas/spark_model.py:169-189`` -> ``elephas/worker.py:39-42``), callbacks included. Same protocol here:
``on_train_begin / on_epoch_begin / on_epoch_end / on_train_end`` with a ``logs`` dict per epoch (``loss``, metrics,
``val_*``), ``model.stop_training`` to end a fit early, ``model.optimizer.learning_rate`` readable and settable.
Batch-level hooks are not called: a training step is one CUDA-graph replay, the host does not see its loss.
"""
import math

import numpy as np


class Callback:
    def __init__(self):
        self.model = None
        self.params = {}

    def set_model(self, model):
        self.model = model

    def set_params(self, params):
        self.params = dict(params)

    def on_train_begin(self, logs=None):
        pass

    def on_train_end(self, logs=None):
        pass

    def on_epoch_begin(self, epoch, logs=None):
        pass

    def on_epoch_end(self, epoch, logs=None):
        pass


class LambdaCallback(Callback):
    def __init__(self, on_epoch_begin=None, on_epoch_end=None, on_train_begin=None, on_train_end=None, **_ignored):
        super().__init__()
        for name, fn in (('on_epoch_begin', on_epoch_begin), ('on_epoch_end', on_epoch_end),
                         ('on_train_begin', on_train_begin), ('on_train_end', on_train_end)):
            if fn is not None:
                setattr(self, name, fn)


def _monitor_op(mode, monitor):
    if mode == 'auto':
        mode = 'max' if ('acc' in monitor or monitor.startswith('fmeasure')) else 'min'
    if mode not in ('min', 'max'):
        raise ValueError("mode must be 'auto', 'min' or 'max'")
    return (lambda a, b: a < b) if mode == 'min' else (lambda a, b: a > b), mode


class EarlyStopping(Callback):
    """Stops when ``monitor`` has not improved by more than ``min_delta`` for ``patience`` epochs."""

    def __init__(self, monitor='val_loss', min_delta=0.0, patience=0, verbose=0, mode='auto', baseline=None,
                 restore_best_weights=False, start_from_epoch=0):
        super().__init__()
        self.monitor, self.min_delta, self.patience, self.verbose = monitor, abs(float(min_delta)), int(patience), verbose
        self.baseline, self.restore_best_weights, self.start_from_epoch = baseline, restore_best_weights, start_from_epoch
        self._better, self.mode = _monitor_op(mode, monitor)
        self.stopped_epoch = 0

    def on_train_begin(self, logs=None):
        self.wait, self.stopped_epoch, self.best_weights = 0, 0, None
        self.best = self.baseline if self.baseline is not None else (math.inf if self.mode == 'min' else -math.inf)

    def on_epoch_end(self, epoch, logs=None):
        current = (logs or {}).get(self.monitor)
        if current is None or epoch < self.start_from_epoch:
            return
        threshold = self.best - self.min_delta if self.mode == 'min' else self.best + self.min_delta
        if self._better(current, threshold):
            self.best, self.wait = current, 0
            if self.restore_best_weights:
                self.best_weights = self.model.get_weights()
            return
        self.wait += 1
        if self.wait >= self.patience and epoch > 0:
            self.stopped_epoch = epoch
            self.model.stop_training = True
            if self.restore_best_weights and self.best_weights is not None:
                self.model.set_weights(self.best_weights)
            if self.verbose:
                print("Epoch %d: early stopping" % (epoch + 1))


class ModelCheckpoint(Callback):
    """Saves the model (Keras-layout HDF5) after every epoch, or only when ``monitor`` improves."""

    def __init__(self, filepath, monitor='val_loss', verbose=0, save_best_only=False, save_weights_only=False,
                 mode='auto', save_freq='epoch'):
        super().__init__()
        if save_freq != 'epoch':
            raise ValueError("ModelCheckpoint: only save_freq='epoch' is supported")
        self.filepath, self.monitor, self.verbose = filepath, monitor, verbose
        self.save_best_only, self.save_weights_only = save_best_only, save_weights_only
        self._better, self.mode = _monitor_op(mode, monitor)
        self.best = math.inf if self.mode == 'min' else -math.inf

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        if self.save_best_only:
            current = logs.get(self.monitor)
            if current is None or not self._better(current, self.best):
                return
            self.best = current
        path = self.filepath.format(epoch=epoch + 1, **logs)
        if self.save_weights_only:
            self.model.save_weights(path)
        else:
            self.model.save(path)
        if self.verbose:
            print("Epoch %d: saving model to %s" % (epoch + 1, path))


class LearningRateScheduler(Callback):
    """``schedule(epoch, lr) -> lr`` (or ``schedule(epoch)``) applied before every epoch."""

    def __init__(self, schedule, verbose=0):
        super().__init__()
        self.schedule, self.verbose = schedule, verbose

    def on_epoch_begin(self, epoch, logs=None):
        lr = float(self.model.optimizer.learning_rate)
        try:
            new = self.schedule(epoch, lr)
        except TypeError:
            new = self.schedule(epoch)
        self.model.optimizer.learning_rate = float(new)
        if self.verbose:
            print("Epoch %d: learning rate %.6g" % (epoch + 1, float(new)))

    def on_epoch_end(self, epoch, logs=None):
        if logs is not None:
            logs['lr'] = float(self.model.optimizer.learning_rate)


class ReduceLROnPlateau(Callback):
    def __init__(self, monitor='val_loss', factor=0.1, patience=10, verbose=0, mode='auto', min_delta=1e-4, cooldown=0,
                 min_lr=0.0):
        super().__init__()
        if factor >= 1.0:
            raise ValueError("ReduceLROnPlateau does not support a factor >= 1.0")
        self.monitor, self.factor, self.patience, self.verbose = monitor, float(factor), int(patience), verbose
        self.min_delta, self.cooldown, self.min_lr = abs(float(min_delta)), int(cooldown), float(min_lr)
        self._better, self.mode = _monitor_op(mode, monitor)

    def on_train_begin(self, logs=None):
        self.best = math.inf if self.mode == 'min' else -math.inf
        self.wait = self.cooldown_counter = 0

    def on_epoch_end(self, epoch, logs=None):
        logs = logs if logs is not None else {}
        logs['lr'] = float(self.model.optimizer.learning_rate)
        current = logs.get(self.monitor)
        if current is None:
            return
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.wait = 0
        threshold = self.best - self.min_delta if self.mode == 'min' else self.best + self.min_delta
        if self._better(current, threshold):
            self.best, self.wait = current, 0
        elif self.cooldown_counter <= 0:
            self.wait += 1
            if self.wait >= self.patience:
                old = float(self.model.optimizer.learning_rate)
                if old > self.min_lr:
                    self.model.optimizer.learning_rate = max(old * self.factor, self.min_lr)
                    if self.verbose:
                        print("Epoch %d: reducing learning rate to %.6g" % (epoch + 1, self.model.optimizer.learning_rate))
                    self.cooldown_counter, self.wait = self.cooldown, 0


class TerminateOnNaN(Callback):
    def on_epoch_end(self, epoch, logs=None):
        loss = (logs or {}).get('loss')
        if loss is not None and not np.isfinite(loss):
            self.model.stop_training = True


class CSVLogger(Callback):
    """One line per epoch (``epoch`` and every entry of the logs) into ``filename``; ``append`` keeps an existing file."""

    def __init__(self, filename, separator=',', append=False):
        super().__init__()
        self.filename, self.separator, self.append = str(filename), separator, bool(append)
        self._keys = None

    def on_train_begin(self, logs=None):
        import os
        keep = self.append and os.path.exists(self.filename) and os.path.getsize(self.filename) > 0
        if keep and self._keys is None:
            with open(self.filename) as fh:
                self._keys = fh.readline().rstrip('\n').split(self.separator)[1:]
        self._fh = open(self.filename, 'a' if keep else 'w')

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        if self._keys is None:
            self._keys = sorted(logs)
            self._fh.write(self.separator.join(['epoch'] + self._keys) + '\n')
        self._fh.write(self.separator.join([str(epoch)] + [repr(float(logs[k])) if k in logs else 'NA' for k in self._keys]) + '\n')
        self._fh.flush()

    def on_train_end(self, logs=None):
        self._fh.close()


class CallbackList:
    """What ``Model.fit`` drives: plain objects that only define ``on_epoch_end`` (the reference's tests pass such
    objects) work as well as :class:`Callback` subclasses."""

    def __init__(self, callbacks, model, params):
        self.callbacks = list(callbacks or [])
        for cb in self.callbacks:
            if hasattr(cb, 'set_model'):
                cb.set_model(model)
            if hasattr(cb, 'set_params'):
                cb.set_params(params)

    def __bool__(self):
        return bool(self.callbacks)

    def _call(self, hook, *args):
        for cb in self.callbacks:
            fn = getattr(cb, hook, None)
            if fn is not None:
                fn(*args)

    def on_train_begin(self):
        self._call('on_train_begin', {})

    def on_train_end(self):
        self._call('on_train_end', {})

    def on_epoch_begin(self, epoch):
        self._call('on_epoch_begin', epoch, {})

    def on_epoch_end(self, epoch, logs):
        self._call('on_epoch_end', epoch, logs)
