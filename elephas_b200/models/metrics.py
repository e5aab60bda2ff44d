"""Keras-named metrics on torch. ``'acc'`` / ``'accuracy'`` resolve against the loss and the
output shape the way ``keras.Model.compile`` does (binary / categorical / sparse categorical).

Reference use: ``['acc']`` (``tests/integration/test_end_to_end.py:40``), ``['mae',
'mean_absolute_percentage_error']`` (``:92``), ``['mae']`` (``tests/test_ml_model.py:151``).
Each metric returns a per-sample vector; callers average.
"""

from . import losses as _losses


def binary_accuracy(y_true, y_pred, threshold=0.5):
    y_true = _losses._match(y_true, y_pred)
    hit = ((y_pred > threshold).to(y_pred.dtype) == y_true).to(y_pred.dtype)
    return hit.flatten(1).mean(dim=1) if hit.dim() > 1 else hit


def categorical_accuracy(y_true, y_pred):
    return (y_true.argmax(dim=-1) == y_pred.argmax(dim=-1)).to(y_pred.dtype)


def sparse_categorical_accuracy(y_true, y_pred):
    idx = y_true.reshape(y_pred.shape[0], -1)[:, 0].long()
    return (idx == y_pred.argmax(dim=-1)).to(y_pred.dtype)


def top_k_categorical_accuracy(y_true, y_pred, k=5):
    k = min(int(k), y_pred.shape[-1])
    top = y_pred.topk(k, dim=-1).indices
    return (top == y_true.argmax(dim=-1, keepdim=True)).any(dim=-1).to(y_pred.dtype)


def sparse_top_k_categorical_accuracy(y_true, y_pred, k=5):
    k = min(int(k), y_pred.shape[-1])
    idx = y_true.reshape(y_pred.shape[0], -1)[:, :1].long()
    return (y_pred.topk(k, dim=-1).indices == idx).any(dim=-1).to(y_pred.dtype)


_BUILTIN = {
    'binary_accuracy': binary_accuracy, 'categorical_accuracy': categorical_accuracy,
    'sparse_categorical_accuracy': sparse_categorical_accuracy,
    'top_k_categorical_accuracy': top_k_categorical_accuracy,
    'sparse_top_k_categorical_accuracy': sparse_top_k_categorical_accuracy,
}


# ---------------------------------------------------------------------------------------------------- metric objects
class Metric:
    """A metric that is not a mean over rows (precision, recall, AUC, RMSE): it accumulates counts over the batches of an
    epoch (``reset_state`` / ``update_state`` / ``result`` as in ``keras.metrics``). Calling it updates the counts and
    returns one value per row, all equal to the result so far - so code that averages per-row vectors keeps working; the
    model replaces the average by ``result()`` at the end of the epoch / evaluation. Models compiled with one train on
    the autograd path."""

    stateful = True
    default_name = 'metric'

    def __init__(self, name=None):
        self.name = name or self.default_name
        self.reset_state()

    @property
    def __name__(self):
        return self.name

    def __call__(self, y_true, y_pred):
        self.update_state(y_true, y_pred)
        return y_pred.new_full((y_pred.shape[0],), self.result())

    def reset_state(self):
        raise NotImplementedError

    def update_state(self, y_true, y_pred):
        raise NotImplementedError

    def result(self):
        raise NotImplementedError


def _flat_pair(y_true, y_pred):
    y_true = _losses._match(y_true, y_pred)
    return y_true.reshape(-1).double(), y_pred.reshape(-1).double()


class _Confusion(Metric):
    """True / false positives and false negatives at one threshold (``top_k`` / ``class_id`` are not supported)."""

    def __init__(self, thresholds=None, name=None, **kwargs):
        if kwargs.get('top_k') is not None or kwargs.get('class_id') is not None:
            raise ValueError("%s: top_k / class_id are not supported" % type(self).__name__)
        if isinstance(thresholds, (list, tuple)):
            if len(thresholds) != 1:
                raise ValueError("%s: one threshold only" % type(self).__name__)
            thresholds = thresholds[0]
        self.threshold = 0.5 if thresholds is None else float(thresholds)
        super().__init__(name)

    def reset_state(self):
        self.tp = self.fp = self.fn = 0.0

    def update_state(self, y_true, y_pred):
        t, p = _flat_pair(y_true, y_pred)
        hit = p > self.threshold
        pos = t > 0.5
        self.tp += float((hit & pos).sum())
        self.fp += float((hit & ~pos).sum())
        self.fn += float((~hit & pos).sum())


class Precision(_Confusion):
    default_name = 'precision'

    def result(self):
        return self.tp / (self.tp + self.fp) if self.tp + self.fp else 0.0


class Recall(_Confusion):
    default_name = 'recall'

    def result(self):
        return self.tp / (self.tp + self.fn) if self.tp + self.fn else 0.0


class AUC(Metric):
    """Area under the ROC (or precision-recall) curve from confusion counts at ``num_thresholds`` evenly spaced
    thresholds, Riemann sum with linear interpolation - the construction of ``keras.metrics.AUC``."""

    default_name = 'auc'

    def __init__(self, num_thresholds=200, curve='ROC', summation_method='interpolation', name=None, **kwargs):
        if curve not in ('ROC', 'PR') or summation_method != 'interpolation':
            raise ValueError("AUC: curve must be 'ROC' or 'PR', summation_method 'interpolation'")
        self.num_thresholds, self.curve = int(num_thresholds), curve
        eps = 1e-7
        import torch
        inner = [(i + 1) / (self.num_thresholds - 1) for i in range(self.num_thresholds - 2)]
        self._thresholds = torch.tensor([0.0 - eps] + inner + [1.0 + eps], dtype=torch.float64)
        super().__init__(name)

    def reset_state(self):
        import torch
        self._counts = torch.zeros(4, self.num_thresholds, dtype=torch.float64)   # tp, fp, tn, fn per threshold

    def update_state(self, y_true, y_pred):
        import torch
        t, p = _flat_pair(y_true, y_pred)
        t, p = t.cpu(), p.cpu()
        hit = p.unsqueeze(0) > self._thresholds.unsqueeze(1)
        pos = (t > 0.5).unsqueeze(0)
        self._counts += torch.stack([(hit & pos).sum(1), (hit & ~pos).sum(1), (~hit & ~pos).sum(1),
                                     (~hit & pos).sum(1)]).double()

    def result(self):
        tp, fp, tn, fn = self._counts
        eps = 1e-7
        recall = (tp + eps) / (tp + fn + eps)
        if self.curve == 'ROC':
            x, y = fp / (fp + tn + eps), recall
        else:
            x, y = recall, (tp + eps) / (tp + fp + eps)
        n = self.num_thresholds
        return float(((x[:n - 1] - x[1:]) * (y[:n - 1] + y[1:]) / 2.0).sum())


class RootMeanSquaredError(Metric):
    default_name = 'root_mean_squared_error'

    def reset_state(self):
        self.total, self.count = 0.0, 0

    def update_state(self, y_true, y_pred):
        t, p = _flat_pair(y_true, y_pred)
        self.total += float(((p - t) ** 2).sum())
        self.count += int(t.numel())

    def result(self):
        return (self.total / self.count) ** 0.5 if self.count else 0.0


_STATEFUL = {'precision': Precision, 'recall': Recall, 'auc': AUC, 'root_mean_squared_error': RootMeanSquaredError,
             'rmse': RootMeanSquaredError, 'Precision': Precision, 'Recall': Recall, 'AUC': AUC,
             'RootMeanSquaredError': RootMeanSquaredError}


def is_stateful(fn):
    return bool(getattr(fn, 'stateful', False))


def resolve(identifier, loss_name, output_dim, custom_objects=None):
    """Return ``(display_name, fn)`` for one entry of ``compile(metrics=[...])``."""
    if isinstance(identifier, Metric):
        import copy
        own = copy.deepcopy(identifier)  # every compiled model (every worker replica) counts for itself
        own.reset_state()
        return own.name, own
    if isinstance(identifier, str) and identifier in _STATEFUL and not (custom_objects and identifier in custom_objects):
        metric = _STATEFUL[identifier]()
        return metric.name, metric
    if callable(identifier):
        return getattr(identifier, '__name__', 'metric'), identifier
    if identifier in ('acc', 'accuracy'):
        if loss_name == 'sparse_categorical_crossentropy':
            fn = sparse_categorical_accuracy
        elif loss_name == 'binary_crossentropy' or output_dim == 1:
            fn = binary_accuracy
        else:
            fn = categorical_accuracy
        return identifier, fn
    if custom_objects and identifier in custom_objects:
        return identifier, custom_objects[identifier]
    if identifier in _BUILTIN:
        return identifier, _BUILTIN[identifier]
    if identifier in _losses._BUILTIN:
        return identifier, _losses._BUILTIN[identifier]
    raise ValueError("Unknown metric: %r" % (identifier,))
