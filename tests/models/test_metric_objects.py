"""Count-based metrics (``models/metrics.py``: Precision, Recall, AUC, RootMeanSquaredError) against scikit-learn, and
their life cycle inside ``fit`` / ``evaluate``."""
import numpy as np
import pytest
import torch
from sklearn import metrics as sk

from elephas_b200.models import Dense, Sequential, metrics
from elephas_b200.models.optimizers import SGD


def _scores(n=2000, seed=0):
    rng = np.random.default_rng(seed)
    y = (rng.random(n) < 0.3).astype(np.float32)
    p = np.clip(0.35 * y + 0.65 * rng.random(n), 0, 1).astype(np.float32)
    return y, p


def test_values_match_scikit_learn_whatever_the_batching():
    y, p = _scores()
    objs = [metrics.Precision(), metrics.Recall(thresholds=0.4), metrics.AUC(num_thresholds=400), metrics.AUC(curve='PR'),
            metrics.RootMeanSquaredError()]
    for lo in range(0, len(y), 333):     # uneven batches: the counts add up
        for m in objs:
            m(torch.tensor(y[lo:lo + 333]).unsqueeze(1), torch.tensor(p[lo:lo + 333]).unsqueeze(1))
    assert objs[0].result() == pytest.approx(sk.precision_score(y, p > 0.5), abs=1e-9)
    assert objs[1].result() == pytest.approx(sk.recall_score(y, p > 0.4), abs=1e-9)
    assert objs[2].result() == pytest.approx(sk.roc_auc_score(y, p), abs=2e-3)
    assert objs[3].result() == pytest.approx(sk.average_precision_score(y, p), abs=2e-2)
    assert objs[4].result() == pytest.approx(float(np.sqrt(np.mean((y - p) ** 2))), rel=1e-6)
    objs[0].reset_state()
    assert objs[0].result() == 0.0
    with pytest.raises(ValueError):
        metrics.Precision(top_k=3)
    with pytest.raises(ValueError):
        metrics.AUC(curve='DET')


def test_fit_and_evaluate_report_epoch_level_results():
    rng = np.random.default_rng(1)
    x = rng.standard_normal((600, 5)).astype(np.float32)
    y = (x @ rng.standard_normal(5) + 0.3 * rng.standard_normal(600) > 0.5).astype(np.float32).reshape(-1, 1)
    torch.manual_seed(0)
    m = Sequential([Dense(8, activation='relu', input_shape=(5,)), Dense(1, activation='sigmoid')])
    mine = metrics.Precision(name='prec')
    m.compile(SGD(learning_rate=0.3), 'binary_crossentropy', ['acc', mine, 'recall', 'auc'])
    assert m.metrics_names == ['loss', 'acc', 'prec', 'recall', 'auc']
    assert m._cpu_step() is None                       # the hand-written CPU step leaves such models to autograd
    h = m.fit(x, y, batch_size=50, epochs=4, validation_split=0.2)
    assert set(h.history) == {'loss', 'acc', 'prec', 'recall', 'auc', 'val_loss', 'val_acc', 'val_prec', 'val_recall', 'val_auc'}
    assert h.history['auc'][-1] > 0.9 and h.history['val_auc'][-1] > 0.9
    loss, acc, prec, rec, auc = m.evaluate(x, y, batch_size=64)
    p = m.predict(x).reshape(-1)
    assert prec == pytest.approx(sk.precision_score(y.reshape(-1), p > 0.5), abs=1e-6)
    assert rec == pytest.approx(sk.recall_score(y.reshape(-1), p > 0.5), abs=1e-6)
    assert auc == pytest.approx(sk.roc_auc_score(y.reshape(-1), p), abs=5e-3)
    assert mine.result() == 0.0                        # the model counts in its own copy of the object
    # evaluating twice gives the same numbers: the counts start from zero each time
    assert m.evaluate(x, y, batch_size=64)[2] == pytest.approx(prec, abs=1e-12)
