"""``pyspark.mllib.evaluation`` stand-ins used by the reference's examples and tests
(``tests/test_ml_model.py:96-98,163-165``): ``MulticlassMetrics`` over ``(prediction, label)`` pairs
and ``RegressionMetrics`` over ``(prediction, observation)`` pairs."""
import numpy as np


def _pairs(rdd_or_list):
    rows = rdd_or_list.collect() if hasattr(rdd_or_list, 'collect') else list(rdd_or_list)
    arr = np.asarray([[float(a), float(b)] for a, b in rows], dtype=np.float64)
    return arr.reshape(-1, 2)


class MulticlassMetrics:
    def __init__(self, predictionAndLabels):
        pl = _pairs(predictionAndLabels)
        self._pred, self._label = pl[:, 0], pl[:, 1]
        self._classes = np.unique(np.concatenate([self._pred, self._label]))

    @property
    def accuracy(self):
        return float((self._pred == self._label).mean()) if len(self._pred) else 0.0

    def precision(self, label=None):
        if label is None:
            return self.accuracy
        sel = self._pred == label
        return float((self._label[sel] == label).mean()) if sel.any() else 0.0

    def recall(self, label=None):
        if label is None:
            return self.accuracy
        sel = self._label == label
        return float((self._pred[sel] == label).mean()) if sel.any() else 0.0

    def fMeasure(self, label=None, beta=1.0):
        if label is None:
            return self.accuracy
        p, r = self.precision(label), self.recall(label)
        return (1 + beta ** 2) * p * r / (beta ** 2 * p + r) if (p + r) else 0.0

    @property
    def weightedPrecision(self):
        return float(sum(self.precision(c) * (self._label == c).mean() for c in self._classes))

    @property
    def weightedRecall(self):
        return float(sum(self.recall(c) * (self._label == c).mean() for c in self._classes))

    def confusionMatrix(self):
        idx = {c: i for i, c in enumerate(self._classes)}
        cm = np.zeros((len(idx), len(idx)))
        for p, l in zip(self._pred, self._label):
            cm[idx[l], idx[p]] += 1
        return cm


class RegressionMetrics:
    def __init__(self, predictionAndObservations):
        po = _pairs(predictionAndObservations)
        self._pred, self._obs = po[:, 0], po[:, 1]

    @property
    def meanSquaredError(self):
        return float(((self._pred - self._obs) ** 2).mean())

    @property
    def rootMeanSquaredError(self):
        return float(np.sqrt(self.meanSquaredError))

    @property
    def meanAbsoluteError(self):
        return float(np.abs(self._pred - self._obs).mean())

    @property
    def r2(self):
        ss_res = float(((self._obs - self._pred) ** 2).sum())
        ss_tot = float(((self._obs - self._obs.mean()) ** 2).sum())
        return 1.0 - ss_res / ss_tot if ss_tot else 0.0

    @property
    def explainedVariance(self):
        return float(((self._pred - self._obs.mean()) ** 2).mean())
