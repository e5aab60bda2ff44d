"""``pyspark.ml.evaluation`` for DataFrames of predictions: the evaluators a pipeline script calls after
``fitted.transform(test_df)`` (the RDD-based ``MulticlassMetrics`` / ``RegressionMetrics`` live in ``evaluation.py``)."""
import numpy as np

from .evaluation import MulticlassMetrics, RegressionMetrics
from .ml import HasLabelCol, Param, Params, keyword_only


class _Evaluator(Params):
    _metrics = ()
    _larger_is_better = ()

    def _declare(self, kwargs, default_metric, prediction_default='prediction'):
        self.predictionCol = Param(self, 'predictionCol', 'prediction column name')
        self.metricName = Param(self, 'metricName', 'metric name: ' + ' | '.join(self._metrics))
        self._setDefault(predictionCol=prediction_default, metricName=default_metric, labelCol='label')
        self._set(**{k: v for k, v in kwargs.items() if v is not None})
        name = self.getOrDefault(self.metricName)
        if name not in self._metrics:
            raise ValueError("metricName must be one of %s, got %r" % (list(self._metrics), name))

    def setMetricName(self, value):
        if value not in self._metrics:
            raise ValueError("metricName must be one of %s, got %r" % (list(self._metrics), value))
        return self._set(metricName=value)

    def getMetricName(self):
        return self.getOrDefault(self.metricName)

    def setPredictionCol(self, value):
        return self._set(predictionCol=value)

    def setLabelCol(self, value):
        return self._set(labelCol=value)

    def isLargerBetter(self):
        return self.getMetricName() in self._larger_is_better

    def evaluate(self, dataset, params=None):
        if params:
            return self.copy(params).evaluate(dataset)
        return float(self._evaluate(dataset))

    def _columns(self, dataset):
        pred = np.asarray(dataset.column_block(self.getOrDefault(self.predictionCol)), dtype=np.float64)
        label = np.asarray(dataset.column_block(self.getLabelCol()), dtype=np.float64).reshape(-1)
        return pred, label


class MulticlassClassificationEvaluator(_Evaluator, HasLabelCol):
    _metrics = ('f1', 'accuracy', 'weightedPrecision', 'weightedRecall')
    _larger_is_better = _metrics

    @keyword_only
    def __init__(self, predictionCol=None, labelCol=None, metricName=None):
        super().__init__()
        self._declare(self._input_kwargs, 'f1')

    def _evaluate(self, dataset):
        pred, label = self._columns(dataset)
        if pred.ndim == 2:   # a probability / score vector per row: the predicted class is its arg-max
            pred = pred.argmax(axis=1)
        from .context import RDD
        pairs = RDD([list(zip(pred.astype(float).tolist(), label.tolist()))], dataset.rdd.context)
        m = MulticlassMetrics(pairs)
        name = self.getMetricName()
        if name == 'accuracy':
            return m.accuracy
        if name == 'weightedPrecision':
            return m.weightedPrecision
        if name == 'weightedRecall':
            return m.weightedRecall
        classes, counts = np.unique(label, return_counts=True)
        return float(sum(m.fMeasure(float(c)) * n for c, n in zip(classes, counts)) / max(len(label), 1))


class RegressionEvaluator(_Evaluator, HasLabelCol):
    _metrics = ('rmse', 'mse', 'r2', 'mae', 'var')
    _larger_is_better = ('r2', 'var')

    @keyword_only
    def __init__(self, predictionCol=None, labelCol=None, metricName=None):
        super().__init__()
        self._declare(self._input_kwargs, 'rmse')

    def _evaluate(self, dataset):
        pred, label = self._columns(dataset)
        from .context import RDD
        m = RegressionMetrics(RDD([list(zip(pred.reshape(-1).tolist(), label.tolist()))], dataset.rdd.context))
        return {'rmse': m.rootMeanSquaredError, 'mse': m.meanSquaredError, 'r2': m.r2, 'mae': m.meanAbsoluteError,
                'var': m.explainedVariance}[self.getMetricName()]


class BinaryClassificationEvaluator(_Evaluator, HasLabelCol):
    """``areaUnderROC`` / ``areaUnderPR`` from a score column: a scalar score, or a vector whose LAST entry is the score of
    the positive class (Spark's ``rawPrediction`` / ``probability`` columns)."""

    _metrics = ('areaUnderROC', 'areaUnderPR')
    _larger_is_better = _metrics

    @keyword_only
    def __init__(self, rawPredictionCol=None, labelCol=None, metricName=None):
        super().__init__()
        kwargs = dict(self._input_kwargs)
        if kwargs.get('rawPredictionCol') is not None:
            kwargs['predictionCol'] = kwargs.pop('rawPredictionCol')
        kwargs.pop('rawPredictionCol', None)
        self._declare(kwargs, 'areaUnderROC', prediction_default='rawPrediction')

    def setRawPredictionCol(self, value):
        return self._set(predictionCol=value)

    def _evaluate(self, dataset):
        score, label = self._columns(dataset)
        score = score[:, -1] if score.ndim == 2 else score
        order = np.argsort(-score, kind='stable')
        score, pos = score[order], (label[order] > 0.5).astype(np.float64)
        last = np.r_[np.flatnonzero(np.diff(score)), len(score) - 1]   # one curve point per distinct score
        tp, fp = np.cumsum(pos)[last], np.cumsum(1.0 - pos)[last]
        total_p, total_n = pos.sum(), len(pos) - pos.sum()
        if total_p == 0 or total_n == 0:
            return 0.0
        if self.getMetricName() == 'areaUnderROC':
            return float(np.trapezoid(np.r_[0.0, tp / total_p], np.r_[0.0, fp / total_n]))
        recall, precision = np.r_[0.0, tp / total_p], np.r_[1.0, tp / (tp + fp)]
        return float(np.trapezoid(precision, recall))
