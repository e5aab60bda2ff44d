"""``pyspark.mllib.regression.LabeledPoint`` stand-in."""
import numpy as np

from .linalg import DenseVector, Vector


class LabeledPoint:
    __slots__ = ('label', 'features')

    def __init__(self, label, features):
        self.label = float(np.asarray(label).reshape(-1)[0])
        self.features = features if isinstance(features, Vector) else DenseVector(features)

    def __repr__(self):
        return "LabeledPoint(%s, %s)" % (self.label, self.features)
