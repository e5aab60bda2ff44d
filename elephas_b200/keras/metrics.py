"""``keras.metrics``: the per-sample functions and the count-based metric classes."""
from ..models.metrics import (AUC, Metric, Precision, Recall, RootMeanSquaredError, binary_accuracy,  # noqa: F401
                              categorical_accuracy, sparse_categorical_accuracy, sparse_top_k_categorical_accuracy,
                              top_k_categorical_accuracy)
from ..models.losses import mean_absolute_error, mean_absolute_percentage_error, mean_squared_error  # noqa: F401
