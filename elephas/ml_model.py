from elephas_b200.ml_model import *  # noqa: F401,F403
from elephas_b200.ml_model import (ElephasEstimator, ElephasTransformer, load_ml_estimator,  # noqa: F401
                                   load_ml_transformer)
