from elephas_b200.ml.params import *  # noqa: F401,F403
