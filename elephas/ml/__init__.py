from elephas_b200.ml import *  # noqa: F401,F403
