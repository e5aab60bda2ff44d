from elephas_b200.ml.adapter import *  # noqa: F401,F403
