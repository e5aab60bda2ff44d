from elephas_b200.mllib.adapter import *  # noqa: F401,F403
