from elephas_b200.mllib import *  # noqa: F401,F403
