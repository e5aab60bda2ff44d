from elephas_b200.parameter import *  # noqa: F401,F403
