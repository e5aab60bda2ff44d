from elephas_b200.parameter.factory import *  # noqa: F401,F403
