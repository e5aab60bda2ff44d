from elephas_b200.parameter.client import *  # noqa: F401,F403
