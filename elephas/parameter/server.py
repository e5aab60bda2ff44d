from elephas_b200.parameter.server import *  # noqa: F401,F403
