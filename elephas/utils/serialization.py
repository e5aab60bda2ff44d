from elephas_b200.utils.serialization import *  # noqa: F401,F403
