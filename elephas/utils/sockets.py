from elephas_b200.utils.sockets import *  # noqa: F401,F403
