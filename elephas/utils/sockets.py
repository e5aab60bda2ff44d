from elephas_b200.utils.sockets import *  # noqa: F401,F403
from elephas_b200.utils.sockets import _receive_all  # noqa: F401  (private helper some callers import by name)
