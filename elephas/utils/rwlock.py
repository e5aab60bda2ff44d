from elephas_b200.utils.rwlock import *  # noqa: F401,F403
