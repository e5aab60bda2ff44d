from elephas_b200.utils.rdd_utils import *  # noqa: F401,F403
