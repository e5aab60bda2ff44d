from elephas_b200.utils.functional_utils import *  # noqa: F401,F403
