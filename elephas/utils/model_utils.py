from elephas_b200.utils.model_utils import *  # noqa: F401,F403
