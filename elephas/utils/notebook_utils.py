from elephas_b200.utils.notebook_utils import *  # noqa: F401,F403
