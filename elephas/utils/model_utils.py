from elephas_b200.utils.model_utils import *  # noqa: F401,F403
from elephas_b200.utils.model_utils import Singleton as _Singleton  # noqa: F401  (the reference's metaclass name)
