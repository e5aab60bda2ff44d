"""Public utility surface. Mirrors ``elephas/utils/__init__.py`` (which star-exports functional_utils,
rdd_utils, serialization, sockets, rwlock); ``model_utils`` and ``notebook_utils`` are imported by
their module path, exactly as in the reference."""
from .functional_utils import *  # noqa: F401,F403
from .functional_utils import add_params, divide_by, get_neutral, subtract_params  # noqa: F401
from .rdd_utils import *  # noqa: F401,F403
from .rdd_utils import (encode_label, from_labeled_point, lp_to_simple_rdd, to_labeled_point,  # noqa: F401
                        to_sharded_rdd, to_simple_rdd)
from .rwlock import *  # noqa: F401,F403
from .rwlock import RWLock  # noqa: F401
from .serialization import *  # noqa: F401,F403
from .serialization import dict_to_model, model_to_dict  # noqa: F401
from .sockets import *  # noqa: F401,F403
from .sockets import determine_master, receive, send  # noqa: F401
