from .adapter import *  # noqa: F401,F403
from .adapter import df_to_simple_rdd, from_data_frame, to_data_frame  # noqa: F401
from .params import *  # noqa: F401,F403
