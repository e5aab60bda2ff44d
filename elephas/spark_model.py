from elephas_b200.spark_model import *  # noqa: F401,F403
from elephas_b200.spark_model import SparkMLlibModel, SparkModel, load_spark_model  # noqa: F401
