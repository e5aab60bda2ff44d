from elephas_b200.worker import *  # noqa: F401,F403
from elephas_b200.worker import AsynchronousSparkWorker, SparkWorker  # noqa: F401
