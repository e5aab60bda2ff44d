"""The PySpark surface Elephas needs, resolved once.

If PySpark is importable (and ``ELEPHAS_B200_FORCE_LOCAL_SPARK`` is unset) every name below is the
real PySpark object, so ``ElephasEstimator`` plugs into real ``pyspark.ml.Pipeline`` objects and
``SparkModel.fit`` accepts real RDDs (their partitions are brought to the driver and handed to the
GPU ranks). Otherwise the local stand-ins of this package are used: same names, same semantics for
the subset Elephas touches, no JVM. The reference hard-depends on PySpark <= 3.3
(``pyproject.toml:15``), which cannot be installed in the B200 image.
"""
import os

HAVE_PYSPARK = False
if not os.environ.get('ELEPHAS_B200_FORCE_LOCAL_SPARK'):
    try:  # pragma: no cover - pyspark is not part of the B200 image
        import pyspark  # noqa: F401
        HAVE_PYSPARK = True
    except Exception:
        HAVE_PYSPARK = False

if HAVE_PYSPARK:  # pragma: no cover
    from pyspark import RDD, SparkConf, SparkContext, keyword_only  # noqa: F401
    from pyspark.broadcast import Broadcast  # noqa: F401
    from pyspark.ml import Estimator, Model, Pipeline, PipelineModel, Transformer  # noqa: F401
    from pyspark.ml.feature import (IndexToString, MinMaxScaler, Normalizer, OneHotEncoder, StandardScaler,  # noqa: F401
                                    StringIndexer, VectorAssembler)
    from pyspark.ml.evaluation import (BinaryClassificationEvaluator, MulticlassClassificationEvaluator,  # noqa: F401
                                       RegressionEvaluator)
    from pyspark.ml.linalg import Vector as MLVector, Vectors as MLVectors  # noqa: F401
    from pyspark.ml.param.shared import HasFeaturesCol, HasLabelCol, HasOutputCol, Param, Params  # noqa: F401
    from pyspark.ml.util import DefaultParamsReadable, DefaultParamsWritable  # noqa: F401
    from pyspark.mllib.evaluation import MulticlassMetrics, RegressionMetrics  # noqa: F401
    from pyspark.mllib.linalg import DenseMatrix, DenseVector, Matrices, Matrix, Vector, Vectors  # noqa: F401
    from pyspark.mllib.regression import LabeledPoint  # noqa: F401
    from pyspark.sql import DataFrame, Row, SparkSession  # noqa: F401
    from pyspark.sql.types import ArrayType, DoubleType, StructField, StructType  # noqa: F401
    PairBlock = ArrayBlock = None
else:
    from .context import ArrayBlock, Broadcast, PairBlock, RDD, SparkConf, SparkContext  # noqa: F401
    from .evaluation import MulticlassMetrics, RegressionMetrics  # noqa: F401
    from .linalg import DenseMatrix, DenseVector, Matrices, Matrix, Vector, Vectors  # noqa: F401
    from .linalg import Vector as MLVector, Vectors as MLVectors  # noqa: F401
    from .ml import (DefaultParamsReadable, DefaultParamsWritable, Estimator, HasFeaturesCol, HasLabelCol,  # noqa: F401
                     HasOutputCol, Model, Param, Params, Pipeline, PipelineModel, StandardScaler, StringIndexer,
                     Transformer, keyword_only)
    from .feature import IndexToString, MinMaxScaler, Normalizer, OneHotEncoder, VectorAssembler  # noqa: F401
    from .ml_evaluation import (BinaryClassificationEvaluator, MulticlassClassificationEvaluator,  # noqa: F401
                                RegressionEvaluator)
    from .regression import LabeledPoint  # noqa: F401
    from .sql import (ArrayType, Column, DataFrame, DoubleType, Row, SparkSession, StructField, StructType,  # noqa: F401
                      functions)


def get_spark_context():
    """The active SparkContext (created on demand), whichever backend is in use."""
    return SparkSession.builder.getOrCreate().sparkContext


def materialize_partitions(rdd):
    """``rdd`` -> list of partitions. Local RDDs hand out their blocks without copying; real PySpark
    RDDs are gathered with ``glom().collect()`` (the single-node analogue of executor-local data)."""
    if hasattr(rdd, 'partitions') and callable(getattr(rdd, 'partitions')):
        return rdd.partitions()
    return rdd.glom().collect()
