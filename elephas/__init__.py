"""Drop-in import paths of the reference package (``import elephas...``): thin re-exports of
``elephas_b200``. Scripts written against Elephas keep their imports; model building moves from
``tensorflow.keras`` to ``elephas_b200.models`` and the Spark context to ``elephas_b200.spark``."""
from elephas_b200 import __version__  # noqa: F401
