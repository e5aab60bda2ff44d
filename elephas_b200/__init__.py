"""elephas_b200: a Blackwell-native data-parallel training / inference framework with the
capabilities and public API of Elephas (reference: arpitjain799/elephas 4.1.0).

Layout
------
``models``     Keras-compatible model layer on torch (Sequential / functional Model, layers, losses,
               metrics, optimizers, ``model_from_json``) and the built-in workload zoo.
``ops``        Python bindings of the hand-written sm_100a kernels in ``csrc/`` plus CPU references.
``parallel``   One-rank-per-GPU runtime: process groups, symmetric (peer) memory, fused collectives,
               the device-resident parameter server.
``utils``      ``functional_utils``, ``rdd_utils``, ``serialization``, ``sockets``, ``rwlock``,
               ``model_utils``, ``notebook_utils`` (reference: ``elephas/utils``).
``parameter``  HTTP / socket / NVLink parameter servers and clients (reference: ``elephas/parameter``).
``spark``      Local stand-ins for the PySpark surface Elephas uses (used when PySpark is absent).
``ml`` / ``mllib``  Spark ML params + adapters (reference: ``elephas/ml``, ``elephas/mllib``).
``io``         Keras-style HDF5 checkpoints (built-in HDF5 subset writer/reader when h5py is absent).

The top-level modules ``spark_model``, ``ml_model`` and ``worker`` mirror the reference's import
paths; the ``elephas`` package in this repository re-exports everything under the original names.
"""
__version__ = "0.1.0"
