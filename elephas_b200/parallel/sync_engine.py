"""Fast synchronous path: native engines + fused delta all-reduce / update (filled in by the GPU
engine; returns None whenever the generic worker path must be used)."""


def try_fit(spark_model, rt, partitions, mine, train_config, sync_rounds):
    if rt.device.type != 'cuda':
        return None
    try:
        from ..ops import engine as _engine
    except Exception:
        return None
    return _engine.try_sync_fit(spark_model, rt, partitions, mine, train_config, sync_rounds)
