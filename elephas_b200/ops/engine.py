"""Native sm_100a engine selection (placeholder until the engine lands)."""


def maybe_create(model):
    return None


def maybe_create_inference(model):
    return None


def try_sync_fit(spark_model, rt, partitions, mine, train_config, sync_rounds):
    return None
