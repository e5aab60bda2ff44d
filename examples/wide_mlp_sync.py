"""BASELINE.json config 4: wide MLP regression, synchronous, frequency='batch' - the bench.py workload at a
size that also runs on a CPU. Launch with torchrun for one rank per GPU."""
import _data  # noqa: F401  (puts the checkout on sys.path)
import numpy as np
from elephas.spark_model import SparkModel
from elephas_b200.models import SGD, zoo
from elephas_b200.parallel import get_runtime
from elephas_b200.spark import SparkContext
from elephas_b200.utils.rdd_utils import to_sharded_rdd

rt = get_runtime()
width = 4096 if rt.device.type == 'cuda' else 256
rng = np.random.default_rng(rt.rank)
x = rng.normal(size=(8192, width)).astype('float32')           # this rank's shard only
y = (x[:, :8].sum(1, keepdims=True) / 4).astype('float32')
model = zoo.wide_mlp(input_dim=width, hidden=width, depth=8, out_dim=1)
model.compile(SGD(learning_rate=1e-3), 'mse', ['mae'])
spark_model = SparkModel(model, mode='synchronous', frequency='batch')
# feature_dtype='bfloat16': the shard is staged once as bf16 in pinned host memory, input_mode='stream': every batch is
# copied host -> device while the previous step computes (half the PCIe bytes of an fp32 shard, no per-step CPU work)
rdd = to_sharded_rdd(SparkContext.getOrCreate(), x, y, feature_dtype='bfloat16' if rt.device.type == 'cuda' else None)
spark_model.fit(rdd, epochs=2, batch_size=1024, log_batch_loss=True,
                input_mode='stream' if rt.device.type == 'cuda' else None)
if rt.is_driver:
    print('epoch losses per rank:', [h['loss'] for h in spark_model.training_histories])
