"""Small CNN through SparkModel in synchronous mode: a model the native Dense engine does not cover.

On a GPU `set_compute_dtype('bfloat16')` puts every convolution, BatchNormalization (+ ReLU), pooling and Dense layer on
this repository's sm_100a kernels (implicit-GEMM convolutions on tcgen05, `ops/cnn.py`), the whole training step is
captured into a CUDA graph over flat weights with the native optimizer kernel, and the weight deltas are averaged by the
fused all-reduce + update kernel over NVLink (per batch with `frequency='batch'`: inside the captured step); on a CPU the
same script runs the eager torch path. BASELINE.json config 5 is the same flow with `zoo.resnet50()`
(`python bench.py --config resnet50`).

    python examples/cnn_sync.py
    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 examples/cnn_sync.py
"""
import numpy as np

from _data import os  # noqa: F401  (puts the repository root on sys.path)
from elephas.spark_model import SparkModel
from elephas.utils.rdd_utils import to_simple_rdd
from elephas_b200.models import SGD, Sequential
from elephas_b200.models.layers import (Activation, BatchNormalization, Conv2D, Dense, GlobalAveragePooling2D,
                                        MaxPooling2D)
from elephas_b200.parallel import get_runtime
from elephas_b200.spark import SparkConf, SparkContext

import torch

rt = get_runtime()
torch.manual_seed(0)  # reproducible initial weights
rng = np.random.default_rng(0)
n, side, classes = (2048 if rt.device.type == 'cuda' else 768), 16, 4  # (smaller on a CPU: the eager path is slow)
labels = rng.integers(0, classes, n)
x = rng.normal(size=(n, side, side, 3)).astype('float32')
for c in range(classes):  # class c brightens one quadrant
    r, q = divmod(c, 2)
    x[labels == c, r * 8:(r + 1) * 8, q * 8:(q + 1) * 8, :] += 1.0
y = np.eye(classes, dtype='float32')[labels]

model = Sequential([Conv2D(16, 3, padding='same', input_shape=(side, side, 3)), BatchNormalization(), Activation('relu'),
                    MaxPooling2D(2), Conv2D(32, 3, padding='same'), BatchNormalization(), Activation('relu'),
                    GlobalAveragePooling2D(), Dense(classes, activation='softmax')])
model.compile(SGD(learning_rate=0.05, momentum=0.9), 'categorical_crossentropy', ['acc'])
if rt.device.type == 'cuda':
    model.set_compute_dtype('bfloat16')

sc = SparkContext(conf=SparkConf().setAppName('cnn_sync').setMaster('local[%d]' % max(rt.world_size, 1)))
spark_model = SparkModel(model, mode='synchronous', frequency='epoch', num_workers=max(rt.world_size, 1))
spark_model.fit(to_simple_rdd(sc, x, y), epochs=6, batch_size=32, verbose=0, validation_split=0.1)
loss, acc = spark_model.evaluate(x[:512], y[:512])
if rt.is_driver:
    print('loss %.4f  acc %.3f  (%s)' % (loss, acc, 'graph-captured bf16 step' if rt.device.type == 'cuda' else 'eager CPU step'))
    assert acc > 0.6
