"""Executed by test_spmd.py under torch.distributed.run (gloo, world_size 2): every rank runs the same
driver program, SparkModel hands partition i to rank i % world."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ['ELEPHAS_B200_DEVICE'] = 'cpu'

from elephas.spark_model import SparkModel  # noqa: E402
from elephas.utils.rdd_utils import to_simple_rdd  # noqa: E402
from elephas_b200.models import SGD, Dense, Sequential  # noqa: E402
from elephas_b200.parallel import get_runtime  # noqa: E402
from elephas_b200.spark import SparkContext  # noqa: E402
from tests.conftest import make_mnist_like  # noqa: E402

mode, ps, freq, port = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
rt = get_runtime()
assert rt.world_size == 2
sc = SparkContext(master='local[2]')
x, y, xt, yt = make_mnist_like(600, 200)
import torch  # noqa: E402
torch.manual_seed(0)
features = len(sys.argv) > 6 and sys.argv[6] == 'features'
fit_extra, seen_epochs = {}, []
if features:
    # the parts of the Keras surface that route around the fused paths: penalties, a constraint, callbacks, class weights
    from elephas_b200.models import callbacks, constraints, regularizers  # noqa: E402
    model = Sequential([Dense(32, input_dim=784, activation='relu', kernel_regularizer=regularizers.l2(1e-4),
                              kernel_constraint=constraints.MaxNorm(4.0)), Dense(10, activation='softmax')])
    fit_extra = {'callbacks': [callbacks.LambdaCallback(on_epoch_end=lambda e, logs: seen_epochs.append(e)),
                               callbacks.EarlyStopping(monitor='loss', patience=3)],
                 'class_weight': {0: 2.0}}
else:
    model = Sequential([Dense(32, input_dim=784, activation='relu'), Dense(10, activation='softmax')])
model.compile(SGD(lr=0.1), 'categorical_crossentropy', ['acc'])
sm = SparkModel(model, mode=mode, frequency=freq, parameter_server_mode=None if ps == 'none' else ps,
                num_workers=2, port=port)
before = sm.master_network.evaluate(xt, yt)
sm.fit(to_simple_rdd(sc, x, y), epochs=2, batch_size=50, verbose=0, validation_split=0.0, shuffle=False, **fit_extra)
pred = sm.predict(xt)
ev = sm.evaluate(xt, yt)
local = sm.master_network.evaluate(xt, yt)
w = sm.master_network.get_weights()
digest = float(sum(np.abs(a).sum() for a in w))
out = {'rank': rt.rank, 'before': before, 'eval': ev, 'local': local, 'digest': digest, 'n_pred': len(pred),
       'argmax': [int(np.argmax(p)) for p in pred[:50]], 'hist': len(sm.training_histories), 'epochs_seen': seen_epochs,
       'max_col_norm': float(np.linalg.norm(w[0], axis=0).max())}
with open(os.path.join(sys.argv[5], 'rank%d.json' % rt.rank), 'w') as fh:
    json.dump(out, fh)
rt.barrier()
rt.shutdown()
