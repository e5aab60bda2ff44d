"""A Keras-style sequence model through SparkModel with the training conveniences a Keras script uses: weight penalties,
a dropout-regularised LSTM, a learning-rate schedule in the optimizer config, early stopping and best-weights checkpoints.

The model reaches the workers as JSON, the optimizer (schedule included) as a config dictionary, the callbacks inside the
fit arguments - as with the reference, whose workers call ``keras.Model.fit(x, y, **train_config)``.

    python examples/lstm_callbacks.py
"""
import os
import tempfile

import numpy as np

from _data import os as _os  # noqa: F401  (puts the repository root on sys.path)
from elephas.spark_model import SparkModel
from elephas.utils.rdd_utils import to_simple_rdd
from elephas_b200.models import (LSTM, Adam, Dense, Embedding, Sequential, callbacks, constraints, load_model, regularizers,
                                 schedules)
from elephas_b200.parallel import get_runtime
from elephas_b200.spark import SparkConf, SparkContext

import torch

rt = get_runtime()
torch.manual_seed(0)
rng = np.random.default_rng(0)
n, steps, vocab = 1536, 12, 40
tokens = rng.integers(1, vocab, size=(n, steps))
labels = (np.sum(tokens < vocab // 4, axis=1) > np.sum(tokens >= 3 * vocab // 4, axis=1)).astype('int64')  # more small than large tokens?
x, y = tokens.astype('float32'), np.eye(2, dtype='float32')[labels]

model = Sequential([
    Embedding(vocab, 16, input_length=steps, embeddings_regularizer=regularizers.l2(1e-5)),
    LSTM(24, dropout=0.1, kernel_regularizer=regularizers.l2(1e-4), recurrent_constraint=constraints.MaxNorm(3.0)),
    Dense(2, activation='softmax', kernel_regularizer='l2'),
])
model.compile(Adam(learning_rate=schedules.ExponentialDecay(0.02, decay_steps=40, decay_rate=0.9)),
              'categorical_crossentropy', ['acc'])

sc = SparkContext(conf=SparkConf().setAppName('lstm_callbacks').setMaster('local[2]'))
best = os.path.join(tempfile.mkdtemp(), 'best.h5')
spark_model = SparkModel(model, mode='synchronous', num_workers=2)
spark_model.fit(to_simple_rdd(sc, x, y), epochs=12, batch_size=32, verbose=0, validation_split=0.15,
                callbacks=[callbacks.EarlyStopping(monitor='val_loss', patience=4, restore_best_weights=True),
                           callbacks.ModelCheckpoint(best, monitor='val_acc', save_best_only=True)])
loss, acc = spark_model.evaluate(x[:512], y[:512])
if rt.is_driver:
    print('loss %.4f  acc %.3f  checkpoint acc %.3f' % (loss, acc, load_model(best).evaluate(x[:512], y[:512])[1]))
    assert acc > 0.7
