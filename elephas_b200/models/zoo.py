"""Built-in workloads: the model shapes the reference ships in its tests and examples plus the two
large configurations named in BASELINE.json.

* ``mnist_mlp``     784-128-128-10, separate Activation/Dropout(0.2) layers (``tests/conftest.py:8-19``,
                    ``examples/mnist_mlp_spark.py:37-45``)
* ``mnist_mlp_functional``  functional-API variant (``tests/conftest.py:31-40``)
* ``boston_mlp``    13-64-64-1 regression, fused activations (``tests/conftest.py:22-28``)
* ``otto_mlp``      93-512-512-512-9 with Dropout(0.5) (``examples/ml_pipeline_otto.py:57-68``)
* ``wide_mlp``      4096-wide x 8 layer regression MLP (BASELINE.json config 4)
* ``resnet50``      ResNet-50-shaped CNN (BASELINE.json config 5)
* ``lstm_classifier`` / ``transformer_classifier``  sequence models (what a Keras user brings besides MLPs and CNNs: the
                    reference accepts any ``model_from_json``-able model, ``elephas/worker.py:30``)
"""
from .layers import (Activation, Add, BatchNormalization, Conv2D, Dense, Dropout, GlobalAveragePooling2D, Input,
                     MaxPooling2D, ZeroPadding2D)
from .model import Model, Sequential


def mnist_mlp(input_dim=784, hidden=128, classes=10, dropout=0.2):
    m = Sequential()
    m.add(Dense(hidden, input_dim=input_dim))
    m.add(Activation('relu'))
    m.add(Dropout(dropout))
    m.add(Dense(hidden))
    m.add(Activation('relu'))
    m.add(Dropout(dropout))
    m.add(Dense(classes))
    m.add(Activation('softmax'))
    return m


def mnist_mlp_functional(input_dim=784, hidden=128, classes=10, dropout=0.2):
    inp = Input(shape=(input_dim,))
    h = Dense(hidden, activation='relu')(inp)
    h = Dropout(dropout)(h)
    h = Dense(hidden, activation='relu')(h)
    h = Dropout(dropout)(h)
    out = Dense(classes, activation='softmax')(h)
    return Model(inputs=inp, outputs=out)


def boston_mlp(input_dim=13, hidden=64):
    m = Sequential()
    m.add(Dense(hidden, activation='relu', input_shape=(input_dim,)))
    m.add(Dense(hidden, activation='relu'))
    m.add(Dense(1, activation='linear'))
    return m


def otto_mlp(input_dim=93, hidden=512, classes=9, dropout=0.5):
    m = Sequential()
    m.add(Dense(hidden, input_shape=(input_dim,)))
    m.add(Activation('relu'))
    m.add(Dropout(dropout))
    for _ in range(2):
        m.add(Dense(hidden))
        m.add(Activation('relu'))
        m.add(Dropout(dropout))
    m.add(Dense(classes))
    m.add(Activation('softmax'))
    return m


def wide_mlp(input_dim=4096, hidden=4096, depth=8, out_dim=1):
    """``depth`` hidden Dense(hidden, relu) layers followed by a linear regression head:
    4096*4096*8 + 4096*8 + 4096 + 1 = 134,254,593 parameters at the default sizes."""
    m = Sequential()
    m.add(Dense(hidden, activation='relu', input_shape=(input_dim,)))
    for _ in range(depth - 1):
        m.add(Dense(hidden, activation='relu'))
    m.add(Dense(out_dim, activation='linear'))
    return m


def _bottleneck(x, filters, stride, project, name):
    f1, f2, f3 = filters, filters, filters * 4
    shortcut = x
    if project:
        shortcut = Conv2D(f3, 1, strides=stride, name=name + '_0_conv')(x)
        shortcut = BatchNormalization(epsilon=1.001e-5, name=name + '_0_bn')(shortcut)
    y = Conv2D(f1, 1, strides=stride, name=name + '_1_conv')(x)
    y = BatchNormalization(epsilon=1.001e-5, name=name + '_1_bn')(y)
    y = Activation('relu', name=name + '_1_relu')(y)
    y = Conv2D(f2, 3, padding='same', name=name + '_2_conv')(y)
    y = BatchNormalization(epsilon=1.001e-5, name=name + '_2_bn')(y)
    y = Activation('relu', name=name + '_2_relu')(y)
    y = Conv2D(f3, 1, name=name + '_3_conv')(y)
    y = BatchNormalization(epsilon=1.001e-5, name=name + '_3_bn')(y)
    y = Add(name=name + '_add')([shortcut, y])
    return Activation('relu', name=name + '_out')(y)


def resnet50(input_shape=(224, 224, 3), classes=1000):
    """ResNet-50 v1 in the layer order / naming of ``keras.applications.ResNet50`` (~25.6 M parameters)."""
    inp = Input(shape=input_shape)
    x = ZeroPadding2D(3, name='conv1_pad')(inp)
    x = Conv2D(64, 7, strides=2, name='conv1_conv')(x)
    x = BatchNormalization(epsilon=1.001e-5, name='conv1_bn')(x)
    x = Activation('relu', name='conv1_relu')(x)
    x = ZeroPadding2D(1, name='pool1_pad')(x)
    x = MaxPooling2D(3, strides=2, name='pool1_pool')(x)
    for stage, (filters, blocks, stride) in enumerate(((64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)), start=2):
        for b in range(blocks):
            x = _bottleneck(x, filters, stride if b == 0 else 1, b == 0, 'conv%d_block%d' % (stage, b + 1))
    x = GlobalAveragePooling2D(name='avg_pool')(x)
    out = Dense(classes, activation='softmax', name='predictions')(x)
    return Model(inputs=inp, outputs=out, name='resnet50')


def lstm_classifier(vocab=20000, maxlen=80, embed=128, units=128, classes=2, bidirectional=False, dropout=0.2):
    """The Keras IMDB example's shape: ``Embedding(mask_zero) -> LSTM (optionally Bidirectional) -> Dense`` over padded
    token sequences (``keras.utils.pad_sequences``; the padding steps are skipped through the mask)."""
    from .layers_extra import LSTM, Bidirectional, Embedding
    rnn = LSTM(units, dropout=dropout, recurrent_dropout=0.0)
    return Sequential([Embedding(vocab, embed, mask_zero=True, input_length=maxlen),
                       Bidirectional(rnn) if bidirectional else rnn,
                       Dense(classes, activation='softmax' if classes > 1 else 'sigmoid')], name='lstm_classifier')


def transformer_classifier(vocab=20000, maxlen=128, embed=64, heads=4, ff=128, blocks=2, classes=2, dropout=0.1):
    """A small pre-softmax transformer encoder for sequence classification (the Keras "text classification with
    Transformer" example): token embedding, ``blocks`` x (self-attention + feed-forward, each with residual and
    LayerNormalization), average pooling, Dense head."""
    from .layers_extra import Embedding, GlobalAveragePooling1D, LayerNormalization
    from .layers_more import MultiHeadAttention
    inp = Input(shape=(maxlen,), name='tokens')
    x = Embedding(vocab, embed, name='token_embedding')(inp)
    for b in range(blocks):
        att = MultiHeadAttention(heads, embed // heads, dropout=dropout, name='block%d_attention' % b)([x, x])
        x = LayerNormalization(epsilon=1e-6, name='block%d_norm1' % b)(Add(name='block%d_add1' % b)([x, att]))
        f = Dense(ff, activation='relu', name='block%d_ff1' % b)(x)
        f = Dropout(dropout, name='block%d_drop' % b)(Dense(embed, name='block%d_ff2' % b)(f))
        x = LayerNormalization(epsilon=1e-6, name='block%d_norm2' % b)(Add(name='block%d_add2' % b)([x, f]))
    x = GlobalAveragePooling1D(name='pool')(x)
    out = Dense(classes, activation='softmax', name='predictions')(x)
    return Model(inputs=inp, outputs=out, name='transformer_classifier')
