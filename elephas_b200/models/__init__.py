"""Keras-compatible model layer on torch plus the built-in workload zoo.

``from elephas_b200.models import Sequential, Dense, ...`` plays the role that
``tensorflow.keras`` plays for the reference (which has no model code of its own).
"""
from . import (activations, constraints, losses, metrics, mixed_precision, optimizers, regularizers,  # noqa: F401
               schedules)
from .layers import (Activation, Add, AveragePooling2D, BatchNormalization, Concatenate, Conv2D, Dense,  # noqa: F401
                     Dropout, Flatten, GlobalAveragePooling2D, Input, InputLayer, Layer, MaxPooling2D, ReLU,
                     Reshape, Softmax, ZeroPadding2D)
from .layers_extra import (ELU, GRU, LSTM, Average, AveragePooling1D, Bidirectional, Conv1D, Conv2DTranspose,  # noqa: F401
                           DepthwiseConv2D, Dot, Embedding, GaussianNoise, GlobalAveragePooling1D, GlobalMaxPooling1D,
                           GlobalMaxPooling2D, LayerNormalization, LeakyReLU, Masking, Maximum, MaxPooling1D, Minimum, Multiply,
                           Permute, RepeatVector, SeparableConv2D, SimpleRNN, SpatialDropout1D, SpatialDropout2D,
                           Subtract, TimeDistributed, UpSampling2D)
from . import callbacks  # noqa: F401
from .layers_more import (AlphaDropout, Attention, Cropping1D, Cropping2D, GaussianDropout, Lambda,  # noqa: F401
                          MultiHeadAttention, Normalization, PReLU, Rescaling, ThresholdedReLU, UpSampling1D, ZeroPadding1D)
from .model import (History, Model, Sequential, load_model, model_from_config, model_from_json,  # noqa: F401
                    to_categorical)
from .optimizers import SGD, Adadelta, Adagrad, Adam, Adamax, AdamW, Nadam, RMSprop  # noqa: F401
from . import utils, zoo  # noqa: F401
