"""``keras.initializers``: initializer objects are their Keras config dictionaries (``{"class_name", "config"}``); the
classes below build those, ``kernel_initializer=HeNormal(seed=1)`` works like the string names do."""
from ..models.layers import _init_config


def _make(class_name):
    def build(**config):
        spec = _init_config(class_name, class_name)
        unknown = set(config) - set(spec['config'])
        if unknown:
            raise TypeError("%s got unexpected argument(s) %s" % (class_name, sorted(unknown)))
        spec['config'].update(config)
        return spec
    build.__name__ = class_name
    build.__doc__ = "Config dictionary of ``keras.initializers.%s``." % class_name
    return build


Zeros, Ones, Constant = _make('Zeros'), _make('Ones'), _make('Constant')
RandomNormal, RandomUniform, TruncatedNormal = _make('RandomNormal'), _make('RandomUniform'), _make('TruncatedNormal')
GlorotUniform, GlorotNormal = _make('GlorotUniform'), _make('GlorotNormal')
HeUniform, HeNormal = _make('HeUniform'), _make('HeNormal')
LecunUniform, LecunNormal = _make('LecunUniform'), _make('LecunNormal')
VarianceScaling, Orthogonal, Identity = _make('VarianceScaling'), _make('Orthogonal'), _make('Identity')
