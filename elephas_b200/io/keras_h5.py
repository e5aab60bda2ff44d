"""Keras-HDF5-compatible model files.

Reference: ``SparkModel.save`` = ``keras.Model.save(file)`` + one extra root attribute
``distributed_config`` (``elephas/spark_model.py:115-125``); ``load_spark_model`` =
``keras.models.load_model`` + reading that attribute back (``:376-381``). The layout reproduced here
is the Keras 2.10 HDF5 format:

* root attributes ``keras_version``, ``backend``, ``model_config`` (the ``to_json()`` payload),
  ``training_config`` (JSON with ``loss``, ``metrics``, ``optimizer_config``);
* group ``model_weights`` with attributes ``layer_names``, ``backend``, ``keras_version`` and one
  sub-group per layer carrying ``weight_names`` and datasets such as ``dense/kernel:0`` (shape
  ``(in, out)``) and ``dense/bias:0``;
* optional group ``optimizer_weights`` (iteration counter and slot variables).
"""
import json
import os

import numpy as np

from . import hdf5
from ..models import model as _model_mod
from ..models import losses as _losses
from ..models import optimizers as _optimizers


def _training_config(model):
    def one_metric(m):
        return m if isinstance(m, str) else getattr(m, '__name__', str(m))

    def metric_names(ms):
        if isinstance(ms, dict):
            return {k: metric_names(v) for k, v in ms.items()}
        if isinstance(ms, (list, tuple)):
            return [metric_names(m) if isinstance(m, (list, tuple)) else one_metric(m) for m in ms]
        return one_metric(ms)

    loss = model.loss
    if isinstance(loss, (list, tuple)):
        loss = [_losses.serialize(l) for l in loss]
    elif isinstance(loss, dict) and 'class_name' not in loss:
        loss = {k: _losses.serialize(v) for k, v in loss.items()}
    else:
        loss = _losses.serialize(loss)
    metrics = model.compiled_metrics._metrics
    return {'loss': loss, 'metrics': metric_names(metrics) if metrics else [], 'weighted_metrics': None,
            'loss_weights': getattr(model, 'loss_weights', None), 'optimizer_config': _optimizers.serialize(model.optimizer)}


def _optimizer_weight_entries(model):
    opt = model.optimizer
    names = ['%s/iter:0' % opt._class_name]
    values = [np.asarray(opt.iterations, dtype=np.int64)]
    params = model.trainable_weights
    layer_of = {}
    for layer in model.layers:
        for widx, (n, w, _) in enumerate(layer._weights):
            layer_of[id(w)] = ('%s/%s' % (layer.name, n), layer, widx)
    for (idx, slot), tensor in sorted(opt._slots.items(), key=lambda kv: (kv[0][0], kv[0][1])):
        if idx < len(params):
            base, layer, widx = layer_of.get(id(params[idx]), ('param_%d' % idx, None, 0))
            names.append('%s/%s/%s:0' % (opt._class_name, base, slot))
            value = tensor.detach()
            if layer is not None and tuple(value.shape) == tuple(params[idx].shape):
                # a slot has the shape of its weight: stored in the Keras layout, exactly like the weight itself
                # (Dense kernels (in, out), Conv2D kernels HWIO), so the file means the same thing to Keras
                value = layer.to_keras_layout(widx, value)
            values.append(np.ascontiguousarray(value.cpu().numpy()))
    return names, values


def build_tree(model, include_optimizer=True, extra_root_attrs=None):
    root = hdf5.Group()
    root.attrs['keras_version'] = _model_mod.KERAS_VERSION.encode('utf8')
    root.attrs['backend'] = _model_mod.BACKEND.encode('utf8')
    root.attrs['model_config'] = json.dumps(
        {'class_name': model._json_class_name, 'config': model.get_config()}).encode('utf8')
    compiled = getattr(model, '_compiled', False)
    if compiled:
        root.attrs['training_config'] = json.dumps(_training_config(model)).encode('utf8')
    _write_weights(model, root.create_group('model_weights'))
    if include_optimizer and compiled and model.optimizer is not None:
        ow = root.create_group('optimizer_weights')
        names, values = _optimizer_weight_entries(model)
        ow.attrs['weight_names'] = np.array([n.encode('utf8') for n in names])
        for n, v in zip(names, values):
            ow.create_dataset(n, v)
    for k, v in (extra_root_attrs or {}).items():
        root.attrs[k] = v
    return root


def _write_weights(model, mw):
    """The weight groups of ``model`` below ``mw`` (``model_weights`` of a full file, the root of a weights-only file)."""
    layers = [l for l in model.layers]
    mw.attrs['layer_names'] = np.array([l.name.encode('utf8') for l in layers]) if layers else np.zeros((0,))
    mw.attrs['backend'] = _model_mod.BACKEND.encode('utf8')
    mw.attrs['keras_version'] = _model_mod.KERAS_VERSION.encode('utf8')
    model._sync_from_engine()
    for layer in layers:
        g = mw.create_group(layer.name)
        names = layer.weight_names()
        g.attrs['weight_names'] = np.array([n.encode('utf8') for n in names]) if names else np.zeros((0,))
        for idx, (name, (_, w, _t)) in enumerate(zip(names, layer._weights)):
            value = layer.to_keras_layout(idx, w.detach()).cpu().numpy()
            g.create_dataset(name, np.asarray(value, dtype=np.float32, order='C'))  # (ascontiguousarray would turn () into (1,))


def save_weights(model, filepath, overwrite=True):
    filepath = str(filepath)
    if os.path.exists(filepath) and not overwrite:
        raise IOError("%s already exists" % filepath)
    root = hdf5.Group()
    _write_weights(model, root)
    hdf5.save_tree(filepath, root)


def load_weights(model, filepath, by_name=False, skip_mismatch=False):
    root = hdf5.load_tree(str(filepath))
    mw = root['model_weights'] if 'model_weights' in root else root
    saved = []
    for lname in _names(mw.attrs.get('layer_names')):
        g = mw[lname]
        saved.append((lname, [np.asarray(g[wn].data) for wn in _names(g.attrs.get('weight_names'))]))
    current = model.get_weights()
    slots, at = {}, 0
    for layer in model.layers:
        slots[id(layer)] = (at, len(layer._weights))
        at += len(layer._weights)

    def assign(layer, values, origin):
        lo, count = slots[id(layer)]
        bad = len(values) != count or any(tuple(v.shape) != tuple(c.shape) for v, c in zip(values, current[lo:lo + count]))
        if bad:
            if skip_mismatch and by_name:
                return
            raise ValueError("layer %r expects weights of shapes %s, %s has %s"
                             % (layer.name, [tuple(c.shape) for c in current[lo:lo + count]], origin,
                                [tuple(v.shape) for v in values]))
        current[lo:lo + count] = values

    if by_name:
        table = dict(saved)
        for layer in model.layers:
            if layer.name in table and (table[layer.name] or layer._weights):
                assign(layer, table[layer.name], 'the file')
    else:
        with_weights = [(n, v) for n, v in saved if v]
        mine = [l for l in model.layers if l._weights]
        if len(with_weights) != len(mine):
            raise ValueError("the file holds weights of %d layers, the model has %d layers with weights"
                             % (len(with_weights), len(mine)))
        for layer, (lname, values) in zip(mine, with_weights):
            assign(layer, values, 'file layer %r' % lname)
    model.set_weights(current)


def save_model(model, filepath, overwrite=True, include_optimizer=True, extra_root_attrs=None):
    filepath = str(filepath)
    if os.path.exists(filepath) and not overwrite:
        raise IOError("%s already exists" % filepath)
    hdf5.save_tree(filepath, build_tree(model, include_optimizer, extra_root_attrs))


def _text(value):
    if isinstance(value, bytes):
        return value.decode('utf8')
    if isinstance(value, np.ndarray) and value.shape == ():
        return _text(value[()])
    return value


def _names(value):
    if value is None:
        return []
    arr = np.asarray(value)
    if arr.size == 0:
        return []
    return [_text(v) for v in arr.ravel().tolist()]


def read_root_attrs(filepath):
    return hdf5.load_tree(str(filepath)).attrs


def load_model(filepath, custom_objects=None, compile=True):
    root = hdf5.load_tree(str(filepath))
    config = json.loads(_text(root.attrs['model_config']))
    model = _model_mod.model_from_config(config, custom_objects)
    mw = root['model_weights']
    by_layer = {}
    for lname in _names(mw.attrs.get('layer_names')):
        g = mw[lname]
        by_layer[lname] = [np.asarray(g[wn].data) for wn in _names(g.attrs.get('weight_names'))]
    weights = []
    for layer in model.layers:
        vals = by_layer.get(layer.name, [])
        if len(vals) != len(layer._weights):
            raise ValueError("layer %r expects %d weights, file has %d" % (layer.name, len(layer._weights), len(vals)))
        weights.extend(vals)
    model.set_weights(weights)
    if compile and 'training_config' in root.attrs:
        tc = json.loads(_text(root.attrs['training_config']))
        def revive(entry):
            return _losses.deserialize(entry, custom_objects) if isinstance(entry, dict) and 'class_name' in entry else entry

        loss = tc.get('loss')
        if isinstance(loss, list):
            loss = [revive(l) for l in loss]
        elif isinstance(loss, dict) and 'class_name' not in loss:
            loss = {k: revive(v) for k, v in loss.items()}
            if len(model._outputs) == 1:   # Keras writes {output name: loss} for one output too
                loss = next(iter(loss.values()))
        else:
            loss = revive(loss)
        model.compile(optimizer=_optimizers.deserialize(tc['optimizer_config'], custom_objects), loss=loss,
                      metrics=tc.get('metrics') or None, loss_weights=tc.get('loss_weights'))
        if 'optimizer_weights' in root:
            _restore_optimizer(model, root['optimizer_weights'])
    return model


def _restore_optimizer(model, group):
    import torch
    opt = model.optimizer
    params = model.trainable_weights
    index_of = {}
    for layer in model.layers:
        for widx, (n, w, _) in enumerate(layer._weights):
            for i, p in enumerate(params):
                if p is w:
                    index_of['%s/%s' % (layer.name, n)] = (i, layer, widx)
    for name in _names(group.attrs.get('weight_names')):
        value = np.asarray(group[name].data)
        parts = name.rsplit(':', 1)[0].split('/')
        if parts[-1] == 'iter':
            opt.iterations = int(np.asarray(value).reshape(-1)[0])
            continue
        slot = parts[-1]
        key = '/'.join(parts[1:-1])
        if key in index_of:
            i, layer, widx = index_of[key]
            value = np.array(value, dtype=np.float32)
            keras_shape = tuple(layer.to_keras_layout(widx, params[i].detach()).shape)
            if value.shape == keras_shape:
                value = np.ascontiguousarray(layer.from_keras_layout(widx, value))  # file: Keras layout, like the weights
            opt._slots[(i, slot)] = torch.from_numpy(value).to(params[i].device)
