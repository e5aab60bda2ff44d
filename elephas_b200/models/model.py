"""Keras-compatible ``Model`` / ``Sequential`` on torch.

Everything the reference asks of a Keras model is here with the same names and semantics:
``to_json`` / ``model_from_json`` (``elephas/worker.py:30,87``, ``elephas/spark_model.py:205,239``),
``compile`` with ``loss`` / ``optimizer`` / ``compiled_metrics._metrics``
(``elephas/spark_model.py:49-54``), ``get_weights`` / ``set_weights`` (Keras order and layouts),
``fit`` / ``train_on_batch`` / ``predict`` / ``evaluate`` (``elephas/worker.py:42,108,123``,
``elephas/spark_model.py:248,291``), ``History.history``, ``save`` / ``load_model`` (Keras HDF5
layout, see ``elephas_b200/io``).

Execution: on a CUDA device an MLP-shaped model (Dense / Activation / Dropout chains, the shape of
every workload the reference ships) trains and predicts through the native sm_100a engine
(``elephas_b200/ops/engine.py``: tcgen05 GEMMs with fused epilogues, fused loss, fused optimizer);
every other model, and every model on the CPU, runs through torch autograd.
"""
import json
import os

import numpy as np
import torch

from . import layers as L
from . import losses as _losses
from . import metrics as _metrics
from . import optimizers as _optimizers

_PREDICT_BLOCK = 256
# fit() keywords that are accepted and have no effect here (data-loader knobs of Keras; options of this package's
# distributed paths that travel in the same train_config dictionary)
_FIT_ACCEPTED = frozenset(('workers', 'use_multiprocessing', 'max_queue_size', 'validation_steps', 'validation_batch_size',
                           'compute_dtype', 'input_mode', 'log_batch_loss', 'keep_context', 'timed_from'))
# layers that hand the mask of their input on unchanged (they keep the [batch, time] axes and look at no neighbours)
_MASK_TRANSPARENT = frozenset(('Dropout', 'SpatialDropout1D', 'Activation', 'ReLU', 'LeakyReLU', 'ELU', 'Softmax', 'Dense',
                               'TimeDistributed', 'LayerNormalization', 'BatchNormalization', 'GaussianNoise'))
KERAS_VERSION = '2.10.0'
BACKEND = 'tensorflow'


def _row_weights(y, sample_weight, class_weight):
    """One weight per row from ``sample_weight`` and / or ``class_weight`` (``{class index: weight}``; the class of a row is
    its integer label or the arg-max of its one-hot row), or None."""
    if sample_weight is None and not class_weight:
        return None
    rows = len(y)
    w = np.ones(rows, dtype=np.float32)
    if sample_weight is not None:
        sample_weight = np.asarray(sample_weight, dtype=np.float32).reshape(-1)
        if len(sample_weight) != rows:
            raise ValueError("sample_weight has %d entries for %d rows" % (len(sample_weight), rows))
        w *= sample_weight
    if class_weight:
        labels = y.detach().cpu().numpy() if isinstance(y, torch.Tensor) else np.asarray(y)
        if labels.ndim > 1 and labels.shape[-1] > 1:
            labels = labels.reshape(rows, -1).argmax(axis=1)
        labels = labels.reshape(rows).astype(np.int64)
        table = np.ones(int(max(labels.max(initial=0), max(class_weight))) + 1, dtype=np.float32)
        for k, v in class_weight.items():
            table[int(k)] = float(v)
        w *= table[labels]
    return w


class _Multi(list):
    """The arrays (or tensors) of a multi-input model: ``fit([xa, xb], y)``. Sliced and indexed together."""


def _rows(x):
    return len(x[0]) if isinstance(x, _Multi) else len(x)


def _take(x, idx):
    return _Multi(a[idx] for a in x) if isinstance(x, _Multi) else x[idx]


def _float(out):
    return [o.float() for o in out] if isinstance(out, (list, tuple)) else out.float()


def _head(out, rows):
    return [o.float()[:rows] for o in out] if isinstance(out, (list, tuple)) else out.float()[:rows]


def _join(blocks):
    """Row blocks of predictions -> one array, or one array per output."""
    if isinstance(blocks[0], (list, tuple)):
        return [torch.cat([b[i] for b in blocks], dim=0).cpu().numpy() for i in range(len(blocks[0]))]
    return torch.cat(blocks, dim=0).cpu().numpy()


def _detach(out):
    return [o.detach() for o in out] if isinstance(out, (list, tuple)) else out.detach()


class _PerOutput:
    """The loss or a metric of ONE output of a model with several outputs, as a function of all targets and all outputs -
    the shape every training / evaluation loop of this module works with."""

    def __init__(self, fn, index):
        self.fn, self.index = fn, index
        self.stateful = bool(getattr(fn, 'stateful', False))

    def __call__(self, ys, outs):
        return self.fn(ys[self.index], outs[self.index])

    def reset_state(self):
        self.fn.reset_state()

    def result(self):
        return self.fn.result()


class History:
    """``History.history`` maps metric name -> list of per-epoch values (Keras semantics)."""

    def __init__(self):
        self.history = {}
        self.epoch = []
        self.params, self.model = {}, None

    def _append(self, epoch, logs):
        self.epoch.append(epoch)
        for k, v in logs.items():
            self.history.setdefault(k, []).append(v)


class _CompiledMetrics:
    """Shim for ``model.compiled_metrics._metrics`` read by ``SparkModel.__init__`` (spark_model.py:52)."""

    def __init__(self, metrics):
        self._metrics = metrics


def _as_tensor(a, device, dtype=torch.float32):
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=dtype if a.is_floating_point() else None)
    arr = np.asarray(a)
    if arr.dtype == object:
        arr = np.stack([np.asarray(r, dtype=np.float32) for r in arr])
    t = torch.from_numpy(np.ascontiguousarray(arr))
    if t.is_floating_point() or t.dtype in (torch.int32, torch.int64, torch.uint8, torch.bool, torch.int16, torch.int8):
        t = t.to(dtype)
    return t.to(device)


class Model:
    """Functional-API model: a DAG of layers between ``inputs`` and ``outputs`` (KerasTensors)."""

    _json_class_name = 'Functional'

    def __init__(self, inputs=None, outputs=None, name=None):
        self.name = name or L.unique_name('model')
        self.layers = []
        self._device = torch.device('cpu')
        self._compiled = False
        self.optimizer = None
        self.stop_training = False
        self._engine = None
        self._engine_key = None
        self.custom_objects = {}
        from . import mixed_precision
        self._compute_dtype = mixed_precision.compute_dtype()  # the global policy at construction time, as in Keras
        if inputs is not None:
            self._init_graph(inputs, outputs)

    # ---------------------------------------------------------------- graph
    def _init_graph(self, inputs, outputs):
        self._wopt_cache = None
        self._inputs = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        self._outputs = list(outputs) if isinstance(outputs, (list, tuple)) else [outputs]
        order, seen = [], set()

        def visit(t):
            key = (id(t.layer), t.node_index)
            if key in seen:
                return
            seen.add(key)
            for src in t.layer.inbound[t.node_index]:
                visit(src)
            order.append(t)

        for o in self._outputs:
            visit(o)
        self._node_order = order
        self._fusion_cache = None
        self._pool_pad_cache = None
        self.layers = []
        for t in order:
            if t.layer not in self.layers:
                self.layers.append(t.layer)
        self._register_with_layers()

    def _register_with_layers(self):
        """Weak back references layer -> model: ``layer.get_weights`` / ``set_weights`` reach the engines through them."""
        import weakref
        for layer in self.layers:
            owners = layer.__dict__.setdefault('_owners', [])
            owners[:] = [ref for ref in owners if ref() is not None and ref() is not self] + [weakref.ref(self)]

    def __setstate__(self, state):   # copies and unpickled models own their (copied) layers
        self.__dict__.update(state)
        self._register_with_layers()

    def _fusion_plan(self):
        """Which nodes collapse into their BatchNormalization producer (computed once per graph):

        * BN -> Activation('relu'), the BN output read by nothing else:                  the BN applies the ReLU
        * BN -> Add([other, BN]) -> Activation('relu'), intermediates read by nothing else: the BN adds ``other`` and
          applies the ReLU (the tail of a residual block); ``other`` must be computed before the BN in node order.

        Returns ``{bn node key: (relu, residual key or None)}, {absorbed node key: node key it aliases}``. The fused
        call computes exactly what the separate layers compute; on B200 in bf16 it is one kernel (``ops/cnn.py``)."""
        plan = getattr(self, '_fusion_cache', None)
        if plan is not None:
            return plan
        key_of = lambda t: (id(t.layer), t.node_index)  # noqa: E731
        order = {key_of(t): i for i, t in enumerate(self._node_order)}
        consumers = {}
        for t in self._node_order:
            for s in t.layer.inbound[t.node_index]:
                consumers.setdefault(key_of(s), []).append(t)
        outputs = {key_of(o) for o in self._outputs}

        def only_consumer(k):
            c = consumers.get(k, [])
            return c[0] if len(c) == 1 and k not in outputs else None

        is_relu = lambda t: isinstance(t.layer, L.Activation) and t.layer.activation_name == 'relu'  # noqa: E731
        fused, alias = {}, {}
        for t in self._node_order:
            if not isinstance(t.layer, L.BatchNormalization):
                continue
            k = key_of(t)
            nxt = only_consumer(k)
            if nxt is None:
                continue
            if is_relu(nxt):
                fused[k] = (True, None)
                alias[key_of(nxt)] = k
            elif isinstance(nxt.layer, L.Add) and len(nxt.layer.inbound[nxt.node_index]) == 2:
                srcs = [key_of(s) for s in nxt.layer.inbound[nxt.node_index]]
                other = srcs[0] if srcs[1] == k else srcs[1]
                after = only_consumer(key_of(nxt))
                # this BN must be the later of the two Add inputs, and not the same tensor twice
                if after is not None and is_relu(after) and other != k and order.get(other, -1) < order[k] \
                        and other not in alias:
                    fused[k] = (True, other)
                    alias[key_of(nxt)] = k
                    alias[key_of(after)] = k
        self._fusion_cache = (fused, alias)
        return self._fusion_cache

    def _pool_pad_plan(self):
        """ZeroPadding2D -> MaxPooling2D(padding='valid'), the padded tensor read by nothing else: the pool takes the
        padding as an argument (zeros that take part in the maximum, exactly like the padded copy would) and the copy is
        never made. Returns ``{pool node key: (padding, key of the pad's input)}, {pad node keys that are skipped}``."""
        plan = getattr(self, '_pool_pad_cache', None)
        if plan is not None:
            return plan
        key_of = lambda t: (id(t.layer), t.node_index)  # noqa: E731
        consumers = {}
        for t in self._node_order:
            for s in t.layer.inbound[t.node_index]:
                consumers.setdefault(key_of(s), []).append(t)
        outputs = {key_of(o) for o in self._outputs}
        pools, skipped = {}, set()
        for t in self._node_order:
            ins = t.layer.inbound[t.node_index]
            if not isinstance(t.layer, L.ZeroPadding2D) or len(ins) != 1 or key_of(t) in outputs:
                continue
            users = consumers.get(key_of(t), [])
            if len(users) == 1 and isinstance(users[0].layer, L.MaxPooling2D) and users[0].layer.padding == 'valid':
                pools[key_of(users[0])] = (t.layer.padding, key_of(ins[0]))
                skipped.add(key_of(t))
        self._pool_pad_cache = (pools, skipped)
        return self._pool_pad_cache

    def _forward(self, xs, training):
        values = {}
        for t, x in zip(self._inputs, xs):
            values[(id(t.layer), t.node_index)] = x
        fuse = os.environ.get('ELEPHAS_B200_FUSE_BN', '1') != '0'
        fused, alias = self._fusion_plan() if fuse else ({}, {})
        pools, skipped = self._pool_pad_plan() if fuse else ({}, set())
        self._activity_terms = []
        masks = {}  # node key -> [batch, time] mask (Embedding(mask_zero=True) / Masking), carried to the recurrent layers
        for t in self._node_order:
            key = (id(t.layer), t.node_index)
            if key in values or key in skipped:
                continue
            if key in pools:
                pads, src = pools[key]
                values[key] = t.layer.call(values[src], training=training, zero_pads=pads)
                continue
            if key in alias:
                values[key] = values[alias[key]]  # absorbed into the BatchNormalization that produced its input
                if alias[key] in masks:
                    masks[key] = masks[alias[key]]
                continue
            ins = [values[(id(s.layer), s.node_index)] for s in t.layer.inbound[t.node_index]]
            if key in fused:
                relu, other = fused[key]
                values[key] = t.layer.call(ins[0], training=training, fuse_relu=relu,
                                           residual=values[other] if other is not None else None)
                continue
            srcs = t.layer.inbound[t.node_index]
            mask = masks.get((id(srcs[0].layer), srcs[0].node_index)) if len(srcs) == 1 else None
            if mask is not None and getattr(t.layer, 'accepts_mask', False):
                values[key] = t.layer.call(ins[0], training=training, mask=mask)
            else:
                values[key] = t.layer.call(ins[0] if len(ins) == 1 else ins, training=training)
            if len(srcs) == 1:
                compute = getattr(t.layer, 'compute_mask', None)
                if compute is not None:
                    mask = compute(ins[0], mask)
                elif type(t.layer).__name__ not in _MASK_TRANSPARENT:
                    mask = None
                if mask is not None:
                    masks[key] = mask
            if t.layer.activity_regularizer is not None:   # (noted in evaluation too: Keras reports the penalties there)
                out = values[key]
                self._activity_terms.append(t.layer.activity_regularizer(out) / out.shape[0])
        outs = [values[(id(o.layer), o.node_index)] for o in self._outputs]
        return outs[0] if len(outs) == 1 else outs

    def __call__(self, x, training=False):
        xs = list(x) if isinstance(x, (list, tuple)) else [x]
        return self._forward(xs, training)

    @property
    def built(self):
        return all(l.built for l in self.layers)

    @property
    def inputs(self):
        return list(self._inputs)

    @property
    def outputs(self):
        return list(self._outputs)

    @property
    def input(self):  # noqa: A003 - the Keras name
        return self._inputs[0] if len(self._inputs) == 1 else list(self._inputs)

    @property
    def output(self):
        return self._outputs[0] if len(self._outputs) == 1 else list(self._outputs)

    def build(self, input_shape=None):
        return None

    @property
    def input_shape(self):
        shapes = [t.shape for t in self._inputs]
        return shapes[0] if len(shapes) == 1 else shapes

    @property
    def output_shape(self):
        shapes = [t.shape for t in self._outputs]
        return shapes[0] if len(shapes) == 1 else shapes

    # -------------------------------------------------------------- weights
    def _weight_entries(self):
        for layer in self.layers:
            for idx, (_, w, _) in enumerate(layer._weights):
                yield layer, idx, w

    @property
    def weights(self):
        return [w for _, _, w in self._weight_entries()]

    @property
    def trainable_weights(self):
        return [w for l in self.layers for w in l.trainable_weights]

    @property
    def non_trainable_weights(self):
        return [w for l in self.layers for w in l.non_trainable_weights]

    @property
    def trainable(self):
        return any(l.trainable for l in self.layers)

    @trainable.setter
    def trainable(self, value):
        """``model.trainable = False`` freezes every layer (a trunk inside a larger graph); recompile afterwards."""
        for layer in self._all_layers():
            layer.trainable = value

    def count_params(self):
        return int(sum(w.numel() for w in self.weights))

    # --------------------------------------------- regularizers / constraints
    def _all_layers(self):
        """The layers of the graph and the layers they own (TimeDistributed / Bidirectional)."""
        seen, todo = [], list(self.layers)
        while todo:
            layer = todo.pop()
            if any(layer is s for s in seen):
                continue
            seen.append(layer)
            todo.extend(v for v in vars(layer).values() if isinstance(v, L.Layer))
        return seen

    def _weight_options(self):
        """``(penalties, constraints, activity)``: ``[(tensor, regularizer)]``, ``[(layer, index, tensor, constraint)]`` and
        whether any layer regularises its output. The walk over the layers is cached (``compile`` drops it); the tensors
        are read from the layers at every use, so ``to()`` needs no invalidation."""
        cached = getattr(self, '_wopt_cache', None)
        if cached is None:
            entries, activity = [], False
            for layer in self._all_layers():
                activity = activity or layer.activity_regularizer is not None
                entries.extend((layer, kind, obj, idx) for kind, obj, idx, _ in layer.weight_options())
            cached = self._wopt_cache = (entries, activity)
        penalties, cons = [], []
        for layer, kind, obj, idx in cached[0]:
            tensor = layer._weights[idx][1]
            if kind == 'regularizer':
                if tensor.requires_grad:
                    penalties.append((tensor, obj))
            else:
                cons.append((layer, idx, tensor, obj))
        return penalties, cons, cached[1]

    def has_weight_options(self):
        """True when a layer carries a regularizer or a constraint: the fused native engines do not apply them, such
        models train on the tape path (eagerly on the CPU, as a captured torch step on the GPU)."""
        penalties, cons, activity = self._weight_options()
        return bool(penalties or cons or activity)

    def _regularization_loss(self):
        """Sum of the weight penalties and of the activity penalties noted by the last training forward, or None."""
        penalties, _, _ = self._weight_options()
        terms = [reg(w) for w, reg in penalties] + list(getattr(self, '_activity_terms', ()))
        self._activity_terms = []
        if not terms:
            return None
        total = terms[0]
        for t in terms[1:]:
            total = total + t
        return total

    @torch.no_grad()
    def _apply_constraints(self):
        _, cons, _ = self._weight_options()
        for layer, idx, tensor, con in cons:
            view = layer.to_keras_layout(idx, tensor)   # a view of the weight in the layout `axis` refers to
            view.copy_(con(view))

    def get_weights(self):
        """List of NumPy arrays in Keras order and Keras layouts (Dense kernels are ``(in, out)``)."""
        self._sync_from_engine()
        return [layer.to_keras_layout(idx, w.detach()).cpu().numpy().copy() for layer, idx, w in self._weight_entries()]

    def set_weights(self, weights):
        entries = list(self._weight_entries())
        if len(entries) != len(weights):
            raise ValueError("set_weights got %d arrays for a model with %d weights" % (len(weights), len(entries)))
        with torch.no_grad():
            for (layer, idx, w), value in zip(entries, weights):
                if isinstance(value, torch.Tensor):
                    src = layer.from_keras_layout(idx, value.detach().cpu().numpy())
                else:
                    src = layer.from_keras_layout(idx, np.asarray(value))
                src = torch.from_numpy(np.asarray(src, dtype=np.float32, order='C'))   # (keeps 0-d weights 0-d)
                if tuple(src.shape) != tuple(w.shape):
                    raise ValueError("weight %s expects shape %s, got %s" % (layer.name, tuple(w.shape), tuple(src.shape)))
                w.copy_(src)
        self._push_to_engine()

    def to(self, device):
        """Move the model (weights and optimizer slots) to ``device``; returns self."""
        device = torch.device(device)
        if device == self._device:
            return self
        self._sync_from_engine()
        # every weight tensor once (wrappers list the tensors of the layers they own under their own name), then every
        # reference to it: the weight lists and whatever attribute a layer keeps it under (kernel, recurrent_kernel, ...)
        everyone, todo = [], list(self.layers)
        while todo:
            layer = todo.pop()
            everyone.append(layer)
            todo.extend(v for v in vars(layer).values() if isinstance(v, L.Layer))
        moved = {}
        for layer in everyone:
            for _, w, _ in layer._weights:
                if id(w) not in moved:
                    moved[id(w)] = (w, w.detach().to(device).requires_grad_(w.requires_grad))
        for layer in everyone:
            layer._weights = [(n, moved[id(w)][1], t) for n, w, t in layer._weights]
            for attr, value in list(vars(layer).items()):
                if isinstance(value, torch.Tensor) and id(value) in moved:
                    setattr(layer, attr, moved[id(value)][1])
        if self.optimizer is not None:
            self.optimizer._slots = {k: v.to(device) for k, v in self.optimizer._slots.items()}
        self._device = device
        self._engine = None
        self._infer_engine = None
        self._torch_step = None
        self._torch_engine_failed = False
        return self

    @property
    def device(self):
        return self._device

    # ---------------------------------------------------------- native engine
    def _native_engine(self, batch_size=None):
        """The sm_100a engine for this model, or None when the model/device does not qualify."""
        if self._device.type != 'cuda' or not self._compiled:
            return None
        from ..ops import engine as _engine_mod
        key = (self._device, id(self.optimizer))
        if self._engine is None or self._engine_key != key:
            self._engine = _engine_mod.maybe_create(self)
            self._engine_key = key
        return self._engine

    def _torch_engine(self):
        """Graph-captured torch step (flat weights + native optimizer kernel) for models without a native engine."""
        if self._device.type != 'cuda' or not self._compiled or getattr(self, '_torch_engine_failed', False):
            return None
        if len(self._inputs) != 1 or len(self._outputs) != 1:
            return None
        te = getattr(self, '_torch_step', None)
        key = (self._device, id(self.optimizer))
        if te is not None and (getattr(self, '_torch_step_key', None) != key or not te.still_valid()):
            te = self._torch_step = None
        if te is None:
            from ..ops import torch_step as _ts
            te = _ts.maybe_create(self)
            if te is None:
                self._torch_engine_failed = True
                return None
            self._torch_step, self._torch_step_key = te, key
        return te

    def _graphed_step(self, xb, yb):
        """Run one training step through the captured graph; False when the model cannot be captured (the eager
        path then takes over for good)."""
        te = self._torch_engine()
        if te is None:
            return False
        try:
            te.train_step(xb, yb)
            return True
        except Exception as exc:  # a layer that synchronises or allocates in a way capture forbids
            import warnings
            warnings.warn("CUDA-graph capture of the training step failed (%s: %s); continuing eagerly"
                          % (type(exc).__name__, exc))
            self._torch_engine_failed = True
            self._torch_step = None
            for p in self.trainable_weights:
                p.grad = None
            return False

    def _sync_from_engine(self):
        if self._engine is not None:
            self._engine.export_weights()
        te = getattr(self, '_torch_step', None)
        if te is not None:
            te.export_optimizer_state()

    def _push_to_engine(self):
        for eng in (self._engine, getattr(self, '_infer_engine', None)):
            if eng is not None:
                eng.import_weights()

    # ---------------------------------------------------------------- config
    def get_config(self):
        names = {}
        layer_cfgs = []
        for layer in self.layers:
            inbound = []
            for node in layer.inbound:
                if node:
                    inbound.append([[s.layer.name, s.node_index, s.tensor_index, {}] for s in node])
            layer_cfgs.append({'class_name': type(layer).__name__, 'config': layer.get_config(), 'name': layer.name,
                               'inbound_nodes': inbound})
            names[layer.name] = layer
        return {'name': self.name, 'layers': layer_cfgs,
                'input_layers': [[t.layer.name, t.node_index, t.tensor_index] for t in self._inputs],
                'output_layers': [[t.layer.name, t.node_index, t.tensor_index] for t in self._outputs]}

    def to_json(self, **kwargs):
        return json.dumps({'class_name': self._json_class_name, 'config': self.get_config(),
                           'keras_version': KERAS_VERSION, 'backend': BACKEND}, **kwargs)

    @classmethod
    def from_config(cls, config, custom_objects=None):
        created = {}
        tensors = {}
        pending = list(config['layers'])
        for spec in pending:
            created[spec['name']] = L.deserialize_layer(spec, custom_objects)
        progress = True
        done = set()
        while len(done) < len(pending) and progress:
            progress = False
            for spec in pending:
                name = spec['name']
                if name in done:
                    continue
                layer = created[name]
                if isinstance(layer, L.InputLayer):
                    t = L.KerasTensor(layer.batch_input_shape, layer)
                    layer.inbound.append([])
                    layer.outputs.append(t)
                    tensors[(name, 0)] = t
                    done.add(name)
                    progress = True
                    continue
                ready = all((src[0], src[1]) in tensors for node in spec['inbound_nodes'] for src in node)
                if not ready:
                    continue
                for node_idx, node in enumerate(spec['inbound_nodes']):
                    ins = [tensors[(src[0], src[1])] for src in node]
                    tensors[(name, node_idx)] = layer(ins if len(ins) > 1 else ins[0])
                done.add(name)
                progress = True
        if len(done) < len(pending):
            raise ValueError("model config has unresolved inbound nodes")
        inputs = [tensors[(n, i)] for n, i, _ in config['input_layers']]
        outputs = [tensors[(n, i)] for n, i, _ in config['output_layers']]
        model = cls(inputs=inputs if len(inputs) > 1 else inputs[0],
                    outputs=outputs if len(outputs) > 1 else outputs[0], name=config.get('name'))
        model.custom_objects = dict(custom_objects or {})
        return model

    # --------------------------------------------------------------- compile
    def compile(self, optimizer='rmsprop', loss=None, metrics=None, **kwargs):
        new_opt = _optimizers.get(optimizer)
        metrics_arg = metrics
        metrics = list(metrics) if metrics else []
        signature = (json.dumps(_optimizers.serialize(new_opt), sort_keys=True, default=str),
                     _losses.canonical_name(loss) if not isinstance(loss, (list, tuple, dict)) or
                     (isinstance(loss, dict) and 'class_name' in loss) else repr(loss),
                     tuple(str(m) for m in metrics) if not isinstance(metrics_arg, dict) else repr(metrics_arg),
                     repr(kwargs.get('loss_weights')))
        # (the device is not part of the signature: ``to()`` drops the engines itself, and a model that was compiled on
        # the host and moved to the GPU by its first fit would otherwise rebuild engine and graphs in its second fit)
        unchanged = self._compiled and getattr(self, '_compile_signature', None) == signature
        same = unchanged and self._engine is not None
        te = getattr(self, '_torch_step', None)
        if te is not None:
            if unchanged and te.still_valid():
                # identical configuration: keep the captured step (flat buffers, CUDA graphs), start from fresh slots
                te.reset_optimizer()
                self._torch_step_key = (self._device, id(new_opt))
            else:
                te.export_optimizer_state()  # the captured step's slots belong to the outgoing optimizer
                self._torch_step = None
        self._torch_engine_failed = False
        self._wopt_cache = None
        self.optimizer = new_opt
        self.loss = loss
        self.loss_weights = kwargs.get('loss_weights')
        self.compiled_metrics = _CompiledMetrics((metrics_arg if isinstance(metrics_arg, dict) else list(metrics)) if metrics else None)
        if len(self._outputs) > 1:
            self._compile_outputs(loss, metrics_arg or [], self.loss_weights)
        else:
            if isinstance(loss, (list, tuple)) and len(loss) == 1:
                loss = self.loss = loss[0]
            self._loss_name = _losses.canonical_name(loss)
            self._loss_fn = _losses.get(loss, self.custom_objects)
            out_shape = self.output_shape
            out_dim = out_shape[-1] if out_shape and not isinstance(out_shape[0], (tuple, list)) else None
            family = _losses.family_name(loss)
            self._metric_fns = [_metrics.resolve(m, family, out_dim, self.custom_objects) for m in metrics]
        self.metrics_names = ['loss'] + [n for n, _ in self._metric_fns]
        self._compiled = True
        self._compile_signature = signature
        if same:
            # identical configuration: keep the native engine (buffers, CUDA graphs), start from fresh slots
            self._engine.reset_optimizer(new_opt)
            self._engine_key = (self._device, id(self.optimizer))
            return
        self._sync_from_engine()
        self._engine = None
        self._infer_engine = None

    def _compile_outputs(self, loss, metrics, loss_weights):
        """Several outputs: one loss per output (one name for all, a list in output order, or a dict by output-layer name),
        optional ``loss_weights``, metrics for all outputs or per output. The total loss is the weighted sum; the reported
        names follow Keras: ``loss``, ``<output>_loss`` ..., ``<output>_<metric>`` ..."""
        names = [t.layer.name for t in self._outputs]

        def per_output(value, what, default=None):
            if isinstance(value, dict) and 'class_name' not in value:
                unknown = set(value) - set(names)
                if unknown:
                    raise ValueError("%s refers to unknown outputs %s (outputs: %s)" % (what, sorted(unknown), names))
                return [value.get(n, default) for n in names]
            if isinstance(value, (list, tuple)):
                if len(value) != len(names):
                    raise ValueError("%s has %d entries for %d outputs" % (what, len(value), len(names)))
                return list(value)
            return [value] * len(names)

        losses = per_output(loss, 'loss')
        if any(l is None for l in losses):
            raise ValueError("every output needs a loss (outputs: %s)" % names)
        weights = [1.0 if w is None else float(w) for w in per_output(loss_weights, 'loss_weights', 1.0)]
        parts = [_PerOutput(_losses.get(l, self.custom_objects), i) for i, l in enumerate(losses)]

        def total(ys, outs):
            value = weights[0] * parts[0](ys, outs)
            for w, part in zip(weights[1:], parts[1:]):
                value = value + w * part(ys, outs)
            return value

        self._loss_fn, self._loss_name = total, 'weighted_sum(%s)' % ', '.join(_losses.canonical_name(l) for l in losses)
        fns = [(n + '_loss', part) for n, part in zip(names, parts)]
        if isinstance(metrics, dict) or (metrics and all(isinstance(m, (list, tuple)) for m in metrics)):
            chosen = per_output(metrics, 'metrics', [])
        else:
            chosen = [list(metrics)] * len(names)
        for i, (n, ms) in enumerate(zip(names, chosen)):
            ms = [ms] if isinstance(ms, str) or callable(ms) else list(ms or [])
            shape = self._outputs[i].shape
            for m in ms:
                label, fn = _metrics.resolve(m, _losses.family_name(losses[i]), shape[-1], self.custom_objects)
                fns.append(('%s_%s' % (n, label), _PerOutput(fn, i)))
        self._metric_fns = fns

    def _host_y(self, y):
        """Targets: one array, or for several outputs a list in output order / a dict by output-layer name."""
        if len(self._outputs) == 1:
            if isinstance(y, (list, tuple)) and len(y) == 1 and isinstance(y[0], (np.ndarray, torch.Tensor)) \
                    and y[0].ndim == len(self._outputs[0].shape):
                y = y[0]   # [array] for the single output (a list holding ONE ROW stays a batch of one row)
            return y if isinstance(y, torch.Tensor) else np.asarray(y)
        if isinstance(y, dict):
            y = [y[t.layer.name] for t in self._outputs]
        if not isinstance(y, (list, tuple)) or len(y) != len(self._outputs):
            raise ValueError("this model has %d outputs: pass a list of %d target arrays" % (len(self._outputs), len(self._outputs)))
        return _Multi(a if isinstance(a, torch.Tensor) else np.asarray(a) for a in y)

    def _device_y(self, y):
        return _Multi(_as_tensor(a, self._device) for a in y) if isinstance(y, _Multi) else _as_tensor(y, self._device)

    def _require_compiled(self):
        if not self._compiled:
            raise RuntimeError("You must compile your model before training/testing. Use `model.compile(optimizer, loss)`.")

    # ------------------------------------------------------------ train step
    def _fused_softmax_ce(self):
        """True when the model ends in softmax and the loss is categorical CE: the gradient is then
        taken through the fused log-softmax form (the same simplification Keras applies)."""
        last = self.layers[-1] if self.layers else None
        ends_softmax = (isinstance(last, (L.Dense, L.Activation)) and last.activation_name == 'softmax') or \
            isinstance(last, L.Softmax)
        return ends_softmax and self._loss_name == 'categorical_crossentropy'

    def set_compute_dtype(self, dtype):
        """Compute precision of models outside the native Dense engine (convolutional models, functional graphs) on CUDA
        devices: ``'bfloat16'`` runs forward / backward in bf16 with fp32 master weights and an fp32 loss (what the Dense
        engine always does) - and puts every Conv2D / BatchNormalization / pooling / Dense layer on the native sm_100a
        kernels of ``ops/cnn.py``; ``'float32'`` (default) keeps the plain torch ops. No counterpart in the reference."""
        if dtype not in ('float32', 'bfloat16'):
            raise ValueError("compute dtype must be 'float32' or 'bfloat16', got %r" % (dtype,))
        self._compute_dtype = dtype

    def _autocast(self, xb):
        xb = xb[0] if isinstance(xb, _Multi) else xb
        on = getattr(self, '_compute_dtype', 'float32') == 'bfloat16' and xb.is_cuda
        return torch.autocast('cuda', dtype=torch.bfloat16, enabled=on)

    def _cpu_step(self):
        """Hand-written forward / backward of Dense chains on the CPU (``ops/cpu_mlp.py``), or None (autograd tape)."""
        if self._device.type != 'cpu':
            return None
        key = (id(self.optimizer), getattr(self, '_compile_signature', None), len(self.layers))
        cached = getattr(self, '_cpu_step_cache', None)
        if cached is None or cached[0] != key:
            from ..ops import cpu_mlp
            cached = self._cpu_step_cache = (key, cpu_mlp.maybe_create(self))
        return cached[1]

    def _torch_train_step(self, xb, yb, wb=None):
        step = self._cpu_step() if wb is None else None
        if step is not None:
            return step(xb, yb)
        params = self.trainable_weights
        for p in params:
            p.grad = None
        with self._autocast(xb):
            out = self(xb, training=True)
        out = _float(out)
        per_sample = self._loss_fn(yb, out)
        if wb is not None:  # Keras' sum-over-batch-size reduction: sum(w * loss) / rows
            per_sample = per_sample * wb.reshape(per_sample.shape[0], *([1] * (per_sample.dim() - 1)))
        loss = per_sample.mean()
        penalty = self._regularization_loss()
        if penalty is not None:
            loss = loss + penalty
        loss.backward()
        grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in params]
        self.optimizer.apply_gradients(params, grads)
        self._apply_constraints()
        with torch.no_grad():
            stats = [float(loss.detach())]
            for _, fn in self._metric_fns:
                stats.append(float(fn(yb, _detach(out)).mean()))
        return stats

    def _tune_cpu_threads(self, rows):
        """CPU only: intra-op thread count fitted to this model's layer sizes (``utils/cpu_threads.py``: the small layers of
        the reference's workloads run many times faster on one thread than on a pool of all cores)."""
        if self._device.type == 'cpu':
            from ..utils import cpu_threads
            cpu_threads.tune_for(self, rows)

    def _host_x(self, x):
        """The ``x`` argument of fit / evaluate / predict: one array, or - for a model with several inputs - a list of
        arrays (or a dict keyed by input-layer name) in the order of ``Model(inputs=[...])``."""
        if len(self._inputs) == 1:
            if isinstance(x, (list, tuple)) and len(x) == 1 and isinstance(x[0], (np.ndarray, torch.Tensor)) \
                    and x[0].ndim == len(self._inputs[0].shape):
                x = x[0]   # Keras also takes [array] for the single input
            return x if isinstance(x, torch.Tensor) else np.asarray(x)
        if isinstance(x, dict):
            x = [x[t.layer.name] for t in self._inputs]
        if not isinstance(x, (list, tuple)) or len(x) != len(self._inputs):
            raise ValueError("this model has %d inputs: pass a list of %d arrays" % (len(self._inputs), len(self._inputs)))
        arrays = _Multi(a if isinstance(a, torch.Tensor) else np.asarray(a) for a in x)
        if len({len(a) for a in arrays}) != 1:
            raise ValueError("the inputs have different numbers of rows: %s" % [len(a) for a in arrays])
        return arrays

    def _device_x(self, x):
        if isinstance(x, _Multi):
            return _Multi(_as_tensor(a, self._device) for a in x)
        xt = _as_tensor(x, self._device)
        return xt.unsqueeze(-1) if xt.dim() == 1 else xt

    def train_on_batch(self, x, y, sample_weight=None, class_weight=None, return_dict=False, **kwargs):
        self._require_compiled()
        row_weight = _row_weights(y, sample_weight, class_weight)
        eng = self._native_engine() if row_weight is None else None
        if eng is not None:
            stats = eng.train_on_batch(x, y)
        else:
            xb = self._device_x(self._host_x(x))
            yb = self._device_y(self._host_y(y))
            self._tune_cpu_threads(_rows(xb))
            te = self._torch_engine() if row_weight is None else None
            if te is not None:
                te.stats.zero_()
            if te is not None and self._graphed_step(xb, yb):
                stats, _ = te.read_stats()
                self.optimizer.iterations = int(te.step.item())
            else:
                wb = torch.as_tensor(row_weight, device=self._device) if row_weight is not None else None
                stats = self._torch_train_step(xb, yb, wb)
        if return_dict:
            return dict(zip(self.metrics_names, stats))
        return stats if len(stats) > 1 else stats[0]

    def test_on_batch(self, x, y, return_dict=False, **kwargs):
        """Loss and metrics of one batch, no update."""
        return self.evaluate(x, y, batch_size=max(len(y), 1), return_dict=return_dict)

    def predict_on_batch(self, x):
        return self.predict(x)

    def get_layer(self, name=None, index=None):
        if index is not None:
            layers = self.layers
            if not -len(layers) <= index < len(layers):
                raise ValueError("Was asked to retrieve layer at index %d but model only has %d layers" % (index, len(layers)))
            return layers[index]
        for layer in self.layers:
            if layer.name == name:
                return layer
        raise ValueError("No such layer: %s" % (name,))

    def fit(self, x=None, y=None, batch_size=None, epochs=1, verbose=0, callbacks=None, validation_split=0.0,
            validation_data=None, shuffle=True, class_weight=None, sample_weight=None, initial_epoch=0, **kwargs):
        self._require_compiled()
        if 'nb_epoch' in kwargs:
            epochs = kwargs.pop('nb_epoch')
        steps_per_epoch = kwargs.pop('steps_per_epoch', None)
        validation_freq = kwargs.pop('validation_freq', 1)
        unknown = set(kwargs) - _FIT_ACCEPTED
        if unknown:
            raise TypeError("fit() got unexpected keyword argument(s) %s" % sorted(unknown))
        batch_size = int(batch_size or 32)
        x = self._host_x(x)
        y = self._host_y(y)
        n_total = _rows(x)
        if class_weight and isinstance(y, _Multi):
            raise ValueError("class_weight is only supported for models with one output")
        row_weight = _row_weights(y[0] if isinstance(y, _Multi) else y, sample_weight, class_weight)
        x_val = y_val = None
        if validation_data is not None:
            x_val, y_val = validation_data[0], validation_data[1]
        elif validation_split and 0.0 < validation_split < 1.0:
            split_at = int(n_total * (1.0 - validation_split))
            x, x_val = _take(x, slice(None, split_at)), _take(x, slice(split_at, None))
            y, y_val = _take(y, slice(None, split_at)), _take(y, slice(split_at, None))
            row_weight = row_weight[:split_at] if row_weight is not None else None
        history = History()
        from .callbacks import CallbackList
        params = {'epochs': epochs, 'batch_size': batch_size, 'samples': _rows(x), 'verbose': verbose,
                  'steps': steps_per_epoch or -(-_rows(x) // batch_size)}
        history.params, history.model = params, self
        cbs = CallbackList(callbacks, self, params)
        self.stop_training = False
        # weighted rows: the loss of the fused engines has no weight operand, the autograd step takes them
        eng = self._native_engine() if row_weight is None and steps_per_epoch is None and validation_freq == 1 else None
        if eng is not None and not cbs:
            eng.fit(x, y, batch_size, epochs, shuffle, history, x_val, y_val, verbose, initial_epoch)
            return history
        cbs.on_train_begin()
        if eng is not None:  # callbacks see every epoch: one engine call per epoch, graphs re-recorded on an lr change
            for epoch in range(initial_epoch, epochs):
                cbs.on_epoch_begin(epoch)
                self._sync_hyper_parameters()
                eng.fit(x, y, batch_size, epoch + 1, shuffle, history, x_val, y_val, verbose, epoch)
                logs = {k: v[-1] for k, v in history.history.items()}
                cbs.on_epoch_end(epoch, logs)
                for k, v in logs.items():  # what a callback adds to the logs ('lr') goes into the history too
                    if k not in history.history:
                        history.history[k] = [v]
                    elif len(history.history[k]) < len(history.epoch):
                        history.history[k].append(v)
                if self.stop_training:
                    break
            cbs.on_train_end()
            return history
        self._tune_cpu_threads(batch_size)
        xt = self._device_x(x)
        yt = self._device_y(y)
        n = _rows(xt)
        wt = torch.as_tensor(row_weight, dtype=torch.float32, device=self._device) if row_weight is not None else None
        names = self.metrics_names
        for epoch in range(initial_epoch, epochs):
            cbs.on_epoch_begin(epoch)
            self._sync_hyper_parameters()
            self._reset_metrics()
            order = torch.randperm(n, device=self._device) if shuffle else torch.arange(n, device=self._device)
            sums = [0.0] * len(names)
            te = self._torch_engine() if wt is None else None
            if te is not None:
                te.stats.zero_()
            graphed_rows = 0
            seen = n if steps_per_epoch is None else min(n, int(steps_per_epoch) * batch_size)
            for start in range(0, seen, batch_size):
                idx = order[start:start + batch_size]
                if te is not None and self._graphed_step(xt[idx], yt[idx]):
                    graphed_rows += len(idx)  # loss / metric sums stay on the device until the epoch ends
                    continue
                te = None
                stats = self._torch_train_step(_take(xt, idx), _take(yt, idx), wt[idx] if wt is not None else None)
                for i, s in enumerate(stats):
                    sums[i] += s * len(idx)
            if graphed_rows:
                te = self._torch_step
                vals, rows = te.read_stats()
                for i, v in enumerate(vals):
                    sums[i] += v * rows
                self.optimizer.iterations = int(te.step.item())
            logs = {k: (s / seen if seen else 0.0) for k, s in zip(names, sums)}
            logs.update(self._stateful_results())
            validate = (epoch + 1) in validation_freq if isinstance(validation_freq, (list, tuple)) \
                else (epoch + 1) % int(validation_freq) == 0
            if x_val is not None and _rows(y_val) and validate:
                val = self.evaluate(x_val, y_val, batch_size=batch_size, verbose=0)
                val = val if isinstance(val, list) else [val]
                logs.update({'val_' + k: v for k, v in zip(names, val)})
            cbs.on_epoch_end(epoch, logs)
            history._append(epoch, logs)
            if verbose:
                print("Epoch %d/%d - " % (epoch + 1, epochs) + " - ".join("%s: %.4f" % kv for kv in logs.items()))
            if self.stop_training:
                break
        cbs.on_train_end()
        return history

    def _sync_hyper_parameters(self):
        """Captured steps bake the optimizer's hyper-parameters into their kernels' arguments: when one was assigned
        since the graphs were recorded (a learning-rate callback), drop the graphs; the next step records new ones."""
        version = getattr(self.optimizer, '_hyper_version', 0)
        if version == getattr(self, '_hyper_seen', 0):
            return
        self._hyper_seen = version
        for holder in (self._engine, getattr(self, '_torch_step', None)):
            if holder is not None:
                holder.invalidate_graphs()

    # ------------------------------------------------------------- inference
    @torch.no_grad()
    def predict(self, x, batch_size=None, verbose=0, **kwargs):
        batch_size = int(batch_size or 32)
        eng = self._native_engine() if self._compiled else None
        if eng is None and self._device.type == 'cuda':
            from ..ops import engine as _engine_mod
            eng = _engine_mod.maybe_create_inference(self)
        if eng is not None:
            return eng.predict(x)
        xt = self._device_x(self._host_x(x))
        if isinstance(xt, _Multi):
            return self._predict_multi(xt)
        self._tune_cpu_threads(_PREDICT_BLOCK)
        outs = []
        # Fixed-size row blocks (the last one zero-padded): every GEMM the library sees has the same
        # shape, so a row's result does not depend on how the caller batched the rows - the
        # "batched == unbatched, bit for bit" contract of tests/test_ml_model.py:354.
        block = _PREDICT_BLOCK
        n = xt.shape[0]
        for start in range(0, n, block):
            xb = xt[start:start + block]
            if xb.shape[0] < block:
                pad = torch.zeros((block - xb.shape[0],) + tuple(xb.shape[1:]), dtype=xb.dtype, device=xb.device)
                with self._autocast(xb):
                    o = self(torch.cat([xb, pad], dim=0), training=False)
                outs.append(_head(o, xb.shape[0]))
            else:
                with self._autocast(xb):
                    o = self(xb, training=False)
                outs.append(_float(o))
        if not outs:
            return np.zeros((0,) + tuple(d for d in self.output_shape[1:]), dtype=np.float32)
        return _join(outs)

    def _predict_multi(self, xt):
        """Several inputs: the same fixed row blocks as above, every input cut and padded alike."""
        outs, block, n = [], _PREDICT_BLOCK, _rows(xt)
        for start in range(0, n, block):
            xb = _take(xt, slice(start, start + block))
            rows = _rows(xb)
            if rows < block:
                xb = _Multi(torch.cat([a, a.new_zeros((block - rows,) + tuple(a.shape[1:]))], dim=0) for a in xb)
            with self._autocast(xb):
                o = self(xb, training=False)
            outs.append(_head(o, rows))
        if not outs:
            return np.zeros((0,) + tuple(d for d in self.output_shape[1:]), dtype=np.float32)
        return _join(outs)

    @torch.no_grad()
    def evaluate(self, x=None, y=None, batch_size=None, verbose=0, sample_weight=None, return_dict=False, **kwargs):
        stats = self._evaluate(x, y, batch_size, sample_weight)
        if return_dict:
            stats = stats if isinstance(stats, list) else [stats]
            return dict(zip(self.metrics_names, stats))
        return stats

    def _evaluate(self, x, y, batch_size, sample_weight=None):
        self._require_compiled()
        batch_size = int(batch_size or 32)
        eng = self._native_engine() if sample_weight is None else None
        if eng is not None:
            stats = eng.evaluate(x, y)
            return stats if len(stats) > 1 else stats[0]
        xt = self._device_x(self._host_x(x))
        yt = self._device_y(self._host_y(y))
        n = _rows(xt)
        sums = [0.0] * len(self.metrics_names)
        self._reset_metrics()
        weights = None
        if sample_weight is not None:
            weights = torch.as_tensor(_row_weights(yt[0] if isinstance(yt, _Multi) else yt, sample_weight, None),
                                      device=self._device)
        step = max(batch_size, 8192)
        self._tune_cpu_threads(min(step, max(n, 1)))
        for start in range(0, n, step):
            xb, yb = _take(xt, slice(start, start + step)), _take(yt, slice(start, start + step))
            with self._autocast(xb):
                out = self(xb, training=False)
            out = _float(out)
            vals = [self._loss_fn(yb, out)] + [fn(yb, out) for _, fn in self._metric_fns]
            if weights is not None:   # the loss only, as in fit: sum(w * loss) / rows
                wb = weights[start:start + step]
                vals[0] = vals[0] * wb.reshape(vals[0].shape[0], *([1] * (vals[0].dim() - 1)))
            for i, v in enumerate(vals):
                sums[i] += float(v.double().sum())
            penalty = self._regularization_loss()   # weight / activity penalties are part of the reported loss, as in Keras
            if penalty is not None:
                sums[0] += float(penalty) * _rows(xb)
        stats = [s / n if n else 0.0 for s in sums]
        final = self._stateful_results()
        stats = [final.get(name, v) for name, v in zip(self.metrics_names, stats)]
        return stats if len(stats) > 1 else stats[0]

    def _reset_metrics(self):
        for _, fn in self._metric_fns:
            if _metrics.is_stateful(fn):
                fn.reset_state()

    def _stateful_results(self):
        """``{name: result}`` of the metrics that accumulate counts instead of averaging rows (``metrics.Metric``)."""
        return {name: fn.result() for name, fn in self._metric_fns if _metrics.is_stateful(fn)}

    # ----------------------------------------------------------------- io
    def save(self, filepath, overwrite=True, include_optimizer=True):
        from ..io.keras_h5 import save_model
        save_model(self, filepath, overwrite=overwrite, include_optimizer=include_optimizer)

    def save_weights(self, filepath, overwrite=True):
        """Weights only, in the layout of ``keras.Model.save_weights`` (one group per layer at the root of the file)."""
        from ..io.keras_h5 import save_weights
        save_weights(self, filepath, overwrite=overwrite)

    def load_weights(self, filepath, by_name=False, skip_mismatch=False):
        """Weights from a file written by ``save_weights`` or ``save`` (ours or Keras'): by position in the layer list, or
        ``by_name`` (only layers whose names match are loaded - the transfer-learning form)."""
        from ..io.keras_h5 import load_weights
        load_weights(self, filepath, by_name=by_name, skip_mismatch=skip_mismatch)

    def summary(self, line_length=None, positions=None, print_fn=print):
        """The table of ``keras.Model.summary``: layer (type), output shape, parameter count, totals."""
        width = int(line_length or 65)
        cols = [int(width * p) for p in (positions or (0.45, 0.85, 1.0))]
        def row(cells):  # noqa: E306
            line = ''
            for text, end in zip(cells, cols):
                line = (line + str(text))[:end - 1].ljust(end)
            return line.rstrip()
        print_fn('Model: "%s"' % self.name)
        print_fn('_' * width)
        print_fn(row(('Layer (type)', 'Output Shape', 'Param #')))
        print_fn('=' * width)
        trainable = 0
        for layer in self.layers:
            shape = layer.outputs[-1].shape if layer.outputs else getattr(layer, 'batch_input_shape', None)
            if isinstance(layer, L.InputLayer):
                shape = [tuple(shape)]
            count = sum(w.numel() for w in layer.weights)
            trainable += sum(w.numel() for w in layer.trainable_weights)
            print_fn(row(('%s (%s)' % (layer.name, type(layer).__name__), shape, count)))
        total = self.count_params()
        print_fn('=' * width)
        print_fn('Total params: {:,}'.format(total))
        print_fn('Trainable params: {:,}'.format(trainable))
        print_fn('Non-trainable params: {:,}'.format(total - trainable))
        print_fn('_' * width)


class Sequential(Model):
    """Linear stack of layers (``tests/conftest.py:8-28`` builds the reference's fixtures this way)."""

    _json_class_name = 'Sequential'

    def __init__(self, layers=None, name=None):
        super().__init__(name=name or L.unique_name('sequential'))
        self._stack = []
        self._tail = None
        self._inputs = []
        self._outputs = []
        self._node_order = []
        for layer in layers or []:
            self.add(layer)

    def add(self, layer):
        if not self._stack:
            if isinstance(layer, L.InputLayer):
                t = L.KerasTensor(layer.batch_input_shape, layer)
                layer.inbound.append([])
                layer.outputs.append(t)
                self._stack.append(layer)
                self._tail = t
                self._init_graph([t], [t])
                return
            if layer.batch_input_shape is None:
                raise ValueError("The first layer of a Sequential model needs input_shape / input_dim")
            inp = L.InputLayer(batch_input_shape=layer.batch_input_shape, name=layer.name + '_input', dtype=layer.dtype)
            t = L.KerasTensor(inp.batch_input_shape, inp)
            inp.inbound.append([])
            inp.outputs.append(t)
            self._stack.append(inp)
            self._tail = t
        self._tail = layer(self._tail)
        self._stack.append(layer)
        self._init_graph([self._stack[0].outputs[0]], [self._tail])

    def pop(self):
        """Removes the last layer (``keras.Sequential.pop``): the usual first step of re-heading a trained stack."""
        if not self.layers_without_input:
            raise TypeError("There are no layers in the model.")
        layer = self._stack.pop()
        self._tail = layer.inbound.pop()[0]
        layer.outputs.pop()
        self._init_graph([self._stack[0].outputs[0]], [self._tail])
        self._engine = self._infer_engine = self._torch_step = self._wopt_cache = None
        return layer

    @property
    def layers_without_input(self):
        return [l for l in self._stack if not isinstance(l, L.InputLayer)]

    def get_config(self):
        return {'name': self.name,
                'layers': [{'class_name': type(l).__name__, 'config': l.get_config()} for l in self._stack]}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        model = cls(name=config.get('name'))
        model.custom_objects = dict(custom_objects or {})
        specs = config['layers'] if isinstance(config, dict) else config
        for spec in specs:
            model.add(L.deserialize_layer(spec, custom_objects))
        return model


def model_from_json(json_string, custom_objects=None):
    """``keras.models.model_from_json``: rebuild an (uncompiled, freshly initialised) model."""
    spec = json.loads(json_string) if isinstance(json_string, (str, bytes)) else json_string
    return model_from_config(spec, custom_objects)


_KERAS3_ONLY_KEYS = frozenset(('lora_rank', 'lora_alpha', 'synchronized', 'optional', 'quantization_config'))


def _keras3_tensors(node):
    """The ``__keras_tensor__`` references inside one Keras 3 inbound node, in call order: positional arguments first
    (nested lists flattened), then the keyword arguments a layer is called with (``value`` / ``key`` of attention)."""
    found = []

    def walk(item):
        if isinstance(item, dict):
            if item.get('class_name') == '__keras_tensor__':
                name, node_index, tensor_index = item['config']['keras_history']
                found.append([name, int(node_index), int(tensor_index), {}])
            else:
                for value in item.values():
                    walk(value)
        elif isinstance(item, (list, tuple)):
            for value in item:
                walk(value)

    walk(node.get('args', []))
    kwargs = node.get('kwargs', {}) or {}
    for key in ('value', 'key'):
        walk(kwargs.get(key))
    return found


def _from_keras3(spec):
    """A model config in the Keras 3 dialect (``module`` / ``registered_name`` entries, ``DTypePolicy`` dictionaries,
    ``batch_shape``, inbound nodes as ``{"args", "kwargs"}``) rewritten into the Keras 2.10 dialect this package reads."""
    import copy
    spec = copy.deepcopy(spec)

    def clean(item):
        if isinstance(item, dict):
            item.pop('module', None), item.pop('registered_name', None)
            for value in item.values():
                clean(value)
        elif isinstance(item, list):
            for value in item:
                clean(value)

    clean(spec)
    cfg = spec['config']
    cfg.pop('dtype', None), cfg.pop('trainable', None), cfg.pop('build_input_shape', None)
    for layer in cfg.get('layers', []):
        lc = layer['config']
        layer.pop('build_config', None)
        layer.setdefault('name', lc.get('name'))
        if isinstance(lc.get('dtype'), dict):
            lc['dtype'] = lc['dtype'].get('config', {}).get('name', 'float32')
        if 'batch_shape' in lc:
            lc['batch_input_shape'] = lc.pop('batch_shape')
        for key in _KERAS3_ONLY_KEYS & set(lc):
            lc.pop(key)
        for inner in ('layer', 'backward_layer'):   # wrappers carry a full layer spec
            if isinstance(lc.get(inner), dict) and isinstance(lc[inner].get('config'), dict):
                sub = {'config': {'layers': [lc[inner]]}}
                lc[inner] = _from_keras3(sub)['config']['layers'][0]
        nodes = layer.get('inbound_nodes')
        if nodes and isinstance(nodes[0], dict):
            layer['inbound_nodes'] = [_keras3_tensors(node) for node in nodes]
    return spec


def _is_keras3(spec):
    if 'module' in spec or 'registered_name' in spec:
        return True
    layers = spec.get('config', {}).get('layers', []) if isinstance(spec.get('config'), dict) else []
    return any('module' in l or isinstance(l.get('config', {}).get('dtype'), dict) for l in layers)


def model_from_config(spec, custom_objects=None):
    if _is_keras3(spec):
        spec = _from_keras3(spec)
    cls_name = spec.get('class_name')
    if cls_name == 'Sequential':
        return Sequential.from_config(spec['config'], custom_objects)
    if cls_name in ('Functional', 'Model'):
        return Model.from_config(spec['config'], custom_objects)
    raise ValueError("Unknown model class %r" % (cls_name,))


def load_model(filepath, custom_objects=None, compile=True):
    from ..io.keras_h5 import load_model as _load
    return _load(filepath, custom_objects=custom_objects, compile=compile)


def to_categorical(y, num_classes=None, dtype='float32'):
    """``keras.utils.to_categorical``."""
    y = np.asarray(y, dtype='int64').ravel()
    if num_classes is None:
        num_classes = int(y.max()) + 1 if y.size else 0
    out = np.zeros((y.shape[0], num_classes), dtype=dtype)
    out[np.arange(y.shape[0]), y] = 1
    return out
