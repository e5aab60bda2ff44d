"""``keras.losses`` objects (``models/losses.py``): default-configured ones are the named functions, the others
(``from_logits``, ``label_smoothing``, ``delta``) compute what Keras computes and survive a model file."""
import numpy as np
import pytest
import torch

from elephas_b200.models import Dense, Sequential, load_model, losses
from elephas_b200.models.optimizers import SGD


def test_default_objects_are_the_named_functions():
    assert losses.canonical_name(losses.CategoricalCrossentropy()) == 'categorical_crossentropy'
    assert losses.get(losses.MeanSquaredError()) is losses.mean_squared_error
    assert losses.canonical_name(losses.Huber(delta=2.0)) == 'Huber(delta=2.0)'
    assert losses.canonical_name({'class_name': 'BinaryCrossentropy', 'config': {'from_logits': True}}) \
        == "BinaryCrossentropy(axis=-1, from_logits=True, label_smoothing=0.0)"
    with pytest.raises(TypeError):
        losses.Huber(gamma=1.0)
    with pytest.raises(ValueError):
        losses.MeanSquaredError(reduction='sum')
    with pytest.raises(ValueError):
        losses.deserialize({'class_name': 'Sideways', 'config': {}})


def test_from_logits_and_label_smoothing_values():
    torch.manual_seed(0)
    logits = torch.randn(16, 5)
    onehot = torch.eye(5)[torch.randint(0, 5, (16,))]
    probs = torch.softmax(logits, dim=1)
    a = losses.CategoricalCrossentropy(from_logits=True)(onehot, logits)
    b = losses.categorical_crossentropy(onehot, probs)
    np.testing.assert_allclose(a.numpy(), b.numpy(), rtol=1e-5, atol=1e-6)
    smooth = losses.CategoricalCrossentropy(label_smoothing=0.1)(onehot, probs)
    target = onehot * 0.9 + 0.02
    np.testing.assert_allclose(smooth.numpy(), -(target * torch.log(probs)).sum(1).numpy(), rtol=1e-5)
    idx = onehot.argmax(1).float().unsqueeze(1)
    s = losses.SparseCategoricalCrossentropy(from_logits=True)(idx, logits)
    np.testing.assert_allclose(s.numpy(), a.numpy(), rtol=1e-5, atol=1e-6)
    z = torch.randn(16, 3) * 4
    y = (torch.rand(16, 3) > 0.5).float()
    c = losses.BinaryCrossentropy(from_logits=True)(y, z)
    d = torch.nn.functional.binary_cross_entropy_with_logits(z, y, reduction='none').mean(1)
    np.testing.assert_allclose(c.numpy(), d.numpy(), rtol=1e-5, atol=1e-6)
    e = losses.BinaryCrossentropy(label_smoothing=0.2)(y, torch.sigmoid(z))
    f = torch.nn.functional.binary_cross_entropy(torch.sigmoid(z).clamp(1e-7, 1 - 1e-7), y * 0.8 + 0.1, reduction='none').mean(1)
    np.testing.assert_allclose(e.numpy(), f.numpy(), rtol=1e-4, atol=1e-6)
    h = losses.Huber(delta=0.5)(torch.zeros(4, 1), torch.tensor([[0.2], [-0.2], [2.0], [-3.0]]))
    np.testing.assert_allclose(h.numpy(), [0.02, 0.02, 0.5 * 2.0 - 0.125, 0.5 * 3.0 - 0.125], rtol=1e-6)


def test_a_logits_model_trains_and_its_loss_object_survives_a_model_file(tmp_path):
    rng = np.random.default_rng(0)
    x = rng.standard_normal((256, 6)).astype(np.float32)
    y = np.eye(3, dtype=np.float32)[(x @ rng.standard_normal((6, 3))).argmax(1)]
    m = Sequential([Dense(16, activation='relu', input_shape=(6,)), Dense(3)])       # no softmax: logits out
    m.compile(SGD(learning_rate=0.2), losses.CategoricalCrossentropy(from_logits=True, label_smoothing=0.05), ['acc'])
    h = m.fit(x, y, batch_size=32, epochs=6)
    assert h.history['loss'][-1] < h.history['loss'][0] and h.history['acc'][-1] > 0.8
    path = str(tmp_path / 'm.h5')
    m.save(path)
    back = load_model(path)
    assert isinstance(back.loss, losses.CategoricalCrossentropy)
    assert back.loss.params == {'from_logits': True, 'label_smoothing': 0.05, 'axis': -1}
    assert back.evaluate(x, y, batch_size=64)[0] == pytest.approx(m.evaluate(x, y, batch_size=64)[0], rel=1e-5)
    # a default-configured object compiles to the same thing as the name
    n = Sequential([Dense(3, activation='softmax', input_shape=(6,))])
    n.compile('sgd', losses.CategoricalCrossentropy())
    assert n._loss_name == 'categorical_crossentropy' and n._loss_fn is losses.categorical_crossentropy


def test_accuracy_follows_the_family_of_a_loss_object():
    from elephas_b200.models import metrics
    sparse = Sequential([Dense(3, input_shape=(4,))])
    sparse.compile('sgd', losses.SparseCategoricalCrossentropy(from_logits=True), ['accuracy'])
    assert sparse._metric_fns[0][1] is metrics.sparse_categorical_accuracy
    binary = Sequential([Dense(2, input_shape=(4,))])
    binary.compile('sgd', losses.BinaryCrossentropy(from_logits=True), ['acc'])
    assert binary._metric_fns[0][1] is metrics.binary_accuracy
    rng = np.random.default_rng(0)
    x, labels = rng.standard_normal((64, 4)).astype(np.float32), rng.integers(0, 3, 64)
    loss, acc = sparse.evaluate(x, labels, batch_size=64)
    assert acc == pytest.approx(float((sparse.predict(x).argmax(1) == labels).mean()))
