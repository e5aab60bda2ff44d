"""``elephas_b200.keras``: a Keras script with the import prefix changed and nothing else."""
import numpy as np


def test_a_keras_script_with_only_its_import_prefix_changed(tmp_path):
    from elephas_b200.keras import callbacks, initializers, optimizers, regularizers
    from elephas_b200.keras.layers import LSTM, Dense, Dropout, Embedding
    from elephas_b200.keras.losses import CategoricalCrossentropy
    from elephas_b200.keras.metrics import AUC
    from elephas_b200.keras.models import Sequential, clone_model, load_model
    from elephas_b200.keras.preprocessing.sequence import pad_sequences
    from elephas_b200.keras.utils import to_categorical

    rng = np.random.default_rng(0)
    seqs = [rng.integers(1, 20, rng.integers(3, 9)).tolist() for _ in range(200)]
    x = pad_sequences(seqs, maxlen=8)
    y = to_categorical([int(np.mean(s) > 10) for s in seqs], 2)

    model = Sequential()
    model.add(Embedding(20, 8, mask_zero=True, input_length=8))
    model.add(LSTM(12, kernel_initializer=initializers.GlorotUniform(seed=1), kernel_regularizer=regularizers.l2(1e-4)))
    model.add(Dropout(0.1))
    model.add(Dense(2, activation='softmax'))
    model.compile(optimizer=optimizers.Adam(learning_rate=0.02), loss=CategoricalCrossentropy(label_smoothing=0.05),
                  metrics=['accuracy', AUC(name='auc')])
    best = str(tmp_path / 'best.h5')
    history = model.fit(x, y, batch_size=20, epochs=10, validation_split=0.2, verbose=0,
                        callbacks=[callbacks.EarlyStopping(monitor='val_loss', patience=5),
                                   callbacks.ModelCheckpoint(best, save_best_only=True)])
    assert history.history['accuracy'][-1] > 0.8 and history.history['auc'][-1] > 0.85
    restored = load_model(best)
    assert restored.predict(x[:4]).shape == (4, 2)
    twin = clone_model(model)
    assert twin.count_params() == model.count_params() and not np.array_equal(twin.get_weights()[-2], model.get_weights()[-2])


def test_namespace_lists_every_layer_and_initializers_check_their_arguments():
    import pytest

    import elephas_b200.keras as keras
    from elephas_b200.models.layers import LAYER_CLASSES
    assert set(LAYER_CLASSES) <= set(keras.layers.__all__)
    assert keras.layers.Conv2D is LAYER_CLASSES['Conv2D'] and keras.Sequential is keras.models.Sequential
    assert keras.initializers.RandomNormal(stddev=0.1)['config']['stddev'] == 0.1
    with pytest.raises(TypeError):
        keras.initializers.Zeros(seed=1)
    assert keras.optimizers.get('adamw').weight_decay == 0.004
    assert keras.constraints.max_norm(3).max_value == 3.0


def test_global_dtype_policy_sets_the_compute_dtype_of_new_models():
    import pytest

    from elephas_b200.keras import mixed_precision
    from elephas_b200.keras.layers import Dense
    from elephas_b200.keras.models import Sequential
    assert mixed_precision.global_policy().name == 'float32'
    before = Sequential([Dense(2, input_shape=(3,))])
    try:
        mixed_precision.set_global_policy('mixed_bfloat16')
        policy = mixed_precision.global_policy()
        assert policy.compute_dtype == 'bfloat16' and policy.variable_dtype == 'float32'
        after = Sequential([Dense(2, input_shape=(3,))])
        assert after._compute_dtype == 'bfloat16' and before._compute_dtype == 'float32'
        assert after.get_weights()[0].dtype == np.float32            # master weights stay fp32
        after.compile('sgd', 'mse')
        after.fit(np.zeros((8, 3), np.float32), np.zeros((8, 2), np.float32), epochs=1, batch_size=4)   # on a CPU: plain fp32
        with pytest.raises(ValueError):
            mixed_precision.set_global_policy('mixed_float16')
    finally:
        mixed_precision.set_global_policy(mixed_precision.Policy('float32'))
    assert Sequential([Dense(2, input_shape=(3,))])._compute_dtype == 'float32'


def test_backend_and_seed_helpers():
    import elephas_b200.keras as keras
    from elephas_b200.keras.layers import Dense
    from elephas_b200.keras.models import Sequential
    keras.backend.clear_session()
    first = Sequential([Dense(2, input_shape=(3,)), Dense(1)])
    assert [l.name for l in first.layers[1:]] == ['dense', 'dense_1']
    keras.backend.clear_session()
    assert Sequential([Dense(2, input_shape=(3,))]).layers[-1].name == 'dense'
    keras.utils.set_random_seed(11)
    a = Sequential([Dense(4, input_shape=(3,))]).get_weights()[0]
    keras.utils.set_random_seed(11)
    b = Sequential([Dense(4, input_shape=(3,))]).get_weights()[0]
    np.testing.assert_array_equal(a, b)
    assert keras.backend.epsilon() == 1e-7 and keras.backend.image_data_format() == 'channels_last'
