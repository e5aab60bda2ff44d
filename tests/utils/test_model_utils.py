import json

from elephas.utils.model_utils import LossModelTypeMapper, ModelType, ModelTypeEncoder, as_enum


def test_builtin_losses():
    m = LossModelTypeMapper()
    assert m.get_model_type('mean_squared_error') == ModelType.REGRESSION
    assert m.get_model_type('mae') == ModelType.REGRESSION
    assert m.get_model_type('categorical_crossentropy') == ModelType.CLASSIFICATION
    assert m.get_model_type('binary_crossentropy') == ModelType.CLASSIFICATION
    assert m.get_model_type('nope') is None


def test_singleton_and_register_by_name():
    LossModelTypeMapper().register_loss('my_loss', ModelType.REGRESSION)
    assert LossModelTypeMapper() is LossModelTypeMapper()
    assert LossModelTypeMapper().get_model_type('my_loss') == ModelType.REGRESSION


def test_register_by_callable():
    def custom_loss(y_true, y_pred):
        return y_pred

    LossModelTypeMapper().register_loss(custom_loss, ModelType.CLASSIFICATION)
    assert LossModelTypeMapper().get_model_type('custom_loss') == ModelType.CLASSIFICATION
    assert LossModelTypeMapper().get_model_type(custom_loss) == ModelType.CLASSIFICATION


def test_enum_json_round_trip():
    s = json.dumps({'model_type': ModelType.REGRESSION, 'x': 1}, cls=ModelTypeEncoder)
    assert '"__enum__": "ModelType.REGRESSION"' in s
    assert json.loads(s, object_hook=as_enum) == {'model_type': ModelType.REGRESSION, 'x': 1}
