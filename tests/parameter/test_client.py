from unittest.mock import patch

import pytest

from elephas.parameter.client import BaseParameterClient, DeviceClient, HttpClient, SocketClient


def test_client_factory_method():
    assert isinstance(BaseParameterClient.get_client('http', 4001), HttpClient)
    with patch('elephas_b200.parameter.client.socket'):
        assert isinstance(BaseParameterClient.get_client('socket', 4002), SocketClient)
    assert isinstance(BaseParameterClient.get_client('nvlink', 4003), DeviceClient)


def test_unknown_client_type():
    with pytest.raises(ValueError):
        BaseParameterClient.get_client('carrier-pigeon')
