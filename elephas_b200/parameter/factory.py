"""Factories that pair a parameter client with its server: ``'http'``, ``'socket'`` and the device store ``'nvlink'``.

Same public surface as the reference (``elephas/parameter/factory.py:6-41``: ``ClientServerFactory.get_factory(name)``,
``create_client`` / ``create_server``, one subclass per transport), built the other way round: a transport is ONE line
that names its two classes, the base class does the rest, and ``get_factory`` looks the name up in a registry that the
subclasses fill in as they are defined (so a user-defined transport registers itself by subclassing, too).
"""
from .client import DeviceClient, HttpClient, SocketClient
from .server import DeviceServer, HttpServer, SocketServer


class ClientServerFactory:
    """``get_factory('socket').create_server(model, port, mode)`` / ``.create_client(port)``."""

    _type = 'base'
    client_class = server_class = None
    _registry = {}

    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        if cls._type != 'base':
            ClientServerFactory._registry[cls._type] = cls

    @classmethod
    def get_factory(cls, _type):
        factory = cls._registry.get(_type)
        if factory is None:
            raise ValueError("Unknown factory type {}".format(_type))
        return factory()

    def _build(self, which, args, kwargs):
        target = getattr(self, which)
        if target is None:
            raise NotImplementedError("%s does not name its %s" % (type(self).__name__, which.replace('_', ' ')))
        return target(*args, **kwargs)

    def create_client(self, *args, **kwargs):
        return self._build('client_class', args, kwargs)

    def create_server(self, *args, **kwargs):
        return self._build('server_class', args, kwargs)


class HttpFactory(ClientServerFactory):
    _type, client_class, server_class = 'http', HttpClient, HttpServer


class SocketFactory(ClientServerFactory):
    _type, client_class, server_class = 'socket', SocketClient, SocketServer


class DeviceFactory(ClientServerFactory):
    _type, client_class, server_class = 'nvlink', DeviceClient, DeviceServer
