"""Client/server factories (reference: ``elephas/parameter/factory.py:6-41``), plus ``'nvlink'``."""
from abc import ABC, abstractmethod

from .client import DeviceClient, HttpClient, SocketClient
from .server import DeviceServer, HttpServer, SocketServer


class ClientServerFactory(ABC):
    _type = 'base'

    @classmethod
    def get_factory(cls, _type):
        for klass in cls.__subclasses__():
            if klass._type == _type:
                return klass()
        raise ValueError("Unknown factory type {}".format(_type))

    @abstractmethod
    def create_client(self, *args, **kwargs):
        pass

    @abstractmethod
    def create_server(self, *args, **kwargs):
        pass


class HttpFactory(ClientServerFactory):
    _type = 'http'

    def create_client(self, *args, **kwargs):
        return HttpClient(*args, **kwargs)

    def create_server(self, *args, **kwargs):
        return HttpServer(*args, **kwargs)


class SocketFactory(ClientServerFactory):
    _type = 'socket'

    def create_client(self, *args, **kwargs):
        return SocketClient(*args, **kwargs)

    def create_server(self, *args, **kwargs):
        return SocketServer(*args, **kwargs)


class DeviceFactory(ClientServerFactory):
    _type = 'nvlink'

    def create_client(self, *args, **kwargs):
        return DeviceClient(*args, **kwargs)

    def create_server(self, *args, **kwargs):
        return DeviceServer(*args, **kwargs)
