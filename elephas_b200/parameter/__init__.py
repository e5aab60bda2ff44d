from .server import *  # noqa: F401,F403
from .client import *  # noqa: F401,F403
from .server import BaseParameterServer, DeviceServer, HttpServer, SocketServer  # noqa: F401
from .client import BaseParameterClient, DeviceClient, HttpClient, SocketClient  # noqa: F401
