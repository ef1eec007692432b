"""Policy server / client of the Pi0 path: the reference's websocket + msgpack-numpy
wire format, byte for byte, in front of the B200 engine."""
from .msgpack_numpy import Packer, Unpacker, pack_array, packb, unpack_array, unpackb  # noqa: F401
from .policy_wrapper import LeRobotPolicyWrapper  # noqa: F401
from .websocket_policy_client import WebsocketPolicyClient  # noqa: F401
from .websocket_policy_server import WebsocketPolicyServer  # noqa: F401
