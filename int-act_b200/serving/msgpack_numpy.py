"""msgpack with NumPy arrays - the wire codec of the policy server (row a2).

Byte-compatible with the reference codec
(/root/reference/packages/policy-server-client/src/policy_server_client/msgpack_numpy.py:21-57): an ndarray travels
as the map {b"__ndarray__": True, b"data": <raw bytes>, b"dtype": dtype.str, b"shape": shape}, a NumPy scalar as
{b"__npgeneric__": True, b"data": item, b"dtype": dtype.str}; void / object / complex dtypes are refused (no pickle
fallback). Unpacking returns a zero-copy, read-only view on the received buffer.
tests/test_serving_cpu.py checks the bytes against frames produced by the reference's own codec (tests/golden/wire.npz).
"""
import functools

import msgpack
import numpy as np


def pack_array(obj):
    if isinstance(obj, (np.ndarray, np.generic)) and obj.dtype.kind in ("V", "O", "c"):
        raise ValueError(f"Unsupported dtype: {obj.dtype}")
    if isinstance(obj, np.ndarray):
        return {b"__ndarray__": True, b"data": obj.tobytes(), b"dtype": obj.dtype.str, b"shape": obj.shape}
    if isinstance(obj, np.generic):
        return {b"__npgeneric__": True, b"data": obj.item(), b"dtype": obj.dtype.str}
    return obj


def unpack_array(obj):
    if b"__ndarray__" in obj:
        return np.ndarray(buffer=obj[b"data"], dtype=np.dtype(obj[b"dtype"]), shape=obj[b"shape"])
    if b"__npgeneric__" in obj:
        return np.dtype(obj[b"dtype"]).type(obj[b"data"])
    return obj


Packer = functools.partial(msgpack.Packer, default=pack_array)
packb = functools.partial(msgpack.packb, default=pack_array)
Unpacker = functools.partial(msgpack.Unpacker, object_hook=unpack_array)
unpackb = functools.partial(msgpack.unpackb, object_hook=unpack_array)
