"""Policy server / client / wrapper host logic (rows a1-a3) on CPU with a fake policy: wire bytes against frames produced
by the reference's own codec (tests/golden/wire.npz), protocol (metadata frame, reset / switch_model / infer replies,
error convention), request batching, and - when /root/reference is mounted - the reference's OWN client talking to
this server."""
import asyncio
import importlib.util
import os
import sys
import threading
from collections import deque

import numpy as np
import pytest
import torch

from int_act_b200.serving import LeRobotPolicyWrapper, WebsocketPolicyClient, WebsocketPolicyServer, msgpack_numpy

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REF_PKG = "/root/reference/packages/policy-server-client/src"


# ---- wire codec ------------------------------------------------------------------------------------------------------
def test_wire_bytes_identical_to_reference_codec():
    d = np.load(os.path.join(G, "wire.npz"))
    act = np.arange(28, dtype=np.float64).reshape(4, 7) / 7.0
    assert msgpack_numpy.packb(act) == d["act_bytes"].tobytes()
    obs = {"observation.images.top": (np.arange(2 * 3 * 3, dtype=np.uint8).reshape(2, 3, 3)),
           "observation.state": np.linspace(0, 1, 8, dtype=np.float32), "task": "put the carrot on the plate",
           "nested": {"agent": {"eef_pos": np.float32(0.5)}}}
    assert msgpack_numpy.packb(obs) == d["obs_bytes"].tobytes()
    assert msgpack_numpy.packb({"reset": True}) == d["reset_bytes"].tobytes()
    assert msgpack_numpy.packb({"status": "reset"}) == d["status_bytes"].tobytes()
    back = msgpack_numpy.unpackb(d["obs_bytes"].tobytes())
    assert np.array_equal(back["observation.images.top"], obs["observation.images.top"])
    assert back["observation.images.top"].flags.writeable is False          # zero-copy view on the frame
    assert back["nested"]["agent"]["eef_pos"] == np.float32(0.5) and back["task"] == obs["task"]
    a2 = msgpack_numpy.unpackb(d["act_bytes"].tobytes())
    assert a2.dtype == np.float64 and a2.shape == (4, 7) and np.array_equal(a2, act)
    with pytest.raises(ValueError):
        msgpack_numpy.packb(np.array([1 + 2j]))
    with pytest.raises(ValueError):
        msgpack_numpy.packb(np.array([object()], dtype=object))


# ---- fakes -----------------------------------------------------------------------------------------------------------
class FakePolicy:
    """Stands in for LeRobotPolicyWrapper: the action encodes the observation so routing can be checked."""

    def __init__(self):
        self.resets, self.switched, self.groups = 0, [], []

    def select_action(self, obs):
        if obs.get("boom"):
            raise RuntimeError("kaboom")
        self.groups.append(1)
        return np.full((4, 7), float(obs["observation.state"][0]), dtype=np.float64)

    def select_action_many(self, obs_list):
        self.groups.append(len(obs_list))
        return [np.full((4, 7), float(o["observation.state"][0]), dtype=np.float64) for o in obs_list]

    def reset(self):
        self.resets += 1

    def switch_model(self, path):
        self.switched.append(path)


class ServerThread:
    def __init__(self, policy, **kw):
        self.server = WebsocketPolicyServer(policy, host="127.0.0.1", port=0, metadata={"model": "pi0", "n": 4}, **kw)
        self.ready = threading.Event()
        self.thread = threading.Thread(target=self._main, daemon=True)

    def _main(self):
        async def go():
            self.loop = asyncio.get_running_loop()
            self.stop = asyncio.Event()
            ready = asyncio.Event()
            task = asyncio.create_task(self.server.run(ready=ready, stop=self.stop))
            await ready.wait()
            self.ready.set()
            await task
        asyncio.run(go())

    def __enter__(self):
        self.thread.start()
        assert self.ready.wait(10)
        return self.server

    def __exit__(self, *a):
        self.loop.call_soon_threadsafe(self.stop.set)
        self.thread.join(5)


def obs_for(v):
    return {"observation.images.top": np.zeros((8, 8, 3), np.uint8), "observation.state": np.full(8, v, np.float32),
            "task": "t"}


# ---- protocol --------------------------------------------------------------------------------------------------------
def test_protocol_roundtrip_and_error_convention():
    pol = FakePolicy()
    with ServerThread(pol) as srv:
        c = WebsocketPolicyClient("127.0.0.1", srv.bound_port)
        assert c.get_server_metadata() == {"model": "pi0", "n": 4}
        assert c.reset() == {"status": "reset"} and pol.resets == 1
        assert c.switch_model("/ckpt/step_1000") == {"status": "model switched"} and pol.switched == ["/ckpt/step_1000"]
        a = c.infer(obs_for(3.0))
        assert isinstance(a, np.ndarray) and a.dtype == np.float64 and a.shape == (4, 7) and np.all(a == 3.0)
        with pytest.raises(RuntimeError, match="Error in inference server:\n(.|\n)*kaboom"):
            c.infer({"boom": True})
        import websockets.exceptions
        with pytest.raises(websockets.exceptions.ConnectionClosed) as ei:   # server closed with 1011 after the traceback
            c.infer(obs_for(1.0))
        assert ei.value.rcvd.code == 1011
        c2 = WebsocketPolicyClient("127.0.0.1", srv.bound_port)          # the server itself keeps serving
        assert np.all(c2.infer(obs_for(5.0)) == 5.0)
        c2.close()


def test_concurrent_requests_are_batched_and_routed():
    pol = FakePolicy()
    n = 6
    with ServerThread(pol, max_batch=4, batch_window_ms=200.0) as srv:
        clients = [WebsocketPolicyClient("127.0.0.1", srv.bound_port) for _ in range(n)]
        out = [None] * n
        ts = [threading.Thread(target=lambda i=i: out.__setitem__(i, clients[i].infer(obs_for(float(i))))) for i in range(n)]
        [t.start() for t in ts]
        [t.join(20) for t in ts]
        for i in range(n):
            assert np.all(out[i] == float(i)), (i, out[i])
        assert sum(pol.groups) == n and max(pol.groups) > 1 and max(pol.groups) <= 4
        assert srv.stats["requests"] == n and srv.stats["max_batch_seen"] == max(pol.groups)
        [c.close() for c in clients]
    pol1 = FakePolicy()
    with ServerThread(pol1) as srv:   # default = the reference's one-at-a-time behaviour
        c = WebsocketPolicyClient("127.0.0.1", srv.bound_port)
        for i in range(3):
            assert np.all(c.infer(obs_for(float(i))) == float(i))
        assert pol1.groups == [1, 1, 1]


@pytest.mark.skipif(not os.path.isdir(REF_PKG), reason="reference tree not mounted")
def test_reference_client_talks_to_this_server():
    sys.path.insert(0, REF_PKG)
    try:
        spec = importlib.util.spec_from_file_location(
            "ref_ws_client", os.path.join(REF_PKG, "policy_server_client", "websocket_policy_client.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    except Exception as e:  # missing optional deps of the reference package
        pytest.skip(f"reference client not importable here: {e}")
    finally:
        sys.path.remove(REF_PKG)
    pol = FakePolicy()
    with ServerThread(pol) as srv:
        c = mod.WebsocketPolicyClient("127.0.0.1", srv.bound_port)
        assert c.get_server_metadata() == {"model": "pi0", "n": 4}
        assert c.reset() == {"status": "reset"}
        a = c.infer(obs_for(2.0))
        assert a.dtype == np.float64 and a.shape == (4, 7) and np.all(a == 2.0)


# ---- wrapper ---------------------------------------------------------------------------------------------------------
class FakeModel:
    class config:
        n_action_steps = 4

    def __init__(self):
        self.queue, self.infers = deque(), []

    def reset(self):
        self.queue.clear()

    def infer_action(self, batch):
        s = batch["observation.state"]
        self.infers.append(s.shape[0])
        return s[:, :1, None].repeat(1, 4, 7) + torch.arange(4.0)[None, :, None]

    def select_action(self, batch):
        if not self.queue:
            self.queue.extend(self.infer_action(batch).transpose(0, 1))
        return self.queue.popleft()


class FakeAdapter:
    def preprocess(self, obs):
        s = torch.as_tensor(np.asarray(obs["observation.state"], dtype=np.float32)).reshape(-1, 8)[:, :7]
        return {"observation.state": s, "task": [obs["task"]] if isinstance(obs["task"], str) else obs["task"]}

    def postprocess(self, actions):
        return actions.double().numpy() * 2.0

    def reset(self):
        pass


def test_wrapper_select_action_semantics():
    m = FakeModel()
    w = LeRobotPolicyWrapper(m, FakeAdapter(), action_step=4)
    a = w.select_action(obs_for(1.0))
    assert a.shape == (4, 7) and a.dtype == np.float64 and np.allclose(a[:, 0], 2.0 * (1.0 + np.arange(4)))
    assert m.infers == [1] and len(m.queue) == 0
    ab = w.select_action({"observation.state": np.stack([np.full(8, 1.0), np.full(8, 5.0)]), "task": ["a", "b"]})
    assert ab.shape == (2, 4, 7) and np.allclose(ab[1, :, 0], 2.0 * (5.0 + np.arange(4)))
    many = w.select_action_many([obs_for(1.0), obs_for(2.0), obs_for(3.0)])
    assert m.infers[-1] == 3 and [x.shape for x in many] == [(4, 7)] * 3
    assert np.allclose(many[2][:, 0], 2.0 * (3.0 + np.arange(4)))
    w2 = LeRobotPolicyWrapper(FakeModel(), FakeAdapter(), action_step=2)   # partial chunk consumption: never batched
    assert not w2.can_batch() and len(w2.select_action_many([obs_for(1.0), obs_for(2.0)])) == 2
    with pytest.raises(ValueError):
        w.switch_model("/x")
    w3 = LeRobotPolicyWrapper(m, FakeAdapter(), model_loader=lambda p: FakeModel())
    w3.switch_model("/ckpt")
    assert w3.model is not m
    w3.reset()
