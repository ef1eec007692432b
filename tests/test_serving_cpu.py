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


def test_two_workers_per_replica_route_every_request_and_keep_control_barriers():
    """workers_per_replica=2: two threads share the one policy (host gathering of the next group overlaps the GPU part of the
    current one); every client still gets its own answer and a reset is still a barrier."""
    pol = FakePolicy()
    n = 10
    with ServerThread(pol, max_batch=4, batch_window_ms=50.0, workers_per_replica=2) as srv:
        clients = [WebsocketPolicyClient("127.0.0.1", srv.bound_port) for _ in range(n)]
        out = [None] * n
        for rnd in range(2):
            ts = [threading.Thread(target=lambda i=i: out.__setitem__(i, clients[i].infer(obs_for(float(i + rnd)))))
                  for i in range(n)]
            [t.start() for t in ts]
            [t.join(20) for t in ts]
            for i in range(n):
                assert np.all(out[i] == float(i + rnd)), (i, out[i])
            assert clients[0].reset() == {"status": "reset"}
        assert sum(pol.groups) == 2 * n and srv.stats["requests"] == 2 * n
        [c.close() for c in clients]


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


# ---- round 2: failure isolation, latency histogram, replicas behind one port, reference-style construction ---------------
def test_one_bad_request_in_a_coalesced_group_only_fails_its_own_connection():
    pol = FakePolicy()
    bad_many = pol.select_action_many

    def many(obs_list):   # like the real wrapper: a malformed observation makes the whole batched call raise
        if any(o.get("boom") for o in obs_list):
            raise RuntimeError("kaboom")
        return bad_many(obs_list)

    pol.select_action_many = many
    with ServerThread(pol, max_batch=4, batch_window_ms=300.0) as srv:
        good = [WebsocketPolicyClient("127.0.0.1", srv.bound_port) for _ in range(2)]
        bad = WebsocketPolicyClient("127.0.0.1", srv.bound_port)
        out, err = [None, None], []

        def run_bad():
            try:
                bad.infer({"boom": True})
            except RuntimeError as e:
                err.append(str(e))

        ts = [threading.Thread(target=lambda i=i: out.__setitem__(i, good[i].infer(obs_for(float(i + 1))))) for i in range(2)]
        ts.append(threading.Thread(target=run_bad))
        [t.start() for t in ts]
        [t.join(20) for t in ts]
        assert np.all(out[0] == 1.0) and np.all(out[1] == 2.0)           # co-batched clients are unharmed
        assert len(err) == 1 and "kaboom" in err[0]
        assert srv.stats["failed_requests"] == 1
        lat = srv.latency_ms()
        assert lat["n"] == 3 and lat["p50"] > 0 and lat["p99"] >= lat["p50"]
        assert np.all(good[0].infer(obs_for(7.0)) == 7.0)                 # and keep their connection
        [c.close() for c in good]


def test_replicas_share_one_port_and_controls_reach_every_replica():
    pols = [FakePolicy() for _ in range(3)]
    import time as _t
    for p in pols:   # make inference slow enough that the other replicas pick up work
        orig = p.select_action
        p.select_action = lambda obs, orig=orig: (_t.sleep(0.05), orig(obs))[1]
    n = 9
    with ServerThread(pols[0], replicas=pols[1:]) as srv:
        clients = [WebsocketPolicyClient("127.0.0.1", srv.bound_port) for _ in range(n)]
        out = [None] * n
        ts = [threading.Thread(target=lambda i=i: out.__setitem__(i, clients[i].infer(obs_for(float(i))))) for i in range(n)]
        [t.start() for t in ts]
        [t.join(20) for t in ts]
        for i in range(n):
            assert np.all(out[i] == float(i))
        assert srv.stats["replicas"] == 3 and sum(srv.stats["per_replica_batches"]) == n
        assert sum(1 for b in srv.stats["per_replica_batches"] if b > 0) >= 2     # more than one GPU did work
        assert clients[0].reset() == {"status": "reset"} and [p.resets for p in pols] == [1, 1, 1]
        assert clients[1].switch_model("/ckpt/a") == {"status": "model switched"}
        assert [p.switched for p in pols] == [["/ckpt/a"]] * 3
        [c.close() for c in clients]


def test_wrapper_reference_constructor_starts_without_a_model_and_loads_on_switch():
    from types import SimpleNamespace

    loaded = []

    class FakePI0:
        @classmethod
        def from_pretrained(cls, pretrained_name_or_path, config=None, **kw):
            loaded.append((pretrained_name_or_path, config))
            m = FakeModel()
            m.model = SimpleNamespace(paligemma_with_expert=SimpleNamespace(
                gemma_expert=SimpleNamespace(lm_head=1), paligemma=SimpleNamespace(language_model=SimpleNamespace(lm_head=1))))
            m.to = lambda a, m=m: m
            m.eval = lambda m=m: m
            return m

    cfg = SimpleNamespace(seed=42, debug=False, gpu_id=0, use_amp=True, use_bf16=True, freeze_lm_head=True,
                          model_cfg=SimpleNamespace(type="pi0"),
                          eval_cfg=SimpleNamespace(action_step=4, port=8000, host="0.0.0.0",
                                                   env_adapter_path="src.experiments.env_adapters.simpler.BridgeSimplerAdapter"))
    w = LeRobotPolicyWrapper(pipeline_cfg=cfg, model_class=FakePI0)
    assert w.model is None and w.env_adapter is None and w.action_step == 4
    with pytest.raises(ValueError, match="Model is not initialized"):
        w.reset()
    with pytest.raises(ValueError, match="Model is not initialized"):
        w.select_action(obs_for(1.0))
    w._initialize_env_adapter = lambda: FakeAdapter()     # the real one needs a CUDA device
    w.switch_model("/ckpt/step_1513")
    assert loaded == [("/ckpt/step_1513", cfg.model_cfg)]
    assert w.model.model.paligemma_with_expert.gemma_expert.lm_head is None            # freeze_lm_head
    assert w.select_action(obs_for(2.0)).shape == (4, 7)
    from int_act_b200.serving.policy_wrapper import resolve_adapter_class
    from int_act_b200 import adapters
    assert resolve_adapter_class("src.experiments.env_adapters.simplerMS3.BridgeSimplerBatchAdapter") is adapters.BridgeSimplerBatchAdapter
    with pytest.raises(ValueError):
        resolve_adapter_class("src.experiments.env_adapters.libero.NoSuchAdapter")


REF_YAML = "/root/reference/config/experiment/simpler/pi0_finetune_bridge_ev.yaml"


@pytest.mark.skipif(not os.path.exists(REF_YAML), reason="reference tree not mounted")
def test_launcher_builds_the_server_from_the_reference_yaml():
    from int_act_b200.pipeline import PipelineConfig, parse_overrides
    from int_act_b200.serve import build_server

    over = parse_overrides(["--eval_cfg.port", "8123", "--use_bf16=False", "--eval_cfg.env_adapter", "BridgeSimplerBatchAdapter"])
    cfg = PipelineConfig.from_yaml(REF_YAML, over)
    assert cfg.eval_cfg.port == 8123 and cfg.use_bf16 is False and cfg.seed == 42
    assert cfg.eval_cfg.env_adapter_path == "src.experiments.env_adapters.simpler.BridgeSimplerBatchAdapter"
    assert cfg.model_cfg.tokenizer_max_length == 72 and cfg.model_cfg.n_action_steps == 4
    assert os.path.exists(cfg.env.dataset_statistics_path)
    server, wrappers, _ = build_server(["--config_path", REF_YAML, "--eval_cfg.port", "0", "--eval_cfg.host", "127.0.0.1",
                                        "--server.max_batch", "4"])
    assert len(wrappers) == 1 and wrappers[0].model is None and wrappers[0].action_step == 4
    # the reference's unmodified flow: connect, reset before any model -> the server reports the error over the wire
    st = ServerThread.__new__(ServerThread)
    st.server, st.ready, st.thread = server, threading.Event(), None
    st.thread = threading.Thread(target=st._main, daemon=True)
    with st as srv:
        c = WebsocketPolicyClient("127.0.0.1", srv.bound_port)
        assert c.get_server_metadata()["model"] == "pi0"
        with pytest.raises(RuntimeError, match="Model is not initialized"):
            c.reset()
