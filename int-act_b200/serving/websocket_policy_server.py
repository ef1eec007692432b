"""Websocket policy server (row a1) - the reference's protocol in front of the B200 engine.

Wire contract (/root/reference/packages/policy-server-client/src/policy_server_client/websocket_policy_server.py:45-91):
websocket, `compression=None`, `max_size=None`; on connect ONE binary frame = msgpack(metadata); then per request a binary
msgpack map: {"new_model_path": str} -> {"status": "model switched"}; {"reset": True} -> {"status": "reset"}; anything
else is an observation -> reply = msgpack(ndarray float64 [action_step, 7] or [B, action_step, 7]) - the bare array.
Any exception in the handler -> the traceback as a TEXT frame, close code 1011, re-raise; a closed connection ends the
handler quietly.

What differs from the reference is scheduling, not bytes:
  * the reference runs `select_action` inside the coroutine, so the event loop (and every other connection) stalls for
    the whole inference. Here requests go through a queue to inference worker threads; requests that arrive while the
    GPU is busy (or within `batch_window_ms`) are coalesced, up to `max_batch`, into ONE engine batch
    (`LeRobotPolicyWrapper.select_action_many`). `max_batch=1` (default) is the reference's one-at-a-time behaviour;
  * the reference scales out with N server processes on N ports (doc/evaluation.md:119-135). `replicas=[wrapper_gpu1,
    ...]` puts N engines (one per GPU of the box) behind ONE port: every replica has its own worker thread pulling
    coalesced groups from the shared queue, so the box's GPUs are load-balanced by whoever is free;
  * `workers_per_replica=2` gives every replica two worker threads: one gathers the next group's frames on the host while the
    other's group is on the GPU (the wrapper serialises the GPU part). Measured with 0.92 MB frames it does not pay (458 vs
    482 requests/s with 64 clients, 508 with 96: the groups get smaller and the single event-loop thread that receives and
    decodes the frames is the cap), so the default stays 1;
  * control messages are barriers: the queue is drained, every in-flight group finishes, then the message is applied to
    EVERY replica, in arrival order;
  * a request that fails (e.g. a malformed observation) only fails its own connection: when a coalesced group raises, its
    requests are re-run one by one so that only the offender gets the traceback frame, as in the reference.
`stats` counts requests / batches and keeps the last 4096 request latencies (queue wait + inference); `latency_ms()`
returns their p50 / p99.
"""
import asyncio
import logging
import threading
import time
import traceback
from collections import deque
from concurrent.futures import Future
from queue import Empty, Queue
from typing import List, Optional

import websockets.asyncio.server
import websockets.exceptions
import websockets.frames

from . import msgpack_numpy


class WebsocketPolicyServer:
    def __init__(self, policy, host: str = "0.0.0.0", port: int = 8000, metadata: Optional[dict] = None, *,
                 max_batch: int = 1, batch_window_ms: float = 0.0, replicas: Optional[List[object]] = None,
                 workers_per_replica: int = 1) -> None:
        self._policies = [policy] + list(replicas or [])
        self._policy = policy
        self._host, self._port = host, port
        self._metadata = metadata or {}
        self._max_batch = max(1, int(max_batch))
        # 2: while one worker's group is on the GPU the other gathers the next group's frames on the host (the wrapper
        # serialises the GPU part itself)
        self._wpr = max(1, int(workers_per_replica))
        self._window_s = max(0.0, batch_window_ms) / 1e3
        self.logger = logging.getLogger("policy_server")
        self.logger.setLevel(logging.INFO)
        self._queue: "Queue[tuple]" = Queue()
        self._workers: List[threading.Thread] = []
        self._stop = threading.Event()
        self._lock = threading.Condition()
        self._inflight = 0          # groups being run by some replica
        self._control_pending = 0   # control messages waiting for the barrier
        self._lat_ms = deque(maxlen=4096)
        self.stats = {"requests": 0, "batches": 0, "max_batch_seen": 0, "replicas": len(self._policies),
                      "per_replica_batches": [0] * len(self._policies), "failed_requests": 0}

    def latency_ms(self) -> dict:
        """p50 / p99 / max of the last (up to 4096) request latencies: enqueue -> result ready, in milliseconds."""
        lat = sorted(self._lat_ms)
        if not lat:
            return {"n": 0}
        return {"n": len(lat), "p50": lat[len(lat) // 2], "p99": lat[min(len(lat) - 1, int(len(lat) * 0.99))],
                "max": lat[-1]}

    # ---- lifecycle ---------------------------------------------------------------------------------------------------
    def serve_forever(self) -> None:
        """Starts the server and runs it forever. This is a blocking call."""
        self.logger.info(f"Starting server on {self._host}:{self._port}")
        asyncio.run(self.run())

    async def run(self, ready: Optional[asyncio.Event] = None, stop: Optional[asyncio.Event] = None):
        self._start_workers()
        try:
            async with websockets.asyncio.server.serve(self._handler, self._host, self._port, compression=None,
                                                       max_size=None) as server:
                self.bound_port = server.sockets[0].getsockname()[1]
                if ready is not None:
                    ready.set()
                if stop is None:
                    await server.serve_forever()
                else:
                    await stop.wait()
        finally:
            self._stop_workers()

    def _start_workers(self):
        if any(w.is_alive() for w in self._workers):
            return
        self._stop.clear()
        self._workers = [threading.Thread(target=self._worker_loop, args=(i,), name=f"pi0-inference-{i}", daemon=True)
                         for i in range(len(self._policies) * self._wpr)]
        for w in self._workers:
            w.start()

    def _stop_workers(self):
        self._stop.set()
        for _ in self._workers:
            self._queue.put(("stop", None, None, 0.0))
        for w in self._workers:
            w.join(timeout=5.0)

    # ---- inference workers: one per replica, each batches what is waiting -----------------------------------------------
    def _worker_loop(self, widx: int):
        idx = widx % len(self._policies)
        policy = self._policies[idx]
        pending = None
        while not self._stop.is_set():
            if pending is None:
                with self._lock:  # a pending control message is a barrier: take no new work until it has run
                    while self._control_pending and not self._stop.is_set():
                        self._lock.wait(timeout=0.05)
                item = self._queue.get()
            else:
                item, pending = pending, None
            kind, payload, fut, t_in = item
            if kind == "stop":
                break
            if kind != "infer":
                self._run_control(kind, payload, fut)
                continue
            with self._lock:  # a control message taken by another worker meanwhile runs first (it is a barrier)
                while self._control_pending and not self._stop.is_set():
                    self._lock.wait(timeout=0.05)
                self._inflight += 1
            group = [(payload, fut, t_in)]
            deadline = time.monotonic() + self._window_s
            while len(group) < self._max_batch:
                try:
                    timeout = deadline - time.monotonic()
                    nxt = self._queue.get_nowait() if timeout <= 0 else self._queue.get(timeout=timeout)
                except Empty:
                    break
                if nxt[0] != "infer":  # a control message / stop is a barrier: run the group first, then the message
                    pending = nxt
                    if nxt[0] in ("switch", "reset"):
                        with self._lock:
                            self._control_pending += 1
                    break
                group.append((nxt[1], nxt[2], nxt[3]))
            with self._lock:
                self.stats["requests"] += len(group)
                self.stats["batches"] += 1
                self.stats["per_replica_batches"][idx] += 1
                self.stats["max_batch_seen"] = max(self.stats["max_batch_seen"], len(group))
            try:
                self._run_group(policy, group)
            finally:
                with self._lock:
                    self._inflight -= 1
                    self._lock.notify_all()
            if pending is not None and pending[0] in ("switch", "reset"):
                k, pl, f, _ = pending
                pending = None
                self._run_control(k, pl, f, counted=True)

    def _finish(self, fut: Future, t_in: float, result=None, exc: Optional[BaseException] = None):
        self._lat_ms.append(1e3 * (time.monotonic() - t_in))
        if exc is not None:
            self.stats["failed_requests"] += 1
            fut.set_exception(exc)
        else:
            fut.set_result(result)

    def _run_group(self, policy, group):
        try:
            if len(group) == 1 or not hasattr(policy, "select_action_many"):
                results = [policy.select_action(o) for o, _, _ in group]
            else:
                results = policy.select_action_many([o for o, _, _ in group])
        except Exception:  # noqa: BLE001 - isolate the offender: one request at a time, each with its own outcome
            for o, f, t in group:
                try:
                    self._finish(f, t, policy.select_action(o))
                except Exception as e:  # noqa: BLE001
                    self._finish(f, t, exc=e)
            return
        for (_, f, t), r in zip(group, results):
            if isinstance(r, BaseException):
                self._finish(f, t, exc=r)
            else:
                self._finish(f, t, r)

    def _run_control(self, kind, payload, fut: Future, counted: bool = False):
        """Barrier: wait until no replica runs a group, then apply the message to every replica."""
        with self._lock:
            if not counted:
                self._control_pending += 1
            while self._inflight > 0 and not self._stop.is_set():
                self._lock.wait(timeout=0.05)
        try:
            fut.set_result(self._control(kind, payload))
        except Exception as e:  # noqa: BLE001
            fut.set_exception(e)
        finally:
            with self._lock:
                self._control_pending -= 1
                self._lock.notify_all()

    def _control(self, kind, payload):
        if kind == "switch":
            for p in self._policies:
                p.switch_model(payload)
                if self._max_batch > 1 and hasattr(p, "warmup"):
                    p.warmup(self._max_batch)
            self.logger.info(f"Loaded new model checkpoint: {payload}")
            return {"status": "model switched"}
        for p in self._policies:
            p.reset()
        return {"status": "reset"}

    async def _submit(self, kind, payload):
        fut: Future = Future()
        self._queue.put((kind, payload, fut, time.monotonic()))
        return await asyncio.wrap_future(fut)

    # ---- one connection ----------------------------------------------------------------------------------------------
    async def _handler(self, websocket):
        self.logger.info(f"Connection from {websocket.remote_address} opened")
        packer = msgpack_numpy.Packer()
        await websocket.send(packer.pack(self._metadata))
        while True:
            try:
                obs = msgpack_numpy.unpackb(await websocket.recv())
                new_model_path = obs.get("new_model_path", None)
                if new_model_path is not None:  # no observation travels with this message
                    await websocket.send(packer.pack(await self._submit("switch", new_model_path)))
                    continue
                if obs.get("reset", False):
                    await websocket.send(packer.pack(await self._submit("reset", None)))
                    continue
                action = await self._submit("infer", obs)
                await websocket.send(packer.pack(action))
            except websockets.exceptions.ConnectionClosed:
                self.logger.info(f"Connection from {websocket.remote_address} closed")
                break
            except Exception:
                await websocket.send(traceback.format_exc())
                await websocket.close(code=websockets.frames.CloseCode.INTERNAL_ERROR,
                                      reason="Internal server error. Traceback included in previous frame.")
                raise
