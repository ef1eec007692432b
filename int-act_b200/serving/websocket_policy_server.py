"""Websocket policy server (row a1) - the reference's protocol in front of the B200 engine.

Wire contract (/root/reference/packages/policy-server-client/src/policy_server_client/websocket_policy_server.py:45-91):
websocket, `compression=None`, `max_size=None`; on connect ONE binary frame = msgpack(metadata); then per request a binary
msgpack map: {"new_model_path": str} -> {"status": "model switched"}; {"reset": True} -> {"status": "reset"}; anything
else is an observation -> reply = msgpack(ndarray float64 [action_step, 7] or [B, action_step, 7]) - the bare array.
Any exception in the handler -> the traceback as a TEXT frame, close code 1011, re-raise; a closed connection ends the
handler quietly.

What differs from the reference is scheduling, not bytes: the reference runs `select_action` inside the coroutine, so
the event loop (and every other connection) stalls for the whole inference. Here requests go to one inference worker
thread through a queue; requests that arrive while the GPU is busy (or within `batch_window_ms`) are coalesced, up to
`max_batch`, into ONE engine batch (`LeRobotPolicyWrapper.select_action_many`). `max_batch=1` (default) is the
reference's one-at-a-time behaviour. Control messages act as barriers: they are executed by the same worker, in
arrival order.
"""
import asyncio
import logging
import threading
import time
import traceback
from concurrent.futures import Future
from queue import Empty, Queue
from typing import Optional

import websockets.asyncio.server
import websockets.exceptions
import websockets.frames

from . import msgpack_numpy


class WebsocketPolicyServer:
    def __init__(self, policy, host: str = "0.0.0.0", port: int = 8000, metadata: Optional[dict] = None, *,
                 max_batch: int = 1, batch_window_ms: float = 0.0) -> None:
        self._policy = policy
        self._host, self._port = host, port
        self._metadata = metadata or {}
        self._max_batch = max(1, int(max_batch))
        self._window_s = max(0.0, batch_window_ms) / 1e3
        self.logger = logging.getLogger("policy_server")
        self.logger.setLevel(logging.INFO)
        self._queue: "Queue[tuple]" = Queue()
        self._worker: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self.stats = {"requests": 0, "batches": 0, "max_batch_seen": 0}

    # ---- lifecycle ---------------------------------------------------------------------------------------------------
    def serve_forever(self) -> None:
        """Starts the server and runs it forever. This is a blocking call."""
        self.logger.info(f"Starting server on {self._host}:{self._port}")
        asyncio.run(self.run())

    async def run(self, ready: Optional[asyncio.Event] = None, stop: Optional[asyncio.Event] = None):
        self._start_worker()
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
            self._stop_worker()

    def _start_worker(self):
        if self._worker is None or not self._worker.is_alive():
            self._stop.clear()
            self._worker = threading.Thread(target=self._worker_loop, name="pi0-inference", daemon=True)
            self._worker.start()

    def _stop_worker(self):
        self._stop.set()
        self._queue.put(("stop", None, None))
        if self._worker is not None:
            self._worker.join(timeout=5.0)

    # ---- inference worker: owns the policy, batches what is waiting ---------------------------------------------------
    def _worker_loop(self):
        pending = None
        while not self._stop.is_set():
            item = pending if pending is not None else self._queue.get()
            pending = None
            kind, payload, fut = item
            if kind == "stop":
                break
            if kind != "infer":
                self._run(fut, lambda: self._control(kind, payload))
                continue
            group = [(payload, fut)]
            deadline = time.monotonic() + self._window_s
            while len(group) < self._max_batch:
                try:
                    timeout = deadline - time.monotonic()
                    nxt = self._queue.get_nowait() if timeout <= 0 else self._queue.get(timeout=timeout)
                except Empty:
                    break
                if nxt[0] != "infer":  # a control message is a barrier: run the group first, then the message
                    pending = nxt
                    break
                group.append((nxt[1], nxt[2]))
            self.stats["requests"] += len(group)
            self.stats["batches"] += 1
            self.stats["max_batch_seen"] = max(self.stats["max_batch_seen"], len(group))
            try:
                if len(group) == 1 or not hasattr(self._policy, "select_action_many"):
                    results = [self._policy.select_action(o) for o, _ in group]
                else:
                    results = self._policy.select_action_many([o for o, _ in group])
                for (_, f), r in zip(group, results):
                    f.set_result(r)
            except Exception as e:  # noqa: BLE001 - every waiting request gets the failure
                for _, f in group:
                    if not f.done():
                        f.set_exception(e)

    def _control(self, kind, payload):
        if kind == "switch":
            self._policy.switch_model(payload)
            self.logger.info(f"Loaded new model checkpoint: {payload}")
            return {"status": "model switched"}
        self._policy.reset()
        return {"status": "reset"}

    @staticmethod
    def _run(fut: Future, fn):
        try:
            fut.set_result(fn())
        except Exception as e:  # noqa: BLE001
            fut.set_exception(e)

    async def _submit(self, kind, payload):
        fut: Future = Future()
        self._queue.put((kind, payload, fut))
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
