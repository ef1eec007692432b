"""Websocket client of the policy server (the counterpart SimplerEnv / ManiSkill3 / LIBERO evaluators hold).
Same calls, frames and error convention as the reference client
(/root/reference/packages/policy-server-client/src/policy_server_client/websocket_policy_client.py:20-81)."""
import logging
import time
from typing import Dict, Tuple

import websockets.sync.client

from . import msgpack_numpy


class WebsocketPolicyClient:
    def __init__(self, host: str, port: int, retry_s: float = 5.0):
        self.host, self.port = host, port
        self.logger = logging.getLogger("websockets.client")
        self._uri = f"ws://{self.host}:{self.port}"
        self._retry_s = retry_s
        self._ws, self._server_metadata = self._wait_for_server()
        self._packer = msgpack_numpy.Packer()

    def get_server_metadata(self) -> Dict:
        return self._server_metadata

    def _wait_for_server(self) -> Tuple[websockets.sync.client.ClientConnection, Dict]:
        self.logger.info(f"Waiting for server at {self._uri}...")
        while True:
            try:
                conn = websockets.sync.client.connect(self._uri, compression=None, max_size=None, ping_timeout=None)
                return conn, msgpack_numpy.unpackb(conn.recv())
            except ConnectionRefusedError:
                self.logger.info("Still waiting for server...")
                time.sleep(self._retry_s)

    def _call(self, payload):
        self._ws.send(self._packer.pack(payload))
        response = self._ws.recv()
        if isinstance(response, str):  # bytes expected; a text frame is the server's traceback
            raise RuntimeError(f"Error in inference server:\n{response}")
        return msgpack_numpy.unpackb(response)

    def infer(self, obs: Dict):
        """Actions for one observation (or one rollout batch)."""
        return self._call(obs)

    def reset(self):
        return self._call({"reset": True})

    def switch_model(self, new_model_path):
        return self._call({"new_model_path": new_model_path})

    def close(self):
        self._ws.close()
