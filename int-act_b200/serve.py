"""Policy-server launcher that consumes the reference's experiment YAML - the flow of /root/reference/src/agent/run.py:45-64
+ src/experiments/policies/policy_wrapper.py:18-113, on the B200 engine:

    python -m int_act_b200.serve --config_path config/experiment/simpler/pi0_finetune_bridge_ev.yaml \
        --eval_cfg.port 8000 --use_bf16 True [--eval_cfg.env_adapter BridgeSimplerBatchAdapter]
        [--server.max_batch 16 --server.batch_window_ms 2 --server.gpus 8 --server.model_path <ckpt dir>]

builds `LeRobotPolicyWrapper(pipeline_cfg=..., model_class=PI0Policy)` - no model yet - and serves the reference's
websocket protocol; the first `{"new_model_path": ...}` request loads the checkpoint, exactly as the reference's
evaluators drive it. Dotted `--a.b value` overrides are the reference's (draccus) command-line form
(doc/evaluation.md:119-152). `--server.*` keys are additions of this server: request coalescing (`max_batch`,
`batch_window_ms`), `gpus` = engines behind this ONE port (the reference's alternative - one process per GPU on N ports,
doc/evaluation.md:119-135 - still works: start N launchers with `--gpu_id i --eval_cfg.port 8000+i`), `model_path` =
load a checkpoint at start-up instead of waiting for the first switch request.
"""
from __future__ import annotations

import logging
import sys
from typing import List, Optional


def build_server(argv: List[str]):
    import torch

    from .pipeline import PipelineConfig, parse_overrides
    from .policy import PI0Policy
    from .serving import LeRobotPolicyWrapper, WebsocketPolicyServer

    over = parse_overrides(argv)
    path = over.pop("config_path", None)
    if path is None:
        raise SystemExit("usage: python -m int_act_b200.serve --config_path <experiment.yaml> [--a.b value ...]")
    srv = {k.split(".", 1)[1]: over.pop(k) for k in list(over) if k.startswith("server.")}
    cfg = PipelineConfig.from_yaml(path, over)
    if cfg.eval_cfg is None:
        raise SystemExit("the experiment config has no eval_cfg section: nothing to serve")
    if cfg.eval_cfg.role != "server":
        raise SystemExit("eval_cfg.role = 'client' (the simulator side) is out of scope of this package; run the "
                         "reference's evaluator against this server")
    if cfg.model_cfg.type != "pi0":
        raise SystemExit(f"model type {cfg.model_cfg.type!r}: this package implements the pi0 path only")
    max_batch = int(srv.get("max_batch", 1))
    gpus = int(srv.get("gpus", 1))
    if gpus > 1 and torch.cuda.device_count() < gpus:
        raise SystemExit(f"--server.gpus {gpus} but only {torch.cuda.device_count()} CUDA devices are visible")
    wrappers = [LeRobotPolicyWrapper(pipeline_cfg=cfg, model_class=PI0Policy, max_batch=max_batch,
                                     device=f"cuda:{cfg.gpu_id + i}") for i in range(gpus)]
    server = WebsocketPolicyServer(policy=wrappers[0], host=cfg.eval_cfg.host, port=cfg.eval_cfg.port,
                                   metadata={"model": "pi0", "action_step": cfg.eval_cfg.action_step},
                                   max_batch=max_batch, batch_window_ms=float(srv.get("batch_window_ms", 0.0)),
                                   workers_per_replica=int(srv.get("workers_per_replica", 1)),
                                   replicas=wrappers[1:])
    model_path = srv.get("model_path")
    if model_path:
        for w in wrappers:
            w.switch_model(model_path)
            if max_batch > 1:
                w.warmup(max_batch)
    return server, wrappers, cfg


def main(argv: Optional[List[str]] = None) -> None:
    logging.basicConfig(level=logging.INFO)
    server, _, _ = build_server(sys.argv[1:] if argv is None else argv)
    server.serve_forever()


if __name__ == "__main__":
    main()
