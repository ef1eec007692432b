"""The slice of the reference's `TrainPipelineConfig` the policy server reads, without draccus / lerobot.

Reference: /root/reference/src/agent/configuration_pipeline.py:95-270 (EnvConfig, EvalConfig, TrainPipelineConfig and
its __post_init__, which derives `eval_cfg.env_adapter_path` from `simulator_name` + `env_adapter`) and the experiment
YAMLs (/root/reference/config/experiment/simpler/pi0_finetune_bridge_ev.yaml: `model_cfg: !include ...json`, `eval_cfg`,
`env`). Same field names and defaults; training-only sections are kept as plain dicts and ignored. Command-line
overrides use the reference's dotted form (`--eval_cfg.port 8001 --use_bf16 False`, doc/evaluation.md:119-152).
"""
from __future__ import annotations

import ast
import os
from types import SimpleNamespace
from typing import Any, Dict, List, Optional

import torch

from .config import PI0Config, load_yaml

EVAL_DEFAULTS = dict(simulator_name="simpler", env_adapter="BridgeSimplerAdapter", task_list=None, n_eval_episode=24,
                     n_video=24, n_parallel_eval=None, recording=True, pretrained_model_path=None,
                     pretrained_model_gradient_step_cnt=None, role="server", host="0.0.0.0", port=8000, action_step=4,
                     language_logic_chain=False, unnorm_key=None)
ENV_DEFAULTS = dict(dataset_statistics_path=None, image_size=(224, 224), action_normalization_type="bound",
                    state_normalization_type="bound")
TOP_DEFAULTS = dict(name=None, seed=42, debug=False, use_torch_compile=True, use_amp=True, use_bf16=True,
                    freeze_lm_head=True, gpu_id=0)


def _ns(d: Dict[str, Any]) -> SimpleNamespace:
    return SimpleNamespace(**d)


def _parse_scalar(text: str):
    low = text.strip().lower()
    if low in ("true", "false"):
        return low == "true"
    if low in ("none", "null"):
        return None
    try:
        return ast.literal_eval(text)
    except (ValueError, SyntaxError):
        return text


def parse_overrides(argv: List[str]) -> Dict[str, Any]:
    """`--a.b value` / `--a.b=value` pairs -> {"a.b": value} (draccus' command-line form)."""
    out, i = {}, 0
    while i < len(argv):
        tok = argv[i]
        if not tok.startswith("--"):
            raise ValueError(f"unexpected argument {tok!r}")
        if "=" in tok:
            k, v = tok[2:].split("=", 1)
            i += 1
        else:
            if i + 1 >= len(argv):
                raise ValueError(f"{tok} needs a value")
            k, v = tok[2:], argv[i + 1]
            i += 2
        out[k] = _parse_scalar(v)
    return out


class PipelineConfig(SimpleNamespace):
    """Attribute bag with the reference's names: .seed .use_bf16 .use_amp .freeze_lm_head .gpu_id .model_cfg (PI0Config)
    .eval_cfg (.env_adapter_path .action_step .host .port ...) .env (.dataset_statistics_path .image_size ...)
    .data (.dataset_stats)."""

    @classmethod
    def from_dict(cls, d: Dict[str, Any], base_dir: str = ".") -> "PipelineConfig":
        top = dict(TOP_DEFAULTS)
        top.update({k: v for k, v in d.items() if k in TOP_DEFAULTS})
        model = d.get("model_cfg") or {}
        model_cfg = model if isinstance(model, PI0Config) else PI0Config.from_dict(model)
        ev = d.get("eval_cfg")
        eval_cfg = None
        if ev is not None:
            e = dict(EVAL_DEFAULTS)
            e.update(ev)
            sim = e["simulator_name"]
            if sim is None:
                raise ValueError("Simulator name is not specified in the config.")
            # configuration_pipeline.py:245-256
            e["env_adapter_path"] = f"src.experiments.env_adapters.{sim}.{e['env_adapter'] or 'BridgeSimplerAdapter'}"
            e["simulator_path"] = f"src.experiments.envs.{sim}.{sim}_evaluator.{sim[:1].upper() + sim[1:]}Evaluator"
            if e["n_parallel_eval"] is not None:
                if sim != "simplerMS3":
                    raise ValueError("n_parallel_eval is only applicable for simplerMS3")
                if e["n_parallel_eval"] <= 1:
                    raise ValueError("n_parallel_eval should be greater than 1")
            eval_cfg = _ns(e)
        env = dict(ENV_DEFAULTS)
        env.update(d.get("env") or {})
        if env["dataset_statistics_path"] and not os.path.isabs(env["dataset_statistics_path"]) \
                and not os.path.exists(env["dataset_statistics_path"]):
            cand = os.path.normpath(os.path.join(base_dir, env["dataset_statistics_path"]))
            if os.path.exists(cand):
                env["dataset_statistics_path"] = cand
        data = d.get("data") or {}
        stats = {k: {s: torch.tensor(v) for s, v in st.items()} for k, st in (data.get("dataset_stats") or {}).items()}
        other = {k: v for k, v in d.items() if k not in TOP_DEFAULTS and k not in ("model_cfg", "eval_cfg", "env", "data")}
        return cls(**top, model_cfg=model_cfg, eval_cfg=eval_cfg, env=_ns(env),
                   data=_ns({**{k: v for k, v in data.items() if k != "dataset_stats"}, "dataset_stats": stats}),
                   extra=other)

    @classmethod
    def from_yaml(cls, path: str, overrides: Optional[Dict[str, Any]] = None, repo_root: Optional[str] = None):
        """`path`: an experiment YAML of the reference (`!include` resolved). `overrides`: dotted keys. Relative paths inside
        the YAML are relative to the reference checkout's root (`repo_root`, default: two levels above `config/`)."""
        d = load_yaml(path)
        for k, v in (overrides or {}).items():
            cur = d
            parts = k.split(".")
            for part in parts[:-1]:
                if cur.get(part) is None:
                    cur[part] = {}
                cur = cur[part]
            cur[parts[-1]] = v
        if repo_root is None:
            ap = os.path.abspath(path)
            repo_root = ap.split(os.sep + "config" + os.sep)[0] if os.sep + "config" + os.sep in ap else os.path.dirname(ap)
        return cls.from_dict(d, base_dir=repo_root)
