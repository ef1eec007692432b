"""`PI0Config`: the constructor contract of the reference policy, without lerobot / draccus.

Accepts every key of /root/reference/config/models/pi0_finetune_bridge.json:1-65 and config/models/hf_pi0.json
(training-only keys `optimizer_*`, `scheduler_*`, `device`, `use_amp`, ... are stored and ignored by inference).
Validation rules follow [mirror] /root/reference/src/model/mvla/configuration_mvla.py:106-124.
`load_yaml` understands the `!include` tag the experiment YAMLs use
(/root/reference/config/experiment/simpler/pi0_finetune_bridge_ev.yaml:3).
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple


@dataclass
class PolicyFeature:
    type: str
    shape: Tuple[int, ...]


def _features(d: Optional[Dict[str, Any]]) -> Dict[str, PolicyFeature]:
    out = {}
    for k, v in (d or {}).items():
        if isinstance(v, PolicyFeature):
            out[k] = v
        else:
            out[k] = PolicyFeature(type=str(v["type"]), shape=tuple(int(s) for s in v["shape"]))
    return out


@dataclass
class PI0Config:
    type: str = "pi0"
    n_obs_steps: int = 1
    normalization_mapping: Dict[str, str] = field(
        default_factory=lambda: {"VISUAL": "IDENTITY", "STATE": "MEAN_STD", "ACTION": "MEAN_STD"})
    input_features: Dict[str, PolicyFeature] = field(default_factory=dict)
    output_features: Dict[str, PolicyFeature] = field(default_factory=dict)
    chunk_size: int = 50
    n_action_steps: int = 50
    max_state_dim: int = 32
    max_action_dim: int = 32
    resize_imgs_with_padding: Optional[Tuple[int, int]] = (224, 224)
    empty_cameras: int = 0
    adapt_to_pi_aloha: bool = False
    use_delta_joint_actions_aloha: bool = False
    tokenizer_max_length: int = 48
    proj_width: int = 1024
    num_steps: int = 10
    use_cache: bool = True
    attention_implementation: str = "eager"  # accepted and ignored: there is one sm_100a path
    freeze_vision_encoder: bool = True
    train_expert_only: bool = False
    train_state_proj: bool = True
    paligemma_pretrained_path: Optional[str] = None  # fork-only key; only None is supported
    device: Optional[str] = None
    use_amp: bool = False
    # the adapters hand over images already in [-1, 1]; the mirrored prepare_images still applies
    # img * 2 - 1. Default keeps the mirrored behaviour.
    images_pre_normalized: bool = False
    extra: Dict[str, Any] = field(default_factory=dict)  # optimizer_*, scheduler_* and any unknown key

    def __post_init__(self):
        self.input_features = _features(self.input_features)
        self.output_features = _features(self.output_features)
        if self.resize_imgs_with_padding is not None:
            self.resize_imgs_with_padding = tuple(int(v) for v in self.resize_imgs_with_padding)
        if self.n_action_steps > self.chunk_size:
            raise ValueError(
                "The chunk size is the upper bound for the number of action steps per model invocation. Got "
                f"{self.n_action_steps} for `n_action_steps` and {self.chunk_size} for `chunk_size`.")
        if self.n_obs_steps != 1:
            raise ValueError(f"Multiple observation steps not handled yet. Got `nobs_steps={self.n_obs_steps}`")
        if self.use_delta_joint_actions_aloha:
            raise NotImplementedError(
                "`use_delta_joint_actions_aloha` is used by pi0 for aloha real models. It is not ported yet in LeRobot.")
        if self.attention_implementation not in ("eager", "fa2", "flex"):
            raise ValueError(f"attention_implementation={self.attention_implementation!r} not in eager/fa2/flex")
        if self.max_state_dim != 32 or self.max_action_dim != 32 or self.proj_width != 1024:
            raise ValueError("the sm_100a kernels are built for max_state_dim = max_action_dim = 32, proj_width = 1024")

    # ---- lerobot PreTrainedConfig-style accessors -------------------------------------------------------
    @property
    def image_features(self) -> Dict[str, PolicyFeature]:
        return {k: f for k, f in self.input_features.items() if f.type == "VISUAL"}

    @property
    def robot_state_feature(self) -> Optional[PolicyFeature]:
        for k, f in self.input_features.items():
            if f.type == "STATE" and k == "observation.state":
                return f
        return None

    @property
    def action_feature(self) -> Optional[PolicyFeature]:
        for k, f in self.output_features.items():
            if f.type == "ACTION" and k == "action":
                return f
        return None

    @property
    def n_cams(self) -> int:
        """Cameras the prefix is built for: every VISUAL input feature, the `empty_camera_<i>` placeholders that
        validate_features adds included."""
        return len(self.image_features)

    def validate_features(self) -> None:
        """lerobot's PI0Config.validate_features: `empty_cameras` placeholder cameras become input features named
        `observation.images.empty_camera_<i>`; they are never in a batch, so prepare_images fills them with -1 images and
        a zero mask ([mirror] modeling_mvla.py:375-383). Idempotent (a config.json that already lists them is fine)."""
        for i in range(int(self.empty_cameras)):
            key = f"observation.images.empty_camera_{i}"
            if key not in self.input_features:
                self.input_features[key] = PolicyFeature(type="VISUAL", shape=(3, 480, 640))

    # ---- construction -------------------------------------------------------------------------------------
    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "PI0Config":
        known = {f for f in cls.__dataclass_fields__ if f != "extra"}
        kw = {k: v for k, v in d.items() if k in known}
        extra = {k: v for k, v in d.items() if k not in known}
        cfg = cls(**kw)
        cfg.extra = extra
        return cfg

    @classmethod
    def from_json(cls, path: str) -> "PI0Config":
        with open(path) as f:
            return cls.from_dict(json.load(f))

    @classmethod
    def from_pretrained(cls, path: str) -> "PI0Config":
        return cls.from_json(os.path.join(path, "config.json") if os.path.isdir(path) else path)

    def to_dict(self) -> Dict[str, Any]:
        d = {k: getattr(self, k) for k in self.__dataclass_fields__ if k not in ("extra", "input_features",
                                                                               "output_features")}
        d["input_features"] = {k: {"type": f.type, "shape": list(f.shape)} for k, f in self.input_features.items()}
        d["output_features"] = {k: {"type": f.type, "shape": list(f.shape)} for k, f in self.output_features.items()}
        d.update(self.extra)
        return d


def bridge_config(**over) -> PI0Config:
    """The INTACT-pi0-finetune-bridge architecture (config/models/pi0_finetune_bridge.json)."""
    d = dict(
        normalization_mapping={"VISUAL": "IDENTITY", "STATE": "IDENTITY", "ACTION": "IDENTITY"},
        input_features={"observation.images.top": {"shape": [3, 224, 224], "type": "VISUAL"},
                        "observation.state": {"shape": [7], "type": "STATE"}},
        output_features={"action": {"type": "ACTION", "shape": [7]}},
        chunk_size=4, n_action_steps=4, tokenizer_max_length=72, num_steps=10, freeze_vision_encoder=False,
        use_amp=True, device="cpu")
    d.update(over)
    return PI0Config.from_dict(d)


def libero_config(**over) -> PI0Config:
    """BASELINE config 5 shape: third-person + wrist camera, horizon 16."""
    d = dict(
        normalization_mapping={"VISUAL": "IDENTITY", "STATE": "IDENTITY", "ACTION": "IDENTITY"},
        input_features={"observation.images.top": {"shape": [3, 224, 224], "type": "VISUAL"},
                        "observation.images.wrist": {"shape": [3, 224, 224], "type": "VISUAL"},
                        "observation.state": {"shape": [8], "type": "STATE"}},
        output_features={"action": {"type": "ACTION", "shape": [7]}},
        chunk_size=16, n_action_steps=16, tokenizer_max_length=72, num_steps=10)
    d.update(over)
    return PI0Config.from_dict(d)


# ---- YAML with !include (draccus-free) ---------------------------------------------------------------------
def load_yaml(path: str) -> Dict[str, Any]:
    import yaml

    base = os.path.dirname(os.path.abspath(path))

    class Loader(yaml.SafeLoader):
        pass

    def include(loader, node):
        target = os.path.join(base, loader.construct_scalar(node))
        if target.endswith(".json"):
            with open(target) as f:
                return json.load(f)
        return load_yaml(target)

    Loader.add_constructor("!include", include)
    with open(path) as f:
        return yaml.load(f, Loader=Loader)
