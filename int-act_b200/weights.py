"""Weights of the Pi0 policy: lerobot / HF-4.48.3 state-dict key layout in, packed device tensors out.

Key layout (roots confirmed by /root/reference/src/agent/trainer.py:110-123,235-236):
  model.paligemma_with_expert.paligemma.vision_tower.vision_model.*          SigLIP-So400m/14
  model.paligemma_with_expert.paligemma.multi_modal_projector.linear.*       projector
  model.paligemma_with_expert.paligemma.language_model.model.*               Gemma-2B
  model.paligemma_with_expert.gemma_expert.model.*                           action expert (Gemma-300M)
  model.{state_proj,action_in_proj,action_out_proj,action_time_mlp_in,action_time_mlp_out}.{weight,bias}

Packing (what the kernels consume, include/pi0b200.h `pi0_weights`):
  * q|k|v rows fused into one [out, in] matrix per layer; gate/up rows interleaved per 256-row tile
    (128 gate rows, then the 128 matching up rows) so one GEMM tile holds both halves of GeGLU;
  * RMSNorm gains stored as fp32 `1 + w`; biases / LayerNorm affine as fp32;
  * Conv2d(3,1152,14,14) flattened to [1152, 588] and zero-padded to K = 592 (TMA row pitch % 16 B);
  * the time half of action_time_mlp_in folded into a per-step bias: the 10 fp32 time values of the Euler loop
    are input-independent ([mirror] modeling_mvla.py:787-803), so `b + W[:, 1024:] . emb(t_k)` is a constant;
  * RoPE cos/sin table computed on the CPU with the reference's formula ([mirror] mllm.py:30-53).
"""
from __future__ import annotations

import ctypes as C
import math
import os
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch

from . import binding as pb

PG = "model.paligemma_with_expert.paligemma."
VT = PG + "vision_tower.vision_model."
LM = PG + "language_model.model."
EX = "model.paligemma_with_expert.gemma_expert.model."
PROJ = "model.paligemma_with_expert.paligemma.multi_modal_projector.linear."

VOCAB = 257152


@dataclass
class Dims:
    """Architecture constants; depth/vocab can be reduced for tests."""
    n_siglip_layers: int = 27
    n_gemma_layers: int = 18
    vocab: int = VOCAB
    vis_hidden: int = 1152
    vis_inter: int = 4304
    hidden: int = 2048
    inter: int = 16384
    exp_hidden: int = 1024
    exp_inter: int = 4096
    heads: int = 8
    head_dim: int = 256
    proj_width: int = 1024
    max_dim: int = 32


def state_dict_shapes(d: Dims) -> Dict[str, tuple]:
    s: Dict[str, tuple] = {}
    s[VT + "embeddings.patch_embedding.weight"] = (d.vis_hidden, 3, 14, 14)
    s[VT + "embeddings.patch_embedding.bias"] = (d.vis_hidden,)
    s[VT + "embeddings.position_embedding.weight"] = (256, d.vis_hidden)
    for i in range(d.n_siglip_layers):
        p = f"{VT}encoder.layers.{i}."
        for ln in ("layer_norm1", "layer_norm2"):
            s[p + ln + ".weight"] = (d.vis_hidden,)
            s[p + ln + ".bias"] = (d.vis_hidden,)
        for pr in ("q_proj", "k_proj", "v_proj", "out_proj"):
            s[p + f"self_attn.{pr}.weight"] = (d.vis_hidden, d.vis_hidden)
            s[p + f"self_attn.{pr}.bias"] = (d.vis_hidden,)
        s[p + "mlp.fc1.weight"] = (d.vis_inter, d.vis_hidden)
        s[p + "mlp.fc1.bias"] = (d.vis_inter,)
        s[p + "mlp.fc2.weight"] = (d.vis_hidden, d.vis_inter)
        s[p + "mlp.fc2.bias"] = (d.vis_hidden,)
    s[VT + "post_layernorm.weight"] = (d.vis_hidden,)
    s[VT + "post_layernorm.bias"] = (d.vis_hidden,)
    s[PROJ + "weight"] = (d.hidden, d.vis_hidden)
    s[PROJ + "bias"] = (d.hidden,)
    s[LM + "embed_tokens.weight"] = (d.vocab, d.hidden)
    for root, hid, inter in ((LM, d.hidden, d.inter), (EX, d.exp_hidden, d.exp_inter)):
        for i in range(d.n_gemma_layers):
            p = f"{root}layers.{i}."
            s[p + "input_layernorm.weight"] = (hid,)
            s[p + "self_attn.q_proj.weight"] = (d.heads * d.head_dim, hid)
            s[p + "self_attn.k_proj.weight"] = (d.head_dim, hid)
            s[p + "self_attn.v_proj.weight"] = (d.head_dim, hid)
            s[p + "self_attn.o_proj.weight"] = (hid, d.heads * d.head_dim)
            s[p + "post_attention_layernorm.weight"] = (hid,)
            s[p + "mlp.gate_proj.weight"] = (inter, hid)
            s[p + "mlp.up_proj.weight"] = (inter, hid)
            s[p + "mlp.down_proj.weight"] = (hid, inter)
        s[root + "norm.weight"] = (hid,)
    s["model.state_proj.weight"] = (d.proj_width, d.max_dim)
    s["model.state_proj.bias"] = (d.proj_width,)
    s["model.action_in_proj.weight"] = (d.proj_width, d.max_dim)
    s["model.action_in_proj.bias"] = (d.proj_width,)
    s["model.action_out_proj.weight"] = (d.max_dim, d.proj_width)
    s["model.action_out_proj.bias"] = (d.max_dim,)
    s["model.action_time_mlp_in.weight"] = (d.proj_width, 2 * d.proj_width)
    s["model.action_time_mlp_in.bias"] = (d.proj_width,)
    s["model.action_time_mlp_out.weight"] = (d.proj_width, d.proj_width)
    s["model.action_time_mlp_out.bias"] = (d.proj_width,)
    return s


def random_state_dict(d: Dims, device="cpu", seed: int = 0, dtype=torch.bfloat16) -> Dict[str, torch.Tensor]:
    """Deterministic random-init weights shared by the oracle and the CUDA path:
    matrices ~ N(0, std) with a fan-in-aware std that keeps activations O(1) through the stack, biases N(0, 0.01),
    norm gains N(0, 0.1) (Gemma uses 1 + w; LayerNorm weights are 1 + N(0, 0.1)). Values are rounded to `dtype`.
    Each tensor has its own generator seed so the values do not depend on the generation order."""
    dev = torch.device(device)
    out = {}
    for idx, (name, shape) in enumerate(state_dict_shapes(d).items()):
        g = torch.Generator(device=dev).manual_seed(seed * 1000003 + idx)
        if name.endswith("embed_tokens.weight"):
            t = torch.empty(shape, dtype=dtype, device=dev).normal_(0.0, 0.02, generator=g)
        elif len(shape) >= 2:
            fan_in = 1
            for v in shape[1:]:
                fan_in *= v
            std = 1.0 / math.sqrt(fan_in)
            if name.endswith("position_embedding.weight"):
                std = 0.02
            if name == "model.action_out_proj.weight":
                # keeps |v_t| < 1 like a trained policy on bound-normalised actions: one bf16 ulp of v_t at
                # |v_t| ~ 3 is 1.6e-2, i.e. above the 1e-2 action tolerance for ANY two bf16 implementations
                std *= 0.25
            t = torch.empty(shape, dtype=torch.float32, device=dev).normal_(0.0, std, generator=g).to(dtype)
        elif "layer_norm" in name and name.endswith(".weight") or name.endswith("post_layernorm.weight"):
            t = (1.0 + torch.empty(shape, dtype=torch.float32, device=dev).normal_(0.0, 0.1, generator=g)).to(dtype)
        elif name.endswith("layernorm.weight") or name.endswith("norm.weight"):
            t = torch.empty(shape, dtype=torch.float32, device=dev).normal_(0.0, 0.1, generator=g).to(dtype)
        else:  # biases
            t = torch.empty(shape, dtype=torch.float32, device=dev).normal_(0.0, 0.01, generator=g).to(dtype)
        out[name] = t
    return out


LM_HEAD = PG + "language_model.lm_head.weight"
NORM_ROOTS = ("normalize_inputs.", "normalize_targets.", "unnormalize_outputs.")


def split_checkpoint(sd: Dict[str, torch.Tensor]):
    """A lerobot checkpoint holds the module weights plus the Normalize / Unnormalize buffers
    (`normalize_inputs.buffer_<key with . -> _>.{mean,std,min,max}`, call sites [mirror] modeling_mvla.py:243-249).
    Returns (weights, buffers) with buffers = {root: {buffer_name: {stat: tensor}}}; the input dict is not modified."""
    weights, buffers = {}, {}
    for k, v in sd.items():
        root = next((r for r in NORM_ROOTS if k.startswith(r + "buffer_")), None)
        if root is None:
            weights[k] = v
            continue
        name, stat = k[len(root) + len("buffer_"):].rsplit(".", 1)
        buffers.setdefault(root[:-1], {}).setdefault(name, {})[stat] = v
    return weights, buffers


def stats_from_buffers(buffers, input_features, output_features):
    """dataset_stats {feature key: {stat: tensor}} from the checkpoint's normalisation buffers; buffers still at their
    +inf placeholder (a checkpoint saved before any statistics were set) are skipped."""
    stats = {}
    for root, feats in (("normalize_inputs", input_features), ("unnormalize_outputs", output_features),
                        ("normalize_targets", output_features)):
        for key in feats:
            b = buffers.get(root, {}).get(key.replace(".", "_"))
            if b and all(torch.isfinite(t.float()).all() for t in b.values()):
                stats.setdefault(key, {}).update({k: t.detach().float().cpu() for k, t in b.items()})
    return stats


def canonical_state_dict(sd: Dict[str, torch.Tensor], strict: bool = False) -> Dict[str, torch.Tensor]:
    """Checks a checkpoint against the key / shape table of the architecture and returns the weights the packer needs.
    * `...language_model.lm_head.weight` stands in for a missing `embed_tokens.weight`: PaliGemma ties the two and
      safetensors keeps only one name of a shared tensor (a checkpoint saved with the head still attached);
    * normalisation buffers are dropped (see split_checkpoint / stats_from_buffers);
    * missing keys and shape mismatches raise with the offending names; unexpected keys raise only when `strict`
      (the reference loads with strict=False, policy_wrapper.py:159-162)."""
    sd, _ = split_checkpoint(sd)
    emb = LM + "embed_tokens.weight"
    if emb not in sd and LM_HEAD in sd:
        sd = dict(sd)
        sd[emb] = sd[LM_HEAD]
    if emb not in sd:
        raise KeyError(f"checkpoint holds neither {emb} nor {LM_HEAD}")
    expected = state_dict_shapes(infer_dims(sd))
    missing = [k for k in expected if k not in sd]
    if missing:
        raise KeyError(f"checkpoint is missing {len(missing)} tensors, e.g. {missing[:4]}")
    bad = [(k, tuple(sd[k].shape), expected[k]) for k in expected if tuple(sd[k].shape) != tuple(expected[k])]
    if bad:
        raise ValueError(f"{len(bad)} tensors have unexpected shapes, e.g. {bad[:3]}")
    extra = [k for k in sd if k not in expected and k != LM_HEAD and not k.endswith("lm_head.weight")]
    if strict and extra:
        raise KeyError(f"unexpected keys in checkpoint (strict=True), e.g. {extra[:4]}")
    return sd


def infer_dims(sd: Dict[str, torch.Tensor]) -> Dims:
    n_sig = 0
    while f"{VT}encoder.layers.{n_sig}.layer_norm1.weight" in sd:
        n_sig += 1
    n_gem = 0
    while f"{LM}layers.{n_gem}.input_layernorm.weight" in sd:
        n_gem += 1
    emb = sd[LM + "embed_tokens.weight"] if LM + "embed_tokens.weight" in sd else sd[LM_HEAD]
    return Dims(n_siglip_layers=n_sig, n_gemma_layers=n_gem, vocab=emb.shape[0])


def euler_times(num_steps: int) -> List[float]:
    """fp32-accumulated time values of the sampling loop ([mirror] modeling_mvla.py:787-803)."""
    dt = torch.tensor(-1.0 / num_steps, dtype=torch.float32)
    time = torch.tensor(1.0, dtype=torch.float32)
    out = []
    while time >= -dt / 2:
        out.append(float(time))
        time = time + dt
    return out


def sinusoidal_time_embedding(times: List[float], dimension: int = 1024, min_period: float = 4e-3,
                              max_period: float = 4.0) -> torch.Tensor:
    """[mirror] modeling_mvla.py:71-89, float64 on the CPU; `times` are the fp32 time values."""
    t = torch.tensor(times, dtype=torch.float32)
    fraction = torch.linspace(0.0, 1.0, dimension // 2, dtype=torch.float64)
    period = min_period * (max_period / min_period) ** fraction
    scaling = 1.0 / period * 2 * math.pi
    x = scaling[None, :] * t[:, None]
    return torch.cat([torch.sin(x), torch.cos(x)], dim=1)


def rope_table(max_pos: int, head_dim: int = 256, max_wavelength: float = 10_000.0):
    """cos/sin of pos / 10000^(2i/d) exactly as [mirror] mllm.py:30-53 computes them (fp32, CPU)."""
    d_half = head_dim // 2
    freq_exponents = (2.0 / head_dim) * torch.arange(d_half, dtype=torch.float32)
    timescale = max_wavelength ** freq_exponents
    pos = torch.arange(max_pos, dtype=torch.int64)
    radians = pos[:, None].to(torch.float32) / timescale[None, :].to(torch.float32)
    return torch.cos(radians).contiguous(), torch.sin(radians).contiguous()


class PackedWeights:
    """Device tensors in the layout of `pi0_weights` plus the ctypes struct pointing at them."""

    def __init__(self, sd: Dict[str, torch.Tensor], device, act_dtype=torch.bfloat16, num_steps: int = 10,
                 max_pos: int = 2048, free_source: bool = False):
        self.device = torch.device(device)
        self.act = act_dtype
        if LM + "embed_tokens.weight" not in sd and LM_HEAD in sd:  # tied head saved under its other name
            sd = dict(sd)
            sd[LM + "embed_tokens.weight"] = sd[LM_HEAD]
        self.dims = infer_dims(sd)
        d = self.dims
        self.keep: List[torch.Tensor] = []
        self.struct = pb.Weights()
        w = self.struct

        def A(t):  # activation-typed matrix on the device
            t = t.detach().to(self.device).to(self.act).contiguous()
            self.keep.append(t)
            return t

        def Fp(t):  # fp32 vector; in the bf16 regime the parameter is bf16-rounded first (model.to(bf16))
            t = t.detach().to(self.device).to(self.act).float().contiguous()
            self.keep.append(t)
            return t

        def P(t):
            return C.c_void_p(t.data_ptr())

        def get(k):
            t = sd[k]
            if free_source:
                del sd[k]
            return t

        pw = get(VT + "embeddings.patch_embedding.weight").detach().to(self.device).to(self.act).reshape(d.vis_hidden, 588)
        pwp = torch.zeros((d.vis_hidden, 592), dtype=self.act, device=self.device)
        pwp[:, :588] = pw
        self.keep.append(pwp)
        w.patch_w = P(pwp)
        w.patch_b = P(Fp(get(VT + "embeddings.patch_embedding.bias")))
        w.pos_emb = P(A(get(VT + "embeddings.position_embedding.weight")))
        for i in range(d.n_siglip_layers):
            p = f"{VT}encoder.layers.{i}."
            lw = w.sig[i]
            lw.ln1_w = P(Fp(get(p + "layer_norm1.weight")))
            lw.ln1_b = P(Fp(get(p + "layer_norm1.bias")))
            lw.wqkv = P(A(torch.cat([get(p + f"self_attn.{n}_proj.weight").to(self.device) for n in "qkv"], dim=0)))
            lw.bqkv = P(Fp(torch.cat([get(p + f"self_attn.{n}_proj.bias").to(self.device) for n in "qkv"], dim=0)))
            lw.wo = P(A(get(p + "self_attn.out_proj.weight")))
            lw.bo = P(Fp(get(p + "self_attn.out_proj.bias")))
            lw.ln2_w = P(Fp(get(p + "layer_norm2.weight")))
            lw.ln2_b = P(Fp(get(p + "layer_norm2.bias")))
            lw.wfc1 = P(A(get(p + "mlp.fc1.weight")))
            lw.bfc1 = P(Fp(get(p + "mlp.fc1.bias")))
            lw.wfc2 = P(A(get(p + "mlp.fc2.weight")))
            lw.bfc2 = P(Fp(get(p + "mlp.fc2.bias")))
        w.post_ln_w = P(Fp(get(VT + "post_layernorm.weight")))
        w.post_ln_b = P(Fp(get(VT + "post_layernorm.bias")))
        w.proj_w = P(A(get(PROJ + "weight")))
        w.proj_b = P(Fp(get(PROJ + "bias")))
        emb = A(get(LM + "embed_tokens.weight"))
        w.embed = P(emb)
        w.vocab = emb.shape[0]
        for root, arr in ((LM, w.gem), (EX, w.exp)):
            for i in range(d.n_gemma_layers):
                p = f"{root}layers.{i}."
                lw = arr[i]
                lw.g1 = P(self._gain(get(p + "input_layernorm.weight")))
                lw.wqkv = P(A(torch.cat([get(p + f"self_attn.{n}_proj.weight").to(self.device) for n in "qkv"], dim=0)))
                lw.wo = P(A(get(p + "self_attn.o_proj.weight")))
                lw.g2 = P(self._gain(get(p + "post_attention_layernorm.weight")))
                wg = get(p + "mlp.gate_proj.weight").to(self.device)
                wu = get(p + "mlp.up_proj.weight").to(self.device)
                K = wg.shape[1]
                wgu = torch.stack([wg.reshape(-1, 128, K), wu.reshape(-1, 128, K)], dim=1).reshape(2 * wg.shape[0], K)
                del wg, wu
                lw.wgu = P(A(wgu))
                del wgu
                lw.wdown = P(A(get(p + "mlp.down_proj.weight")))
        w.exp_norm_g = P(self._gain(get(EX + "norm.weight")))
        w.state_w = P(A(get("model.state_proj.weight")))
        w.state_b = P(Fp(get("model.state_proj.bias")))
        w.action_in_w = P(A(get("model.action_in_proj.weight")))
        w.action_in_b = P(Fp(get("model.action_in_proj.bias")))
        w_in = get("model.action_time_mlp_in.weight").detach().to(self.device).to(self.act)
        b_in = get("model.action_time_mlp_in.bias").detach().to(self.device).to(self.act).float()
        w.time_in_wa = P(A(w_in[:, : d.proj_width]))
        temb = sinusoidal_time_embedding(euler_times(num_steps), d.proj_width).to(self.device)
        temb = temb.to(self.act).double()  # time_emb.type(dtype): float64 -> activation dtype in one rounding
        tb = (b_in.double()[None, :] + temb @ w_in[:, d.proj_width:].double().t()).float().contiguous()
        self.keep.append(tb)
        self.time_bias = tb
        w.time_bias = P(tb)
        w.time_out_w = P(A(get("model.action_time_mlp_out.weight")))
        w.time_out_b = P(Fp(get("model.action_time_mlp_out.bias")))
        w.action_out_w = P(A(get("model.action_out_proj.weight")))
        w.action_out_b = P(Fp(get("model.action_out_proj.bias")))
        cos, sin = rope_table(max_pos)
        self.rope_cos = cos.to(self.device)
        self.rope_sin = sin.to(self.device)
        w.rope_cos = P(self.rope_cos)
        w.rope_sin = P(self.rope_sin)
        w.max_pos = max_pos
        self.num_steps = num_steps

    def _gain(self, t):
        g = (1.0 + t.detach().to(self.device).to(self.act).float()).contiguous()
        self.keep.append(g)
        return g

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in self.keep)


def load_safetensors(path: str, device="cpu") -> Dict[str, torch.Tensor]:
    """Reads `<path>/model.safetensors` (what lerobot's from_pretrained reads;
    /root/reference/src/experiments/policies/policy_wrapper.py:159-162)."""
    from safetensors.torch import load_file

    f = os.path.join(path, "model.safetensors") if os.path.isdir(path) else path
    return load_file(f, device=str(device))


def save_safetensors(sd: Dict[str, torch.Tensor], path: str):
    from safetensors.torch import save_file

    os.makedirs(path, exist_ok=True)
    save_file({k: v.contiguous() for k, v in sd.items()}, os.path.join(path, "model.safetensors"))
