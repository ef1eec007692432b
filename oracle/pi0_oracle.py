"""CPU oracle of the Pi0 `infer_action` hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module; the product package (int-act_b200/) never does.

What it is: a self-contained, pure-PyTorch restatement of the algorithm the reference runs for
`PI0Policy.select_action -> PI0FlowMatching.sample_actions`. The reference's own
implementation lives in the un-vendored submodule `lerobot` (IrvingF7/lerobot, branch `pi0`, commit
35f6e02315dcb3fa25f3a740478265c6c793a95e) on top of `transformers==4.48.3`; neither is present in
/root/reference or importable here, so every function below cites the in-tree derivative ("[mirror]",
/root/reference/src/model/mvla/*.py) that carries the same arithmetic.

PARITY STATUS: the integer/bool/indexing helpers (masks, position ids) and the closed-form float helpers
(RoPE, sinusoidal time embedding, pad_vector, resize_with_pad, the fp32 time sequence) are PINNED: they are
checked bit-exactly against the reference's own functions executed from /root/reference by
tests/golden/make_golden.py (fixtures in tests/golden/*.npz). The neural-network forward (SigLIP, Gemma,
expert) is "PARITY UNPINNED": the reference holds no test, golden vector or runnable implementation for it
(SURVEY.md F1/F5), so it is a restatement of the published architecture anchored on the mirror's call sites.

Two dtype regimes:
  * "fp32": all parameters and activations fp32 (the shipped eval regime, --use_bf16 False).
  * "bf16": parameters bf16, torch.autocast(bf16) semantics emulated on the CPU: every Linear/matmul/conv
    takes bf16-rounded inputs, accumulates in fp32 and rounds its output to bf16 once; LayerNorm / softmax
    run in fp32; RMSNorm, RoPE, GELU, SiLU compute in fp32 and round to bf16; x_t/time stay fp32.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
BIG_NEG = -2.3819763e38  # [mirror] mllm.py:301


@dataclass
class OracleDims:
    """Architecture constants ([mirror] mllm.py:86-110, expert.py:70-93)."""
    n_cams: int = 1
    lang_len: int = 72
    horizon: int = 4
    max_state_dim: int = 32
    max_action_dim: int = 32
    num_steps: int = 10
    n_siglip_layers: int = 27
    n_gemma_layers: int = 18
    # SigLIP-So400m/14
    vis_hidden: int = 1152
    vis_inter: int = 4304
    vis_heads: int = 16
    patch: int = 14
    image: int = 224
    # Gemma-2B / expert
    hidden: int = 2048
    inter: int = 16384
    exp_hidden: int = 1024
    exp_inter: int = 4096
    heads: int = 8
    head_dim: int = 256
    proj_width: int = 1024
    eps: float = 1e-6

    @property
    def n_patches(self):
        return (self.image // self.patch) ** 2

    @property
    def prefix_len(self):
        return self.n_cams * self.n_patches + self.lang_len


PG = "model.paligemma_with_expert.paligemma."
VT = PG + "vision_tower.vision_model."
LM = PG + "language_model.model."
EX = "model.paligemma_with_expert.gemma_expert.model."


# ------------------------------------------------------------------------------------------------------
# pinned helpers (bit-exact against the reference's own functions; see tests/golden/make_golden.py)
# ------------------------------------------------------------------------------------------------------
def make_att_2d_masks(pad_masks: Tensor, att_masks: Tensor) -> Tensor:
    """[mirror] modeling_mvla.py:98-128. mask[b,i,j] = cum[j] <= cum[i] and pad[i] and pad[j]."""
    cumsum = torch.cumsum(att_masks, dim=1)
    att_2d = cumsum[:, None, :] <= cumsum[:, :, None]
    pad_2d = pad_masks[:, None, :] * pad_masks[:, :, None]
    return att_2d & pad_2d


def prefix_position_ids(pad_masks: Tensor) -> Tensor:
    """[mirror] modeling_mvla.py:766: cumsum(pad) - 1."""
    return torch.cumsum(pad_masks, dim=1) - 1


def suffix_position_ids(prefix_pad_masks: Tensor, suffix_pad_masks: Tensor) -> Tensor:
    """[mirror] modeling_mvla.py:820-821 pattern: prefix_offsets + cumsum(suffix_pad) - 1."""
    prefix_offsets = torch.sum(prefix_pad_masks, dim=-1)[:, None]
    return prefix_offsets + torch.cumsum(suffix_pad_masks, dim=1) - 1


def create_sinusoidal_pos_embedding(time: Tensor, dimension: int, min_period: float, max_period: float) -> Tensor:
    """[mirror] modeling_mvla.py:71-89 (float64 on cpu/cuda)."""
    fraction = torch.linspace(0.0, 1.0, dimension // 2, dtype=torch.float64, device=time.device)
    period = min_period * (max_period / min_period) ** fraction
    scaling_factor = 1.0 / period * 2 * math.pi
    sin_input = scaling_factor[None, :] * time[:, None]
    return torch.cat([torch.sin(sin_input), torch.cos(sin_input)], dim=1)


def apply_rope(x: Tensor, positions: Tensor, max_wavelength: float = 10_000) -> Tensor:
    """[mirror] mllm.py:30-53. x [B,L,H,D], positions [B,L]; fp32 inside, half-split rotation."""
    d_half = x.shape[-1] // 2
    dtype = x.dtype
    x = x.to(torch.float32)
    freq_exponents = (2.0 / x.shape[-1]) * torch.arange(d_half, dtype=torch.float32, device=x.device)
    timescale = max_wavelength ** freq_exponents
    radians = positions[..., None].to(torch.float32) / timescale[None, None, :].to(torch.float32)
    radians = radians[..., None, :]
    sin, cos = torch.sin(radians), torch.cos(radians)
    x1, x2 = x.split(d_half, dim=-1)
    res = torch.empty_like(x)
    res[..., :d_half] = x1 * cos - x2 * sin
    res[..., d_half:] = x2 * cos + x1 * sin
    return res.to(dtype)


def pad_vector(vector: Tensor, new_dim: int) -> Tensor:
    """[mirror] modeling_mvla.py:153-164."""
    if vector.shape[-1] == new_dim:
        return vector
    shape = list(vector.shape)
    cur = shape[-1]
    shape[-1] = new_dim
    out = torch.zeros(*shape, dtype=vector.dtype, device=vector.device)
    out[..., :cur] = vector
    return out


def resize_with_pad(img: Tensor, width: int, height: int, pad_value: float = -1) -> Tensor:
    """[mirror] modeling_mvla.py:131-150 (bilinear, pad left/top)."""
    cur_h, cur_w = img.shape[2:]
    ratio = max(cur_w / width, cur_h / height)
    rh, rw = int(cur_h / ratio), int(cur_w / ratio)
    resized = F.interpolate(img, size=(rh, rw), mode="bilinear", align_corners=False)
    ph, pw = max(0, int(height - rh)), max(0, int(width - rw))
    return F.pad(resized, (pw, 0, ph, 0), value=pad_value)


def euler_times(num_steps: int) -> List[float]:
    """The fp32-accumulated time values of the sampling loop ([mirror] modeling_mvla.py:787-803)."""
    dt = torch.tensor(-1.0 / num_steps, dtype=torch.float32)
    time = torch.tensor(1.0, dtype=torch.float32)
    out = []
    while time >= -dt / 2:
        out.append(float(time))
        time = time + dt
    return out


# ------------------------------------------------------------------------------------------------------
# the model
# ------------------------------------------------------------------------------------------------------
class Pi0Oracle:
    def __init__(self, dims: OracleDims, state_dict: Dict[str, Tensor], mode: str = "fp32",
                 images_pre_normalized: bool = False):
        assert mode in ("fp32", "bf16")
        self.d = dims
        self.mode = mode
        self.images_pre_normalized = images_pre_normalized
        self.act = torch.float32 if mode == "fp32" else torch.bfloat16
        # parameters: fp32 regime keeps them as given (cast to fp32); bf16 regime rounds them to bf16 (model.to(bf16))
        self.w = {k: (v.detach().to("cpu").to(self.act).float() if v.is_floating_point() else v.detach().to("cpu"))
                  for k, v in state_dict.items()}
        self.trace: Optional[Dict[str, Tensor]] = None

    # -- rounding helpers ---------------------------------------------------------------------------
    def r(self, x: Tensor) -> Tensor:
        """Round to the activation dtype, keep fp32 storage."""
        return x if self.mode == "fp32" else x.to(torch.bfloat16).float()

    def lin(self, x: Tensor, wname: str, bname: Optional[str] = None) -> Tensor:
        """nn.Linear under the regime: inputs rounded to the activation type, fp32 accumulate, one rounding."""
        b = self.w[bname] if bname is not None else None
        return self.r(F.linear(self.r(x), self.w[wname], b))

    def _t(self, name: str, x: Tensor):
        if self.trace is not None:
            self.trace[name] = x.detach().clone()

    # -- SigLIP + projector: [mirror] mllm.py:182-187 -> HF 4.48.3 get_image_features ---------
    def layer_norm(self, x: Tensor, prefix: str) -> Tensor:
        return F.layer_norm(x, (x.shape[-1],), self.w[prefix + ".weight"], self.w[prefix + ".bias"], self.d.eps)

    def embed_image(self, pixels: Tensor) -> Tensor:
        d = self.d
        B = pixels.shape[0]
        conv_w = self.w[VT + "embeddings.patch_embedding.weight"]
        x = self.r(F.conv2d(self.r(pixels.float()), conv_w, self.w[VT + "embeddings.patch_embedding.bias"],
                            stride=d.patch))
        x = x.flatten(2).transpose(1, 2)  # [B, 256, 1152]
        x = self.r(x + self.w[VT + "embeddings.position_embedding.weight"][None])
        self._t("siglip.embed", x)
        hd = d.vis_hidden // d.vis_heads
        scale = hd ** -0.5
        for i in range(d.n_siglip_layers):
            p = f"{VT}encoder.layers.{i}."
            h = self.layer_norm(x, p + "layer_norm1")
            q = self.lin(h, p + "self_attn.q_proj.weight", p + "self_attn.q_proj.bias")
            k = self.lin(h, p + "self_attn.k_proj.weight", p + "self_attn.k_proj.bias")
            v = self.lin(h, p + "self_attn.v_proj.weight", p + "self_attn.v_proj.bias")
            q = q.view(B, -1, d.vis_heads, hd).transpose(1, 2)
            k = k.view(B, -1, d.vis_heads, hd).transpose(1, 2)
            v = v.view(B, -1, d.vis_heads, hd).transpose(1, 2)
            att = self.r(self.r(torch.matmul(q, k.transpose(2, 3))) * scale)
            probs = self.r(torch.softmax(att, dim=-1, dtype=torch.float32))
            o = self.r(torch.matmul(probs, v)).transpose(1, 2).reshape(B, -1, d.vis_hidden)
            o = self.lin(o, p + "self_attn.out_proj.weight", p + "self_attn.out_proj.bias")
            x = self.r(x + o)
            h = self.layer_norm(x, p + "layer_norm2")
            h = self.lin(h, p + "mlp.fc1.weight", p + "mlp.fc1.bias")
            h = self.r(F.gelu(h, approximate="tanh"))
            h = self.lin(h, p + "mlp.fc2.weight", p + "mlp.fc2.bias")
            x = self.r(x + h)
            self._t(f"siglip.layer{i}", x)
        x = self.layer_norm(x, VT + "post_layernorm")
        x = self.lin(x, PG + "multi_modal_projector.linear.weight", PG + "multi_modal_projector.linear.bias")
        x = self.r(x / (d.hidden ** 0.5))  # HF 4.48.3 get_image_features
        return x

    # -- prefix: [mirror] modeling_mvla.py:582-624,636-641 (metaquery lines 625-634 are MVLA-only) ----------------
    def embed_prefix(self, images: List[Tensor], img_masks: List[Tensor], lang_tokens: Tensor, lang_masks: Tensor):
        d = self.d
        embs, pads = [], []
        for img, m in zip(images, img_masks):
            e = self.embed_image(img)
            e = self.r(e)  # .to(bf16)
            scale = torch.tensor(d.hidden ** 0.5, dtype=self.act).float()
            e = self.r(e * scale)
            embs.append(e)
            pads.append(m[:, None].expand(e.shape[0], e.shape[1]))
        lang = self.w[LM + "embed_tokens.weight"][lang_tokens]
        lang = self.r(lang * math.sqrt(d.hidden))
        embs.append(lang)
        pads.append(lang_masks)
        embs = torch.cat(embs, dim=1)
        pad_masks = torch.cat(pads, dim=1).bool()
        att_masks = torch.zeros_like(pad_masks)
        self._t("prefix.embs", embs)
        return embs, pad_masks, att_masks

    # -- shared Gemma layer pieces: [mirror] mllm.py:206-245, expert.py:128-179,387-409 -----------------------------
    def rms_norm(self, x: Tensor, wname: str) -> Tensor:
        xf = x.float()
        out = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.d.eps)
        return self.r(out * (1.0 + self.w[wname]))

    def attention(self, mask: Tensor, q: Tensor, k: Tensor, v: Tensor) -> Tensor:
        """[mirror] mllm.py:265-317. q [B,T,8,256], k,v [B,Tk,1,256], mask bool [B,T,Tk]."""
        d = self.d
        B, T = q.shape[:2]
        qh = q.transpose(1, 2)  # [B,8,T,256]
        kh = k.transpose(1, 2)  # [B,1,Tk,256]
        vh = v.transpose(1, 2)
        att = self.r(torch.matmul(qh, kh.transpose(2, 3)))
        att = self.r(att * d.head_dim ** -0.5)
        att = torch.where(mask[:, None, :, :], att, torch.tensor(BIG_NEG))
        probs = self.r(torch.softmax(att, dim=-1))
        out = self.r(torch.matmul(probs, vh))  # [B,8,T,256]
        return out.permute(0, 2, 1, 3).reshape(B, T, d.heads * d.head_dim)

    def gemma_layer(self, root: str, i: int, x: Tensor, pos: Tensor, mask: Tensor, cache: Optional[dict],
                    fill: bool, skip_tail: bool = False):
        d = self.d
        p = f"{root}layers.{i}."
        B, T = x.shape[:2]
        h = self.rms_norm(x, p + "input_layernorm.weight")
        q = self.lin(h, p + "self_attn.q_proj.weight").view(B, T, d.heads, d.head_dim)
        k = self.lin(h, p + "self_attn.k_proj.weight").view(B, T, 1, d.head_dim)
        v = self.lin(h, p + "self_attn.v_proj.weight").view(B, T, 1, d.head_dim)
        q = self.r(apply_rope(q, pos))
        k = self.r(apply_rope(k, pos))
        if fill:
            cache[i] = {"key_states": k, "value_states": v}
            if skip_tail:
                return x
        else:
            k = torch.cat([cache[i]["key_states"], k], dim=1)
            v = torch.cat([cache[i]["value_states"], v], dim=1)
        att = self.attention(mask, q, k, v)
        self._t(f"{root}layer{i}.att", att)
        o = self.lin(att, p + "self_attn.o_proj.weight")
        y = self.r(o + x)
        h = self.rms_norm(y, p + "post_attention_layernorm.weight")
        g = self.r(F.gelu(self.lin(h, p + "mlp.gate_proj.weight"), approximate="tanh"))
        u = self.lin(h, p + "mlp.up_proj.weight")
        m = self.lin(self.r(g * u), p + "mlp.down_proj.weight")
        out = self.r(m + y)
        self._t(f"{root}layer{i}.out", out)
        return out

    # -- suffix: [mirror] modeling_mvla.py:643-694 ------------------------------------------------------
    def embed_suffix(self, state: Tensor, noisy_actions: Tensor, timestep: Tensor):
        d = self.d
        state_emb = self.lin(state.float(), "model.state_proj.weight", "model.state_proj.bias")
        time_emb = create_sinusoidal_pos_embedding(timestep, d.proj_width, 4e-3, 4.0)
        time_emb = self.r(time_emb.to(torch.float32)) if self.mode == "fp32" else time_emb.to(torch.bfloat16).float()
        action_emb = self.lin(noisy_actions.float(), "model.action_in_proj.weight", "model.action_in_proj.bias")
        time_emb = time_emb[:, None, :].expand_as(action_emb)
        at = torch.cat([action_emb, time_emb], dim=2)
        at = self.lin(at, "model.action_time_mlp_in.weight", "model.action_time_mlp_in.bias")
        at = self.r(F.silu(at))
        at = self.lin(at, "model.action_time_mlp_out.weight", "model.action_time_mlp_out.bias")
        embs = torch.cat([state_emb[:, None, :], at], dim=1)
        B, S = embs.shape[:2]
        pad = torch.ones(B, S, dtype=torch.bool)
        att = torch.tensor([1] + [1] + [0] * (d.horizon - 1), dtype=torch.int64)[None].expand(B, S)
        return embs, pad, att

    # -- sample_actions: [mirror] modeling_mvla.py:753-804 + denoise_step :806-839 ---------------
    @torch.no_grad()
    def sample_actions(self, images: List[Tensor], img_masks: List[Tensor], lang_tokens: Tensor, lang_masks: Tensor,
                       state: Tensor, noise: Tensor, trace: Optional[dict] = None, skip_dead_work: bool = True):
        d = self.d
        self.trace = trace
        B = state.shape[0]
        if not self.images_pre_normalized:
            images = [img.float() * 2.0 - 1.0 for img in images]  # prepare_images, modeling_mvla.py:366-367
        state = pad_vector(state.float(), d.max_state_dim)
        prefix_embs, prefix_pad, prefix_att = self.embed_prefix(images, img_masks, lang_tokens, lang_masks.bool())
        prefix_2d = make_att_2d_masks(prefix_pad, prefix_att.long())
        prefix_pos = prefix_position_ids(prefix_pad.long())
        cache: dict = {}
        x = prefix_embs
        for i in range(d.n_gemma_layers):
            # the prefill output is discarded by sample_actions: the last layer only contributes its K/V
            x = self.gemma_layer(LM, i, x, prefix_pos, prefix_2d, cache, fill=True,
                                 skip_tail=skip_dead_work and i == d.n_gemma_layers - 1)
        self._t("cache.k0", cache[0]["key_states"])
        self._t("cache.v0", cache[0]["value_states"])

        dt = torch.tensor(-1.0 / d.num_steps, dtype=torch.float32)
        x_t = noise.clone().float()
        time = torch.tensor(1.0, dtype=torch.float32)
        step = 0
        while time >= -dt / 2:
            v_t = self.denoise_step(state, prefix_pad, cache, x_t, time.expand(B), step)
            # dt (0-dim fp32) * v_t: result takes v_t's dtype (bf16 in the bf16 regime), then fp32 accumulate
            x_t = x_t + self.r(dt * v_t)
            time = time + dt
            step += 1
        self.trace = None
        return x_t

    def denoise_step(self, state, prefix_pad, cache, x_t, timestep, step=0):
        d = self.d
        suffix_embs, suffix_pad, suffix_att = self.embed_suffix(state, x_t, timestep)
        B, S = suffix_pad.shape
        P = prefix_pad.shape[1]
        prefix_2d = prefix_pad[:, None, :].expand(B, S, P)
        suffix_2d = make_att_2d_masks(suffix_pad, suffix_att)
        full_2d = torch.cat([prefix_2d, suffix_2d], dim=2)
        pos = suffix_position_ids(prefix_pad.long(), suffix_pad.long())
        x = suffix_embs
        if step == 0:
            self._t("suffix.embs0", x)
        for i in range(d.n_gemma_layers):
            x = self.gemma_layer(EX, i, x, pos, full_2d, cache, fill=False)
        x = self.rms_norm(x, EX + "norm.weight")
        x = x[:, -d.horizon:]
        v = self.lin(x, "model.action_out_proj.weight", "model.action_out_proj.bias")
        if step == 0:
            self._t("v0", v)
        return v

    # -- the chunk-level entry north_star names ----------------------------------------------------------------
    def infer_action(self, images, img_masks, lang_tokens, lang_masks, state, noise, action_dim: int = 7):
        """actions[:, :, :action_dim] of sample_actions ([mirror] modeling_mvla.py:291-297); identity un-normalisation."""
        return self.sample_actions(images, img_masks, lang_tokens, lang_masks, state, noise)[:, :, :action_dim]
