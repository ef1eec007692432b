"""PyTorch-eager GPU baseline of the Pi0 `sample_actions` path -- TEST / BENCH INFRASTRUCTURE, NOT PRODUCT CODE.

What INT-ACT actually executes on a GPU is PyTorch eager under `torch.autocast(bfloat16)` + `inference_mode`
(/root/reference/src/experiments/policies/policy_wrapper.py:121-139) with `"attention_implementation": "eager"`
(/root/reference/config/models/pi0_finetune_bridge.json:50). The lerobot module that holds that code is not in the tree
(SURVEY.md F1), so this file restates the SAME op sequence with stock torch ops on the device - one `F.linear` per
projection (cuBLAS), `F.conv2d` patch embedding (cuDNN), `F.layer_norm`, materialised KV-head expansion, fp32 upcast of
Q / K that autocast re-casts, `torch.where` + `softmax`, `torch.cat` of the cached prefix K / V every layer of every step,
the Python `while time >= -dt / 2` loop with its device-tensor comparison - under REAL `torch.autocast`, not the
rounding emulation of pi0_oracle.py. Used by `bench.py --impl eager` / the `gpu_eager` record (the "reference-equivalent
GPU" number of SURVEY.md 8d) and by tests/test_eager_gpu.py, which checks it against the CPU oracle: that comparison also
validates the oracle's autocast emulation against the real thing. Only tests/ and bench.py import this module.

Call sites followed: [mirror] /root/reference/src/model/mvla/modeling_mvla.py:582-624,636-694,753-839;
mllm.py:30-53,182-191,206-245,265-317; expert.py:128-179,387-409; HF 4.48.3 SigLIP / Gemma block semantics per SURVEY.md 9.2.
"""
from __future__ import annotations

import math
from typing import Dict, List

import torch
import torch.nn.functional as F

from .pi0_oracle import BIG_NEG, EX, LM, PG, VT, OracleDims, create_sinusoidal_pos_embedding, make_att_2d_masks

Tensor = torch.Tensor


def _rope(x: Tensor, positions: Tensor, max_wavelength: float = 10_000.0) -> Tensor:
    """[mirror] mllm.py:30-53 on the device: fp32 trig, half-split rotation, cast back."""
    d_half = x.shape[-1] // 2
    dtype = x.dtype
    x = x.to(torch.float32)
    freq = (2.0 / x.shape[-1]) * torch.arange(d_half, dtype=torch.float32, device=x.device)
    radians = positions[..., None].to(torch.float32) / (max_wavelength ** freq)[None, None, :]
    radians = radians[..., None, :]
    sin, cos = torch.sin(radians), torch.cos(radians)
    x1, x2 = x.split(d_half, dim=-1)
    out = torch.empty_like(x)
    out[..., :d_half] = x1 * cos - x2 * sin
    out[..., d_half:] = x2 * cos + x1 * sin
    return out.to(dtype)


class Pi0Eager:
    """Eager restatement with parameters resident on `device` in `param_dtype` (the wrapper's `model.to(dtype)`)."""

    def __init__(self, dims: OracleDims, state_dict: Dict[str, Tensor], device, use_bf16: bool = True,
                 images_pre_normalized: bool = False):
        self.d = dims
        self.device = torch.device(device)
        self.use_bf16 = use_bf16
        self.images_pre_normalized = images_pre_normalized
        pdt = torch.bfloat16 if use_bf16 else torch.float32
        self.w = {k: (v.detach().to(self.device).to(pdt) if v.is_floating_point() else v.detach().to(self.device))
                  for k, v in state_dict.items()}
        self.act = pdt

    def _autocast(self):
        # policy_wrapper.py:130: torch.autocast(device_type="cuda", dtype=bfloat16 if use_bf16 else float32)
        return torch.autocast(device_type=self.device.type, dtype=torch.bfloat16 if self.use_bf16 else torch.float32,
                              enabled=self.use_bf16 or self.device.type == "cuda")

    def lin(self, x, name, bias=True):
        return F.linear(x, self.w[name + ".weight"], self.w[name + ".bias"] if bias else None)

    def ln(self, x, name):
        return F.layer_norm(x, (x.shape[-1],), self.w[name + ".weight"], self.w[name + ".bias"], self.d.eps)

    def rms(self, x, name):
        xf = x.float()
        out = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.d.eps)
        return (out * (1.0 + self.w[name].float())).type_as(x)

    # ---- SigLIP tower + projector (HF 4.48.3 get_image_features) ------------------------------------------------------
    def embed_image(self, pixels: Tensor) -> Tensor:
        d = self.d
        B = pixels.shape[0]
        x = F.conv2d(pixels, self.w[VT + "embeddings.patch_embedding.weight"], self.w[VT + "embeddings.patch_embedding.bias"],
                     stride=d.patch)
        x = x.flatten(2).transpose(1, 2)
        x = x + self.w[VT + "embeddings.position_embedding.weight"][None]
        hd = d.vis_hidden // d.vis_heads
        for i in range(d.n_siglip_layers):
            p = f"{VT}encoder.layers.{i}."
            res = x
            h = self.ln(x, p + "layer_norm1")
            q = self.lin(h, p + "self_attn.q_proj").view(B, -1, d.vis_heads, hd).transpose(1, 2)
            k = self.lin(h, p + "self_attn.k_proj").view(B, -1, d.vis_heads, hd).transpose(1, 2)
            v = self.lin(h, p + "self_attn.v_proj").view(B, -1, d.vis_heads, hd).transpose(1, 2)
            att = torch.matmul(q, k.transpose(2, 3)) * hd ** -0.5
            att = F.softmax(att, dim=-1, dtype=torch.float32).to(q.dtype)
            o = torch.matmul(att, v).transpose(1, 2).reshape(B, -1, d.vis_hidden)
            x = res + self.lin(o, p + "self_attn.out_proj")
            res = x
            h = self.ln(x, p + "layer_norm2")
            h = self.lin(F.gelu(self.lin(h, p + "mlp.fc1"), approximate="tanh"), p + "mlp.fc2")
            x = res + h
        x = self.ln(x, VT + "post_layernorm")
        x = self.lin(x, PG + "multi_modal_projector.linear")
        return x / (d.hidden ** 0.5)

    def embed_prefix(self, images, img_masks, lang_tokens, lang_masks):
        d = self.d
        embs, pads = [], []
        for img, m in zip(images, img_masks):
            e = self.embed_image(img).to(torch.bfloat16 if self.use_bf16 else torch.float32)
            e = e * torch.tensor(e.shape[-1] ** 0.5, dtype=e.dtype, device=e.device)
            embs.append(e)
            pads.append(m[:, None].expand(e.shape[0], e.shape[1]))
        lang = F.embedding(lang_tokens, self.w[LM + "embed_tokens.weight"])
        embs.append(lang * math.sqrt(lang.shape[-1]))
        pads.append(lang_masks)
        embs = torch.cat(embs, dim=1)
        pad = torch.cat(pads, dim=1).bool()
        return embs, pad, torch.zeros_like(pad)

    # ---- attention exactly as the mirror's eager_attention_forward lays it out ---------------------------------------
    def attention(self, mask, q, k, v):
        d = self.d
        B, Tk = k.shape[0], k.shape[1]
        k = k[:, :, :, None, :].expand(B, Tk, 1, d.heads, d.head_dim).reshape(B, Tk, d.heads, d.head_dim)
        v = v[:, :, :, None, :].expand(B, Tk, 1, d.heads, d.head_dim).reshape(B, Tk, d.heads, d.head_dim)
        q = q.to(torch.float32).transpose(1, 2)
        k = k.to(torch.float32).transpose(1, 2)
        att = torch.matmul(q, k.transpose(2, 3))
        att *= d.head_dim ** -0.5
        att = torch.where(mask[:, None, :, :], att, BIG_NEG)
        probs = F.softmax(att, dim=-1).to(v.dtype)
        out = torch.matmul(probs, v.permute(0, 2, 1, 3)).permute(0, 2, 1, 3)
        return out.reshape(B, -1, d.heads * d.head_dim)

    def gemma_layer(self, root, i, x, pos, mask, cache, fill):
        d = self.d
        p = f"{root}layers.{i}."
        B, T = x.shape[:2]
        h = self.rms(x, p + "input_layernorm.weight").to(self.act)
        q = self.lin(h, p + "self_attn.q_proj", False).view(B, T, d.heads, d.head_dim)
        k = self.lin(h, p + "self_attn.k_proj", False).view(B, T, 1, d.head_dim)
        v = self.lin(h, p + "self_attn.v_proj", False).view(B, T, 1, d.head_dim)
        q, k = _rope(q, pos), _rope(k, pos)
        if fill:
            cache[i] = {"key_states": k, "value_states": v}
        else:
            k = torch.cat([cache[i]["key_states"], k], dim=1)
            v = torch.cat([cache[i]["value_states"], v], dim=1)
        att = self.attention(mask, q, k, v).to(self.act)
        y = self.lin(att, p + "self_attn.o_proj", False) + x
        h = self.rms(y, p + "post_attention_layernorm.weight")
        m = self.lin(F.gelu(self.lin(h, p + "mlp.gate_proj", False), approximate="tanh") * self.lin(h, p + "mlp.up_proj", False),
                     p + "mlp.down_proj", False)
        return m + y

    def embed_suffix(self, state, x_t, timestep):
        d = self.d
        state_emb = self.lin(state, "model.state_proj")
        time_emb = create_sinusoidal_pos_embedding(timestep, d.proj_width, 4e-3, 4.0).type(self.act)
        action_emb = self.lin(x_t, "model.action_in_proj")
        time_emb = time_emb[:, None, :].expand_as(action_emb)
        at = self.lin(torch.cat([action_emb, time_emb], dim=2), "model.action_time_mlp_in")
        at = self.lin(F.silu(at), "model.action_time_mlp_out")
        embs = torch.cat([state_emb[:, None, :].to(at.dtype), at], dim=1)
        B, S = embs.shape[:2]
        pad = torch.ones(B, S, dtype=torch.bool, device=embs.device)
        att = torch.tensor([1] + [1] + [0] * (d.horizon - 1), dtype=torch.int64, device=embs.device)[None].expand(B, S)
        return embs, pad, att

    def denoise_step(self, state, prefix_pad, cache, x_t, timestep):
        d = self.d
        embs, pad, att = self.embed_suffix(state, x_t, timestep)
        B, S = pad.shape
        P = prefix_pad.shape[1]
        full = torch.cat([prefix_pad[:, None, :].expand(B, S, P), make_att_2d_masks(pad, att)], dim=2)
        pos = torch.sum(prefix_pad, dim=-1)[:, None] + torch.cumsum(pad, dim=1) - 1
        x = embs
        for i in range(d.n_gemma_layers):
            x = self.gemma_layer(EX, i, x, pos, full, cache, fill=False)
        x = self.rms(x, EX + "norm.weight")[:, -d.horizon:].to(torch.float32)
        return self.lin(x, "model.action_out_proj")

    @torch.inference_mode()
    def sample_actions(self, images: List[Tensor], img_masks: List[Tensor], lang_tokens: Tensor, lang_masks: Tensor,
                       state: Tensor, noise: Tensor) -> Tensor:
        """All inputs already on the device (policy_wrapper.py:125-128). Returns x_t fp32 [B, H, 32]."""
        d = self.d
        with self._autocast():
            if not self.images_pre_normalized:
                images = [img * 2.0 - 1.0 for img in images]
            st = state.float()
            if st.shape[-1] < d.max_state_dim:
                st = F.pad(st, (0, d.max_state_dim - st.shape[-1]))
            embs, pad, att = self.embed_prefix(images, img_masks, lang_tokens, lang_masks.bool())
            mask = make_att_2d_masks(pad, att.long())
            pos = torch.cumsum(pad.long(), dim=1) - 1
            cache: dict = {}
            x = embs
            for i in range(d.n_gemma_layers):
                x = self.gemma_layer(LM, i, x, pos, mask, cache, fill=True)
            B = st.shape[0]
            dt = torch.tensor(-1.0 / d.num_steps, dtype=torch.float32, device=self.device)
            x_t = noise.clone().float()
            time = torch.tensor(1.0, dtype=torch.float32, device=self.device)
            while time >= -dt / 2:  # one device -> host sync per step, as in the reference loop
                v_t = self.denoise_step(st, pad, cache, x_t, time.expand(B))
                x_t = x_t + dt * v_t
                time = time + dt
        return x_t
