"""Shared builders for the parity tests: seeded synthetic observations and the oracle wired
to the same state dict as the CUDA engine."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle.pi0_oracle import OracleDims, Pi0Oracle  # noqa: E402  (tests may use the oracle)


def synthetic_inputs(B, n_cams=1, lang_len=72, horizon=4, vocab=257152, n_valid_text=None, seed=1234,
                     pre_normalized=False):
    """images uniform in [0,1] (or [-1,1] when pre_normalized), state N(0,0.5) clipped to [-1,1] (7 dims),
    tokens: bos=2, ids uniform in [3, vocab), last valid = 108, pad 0; noise N(0,1) with manual_seed(42)."""
    g = torch.Generator().manual_seed(seed)
    images = []
    for _ in range(n_cams):
        img = torch.rand((B, 3, 224, 224), generator=g)
        images.append(img * 2 - 1 if pre_normalized else img)
    img_masks = [torch.ones(B, dtype=torch.bool) for _ in range(n_cams)]
    state = (torch.randn((B, 7), generator=g) * 0.5).clamp(-1, 1)
    tokens = torch.zeros((B, lang_len), dtype=torch.int64)
    masks = torch.zeros((B, lang_len), dtype=torch.bool)
    for b in range(B):
        n = n_valid_text[b % len(n_valid_text)] if isinstance(n_valid_text, (list, tuple)) else (n_valid_text or lang_len)
        n = max(2, min(lang_len, n))
        tokens[b, 0] = 2
        tokens[b, 1:n] = torch.randint(3, vocab, (n - 1,), generator=g)
        tokens[b, n - 1] = 108
        masks[b, :n] = True
    noise = torch.normal(0.0, 1.0, (B, horizon, 32), generator=torch.Generator().manual_seed(42))
    return images, img_masks, tokens, masks, state, noise


def oracle_for(sd, dims, n_cams, lang_len, horizon, mode, num_steps=10, pre_normalized=False):
    od = OracleDims(n_cams=n_cams, lang_len=lang_len, horizon=horizon, num_steps=num_steps,
                    n_siglip_layers=dims.n_siglip_layers, n_gemma_layers=dims.n_gemma_layers)
    return Pi0Oracle(od, {k: v.cpu() for k, v in sd.items()}, mode=mode, images_pre_normalized=pre_normalized)
