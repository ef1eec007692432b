"""CPU: pins the oracle (and the host-side helpers of the product package) against golden vectors produced by the
reference's OWN functions (tests/golden/make_golden.py, executed in the build container from /root/reference)."""
import os

import numpy as np
import torch

from oracle import pi0_oracle as orc
from int_act_b200 import weights as W
from int_act_b200.policy import resize_with_pad

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(G, name), allow_pickle=False)


def test_masks_and_position_ids_bit_exact():
    z = _load("masks.npz")
    pad = torch.from_numpy(z["pad"])
    B, P = pad.shape
    H = int(z["H"])
    att = torch.zeros(B, P, dtype=torch.int64)
    m = orc.make_att_2d_masks(pad, att)
    assert np.array_equal(np.packbits(m.numpy(), axis=-1), z["prefix_2d"])
    assert np.array_equal(orc.prefix_position_ids(pad.long()).numpy(), z["prefix_pos"])
    spad = torch.ones(B, 1 + H, dtype=torch.bool)
    satt = torch.tensor([1, 1] + [0] * (H - 1), dtype=torch.int64)[None].expand(B, 1 + H)
    s2d = orc.make_att_2d_masks(spad, satt)
    assert np.array_equal(s2d.numpy(), z["suffix_2d"])
    full = torch.cat([pad[:, None, :].expand(B, 1 + H, P), s2d], dim=2)
    assert np.array_equal(np.packbits(full.numpy(), axis=-1), z["full_2d"])
    assert np.array_equal(orc.suffix_position_ids(pad.long(), spad.long()).numpy(), z["suffix_pos"])
    blk = orc.make_att_2d_masks(torch.from_numpy(z["pad2"]), torch.from_numpy(z["att2"]))
    assert np.array_equal(blk.numpy(), z["blk_2d"])


def test_time_embedding_and_euler_times_bit_exact():
    z = _load("time_emb.npz")
    times = orc.euler_times(10)
    assert np.array_equal(np.array(times, dtype=np.float32), z["times"])
    assert W.euler_times(10) == times
    emb = orc.create_sinusoidal_pos_embedding(torch.from_numpy(z["times"]), 1024, 4e-3, 4.0)
    assert emb.dtype == torch.float64 and np.array_equal(emb.numpy(), z["emb"])
    assert np.array_equal(W.sinusoidal_time_embedding(times, 1024).numpy(), z["emb"])


def test_rope_bit_exact_and_table_matches():
    z = _load("rope.npz")
    pos = torch.from_numpy(z["pos"])
    assert np.array_equal(orc.apply_rope(torch.from_numpy(z["x"]), pos).numpy(), z["y"])
    yb = orc.apply_rope(torch.from_numpy(z["xb"]).to(torch.bfloat16), pos)
    assert np.array_equal(yb.float().numpy(), z["yb"])
    # the product's cos/sin table reproduces the rotation bit-exactly (the kernel multiplies by these entries)
    cos, sin = W.rope_table(512)
    x = torch.from_numpy(z["x"])
    c, s = cos[pos][:, :, None, :], sin[pos][:, :, None, :]
    x1, x2 = x[..., :128], x[..., 128:]
    y = torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1)
    assert np.array_equal(y.numpy(), z["y"])


def test_pad_vector_and_resize_with_pad():
    z = _load("prepare.npz")
    assert np.array_equal(orc.pad_vector(torch.from_numpy(z["v"]), 32).numpy(), z["v32"])
    img = torch.from_numpy(z["img"])
    for fn in (orc.resize_with_pad, resize_with_pad):
        assert np.array_equal(fn(img, 32, 32, pad_value=-1).numpy(), z["img_rp"])
        assert np.array_equal(fn(img, 64, 48, pad_value=-1).numpy(), z["img_same"])


def test_oracle_runs_and_is_deterministic_tiny():
    """One tiny forward of the oracle on the CPU (1 SigLIP layer, 1 Gemma layer, 2 Euler steps)."""
    from pi0_helpers import oracle_for, synthetic_inputs

    dims = W.Dims(n_siglip_layers=1, n_gemma_layers=1, vocab=256)
    sd = W.random_state_dict(dims, seed=5)
    inp = synthetic_inputs(1, vocab=256, n_valid_text=[6])
    outs = []
    for mode in ("fp32", "bf16"):
        o = oracle_for(sd, dims, 1, 72, 4, mode, num_steps=2)
        a = o.sample_actions(*inp)
        b = o.sample_actions(*inp)
        assert torch.equal(a, b) and a.shape == (1, 4, 32) and torch.isfinite(a).all()
        outs.append(a)
    # the two regimes agree to bf16 accuracy but are not the same computation
    assert (outs[0] - outs[1]).abs().max().item() < 0.25
    # skipping the dead tail of the last prefill layer does not change the result
    o = oracle_for(sd, dims, 1, 72, 4, "fp32", num_steps=2)
    assert torch.equal(o.sample_actions(*inp, skip_dead_work=False), outs[0])
