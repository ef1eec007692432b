"""GPU parity of the whole hot path through the C ABI (pi0_infer) against the CPU oracle on identical random-init
weights, inputs and noise. Tolerances are BASELINE.json's: max-abs action error <= 1e-2 in bf16, <= 1e-4 in the
fp32 check mode; pad mask / position ids bit-exact."""
import pytest
import torch

from int_act_b200 import binding as pb
from int_act_b200.engine import Pi0Engine
from int_act_b200.weights import Dims, PackedWeights, random_state_dict
from pi0_helpers import oracle_for, synthetic_inputs

pytestmark = pytest.mark.gpu

TOL = {"bf16": 1e-2, "fp32": 1e-4}


def _run(cuda, dims, mode, B, n_cams=1, lang_len=72, horizon=4, n_valid_text=None, graph=False, trace=None,
         pre_normalized=False, seed=0):
    act = torch.bfloat16 if mode == "bf16" else torch.float32
    sd = random_state_dict(dims, device=cuda, seed=seed, dtype=torch.bfloat16)
    packed = PackedWeights(sd, cuda, act)
    eng = Pi0Engine(packed, n_cams=n_cams, lang_len=lang_len, horizon=horizon, max_batch=B, dtype=mode,
                    use_cuda_graph=graph, images_pre_normalized=pre_normalized)
    inp = synthetic_inputs(B, n_cams, lang_len, horizon, dims.vocab, n_valid_text, pre_normalized=pre_normalized)
    images, img_masks, tokens, masks, state, noise = inp
    out = eng.sample_actions([i.to(cuda) for i in images], [m.to(cuda) for m in img_masks], tokens.to(cuda),
                             masks.to(cuda), state.to(cuda), noise.to(cuda))
    torch.cuda.synchronize()
    orc = oracle_for(sd, dims, n_cams, lang_len, horizon, mode, pre_normalized=pre_normalized)
    ref = orc.sample_actions(images, img_masks, tokens, masks, state, noise, trace=trace)
    return eng, out.cpu(), ref, orc, inp


@pytest.mark.parametrize("mode", ["bf16", "fp32"])
def test_small_depth_parity(cuda, mode):
    dims = Dims(n_siglip_layers=2, n_gemma_layers=2, vocab=2048)
    trace = {}
    eng, out, ref, orc, inp = _run(cuda, dims, mode, B=2, n_valid_text=[12, 72], trace=trace)
    # integer work: bit-exact
    pad = eng.debug_buffer("pad").view(2, 328).bool().cpu()
    exp_pad = torch.cat([torch.ones(2, 256, dtype=torch.bool), inp[3]], dim=1)
    assert torch.equal(pad, exp_pad)
    pos = eng.debug_buffer("prefix_pos").view(2, 328).cpu()
    assert torch.equal(pos.long(), torch.cumsum(exp_pad.long(), 1) - 1)
    spos = eng.debug_buffer("suffix_pos").view(2, 5).cpu()
    assert torch.equal(spos.long(), exp_pad.sum(1)[:, None] + torch.arange(5)[None])
    # prefix embedding and layer-0 K cache (valid rows) against the oracle trace
    embs = eng.debug_buffer("prefix.x")  # overwritten in place by the layers: only check finite
    assert torch.isfinite(embs.float()).all()
    k0 = eng.debug_buffer("kcache").float().cpu()[: 2 * 333 * 256].view(2, 333, 256)[:, :328]
    ref_k0 = trace["cache.k0"].view(2, 328, 256)
    valid = exp_pad[:, :, None].expand_as(ref_k0)
    kerr = ((k0 - ref_k0).abs() * valid).max().item()
    assert kerr <= (6e-2 if mode == "bf16" else 1e-4), kerr
    err = (out - ref).abs().max().item()
    print(f"\n[small depth {mode}] k0 err {kerr:.3e}  action err {err:.3e}  |ref|max {ref.abs().max().item():.3f} "
          f"|ref-noise|max {(ref - inp[5]).abs().max().item():.3f}")
    assert err <= TOL[mode], err
    assert (ref - inp[5]).abs().max().item() > 0.3  # the model moved the noise: the comparison is not vacuous


def test_cuda_graph_replay_matches_eager_launches(cuda):
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=3)
    packed = PackedWeights(sd, cuda, torch.bfloat16)
    inp = synthetic_inputs(3, vocab=dims.vocab, n_valid_text=[5, 30, 72])
    args = ([i.to(cuda) for i in inp[0]], [m.to(cuda) for m in inp[1]], inp[2].to(cuda), inp[3].to(cuda),
            inp[4].to(cuda), inp[5].to(cuda))
    e0 = Pi0Engine(packed, max_batch=3, use_cuda_graph=False)
    e1 = Pi0Engine(packed, max_batch=3, use_cuda_graph=True)
    a = e0.sample_actions(*args)
    b = e1.sample_actions(*args)
    c = e1.sample_actions(*args)  # second call = pure replay
    torch.cuda.synchronize()
    assert torch.equal(a, b) and torch.equal(b, c)
    assert e1.launches_per_infer[3] > 50


def test_rollout_batch_split_attention_parity(cuda):
    """Batch >= 4: the denoise attention runs as the two wide launches of attn_split.cu (ragged text lengths so the
    pad mask matters), tcgen05 GEMMs for the suffix rows."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    eng, out, ref, _, _ = _run(cuda, dims, "bf16", B=6, n_valid_text=[72, 9, 33, 2, 50, 17])
    assert (out - ref).abs().max().item() <= TOL["bf16"]
    # B > 12: one CTA pair per sample with K / V^T staged in shared memory (attn_batch.cu)
    eng, out, ref, _, _ = _run(cuda, dims, "bf16", B=14, n_valid_text=[72, 9, 33, 2, 50, 17, 64])
    assert (out - ref).abs().max().item() <= TOL["bf16"]


def test_rollout_batch_fused_prefix_attention_parity(cuda):
    """Batch 16 (>= 2 waves of 128-row tiles): the prefix pass runs the fused tcgen05 attention (attn_prefill.cu) and the
    CTA-pair GEMMs; ragged text lengths so the key mask matters, 3 Gemma layers so a KV-only last layer follows a full
    one."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=3, vocab=1024)
    eng, out, ref, _, _ = _run(cuda, dims, "bf16", B=16, n_valid_text=[72, 9, 33, 2, 50, 17, 64, 1])
    assert (out - ref).abs().max().item() <= TOL["bf16"]


def test_two_cameras_long_horizon(cuda):
    """LIBERO-shaped (BASELINE config 5): 2 cameras, horizon 16 -> P = 584, S = 17."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    eng, out, ref, _, _ = _run(cuda, dims, "bf16", B=2, n_cams=2, horizon=16, n_valid_text=[72, 9])
    assert out.shape == (2, 16, 32)
    assert (out - ref).abs().max().item() <= TOL["bf16"]


def test_rephrase_prefix_and_prenormalized(cuda):
    """BASELINE config 4 shape (64 text tokens, P = 320) with images handed over already in [-1, 1]."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    eng, out, ref, _, _ = _run(cuda, dims, "bf16", B=1, lang_len=64, pre_normalized=True)
    assert (out - ref).abs().max().item() <= TOL["bf16"]


def test_full_depth_bf16_parity(cuda):
    """The full bridge architecture (27 SigLIP + 18 Gemma + 18 expert layers, 10 Euler steps), batch 1, graph path."""
    dims = Dims(vocab=4096)
    eng, out, ref, _, _ = _run(cuda, dims, "bf16", B=1, n_valid_text=[12], graph=True)
    err = (out - ref).abs().max().item()
    print(f"\n[full depth bf16] max-abs action error vs oracle: {err:.3e}; |ref|max {ref.abs().max().item():.3f}; "
          f"launches/infer {eng.launches_per_infer[1]}")
    assert err <= TOL["bf16"], err


def test_missing_library_or_bad_args_fail_loudly(cuda):
    dims = Dims(n_siglip_layers=1, n_gemma_layers=1, vocab=256)
    sd = random_state_dict(dims, device=cuda, seed=1)
    packed = PackedWeights(sd, cuda, torch.bfloat16)
    eng = Pi0Engine(packed, max_batch=1, use_cuda_graph=False)
    inp = synthetic_inputs(2, vocab=256)
    with pytest.raises(pb.Pi0Error):
        eng.sample_actions([i.to(cuda) for i in inp[0]], [m.to(cuda) for m in inp[1]], inp[2].to(cuda),
                           inp[3].to(cuda), inp[4].to(cuda), inp[5].to(cuda))
    with pytest.raises(pb.Pi0Error):
        Pi0Engine(packed, max_batch=1, dtype="fp32")
