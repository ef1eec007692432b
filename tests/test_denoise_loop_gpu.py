"""The whole-loop persistent denoise kernel (int-act_b200/csrc/denoise_loop.cu, flag-in-data exchanges) against
(a) the round-1 per-step persistent kernel (grid barriers) - bit for bit in layers-only mode, since both follow the same
arithmetic order - and (b) the CPU oracle at BASELINE.json's bf16 tolerance. PI0_MEGA selects the path at ctx creation:
0 = per-op launches, 1 = round-1 kernel, 2 = new kernel per step (layers only), 3 / unset = whole Euler loop in one launch."""
import os

import pytest
import torch

from int_act_b200.engine import Pi0Engine
from int_act_b200.weights import Dims, PackedWeights, random_state_dict
from pi0_helpers import oracle_for, synthetic_inputs

pytestmark = pytest.mark.gpu


def _engine(packed, B, mega, graph=False, **kw):
    old = os.environ.get("PI0_MEGA")
    os.environ["PI0_MEGA"] = str(mega)
    try:
        return Pi0Engine(packed, max_batch=B, use_cuda_graph=graph, **kw)
    finally:
        if old is None:
            os.environ.pop("PI0_MEGA", None)
        else:
            os.environ["PI0_MEGA"] = old


def _args(cuda, inp):
    return ([i.to(cuda) for i in inp[0]], [m.to(cuda) for m in inp[1]], inp[2].to(cuda), inp[3].to(cuda), inp[4].to(cuda),
            inp[5].to(cuda))


@pytest.mark.parametrize("B,n_valid", [(1, [12]), (2, [72, 9]), (3, [5, 30, 72])])
def test_layers_only_mode_is_bitwise_the_round1_kernel(cuda, B, n_valid):
    dims = Dims(n_siglip_layers=1, n_gemma_layers=3, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=5)
    packed = PackedWeights(sd, cuda, torch.bfloat16)
    inp = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=n_valid)
    args = _args(cuda, inp)
    old = _engine(packed, B, 1).sample_actions(*args)
    new = _engine(packed, B, 2).sample_actions(*args)
    torch.cuda.synchronize()
    assert torch.isfinite(new).all()
    assert torch.equal(old, new), (old - new).abs().max().item()


@pytest.mark.parametrize("B,n_valid", [(1, [12]), (2, [72, 9]), (3, [5, 30, 72])])
def test_whole_loop_launch_matches_oracle_and_per_op_path(cuda, B, n_valid):
    dims = Dims(n_siglip_layers=1, n_gemma_layers=3, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=5)
    packed = PackedWeights(sd, cuda, torch.bfloat16)
    inp = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=n_valid)
    args = _args(cuda, inp)
    e_loop = _engine(packed, B, 3)
    loop = e_loop.sample_actions(*args)
    per_op = _engine(packed, B, 0).sample_actions(*args)
    torch.cuda.synchronize()
    ref = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*inp)
    err = (loop.cpu() - ref).abs().max().item()
    dif = (loop - per_op).abs().max().item()
    print(f"\n[denoise loop B={B}] vs oracle {err:.3e}; vs per-op launches {dif:.3e}")
    assert err <= 1e-2, err
    assert dif <= 1e-2, dif
    # the last layer's residual stream is still written to the workspace (debug buffer) by the loop kernel
    assert torch.isfinite(e_loop.debug_buffer("suffix.x").float()[: B * 5 * 1024]).all()


def test_graph_replays_advance_the_epoch_and_stay_deterministic(cuda):
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=7)
    packed = PackedWeights(sd, cuda, torch.bfloat16)
    inp = synthetic_inputs(1, vocab=dims.vocab, n_valid_text=[20])
    args = _args(cuda, inp)
    eng = _engine(packed, 1, 3, graph=True)
    outs = [eng.sample_actions(*args) for _ in range(6)]
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(outs[0], o)
    # one launch for the whole denoise loop: the batch-1 graph holds far fewer nodes than 10 x 7 per-step launches
    eager = _engine(packed, 1, 0, graph=True)
    eager.sample_actions(*args)
    assert eng.launches_per_infer[1] <= eager.launches_per_infer[1] - 60
    epoch = eng.debug_buffer("loop.epoch", dtype=torch.int32)[0].item()
    assert epoch > 0


@pytest.mark.parametrize("B,n_valid", [(4, [72, 9, 33, 2]), (6, [72, 9, 33, 2, 50, 17]), (9, [5, 30, 72, 64, 1, 12, 40, 72, 9])])
def test_persistent_kernel_with_chunked_operand_staging_for_batch_4_to_9(cuda, B, n_valid):
    """M = 5 B suffix rows in (16, 48]: the persistent per-step kernel runs two / three m16 tiles with the activation operand
    staged one 1024-k slice at a time (round 1 fell back to ~90 per-op launches per step from batch 4 on). Checked against
    the per-op launches and the oracle."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=3, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=5)
    packed = PackedWeights(sd, cuda, torch.bfloat16)
    inp = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=n_valid)
    args = _args(cuda, inp)
    os.environ["PI0_MEGA_MAX_ROWS"] = "48"     # the default stops at 32 rows (batch 6): 48 rows measured slower than per-op
    try:
        e_mega = _engine(packed, B, 1)
    finally:
        os.environ.pop("PI0_MEGA_MAX_ROWS", None)
    assert e_mega.path_report(B)["denoise"] == "persistent_layers_kernel_per_step"
    mega = e_mega.sample_actions(*args)
    per_op = _engine(packed, B, 0).sample_actions(*args)
    torch.cuda.synchronize()
    ref = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*inp)
    err = (mega.cpu() - ref).abs().max().item()
    dif = (mega - per_op).abs().max().item()
    print(f"\n[persistent kernel B={B}] vs oracle {err:.3e}; vs per-op launches {dif:.3e}")
    assert err <= 1e-2, err
    assert dif <= 1e-2, dif


def _fused_engine(packed, B, on, **kw):
    old = os.environ.get("PI0_FUSED")
    os.environ["PI0_FUSED"] = "1" if on else "0"
    try:
        return Pi0Engine(packed, max_batch=B, use_cuda_graph=kw.pop("graph", False), **kw)
    finally:
        if old is None:
            os.environ.pop("PI0_FUSED", None)
        else:
            os.environ["PI0_FUSED"] = old


@pytest.mark.parametrize("B,n_valid", [(7, [72, 9, 33, 2, 50, 17, 64]), (16, [72, 9, 33, 2, 50, 17, 64, 1]),
                                       (64, [72, 9, 33, 2, 50, 17, 64, 1, 12, 40, 72])])
def test_fused_tcgen05_denoise_step_kernel_matches_per_op_path_and_oracle(cuda, B, n_valid):
    """Rollout batches: the 18 expert layers of an Euler step as ONE persistent tcgen05 launch (split-K partial GEMMs, reduce
    phases carrying RoPE / residual + norm / GeGLU, in-kernel batch attention) against the per-op launches and the oracle."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=3, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=5)
    packed = PackedWeights(sd, cuda, torch.bfloat16)
    inp = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=n_valid)
    args = _args(cuda, inp)
    e_f = _fused_engine(packed, B, True)
    assert e_f.path_report(B)["denoise"] == "fused_tcgen05_layers_kernel_per_step"
    fused = e_f.sample_actions(*args)
    e_p = _fused_engine(packed, B, False)
    assert e_p.path_report(B)["denoise"] == "tcgen05_per_op"
    per_op = e_p.sample_actions(*args)
    torch.cuda.synchronize()
    assert torch.isfinite(fused).all()
    rows = list(range(B)) if B <= 16 else [0, 1, 31, 63]
    sub = ([i[rows] for i in inp[0]], [m[rows] for m in inp[1]], inp[2][rows], inp[3][rows], inp[4][rows], inp[5][rows])
    ref = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*sub)
    err = (fused.cpu()[rows] - ref).abs().max().item()
    dif = (fused - per_op).abs().max().item()
    print(f"\n[fused denoise step B={B}] vs oracle {err:.3e}; vs per-op launches {dif:.3e}; launches "
          f"{e_f.lib.pi0_launch_count(0)}")
    assert err <= 1e-2, err
    assert dif <= 1e-2, dif
