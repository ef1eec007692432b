"""The PyTorch-eager GPU baseline (oracle/pi0_eager.py: stock torch ops under REAL torch.autocast, the op sequence the
reference executes) against the CPU oracle (oracle/pi0_oracle.py: rounding emulation). Agreement here means (a) bench.py's
`gpu_eager` number times the same arithmetic as the oracle the CUDA path is checked against, and (b) the oracle's
emulation of autocast(bf16) rounding points matches what torch.autocast really does."""
import pytest
import torch

from int_act_b200.weights import Dims, random_state_dict
from oracle.pi0_eager import Pi0Eager
from oracle.pi0_oracle import OracleDims
from pi0_helpers import oracle_for, synthetic_inputs


def _pair(dims, dev, mode, B=2, n_cams=1, horizon=4, n_valid=(12, 72)):
    sd = random_state_dict(dims, device="cpu", seed=2)
    inp = synthetic_inputs(B, n_cams, 72, horizon, dims.vocab, list(n_valid))
    od = OracleDims(n_cams=n_cams, lang_len=72, horizon=horizon, n_siglip_layers=dims.n_siglip_layers,
                    n_gemma_layers=dims.n_gemma_layers)
    eager = Pi0Eager(od, sd, dev, use_bf16=(mode == "bf16"))
    args = ([i.to(dev) for i in inp[0]], [m.to(dev) for m in inp[1]], inp[2].to(dev), inp[3].to(dev), inp[4].to(dev),
            inp[5].to(dev))
    out = eager.sample_actions(*args).float().cpu()
    ref = oracle_for(sd, dims, n_cams, 72, horizon, mode).sample_actions(*inp, skip_dead_work=False)
    return out, ref, inp


def test_eager_fp32_on_cpu_is_the_oracle():
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=512)
    out, ref, inp = _pair(dims, torch.device("cpu"), "fp32")
    assert (ref - inp[5]).abs().max().item() > 0.05
    assert (out - ref).abs().max().item() <= 2e-5


def test_eager_real_autocast_bf16_on_cpu_matches_the_oracle_emulation():
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=512)
    out, ref, _ = _pair(dims, torch.device("cpu"), "bf16")
    err = (out - ref).abs().max().item()
    print(f"\n[eager autocast(bf16) on CPU vs oracle emulation] max-abs {err:.3e}")
    assert err <= 1e-2, err


@pytest.mark.gpu
@pytest.mark.parametrize("mode,tol", [("bf16", 1e-2), ("fp32", 1e-4)])
def test_eager_on_gpu_matches_oracle(cuda, mode, tol):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dims = Dims(n_siglip_layers=2, n_gemma_layers=2, vocab=512)
    out, ref, _ = _pair(dims, cuda, mode)
    err = (out - ref).abs().max().item()
    print(f"\n[eager {mode} on GPU vs CPU oracle] max-abs {err:.3e}")
    assert err <= tol, err
