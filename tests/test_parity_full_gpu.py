"""Round-2 parity cases the round-1 verdict asked for: FULL-DEPTH (27 SigLIP + 18 Gemma + 18 expert layers) parity of the
kernels the benchmarked configurations actually run - batch 64 (CTA-pair tcgen05 GEMMs, fused prefix attention, batch
decode attention, cluster split-K), fp32 check mode, LIBERO- and rephrase-shaped prefixes - plus the config surface of
SURVEY.md 8f-4 (MEAN_STD normalisation, a masked empty camera, three cameras, pi-aloha maps, odd tokenizer_max_length)
through `PI0Policy`. Tolerances are BASELINE.json's (bf16 <= 1e-2, fp32 <= 1e-4 max-abs on the action chunk)."""
import pytest
import torch

from int_act_b200.config import PI0Config, bridge_config
from int_act_b200.engine import Pi0Engine
from int_act_b200.policy import PI0Policy
from int_act_b200.weights import Dims, PackedWeights, random_state_dict
from pi0_helpers import oracle_for, synthetic_inputs

pytestmark = pytest.mark.gpu


def _engine_out(cuda, sd, mode, B, inp, n_cams=1, lang_len=72, horizon=4, graph=True):
    act = torch.bfloat16 if mode == "bf16" else torch.float32
    eng = Pi0Engine(PackedWeights(sd, cuda, act), n_cams=n_cams, lang_len=lang_len, horizon=horizon, max_batch=B,
                    dtype=mode, use_cuda_graph=graph)
    out = eng.sample_actions([i.to(cuda) for i in inp[0]], [m.to(cuda) for m in inp[1]], inp[2].to(cuda), inp[3].to(cuda),
                             inp[4].to(cuda), inp[5].to(cuda))
    torch.cuda.synchronize()
    return eng, out.cpu()


def _subset(inp, rows):
    return ([i[rows] for i in inp[0]], [m[rows] for m in inp[1]], inp[2][rows], inp[3][rows], inp[4][rows], inp[5][rows])


def test_full_depth_batch64_parity_on_a_subsample(cuda):
    """The headline configuration: batch 64, bf16, CUDA graph. Observations 0, 1, 31, 63 (ragged text lengths, so the key
    mask matters) are checked against the oracle run on those four alone - rows are independent, and a wrong row mapping
    in any batched kernel would show up here."""
    dims = Dims(vocab=4096)
    sd = random_state_dict(dims, device=cuda, seed=0)
    B = 64
    n_valid = [72, 9, 33, 2, 50, 17, 64, 1, 12, 40, 72]
    inp = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=n_valid)
    eng, out = _engine_out(cuda, sd, "bf16", B, inp)
    rows = [0, 1, 31, 63]
    ref = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*_subset(inp, rows))
    err = (out[rows] - ref).abs().max().item()
    moved = (ref - inp[5][rows]).abs().max().item()
    print(f"\n[full depth bf16, batch 64, rows {rows}] max-abs action error {err:.3e}; |ref - noise|max {moved:.3f}; "
          f"launches/infer {eng.launches_per_infer[B]}")
    assert moved > 0.3          # the model moved the noise: not a vacuous comparison
    assert err <= 1e-2, err
    assert torch.isfinite(out).all()


def test_full_depth_fp32_check_mode(cuda):
    """fp32 check mode at full depth (the shipped eval regime of the reference is --use_bf16 False): <= 1e-4."""
    dims = Dims(vocab=4096)
    sd = random_state_dict(dims, device=cuda, seed=0)
    inp = synthetic_inputs(1, vocab=dims.vocab, n_valid_text=[12])
    _, out = _engine_out(cuda, sd, "fp32", 1, inp, graph=False)
    ref = oracle_for(sd, dims, 1, 72, 4, "fp32").sample_actions(*inp)
    err = (out - ref).abs().max().item()
    print(f"\n[full depth fp32] max-abs action error vs oracle: {err:.3e}")
    assert (ref - inp[5]).abs().max().item() > 0.3
    assert err <= 1e-4, err


@pytest.mark.parametrize("name,n_cams,lang_len,horizon,B,n_valid", [
    ("libero", 2, 72, 16, 2, [72, 9]),        # BASELINE config 5 shape: P = 584, S = 17
    ("rephrase", 1, 64, 4, 2, [64, 20]),      # BASELINE config 4 shape: P = 320
])
def test_full_depth_other_baseline_shapes(cuda, name, n_cams, lang_len, horizon, B, n_valid):
    dims = Dims(vocab=4096)
    sd = random_state_dict(dims, device=cuda, seed=0)
    inp = synthetic_inputs(B, n_cams, lang_len, horizon, dims.vocab, n_valid)
    _, out = _engine_out(cuda, sd, "bf16", B, inp, n_cams, lang_len, horizon)
    ref = oracle_for(sd, dims, n_cams, lang_len, horizon, "bf16").sample_actions(*inp)
    err = (out - ref).abs().max().item()
    print(f"\n[full depth bf16 {name}] max-abs action error vs oracle: {err:.3e}")
    assert (ref - inp[5]).abs().max().item() > 0.3
    assert err <= 1e-2, err


@pytest.mark.parametrize("n_cams,horizon,B,n_valid", [(2, 16, 3, [72, 9, 40]), (3, 50, 2, [72, 30])])
def test_long_key_attention_kernel_in_the_engine(cuda, monkeypatch, n_cams, horizon, B, n_valid):
    """LIBERO (P = 584, 17 suffix tokens) and 3-camera / 50-step (P = 840, 51 suffix tokens) shapes with the long-key tcgen05
    attention forced on at a small batch (it is picked from 16 (sample, tile) items on): against the oracle and against the
    fallback chain it replaces."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=3, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=3)
    inp = synthetic_inputs(B, n_cams, 72, horizon, dims.vocab, n_valid)
    monkeypatch.setenv("PI0_ATTN_LONG_MIN_ITEMS", "1")
    monkeypatch.setenv("PI0_ATTN_LONG_PREFIX_MIN_ITEMS", "1")
    eng, out = _engine_out(cuda, sd, "bf16", B, inp, n_cams, 72, horizon, graph=False)
    rep = eng.path_report(B)
    assert rep["prefill_attention"] == "fused_tcgen05_long" and rep["denoise_attention"] == "fused_tcgen05_long", rep
    monkeypatch.setenv("PI0_ATTN_LONG", "0")
    eng2, out2 = _engine_out(cuda, sd, "bf16", B, inp, n_cams, 72, horizon, graph=False)
    assert eng2.path_report(B)["prefill_attention"] == "gemm_softmax_gemm"
    ref = oracle_for(sd, dims, n_cams, 72, horizon, "bf16").sample_actions(*inp)
    err, dif = (out - ref).abs().max().item(), (out - out2).abs().max().item()
    print(f"\n[long attention n_cams={n_cams} H={horizon}] vs oracle {err:.3e}; vs fallback chain {dif:.3e}")
    assert err <= 1e-2, err
    assert dif <= 1e-2, dif


def test_batch_above_96_takes_the_blocked_attention_kernel_in_the_denoise_step(cuda):
    """Bridge shape, 100 observations per GPU (C4 runs 256): the denoise attention switches from the batch kernel to the blocked
    tcgen05 kernel; rows 0, 50, 99 against the oracle."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=4)
    B = 100
    inp = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=[72, 9, 33, 2, 50, 17, 64])
    eng, out = _engine_out(cuda, sd, "bf16", B, inp, graph=False)
    assert eng.path_report(B)["denoise_attention"] == "fused_tcgen05_long"
    rows = [0, 50, 99]
    ref = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*_subset(inp, rows))
    err = (out[rows] - ref).abs().max().item()
    assert err <= 1e-2, err


def test_fused_rmsnorm_of_the_rollout_denoise_path_against_the_separate_norm_kernels(cuda, monkeypatch):
    """Batch 16 (per-op tcgen05 denoise path): RMSNorms fused into the q|k|v / gate-up GEMMs (default up to 64 suffix rows; the
    limit is lifted here) against PI0_NORM_FUSE=0 and the oracle; the fused path launches 2 x layers - 1 fewer kernels per
    Euler step."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=3, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=6)
    B = 16
    inp = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=[72, 9, 33, 2, 50, 17, 64, 1])
    monkeypatch.setenv("PI0_NORM_FUSE_MAX_ROWS", "512")
    eng, out = _engine_out(cuda, sd, "bf16", B, inp)
    monkeypatch.setenv("PI0_NORM_FUSE", "0")
    eng0, out0 = _engine_out(cuda, sd, "bf16", B, inp)
    ref = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*inp)
    err, err0, dif = (out - ref).abs().max().item(), (out0 - ref).abs().max().item(), (out - out0).abs().max().item()
    saved = eng0.launches_per_infer[B] - eng.launches_per_infer[B]
    print(f"\n[fused rmsnorm B={B}] vs oracle {err:.3e} (separate norms {err0:.3e}); fused vs separate {dif:.3e}; "
          f"{saved} launches fewer per infer")
    assert err <= 1e-2 and dif <= 1e-2
    assert saved == 10 * (2 * dims.n_gemma_layers - 1)


def test_state_row_and_vacuity_guards(cuda):
    """Zeroing the state moves the action by only ~1e-2 (the tolerance), so a broken state_proj path could hide behind the
    action check: the state row of the suffix embedding is compared directly."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=0)
    inp = synthetic_inputs(2, vocab=dims.vocab, n_valid_text=[12, 72])
    eng, out = _engine_out(cuda, sd, "bf16", 2, inp, graph=False)
    trace = {}
    ref = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*inp, trace=trace)
    assert (ref - inp[5]).abs().max().item() > 0.3
    assert (out - ref).abs().max().item() <= 1e-2
    suf = eng.debug_buffer("suffix.embs").float().cpu()[: 2 * 5 * 1024].view(2, 5, 1024)
    state_row = trace["suffix.embs0"][:, 0]
    assert state_row.abs().max().item() > 0.05
    assert (suf[:, 0] - state_row).abs().max().item() <= 2e-2 * max(1.0, state_row.abs().max().item())


# ---- config surface through PI0Policy -----------------------------------------------------------------
def _policy_vs_oracle(cuda, cfg, dims, sd, batch, oracle_args, stats=None, **okw):
    pol = PI0Policy(cfg, stats, state_dict=sd, device=cuda, max_batch=batch["observation.state"].shape[0])
    pol.to(torch.bfloat16)
    out = pol.infer_action({k: (v.to(cuda) if torch.is_tensor(v) else v) for k, v in batch.items()},
                           noise=oracle_args[5].to(cuda)).cpu()
    n_cams = len(oracle_args[0])
    ref = oracle_for(sd, dims, n_cams, cfg.tokenizer_max_length, cfg.n_action_steps, "bf16").sample_actions(*oracle_args)
    return out, ref[:, :, : out.shape[-1]]


def test_mean_std_normalisation_through_the_policy(cuda):
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=1)
    cfg = bridge_config(normalization_mapping={"VISUAL": "IDENTITY", "STATE": "MEAN_STD", "ACTION": "MEAN_STD"})
    g = torch.Generator().manual_seed(3)
    stats = {"observation.state": {"mean": torch.randn(7, generator=g) * 0.2, "std": torch.rand(7, generator=g) + 0.5},
             "action": {"mean": torch.randn(7, generator=g) * 0.3, "std": torch.rand(7, generator=g) + 0.5}}
    inp = synthetic_inputs(2, vocab=dims.vocab, n_valid_text=[12, 40])
    raw_state = inp[4] * 1.5 + 0.3
    batch = {"observation.images.top": inp[0][0], "observation.state": raw_state, "lang_tokens": inp[2], "lang_masks": inp[3]}
    norm_state = (raw_state - stats["observation.state"]["mean"]) / (stats["observation.state"]["std"] + 1e-8)
    out, ref = _policy_vs_oracle(cuda, cfg, dims, sd, batch, (inp[0], inp[1], inp[2], inp[3], norm_state, inp[5]), stats)
    ref = ref * stats["action"]["std"] + stats["action"]["mean"]          # lerobot Unnormalize, MEAN_STD
    assert (out - ref).abs().max().item() <= 1e-2 * float(stats["action"]["std"].max()) + 1e-6


def test_fp32_regime_through_the_policy(cuda):
    """`--use_bf16 False` (the reference's shipped eval flag): PI0Policy.to(torch.float32) -> fp32 weights and activations, GEMMs
    as split-bf16 products on the tensor cores (gemm_x3.cu) and the one-pass fp32 kernel for the denoise steps; <= 1e-4 against
    the fp32 oracle at batch 1 and at a batch that leaves the skinny kernel (M = 70 suffix rows)."""
    dims = Dims(n_siglip_layers=2, n_gemma_layers=3, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=8)
    cfg = bridge_config()
    for B, n_valid in [(1, [33]), (14, [72, 9, 33, 2, 50, 17, 64])]:
        inp = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=n_valid)
        batch = {"observation.images.top": inp[0][0], "observation.state": inp[4], "lang_tokens": inp[2], "lang_masks": inp[3]}
        pol = PI0Policy(cfg, None, state_dict=sd, device=cuda, max_batch=B)
        pol.to(torch.float32)
        out = pol.infer_action({k: (v.to(cuda) if torch.is_tensor(v) else v) for k, v in batch.items()},
                               noise=inp[5].to(cuda)).cpu()
        assert pol._engine.path_report(B)["gemm"] == "tcgen05_split_bf16x3_fp32"
        ref = oracle_for(sd, dims, 1, cfg.tokenizer_max_length, cfg.n_action_steps, "fp32").sample_actions(*inp)
        err = (out - ref[:, :, : out.shape[-1]]).abs().max().item()
        print(f"\n[fp32 regime through the policy, B={B}] max-abs action error vs the fp32 oracle: {err:.3e}")
        assert err <= 1e-4, err


def test_masked_empty_camera_and_three_cameras(cuda):
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=2)
    # (a) one real camera + one empty placeholder (empty_cameras = 1): the placeholder is a -1 image with mask 0
    cfg = bridge_config(empty_cameras=1)
    inp = synthetic_inputs(2, vocab=dims.vocab, n_valid_text=[12, 72])
    batch = {"observation.images.top": inp[0][0], "observation.state": inp[4], "lang_tokens": inp[2], "lang_masks": inp[3]}
    empty_img = torch.zeros_like(inp[0][0])      # the oracle maps [0,1] -> [-1,1]: 0 becomes the reference's -1 fill
    args = ([inp[0][0], empty_img], [inp[1][0], torch.zeros(2, dtype=torch.bool)], inp[2], inp[3], inp[4], inp[5])
    out, ref = _policy_vs_oracle(cuda, cfg, dims, sd, batch, args)
    assert (out - ref).abs().max().item() <= 1e-2
    one_cam = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*inp)[:, :, :7]
    assert (ref - one_cam).abs().max().item() > 1e-3       # the masked camera still shifts the position ids: not a no-op
    # (b) three cameras, the stock hf_pi0.json shape (P = 3 * 256 + 48 = 816)
    cfg3 = PI0Config.from_dict({
        **bridge_config().to_dict(), "tokenizer_max_length": 48,
        "input_features": {"observation.images.base_0_rgb": {"shape": [3, 224, 224], "type": "VISUAL"},
                           "observation.images.left_wrist_0_rgb": {"shape": [3, 224, 224], "type": "VISUAL"},
                           "observation.images.right_wrist_0_rgb": {"shape": [3, 224, 224], "type": "VISUAL"},
                           "observation.state": {"shape": [7], "type": "STATE"}}})
    inp3 = synthetic_inputs(2, 3, 48, 4, dims.vocab, [48, 7])
    batch3 = {"observation.images.base_0_rgb": inp3[0][0], "observation.images.left_wrist_0_rgb": inp3[0][1],
              "observation.images.right_wrist_0_rgb": inp3[0][2], "observation.state": inp3[4], "lang_tokens": inp3[2],
              "lang_masks": inp3[3]}
    out3, ref3 = _policy_vs_oracle(cuda, cfg3, dims, sd, batch3, inp3)
    assert (out3 - ref3).abs().max().item() <= 1e-2


def test_odd_tokenizer_max_length_is_padded_internally(cuda):
    """tokenizer_max_length = 45 (not a multiple of 8): the engine pads the text block with masked positions."""
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    sd = random_state_dict(dims, device=cuda, seed=0)
    inp = synthetic_inputs(2, 1, 45, 4, dims.vocab, [45, 6])
    _, out = _engine_out(cuda, sd, "bf16", 2, inp, 1, 45, 4, graph=False)
    ref = oracle_for(sd, dims, 1, 45, 4, "bf16").sample_actions(*inp)
    assert (out - ref).abs().max().item() <= 1e-2


def test_path_report_names_the_variant_every_phase_takes(cuda):
    dims = Dims(n_siglip_layers=1, n_gemma_layers=1, vocab=256)
    packed = PackedWeights(random_state_dict(dims, device=cuda, seed=0), cuda, torch.bfloat16)
    eng = Pi0Engine(packed, max_batch=64)
    b1, b64 = eng.path_report(1), eng.path_report(64)
    assert b1["denoise"] == "persistent_layers_kernel_per_step" and b1["prefill_attention"] == "gemm_softmax_gemm"
    assert b64["denoise"] == "tcgen05_per_op" and b64["denoise_attention"] == "batch_cta_pair_per_sample"
    assert b64["prefill_attention"] == "fused_tcgen05_long" and b64["siglip_attention"] == "fused_tcgen05"
    lib = Pi0Engine(packed, n_cams=2, horizon=16, max_batch=64).path_report(64)      # LIBERO shape: P = 584 > 336
    assert lib["prefix_tokens"] == 584 and lib["prefill_attention"] == "fused_tcgen05_long"
    assert lib["denoise_attention"] == "fused_tcgen05_long"
    f32 = Pi0Engine(PackedWeights(random_state_dict(dims, device=cuda, seed=0), cuda, torch.float32), dtype="fp32")
    assert f32.path_report(1)["gemm"] == "tcgen05_split_bf16x3_fp32"


@pytest.mark.parametrize("B,n_cams,horizon,mega", [(1, 1, 4, 1), (1, 1, 4, 3), (3, 1, 4, 1), (6, 1, 4, 1), (16, 1, 4, 1),
                                                     (2, 2, 16, 1), (64, 1, 4, 1)])
def test_no_kernel_writes_outside_its_workspace_buffer(cuda, monkeypatch, B, n_cams, horizon, mega):
    """Bounds check of our own (compute-sanitizer is closed on this GPU pool, profiles/r02_sanitizer_summary.txt): with
    PI0_GUARD=1 every workspace buffer is followed by a 0xA5 guard zone; after a full pi0_infer through each kernel family
    (persistent denoise kernel, whole-loop kernel, split / batch decode attention, fused prefix attention, CTA-pair and
    split-K GEMMs, LIBERO shapes) all zones must be intact - and a deliberate overrun must be caught."""
    monkeypatch.setenv("PI0_GUARD", "1")
    monkeypatch.setenv("PI0_MEGA", str(mega))
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=512)
    sd = random_state_dict(dims, device=cuda, seed=0)
    inp = synthetic_inputs(B, n_cams, 72, horizon, dims.vocab, [72, 9, 33])
    eng, out = _engine_out(cuda, sd, "bf16", B, inp, n_cams, 72, horizon, graph=False)
    assert torch.isfinite(out).all()
    eng.check_guards()
    if B == 1 and mega == 1:   # the check has teeth: write one byte past the end of a buffer
        import ctypes as C
        n = C.c_int64()
        p = eng.lib.pi0_debug_buffer(eng.ctx, b"suffix.q", C.byref(n))
        off = p - eng.workspace.data_ptr()
        eng.workspace[off + n.value * 2] = 7
        with pytest.raises(Exception, match="suffix.q"):
            eng.check_guards()


def test_unshrunk_action_out_proj_error_is_reported_next_to_the_tolerance(cuda):
    """`random_state_dict` shrinks action_out_proj by 0.25 so that |v_t| < 1 like a trained policy on bound-normalised actions
    (one bf16 ulp of v_t at |v_t| ~ 3 is 1.6e-2 > the 1e-2 tolerance for ANY two bf16 implementations). This case undoes the
    shrink, reports the error, and bounds it RELATIVE to the velocity scale: the kernels are as accurate as before, only the
    quantity is 4x larger."""
    dims = Dims(vocab=4096)
    sd = random_state_dict(dims, device=cuda, seed=0)
    sd["model.action_out_proj.weight"] = (sd["model.action_out_proj.weight"].float() * 4.0).to(torch.bfloat16)
    inp = synthetic_inputs(1, vocab=dims.vocab, n_valid_text=[12])
    _, out = _engine_out(cuda, sd, "bf16", 1, inp)
    ref = oracle_for(sd, dims, 1, 72, 4, "bf16").sample_actions(*inp)
    err = (out - ref).abs().max().item()
    moved = (ref - inp[5]).abs().max().item()
    print(f"\n[full depth bf16, un-shrunk action_out_proj] max-abs action error {err:.3e} for |actions - noise|max {moved:.3f} "
          f"(relative {err / moved:.2e}; shrunk weights: 4.6e-3 for 0.64)")
    assert err / moved <= 1.5e-2
