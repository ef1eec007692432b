"""CPU: host logic of the product package (config contract, weight packing, policy queue, C-ABI symbols)."""
import ctypes
import os
import re

import pytest
import torch

from int_act_b200 import binding as pb
from int_act_b200 import weights as W
from int_act_b200.config import PI0Config, bridge_config, libero_config, load_yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# the reference's shipped model config (config/models/pi0_finetune_bridge.json), restated key for key
BRIDGE_JSON = {
    "type": "pi0", "n_obs_steps": 1, "device": "cpu", "use_amp": True,
    "normalization_mapping": {"VISUAL": "IDENTITY", "STATE": "IDENTITY", "ACTION": "IDENTITY"},
    "input_features": {"observation.images.top": {"shape": [3, 224, 224], "type": "VISUAL"},
                       "observation.state": {"shape": [7], "type": "STATE"}},
    "output_features": {"action": {"type": "ACTION", "shape": [7]}},
    "chunk_size": 4, "n_action_steps": 4, "max_state_dim": 32, "max_action_dim": 32,
    "resize_imgs_with_padding": [224, 224], "empty_cameras": 0, "adapt_to_pi_aloha": False,
    "use_delta_joint_actions_aloha": False, "tokenizer_max_length": 72, "proj_width": 1024, "num_steps": 10,
    "use_cache": True, "attention_implementation": "eager", "freeze_vision_encoder": False,
    "train_expert_only": False, "train_state_proj": True, "paligemma_pretrained_path": None,
    "optimizer_lr": 5e-5, "optimizer_betas": [0.9, 0.999], "optimizer_eps": 1e-8, "optimizer_weight_decay": 0,
    "scheduler_warmup_steps": 200, "scheduler_decay_steps": 30000, "scheduler_decay_lr": 2.5e-06,
}


def test_config_accepts_reference_keys_and_validates():
    cfg = PI0Config.from_dict(BRIDGE_JSON)
    assert cfg.n_cams == 1 and cfg.tokenizer_max_length == 72 and cfg.n_action_steps == 4
    assert cfg.action_feature.shape == (7,) and cfg.robot_state_feature.shape == (7,)
    assert cfg.extra["optimizer_lr"] == 5e-5 and cfg.resize_imgs_with_padding == (224, 224)
    assert PI0Config.from_dict(cfg.to_dict()).to_dict() == cfg.to_dict()
    assert bridge_config().to_dict()["input_features"] == cfg.to_dict()["input_features"]
    assert libero_config().n_cams == 2
    with pytest.raises(ValueError):
        PI0Config.from_dict({**BRIDGE_JSON, "n_action_steps": 8})
    with pytest.raises(ValueError):
        PI0Config.from_dict({**BRIDGE_JSON, "n_obs_steps": 2})
    with pytest.raises(NotImplementedError):
        PI0Config.from_dict({**BRIDGE_JSON, "use_delta_joint_actions_aloha": True})


def test_yaml_include(tmp_path):
    import json

    (tmp_path / "models").mkdir()
    (tmp_path / "models" / "m.json").write_text(json.dumps(BRIDGE_JSON))
    (tmp_path / "exp.yaml").write_text("seed: 42\nmodel_cfg: !include models/m.json\neval_cfg:\n  port: 8000\n")
    y = load_yaml(str(tmp_path / "exp.yaml"))
    assert y["seed"] == 42 and y["model_cfg"]["tokenizer_max_length"] == 72 and y["eval_cfg"]["port"] == 8000
    assert PI0Config.from_dict(y["model_cfg"]).num_steps == 10


def test_state_dict_layout_and_param_counts():
    shapes = W.state_dict_shapes(W.Dims())
    n = lambda pref: sum(int(torch.tensor(s).prod()) for k, s in shapes.items() if k.startswith(pref))
    assert n(W.VT) + n(W.PROJ) == 414_441_856 - 0 or abs(n(W.VT) + n(W.PROJ) - 414.4e6) < 0.5e6  # SigLIP + projector
    gem = n(W.LM) - 257152 * 2048
    assert abs(gem - 1981.8e6) < 0.5e6          # 1 981.8 M Gemma params without the table
    assert abs(n(W.EX) - 311.4e6) < 0.5e6       # 311.4 M expert params
    assert "model.action_time_mlp_in.weight" in shapes and shapes["model.action_time_mlp_in.weight"] == (1024, 2048)


def test_packing_on_cpu_matches_definitions():
    dims = W.Dims(n_siglip_layers=1, n_gemma_layers=1, vocab=32)
    sd = W.random_state_dict(dims, seed=7)
    again = W.random_state_dict(dims, seed=7)
    assert all(torch.equal(sd[k], again[k]) for k in sd)
    pk = W.PackedWeights(dict(sd), "cpu", torch.bfloat16)
    assert pk.dims.n_siglip_layers == 1 and pk.dims.n_gemma_layers == 1 and pk.struct.vocab == 32
    # GeGLU interleave: rows [0,128) of tile t are gate rows t*128.., rows [128,256) the matching up rows
    p = f"{W.LM}layers.0.mlp."
    wgu = next(t for t in pk.keep if t.shape == (2 * 16384, 2048))
    assert torch.equal(wgu[256 * 3: 256 * 3 + 128], sd[p + "gate_proj.weight"][128 * 3: 128 * 4])
    assert torch.equal(wgu[256 * 3 + 128: 256 * 4], sd[p + "up_proj.weight"][128 * 3: 128 * 4])
    # time bias == b + W[:, 1024:] . bf16(time_emb(t_k))
    w_in = sd["model.action_time_mlp_in.weight"].double()
    temb = W.sinusoidal_time_embedding(W.euler_times(10)).to(torch.bfloat16).double()
    exp = sd["model.action_time_mlp_in.bias"].double()[None] + temb @ w_in[:, 1024:].t()
    assert torch.allclose(pk.time_bias.double(), exp, atol=1e-6)
    assert pk.time_bias.shape == (10, 1024)


def test_header_symbols_all_exported_and_bound():
    """Every function include/pi0b200.h declares is exported by libpi0b200.so and typed in the binding."""
    hdr = open(os.path.join(ROOT, "include", "pi0b200.h")).read()
    declared = set(re.findall(r"\b(pi0_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"pi0_model_cfg", "pi0_weights", "pi0_ctx", "pi0_gemm_desc"}
    assert declared == set(pb.SIGNATURES), declared ^ set(pb.SIGNATURES)
    assert os.path.exists(pb.LIB_PATH), "build the library first: python int-act_b200/build.py"
    handle = ctypes.CDLL(pb.LIB_PATH)
    for name in declared:
        assert hasattr(handle, name), name
    lib = pb.lib()
    assert lib.pi0_abi_version() == pb.ABI_VERSION
    # host-only entry points work without a GPU
    cfg = pb.ModelCfg(n_cams=1, lang_len=72, horizon=4, max_state_dim=32, max_action_dim=32, num_steps=10, dtype=0,
                      max_batch=2, num_sms=148, n_siglip_layers=27, n_gemma_layers=18)
    assert lib.pi0_workspace_bytes(ctypes.byref(cfg)) > 0
    cfg.lang_len = 70
    assert lib.pi0_workspace_bytes(ctypes.byref(cfg)) == 0
    assert b"lang_len" in lib.pi0_last_error()
    assert ctypes.sizeof(pb.Weights) == 8 * (3 + 27 * 12 + 4 + 2 + 18 * 6 * 2 + 1 + 10 + 2) + 8


def test_policy_fails_loudly_without_cuda():
    from int_act_b200.policy import PI0Policy, SyntheticTokenizer

    tok = SyntheticTokenizer(1000)
    ids, mask = tok(["put the carrot on the plate\n", "x\n"], 72)
    assert ids.shape == (2, 72) and ids[0, 0] == 2 and ids[0, int(mask[0].sum()) - 1] == 108 and mask[0].sum() == 8
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    pol = PI0Policy(bridge_config(), dims=W.Dims(n_siglip_layers=1, n_gemma_layers=1, vocab=1000), tokenizer=tok)
    batch = {"observation.images.top": torch.zeros(1, 3, 224, 224), "observation.state": torch.zeros(1, 7),
             "task": ["pick"]}
    with pytest.raises(pb.Pi0Error):
        pol.select_action(batch)


def test_bench_batch1_roofline_per_phase():
    """bench.py's batch-1 bound: per phase the slower of tensor time and HBM time; the 328-token prefill is tensor-bound,
    the 5-row denoise steps are weight streaming, and the FLOP split adds up to flops_per_chunk."""
    import importlib.util
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(root, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        spec.loader.exec_module(mod)
    finally:
        sys.argv = argv
    peaks = {"bf16_tflops": 1600.0, "hbm_gbs": 6500.0}
    total, ph = mod.roofline_ms_b1(1, 72, 4, peaks)
    assert ph["prefill"]["tensor_ms"] > ph["prefill"]["hbm_ms"]
    assert ph["denoise"]["hbm_ms"] > 10 * ph["denoise"]["tensor_ms"]
    assert abs(total - sum(max(v.values()) for v in ph.values())) < 1e-9
    flops = sum(v["tensor_ms"] for v in ph.values()) * 1e-3 * peaks["bf16_tflops"] * 1e12
    assert abs(flops - mod.flops_per_chunk(1, 72, 4)) <= 1e-6 * flops
    by = sum(v["hbm_ms"] for v in ph.values()) * 1e-3 * peaks["hbm_gbs"] * 1e9
    assert abs(by - mod.bytes_per_chunk_b1(1, 72, 4)) <= 1e-6 * by


# ---- round 2: checkpoint canonicalisation, normalisation buffers, empty cameras, pi-aloha maps ---------------------------
REF_BRIDGE_JSON = "/root/reference/config/models/pi0_finetune_bridge.json"
REF_HF_JSON = "/root/reference/config/models/hf_pi0.json"


@pytest.mark.skipif(not os.path.exists(REF_BRIDGE_JSON), reason="reference tree not mounted")
def test_reference_config_files_load_and_match_the_restated_keys():
    import json

    ref = json.load(open(REF_BRIDGE_JSON))
    assert ref == BRIDGE_JSON, {k for k in set(ref) | set(BRIDGE_JSON) if ref.get(k) != BRIDGE_JSON.get(k)}
    cfg = PI0Config.from_json(REF_BRIDGE_JSON)
    assert cfg.n_cams == 1 and cfg.tokenizer_max_length == 72
    hf = PI0Config.from_json(REF_HF_JSON)   # the stock lerobot pi0 config: 3 cameras, MEAN_STD state / action
    hf.validate_features()
    assert hf.n_cams == 3 + hf.empty_cameras and hf.normalization_mapping["STATE"] == "MEAN_STD"


def test_validate_features_adds_empty_camera_placeholders_once():
    cfg = bridge_config(empty_cameras=2)
    assert cfg.n_cams == 1
    cfg.validate_features()
    cfg.validate_features()   # idempotent
    assert cfg.n_cams == 3 and "observation.images.empty_camera_1" in cfg.image_features
    again = PI0Config.from_dict(cfg.to_dict())   # a config.json that already lists the placeholders
    again.validate_features()
    assert again.n_cams == 3


def test_checkpoint_with_tied_lm_head_and_mean_std_buffers(tmp_path):
    from int_act_b200.policy import PI0Policy

    dims = W.Dims(n_siglip_layers=1, n_gemma_layers=1, vocab=64)
    sd = W.random_state_dict(dims, seed=4)
    emb = sd.pop(W.LM + "embed_tokens.weight")          # saved with the PaliGemma head still tied: only lm_head survives
    sd[W.LM_HEAD] = emb
    sd["model.paligemma_with_expert.gemma_expert.lm_head.weight"] = torch.zeros(8, 1024, dtype=torch.bfloat16)
    mean, std = torch.arange(7.0), torch.arange(7.0) * 0.1 + 0.5
    for root in ("normalize_inputs.buffer_observation_state", "unnormalize_outputs.buffer_action",
                 "normalize_targets.buffer_action"):
        sd[root + ".mean"], sd[root + ".std"] = mean.clone(), std.clone()
    ok = W.canonical_state_dict(sd)
    assert torch.equal(ok[W.LM + "embed_tokens.weight"], emb) and not any(k.startswith("normalize") for k in ok)
    assert W.infer_dims(ok).vocab == 64
    with pytest.raises(KeyError, match="unexpected keys"):
        W.canonical_state_dict({**sd, "model.bogus.weight": torch.zeros(1)}, strict=True)
    broken = dict(sd)
    del broken["model.state_proj.bias"]
    with pytest.raises(KeyError, match="missing 1 tensors"):
        W.canonical_state_dict(broken)
    wrong = {**sd, "model.state_proj.bias": torch.zeros(3)}
    with pytest.raises(ValueError, match="unexpected shapes"):
        W.canonical_state_dict(wrong)
    # end to end through from_pretrained: the statistics come out of the checkpoint, as in lerobot
    W.save_safetensors(sd, str(tmp_path))
    cfg = bridge_config(normalization_mapping={"VISUAL": "IDENTITY", "STATE": "MEAN_STD", "ACTION": "MEAN_STD"})
    pol = PI0Policy.from_pretrained(str(tmp_path), config=cfg)
    x = torch.ones(2, 7)
    got = pol.normalize_inputs({"observation.state": x})["observation.state"]
    assert torch.allclose(got, (x - mean) / (std + 1e-8))
    back = pol.unnormalize_outputs({"action": torch.ones(2, 4, 7)})["action"]
    assert torch.allclose(back, torch.ones(2, 4, 7) * std + mean)
    with pytest.raises(ValueError, match="needs dataset_stats"):   # MEAN_STD without statistics anywhere: loud
        PI0Policy(cfg, None, state_dict=ok)


def test_pi_aloha_maps_match_the_reference_functions():
    import numpy as np

    from int_act_b200.policy import pi_aloha_decode_state, pi_aloha_encode_actions

    d = np.load(os.path.join(ROOT, "tests", "golden", "aloha.npz"))
    s_in = torch.from_numpy(d["state_in"])
    keep = s_in.clone()
    assert torch.equal(pi_aloha_decode_state(s_in), torch.from_numpy(d["state_out"]))
    assert torch.equal(s_in, keep)   # the caller's tensor is not modified (the reference mutates in place)
    assert torch.equal(pi_aloha_encode_actions(torch.from_numpy(d["act_in"])), torch.from_numpy(d["act_out"]))


def test_hf_tokenizer_known_answer_when_the_files_are_cached():
    """prepare_language ([mirror] modeling_mvla.py:387-407): "\\n" appended, bos first, right padding to
    tokenizer_max_length, truncation. The PaliGemma SentencePiece files are not on this box (no network), so
    this known-answer test SKIPS here; on a machine with the tokenizer cached it pins the ids the synthetic stand-in
    only imitates (bos = 2, "\\n" = 108, pad = 0)."""
    from int_act_b200.policy import HFTokenizer, PI0Policy

    try:
        tok = HFTokenizer()
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"google/paligemma-3b-pt-224 tokenizer files not cached: {type(e).__name__}")
    pol = PI0Policy(bridge_config(), tokenizer=tok)
    ids, mask = pol.prepare_language({"task": ["put the carrot on the plate", "stack the cube\n"]})
    assert ids.shape == (2, 72) and mask.shape == (2, 72) and mask.dtype == torch.bool
    n = mask.sum(1)
    for b in range(2):
        assert ids[b, 0].item() == 2                          # bos prepended by the tokenizer
        assert ids[b, n[b] - 1].item() == 108                 # the trailing "\n"
        assert mask[b, : n[b]].all() and not mask[b, n[b]:].any() and (ids[b, n[b]:] == 0).all()   # right padding
    assert n[0].item() > n[1].item() >= 4                     # bos + sentencepiece pieces + "\n"
    long = pol.prepare_language({"task": "word " * 200})
    assert long[0].shape == (1, 72) and long[1].all()         # truncation=True (the fork's addition, modeling_mvla.py:402)


def test_synthetic_tokenizer_keeps_the_call_shape():
    from int_act_b200.policy import PI0Policy, SyntheticTokenizer

    pol = PI0Policy(bridge_config(), tokenizer=SyntheticTokenizer())
    ids, mask = pol.prepare_language({"task": "put the carrot on the plate"})
    assert ids.shape == (1, 72) and ids[0, 0] == 2 and ids[0, mask.sum() - 1] == 108 and mask.sum() == 8
