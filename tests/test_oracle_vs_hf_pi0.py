"""END-TO-END cross-check of the oracle (oracle/pi0_oracle.py) against the Hugging Face-native Pi0
(`transformers.models.pi0.modeling_pi0.PI0ForConditionalGeneration`, the port that loads `lerobot/pi0_base`) installed in
this image: the SAME random weights (lerobot key layout mapped onto the HF module tree), the same observation, the same
noise -> `sample_actions` in fp32 on the CPU.

Why it matters: the reference's own Pi0 lives in the un-vendored lerobot submodule, so the oracle's neural forward is
"parity unpinned" against the reference. The HF port is an INDEPENDENT implementation of the same model
by other authors; this test runs it through everything the block-wise test (test_oracle_vs_hf_blocks.py) could not reach:
embed_prefix with the image-token scatter, the prefix pad mask / position ids, the KV cache hand-over from the VLM to the
action expert, embed_suffix (state_proj, action_in_proj, sinusoidal time embedding, action_time_mlp), the block mask of the
joint attention (state | actions), the final norm + action_out_proj and the 10-step Euler loop.

Intentional differences between the two, all far below the tolerance: HF uses t_k = 1 + k * dt (the reference accumulates
t += dt in fp32: 0.79999995 instead of 0.8, modeling_mvla.py:787-803), computes the time embedding in fp32 (reference: fp64,
modeling_mvla.py:71-89), leaves the image features unscaled where the reference divides and re-multiplies by sqrt(2048)
(mllm.py:182-187 + modeling_mvla.py:600-606), and scales the language embedding inside `embed_tokens`.
It is not the reference - the status stays "unpinned" - but a disagreement here would have been a wrong oracle."""
import pytest
import torch

from int_act_b200.weights import Dims, random_state_dict
from oracle.pi0_oracle import EX, LM, PG, VT, OracleDims, Pi0Oracle
from pi0_helpers import synthetic_inputs

transformers = pytest.importorskip("transformers")


def _hf_model(n_sig, n_gem, vocab, horizon):
    try:
        from transformers import PI0Config, PI0ForConditionalGeneration
    except ImportError:
        pytest.skip("this transformers build has no native Pi0")
    common = {"hidden_activation": "gelu_pytorch_tanh", "rms_norm_eps": 1e-6, "num_attention_heads": 8,
              "num_key_value_heads": 1, "head_dim": 256, "vocab_size": vocab, "num_hidden_layers": n_gem}
    cfg = PI0Config(
        vlm_config={"model_type": "paligemma", "projection_dim": 2048, "image_token_id": vocab, "vocab_size": vocab,
                    "text_config": {"model_type": "gemma", "hidden_size": 2048, "intermediate_size": 16384, **common},
                    "vision_config": {"model_type": "siglip_vision_model", "hidden_size": 1152, "intermediate_size": 4304,
                                      "patch_size": 14, "image_size": 224, "num_hidden_layers": n_sig,
                                      "num_attention_heads": 16, "vision_use_head": False,
                                      "hidden_act": "gelu_pytorch_tanh", "layer_norm_eps": 1e-6}},
        dit_config={"model_type": "gemma", "hidden_size": 1024, "intermediate_size": 4096, **common},
        chunk_size=horizon, num_inference_steps=10)
    cfg._attn_implementation = "eager"
    return PI0ForConditionalGeneration(cfg).eval().float()


def _to_hf_keys(sd):
    out = {}
    for k, v in sd.items():
        if k.startswith(VT):
            out["model.vlm.vision_tower.vision_model." + k[len(VT):]] = v
        elif k.startswith(PG + "multi_modal_projector."):
            out["model.vlm." + k[len(PG):]] = v
        elif k.startswith(LM):
            out["model.vlm.language_model." + k[len(LM):]] = v
        elif k.startswith(EX):
            out["model.dit." + k[len(EX):]] = v
        elif k.startswith("model.action_out_proj."):
            out[k[len("model."):]] = v
        else:
            out["embed_action_time." + k[len("model."):]] = v
    return out


@pytest.mark.parametrize("n_cams,horizon,n_valid", [(1, 4, [12, 72]), (2, 6, [72, 5])])
def test_sample_actions_matches_the_hf_native_pi0_end_to_end(n_cams, horizon, n_valid):
    n_sig, n_gem, vocab, B, L = 1, 2, 512, 2, 72
    dims = Dims(n_siglip_layers=n_sig, n_gemma_layers=n_gem, vocab=vocab)
    sd = {k: v.float() for k, v in random_state_dict(dims, device="cpu", seed=5).items()}
    hf = _hf_model(n_sig, n_gem, vocab, horizon)
    missing, unexpected = hf.load_state_dict(_to_hf_keys(sd), strict=False)
    assert not unexpected and missing == ["model.dit.embed_tokens.weight"], (missing, unexpected)   # the expert has no embedding
    images, img_masks, tokens, masks, state, noise = synthetic_inputs(B, n_cams, L, horizon, vocab, n_valid)
    ids = torch.cat([torch.full((B, 256 * n_cams), vocab, dtype=torch.long), tokens], dim=1)       # image placeholders + text
    attn = torch.cat([torch.ones(B, 256 * n_cams, dtype=torch.long), masks.long()], dim=1)
    pixels = torch.stack([img * 2 - 1 for img in images], dim=1)                                     # prepare_images: [0,1] -> [-1,1]
    st = torch.zeros(B, 32)
    st[:, :7] = state
    with torch.no_grad():
        out = hf.sample_actions(state=st, input_ids=ids, pixel_values=pixels, noise=noise.clone(), attention_mask=attn,
                                pixel_attention_mask=torch.ones(B, n_cams, dtype=torch.bool))
    orc = Pi0Oracle(OracleDims(n_cams=n_cams, lang_len=L, horizon=horizon, n_siglip_layers=n_sig, n_gemma_layers=n_gem),
                    sd, mode="fp32")
    ref = orc.sample_actions(images, img_masks, tokens, masks, state, noise, skip_dead_work=False)
    err = (out - ref).abs().max().item()
    moved = (ref - noise).abs().max().item()
    print(f"\n[oracle vs HF-native Pi0, {n_cams} cam, H={horizon}] max-abs {err:.3e}; |ref - noise|max {moved:.3f}")
    assert moved > 0.3           # the model moved the noise: not a vacuous comparison
    assert err <= 1e-4, err      # north_star's fp32 tolerance
