"""Independent cross-check of the neural oracle (oracle/pi0_oracle.py) against the Hugging Face modules the reference's
path is built from (SigLIP vision tower, Gemma decoder layer), instantiated from the transformers package installed
here with the SAME random weights. This is NOT the reference (its Pi0 lives in the un-vendored lerobot submodule on
transformers 4.48.3; the installed version is 5.x) - it pins the block semantics SURVEY.md 9.2 lists as unchanged:
SigLIP embeddings / pre-LN MHA with bias / gelu-tanh MLP / post-LN, GemmaRMSNorm (1 + w), half-split RoPE with
theta = 10000, 1/sqrt(head_dim) scaling, MQA with one KV head, GeGLU MLP. CPU, fp32."""
import pytest
import torch

from int_act_b200.weights import Dims, random_state_dict
from oracle.pi0_oracle import LM, VT, OracleDims, Pi0Oracle

transformers = pytest.importorskip("transformers")


def _sub(sd, prefix):
    return {k[len(prefix):]: v.float() for k, v in sd.items() if k.startswith(prefix)}


def test_siglip_tower_matches_hf_module():
    from transformers import SiglipVisionConfig, SiglipVisionModel

    dims = Dims(n_siglip_layers=2, n_gemma_layers=1, vocab=64)
    sd = {k: v.float() for k, v in random_state_dict(dims, device="cpu", seed=3).items()}
    cfg = SiglipVisionConfig(hidden_size=1152, intermediate_size=4304, num_hidden_layers=2, num_attention_heads=16,
                             image_size=224, patch_size=14, hidden_act="gelu_pytorch_tanh", layer_norm_eps=1e-6,
                             vision_use_head=False)
    cfg._attn_implementation = "eager"
    hf = SiglipVisionModel(cfg).eval()
    missing, unexpected = hf.vision_model.load_state_dict(_sub(sd, VT), strict=False)
    assert not unexpected and all("head" in m for m in missing), (missing, unexpected)
    pixels = torch.rand(2, 3, 224, 224, generator=torch.Generator().manual_seed(0)) * 2 - 1
    with torch.no_grad():
        out = hf(pixel_values=pixels, output_hidden_states=True)
    orc = Pi0Oracle(OracleDims(n_siglip_layers=2, n_gemma_layers=1), sd, mode="fp32")
    orc.trace = {}
    orc.embed_image(pixels)
    ref_last_layer = out.hidden_states[-1]                       # before post_layernorm
    got = orc.trace["siglip.layer1"]
    assert (got - ref_last_layer).abs().max().item() <= 2e-4 * max(1.0, ref_last_layer.abs().max().item())
    got_ln = orc.layer_norm(got, VT + "post_layernorm")
    assert (got_ln - out.last_hidden_state).abs().max().item() <= 2e-4 * max(1.0, out.last_hidden_state.abs().max().item())
    assert (orc.trace["siglip.embed"] - out.hidden_states[0]).abs().max().item() <= 1e-5


def test_gemma_layer_matches_hf_module():
    from transformers import GemmaConfig
    from transformers.models.gemma.modeling_gemma import GemmaDecoderLayer, GemmaRotaryEmbedding

    dims = Dims(n_siglip_layers=1, n_gemma_layers=1, vocab=64)
    sd = {k: v.float() for k, v in random_state_dict(dims, device="cpu", seed=4).items()}
    cfg = GemmaConfig(hidden_size=2048, intermediate_size=16384, num_hidden_layers=1, num_attention_heads=8,
                      num_key_value_heads=1, head_dim=256, hidden_activation="gelu_pytorch_tanh", rms_norm_eps=1e-6,
                      rope_theta=10000.0, attention_bias=False, vocab_size=64, max_position_embeddings=512)
    cfg._attn_implementation = "eager"
    layer = GemmaDecoderLayer(cfg, layer_idx=0).eval()
    layer.load_state_dict(_sub(sd, LM + "layers.0."), strict=True)
    rope = GemmaRotaryEmbedding(cfg)
    B, T = 2, 40
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, T, 2048, generator=g)
    pad = torch.ones(B, T, dtype=torch.bool)
    pad[0, 31:] = False                                           # right padding, like the language tokens
    pos = torch.cumsum(pad.long(), 1) - 1                         # modeling_mvla.py:765-766
    mask = pad[:, None, :] & pad[:, :, None]                      # att_masks = 0 everywhere: bidirectional over valid
    add = torch.zeros(B, 1, T, T).masked_fill(~mask[:, None], torch.finfo(torch.float32).min)
    with torch.no_grad():
        cos, sin = rope(x, pos)
        ref = layer(x, attention_mask=add, position_ids=pos, position_embeddings=(cos, sin))
    ref = ref[0] if isinstance(ref, tuple) else ref
    orc = Pi0Oracle(OracleDims(n_siglip_layers=1, n_gemma_layers=1), sd, mode="fp32")
    got = orc.gemma_layer(LM, 0, x, pos, mask, {}, fill=True)
    valid = pad[:, :, None].expand_as(ref)                        # pad query rows are never consumed downstream
    err = ((got - ref).abs() * valid).max().item()
    assert err <= 2e-4 * max(1.0, ref.abs().max().item()), err
