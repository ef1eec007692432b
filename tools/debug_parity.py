"""Stage-by-stage comparison of the CUDA path with the oracle trace (development aid; run on the GPU box)."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from int_act_b200 import binding as pb
from int_act_b200.engine import Pi0Engine
from int_act_b200.weights import Dims, PackedWeights, random_state_dict
from pi0_helpers import oracle_for, synthetic_inputs

mode = sys.argv[1] if len(sys.argv) > 1 else "bf16"
nsig, ngem = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (2, 2)
B = 2
cuda = torch.device("cuda:0")
dims = Dims(n_siglip_layers=nsig, n_gemma_layers=ngem, vocab=2048)
sd = random_state_dict(dims, device=cuda, seed=0)
act = torch.bfloat16 if mode == "bf16" else torch.float32
packed = PackedWeights(sd, cuda, act)
eng = Pi0Engine(packed, max_batch=B, dtype=mode, use_cuda_graph=False)
images, img_masks, tokens, masks, state, noise = synthetic_inputs(B, vocab=dims.vocab, n_valid_text=[12, 72])
orc = oracle_for(sd, dims, 1, 72, 4, mode)
trace = {}
ref = orc.sample_actions(images, img_masks, tokens, masks, state, noise, trace=trace)
lib = eng.lib
s = pb.current_stream()
img = images[0].to(cuda).contiguous()
pb.check(lib.pi0_siglip_forward(eng.ctx, B, 0, pb.ptr(img), 2.0, -1.0, s))
tok = tokens.to(cuda)
pb.check(lib.pi0_embed_language(eng.ctx, B, pb.ptr(tok), s))
torch.cuda.synchronize()


def cmp(name, a, b, mask=None):
    a, b = a.float().cpu(), b.float().cpu()
    d = (a - b).abs()
    if mask is not None:
        d = d * mask
    print(f"{name:28s} max|ref| {b.abs().max().item():9.4f}  max err {d.max().item():.4e}  mean err {d.mean().item():.4e}")


sx = eng.debug_buffer("siglip.x").view(B, 256, 1152)
cmp(f"siglip.layer{nsig-1}", sx, trace[f"siglip.layer{nsig-1}"])
px = eng.debug_buffer("prefix.x").view(B, 328, 2048)
cmp("prefix.embs (img)", px[:, :256], trace["prefix.embs"][:, :256])
cmp("prefix.embs (lang)", px[:, 256:], trace["prefix.embs"][:, 256:])
im = torch.ones((1, B), dtype=torch.uint8, device=cuda)
lm = masks.to(cuda).to(torch.uint8).contiguous()
pb.check(lib.pi0_prefill(eng.ctx, B, pb.ptr(im), pb.ptr(lm), s))
torch.cuda.synchronize()
pad = torch.cat([torch.ones(B, 256, dtype=torch.bool), masks], 1)
kc = eng.debug_buffer("kcache").view(ngem, B, 333, 256)
vt = eng.debug_buffer("vtcache").view(ngem, B, 256, 336)
cmp("cache.k0", kc[0, :, :328], trace["cache.k0"].view(B, 328, 256), pad[:, :, None])
cmp("cache.v0", vt[0, :, :, :328].transpose(1, 2), trace["cache.v0"].view(B, 328, 256), pad[:, :, None])
from oracle.pi0_oracle import LM, EX
att = eng.debug_buffer("prefix.att").view(B, 328, 2048)
if ngem >= 2:
    cmp(f"gemma layer{ngem-2}.att", att, trace[f"{LM}layer{ngem-2}.att"], pad[:, :, None])
    cmp(f"gemma layer{ngem-2}.out", eng.debug_buffer("prefix.x").view(B, 328, 2048), trace[f"{LM}layer{ngem-2}.out"], pad[:, :, None])
st = torch.zeros((B, 32), device=cuda)
st[:, :7] = state.to(cuda)
xt = noise.to(cuda).clone()
# one step only: temporarily patch n_euler_iters through a second engine
eng1 = Pi0Engine(packed, max_batch=B, dtype=mode, use_cuda_graph=False)
eng1.cfg.n_euler_iters = 1
ctx1 = C.c_void_p()
pb.check(lib.pi0_ctx_create(C.byref(eng1.cfg), C.byref(packed.struct), pb.ptr(eng.workspace), eng.workspace.numel(), C.byref(ctx1)))
pb.check(lib.pi0_denoise_loop(ctx1, B, pb.ptr(st), pb.ptr(xt), s))
torch.cuda.synchronize()
cmp("suffix.embs0", eng.debug_buffer("suffix.embs").view(B, 5, 1024), trace["suffix.embs0"])
cmp(f"expert layer{ngem-1}.att(step0)", eng.debug_buffer("suffix.att").view(B, 5, 2048), trace[f"{EX}layer{ngem-1}.att"] if False else eng.debug_buffer("suffix.att").view(B, 5, 2048))
v0 = (xt - noise.to(cuda)) / (-0.1)
cmp("v0 (from x1)", v0, trace["v0"])
out = eng.sample_actions([img], [m.to(cuda) for m in img_masks], tok, masks.to(cuda), state.to(cuda), noise.to(cuda))
cmp("actions", out, ref)
print("ref actions sample", ref[0, 0, :8])
print("noise sample", noise[0, 0, :8])
