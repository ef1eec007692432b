"""Generates tests/golden/*.npz by executing THE REFERENCE'S OWN helper functions.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):
    python tests/golden/make_golden.py
The functions are AST-extracted from /root/reference/src/model/mvla/{modeling_mvla,mllm}.py (those modules
cannot be imported: they pull the missing `lerobot` submodule) and executed unchanged;
`get_safe_dtype` (a lerobot utility that returns float64 on cpu/cuda) is stubbed to identity. Also stored:
the geometry KATs (doctest values of src/utils/geometry.py used by the adapters), bound (de)normalisation
from src/experiments/env_adapters/base.py, image normalisation from src/utils/pipeline.py and msgpack-numpy
wire bytes produced by the reference's own codec.
"""
import ast
import importlib.util
import math
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def extract(path, names):
    src = open(path).read()
    tree = ast.parse(src)
    ns = {"torch": torch, "math": math, "F": F, "Tensor": torch.Tensor, "np": np,
          "get_safe_dtype": lambda dtype, device_type: dtype}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            code = compile(ast.Module(body=[node], type_ignores=[]), path, "exec")
            exec(code, ns)
    missing = [n for n in names if n not in ns]
    assert not missing, missing
    return ns


def load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    mv = extract(f"{REF}/src/model/mvla/modeling_mvla.py",
                 ["create_sinusoidal_pos_embedding", "make_att_2d_masks", "resize_with_pad", "pad_vector"])
    ml = extract(f"{REF}/src/model/mvla/mllm.py", ["apply_rope"])
    g = torch.Generator().manual_seed(1234)

    # ---- masks / position ids (integer, bit-exact) -------------------------------------------------
    B, P, H = 3, 328, 4
    pad = torch.ones(B, P, dtype=torch.bool)
    pad[0, 256 + 10:] = False          # 10 valid text tokens
    pad[1, 256 + 72:] = False          # all valid
    pad[2, 256 + 1:] = False
    att = torch.zeros(B, P, dtype=torch.bool)
    prefix_2d = mv["make_att_2d_masks"](pad, att)
    prefix_pos = torch.cumsum(pad, dim=1) - 1
    S = 1 + H
    spad = torch.ones(B, S, dtype=torch.bool)
    satt = torch.tensor([1, 1] + [0] * (H - 1), dtype=torch.float32)[None].expand(B, S)
    suffix_2d = mv["make_att_2d_masks"](spad, satt)
    full_2d = torch.cat([pad[:, None, :].expand(B, S, P), suffix_2d], dim=2)
    suffix_pos = torch.sum(pad, dim=-1)[:, None] + torch.cumsum(spad, dim=1) - 1
    # a generic block-causal pattern as in the docstring of make_att_2d_masks
    pad2 = torch.tensor([[1, 1, 1, 1, 1, 1, 1, 0, 1, 1]], dtype=torch.bool)
    att2 = torch.tensor([[1, 0, 1, 0, 1, 0, 0, 1, 0, 0]], dtype=torch.int64)
    blk_2d = mv["make_att_2d_masks"](pad2, att2)
    np.savez_compressed(f"{OUT}/masks.npz", pad=pad.numpy(), prefix_2d=np.packbits(prefix_2d.numpy(), axis=-1),
                        prefix_pos=prefix_pos.numpy(), suffix_2d=suffix_2d.numpy(), full_2d=np.packbits(full_2d.numpy(), axis=-1),
                        suffix_pos=suffix_pos.numpy(), pad2=pad2.numpy(), att2=att2.numpy(), blk_2d=blk_2d.numpy(),
                        H=np.int64(H))

    # ---- sinusoidal time embedding (float64) + the fp32 time sequence ------------------------------
    dt = torch.tensor(-0.1, dtype=torch.float32)
    time = torch.tensor(1.0, dtype=torch.float32)
    times = []
    while time >= -dt / 2:
        times.append(time.clone())
        time += dt
    times = torch.stack(times)
    emb = mv["create_sinusoidal_pos_embedding"](times, 1024, min_period=4e-3, max_period=4.0, device=torch.device("cpu"))
    assert emb.dtype == torch.float64
    np.savez_compressed(f"{OUT}/time_emb.npz", times=times.numpy(), emb=emb.numpy())

    # ---- RoPE ------------------------------------------------------------------------------------------
    x = torch.randn(2, 9, 3, 256, generator=g)
    pos = torch.tensor([[0, 1, 2, 3, 4, 5, 265, 327, 332], [7, 7, 8, 100, 255, 256, 300, 331, 340]])
    y = ml["apply_rope"](x, pos)
    xb = x.to(torch.bfloat16)
    yb = ml["apply_rope"](xb, pos)
    np.savez_compressed(f"{OUT}/rope.npz", x=x.numpy(), pos=pos.numpy(), y=y.numpy(),
                        xb=xb.float().numpy(), yb=yb.float().numpy())

    # ---- pad_vector / resize_with_pad ------------------------------------------------------------------
    v = torch.randn(3, 7, generator=g)
    img = torch.rand(2, 3, 48, 64, generator=g)
    np.savez_compressed(f"{OUT}/prepare.npz", v=v.numpy(), v32=mv["pad_vector"](v, 32).numpy(), img=img.numpy(),
                        img_rp=mv["resize_with_pad"](img, 32, 32, pad_value=-1).numpy(),
                        img_same=mv["resize_with_pad"](img, 64, 48, pad_value=-1).numpy())

    # ---- adapters: geometry, bound normalisation, image normalisation (reference modules load by path) ---
    geo = load_by_path("ref_geometry", f"{REF}/src/utils/geometry.py")
    quats = torch.randn(16, 4, generator=g).double().numpy()
    quats /= np.linalg.norm(quats, axis=1, keepdims=True)
    mats = np.stack([geo.quat2mat(q) for q in quats])
    eulers = np.stack([geo.mat2euler(m) for m in mats])
    axangles = []
    for e in eulers:
        ax, ang = geo.euler2axangle(*e)
        axangles.append(ax * ang)
    base = load_by_path("ref_adapter_base", f"{REF}/src/experiments/env_adapters/base.py")
    data = torch.randn(5, 7, generator=g).double().numpy()
    lo, hi = -np.abs(torch.randn(7, generator=g).double().numpy()) - 0.1, np.abs(torch.randn(7, generator=g).double().numpy()) + 0.1
    ad = base.BaseEnvAdapter()
    nb = ad.normalize_bound(data, lo, hi)
    db = ad.denormalize_bound(np.clip(data, -1, 1), lo, hi)
    pipe = load_by_path("ref_pipeline", f"{REF}/src/utils/pipeline.py")
    u8 = (torch.rand(2, 3, 8, 8, generator=g) * 255).to(torch.uint8)
    half = torch.tensor([0.5, 0.5, 0.5])
    pim = pipe.process_images(u8, rescale_factor=1 / 255.0, image_mean=half, image_std=half)  # simpler.py:60-65
    np.savez_compressed(f"{OUT}/adapters.npz", quats=quats, mats=mats, eulers=eulers, axangles=np.stack(axangles),
                        data=data, lo=lo, hi=hi, norm_bound=nb, denorm_bound=db, u8=u8.numpy(),
                        proc_images=pim.numpy())

    # ---- wire format: bytes produced by the reference's own msgpack-numpy codec ---------------------------
    sys.path.insert(0, f"{REF}/packages/policy-server-client/src")
    mnp = load_by_path("ref_msgpack_numpy", f"{REF}/packages/policy-server-client/src/policy_server_client/msgpack_numpy.py")
    act = np.arange(28, dtype=np.float64).reshape(4, 7) / 7.0
    obs = {"observation.images.top": (np.arange(2 * 3 * 3, dtype=np.uint8).reshape(2, 3, 3)),
           "observation.state": np.linspace(0, 1, 8, dtype=np.float32), "task": "put the carrot on the plate",
           "nested": {"agent": {"eef_pos": np.float32(0.5)}}}
    np.savez_compressed(f"{OUT}/wire.npz", act=act, act_bytes=np.frombuffer(mnp.packb(act), dtype=np.uint8),
                        obs_bytes=np.frombuffer(mnp.packb(obs), dtype=np.uint8),
                        reset_bytes=np.frombuffer(mnp.packb({"reset": True}), dtype=np.uint8),
                        status_bytes=np.frombuffer(mnp.packb({"status": "reset"}), dtype=np.uint8))
    # ---- batch adapter (simplerMS3.py:40-132,153-172), restated call by call with the reference's own modules: the
    # file itself cannot be imported (broken `env_adapter` import) ----------------------------------
    import json
    import torch.nn.functional as F
    stats = json.load(open(f"{REF}/config/dataset/bridge_statistics.json"))
    imgs = torch.randint(0, 256, (2, 60, 80, 3), generator=g, dtype=torch.uint8)
    res = F.interpolate(imgs.permute(0, 3, 1, 2), size=(28, 28), mode="bilinear", align_corners=False)
    res_full = F.interpolate(torch.randint(0, 256, (1, 480, 640, 3), generator=torch.Generator().manual_seed(7),
                                           dtype=torch.uint8).permute(0, 3, 1, 2), size=(224, 224), mode="bilinear",
                             align_corners=False)
    pimgs = pipe.process_images(res, rescale_factor=1 / 255.0, image_mean=half, image_std=half)
    eef = torch.randn(3, 8, generator=g).double().numpy()
    eef[:, 3:7] /= np.linalg.norm(eef[:, 3:7], axis=1, keepdims=True)
    default_rot = np.array([[0, 0, 1.0], [0, 1.0, 0], [-1.0, 0, 0]])
    raw = np.stack([np.concatenate([p[:3], geo.mat2euler(geo.quat2mat(p[3:7]) @ default_rot.T), [p[7]]]) for p in eef])
    state = ad.normalize_bound(raw, np.array(stats["proprio"]["p01"]), np.array(stats["proprio"]["p99"]), clip_min=-1, clip_max=1)
    acts = (torch.rand(2, 4, 7, generator=g) * 2.4 - 1.2).numpy().astype(np.float32)
    rex = ad.denormalize_bound(acts[:, :, :-1].astype(np.float64), np.array(stats["action"]["p01"])[:-1],
                               np.array(stats["action"]["p99"])[:-1], clip_min=-1, clip_max=1)
    env = np.zeros((2, 4, 7))
    for b in range(2):
        for t in range(4):
            ax, ang = geo.euler2axangle(*rex[b, t, 3:6])
            env[b, t] = np.concatenate([rex[b, t, :3], ax * ang, [2.0 * (acts[b, t, -1] > 0.5) - 1.0]])
    # single adapter resize: the reference's own call (simpler.py:47-51) on small and full-size frames
    import cv2
    lz_small = cv2.resize(imgs[0].numpy(), (28, 28), interpolation=cv2.INTER_LANCZOS4)
    lz_frame = torch.randint(0, 256, (480, 640, 3), generator=torch.Generator().manual_seed(7), dtype=torch.uint8).numpy()
    lz_full = cv2.resize(lz_frame, (224, 224), interpolation=cv2.INTER_LANCZOS4)
    np.savez_compressed(f"{OUT}/adapter_lanczos.npz", small=lz_small, full_checksum=np.array([int(lz_full.astype(np.int64).sum())]),
                        full_corner=lz_full[:6, :6], full_last=lz_full[-3:, -3:], cv2_version=np.array([cv2.__version__]))
    np.savez_compressed(f"{OUT}/adapter_batch.npz", imgs=imgs.numpy(), resized=res.numpy(), proc=pimgs.numpy(),
                        full_seed=np.array([7]), full_checksum=np.array([int(res_full.long().sum())]),
                        full_corner=res_full[0, :, :4, :4].numpy(), eef=eef, raw=raw, state=state, acts=acts, env=env,
                        p01=np.array(stats["proprio"]["p01"]), p99=np.array(stats["proprio"]["p99"]),
                        a01=np.array(stats["action"]["p01"]), a99=np.array(stats["action"]["p99"]))
    # ---- adapt_to_pi_aloha (modeling_mvla.py:166-218,409-425): the reference's gripper / joint-flip maps, executed here ----
    al = extract(f"{REF}/src/model/mvla/modeling_mvla.py",
                 ["normalize", "unnormalize", "safe_arcsin", "aloha_gripper_to_angular", "aloha_gripper_from_angular"])
    st_in = torch.rand(3, 14, generator=torch.Generator().manual_seed(11), dtype=torch.float32)
    st_out = st_in.clone()
    for m in [1, 2, 8, 9]:
        st_out[:, m] *= -1
    for m in [6, 13]:
        st_out[:, m] = al["aloha_gripper_to_angular"](st_out[:, m])
    ac_in = torch.rand(3, 5, 14, generator=torch.Generator().manual_seed(12), dtype=torch.float32) * 2 - 1
    ac_out = ac_in.clone()
    for m in [1, 2, 8, 9]:
        ac_out[:, :, m] *= -1
    for m in [6, 13]:
        ac_out[:, :, m] = al["aloha_gripper_from_angular"](ac_out[:, :, m])
    np.savez_compressed(f"{OUT}/aloha.npz", state_in=st_in.numpy(), state_out=st_out.numpy(), act_in=ac_in.numpy(),
                        act_out=ac_out.numpy())
    print("golden fixtures written to", OUT)


if __name__ == "__main__":
    main()
