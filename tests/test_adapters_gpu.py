"""GPU adapters (csrc/adapters.cu through the C ABI) against the adapter oracle and the reference-generated vectors.
Bit-exact for the uint8 resize + fp32 normalisation; fp64 geometry within 1e-12 (device libm vs host libm)."""
import os

import numpy as np
import pytest
import torch

from oracle import adapters_oracle as ao

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _stats(d):
    return {"proprio": {"p01": d["p01"].tolist(), "p99": d["p99"].tolist(), "mean": d["p01"].tolist(), "std": (d["p99"] - d["p01"]).tolist()},
            "action": {"p01": d["a01"].tolist(), "p99": d["a99"].tolist(), "mean": d["a01"].tolist(), "std": (d["a99"] - d["a01"]).tolist()}}


def _adapter(cls, stats, image_size=(224, 224), bf16=False, **kw):
    from int_act_b200 import adapters as ad

    cfg = ad.adapter_config("unused", image_size=image_size, use_bf16=bf16, **kw)
    return getattr(ad, cls)(cfg, dataset_statistics=stats)


def test_batch_preprocess_matches_reference_vectors(cuda):
    d = np.load(os.path.join(G, "adapter_batch.npz"))
    a = _adapter("BridgeSimplerBatchAdapter", _stats(d), image_size=(28, 28))
    out = a.preprocess({"observation.images.top": d["imgs"], "observation.state": d["eef"], "task": ["a", "b"]})
    assert out["task"] == ["a", "b"]
    assert np.array_equal(out["observation.images.top"].cpu().numpy(), d["proc"])          # bit-exact
    st = a.proprio(torch.from_numpy(d["eef"]).cuda()).cpu().numpy()
    assert np.abs(st - d["state"].astype(np.float32)).max() <= 1e-6                        # fp64 math, f32 store


@pytest.mark.parametrize("hw", [(480, 640), (224, 224), (100, 37), (512, 512)])
def test_resize_bit_exact_vs_oracle_full_size(cuda, hw):
    d = np.load(os.path.join(G, "adapter_batch.npz"))
    a = _adapter("BridgeSimplerBatchAdapter", _stats(d))
    frames = torch.randint(0, 256, (3, hw[0], hw[1], 3), generator=torch.Generator().manual_seed(3), dtype=torch.uint8)
    got = a.resize_normalize(frames.cuda()).cpu().numpy()
    ref = ao.process_images(np.stack([ao.resize_u8_bilinear(f, 224, 224) for f in frames.numpy()]).transpose(0, 3, 1, 2))
    assert np.array_equal(got, ref)


def test_proprio_and_postprocess_vs_oracle(cuda):
    d = np.load(os.path.join(G, "adapter_batch.npz"))
    stats = _stats(d)
    g = torch.Generator().manual_seed(11)
    eef = torch.randn(257, 8, generator=g, dtype=torch.float64)
    eef[0, 3:7] = torch.tensor([1.0, 0.0, 0.0, 0.0], dtype=torch.float64)       # identity rotation
    eef[1, 3:7] = torch.tensor([0.5, 0.5, 0.5, 0.5], dtype=torch.float64)
    eef[2, 3:7] = torch.tensor([0.7071067811865476, 0.0, 0.7071067811865476, 0.0], dtype=torch.float64)  # gimbal lock
    for norm in ("bound", "gaussian"):
        a = _adapter("BridgeSimplerBatchAdapter", stats, state_normalization_type=norm, action_normalization_type=norm)
        raw = ao.bridge_preprocess_proprio(eef.numpy())
        want = ao.normalize_bound(raw, d["p01"], d["p99"]) if norm == "bound" else \
            ao.normalize_gaussian(raw, d["p01"], d["p99"] - d["p01"])
        got = a.proprio(eef.cuda()).cpu().numpy()
        assert np.abs(got - want).max() <= 2e-6, norm
        # float32 input is widened to float64 on the device
        got32 = a.proprio(eef.float().cuda()).cpu().numpy()
        assert np.abs(got32 - want).max() <= 1e-4
        acts = (torch.rand(64, 4, 7, generator=g) * 2.4 - 1.2).float()
        acts[0, 0, 3:6] = 0.0  # zero rotation -> identity branch of quat2axangle... after de-normalisation not zero; keep
        env = a.postprocess(acts.cuda())
        assert env.dtype == np.float64 and env.shape == (64, 4, 7)
        st = {"action": {"p01": d["a01"], "p99": d["a99"], "mean": d["a01"], "std": d["a99"] - d["a01"]}}
        want_env = ao.bridge_postprocess(acts.numpy(), st, norm=norm)
        assert np.abs(env - want_env).max() <= 1e-12, norm
        assert set(np.unique(env[..., 6])) <= {-1.0, 1.0}
    # reference-generated vector
    a = _adapter("BridgeSimplerBatchAdapter", stats)
    assert np.abs(a.postprocess(d["acts"]) - d["env"]).max() <= 1e-12
    # bf16 adapter dtype: state carries the f64 -> bf16 rounding of torch.as_tensor(proprio, dtype=bfloat16)
    ab = _adapter("BridgeSimplerBatchAdapter", stats, bf16=True)
    got = ab.proprio(torch.from_numpy(d["eef"]).cuda()).cpu()
    assert torch.equal(got, torch.as_tensor(d["state"], dtype=torch.bfloat16).float())


@pytest.mark.parametrize("hw", [(480, 640), (224, 224), (100, 37), (512, 512)])
def test_lanczos4_bit_exact_vs_oracle_and_opencv(cuda, hw):
    d = np.load(os.path.join(G, "adapter_batch.npz"))
    a = _adapter("BridgeSimplerAdapter", _stats(d))
    frames = torch.randint(0, 256, (2, hw[0], hw[1], 3), generator=torch.Generator().manual_seed(4), dtype=torch.uint8)
    got = a.lanczos4_normalize(frames.cuda()).cpu().numpy()
    ref = ao.process_images(np.stack([ao.resize_u8_lanczos4(f, 224, 224) for f in frames.numpy()]).transpose(0, 3, 1, 2))
    assert np.array_equal(got, ref)
    try:
        import cv2
    except ImportError:
        return
    ref2 = np.stack([cv2.resize(f, (224, 224), interpolation=cv2.INTER_LANCZOS4) for f in frames.numpy()])
    assert np.array_equal(got, ao.process_images(ref2.transpose(0, 3, 1, 2)))


def test_single_adapter_matches_reference_calls(cuda):
    d = np.load(os.path.join(G, "adapter_batch.npz"))
    lz = np.load(os.path.join(G, "adapter_lanczos.npz"))
    a = _adapter("BridgeSimplerAdapter", _stats(d))
    frame = torch.randint(0, 256, (480, 640, 3), generator=torch.Generator().manual_seed(7), dtype=torch.uint8).numpy()
    eef = d["eef"][0].astype(np.float32)
    out = a.preprocess({"observation.images.top": frame, "observation.state": {"agent": {"eef_pos": eef}}, "task": "t"})
    img = out["observation.images.top"].cpu().numpy()
    # undo the normalisation exactly: (x * (1/255) - 0.5) / 0.5 is injective on 0..255
    lut = ao.process_images(np.arange(256, dtype=np.uint8).reshape(1, 1, 1, 256))[0, 0, 0]
    u8 = np.searchsorted(lut, img).astype(np.uint8)
    assert np.array_equal(lut[u8], img)
    assert int(u8.astype(np.int64).sum()) == int(lz["full_checksum"][0])           # the frame cv2 resized at fixture time
    assert np.array_equal(u8[0].transpose(1, 2, 0)[:6, :6], lz["full_corner"])
    assert out["task"] == ["t"] and tuple(out["observation.state"].shape) == (1, 7)
    env = a.postprocess(torch.from_numpy(d["acts"][0]).cuda())
    assert env.shape == (4, 7) and np.abs(env - d["env"][0]).max() <= 1e-12
