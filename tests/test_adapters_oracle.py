"""The adapter oracle (oracle/adapters_oracle.py) against vectors produced by the reference's own modules
(tests/golden/make_golden.py: geometry.py, env_adapters/base.py, utils/pipeline.py, torch's uint8 bilinear resize), and the
product-side resize tables against the oracle's. CPU only."""
import os

import numpy as np

from oracle import adapters_oracle as ao

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_geometry_against_reference_vectors():
    d = np.load(os.path.join(G, "adapters.npz"))
    for q, m, e, a in zip(d["quats"], d["mats"], d["eulers"], d["axangles"]):
        assert np.array_equal(ao.quat2mat(q), m)
        assert np.array_equal(np.array(ao.mat2euler_sxyz(m)), e)
        assert np.array_equal(ao.euler2axangle_sxyz(*e), a)
    assert np.array_equal(ao.normalize_bound(d["data"], d["lo"], d["hi"]), d["norm_bound"])
    assert np.array_equal(ao.denormalize_bound(np.clip(d["data"], -1, 1), d["lo"], d["hi"]), d["denorm_bound"])
    assert np.array_equal(ao.process_images(d["u8"]), d["proc_images"])


def test_batch_adapter_against_reference_vectors():
    d = np.load(os.path.join(G, "adapter_batch.npz"))
    stats = {"proprio": {"p01": d["p01"], "p99": d["p99"]}, "action": {"p01": d["a01"], "p99": d["a99"]}}
    # uint8 bilinear resize: bit-exact with F.interpolate(uint8) (integer work)
    res = np.stack([ao.resize_u8_bilinear(im, 28, 28) for im in d["imgs"]]).transpose(0, 3, 1, 2)
    assert np.array_equal(res, d["resized"])
    imgs, state = ao.bridge_batch_preprocess(d["imgs"], d["eef"], stats, image_size=(28, 28))
    assert np.array_equal(imgs, d["proc"])
    assert np.array_equal(ao.bridge_preprocess_proprio(d["eef"]), d["raw"])
    assert np.array_equal(state, d["state"])
    assert np.array_equal(ao.bridge_postprocess(d["acts"], stats), d["env"])
    # full-size frame (480x640 -> 224x224): checksum + corner of the reference's output
    x = np.random.default_rng(0)  # noqa: F841  (the frame itself is regenerated from the torch seed below)
    import torch

    frame = torch.randint(0, 256, (1, 480, 640, 3), generator=torch.Generator().manual_seed(int(d["full_seed"][0])),
                          dtype=torch.uint8).numpy()[0]
    r = ao.resize_u8_bilinear(frame, 224, 224)
    assert int(r.astype(np.int64).sum()) == int(d["full_checksum"][0])
    assert np.array_equal(r.transpose(2, 0, 1)[:, :4, :4], d["full_corner"])


def test_lanczos4_against_opencv_vectors():
    """The single adapter's resize (simpler.py:47-51): the integer restatement of OpenCV's uint8 INTER_LANCZOS4 against
    frames resized by cv2 itself when the fixtures were generated (and against the installed cv2, when there is one)."""
    import torch

    d = np.load(os.path.join(G, "adapter_lanczos.npz"))
    imgs = np.load(os.path.join(G, "adapter_batch.npz"))["imgs"]
    assert np.array_equal(ao.resize_u8_lanczos4(imgs[0], 28, 28), d["small"])
    frame = torch.randint(0, 256, (480, 640, 3), generator=torch.Generator().manual_seed(7), dtype=torch.uint8).numpy()
    r = ao.resize_u8_lanczos4(frame, 224, 224)
    assert int(r.astype(np.int64).sum()) == int(d["full_checksum"][0])
    assert np.array_equal(r[:6, :6], d["full_corner"]) and np.array_equal(r[-3:, -3:], d["full_last"])
    try:
        import cv2
    except ImportError:
        return
    rng = np.random.default_rng(2)
    for (H, W, oh, ow) in ((48, 64, 22, 22), (100, 37, 64, 48), (256, 256, 224, 224)):
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        assert np.array_equal(ao.resize_u8_lanczos4(img, oh, ow), cv2.resize(img, (ow, oh), interpolation=cv2.INTER_LANCZOS4))


def test_product_resize_tables_match_oracle():
    from int_act_b200.adapters import lanczos4_tables, u8_bilinear_tables

    for n_in, n_out in ((640, 224), (480, 224), (224, 224), (37, 64)):
        a, b = lanczos4_tables(n_in, n_out), ao.lanczos4_tables(n_in, n_out)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])

    for n_in, n_out in ((640, 224), (480, 224), (224, 224), (37, 224), (512, 224), (60, 28)):
        a, b = u8_bilinear_tables(n_in, n_out), ao.u8_bilinear_tables(n_in, n_out)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2]
