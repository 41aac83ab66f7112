"""CPU tests of the Keras cGAN path: HDF5/.keras reader, host mirror, and the oracle's anchors
(oracle/keras_cgan_oracle.py header: parity for this path is UNPINNED — no TensorFlow anywhere —
so these tests pin the layer semantics by independent direct summation and by the plausibility of
the reference's own trained weights)."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from cid_b200 import keras_cgan as kc
from cid_b200.keras_h5 import H5Error, read_h5_datasets
from oracle import keras_cgan_oracle as ko

REF_ARCHIVE = Path("/root/reference/backend/weights/cgan_epoch_500.keras")


@pytest.fixture(scope="module")
def trained(golden_dir):
    z = np.load(golden_dir / "keras_cgan_epoch500.npz")
    return {k: z[k] for k in ko.PARAM_SHAPES}, z["face_u8"], z["oracle_out_u8"]


def _psnr(a, b):
    return 10 * np.log10(255.0 ** 2 / ((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean())


def test_shapes_agree_between_product_oracle_and_fixture(trained):
    var, _, _ = trained
    assert kc.KERAS_CGAN_SHAPES == ko.PARAM_SHAPES
    assert {k: v.shape for k, v in var.items()} == ko.PARAM_SHAPES
    assert sum(v.size for v in var.values()) == 530755  # SURVEY.md §2 row 4 (config.json's model)


def test_conv_k4s2_same_matches_direct_summation():
    p = ko.init_params(3)
    x = np.random.default_rng(0).uniform(-1, 1, (8, 12, 3)).astype(np.float32)
    a = ko.naive_conv_k4s2_same(x, p["conv2d.kernel"], p["conv2d.bias"])
    b = F.conv2d(torch.from_numpy(x).permute(2, 0, 1)[None], ko._conv_w(p, "conv2d"),
                 torch.from_numpy(p["conv2d.bias"]), stride=2, padding=1)[0].permute(1, 2, 0).numpy()
    assert a.shape == (4, 6, 64) and np.abs(a - b).max() < 1e-5


def test_convt_k4s2_same_matches_direct_summation():
    p = ko.init_params(4)
    x = np.random.default_rng(1).uniform(-1, 1, (4, 6, 128)).astype(np.float32)
    a = ko.naive_convt_k4s2_same(x, p["conv2d_transpose_1.kernel"], p["conv2d_transpose_1.bias"])
    b = F.conv_transpose2d(torch.from_numpy(x).permute(2, 0, 1)[None], ko._convt_w(p, "conv2d_transpose_1"),
                           torch.from_numpy(p["conv2d_transpose_1.bias"]), stride=2, padding=1)[0].permute(1, 2, 0).numpy()
    assert a.shape == (8, 12, 64) and np.abs(a - b).max() < 1e-4


def test_oracle_output_on_trained_weights_is_stable(trained):
    var, face, stored = trained
    out = ko.enhance_u8(var, face)
    d = np.abs(out.astype(np.int32) - stored.astype(np.int32))
    assert d.max() <= 1 and (d != 0).mean() < 1e-3  # other torch CPU builds may flip a truncation boundary


def test_trained_weights_plausibility_anchors_the_semantics(trained):
    """The reference's trained generator must reproduce its (noisy) input closely under the right layer
    semantics, and every plausible mis-reading of Keras must do visibly worse."""
    var, face, _ = trained
    good = _psnr(ko.enhance_u8(var, face), face)
    assert good > 28.0

    def variant(**changes):
        q = dict(var)
        q.update(changes)
        return _psnr(ko.enhance_u8(q, face), face)

    bn = "batch_normalization_2"
    assert variant(**{bn + ".moving_mean": var[bn + ".moving_variance"],
                      bn + ".moving_variance": var[bn + ".moving_mean"]}) < 12.0      # wrong BN variable order
    assert variant(**{"batch_normalization.gamma": var["batch_normalization.beta"],
                      "batch_normalization.beta": var["batch_normalization.gamma"]}) < 12.0
    assert variant(**{"conv2d_transpose_1.kernel": var["conv2d_transpose_1.kernel"][::-1, ::-1].copy()}) < good - 2.0
    assert variant(**{"conv2d_1.kernel": var["conv2d_1.kernel"][::-1, ::-1].copy()}) < good - 2.0   # flipped taps
    assert variant(**{"conv2d.kernel": np.roll(var["conv2d.kernel"], 1, axis=0)}) < good - 2.0       # wrong pad side


def test_enhance_u8_pads_and_crops_like_the_route():
    p = ko.init_params(0)
    img = np.random.default_rng(2).integers(0, 256, (1, 30, 27, 3), dtype=np.uint8)
    out = ko.enhance_u8(p, img)
    assert out.shape == img.shape and out.dtype == np.uint8
    assert ko.get_padding(27, 30) == (0, 1, 1, 1)


def test_host_mirror_raises_without_cuda():
    net = kc.KerasCGANGenerator(ko.init_params(0), device="cpu")
    assert len(net.inputs) == 1  # the route's single-input branch (backend/app.py:446-458)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        net.predict(np.zeros((1, 8, 8, 3), np.float32))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        net.forward_u8(torch.zeros((1, 8, 8, 3), dtype=torch.uint8))


def test_check_config_accepts_only_the_reference_generator():
    layers = [{"class_name": c, "config": dict(k)} for c, k in kc._EXPECTED_LAYERS]
    kc.check_config({"config": {"layers": layers}})
    bad = json.loads(json.dumps({"config": {"layers": layers}}))
    bad["config"]["layers"][3]["config"]["strides"] = [1, 1]
    with pytest.raises(RuntimeError, match="strides"):
        kc.check_config(bad)
    with pytest.raises(RuntimeError, match="layers"):
        kc.check_config({"config": {"layers": layers[:-1]}})


def test_h5_reader_rejects_garbage():
    with pytest.raises(H5Error):
        read_h5_datasets(b"not an hdf5 file at all")


@pytest.mark.skipif(not REF_ARCHIVE.exists(), reason="reference archive only exists in the build container")
def test_reader_on_the_reference_archive_matches_fixture(trained):
    var, _, _ = trained
    config, datasets = kc.read_keras_archive(REF_ARCHIVE)
    kc.check_config(config)
    got = kc.variables_from_h5(datasets)
    assert set(got) == set(var)
    for k in var:
        assert np.array_equal(got[k], var[k]), k
