"""Host-side logic of the drop-in boundary (no GPU): module surface, checkpoint loading rules,
padding, sharding, the C-ABI library's exported symbols, and the loud failure without CUDA."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest
import torch

import cid_b200
from cid_b200 import DenoiseGenerator, STATE_DICT_SHAPES, get_padding, load_state_safely
from cid_b200 import lib as cid_lib
from cid_b200.loader import pick_state
from cid_b200.sharding import shard_range, shard_sizes
from oracle import denoise_oracle as orc

ROOT = Path(__file__).resolve().parent.parent


def test_state_dict_surface_matches_reference():
    """Same 24 keys, shapes and order as DenoiseGenerator in backend/app.py:42-78."""
    net = DenoiseGenerator()
    sd = net.state_dict()
    assert list(sd) == [
        "down1.0.weight", "down1.0.bias", "down1.2.weight", "down1.2.bias",
        "down2.0.weight", "down2.0.bias", "down2.2.weight", "down2.2.bias",
        "bottleneck.0.weight", "bottleneck.0.bias", "bottleneck.2.weight", "bottleneck.2.bias",
        "up2.weight", "up2.bias", "upconv2.0.weight", "upconv2.0.bias", "upconv2.2.weight", "upconv2.2.bias",
        "up1.weight", "up1.bias", "upconv1.0.weight", "upconv1.0.bias", "upconv1.2.weight", "upconv1.2.bias"]
    for k, v in sd.items():
        assert tuple(v.shape) == STATE_DICT_SHAPES[k] == orc.state_dict_shapes()[k]
        assert v.dtype == torch.float32
    assert net.eval() is net                       # used as PT_MODELS[model].eval() (app.py:423)
    net.load_state_dict(orc.init_state_dict(0), strict=True)


def test_live_reference_state_dict_is_interchangeable():
    from oracle.ref_import import load_reference_app, reference_available

    if not reference_available():
        pytest.skip("/root/reference is not on this machine")
    app = load_reference_app()
    ref = app.DenoiseGenerator()
    ours = DenoiseGenerator()
    assert list(ref.state_dict()) == list(ours.state_dict())
    ours.load_state_dict(ref.state_dict(), strict=True)
    ref.load_state_dict(ours.state_dict(), strict=True)


def test_cpu_tensor_raises_no_fallback():
    net = DenoiseGenerator().eval()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        net(torch.zeros(1, 3, 8, 8))
    with pytest.raises(RuntimeError):
        net.forward_u8(torch.zeros(1, 8, 8, 3, dtype=torch.uint8))
    with pytest.raises(RuntimeError, match=r"expected input \[B,3,H,W\]"):
        net(torch.zeros(1, 4, 8, 8))


@pytest.mark.parametrize("wrap", [None, "generator", "state_dict", "G"])
@pytest.mark.parametrize("prefix", ["", "module."])
def test_load_state_safely_rules(tmp_path, wrap, prefix):
    """Sub-dict candidates and the DataParallel prefix (backend/app.py:257-274)."""
    sd = orc.init_state_dict(5)
    stored = {prefix + k: v for k, v in sd.items()}
    ckpt = stored if wrap is None else {wrap: stored, "epoch": 3, "metric_history": {"psnr": [1.0, 2.0]}}
    path = tmp_path / "ckpt.pth"
    torch.save(ckpt, path)
    net = DenoiseGenerator()
    net.train()
    load_state_safely(net, str(path))
    assert not net.training
    for k, v in net.state_dict().items():
        assert torch.equal(v, sd[k])


def test_load_state_safely_is_non_strict_and_raises_on_missing_file(tmp_path):
    sd = orc.init_state_dict(5)
    partial = {k: v for k, v in sd.items() if k.startswith("down1")}
    partial["not.a.key"] = torch.zeros(3)
    torch.save({"generator": partial}, tmp_path / "p.pth")
    net = DenoiseGenerator()
    before = net.state_dict()["up1.bias"].clone()
    load_state_safely(net, str(tmp_path / "p.pth"))
    assert torch.equal(net.state_dict()["down1.0.weight"], sd["down1.0.weight"])
    assert torch.equal(net.state_dict()["up1.bias"], before)
    with pytest.raises(Exception):                 # caller catches and keeps random init (app.py:333-336)
        load_state_safely(net, str(tmp_path / "missing.pth"))


def test_pick_state_prefers_first_candidate():
    a, b = {"x": torch.zeros(1)}, {"x": torch.ones(1)}
    assert pick_state({"state_dict": b, "generator": a})["x"].item() == 0.0


def test_get_padding_matches_golden(golden_dir):
    for w, h, l, t, r, b in np.load(golden_dir / "get_padding.npz")["table"]:
        assert get_padding((int(w), int(h)), 4, 1) == (l, t, r, b)
    from PIL import Image
    assert get_padding(Image.new("RGB", (130, 127)), 4, 1) == (1, 0, 1, 1)   # SURVEY.md §4 probed case


def test_shard_ranges_partition():
    for n in (0, 1, 7, 256, 2048, 2049):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
                assert a1 == b0
            sizes = shard_sizes(n, world)
            assert sum(sizes) == n and max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _header_symbols():
    text = (ROOT / "include" / "cid_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cid_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(built_library):
    """The C-ABI .so loads on a CPU-only machine and exports everything include/cid_b200.h declares."""
    syms = _header_symbols()
    assert len(syms) >= 15
    so = ctypes.CDLL(str(built_library))
    for s in syms:
        assert hasattr(so, s), f"{s} declared in include/cid_b200.h but not exported"
    assert sorted(cid_lib.EXPORTED_SYMBOLS) == syms   # the ctypes binding mirrors the header 1:1
    lib = cid_lib.load()
    assert lib.cid_version() == cid_lib.ABI_VERSION


def test_library_fails_loudly_without_gpu(built_library):
    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU-only container")
    lib = cid_lib.load()
    h = ctypes.c_void_p()
    assert lib.cid_create(0, ctypes.byref(h)) != 0
    assert b"no CPU fallback" in lib.cid_last_error()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cid_lib.check(1, "cid_create")


def test_product_package_never_imports_the_oracle():
    pkg = ROOT / "celebrity-image-denoiser_b200"
    for f in pkg.rglob("*.py"):
        assert "oracle" not in f.read_text(), f"{f} references the oracle"
    for f in list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")) + list(pkg.rglob("*.h")):
        assert "oracle" not in f.read_text()


def test_public_surface():
    for name in ("DenoiseGenerator", "load_state_safely", "get_padding", "MODEL_CFG", "enhance_image",
                 "enhance_batch_u8", "enhance_batch_host", "patch_app"):
        assert hasattr(cid_b200, name)
    assert cid_b200.MODEL_CFG["denoise"]["pad_divisor"] == 4 and cid_b200.MODEL_CFG["denoise"]["scale"] == 1


def test_micro_batcher_groups_concurrent_requests_by_shape():
    """Route-level micro-batching (backend/app.py:401-406 builds a batch of one per request): requests
    queued together run as one forward per image size, every caller gets its own image's result, errors propagate."""
    import threading

    from cid_b200 import MicroBatcher
    from cid_b200.batcher import _Request, group_by_shape

    reqs = [_Request(np.zeros((h, w, 3), np.uint8)) for h, w in [(8, 8), (8, 12), (8, 8), (8, 8), (8, 12), (8, 8)]]
    groups = group_by_shape(reqs, max_batch=3)
    assert [[r.image.shape[:2] for r in g] for g in groups] == [[(8, 8)] * 3, [(8, 12)] * 2, [(8, 8)]]

    gate = threading.Event()
    seen = []

    def run_batch(host_in):                      # stands in for H2D -> cid_forward_u8_nhwc -> D2H
        gate.wait(5)
        seen.append(tuple(host_in.shape))
        if host_in.shape[1] == 5:
            raise RuntimeError("cid_forward_u8_nhwc: boom")
        return 255 - host_in

    rng = np.random.default_rng(0)
    with MicroBatcher(run_batch=run_batch, device="cpu", max_batch=4, max_wait_ms=200.0) as mb:
        imgs = [rng.integers(0, 256, (6, 7, 3), dtype=np.uint8) for _ in range(6)] + \
               [rng.integers(0, 256, (9, 4, 3), dtype=np.uint8) for _ in range(2)]
        futs = [mb.submit(im) for im in imgs]
        bad = mb.submit(np.zeros((5, 5, 3), np.uint8))
        gate.set()
        for im, f in zip(imgs, futs):
            assert np.array_equal(f.result(10), 255 - im)
        with pytest.raises(RuntimeError, match="boom"):
            bad.result(10)
        with pytest.raises(ValueError):
            mb.submit(np.zeros((4, 4), np.uint8))
    assert sum(mb.batches) == 8 and len(mb.batches) < 8 and max(mb.batches) <= 4     # batched, bounded by max_batch
    assert all(s[1:] in ((6, 7, 3), (9, 4, 3), (5, 5, 3)) for s in seen)
    with pytest.raises(RuntimeError):
        mb.submit(imgs[0])                         # closed
