"""GPU bring-up diagnostic: per-layer comparison of the CUDA path against the oracle.

    python tools/gpu_debug.py [B H W [gain]]

Prints, for every intermediate buffer, max-abs error relative to the oracle activation range.
(Test tooling: imports oracle/; not part of the product.)
"""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import numpy as np
import torch

from oracle import denoise_oracle as orc
from cid_b200.generator import DenoiseGenerator


def nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous()


def main():
    args = [float(a) for a in sys.argv[1:]]
    B, H, W = (int(args[0]), int(args[1]), int(args[2])) if len(args) >= 3 else (2, 32, 32)
    gain = args[3] if len(args) >= 4 else 2.0
    torch.manual_seed(0)
    sd = orc.init_state_dict(0, gain)
    x = (torch.rand(B, 3, H, W) * 2 - 1)
    inter = {}
    with torch.no_grad():
        y_ref = orc.forward(sd, x, inter)
    net = DenoiseGenerator()
    net.load_state_dict(sd, strict=True)
    net = net.to("cuda").eval()
    import os
    for opt in ("regime", "graph"):
        if os.environ.get("CID_" + opt.upper()):
            net.set_option(opt, int(os.environ["CID_" + opt.upper()]))
    t0 = time.time()
    y = net(x.cuda())
    torch.cuda.synchronize()
    print(f"forward ok in {time.time() - t0:.3f}s  B={B} H={H} W={W} gain={gain}")
    ref_bufs = {
        "a0": nhwc(inter["a0"]),
        "cat1": torch.cat([nhwc(inter["d1"]), nhwc(inter["e1"])], dim=3),
        "p1": nhwc(inter["p1"]),
        "a2": nhwc(inter["a2"]),
        "cat2": torch.cat([nhwc(inter["d2"]), nhwc(inter["e2"])], dim=3),
        "p2": nhwc(inter["p2"]),
        "b0": nhwc(inter["b0"]),
        "b1": nhwc(inter["b1"]),
        "u2a": nhwc(inter["u2a"]),
        "u2b": nhwc(inter["u2b"]),
        "u1a": nhwc(inter["u1a"]),
    }
    order = ["a0", "cat1", "p1", "a2", "cat2", "p2", "b0", "b1", "u2a", "u2b", "u1a"]
    for name in order:
        got = net.debug_activation(name, (B, 3, H, W))
        ref = ref_bufs[name]
        if name in ("cat1", "cat2"):
            c = ref.shape[3] // 2
            for part, sl in (("d", slice(0, c)), ("e", slice(c, 2 * c))):
                d = (got[..., sl] - ref[..., sl]).abs()
                print(f"  {name}[{part}] max|ref|={ref[..., sl].abs().max():.4f} maxerr={d.max():.5f} "
                      f"rel={d.max() / ref[..., sl].abs().max():.5f} meanerr={d.mean():.6f}")
        else:
            d = (got - ref).abs()
            print(f"  {name:5s} max|ref|={ref.abs().max():.4f} maxerr={d.max():.5f} rel={d.max() / ref.abs().max():.5f} "
                  f"meanerr={d.mean():.6f}")
    d = (y.cpu() - y_ref).abs()
    print(f"  out   max|ref|={y_ref.abs().max():.4f} maxerr={d.max():.5f} psnr={orc.psnr(y.cpu(), y_ref):.2f} dB")
    ok = bool(d.max() <= 2e-2)
    print("PARITY", "OK" if ok else "FAIL")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
