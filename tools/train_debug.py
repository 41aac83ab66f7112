"""GPU bring-up of the training step: compare every debug buffer and every parameter gradient of
cid_train_d_backward / cid_train_g_backward (no optimizer step in between) with the instrumented oracle."""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200.training import DenoiseDiscriminator, DenoiseGANTrainStep  # noqa: E402
from cid_b200 import DenoiseGenerator  # noqa: E402
from oracle import train_oracle as to  # noqa: E402


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    den = b.norm().item()
    cos = (a @ b).item() / max(a.norm().item() * den, 1e-300)
    return (a - b).norm().item() / max(den, 1e-300), cos, den


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=4)
    ap.add_argument("--H", type=int, default=32)
    ap.add_argument("--W", type=int, default=32)
    ap.add_argument("--seed", type=int, default=7)
    a = ap.parse_args()
    og, od = to.make_models(a.seed)
    noisy, clean = to.synthetic_pair(11, a.B, a.H, a.W)
    ref = to.instrumented_gradients(og, od, noisy, clean)
    g, d = DenoiseGenerator(), DenoiseDiscriminator()
    g.load_state_dict(og.state_dict()); d.load_state_dict(od.state_dict())
    tr = DenoiseGANTrainStep(g, d, device="cuda:0")
    nz, cl = noisy.cuda().contiguous(), clean.cuda().contiguous()
    tr.d_backward(nz, cl)
    torch.cuda.synchronize()
    print("== D step ==")
    for k, v in tr.grads("discriminator").items():
        e, c, n = rel(v.cpu().reshape(ref["d_grads"][k].shape), ref["d_grads"][k])
        print(f"  d_grad {k:22s} rel {e:9.3e} cos {c:8.5f} |ref| {n:9.3e}")
    tr.g_backward(nz, cl)
    torch.cuda.synchronize()
    L = tr.losses()
    print("losses gpu", L, "\n       ref", {k: ref[k] for k in ("d_loss", "g_loss", "content", "adv")})
    print("== G step buffers ==")
    order = ["y", "d_a0", "d_z1", "d_a1", "d_z2", "d_a2", "d_z3", "d_a3", "d_g3", "d_g2", "d_g1", "d_g0", "gx", "gz", "dz10",
             "gcat1", "dz8", "dz7", "gcat2", "dz5", "dz4", "gp2", "dz3", "dz2", "gp1", "dz1", "dz0"]
    for name in order:
        got = tr.debug_buffer(name, a.B, a.H, a.W)
        want = ref["acts"][name]
        if name == "gcat1" or name == "gcat2":  # the skip half is only the conv's data gradient on both sides
            pass
        e, c, n = rel(got, want)
        print(f"  {name:6s} rel {e:9.3e} cos {c:8.5f} |ref| {n:9.3e}")
    print("== G grads ==")
    for k, v in tr.grads("generator").items():
        e, c, n = rel(v.cpu().reshape(ref["g_grads"][k].shape), ref["g_grads"][k])
        print(f"  g_grad {k:22s} rel {e:9.3e} cos {c:8.5f} |ref| {n:9.3e}")
    print("launches", tr.kernel_launches())


if __name__ == "__main__":
    main()
