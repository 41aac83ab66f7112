"""The training-step oracle (oracle/train_oracle.py) against golden vectors minted from the reference's own classes
(oracle/make_golden_train.py -> tests/golden/train_step.npz), and the host-side training state table."""
from pathlib import Path

import numpy as np
import torch

from oracle import train_oracle as to

GOLDEN = Path(__file__).parent / "golden" / "train_step.npz"


def test_oracle_matches_reference_iterations():
    gold = np.load(GOLDEN)
    torch.manual_seed(7)
    g, d = to.DenoiseGenerator(), to.DenoiseDiscriminator()
    assert sorted(k for k, _ in d.named_parameters()) == list(gold["d_keys"])
    assert sorted(k for k, _ in g.named_parameters()) == list(gold["g_keys"])
    g_opt, d_opt = to.make_optimizers(g, d)
    noisy, clean = to.synthetic_pair(11, 4, 32, 32)
    for it in range(2):
        r = to.train_iteration(g, d, g_opt, d_opt, noisy, clean)
        losses = np.array([r["d_loss"], r["g_loss"], r["content"], r["adv"]])
        np.testing.assert_allclose(losses, gold[f"it{it}_losses"], rtol=1e-5, atol=1e-7)
        dn = np.array([float(r["d_grads"][k].norm()) for k in sorted(r["d_grads"])])
        gn = np.array([float(r["g_grads"][k].norm()) for k in sorted(r["g_grads"])])
        np.testing.assert_allclose(dn, gold[f"it{it}_d_gradnorm"], rtol=2e-4, atol=1e-8)
        np.testing.assert_allclose(gn, gold[f"it{it}_g_gradnorm"], rtol=2e-4, atol=1e-8)
        np.testing.assert_allclose(r["d_grads"]["model.0.weight"].numpy(), gold[f"it{it}_d_grad_model0"], rtol=1e-3, atol=1e-7)
        np.testing.assert_allclose(r["g_grads"]["upconv1.2.weight"].numpy(), gold[f"it{it}_g_grad_upconv1_2"], rtol=1e-3, atol=1e-7)
    gs = np.array([float(p.double().sum()) for _, p in sorted(g.named_parameters())])
    ds = np.array([float(p.double().sum()) for _, p in sorted(d.named_parameters())])
    np.testing.assert_allclose(gs, gold["g_param_sums"], rtol=1e-5, atol=1e-4)
    np.testing.assert_allclose(ds, gold["d_param_sums"], rtol=1e-5, atol=1e-4)
    np.testing.assert_allclose(d.model[3].running_mean.numpy(), gold["d_running_mean3"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(d.model[9].running_var.numpy(), gold["d_running_var9"], rtol=1e-4, atol=1e-6)


def test_parameter_counts_match_survey():
    g, d = to.make_models(0)
    assert sum(p.numel() for p in g.parameters()) == 1_827_587
    assert sum(p.numel() for p in d.parameters()) == 260_929


def test_host_mirror_has_reference_keys():
    from cid_b200.training import DenoiseDiscriminator
    ref = to.DenoiseDiscriminator().state_dict()
    mine = DenoiseDiscriminator().state_dict()
    assert list(ref) == list(mine)
    assert all(ref[k].shape == mine[k].shape for k in ref)


def test_bf16_storage_sensitivity():
    """Why the discriminator's gradient tolerance is loose: round the discriminator's stored activations (and weights) to
    bf16 inside the ORACLE and its own weight gradients move by 10-40 % (cosine 0.9-0.995), while the losses move by
    < 1e-3 - the LeakyReLU on/off pattern under a global average pool is what carries the gradient."""
    import torch.nn as nn
    g, d = to.make_models(7)
    noisy, clean = to.synthetic_pair(11, 4, 32, 32)
    ref = to.instrumented_gradients(g, d, noisy, clean)
    q = lambda t: t + (t.bfloat16().float() - t).detach()      # straight-through bf16 rounding
    m = d.model
    def disc_bf16(x):
        a0 = q(m[1](m[0](x)))
        a1 = q(m[4](m[3](nn.functional.conv2d(a0, q(m[2].weight), m[2].bias, stride=2, padding=1))))
        a2 = q(m[7](m[6](nn.functional.conv2d(a1, q(m[5].weight), m[5].bias, padding=1))))
        a3 = q(m[10](m[9](nn.functional.conv2d(a2, q(m[8].weight), m[8].bias, stride=2, padding=1))))
        return m[13](m[12](m[11](a3))).view(-1)
    bce = nn.BCELoss()
    d.zero_grad()
    with torch.enable_grad():
        y = ref["acts"]["y"]
        loss = bce(disc_bf16(clean), torch.ones(4)) + bce(disc_bf16(y), torch.zeros(4))
        loss.backward()
    assert abs(float(loss) - ref["d_loss"]) < 2e-3
    w = dict(d.named_parameters())["model.2.weight"].grad
    r = ref["d_grads"]["model.2.weight"]
    rel = float((w - r).norm() / r.norm())
    assert 0.03 < rel < 0.6, rel
