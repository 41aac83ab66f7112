"""Parity of the CUDA training step (C-ABI cid_train_*) against the oracle (oracle/train_oracle.py, pinned to the
reference's classes by tests/test_train_oracle.py).

Tolerances. The path computes in bf16 (fp32 accumulate, fp32 master weights); the reference trains in fp32. Losses and
the forward agree to ~1e-3. Gradients carry the bf16 round-off of every stored activation, and one effect dominates:
an activation whose stored value differs by its last bf16 bit can land on the other side of zero, which flips the
ReLU / LeakyReLU derivative of that element (relative gradient error ~ sqrt(fraction flipped)). It averages out in the
generator's weight gradients (sums over >= 10^3 pixels: a few 1e-3 .. 5e-2 relative, cosine >= 0.997). The
discriminator is the hard case: after the global average pool every pixel of a (sample, channel) plane receives the
SAME upstream gradient, so after BatchNorm's mean subtraction the signal IS the LeakyReLU on/off pattern; there a
2 % forward difference (bf16 after three BatchNorm layers on smooth, DC-dominated images) moves the gradient by
~15-30 % (cosine 0.95-0.99). tests/test_train_oracle.py::test_bf16_storage_sensitivity shows the same figure for the
oracle itself with bf16-rounded activations, i.e. it is the precision, not the kernels - and
test_discriminator_backward_kernels_given_their_own_forward shows it from the other side: with the forward held fixed
(the GPU pass's own stored tensors), every backward stage of the discriminator agrees with fp32 torch to 0.3 %.
Conv biases that feed a BatchNorm have a mathematically zero gradient (1e-8 round-off in the reference, ~1e-4..1e-3 of
bf16 round-off here; Adam turns either into a +-lr random walk that BatchNorm cancels): excluded from the comparison.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from cid_b200 import DenoiseGenerator
from cid_b200.training import DenoiseDiscriminator, DenoiseGANTrainStep
from oracle import train_oracle as to

ZERO_GRAD_KEYS = {"model.2.bias", "model.5.bias", "model.8.bias"}  # conv bias followed by BatchNorm


def _rel_cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm()), float((a @ b) / (a.norm() * b.norm()))


def _make(seed, B, H, W):
    og, od = to.make_models(seed)
    noisy, clean = to.synthetic_pair(seed + 4, B, H, W)
    g, d = DenoiseGenerator(), DenoiseDiscriminator()
    g.load_state_dict(og.state_dict())
    d.load_state_dict(od.state_dict())
    tr = DenoiseGANTrainStep(g, d, device="cuda:0")
    return og, od, tr, noisy, clean


@pytest.mark.parametrize("B,H,W", [(4, 32, 32), (3, 40, 24), (2, 16, 16), (5, 64, 48)])
def test_gradients_match_oracle(B, H, W):
    og, od, tr, noisy, clean = _make(7, B, H, W)
    ref = to.instrumented_gradients(og, od, noisy, clean)
    nz, cl = noisy.cuda().contiguous(), clean.cuda().contiguous()
    tr.d_backward(nz, cl)
    d_grads = tr.grads("discriminator")
    tr.g_backward(nz, cl)  # no optimizer step in between: same discriminator as the oracle's instrumented pass
    g_grads = tr.grads("generator")
    L = tr.losses()
    for k in ("d_loss", "g_loss", "content", "adv"):
        assert abs(L[k] - ref[k]) <= 2e-3 * abs(ref[k]) + 1e-5, (k, L[k], ref[k])
    y = tr.debug_buffer("y", B, H, W)
    assert float((y - ref["acts"]["y"]).abs().max()) <= 2e-2  # the forward's north-star tolerance
    for k, want in ref["g_grads"].items():
        rel, cos = _rel_cos(g_grads[k].cpu().reshape(want.shape), want)
        assert cos >= 0.997 and rel <= 8e-2, ("generator", k, rel, cos)
    for k, want in ref["d_grads"].items():
        if k in ZERO_GRAD_KEYS:
            # sum of bf16-stored dz values whose exact sum is zero: pure round-off, ~sqrt(pixels) * |dz| * 2^-9
            assert float(d_grads[k].abs().max()) <= 1e-2
            continue
        rel, cos = _rel_cos(d_grads[k].cpu().reshape(want.shape), want)
        assert cos >= 0.93, ("discriminator", k, rel, cos)
    # head-side gradients do not pass through a flipped LeakyReLU pattern: tight
    for k in ("model.12.weight", "model.12.bias", "model.9.weight"):
        rel, cos = _rel_cos(d_grads[k].cpu().reshape(ref["d_grads"][k].shape), ref["d_grads"][k])
        assert rel <= 5e-2, (k, rel, cos)
    tr.close()


@pytest.mark.parametrize("B,H,W", [(2, 24, 20), (4, 36, 44)])
def test_generator_gradients_with_partial_tiles_in_both_dimensions(B, H, W):
    """W % 8 = 4 and H % 16 = 8 / 4: every kernel of the step (tile 8 x 16 pixels) clips a tile in x AND in y at every level,
    including the fused 3-channel weight-gradient kernel's source windows. Generator side only: the discriminator's
    end-to-end tolerance is precision-limited at such small batches (its kernels are checked at these shapes by
    test_discriminator_backward_kernels_given_their_own_forward)."""
    og, od, tr, noisy, clean = _make(9, B, H, W)
    ref = to.instrumented_gradients(og, od, noisy, clean)
    nz, cl = noisy.cuda().contiguous(), clean.cuda().contiguous()
    tr.d_backward(nz, cl)
    tr.g_backward(nz, cl)
    g_grads = tr.grads("generator")
    L = tr.losses()
    for k in ("d_loss", "g_loss", "content", "adv"):
        assert abs(L[k] - ref[k]) <= 2e-3 * abs(ref[k]) + 1e-5, (k, L[k], ref[k])
    for k, want in ref["g_grads"].items():
        rel, cos = _rel_cos(g_grads[k].cpu().reshape(want.shape), want)
        assert cos >= 0.997 and rel <= 8e-2, ("generator", k, rel, cos)
    tr.close()


def test_activation_gradients_match_oracle():
    B, H, W = 4, 32, 32
    og, od, tr, noisy, clean = _make(3, B, H, W)
    ref = to.instrumented_gradients(og, od, noisy, clean)
    nz, cl = noisy.cuda().contiguous(), clean.cuda().contiguous()
    tr.d_backward(nz, cl)
    tr.g_backward(nz, cl)
    for name in ("d_a0", "d_z1", "d_a1", "d_z2", "d_a2", "d_z3", "d_a3"):
        rel, cos = _rel_cos(tr.debug_buffer(name, B, H, W), ref["acts"][name])
        assert rel <= 5e-2, (name, rel)
    for name in ("gz", "dz10", "gcat1", "dz8", "dz7", "gcat2", "dz5", "dz4", "gp2", "dz3", "dz2", "gp1", "dz1", "dz0"):
        rel, cos = _rel_cos(tr.debug_buffer(name, B, H, W), ref["acts"][name])
        assert cos >= 0.99, (name, rel, cos)
    tr.close()


@pytest.mark.parametrize("B,H,W", [(4, 32, 32), (2, 24, 20), (4, 36, 44)])
def test_discriminator_backward_kernels_given_their_own_forward(B, H, W):
    """(The second and third shapes leave partial tiles in BOTH dimensions at every level: W % 8 = 4, H % 16 = 8 / 4.)
    Separates precision from kernels for the discriminator (whose end-to-end gradient tolerance is loose because bf16
    forward differences flip LeakyReLU signs): every backward stage is recomputed in fp32 torch FROM THE TENSORS THE GPU
    PASS ITSELF STORED (its fp32 pre-BatchNorm z, its bf16 activations, its own upstream gradient) and compared with
    what the next kernel produced. With the forward held fixed only the backward kernels' own round-off is left
    (bf16 weights in the data-gradient convs, bf16 storage of each result): tight."""
    F = torch.nn.functional
    og, od, tr, noisy, clean = _make(5, B, H, W)
    tr.d_backward(noisy.cuda().contiguous(), clean.cuda().contiguous())  # buffers now hold the pass on G(noisy), target 0
    m = od.model
    nchw = lambda name: tr.debug_buffer(name, B, H, W).permute(0, 3, 1, 2).contiguous()
    z = {l: nchw(f"d_z{l}") for l in (1, 2, 3)}
    a0 = nchw("d_a0")
    g = {l: nchw(f"d_g{l}") for l in (0, 1, 2, 3)}
    bn_of = {1: m[3], 2: m[6], 3: m[9]}

    def bn_lrelu(l):
        zt = z[l].clone().requires_grad_(True)
        bn = bn_of[l]
        return F.leaky_relu(F.batch_norm(zt, None, None, bn.weight.detach(), bn.bias.detach(), training=True, eps=bn.eps), 0.2), zt

    got = {}
    with torch.enable_grad():
        # layer 3: average pool, 1x1 conv, sigmoid, BCE(mean) against zeros, back through LeakyReLU + train-mode BatchNorm
        y3, z3 = bn_lrelu(3)
        pred = torch.sigmoid(F.conv2d(y3.mean((2, 3), keepdim=True), m[12].weight.detach(), m[12].bias.detach())).view(-1)
        F.binary_cross_entropy(pred, torch.zeros(B)).backward()
        got[3] = z3.grad
        # layers 2, 1: the conv's data gradient of the GPU's OWN dz above, then LeakyReLU + BatchNorm backward on its own z
        for l, conv, stride in ((2, m[8], 2), (1, m[5], 1)):
            y, zt = bn_lrelu(l)
            y.backward(torch.nn.grad.conv2d_input(y.shape, conv.weight.detach(), g[l + 1], stride=stride, padding=1))
            got[l] = zt.grad
    # layer 0: data gradient of conv1 (stride 2) times the LeakyReLU derivative of the stored activation
    dA0 = torch.nn.grad.conv2d_input(a0.shape, m[2].weight.detach(), g[1], stride=2, padding=1)
    got[0] = dA0 * torch.where(a0 > 0, 1.0, 0.2)
    for l in (3, 2, 1, 0):
        rel, cos = _rel_cos(g[l], got[l])
        assert rel <= 8e-3 and cos >= 0.9999, (l, rel, cos)   # measured: rel 0.17-0.29 %, cosine >= 0.999996
    tr.close()


def test_full_iterations_track_the_reference_loop():
    """Two complete iterations (D step, Adam, G step, Adam) against oracle.train_iteration: losses, the first Adam
    update (= -lr * g / (|g| + eps): the sign of every gradient element) and the BatchNorm running statistics."""
    B, H, W = 4, 32, 32
    og, od, tr, noisy, clean = _make(11, B, H, W)
    g0 = {k: v.detach().clone() for k, v in og.state_dict().items()}
    g_opt, d_opt = to.make_optimizers(og, od)
    nz, cl = noisy.cuda().contiguous(), clean.cuda().contiguous()
    lr = 1e-4
    for it in range(2):
        ref = to.train_iteration(og, od, g_opt, d_opt, noisy, clean)
        L = tr.step(nz, cl)
        for k in ("d_loss", "g_loss", "content", "adv"):
            assert abs(L[k] - ref[k]) <= 5e-3 * abs(ref[k]) + 1e-5, (it, k, L[k], ref[k])
    g, d = DenoiseGenerator(), DenoiseDiscriminator()
    tr.export_state_dicts(g, d)
    gsd, dsd = g.state_dict(), d.state_dict()
    moved, agree = 0, 0
    for k, want in og.state_dict().items():
        got = gsd[k].cpu()
        assert float((got - want).abs().max()) <= 4.2 * lr, k       # two Adam steps move a weight by at most ~2 lr each
        moved += want.numel()
        agree += int(((got - g0[k]).sign() == (want - g0[k]).sign()).sum())
    assert agree / moved >= 0.97, agree / moved                     # same update direction element by element
    for k in ("model.3.running_mean", "model.3.running_var", "model.6.running_mean", "model.9.running_var"):
        np.testing.assert_allclose(dsd[k].cpu().numpy(), od.state_dict()[k].numpy(), rtol=3e-2, atol=2e-3)
    assert int(dsd["model.3.num_batches_tracked"]) == int(od.state_dict()["model.3.num_batches_tracked"]) == 6
    tr.close()


def test_errors_are_loud():
    g, d = DenoiseGenerator(), DenoiseDiscriminator()
    tr = DenoiseGANTrainStep(g, d, device="cuda:0")
    x = torch.zeros(1, 3, 32, 32, device="cuda:0")
    with pytest.raises(RuntimeError):
        tr.d_backward(x, x)                     # batch 1: train-mode BatchNorm needs more than one sample
    x = torch.zeros(2, 3, 30, 32, device="cuda:0")
    with pytest.raises(RuntimeError):
        tr.d_backward(x, x)                     # not a multiple of 4
    with pytest.raises(RuntimeError):
        tr.d_backward(x.cpu(), x.cpu())         # no CPU path
    with pytest.raises(RuntimeError):
        d(x)                                    # the host mirror has no PyTorch forward
    tr.close()


def test_many_tiles_per_cta_match_oracle():
    """8 x 128 x 128: 1024 pixel tiles, i.e. every persistent CTA walks many tiles (the weight-gradient kernel's
    stage ring wraps several times, split-K slices are uneven) - still against the oracle's autograd."""
    B, H, W = 8, 128, 128
    og, od, tr, noisy, clean = _make(5, B, H, W)
    ref = to.instrumented_gradients(og, od, noisy, clean)
    nz, cl = noisy.cuda().contiguous(), clean.cuda().contiguous()
    tr.d_backward(nz, cl)
    d_grads = tr.grads("discriminator")
    tr.g_backward(nz, cl)
    g_grads = tr.grads("generator")
    L = tr.losses()
    for k in ("d_loss", "g_loss"):
        assert abs(L[k] - ref[k]) <= 2e-3 * abs(ref[k]) + 1e-5, (k, L[k], ref[k])
    for k, want in ref["g_grads"].items():
        rel, cos = _rel_cos(g_grads[k].cpu().reshape(want.shape), want)
        assert cos >= 0.997 and rel <= 8e-2, ("generator", k, rel, cos)
    for k, want in ref["d_grads"].items():
        if k in ZERO_GRAD_KEYS:
            continue
        rel, cos = _rel_cos(d_grads[k].cpu().reshape(want.shape), want)
        assert cos >= 0.93, ("discriminator", k, rel, cos)
    tr.close()


def test_full_size_gradient_is_the_loss_derivative():
    """BASELINE config #5 size (batch 64 @ 128x128), where the CPU oracle is too slow: a size-independent property.
    Move the generator's weights by -eps * grad; the generator loss, re-evaluated by the same path, must drop by
    eps * |grad|^2 to first order (0.6 .. 1.1 of it: curvature only ever reduces the drop at this step size)."""
    B, H, W = 64, 128, 128
    og, od, tr, noisy, clean = _make(9, B, H, W)
    nz, cl = noisy.cuda().contiguous(), clean.cuda().contiguous()
    tr.d_backward(nz, cl)
    tr.g_backward(nz, cl)
    L0 = tr.losses()["g_loss"]
    g = tr.section("generator", 1).clone()
    gg = float((g.double() ** 2).sum())
    assert gg > 0 and np.isfinite(gg)
    drop = 0.02 * L0
    eps = drop / gg
    tr.section("generator", 0).sub_(eps * g)
    tr.lib.cid_train_sync_weights(tr.handle, None)
    tr.d_backward(nz, cl)
    tr.g_backward(nz, cl)
    L1 = tr.losses()["g_loss"]
    ratio = (L0 - L1) / drop
    assert 0.6 <= ratio <= 1.1, (L0, L1, ratio)
    # and the step is deterministic up to the atomics' summation order
    g2 = tr.section("generator", 1).clone()
    tr.d_backward(nz, cl)
    tr.g_backward(nz, cl)
    g3 = tr.section("generator", 1)
    assert float((g2 - g3).norm() / g2.norm()) < 1e-3
    tr.close()


def test_fused_allreduce_adam_two_replicas_one_gpu():
    """cid_train_set_peers: two replicas (here both on cuda:0, each launch on its own stream so that the in-kernel
    barriers can meet) average their gradients inside the optimiser kernel. Against torch.optim.Adam on the mean
    gradient, both replicas bit-identical to each other, three steps."""
    from cid_b200 import lib as cid_lib
    lib = cid_lib.load()
    g, d = DenoiseGenerator(), DenoiseDiscriminator()
    reps = [DenoiseGANTrainStep(g, d, device="cuda:0") for _ in range(2)]
    flags = [torch.zeros(lib.cid_train_peer_flag_ints(), dtype=torch.int32, device="cuda:0") for _ in range(2)]
    for r, tr in enumerate(reps):
        tr.bind_peers_raw(2, r, [t.state.data_ptr() for t in reps], [f.data_ptr() for f in flags], keep_alive=flags)
        assert tr.peer_fused
    streams = [torch.cuda.Stream(device="cuda:0") for _ in range(2)]
    gen = torch.Generator().manual_seed(11)
    for which, net in ((1, "discriminator"), (0, "generator")):
        w0 = reps[0].section(net, 0).clone()
        ref_w = torch.nn.Parameter(w0.double().cpu())
        opt = torch.optim.Adam([ref_w], lr=1e-4, betas=(0.9, 0.999), eps=1e-8)
        for step in range(3):
            grads = [torch.randn(w0.numel(), generator=gen) * (10.0 ** -step) for _ in range(2)]
            for tr, gr in zip(reps, grads):
                tr.section(net, 1).copy_(gr.cuda())
            torch.cuda.synchronize()
            for tr, st in zip(reps, streams):
                with torch.cuda.stream(st):
                    tr.adam(which)
            torch.cuda.synchronize()
            reps[0].losses()                      # raises if a barrier timed out
            ref_w.grad = ((grads[0].float() + grads[1].float()) * 0.5).double()
            opt.step()
            a, b = reps[0].section(net, 0), reps[1].section(net, 0)
            assert torch.equal(a, b), "replicas diverged"
            assert torch.equal(reps[0].section(net, 2), reps[1].section(net, 2))
            err = (a.cpu().double() - ref_w.detach()).abs().max().item()
            assert err < 2e-6, (net, step, err)
            # the gradient section keeps the rank-local gradients
            assert torch.equal(reps[1].section(net, 1).cpu(), grads[1])
    for tr in reps:
        tr.close()
