"""Kernel-level parity of the CW-L2 update kernels, the PGD random start and the remaining elementwise edge cases
(SURVEY.md rows a7, a8, a10): advb200_cw_init_f32 / advb200_cw_step_f32 / advb200_cw_finalize_f32 are driven through the C
ABI with gradients torch computes for the tiny linear classifier of tests/golden/tiny_linear.npz — the fixture the
reference's own cw_l2_attack outputs were generated on — and compared per step with the oracle's written-out Adam
(oracle/attacks_ref.adam_update, pinned to torch.optim.Adam through the reference) and at the end with the reference's
x_adv and success flags."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import attacks_ref, weights  # noqa: E402


class CwDriver:
    """The loop of advb200.attack.cw_l2_attack with the model's gradient supplied by the test."""

    def __init__(self, x, dev):
        from advb200 import _lib
        self._lib, self.lib = _lib, _lib.load()
        self.x = x.to(dev).contiguous()
        self.n, self.per = x.shape[0], x[0].numel()
        self.parts = int(self.lib.advb200_cw_parts(self.per))
        e = lambda: torch.empty_like(self.x)
        self.x0, self.w, self.m, self.v, self.x_adv, self.best_adv = e(), e(), e(), e(), e(), e()
        self.best_l2 = torch.empty(self.n, device=dev)
        self.best_success = torch.empty(self.n, dtype=torch.uint8, device=dev)
        self.l2_now = torch.empty(self.n, device=dev)
        self.pa = torch.empty(self.n * self.parts + (self.n + 3) // 4 + 1, device=dev)
        self.pb = torch.empty_like(self.pa)
        p = _lib.ptr
        _lib.check(self.lib.advb200_cw_init_f32(p(self.x), p(self.x0), p(self.w), p(self.m), p(self.v), p(self.x_adv),
                                                p(self.best_adv), p(self.best_l2), p(self.best_success), p(self.pa), self.n,
                                                self.per, _lib.stream_ptr()), "cw_init")

    def step(self, grad, success, t, lr):
        p = self._lib.ptr
        self._lib.check(self.lib.advb200_cw_step_f32(
            p(self.w), p(self.m), p(self.v), p(self.x0), p(grad), p(self.x_adv), p(self.best_adv), p(self.best_l2),
            p(self.best_success), p(success), p(self.pa), p(self.pb), p(self.l2_now), self.n, self.per,
            lr / (1.0 - 0.9 ** t), math.sqrt(1.0 - 0.999 ** t), 0.9, 0.999, 1e-8, self._lib.stream_ptr()), "cw_step")
        self.pa, self.pb = self.pb, self.pa

    def finalize(self):
        p = self._lib.ptr
        out = torch.empty_like(self.x)
        self._lib.check(self.lib.advb200_cw_finalize_f32(p(self.x_adv), p(self.best_adv), p(self.best_success), p(out), self.n,
                                                         self.per, self._lib.stream_ptr()), "cw_finalize")
        return out


def margin(logits, y, kappa, targeted):
    onehot = torch.nn.functional.one_hot(y, logits.shape[1]).to(logits.dtype)
    real = (logits * onehot).sum(1)
    other = (logits - 1e4 * onehot).amax(1)
    f = torch.clamp(other - real + kappa, min=0) if targeted else torch.clamp(real - other + kappa, min=0)
    succ = (logits.argmax(1) == y) if targeted else (logits.argmax(1) != y)
    return f, succ


@pytest.mark.parametrize("targeted", [False, True])
def test_cw_kernels_step_by_step_and_final_vs_reference(golden_dir, targeted):
    dev = torch.device("cuda:0")
    g = np.load(golden_dir / "tiny_linear.npz")
    mean, std = weights.mean_std()
    x, wt = torch.from_numpy(g["x"]), torch.from_numpy(g["wt"]).to(dev)
    y = torch.from_numpy(g["yt"] if targeted else g["y"]).to(dev)
    c, kappa, steps, lr = 2.0, (0.0 if targeted else 0.1), 25, 0.05          # the fixture's settings (oracle/gen_golden.py)
    d = CwDriver(x, dev)
    # a8: tanh-space initialisation (attack.py:95-99): w0 = atanh((x0(1-2e-6)+1e-6)*2-1), x(w0) ~ x0
    x0 = x.clamp(0, 1)
    t0 = (x0 * (1.0 - 2e-6) + 1e-6) * 2.0 - 1.0
    w0 = 0.5 * (torch.log1p(t0) - torch.log1p(-t0))
    np.testing.assert_allclose(d.w.cpu().numpy(), w0.numpy(), rtol=2e-6, atol=2e-6)
    assert torch.equal(d.x0.cpu(), x0) and torch.equal(d.best_adv.cpu(), x0)
    np.testing.assert_allclose(d.x_adv.cpu().numpy(), (0.5 * (torch.tanh(w0) + 1)).numpy(), atol=2e-7)
    assert bool(torch.isinf(d.best_l2).all()) and int(d.best_success.sum()) == 0 and float(d.m.abs().max()) == 0
    worst = 0.0
    for t in range(1, steps + 1):
        w_prev, m_prev, v_prev = d.w.clone(), d.m.clone(), d.v.clone()
        best_prev, bs_prev = d.best_l2.clone(), d.best_success.clone()
        # gradient of c*f w.r.t. the image at the kernel's own iterate (torch autograd on the tiny linear model)
        xi = d.x_adv.clone().requires_grad_(True)
        logits = attacks_ref.normalize(xi, mean.to(dev), std.to(dev)).flatten(1) @ wt
        f, succ = margin(logits, y, kappa, targeted)
        (gx,) = torch.autograd.grad((c * f).sum(), xi)
        l2_cur = (d.x_adv - d.x0).flatten(1).pow(2).sum(1)
        d.step(gx.contiguous(), succ.to(torch.uint8), t, lr)
        # a10: the oracle's written-out Adam on the gradient w.r.t. w of  l2 + c*f  at the same (w, m, v)
        wv = w_prev.clone().requires_grad_(True)
        xa = 0.5 * (torch.tanh(wv) + 1.0)
        lo = attacks_ref.normalize(xa, mean.to(dev), std.to(dev)).flatten(1) @ wt
        fo, _ = margin(lo, y, kappa, targeted)
        (gw,) = torch.autograd.grad(((xa - d.x0).flatten(1).pow(2).sum(1) + c * fo).sum(), wv)
        w_ref, m_ref, v_ref = attacks_ref.adam_update(w_prev, m_prev, v_prev, gw, t, lr)
        for got, want in ((d.w, w_ref), (d.m, m_ref), (d.v, v_ref)):
            err = float((got - want).abs().max() / want.abs().max().clamp_min(1e-30))
            worst = max(worst, err)
            assert err < 2e-6, (t, err)
        np.testing.assert_allclose(d.x_adv.cpu().numpy(), (0.5 * (torch.tanh(d.w) + 1)).cpu().numpy(), atol=2e-7)
        # a9/a11: per-image L2 of the iterate the gradient was taken at, improved = success & (l2 < best_l2)
        np.testing.assert_allclose(d.l2_now.cpu().numpy(), l2_cur.cpu().numpy(), rtol=2e-6)
        improved = succ & (d.l2_now < best_prev)
        assert torch.equal(d.best_success.bool(), bs_prev.bool() | improved)
        assert torch.equal(d.best_l2, torch.where(improved, d.l2_now, best_prev))
    final = d.finalize().cpu()
    want_x, want_s = (g["cwt_x"], g["cwt_s"]) if targeted else (g["cwu_x"], g["cwu_s"])
    assert np.array_equal(d.best_success.bool().cpu().numpy(), want_s)
    np.testing.assert_allclose(final.numpy(), want_x, rtol=0, atol=5e-6)       # the REFERENCE's cw_l2_attack output
    print(f"CW kernels ({'targeted' if targeted else 'untargeted'}): worst per-step Adam state error {worst:.1e}; "
          f"final x_adv max |err| vs the reference {float(np.abs(final.numpy() - want_x).max()):.1e}")


def test_cw_step_full_size_vector_path_vs_oracle():
    """224x224 images (49 blocks per image, 128-bit path), three steps on random gradients, N = 3; also 4-byte-aligned
    views (scalar path) must give the same bits."""
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(31)
    x = torch.rand(3, 3, 224, 224, generator=gen)
    d = CwDriver(x, dev)
    succ = torch.tensor([1, 0, 1], dtype=torch.uint8, device=dev)
    best_l2 = torch.full((3,), float("inf"), device=dev)
    best_adv = d.x0.clone()
    for t in range(1, 4):
        gx = (torch.randn(3, 3, 224, 224, generator=gen) * 0.05).to(dev)
        w_prev, m_prev, v_prev, xa_prev = d.w.clone(), d.m.clone(), d.v.clone(), d.x_adv.clone()
        d.step(gx, succ, t, 0.01)
        tw = torch.tanh(w_prev)
        gw = (2.0 * (0.5 * (tw + 1.0) - d.x0) + gx) * 0.5 * (1.0 - tw * tw)
        w_ref, m_ref, v_ref = attacks_ref.adam_update(w_prev, m_prev, v_prev, gw, t, 0.01)
        assert float((d.w - w_ref).abs().max()) < 2e-6 and float((d.m - m_ref).abs().max()) < 1e-7
        assert float(((d.v - v_ref).abs() / v_ref.abs().clamp_min(1e-12)).max()) < 1e-5
        np.testing.assert_allclose(d.l2_now.cpu().numpy(), (xa_prev - d.x0).flatten(1).double().pow(2).sum(1).cpu().numpy(),
                                   rtol=2e-6, atol=1e-9)
        improved = succ.bool() & (d.l2_now < best_l2)                      # attack.py:133-137
        best_l2 = torch.where(improved, d.l2_now, best_l2)
        best_adv = torch.where(improved.view(-1, 1, 1, 1), xa_prev, best_adv)
        assert torch.equal(d.best_adv, best_adv) and torch.equal(d.best_l2, best_l2)     # only improved images copy
    # misaligned state arrays: same results through the scalar path
    d2 = CwDriver(x, dev)
    from advb200 import _lib
    lib = _lib.load()
    per, n = d2.per, d2.n
    shifted = lambda t: torch.cat([torch.zeros(1, device=dev), t.flatten()])[1:].view_as(t)
    w2, m2, v2, xa2, ba2 = (shifted(t) for t in (d2.w, d2.m, d2.v, d2.x_adv, d2.best_adv))
    gx = (torch.randn(3, 3, 224, 224, generator=gen) * 0.05).to(dev)
    wa, ma, va = d2.w.clone(), d2.m.clone(), d2.v.clone()
    xaa, baa = d2.x_adv.clone(), d2.best_adv.clone()
    p = _lib.ptr
    for (w_, m_, v_, xa_, ba_, pa_, pb_) in ((w2, m2, v2, xa2, ba2, d2.pa.clone(), d2.pb.clone()),
                                            (wa, ma, va, xaa, baa, d2.pa.clone(), d2.pb.clone())):
        bl2, bs = d2.best_l2.clone(), d2.best_success.clone()
        _lib.check(lib.advb200_cw_step_f32(p(w_), p(m_), p(v_), p(d2.x0), p(gx), p(xa_), p(ba_), p(bl2), p(bs), p(succ),
                                           p(pa_), p(pb_), None, n, per, 0.1, 0.0316, 0.9, 0.999, 1e-8, _lib.stream_ptr()),
                   "cw_step")
    assert torch.equal(w2, wa) and torch.equal(m2, ma) and torch.equal(v2, va) and torch.equal(xa2, xaa)


@pytest.mark.parametrize("shape", [(4, 3, 224, 224), (2, 3, 7, 9), (1, 1, 1, 1)])
def test_random_start_bit_exact(shape):
    """a7: x_adv = clamp(x + U(-eps,eps), 0, 1) (attack.py:50-51) on the noise tensor torch drew."""
    from advb200 import _lib
    dev = torch.device("cuda:0")
    eps = 8 / 255
    x = torch.rand(shape, generator=torch.Generator().manual_seed(2))
    noise = weights.pgd_noise(shape, eps, 9)
    want = (x + noise).clamp(0.0, 1.0)
    out = torch.empty(shape, device=dev)
    xd, nd = x.to(dev), noise.to(dev)                   # (kept alive: the C ABI takes raw pointers)
    _lib.check(_lib.load().advb200_random_start_f32(_lib.ptr(out), _lib.ptr(xd), _lib.ptr(nd), out.numel(),
                                                    _lib.stream_ptr()), "random_start")
    assert torch.equal(out.cpu(), want)
    assert float(out.min()) >= 0 and float(out.max()) <= 1


def test_pgd_with_injected_noise_equals_seeded_random_start():
    """pgd_linf_attack(noise=...) (the hook the config-5 parity tests use to share the reference's start noise) is the
    same computation as the default random start fed the same tensor; the reference signature is unchanged without it."""
    import advb200
    dev = torch.device("cuda:0")
    nm = advb200.NativeResNet50(weights.seeded_resnet50(0).to(dev), precision="bf16")
    mean, std = weights.mean_std()
    x = weights.synthetic_images(2).to(dev)
    y = torch.tensor([713, 713], device=dev)
    eps = 8 / 255
    torch.manual_seed(3)
    a = advb200.pgd_linf_attack(nm, x, y, eps=eps, alpha=2 / 255, steps=2, mean=mean, std=std)
    torch.manual_seed(3)
    noise = torch.empty_like(x).uniform_(-eps, eps)
    b = advb200.pgd_linf_attack(nm, x, y, eps=eps, alpha=2 / 255, steps=2, mean=mean, std=std, noise=noise)
    assert torch.equal(a, b)
    with pytest.raises(ValueError):
        advb200.pgd_linf_attack(nm, x, y, eps=eps, alpha=2 / 255, steps=1, mean=mean, std=std, noise=noise[:1])


def test_median_and_smoothing_with_a_misaligned_output_plane():
    """a 16-byte-aligned input with a 4-byte-aligned output view must not take the 128-bit store."""
    from advb200 import _lib
    from oracle import defenses_ref
    dev = torch.device("cuda:0")
    lib = _lib.load()
    x = torch.rand(2, 3, 8, 8, generator=torch.Generator().manual_seed(4))
    xd = x.to(dev)
    obuf = torch.zeros(1 + x.numel(), device=dev)
    out = obuf[1:].view(2, 3, 8, 8)
    assert xd.data_ptr() % 16 == 0 and out.data_ptr() % 16 == 4
    _lib.check(lib.advb200_median3x3_f32(_lib.ptr(xd), _lib.ptr(out), 6, 8, 8, _lib.stream_ptr()), "median")
    assert np.array_equal(out.cpu().numpy(), defenses_ref.median3x3(x.numpy()))
    _lib.check(lib.advb200_smooth3x3_f32(_lib.ptr(xd), _lib.ptr(out), 6, 8, 8, _lib.stream_ptr()), "smooth")
    assert np.array_equal(out.cpu().numpy(), defenses_ref.smooth3x3(x.numpy()))
    _lib.check(lib.advb200_defend_input_f32(_lib.ptr(xd), _lib.ptr(out), 6, 8, 8, 16, _lib.stream_ptr()), "defend")
    assert np.array_equal(out.cpu().numpy(), defenses_ref.defend_input(x.numpy()))
    torch.cuda.synchronize()


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_native_graph_loop_equals_python_loop(precision):
    """advb200_pgd_steps (captured CUDA graph replayed per step; plain native loop) == the Python loop over
    advb200_input_grad + advb200_linf_step_f32, bit for bit, for a full chunk, a tail chunk and a repeated call."""
    import advb200
    from advb200 import attack as A
    dev = torch.device("cuda:0")
    nm = advb200.NativeResNet50(weights.conditioned_resnet50(0).to(dev), precision=precision)
    nm.max_chunk = 4
    mean, std = weights.mean_std()
    x = weights.mixed_images(6, seed=5).to(dev)                    # chunks of 4 + 2
    y = nm.forward(x, mean, std, want_preds=True)["preds"]
    eps = 4 / 255
    noise = weights.pgd_noise(x.shape, eps, 3).to(dev)
    kw = dict(eps=eps, alpha=2 / 255, steps=3, mean=mean, std=std, noise=noise)
    outs = {}
    for mode, (native, graph) in {"python": (False, False), "native": (True, False), "graph": (True, True)}.items():
        A._NATIVE_LOOP, A._USE_GRAPH = native, graph
        try:
            outs[mode] = advb200.pgd_linf_attack(nm, x, y, **kw)
            if mode == "graph":
                again = advb200.pgd_linf_attack(nm, x, y, **kw)      # second call replays the cached graphs
                assert torch.equal(again, outs[mode])
        finally:
            A._NATIVE_LOOP, A._USE_GRAPH = True, True
    assert torch.equal(outs["native"], outs["python"]) and torch.equal(outs["graph"], outs["python"])
    assert not torch.equal(outs["python"], x)
    assert bool(((outs["graph"] <= x + eps) & (outs["graph"] >= x - eps)).all())
