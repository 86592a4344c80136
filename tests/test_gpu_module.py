"""the `nn.Module` face of the native model, for callers that build the graph
themselves — `logits = model(normalize_batch(x, mean, std)); g = torch.autograd.grad(loss, x)` as in
visualized_attacks.py:337-358 and defense_experiments.py:458-479.  The checker is torch autograd through the
torch model (fp32, TF32 off) and the native input_grad path that the other parity tests pin to the oracle."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import attacks_ref, weights  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def torch_model(dev):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return weights.seeded_resnet50(0).to(dev)


@pytest.fixture(scope="module")
def modules(torch_model):
    import advb200
    return {p: advb200.NativeResNet50Module(torch_model, precision=p) for p in ("fp32", "bf16")}


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm())


def caller_side_gradient(model, x, y, mean, std):
    """What the reference's own trajectory code does (visualized_attacks.py:344-349)."""
    import advb200
    xi = x.clone().detach().requires_grad_(True)
    logits = model(advb200.normalize_batch(xi, mean, std))
    loss = F.cross_entropy(logits, y)
    (g,) = torch.autograd.grad(loss, xi, only_inputs=True)
    return logits.detach(), g


def test_module_fp32_gradient_matches_torch_autograd(modules, torch_model, dev):
    mean, std = (t.to(dev) for t in weights.mean_std())
    x = weights.synthetic_images(3).to(dev)
    y = torch_model(attacks_ref.normalize(x, mean, std)).argmax(1)
    lref, gref = caller_side_gradient(torch_model, x, y, mean, std)
    lgot, ggot = caller_side_gradient(modules["fp32"], x, y, mean, std)
    assert rel(lgot, lref) < 1e-4
    # same bound as the input_grad test: the reference's own fp32 noise floor on this random-init model
    assert rel(ggot, gref) < 5e-2
    big = gref.abs() > 1e-6 * gref.abs().max()
    assert float((torch.sign(ggot[big]) == torch.sign(gref[big])).float().mean()) > 0.99


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_module_gradient_equals_native_input_grad(modules, dev, precision):
    """The autograd route (forward, torch CE on the logits, advb200_backward) and the one-call route
    (advb200_input_grad with its own CE kernel) run the same conv chain: same logits bit for bit, gradients
    equal up to the fp32 rounding of dlogits."""
    mean, std = (t.to(dev) for t in weights.mean_std())
    mod = modules[precision]
    x = weights.synthetic_images(4).to(dev)
    y = mod.native.forward(x, mean, std, want_preds=True)["preds"]
    logits, g = caller_side_gradient(mod, x, y, mean, std)
    out = mod.native.input_grad(x, mean, std, y, loss="ce", want_logits=True)
    assert torch.equal(logits, out["logits"])
    assert rel(g, out["grad"]) < (2e-2 if precision == "bf16" else 1e-4)
    big = out["grad"].abs() > 1e-3 * out["grad"].abs().max()
    assert float((torch.sign(g[big]) == torch.sign(out["grad"][big])).float().mean()) > 0.995


def test_module_recomputes_when_its_forward_state_was_overwritten(modules, dev):
    mean, std = (t.to(dev) for t in weights.mean_std())
    mod = modules["bf16"]
    x = weights.synthetic_images(4).to(dev)
    y = torch.tensor([1, 2, 3, 4], device=dev)
    _, g_fresh = caller_side_gradient(mod, x, y, mean, std)
    # two graphs alive at once: the second forward overwrites the first one's state
    import advb200
    xa = x.clone().requires_grad_(True)
    xb = x.flip(0).clone().requires_grad_(True)
    la = F.cross_entropy(mod(advb200.normalize_batch(xa, mean, std)), y)
    lb = F.cross_entropy(mod(advb200.normalize_batch(xb, mean, std)), y.flip(0))
    (ga,) = torch.autograd.grad(la, xa)
    (gb,) = torch.autograd.grad(lb, xb)
    assert torch.equal(ga, g_fresh)
    assert torch.equal(gb.flip(0), g_fresh)
    # chunked batches (n > max_chunk) take the recompute route chunk by chunk
    old = mod.native.max_chunk
    try:
        mod.native.max_chunk = 3
        _, g_chunked = caller_side_gradient(mod, x, y, mean, std)
    finally:
        mod.native.max_chunk = old
    # CE 'mean' divides by the caller's N in torch, so chunking does not change the scale
    assert rel(g_chunked, g_fresh) < 2e-2


def test_module_inference_and_errors(modules, dev):
    mean, std = (t.to(dev) for t in weights.mean_std())
    mod = modules["bf16"]
    x = weights.synthetic_images(2).to(dev)
    with torch.no_grad():
        a = mod(attacks_ref.normalize(x, mean, std))
    b = mod.native.forward(x, mean, std)["logits"]
    assert a.shape == (2, 1000) and a.dtype == torch.float32
    assert rel(a, b) < 1e-2           # normalisation outside vs fused: same formula, same rounding
    assert mod.eval() is mod
    with pytest.raises(RuntimeError):
        mod.train()
    with pytest.raises(ValueError):
        mod(torch.zeros(3, 224, 224, device=dev))
    with pytest.raises(ValueError):
        mod(torch.zeros(1, 3, 224, 224))          # CPU tensor: no fallback
    # backward without forward state is refused by the C ABI, not silently wrong
    nm = mod.native
    nm.forward(x, mean, std, stop_at_layer3=True, want_logits=False, want_score=True)
    with pytest.raises((RuntimeError, ValueError)):
        nm.backward(torch.zeros(2, 1000, device=dev), std, 2)


def test_reference_style_pgd_loop_through_the_module(modules, dev):
    """visualized_attacks.py:337-358 re-implements PGD around `model(normalize_batch(x))`; run through the module it
    must land where advb200.pgd_linf_attack lands (same kernels underneath, CE gradient computed by torch)."""
    import advb200
    mean, std = (t.to(dev) for t in weights.mean_std())
    mod = modules["bf16"]
    x = weights.synthetic_images(4).to(dev)
    y = mod.native.forward(x, mean, std, want_preds=True)["preds"]
    eps, alpha, steps = 8 / 255, 2 / 255, 3
    x_adv = x.clone()
    same = []
    for k in range(steps):
        _, g = caller_side_gradient(mod, x_adv, y, mean, std)
        x_adv = x_adv + alpha * g.sign()
        x_adv = torch.max(torch.min(x_adv, x + eps), x - eps)
        x_adv = torch.clamp(x_adv, 0.0, 1.0)
        want = advb200.pgd_linf_attack(mod.native, x, y, eps=eps, alpha=alpha, steps=k + 1, mean=mean, std=std,
                                       random_start=False)
        same.append(float((x_adv == want).float().mean()))
    print("pixels identical to pgd_linf_attack after 1..3 caller-side steps:", [round(v, 5) for v in same])
    # step 1 differs only where the last-bit difference between torch's CE gradient and the native CE kernel flips
    # the sign of a near-zero input gradient; the model is chaotic in its input, so the difference compounds later
    # (the oracle-vs-native PGD test sees the same growth)
    assert same[0] > 0.97
    assert float((x_adv - x).abs().max()) <= eps + 1e-6
    assert float(x_adv.min()) >= 0.0 and float(x_adv.max()) <= 1.0
    fooled_loop = mod.native.forward(x_adv, mean, std, want_preds=True)["preds"] != y
    fooled_api = mod.native.forward(want, mean, std, want_preds=True)["preds"] != y
    assert abs(int(fooled_loop.sum()) - int(fooled_api.sum())) <= 1
