"""SURVEY section 8(f) rank 4 (VGG-19 part): the forward-only transfer target of blackbox_transfer.py on the native conv
kernel.  Checker: torchvision's VGG-19 (seeded random init, fp32, TF32 off) on the same inputs; tolerance = the bf16
bar of the north star (1e-2 relative on logits would be per element; the bound asserted is relative L2)."""
import pytest
import torch
import torchvision

pytestmark = pytest.mark.gpu

from oracle import attacks_ref, weights  # noqa: E402


@pytest.fixture(scope="module")
def vgg():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)
    m = torchvision.models.vgg19(weights=None).eval().cuda()
    with torch.no_grad():                      # non-zero biases so the bias path is exercised
        for p in m.parameters():
            if p.ndim == 1:
                p.normal_(0, 0.05)
    return m


def test_vgg19_logits_match_torchvision(vgg):
    from advb200.transfer import NativeVGG19
    nm = NativeVGG19(vgg)
    mean, std = (t.cuda() for t in weights.mean_std())
    x = weights.synthetic_images(6).cuda()
    with torch.no_grad():
        want = vgg(attacks_ref.normalize(x, mean, std))
    got = nm.forward(x, mean, std)
    rel = float((got - want).norm() / want.norm())
    print(f"VGG-19 logits rel-L2 vs torchvision fp32: {rel:.3e}")
    assert got.shape == (6, 1000) and rel < 2e-2
    # same through the model(normalize_batch(x)) call shape, and per image (tile contents do not depend on the batch)
    assert torch.equal(nm(attacks_ref.normalize(x, mean, std)), got)
    one = nm.forward(x[4:5], mean, std)
    assert torch.equal(one, got[4:5])
    top_w, top_g = want.topk(5, 1).indices, got.argmax(1)
    assert all(int(top_g[i]) in top_w[i].tolist() for i in range(6))


def test_vgg19_chunking_errors_and_transfer_rate(vgg):
    from advb200.transfer import NativeVGG19, transfer_success_rate
    nm = NativeVGG19(vgg)
    mean, std = (t.cuda() for t in weights.mean_std())
    x = weights.synthetic_images(5).cuda()
    full = nm.forward(x, mean, std)
    nm.max_chunk = 2
    assert torch.equal(nm.forward(x, mean, std), full)
    y = full.argmax(1)
    assert transfer_success_rate(nm, x, y, mean, std) == 0.0
    assert transfer_success_rate(nm, x, (y + 1) % 1000, mean, std) == 1.0
    with pytest.raises(ValueError):
        nm.forward(x.cpu(), mean, std)
    with pytest.raises(ValueError):
        nm.forward(x[0], mean, std)
    with pytest.raises(TypeError):
        NativeVGG19(torchvision.models.resnet18(weights=None).eval().cuda())
    with pytest.raises(ValueError):
        NativeVGG19(torchvision.models.vgg19(weights=None).cuda().train())
