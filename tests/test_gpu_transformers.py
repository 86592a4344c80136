"""Forward-only transformer transfer targets (blackbox_transfer.py:36-54, 233-251): the native
ViT-B/16 and Swin-T against torchvision's own modules in fp32 (TF32 off) on seeded random-init weights and the fixture
images — logits to 1e-2 relative L2 (the bf16 bar of the north star), identical predictions wherever the reference's top-2
margin exceeds the logit error, results independent of the batch an image rides in, and the reference's error behaviour."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import weights  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return torch.device("cuda:0")


def seeded(kind, dev):
    import torchvision
    if kind == "vit":
        torch.manual_seed(21)
        m = torchvision.models.vit_b_16(weights=None).eval()
        with torch.no_grad():                      # torchvision zero-initialises the ViT head and the class token
            m.heads.head.weight.normal_(std=0.02)
            m.heads.head.bias.normal_(std=0.02)
            m.class_token.normal_(std=0.02)
    else:
        torch.manual_seed(22)
        m = torchvision.models.swin_t(weights=None).eval()
        with torch.no_grad():                      # relative-position tables and LayerNorm parameters off their init values
            for name, p in m.named_parameters():
                if "relative_position_bias_table" in name:
                    p.normal_(std=0.5)
                elif "norm" in name and name.endswith("weight"):
                    p.uniform_(0.5, 1.5)
                elif "norm" in name and name.endswith("bias"):
                    p.uniform_(-0.2, 0.2)
    if kind == "vit":
        with torch.no_grad():
            for name, p in m.named_parameters():
                if "ln" in name and name.endswith("weight"):
                    p.uniform_(0.5, 1.5)
                elif "ln" in name and name.endswith("bias"):
                    p.uniform_(-0.2, 0.2)
    return m.to(dev)


@pytest.mark.parametrize("kind", ["vit", "swin"])
def test_logits_vs_torchvision_fp32(kind, dev):
    from advb200.transformers import NativeSwinT, NativeViTB16
    from oracle import attacks_ref
    model = seeded(kind, dev)
    nm = (NativeViTB16 if kind == "vit" else NativeSwinT)(model)
    mean, std = weights.mean_std()
    x = weights.mixed_images(6, seed=5).to(dev)
    with torch.no_grad():
        ref = model(attacks_ref.normalize(x, mean.to(dev), std.to(dev)))
    got = nm.forward(x, mean, std)
    rel = float((got.double() - ref.double()).norm() / ref.double().norm())
    top2 = ref.topk(2, 1).values
    margin = top2[:, 0] - top2[:, 1]
    err = (got - ref).abs().max(1).values
    clear = margin > 2 * err
    print(f"[{kind}] logits rel-L2 vs torchvision fp32 {rel:.2e}; max |err| {float(err.max()):.2e} (logit range {float(ref.min()):.2f}..{float(ref.max()):.2f}); "
          f"predictions equal {int((got.argmax(1) == ref.argmax(1)).sum())}/6 ({int(clear.sum())} with a top-2 margin above twice the error)")
    assert rel < 1e-2
    assert torch.equal(got.argmax(1)[clear], ref.argmax(1)[clear])
    assert not bool(torch.isnan(got).any())
    # an image's logits do not depend on the batch it rides in; the call shape of blackbox_transfer.py:241 works too
    one = torch.cat([nm.forward(x[i:i + 1], mean, std) for i in (0, 3)])
    assert torch.equal(one, got[[0, 3]])
    again = nm(attacks_ref.normalize(x[:2], mean.to(dev), std.to(dev)))
    assert float((again - got[:2]).abs().max()) < 2e-2 * float(ref.abs().max())
    # the tensor-core attention kernel against the CUDA-core one (same arithmetic up to the bf16 rounding of P)
    from advb200 import _lib
    _lib.load().advb200_test_attention_simt(1)
    try:
        simt = nm.forward(x, mean, std)
    finally:
        _lib.load().advb200_test_attention_simt(0)
    rel_s = float((simt.double() - ref.double()).norm() / ref.double().norm())
    rel_k = float((simt.double() - got.double()).norm() / got.double().norm())
    print(f"[{kind}] CUDA-core attention: rel-L2 vs torchvision {rel_s:.2e}; tensor-core vs CUDA-core kernels {rel_k:.2e}")
    # (two bf16 paths that are each ~8e-3 from the fp32 reference are up to twice that from each other)
    assert rel_s < 1e-2 and rel_k < 1.5e-2 and not torch.equal(simt, got)


def test_error_behaviour(dev):
    import torchvision
    from advb200.transformers import NativeSwinT, NativeViTB16, transfer_target
    mean, std = weights.mean_std()
    vit = seeded("vit", dev)
    nm = NativeViTB16(vit)
    with pytest.raises(ValueError):
        nm.forward(torch.zeros(1, 3, 64, 64, device=dev), mean, std)
    with pytest.raises(ValueError):
        nm.forward(torch.zeros(1, 3, 224, 224), mean, std)
    with pytest.raises(TypeError):
        NativeViTB16(torchvision.models.resnet18(weights=None).eval().to(dev))
    with pytest.raises(TypeError):
        NativeSwinT(vit)
    with pytest.raises(ValueError):
        NativeViTB16(torchvision.models.vit_b_16(weights=None).to(dev).train())
    assert isinstance(transfer_target(vit), NativeViTB16)
    assert nm.forward(torch.zeros(0, 3, 224, 224, device=dev), mean, std).shape == (0, 1000)
