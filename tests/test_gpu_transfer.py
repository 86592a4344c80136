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
    # the class the native model picks is (within the bf16 error) the checker's best class
    picked = want.gather(1, got.argmax(1, keepdim=True)).squeeze(1)
    spread = want.max(1).values - want.min(1).values
    assert bool((want.max(1).values - picked <= 0.05 * spread).all())


def test_vgg19_chunking_errors_and_transfer_rate(vgg):
    from advb200.transfer import NativeVGG19, transfer_success_rate
    nm = NativeVGG19(vgg)
    mean, std = (t.cuda() for t in weights.mean_std())
    x = weights.synthetic_images(5).cuda()
    full = nm.forward(x, mean, std)
    nm.max_chunk = 2
    assert torch.equal(nm.forward(x, mean, std), full)
    assert transfer_success_rate(nm, x, x, mean, std) == 0.0
    flipped = transfer_success_rate(nm, x, 1.0 - x, mean, std)
    want = float((full.argmax(1) != nm.forward(1.0 - x, mean, std).argmax(1)).float().mean())
    assert flipped == want
    with pytest.raises(ValueError):
        nm.forward(x.cpu(), mean, std)
    with pytest.raises(ValueError):
        nm.forward(x[0], mean, std)
    with pytest.raises(TypeError):
        NativeVGG19(torchvision.models.resnet18(weights=None).eval().cuda())
    with pytest.raises(ValueError):
        NativeVGG19(torchvision.models.vgg19(weights=None).cuda().train())


def test_blackbox_transfer_cli_table(capsys):
    """cli/blackbox_transfer.py end to end on synthetic images: the printed table has the reference's format
    (blackbox_transfer.py:316-340) and its VGG-19 column equals the API computation."""
    import importlib.util
    from pathlib import Path

    import advb200
    from advb200.transfer import NativeVGG19, transfer_counts
    root = Path(__file__).resolve().parent.parent / "image-recognition-adversarial-example-attack_b200"
    spec = importlib.util.spec_from_file_location("advb200_cli_blackbox", root / "cli" / "blackbox_transfer.py")
    cli = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(cli)
    cli.main(["--synthetic", "6", "--weights", "random", "--seed", "0", "--attacks", "fgsm", "pgd", "--steps", "3",
              "--batch", "4"])
    lines = capsys.readouterr().out.strip().splitlines()
    assert lines[0] == "Attack/Model\tVGG19\tViT\tSwin"
    assert [l.split("\t")[0] for l in lines[1:]] == ["FGSM", "PGD", "CW"]
    assert lines[3] == "CW\t0.0%\t0.0%\t0.0%"                   # not requested: 0 of 0, as in the reference
    for l in lines[1:3]:
        assert all(c.endswith("%") for c in l.split("\t")[1:])     # all three target columns are filled

    dev = torch.device("cuda")
    src, vgg_m = cli.load_models("random", 0, dev)
    source, target = advb200.native_model(src), NativeVGG19(vgg_m)
    mean = torch.tensor([0.485, 0.456, 0.406], device=dev)
    std = torch.tensor([0.229, 0.224, 0.225], device=dev)
    x = torch.rand(6, 3, 224, 224, generator=torch.Generator().manual_seed(1234)).to(dev)
    succ = 0
    for lo in (0, 4):
        xb = x[lo:lo + 4]
        y = source.forward(xb, mean, std, want_logits=False, want_preds=True)["preds"]
        succ += transfer_counts(target, xb, advb200.fgsm_attack(source, xb, y, eps=8 / 255, mean=mean, std=std), mean, std)[0]
    assert lines[1].split("\t")[1] == f"{100.0 * succ / 6:.1f}%"
