"""`load_image`'s Resize(256) -> CenterCrop(224) -> ToTensor and `--save_adv`'s ToPILImage
on the device.  Byte work: bit-exact against the golden vectors the reference's own transform produced on PIL
(tests/golden/preprocess_cases.npz) and against the CPU oracle (oracle/preprocess_ref.py) on further sizes."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import preprocess_ref  # noqa: E402
from oracle.gen_golden_preprocess import SHAPES, make_image  # noqa: E402


@pytest.fixture(scope="module")
def pp():
    from advb200 import preprocess
    return preprocess


@pytest.mark.parametrize("shape", SHAPES + ["example"])
def test_preprocess_bit_exact_vs_reference_golden(pp, golden_dir, shape):
    g = np.load(golden_dir / "preprocess_cases.npz")
    if shape == "example":
        img, want = g["example_crop_in"], g["example_crop_out"]
    else:
        img, want = make_image(*shape), g["out_%dx%d" % shape]
    got = pp.get_transform("cuda:0")(img)
    assert got.shape == (3, 224, 224) and got.dtype == torch.float32 and got.is_cuda
    assert np.array_equal(got.cpu().numpy(), want.astype(np.float32) / np.float32(255))


@pytest.mark.parametrize("shape", [(225, 3000), (4000, 230), (1080, 1920), (224, 225), (300, 300), (999, 1001)])
def test_preprocess_bit_exact_vs_oracle(pp, shape):
    img = make_image(*shape)
    want = preprocess_ref.load_image_tensor(img)[0]
    got = pp.get_transform("cuda:0")(torch.from_numpy(img)).cpu().numpy()
    assert np.array_equal(got, want)


def test_load_image_and_to_pil_round_trip(pp, tmp_path, golden_dir):
    from PIL import Image
    g = np.load(golden_dir / "preprocess_cases.npz")
    path = tmp_path / "example.png"                                  # PNG: lossless, so the decoded pixels are known
    Image.fromarray(g["example_crop_in"]).save(path)
    x = pp.load_image(path, torch.device("cuda:0"))
    assert x.shape == (1, 3, 224, 224)
    assert np.array_equal(x[0].cpu().numpy(), g["example_crop_out"].astype(np.float32) / np.float32(255))
    # --save_adv: ToPILImage truncates x*255; k/255 * 255 can land just below k, so compare with the oracle's rule
    img = pp.to_pil_image(x[0])
    assert np.array_equal(np.asarray(img), preprocess_ref.to_pil_u8(x[0].cpu().numpy()))
    y = torch.rand(3, 37, 53, generator=torch.Generator().manual_seed(1)).cuda()
    assert np.array_equal(np.asarray(pp.to_pil_image(y)), preprocess_ref.to_pil_u8(y.cpu().numpy()))


def test_preprocess_errors(pp):
    tf = pp.get_transform("cuda:0")
    with pytest.raises(ValueError):
        tf(np.zeros((100, 100), np.uint8))
    with pytest.raises(ValueError):
        tf(np.zeros((100, 100, 3), np.float32))
    with pytest.raises(ValueError):
        pp.NativeTransform("cuda:0", size=256, crop=300)(np.zeros((100, 1000, 3), np.uint8))   # resized 256 x 2560 < crop
    with pytest.raises(RuntimeError):
        pp.NativeTransform("cpu")
    with pytest.raises(ValueError):
        pp.to_pil_image(torch.zeros(3, 4, 4))


def test_cli_resnet_end_to_end_with_native_preprocessing(pp, tmp_path, golden_dir, capsys):
    """cli/ResNet.py on an image FILE: PIL decode -> device Resize/CenterCrop/ToTensor -> native FGSM -> --save_adv.
    Its printed top-5 lines and the saved picture equal what the API gives on the same preprocessed tensor."""
    import importlib.util
    import sys
    from pathlib import Path

    from PIL import Image

    import advb200
    from oracle import weights
    root = Path(__file__).resolve().parent.parent / "image-recognition-adversarial-example-attack_b200"
    spec = importlib.util.spec_from_file_location("advb200_cli_resnet", root / "cli" / "ResNet.py")
    cli = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(cli)
    g = np.load(golden_dir / "preprocess_cases.npz")
    src, dst = tmp_path / "in.png", tmp_path / "out" / "adv.png"
    Image.fromarray(g["example_crop_in"]).save(src)
    cli.main([str(src), "--attack", "fgsm", "--eps", str(8 / 255), "--topk", "3", "--weights", "random", "--seed", "0",
              "--save_adv", str(dst)])
    lines = capsys.readouterr().out.strip().splitlines()
    assert lines[0] == f"Image: {src}" and lines[1] == "Clean:" and lines[5] == "Adversarial (fgsm):" and len(lines) == 9

    dev = torch.device("cuda:0")
    nm = advb200.native_model(weights.seeded_resnet50(0).to(dev))
    mean, std = weights.mean_std()
    x = pp.load_image(src, dev)
    logits = nm.forward(x, mean, std)["logits"]
    x_adv = advb200.fgsm_attack(nm, x, logits.argmax(1), eps=8 / 255, mean=mean, std=std)

    def fmt(lg):
        p, i = torch.softmax(lg, 1).topk(3, 1)
        return [f"Top {r}: {idx} (class {idx}), prob = {pv:.4f}" for r, (pv, idx) in
                enumerate(zip(p[0].cpu().tolist(), i[0].cpu().tolist()), start=1)]
    strip = lambda ls: [l.split(": ", 1)[1].split(" (class")[1] for l in ls]       # labels may be category names
    assert strip(lines[2:5]) == strip(fmt(logits))
    assert strip(lines[6:9]) == strip(fmt(nm.forward(x_adv, mean, std)["logits"]))
    assert np.array_equal(np.asarray(Image.open(dst)), preprocess_ref.to_pil_u8(x_adv[0].cpu().numpy()))
