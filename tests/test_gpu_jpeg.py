"""the JPEG defense (`_jpeg_compress_batch`, defense_experiments.py:188-207) on the
device.  Integer/byte work, so the bar is bit-exact: against the golden vectors the reference's own PIL round trip
produced (tests/golden/jpeg_cases.npz) and against the CPU oracle (oracle/jpeg_ref.py) on seeded inputs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import defenses_ref, jpeg_ref, weights  # noqa: E402

from test_oracle_golden import JPEG_CASES, jpeg_case  # noqa: E402


@pytest.fixture(scope="module")
def adv():
    import advb200
    return advb200


@pytest.mark.parametrize("name", JPEG_CASES)
def test_jpeg_bit_exact_vs_reference_golden(adv, golden_dir, name):
    x, q, want = jpeg_case(np.load(golden_dir / "jpeg_cases.npz"), name)
    got = adv._jpeg_compress_batch(torch.from_numpy(x).cuda(), q).cpu().numpy()
    assert np.array_equal(got, want.astype(np.float32) / np.float32(255))


@pytest.mark.parametrize("shape,q", [((3, 3, 48, 300), 75), ((2, 3, 33, 517), 60), ((1, 3, 8, 8), 75), ((1, 3, 15, 9), 90),
                                     ((4, 3, 224, 224), 75), ((1, 3, 256, 256), 10), ((1, 3, 2, 3), 100), ((2, 3, 31, 32), 1)])
def test_jpeg_bit_exact_vs_oracle(adv, shape, q):
    g = torch.Generator().manual_seed(sum(shape) + q)
    x = torch.rand(shape, generator=g) * 1.2 - 0.1            # includes values the clamp has to handle
    if shape[2] >= 16:                                          # low-frequency content so that AC terms survive
        x = torch.nn.functional.avg_pool2d(x, 5, 1, 2) * 1.5 - 0.2
    want = jpeg_ref.jpeg_compress_batch(x.numpy(), q)
    got = adv._jpeg_compress_batch(x.cuda(), q).cpu().numpy()
    assert np.array_equal(got, want)


def test_defend_input_with_jpeg_switch(adv, golden_dir):
    from advb200 import defenses
    g = np.load(golden_dir / "jpeg_cases.npz")
    x = torch.from_numpy(g["defend_x"]).cuda()
    defenses._DEFENSE_USE_JPEG, defenses._DEFENSE_JPEG_QUALITY = True, 75
    try:
        got = adv.defend_input(x).cpu().numpy()
    finally:
        defenses._DEFENSE_USE_JPEG = False
    assert np.array_equal(got, g["defend_jpeg_out"])
    assert np.array_equal(got, defenses_ref.defend_input(g["defend_x"], use_jpeg=True, jpeg_quality=75))
    assert np.array_equal(adv.defend_input(x).cpu().numpy(), defenses_ref.defend_input(g["defend_x"]))   # switch off again


def test_jpeg_full_batch_properties_and_errors(adv):
    """N=256 at 224x224 (BASELINE configs[1] shape): every value is k/255; each image is independent of the batch it
    rides in; a sample of the batch equals the oracle."""
    x = weights.synthetic_images(32).cuda().repeat(8, 1, 1, 1)
    x = (x + 0.1 * torch.rand(x.shape, generator=torch.Generator().manual_seed(5)).cuda()).clamp(0, 1)
    y = adv._jpeg_compress_batch(x, 75)
    k = (y * 255).round().cpu().numpy().astype(np.float32)
    assert np.array_equal(k / np.float32(255), y.cpu().numpy())      # true division, as ToTensor does on the host
    assert float(y.min()) >= 0 and float(y.max()) <= 1
    for i in (0, 100, 255):
        assert torch.equal(adv._jpeg_compress_batch(x[i:i + 1], 75)[0], y[i])
    idx = [3, 77, 200]
    assert np.array_equal(y[idx].cpu().numpy(), jpeg_ref.jpeg_compress_batch(x[idx].cpu().numpy(), 75))
    assert adv._jpeg_compress_batch(x[:0], 75).shape == (0, 3, 224, 224)
    with pytest.raises(ValueError):
        adv._jpeg_compress_batch(x[0], 75)
    with pytest.raises(ValueError):
        adv._jpeg_compress_batch(x.cpu(), 75)
    with pytest.raises(ValueError):
        adv._jpeg_compress_batch(x[:1], 0)
