"""advb200 — B200-native drop-in for the attack hot path of
picklerick422/Image-recognition-adversarial-example-attack (attack.py + the defenses of
defense_experiments.py).  Python here is only the host-side mirror of the reference's call
surface; all arithmetic runs in libadvb200.so (hand-written sm_100a CUDA behind a C ABI)."""
from .attack import CWResult, cw_l2_attack, fgsm_attack, normalize_batch, pgd_linf_attack  # noqa: F401
from .defenses import (  # noqa: F401
    _defense_median_smoothing, _defense_quantization, _defense_smoothing, _jpeg_compress_batch, _layer3_features, calibrate_feature_threshold,
    calibrate_threshold_from_scores, defend_input, evaluate_defenses, evaluate_defenses_batched,
    is_adversarial_by_feature, run_attack, COUNTER_KEYS,
)
from .model import NativeResNet50, NativeResNet50Module, get_default_precision, native_model, set_default_precision  # noqa: F401
from .pipeline import HostBatchPipeline  # noqa: F401
from . import preprocess  # noqa: F401
from .preprocess import get_transform, load_image  # noqa: F401
