"""silent_speech_b200 -- B200-native (sm_100a) implementation of the Silent-Speech
batched clip-classification hot path: landmark features, mouth-ROI CNN, 2-layer BiGRU,
attention-pooled word head and its train step, behind the C ABI of include/ssb200.h.

The package never falls back to eager PyTorch or the CPU: if libssb200.so is missing
or the tensors are not on a B200 the calls raise."""
from .landmarks import FIXED_IDXS, crop_rois, extract_feature, extract_features  # noqa: F401
from .model import BiGRUClassifier, load_classifier  # noqa: F401
from .dataset import ClipStore, scan_clips, split_by_label  # noqa: F401
from .legacy import GRUClassifier, MouthOpenDetector, load_word_model  # noqa: F401
from .streaming import StreamScheduler  # noqa: F401
from .live import ClipRecorder, classify_camera_clips, clips_from_camera, evaluate, top_confusions  # noqa: F401

__all__ = ["BiGRUClassifier", "load_classifier", "extract_features", "extract_feature", "crop_rois", "FIXED_IDXS",
           "ClipStore", "scan_clips", "split_by_label", "StreamScheduler", "GRUClassifier", "MouthOpenDetector",
           "load_word_model", "ClipRecorder", "classify_camera_clips", "clips_from_camera", "evaluate", "top_confusions"]
