"""Experiment 8c: `overlap_input.inputs()` (8c/overlap_input.py:76-124) for the B200 backend -- the handles main.py
evaluates with `.eval()` or fetches jointly through the session; the pipeline itself is magic_b200/overlap_input.py."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from magic_b200 import overlap_input as _b200_input  # noqa: E402

try:
    from .constants import FLAGS
except ImportError:
    from constants import FLAGS


def inputs(normalize=False, reshape=False, rotation=False):
    """(images, labels): images.eval() -> a batch [BATCH_SIZE, IMAGE_SIZE * IMAGE_SIZE * NUM_LAYERS] that stays in HBM
    (np.asarray() materialises it), labels.eval() -> its overlap areas [BATCH_SIZE]."""
    return _b200_input.inputs(normalize=normalize, reshape=reshape, rotation=rotation, flags=FLAGS)
