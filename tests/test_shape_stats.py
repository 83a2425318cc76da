"""CPU checks of the integer restatements the generator / dataset kernels are tested against (tests/shape_stats.py), on
the reference's own images and labels (tests/golden/overlap_micro_100.npz = overlap_micro.zip, overlap_areas.npy =
2a/OVERLAP_AREAS): the label search reproduces the stored max-overlap labels, the Q14 rotation agrees with a float32
restatement of tf.contrib.image.rotate, and the operand-stage density of real batches is what the benchmark's
generated batches are held to (tests/test_generator.py, GPU)."""
import os

import numpy as np

from tests.shape_stats import (GOLDEN, max_overlap_label, nn_rotate, real_batch_planes, real_images, shape_stats,
                               stage_density, tf_rotate_float)


def test_label_search_reproduces_the_reference_labels():
    """README.md:19-27: the label is "the maximum possible overlap area between the two shapes".  The generator's
    search (8 rotations with aligned centroids + 4 shifts of the best one) must come close to the stored labels of
    the real pairs and never exceed what an exhaustive search could give (the stored label, up to resampling)."""
    L, K = real_images()
    y = np.load(os.path.join(GOLDEN, "overlap_areas.npy"))[:100]
    got = np.array([max_overlap_label(L[i], K[i]) for i in range(100)], dtype=np.float64)
    ratio = got / y
    assert ratio.mean() > 0.88 and np.median(ratio) > 0.93, (ratio.mean(), np.median(ratio))
    assert ratio.max() < 1.05, ratio.max()                 # nearest-neighbour resampling may gain a few pixels
    assert abs(got.mean() - y.mean()) < 0.12 * y.mean(), (got.mean(), y.mean())


def test_q14_rotation_agrees_with_tf_contrib_rotate():
    """The N1 / N2 kernels rotate the lock with a Q14 integer inverse map; tf.contrib.image.rotate (8c/utils.py:337-338)
    uses float32 projective transforms.  Bound the divergence on the real locks: pixels that differ, as a fraction of
    the lock's area, stay below 0.5 % per image (measured: mean 0.02 %, max 0.14 %)."""
    L, _ = real_images()
    rng = np.random.default_rng(0)
    worst = 0.0
    for i in range(16):
        for th in rng.uniform(0.0, 2.0 * np.pi, 6):
            a = nn_rotate(L[i], int(round(np.cos(th) * 16384)), int(round(np.sin(th) * 16384)))
            b = tf_rotate_float(L[i], float(th))
            worst = max(worst, float((a != b).sum()) / float(L[i].sum()))
    assert worst < 0.005, worst


def test_real_batches_fill_the_first_layer_operand():
    """What the zero-stage skipping of the first-layer GEMMs can save on the reference's data: 94 % of the forward
    GEMM's 128 x 64 operand stages and 96-97 % of the weight gradient's hold a set pixel."""
    fwd, wg = stage_density(real_batch_planes(512, seed=0), 512)
    assert 0.92 < fwd < 0.96 and 0.95 < wg < 0.985, (fwd, wg)


def test_real_shape_statistics():
    """The numbers quoted in kernels.cu / SURVEY 8.4 for the real images."""
    L, K = real_images()
    area, h, w, cy, cx = shape_stats(np.concatenate([L, K]))
    assert (area.min(), area.max()) == (710, 7213) and abs(area.mean() - 3401.7) < 0.1
    assert abs(np.median(h) - 98.5) <= 1 and abs(np.median(w) - 98) <= 1
    occ = np.concatenate([L, K]).mean(0)
    assert abs((occ > 0).mean() - 0.83175) < 1e-6
