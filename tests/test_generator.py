"""On-device shape-pair generator (replaces 8c/overlap_input.py): a numpy restatement of its integer
rotation and label arithmetic must reproduce the rotated plane and the label bit for bit from the exported lock/key
planes; the shape statistics must look like the reference's data (tests/golden/overlap_micro_16.npz)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from tests.shape_stats import max_overlap_label, nn_rotate, real_batch_planes, real_images, shape_stats, stage_density


def _generate(side: int, B: int, seed: int, step: int):
    from magic_b200.engine import Engine, EngineConfig
    eng = Engine(EngineConfig(D=side * side, encoder_neurons=(64,), decoder_neurons=(64, 64), latent=8, batch=B))
    eng.generate_batch(seed, step)
    X, y = eng.export_batch()
    rot = eng.export_rotation().cpu().numpy()
    X = X.cpu().numpy().reshape(B, side, side, 3)
    y = y.cpu().numpy()
    return eng, X[..., 0], X[..., 1], X[..., 2], y, rot


@pytest.mark.parametrize("side,B", [(200, 48), (40, 37), (256, 5)])
def test_rotation_and_label_are_bit_exact(side, B):
    eng, lock, rotp, key, y, rot = _generate(side, B, seed=3, step=7)
    assert set(np.unique(lock)) <= {0.0, 1.0} and set(np.unique(key)) <= {0.0, 1.0} and set(np.unique(rotp)) <= {0.0, 1.0}
    for i in range(B):
        cs, sn = int(rot[i, 0]), int(rot[i, 1])
        assert abs(cs * cs + sn * sn - 16384 * 16384) < 4 * 16384          # a unit vector in Q14
        np.testing.assert_array_equal(rotp[i], nn_rotate(lock[i], cs, sn), err_msg="rot plane, pair %d" % i)
        best = max_overlap_label(lock[i], key[i])
        assert int(y[i]) == best, (i, y[i], best)
    eng.close()


def test_batches_depend_on_seed_step_and_are_reproducible():
    eng, lock_a, _, _, y_a, _ = _generate(200, 8, seed=1, step=0)
    eng.generate_batch(1, 1)
    lock_b = eng.export_batch()[0].cpu().numpy().reshape(8, 200, 200, 3)[..., 0]
    eng.generate_batch(1, 0)
    lock_c = eng.export_batch()[0].cpu().numpy().reshape(8, 200, 200, 3)[..., 0]
    assert not np.array_equal(lock_a, lock_b)
    np.testing.assert_array_equal(lock_a, lock_c)
    eng.close()


def test_statistics_resemble_the_reference_data():
    """Shape statistics against the reference's own images (overlap_micro.zip, committed bit-packed as
    tests/golden/overlap_micro_100.npz): area, bounding box, centroid spread, canvas occupancy, label / area ratio."""
    L, K = real_images()
    r_area, r_h, r_w, r_cy, r_cx = shape_stats(np.concatenate([L, K]))
    n = 512
    eng, lock, rotp, key, y, rot = _generate(200, n, seed=0, step=0)
    imgs = np.concatenate([lock, key]) > 0
    area, h, w, cy, cx = shape_stats(imgs)
    q = [10, 50, 90]
    # area: 710..7213 px in the real data, mean 3402
    assert abs(area.mean() - r_area.mean()) < 0.15 * r_area.mean(), (area.mean(), r_area.mean())
    assert area.min() > 300 and area.max() < 12000, (area.min(), area.max())
    assert np.all(np.abs(np.percentile(area, q) - np.percentile(r_area, q)) < 0.35 * np.percentile(r_area, q))
    # bounding box (real: 10 / 50 / 90 % = 75 / 98 / 109 px high) and centroids (real: 57 / 101 / 140)
    for mine, real in ((h, r_h), (w, r_w)):
        assert np.all(np.abs(np.percentile(mine, q) - np.percentile(real, q)) < 16), (np.percentile(mine, q), np.percentile(real, q))
    for mine, real in ((cy, r_cy), (cx, r_cx)):
        assert np.all(np.abs(np.percentile(mine, q) - np.percentile(real, q)) < 12), (np.percentile(mine, q), np.percentile(real, q))
    # occupancy: the shapes reach the canvas edge (real: 83 % of all pixels are set in at least one of 200 images) and
    # the per-row / per-column occupancy profiles have the real data's shape
    occ, r_occ = imgs[:200].mean(0), np.concatenate([L, K]).mean(0)
    assert abs((occ > 0).mean() - (r_occ > 0).mean()) < 0.10, ((occ > 0).mean(), (r_occ > 0).mean())
    full = imgs.mean(0)
    for ax in (0, 1):
        assert np.corrcoef(full.mean(ax), r_occ.mean(ax))[0, 1] > 0.95
    # rotation keeps the area unless it moves part of the shape off the canvas; labels are overlap areas
    ar, al, ak = rotp.reshape(n, -1).sum(1), lock.reshape(n, -1).sum(1), key.reshape(n, -1).sum(1)
    assert np.all(ar <= al + 0.05 * al + 8)
    assert np.all(y <= np.minimum(al, ak)) and np.all(y >= 0)
    real_y = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "overlap_areas.npy"))
    ratio = y / np.minimum(al, ak)                         # real: label / min(area) 10 / 50 / 90 % = 0.63 / 0.81 / 0.98
    assert 0.55 < np.median(ratio) < 0.95, np.median(ratio)
    assert 0.6 * real_y.mean() < y.mean() < 1.4 * real_y.mean(), (y.mean(), real_y.mean())      # real mean 1972.8
    ang = np.arctan2(rot[:, 1], rot[:, 0])
    assert ang.min() < -2.0 and ang.max() > 2.0                          # angles cover the circle (U[0, 2 pi))
    eng.close()


def test_operand_stage_density_matches_real_batches():
    """The first layer's GEMMs skip operand stages without a set pixel, so the benchmark's generated batches must be as
    dense as batches of the reference's images: magic_input_density of a generated 4096-pair batch within +-0.05 of a
    4096-pair batch sampled from overlap_micro (0.94 forward / 0.96 weight gradient)."""
    from magic_b200.engine import Engine, EngineConfig
    B = 4096
    want = stage_density(real_batch_planes(B, seed=0), B)
    eng = Engine(EngineConfig.experiment("8d", batch=B, precision="bf16"))
    eng.generate_batch(0, 0)
    got = eng.input_density()
    eng.close()
    assert abs(got[0] - want[0]) < 0.05 and abs(got[1] - want[1]) < 0.05, (got, want)


def test_generated_batch_trains():
    """The packed lock bitmap the BCE epilogue reads and the bf16 planes describe the same image: a step on a
    generated batch equals a step on the same batch re-loaded through the interleaved fp32 path."""
    from magic_b200.engine import Engine, EngineConfig
    cfg = EngineConfig(D=1600, encoder_neurons=(120, 72), decoder_neurons=(88, 104), latent=20, batch=64, precision="fp32")
    eng = Engine(cfg)
    eng.generate_batch(5, 2)
    X, y = eng.export_batch()
    eps = torch.randn(3, 64, 20, generator=torch.Generator().manual_seed(0))
    l1, _ = eng.step(labels=y, eps3=eps, apply=False)
    l1 = l1.cpu().numpy().copy()
    g1 = eng.get_grad(1, "decoder_out_mean_weights").copy()
    eng.load_interleaved(X)
    l2, _ = eng.step(labels=y, eps3=eps, apply=False)
    np.testing.assert_allclose(l1, l2.cpu().numpy(), rtol=1e-6)
    g2 = eng.get_grad(1, "decoder_out_mean_weights")      # split-K red.add order differs run to run: fp32 round-off
    assert np.linalg.norm(g1 - g2) <= 1e-5 * np.linalg.norm(g2)
    eng.close()


def test_zero_stage_maps_match_the_exported_batch():
    """magic_input_density counts the 128 x 64 first-layer operand stages with a set pixel from the engine's
    zero-stage maps (the data its GEMMs skip by); recount them from the exported batch.  Forward: tiles of 128 packed
    rows [rot; lock; key] x 64 pixels (bias column D included); weight gradient: tiles of 128 pixels x 64 rows of
    the row ranges [0, 2B) and [B, 3B)."""
    from magic_b200.engine import Engine, EngineConfig
    B = 200                                    # not a multiple of 64 or 128: ragged last tiles
    eng = Engine(EngineConfig.experiment("8c", batch=B, precision="bf16"))
    eng.generate_batch(5, 2)
    dens = eng.input_density()
    X, _ = eng.export_batch()
    D = eng.D
    Xi = X.view(B, D, 3)
    planes = torch.cat([Xi[:, :, 1], Xi[:, :, 0], Xi[:, :, 2]], 0) > 0          # rows [rot; lock; key]
    kb = (D + 1 + 63) // 64
    rows_pad = (3 * B + 127) // 128 * 128
    P = torch.zeros(rows_pad, kb * 64, dtype=torch.bool, device=X.device)
    P[:3 * B, :D] = planes
    P[:3 * B, D] = True
    fwd = P.view(rows_pad // 128, 128, kb, 64).any(3).any(1)
    want_fwd = float(fwd.sum()) / (fwd.shape[0] * kb)
    tm = (D + 1 + 127) // 128
    kbw = (2 * B + 63) // 64
    tot = 0
    for g in range(2):
        Q = torch.zeros(kbw * 64, tm * 128, dtype=torch.bool, device=X.device)
        Q[:2 * B, :kb * 64] = P[g * B:g * B + 2 * B]
        tot += int(Q.view(kbw, 64, tm, 128).any(3).any(1).sum())
    want_wg = tot / (2.0 * tm * kbw)
    assert abs(dens[0] - want_fwd) < 1e-12 and abs(dens[1] - want_wg) < 1e-12, (dens, want_fwd, want_wg)
    assert 0.5 < dens[0] <= 1.0
    eng.close()
