"""The two random sources of the engine that replace unseeded TensorFlow ops, validated statistically (SURVEY 7.2:
"mean / var / KS, not bit-wise"):

* A2  `xavier_init` (8c/utils.py:346-353): tf.random_uniform(-b, b), b = sqrt(6 / (fan_in + fan_out)), zero biases
  (8c/vae.py:119) -- read back with magic_get_param right after magic_create;
* the N(0,1) noise of the reparameterisation (tf.random_normal, 8c/vae.py:367): three independent draws per step, one
  per stream, `normal_from_key` in kernels.cu (splitmix64 counter + Box-Muller) -- read through magic_export_latents
  with the heads zeroed, so that z = 0 + sqrt(exp(0)) * eps = eps exactly.
"""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ks_uniform(u: np.ndarray) -> float:
    u = np.sort(u)
    n = len(u)
    return float(max(np.max(np.arange(1, n + 1) / n - u), np.max(u - np.arange(0, n) / n)))


def _ks_normal(x: np.ndarray) -> float:
    from scipy.special import ndtr
    return _ks_uniform(ndtr(np.asarray(x, dtype=np.float64)))


def test_xavier_init_matches_the_reference_bounds():
    from magic_b200.engine import Engine, EngineConfig
    eng = Engine(EngineConfig(D=1600, encoder_neurons=(120, 72), decoder_neurons=(88, 104), latent=20, batch=32, seed=3))
    eng2 = Engine(EngineConfig(D=1600, encoder_neurons=(120, 72), decoder_neurons=(88, 104), latent=20, batch=32, seed=4))
    seen = []
    for name, shape in eng.param_shapes.items():
        w = eng.get_param(name).astype(np.float64)
        if len(shape) == 1:
            assert not w.any(), name                                    # tf.zeros, 8c/vae.py:119
            continue
        fan_in, fan_out = shape
        b = math.sqrt(6.0 / (fan_in + fan_out))                         # 8c/utils.py:349-350
        assert -b <= w.min() and w.max() < b, (name, w.min(), w.max(), b)
        n = w.size
        assert abs(w.mean()) < 5.0 * b / math.sqrt(3.0 * n), (name, w.mean())
        assert abs(w.var() / (b * b / 3.0) - 1.0) < 5.0 * math.sqrt(0.8 / n) + 1e-3, (name, w.var(), b * b / 3.0)
        assert w.max() > b * (1.0 - 20.0 / n) - 1e-9 or n < 2000          # the range is actually used up to its ends
        if n >= 5000:
            assert _ks_uniform((w.ravel() + b) / (2 * b)) < 1.63 / math.sqrt(n), name     # alpha = 0.01
            # rows / columns are not correlated copies of each other
            assert abs(np.corrcoef(w[0], w[1])[0, 1]) < 5.0 / math.sqrt(w.shape[1])
        assert not np.array_equal(w, eng2.get_param(name)), name         # the seed matters
        seen.append(w.ravel()[:1000])
    # different tensors draw from different streams
    for i in range(len(seen)):
        for j in range(i + 1, len(seen)):
            m = min(len(seen[i]), len(seen[j]))
            assert abs(np.corrcoef(seen[i][:m], seen[j][:m])[0, 1]) < 0.2
    eng.close(); eng2.close()


def _draw(eng, steps=1):
    """eps of `steps` consecutive steps: heads zeroed -> z == eps.  [steps, 3, B, L]"""
    out = []
    for _ in range(steps):
        eng.step()                      # d_eps3 = NULL: the internal counter-based draws; Adam moves the heads' weights by
        out.append(eng.export_latents().cpu().numpy().astype(np.float64))   # ~lr per step, far below the test's resolution
    return np.stack(out)


def test_device_normal_draws_are_standard_normal_and_independent():
    from magic_b200.engine import Engine, EngineConfig
    B, L = 512, 64
    cfg = dict(D=1600, encoder_neurons=(120, 72), decoder_neurons=(88, 104), latent=L, batch=B, lr=(1e-12, 1e-12),
               metric="sqdiff")                 # no batch-axis all-reduce: the rank-1 engine below draws without a communicator
    eng = Engine(EngineConfig(seed=11, **cfg))
    for name, shape in eng.param_shapes.items():
        if name.startswith("encoder_out_"):
            eng.set_param(name, np.zeros(shape, dtype=np.float32))
    eng.generate_batch(0, 0)
    z = _draw(eng, steps=3)                                              # [3 steps, 3 streams, B, L]
    n = z[0, 0].size
    for s in range(3):
        for k in range(3):
            x = z[s, k].ravel()
            assert abs(x.mean()) < 5.0 / math.sqrt(n), (s, k, x.mean())
            assert abs(x.var() - 1.0) < 5.0 * math.sqrt(2.0 / n), (s, k, x.var())
            assert abs((x ** 3).mean()) < 5.0 * math.sqrt(15.0 / n)      # skewness
            assert abs((x ** 4).mean() - 3.0) < 5.0 * math.sqrt(96.0 / n)   # kurtosis
            assert _ks_normal(x) < 1.63 / math.sqrt(n), (s, k)           # alpha = 0.01
            assert np.abs(x).max() > 3.5                                 # tails are populated
    # independence: between the three streams of a step, between consecutive steps, along the batch and latent axes
    flat = z.reshape(9, -1)
    c = np.corrcoef(flat)
    off = c[~np.eye(9, dtype=bool)]
    assert np.abs(off).max() < 5.0 / math.sqrt(n), np.abs(off).max()
    x = z[0, 0]
    assert abs(np.corrcoef(x[:-1].ravel(), x[1:].ravel())[0, 1]) < 5.0 / math.sqrt(n)
    assert abs(np.corrcoef(x[:, :-1].ravel(), x[:, 1:].ravel())[0, 1]) < 5.0 / math.sqrt(n)
    # another seed and another rank give other draws; the same seed reproduces them
    eng_b = Engine(EngineConfig(seed=12, **cfg))
    eng_r = Engine(EngineConfig(seed=11, rank=1, world=2, batch_global=2 * B, **cfg))       # no communicator needed to draw
    eng_c = Engine(EngineConfig(seed=11, **cfg))
    for e in (eng_b, eng_r, eng_c):
        for name, shape in e.param_shapes.items():
            if name.startswith("encoder_out_"):
                e.set_param(name, np.zeros(shape, dtype=np.float32))
        e.generate_batch(0, 0)
    # forward() draws with the current step counter (0 for all of these)
    from magic_b200 import _lib
    zs = []
    for e in (eng_b, eng_r, eng_c):
        e.forward(_lib.FWD_DISTANCE)
        zs.append(e.export_latents().cpu().numpy().astype(np.float64))
    np.testing.assert_array_equal(zs[2], z[0])                           # same (seed, step, rank) -> same draws
    for other in zs[:2]:
        assert abs(np.corrcoef(other.ravel(), z[0].ravel())[0, 1]) < 5.0 / math.sqrt(3 * n)
    for e in (eng, eng_b, eng_r, eng_c):
        e.close()
