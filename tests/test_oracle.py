"""CPU tests of the oracle (oracle/tango_oracle.py) against everything the reference leaves behind that can pin
it: the label fixture and README numbers (KAT-1), the step-0 loss statistics printed in
{8c,8d,8e}/nohup.out (KAT-2), analytic checks (KAT-4), and the committed golden step (tests/golden/)."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import tango_oracle as orc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ------------------------------------------------------------------ KAT-1: label fixture vs README.md:44-47
def test_kat1_label_fixture_reproduces_readme_table():
    y = np.load(os.path.join(GOLD, "overlap_areas.npy"))
    assert y.shape == (2000,) and y.dtype == np.int64
    assert y[:10].tolist() == [2079, 839, 1365, 831, 1493, 490, 1692, 1369, 2180, 957]
    assert (y.min(), y.max()) == (296, 6426)
    y50 = y[:50].astype(np.float64)                       # 2a/constants.py:29 uses the first 50 labels
    # README 2d: guess 0 -> 5.8e+6 ; 2b: guess 500 -> 3.95e+6   (MSE reduction of 8c/vae.py:307)
    assert abs(np.mean(y50 ** 2) - 5800790.58) < 0.01
    assert abs(np.mean((500.0 - y50) ** 2) - 3958730.58) < 0.01
    # README 2c: guess U[350,2000) -> 2.50e+6 (expectation in closed form: var(u) + E(mean(u)-y)^2)
    a, b = 350.0, 2000.0
    exp_2c = np.mean((b - a) ** 2 / 12.0 + ((a + b) / 2.0 - y50) ** 2)
    assert abs(exp_2c - 2.50e6) / 2.50e6 < 0.01
    # the oracle's training_loss is that same reduction
    t = torch.tensor(y50)
    assert abs(float(((torch.zeros(50, dtype=torch.float64) - t) ** 2).mean()) - 5800790.58) < 0.01


def test_real_pairs_fixture_statistics():
    d = np.load(os.path.join(GOLD, "overlap_micro_16.npz"))
    assert d["lock_bits"].shape == (16, 5000) and d["key_bits"].shape == (16, 5000)
    al, ak = d["area_lock"], d["area_key"]
    both = np.concatenate([al, ak])
    assert both.min() == 710 and both.max() == 7213          # SURVEY 2.1 (measured on overlap_micro.zip)
    assert abs(both.mean() / 40000.0 - 0.085) < 0.005
    lock0 = np.unpackbits(d["lock_bits"][0])[:40000]
    assert int(lock0.sum()) == int(al[0])
    # label <= min(area_L, area_K) for (almost) every real pair: the label is an overlap area
    y = np.load(os.path.join(GOLD, "overlap_areas.npy"))[:100]
    assert np.sum(y <= np.minimum(al, ak)) >= 99


# ------------------------------------------------------------------ KAT-2: step-0 statistics of the shipped logs
@pytest.mark.parametrize("exp,log_rec,log_kl,log_def", [
    ("8c", 27734.2, 1.72452, 470.783),        # 8c/nohup.out:27
    ("8d", 27749.8, 16.012, 4471.38),         # 8d/nohup.out:27
])
def test_kat2_step0_statistics(exp, log_rec, log_kl, log_def):
    cfg = orc.OracleConfig.experiment(exp)
    torch.manual_seed(0)
    st = orc.OracleState(cfg, seed=0, dtype=torch.float32)
    B = 16
    xl, xr, xk, y, eps = orc.synthetic_batch(B, cfg.latent, seed=1, dtype=torch.float32)
    out = orc.forward(cfg, st.params, xl, xr, xk, y, eps)
    rec, kl, de = float(out["rec_loss"]), float(out["lat_loss"]), float(out["def_loss"])
    assert abs(rec - 40000 * math.log(2.0)) / (40000 * math.log(2.0)) < 0.02
    assert abs(rec - log_rec) / log_rec < 0.02
    # KL and deformation depend on the (unseeded) init and data: order of magnitude, like the logs
    assert 0.3 * log_kl < kl < 3.0 * log_kl, (kl, log_kl)
    assert 0.5 * log_def < de < 2.0 * log_def, (de, log_def)
    assert abs(float(out["cost"]) - (rec + kl + de)) < 1e-2 * (rec + kl + de)


# ------------------------------------------------------------------ KAT-4: analytic checks
MINI = orc.OracleConfig(encoder_neurons=[24, 16], decoder_neurons=[12, 20], latent=5, n_input=64)


@pytest.mark.parametrize("variant", ["cosine", "cosine-recip", "sqdiff-recip-elu"])
def test_kat4_autograd_matches_finite_differences(variant):
    cfg = orc.OracleConfig(encoder_neurons=[24, 16], decoder_neurons=[12, 20], latent=5, n_input=64)
    if variant == "cosine-recip":
        cfg.reciprocal = True
    if variant == "sqdiff-recip-elu":
        cfg.metric, cfg.reciprocal, cfg.act, cfg.def_weight = "sqdiff", True, "elu", 100.0
    st = orc.OracleState(cfg, seed=5)
    xl, xr, xk, y, eps = orc.synthetic_batch(6, cfg.latent, seed=6, n_input=cfg.n_input)
    if cfg.reciprocal:
        y = y * 1e-3           # keep (1/d - y) well scaled for the finite difference
    out, g1, g2 = st.grads(xl, xr, xk, y, eps)
    rng = np.random.RandomState(0)
    for which, key, g in ((1, "cost", g1), (2, "training_loss", g2)):
        for name in ("encoder_h0_weights", "encoder_out_log_sigma_weights", "encoder_b1_biases") + \
                (("decoder_h1_weights", "decoder_out_mean_biases") if which == 1 else ()):
            p = st.params[name]
            for _ in range(3):
                idx = tuple(int(rng.randint(s)) for s in p.shape)
                h = 1e-6
                old = float(p[idx])
                p[idx] = old + h
                fp = float(orc.forward(cfg, st.params, xl, xr, xk, y, eps)[key])
                p[idx] = old - h
                fm = float(orc.forward(cfg, st.params, xl, xr, xk, y, eps)[key])
                p[idx] = old
                fd = (fp - fm) / (2 * h)
                an = float(g[name][idx])
                # truncation O(h^2) + float64 round-off of the loss value itself (|f| eps / h)
                tol = 1e-5 * max(1.0, abs(an), abs(fd)) + 8e-16 * max(abs(fp), abs(fm)) / h
                assert abs(fd - an) <= tol, (variant, which, name, idx, fd, an)


def test_kat4_tf_literal_rec_loss_equals_stable_form():
    """8c/vae.py:260-261 writes -sum log(p^x (1-p)^(1-x)); for binary x that is softplus(a) - x a."""
    st = orc.OracleState(MINI, seed=7)
    xl, xr, xk, y, eps = orc.synthetic_batch(6, MINI.latent, seed=8, n_input=MINI.n_input)
    a = orc.forward(MINI, st.params, xl, xr, xk, y, eps, rec_mode="tf_literal")
    b = orc.forward(MINI, st.params, xl, xr, xk, y, eps, rec_mode="stable")
    assert abs(float(a["rec_loss"]) - float(b["rec_loss"])) <= 1e-12 * abs(float(b["rec_loss"]))
    assert abs(float(a["cost"]) - float(b["cost"])) <= 1e-12 * abs(float(b["cost"]))


def test_tf_literal_overflows_where_the_reference_would():
    """fp32 sigmoid saturation -> log(0) = -inf -> NaN/inf cost (the caller's NaN guard, 8c/main.py:51-52)."""
    p = torch.tensor([1.0], dtype=torch.float32)            # sigmoid saturated to exactly 1 on a 0 pixel
    x = torch.tensor([0.0], dtype=torch.float32)
    v = -torch.log(torch.pow(p, x) * torch.pow(1.0 - p, 1.0 - x))
    assert torch.isinf(v).all()


def test_cosine_distance_normalises_over_the_batch_axis():
    a = torch.tensor([[3.0, 1.0], [4.0, 1.0]], dtype=torch.float64)
    b = torch.tensor([[1.0, 2.0], [1.0, 2.0]], dtype=torch.float64)
    d = orc.cosine_distance(a, b)
    # columns of a: norms 5, sqrt2; columns of b: sqrt2, sqrt8
    exp0 = 3 / 5 * 1 / math.sqrt(2) + 1 / math.sqrt(2) * 2 / math.sqrt(8)
    exp1 = 4 / 5 * 1 / math.sqrt(2) + 1 / math.sqrt(2) * 2 / math.sqrt(8)
    assert torch.allclose(d, torch.tensor([exp0, exp1], dtype=torch.float64), atol=1e-14)
    z = torch.zeros(2, 2, dtype=torch.float64)               # clamp at 1e-12: no NaN
    assert torch.isfinite(orc.cosine_distance(z, b)).all()


def test_adam_is_the_tf1_formula_not_torch_optim():
    cfg = orc.OracleConfig(encoder_neurons=[8], decoder_neurons=[6, 6], latent=3, n_input=16)
    st = orc.OracleState(cfg, seed=9)
    name = "decoder_h2_weights"
    g = {k: torch.full_like(st.params[k], 0.5) for k in st.names1}
    g2 = {k: torch.full_like(st.params[k], -0.25) for k in st.names2}
    p0 = {k: v.clone() for k, v in st.params.items()}
    st.apply(g, g2)
    lr_t = 1e-4 * math.sqrt(1 - 0.999) / (1 - 0.9)
    m, v = 0.1 * 0.5, 0.001 * 0.25
    exp = p0[name] - lr_t * m / (math.sqrt(v) + 1e-8)
    assert torch.allclose(st.params[name], exp, rtol=0, atol=1e-15)
    # encoder variable: both optimisers subtract (W <- W - d1 - d2)
    en = "encoder_h0_weights"
    lr_t2 = 1e-6 * math.sqrt(1 - 0.999) / (1 - 0.9)
    m2, v2 = 0.1 * -0.25, 0.001 * 0.0625
    exp_e = p0[en] - lr_t * m / (math.sqrt(v) + 1e-8) - lr_t2 * m2 / (math.sqrt(v2) + 1e-8)
    assert torch.allclose(st.params[en], exp_e, rtol=0, atol=1e-15)
    # dead variables never move (8c/vae.py:141-142,147)
    for k in orc.DEAD_PARAMS:
        assert torch.equal(st.params[k], p0[k])


def test_param_shapes_counts_match_survey():
    n1 = {"8c": 41073040, "8d": 41343400, "8e": 44047000, "9a": 41046004, "11a": 40662160}
    n2 = {"8c": 20772040, "8d": 20952400, "8e": 22756000, "9a": 20754004, "11a": 20331160}
    for exp in n1:
        cfg = orc.OracleConfig.experiment(exp)
        shp = orc.param_shapes(cfg)
        c1 = sum(int(np.prod(s)) for k, s in shp.items() if k not in orc.DEAD_PARAMS)
        c2 = sum(int(np.prod(s)) for k, s in shp.items() if orc.is_encoder_param(k))
        assert (c1, c2) == (n1[exp], n2[exp]), exp


def test_interleave_roundtrip_and_channel_order():
    xl, xr, xk, _, _ = orc.synthetic_batch(3, 2, seed=0, n_input=64)
    X = orc.interleave(xl, xr, xk)
    assert X.shape == (3, 192)
    assert torch.equal(X[:, 0::3], xl) and torch.equal(X[:, 1::3], xr) and torch.equal(X[:, 2::3], xk)
    a, b, c = orc.deinterleave(X)
    assert torch.equal(a, xl) and torch.equal(b, xr) and torch.equal(c, xk)


# ------------------------------------------------------------------ committed golden step
def test_golden_step_is_reproduced_bit_for_bit():
    d = np.load(os.path.join(GOLD, "oracle_step_mini.npz"))
    cfg = orc.OracleConfig(encoder_neurons=[24, 16], decoder_neurons=[12, 20], latent=5, n_input=64)
    params = {k[3:]: torch.tensor(d[k]) for k in d.files if k.startswith("p0_")}
    st = orc.OracleState(cfg, params=params)
    t = lambda k: torch.tensor(d[k])
    out, g1, g2 = st.step(t("xl"), t("xr"), t("xk"), t("y"), t("eps"))
    np.testing.assert_allclose([float(out[k]) for k in orc.LOSS_KEYS], d["losses"], rtol=1e-12)
    np.testing.assert_allclose(out["distance"].numpy(), d["distance"], rtol=1e-12)
    for k, v in g1.items():
        np.testing.assert_allclose(v.numpy(), d["g1_" + k], rtol=1e-9, atol=1e-14)
    for k, v in g2.items():
        np.testing.assert_allclose(v.numpy(), d["g2_" + k], rtol=1e-9, atol=1e-14)
    for k, v in st.params.items():
        np.testing.assert_allclose(v.numpy(), d["p1_" + k], rtol=1e-12, atol=1e-15)
