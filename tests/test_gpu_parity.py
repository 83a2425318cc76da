"""Parity of the CUDA engine (through the C ABI) against the float64 oracle on identical weights, inputs,
labels and noise.  Tolerances are BASELINE.json's: 1e-4 (fp32 mode) and 2e-2 (bf16 mode), applied as relative
L2 error per tensor (tests/util.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import tango_oracle as orc
from tests.util import TOL, engine_config_from_oracle, make_pair, oracle_config, rel_l2

pytestmark = pytest.mark.gpu

MINI = dict(encoder_neurons=[120, 72], decoder_neurons=[88, 104], n_input=1600)

CASES = [
    # name, oracle config kwargs, batch
    ("mini-cos", dict(latent=20, **MINI), 64),
    ("mini-cos-recip", dict(latent=2, reciprocal=True, **MINI), 48),
    ("mini-sqdiff-elu", dict(latent=80, act="elu", metric="sqdiff", reciprocal=True, def_weight=100.0,
                             lr=(1e-6, 1e-8), **MINI), 24),
    ("mini-ragged", dict(latent=7, encoder_neurons=[100], decoder_neurons=[60, 52], n_input=400), 37),
]


# Tolerances: BASELINE.json's north_star, 1e-4 (fp32 mode) / 2e-2 (bf16 mode) relative error, asserted as they stand
# for every loss, the distances and every gradient tensor of both optimisers.  One exception, and only for the
# configurations that take 1/distance (9a/vae.py:301, 11a/vae.py:306-311): training_loss, distance and the gradients of
# training_loss (optimiser 2) sit behind a pole that amplifies ANY rounding, the reference's own fp32 arithmetic
# included (with L = 2 a cosine sum is the sum of two terms of either sign).  For those tensors alone the bound is
# max(tol, AMP x floor), floor = the error of the oracle itself run in float32 and AMP = the operand precision of the
# mode relative to fp32 (hi+lo bf16 pairs 2^-17, bf16 2^-9; x2).  profiles/r02_parity_table.txt lists every tensor's
# achieved error, bound and floor; the cost-side gradients and the other four losses of 9a / 11a meet the plain tolerance.
AMP = {"fp32": 2.0 * 2 ** 7, "bf16": 2.0 * 2 ** 15}


def _allowed(ocfg, key, tol, amp, floor):
    behind_pole = ocfg.reciprocal and (key in ("training_loss", "distance") or (isinstance(key, tuple) and key[0] == 2))
    return max(tol, amp * floor) if behind_pole else tol


def _log_table(title, report):
    """Per-tensor achieved error / bound / float32-oracle floor; appended to $MAGIC_PARITY_TABLE (profiles/r02_parity_table.txt)."""
    path = os.environ.get("MAGIC_PARITY_TABLE")
    lines = ["== %s" % title] + ["%-46s err %.3e  bound %.3e  f32-oracle floor %.3e%s" % (
        k if isinstance(k, str) else "g%d %s" % k, v[0], v[1], v[2], "" if v[0] <= v[1] else "   <-- FAIL") for k, v in report.items()]
    if path:
        with open(path, "a") as f:
            f.write("\n".join(lines) + "\n")
    return "\n".join(lines)


def _compare_step(ocfg, batch, precision, title=None, scale_params=None):
    state, eng, (xl, xr, xk, y, eps) = make_pair(ocfg, batch, precision)
    if scale_params:
        for k, f in scale_params.items():
            state.params[k] = state.params[k] * f
        eng.set_params({k: v.to(torch.float32).numpy() for k, v in state.params.items() if k not in orc.DEAD_PARAMS})
    tol, amp = TOL[precision], AMP[precision]
    out, g1, g2 = state.grads(xl, xr, xk, y, eps)
    s32 = orc.OracleState(ocfg, params=state.params, dtype=torch.float32)
    f32 = lambda t: t.to(torch.float32)
    out32, g1_32, g2_32 = s32.grads(f32(xl), f32(xr), f32(xk), f32(y), f32(eps))
    eng.load_interleaved(orc.interleave(xl, xr, xk).to(torch.float32))
    losses, dist = eng.step(labels=y.to(torch.float32), eps3=eps.to(torch.float32), apply=False, want_distance=True)
    losses = losses.cpu().numpy()
    report = {}
    for i, k in enumerate(orc.LOSS_KEYS):
        ref = float(out[k])
        floor = abs(float(out32[k]) - ref) / max(abs(ref), 1e-30)
        err = abs(losses[i] - ref) / max(abs(ref), 1e-30)
        report[k] = (err, _allowed(ocfg, k, tol, amp, floor), floor)
    floor = rel_l2(out32["distance"].numpy(), out["distance"].numpy())
    report["distance"] = (rel_l2(dist.cpu().numpy(), out["distance"].numpy()), _allowed(ocfg, "distance", tol, amp, floor), floor)
    for which, g, g32 in ((1, g1, g1_32), (2, g2, g2_32)):
        for name, ref in g.items():
            floor = rel_l2(g32[name].numpy(), ref.numpy())
            report[(which, name)] = (rel_l2(eng.get_grad(which, name), ref.numpy()), _allowed(ocfg, (which, name), tol, amp, floor), floor)
    table = _log_table("%s  B=%d  %s" % (title or "L=%d %s" % (ocfg.latent, ocfg.metric), batch, precision), report)
    bad = {k: v for k, v in report.items() if not v[0] <= v[1]}
    assert not bad, "mismatch:\n" + table
    eng.close()
    return report


@pytest.mark.parametrize("name,kw,batch", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_losses_and_gradients(name, kw, batch, precision):
    _compare_step(oracle_config(**kw), batch, precision, title=name)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_full_size_8c(precision):
    """Experiment 8c as written (D=40000, 500x4, L=20), small batch so the float64 oracle finishes in seconds."""
    _compare_step(orc.OracleConfig.experiment("8c"), 12, precision, title="8c full geometry")


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_benchmark_config_8d_at_full_batch(precision):
    """bench.py's own workload: experiment 8d (L=200) at 4096 pairs -- the tile counts, split-K choices and
    persistent-loop trip counts of the measured run, checked against the float64 oracle (about 10 s of CPU each)."""
    _compare_step(orc.OracleConfig.experiment("8d"), 4096, precision, title="8d benchmark configuration")


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("exp,batch", [("8e", 24), ("9a", 48), ("11a", 24)])
def test_full_size_other_baseline_configs(exp, batch, precision):
    """BASELINE.json configs 3-5 at their real geometry (D = 40000): 8e L = 2000 (4000-wide head GEMM, K = 2000 decoder
    input), 9a L = 2 (heads that are almost all padding, reciprocal cosine), 11a (2 x 500 ELU encoder, L = 80,
    reciprocal squared difference, deformation weight 100) -- small batches so the float64 oracle finishes in seconds."""
    _compare_step(orc.OracleConfig.experiment(exp), batch, precision, title="%s full geometry" % exp)


def test_saturated_logits_bf16():
    """A trained decoder is confident: |logit| ~ 10-15 on most pixels.  The fast BCE epilogue forms sigmoid(s) as
    0.5 + (u - 0.5) from approximate ex2 / rcp; check its gradient where that subtraction cancels (all other parity
    cases use Xavier-initialised weights, |logit| < 1)."""
    ocfg = oracle_config(latent=20, **MINI)
    _compare_step(ocfg, 64, "bf16", title="saturated logits (decoder_out x40)", scale_params={"decoder_out_mean_weights": 40.0})


def _ten_steps(ocfg, batch, precision, title, steps=10, names=None):
    """Both Adam optimisers (TF1 formula, W <- W - d1 - d2) for `steps` steps on one batch; returns the per-tensor
    relative L2 error of the accumulated UPDATE after the first and after the last step.  The first Adam steps move
    every weight by ~lr regardless of |g|, so the update vector (not the parameter, which barely moves) is the
    sensitive quantity; what bounds the agreement is the sign of gradients that are zero to working precision."""
    state, eng, (xl, xr, xk, y, eps) = make_pair(ocfg, batch, precision)
    X = orc.interleave(xl, xr, xk).to(torch.float32)
    p0 = {k: v.clone() for k, v in state.params.items()}
    names = names or eng.param_names
    errs = {}
    for it in range(steps):
        out, _, _ = state.step(xl, xr, xk, y, eps)
        eng.load_interleaved(X)
        losses, _ = eng.step(labels=y.to(torch.float32), eps3=eps.to(torch.float32))
        l = losses.cpu().numpy()
        assert abs(l[0] - float(out["cost"])) <= max(TOL[precision], 1e-3) * abs(float(out["cost"])), (it, l[0])
        if it in (0, steps - 1):
            for name in names:
                upd_ref = (state.params[name] - p0[name]).numpy()
                upd = eng.get_param(name).astype(np.float64) - p0[name].numpy()
                if np.linalg.norm(upd_ref) == 0:
                    assert np.linalg.norm(upd) == 0, name
                    continue
                cos = float((upd * upd_ref).sum() / (np.linalg.norm(upd) * np.linalg.norm(upd_ref)))
                errs[(it + 1, name)] = (rel_l2(upd, upd_ref), cos)
    assert eng.step_count == steps
    eng.close()
    path = os.environ.get("MAGIC_PARITY_TABLE")
    if path:
        with open(path, "a") as f:
            f.write("== update after n steps: %s  B=%d  %s\n" % (title, batch, precision))
            for (n, name), (e, c) in errs.items():
                f.write("step %2d  %-40s rel-L2 of the update %.3e  cosine %.6f\n" % (n, name, e, c))
    return errs


# bound on the relative L2 error of the accumulated update (measured, profiles/r02_parity_table.txt: fp32 mode <= 0.5 %,
# bf16 mode 6-9.5 % after the first step and 1-4 % after ten): the first Adam steps are sign-like, so what is compared is
# the sign of gradients that are zero to working precision
UPDATE_TOL = {"fp32": 0.01, "bf16": 0.12}


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_ten_steps_track_the_oracle(precision):
    errs = _ten_steps(oracle_config(latent=20, **MINI), 64, precision, "mini-cos")
    for (n, name), (e, c) in errs.items():
        assert e <= UPDATE_TOL[precision] and c >= 1.0 - UPDATE_TOL[precision] ** 2, (n, name, e, c)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_full_size_8c_parameters_after_one_and_ten_steps(precision):
    """Experiment 8c at its real geometry: parameters after 1 and after 10 optimiser steps (SURVEY 8.3 parity procedure)."""
    names = ["encoder_h0_weights", "encoder_h3_weights", "encoder_out_mean_weights", "encoder_b0_biases", "decoder_h1_weights",
             "decoder_out_mean_weights", "decoder_out_mean_biases"]
    errs = _ten_steps(orc.OracleConfig.experiment("8c"), 12, precision, "8c full geometry", names=names)
    for (n, name), (e, c) in errs.items():
        assert e <= UPDATE_TOL[precision] and c >= 1.0 - UPDATE_TOL[precision] ** 2, (n, name, e, c)


def test_inference_surface():
    """get_predictions / transform / reconstruct / generate (8c/vae.py:407-434)."""
    from magic_b200 import _lib
    ocfg = oracle_config(latent=20, **MINI)
    state, eng, (xl, xr, xk, y, eps) = make_pair(ocfg, 64, "fp32")
    eng.load_interleaved(orc.interleave(xl, xr, xk).to(torch.float32))
    out = orc.forward(ocfg, state.params, xl, xr, xk, y, eps)
    e32 = eps.to(torch.float32)
    assert rel_l2(eng.forward(_lib.FWD_DISTANCE, e32).cpu().numpy(), out["distance"].numpy()) <= 1e-4
    assert rel_l2(eng.forward(_lib.FWD_ZMEAN, e32).cpu().numpy(), out["z_mean"].numpy()) <= 1e-4
    rec = torch.sigmoid(out["logits"]).numpy()
    assert rel_l2(eng.forward(_lib.FWD_RECON, e32).cpu().numpy(), rec) <= 1e-4
    z = torch.randn(64, 20, dtype=torch.float64)
    ref = torch.sigmoid(orc.decode_logits(ocfg, state.params, z)).numpy()
    assert rel_l2(eng.decode(z.to(torch.float32)).cpu().numpy(), ref) <= 1e-4
    eng.close()


def test_host_partial_fit_matches_device_path():
    ocfg = oracle_config(latent=20, **MINI)
    state, eng, (xl, xr, xk, y, eps) = make_pair(ocfg, 64, "fp32")
    X = orc.interleave(xl, xr, xk).to(torch.float32).numpy()
    out, _, _ = state.step(xl, xr, xk, y, eps)
    losses, dist = eng.partial_fit_host(X, y.numpy().astype(np.float32), eps.numpy().astype(np.float32), True)
    for i, k in enumerate(orc.LOSS_KEYS):
        assert abs(losses[i] - float(out[k])) <= 1e-4 * abs(float(out[k])), k
    assert rel_l2(dist, out["distance"].numpy()) <= 1e-4
    eng.close()


def test_uint8_feed_equals_float_feed():
    """A uint8 / bool X is a legal argument of the same partial_fit (feed_dict casts to the placeholder dtype,
    8c/vae.py:395): same step, a quarter of the PCIe bytes."""
    ocfg = oracle_config(latent=20, **MINI)
    _, eng, (xl, xr, xk, y, eps) = make_pair(ocfg, 64, "fp32", seed=0)
    _, eng2, _ = make_pair(ocfg, 64, "fp32", seed=0)
    X = orc.interleave(xl, xr, xk).to(torch.float32).numpy()
    yn, en = y.numpy().astype(np.float32), eps.numpy().astype(np.float32)
    la, da = eng.partial_fit_host(X, yn, en, True)
    lb, db = eng2.partial_fit_host((X * 255).astype(np.uint8), yn, en, True)
    np.testing.assert_allclose(la, lb, rtol=1e-5)
    np.testing.assert_allclose(da, db, rtol=1e-4, atol=1e-7)
    eng2.load_interleaved(torch.as_tensor(X > 0))                 # bool tensor through the device loader
    eng.load_interleaved(X)
    l3, _ = eng2.step(labels=yn, eps3=en, apply=False)
    l4, _ = eng.step(labels=yn, eps3=en, apply=False)
    np.testing.assert_allclose(l3.cpu().numpy(), l4.cpu().numpy(), rtol=1e-5)
    eng.close(); eng2.close()


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_packed_first_layer_operand_matches_bf16_planes(precision, monkeypatch):
    """The first layer's A operand is read as 1 bit per pixel and expanded inside the GEMM, and operand stages
    without a set pixel are skipped (default); "skip_zero_stages" = 0 processes every stage; with MAGIC_BITS_A=0 the
    same GEMMs read bf16 planes through TMA.  Same products, exact zeros left out: losses, the batch the generator
    exports and every gradient must agree to fp32 summation order (tile widths / split-K slices differ)."""
    from magic_b200.engine import Engine
    from tests.util import engine_config_from_oracle
    ocfg = oracle_config(latent=20, encoder_neurons=[120, 72], decoder_neurons=[88, 104], n_input=1600)
    runs = []
    for flag, skip in (("1", 1), ("1", 0), ("0", 1)):
        monkeypatch.setenv("MAGIC_BITS_A", flag)
        eng = Engine(engine_config_from_oracle(ocfg, 300, precision, seed=5))
        eng.set_option("skip_zero_stages", skip)
        eng.generate_batch(3, 7)
        dens = eng.input_density()
        losses, dist = eng.step(apply=False, want_distance=True)
        X, y = eng.export_batch()
        grads = {(w, n): eng.get_grad(w, n) for w in (1, 2) for n in eng.param_names if eng.param_opts[n] & w}
        runs.append((losses.cpu().numpy(), dist.cpu().numpy(), X.cpu().numpy(), y.cpu().numpy(), grads, dens))
        eng.close()
    a = runs[0]
    assert set(np.unique(a[2])) <= {0.0, 1.0} and a[2].sum() > 0
    assert 0.0 < a[5][0] <= 1.0 and 0.0 < a[5][1] <= 1.0, a[5]
    assert runs[2][5] == (1.0, 1.0)                                  # no packed rows, no map
    # fp32 sums taken in another order differ in the last bits; in the bf16 mode such a difference can flip the
    # rounding of a stored activation (2^-9 relative), so the modes agree to a tenth of the mode's tolerance
    tol = TOL[precision] / 10
    for b in runs[1:]:
        assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])          # the same generated batch
        np.testing.assert_allclose(a[0], b[0], rtol=tol)
        assert rel_l2(a[1], b[1]) < tol
        for k in a[4]:
            assert rel_l2(a[4][k], b[4][k]) < tol, k


def test_non_binary_input_is_rejected():
    """The engine keeps images as bits.  The reference's placeholder would take any float X, but its pipeline only ever
    feeds {0,1} (8c/overlap_input.py:117); an array feed with a grey pixel must raise, not be thresholded silently."""
    from magic_b200._lib import MagicError
    ocfg = oracle_config(latent=20, **MINI)
    _, eng, (xl, xr, xk, y, eps) = make_pair(ocfg, 64, "bf16", seed=0)
    X = orc.interleave(xl, xr, xk).to(torch.float32).numpy()
    eng.load_interleaved(X)                                     # binary: accepted
    Xg = X.copy()
    Xg[3, 1234] = 0.5
    with pytest.raises(MagicError, match="other than 0 and 1"):
        eng.load_interleaved(Xg)
    with pytest.raises(MagicError, match="other than 0 and 1"):
        eng.partial_fit_host(Xg, y.numpy().astype(np.float32))
    X8 = (X * 255).astype(np.uint8)
    eng.load_interleaved(X8)                                    # {0,255}: accepted
    X8[5, 77] = 128
    with pytest.raises(MagicError, match="other than 0 and 1"):
        eng.load_interleaved(X8)
    losses, _ = eng.partial_fit_host(X, y.numpy().astype(np.float32))      # the engine is still usable afterwards
    assert np.all(np.isfinite(losses))
    eng.close()


def test_zero_stage_skipping_on_sparse_images_at_full_geometry():
    """set_option("skip_zero_stages", 1) at the 8c geometry on images that leave most of the canvas empty (a 32 x 32
    patch): the first-layer GEMMs skip the empty operand stages and must reproduce the default (every stage multiplied)
    path up to fp32 summation order."""
    from magic_b200.engine import Engine, EngineConfig
    B = 160
    rng = np.random.default_rng(0)
    img = np.zeros((B, 200, 200, 3), dtype=np.float32)
    img[:, 64:96, 96:128, :] = (rng.random((B, 32, 32, 3)) < 0.5)
    X = img.reshape(B, -1)
    y = rng.uniform(300, 6000, B).astype(np.float32)
    eps = rng.normal(size=(3, B, 20)).astype(np.float32)
    out = []
    for skip in (0, 1):
        eng = Engine(EngineConfig.experiment("8c", batch=B, precision="bf16", seed=2))
        eng.set_option("skip_zero_stages", skip)
        eng.load_interleaved(X)
        dens = eng.input_density()
        losses, dist = eng.step(labels=y, eps3=eps, apply=False, want_distance=True)
        out.append((losses.cpu().numpy().copy(), dist.cpu().numpy().copy(), dens,
                    {(w, n): eng.get_grad(w, n) for w, n in ((1, "encoder_h0_weights"), (2, "encoder_h0_weights"),
                                                              (1, "encoder_b0_biases"), (1, "decoder_out_mean_weights"))}))
        eng.close()
    assert out[1][2][0] < 0.25 and out[1][2][1] < 0.25, out[1][2]          # almost every stage is empty
    tol = TOL["bf16"] / 10
    np.testing.assert_allclose(out[0][0], out[1][0], rtol=tol)
    assert rel_l2(out[0][1], out[1][1]) < tol
    for k in out[0][3]:
        assert rel_l2(out[0][3][k], out[1][3][k]) < tol, k


CHAIN_CASES = [
    # name, engine config, batch, exact
    ("8d-4096", dict(exp="8d"), 4096, True),              # the benchmark shape: per-layer plans = 256-wide tiles, no split-K
    ("8e-2048", dict(exp="8e"), 2048, False),             # heads too deep for the chain (K = 4000): hidden layers only
    ("11a-4096", dict(exp="11a"), 4096, False),           # E = 2, ELU: no forward chain (one hidden layer), dgrad chain of two
    ("small-D-8192", dict(D=1600, encoder_neurons=(500, 500, 500, 500), decoder_neurons=(500, 500), latent=20), 8192, False),
]                                                          # 192 / 256 slabs: more than one slab per CTA


@pytest.mark.parametrize("name,kw,batch,exact", CHAIN_CASES, ids=[c[0] for c in CHAIN_CASES])
def test_hidden_layer_chains(name, kw, batch, exact, monkeypatch):
    """chain_sm100.cuh runs the hidden layers (forward) and their dgrads in one launch each, the activation slab resident
    in shared memory.  Same MMAs in the same order and the same epilogue arithmetic as the per-layer launches: with
    MAGIC_CHAIN=0 (every layer its own launch) losses, distances and every gradient must come out the same -- to the
    last bits of the atomically summed scalars where the per-layer plans are not split along K, to a tenth of the mode's
    tolerance otherwise.  Level 2 = forward and dgrad chains (the engine's default, level 1, is the forward chain only and
    is what every other test in this file runs)."""
    from magic_b200.engine import Engine, EngineConfig
    runs = []
    for chain in ("2", "0"):
        monkeypatch.setenv("MAGIC_CHAIN", chain)
        kw2 = dict(kw)
        cfg = EngineConfig.experiment(kw2.pop("exp"), batch=batch, precision="bf16", seed=3) if "exp" in kw2 \
            else EngineConfig(batch=batch, precision="bf16", seed=3, **kw2)
        eng = Engine(cfg)
        eng.generate_batch(5, 1)
        eps = torch.randn(3, batch, cfg.latent, generator=torch.Generator().manual_seed(1))
        losses, dist = eng.step(eps3=eps, apply=False, want_distance=True)
        names = [n for n in eng.param_names if n.startswith("encoder_") and not n.endswith("h0_weights")]
        grads = {(w, n): eng.get_grad(w, n) for w in (1, 2) for n in names}
        grads[(1, "encoder_h0_weights")] = eng.get_grad(1, "encoder_h0_weights")
        runs.append((losses.cpu().numpy().copy(), dist.cpu().numpy().copy(), grads, int(eng.stats().launches_last_step)))
        eng.close()
    on, off = runs
    assert on[3] < off[3], (on[3], off[3])                 # fewer launches with the chains
    if exact:
        # same products in the same order inside every GEMM; what differs run to run anyway is the order of the atomic
        # adds behind the loss scalars, the head-bias sums and the batch-axis statistics (last bits)
        np.testing.assert_allclose(on[0], off[0], rtol=2e-6)
        assert rel_l2(on[1], off[1]) < 2e-6
        for k in on[2]:
            # bf16 gradients downstream of those sums: a last-bit change upstream flips a few bf16 roundings (measured
            # 4.5e-5 on optimiser 2's encoder_h1_weights; one flipped bf16 rounding is 4e-3 of that element)
            assert rel_l2(on[2][k], off[2][k]) < 2e-4, k
    else:
        tol = TOL["bf16"] / 10
        np.testing.assert_allclose(on[0], off[0], rtol=tol)
        assert rel_l2(on[1], off[1]) < 10 * tol                                  # (distances of reciprocal configs sit behind a pole)
        for k in on[2]:
            assert rel_l2(on[2][k], off[2][k]) < (10 * tol if k[0] == 2 else tol), k


ADAM_FUSE_CASES = [
    # name, experiment, batch, precision
    ("8d-4096-bf16", "8d", 4096, "bf16"),       # the benchmark shape: 592 of 626 tile pairs updated in the epilogue
    ("8c-300-bf16", "8c", 300, "bf16"),         # ragged batch (K = 600 rows per group)
    ("8d-1024-fp32", "8d", 1024, "fp32"),       # split-bf16 mode: three precision segments, hi + lo shadow weights
]


@pytest.mark.parametrize("name,exp,batch,precision", ADAM_FUSE_CASES, ids=[c[0] for c in ADAM_FUSE_CASES])
def test_first_layer_adam_in_the_wgrad_epilogue(name, exp, batch, precision):
    """Single GPU: the first layer's dual Adam (8c/vae.py:314-321) runs in the epilogue of its weight-gradient GEMM
    (EPI_ADAM, gemm_sm100.cuh).  Same arithmetic on the same accumulators as storing the gradient and updating in
    adam_dual_kernel: after four steps on fresh batches (both input buffers, growing bias correction) every parameter and
    both optimisers' moments must agree with the option off as closely as two runs with the option off agree."""
    from magic_b200.engine import Engine, EngineConfig
    runs = []
    for fuse in (1, 0, 0):
        eng = Engine(EngineConfig.experiment(exp, batch=batch, precision=precision, seed=9))
        eng.set_option("fuse_first_layer_adam", fuse)          # raises if the handle could not build the fused plans
        eps_gen = torch.Generator().manual_seed(4)
        losses = []
        for step in range(4):
            eng.generate_batch(7, step)
            eps = torch.randn(3, batch, eng.L, generator=eps_gen)
            losses.append(eng.step(eps3=eps)[0].cpu().numpy().copy())
        params = {n: eng.get_param(n) for n in eng.param_names}
        slots = {(w, s): eng.get_slot(w, s, "encoder_h0_weights") for w in (1, 2) for s in (0, 1)}
        slots[("b", 1, 0)] = eng.get_slot(1, 0, "encoder_b0_biases")
        runs.append((np.stack(losses), params, slots))
        eng.close()
    on, off, off2 = runs
    # Two runs of ONE setting already differ: the batch-axis statistics are summed with atomics, their last bits move
    # the gradients' last bits, and Adam's first steps turn a near-zero gradient's sign into +-lr (measured on 8d / 4096:
    # 1.9e-4 of the first layer's norm after four steps).  The fused path must stay inside that run-to-run band (x10, floor
    # 1e-4: a different FMA contraction of the same expression in the two kernels is enough to leave a x5 band; a wrong
    # update -- a skipped or doubled tile -- is 1e-2).
    rel = lambda a, b: float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-30)))
    assert rel(on[0], off[0]) <= max(5.0 * rel(off2[0], off[0]), 1e-4), (rel(on[0], off[0]), rel(off2[0], off[0]))
    for n in on[1]:
        floor = rel_l2(off2[1][n], off[1][n])
        assert rel_l2(on[1][n], off[1][n]) <= max(10.0 * floor, 1e-4), (n, floor)
    for k in on[2]:
        floor = rel_l2(off2[2][k], off[2][k])
        assert rel_l2(on[2][k], off[2][k]) <= max(10.0 * floor, 1e-4), (k, floor)
