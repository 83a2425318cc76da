"""The reference-facing Python surface on the GPU: `TangoEncoder` of an experiment directory (constants.py / vae.py /
main.py, SURVEY.md 8.2 B-py) against the oracle's `partial_fit`, plus the error behaviour main.py relies on."""
import importlib.util
import io
import os
import sys

import numpy as np
import pytest
import torch

from oracle import tango_oracle as orc
from tests.util import rel_l2

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_experiment(exp):
    """Import experiments/<exp>/{constants,vae}.py the way the reference runs them: as top-level modules from inside
    the directory."""
    d = os.path.join(ROOT, "experiments", exp)
    mods = {}
    for name in ("constants", "vae"):
        sys.modules.pop(name, None)
    sys.path.insert(0, d)
    try:
        for name in ("constants", "vae"):
            spec = importlib.util.spec_from_file_location(name, os.path.join(d, name + ".py"))
            m = importlib.util.module_from_spec(spec)
            sys.modules[name] = m
            spec.loader.exec_module(m)
            mods[name] = m
    finally:
        sys.path.remove(d)
        for name in ("constants", "vae"):
            sys.modules.pop(name, None)
    return mods["constants"].FLAGS, mods["vae"].TangoEncoder


@pytest.mark.parametrize("exp", ["8c", "9a", "11a"])
def test_partial_fit_matches_the_oracle_at_the_experiment_geometry(exp, capsys):
    FLAGS, TangoEncoder = load_experiment(exp)
    B = 8
    vae = TangoEncoder(sess=None, batch_size=B, precision="tf32", seed=0)
    printed = capsys.readouterr().out
    assert "Connecting network layer [40000->500]." in printed and "Adding 500 biases." in printed   # 8c/nohup.out:19-26
    assert vae.latent_dimensions == orc.OracleConfig.experiment(exp).latent and vae.batch_size == B
    ocfg = orc.OracleConfig.experiment(exp)
    from tests.util import well_conditioned_seed
    seed = well_conditioned_seed(ocfg, B, 0.02)
    state = orc.OracleState(ocfg, seed=seed)
    vae.set_params({k: v.to(torch.float32).numpy() for k, v in state.params.items() if k not in orc.DEAD_PARAMS})
    xl, xr, xk, y, eps = orc.synthetic_batch(B, ocfg.latent, seed=seed + 1)
    X = orc.interleave(xl, xr, xk).to(torch.float32).numpy()
    want = state.partial_fit(torch.tensor(X, dtype=torch.float64), y, eps)
    got = vae.partial_fit(X, overlap_areas=y.numpy(), eps=eps.to(torch.float32))
    assert len(got) == (6 if exp == "11a" else 5)                      # 11a/vae.py:399-411 returns distance too
    for g, w, k in zip(got[:5], want, orc.LOSS_KEYS):
        assert isinstance(g, float)
        # 1/distance (9a, 11a) amplifies any rounding near its pole: looser on the metric loss there
        tol = 1e-2 if (ocfg.reciprocal and k == "training_loss") else 2e-4
        assert abs(g - w) <= tol * abs(w), (k, g, w)
    # inference surface on the updated weights
    out = orc.forward(ocfg, state.params, xl, xr, xk, y, eps)
    e32 = eps.to(torch.float32)
    assert rel_l2(vae.transform(X, overlap_areas=y.numpy()), out["z_mean"].numpy()) <= 1e-3
    rec = vae.reconstruct(X, overlap_areas=y.numpy(), eps=e32)
    assert rec.shape == (B, 40000) and rel_l2(rec, torch.sigmoid(out["logits"]).numpy()) <= 1e-3
    z = np.random.RandomState(0).normal(size=(B, ocfg.latent))
    gen = vae.generate(z)
    assert rel_l2(gen, torch.sigmoid(orc.decode_logits(ocfg, state.params, torch.tensor(z))).numpy()) <= 1e-3
    assert vae.generate([0.0] * ocfg.latent).shape == (B, 40000)       # 8c/main.py:155 style call
    vae.close()


def importlib_vae(exp):
    import importlib
    return importlib.import_module("experiments.%s.vae" % exp)


def test_error_behaviour():
    FLAGS, TangoEncoder = load_experiment("8c")
    vae = TangoEncoder(sess=None, batch_size=4, verbose=False)
    with pytest.raises(ValueError):
        vae.partial_fit(np.zeros((3, 120000), np.float32), overlap_areas=np.zeros(3, np.float32))   # static batch size
    with pytest.raises(ValueError):
        vae.partial_fit(np.zeros((4, 120000), np.float32), overlap_areas=np.zeros(5, np.float32))
    vae.close()

    class BadFlags(FLAGS):
        DISTANCE_METRIC = "Manhattan"
    with pytest.raises(ValueError, match="Invalid distance metric"):                                # 8c/vae.py:305
        TangoEncoder(sess=None, flags=BadFlags, batch_size=4, verbose=False)
    # NaN is reported through cost, for the caller's guard (8c/main.py:51-52)
    vae = TangoEncoder(sess=None, batch_size=4, verbose=False)
    X = np.zeros((4, 120000), np.float32)
    cost = vae.partial_fit(X, overlap_areas=np.array([np.nan, 1, 2, 3], np.float32))
    assert np.isnan(cost[1]) and not np.isnan(cost[0])              # labels only reach the metric loss
    vae.close()


def test_main_train_runs_like_the_reference_driver(capsys):
    """experiments/8c/main.py: `train(sess, batch_size, training_epochs, display_step)` (8c/main.py:33) for one epoch at
    batch 128 (75 steps): the per-step loss line in the reference's format, the `mse:` line every 24 steps, the per-epoch
    cost line; reconstruction loss falls from D ln 2 (KAT-2 / KAT-3: 8c/nohup.out:27 -> :5039)."""
    import importlib
    main = importlib.import_module("experiments.8c.main")
    with main.Session() as sess:
        vae = main.train(sess, batch_size=128, training_epochs=1, display_step=1, verbose=False)
    out = capsys.readouterr().out
    lines = [l for l in out.splitlines() if l.startswith("Epoch: (")]
    assert len(lines) == 75 and "TRAINING LOSS:" in lines[0] and "[Reconstruction loss:" in lines[0]
    first = float(lines[0].split("Reconstruction loss: ")[1].split(";")[0])
    last = float(lines[-1].split("Reconstruction loss: ")[1].split(";")[0])
    assert abs(first - 40000 * np.log(2)) < 0.03 * 40000 * np.log(2)
    assert last < 0.7 * first, (first, last)
    assert out.count("mse: ") == 4                                       # i = 0, 24, 48, 72
    assert "Epoch: 0001 cost= " in out                                   # 8c/main.py:86-88
    assert vae.state_dict()["step"] == 75
    vae.close()


def test_reference_call_sequence_runs_unchanged(capsys):
    """The calls 8c/main.py:33-94 makes, in its order and with its expressions, against experiments/8c with
    BACKEND = 'b200': module-level `overlap_input.inputs(normalize=True, reshape=True, rotation=FLAGS.ROTATE)`,
    `images_batch.eval()` / `labels_batch.eval()`, `vae.partial_fit(batch_xs, overlap_areas=...)` -> 5 floats,
    `np.isnan(cost)`, `vae.get_predictions(...)`, `((predictions - areas) ** 2).mean(axis=None)`; then 11a's joint
    `sess.run([images_batch, labels_batch])` and 6-tuple (11a/main.py:65-68,94-105)."""
    import importlib
    from magic_b200.tango import Session
    overlap_input = importlib.import_module("experiments.8c.overlap_input")
    FLAGS, TangoEncoder = load_experiment("8c")
    images_batch, labels_batch = overlap_input.inputs(normalize=True, reshape=True, rotation=FLAGS.ROTATE)
    sess = Session()
    vae = TangoEncoder(sess=sess, verbose=False)
    assert vae.batch_size == FLAGS.BATCH_SIZE == 100
    costs = []
    for i in range(3):
        batch_xs = images_batch.eval()
        overlap_areas = labels_batch.eval()
        cost, training_loss, rec_loss, lat_loss, def_loss = vae.partial_fit(batch_xs, overlap_areas=overlap_areas)
        assert not np.isnan(cost) and isinstance(cost, float)
        costs.append(cost)
        if i % 24 == 0:
            accuracy_testing_images = images_batch.eval()
            accuracy_testing_overlap_areas = labels_batch.eval()
            predictions = vae.get_predictions(accuracy_testing_images, overlap_areas=accuracy_testing_overlap_areas)
            mse = ((predictions - accuracy_testing_overlap_areas) ** 2).mean(axis=None)
            assert predictions.shape == (100,) and np.isfinite(mse)
    assert costs[-1] < costs[0]
    # a caller that insists on arrays gets them, in the reference's layout, and they train the same way
    X = np.asarray(batch_xs)
    y = np.asarray(overlap_areas)
    assert X.shape == (100, 200 * 200 * 3) and X.dtype == np.float32 and set(np.unique(X)) <= {0.0, 1.0}
    assert y.shape == (100,) and np.all(y <= X.reshape(100, 40000, 3)[:, :, 0].sum(1))
    assert len(vae.partial_fit(X, overlap_areas=y)) == 5
    vae.close()
    # 11a: joint fetch, 6-tuple, inverted predictions
    overlap_input = importlib.import_module("experiments.11a.overlap_input")
    FLAGS, TangoEncoder = load_experiment("11a")
    images_batch, labels_batch = overlap_input.inputs(normalize=True, reshape=True, rotation=FLAGS.ROTATE)
    with Session() as sess:
        vae = TangoEncoder(sess=sess, verbose=False)
        batch_xs, overlap_areas = sess.run([images_batch, labels_batch])
        cost, training_loss, rec_loss, lat_loss, def_loss, distance = vae.partial_fit(batch_xs, overlap_areas=overlap_areas)
        assert distance.shape == (24,) and np.isfinite(cost)
        imgs2, areas2 = sess.run([images_batch, labels_batch])
        assert imgs2 is not batch_xs                                     # every run() dequeues a new batch
        predictions = 1. / vae.get_predictions(imgs2, overlap_areas=areas2)
        assert np.isfinite((np.array(predictions - areas2) ** 2).mean(axis=None))
        vae.close()
    with pytest.raises(NotImplementedError):
        overlap_input.inputs(normalize=False, reshape=True, rotation=True)


def test_cosine_distance_is_a_real_call():
    """`from vae import cosine_distance` (8c/vae.py:438-452): batch-axis l2-normalisation, sum over the latent axis."""
    cd = importlib_vae("8c").cosine_distance
    rng = np.random.default_rng(0)
    for B, L in ((100, 20), (7, 200), (513, 3)):
        a, b = rng.normal(size=(B, L)).astype(np.float32), rng.normal(size=(B, L)).astype(np.float32)
        na = a / np.sqrt(np.maximum((a.astype(np.float64) ** 2).sum(0), 1e-12))
        nb = b / np.sqrt(np.maximum((b.astype(np.float64) ** 2).sum(0), 1e-12))
        np.testing.assert_allclose(cd(a, b), (na * nb).sum(1), rtol=2e-5, atol=1e-6)
    z = np.zeros((4, 3), dtype=np.float32)
    assert np.array_equal(cd(z, z), np.zeros(4, dtype=np.float32))       # the 1e-12 clamp, not a NaN
    with pytest.raises(ValueError):
        cd(np.zeros((4, 3)), np.zeros((4, 2)))


def test_checkpoint_roundtrip():
    FLAGS, TangoEncoder = load_experiment("11a")
    a = TangoEncoder(sess=None, batch_size=16, verbose=False, seed=1)
    b0 = a.next_device_batch(0)
    a.partial_fit(b0, b0.labels)
    sd = a.state_dict()
    b = TangoEncoder(sess=None, batch_size=16, verbose=False, seed=2)
    b.load_state_dict(sd)
    X, y = b0.eval(), b0.labels.eval()
    eps = torch.randn(3, 16, 80, generator=torch.Generator().manual_seed(3))
    ra = a.partial_fit(X, y, eps=eps)
    rb = b.partial_fit(X, y, eps=eps)
    np.testing.assert_allclose(ra[:5], rb[:5], rtol=1e-5)
    np.testing.assert_allclose(ra[5], rb[5], rtol=1e-4)
    for n in ("encoder_h0_weights", "decoder_out_mean_biases"):
        pa, pb = a.get_params()[n], b.get_params()[n]
        assert np.linalg.norm(pa - pb) <= 1e-6 * np.linalg.norm(pa)
    a.close(); b.close()


def test_read_losses_returns_the_step_s_losses_without_draining_the_step():
    """magic_read_losses (partial_fit's way to its return values, 8c/vae.py:395-405): the five losses of the step just
    enqueued, equal to what the step wrote to its device buffer -- and available while the backward pass is still in
    flight (checked through a CUDA event recorded behind the step)."""
    from magic_b200 import _lib
    from magic_b200.engine import Engine, EngineConfig
    eng = Engine(EngineConfig.experiment("8d", batch=2048, precision="bf16", seed=1))
    with pytest.raises(_lib.MagicError, match="no step"):
        eng.read_losses()
    early = 0
    for step in range(6):
        eng.generate_batch(3, step)
        losses_dev, _ = eng.step()
        done = torch.cuda.Event()
        done.record()
        got = eng.read_losses()
        early += 0 if done.query() else 1                 # the step's tail had not finished when the losses were back
        want = losses_dev.cpu().numpy()
        np.testing.assert_array_equal(got, want)
        assert np.all(np.isfinite(got)) and got[0] > 0
    assert early >= 1, early                              # (a loaded box may finish a step before the host looks: 6 of 6 when idle)
    # steps that do not apply the update publish their losses the same way
    eng.generate_batch(3, 99)
    losses_dev, _ = eng.step(apply=False)
    np.testing.assert_array_equal(eng.read_losses(), losses_dev.cpu().numpy())
    eng.close()
