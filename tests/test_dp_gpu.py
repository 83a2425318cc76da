"""Two-GPU data-parallel step (NCCL inside libmagic_b200.so) against the full-batch float64 oracle: each rank holds
half the batch; losses are global means, gradients are all-reduced sums, parameters stay identical on both ranks.
Needs 2 GPUs (`gpurun --gpus 2`); skipped on a single-GPU box."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MINI = dict(encoder_neurons=[120, 72], decoder_neurons=[88, 104], n_input=1600)
B_GLOBAL = 64


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, kw, precision, q):
    import torch.distributed as dist
    from oracle import tango_oracle as orc
    from magic_b200.engine import Engine
    from tests.util import engine_config_from_oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ocfg = orc.OracleConfig(**MINI, **kw)
        state = orc.OracleState(ocfg, seed=0)
        xl, xr, xk, y, eps = orc.synthetic_batch(B_GLOBAL, ocfg.latent, seed=1, n_input=ocfg.n_input)
        n = B_GLOBAL // world
        sl = slice(rank * n, (rank + 1) * n)
        eng = Engine(engine_config_from_oracle(ocfg, n, precision, rank=rank, world=world, batch_global=B_GLOBAL, seed=0))
        eng.comm_init_from_process_group()
        eng.set_params({k: v.to(torch.float32).numpy() for k, v in state.params.items() if k not in orc.DEAD_PARAMS})
        X = orc.interleave(xl[sl], xr[sl], xk[sl]).to(torch.float32)
        res = {}
        eng.load_interleaved(X)
        losses, d = eng.step(labels=y[sl].to(torch.float32), eps3=eps[:, sl].to(torch.float32), apply=False, want_distance=True)
        res["losses"] = losses.cpu().numpy().copy()
        res["distance"] = d.cpu().numpy().copy()
        res["g1"] = {k: eng.get_grad(1, k) for k in ("encoder_h0_weights", "encoder_out_mean_weights", "decoder_out_mean_weights", "decoder_b1_biases")}
        res["g2"] = {k: eng.get_grad(2, k) for k in ("encoder_h0_weights", "encoder_out_log_sigma_weights", "encoder_b1_biases")}
        for _ in range(3):
            eng.load_interleaved(X)
            eng.step(labels=y[sl].to(torch.float32), eps3=eps[:, sl].to(torch.float32))
        # a 4th evaluation on the updated weights: checks the bf16 shadows every rank received from the owners
        eng.load_interleaved(X)
        l4, _ = eng.step(labels=y[sl].to(torch.float32), eps3=eps[:, sl].to(torch.float32), apply=False)
        res["losses4"] = l4.cpu().numpy().copy()
        res["p"] = eng.get_param("encoder_h1_weights")
        res["p_out"] = eng.get_param("decoder_out_mean_weights")
        res["fused"] = os.environ.get("MAGIC_DP_MODE", "fused")
        eng.sync()
        q.put((rank, res))
        eng.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name,kw", [("cosine", dict(latent=20)),
                                     ("sqdiff-recip-elu", dict(latent=80, act="elu", metric="sqdiff", reciprocal=True,
                                                               def_weight=100.0, lr=(1e-6, 1e-8)))])
def test_two_gpu_step_matches_full_batch_oracle(name, kw):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from oracle import tango_oracle as orc
    from tests.util import rel_l2
    world, precision, tol = 2, "fp32", 1e-4
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, kw, precision, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    ocfg = orc.OracleConfig(**MINI, **kw)
    state = orc.OracleState(ocfg, seed=0)
    xl, xr, xk, y, eps = orc.synthetic_batch(B_GLOBAL, ocfg.latent, seed=1, n_input=ocfg.n_input)
    out, g1, g2 = state.grads(xl, xr, xk, y, eps)
    n = B_GLOBAL // world
    for rank in range(world):
        r = results[rank]
        for i, k in enumerate(orc.LOSS_KEYS):
            t = 1e-2 if (ocfg.reciprocal and k == "training_loss") else tol
            assert abs(r["losses"][i] - float(out[k])) <= t * abs(float(out[k])), (rank, k, r["losses"][i], float(out[k]))
        assert rel_l2(r["distance"], out["distance"].numpy()[rank * n:(rank + 1) * n]) <= (1e-2 if ocfg.reciprocal else tol)
        for k, v in r["g1"].items():
            assert rel_l2(v, g1[k].numpy()) <= tol, (rank, 1, k)
        for k, v in r["g2"].items():
            assert rel_l2(v, g2[k].numpy()) <= (2e-2 if ocfg.reciprocal else tol), (rank, 2, k)
    np.testing.assert_array_equal(results[0]["p"], results[1]["p"])       # replicas stay bit-identical
    # three optimiser steps, then the losses of a 4th forward, against the oracle stepping on the full batch
    p0 = {k: v.clone() for k, v in state.params.items()}
    for _ in range(3):
        state.step(xl, xr, xk, y, eps)
    out4 = orc.forward(ocfg, state.params, xl, xr, xk, y, eps)
    for rank in range(world):
        for i, k in enumerate(orc.LOSS_KEYS):
            t = 2e-2 if (ocfg.reciprocal and k == "training_loss") else 2e-3
            assert abs(results[rank]["losses4"][i] - float(out4[k])) <= t * abs(float(out4[k])), (rank, k)
    for name, key in (("encoder_h1_weights", "p"), ("decoder_out_mean_weights", "p_out")):
        upd_ref = (state.params[name] - p0[name]).numpy()
        upd = results[0][key].astype(np.float64) - p0[name].numpy()
        assert rel_l2(upd, upd_ref) <= 0.05, name
