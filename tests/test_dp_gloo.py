"""World-size-2 CPU test (gloo) of the data-parallel decomposition the engine's step uses (oracle/dp_reference.py):
two ranks, each holding half the batch, all-reducing the cosine batch statistics and the two gradients, must
reproduce the full-batch oracle.  Also covers the single-process fold (all_reduce = identity at world 1)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import dp_reference as dp
from oracle import tango_oracle as orc

VARIANTS = {
    "cosine": dict(latent=6),
    "cosine-recip": dict(latent=3, reciprocal=True),
    "sqdiff-recip-elu": dict(latent=8, act="elu", metric="sqdiff", reciprocal=True, def_weight=100.0, lr=(1e-6, 1e-8)),
}
GEOM = dict(encoder_neurons=[24, 16], decoder_neurons=[12, 20], n_input=64)
B_GLOBAL = 8


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _inputs(cfg):
    state = orc.OracleState(cfg, seed=11)
    xl, xr, xk, y, eps = orc.synthetic_batch(B_GLOBAL, cfg.latent, seed=12, n_input=cfg.n_input)
    if cfg.reciprocal:
        y = y * 1e-3
    return state, xl, xr, xk, y, eps


def _worker(rank, world, port, variant, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        cfg = orc.OracleConfig(**GEOM, **VARIANTS[variant])
        state, xl, xr, xk, y, eps = _inputs(cfg)

        def all_reduce(t):
            t = t.clone()
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            return t

        sh = lambda t: dp.shard(t, rank, world)
        losses, d, g1, g2 = dp.rank_step_grads(cfg, state.params, sh(xl), sh(xr), sh(xk), sh(y),
                                               dp.shard(eps, rank, world, dim=1), B_GLOBAL, all_reduce)
        q.put((rank, {k: float(v) for k, v in losses.items()}, d.numpy(), {k: v.numpy() for k, v in g1.items()},
               {k: v.numpy() for k, v in g2.items()}))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("variant", sorted(VARIANTS))
def test_two_rank_decomposition_equals_full_batch_oracle(variant):
    import numpy as np
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, variant, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cfg = orc.OracleConfig(**GEOM, **VARIANTS[variant])
    state, xl, xr, xk, y, eps = _inputs(cfg)
    out, g1, g2 = state.grads(xl, xr, xk, y, eps)
    for rank, losses, d, r1, r2 in results:
        for k in orc.LOSS_KEYS:
            assert abs(losses[k] - float(out[k])) <= 1e-10 * abs(float(out[k])), (rank, k)
        n = B_GLOBAL // world
        np.testing.assert_allclose(d, out["distance"].numpy()[rank * n:(rank + 1) * n], rtol=1e-10)
        for k, v in g1.items():
            np.testing.assert_allclose(r1[k], v.numpy(), rtol=1e-8, atol=1e-12 * float(v.abs().max() + 1))
        for k, v in g2.items():
            np.testing.assert_allclose(r2[k], v.numpy(), rtol=1e-8, atol=1e-12 * float(v.abs().max() + 1))


def test_world_one_fold_is_the_oracle():
    import numpy as np
    cfg = orc.OracleConfig(**GEOM, **VARIANTS["cosine"])
    state, xl, xr, xk, y, eps = _inputs(cfg)
    losses, d, r1, r2 = dp.rank_step_grads(cfg, state.params, xl, xr, xk, y, eps, B_GLOBAL, lambda t: t)
    out, g1, g2 = state.grads(xl, xr, xk, y, eps)
    assert abs(float(losses["cost"]) - float(out["cost"])) <= 1e-12 * abs(float(out["cost"]))
    for k, v in g2.items():
        np.testing.assert_allclose(r2[k].numpy(), v.numpy(), rtol=1e-8, atol=1e-12 * float(v.abs().max() + 1))
