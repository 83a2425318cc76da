"""Data-parallel restatement of the oracle step  --  TEST INFRASTRUCTURE ONLY (see tango_oracle.py's header).

The reference is single-device ("no collective exists in the reference"); the engine shards the batch
over ranks.  This file states, on the CPU and with an injected `all_reduce`, exactly the decomposition the engine's
step uses, so a world-size-2 gloo test can prove that it reproduces the full-batch oracle:

  * every batch mean (8c/vae.py:267,278,288,292,307) divides by the GLOBAL batch, so local gradients SUM over ranks;
  * cosine metric (8c/vae.py:438-452): forward all-reduce of the 2L batch-axis sums of squares, backward all-reduce of
    the L sums  S_j = sum_i g_i a^_ij b^_ij;  then  dL/da_ij = (g_i b^_ij - a^_ij S_j) / c_j  (zero normalisation term
    where the 1e-12 clamp is active);
  * g1 = d cost and g2 = d training_loss are all-reduced separately (two optimisers).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Callable, Dict, Tuple

import torch

from . import tango_oracle as orc


def rank_step_grads(cfg: orc.OracleConfig, params: Dict[str, torch.Tensor], x_lock, x_rot, x_key, labels, eps3,
                    batch_global: int, all_reduce: Callable[[torch.Tensor], torch.Tensor]):
    """One rank's share of a step.  `all_reduce(t)` must return the element-wise sum of t over all ranks.
    Returns (losses dict of GLOBAL means, distance[B_local], g1, g2) with g1/g2 already all-reduced."""
    names1 = [k for k in params if k not in orc.DEAD_PARAMS]
    names2 = [k for k in params if orc.is_encoder_param(k)]
    leaves = OrderedDict((k, v.detach().clone().requires_grad_(True)) for k, v in params.items())
    inv_bg = 1.0 / float(batch_global)

    mu, lv, z = orc.encode(cfg, leaves, x_lock, eps3[0])
    _, _, z_rot = orc.encode(cfg, leaves, x_rot, eps3[1])
    _, _, z_key = orc.encode(cfg, leaves, x_key, eps3[2])
    a = orc.decode_logits(cfg, leaves, z)
    rec = (torch.nn.functional.softplus(a) - x_lock * a).sum() * inv_bg
    kl = (-0.5 * (1.0 + lv - mu * mu - torch.exp(lv))).sum() * inv_bg
    de = cfg.def_weight * ((z - z_rot) ** 2).sum() * inv_bg
    cost_local = rec + kl + de
    g1 = torch.autograd.grad(cost_local, [leaves[k] for k in names1], retain_graph=True)

    zl, zk = z.detach(), z_key.detach()
    if cfg.metric == "cosine":
        colsq = all_reduce(torch.stack([(zl * zl).sum(0), (zk * zk).sum(0)]))          # [2, L] over the global batch
        ra = torch.rsqrt(torch.clamp(colsq[0], min=1e-12))
        rb = torch.rsqrt(torch.clamp(colsq[1], min=1e-12))
        ah, bh = zl * ra, zk * rb
        raw = (ah * bh).sum(1)
    else:
        raw = ((zl - zk) ** 2).sum(1)
    dist = 1.0 / raw if cfg.reciprocal else raw
    err = dist - labels
    train_local = (err * err).sum() * inv_bg
    dfd = -(dist * dist) if cfg.reciprocal else torch.ones_like(dist)
    g = 2.0 * err * inv_bg * dfd                                                        # d training_loss / d raw_i
    if cfg.metric == "cosine":
        S = all_reduce((g[:, None] * ah * bh).sum(0))                                   # [L]
        da = (g[:, None] * bh - torch.where(colsq[0] > 1e-12, ah * S, torch.zeros_like(ah))) * ra
        db = (g[:, None] * ah - torch.where(colsq[1] > 1e-12, bh * S, torch.zeros_like(bh))) * rb
    else:
        da = g[:, None] * 2.0 * (zl - zk)
        db = -da
    g2 = torch.autograd.grad([z, z_key], [leaves[k] for k in names2], grad_outputs=[da, db])

    g1 = [all_reduce(t) for t in g1]
    g2 = [all_reduce(t) for t in g2]
    sums = all_reduce(torch.stack([rec.detach(), kl.detach(), de.detach(), train_local.detach()]))
    losses = dict(cost=sums[0] + sums[1] + sums[2], training_loss=sums[3], rec_loss=sums[0], lat_loss=sums[1],
                  def_loss=sums[2])
    return losses, dist, dict(zip(names1, g1)), dict(zip(names2, g2))


def shard(t: torch.Tensor, rank: int, world: int, dim: int = 0) -> torch.Tensor:
    n = t.shape[dim] // world
    return t.narrow(dim, rank * n, n)
