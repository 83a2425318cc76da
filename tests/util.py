"""Shared helpers of the parity tests: build an oracle state and a CUDA engine from the SAME fp32 weights,
inputs, labels and noise, and compare tensor by tensor."""
from __future__ import annotations

import numpy as np
import torch

from oracle import tango_oracle as orc

# Tolerances of BASELINE.json's north_star, as relative L2 error per tensor (||a-b|| / ||b||):
TOL = {"fp32": 1e-4, "bf16": 2e-2}


def rel_l2(a, b) -> float:
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / den) if den > 0 else float(np.linalg.norm(a))


def oracle_config(**kw) -> orc.OracleConfig:
    return orc.OracleConfig(**kw)


def engine_config_from_oracle(ocfg: orc.OracleConfig, batch: int, precision: str, **kw):
    from magic_b200.engine import EngineConfig
    return EngineConfig(D=ocfg.n_input, encoder_neurons=tuple(ocfg.encoder_neurons),
                        decoder_neurons=tuple(ocfg.decoder_neurons), latent=ocfg.latent, act=ocfg.act,
                        metric=ocfg.metric, reciprocal=ocfg.reciprocal, def_weight=ocfg.def_weight, lr=tuple(ocfg.lr),
                        beta1=ocfg.beta1, beta2=ocfg.beta2, eps=ocfg.eps, precision=precision, batch=batch, **kw)


def well_conditioned_seed(ocfg: orc.OracleConfig, batch: int, margin: float) -> int:
    """9a/11a take 1/distance (9a/vae.py:301): a pair whose distance is ~0 sits on a pole where no finite
    precision agrees with another.  Pick the first seed whose batch keeps every |distance| at least `margin`
    times the typical magnitude of the terms it sums (the pole itself is not a property of the kernels)."""
    if not ocfg.reciprocal:
        return 0
    for seed in range(200):
        state = orc.OracleState(ocfg, seed=seed)
        xl, xr, xk, y, eps = orc.synthetic_batch(batch, ocfg.latent, seed=seed + 1, n_input=ocfg.n_input)
        out = orc.forward(ocfg, state.params, xl, xr, xk, y, eps)
        raw = 1.0 / out["distance"]
        if ocfg.metric == "cosine":
            a, b = out["z"], out["z_key"]
            na = a * torch.rsqrt((a * a).sum(0, keepdim=True))
            nb = b * torch.rsqrt((b * b).sum(0, keepdim=True))
            scale = (na * nb).abs().sum(1)
        else:
            scale = raw.abs().mean().expand_as(raw)
        if float((raw.abs() / scale).min()) >= margin:
            return seed
    raise RuntimeError("no well-conditioned seed found")


def make_pair(ocfg: orc.OracleConfig, batch: int, precision: str, seed: int = None):
    """(oracle state, engine, inputs) sharing bit-identical fp32 weights."""
    from magic_b200.engine import Engine
    if seed is None:
        seed = well_conditioned_seed(ocfg, batch, 0.02 if precision == "fp32" else 0.2)
    state = orc.OracleState(ocfg, seed=seed)
    eng = Engine(engine_config_from_oracle(ocfg, batch, precision, seed=seed))
    eng.set_params({k: v.to(torch.float32).numpy() for k, v in state.params.items() if k not in orc.DEAD_PARAMS})
    xl, xr, xk, y, eps = orc.synthetic_batch(batch, ocfg.latent, seed=seed + 1, n_input=ocfg.n_input)
    return state, eng, (xl, xr, xk, y, eps)
