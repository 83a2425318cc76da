"""CPU oracle for the magic autoencoder hot path  --  TEST INFRASTRUCTURE ONLY.

This file is a float64 (or float32, for the timed CPU baseline) PyTorch-CPU restatement
of the reference's `TangoEncoder` graph.  It is the *checker* for the CUDA engine; only
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import it.  The product path (`magic_b200/`) never does.

PARITY UNPINNED: the reference is Python-2 / TensorFlow-1.x and cannot be imported or
run in this environment (no tensorflow, no python2, no network), and it ships no test,
golden vector or fixture of the step.  The oracle is therefore pinned only by
  KAT-1  the label fixture 2a/OVERLAP_AREAS reproducing README.md:44-47,
  KAT-2  step-0 loss statistics printed in {8c,8d,8e}/nohup.out:27,
  KAT-4  float64 finite differences vs autograd, tf_literal vs stable rec-loss,
all encoded in tests/test_oracle.py.

Reference sites restated here (paths relative to /root/reference):
  layout / de-interleave ........ 8c/vae.py:166-186, 8c/utils.py:315-316
  encoder x3, shared weights .... 8c/vae.py:203-214, 329-375
  decoder ....................... 8c/vae.py:221-226
  rec / KL / def / cost ......... 8c/vae.py:260-292   (def weight 100: 11a/vae.py:293)
  metric distance ............... 8c/vae.py:300-307, 438-452; reciprocal 9a/vae.py:301-303,
                                  11a/vae.py:306-311
  two Adam optimisers ........... 8c/vae.py:314-321  (TF1 ApplyAdam formula)
  xavier init ................... 8c/utils.py:346-353
  variable creation order ....... 8c/vae.py:89-147
"""
from __future__ import annotations

import math
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch

IMAGE_SIZE = 200            # 8c/constants.py:32
D_PIXELS = IMAGE_SIZE * IMAGE_SIZE


@dataclass
class OracleConfig:
    """Hyper-parameters that the reference hard-codes in vae.py / constants.py."""
    encoder_neurons: List[int] = field(default_factory=lambda: [500, 500, 500, 500])  # 8c/vae.py:30
    decoder_neurons: List[int] = field(default_factory=lambda: [500, 500])            # 8c/vae.py:36
    latent: int = 20                         # 8c/vae.py:49
    act: str = "tanh"                        # 8c/vae.py:56 ; "elu" 11a/vae.py:58
    metric: str = "cosine"                   # 8c/constants.py:60 ; "sqdiff" 11a/constants.py:60
    reciprocal: bool = False                 # 9a/vae.py:301,303
    def_weight: float = 10.0                 # 8c/vae.py:287 ; 100 at 11a/vae.py:293
    lr: Tuple[float, float] = (1e-4, 1e-6)   # 8c/vae.py:63
    beta1: float = 0.9                       # tf.train.AdamOptimizer defaults
    beta2: float = 0.999
    eps: float = 1e-8
    n_input: int = D_PIXELS                  # 8c/vae.py:81

    @staticmethod
    def experiment(name: str) -> "OracleConfig":
        """The five BASELINE.json configs."""
        if name == "8c":
            return OracleConfig(latent=20)
        if name == "8d":
            return OracleConfig(latent=200)
        if name == "8e":
            return OracleConfig(latent=2000)
        if name == "9a":
            return OracleConfig(latent=2, reciprocal=True)
        if name == "11a":
            return OracleConfig(encoder_neurons=[500, 500], latent=80, act="elu",
                                metric="sqdiff", reciprocal=True, def_weight=100.0,
                                lr=(1e-6, 1e-8))
        raise KeyError(name)


# --------------------------------------------------------------------------- parameters

def param_shapes(cfg: OracleConfig) -> "OrderedDict[str, Tuple[int, ...]]":
    """Names and shapes in the reference's creation order (8c/vae.py:89-147)."""
    enc, dec, L, D = cfg.encoder_neurons, cfg.decoder_neurons, cfg.latent, cfg.n_input
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    s["encoder_h0_weights"] = (D, enc[0])
    for i in range(len(enc) - 1):
        s["encoder_h%d_weights" % (i + 1)] = (enc[i], enc[i + 1])
    s["encoder_out_mean_weights"] = (enc[-1], L)
    s["encoder_out_log_sigma_weights"] = (enc[-1], L)
    s["encoder_b0_biases"] = (enc[0],)
    for i in range(1, len(enc)):
        s["encoder_b%d_biases" % i] = (enc[i],)
    s["encoder_out_mean_biases"] = (L,)
    s["encoder_out_log_sigma_biases"] = (L,)
    s["decoder_h1_weights"] = (L, dec[0])
    s["decoder_h2_weights"] = (dec[0], dec[1])
    s["decoder_out_mean_weights"] = (dec[1], D)
    s["decoder_out_log_sigma_weights"] = (dec[1], D)        # dead: 8c/vae.py:141-142
    s["decoder_b1_biases"] = (dec[0],)
    s["decoder_b2_biases"] = (dec[1],)
    s["decoder_out_mean_biases"] = (D,)
    s["decoder_out_log_sigma_biases"] = (D,)                # dead: 8c/vae.py:147
    return s


DEAD_PARAMS = ("decoder_out_log_sigma_weights", "decoder_out_log_sigma_biases")


def is_encoder_param(name: str) -> bool:
    return name.startswith("encoder_")


def init_params(cfg: OracleConfig, seed: int = 0, dtype=torch.float64) -> "OrderedDict[str, torch.Tensor]":
    """Xavier-uniform weights (8c/utils.py:346-353), zero biases (8c/vae.py:119).

    Values are drawn in float32 (as the reference does) and then widened, so the CUDA
    engine and the oracle can start from bit-identical fp32 weights."""
    g = torch.Generator().manual_seed(seed)
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for name, shp in param_shapes(cfg).items():
        if len(shp) == 2:
            bound = math.sqrt(6.0 / (shp[0] + shp[1]))
            w = (torch.rand(shp, generator=g, dtype=torch.float32) * 2.0 - 1.0) * bound
        else:
            w = torch.zeros(shp, dtype=torch.float32)
        out[name] = w.to(dtype)
    return out


# --------------------------------------------------------------------------- forward

def deinterleave(X: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """[B, 120000] channel-last -> lock, rot, key planes [B, 40000] (8c/vae.py:166-179)."""
    B = X.shape[0]
    x = X.reshape(B, -1, 3)
    return x[:, :, 0].contiguous(), x[:, :, 1].contiguous(), x[:, :, 2].contiguous()


def _act(cfg: OracleConfig, x: torch.Tensor) -> torch.Tensor:
    if cfg.act == "tanh":
        return torch.tanh(x)
    if cfg.act == "elu":
        return torch.nn.functional.elu(x)
    raise ValueError("Invalid activation: " + str(cfg.act))


def encode(cfg: OracleConfig, p: Dict[str, torch.Tensor], x: torch.Tensor, eps: torch.Tensor):
    """get_latent_representation (8c/vae.py:329-375): returns (z_mean, z_log_sigma_sq, z)."""
    h = _act(cfg, x @ p["encoder_h0_weights"] + p["encoder_b0_biases"])
    for i in range(1, len(cfg.encoder_neurons)):
        h = _act(cfg, h @ p["encoder_h%d_weights" % i] + p["encoder_b%d_biases" % i])
    mu = h @ p["encoder_out_mean_weights"] + p["encoder_out_mean_biases"]
    lv = h @ p["encoder_out_log_sigma_weights"] + p["encoder_out_log_sigma_biases"]
    z = mu + torch.sqrt(torch.exp(lv)) * eps
    return mu, lv, z


def decode_logits(cfg: OracleConfig, p: Dict[str, torch.Tensor], z: torch.Tensor) -> torch.Tensor:
    """Decoder up to the pre-sigmoid activation (8c/vae.py:221-226)."""
    d1 = _act(cfg, z @ p["decoder_h1_weights"] + p["decoder_b1_biases"])
    d2 = _act(cfg, d1 @ p["decoder_h2_weights"] + p["decoder_b2_biases"])
    return d2 @ p["decoder_out_mean_weights"] + p["decoder_out_mean_biases"]


def cosine_distance(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """8c/vae.py:438-452.  tf.nn.l2_normalize(x, 0): x * rsqrt(max(sum_over_axis0(x^2), 1e-12)),
    i.e. every latent COLUMN is normalised over the BATCH axis; then summed over features."""
    na = a * torch.rsqrt(torch.clamp((a * a).sum(0, keepdim=True), min=1e-12))
    nb = b * torch.rsqrt(torch.clamp((b * b).sum(0, keepdim=True), min=1e-12))
    return (na * nb).sum(1)


def forward(cfg: OracleConfig, p: Dict[str, torch.Tensor],
            x_lock: torch.Tensor, x_rot: torch.Tensor, x_key: torch.Tensor,
            labels: torch.Tensor, eps3: torch.Tensor, rec_mode: str = "stable") -> Dict[str, torch.Tensor]:
    """The whole loss graph (8c/vae.py:203-307).  eps3 is [3, B, L] in stream order
    (lock, rot, key): three independent draws, 8c/vae.py:367.

    rec_mode "tf_literal" evaluates log(p^x * (1-p)^(1-x)) exactly as 8c/vae.py:260-261;
    "stable" evaluates the algebraically equal softplus(a) - x*a."""
    mu, lv, z = encode(cfg, p, x_lock, eps3[0])
    _, _, z_rot = encode(cfg, p, x_rot, eps3[1])
    _, _, z_key = encode(cfg, p, x_key, eps3[2])
    a = decode_logits(cfg, p, z)
    if rec_mode == "tf_literal":
        pr = torch.sigmoid(a)
        rec_b = -torch.log(torch.pow(pr, x_lock) * torch.pow(1.0 - pr, 1.0 - x_lock)).sum(1)
    elif rec_mode == "stable":
        rec_b = (torch.nn.functional.softplus(a) - x_lock * a).sum(1)
    else:
        raise ValueError(rec_mode)
    kl_b = -0.5 * (1.0 + lv - mu * mu - torch.exp(lv)).sum(1)            # 8c/vae.py:275-277
    def_b = cfg.def_weight * ((z - z_rot) ** 2).sum(1)                   # 8c/vae.py:287
    cost = (rec_b + kl_b + def_b).mean()                                 # 8c/vae.py:292
    if cfg.metric == "cosine":
        dist = cosine_distance(z, z_key)
    elif cfg.metric == "sqdiff":
        dist = ((z - z_key) ** 2).sum(1)
    else:
        raise ValueError("Invalid distance metric: " + str(cfg.metric))  # 8c/vae.py:305
    if cfg.reciprocal:
        dist = 1.0 / dist                                                # 9a/vae.py:301,303
    training_loss = ((dist - labels) ** 2).mean()                        # 8c/vae.py:307
    return dict(cost=cost, training_loss=training_loss, rec_loss=rec_b.mean(), lat_loss=kl_b.mean(),
                def_loss=def_b.mean(), distance=dist, z_mean=mu, z_log_sigma_sq=lv, z=z,
                z_rot=z_rot, z_key=z_key, logits=a)


LOSS_KEYS = ("cost", "training_loss", "rec_loss", "lat_loss", "def_loss")   # partial_fit order, 8c/vae.py:405


# --------------------------------------------------------------------------- training step

class OracleState:
    """Parameters plus the two optimisers' slots (m, v, beta powers), SURVEY 3.3 / A12."""

    def __init__(self, cfg: OracleConfig, params: Optional[Dict[str, torch.Tensor]] = None,
                 seed: int = 0, dtype=torch.float64):
        self.cfg = cfg
        self.dtype = dtype
        src = params if params is not None else init_params(cfg, seed, dtype)
        self.params: "OrderedDict[str, torch.Tensor]" = OrderedDict(
            (k, torch.as_tensor(v).detach().clone().to(dtype)) for k, v in src.items())
        self.names1 = [k for k in self.params if k not in DEAD_PARAMS]            # optimiser 1: cost
        self.names2 = [k for k in self.params if is_encoder_param(k)]             # optimiser 2: metric
        self.m = [{k: torch.zeros_like(self.params[k]) for k in self.names1},
                  {k: torch.zeros_like(self.params[k]) for k in self.names2}]
        self.v = [{k: torch.zeros_like(self.params[k]) for k in self.names1},
                  {k: torch.zeros_like(self.params[k]) for k in self.names2}]
        self.t = 0

    def grads(self, x_lock, x_rot, x_key, labels, eps3, rec_mode="stable"):
        """Both gradients at the pre-step weights: g1 = d cost, g2 = d training_loss."""
        leaves = OrderedDict((k, v.detach().clone().requires_grad_(True)) for k, v in self.params.items())
        out = forward(self.cfg, leaves, x_lock, x_rot, x_key, labels, eps3, rec_mode)
        g1 = torch.autograd.grad(out["cost"], [leaves[k] for k in self.names1], retain_graph=True)
        g2 = torch.autograd.grad(out["training_loss"], [leaves[k] for k in self.names2])
        res = {k: (v.detach() if isinstance(v, torch.Tensor) else v) for k, v in out.items()}
        return res, dict(zip(self.names1, g1)), dict(zip(self.names2, g2))

    def apply(self, g1: Dict[str, torch.Tensor], g2: Dict[str, torch.Tensor]) -> None:
        """TF1 ApplyAdam, twice:  W <- W - d1 - d2  (updates commute)."""
        cfg = self.cfg
        self.t += 1
        for which, (names, g) in enumerate(((self.names1, g1), (self.names2, g2))):
            lr_t = cfg.lr[which] * math.sqrt(1.0 - cfg.beta2 ** self.t) / (1.0 - cfg.beta1 ** self.t)
            for k in names:
                m, v = self.m[which][k], self.v[which][k]
                m.mul_(cfg.beta1).add_(g[k], alpha=1.0 - cfg.beta1)
                v.mul_(cfg.beta2).addcmul_(g[k], g[k], value=1.0 - cfg.beta2)
                self.params[k].sub_(lr_t * m / (torch.sqrt(v) + cfg.eps))

    def step(self, x_lock, x_rot, x_key, labels, eps3, rec_mode="stable"):
        out, g1, g2 = self.grads(x_lock, x_rot, x_key, labels, eps3, rec_mode)
        self.apply(g1, g2)
        return out, g1, g2

    def partial_fit(self, X: torch.Tensor, overlap_areas: torch.Tensor, eps3: torch.Tensor):
        """Same signature/return as TangoEncoder.partial_fit (8c/vae.py:379-405)."""
        xl, xr, xk = deinterleave(torch.as_tensor(X).to(self.dtype))
        out, _, _ = self.step(xl, xr, xk, torch.as_tensor(overlap_areas).to(self.dtype), eps3.to(self.dtype))
        return tuple(float(out[k]) for k in LOSS_KEYS)


# --------------------------------------------------------------------------- synthetic inputs

def synthetic_batch(B: int, L: int, seed: int = 0, dtype=torch.float64, n_input: int = D_PIXELS):
    """Small CPU generator used by tests: random filled discs/boxes on the 200x200 canvas with
    the fill fraction of the real data (~8.5 %), a rotated copy, a second shape,
    integer-valued labels in the range of 2a/OVERLAP_AREAS, eps ~ N(0,1) [3,B,L]."""
    g = torch.Generator().manual_seed(seed)
    side = int(round(math.sqrt(n_input)))
    yy, xx = torch.meshgrid(torch.arange(side), torch.arange(side), indexing="ij")

    def blob():
        cx, cy = (torch.rand(2, generator=g) * 0.5 + 0.25) * side
        r = (0.10 + 0.12 * torch.rand(1, generator=g)) * side
        if torch.rand(1, generator=g) < 0.5:
            m = ((xx - cx) ** 2 + (yy - cy) ** 2) <= r * r
        else:
            m = ((xx - cx).abs() <= r) & ((yy - cy).abs() <= 0.6 * r)
        return m.reshape(-1)

    lock = torch.stack([blob() for _ in range(B)])
    key = torch.stack([blob() for _ in range(B)])
    rot = torch.stack([lock[i].reshape(side, side).rot90(int(torch.randint(1, 4, (1,), generator=g))).reshape(-1)
                       for i in range(B)])
    labels = torch.randint(296, 6427, (B,), generator=g).to(dtype)
    eps3 = torch.randn(3, B, L, generator=g, dtype=torch.float64).to(dtype)
    return lock.to(dtype), rot.to(dtype), key.to(dtype), labels, eps3


def interleave(x_lock, x_rot, x_key) -> torch.Tensor:
    """Inverse of deinterleave: the [B,120000] tensor main.py feeds (8c/overlap_input.py:195,121)."""
    return torch.stack([x_lock, x_rot, x_key], dim=2).reshape(x_lock.shape[0], -1)
