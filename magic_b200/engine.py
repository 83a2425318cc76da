"""Thin PyTorch host over the C-ABI handle: torch supplies device memory, streams and (for data parallel
start-up) the process group; every kernel lives in libmagic_b200.so."""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import MagicConfig, MagicError, MagicStats, check

LOSS_KEYS = ("cost", "training_loss", "rec_loss", "lat_loss", "def_loss")   # 8c/vae.py:405


@dataclass
class EngineConfig:
    """Hyper-parameters of TangoEncoder.__init__ (8c/vae.py:30-66) + batch/precision/DP settings."""
    D: int = 40000
    encoder_neurons: Sequence[int] = (500, 500, 500, 500)
    decoder_neurons: Sequence[int] = (500, 500)
    latent: int = 20
    act: str = "tanh"
    metric: str = "cosine"
    reciprocal: bool = False
    def_weight: float = 10.0
    lr: Tuple[float, float] = (1e-4, 1e-6)
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    precision: str = "bf16"          # "bf16" | "fp32"
    batch: int = 100                 # local (per-rank) batch
    batch_global: Optional[int] = None
    rank: int = 0
    world: int = 1
    seed: int = 0

    @staticmethod
    def experiment(name: str, **kw) -> "EngineConfig":
        """The BASELINE.json configs."""
        base = {
            "8c": dict(latent=20),
            "8d": dict(latent=200),
            "8e": dict(latent=2000),
            "9a": dict(latent=2, reciprocal=True),
            "11a": dict(encoder_neurons=(500, 500), latent=80, act="elu", metric="sqdiff", reciprocal=True,
                        def_weight=100.0, lr=(1e-6, 1e-8)),
        }[name]
        base.update(kw)
        return EngineConfig(**base)

    def to_c(self) -> MagicConfig:
        c = MagicConfig()
        c.D = int(self.D)
        enc = list(self.encoder_neurons)
        dec = list(self.decoder_neurons)
        assert len(enc) > 0                      # 8c/vae.py:31
        assert len(dec) == 2                     # 8c/vae.py:37
        if len(enc) > _lib.MAX_ENC_LAYERS:
            raise ValueError("at most %d encoder layers" % _lib.MAX_ENC_LAYERS)
        c.n_enc = len(enc)
        for i, w in enumerate(enc):
            c.enc[i] = int(w)
        c.dec[0], c.dec[1] = int(dec[0]), int(dec[1])
        c.L = int(self.latent)
        c.act = {"tanh": _lib.ACT_TANH, "elu": _lib.ACT_ELU}[self.act]
        if self.metric not in ("cosine", "sqdiff"):
            raise ValueError("Invalid distance metric: " + str(self.metric))       # 8c/vae.py:305
        c.metric = {"cosine": _lib.METRIC_COSINE, "sqdiff": _lib.METRIC_SQDIFF}[self.metric]
        c.reciprocal = int(bool(self.reciprocal))
        c.def_weight = float(self.def_weight)
        c.lr1, c.lr2 = float(self.lr[0]), float(self.lr[1])
        c.beta1, c.beta2, c.eps = float(self.beta1), float(self.beta2), float(self.eps)
        c.precision = {"bf16": _lib.PRECISION_BF16, "fp32": _lib.PRECISION_FP32}[self.precision]
        c.batch_local = int(self.batch)
        c.batch_global = int(self.batch_global if self.batch_global is not None else self.batch * self.world)
        c.rank, c.world = int(self.rank), int(self.world)
        c.seed = int(self.seed) & 0xFFFFFFFFFFFFFFFF
        return c


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class Engine:
    """One handle = one model replica on one GPU (cuda:<current device>), driven by one thread."""

    def __init__(self, cfg: EngineConfig, device: Optional[torch.device] = None):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise MagicError("magic_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
        self.cfg = cfg
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        torch.cuda.set_device(self.device)
        self.stream = torch.cuda.current_stream(self.device)
        self._h = C.c_void_p()
        ccfg = cfg.to_c()
        check(self.lib.magic_create(C.byref(ccfg), C.c_void_p(self.stream.cuda_stream), C.byref(self._h)))
        self.B, self.D, self.L = int(cfg.batch), int(cfg.D), int(cfg.latent)
        n = C.c_int32()
        check(self.lib.magic_param_count(self._h, C.byref(n)))
        self.param_names: List[str] = []
        self.param_shapes: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
        self.param_opts: Dict[str, int] = {}
        for i in range(n.value):
            name = C.c_char_p()
            shape = (C.c_int64 * 2)()
            opts = C.c_int32()
            check(self.lib.magic_param_info(self._h, i, C.byref(name), C.byref(shape), C.byref(opts)))
            nm = name.value.decode()
            self.param_names.append(nm)
            self.param_shapes[nm] = (shape[0],) if shape[1] == 0 else (shape[0], shape[1])
            self.param_opts[nm] = opts.value
        self._losses = torch.zeros(8, dtype=torch.float32, device=self.device)
        self._dist = torch.zeros(self.B, dtype=torch.float32, device=self.device)

    # ------------------------------------------------------------------ lifetime
    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.magic_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self) -> None:
        check(self.lib.magic_sync(self._h))

    def stats(self) -> MagicStats:
        s = MagicStats()
        check(self.lib.magic_get_stats(self._h, C.byref(s)))
        return s

    # ------------------------------------------------------------------ parameters / state
    def _index(self, name: str) -> int:
        return self.param_names.index(name)

    def get_param(self, name: str) -> np.ndarray:
        out = np.empty(self.param_shapes[name], dtype=np.float32)
        check(self.lib.magic_get_param(self._h, self._index(name), out.ctypes.data_as(C.c_void_p)))
        return out

    def set_param(self, name: str, value) -> None:
        a = np.ascontiguousarray(np.asarray(value, dtype=np.float32))
        if tuple(a.shape) != tuple(self.param_shapes[name]):
            raise ValueError("shape mismatch for %s: %s vs %s" % (name, a.shape, self.param_shapes[name]))
        check(self.lib.magic_set_param(self._h, self._index(name), a.ctypes.data_as(C.c_void_p)))

    def get_params(self) -> "OrderedDict[str, np.ndarray]":
        return OrderedDict((n, self.get_param(n)) for n in self.param_names)

    def set_params(self, params) -> None:
        for n in self.param_names:
            if n in params:
                v = params[n]
                self.set_param(n, v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v)

    def get_grad(self, which: int, name: str) -> np.ndarray:
        out = np.empty(self.param_shapes[name], dtype=np.float32)
        check(self.lib.magic_get_grad(self._h, which, self._index(name), out.ctypes.data_as(C.c_void_p)))
        return out

    def get_slot(self, which: int, slot: int, name: str) -> np.ndarray:
        out = np.empty(self.param_shapes[name], dtype=np.float32)
        check(self.lib.magic_get_slot(self._h, which, slot, self._index(name), out.ctypes.data_as(C.c_void_p)))
        return out

    def set_slot(self, which: int, slot: int, name: str, value) -> None:
        a = np.ascontiguousarray(np.asarray(value, dtype=np.float32))
        check(self.lib.magic_set_slot(self._h, which, slot, self._index(name), a.ctypes.data_as(C.c_void_p)))

    @property
    def step_count(self) -> int:
        t = C.c_int64()
        check(self.lib.magic_get_step(self._h, C.byref(t)))
        return t.value

    @step_count.setter
    def step_count(self, t: int) -> None:
        check(self.lib.magic_set_step(self._h, int(t)))

    def state_dict(self) -> dict:
        """Parameters + both optimisers' slots + step (no precedent in 7c+)."""
        sd = {"step": self.step_count, "params": self.get_params(), "slots": {}}
        for which in (1, 2):
            for n in self.param_names:
                if self.param_opts[n] & which:
                    sd["slots"][(which, "m", n)] = self.get_slot(which, _lib.SLOT_M, n)
                    sd["slots"][(which, "v", n)] = self.get_slot(which, _lib.SLOT_V, n)
        return sd

    def load_state_dict(self, sd: dict) -> None:
        self.set_params(sd["params"])
        for (which, kind, n), v in sd["slots"].items():
            self.set_slot(which, _lib.SLOT_M if kind == "m" else _lib.SLOT_V, n, v)
        self.step_count = sd["step"]

    # ------------------------------------------------------------------ inputs
    def _dev_f32(self, x, shape) -> torch.Tensor:
        t = x if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x))
        t = t.to(device=self.device, dtype=torch.float32, non_blocking=True).contiguous()
        if tuple(t.shape) != tuple(shape):
            raise ValueError("expected shape %s, got %s" % (tuple(shape), tuple(t.shape)))
        return t

    def load_interleaved(self, X) -> None:
        """X: [B, 3*D], channel-interleaved as main.py feeds it (8c/vae.py:166); float (any width) or uint8/bool."""
        probe = X if isinstance(X, torch.Tensor) else np.asarray(X)
        if probe.dtype in (torch.uint8, torch.bool, np.uint8, np.bool_):
            t = probe if isinstance(probe, torch.Tensor) else torch.as_tensor(probe)
            t = t.to(device=self.device, non_blocking=True).contiguous().view(torch.uint8)
            if tuple(t.shape) != (self.B, 3 * self.D):
                raise ValueError("expected shape %s, got %s" % ((self.B, 3 * self.D), tuple(t.shape)))
            check(self.lib.magic_load_interleaved_u8(self._h, C.c_void_p(t.data_ptr())))
        else:
            t = self._dev_f32(X, (self.B, 3 * self.D))
            check(self.lib.magic_load_interleaved(self._h, C.c_void_p(t.data_ptr())))
        self._keep = t

    def load_planar_u8(self, lock: torch.Tensor, rot: torch.Tensor, key: torch.Tensor) -> None:
        ts = [x.to(device=self.device, dtype=torch.uint8).contiguous() for x in (lock, rot, key)]
        for x in ts:
            if tuple(x.shape) != (self.B, self.D):
                raise ValueError("planes must be [B, D] uint8")
        check(self.lib.magic_load_planar_u8(self._h, *[C.c_void_p(x.data_ptr()) for x in ts]))
        self._keep = ts

    def load_pairs_u8(self, lock: torch.Tensor, key: torch.Tensor, seed: int, step: int) -> None:
        """Dataset pairs (uint8 [B, D]) -> the engine's packed input rows; the rotated lock is produced on the device."""
        ts = [x.to(device=self.device, dtype=torch.uint8).contiguous() for x in (lock, key)]
        for x in ts:
            if tuple(x.shape) != (self.B, self.D):
                raise ValueError("planes must be [B, D] uint8")
        check(self.lib.magic_load_pairs_u8(self._h, C.c_void_p(ts[0].data_ptr()), C.c_void_p(ts[1].data_ptr()),
                                           C.c_uint64(seed), C.c_uint64(step)))
        self._keep = ts

    @property
    def words_per_row(self) -> int:
        """32-bit words of one bit-packed plane row (pixel p = bit p & 31 of word p >> 5)."""
        return (self.D + 31) // 32

    def load_pairs_bits(self, lock_bits: torch.Tensor, key_bits: torch.Tensor, seed: int, step: int) -> None:
        """Bit-packed dataset pairs already on the device (int32/uint32 [B, words_per_row]) -> packed input rows."""
        ts = [x.to(device=self.device).contiguous() for x in (lock_bits, key_bits)]
        for x in ts:
            if x.dim() != 2 or x.shape[0] != self.B or x.shape[1] < self.words_per_row or x.element_size() != 4:
                raise ValueError("packed planes must be [B, >= %d] 32-bit words" % self.words_per_row)
        check(self.lib.magic_load_pairs_bits(self._h, C.c_void_p(ts[0].data_ptr()), C.c_void_p(ts[1].data_ptr()),
                                             C.c_int32(ts[0].shape[1]), C.c_uint64(seed), C.c_uint64(step)))
        self._keep = ts

    def feed_pairs_bits(self, lock_bits, key_bits, labels, seed: int, step: int, activate: bool = True) -> None:
        """Bit-packed pairs in HOST memory (numpy or CPU torch, 32-bit words [B, words_per_row]; pinned keeps the
        copy asynchronous) -> packed input rows, on the engine's input stream.  activate=False leaves the batch staged
        until swap_batch().  The arrays must not change until the step that consumes them has been waited for."""
        arrs = []
        for x in (lock_bits, key_bits):
            x = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x))
            if x.is_cuda or x.dim() != 2 or x.shape[0] != self.B or x.shape[1] < self.words_per_row \
                    or x.element_size() != 4 or not x.is_contiguous():
                raise ValueError("packed planes must be contiguous host arrays [B, >= %d] of 32-bit words" % self.words_per_row)
            arrs.append(x)
        lab = None
        if labels is not None:
            lab = labels if isinstance(labels, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(labels, dtype=np.float32))
            if lab.is_cuda or lab.dtype != torch.float32 or tuple(lab.shape) != (self.B,) or not lab.is_contiguous():
                raise ValueError("labels must be a contiguous host float32 [B]")
        check(self.lib.magic_feed_pairs_bits_host(
            self._h, C.c_void_p(arrs[0].data_ptr()), C.c_void_p(arrs[1].data_ptr()),
            C.c_void_p(lab.data_ptr()) if lab is not None else None, C.c_int32(arrs[0].shape[1]),
            C.c_uint64(seed), C.c_uint64(step), C.c_int32(1 if activate else 0)))
        self._keep_feed = (getattr(self, "_keep_feed", ()) + ((arrs, lab),))[-2:]     # the two batches in flight

    def feed_pairs_bits_gather(self, lock_all: torch.Tensor, key_all: torch.Tensor, idx: torch.Tensor, labels, seed: int,
                               step: int, activate: bool = True) -> None:
        """Rows `idx` ([B] int32, host) of a packed data set that lives in PINNED host memory ([n, words] int32 tensors,
        words % 4 == 0) -> packed input rows; the gather runs on the GPU, reading the rows over PCIe
        (magic_feed_pairs_bits_gather).  idx / labels must not change until the step that consumes them has been waited for."""
        for x in (lock_all, key_all):
            if not isinstance(x, torch.Tensor) or x.is_cuda or not x.is_pinned() or x.dim() != 2 or x.element_size() != 4 \
                    or not x.is_contiguous() or x.shape != lock_all.shape or x.shape[1] < self.words_per_row or x.shape[1] % 4:
                raise ValueError("the data set must be two pinned, contiguous host tensors [n, >= %d words, a multiple of 4] "
                                 "of 32-bit words" % self.words_per_row)
        if idx.is_cuda or idx.dtype != torch.int32 or tuple(idx.shape) != (self.B,) or not idx.is_contiguous():
            raise ValueError("idx must be a contiguous host int32 [B]")
        lab = None
        if labels is not None:
            lab = labels if isinstance(labels, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(labels, dtype=np.float32))
            if lab.is_cuda or lab.dtype != torch.float32 or tuple(lab.shape) != (self.B,) or not lab.is_contiguous():
                raise ValueError("labels must be a contiguous host float32 [B]")
        check(self.lib.magic_feed_pairs_bits_gather(
            self._h, C.c_void_p(lock_all.data_ptr()), C.c_void_p(key_all.data_ptr()), C.c_int32(lock_all.shape[1]),
            C.c_int64(lock_all.shape[0]), C.c_void_p(idx.data_ptr()), C.c_void_p(lab.data_ptr()) if lab is not None else None,
            C.c_uint64(seed), C.c_uint64(step), C.c_int32(1 if activate else 0)))
        self._keep_feed = (getattr(self, "_keep_feed", ()) + (((lock_all, key_all, idx), lab),))[-2:]

    def generate_batch(self, seed: int, step: int) -> None:
        check(self.lib.magic_generate_batch(self._h, C.c_uint64(seed), C.c_uint64(step)))

    def export_batch(self) -> Tuple[torch.Tensor, torch.Tensor]:
        x = torch.empty(self.B, 3 * self.D, dtype=torch.float32, device=self.device)
        y = torch.empty(self.B, dtype=torch.float32, device=self.device)
        check(self.lib.magic_export_batch(self._h, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr())))
        return x, y

    def set_option(self, name: str, value: int) -> None:
        """Run-time switches of the engine: "skip_zero_stages" (default 0; takes effect with the next loaded batch); see
        include/magic_b200.h."""
        check(self.lib.magic_set_option(self._h, name.encode(), int(value)))

    def input_density(self) -> Tuple[float, float]:
        """Fraction of 128 x 64 first-layer operand stages of the loaded batch that hold a set pixel:
        (forward GEMM, weight-gradient GEMM).  Synchronises."""
        out = (C.c_double * 2)()
        check(self.lib.magic_input_density(self._h, out))
        return float(out[0]), float(out[1])

    def swap_batch(self) -> None:
        """Make the previously loaded batch (still resident in HBM) the loaded one again."""
        check(self.lib.magic_swap_batch(self._h))

    def export_rotation(self) -> torch.Tensor:
        """[B,2] int32: Q14 (cos, sin) of the rotation the generator applied to each lock."""
        r = torch.empty(self.B, 2, dtype=torch.int32, device=self.device)
        check(self.lib.magic_export_rotation(self._h, C.c_void_p(r.data_ptr())))
        return r

    def export_latents(self) -> torch.Tensor:
        """[3, B, L] float32: the sampled z of the last step / forward call, streams lock, rot, key."""
        z = torch.empty(3, self.B, self.L, dtype=torch.float32, device=self.device)
        check(self.lib.magic_export_latents(self._h, C.c_void_p(z.data_ptr())))
        return z

    # ------------------------------------------------------------------ step / inference
    def step(self, labels=None, eps3=None, apply: bool = True, want_distance: bool = False):
        """One optimiser step on the loaded batch; returns (losses[5] device tensor, distance or None).
        Asynchronous: read the tensors (or call sync()) to wait."""
        lab = None if labels is None else self._dev_f32(labels, (self.B,))
        eps = None if eps3 is None else self._dev_f32(eps3, (3, self.B, self.L))
        fn = self.lib.magic_step if apply else self.lib.magic_step_grads_only
        check(fn(self._h, C.c_void_p(_ptr(lab)), C.c_void_p(_ptr(eps)), C.c_void_p(self._losses.data_ptr()),
                 C.c_void_p(self._dist.data_ptr()) if want_distance else None))
        self._keep2 = (lab, eps)
        return self._losses[:5], (self._dist if want_distance else None)

    def read_losses(self) -> np.ndarray:
        """The last step's five losses as a host array (magic_read_losses): waits for the point of the step where they are
        final -- the end of the forward pass on one rank -- not for the backward pass and the update behind it."""
        out = np.empty(5, dtype=np.float32)
        check(self.lib.magic_read_losses(self._h, out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    def partial_fit_host(self, X: np.ndarray, labels: np.ndarray, eps3: Optional[np.ndarray] = None,
                         want_distance: bool = False):
        """Host-buffer partial_fit through magic_partial_fit_host (copies + step + sync)."""
        u8 = X.dtype in (np.uint8, np.bool_)
        X = np.ascontiguousarray(X) if u8 else np.ascontiguousarray(X, dtype=np.float32)
        labels = np.ascontiguousarray(labels, dtype=np.float32)
        if X.shape != (self.B, 3 * self.D) or labels.shape != (self.B,):
            raise ValueError("partial_fit expects X[%d,%d] and overlap_areas[%d]" % (self.B, 3 * self.D, self.B))
        losses = np.empty(5, dtype=np.float32)
        dist = np.empty(self.B, dtype=np.float32) if want_distance else None
        e = None if eps3 is None else np.ascontiguousarray(eps3, dtype=np.float32)
        fn = self.lib.magic_partial_fit_host_u8 if u8 else self.lib.magic_partial_fit_host
        check(fn(
            self._h, X.ctypes.data_as(C.c_void_p), labels.ctypes.data_as(C.c_void_p),
            None if e is None else e.ctypes.data_as(C.c_void_p), losses.ctypes.data_as(C.c_void_p),
            None if dist is None else dist.ctypes.data_as(C.c_void_p)))
        return losses, dist

    def forward(self, what: int, eps3=None) -> torch.Tensor:
        shape = {_lib.FWD_DISTANCE: (self.B,), _lib.FWD_ZMEAN: (self.B, self.L), _lib.FWD_RECON: (self.B, self.D)}[what]
        out = torch.empty(shape, dtype=torch.float32, device=self.device)
        eps = None if eps3 is None else self._dev_f32(eps3, (3, self.B, self.L))
        check(self.lib.magic_forward(self._h, what, C.c_void_p(_ptr(eps)), C.c_void_p(out.data_ptr())))
        self._keep2 = eps
        return out

    def decode(self, z) -> torch.Tensor:
        zt = self._dev_f32(z, (self.B, self.L))
        out = torch.empty(self.B, self.D, dtype=torch.float32, device=self.device)
        check(self.lib.magic_decode(self._h, C.c_void_p(zt.data_ptr()), C.c_void_p(out.data_ptr())))
        self._keep2 = zt
        return out

    # ------------------------------------------------------------------ profiling
    def profile(self, on: bool) -> None:
        check(self.lib.magic_profile_enable(self._h, int(on)))

    def profile_report(self) -> Dict[str, Tuple[float, int, float]]:
        """tag -> (ms summed, launches, tensor flops summed) since the last report."""
        buf = C.create_string_buffer(1 << 16)
        check(self.lib.magic_profile_report(self._h, buf, len(buf)))
        out = {}
        for line in buf.value.decode().splitlines():
            tag, ms, n, fl = line.split()
            out[tag] = (float(ms), int(n), float(fl))
        return out

    # ------------------------------------------------------------------ data parallel
    def comm_init_from_process_group(self) -> None:
        """Create the library's NCCL communicator: rank 0 makes the unique id, torch.distributed carries it."""
        import torch.distributed as dist
        ident = torch.zeros(128, dtype=torch.uint8)
        if dist.get_rank() == 0:
            buf = (C.c_char * 128)()
            check(self.lib.magic_comm_unique_id(buf))
            ident = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        if dist.get_backend() == "nccl":
            ident = ident.to(self.device)
        dist.broadcast(ident, src=0)
        raw = bytes(ident.cpu().numpy().tobytes())
        check(self.lib.magic_comm_init(self._h, C.c_char_p(raw)))
        # peer-memory handles for the fused reduce-scatter / all-gather (include/magic_b200.h: magic_comm_export_peer)
        blob = (C.c_char * 256)()
        check(self.lib.magic_comm_export_peer(self._h, blob))
        mine = torch.frombuffer(bytearray(blob.raw), dtype=torch.uint8).clone()
        if dist.get_backend() == "nccl":
            mine = mine.to(self.device)
        allb = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
        dist.all_gather(allb, mine)
        joined = b"".join(bytes(t.cpu().numpy().tobytes()) for t in allb)
        check(self.lib.magic_comm_import_peers(self._h, C.c_char_p(joined)))
