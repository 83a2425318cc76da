"""`TangoEncoder` on the B200 engine: the host-side mirror of the reference's model class.

The reference's boundary for this path is the Python surface of an experiment directory:
`constants.FLAGS`, `vae.TangoEncoder(sess)`, and the five methods `main.py` calls.  This module re-creates that
surface -- same names, argument meaning, return tuples and error behaviour -- over the C-ABI library, so an
experiment directory switches backend with one flag (`FLAGS.BACKEND = 'b200'`, see INTEGRATION.md and
`experiments/`).  Nothing here computes: every number comes out of libmagic_b200.so.

Reference sites: constructor 8c/vae.py:22-66 (hyper-parameters), :89-147 (variables, connection_info/bias_info
prints from 8c/utils.py:304-309), partial_fit :379-405 (6-tuple variant 11a/vae.py:385-411), get_predictions
:407-408, transform :410-414, generate :416-428, reconstruct :430-434, metric dispatch :300-305.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .engine import Engine, EngineConfig


class DISTANCE_METRICS:                       # 8c/constants.py:5-7
    COSINE_DISTANCE = "Cosine distance"
    SQUARED_DIFFERENCE = "Squared difference"


class DeviceBatch:
    """Handle onto a batch that already lives in the engine's HBM buffers (made by the on-device generator).
    `partial_fit(batch, batch.labels)` consumes it without leaving the GPU; `.eval()` materialises numpy arrays
    for callers that insist on the reference's `images_batch.eval()` protocol (8c/main.py:44-45)."""

    def __init__(self, model: "TangoEncoder", seed: int, step: int):
        self.model, self.seed, self.step = model, seed, step
        self.labels = DeviceLabels(self)

    def _ensure_loaded(self) -> None:
        m = self.model
        if m._loaded_key != (self.seed, self.step):
            m.engine.generate_batch(self.seed, self.step)
            m._loaded_key = (self.seed, self.step)

    def eval(self) -> np.ndarray:
        self._ensure_loaded()
        return self.model.engine.export_batch()[0].cpu().numpy()

    def __array__(self, dtype=None, copy=None):
        a = self.eval()
        return a if dtype is None else a.astype(dtype)


class ResidentBatch(DeviceBatch):
    """A dataset batch that a loader call has already placed in the engine's input buffer
    (magic_b200/overlap_input.py: `device_batches`).  `labels` is the matching numpy array of overlap areas."""

    def __init__(self, model: "TangoEncoder", labels: np.ndarray):
        self.model, self.labels = model, labels
        self._token = object()
        model._loaded_key = self._token

    def _ensure_loaded(self) -> None:
        if self.model._loaded_key is not self._token:
            raise RuntimeError("this dataset batch is no longer resident: a later batch has replaced it")


class DeviceLabels:
    def __init__(self, batch: DeviceBatch):
        self.batch = batch

    def eval(self) -> np.ndarray:
        self.batch._ensure_loaded()
        return self.batch.model.engine.export_batch()[1].cpu().numpy()

    def __array__(self, dtype=None, copy=None):
        a = self.eval()
        return a if dtype is None else a.astype(dtype)


class TangoEncoder(object):
    """Drop-in for `vae.TangoEncoder` (8c/vae.py:13).  Experiment directories subclass it and override the class
    attributes exactly where the reference hard-codes them in `__init__` (8c/vae.py:30,36,49,56,63)."""

    encoder_neurons: Sequence[int] = (500, 500, 500, 500)     # 8c/vae.py:30
    decoder_neurons: Sequence[int] = (500, 500)               # 8c/vae.py:36
    latent_dimensions: int = 20                               # 8c/vae.py:49
    transfer_fct: str = "tanh"                                # 8c/vae.py:56 ; "elu" 11a/vae.py:58
    learning_rates: Tuple[float, float] = (1e-4, 1e-6)        # 8c/vae.py:63
    deformation_weight: float = 10.0                          # 8c/vae.py:287 ; 100 at 11a/vae.py:293
    reciprocal_distance: bool = False                         # 9a/vae.py:301,303 ; 11a/vae.py:306-311
    returns_distance: bool = False                            # 11a/vae.py:399-411: partial_fit returns 6 values

    def __init__(self, sess=None, flags=None, precision: Optional[str] = None, seed: int = 0,
                 batch_size: Optional[int] = None, rank: int = 0, world: int = 1, device=None, verbose: bool = True):
        self.sess = sess                                      # accepted and ignored: there is no session
        if flags is None:
            raise ValueError("TangoEncoder needs the experiment's FLAGS (constants.FLAGS)")
        self.flags = flags
        assert len(self.encoder_neurons) > 0                  # 8c/vae.py:31
        assert len(self.decoder_neurons) == 2                 # 8c/vae.py:37
        self.num_input_neurons = flags.IMAGE_SIZE * flags.IMAGE_SIZE * flags.NUM_LAYERS      # 8c/vae.py:45
        if flags.NUM_LAYERS != 3:
            raise ValueError("the B200 backend implements the three-stream model (FLAGS.ROTATE = True, NUM_LAYERS = 3)")
        self.batch_size = int(batch_size if batch_size is not None else flags.BATCH_SIZE)    # 8c/vae.py:66
        metric = getattr(flags, "DISTANCE_METRIC", DISTANCE_METRICS.COSINE_DISTANCE)
        if metric == DISTANCE_METRICS.COSINE_DISTANCE:
            m = "cosine"
        elif metric == DISTANCE_METRICS.SQUARED_DIFFERENCE:
            m = "sqdiff"
        else:
            raise ValueError("Invalid distance metric: " + str(metric))                      # 8c/vae.py:305
        precision = precision or getattr(flags, "PRECISION", "bf16")
        precision = {"tf32": "fp32", "fp32": "fp32", "bf16": "bf16"}[precision]
        D = flags.IMAGE_SIZE * flags.IMAGE_SIZE                                               # 8c/vae.py:81
        self.config = EngineConfig(
            D=D, encoder_neurons=tuple(self.encoder_neurons), decoder_neurons=tuple(self.decoder_neurons),
            latent=int(self.latent_dimensions), act=self.transfer_fct, metric=m, reciprocal=bool(self.reciprocal_distance),
            def_weight=float(self.deformation_weight), lr=tuple(self.learning_rates), precision=precision,
            batch=self.batch_size, rank=rank, world=world, seed=seed)
        self.engine = Engine(self.config, device=device)
        self._loaded_key = None
        self._gen_step = 0
        from . import overlap_input as _oi
        _oi.bind_model(self)                                   # inputs() handles made before the model now evaluate on it
        if verbose:                                            # the reference's build log, 8c/nohup.out:19-26
            for name, shape in self.engine.param_shapes.items():
                if len(shape) == 2:
                    print("Connecting network layer [" + str(shape[0]) + "->" + str(shape[1]) + "].")
                else:
                    print("Adding " + str(shape[0]) + " biases.")

    # ------------------------------------------------------------------ inputs
    def _load(self, X) -> None:
        if isinstance(X, DeviceBatch):
            X._ensure_loaded()
            return
        self._loaded_key = None
        if isinstance(X, torch.Tensor):
            self.engine.load_interleaved(X)
            return
        X = np.asarray(X)
        if X.shape != (self.batch_size, self.num_input_neurons):
            # static batch size, like the reference's placeholders and its eps of shape (batch_size, n_z), 8c/vae.py:367
            raise ValueError("Cannot feed value of shape %s for input of shape (%d, %d)" %
                             (X.shape, self.batch_size, self.num_input_neurons))
        self.engine.load_interleaved(X)

    def _labels(self, overlap_areas):
        if isinstance(overlap_areas, DeviceLabels):
            overlap_areas.batch._ensure_loaded()
            return None                                        # the engine uses the generated labels
        a = np.asarray(overlap_areas, dtype=np.float32).reshape(-1)
        if a.shape[0] != self.batch_size:
            raise ValueError("overlap_areas must have %d entries" % self.batch_size)
        return a

    def next_device_batch(self, seed: int = 0) -> DeviceBatch:
        """`overlap_input.inputs()` for the 'b200' backend: the next synthetic batch, produced on the GPU."""
        b = DeviceBatch(self, seed, self._gen_step)
        self._gen_step += 1
        return b

    # ------------------------------------------------------------------ the reference's methods
    def partial_fit(self, X, overlap_areas, eps=None):
        """Train on one mini-batch (8c/vae.py:379-405).  Returns (cost, training_loss, rec_loss, lat_loss,
        def_loss) as Python floats, plus distance[B] when the experiment is 10b/11a-style.  `eps` ([3,B,L],
        streams lock/rot/key) replaces the internal N(0,1) draws; parity tests use it."""
        labels = self._labels(overlap_areas)
        self._load(X)
        losses, dist = self.engine.step(labels=labels, eps3=eps, want_distance=self.returns_distance)
        # the cost is final after the forward pass: wait for that, not for the rest of the step (8c/main.py reads it every
        # step; the host then prepares the next call while the backward pass and both Adams are still running)
        out = tuple(float(v) for v in self.engine.read_losses())
        if self.returns_distance:
            return out + (dist.cpu().numpy(),)
        return out

    def get_predictions(self, X, overlap_areas=None, eps=None):
        """distance[B] on sampled latents (8c/vae.py:407-408)."""
        self._load(X)
        return self.engine.forward(_lib.FWD_DISTANCE, eps3=eps).cpu().numpy()

    def transform(self, X, overlap_areas=None):
        """z_mean[B, L] of the lock stream (8c/vae.py:410-414)."""
        self._load(X)
        return self.engine.forward(_lib.FWD_ZMEAN).cpu().numpy()

    def reconstruct(self, X, overlap_areas=None, eps=None):
        """x_reconstructed_mean[B, D] (8c/vae.py:430-434)."""
        self._load(X)
        return self.engine.forward(_lib.FWD_RECON, eps3=eps).cpu().numpy()

    def generate(self, z_mu=None):
        """Decoder-only forward from a fed z (8c/vae.py:416-428).  The reference feeds the `z` tensor itself, whose
        static shape is (batch_size, n_z): callers pass `[[x, y]] * vae.batch_size` (8c/main.py:155)."""
        if z_mu is None:
            z_mu = np.random.normal(size=(self.batch_size, self.latent_dimensions))
        z = np.asarray(z_mu, dtype=np.float32)
        if z.ndim == 1:
            z = np.tile(z[None, :], (self.batch_size, 1))
        if z.shape != (self.batch_size, self.latent_dimensions):
            raise ValueError("Cannot feed value of shape %s for z of shape (%d, %d)" %
                             (z.shape, self.batch_size, self.latent_dimensions))
        return self.engine.decode(z).cpu().numpy()

    # ------------------------------------------------------------------ additions (no precedent in 7c+)
    def get_params(self):
        return self.engine.get_params()

    def set_params(self, params) -> None:
        self.engine.set_params(params)

    def state_dict(self):
        return self.engine.state_dict()

    def load_state_dict(self, sd) -> None:
        self.engine.load_state_dict(sd)

    def close(self) -> None:
        self.engine.close()


def cosine_distance(a, b):
    """`cosine_distance(a, b)` of 8c/vae.py:438-452: both [B, L] arrays l2-normalised over axis 0 (the batch axis, as
    tf.nn.l2_normalize(x, 0) does) and the products summed over the latent axis -> [B].  Evaluated on the GPU by
    magic_cosine_distance (inside a step the latent kernels compute the same quantity); there is no host path."""
    import ctypes as C
    a = np.ascontiguousarray(a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else a, dtype=np.float32)
    b = np.ascontiguousarray(b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else b, dtype=np.float32)
    if a.ndim != 2 or a.shape != b.shape:
        raise ValueError("cosine_distance expects two arrays of the same shape [B, L]")
    out = np.empty(a.shape[0], dtype=np.float32)
    lib = _lib.load()
    _lib.check(lib.magic_cosine_distance(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), a.shape[0], a.shape[1],
                                         out.ctypes.data_as(C.c_void_p)))
    return out


class Session(object):
    """Stand-in for the `tf.Session` an experiment's main.py opens (8c/main.py:96, 11a/main.py:136) when
    FLAGS.BACKEND == 'b200': a context manager whose `run(fetches)` evaluates input handles JOINTLY -- images and
    labels of ONE batch, as `sess.run([images_batch, labels_batch])` does in 11a/main.py:65."""

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def run(self, fetches, feed_dict=None):
        single = not isinstance(fetches, (list, tuple))
        items = [fetches] if single else list(fetches)
        pipes = {id(f.pipeline): f.pipeline for f in items if hasattr(f, "pipeline")}
        for p in pipes.values():
            p.advance()                                   # one dequeue per run() call, shared by every fetch of it
        out = [f._current() if hasattr(f, "pipeline") else f for f in items]
        return out[0] if single else out

    def close(self):
        pass
