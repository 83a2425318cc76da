"""Row N2: the overlap dataset path -- `{N}_L.png` / `{N}_K.png` pairs + OVERLAP_AREAS -> device batches.

Host-side mirror of the reference's input pipeline for 8c-11a (8c/overlap_input.py):

  * `read_png`      <- `decode_mshapes` (8c/overlap_input.py:41-73): whole file -> uint8 image.  The reference calls
                       tf.image.decode_png; here a small pure-Python decoder (zlib inflate + numpy un-filtering) for
                       the non-interlaced 8-bit grey / grey+alpha / RGB / RGBA / palette files the dataset contains.
  * `OverlapDataset`<- `_inputs_with_rotation` (8c/overlap_input.py:127-211): file lists `%d_L.png` / `%d_K.png`
                       (lines 158-165), labels from the OVERLAP_AREAS pickle (lines 146-149), pixels normalised to
                       {0,1} (`normalize=True`, line 110-112; the files hold {0,255}).
  * `batches`       <- `tf.train.shuffle_batch` (8c/overlap_input.py:316-322) + the JOINT dequeue of images and
                       labels that 10a/main.py:64 introduced (8c-9b dequeue them separately and pair a batch of
                       images with another batch's labels).  Shuffling is a fresh permutation per
                       epoch; the last ragged batch of an epoch is dropped (the graph has a static batch size,
                       8c/vae.py:367).
  * the random rotation of the lock (8c/overlap_input.py:181-182, 8c/utils.py:329-340) happens on the device:
                       `magic_load_pairs_u8` / `magic_load_pairs_bits` (include/magic_b200.h).
  * `device_batches`<- the QueueRunner threads + the two `.eval()` dequeues of the training loop (8c/main.py:40-45):
                       the dataset is kept bit-packed (10 KB per pair), batch n+1 is gathered into pinned memory and
                       sent with `magic_feed_pairs_bits_host` while step n runs.

Nothing here touches TensorFlow queues, threads or summaries (out of scope)."""
from __future__ import annotations

import io
import os
import pickle
import struct
import zipfile
import zlib
from typing import Iterator, Optional, Sequence, Tuple, Union

import numpy as np

_PNG_MAGIC = b"\x89PNG\r\n\x1a\n"
_CHANNELS = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}      # colour type -> samples per pixel


class PngError(ValueError):
    pass


def _unfilter(raw: np.ndarray, height: int, stride: int, bpp: int) -> np.ndarray:
    """PNG filter types 0-4 (None, Sub, Up, Average, Paeth), one filter byte per scanline."""
    out = np.zeros((height, stride), dtype=np.uint8)
    prev = np.zeros(stride, dtype=np.int32)
    pos = 0
    for y in range(height):
        ft = int(raw[pos])
        line = raw[pos + 1:pos + 1 + stride].astype(np.int32)
        pos += 1 + stride
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        elif ft in (1, 3, 4):
            cur = np.zeros(stride, dtype=np.int32)
            for x in range(stride):               # sequential along the line by definition of these filters
                a = cur[x - bpp] if x >= bpp else 0
                b = prev[x]
                if ft == 1:
                    pred = a
                elif ft == 3:
                    pred = (a + b) >> 1
                else:
                    c = prev[x - bpp] if x >= bpp else 0
                    p = a + b - c
                    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
                    pred = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[x] = (line[x] + pred) & 255
        else:
            raise PngError("bad PNG filter type %d" % ft)
        out[y] = cur
        prev = cur
    return out


def read_png(src: Union[str, bytes, os.PathLike]) -> np.ndarray:
    """Decode a non-interlaced 8-bit PNG into uint8 [H, W] (grey / palette index expanded to luminance-free first
    channel semantics the dataset needs: for colour images the FIRST channel is returned, as the reference reshapes
    the decoded tensor to [IMAGE_SIZE, IMAGE_SIZE, 1], 8c/overlap_input.py:68)."""
    data = src if isinstance(src, (bytes, bytearray)) else open(src, "rb").read()
    if data[:8] != _PNG_MAGIC:
        raise PngError("not a PNG file")
    pos, idat, palette = 8, [], None
    width = height = depth = ctype = interlace = None
    while pos + 8 <= len(data):
        length, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + length]
        if len(body) != length:
            raise PngError("truncated PNG chunk")
        crc = struct.unpack(">I", data[pos + 8 + length:pos + 12 + length])[0]
        if zlib.crc32(tag + body) & 0xFFFFFFFF != crc:
            raise PngError("PNG chunk CRC mismatch in %r" % tag)
        pos += 12 + length
        if tag == b"IHDR":
            width, height, depth, ctype, _comp, _filt, interlace = struct.unpack(">IIBBBBB", body)
        elif tag == b"PLTE":
            palette = np.frombuffer(body, dtype=np.uint8).reshape(-1, 3)
        elif tag == b"IDAT":
            idat.append(body)
        elif tag == b"IEND":
            break
    if width is None or not idat:
        raise PngError("PNG without IHDR/IDAT")
    if depth != 8 or interlace != 0 or ctype not in _CHANNELS:
        raise PngError("unsupported PNG (depth %s, colour type %s, interlace %s)" % (depth, ctype, interlace))
    ch = _CHANNELS[ctype]
    stride = width * ch
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), dtype=np.uint8)
    if raw.size != height * (stride + 1):
        raise PngError("PNG data size mismatch")
    img = _unfilter(raw, height, stride, ch).reshape(height, width, ch)
    if ctype == 3:
        if palette is None:
            raise PngError("palette PNG without PLTE")
        return palette[img[:, :, 0], 0].copy()
    return img[:, :, 0].copy()


def load_overlap_areas(path: Union[str, os.PathLike]) -> np.ndarray:
    """OVERLAP_AREAS: a Python-2 pickle of max-overlap pixel counts (8c/overlap_input.py:146-149), or a .npy copy."""
    if str(path).endswith(".npy"):
        return np.load(path).astype(np.float32)
    with open(path, "rb") as f:
        return np.asarray(pickle.load(f, encoding="latin1"), dtype=np.float32)


def pack_planes(planes: np.ndarray) -> np.ndarray:
    """uint8/bool [N, D] (non-zero = set) -> uint32 [N, ceil(D/32)], pixel p = bit (p & 31) of word p >> 5."""
    planes = np.asarray(planes)
    n, d = planes.shape
    by = np.packbits(planes != 0, axis=1, bitorder="little")
    wpr = (d + 31) // 32
    if by.shape[1] != 4 * wpr:
        by = np.concatenate([by, np.zeros((n, 4 * wpr - by.shape[1]), dtype=np.uint8)], axis=1)
    return np.ascontiguousarray(by).view("<u4")


def unpack_planes(words: np.ndarray, d: int) -> np.ndarray:
    """Inverse of pack_planes: uint32 [N, W] -> uint8 {0,1} [N, d]."""
    by = np.ascontiguousarray(np.asarray(words).astype("<u4")).view(np.uint8)
    return np.unpackbits(by, axis=1, bitorder="little")[:, :d]


class OverlapDataset:
    """Lock/key planes of the first `num_examples` pairs, uint8 {0,1}, [N, D] each, and their labels [N]."""

    def __init__(self, data: Union[str, os.PathLike], overlap_areas: Union[str, os.PathLike, Sequence[float]],
                 num_examples: Optional[int] = None, image_size: int = 200):
        areas = load_overlap_areas(overlap_areas) if isinstance(overlap_areas, (str, os.PathLike)) \
            else np.asarray(overlap_areas, dtype=np.float32)
        zf = zipfile.ZipFile(data) if str(data).endswith(".zip") else None
        names = {os.path.basename(n): n for n in zf.namelist()} if zf else None

        def blob(name):
            if zf:
                if name not in names:
                    raise FileNotFoundError(name)
                return zf.read(names[name])
            return open(os.path.join(str(data), name), "rb").read()

        n = int(num_examples if num_examples is not None else len(areas))
        if n > len(areas):
            raise ValueError("more examples requested than labels in OVERLAP_AREAS")
        D = image_size * image_size
        self.lock = np.empty((n, D), dtype=np.uint8)
        self.key = np.empty((n, D), dtype=np.uint8)
        for i in range(n):                                   # '%d_L.png' / '%d_K.png', 8c/overlap_input.py:162-163
            for kind, dst in (("L", self.lock), ("K", self.key)):
                img = read_png(blob("%d_%s.png" % (i, kind)))
                if img.shape != (image_size, image_size):
                    raise ValueError("%d_%s.png is %s, expected %dx%d" % (i, kind, img.shape, image_size, image_size))
                dst[i] = (img.reshape(-1) > 0)
        self.labels = areas[:n].copy()
        self.image_size = image_size
        self.lock_bits = pack_planes(self.lock)
        self.key_bits = pack_planes(self.key)

    @classmethod
    def from_arrays(cls, lock: np.ndarray, key: np.ndarray, overlap_areas: Sequence[float], image_size: int = 200) -> "OverlapDataset":
        """A dataset from planes already in memory: lock / key [N, D] (any non-zero value = set pixel), labels [N]."""
        self = cls.__new__(cls)
        self.lock = np.ascontiguousarray(np.asarray(lock).reshape(len(lock), -1) != 0).astype(np.uint8)
        self.key = np.ascontiguousarray(np.asarray(key).reshape(len(key), -1) != 0).astype(np.uint8)
        self.labels = np.asarray(overlap_areas, dtype=np.float32)[:len(self.lock)].copy()
        if self.lock.shape != self.key.shape or self.lock.shape[1] != image_size * image_size or len(self.labels) != len(self.lock):
            raise ValueError("lock / key must be [N, %d] with N labels" % (image_size * image_size))
        self.image_size = image_size
        self.lock_bits = pack_planes(self.lock)
        self.key_bits = pack_planes(self.key)
        return self

    def __len__(self) -> int:
        return len(self.labels)

    def index_batches(self, batch_size: int, epochs: int = 1, shuffle: bool = True, seed: int = 0,
                      steps: Optional[int] = None) -> Iterator[np.ndarray]:
        """The example indices of every batch: a fresh permutation per epoch, the ragged tail dropped.  With `steps`
        the iterator instead draws that many batches with replacement (any batch size: the reference's
        shuffle_batch queue likewise keeps serving examples for as long as the loop dequeues, 8c/overlap_input.py:297-340)."""
        rng = np.random.default_rng(seed)
        n = len(self)
        if steps is not None:
            for _ in range(steps):
                yield rng.integers(0, n, batch_size)
            return
        if batch_size > n:
            raise ValueError("batch_size %d exceeds the %d loaded examples" % (batch_size, n))
        for _ in range(epochs):
            order = rng.permutation(n) if shuffle else np.arange(n)
            for s in range(0, n - batch_size + 1, batch_size):
                yield order[s:s + batch_size]

    def batches(self, batch_size: int, epochs: int = 1, shuffle: bool = True,
                seed: int = 0) -> Iterator[Tuple[np.ndarray, np.ndarray, np.ndarray]]:
        """Yields (lock[B, D] uint8, key[B, D] uint8, labels[B] float32), images and labels drawn JOINTLY."""
        for idx in self.index_batches(batch_size, epochs, shuffle, seed):
            yield self.lock[idx], self.key[idx], self.labels[idx]


def device_batches(vae, dataset: OverlapDataset, epochs: int = 1, shuffle: bool = True, seed: int = 0,
                   packed: bool = True, steps: Optional[int] = None, gather: str = "auto"):
    """`overlap_input.inputs(normalize=True, reshape=True, rotation=True)` for the B200 backend: yields handles
    (magic_b200/tango.py: ResidentBatch) whose planes are already in HBM; `batch.labels` are the matching overlap
    areas, so `vae.partial_fit(batch, overlap_areas=batch.labels)` trains on a consistent (images, labels) pair.

    packed=True (default): batches travel bit-packed and one step ahead -- while the caller trains on batch n,
    batch n+1 is already crossing PCIe / being unpacked on the engine's input stream (`magic_feed_pairs_bits_host`
    with activate=0, then `magic_swap_batch`).  packed=False sends uint8 planes with `magic_load_pairs_u8`.

    gather: where the rows of a batch are picked out of the data set.  "device" pins the packed data set once and lets the
    GPU read the rows of each batch over PCIe (`magic_feed_pairs_bits_gather`): per step the host handles 16 KB of indices
    and labels.  "host" gathers them into a pinned slot on worker threads and copies the slot (copy engine: faster per byte,
    but the host must keep up).  "auto" = "host" when this rank has at least six host threads to itself, else "device"."""
    import torch
    from .tango import ResidentBatch
    eng = vae.engine
    if not packed:
        step = 0
        for idx in dataset.index_batches(eng.B, epochs=epochs, shuffle=shuffle, seed=seed, steps=steps):
            lock, key, labels = dataset.lock[idx], dataset.key[idx], dataset.labels[idx]
            eng.load_pairs_u8(torch.from_numpy(np.ascontiguousarray(lock)), torch.from_numpy(np.ascontiguousarray(key)),
                              seed, step)
            step += 1
            yield ResidentBatch(vae, np.ascontiguousarray(labels, dtype=np.float32))
        return

    wpr = dataset.lock_bits.shape[1]
    src = [torch.from_numpy(dataset.lock_bits.view(np.int32)), torch.from_numpy(dataset.key_bits.view(np.int32)),
           torch.from_numpy(dataset.labels)]
    pin = torch.cuda.is_available()
    if gather not in ("auto", "device", "host"):
        raise ValueError("gather must be 'auto', 'device' or 'host'")
    local_world = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1") or 1))
    host_threads = (os.cpu_count() or 8) // local_world - 1              # what this rank can spend on its gather
    if gather == "auto":
        # measured on B200 boxes: GPU-side reads of pinned memory reach ~18 GB/s (a 4096-pair batch in 2.3 ms), the copy engine
        # 50 GB/s -- the host gather wins whenever the host keeps up (7 threads: yes; 3 threads, 8 ranks on 32 cores: 5.2 ms)
        gather = "host" if host_threads >= 6 else "device"
    if gather != "host" and pin:
        pinned = getattr(dataset, "_pinned_bits", None)
        if pinned is None:
            try:
                wpr4 = (wpr + 3) // 4 * 4                     # rows padded to 16 bytes: the gather moves 16-byte pieces
                pinned = tuple(torch.zeros((len(dataset.labels), wpr4), dtype=torch.int32).pin_memory() for _ in range(2))
                pinned[0][:, :wpr] = src[0]                   # once per data set (10 KB per pair)
                pinned[1][:, :wpr] = src[1]
                dataset._pinned_bits = pinned
            except RuntimeError:
                if gather == "device":
                    raise
                pinned = None
        if pinned is not None:
            yield from _device_gather_batches(vae, dataset, pinned, epochs, shuffle, seed, steps)
            return
    elif gather == "device":
        raise ValueError("gather='device' needs CUDA")
    # three pinned slots: batch n is being trained on, batch n+1 is crossing PCIe / staged on the device, batch n+2 is
    # being gathered from the dataset by a worker thread (index_select releases the GIL) while the caller waits for
    # the losses of step n -- the host gather never sits between two steps
    ring = [(torch.empty((eng.B, wpr), dtype=torch.int32, pin_memory=pin), torch.empty((eng.B, wpr), dtype=torch.int32, pin_memory=pin),
             torch.empty((eng.B,), dtype=torch.float32, pin_memory=pin)) for _ in range(3)]
    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(max_workers=1)
    # The gather itself is split over a few threads of its own: launchers such as torchrun set OMP_NUM_THREADS=1, which
    # makes a single index_select of 41 MB take longer than a training step (measured at 2 GPUs: 4.0 vs 2.7 ms per step).
    n_gather = max(1, min(8, host_threads))
    gpool = ThreadPoolExecutor(max_workers=n_gather) if n_gather > 1 else None

    def gather(n, idx):
        slot = ring[n % 3]                                    # its previous batch (n - 3) was consumed two steps ago
        it = torch.from_numpy(np.ascontiguousarray(idx, dtype=np.int64))

        def part(lo, hi):
            for dst, s_ in zip(slot, src):
                torch.index_select(s_, 0, it[lo:hi], out=dst[lo:hi])
        if gpool is None:
            part(0, len(it))
        else:
            step_ = (len(it) + n_gather - 1) // n_gather
            for f in [gpool.submit(part, lo, min(len(it), lo + step_)) for lo in range(0, len(it), step_)]:
                f.result()
        return n

    def feed(n, activate):
        slot = ring[n % 3]
        eng.feed_pairs_bits(slot[0], slot[1], slot[2], seed, n, activate=activate)
        return slot[2].numpy().copy()

    order = dataset.index_batches(eng.B, epochs=epochs, shuffle=shuffle, seed=seed, steps=steps)
    try:
        idx = next(order, None)
        if idx is None:
            return
        gather(0, idx)
        labels = feed(0, True)
        idx = next(order, None)
        pending = pool.submit(gather, 1, idx) if idx is not None else None
        n = 0
        while True:
            staged = None
            if pending is not None:
                pending.result()
                staged = feed(n + 1, False)
                idx = next(order, None)
                pending = pool.submit(gather, n + 2, idx) if idx is not None else None
            batch = ResidentBatch(vae, labels)
            yield batch
            if staged is None:
                return
            if vae._loaded_key is not batch._token:
                raise RuntimeError("another load replaced the engine's input buffers while iterating the dataset")
            eng.swap_batch()
            labels, n = staged, n + 1
    finally:
        pool.shutdown(wait=True)
        if gpool is not None:
            gpool.shutdown(wait=True)


def _device_gather_batches(vae, dataset: OverlapDataset, pinned, epochs, shuffle, seed, steps):
    """device_batches with the gather on the GPU: one batch ahead, three small pinned slots of (indices, labels)."""
    import torch
    from .tango import ResidentBatch
    eng = vae.engine
    ring = [(torch.empty((eng.B,), dtype=torch.int32, pin_memory=True), torch.empty((eng.B,), dtype=torch.float32, pin_memory=True))
            for _ in range(3)]
    labels_all = dataset.labels

    def feed(n, idx, activate):
        slot = ring[n % 3]                                    # its previous batch (n - 3) was consumed two steps ago
        slot[0].numpy()[:] = idx
        slot[1].numpy()[:] = labels_all[idx]
        eng.feed_pairs_bits_gather(pinned[0], pinned[1], slot[0], slot[1], seed, n, activate=activate)
        return slot[1].numpy().copy()

    order = dataset.index_batches(eng.B, epochs=epochs, shuffle=shuffle, seed=seed, steps=steps)
    idx = next(order, None)
    if idx is None:
        return
    labels = feed(0, idx, True)
    n = 0
    while True:
        idx = next(order, None)
        staged = feed(n + 1, idx, False) if idx is not None else None
        batch = ResidentBatch(vae, labels)
        yield batch
        if staged is None:
            return
        if vae._loaded_key is not batch._token:
            raise RuntimeError("another load replaced the engine's input buffers while iterating the dataset")
        eng.swap_batch()
        labels, n = staged, n + 1


# ------------------------------------------------------------------ `overlap_input.inputs()` for the B200 backend
# (8c/overlap_input.py:76-124, called at module level by 8c/main.py:25 and consumed as `images_batch.eval()` /
# `labels_batch.eval()` at 8c/main.py:44-45 or jointly as `sess.run([images_batch, labels_batch])` at 11a/main.py:65.)
_PIPELINES: list = []


class InputPipeline:
    """What the reference's queue runners + shuffle_batch are for one experiment: an endless source of batches.  With a
    dataset under FLAGS.DATA_DIR (the '%d_L.png' / '%d_K.png' pairs + OVERLAP_AREAS) batches come from
    `device_batches` (row N2), otherwise from the on-device generator (row N1).  Bound to the model when the
    TangoEncoder is constructed -- main.py calls inputs() before the model exists, exactly like the reference."""

    def __init__(self, flags, seed: int = 0):
        self.flags, self.seed = flags, seed
        self.model = None
        self.batch = None
        self._iter = None
        self._dataset = None
        d = getattr(flags, "DATA_DIR", None)
        if d and (os.path.isdir(str(d)) or str(d).endswith(".zip")) and os.path.exists(str(d)):
            areas = os.path.join(str(d), "OVERLAP_AREAS") if os.path.isdir(str(d)) else getattr(flags, "OVERLAP_AREAS", None)
            if areas and os.path.exists(areas):
                self._dataset = (str(d), areas)

    @property
    def source(self) -> str:
        return "dataset " + self._dataset[0] if self._dataset else "on-device generator"

    def bind(self, model) -> None:
        self.model, self.batch, self._iter = model, None, None

    def advance(self):
        """Dequeue the next batch (images and labels together)."""
        m = self.model
        if m is None:
            raise RuntimeError("overlap_input.inputs(): no TangoEncoder has been constructed yet (the handles evaluate on its engine)")
        if self._dataset is None:
            self.batch = m.next_device_batch(self.seed)
            return self.batch
        if self._iter is None:
            ds = OverlapDataset(self._dataset[0], self._dataset[1], image_size=self.flags.IMAGE_SIZE) \
                if not isinstance(self._dataset, OverlapDataset) else self._dataset
            big = m.engine.B > len(ds)
            self._iter = device_batches(m, ds, epochs=1 << 40, seed=self.seed, steps=(1 << 60) if big else None)
        self.batch = next(self._iter)
        return self.batch


class _Handle:
    def __init__(self, pipeline: InputPipeline):
        self.pipeline = pipeline


class ImagesHandle(_Handle):
    """`images_batch`: .eval() dequeues a NEW batch and returns a handle onto it that `partial_fit`, `get_predictions`
    and `reconstruct` accept without the images leaving HBM; np.asarray(handle) materialises [B, 120000] float32."""

    def eval(self, session=None):
        return self.pipeline.advance()

    def _current(self):
        return self.pipeline.batch


class LabelsHandle(_Handle):
    """`labels_batch`: .eval() returns the overlap areas of the batch the last images_batch.eval() dequeued (the
    reference's two .eval() calls dequeue twice and so pair images with the labels of another batch, 8c/main.py:44-45;
    10a/main.py:64 fixed that with a joint fetch -- here both forms yield a consistent pair)."""

    def eval(self, session=None):
        if self.pipeline.batch is None:
            self.pipeline.advance()
        return self._current()

    def _current(self):
        return self.pipeline.batch.labels


def inputs(normalize: bool = False, reshape: bool = False, rotation: bool = False, flags=None, seed: int = 0):
    """Returns (images, labels) handles.  The B200 backend implements the configuration every experiment from 7c on
    uses: normalize=True (pixels {0,1}), reshape=True ([B, 120000] rows), rotation=True (three planes)."""
    if flags is None:
        raise ValueError("inputs() needs the experiment's FLAGS")
    if not getattr(flags, "DATA_DIR", None):
        raise ValueError("Please supply a data_dir")                      # 8c/overlap_input.py:98-99
    if not (normalize and reshape and rotation):
        raise NotImplementedError("the B200 backend feeds the three-stream model: inputs(normalize=True, reshape=True, rotation=True)")
    p = InputPipeline(flags, seed)
    _PIPELINES.append(p)
    del _PIPELINES[:-8]                                                   # pipelines of long-gone experiments
    return ImagesHandle(p), LabelsHandle(p)


def bind_model(model) -> None:
    """Called by TangoEncoder.__init__: pipelines created before the model existed now evaluate on it."""
    for p in _PIPELINES:
        stale = p.model is None or not getattr(getattr(p.model.engine, "_h", None), "value", None)     # never bound / closed
        if stale or p.flags is model.flags:
            p.bind(model)
