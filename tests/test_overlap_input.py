"""Row N2: the PNG dataset path (magic_b200/overlap_input.py, magic_load_pairs_u8).

CPU: the PNG decoder against the committed dataset samples (tests/golden/png, byte copies of overlap_micro.zip) and
their bit-packed golden decode (overlap_micro_16.npz, made with PIL in the build container), against a small PNG
encoder that exercises every filter and colour type, and the dataset/batching logic of 8c/overlap_input.py.
GPU: dataset pairs -> device planes, rotated lock reproduced bit for bit from the exported (cos, sin), and a
partial_fit on (images, labels) drawn jointly."""
import os
import struct
import zlib

import numpy as np
import pytest

from magic_b200 import overlap_input as oi

HERE = os.path.dirname(os.path.abspath(__file__))
PNG_DIR = os.path.join(HERE, "golden", "png")


# --------------------------------------------------------------------------- a minimal PNG writer (test only)
def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)


def write_png(img: np.ndarray, ctype: int, filters, palette=None) -> bytes:
    """img uint8 [H, W, C]; one filter type per scanline taken cyclically from `filters`."""
    h, w, ch = img.shape
    stride = w * ch
    lines = img.reshape(h, stride).astype(np.int32)
    raw = bytearray()
    prev = np.zeros(stride, dtype=np.int32)
    for y in range(h):
        ft = filters[y % len(filters)]
        cur = lines[y]
        out = np.zeros(stride, dtype=np.int32)
        for x in range(stride):
            a = cur[x - ch] if x >= ch else 0
            b = prev[x]
            c = prev[x - ch] if x >= ch else 0
            pred = [0, a, b, (a + b) >> 1, _paeth(a, b, c)][ft]
            out[x] = (cur[x] - pred) & 255
        raw.append(ft)
        raw.extend(out.astype(np.uint8).tobytes())
        prev = cur

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)

    blob = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, ctype, 0, 0, 0))
    if palette is not None:
        blob += chunk(b"PLTE", palette.astype(np.uint8).tobytes())
    comp = zlib.compress(bytes(raw), 6)
    blob += chunk(b"IDAT", comp[:len(comp) // 2]) + chunk(b"IDAT", comp[len(comp) // 2:])   # split IDAT is legal
    return blob + chunk(b"IEND", b"")


# --------------------------------------------------------------------------- CPU tests
def test_decoder_matches_the_golden_decode_of_the_dataset_samples():
    g = np.load(os.path.join(HERE, "golden", "overlap_micro_16.npz"))
    for i in range(4):
        for kind, bits in (("L", g["lock_bits"]), ("K", g["key_bits"])):
            img = oi.read_png(os.path.join(PNG_DIR, "%d_%s.png" % (i, kind)))
            assert img.shape == (200, 200) and img.dtype == np.uint8
            assert set(np.unique(img)) <= {0, 255}
            assert np.array_equal(np.packbits(img.reshape(-1) > 0), bits[i])


@pytest.mark.parametrize("ctype,ch", [(0, 1), (4, 2), (2, 3), (6, 4)])
def test_every_filter_type_round_trips(ctype, ch):
    rng = np.random.default_rng(ctype)
    img = rng.integers(0, 256, size=(13, 17, ch), dtype=np.uint8)
    for filters in ([0], [1], [2], [3], [4], [0, 1, 2, 3, 4]):
        out = oi.read_png(write_png(img, ctype, filters))
        assert np.array_equal(out, img[:, :, 0]), (ctype, filters)


def test_palette_images_and_errors():
    rng = np.random.default_rng(5)
    idx = rng.integers(0, 4, size=(9, 11, 1), dtype=np.uint8)
    pal = np.array([[0, 0, 0], [255, 0, 0], [7, 8, 9], [200, 1, 2]], dtype=np.uint8)
    out = oi.read_png(write_png(idx, 3, [4], palette=pal))
    assert np.array_equal(out, pal[idx[:, :, 0], 0])
    blob = bytearray(write_png(idx, 0, [0]))
    blob[40] ^= 0x55                                                  # flips a byte inside IDAT: CRC must catch it
    with pytest.raises(oi.PngError):
        oi.read_png(bytes(blob))
    with pytest.raises(oi.PngError):
        oi.read_png(b"not a png at all")


def test_dataset_and_joint_batches():
    areas = np.load(os.path.join(HERE, "golden", "overlap_areas.npy"))
    ds = oi.OverlapDataset(PNG_DIR, areas, num_examples=4)
    assert len(ds) == 4 and ds.lock.shape == (4, 40000) and ds.key.dtype == np.uint8
    assert set(np.unique(ds.lock)) <= {0, 1}
    g = np.load(os.path.join(HERE, "golden", "overlap_micro_16.npz"))
    assert list(ds.lock.sum(1)) == list(g["area_lock"][:4]) and list(ds.key.sum(1)) == list(g["area_key"][:4])
    assert np.array_equal(ds.labels, areas[:4].astype(np.float32))
    # images and labels stay paired under shuffling (the 10a/main.py:64 fix), epochs re-permute, ragged tail is dropped
    seen = []
    for lock, key, y in ds.batches(3, epochs=2, shuffle=True, seed=1):
        assert lock.shape == (3, 40000) and y.shape == (3,)
        for r in range(3):
            i = int(np.flatnonzero((ds.lock == lock[r]).all(1))[0])
            assert np.array_equal(ds.key[i], key[r]) and ds.labels[i] == y[r]
            seen.append(i)
    assert len(seen) == 6
    again = [y.copy() for _, _, y in ds.batches(3, epochs=2, shuffle=True, seed=1)]
    first = [y.copy() for _, _, y in ds.batches(3, epochs=2, shuffle=True, seed=1)]
    assert all(np.array_equal(a, b) for a, b in zip(again, first))   # deterministic in the seed
    with pytest.raises(ValueError):
        next(ds.batches(5))
    with pytest.raises(ValueError):
        oi.OverlapDataset(PNG_DIR, areas[:2], num_examples=4)
    with pytest.raises(FileNotFoundError):
        oi.OverlapDataset(PNG_DIR, areas, num_examples=5)


def test_bit_packing_layout_and_round_trip():
    rng = np.random.default_rng(11)
    for d in (40000, 45, 64, 1):
        x = (rng.random((3, d)) < 0.4).astype(np.uint8)
        w = oi.pack_planes(x * 255)                                  # any non-zero byte is a set pixel
        assert w.dtype == np.dtype("<u4") and w.shape == (3, (d + 31) // 32)
        assert np.array_equal(oi.unpack_planes(w, d), x)
        for p in rng.integers(0, d, size=8):
            assert ((int(w[1, p >> 5]) >> (int(p) & 31)) & 1) == x[1, p]      # pixel p = bit p & 31 of word p >> 5
    areas = np.load(os.path.join(HERE, "golden", "overlap_areas.npy"))
    ds = oi.OverlapDataset(PNG_DIR, areas, num_examples=4)
    assert ds.lock_bits.shape == (4, 1250) and np.array_equal(oi.unpack_planes(ds.key_bits, 40000), ds.key)
    got = [idx.copy() for idx in ds.index_batches(2, epochs=2, shuffle=True, seed=3)]
    assert len(got) == 4 and sorted(np.concatenate(got[:2])) == [0, 1, 2, 3] == sorted(np.concatenate(got[2:]))


# --------------------------------------------------------------------------- GPU tests
@pytest.mark.gpu
def test_dataset_pairs_on_device_and_training_step():
    import torch
    from magic_b200.tango import TangoEncoder
    from tests.test_generator import nn_rotate

    class FLAGS:                                            # an experiment's constants.FLAGS, reduced hidden sizes
        IMAGE_SIZE, NUM_LAYERS, BATCH_SIZE = 200, 3, 4
        NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN = 4
        BACKEND, PRECISION = "b200", "bf16"

    class Model(TangoEncoder):
        encoder_neurons, decoder_neurons, latent_dimensions = (64, 64), (64, 64), 8

    areas = np.load(os.path.join(HERE, "golden", "overlap_areas.npy"))
    ds = oi.OverlapDataset(PNG_DIR, areas, num_examples=4)
    vae = Model(flags=FLAGS, verbose=False)
    it = oi.device_batches(vae, ds, epochs=3, shuffle=True, seed=7)
    batch = next(it)
    X, _ = vae.engine.export_batch()
    X = X.cpu().numpy().reshape(4, 200, 200, 3)
    rot = vae.engine.export_rotation().cpu().numpy()
    order = [int(np.flatnonzero((ds.lock == X[r, :, :, 0].reshape(-1).astype(np.uint8)).all(1))[0]) for r in range(4)]
    assert sorted(order) == [0, 1, 2, 3]
    for r, i in enumerate(order):
        assert np.array_equal(X[r, :, :, 2].reshape(-1), ds.key[i])                 # key plane, same pair
        assert batch.labels[r] == ds.labels[i]                                       # and its label
        want = nn_rotate(ds.lock[i].reshape(200, 200), int(rot[r, 0]), int(rot[r, 1]))
        assert np.array_equal(X[r, :, :, 1].astype(np.uint8), want)                  # rotated lock, bit for bit
    assert len({(int(a), int(b)) for a, b in rot}) > 1                               # per-pair angles differ
    out = vae.partial_fit(batch, overlap_areas=batch.labels)
    assert len(out) == 5 and all(np.isfinite(float(v)) for v in out)
    assert abs(out[2] - 40000 * np.log(2.0)) / (40000 * np.log(2.0)) < 0.05          # rec ~ D ln 2 at initialisation
    nxt = next(it)
    with pytest.raises(RuntimeError):
        vae.partial_fit(batch, overlap_areas=batch.labels)                            # replaced by the later load
    out2 = vae.partial_fit(nxt, overlap_areas=nxt.labels)
    assert np.isfinite(out2[0])
    torch.cuda.synchronize()


@pytest.mark.gpu
def test_packed_feed_equals_the_uint8_loader_and_pipelines_one_batch_ahead():
    import torch
    from magic_b200.tango import TangoEncoder

    class FLAGS:
        IMAGE_SIZE, NUM_LAYERS, BATCH_SIZE = 200, 3, 2
        NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN = 4
        BACKEND, PRECISION = "b200", "bf16"

    class Model(TangoEncoder):
        encoder_neurons, decoder_neurons, latent_dimensions = (64, 64), (64, 64), 8

    areas = np.load(os.path.join(HERE, "golden", "overlap_areas.npy"))
    ds = oi.OverlapDataset(PNG_DIR, areas, num_examples=4)
    vae = Model(flags=FLAGS, verbose=False)
    eng = vae.engine

    def loaded():
        X, y = eng.export_batch()
        return X.cpu().numpy(), eng.export_rotation().cpu().numpy(), y.cpu().numpy()

    # 1. the three loaders agree bit for bit: uint8 planes, packed words on the device, packed words from the host
    idx = np.array([2, 0])
    eng.load_pairs_u8(torch.from_numpy(ds.lock[idx]), torch.from_numpy(ds.key[idx]), 5, 9)
    X8, R8, _ = loaded()
    eng.load_pairs_bits(torch.from_numpy(ds.lock_bits[idx].view(np.int32)), torch.from_numpy(ds.key_bits[idx].view(np.int32)), 5, 9)
    Xd, Rd, _ = loaded()
    eng.feed_pairs_bits(ds.lock_bits[idx], ds.key_bits[idx], ds.labels[idx], 5, 9)
    Xh, Rh, yh = loaded()
    assert np.array_equal(X8, Xd) and np.array_equal(X8, Xh) and np.array_equal(R8, Rd) and np.array_equal(R8, Rh)
    assert np.array_equal(yh, ds.labels[idx]) and X8.sum() > 0
    out_a = vae.engine.step(labels=ds.labels[idx])[0].cpu().numpy()                # explicit labels
    eng.feed_pairs_bits(ds.lock_bits[idx], ds.key_bits[idx], ds.labels[idx], 5, 9)
    # 2. a staged batch does not disturb the loaded one until swap_batch
    other = np.array([1, 3])
    eng.feed_pairs_bits(ds.lock_bits[other], ds.key_bits[other], ds.labels[other], 5, 10, activate=False)
    assert np.array_equal(loaded()[0], X8)
    eng.swap_batch()
    Xo, _, yo = loaded()
    assert np.array_equal(yo, ds.labels[other]) and not np.array_equal(Xo, X8)
    assert np.array_equal(Xo.reshape(2, 40000, 3)[:, :, 0].astype(np.uint8), ds.lock[other])
    assert np.isfinite(out_a).all()
    # 3. the iterator: 2 epochs x 2 batches, each batch pairs images with their own labels, one batch staged ahead
    #    -- with the rows of a batch gathered by the GPU from the pinned data set (the default) and by the host
    per_mode = {}
    for mode in ("device", "host"):
        seen, rots = [], []
        for batch in oi.device_batches(vae, ds, epochs=2, shuffle=True, seed=4, gather=mode):
            X, R, y = loaded()
            assert np.array_equal(y, batch.labels)
            lock = X.reshape(2, 40000, 3)[:, :, 0].astype(np.uint8)
            for r in range(2):
                i = int(np.flatnonzero((ds.lock == lock[r]).all(1))[0])
                assert ds.labels[i] == batch.labels[r]
                assert np.array_equal(X.reshape(2, 40000, 3)[r, :, 2].astype(np.uint8), ds.key[i])
                seen.append(i)
            rots.append(R.copy())
            out = vae.partial_fit(batch, overlap_areas=batch.labels)
            assert np.isfinite(out[0])
        assert sorted(seen[:4]) == [0, 1, 2, 3] == sorted(seen[4:])
        per_mode[mode] = (seen, np.stack(rots))
    assert per_mode["device"][0] == per_mode["host"][0]                            # same order, same rotations
    assert np.array_equal(per_mode["device"][1], per_mode["host"][1])
    with pytest.raises(Exception, match="pinned"):                                 # the device gather reads pinned memory only
        eng.feed_pairs_bits_gather(torch.zeros((4, 1252), dtype=torch.int32), torch.zeros((4, 1252), dtype=torch.int32),
                                   torch.zeros(2, dtype=torch.int32), None, 0, 0)
    # a foreign load between two batches is reported, not silently trained on
    it = oi.device_batches(vae, ds, epochs=2, shuffle=False)
    next(it)
    eng.load_pairs_u8(torch.from_numpy(ds.lock[idx]), torch.from_numpy(ds.key[idx]), 5, 9)
    vae._loaded_key = object()
    with pytest.raises(RuntimeError):
        next(it)
    torch.cuda.synchronize()
