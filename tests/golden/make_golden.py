"""Regenerates the fixtures in this directory from the reference tree (run in the build container, where
/root/reference exists; the GPU box only ever reads the committed outputs).

    python tests/golden/make_golden.py

overlap_areas.npy      2a/OVERLAP_AREAS (pickle of 2000 int64 max-overlap labels) re-saved as a plain .npy
overlap_micro_16.npz   the first 16 lock/key PNG pairs of overlap_micro.zip, 200x200, bit-packed ({0,255} -> 1 bit/px),
                       plus the pixel areas of all 100 locks and keys (generator statistics)
overlap_micro_100.npz  all 100 lock/key pairs of overlap_micro.zip, bit-packed the same way (np.packbits, MSB first): the
                       real-data batches of bench.py and of the generator-statistics tests are sampled from these
png/{0..3}_{L,K}.png   the first four lock/key files of overlap_micro.zip, byte for byte (dataset samples, ~0.8 kB each):
                       fixtures of the PNG loader path (row N2, tests/test_overlap_input.py)
oracle_step_mini.npz   inputs and float64 outputs of one oracle step on a reduced geometry (losses, distance, a few
                       gradient tensors): pins the oracle against accidental edits of oracle/tango_oracle.py
"""
import io
import os
import pickle
import sys
import zipfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def labels():
    with open(os.path.join(REF, "2a", "OVERLAP_AREAS"), "rb") as f:
        y = np.asarray(pickle.load(f, encoding="latin1"), dtype=np.int64)
    assert y.shape == (2000,)
    np.save(os.path.join(HERE, "overlap_areas.npy"), y)
    return y


def micro():
    from PIL import Image
    z = zipfile.ZipFile(os.path.join(REF, "overlap_micro.zip"))
    names = z.namelist()
    locks, keys, area_l, area_k = [], [], [], []
    for i in range(100):
        for kind, dst, areas in (("L", locks, area_l), ("K", keys, area_k)):
            cand = [n for n in names if n.endswith("/%d_%s.png" % (i, kind)) or n == "%d_%s.png" % (i, kind)]
            assert len(cand) == 1, (i, kind, cand)
            img = np.asarray(Image.open(io.BytesIO(z.read(cand[0]))).convert("L"))
            assert img.shape == (200, 200) and set(np.unique(img)) <= {0, 255}
            bits = img > 0
            areas.append(int(bits.sum()))
            dst.append(np.packbits(bits.reshape(-1)))
    np.savez_compressed(os.path.join(HERE, "overlap_micro_16.npz"), lock_bits=np.stack(locks[:16]), key_bits=np.stack(keys[:16]),
                        area_lock=np.asarray(area_l), area_key=np.asarray(area_k))
    # all 100 pairs: the real-data leg of bench.py and the generator-statistics tests sample their batches from these
    np.savez_compressed(os.path.join(HERE, "overlap_micro_100.npz"), lock_bits=np.stack(locks), key_bits=np.stack(keys))


def png_fixtures():
    z = zipfile.ZipFile(os.path.join(REF, "overlap_micro.zip"))
    os.makedirs(os.path.join(HERE, "png"), exist_ok=True)
    for i in range(4):
        for kind in ("L", "K"):
            cand = [n for n in z.namelist() if os.path.basename(n) == "%d_%s.png" % (i, kind)]
            assert len(cand) == 1
            with open(os.path.join(HERE, "png", "%d_%s.png" % (i, kind)), "wb") as f:
                f.write(z.read(cand[0]))


def oracle_step():
    import torch
    from oracle import tango_oracle as orc
    cfg = orc.OracleConfig(encoder_neurons=[24, 16], decoder_neurons=[12, 20], latent=5, n_input=64)
    st = orc.OracleState(cfg, seed=3)
    xl, xr, xk, y, eps = orc.synthetic_batch(6, cfg.latent, seed=4, n_input=cfg.n_input)
    params0 = {k: v.numpy().copy() for k, v in st.params.items()}
    out, g1, g2 = st.step(xl, xr, xk, y, eps)
    d = {"p0_" + k: v for k, v in params0.items()}
    d.update({"p1_" + k: v.numpy() for k, v in st.params.items()})
    d.update({"g1_" + k: v.numpy() for k, v in g1.items()})
    d.update({"g2_" + k: v.numpy() for k, v in g2.items()})
    d.update(xl=xl.numpy(), xr=xr.numpy(), xk=xk.numpy(), y=y.numpy(), eps=eps.numpy(),
             losses=np.asarray([float(out[k]) for k in orc.LOSS_KEYS]), distance=out["distance"].numpy())
    np.savez_compressed(os.path.join(HERE, "oracle_step_mini.npz"), **d)


if __name__ == "__main__":
    labels()
    micro()
    png_fixtures()
    oracle_step()
    print("golden fixtures written to", HERE)
