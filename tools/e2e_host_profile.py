"""Where does the host spend a step of the end-to-end loop (bench.py leg 2)?  Wall-clock per phase of
`for b in device_batches(...): vae.partial_fit(b, b.labels)` on the real images, one GPU.
    python tools/e2e_host_profile.py [steps]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from magic_b200.overlap_input import OverlapDataset, device_batches  # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    torch.cuda.set_device(0)
    cfg, eng, vae = bench.make_engine("8d", 4096, "bf16", 0, 1, torch.device("cuda", 0))
    g = np.load(os.path.join(ROOT, "tests", "golden", "overlap_micro_100.npz"))
    Lr = np.unpackbits(g["lock_bits"], axis=1)[:, :cfg.D]
    Kr = np.unpackbits(g["key_bits"], axis=1)[:, :cfg.D]
    areas = np.load(os.path.join(ROOT, "tests", "golden", "overlap_areas.npy"))[:len(Lr)].astype(np.float32)
    ds = OverlapDataset.from_arrays(Lr, Kr, areas, image_size=int(round(cfg.D ** 0.5)))
    it = device_batches(vae, ds, seed=0, steps=steps + 8)
    for _ in range(5):
        b = next(it)
        vae.partial_fit(b, overlap_areas=b.labels)
    torch.cuda.synchronize()
    acc = {"next(batch)": 0.0, "labels+load": 0.0, "step enqueue": 0.0, "read_losses wait": 0.0}
    t_all = time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        b = next(it)
        t1 = time.perf_counter()
        labels = vae._labels(b.labels)
        vae._load(b)
        t2 = time.perf_counter()
        eng.step(labels=labels)
        t3 = time.perf_counter()
        eng.read_losses()
        t4 = time.perf_counter()
        for k, d in zip(acc, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
            acc[k] += d
    torch.cuda.synchronize()
    total = (time.perf_counter() - t_all) / steps * 1e3
    print("e2e host profile, %d steps: %.3f ms per step" % (steps, total))
    for k, v in acc.items():
        print("  %-18s %.3f ms" % (k, v / steps * 1e3))


if __name__ == "__main__":
    main()
