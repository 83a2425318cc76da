"""How many (tile, k-block) stages of the first layer's GEMMs are entirely zero on generated batches."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from magic_b200.engine import Engine, EngineConfig

B = 4096
eng = Engine(EngineConfig.experiment("8d", batch=B, precision="bf16"))
eng.generate_batch(0, 0)
X, y = eng.export_batch()
D = eng.D
Xi = X.view(B, D, 3)
planes = torch.stack([Xi[:, :, 1], Xi[:, :, 0], Xi[:, :, 2]], 0).reshape(3 * B, D) > 0     # rows [rot; lock; key]
print("fill fraction per stream (rot, lock, key):", [float(planes[i * B:(i + 1) * B].float().mean()) for i in range(3)])
Dp = (D + 1 + 63) // 64 * 64
P = torch.zeros(3 * B, Dp, dtype=torch.bool, device=X.device)
P[:, :D] = planes
P[:, D] = True
for kw in (64, 32, 16):
    # forward: tiles of 128 batch rows x kw pixels
    t = P.view(3 * B // 128, 128, Dp // kw, kw).any(3).any(1)
    print("fwd  : 128 rows x %2d px blocks nonzero: %.3f" % (kw, float(t.float().mean())))
# wgrad: tiles of 128 pixels x 64 batch rows, groups [0,2B) and [B,3B)
Dw = (D + 1 + 127) // 128 * 128
Q = torch.zeros(3 * B, Dw, dtype=torch.bool, device=X.device)
Q[:, :Dp] = P
for kb in (64, 32, 16):
    for g in range(2):
        rows = Q[g * B:g * B + 2 * B]
        t = rows.view(2 * B // kb, kb, Dw // 128, 128).any(3).any(1)
        print("wgrad: group %d, 128 px x %2d batch rows nonzero: %.3f ; tiles with any: %.3f" % (g, kb, float(t.float().mean()), float(t.any(0).float().mean())))
# per-pixel union over the whole batch
u = planes.any(0).view(200, 200)
print("pixels ever set: %.3f ; rows ever set: %d..%d" % (float(u.float().mean()), int(u.any(1).nonzero().min()), int(u.any(1).nonzero().max())))
