"""Where the host-fed loop's time goes (8d, 4096 pairs): the step alone with a per-step loss read-back, the packed
feed alone (H2D + rotation kernel), and both together as bench.py's e2e leg runs them."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from magic_b200.engine import Engine, EngineConfig
from magic_b200.overlap_input import pack_planes

B = 4096
eng = Engine(EngineConfig.experiment("8d", batch=B, precision="bf16"))
eng.generate_batch(0, 0)
X, y = eng.export_batch()
Xi = X.view(B, eng.D, 3).cpu()
lock_b = torch.from_numpy(pack_planes(Xi[:, :, 0].numpy()).view(np.int32)).pin_memory()
key_b = torch.from_numpy(pack_planes(Xi[:, :, 2].numpy()).view(np.int32)).pin_memory()
y_host = y.cpu().pin_memory()
h_loss = torch.empty(5, dtype=torch.float32, pin_memory=True)
eng.generate_batch(0, 1)
K = 60

def timed(fn, n=K):
    for _ in range(5):
        fn(0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        fn(i)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3

def step_only(i):
    eng.swap_batch(); eng.step()
def step_sync(i):
    eng.swap_batch(); ld, _ = eng.step(); h_loss.copy_(ld, non_blocking=True); torch.cuda.current_stream().synchronize()
def feed_only(i):
    eng.feed_pairs_bits(lock_b, key_b, y_host, 1, i, activate=False)
def feed_step(i):
    eng.feed_pairs_bits(lock_b, key_b, y_host, 1, i + 1, activate=False)
    ld, _ = eng.step(); h_loss.copy_(ld, non_blocking=True); torch.cuda.current_stream().synchronize(); eng.swap_batch()
def gen_step(i):
    eng.generate_batch(0, 100 + i); eng.step()
def gen_only(i):
    eng.generate_batch(0, 100 + i)

print("step, resident inputs, no sync      %.4f ms" % timed(step_only))
print("step + loss read-back every step    %.4f ms" % timed(step_sync))
print("packed feed alone (H2D + rotation)  %.4f ms" % timed(feed_only))
eng.feed_pairs_bits(lock_b, key_b, y_host, 1, 0)
print("feed one ahead + step + read-back   %.4f ms" % timed(feed_step))
print("generator alone                     %.4f ms" % timed(gen_only))
print("generator + step                    %.4f ms" % timed(gen_step))
eng.profile(True)
for i in range(10):
    feed_step(i)
rep = eng.profile_report()
for tag, (ms, n, _f) in sorted(rep.items(), key=lambda kv: -kv[1][0]):
    if tag.startswith("input"):
        print("  %-22s %.4f ms per launch (%d)" % (tag, ms / n, n))
eng.close()
