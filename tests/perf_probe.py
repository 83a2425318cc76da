import os, sys, time, torch
sys.path.insert(0, '/root/repo')
from magic_b200.engine import Engine, EngineConfig
exp = sys.argv[1] if len(sys.argv) > 1 else "8d"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
if B <= 0:
    B = {"8c": 100, "8d": 4096, "8e": 2048, "9a": 8192, "11a": 4096}.get(exp, 4096)
prec = sys.argv[3] if len(sys.argv) > 3 else "bf16"
eng = Engine(EngineConfig.experiment(exp, batch=B, precision=prec))
st = eng.stats()
print("hbm GB", st.hbm_bytes/1e9, "flops/pair", st.model_flops_per_pair, "issued flops/step", st.gemm_flops_per_step)
for i in range(3):
    eng.generate_batch(0, i); l, _ = eng.step()
print("losses", l.cpu().numpy(), "launches", eng.stats().launches_last_step)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = int(os.environ.get("PROBE_STEPS", "30"))
torch.cuda.synchronize(); e0.record()
for i in range(K):
    eng.generate_batch(0, 3 + i); l, _ = eng.step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
print("%s B=%d %s: %.3f ms/step, %.0f pairs/s, %.1f model TFLOP/s" % (exp, B, prec, ms, B / ms * 1e3, B * st.model_flops_per_pair / ms * 1e-9))
print("losses", l.cpu().numpy())
eng.profile(True)
for i in range(5):
    eng.generate_batch(0, 100 + i); eng.step()
rep = eng.profile_report()
tot = sum(v[0] for v in rep.values())
for k, v in sorted(rep.items(), key=lambda kv: -kv[1][0]):
    print("%-28s %8.3f ms/step  n=%3d  %7.1f TFLOP/s issued" % (k, v[0] / 5, v[1] // 5, v[2] / v[0] * 1e-9 if v[0] > 0 else 0))
print("sum of kernels %.3f ms/step" % (tot / 5))
