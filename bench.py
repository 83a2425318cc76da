#!/usr/bin/env python
"""bench.py -- shape-pair train throughput (pairs/s) of the magic autoencoder step on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--exp 8d] [--batch B_per_gpu] [--precision bf16|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the CPU restatement of the reference step on the host cores

One "step" = TangoEncoder.partial_fit on one batch: forward of three encoder streams + decoder, both backward
passes (cost, training_loss), both Adam updates (8c/vae.py:379-405).  Default workload at N=1 is BASELINE.json
configs[1]: experiment 8d (encoder 500x4, latent 200, cosine metric) at batch 4096 per GPU; with N GPUs the
per-GPU batch stays fixed (weak scaling) and the gradient exchange runs inside the step.

Printed line (rank 0):
  value     pairs/s of the training loop with a NEW batch every step, produced on the device by the generator (row
            N1) inside the timed region.  The generator's shapes are fitted to the reference's images
            (tests/test_generator.py holds the operand-stage density of a generated batch within +-0.05 of real
            batches: 0.94 / 0.96); every operand stage is multiplied (zero-stage skipping is off by default).
  roofline  the dominant kernel: achieved = EXECUTED tensor FLOPs (algorithmic FLOPs x the fraction of operand stages
            that are multiplied) / its average launch time; frac <= 1.  The dict also carries what explains `value`:
            resident_value (two batches kept in HBM and alternated: no generator), zero_skip_value (the optional
            skipping of empty operand stages switched on), density, step_frac_of_peak (model FLOPs of the whole step / sustained bf16 peak),
            issued_over_algorithmic, precision_modes (the 1e-4 "fp32" mode's throughput), configs (the other
            BASELINE.json configs at this N), dp_parity (N > 1: a small data-parallel step against the CPU oracle).
  e2e       the same metric through the public API with HOST inputs: OverlapDataset (the reference's real image
            pairs, tests/golden/overlap_micro_100.npz) -> device_batches (the packed data set sits in pinned host memory; the
            rows of each batch cross PCIe one batch ahead, read by the GPU) -> TangoEncoder.partial_fit -> five Python
            floats per step.
            e2e.float32_feed = partial_fit(X[B,120000] float32) exactly as the reference's main.py feeds it.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "shape-pair train throughput"
UNIT = "pairs/s"
DEFAULT_BATCH = {"8c": 100, "8d": 4096, "8e": 2048, "9a": 8192, "11a": 4096}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--exp", default="8d", choices=sorted(DEFAULT_BATCH))
    ap.add_argument("--batch", type=int, default=0, help="pairs per GPU per step (0 = the config's)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="host-fed steps to time (0 = min(steps, 8))")
    ap.add_argument("--cpu-sample", type=int, default=0, help="pairs per CPU-baseline step (0 = the full per-GPU batch)")
    ap.add_argument("--no-extra-legs", action="store_true", help="skip the fp32-mode, other-config and dp-parity legs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


def workload_name(exp, batch, n):
    return "experiment %s: TangoEncoder.partial_fit, batch %d per GPU x %d GPU(s)" % (exp, batch, n)


# --------------------------------------------------------------------------- CPU legs (oracle as the timed baseline)
def cpu_step_timer(exp: str, sample: int, steps: int, warmup: int, budget_s: float):
    """Times the float32 CPU restatement of the reference step (oracle/tango_oracle.py) on all host threads.
    Returns (pairs_per_s from the MEDIAN step, median ms_per_step, cores, sample, steps timed)."""
    import torch
    from oracle import tango_oracle as orc
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.set_num_threads(cores)
    cfg = orc.OracleConfig.experiment(exp)
    state = orc.OracleState(cfg, seed=0, dtype=torch.float32)
    batch = orc.synthetic_batch(sample, cfg.latent, seed=1, dtype=torch.float32)
    t0 = time.perf_counter()
    state.step(*batch)
    first = time.perf_counter() - t0
    # keep the whole run inside the budget: fewer timed steps first (never below 3), then a smaller sample
    while steps > 3 and first * (steps + warmup) > budget_s:
        steps -= 1
    while sample > 32 and first * (steps + warmup) > budget_s:
        sample //= 2
        batch = orc.synthetic_batch(sample, cfg.latent, seed=1, dtype=torch.float32)
        t0 = time.perf_counter()
        state.step(*batch)
        first = time.perf_counter() - t0
    for _ in range(max(0, warmup - 1)):
        state.step(*batch)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        state.step(*batch)
        times.append(time.perf_counter() - t0)
    med = sorted(times)[len(times) // 2]
    return sample / med, med * 1e3, cores, sample, steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = args.batch or DEFAULT_BATCH[args.exp]
    # the full batch, 3 warm-up + >= 10 timed steps, median (the budget trims steps, then the
    # sample, only if a step is so slow that the run would not end within a few minutes)
    v, ms, cores, sample, steps = cpu_step_timer(args.exp, args.cpu_sample or batch, max(args.steps, 3), max(1, args.warmup), 300.0)
    what = ("float32 torch-CPU restatement of %s/vae.py (oracle/tango_oracle.py), full step (fwd, 2 bwd, 2 Adam) on %s, "
            "median of %d steps, %d threads"
            % (args.exp, "the full %d-pair batch" % batch if sample == batch else "a %d-pair sample of the %d-pair batch" % (sample, batch),
               steps, cores))
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.exp, batch, args.gpus), "experiment": args.exp,
                   "note": "reference TF-1/py2 graph cannot run here (no tensorflow); CPU oracle port timed instead"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": what},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- clocks
CLOCK_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
               "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
               "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples SM clock, power and clock-event reasons of one GPU every 20 ms through NVML (the same counters
    as B200_PROFILING.md's nvidia-smi clocks line) on a thread while the timed region runs."""

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.rows = []
        self.max_mhz = None
        self._stop = False
        self._thr = None
        self._nv = None

    def start(self):
        try:
            import threading
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.idx
            if vis:
                try:
                    idx = int(vis.split(",")[self.idx])
                except Exception:
                    pass
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self._nv = nv

            def loop():
                while not self._stop:
                    try:
                        sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                        pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                        rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        self.rows.append((float(sm), pw, int(rs)))
                    except Exception:
                        pass
                    time.sleep(0.02)

            self._thr = threading.Thread(target=loop, daemon=True)
            self._thr.start()
        except Exception:
            self._thr = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        if self._thr is None:
            return out
        self._stop = True
        self._thr.join(timeout=2)
        nv = self._nv
        masks = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        if self.rows:
            pw = [r[1] for r in self.rows]
            thr = 0.5 * (min(pw) + max(pw))
            load = sorted(r[0] for r in self.rows if r[1] >= thr) or sorted(r[0] for r in self.rows)
            reasons = sorted(k for k, m in masks.items() if any(r[2] & m for r in self.rows))
            out.update(sm_mhz=load[len(load) // 2], reasons=reasons, samples=len(self.rows), power_w_max=max(pw))
        return out


# --------------------------------------------------------------------------- roofline bookkeeping
def algorithmic_work(cfg, B, early_adam=False, fused_wgrad=False):
    """tag -> ('tensor', FLOPs per launch, FLOPs of it that zero-stage skipping applies to) or ('hbm', bytes per
    launch, 0): unpadded, un-split."""
    D, H0, Hl = cfg.D, cfg.encoder_neurons[0], cfg.decoder_neurons[1]
    enc = list(cfg.encoder_neurons)
    n_enc = D * enc[0] + enc[0] + sum(a * b + b for a, b in zip(enc[:-1], enc[1:])) + 2 * (enc[-1] * cfg.latent + cfg.latent)
    dec = list(cfg.decoder_neurons)
    n_dec = cfg.latent * dec[0] + dec[0] + dec[0] * dec[1] + dec[1] + dec[1] * D + D
    n_early = (dec[1] * D + D) if fused_wgrad else n_dec
    # weight gradients of the small layers; on one GPU they are extra problems of the first-layer launch
    small_wg = 2.0 * B * (dec[0] * dec[1] + cfg.latent * dec[0]) + 2.0 * 4 * B * (
        enc[-1] * 2 * cfg.latent + sum(a * b for a, b in zip(enc[:-1], enc[1:])))
    fwd0, wg0 = 2.0 * 3 * B * D * H0, 2.0 * 4 * B * D * H0
    return {
        "gemm.fwd_enc0": ("tensor", fwd0, fwd0),
        "gemm.wgrad_enc0": ("tensor", wg0 + (small_wg if fused_wgrad else 0.0), wg0),
        "gemm.fwd_dec_out_bce": ("tensor", 2.0 * B * Hl * D, 0.0),
        "gemm.wgrad_dec_out": ("tensor", 2.0 * B * Hl * D, 0.0),
        "gemm.dgrad_dec_out": ("tensor", 2.0 * B * Hl * D, 0.0),
        # fused dual Adam: encoder params read p,g1,g2,m1,v1,m2,v2 + write p,m1,v1,m2,v2 (48 B) + bf16 shadow (2 B);
        # decoder params read p,g1,m1,v1 + write p,m1,v1 (28 B) + shadow (2 B)
        # (single GPU: the decoder's output layer runs as its own launch "adam.decoder" under the encoder backward)
        "adam.dual": ("hbm", 50.0 * n_enc + 30.0 * (n_dec - n_early if early_adam else n_dec), 0.0),
        "adam.decoder": ("hbm", 30.0 * n_early, 0.0),
        # generator writes three packed planes (1 bit per pixel)
        "input.generate": ("hbm", 3.0 * B * D / 8, 0.0),
    }


def make_engine(exp, B, precision, rank, world, dev):
    """The experiment directory's own model class (experiments/<exp>/vae.py: TangoEncoder, the drop-in for the
    reference's <exp>/vae.py) -> (engine config, its engine, the model)."""
    import importlib
    mod = importlib.import_module("experiments.%s.vae" % exp)
    vae = mod.TangoEncoder(sess=None, precision=precision, batch_size=B, rank=rank, world=world, device=dev, seed=0,
                           verbose=False)
    if world > 1:
        vae.engine.comm_init_from_process_group()
    return vae.config, vae.engine, vae


class Timer:
    """CUDA-event timing of a loop on the current stream, barrier + synchronize on both sides, max over ranks."""

    def __init__(self, world, dev):
        self.world, self.dev = world, dev

    def barrier(self):
        import torch
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, ms: float) -> float:
        import torch
        if self.world == 1:
            return ms
        import torch.distributed as dist
        t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def time(self, fn, n: int, wall: bool = False) -> float:
        """ms for n calls of fn(i); with wall=True the larger of the device and the host-side time (host-fed loops)."""
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        t0 = time.perf_counter()
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        self.barrier()
        ms = e0.elapsed_time(e1)
        if wall:
            ms = max(ms, (time.perf_counter() - t0) * 1e3)
        return self.max_over_ranks(ms)


def generator_loop(eng, tm, n, warm, step0):
    """ms per step of: generate a new batch (side stream, under the previous step's backward) + step."""
    for i in range(warm):
        eng.generate_batch(0, step0 + i)
        eng.step()

    def one(i):
        eng.generate_batch(0, step0 + warm + i)
        eng.step()
    return tm.time(one, n) / n


def dp_parity_check(rank, world, dev):
    """N > 1: one data-parallel step of a reduced geometry (fused reduce-scatter path, fp32 mode) against the
    full-batch float64 CPU oracle, on the very process group the timed run uses.  Returns {max_rel_err, tol, ...}
    (rank 0's numbers; every rank takes part).  bench.py's CPU legs may use the oracle (test infrastructure)."""
    import numpy as np
    import torch
    from magic_b200.engine import Engine, EngineConfig
    from oracle import tango_oracle as orc
    ocfg = orc.OracleConfig(encoder_neurons=[120, 72], decoder_neurons=[88, 104], n_input=1600, latent=20)
    Bg = 32 * world
    state = orc.OracleState(ocfg, seed=0)
    xl, xr, xk, y, eps = orc.synthetic_batch(Bg, ocfg.latent, seed=1, n_input=ocfg.n_input)
    n = Bg // world
    sl = slice(rank * n, (rank + 1) * n)
    eng = Engine(EngineConfig(D=ocfg.n_input, encoder_neurons=tuple(ocfg.encoder_neurons), decoder_neurons=tuple(ocfg.decoder_neurons),
                              latent=ocfg.latent, precision="fp32", batch=n, batch_global=Bg, rank=rank, world=world, seed=0),
                 device=dev)
    eng.comm_init_from_process_group()
    eng.set_params({k: v.to(torch.float32).numpy() for k, v in state.params.items() if k not in orc.DEAD_PARAMS})
    eng.load_interleaved(orc.interleave(xl[sl], xr[sl], xk[sl]).to(torch.float32))
    losses, _ = eng.step(labels=y[sl].to(torch.float32), eps3=eps[:, sl].to(torch.float32), apply=False)
    got_l = losses.cpu().numpy().astype(np.float64)
    names1 = ("encoder_h0_weights", "decoder_out_mean_weights", "decoder_b1_biases")
    names2 = ("encoder_h0_weights", "encoder_out_log_sigma_weights")
    got1 = {k: eng.get_grad(1, k).astype(np.float64) for k in names1}     # collective: every rank calls
    got2 = {k: eng.get_grad(2, k).astype(np.float64) for k in names2}
    eng.close()
    if rank != 0:
        return None
    out, g1, g2 = state.grads(xl, xr, xk, y, eps)
    errs = {}
    for i, k in enumerate(orc.LOSS_KEYS):
        errs["loss." + k] = abs(got_l[i] - float(out[k])) / abs(float(out[k]))
    for which, got, ref in ((1, got1, g1), (2, got2, g2)):
        for k, v in got.items():
            r = ref[k].numpy()
            errs["g%d.%s" % (which, k)] = float(np.linalg.norm(v - r) / np.linalg.norm(r))
    worst = max(errs, key=errs.get)
    return {"max_rel_err": errs[worst], "worst": worst, "tol": 1e-4, "ok": bool(errs[worst] <= 1e-4), "world": world,
            "what": "one fp32-mode data-parallel step (fused reduce-scatter path) of a reduced geometry, global batch %d, "
                    "vs the full-batch float64 CPU oracle: 5 losses + 5 gradient tensors, relative L2" % Bg}


# --------------------------------------------------------------------------- the B200 arm
PER_BATCH_LAUNCHES = 1     # generate_pairs (zero-stage maps are off by default)


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE line, the JSON: whatever native libraries print there (NCCL's version banner at
    # communicator creation) goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch N>1 with: python -m torch.distributed.run --nnodes=1 --nproc-per-node %d "
                         "--master-addr 127.0.0.1 --master-port P bench.py --gpus %d ..." % (args.gpus, args.gpus))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the engine has no CPU path (use --impl reference for the CPU leg)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    tm = Timer(world, dev)
    pk = peaks()
    extra = not args.no_extra_legs

    # ---- N > 1: prove the data-parallel path before timing it (the driver's GPU test box has one GPU)
    dp_parity = None
    if world > 1 and extra:
        dp_parity = dp_parity_check(rank, world, dev)

    B = args.batch or DEFAULT_BATCH[args.exp]
    cfg, eng, args._vae = make_engine(args.exp, B, args.precision, rank, world, dev)
    st = eng.stats()
    K, W = args.steps, max(args.warmup, 3)

    # ---- leg 1 (headline `value`): a NEW generated batch every step, generator and per-batch map / list / work-order
    # kernels inside the timed region.  N > 1: the first few dozen steps of a process group run slow (lazily mapped
    # CUDA-IPC peer memory is touched page by page, NCCL brings up its channels: 2.29 vs 1.90 ms per step at 8 GPUs,
    # profiles/r01p_bench_n8*.json); those settling steps run on top of the W warm-up steps, untimed.
    settle = 40 if world > 1 else 0
    for i in range(W + settle):
        eng.generate_batch(0, i)
        eng.step()
    sampler = ClockSampler(local)
    sampler.start()

    def head(i):
        eng.generate_batch(0, 1000 + i)
        head.losses, _ = eng.step()
    ms_total = tm.time(head, K)
    clocks = sampler.stop()
    launches_per_step = int(eng.stats().launches_last_step) + PER_BATCH_LAUNCHES
    final_losses = head.losses.cpu().numpy().tolist()
    ms_step = ms_total / K
    value = B * world * K / (ms_total * 1e-3)
    density = eng.input_density()

    # ---- leg 1b: the same loop with a CUDA-event pair around every launch (engine profile mode): per-kernel device
    # times for the roofline and the shares.  Kept out of the headline loop because an event between two kernels
    # disables their programmatic dependent launch overlap.
    Kp = max(3, min(K, 40))
    eng.profile(True)
    eng.generate_batch(0, 5000)
    eng.step()
    eng.profile_report()          # warm the event pool, drop the warm-up record

    def prof(i):
        eng.generate_batch(0, 5001 + i)
        eng.step()
    ms_step_events = tm.time(prof, Kp) / Kp
    rep = eng.profile_report()
    eng.profile(False)

    # ---- leg 1c: two batches resident in HBM, alternated (no generator, no per-batch kernels inside the timed region)
    Kr = max(3, min(K, 60))
    eng.generate_batch(0, 7000)
    eng.generate_batch(0, 7001)
    for i in range(3):
        eng.swap_batch()
        eng.step()

    def resident(i):
        eng.swap_batch()
        eng.step()
    ms_res = tm.time(resident, Kr) / Kr

    # ---- leg 1d: the headline loop with zero-stage skipping switched on (per-batch maps, lists and work order + skipped
    # stages; off by default because the reference's images fill 94-97 % of the stages)
    eng.set_option("skip_zero_stages", 1)
    ms_skip = generator_loop(eng, tm, Kr, 3, 8000)
    eng.set_option("skip_zero_stages", 0)

    # ---- roofline of the dominant kernel family (by summed device time in the timed region)
    fused_wgrad = world == 1 and os.environ.get("MAGIC_FUSE_WGRAD", "1") != "0"
    work = algorithmic_work(cfg, B, early_adam=(world == 1 and os.environ.get("MAGIC_EARLY_ADAM", "1") != "0"),
                            fused_wgrad=fused_wgrad)
    skipping = os.environ.get("MAGIC_SKIP_ZERO", "0") != "0" and os.environ.get("MAGIC_BITS_A", "1") != "0"
    roof = None
    shares = {}
    tot_ms = sum(v[0] for v in rep.values()) or 1.0
    ranked = sorted(rep.items(), key=lambda kv: -kv[1][0])
    for tag, (ms, n, _issued) in ranked:
        shares[tag] = round(ms / tot_ms, 4)
    for tag, (ms, n, _issued) in ranked:
        # side-stream kernels (generator, per-batch maps, the decoder's early Adam) overlap the step: their event times
        # are not exclusive, so the dominant kernel is taken from the step's own stream
        if tag not in work or n == 0 or tag.startswith("input.") or tag == "adam.decoder":
            continue
        kind, amount, skippable = work[tag]
        avg_s = ms / n * 1e-3
        traffic, traffic_src = None, None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            ent = tj.get("%s/%d/%s" % (args.exp, B, args.precision), {}).get(tag) if world == 1 else None
            if ent:
                traffic, traffic_src = ent["bytes"], ent["source"]
        except Exception:
            traffic = None
        if kind == "tensor":
            f = (density[1] if tag == "gemm.wgrad_enc0" else density[0]) if (skipping and skippable) else 1.0
            executed = amount - skippable * (1.0 - f)
            ach, peak, unit = executed / avg_s * 1e-12, pk["tf_sustained"], "TFLOP/s"
        else:
            f, executed = 1.0, amount
            ach, peak, unit = amount / avg_s * 1e-9, pk["hbm"], "GB/s"
        roof = {"kernel": tag, "bound": "tensor" if kind == "tensor" else "hbm", "achieved": round(ach, 2), "peak": peak,
                "unit": unit, "frac": round(ach / peak, 4), "traffic": traffic,
                "traffic_unit": "dram bytes per launch (ncu)", "traffic_source": traffic_src,
                "avg_launch_ms": round(ms / n, 5), "launches_timed": n,
                "measured_over": "%d steps of the headline loop with an event pair around every launch" % Kp,
                "peak_source": pk["source"] + (" (sustained)" if kind == "tensor" else ""),
                "algorithmic_per_launch": amount, "executed_per_launch": executed, "executed_fraction_of_main_problem": round(f, 4),
                "dense_equivalent_frac": round(amount / avg_s * (1e-12 if kind == "tensor" else 1e-9) / peak, 4),
                "share_of_step": shares[tag],
                "peak_burst": pk["tf_burst"] if kind == "tensor" else None,
                "frac_of_burst_peak": round(ach / pk["tf_burst"], 4) if kind == "tensor" else None,
                "note_peak": "peak = cuBLAS bf16 8192^3 back to back for seconds (power-capped clocks), as the contract asks for a "
                             "kernel timed inside a long step; this launch sits between HBM-bound and latency-bound kernels of the "
                             "step, i.e. with power headroom, so frac can pass 1.0 -- frac_of_burst_peak (cuBLAS best of 10) is the "
                             "utilisation of the tensor pipe proper",
                "note": "achieved = executed FLOPs / launch time.  Zero-stage skipping is off by default (every stage is "
                        "multiplied: executed = algorithmic FLOPs of SURVEY 8.4); with MAGIC_SKIP_ZERO=1 executed counts only the "
                        "fraction of the main problem's 128x64 operand stages that hold a set pixel"}
        break
    model_tflops = value / world * st.model_flops_per_pair * 1e-12
    if roof is None:
        roof = {"kernel": None, "bound": "tensor", "achieved": round(model_tflops, 2), "peak": pk["tf_sustained"],
                "unit": "TFLOP/s", "frac": round(model_tflops / pk["tf_sustained"], 4), "traffic": None}
    roof.update({
        "step_model_tflops_per_gpu": round(model_tflops, 1),
        "step_frac_of_peak": round(model_tflops / pk["tf_sustained"], 4),
        "step_frac_of_burst_peak": round(model_tflops / pk["tf_burst"], 4),
        "issued_over_algorithmic": round(st.gemm_flops_per_step / (st.model_flops_per_pair * B), 4),
        "density_fwd": round(density[0], 4), "density_wgrad": round(density[1], 4),
        "resident_value": round(B * world / (ms_res * 1e-3), 1), "resident_ms_per_step": round(ms_res, 4),
        "zero_skip_value": round(B * world / (ms_skip * 1e-3), 1), "zero_skip_ms_per_step": round(ms_skip, 4),
        "ms_per_step_with_per_launch_events": round(ms_step_events, 4),
        "legs": "value: new generated batch every step (generator timed), every operand stage multiplied; resident_value: two "
                "batches kept in HBM and alternated, %d steps; zero_skip_value: the headline loop with skip_zero_stages = 1 "
                "(per-batch maps + skipped empty stages), %d steps" % (Kr, Kr),
    })
    if dp_parity is not None:
        roof["dp_parity"] = dp_parity

    # ---- leg 2: end to end through the public API, HOST inputs: the reference's real image pairs (OverlapDataset) ->
    # device_batches (the batch's rows read from the pinned packed data set over PCIe, one batch ahead) -> TangoEncoder.partial_fit
    e2e = None
    if not args.no_e2e:
        from magic_b200.overlap_input import OverlapDataset, device_batches
        g = np.load(os.path.join(ROOT, "tests", "golden", "overlap_micro_100.npz"))
        side = int(round(cfg.D ** 0.5))
        Lr = np.unpackbits(g["lock_bits"], axis=1)[:, :cfg.D]
        Kr_ = np.unpackbits(g["key_bits"], axis=1)[:, :cfg.D]
        areas = np.load(os.path.join(ROOT, "tests", "golden", "overlap_areas.npy"))[:len(Lr)].astype(np.float32)
        ds = OverlapDataset.from_arrays(Lr, Kr_, areas, image_size=side)
        vae = args._vae
        Ke = args.e2e_steps or max(8, min(K, 50))
        it = device_batches(vae, ds, seed=rank, steps=Ke + 4)
        for _ in range(3):
            b = next(it)
            vae.partial_fit(b, overlap_areas=b.labels)
        real_density = eng.input_density()

        def fed(i):
            b = next(it)
            fed.out = vae.partial_fit(b, overlap_areas=b.labels)      # returns this step's losses as Python floats
        ms_fed = tm.time(fed, Ke, wall=True)
        wpr = ds.lock_bits.shape[1]
        e2e = {"value": B * world * Ke / (ms_fed * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(2 * B * ((wpr + 3) // 4 * 4) * 4 + 2 * B * 4), "d2h_bytes_per_step": 20,
               "steps": Ke, "ms_per_step": ms_fed / Ke,
               "api": "OverlapDataset(real overlap_micro pairs, sampled with replacement) -> device_batches one batch ahead: rows gathered "
                      "into a pinned slot by host threads + magic_feed_pairs_bits_host (copy engine) when this rank has >= 6 host "
                      "threads, else magic_feed_pairs_bits_gather (the GPU reads the rows from the pinned data set over PCIe) "
                      "-> TangoEncoder.partial_fit "
                      "(magic_step; rotated lock made on the device) -> 5 Python floats every step (magic_read_losses: the call returns once "
                      "the forward pass has produced them, the backward pass and the update overlap the host's next call; the "
                      "timed region ends with a device synchronize)",
               "data": "real: tests/golden/overlap_micro_100.npz + OVERLAP_AREAS[:100]",
               "note": "may exceed `value`: value's loop runs the on-device generator (shapes, rotation, max-overlap label search: a "
                       "0.7 ms kernel under every step), this loop copies 41 MB and runs the lighter rotation / layout kernel",
               "input_density": [round(real_density[0], 4), round(real_density[1], 4)]}
        # the reference-signature call: partial_fit(X[B,120000] float32) from pinned host memory (PCIe-bound)
        X_dev, y_dev = eng.export_batch()
        X_host = torch.empty((B, 3 * cfg.D), dtype=torch.float32, pin_memory=True)
        y_host = torch.empty((B,), dtype=torch.float32, pin_memory=True)
        X_host.copy_(X_dev)
        y_host.copy_(y_dev)
        del X_dev
        torch.cuda.synchronize()
        Xn, yn = X_host.numpy(), y_host.numpy()
        eng.partial_fit_host(Xn, yn)
        Kf = 4
        ms_f32 = tm.time(lambda i: eng.partial_fit_host(Xn, yn), Kf, wall=True)
        e2e["float32_feed"] = {"value": B * world * Kf / (ms_f32 * 1e-3), "unit": UNIT,
                               "h2d_bytes_per_step": int(Xn.nbytes + yn.nbytes), "d2h_bytes_per_step": 20, "steps": Kf,
                               "ms_per_step": ms_f32 / Kf,
                               "api": "magic_partial_fit_host(X[B,120000] fp32 pinned, overlap_areas[B]) -> 5 losses: the reference's "
                                      "partial_fit(X) signature unchanged, 480 KB per pair, PCIe-bound (%.0f GB/s)"
                                      % ((Xn.nbytes + yn.nbytes) / (ms_f32 / Kf * 1e-3) * 1e-9)}
        del X_host
    eng.close()

    # ---- extra legs (short): the fp32-equivalent mode on the same workload, and the other BASELINE.json configs
    precision_modes, configs = {}, {}
    if extra:
        def short_leg(exp, b, precision, steps=40):
            c2, e2, _ = make_engine(exp, b, precision, rank, world, dev)
            s2 = e2.stats()
            ms = generator_loop(e2, tm, steps, 8 + settle, 0)
            e2.close()
            tf = b / (ms * 1e-3) * s2.model_flops_per_pair * 1e-12
            return {"value": round(b * world / (ms * 1e-3), 1), "unit": UNIT, "ms_per_step": round(ms, 4), "steps": steps,
                    "batch_per_gpu": b, "n_gpus": world, "precision": precision,
                    "model_tflops_per_gpu": round(tf, 1), "frac_of_bf16_sustained_peak": round(tf / pk["tf_sustained"], 4),
                    "issued_over_algorithmic": round(s2.gemm_flops_per_step / (s2.model_flops_per_pair * b), 4)}
        if args.precision == "bf16":
            # the 1e-4 mode: bf16x3 split operands (hi*hi + hi*lo + lo*hi; binary inputs need two terms), fp32 accumulate.
            # Its peak: the same tensor pipe issues 2-3 bf16 MMAs per algorithmic one, so the bf16 peak / issued ratio; the
            # TF32 cuBLAS peak (SURVEY 8.4 recipe: torch.matmul 8192^3, allow_tf32) is measured beside it for reference.
            leg = short_leg(args.exp, B, "fp32", steps=30)
            try:
                torch.backends.cuda.matmul.allow_tf32 = True
                a = torch.randn(8192, 8192, device=dev)
                b_ = torch.randn(8192, 8192, device=dev)
                best = 1e9
                for _ in range(12):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    torch.matmul(a, b_)
                    e1.record()
                    torch.cuda.synchronize()
                    best = min(best, e0.elapsed_time(e1))
                del a, b_
                leg["tf32_cublas_peak_tflops"] = round(2 * 8192 ** 3 / (best * 1e-3) * 1e-12, 1)
                leg["frac_of_tf32_peak"] = round(leg["model_tflops_per_gpu"] / leg["tf32_cublas_peak_tflops"], 4)
            except Exception as ex:       # the peak is context, not a result
                leg["tf32_cublas_peak_tflops"] = None
                leg["tf32_peak_error"] = str(ex)[:80]
            leg["issued_frac_of_bf16_sustained_peak"] = round(leg["frac_of_bf16_sustained_peak"] * leg["issued_over_algorithmic"], 4)
            precision_modes["fp32"] = leg
        others = [e for e in ("8e", "9a", "11a") if e != args.exp] + (["8c"] if world == 1 and args.exp != "8c" else [])
        for e in others:
            configs["%s@%d/gpu" % (e, DEFAULT_BATCH[e])] = short_leg(e, DEFAULT_BATCH[e], "bf16")
        roof["precision_modes"] = precision_modes
        roof["configs"] = configs

    # ---- CPU baseline beside it (rank 0, N=1 only): the full batch unless a step is too slow for the time budget
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, ms, cores, sample, nst = cpu_step_timer(args.exp, args.cpu_sample or B, 5, 1, 45.0)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "median of %d full steps of the float32 torch-CPU oracle on %s (%.0f ms/step, %d threads)"
                         % (nst, "the full %d-pair batch" % B if sample == B else "a %d-pair sample of the batch" % sample, ms, cores)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.precision == "bf16" else "bf16x3 (fp32-equivalent split, fp32 accumulate)",
            "data": "synthetic",
            "config": {"workload": workload_name(args.exp, B, world), "experiment": args.exp,
                       "global_batch": B * world, "encoder": list(cfg.encoder_neurons), "latent": cfg.latent,
                       "metric_loss": cfg.metric + ("-reciprocal" if cfg.reciprocal else ""),
                       "parallelism": "dp%d" % world,
                       "l2": "no flush needed: %.0f MB of weights, gradients and Adam state streamed per step > 126 MB L2"
                             % ((50.0 * st.n_params2 + 30.0 * (st.n_params1 - st.n_params2)) / 1e6),
                       "settling_steps": settle,
                       "inputs": "a new generated batch every step, inside the timed region"},
            "roofline": roof, "kernel_time_shares": shares,
            "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks,
            "gpu_launches": launches_per_step * K, "gpu_launches_per_step": launches_per_step,
            "final_losses": final_losses,
        }
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
