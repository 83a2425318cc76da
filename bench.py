#!/usr/bin/env python
"""bench.py -- shape-pair train throughput (pairs/s) of the magic autoencoder step on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--exp 8d] [--batch B_per_gpu] [--precision bf16|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the CPU restatement of the reference step on the host cores

One "step" = TangoEncoder.partial_fit on one batch: forward of three encoder streams + decoder, both backward
passes (cost, training_loss), both Adam updates (8c/vae.py:379-405).  Default workload at N=1 is BASELINE.json
configs[1]: experiment 8d (encoder 500x4, latent 200, cosine metric) at batch 4096 per GPU; with N GPUs the
per-GPU batch stays fixed (weak scaling) and gradients are all-reduced over NCCL inside the step.

Printed line (rank 0): value = pairs/s of the step with its inputs resident in HBM (two generated batches kept in
the engine's input buffers and alternated; `with_generator` = the loop that also generates a new batch every step).
The images are binary and mostly empty, the engine keeps them packed and its first-layer GEMMs skip operand stages
without a set pixel (bit-identical results); `dense_inputs` = the same loop with that switched off (every stage
multiplied), `input_density` = the fraction of stages that hold a set pixel;
e2e = pairs/s of the training loop fed from pinned HOST memory, H2D of every step's inputs + step + D2H of the
losses inside the timed region: headline = the dataset path (bit-packed lock/key pairs + labels through
magic_feed_pairs_bits_host, one batch ahead; the rotated lock is made on the device as the reference's pipeline makes
it on the CPU); e2e.float32_feed = magic_partial_fit_host with X[B,120000] fp32 exactly as main.py feeds partial_fit
(PCIe-bound: 480 KB per pair), e2e.uint8_feed = the same array as uint8.
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
    ap.add_argument("--cpu-sample", type=int, default=256, help="pairs per CPU-baseline step")
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
    Returns (pairs_per_s, ms_per_step, cores, sample)."""
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
    # keep the whole run inside the budget: shrink the sample if the first step says it will not fit
    while sample > 32 and first * (steps + warmup) > budget_s:
        sample //= 2
        batch = orc.synthetic_batch(sample, cfg.latent, seed=1, dtype=torch.float32)
        t0 = time.perf_counter()
        state.step(*batch)
        first = time.perf_counter() - t0
    for _ in range(max(0, warmup - 1)):
        state.step(*batch)
    t0 = time.perf_counter()
    for _ in range(steps):
        state.step(*batch)
    dt = time.perf_counter() - t0
    return sample * steps / dt, dt / steps * 1e3, cores, sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = args.batch or DEFAULT_BATCH[args.exp]
    v, ms, cores, sample = cpu_step_timer(args.exp, args.cpu_sample, args.steps, max(1, args.warmup), 240.0)
    what = ("float32 torch-CPU restatement of %s/vae.py (oracle/tango_oracle.py), full step "
            "(fwd, 2 bwd, 2 Adam) on a %d-pair sample of the batch, %d threads" % (args.exp, sample, cores))
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
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
    """tag -> ('tensor', FLOPs per launch) or ('hbm', bytes per launch): unpadded, un-split."""
    D, H0, Hl = cfg.D, cfg.encoder_neurons[0], cfg.decoder_neurons[1]
    enc = list(cfg.encoder_neurons)
    n_enc = D * enc[0] + enc[0] + sum(a * b + b for a, b in zip(enc[:-1], enc[1:])) + 2 * (enc[-1] * cfg.latent + cfg.latent)
    dec = list(cfg.decoder_neurons)
    n_dec = cfg.latent * dec[0] + dec[0] + dec[0] * dec[1] + dec[1] + dec[1] * D + D
    n_early = (dec[1] * D + D) if fused_wgrad else n_dec
    # weight gradients of the small layers; on one GPU in bf16 mode they are extra problems of the first-layer launch
    small_wg = 2.0 * B * (dec[0] * dec[1] + cfg.latent * dec[0]) + 2.0 * 4 * B * (
        enc[-1] * 2 * cfg.latent + sum(a * b for a, b in zip(enc[:-1], enc[1:])))
    return {
        "gemm.fwd_enc0": ("tensor", 2.0 * 3 * B * D * H0),
        "gemm.wgrad_enc0": ("tensor", 2.0 * 4 * B * D * H0 + (small_wg if fused_wgrad else 0.0)),
        "gemm.fwd_dec_out_bce": ("tensor", 2.0 * B * Hl * D),
        "gemm.wgrad_dec_out": ("tensor", 2.0 * B * Hl * D),
        "gemm.dgrad_dec_out": ("tensor", 2.0 * B * Hl * D),
        # fused dual Adam: encoder params read p,g1,g2,m1,v1,m2,v2 + write p,m1,v1,m2,v2 (48 B) + bf16 shadow (2 B);
        # decoder params read p,g1,m1,v1 + write p,m1,v1 (28 B) + shadow (2 B)
        # (single GPU: the decoder range runs as its own launch "adam.decoder" under the encoder backward)
        # (with the small wgrads folded into the last launch only the decoder's output layer is final early)
        "adam.dual": ("hbm", 50.0 * n_enc + 30.0 * (n_dec - n_early if early_adam else n_dec)),
        "adam.decoder": ("hbm", 30.0 * n_early),
        # generator writes three packed planes (1 bit per pixel)
        "input.generate": ("hbm", 3.0 * B * D / 8),
    }


# --------------------------------------------------------------------------- the B200 arm
def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from magic_b200.engine import Engine, EngineConfig

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

    B = args.batch or DEFAULT_BATCH[args.exp]
    cfg = EngineConfig.experiment(args.exp, batch=B, precision=args.precision, rank=rank, world=world, seed=0)
    eng = Engine(cfg, device=dev)
    if world > 1:
        eng.comm_init_from_process_group()
    st = eng.stats()
    K, W = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- leg 1 (headline `value`): the hot path with its inputs resident in HBM.  Two generated batches (983 MB of
    # planes each at 8d/4096: together 15x the L2) stay in the engine's two input buffers and are alternated with
    # magic_swap_batch, so every step reads a batch that the previous step evicted from L2.  No events inside.
    eng.generate_batch(0, 0)
    eng.generate_batch(0, 1)
    # N > 1: the first few dozen steps of a process group run slow (lazily mapped CUDA-IPC peer memory is touched page by
    # page, NCCL brings up its channels): 8 GPUs measured 2.29 ms per step over steps 6..105 and 1.90 ms over steps
    # 41..240 (profiles/r01p_bench_n8*.json).  Those settling steps are run on top of the W warm-up steps, untimed.
    settle = 40 if world > 1 else 0
    for i in range(W + settle):
        eng.swap_batch()
        eng.step()
    sampler = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    e0.record()
    for i in range(K):
        eng.swap_batch()
        losses, _ = eng.step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches_per_step = int(eng.stats().launches_last_step)
    final_losses = losses.cpu().numpy().tolist()
    ms_step = ms_total / K
    value = B * world * K / (ms_total * 1e-3)

    # ---- leg 1b: the same loop with a CUDA-event pair around every launch (engine profile mode): per-kernel device
    # times for the roofline and the shares.  Kept out of the headline loop because an event between two kernels
    # disables their programmatic dependent launch overlap.
    Kp = max(3, min(K, 40))
    eng.profile(True)
    eng.swap_batch()
    eng.step()
    eng.profile_report()          # warm the event pool, drop the warm-up record
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e2.record()
    for i in range(Kp):
        eng.swap_batch()
        eng.step()
    e3.record()
    barrier()
    ms_step_events = max_over_ranks(e2.elapsed_time(e3)) / Kp
    rep = eng.profile_report()
    eng.profile(False)

    # ---- leg 1c: the generator inside the loop (row N1): a NEW batch every step, produced on the device by the
    # engine's generator on its own stream under the previous step's backward pass
    Kg = max(3, min(K, 100))
    for i in range(3):
        eng.generate_batch(0, 10 + i)
        eng.step()
    e8, e9 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e8.record()
    for i in range(Kg):
        eng.generate_batch(0, 100 + i)
        eng.step()
    e9.record()
    barrier()
    ms_gen = max_over_ranks(e8.elapsed_time(e9)) / Kg
    with_generator = {"value": B * world / (ms_gen * 1e-3), "unit": UNIT, "ms_per_step": ms_gen, "steps": Kg,
                      "note": "new generated batch every step inside the timed region (generator kernel + step)"}

    # ---- leg 1d: the same resident-input loop with every operand stage processed (no zero-stage skipping)
    eng.swap_batch()
    density = eng.input_density()
    Kd = max(3, min(K, 60))
    eng.set_option("skip_zero_stages", 0)
    for i in range(3):
        eng.swap_batch()
        eng.step()
    e10, e11 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e10.record()
    for i in range(Kd):
        eng.swap_batch()
        eng.step()
    e11.record()
    barrier()
    ms_dense = max_over_ranks(e10.elapsed_time(e11)) / Kd
    eng.set_option("skip_zero_stages", 1)
    dense_inputs = {"value": B * world / (ms_dense * 1e-3), "unit": UNIT, "ms_per_step": ms_dense, "steps": Kd,
                    "note": "skip_zero_stages = 0: every 128x64 operand stage of the first layer is expanded and multiplied "
                            "(what an input without empty regions would cost)"}

    # ---- roofline of the dominant kernel family (by summed device time in the timed region)
    pk = peaks()
    fused_wgrad = world == 1 and args.precision == "bf16" and os.environ.get("MAGIC_FUSE_WGRAD", "1") != "0"
    work = algorithmic_work(cfg, B, early_adam=(world == 1 and os.environ.get("MAGIC_EARLY_ADAM", "1") != "0"),
                            fused_wgrad=fused_wgrad)
    roof = None
    shares = {}
    tot_ms = sum(v[0] for v in rep.values()) or 1.0
    for tag, (ms, n, _issued) in sorted(rep.items(), key=lambda kv: -kv[1][0]):
        shares[tag] = round(ms / tot_ms, 4)
    for tag, (ms, n, _issued) in sorted(rep.items(), key=lambda kv: -kv[1][0]):
        # the generator runs on its own stream, spread under the backward pass at one CTA per SM: its event time is
        # not exclusive, so the dominant kernel is taken from the step's own stream
        if tag in work and n > 0 and tag not in ("input.generate", "adam.decoder"):   # side-stream kernels overlap
            kind, amount = work[tag]
            avg_s = ms / n * 1e-3
            if kind == "tensor":
                ach, peak, unit = amount / avg_s * 1e-12, pk["tf_sustained"], "TFLOP/s"
            else:
                ach, peak, unit = amount / avg_s * 1e-9, pk["hbm"], "GB/s"
            traffic, traffic_src = None, None
            try:
                tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
                ent = tj.get("%s/%d/%s" % (args.exp, B, args.precision), {}).get(tag) if world == 1 else None
                if ent:
                    traffic, traffic_src = ent["bytes"], ent["source"]
            except Exception:
                traffic = None
            roof = {"kernel": tag, "bound": kind, "achieved": round(ach, 2), "peak": peak, "unit": unit,
                    "frac": round(ach / peak, 4), "traffic": traffic, "traffic_unit": "dram bytes per launch (ncu)", "traffic_source": traffic_src, "avg_launch_ms": round(ms / n, 5),
                    "launches_timed": n, "measured_over": "%d steps with an event pair around every launch, run right "
                    "after the timed region on the same inputs" % Kp, "peak_source": pk["source"] + (" (sustained)" if kind == "tensor" else ""),
                    "algorithmic_per_launch": amount, "share_of_step": shares[tag]}
            if tag in ("gemm.wgrad_enc0", "gemm.fwd_enc0") and os.environ.get("MAGIC_SKIP_ZERO", "1") != "0" \
                    and os.environ.get("MAGIC_BITS_A", "1") != "0":
                f = density[1] if tag == "gemm.wgrad_enc0" else density[0]
                roof["executed_fraction"] = round(f, 4)
                roof["achieved_executed"] = round(ach * f, 2)
                roof["frac_executed"] = round(ach * f / peak, 4)
                roof["note"] = ("achieved counts the algorithmic (dense) FLOPs of SURVEY 8.4; %.1f %% of this GEMM's operand "
                                "stages hold a set pixel and are multiplied, the rest are skipped as exact zeros, so the "
                                "tensor pipe sustains achieved_executed (a fraction > 1 of the dense roofline is the "
                                "sparsity of the images, not a faster pipe)" % (100 * f))
            break
    model_tflops = value / world * st.model_flops_per_pair * 1e-12

    # ---- leg 2: end to end through the host-buffer C-ABI call, pinned host inputs
    e2e = None
    if not args.no_e2e:
        Ke = args.e2e_steps or min(K, 8)
        X_dev, y_dev = eng.export_batch()
        X_host = torch.empty((B, 3 * cfg.D), dtype=torch.float32, pin_memory=True)
        y_host = torch.empty((B,), dtype=torch.float32, pin_memory=True)
        X_host.copy_(X_dev)
        y_host.copy_(y_dev)
        del X_dev
        torch.cuda.synchronize()
        Xn, yn = X_host.numpy(), y_host.numpy()
        for _ in range(2):
            eng.partial_fit_host(Xn, yn)
        barrier()
        t0 = time.perf_counter()
        e4, e5 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e4.record()
        for _ in range(Ke):
            hl, _ = eng.partial_fit_host(Xn, yn)      # H2D X+labels, step, D2H losses, stream sync
        e5.record()
        torch.cuda.synchronize()
        ms_e2e = max_over_ranks(max(e4.elapsed_time(e5), (time.perf_counter() - t0) * 1e3))
        f32_feed = {"value": B * world * Ke / (ms_e2e * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": int(Xn.nbytes + yn.nbytes), "d2h_bytes_per_step": 20,
                    "steps": Ke, "ms_per_step": ms_e2e / Ke,
                    "api": "magic_partial_fit_host(X[B,120000] fp32 pinned, overlap_areas[B]) -> 5 losses",
                    "note": "the reference's partial_fit(X) signature unchanged: 480 KB per pair, PCIe-bound (%.0f GB/s)"
                            % ((Xn.nbytes + yn.nbytes) / (ms_e2e / Ke * 1e-3) * 1e-9)}
        e2e = {}

        # same call with the compact feed the same signature accepts: uint8 X (feed_dict casts to the placeholder dtype)
        X8 = torch.empty((B, 3 * cfg.D), dtype=torch.uint8, pin_memory=True)
        X8.copy_(X_host.to(torch.uint8))
        X8n = X8.numpy()
        for _ in range(2):
            eng.partial_fit_host(X8n, yn)
        barrier()
        e6, e7 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e6.record()
        for _ in range(Ke):
            eng.partial_fit_host(X8n, yn)
        e7.record()
        torch.cuda.synchronize()
        ms_u8 = max_over_ranks(max(e6.elapsed_time(e7), (time.perf_counter() - t0) * 1e3))
        e2e["uint8_feed"] = {"value": B * world * Ke / (ms_u8 * 1e-3), "unit": UNIT, "ms_per_step": ms_u8 / Ke,
                             "h2d_bytes_per_step": int(X8n.nbytes + yn.nbytes), "d2h_bytes_per_step": 20,
                             "api": "magic_partial_fit_host_u8(X[B,120000] uint8 pinned, overlap_areas[B]) -> 5 losses"}
        # the dataset path (row N2): bit-packed lock/key pairs + labels from pinned host memory, the rotated lock made on
        # the device; batch n+1 is sent on the input stream while step n runs, losses read back every step
        from magic_b200.overlap_input import pack_planes
        Xi = X_host.view(B, cfg.D, 3)
        lock_b = torch.from_numpy(pack_planes(Xi[:, :, 0].numpy()).view(np.int32)).pin_memory()
        key_b = torch.from_numpy(pack_planes(Xi[:, :, 2].numpy()).view(np.int32)).pin_memory()
        h_loss = torch.empty(5, dtype=torch.float32, pin_memory=True)

        def packed_steps(n_steps, s0):
            eng.feed_pairs_bits(lock_b, key_b, y_host, 1, s0)
            for i in range(n_steps):
                eng.feed_pairs_bits(lock_b, key_b, y_host, 1, s0 + i + 1, activate=False)
                ld, _ = eng.step()
                h_loss.copy_(ld, non_blocking=True)
                torch.cuda.current_stream().synchronize()           # the caller reads the losses of every step
                eng.swap_batch()
        Kp8 = max(Ke, min(K, 50))
        packed_steps(3, 0)
        barrier()
        e8, e9 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e8.record()
        packed_steps(Kp8, 100)
        e9.record()
        torch.cuda.synchronize()
        ms_pk = max_over_ranks(max(e8.elapsed_time(e9), (time.perf_counter() - t0) * 1e3))
        # headline = the training loop a user of the drop-in runs (experiments/*/main.py with a dataset,
        # INTEGRATION.md section 2): input pipeline + partial_fit; the two array-fed forms of partial_fit beside it
        e2e = {
            "value": B * world * Kp8 / (ms_pk * 1e-3), "unit": UNIT,
            "h2d_bytes_per_step": int(lock_b.numel() * 4 + key_b.numel() * 4 + y_host.numel() * 4), "d2h_bytes_per_step": 20,
            "steps": Kp8, "ms_per_step": ms_pk / Kp8,
            "api": "magic_feed_pairs_bits_host(lock/key bits [B,%d] u32 pinned, overlap_areas[B]) one batch ahead + magic_step "
                   "-> 5 losses read back every step (rotated lock made on the device; OverlapDataset / device_batches path)"
                   % lock_b.shape[1],
            "float32_feed": f32_feed, "uint8_feed": e2e["uint8_feed"]}
        del X8, X_host

    # ---- CPU baseline beside it (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, ms, cores, sample = cpu_step_timer(args.exp, args.cpu_sample, 8, 2, 40.0)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "8 full steps of the float32 torch-CPU oracle on a %d-pair sample of the batch "
                         "(%.0f ms/step, %d threads)" % (sample, ms, cores)}

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
                       "l2": "operands per step (weights, gradients and Adam state: %.0f MB streamed per step; packed input rows "
                             "%.0f MB per batch) exceed the 126 MB L2; no flush needed"
                             % ((50.0 * st.n_params2 + 30.0 * (st.n_params1 - st.n_params2)) / 1e6, 3 * B * cfg.D / 8 / 1e6),
                       "settling_steps": settle,
                       "zero_stage_skipping": "on: first-layer operand stages (128x64 pixels) without a set pixel add exact "
                                              "zeros and are skipped, results bit-identical; `dense_inputs` is the same loop with "
                                              "every stage multiplied, `input_density` the fraction of stages that are not empty",
                       "inputs": "two generated batches resident in HBM, alternated every step (magic_swap_batch); "
                                 "`with_generator` times the loop that generates a new batch every step"},
            "model_tflops_per_gpu": round(model_tflops, 1),
            "frac_of_bf16_sustained_peak": round(model_tflops / pk["tf_sustained"], 4),
            "with_generator": with_generator,
            "dense_inputs": dense_inputs,
            "input_density": {"forward": round(density[0], 4), "weight_gradient": round(density[1], 4),
                              "what": "fraction of the first layer's 128x64 operand stages with a set pixel (generated batch)"},
            "ms_per_step_with_per_launch_events": ms_step_events,
            "roofline": roof, "kernel_time_shares": shares,
            "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks,
            "gpu_launches": launches_per_step * K, "gpu_launches_per_step": launches_per_step,
            "final_losses": final_losses,
        }
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    eng.close()
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
