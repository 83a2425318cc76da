"""Per-tag time table of the data-parallel step (rank 0's view), under torchrun:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29512 tools/dp_probe.py"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg, eng, vae = bench.make_engine("8d", 4096, "bf16", rank, world, dev)
    for i in range(8):
        eng.generate_batch(0, i)
        eng.step()
    K = 40
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for i in range(K):
        eng.generate_batch(0, 8 + i)
        eng.step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    eng.profile(True)
    for i in range(3):
        eng.generate_batch(0, 100 + i)
        eng.step()
    eng.profile_report()
    n = 10
    for i in range(n):
        eng.generate_batch(0, 200 + i)
        eng.step()
    rep = eng.profile_report()
    if world > 1:
        dist.barrier()
    if rank == 0:
        print("world %d: %.3f ms/step (%.0f pairs/s over all ranks)" % (world, ms, 4096 * world / ms * 1e3))
        tot = sum(v[0] for v in rep.values())
        for k, v in sorted(rep.items(), key=lambda kv: -kv[1][0]):
            print("%-28s %8.3f ms/step  n=%3d" % (k, v[0] / n, v[1] // n))
        print("sum of tags %.3f ms/step" % (tot / n))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
