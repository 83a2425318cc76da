#!/usr/bin/env python
"""One training step out of an ncu launch list (gpu__time_duration of every launch of a bench.py run):

    python tools/launch_list_step.py profiles/<list>.csv [n] > profiles/<list>_step.txt

prints the kernels of the n-th step (default 6: a step of the timed loop, zero-stage skipping on; the list's last
steps belong to the `dense_inputs` leg) with their durations and shares.  A step ends with the large dual-Adam launch."""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 6
hdr, out = None, []
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        out.append(dict(zip(hdr, r)))
ends = [i for i, d in enumerate(out) if "adam_dual" in d["Kernel Name"] and int(d["Grid Size"].strip("()").split(",")[0]) > 1000]
lo, hi = ends[n - 1] + 1, ends[n] + 1
step = [d for d in out[lo:hi] if not any(k in d["Kernel Name"] for k in ("generate_pairs", "kmask", "klist", "schedule", "pairs_bits", "touch_rows"))]
tot = sum(float(d["Metric Value"]) for d in step) / 1000
print("step %d of %s: %d launches, %.1f us summed (serialised, cold caches: shares, not absolutes)" % (n, sys.argv[1], len(step), tot))
for d in step:
    name = re.sub(r"\(magic::.*|\(uint4.*|\(const float.*", "", d["Kernel Name"]).replace("void ", "").replace("magic::", "")
    name = name.replace("(bool)", "").replace("(int)", "")
    t = float(d["Metric Value"]) / 1000
    print("  %-44s %-14s %8.1f us  %5.1f %%" % (name[:44], d["Grid Size"], t, 100 * t / tot))
