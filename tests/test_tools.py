"""tools/launch_list_step.py on the committed ncu launch list: the extract kept under profiles/ is what the tool
prints, its shares add up, and the three largest launches of a step are GEMMs."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_launch_list_step_reproduces_the_committed_extract():
    lst = os.path.join("profiles", "r01q_launches_one_step_head.csv")
    out = subprocess.run([sys.executable, os.path.join("tools", "launch_list_step.py"), lst, "6"], cwd=ROOT,
                         capture_output=True, text=True, check=True).stdout
    assert out == open(os.path.join(ROOT, "profiles", "r01q_launches_one_step_head_step.txt")).read()
    lines = out.splitlines()
    assert lines[0].startswith("step 6 of %s: 28 launches" % lst)
    shares = [float(l.split()[-2]) for l in lines[1:]]
    assert len(shares) == 28 and abs(sum(shares) - 100.0) < 1.0
    # the two packed first-layer GEMMs and the BCE GEMM are the three largest launches of the step
    top = sorted(lines[1:], key=lambda l: -float(l.split()[-2]))[:3]
    assert all("gemm_tcgen05_kernel" in l for l in top)
