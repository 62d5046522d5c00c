"""Worker: the per-rank before/after example (reference docs/poc_rocmaware.png workflow) on CPU ranks."""
import os
import runpy
import sys

outdir = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = ["poc_per_rank_plots.py", "--cpu", "--size", "48", "--nt", "60", "--outdir", outdir]
runpy.run_path(os.path.join(root, "examples", "poc_per_rank_plots.py"), run_name="__main__")
rank = int(os.environ["RANK"])
for tag in ("before", "after"):
    assert os.path.getsize(os.path.join(outdir, f"poc_rank{rank}_{tag}.png")) > 1000
print(f"rank {rank} ok")
