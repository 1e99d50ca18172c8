"""One warm-up loop + one measured loop of a bench workload — the command profiled under ncu."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import icp_b200

name = sys.argv[1] if len(sys.argv) > 1 else "horse"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else None
m, p, max_iter, check, desc = bench.get_workload(name)
ctx = icp_b200.Context(0)
ctx.set_model(m); ctx.set_scene(p)
for _ in range(2):
    ctx.reset_scene()
    it, hist = ctx.run(iters or max_iter, 1e-5, check)
st = ctx.stats()
print(desc, "iterations", it, "dev ms", st["ms_total"], "filter ms", st["ms_filter"], "splits", st["n_splits"])
