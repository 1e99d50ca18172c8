"""Per-kernel event times of one loop (ICP_PROFILE=1) on a bench workload, with the HBM view of the linear passes:
K2+K4 move 52 + 88 = 140 algorithmic bytes per scene point.  usage: prof_tail.py <workload> [iters]"""
import os, sys, json
os.environ["ICP_PROFILE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, icp_b200
name = sys.argv[1] if len(sys.argv) > 1 else "horse"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else None
m, p, max_iter, check, desc = bench.get_workload(name, iters)
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
for mode in ("fused", "split"):
    os.environ["ICP_ALIGN"] = mode
    ctx = icp_b200.Context(0)
    ctx.set_model(m); ctx.set_scene(p)
    for _ in range(2):
        ctx.reset_scene()
        it, hist = ctx.run(max_iter, 1e-5, check)
    st = ctx.stats()
    al = st["ms_align"] / it * 1e3
    gbs = len(p) * 140 / (al * 1e-6) / 1e9
    print(f"{desc} [{mode}]: {it} it; per iteration us: filter {st['ms_filter']/it*1e3:.1f} refine {st['ms_refine']/it*1e3:.1f} "
          f"exact {st['ms_exact']/it*1e3:.1f} align {al:.1f}  -> K2+K4 {gbs:.0f} GB/s algorithmic = {gbs/peak*100:.1f}% of measured HBM copy peak "
          f"({peak:.0f} GB/s); queued/pass {st['rerank_full']/max(st['nn_calls'],1):.0f}", flush=True)
    ctx.close()
