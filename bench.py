#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native ICP hot path (contract: see DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A *step* is one full find_corresponding loop (reference src/cpu.cc:55-79) over one cloud pair.
Default workload = BASELINE.json configs[1]: horse_ref vs horse_tr1, max_iter 30 (48 485 x 48 485 points,
runs all 30 iterations).  metric = nearest-neighbour distance evaluations per second (np*nm per
iteration); ICP iterations/s is reported next to it.

  value     device-timed (CUDA events on the library's stream), clouds already resident in HBM
  e2e       same loop through the C-ABI with HOST buffers: set_model + set_scene (H2D, Morton sort),
            run, get_scene (D2H) inside the timed region, wall clock
  roofline  the FP32 filter kernel (nn_filter_kernel): 8 algorithmic flops per evaluation against the
            FP32 FMA peak  SMs x 128 lanes x 2 x f_max
  cpu_baseline / --impl reference : the CPU oracle (restatement of the reference's cpu.cc; the reference
            itself needs Eigen, which is not installed) timed on the host cores on a bounded sample.

N > 1 (torchrun): the scene is N horse-sized shards (rank r holds one), the model is replicated, the
18 moment sums and the residual are all-reduced through NCCL every iteration -> weak scaling.
"""
import argparse
import json
import lzma
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "iterative-closest-point_b200"))
sys.path.insert(0, ROOT)

WORKLOADS = {   # name -> (model cloud, scene cloud, max_iter, check_sizes)
    "cow": ("cow_ref", "cow_tr1", 30, True),
    "horse": ("horse_ref", "horse_tr1", 30, True),
    "horse_tr2": ("horse_ref", "horse_tr2", 30, True),
    "bunny": ("bun000", "bun045", 30, False),
}
SYNTHETIC = {"syn64k": 1 << 16, "syn256k": 1 << 18, "syn1m": 1 << 20, "syn2m": 1 << 21, "syn4m": 1 << 22}


def load_cloud(name):
    """Parses tests/golden/clouds/<name>.txt.xz like src/load.cc (header dropped, first 3 columns)."""
    with lzma.open(os.path.join(ROOT, "tests", "golden", "clouds", name + ".txt.xz"), "rt") as f:
        lines = f.read().splitlines()[1:]
    return np.array([[float(x) for x in l.split(",")[:3]] for l in lines], dtype=np.float64)


def synthetic(n, seed, angle_deg=5.0):
    """model ~ U[-1,1]^3, scene = R m + t + 1e-3 noise, pseudo-randomly permuted."""
    rng = np.random.default_rng(seed)
    m = rng.uniform(-1.0, 1.0, (n, 3))
    ax = np.array([1.0, 2.0, 3.0]) / np.sqrt(14.0)
    a = np.deg2rad(angle_deg)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    R = np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * (K @ K)
    p = m @ R.T + np.array([0.05, -0.03, 0.02]) + 1e-3 * rng.standard_normal((n, 3))
    return m, np.ascontiguousarray(p[rng.permutation(n)])


def get_workload(name, max_iter=None):
    m, p, it, chk, desc = _get_workload(name)
    if max_iter:
        it = max_iter
        desc = desc.split(", max_iter")[0].split(", 10 iterations")[0] + f", max_iter {it}"
    return m, p, it, chk, desc


def _get_workload(name):
    if name in WORKLOADS:
        mn, pn, it, chk = WORKLOADS[name]
        return load_cloud(mn), load_cloud(pn), it, chk, f"{mn} vs {pn}, max_iter {it}"
    if name.startswith("batch"):     # reference arm / CPU baseline of the batched workload: its first pair
        n = {"batch64k": 1 << 16, "batch16k": 1 << 14}[name]
        m, p = synthetic(n, 1000, angle_deg=2.0)
        return m, p, 30, True, f"first pair of the batch: synthetic {n} points, rigid 2deg + 1e-3 noise, max_iter 30"
    n = SYNTHETIC[name]
    m, p = synthetic(n, 1000 + int(np.log2(n)))
    return m, p, 10, True, f"synthetic 2^{int(np.log2(n))} points, rigid 5deg + 1e-3 noise, 10 iterations"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows[-3:]]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons, "samples": len(rows),
                "power_w_max": max(float(r[2]) for r in rows)}


def fp32_peak_tflops():
    mhz, src = 1965.0, "nominal 148 SMs x 128 lanes x 2 x 1965 MHz (no MEASURED_PEAKS.json)"
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        mhz = float(pk["sm_max_mhz"])
        src = "148 SMs x 128 FP32 lanes x 2 x sm_max_mhz of MEASURED_PEAKS.json (the file holds no FP32 figure)"
    except Exception:
        pass
    return 148 * 128 * 2 * mhz * 1e6 / 1e12, src


def host_threads():
    """All host cores this process may use (torchrun exports OMP_NUM_THREADS=1, which is not what we want)."""
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_port_throughput(m, p, iters, threads=0):
    """Times the oracle's loop (restated src/cpu.cc:55-79, x*x + sqrt distance form) -> evals/s."""
    from oracle import oracle as O
    t0 = time.perf_counter()
    r = O.icp_run(m, p, iters, nn_mode=O.NN_MULSQRT, allow_unequal=True, nthreads=threads)
    dt = time.perf_counter() - t0
    return r["iters"] * len(m) * len(p) / dt, r["iters"], dt, (threads or O.num_threads())


def ref_source_throughput(m, p, max_pairs=2.0e7):
    """Extra context for cpu_baseline: the reference's OWN src/cpu.cc (oracle/_ref, compiled against the Eigen stand-in;
    single-threaded like the reference) on a small sample of CPU::closest_matrix.  Not the timed baseline — a
    stand-in Eigen says little about real Eigen's speed — but it is the reference's source text running."""
    so = os.path.join(ROOT, "oracle", "_ref", "libref_cpu.so")
    if not os.path.exists(so):
        return None
    import ctypes as C
    try:
        L = C.CDLL(so)
        dp = C.POINTER(C.c_double)
        L.ref_closest_matrix.argtypes = [C.c_long, C.c_long, dp, dp, dp]
        n_scene = int(max(1, min(len(p), max_pairs // len(m))))
        ps = np.ascontiguousarray(p[:n_scene])
        Y = np.empty_like(ps)
        t0 = time.perf_counter()
        L.ref_closest_matrix(len(ps), len(m), ps.ctypes.data_as(dp), m.ctypes.data_as(dp), Y.ctypes.data_as(dp))
        dt = time.perf_counter() - t0
        return {"value": len(ps) * len(m) / dt, "unit": "evals/s", "cores": 1, "kind": "reference source (src/cpu.cc) + Eigen stand-in",
                "sample": f"CPU::closest_matrix, {len(ps)} scene x {len(m)} model points"}
    except Exception as e:      # never let the extra line break the bench
        return {"error": str(e)}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port; Eigen is absent so cpu.cc cannot be
    compiled) on the host cores, all threads, on a bounded sample of the same workload."""
    if rank != 0:
        return
    m, p, max_iter, _, desc = get_workload(args.workload, args.max_iter)
    sample_iters = 2 if len(m) * len(p) > 1e9 else max_iter
    if len(m) * len(p) > 1e11:      # synthetic >= 2^20: sample 4096 scene points against the full model
        p = p[:4096]
        sample_iters = 1
    vals = []
    for i in range(args.warmup + args.steps):
        v, it, dt, cores = cpu_port_throughput(m, p, sample_iters, threads=host_threads())
        if i >= args.warmup:
            vals.append((v, it, dt))
    value = float(np.mean([v for v, _, _ in vals]))
    ms = float(np.mean([dt for _, _, dt in vals])) * 1e3
    sample = f"{vals[-1][1]} ICP iteration(s) of {desc} ({len(p)} scene x {len(m)} model points) per step"
    line = {"impl": "reference", "metric": "nn_distance_evals_per_sec", "value": value, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "reference clouds",
            "icp_iterations_per_sec": float(np.mean([it / dt for _, it, dt in vals])),
            "config": {"workload": args.workload, "description": desc},
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": cores, "kind": "port", "sample": sample,
                             "reference_source": ref_source_throughput(m, p)},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


NN_DESC = {"brute": "exact brute force (every pair evaluated)",
           "tree": "exact bounding-box hierarchy (ICP_NN_TREE): identical indices; evals/s counts the np*nm pairs a brute-force "
                   "search would evaluate, not executed work - compare icp_iterations_per_sec"}


def run_ours(args, rank, world, local_rank):
    import torch
    import icp_b200
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if args.workload.startswith('batch'):
        return run_batch(args, rank, world, local_rank, torch, dist)
    m, p, max_iter, check, desc = get_workload(args.workload, args.max_iter)
    if world > 1:
        check = False        # global scene = `world` shards of np points; the model stays nm points

    if world > 1 and args.scaling == "strong":      # one global scene, contiguous block per rank
        lo, hi = icp_b200.shard_range(len(p), rank, world)
        p_global_n = len(p)
        p = np.ascontiguousarray(p[lo:hi])
    else:
        p_global_n = len(p) * world
    ctx = icp_b200.Context(local_rank)
    comm_kind = "none"
    if world > 1:
        uid = [icp_b200.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], rank, world)
        comm_kind = "nccl all-reduce of 18+1 doubles per iteration"
        if os.environ.get("ICP_COMM", "peer") != "nccl":
            handles = [None] * world
            dist.all_gather_object(handles, ctx.comm_peer_handle())
            ctx.comm_peer_attach(handles, rank, world)
            dist.barrier()
            comm_kind = "in-kernel peer-mailbox exchange of 18+1 doubles per iteration over NVLink (no collective launch)"
    ctx.set_model(m)
    ctx.set_scene(p)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        flush.fill_(1)                       # flush L2 between timed iterations
        torch.cuda.synchronize()
        ctx.reset_scene()
        it, hist = ctx.run(max_iter, 1e-5, check)
        st = ctx.stats()
        return it, hist, st

    for _ in range(max(args.warmup, 3)):
        it, hist, st = step()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    t_begin = time.perf_counter()
    dev_ms = filt_ms = 0.0
    launches = iters_total = 0
    for _ in range(args.steps):
        it, hist, st = step()
        dev_ms += st["ms_total"]
        filt_ms += st["ms_filter"]
        launches += st["kernels_launched"] + 1     # + scene restore
        iters_total += it
    barrier()
    t_end = time.perf_counter()
    clocks = sampler.stop(t_begin, t_end) if sampler else None

    # ---- end to end through the C-ABI with host buffers --------------------------------------------
    # Host buffers in page-locked memory, as the contract asks (inputs copied from pinned host memory, the result read
    # back into a pinned buffer); if pinning fails the pageable arrays are used and the line says so.
    ctx2 = ctx
    host_kind = "pinned"
    try:
        keep = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (m, p)]
        keep.append(torch.empty((len(p), 3), dtype=torch.float64).pin_memory())
        m_h, p_h, out_h = (t.numpy() for t in keep)
    except Exception as e:                     # noqa: BLE001 - any failure: fall back, never lose the bench line
        sys.stderr.write(f"[bench] pinning host buffers failed ({e}); e2e uses pageable memory\n")
        m_h, p_h, out_h, host_kind = m, p, None, "pageable"
    e2e_t = []
    e2e_reps = min(args.steps, 10)
    for i in range(2 + e2e_reps):
        barrier()
        t0 = time.perf_counter()
        ctx2.set_model(m_h)
        ctx2.set_scene(p_h)
        it2, hist2 = ctx2.run(max_iter, 1e-5, check)
        out = ctx2.get_scene(out_h)
        torch.cuda.synchronize()
        e2e_t.append(time.perf_counter() - t0)
    e2e_s = float(np.mean(e2e_t[2:]))

    tim = torch.tensor([dev_ms, e2e_s], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(tim, op=dist.ReduceOp.MAX)
    dev_ms, e2e_s = tim.tolist()
    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return

    pairs_per_iter = len(m) * p_global_n
    evals = pairs_per_iter * (iters_total / args.steps)
    ms_per_step = dev_ms / args.steps
    value = evals / (ms_per_step * 1e-3)
    peak, peak_src = fp32_peak_tflops()
    filt_ms_per_launch = filt_ms / max(iters_total, 1)
    achieved = 8.0 * len(m) * len(p) / (filt_ms_per_launch * 1e-3) / 1e12
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "nn_filter_traffic.json")))
    except Exception:
        pass
    cpu = None
    if world == 1 and not args.skip_cpu_baseline:      # the CPU baseline is reported at N = 1 only
        cpu_iters = 8 if len(m) * len(p) < 5e9 else 1
        pc = p if len(m) * len(p) < 1e11 else p[:4096]
        cv, cit, cdt, cores = cpu_port_throughput(m, pc, cpu_iters, threads=host_threads())
        cpu = {"value": cv, "unit": "evals/s", "cores": cores, "kind": "port",
               "sample": f"{cit} ICP iteration(s), {len(pc)} scene x {len(m)} model points, oracle x*x+sqrt form",
               "reference_source": ref_source_throughput(m, p)}
    line = {
        "metric": "nn_distance_evals_per_sec", "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": args.scaling if world > 1 else "weak",
        "vs_baseline": None, "dtype": "f32 filter + f64 re-rank/alignment",
        "data": "reference clouds (tests/golden/clouds)" if args.workload in WORKLOADS else "synthetic",
        "icp_iterations_per_sec": (iters_total / args.steps) / (ms_per_step * 1e-3),
        "iterations_per_step": iters_total / args.steps, "final_error_value": float(hist[-1]),
        "config": {"workload": args.workload, "description": desc, "np_per_gpu": len(p), "nm": len(m),
                   "sharding": "1 GPU" if world == 1 else f"scene = {world} shards of ~{len(p)} points "
                   f"({p_global_n} in total), model replicated, {comm_kind}",
                   "l2": "flushed between steps (256 MB write)", "model_splits": st["n_splits"], "nn": NN_DESC[args.nn]},
        "roofline": None if args.nn == "tree" else {"bound": "fp32", "kernel": "nn_filter_kernel", "achieved": achieved, "peak": peak,
                     "unit": "TFLOP/s", "frac": achieved / peak, "peak_source": peak_src,
                     "algorithmic_flops_per_eval": 8, "evals_per_launch": len(m) * len(p),
                     "avg_launch_ms": filt_ms_per_launch,
                     "traffic": traffic["dram_bytes_per_launch"] if traffic else None},
        "cpu_baseline": cpu,
        "e2e": {"value": evals / e2e_s, "unit": "evals/s", "ms_per_step": e2e_s * 1e3,
                "h2d_bytes_per_step": (len(m) + len(p)) * 24, "d2h_bytes_per_step": len(p) * 24 + it2 * 8 + 64,
                "host_buffers": host_kind},
        "gpu_launches": launches, "clocks": clocks,
        "rerank": {"full_fp64_scans_per_pass": st["rerank_full"] / max(st["nn_calls"], 1), "scene_points": len(p)},
        "wall_ms_per_step_incl_flush": (t_end - t_begin) * 1e3 / args.steps,
    }
    print(json.dumps(line), flush=True)
    if dist:
        dist.destroy_process_group()


def run_batch(args, rank, world, local_rank, torch, dist):
    """BASELINE config 5: many independent synthetic pairs, dealt round-robin to the GPUs (pair-per-GPU, nothing is
    exchanged); on each GPU `lanes` contexts run concurrently (icp_batch_run).  A step registers every pair once,
    through the C-ABI with HOST buffers (the batch entry point has no device-resident form), so value == e2e."""
    import icp_b200
    n = {"batch64k": 1 << 16, "batch16k": 1 << 14}[args.workload]
    max_iter = args.max_iter or 30
    mine = list(range(rank, args.pairs, world))
    models, scenes = [], []
    for k in mine:
        m, p = synthetic(n, 1000 + k, angle_deg=2.0 + 6.0 * ((k * 37) % 101) / 100.0)
        models.append(m)
        scenes.append(p)

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        return icp_b200.batch_run(models, scenes, max_iter=max_iter, device=local_rank, lanes=args.lanes, want_scene=True)

    for _ in range(max(1, min(args.warmup, 3))):
        res = step()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = step()
    torch.cuda.synchronize()
    dt_local = time.perf_counter() - t0
    barrier()
    t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1) if sampler else None
    evals_local = float(np.sum(res["iters"].astype(np.float64))) * n * n
    tim = torch.tensor([dt_local], dtype=torch.float64, device="cuda")
    tot = torch.tensor([evals_local, float(np.sum(res["iters"]))], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(tim, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    if rank == 0:
        sec = tim.item() / args.steps
        evals, iters = tot.tolist()
        line = {"metric": "nn_distance_evals_per_sec", "value": evals / sec, "unit": "evals/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(1, min(args.warmup, 3)), "ms_per_step": sec * 1e3, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32 filter + f64 re-rank/alignment", "data": "synthetic",
                "pairs_per_sec": args.pairs / sec, "icp_iterations_per_sec": iters / sec,
                "config": {"workload": args.workload, "pairs": args.pairs, "points_per_cloud": n, "max_iter": max_iter,
                           "lanes_per_gpu": args.lanes, "nn": NN_DESC[args.nn], "sharding": f"pair-per-GPU, {world} GPU(s), no communication",
                           "l2": "each pair's clouds are uploaded from host memory inside the step (inputs > L2 in aggregate)"},
                "roofline": None, "cpu_baseline": None,
                "e2e": {"value": evals / sec, "unit": "evals/s", "h2d_bytes_per_step": args.pairs * 2 * n * 24,
                        "d2h_bytes_per_step": args.pairs * (n * 24 + 13 * 8 + 12)},
                "gpu_launches": int(iters) * 4 + args.pairs * 12, "clocks": clocks,
                "mean_iterations_per_pair": iters / args.pairs}
        print(json.dumps(line), flush=True)
    if dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="horse", choices=list(WORKLOADS) + list(SYNTHETIC) + ["batch64k", "batch16k"])
    ap.add_argument("--max-iter", type=int, default=None, help="override the workload's max_iter")
    ap.add_argument("--pairs", type=int, default=64, help="batch workloads: cloud pairs in total (dealt round-robin to the ranks)")
    ap.add_argument("--lanes", type=int, default=3, help="batch workloads: concurrent contexts per GPU")
    ap.add_argument("--nn", default="brute", choices=["brute", "tree"],
                    help="brute: the exact brute-force search BASELINE.json's metric is quoted on (default); tree: the exact "
                         "bounding-box-hierarchy search (SURVEY 8f-4) - same results, `value` then counts brute-force-EQUIVALENT pairs")
    ap.add_argument("--skip-cpu-baseline", action="store_true", help="side runs only: leave cpu_baseline null (saves 10-30 s)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = every rank holds a full copy of the scene as its shard (per-GPU work fixed); "
                         "strong = the one scene is cut into N contiguous blocks")
    args = ap.parse_args()
    if args.nn == "tree":
        os.environ["ICP_NN"] = "tree"        # read by icp_create (every context, batch lanes included)
    elif os.environ.get("ICP_NN") == "tree":
        args.nn = "tree"
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
