#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native ICP hot path (contract: see DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A *step* is one find_corresponding loop (reference src/cpu.cc:55-79) over one cloud pair.
Default workload = north_star's own target, BASELINE.json configs[3]: the synthetic 2^22 x 2^22 cloud pair of
SURVEY §8d (splitmix64 / Box-Muller recipe, icp_b200/synth.py), exact brute force, max_iter 2 per step, **strong
scaling**: the ONE scene is cut into N contiguous blocks (rank r holds block r), the model is replicated — the
reference's own answer to "more points" is the slab loop of src/GPU/compute.cu:156-235.  The round-1 headline
(configs[1], horse_ref x horse_tr1, 30 iterations) is measured in the same run and reported under `secondary.horse`.
metric = nearest-neighbour distance evaluations per second (np*nm per iteration); ICP iterations/s next to it.

  value     device-timed (CUDA events on the library's stream), clouds already resident in HBM
  e2e       same loop through the C-ABI with pinned HOST buffers: set_model + set_scene (H2D, Morton sort),
            run, get_scene (D2H) inside the timed region, wall clock
  roofline  the FP32 filter kernel (nn_filter_kernel): 8 algorithmic flops per evaluation against the
            FP32 FMA peak  SMs x 128 lanes x 2 x f_max
  parity    N > 1 only: BEFORE timing, the communicator that will be timed runs horse_ref x horse_tr2 cut into N blocks
            and must reproduce the golden vectors (iteration count, error history 1e-5, per-shard indices bit-exact,
            bit-identical transform on every rank); a failure exits non-zero.
  cpu_baseline / --impl reference : the CPU oracle (restatement of the reference's cpu.cc; the reference
            itself needs Eigen, which is not installed) timed on the host cores on a bounded sample.
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
SYN_ITERS = {"syn64k": 10, "syn256k": 10, "syn1m": 2, "syn2m": 2, "syn4m": 2}      # iterations per step (fixed)
DEFAULT_WORKLOAD = "syn4m"


def load_cloud(name):
    """Parses tests/golden/clouds/<name>.txt.xz like src/load.cc (header dropped, first 3 columns)."""
    with lzma.open(os.path.join(ROOT, "tests", "golden", "clouds", name + ".txt.xz"), "rt") as f:
        lines = f.read().splitlines()[1:]
    return np.array([[float(x) for x in l.split(",")[:3]] for l in lines], dtype=np.float64)


def get_workload(name, max_iter=None):
    m, p, it, chk, desc = _get_workload(name)
    if max_iter:
        it = max_iter
        desc = desc.split(", max_iter")[0] + f", max_iter {it}"
    return m, p, it, chk, desc


def _get_workload(name):
    from icp_b200 import synth
    if name in WORKLOADS:
        mn, pn, it, chk = WORKLOADS[name]
        return load_cloud(mn), load_cloud(pn), it, chk, f"{mn} vs {pn}, max_iter {it}"
    if name.startswith("batch"):     # reference arm / CPU baseline of the batched workload: its first pair
        n = {"batch64k": 1 << 16, "batch16k": 1 << 14}[name]
        m, p = synth.config5_pair(0, n)
        return m, p, 30, True, f"first pair of the batch: synthetic {n} points (SURVEY 8d config 5), max_iter 30"
    n = SYNTHETIC[name]
    m, p = synth.config4(n)
    return m, p, SYN_ITERS[name], True, (f"synthetic 2^{int(np.log2(n))} x 2^{int(np.log2(n))} points (SURVEY 8d config 4: splitmix64 "
                                         f"model, rigid 5deg + 1e-3 Box-Muller noise, permuted), max_iter {SYN_ITERS[name]}")


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


def filter_traffic(workload, nn):
    """dram__bytes_read + dram__bytes_write of ONE nn_filter_kernel launch from a committed `ncu --set full` capture of this
    workload (profiles/nn_filter_traffic.json: {"workloads": {name: bytes}}); None where no capture covers the workload."""
    if nn != "brute":
        return None
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "nn_filter_traffic.json")))
        return t.get("workloads", {}).get(workload)
    except Exception:
        return None


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


NN_DESC = {"brute": "exact brute force (every pair evaluated)",
           "tree": "exact bounding-box hierarchy (ICP_NN_TREE): identical indices; evals/s counts the np*nm pairs a brute-force "
                   "search would evaluate, not executed work - compare icp_iterations_per_sec"}
L2_NOTE = "flushed between steps (256 MB write)"


def workload_config(args, m, p_global_n, max_iter, desc):
    """`config` of the JSON line: what is measured, the same in both arms (sharding and launch geometry go under `layout`)."""
    return {"workload": args.workload, "description": desc, "np": int(p_global_n), "nm": int(len(m)), "max_iter": int(max_iter),
            "nn": NN_DESC[args.nn], "l2": L2_NOTE}


def cpu_sample(m, p, max_iter):
    """Bounded CPU sample of a workload: (scene sample, iterations, description)."""
    if len(m) * len(p) > 1e11:          # synthetic >= 2^20: 4096 scene points against the full model, one iteration
        return p[:4096], 1, "4096-point scene sample"
    if len(m) * len(p) > 1e9:
        return p, min(2, max_iter), "full clouds"
    return p, max_iter, "full clouds"


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port; Eigen is absent so cpu.cc cannot be
    compiled) on the host cores, all threads, on a bounded sample of the same workload."""
    if rank != 0:
        return
    m, p, max_iter, _, desc = get_workload(args.workload, args.max_iter)
    ps, sample_iters, what = cpu_sample(m, p, max_iter)
    vals = []
    for i in range(args.warmup + args.steps):
        v, it, dt, cores = cpu_port_throughput(m, ps, sample_iters, threads=host_threads())
        if i >= args.warmup:
            vals.append((v, it, dt))
    value = float(np.mean([v for v, _, _ in vals]))
    ms = float(np.mean([dt for _, _, dt in vals])) * 1e3
    sample = f"{vals[-1][1]} ICP iteration(s) per step, {what} ({len(ps)} scene x {len(m)} model points), oracle x*x+sqrt form, OpenMP"
    line = {"impl": "reference", "metric": "nn_distance_evals_per_sec", "value": value, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
            "data": "reference clouds (tests/golden/clouds)" if args.workload in WORKLOADS else "synthetic",
            "icp_iterations_per_sec": float(np.mean([it / dt for _, it, dt in vals])) * (len(ps) / len(p)),
            "config": workload_config(args, m, len(p), max_iter, desc),
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": cores, "kind": "port", "sample": sample,
                             "reference_source": ref_source_throughput(m, p)},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---- the GPU arm ---------------------------------------------------------------------------------------------

class Harness:
    """One context per rank (+ communicator when world > 1) and the timing helpers shared by the main and secondary lines."""

    def __init__(self, args, rank, world, local_rank, torch, dist):
        import icp_b200
        self.icp, self.args, self.rank, self.world, self.local, self.torch, self.dist = icp_b200, args, rank, world, local_rank, torch, dist
        self.ctx = icp_b200.Context(local_rank)
        self.comm_kind = "none"
        if world > 1:
            uid = [icp_b200.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            self.ctx.comm_init(uid[0], rank, world)
            self.comm_kind = "nccl all-reduce of 18+1 doubles per iteration"
            if os.environ.get("ICP_COMM", "peer") != "nccl":
                handles = [None] * world
                dist.all_gather_object(handles, self.ctx.comm_peer_handle())
                self.ctx.comm_peer_attach(handles, rank, world)
                dist.barrier()
                self.comm_kind = "in-kernel peer-mailbox exchange of 18+1 doubles per iteration over NVLink (no collective launch)"
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce_max(self, vals):
        t = self.torch.tensor(vals, dtype=self.torch.float64, device="cuda")
        if self.dist:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()

    def shard(self, p, scaling):
        """-> (this rank's scene block, global scene size)."""
        if self.world > 1 and scaling == "strong":
            lo, hi = self.icp.shard_range(len(p), self.rank, self.world)
            return np.ascontiguousarray(p[lo:hi]), len(p)
        return p, len(p) * self.world

    def parity_gate(self):
        """horse_ref x horse_tr2 (8 iterations, the knife-edge threshold case) cut into `world` blocks, on the
        communicator that is about to be timed, against tests/golden/horse_tr2.npz.  Loop being sharded: src/cpu.cc:55-79."""
        torch, dist, ctx = self.torch, self.dist, self.ctx
        m, p = load_cloud("horse_ref"), load_cloud("horse_tr2")
        g = np.load(os.path.join(ROOT, "tests", "golden", "horse_tr2.npz"))
        lo, hi = self.icp.shard_range(len(p), self.rank, self.world)
        ctx.set_model(m)
        ctx.set_scene(p[lo:hi])
        it, hist = ctx.run(30, 1e-5, True)
        idx, new_p = ctx.indices(), ctx.get_scene()
        s, R, t = ctx.transform()
        checks = [it == int(g["iters"]),
                  len(hist) == int(g["iters"]) and bool(np.allclose(hist, g["err"][:it], rtol=1e-5, atol=1e-18)),
                  bool(np.array_equal(idx, g["idx_last"][lo:hi])),
                  bool(np.abs(new_p - g["new_p"][lo:hi]).max() < 1e-9),
                  bool(np.allclose(R, g["R"][-1], rtol=1e-5, atol=1e-9) and abs(s - g["s"][-1]) <= 1e-5 * abs(g["s"][-1]))]
        flags = torch.tensor([int(c) for c in checks], device="cuda")
        xf = torch.tensor(np.concatenate([[s], R.ravel(), t, hist]), device="cuda")
        lo_xf, hi_xf = xf.clone(), xf.clone()
        if dist:
            dist.all_reduce(flags, op=dist.ReduceOp.MIN)
            dist.all_reduce(lo_xf, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi_xf, op=dist.ReduceOp.MAX)
        same = bool(torch.equal(lo_xf, hi_xf))
        f = [bool(x) for x in flags.tolist()]
        return {"case": "horse_tr2", "world": self.world, "ok": all(f) and same, "iterations": it, "golden_iterations": int(g["iters"]),
                "iteration_count": f[0], "error_history_1e-5": f[1], "indices_bit_exact_per_shard": f[2], "new_p_1e-9": f[3],
                "s_R_1e-5": f[4], "transform_and_history_bit_identical_on_all_ranks": same,
                "exchange": os.environ.get("ICP_COMM", "peer"), "nn": self.args.nn}

    def timed_loop(self, m, p_shard, max_iter, check, steps, warmup, sample_clocks):
        """Device-timed steps with the clouds resident.  -> dict (times are this rank's; reduce with reduce_max)."""
        ctx, torch = self.ctx, self.torch
        ctx.set_model(m)
        ctx.set_scene(p_shard)

        def step():
            self.flush.fill_(1)                  # flush L2 between timed iterations
            torch.cuda.synchronize()
            ctx.reset_scene()
            it, hist = ctx.run(max_iter, 1e-5, check)
            return it, hist, ctx.stats()

        for _ in range(max(warmup, 3)):
            it, hist, st = step()
        self.barrier()
        sampler = ClockSampler(self.local) if (sample_clocks and self.rank == 0) else None
        t_begin = time.perf_counter()
        dev_ms = filt_ms = 0.0
        launches = iters_total = 0
        for _ in range(steps):
            it, hist, st = step()
            dev_ms += st["ms_total"]
            filt_ms += st["ms_filter"]
            launches += st["kernels_launched"] + 1     # + scene restore
            iters_total += it
        self.barrier()
        t_end = time.perf_counter()
        clocks = sampler.stop(t_begin, t_end) if sampler else None
        return {"dev_ms": dev_ms, "filt_ms": filt_ms, "launches": launches, "iters_total": iters_total, "hist": hist, "st": st,
                "clocks": clocks, "wall_s": t_end - t_begin}

    def e2e_loop(self, m, p_shard, max_iter, check, reps):
        """The same loop through the C-ABI with page-locked HOST buffers, wall clock.  -> (seconds per step, iterations, kind)."""
        ctx, torch = self.ctx, self.torch
        host_kind = "pinned"
        try:
            keep = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (m, p_shard)]
            keep.append(torch.empty((len(p_shard), 3), dtype=torch.float64).pin_memory())
            m_h, p_h, out_h = (t.numpy() for t in keep)
        except Exception as e:                     # noqa: BLE001 - any failure: fall back, never lose the bench line
            sys.stderr.write(f"[bench] pinning host buffers failed ({e}); e2e uses pageable memory\n")
            m_h, p_h, out_h, host_kind = m, p_shard, None, "pageable"
        ts, it2 = [], 0
        for i in range(1 + reps):
            self.barrier()
            t0 = time.perf_counter()
            ctx.set_model(m_h)
            ctx.set_scene(p_h)
            it2, _ = ctx.run(max_iter, 1e-5, check)
            ctx.get_scene(out_h)
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        return float(np.mean(ts[1:])), it2, host_kind

    def measure(self, workload, m, p, max_iter, check, steps, warmup, e2e_reps, scaling, sample_clocks):
        """One workload at this world size -> dict of the reduced numbers (valid on every rank)."""
        p_shard, p_global_n = self.shard(p, scaling)
        # the reference's size check (src/cpu.cc:42-53) is on the GLOBAL scene (the library sums np over the communicator);
        # the weak form's N copies of the scene cannot pass it
        check = check and (self.world == 1 or scaling == "strong")
        r = self.timed_loop(m, p_shard, max_iter, check, steps, warmup, sample_clocks)
        e2e_s, it2, host_kind = self.e2e_loop(m, p_shard, max_iter, check, e2e_reps)
        dev_ms, e2e_s, filt_ms = self.reduce_max([r["dev_ms"], e2e_s, r["filt_ms"]])
        pairs_per_iter = len(m) * p_global_n
        iters_per_step = r["iters_total"] / steps
        evals = pairs_per_iter * iters_per_step
        ms_per_step = dev_ms / steps
        peak, peak_src = fp32_peak_tflops()
        filt_ms_per_launch = filt_ms / max(r["iters_total"], 1)
        achieved = 8.0 * len(m) * len(p_shard) / (max(filt_ms_per_launch, 1e-12) * 1e-3) / 1e12
        roofline = None
        if self.args.nn != "tree" and filt_ms > 0:
            roofline = {"bound": "fp32", "kernel": "nn_filter_kernel", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                        "frac": achieved / peak, "peak_source": peak_src, "algorithmic_flops_per_eval": 8,
                        "evals_per_launch": len(m) * len(p_shard), "avg_launch_ms": filt_ms_per_launch,
                        "traffic": filter_traffic(workload, self.args.nn) if self.world == 1 else None}
        return {"value": evals / (ms_per_step * 1e-3), "ms_per_step": ms_per_step, "iters_per_step": iters_per_step,
                "evals": evals, "roofline": roofline, "np_per_gpu": len(p_shard), "p_global_n": p_global_n,
                "e2e": {"value": evals / e2e_s, "unit": "evals/s", "ms_per_step": e2e_s * 1e3,
                        "h2d_bytes_per_step": (len(m) + len(p_shard)) * 24, "d2h_bytes_per_step": len(p_shard) * 24 + it2 * 8 + 64,
                        "host_buffers": host_kind, "steps": e2e_reps},
                "launches": r["launches"], "clocks": r["clocks"], "hist": r["hist"], "st": r["st"],
                "wall_ms_per_step_incl_flush": r["wall_s"] * 1e3 / steps}


def run_ours(args, rank, world, local_rank):
    import torch
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
    H = Harness(args, rank, world, local_rank, torch, dist)

    # ---- parity gate (N > 1): before anything is timed ------------------------------------------------------
    parity = None
    if world > 1 and not args.skip_parity:
        parity = H.parity_gate()
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"parity": parity, "error": "sharded parity check failed; nothing was timed"}), flush=True)
            dist.destroy_process_group()
            sys.exit(1)

    m, p, max_iter, check, desc = get_workload(args.workload, args.max_iter)
    big = len(m) * len(p) > 1e12
    e2e_reps = max(1, min(args.steps, 3 if big else 10))
    R = H.measure(args.workload, m, p, max_iter, check, args.steps, args.warmup, e2e_reps, args.scaling, True)

    # ---- secondary: the round-1 headline (configs[1]) in the same run ---------------------------------------
    secondary = None
    if args.workload == DEFAULT_WORKLOAD and not args.no_secondary:
        hm, hp, hit, hchk, hdesc = get_workload("horse")
        secondary = {}
        forms = [("horse", "strong")] if world == 1 else [("horse_strong", "strong"), ("horse_weak", "weak")]
        for key, sc in forms:
            S = H.measure("horse", hm, hp, hit, hchk, 10, 3, 5, sc, False)
            secondary[key] = {"value": S["value"], "unit": "evals/s", "ms_per_step": S["ms_per_step"], "steps": 10,
                              "iterations_per_step": S["iters_per_step"], "e2e": S["e2e"]["value"],
                              "e2e_ms_per_step": S["e2e"]["ms_per_step"], "roofline_frac": S["roofline"]["frac"] if S["roofline"] else None,
                              "filter_avg_launch_ms": S["roofline"]["avg_launch_ms"] if S["roofline"] else None,
                              "np_per_gpu": S["np_per_gpu"], "scaling": sc if world > 1 else None, "description": hdesc +
                              ("" if world == 1 else (": the one scene cut into N blocks" if sc == "strong" else
                                                      ": every rank holds a full COPY of the scene (kernel-scaling probe, not a sharded scene)"))}
    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return

    cpu = None
    if world == 1 and not args.skip_cpu_baseline:      # the CPU baseline is reported at N = 1 only
        ps, cpu_iters, what = cpu_sample(m, p, min(max_iter, 8))
        cv, cit, cdt, cores = cpu_port_throughput(m, ps, cpu_iters, threads=host_threads())
        cpu = {"value": cv, "unit": "evals/s", "cores": cores, "kind": "port",
               "sample": f"{cit} ICP iteration(s), {what} ({len(ps)} scene x {len(m)} model points), oracle x*x+sqrt form, OpenMP",
               "reference_source": ref_source_throughput(m, p)}
    st = R["st"]
    line = {
        "metric": "nn_distance_evals_per_sec", "value": R["value"], "unit": "evals/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": R["ms_per_step"], "higher_is_better": True,
        "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f32 filter + f64 re-rank/alignment",
        "data": "reference clouds (tests/golden/clouds)" if args.workload in WORKLOADS else "synthetic",
        "icp_iterations_per_sec": R["iters_per_step"] / (R["ms_per_step"] * 1e-3),
        "iterations_per_step": R["iters_per_step"], "final_error_value": float(R["hist"][-1]),
        "config": workload_config(args, m, R["p_global_n"], max_iter, desc),
        "layout": {"np_per_gpu": R["np_per_gpu"],
                   "sharding": "1 GPU" if world == 1 else (
                       f"scene of {R['p_global_n']} points cut into {world} contiguous blocks of ~{R['np_per_gpu']}" if args.scaling == "strong"
                       else f"every rank holds a full copy of the scene as its shard ({R['p_global_n']} points in total)") +
                   ("" if world == 1 else f", model replicated, {H.comm_kind}"),
                   "model_splits": st["n_splits"]},
        "roofline": R["roofline"],
        "cpu_baseline": cpu,
        "e2e": R["e2e"],
        "gpu_launches": R["launches"], "clocks": R["clocks"],
        "rerank": {"full_fp64_scans_per_pass": st["rerank_full"] / max(st["nn_calls"], 1), "scene_points": R["np_per_gpu"]},
        "wall_ms_per_step_incl_flush": R["wall_ms_per_step_incl_flush"],
        "parity": parity, "secondary": secondary,
    }
    print(json.dumps(line), flush=True)
    if dist:
        dist.destroy_process_group()


def run_batch(args, rank, world, local_rank, torch, dist):
    """BASELINE config 5: many independent synthetic pairs, dealt round-robin to the GPUs (pair-per-GPU, nothing is
    exchanged); on each GPU `lanes` contexts run concurrently (icp_batch_run).  A step registers every pair once,
    through the C-ABI with HOST buffers (the batch entry point has no device-resident form), so value == e2e."""
    import icp_b200
    from icp_b200 import synth
    n = {"batch64k": 1 << 16, "batch16k": 1 << 14}[args.workload]
    max_iter = args.max_iter or 30
    mine = list(range(rank, args.pairs, world))
    models, scenes = [], []
    for k in mine:
        m, p = synth.config5_pair(k, n)
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
    launches = 0
    for _ in range(args.steps):
        res = step()
        launches += int(res.get("kernels_launched", 0))
    torch.cuda.synchronize()
    dt_local = time.perf_counter() - t0
    barrier()
    t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1) if sampler else None
    evals_local = float(np.sum(res["iters"].astype(np.float64))) * n * n
    tim = torch.tensor([dt_local], dtype=torch.float64, device="cuda")
    tot = torch.tensor([evals_local, float(np.sum(res["iters"])), float(launches)], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(tim, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    if rank == 0:
        sec = tim.item() / args.steps
        evals, iters, launches_all = tot.tolist()
        cpu = None
        if world == 1 and not args.skip_cpu_baseline:       # the port on the batch's first pair, two iterations
            cv, cit, cdt, cores = cpu_port_throughput(models[0], scenes[0], 2, threads=host_threads())
            cpu = {"value": cv, "unit": "evals/s", "cores": cores, "kind": "port",
                   "sample": f"{cit} ICP iterations of pair 0 ({n} x {n} points), oracle x*x+sqrt form, OpenMP"}
        line = {"metric": "nn_distance_evals_per_sec", "value": evals / sec, "unit": "evals/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(1, min(args.warmup, 3)), "ms_per_step": sec * 1e3, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32 filter + f64 re-rank/alignment", "data": "synthetic",
                "pairs_per_sec": args.pairs / sec, "icp_iterations_per_sec": iters / sec,
                "config": {"workload": args.workload, "pairs": args.pairs, "points_per_cloud": n, "max_iter": max_iter,
                           "lanes_per_gpu": args.lanes, "nn": NN_DESC[args.nn], "sharding": f"pair-per-GPU, {world} GPU(s), no communication",
                           "l2": "each pair's clouds are uploaded from host memory inside the step (inputs > L2 in aggregate)"},
                "roofline": None, "cpu_baseline": cpu,
                "e2e": {"value": evals / sec, "unit": "evals/s", "h2d_bytes_per_step": args.pairs * 2 * n * 24,
                        "d2h_bytes_per_step": args.pairs * (n * 24 + 13 * 8 + 12)},
                "gpu_launches": int(launches_all), "clocks": clocks,
                "mean_iterations_per_pair": iters / args.pairs}
        print(json.dumps(line), flush=True)
    if dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=list(WORKLOADS) + list(SYNTHETIC) + ["batch64k", "batch16k"])
    ap.add_argument("--max-iter", type=int, default=None, help="override the workload's max_iter")
    ap.add_argument("--pairs", type=int, default=64, help="batch workloads: cloud pairs in total (dealt round-robin to the ranks)")
    ap.add_argument("--lanes", type=int, default=3, help="batch workloads: concurrent contexts per GPU")
    ap.add_argument("--nn", default="brute", choices=["brute", "tree"],
                    help="brute: the exact brute-force search BASELINE.json's metric is quoted on (default); tree: the exact "
                         "bounding-box-hierarchy search (SURVEY 8f-4) - same results, `value` then counts brute-force-EQUIVALENT pairs")
    ap.add_argument("--skip-cpu-baseline", action="store_true", help="side runs only: leave cpu_baseline null (saves 10-30 s)")
    ap.add_argument("--skip-parity", action="store_true", help="side runs only: N > 1 without the sharded-parity gate")
    ap.add_argument("--no-secondary", action="store_true", help="side runs only: no secondary.horse block")
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="N > 1: strong (default) = the one scene is cut into N contiguous blocks; weak = every rank holds a full "
                         "copy of the scene as its shard (per-GPU work fixed; a kernel-scaling probe)")
    args = ap.parse_args()
    if args.nn == "tree":
        os.environ["ICP_NN"] = "tree"        # read by icp_create (every context, batch lanes included)
    elif os.environ.get("ICP_NN") == "tree":
        args.nn = "tree"
    else:
        os.environ["ICP_NN"] = "brute"       # the metric counts executed brute-force pairs: no automatic tree on small clouds
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
