"""Multi-GPU parity check (one process per GPU, torchrun): the scene of a reference cloud pair is cut into
WORLD contiguous blocks, the model is replicated, and the sharded loop must reproduce the oracle's iteration
count, error values (1e-5 relative, north_star's tolerance) and correspondence indices (bit-exact).
Run for all exchange paths: in-kernel peer mailboxes (split kernels = default, and the cooperative fused kernel)
and the NCCL all-reduce.

  torchrun --nproc-per-node N tools/dist_check.py [case ...]      (ICP_NN=brute|tree restricts the search modes)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "iterative-closest-point_b200"))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    import bench
    import icp_b200
    # usage: dist_check.py [case ...]   (default: cow_tr1 horse_tr2); ICP_NN=brute|tree restricts the search modes (default both)
    cases = [a for a in sys.argv[1:]] or ["cow_tr1", "horse_tr2"]
    nns = [os.environ["ICP_NN"]] if os.environ.get("ICP_NN") in ("brute", "tree") else ["brute", "tree"]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    for case in cases:
        mname, pname = {"cow_tr1": ("cow_ref", "cow_tr1"), "horse_tr2": ("horse_ref", "horse_tr2")}[case]
        m, p = bench.load_cloud(mname), bench.load_cloud(pname)
        g = np.load(os.path.join(ROOT, "tests", "golden", case + ".npz"))
        lo, hi = icp_b200.shard_range(len(p), rank, world)
        for nn in nns:
            os.environ["ICP_NN"] = nn                      # read by icp_create
            for mode in ("peer", "peer-fused", "nccl"):
                os.environ["ICP_ALIGN"] = "fused" if mode == "peer-fused" else "split"
                ctx = icp_b200.Context(local)
                uid = [icp_b200.comm_unique_id() if rank == 0 else None]
                dist.broadcast_object_list(uid, src=0)
                ctx.comm_init(uid[0], rank, world)
                if mode.startswith("peer"):
                    handles = [None] * world
                    dist.all_gather_object(handles, ctx.comm_peer_handle())
                    ctx.comm_peer_attach(handles, rank, world)
                    dist.barrier()
                ctx.set_model(m)
                ctx.set_scene(p[lo:hi])
                ctx.run(30, 1e-5, True)                    # warm-up (first-use set-up, tree build)
                ctx.reset_scene()
                it, hist = ctx.run(30, 1e-5, True)
                idx = ctx.indices()
                new_p = ctx.get_scene()
                s, R, t = ctx.transform()
                good = (it == int(g["iters"]) and np.allclose(hist, g["err"][:it], rtol=1e-5, atol=1e-18)
                        and np.array_equal(idx, g["idx_last"][lo:hi]) and np.abs(new_p - g["new_p"][lo:hi]).max() < 1e-9)
                flag = torch.tensor([int(good)], device="cuda")
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                # every rank must hold the bit-identical transform and error history
                xf = torch.tensor(np.concatenate([[s], R.ravel(), t, hist]), device="cuda")
                lo_xf, hi_xf = xf.clone(), xf.clone()
                dist.all_reduce(lo_xf, op=dist.ReduceOp.MIN)
                dist.all_reduce(hi_xf, op=dist.ReduceOp.MAX)
                same = bool(torch.equal(lo_xf, hi_xf))
                ms = torch.tensor([ctx.stats()["ms_total"]], device="cuda")
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                if rank == 0:
                    print(f"dist_check {case} world={world} nn={nn} mode={mode}: iterations {it} (golden {int(g['iters'])}), "
                          f"parity {'ok' if flag.item() else 'FAILED'}, transform + error history identical on all ranks: {same}, "
                          f"device ms {ms.item():.3f} (slowest rank, second run)", flush=True)
                ok = ok and bool(flag.item()) and same
                ctx.close()
                dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
