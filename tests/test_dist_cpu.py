"""world_size-2 `gloo` test of the multi-GPU exchange protocol, on CPU.

The CUDA path all-reduces 18 doubles of moments (n, sum a, sum b, sum a b^T, sum |a|^2, sum |b|^2 with
a = p - c0, b = y - c0) and 1 double of residual per iteration.  Here each rank builds those sums for its
contiguous scene shard with numpy, all-reduces them over gloo, and derives s, R, t exactly as K3 does;
the result must equal the oracle's unsharded find_alignment.  Also checks the shard partition.
"""
import os
import sys

import numpy as np
import pytest

from conftest import PKG_DIR, ROOT


def _moments(p, y, c0):
    a, b = p - c0, y - c0
    return np.concatenate([[len(p)], a.sum(0), b.sum(0), (a[:, :, None] * b[:, None, :]).sum(0).ravel(),
                           [(a * a).sum()], [(b * b).sum()]])


def _solve_from_moments(mom, c0):
    """Mirror of horn_solve (csrc/align_kernels.cu) using the oracle's Horn routines."""
    from oracle import oracle as O
    n = mom[0]
    mua, mub = mom[1:4] / n, mom[4:7] / n
    S = mom[7:16].reshape(3, 3) - n * np.outer(mua, mub)
    _, R, _, _ = O.horn(S)
    d_caps = mom[17] - n * (mub @ mub)
    sp = mom[16] - n * (mua @ mua)
    s = np.sqrt(d_caps / sp)
    t = (mub + c0) - s * R @ (mua + c0)
    return s, R, t


def _worker(rank, world, port, q):
    sys.path[:0] = [ROOT, PKG_DIR, os.path.join(ROOT, "tests")]
    import torch
    import torch.distributed as dist
    import icp_b200
    from oracle import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(0)                      # same data on every rank (model replicated)
    m = rng.uniform(-1, 1, (3001, 3))
    p = m[rng.permutation(3001)] * 1.01 + 0.003 + 1e-4 * rng.standard_normal((3001, 3))
    idx, Y = O.closest_matrix(p, m)
    c0 = 0.5 * (m.min(0) + m.max(0))
    lo, hi = icp_b200.shard_range(len(p), rank, world)
    mom = torch.from_numpy(_moments(p[lo:hi], Y[lo:hi], c0))
    assert mom.numel() == icp_b200.MOMENT_DOUBLES
    dist.all_reduce(mom)                                 # the 18-double exchange
    s, R, t = _solve_from_moments(mom.numpy(), c0)
    q_local = p[lo:hi] @ (s * R).T + t
    res = torch.tensor([((Y[lo:hi] - q_local) ** 2).sum()], dtype=torch.float64)
    dist.all_reduce(res)                                 # the 1-double exchange
    ro = O.find_alignment(p, Y)
    ok = (abs(s - ro["s"]) < 1e-12 and np.abs(R - ro["R"]).max() < 1e-12 and np.abs(t - ro["t"]).max() < 1e-12
          and abs(res.item() - ro["err"]) <= 1e-9 * ro["err"] and int(mom[0].item()) == len(p))
    q.put((rank, bool(ok), (lo, hi)))
    dist.destroy_process_group()


def test_two_rank_exchange_reproduces_unsharded_alignment():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    out = sorted(q.get(timeout=120) for _ in procs)
    for pr in procs:
        pr.join(60)
    assert all(ok for _, ok, _ in out)
    assert out[0][2] == (0, 1501) and out[1][2] == (1501, 3001)


def test_shard_range_partitions():
    import icp_b200
    for n in (1, 7, 48485, 1 << 22):
        for world in (1, 2, 4, 8):
            edges = [icp_b200.shard_range(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1


def test_peer_mailbox_protocol_model():
    """Host model of peer_allreduce (csrc/align_kernels.cu): every rank posts its values into slot [q & 1][rank]
    of EVERY rank's mailbox, publishes q in the slot's flag, waits for the flags of all ranks in its own mailbox,
    then adds the slots in rank order.  Threads with random stalls stand in for GPUs; the model checks what the
    kernel relies on: sums are right and bit-identical on all ranks over hundreds of back-to-back exchanges with
    only TWO buffers per slot (a rank can be at most one exchange ahead of the slowest peer) and flags that
    are never reset (monotonic sequence numbers)."""
    import random
    import threading
    import time

    world, rounds, n = 4, 300, 18
    data = np.zeros((world, 2, world, n))          # data[owner][parity][source]
    flag = np.zeros((world, 2, world), dtype=np.int64)
    results = [[None] * rounds for _ in range(world)]
    rng = np.random.default_rng(0)
    contrib = rng.standard_normal((rounds, world, n))
    errors = []

    def rank_fn(r):
        rnd = random.Random(r)
        for i in range(rounds):
            q = i + 1
            par = q & 1
            if rnd.random() < 0.2:
                time.sleep(rnd.random() * 2e-3)     # a slow rank
            for dst in range(world):                # post, then publish
                data[dst, par, r, :] = contrib[i, r]
                flag[dst, par, r] = q
            t0 = time.time()
            while not all(flag[r, par, src] == q for src in range(world)):
                if time.time() - t0 > 20:
                    errors.append((r, q, "timeout"))
                    return
                time.sleep(0)
            acc = np.zeros(n)
            for src in range(world):                # rank order: identical everywhere
                acc = acc + data[r, par, src]
            results[r][i] = acc

    th = [threading.Thread(target=rank_fn, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errors, errors
    for i in range(rounds):
        want = np.zeros(n)
        for src in range(world):
            want = want + contrib[i, src]
        for r in range(world):
            assert np.array_equal(results[r][i], want), (r, i)
