"""CPU model of how the chunk minima of one scene point travel from nn_filter_kernel to nn_refine_kernel.

The filter tracks, per (CTA, scene tile) segment, three values per scene point over the segment's model chunks:
`best` (smallest chunk minimum), `best_chunk` (its chunk) and `runner` (smallest chunk minimum of any OTHER chunk,
equal to `best` when two chunks tie) — csrc/nn_kernels.cu, the block after the q loop.  nn_refine merges the
segments' triples of a tile ("slots") with 8 lanes striding over the slots and a butterfly over the lanes.  Whatever
the partition of the chunks into slots — the stream-K ranges, the build-time variants, ICP_NN_MIN_UNITS = 1 (one
chunk per slot) — the merged triple must be the one a single pass over all chunks gives, because the decision
`runner > best + tau` ("every possible argmin lies in the best chunk") is taken on it.
"""
import numpy as np
import pytest

INF = np.float32(np.inf)
RF = 8


def filter_segment(cm, c_lo, c_hi):
    """nn_filter_kernel's per-chunk update, chunks c_lo..c_hi-1 in order."""
    best, runner, bchunk = INF, INF, 0
    for c in range(c_lo, c_hi):
        tm = cm[c]
        lt = tm < best
        runner = min(runner, max(best, tm))
        bchunk = c if lt else bchunk
        best = min(best, tm)
    return best, runner, bchunk


def refine_merge(slots):
    """nn_refine_kernel: lane `sub` folds slots sub, sub+RF, ...; then the group's butterfly."""
    b = [INF] * RF
    rn = [INF] * RF
    ch = [0] * RF
    for sub in range(RF):
        for s in range(sub, len(slots), RF):
            sb, sr, sc = slots[s]
            if sb < b[sub]:
                rn[sub] = min(min(rn[sub], b[sub]), sr)
                b[sub] = sb
                ch[sub] = sc
            else:
                rn[sub] = min(rn[sub], min(sr, sb))
    best = min(b)
    eq = [i for i in range(RF) if b[i] == best]
    src = eq[0]
    runner = [rn[i] if i == src else b[i] for i in range(RF)]
    if len(eq) > 1 and best < INF:
        runner = [best] * RF
    return best, min(runner), ch[src]


def reference(cm):
    order = np.argsort(cm, kind="stable")
    best = cm[order[0]]
    runner = cm[order[1]] if len(cm) > 1 else INF
    return best, runner, int(order[0])


def random_partition(rng, n_chunks):
    k = int(rng.integers(1, min(n_chunks, 60) + 1))
    cuts = sorted(rng.choice(np.arange(1, n_chunks), size=k - 1, replace=False)) if k > 1 else []
    edges = [0] + [int(c) for c in cuts] + [n_chunks]
    return list(zip(edges[:-1], edges[1:]))


@pytest.mark.parametrize("seed", range(6))
def test_any_partition_gives_the_single_pass_triple(seed):
    rng = np.random.default_rng(seed)
    for _ in range(300):
        n_chunks = int(rng.integers(1, 200))
        kind = rng.integers(0, 4)
        if kind == 0:
            cm = rng.standard_normal(n_chunks).astype(np.float32)
        elif kind == 1:                                   # many exact ties
            cm = rng.integers(0, 4, n_chunks).astype(np.float32)
        elif kind == 2:                                   # sentinel chunks (padding): huge values, never the best
            cm = np.where(rng.random(n_chunks) < 0.3, np.float32(3e36), rng.standard_normal(n_chunks)).astype(np.float32)
        else:                                             # near ties one ulp apart
            base = np.float32(rng.standard_normal())
            cm = np.array([np.nextafter(base, INF) if rng.random() < 0.5 else base for _ in range(n_chunks)], np.float32)
        parts = [random_partition(rng, n_chunks)] if n_chunks > 1 else [[(0, 1)]]
        parts.append([(c, c + 1) for c in range(n_chunks)])          # ICP_NN_MIN_UNITS = 1: one chunk per slot
        parts.append([(0, n_chunks)])                                # a single slot
        want = reference(cm)
        for part in parts:
            slots = [filter_segment(cm, lo, hi) for lo, hi in part]
            best, runner, chunk = refine_merge(slots)
            assert best == want[0] and runner == want[1], (cm, part)
            if runner > best:                                        # unique minimum: the chunk is decided
                assert chunk == want[2]
            else:                                                    # tie: the point is queued, any tied chunk will do
                assert cm[chunk] == best
