"""CPU model of the filter's work decomposition (csrc/nn_kernels.cu "stream-K", icp_api.cu choose_splits()).

The np_pad/2048 scene tiles x nm_pad/128 model chunks form one linear sequence of T chunk-units; CTA b owns
[b*T/n, (b+1)*T/n), writes one partial result per (CTA, scene tile) segment into slot = b - first CTA of that tile, and
nn_refine_kernel recomputes the slot count of a tile from the same integers.  This restates that integer arithmetic
and asserts, over small exhaustive sweeps and the sizes of BASELINE.json's configs, what the kernels rely on:

  * the ranges partition [0, T): every (scene tile, model chunk) pair is evaluated exactly once;
  * the slots a tile receives are exactly 0 .. n_slots-1, each written once, with n_slots as nn_refine derives it
    and n_slots <= n_splits (the scratch rows choose_splits() sizes);
  * a segment only ever stages model tiles inside the padded model;
  * the 32-bit loop state of the kernel cannot overflow at 2^22 x 2^22 points.
"""
import numpy as np
import pytest

NN_TILE, NN_G, NN_TM = 2048, 128, 512      # round 2 defaults: 16 scene points per thread, 128-point chunks
CPM = NN_TM // NN_G
SMS = 148
MIN_CTAS = 2          # resident filter CTAs per SM (NN_MIN_CTAS)


def pad(n, q):
    return (n + q - 1) // q * q


def choose_splits(np_, nm, forced=0, sm_count=SMS):
    """icp_api.cu choose_splits(): (n_cta, n_splits)."""
    n_tiles = pad(np_, NN_TILE) // NN_TILE
    cpt = pad(nm, NN_TM) // NN_G
    total = n_tiles * cpt
    wave = sm_count * MIN_CTAS
    per_cta = 65536 // NN_G
    waves = max(1, (total + per_cta * wave - 1) // (per_cta * wave))
    n_cta = wave * waves
    cap = max(total // CPM, min(total, sm_count))      # one staged tile per CTA, unless that leaves SMs without work
    n_cta = max(min(n_cta, cap), 1)
    if forced > 0:                                     # any grid up to one CTA per unit must work (the kernels are generic in it)
        n_cta = min(forced, total)
    return n_cta, (n_cta + n_tiles - 1) // n_tiles + 1


def cta_of_unit(x, T, n):
    return ((x + 1) * n - 1) // T


def segments(b, n_cta, n_tiles, cpt):
    """The (tile, c_lo, c_hi, slot) sequence CTA b walks in nn_filter_kernel."""
    T = n_tiles * cpt
    begin, end = b * T // n_cta, (b + 1) * T // n_cta
    tile, c_lo, remaining = begin // cpt, begin % cpt, end - begin
    assert remaining < 2 ** 31 and tile < 2 ** 31
    while remaining > 0:
        c_hi = min(cpt, c_lo + remaining)
        slot = b - cta_of_unit(tile * cpt, T, n_cta)
        yield tile, c_lo, c_hi, slot
        remaining -= c_hi - c_lo
        c_lo = 0
        tile += 1


def slot_chunks(tile, slot, T, n_cta, cpt):
    """nn_kernels.cu slot_chunks(): the chunk range a slot of a scene tile covers (what pruned_rescan scans)."""
    x0 = tile * cpt
    b = cta_of_unit(x0, T, n_cta) + slot
    begin, end = b * T // n_cta, (b + 1) * T // n_cta
    return max(begin, x0) - x0, min(end, x0 + cpt) - x0


def check(np_, nm, forced=0):
    n_cta, n_splits = choose_splits(np_, nm, forced)
    n_tiles = pad(np_, NN_TILE) // NN_TILE
    cpt = pad(nm, NN_TM) // NN_G
    T = n_tiles * cpt
    covered = np.zeros((n_tiles, cpt), np.int32)
    slots = [[] for _ in range(n_tiles)]
    for b in range(n_cta):
        for tile, c_lo, c_hi, slot in segments(b, n_cta, n_tiles, cpt):
            assert 0 <= c_lo < c_hi <= cpt and tile < n_tiles
            covered[tile, c_lo:c_hi] += 1
            slots[tile].append(slot)
            mt0, mt1 = c_lo // CPM, (c_hi + CPM - 1) // CPM
            assert mt1 * NN_TM <= pad(nm, NN_TM)                  # bulk copies stay inside mhat
            assert b == cta_of_unit(tile * cpt + c_lo, T, n_cta)  # cta_of_unit() inverts the range formula
            assert slot_chunks(tile, slot, T, n_cta, cpt) == (c_lo, c_hi)   # the rescan of undecided points prunes by slot
    assert (covered == 1).all()
    for tile in range(n_tiles):
        x0 = tile * cpt
        n_slots = cta_of_unit(x0 + cpt - 1, T, n_cta) - cta_of_unit(x0, T, n_cta) + 1    # nn_refine_kernel
        assert sorted(slots[tile]) == list(range(n_slots)), (np_, nm, forced, tile)
        assert n_slots <= n_splits
    return n_cta, n_splits


def test_small_sweep_exhaustive():
    for np_ in (1, 1023, 1024, 1025, 2903, 5000, 9000):
        for nm in (1, 63, 64, 65, 511, 512, 513, 2903, 4097, 20000):
            check(np_, nm)
            for forced in (1, 2, 3, 5, 7, 16, 37, 148, 592, 593, 5000):
                check(np_, nm, forced)


@pytest.mark.parametrize("np_,nm", [(2903, 2903), (48485, 48485), (40097, 40256), (40256, 40097), (65536, 65536),
                                    (262144, 262144), (6061, 48485), (131072, 1048576)])
def test_config_sizes(np_, nm):
    n_cta, n_splits = check(np_, nm)
    total = (pad(np_, NN_TILE) // NN_TILE) * (pad(nm, NN_TM) // NN_G)
    assert n_cta % (SMS * MIN_CTAS) == 0 or n_cta == max(total // CPM, min(total, SMS))
    # whole waves, equal work: ranges differ by at most one chunk
    T = (pad(np_, NN_TILE) // NN_TILE) * (pad(nm, NN_TM) // NN_G)
    sizes = {(b + 1) * T // n_cta - b * T // n_cta for b in range(n_cta)}
    assert max(sizes) - min(sizes) <= 1


def test_largest_configuration_fits_the_kernel_state():
    # 2^22 x 2^22 (BASELINE config 4): checked arithmetically, the full walk is 9e7 units
    np_ = nm = 1 << 22
    n_cta, n_splits = choose_splits(np_, nm)
    n_tiles, cpt = pad(np_, NN_TILE) // NN_TILE, nm // NN_G
    T = n_tiles * cpt
    assert (n_tiles, cpt) == (2048, 32768) and n_cta % (SMS * MIN_CTAS) == 0
    assert T // n_cta + 1 < 2 ** 31 and n_cta < 2 ** 31
    assert 450 <= T / n_cta <= 550                                # ranges of ~64 K model points x one scene tile
    assert n_splits * pad(np_, NN_TILE) * 4 < 2 ** 40             # scratch rows: a few hundred MB at most
    rng = np.random.default_rng(0)
    for b in list(rng.integers(0, n_cta, 200)) + [0, n_cta - 1]:
        segs = list(segments(int(b), n_cta, n_tiles, cpt))
        assert sum(c_hi - c_lo for _, c_lo, c_hi, _ in segs) == (b + 1) * T // n_cta - b * T // n_cta
        assert all(0 <= s < n_splits for *_, s in segs)


@pytest.mark.parametrize("r,g,ctas", [(8, 64, 4), (8, 128, 4), (12, 64, 3), (12, 128, 3), (16, 256, 2)])
def test_variant_geometries(monkeypatch, r, g, ctas):
    """The build-time variants of the filter (icp_common.cuh ICP_NN_R / _G / _MIN_CTAS, `make variants`) run the same
    integer arithmetic with other constants: 128-point chunks, 1 536- and 2 048-point scene tiles, 3 or 2 CTAs per SM."""
    import sys
    mod = sys.modules[__name__]
    monkeypatch.setattr(mod, "NN_TILE", 128 * r)
    monkeypatch.setattr(mod, "NN_G", g)
    monkeypatch.setattr(mod, "CPM", NN_TM // g)
    monkeypatch.setattr(mod, "MIN_CTAS", ctas)
    for np_, nm in [(1, 1), (1535, 513), (2903, 2903), (9000, 20000), (48485, 48485), (40097, 40256), (65536, 65536)]:
        n_cta, _ = check(np_, nm)
        assert n_cta <= SMS * ctas or n_cta % (SMS * ctas) == 0
    for forced in (1, 3, 37, 444, 5000):
        check(5000, 4097, forced)


def test_small_clouds_use_every_sm_they_can():
    """A CTA normally owns at least one staged model tile, but never so few CTAs that SMs stay idle while chunk-units
    remain (round 2: cow's filter 35 -> 25 us).  cow = 2 scene tiles x 24 chunks = 48 units -> 48 CTAs of one unit."""
    n_cta, n_splits = check(2903, 2903)
    assert n_cta == 48 and n_splits == 25
    n_cta, _ = check(8192, 8192)                 # 4 tiles x 64 chunks = 256 units: one CTA per SM, not 64
    assert n_cta == SMS
    n_cta, _ = check(48485, 48485)               # horse: one full wave, ~31 chunks per CTA
    assert n_cta == SMS * MIN_CTAS
    for np_ in (1, 1023, 1025, 2903, 5000, 9000):
        for nm in (1, 64, 65, 513, 2903, 4097):
            check(np_, nm)
