"""Static audit of the machine code in libicp_b200.so (runs without a GPU: cuobjdump reads the sm_100a cubin).

DESIGN.md §4 makes claims about the code the kernels execute; this file pins the ones that can be read off the
SASS, so that a change of compiler flags or of the source cannot silently lose them:

  * nn_filter_kernel (K1, the kernel `roofline` refers to): 255 registers without a spill (2 CTAs of 128 threads per
    SM), model tiles arrive by TMA bulk copy completing on an mbarrier (UBLKCP / SYNCS), the inner loop is 4 LDS.128 +
    96 FFMA2 + 32 FMNMX3 per 64 pairs (16 scene points per thread) with nothing spilled inside it, and neither the FP64
    pipe nor the tensor cores are used;
  * moments_kernel / transform_kernel: the peer exchange is plain system-scope stores / loads / fences next to the
    reduction (the evidence B200_PROFILING.md asks for of a compute + exchange kernel);
  * the FP64 decision code keeps the reference's operation order: `dist2_64` must not be contracted into FMAs
    (checked on nn_refine_kernel, whose only legitimate DFMAs are the ones of the software square root).
"""
import os
import re
import shutil
import subprocess

import pytest

from conftest import PKG_DIR

LIB = os.path.join(PKG_DIR, "lib", "libicp_b200.so")
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"

pytestmark = pytest.mark.skipif(not (os.path.exists(LIB) and os.path.exists(CUOBJDUMP)),
                                reason="needs the built library and cuobjdump")

_INSTR = re.compile(r"^\s+/\*([0-9a-f]{4,})\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)(.*?);")


@pytest.fixture(scope="module")
def sass():
    """kernel short name -> list of (address, opcode, operand text)."""
    out = subprocess.run([CUOBJDUMP, "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = {}, None
    for line in out.splitlines():
        if "Function :" in line:
            mangled = line.split("Function :")[1].strip()
            m = re.search(r"\d([a-z_]+_kernel)E", mangled)
            cur = kernels.setdefault(m.group(1) if m else mangled, [])
            continue
        m = _INSTR.match(line)
        if m and cur is not None:
            cur.append((int(m.group(1), 16), m.group(2), m.group(3)))
    return kernels


@pytest.fixture(scope="module")
def resources():
    out = subprocess.run([CUOBJDUMP, "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    res, name = {}, None
    for line in out.splitlines():
        m = re.search(r"Function .*?\d([a-z_]+_kernel)E", line)
        if m:
            name = m.group(1)
        r = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if r and name:
            res[name] = dict(zip(("reg", "stack", "shared", "local"), map(int, r.groups())))
            name = None
    return res


def count(instrs, prefix):
    return sum(1 for _, op, _ in instrs if op.startswith(prefix))


def test_library_is_sm100a_only():
    out = subprocess.run([CUOBJDUMP, "-lelf", LIB], capture_output=True, text=True, check=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_filter_resources(resources):
    r = resources["nn_filter_kernel"]
    assert r["reg"] <= 255            # 2 CTAs x 128 threads x 255 registers fit the 64 K registers of an SM
    assert r["local"] == 0
    assert r["stack"] == 0            # nothing spilled anywhere in the kernel
    # the streaming passes and the refine kernel run 4 CTAs of 256 threads per SM (64 registers), K4 three (<= 85)
    for k, lim in (("moments_kernel", 64), ("nn_refine_kernel", 64), ("transform_kernel", 85)):
        assert resources[k]["reg"] <= lim and resources[k]["local"] == 0, (k, resources[k])


def _inner_loop(instrs):
    """The q loop of nn_filter_kernel: the innermost backward branch whose body holds the FFMA2s."""
    addr = {a: i for i, (a, _, _) in enumerate(instrs)}
    best = None
    for i, (a, op, args) in enumerate(instrs):
        t = re.search(r"0x([0-9a-f]+)", args) if op.startswith("BRA") else None
        if t and int(t.group(1), 16) < a and int(t.group(1), 16) in addr:
            body = instrs[addr[int(t.group(1), 16)]:i + 1]
            if count(body, "FFMA2") and (best is None or len(body) < len(best)):
                best = body
    assert best is not None, "no backward branch closes a loop with FFMA2"
    return best


def test_filter_inner_loop_mix(sass):
    k = sass["nn_filter_kernel"]
    loop = _inner_loop(k)
    lds, ffma2, fmnmx3 = count(loop, "LDS.128"), count(loop, "FFMA2"), count(loop, "FMNMX3")
    # per LDS.128 quadruple (four model points x NN_R = 16 scene points = 64 pairs): 96 FFMA2 + 32 FMNMX3
    assert lds % 4 == 0 and lds >= 4
    assert ffma2 == 24 * lds and fmnmx3 == 8 * lds, (lds, ffma2, fmnmx3)
    # nothing else of substance: loop bookkeeping only (address, counter, compare, branch), no spills, no global memory
    other = len(loop) - lds - ffma2 - fmnmx3
    assert other <= 6, [op for _, op, _ in loop if not op.startswith(("LDS.128", "FFMA2", "FMNMX3"))]
    for bad in ("LDL", "STL", "LDG", "STG", "BAR", "FMNMX.", "FFMA."):
        assert count(loop, bad) == 0, bad
    # the scene coordinate enters FFMA2 as a scalar-broadcast operand: no duplicated register pairs per point
    for _, op, args in loop:
        if op == "FFMA2":
            assert args.split(",")[1].strip().replace(".reuse", "").endswith(".F32"), args


def test_filter_uses_tma_and_no_fp64_or_tensor_pipe(sass):
    k = sass["nn_filter_kernel"]
    assert count(k, "UBLKCP") >= 1                      # cp.async.bulk global -> shared
    assert count(k, "SYNCS") >= 2                       # mbarrier expect_tx / try_wait
    for pipe in ("DFMA", "DADD", "DMUL", "HMMA", "IMMA", "UTCHMMA", "UTCMMA", "UTCQMMA"):
        assert count(k, pipe) == 0, pipe


def test_peer_exchange_is_in_the_reduction_kernels(sass):
    for name in ("moments_kernel", "transform_kernel", "align_fused_kernel"):
        k = sass[name]
        assert count(k, "STG.E.64.STRONG.SYS") >= 1, name      # post into the peers' mailboxes over NVLink
        assert count(k, "LDG.E.64.STRONG.SYS") >= 1, name      # acquire-spin on / read of the own mailbox
        assert count(k, "MEMBAR.SC.SYS") + count(k, "MEMBAR.ALL.SYS") >= 1, name


def test_fp64_decision_is_not_contracted(sass):
    """dist2_64 = (dx*dx + dy*dy) + dz*dz with every product and sum rounded separately (src/cpu.cc:16-19).  Every call
    site is followed by one software square root (one MUFU.RSQ64H and at most six Newton DFMAs each): per site the
    uncontracted form leaves 3 DMUL + 5 DADD (three subtractions, two sums); a contracted one would leave ONE DMUL, THREE
    DADD and two more DFMAs."""
    for name in ("nn_refine_kernel", "nn_exact_kernel", "nn_tree_kernel", "nn_wtree_kernel"):
        k = sass[name]
        sites = count(k, "MUFU.RSQ64H")
        assert sites >= 1, name
        assert count(k, "DFMA") <= 6 * sites, (name, count(k, "DFMA"), sites)
        assert count(k, "DMUL") >= 3 * sites and count(k, "DADD") >= 5 * (sites - 1) - 4, (name, count(k, "DMUL"), count(k, "DADD"), sites)


def test_tree_and_filter_publish_progress_to_host(sass):
    # the first kernel of every iteration stores the loop progress into mapped host memory (icp_run's look-ahead)
    for name in ("nn_filter_kernel", "nn_wtree_kernel", "nn_tree_kernel"):
        assert count(sass[name], "STG.E.STRONG.SYS") >= 1, name
