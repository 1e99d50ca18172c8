"""Multi-GPU path on real devices.  test_peer_mailbox_self_exchange needs one GPU (a world of one rank posts
to and reads from its own mailbox: the whole in-kernel exchange protocol runs); test_two_gpu_sharded_parity
needs two and is skipped otherwise (run it with `gpurun --gpus 2`)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("align", ["split", "fused"])
def test_peer_mailbox_self_exchange(clouds, monkeypatch, align):
    import icp_b200
    monkeypatch.setenv("ICP_ALIGN", align)
    m, p = clouds("cow_ref"), clouds("cow_tr1")
    plain = icp_b200.Context(0)
    plain.set_model(m)
    plain.set_scene(p)
    it0, hist0 = plain.run(30)
    ctx = icp_b200.Context(0)
    ctx.comm_init(icp_b200.comm_unique_id(), 0, 1)
    ctx.comm_peer_attach([ctx.comm_peer_handle()], 0, 1)
    ctx.set_model(m)
    ctx.set_scene(p)
    for _ in range(3):                     # sequence numbers keep increasing across runs
        ctx.reset_scene()
        it, hist = ctx.run(30)
        assert it == it0 == 7
        assert np.array_equal(hist, hist0)  # one rank: the same sums in the same order
    assert np.array_equal(ctx.get_scene(), plain.get_scene())
    a = ctx.find_alignment()
    b = plain.find_alignment()
    assert a["err"] == b["err"] and np.array_equal(a["R"], b["R"])
    ctx.close()
    plain.close()


@pytest.mark.parametrize("nn", ["brute", "tree"])
def test_two_gpu_sharded_parity(nn):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, ICP_NN=nn)
    for case in ("cow_tr1", "horse_tr2"):
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                            "--master-addr", "127.0.0.1", "--master-port", "29533",
                            os.path.join(ROOT, "tools", "dist_check.py"), case], capture_output=True, text=True, timeout=600,
                           env=env)
        assert r.returncode == 0, r.stdout + r.stderr
        assert "FAILED" not in r.stdout
