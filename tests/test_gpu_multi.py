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


# ---- one process, several devices: icp_group_* (SURVEY 8b icp_create(ctx, devices, ndev)) ---------------------------

def test_group_of_one_device_is_a_plain_context(clouds):
    import icp_b200
    m, p = clouds("cow_ref"), clouds("cow_tr1")
    plain = icp_b200.Context(0)
    plain.set_model(m)
    plain.set_scene(p)
    it0, hist0 = plain.run(30)
    g = icp_b200.Group([0])
    g.set_model(m)
    g.set_scene(p)
    it, hist = g.run(30)
    assert it == it0 == 7 and np.array_equal(hist, hist0)
    assert np.array_equal(g.get_scene(), plain.get_scene()) and np.array_equal(g.indices(), plain.indices())
    assert g.stats()["np"] == len(p) and g.transform()[0] == plain.transform()[0]
    g.reset_scene()
    idx, Y = g.closest()
    plain.reset_scene()
    idx0, Y0 = plain.closest()
    assert np.array_equal(idx, idx0) and np.array_equal(Y, Y0)
    g.close()
    plain.close()
    with pytest.raises(icp_b200.IcpError):
        icp_b200.Group([0, 0])                      # a device listed twice: two spinning exchanges would share one GPU
    with pytest.raises(icp_b200.IcpError):
        icp_b200.Group([])


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("nn", ["brute", "tree"])
@pytest.mark.parametrize("ndev", [2, 3, 4, 8])
def test_group_sharded_parity(clouds, golden, nn, ndev):
    # the sharded loop inside ONE process against the golden vectors: iteration count, error history (1e-5), indices of
    # the last iteration bit-exact, new_p, transform; plus the collective one-shot calls against a single context
    import icp_b200
    if _n_gpus() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    for case, mname, pname in (("cow_tr1", "cow_ref", "cow_tr1"), ("horse_tr2", "horse_ref", "horse_tr2")):
        m, p, gd = clouds(mname), clouds(pname), golden(case)
        g = icp_b200.Group(list(range(ndev)))
        g.set_nn_mode(nn)
        g.set_model(m)
        g.set_scene(p)
        idx0, Y0 = g.closest()
        assert np.array_equal(idx0, gd["idx_first"])
        fa = g.find_alignment()
        one = icp_b200.Context(0)
        one.set_model(m)
        one.set_scene(p)
        one.closest(want_idx=False, want_Y=False)
        fb = one.find_alignment()
        assert abs(fa["s"] - fb["s"]) <= 1e-12 * abs(fb["s"]) and np.abs(fa["R"] - fb["R"]).max() < 1e-12
        assert abs(fa["err"] - fb["err"]) <= 1e-10 * abs(fb["err"])
        mu_p, mu_y = g.centroid()
        mu_p1, mu_y1 = one.centroid()
        assert np.abs(mu_p - mu_p1).max() < 1e-12 and np.abs(mu_y - mu_y1).max() < 1e-12
        assert abs(g.err_compute(False) - one.err_compute(False)) <= 1e-10 * abs(fb["err"])
        one.close()
        for rep in range(2):                         # exchange numbers keep counting across runs
            g.reset_scene()
            it, hist = g.run(30)
            assert it == int(gd["iters"])
            assert np.allclose(hist, gd["err"][:it], rtol=1e-5, atol=1e-18)
            assert np.array_equal(g.indices(), gd["idx_last"])
            assert np.abs(g.get_scene() - gd["new_p"]).max() < 1e-9
            s, R, t = g.transform()
            assert abs(s - gd["s"][-1]) <= 1e-5 * abs(gd["s"][-1]) and np.allclose(R, gd["R"][-1], rtol=1e-5, atol=1e-9)
        st = g.stats()
        assert st["np"] == len(p) and st["nn_pairs"] >= len(p) * len(m) * it
        g.close()


def test_group_batch_run_deals_pairs_to_devices():
    import icp_b200
    from conftest import synthetic_pair
    from oracle import oracle as O
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    models, scenes = [], []
    for k in range(5):
        m, p, _ = synthetic_pair(1200 + 101 * k, seed=300 + k, angle_deg=2.0 + k)
        models.append(m)
        scenes.append(p)
    res = icp_b200.batch_run(models, scenes, max_iter=20, lanes=2, devices=[0, 1])
    assert res["kernels_launched"] > 0
    for k in range(5):
        ro = O.icp_run(models[k], scenes[k], 20)
        assert res["iters"][k] == ro["iters"], k
        assert abs(res["err"][k] - ro["err"][-1]) <= 1e-5 * abs(ro["err"][-1]) + 1e-18
        assert np.abs(res["new_p"][k] - ro["new_p"]).max() < 1e-9
