"""N > 1 host logic on CPU: two gloo ranks run the sharded SVN iteration.  The compute stand-in is the
ORACLE (tests may use it); what is under test is slam_sam_b200.distributed -- ownership, the
zero-padded all-reduce and that all ranks end bit-identical to the single-process result."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "ndt_golden.npz")


class OracleBackedHandle:
    """Duck-types the two Handle methods distributed.svn_step needs, on top of the oracle."""

    def __init__(self, gold):
        import oracle_lib as O

        self.svn = O.OracleSvn(resolution=2.0, min_pts=3, K=8, use_gn=True, kernel_h=1.0, step_size=1.0)
        self.svn.set_target(gold["target"]); self.svn.set_source(gold["source"])

    def svn_stage1_local(self, poses, rank, world):
        rec = np.zeros((len(poses), 43))
        for k in range(rank, len(poses), world):
            s, g, H = self.svn.particle_derivatives(poses[k])
            rec[k, 0] = s; rec[k, 1:7] = g; rec[k, 7:] = H.reshape(36)
        return rec

    def svn_finish_step(self, poses, rec):
        import oracle_lib as O

        u = self.svn.stage2(poses, rec[:, 1:7], rec[:, 7:].reshape(-1, 6, 6))
        new = np.array(poses, float).copy()
        L = O.lib()
        for k in range(len(poses)):
            M = np.zeros((4, 4)); L.orc_pose_expmap(O._d(np.ascontiguousarray(u[k])), O._d(M))
            new[k] = poses[k] @ M
        return new, u


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist

    from slam_sam_b200 import distributed as D

    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    gold = dict(np.load(GOLD))
    h = OracleBackedHandle(gold)
    P = gold["svn_particles0"]
    assert list(D.local_indices(len(P), rank, world)) == [k for k in range(len(P)) if D.owner(k, world) == rank]
    P1, u = D.svn_step(h, P, rank, world)
    P2, u2 = D.svn_step(h, P1, rank, world)
    t = D.max_over_ranks(1.0 + rank); s = D.sum_over_ranks(1.0 + rank)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), P1=P1, u=u, P2=P2, t=t, s=s)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_svn_step_matches_single_process(tmp_path):
    world = 2
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0 = np.load(tmp_path / "r0.npz"); r1 = np.load(tmp_path / "r1.npz")
    for k in ("P1", "u", "P2"):
        assert np.array_equal(r0[k], r1[k])                      # replicas stay bit-identical
    assert float(r0["t"]) == 2.0 and float(r0["s"]) == 3.0       # max / sum over ranks
    gold = dict(np.load(GOLD))
    assert np.array_equal(r0["u"], gold["svn_step_gn1_updates"])  # == the single-process oracle step
    assert np.abs(r0["P1"] - gold["svn_step_gn1_particles1"]).max() <= 1e-12
