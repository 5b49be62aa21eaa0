"""Multi-GPU plumbing: one process per GPU over torch.distributed (NCCL on the box, gloo in CPU tests).

The path shards at two granularities and nowhere else:
  * SVN particles: particle k is evaluated on rank k % world (stage 1, svn_ndt_impl.hpp:758-781);
    stage 2 couples all particles (:795-808).  Product path (SvnComm -> ndtb_svn_align_sharded): the
    last block of a particle's derivative kernel stores the particle's 43-double record into every
    rank's window over NVLink and releases a flag; every rank's Stein kernel acquires the K flags and
    updates all K particles redundantly, so poses never travel and replicas stay bit-identical.
    svn_step / svn_align below are the host-staged form of the same split (stage 1 local, ONE
    all-reduce of a zero-padded K x 43 buffer, Stein update) -- used by the CPU (gloo) tests of the
    sharding logic and by callers that want to move the records themselves.
  * replay batches: independent alignments against a replicated map, scan i on rank i % world, no
    data-path collective at all.
A single alignment is NOT split across GPUs (one evaluation is ~10 us of work): replicas only.
"""
import numpy as np


def owner(k, world):
    return k % world


def local_indices(n, rank, world):
    """Indices of the work items (particles / scans) rank `rank` owns."""
    return np.arange(rank, n, world)


def allreduce_records(rec, device=None):
    """Sum a K x 43 record buffer over all ranks (zero rows for particles a rank does not own)."""
    import torch
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size() == 1:
        return rec
    t = torch.from_numpy(np.ascontiguousarray(rec))
    if dist.get_backend() == "nccl":
        t = t.cuda(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


def svn_step(handle, poses, rank, world, device=None):
    """One SVN iteration with particles sharded over `world` ranks.  Returns (new_poses, updates)."""
    rec = handle.svn_stage1_local(poses, rank, world)
    rec = allreduce_records(rec, device)
    return handle.svn_finish_step(poses, rec)


def svn_align(handle, prior, particles, rank, world, max_iterations=50, stop_threshold=1e-4, device=None):
    """align(source, prior) of svn_ndt (svn_ndt_impl.hpp:675-964) with the particles sharded over `world` ranks:
    the loop of ndtb_svn_align, stage 1 local + one all-reduce per iteration.  Returns (mean pose, particles,
    iterations, converged); every rank returns bit-identical values."""
    from . import ndtb

    P = np.ascontiguousarray(particles, np.float64).copy()
    mean_cur = np.ascontiguousarray(prior, np.float64).reshape(4, 4)
    it, converged = 0, False
    for it in range(1, max_iterations + 1):
        mean_last = mean_cur
        P, _ = svn_step(handle, P, rank, world, device)
        mean_cur = ndtb.svn_mean_pose(P, prior)
        if ndtb.svn_pose_distance(mean_last, mean_cur) < stop_threshold:
            converged = True
            break
    return mean_cur, P, it, converged


class SvnComm:
    """Communicator of the device-resident sharded SVN align (include/ndtb.h, ndtb_comm_*): the record windows of all ranks are
    mapped into every rank (CUDA IPC, peer access over NVLink); torch.distributed only carries the 64-byte window handles once.
    After that `align` is the library's collective ndtb_svn_align: no NCCL call and no host copy per iteration."""

    def __init__(self, handle, rank, world, device=None):
        import torch
        import torch.distributed as dist

        self.h, self.rank, self.world = handle, rank, world
        mine = torch.from_numpy(handle.comm_window_handle().copy())
        if dist.get_backend() == "nccl":
            mine = mine.cuda(device)
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        handle.comm_connect(rank, world, np.stack([t.cpu().numpy() for t in allh]))
        dist.barrier()   # every window is zeroed and every rank's stamps restarted before the first collective align

    def align(self, prior, particles):
        from . import ndtb

        h = self.h
        prior = np.ascontiguousarray(prior, np.float64).reshape(16)
        init = np.ascontiguousarray(particles, np.float64)
        K = init.shape[0]
        out = np.zeros((K, 4, 4)); r = ndtb.SvnResult()
        h.check(h.L.ndtb_svn_align_sharded(h.h, ndtb._ptr(prior, ndtb.c_d), ndtb._ptr(init, ndtb.c_d), 0, ndtb.C.byref(r), ndtb._ptr(out, ndtb.c_d)))
        return dict(final_pose=np.array(r.final_pose).reshape(4, 4), final_covariance=np.array(r.final_covariance).reshape(6, 6),
                    converged=bool(r.converged), iterations=r.iterations, particles=out)

    def close(self):
        import torch.distributed as dist

        dist.barrier()
        self.h.comm_close()


def max_over_ranks(value, device=None):
    import torch
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda(device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device=None):
    import torch
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
