"""One process per GPU over NCCL (the production multi-GPU path): needs >= 2 GPUs,
skipped otherwise. Each rank reconstructs its block; every rank checks its result
against the in-process multi-rank emulation of the oracle."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NPROC = {2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}
# small: plain host path; large with 2 slabs: the slab-pipelined host path of parallel plans
# "nccl": exchange #1 by NCCL send/recv (WENOEXT_P2P=0) instead of the one-sided push over
# peer memory that is the default with one process per GPU
MESHES = {"small": ((24, 16, 16), None), "large": ((64, 32, 32), "2"), "nccl": ((24, 16, 16), None)}


def _worker(rank, world, port, q, mesh):
    import torch
    import torch.distributed as dist

    import wenoext_b200 as wx
    from tests import helpers as H
    from tests.oracle_binding import Oracle, correction_multirank

    N, slabs = MESHES[mesh]
    if slabs:
        os.environ["WENOEXT_E2E_SLABS"] = slabs
    if mesh == "nccl":
        os.environ["WENOEXT_P2P"] = "0"
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        cases = [wx.BoxCase(N, 2, nProc=NPROC[world], myProc=r) for r in range(world)]
        wx.exchange_halo_lists(cases)  # every process builds all ranks: the oracle needs them
        mine = cases[rank]
        plan = wx.Plan(mine.desc, device=rank)
        uid = [wx.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        plan.attach_nccl(uid[0], rank, world)
        vfs = [H.vector_field(c.C, seed=4) for c in cases]
        fls = [H.face_flux(c.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.03) for c in cases]
        errs = []
        for scheme, bounded in ((0, 0), (0, 1), (1, 1)):
            ref = correction_multirank(Oracle("c"), cases, scheme, bounded, vfs, fls)[rank]
            for _ in range(2):  # twice: buffers and events are reused across calls
                got = plan.correction(scheme, bounded, vfs[rank], fls[rank])
            errs.append(H.normwise_err(np.asarray(got).reshape(ref.shape), ref))
            # page-locked buffers from the library: the faces that boundary cells finalise late are
            # written into the result by the device (zero-copy) instead of scattered by the host;
            # and the flux uploaded once for the "time step"
            pv, po = wx.pinned_empty(vfs[rank].shape), wx.pinned_empty((mine.nFaces, 3))
            pv[...] = vfs[rank]
            po[...] = np.nan
            plan.set_face_flux(fls[rank])
            plan.correction(scheme, bounded, pv, None, out=po)
            plan.set_face_flux(None)
            errs.append(H.normwise_err(np.asarray(po).reshape(ref.shape), ref))
            dvf = torch.from_numpy(vfs[rank]).cuda()
            dfl = torch.from_numpy(fls[rank]).cuda()
            out = torch.empty((mine.nFaces, 3), dtype=torch.float64, device="cuda")
            plan.correction(scheme, bounded, dvf, dfl, out=out)
            errs.append(H.normwise_err(out.cpu().numpy().reshape(ref.shape), ref))
        # scalar field (value stride 1: four cells per 32-byte sector at the halo boundary)
        svfs = [np.ascontiguousarray(v[:, 0]) for v in vfs]
        ref = correction_multirank(Oracle("c"), cases, 0, 1, svfs, fls)[rank]
        out = torch.empty((mine.nFaces,), dtype=torch.float64, device="cuda")
        for _ in range(2):
            plan.correction(0, 1, torch.from_numpy(svfs[rank]).cuda(), torch.from_numpy(fls[rank]).cuda(), out=out)
        errs.append(H.normwise_err(out.cpu().numpy().reshape(ref.shape), ref))
        q.put((rank, errs))
        dist.barrier()
        plan.destroy()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mesh", ["small", "large", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_ranks_match_oracle(world, mesh):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, mesh)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r in range(world):
        assert max(got[r]) <= 1e-12, (r, got[r])
