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
        # wenoext_div on a parallel plan: host buffers (pipelined, face field stays on the device) against
        # device buffers, same bits
        rng = np.random.default_rng(11 + rank)
        bnd = rng.uniform(-1, 1, (mine.nFaces - mine.nInternalFaces, 3))
        vol = rng.uniform(0.5, 1.5, mine.nCells) * 1e-3
        tt = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()  # noqa: E731
        dd = plan.div(0, 1, tt(vfs[rank]), tt(fls[rank]), tt(bnd), tt(vol)).cpu().numpy()
        dh = np.asarray(plan.div(0, 1, vfs[rank], fls[rank], bnd, vol))
        errs.append(0.0 if np.array_equal(dd, dh) else 1.0)
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


def _timeout_worker(rank, world, port, q):
    """rank 1 arrives 3 s late; rank 0 waits 0.5 s for its halo (WENOEXT_HALO_TIMEOUT_S)"""
    import time

    import torch
    import torch.distributed as dist

    import wenoext_b200 as wx
    from tests import helpers as H
    from tests.oracle_binding import Oracle

    os.environ["WENOEXT_HALO_TIMEOUT_S"] = "0.5" if rank == 0 else "60"
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        cases = [wx.BoxCase((24, 16, 16), 2, nProc=NPROC[world], myProc=r) for r in range(world)]
        wx.exchange_halo_lists(cases)
        mine = cases[rank]
        plan = wx.Plan(mine.desc, device=rank)
        uid = [wx.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        plan.attach_nccl(uid[0], rank, world)
        vf = H.vector_field(mine.C, seed=4)
        fl = H.face_flux(mine.Sf, U=(1.0, -1.0, 0.5))
        dist.barrier()
        torch.cuda.synchronize()
        res = {"rank": rank}
        if rank == 1:
            time.sleep(3.0)
            plan.correction(0, 1, vf, fl)
        else:
            try:
                plan.correction(0, 1, vf, fl)
                res["first"] = "no error"
            except wx.WenoextError as e:
                res["first"] = str(e)
            try:  # sticky: the plan's results can no longer be trusted
                plan.correction(0, 1, vf, fl)
                res["second"] = "no error"
            except wx.WenoextError as e:
                res["second"] = str(e)
            # the CUDA context is alive: a fresh serial plan on the same device still works
            serial = wx.BoxCase((8, 8, 8), 2, useClassCache=False)
            p2 = wx.Plan(serial.desc, device=rank)
            svf, sfl = H.vector_field(serial.C, seed=1), H.rotating_flux(serial)
            got = p2.correction(0, 1, svf, sfl)
            res["serial_err"] = H.normwise_err(np.asarray(got), Oracle("c").correction(serial, 0, 1, svf, sfl))
            p2.destroy()
        q.put(res)
        q.close()
        q.join_thread()  # flush before the hard exit below
    finally:
        os._exit(0)  # skip the destructors: the two ranks are not at the same point of the exchange


def test_halo_timeout_is_an_error_code_not_a_trap():
    """A halo that does not arrive within WENOEXT_HALO_TIMEOUT_S makes the call return
    WENOEXT_ERR_TIMEOUT (sticky for the plan); the CUDA context stays usable."""
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_timeout_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in procs:
        r = q.get(timeout=120)
        got[r["rank"]] = r
    for p in procs:
        p.join(timeout=60)
    r0 = got[0]
    assert "wenoext error 5" in r0["first"] and "did not arrive" in r0["first"], r0
    assert "wenoext error 5" in r0["second"], r0
    assert r0["serial_err"] <= 1e-12
