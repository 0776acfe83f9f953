"""N > 1 host logic on CPU: two processes (torch.distributed, gloo, 127.0.0.1).

Each rank builds its own block of the decomposed box, ships its halo request lists
to the owner ranks (distributeStencils, WENOBase.C:334-407), and runs the two
exchange steps of the hot path (collectData WENOCoeff.C:224-296, swapData
WENOUpwindFit.C:326-385) around the oracle. Rank 0 checks the lists and the result
against the in-process emulation, which test_oracle_cpu.py ties to the serial run."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import wenoext_b200 as wx
from tests import helpers as H
from tests.oracle_binding import Oracle, _send_list, correction_distributed, correction_multirank

N, NPROC = (8, 6, 6), (2, 1, 1)
CASES = [(0, 0), (0, 1), (1, 1)]  # (scheme, bounded)


def _fields(case):
    lo, hi = np.zeros(3), np.ones(3)
    X = (case.C - lo) / (hi - lo)
    return H.vector_field(X, seed=4), H.face_flux(case.Sf, U=(1.0, -1.0, 0.5), zero_frac=0.03)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        case = wx.BoxCase(N, 2, nProc=NPROC, myProc=rank)
        wx.exchange_halo_lists_distributed(case)
        sends = [_send_list(case, p).tolist() for p in range(world)]
        orc = Oracle("c")
        vf, flux = _fields(case)
        outs = [correction_distributed(orc, case, s, b, vf, flux) for s, b in CASES]
        q.put((rank, sends, [np.asarray(o) for o in outs]))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_equal_in_process():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in procs:
        r, sends, outs = q.get(timeout=240)
        got[r] = (sends, outs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # in-process emulation of the same two ranks
    cases = [wx.BoxCase(N, 2, nProc=NPROC, myProc=r) for r in range(2)]
    wx.exchange_halo_lists(cases)
    orc = Oracle("c")
    fl = [_fields(c) for c in cases]
    for r in range(2):
        assert got[r][0] == [_send_list(cases[r], p).tolist() for p in range(2)]
    for i, (s, b) in enumerate(CASES):
        ref = correction_multirank(orc, cases, s, b, [f[0] for f in fl], [f[1] for f in fl])
        for r in range(2):
            np.testing.assert_array_equal(got[r][1][i], ref[r])
