"""Launch-geometry sweep of the fused cell kernel on one case (developer tool).
Usage: python tools/tune.py N ORDER NCOMP "cellsPerWarp,warps,ringSteps,smemKB" ...   (0 = auto)"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import wenoext_b200 as wx  # noqa: E402
from bench import make_fields  # noqa: E402

n, order, ncomp = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
t0 = time.time()
case = wx.BoxCase((n, n, n), order)
print(f"setup {time.time() - t0:.1f}s cells {case.nCells}", flush=True)
vf_h, phi_h = make_fields(case, ncomp)
vf = torch.from_numpy(vf_h).cuda()
phi = torch.from_numpy(phi_h).cuda()
ref = None
for cfg in sys.argv[4:]:
    vals = [int(x) for x in cfg.split(",")] + [0, 0, 0]
    G, groups, slots, smem = vals[:4]
    for k, v in (("WENOEXT_CELLS", G), ("WENOEXT_WARPS", groups), ("WENOEXT_SLOTS", slots),
                 ("WENOEXT_SMEM_KB", smem)):
        if v > 0:
            os.environ[k] = str(v)
        else:
            os.environ.pop(k, None)
    plan = wx.Plan(case.desc)
    out = torch.empty((plan.nFaces, ncomp), dtype=torch.float64, device="cuda")
    for _ in range(3):
        plan.correction(0, 1, vf, phi, out=out)
    plan.enable_kernel_timing(True)
    for _ in range(10):
        plan.correction(0, 1, vf, phi, out=out)
    ms, k = plan.kernel_time()
    ms /= k
    gbs = plan.alg_bytes(ncomp) / ms / 1e6
    o = out.cpu().numpy()
    if ref is None:
        ref = o
    err = np.abs(o - ref).max()
    print(f"cfg cellsPerWarp={G} warps={groups} ringSteps={slots} smemKB={smem}: {ms:.4f} ms  {gbs:.0f} GB/s  "
          f"frac {gbs / 6547.8:.3f}  maxdiff_vs_first {err:.2e}", flush=True)
    plan.destroy()
