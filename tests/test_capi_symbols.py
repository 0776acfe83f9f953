"""The C-ABI library loads and exports every function include/wenoext_b200.h declares, the
ctypes binding covers exactly that set, and without a GPU the compute entry points fail
loudly instead of falling back to anything."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import wenoext_b200 as wx
from wenoext_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions(header="wenoext_b200.h"):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"^(?:int|const char\*)\s+(wenoext_\w+)\s*\(", txt, flags=re.M)))


def test_every_declared_symbol_is_exported_and_bound():
    for header, path, table in (("wenoext_b200.h", _capi.LIB_PATH, _capi.SYMBOLS),
                                ("wenoext_setup.h", _capi.SETUP_LIB_PATH, _capi.SETUP_SYMBOLS)):
        names = declared_functions(header)
        assert len(names) >= 10
        raw = C.CDLL(path)
        for nm in names:
            assert hasattr(raw, nm), f"{nm} declared in {header} but not exported by {path}"
        assert sorted(table) == names
    wx.lib().product()  # sets restype/argtypes for all of them, raises on a missing one
    wx.lib().setup()


def test_setup_library_has_no_cuda_and_cpu_users_do_not_map_the_product_library():
    """The descriptor producers live in libwenoext_setup.so (plain C++): building a case and running the
    CPU checkers must not load the GPU library (the reference arm mapped the product .so)."""
    import subprocess
    import sys
    code = (
        "import wenoext_b200 as wx, numpy as np\n"
        "from wenoext_b200 import _capi\n"
        "from tests.oracle_binding import Oracle\n"
        "c = wx.BoxCase((5, 5, 5), 2)\n"
        "Oracle('c').correction(c, 0, 1, np.zeros(c.nCells), np.zeros(c.nFaces))\n"
        "maps = open('/proc/self/maps').read()\n"
        "assert 'libwenoext_setup.so' in maps\n"
        "assert 'libwenoext_b200.so' not in maps and not _capi.product_loaded()\n"
        "assert 'libcudart' not in maps\n")
    subprocess.run([sys.executable, "-c", code], check=True, cwd=ROOT)


def test_struct_layouts_match_the_header(tmp_path):
    """sizeof of every public struct as gcc lays it out from the header == the ctypes mirror"""
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "wenoext_setup.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(wenoext_patch),sizeof(wenoext_plan_desc),sizeof(wenoext_plan_info),sizeof(wenoext_box_spec),'
                   'sizeof(wenoext_setup_params),sizeof(wenoext_mesh_desc),sizeof(int),sizeof(wenoext_patch64),'
                   'sizeof(wenoext_plan_desc64),sizeof(wenoext_ls_batch));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(t) for t in (_capi.Patch, _capi.PlanDesc, _capi.PlanInfo, _capi.BoxSpec, _capi.SetupParams,
                                  _capi.MeshDesc, C.c_int, _capi.Patch64, _capi.PlanDesc64, _capi.LsBatch)]
    assert got == want


def test_no_gpu_means_loud_failure():
    if wx.lib().wenoext_device_count() > 0:
        pytest.skip("a GPU is present")
    case = wx.BoxCase((4, 4, 4), 2)
    with pytest.raises(wx.WenoextError):
        wx.Plan(case.desc)
    z = np.zeros(9)
    rows = np.array([1], dtype=np.int32)
    start = np.array([0, 9], dtype=np.int64)
    zi = np.zeros(1, dtype=np.int32)
    rc = wx.lib().wenoext_pinv_batched(1, 9, rows.ctypes.data_as(_capi.i32p), start.ctypes.data_as(_capi.i64p),
                                       z.ctypes.data_as(_capi.f64p), 1e-5, z.ctypes.data_as(_capi.f64p),
                                       zi.ctypes.data_as(_capi.i32p), None, zi.ctypes.data_as(_capi.i32p), 0)
    assert rc != 0
    with pytest.raises(wx.WenoextError):
        wx.BoxCase((4, 4, 4), 2, useClassCache=False, gpuPinv=True)


def test_plan_create_rejects_out_of_range_ids():
    """Every index the kernels follow unchecked is validated once at plan creation: a bad
    descriptor or a stale cache fails with a message, with or without a GPU."""
    case = wx.BoxCase((5, 5, 5), 2, useClassCache=False)
    d = case.desc

    def expect(msg):
        with pytest.raises(wx.WenoextError, match=msg):
            wx.Plan(d)

    e = d.stencilEntryStart[0] + 1
    keep = d.stencilsID[e]
    d.stencilsID[e] = case.nCells + 7
    expect("stencil cell id out of range")
    d.stencilsID[e] = keep
    keep = d.owner[3]
    d.owner[3] = -2
    expect("owner out of range")
    d.owner[3] = keep
    keep = d.stencilMatrix[0]
    d.stencilMatrix[0] = d.nUniqueMatrices
    expect("stencilMatrix out of range")
    d.stencilMatrix[0] = keep
    keep = d.cellFaces[0]
    d.cellFaces[0] = case.nFaces
    expect("cellFaces out of range")
    d.cellFaces[0] = keep
    keep = d.patches[0].start
    d.patches[0].start = 0
    expect("outside the boundary faces")
    d.patches[0].start = keep
    # parallel lists
    par = wx.BoxCase((6, 4, 4), 2, nProc=(2, 1, 1), myProc=0)
    other = wx.BoxCase((6, 4, 4), 2, nProc=(2, 1, 1), myProc=1)
    wx.exchange_halo_lists([par, other])
    dp = par.desc
    keep = dp.receiveHaloSize[1]
    dp.receiveHaloSize[1] = 1
    with pytest.raises(wx.WenoextError, match="halo id beyond receiveHaloSize"):
        wx.Plan(dp)
    dp.receiveHaloSize[1] = keep
    keep = dp.sendHaloCellID[0]
    dp.sendHaloCellID[0] = par.nCells
    with pytest.raises(wx.WenoextError, match="sendHaloCellID out of range"):
        wx.Plan(dp)
    dp.sendHaloCellID[0] = keep
    if wx.lib().wenoext_device_count() == 0:
        with pytest.raises(wx.WenoextError, match="no CUDA device"):  # the intact descriptor gets that far
            wx.Plan(d)


def test_labels64_descriptor_is_narrowed_and_checked():
    """wenoext_plan_create_labels64 (WM_LABEL_SIZE=64 builds): a value that does not fit 32 bits is refused
    with the array's name; an intact descriptor reaches the same checks as the 32-bit one."""
    case = wx.BoxCase((5, 5, 5), 2, useClassCache=False)
    d64, keep = wx.labels64(case.desc)
    sid = next(a for a in keep if isinstance(a, np.ndarray) and a.size == case.desc.stencilEntryStart[case.desc.nStencils])
    old = sid[1]
    sid[1] = 1 << 40
    with pytest.raises(wx.WenoextError, match="stencilsID holds a value that does not fit 32 bits"):
        wx.Plan(d64)
    sid[1] = case.nCells + 3  # fits, but is no cell: the id validation of the 32-bit path sees it
    with pytest.raises(wx.WenoextError, match="stencil cell id out of range"):
        wx.Plan(d64)
    sid[1] = old
    d64.nCells = 1 << 33
    with pytest.raises(wx.WenoextError, match="nCells out of range"):
        wx.Plan(d64)
    d64.nCells = case.nCells
    if wx.lib().wenoext_device_count() == 0:
        with pytest.raises(wx.WenoextError, match="no CUDA device"):
            wx.Plan(d64)


def test_ls_batch_is_validated_before_any_device_work():
    """wenoext_ls_pinv_batched checks its index lists on the host (bad input fails with a message, with or without a
    GPU); an intact batch then needs the device and fails loudly without one."""
    f64, i32, i64 = _capi.f64p, _capi.i32p, _capi.i64p
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], dtype=np.float64)
    tri = np.array([[0, 2, 1], [0, 1, 3], [0, 3, 2], [1, 2, 3]], dtype=np.int32)
    triStart = np.array([0, 4], dtype=np.int64)
    cen, vol = np.array([[0.25, 0.25, 0.25]]), np.array([1.0 / 6.0])
    J, ref, det = np.eye(3).ravel().copy(), np.zeros(3), np.ones(1)
    fc = np.zeros(1, dtype=np.int32)
    volI = np.zeros(27)
    pf, pc = np.zeros(1, dtype=np.int32), np.zeros(1, dtype=np.int32)
    b = _capi.LsBatch()
    b.polOrder, b.nDvt = 2, 9
    for k in range(3):
        b.dim[k] = 2
    b.nPoints, b.nSrcCells, b.nFrames, b.nPairs, b.nMat = 4, 1, 1, 1, 0
    p = lambda a, t: a.ctypes.data_as(t)  # noqa: E731
    b.points, b.triStart, b.triPoints, b.cellCentres, b.cellVolumes = p(pts, f64), p(triStart, i64), p(tri, i32), p(cen, f64), p(vol, f64)
    b.JInv, b.refPoint, b.refDet, b.frameCell, b.volI = p(J, f64), p(ref, f64), p(det, f64), p(fc, i32), p(volI, f64)
    b.pairFrame, b.pairCell = p(pf, i32), p(pc, i32)
    rows = np.empty((1, 9))
    call = lambda: wx.lib().wenoext_ls_pinv_batched(C.byref(b), 1e-5, None, None, None, None, p(rows, f64), None, 0)  # noqa: E731
    tri[1, 2] = 9
    assert call() == 1 and b"triPoints out of range" in wx.lib().wenoext_last_error()
    tri[1, 2] = 3
    pc[0] = 5
    assert call() == 1 and b"pair out of range" in wx.lib().wenoext_last_error()
    pc[0] = 0
    b.polOrder = 9
    assert call() == 1 and b"order" in wx.lib().wenoext_last_error()
    b.polOrder = 2
    if wx.lib().wenoext_device_count() == 0:
        assert call() != 0  # no CPU fallback
    else:
        assert call() == 0 and np.isfinite(rows).all()
