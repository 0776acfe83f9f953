"""constant/WENOBase<order>/ cache: writer + reader of the eleven files of
WENOBase::writeList / readList (libWENOEXT/WENOBase/WENOBase.C:1196-1312).

Modelled on the reference's own round-trip test tests/src/WENOBaseIO-Test.C:114-187 (write
the lists, read them back, require equality of every list), plus byte-level known answers of
the stream framing and a check of the padded matrix blobs against the real Blaze container.
"""
import ctypes as C
import os
import struct

import numpy as np
import pytest

import wenoext_b200 as wx
from tests import helpers as H
from tests.oracle_binding import Oracle
from tests.test_box_fast import _arrays

FILES = ["sendProcList", "receiveProcList", "DimLists", "StencilIDs", "CellToProcMap",
         "Pseudoinverses", "sendHaloCellIDList", "receiveHaloSize", "B", "intBasTrans", "refFacAr"]


def _stencil_matrices(a):
    """per (cell, stencil) matrix, pool order removed"""
    return [None if u < 0 else a["md"][a["ms"][u]:a["ms"][u + 1]] for u in a["sm"]]


def _same_plan(c0, c1):
    a, b = _arrays(c0), _arrays(c1)
    d0, d1 = c0.desc, c1.desc
    for k in ("polOrder", "nDvt", "nCells", "nFaces", "nInternalFaces", "nStencils", "nProcs", "myProc",
              "nUniqueMatrices"):
        assert getattr(d0, k) == getattr(d1, k), k
    assert list(d0.dim) == list(d1.dim)
    assert c0.patches == c1.patches
    for k in ("owner", "neighbour", "cfs", "cf", "css", "ses", "sid", "B", "ibt", "rfa"):
        assert np.array_equal(a[k], b[k]), k
    if "c2p" in a:
        assert np.array_equal(a["c2p"], b["c2p"])
    else:  # a serial plan has no map: the cache carries -1 (local) / -4 (deleted)
        assert set(np.unique(b["c2p"])) <= {-1, -4}
        assert np.array_equal(b["c2p"] == -4, a["sid"] == -4)
    for ma, mb in zip(_stencil_matrices(a), _stencil_matrices(b)):
        assert (ma is None) == (mb is None)
        if ma is not None:
            assert np.array_equal(ma, mb)  # bitwise: blobs are copied, not re-derived
    np_ = d0.nProcs
    for k in ("sendProcList", "receiveProcList", "receiveHaloSize"):
        assert np.array_equal(wx._np_view(getattr(d0, k), (np_,), np.int32),
                              wx._np_view(getattr(d1, k), (np_,), np.int32)), k
    s0 = wx._np_view(d0.sendHaloStart, (np_ + 1,), np.int64)
    s1 = wx._np_view(d1.sendHaloStart, (np_ + 1,), np.int64)
    assert np.array_equal(s0, s1)
    assert np.array_equal(wx._np_view(d0.sendHaloCellID, (int(s0[-1]),), np.int32),
                          wx._np_view(d1.sendHaloCellID, (int(s1[-1]),), np.int32))


CASES = {
    "hex3d": dict(n=(6, 5, 4), order=2),
    "hex3d_order3_generic": dict(n=(7, 7, 7), order=3, useClassCache=False),
    "cavity2d": dict(n=(8, 8, 1), order=2, hi=(0.1, 0.1, 0.01), sides=(0, 0, 0, 0, 3, 3)),
    "tets_bestConditioned": dict(n=(3, 3, 3), order=2, tetSplit=6, bestConditioned=True, useClassCache=False),
    "cyclic": dict(n=(6, 6, 6), order=2, sides=(2, 2, 0, 0, 0, 0)),
}


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("labelBytes,simd", [(4, 8), (8, 4), (4, 1)])
def test_round_trip(tmp_path, name, labelBytes, simd):
    case = wx.BoxCase(**CASES[name])
    d = tmp_path / f"WENOBase{case.order}"
    case.write_cache(d, labelBytes=labelBytes, simdDoubles=simd)
    assert sorted(os.listdir(d)) == sorted(FILES)
    back = wx.CachedCase(d, mesh=case.mesh_desc())  # label size detected
    assert back.order == case.order
    _same_plan(case, back)
    # explicit label size, and the wrong one must fail loudly rather than mis-read
    wx.CachedCase(d, mesh=case.mesh_desc(), labelBytes=labelBytes)
    with pytest.raises(wx.WenoextError):
        wx.CachedCase(d, mesh=case.mesh_desc(), labelBytes=12 - labelBytes)


def test_round_trip_two_ranks(tmp_path):
    """processor directories: one cache per rank, halo lists included (WENOBase.H:122-138)"""
    cases = [wx.BoxCase((8, 6, 6), 2, nProc=(2, 1, 1), myProc=r) for r in range(2)]
    wx.exchange_halo_lists(cases)
    for r, case in enumerate(cases):
        d = tmp_path / f"processor{r}_WENOBase2"
        case.write_cache(d)
        back = wx.CachedCase(d, mesh=case.mesh_desc(), myProc=r)
        _same_plan(case, back)
        assert back.desc.nProcs == 2 and back.desc.receiveHaloSize[1 - r] > 0


def test_mesh_from_arrays_derives_cells(tmp_path):
    """without mesh.cells() the reader derives it as primitiveMesh::calcCells does"""
    case = wx.BoxCase((4, 4, 4), 2)
    d = tmp_path / "WENOBase2"
    case.write_cache(d)
    back = wx.CachedCase(d, owner=case.owner, neighbour=case.neighbour, patches=case.patches)
    a, b = _arrays(case), _arrays(back)
    assert np.array_equal(a["cfs"], b["cfs"])
    for c in range(case.nCells):
        fa = a["cf"][a["cfs"][c]:a["cfs"][c + 1]]
        fb = b["cf"][b["cfs"][c]:b["cfs"][c + 1]]
        assert sorted(fa) == sorted(fb)
        own = [f for f in fb if case.owner[f] == c]
        assert list(fb) == sorted(own) + sorted(set(fb) - set(own))


def test_mismatching_mesh_is_rejected(tmp_path):
    case = wx.BoxCase((4, 4, 4), 2)
    other = wx.BoxCase((5, 4, 4), 2)
    d = tmp_path / "WENOBase2"
    case.write_cache(d)
    with pytest.raises(wx.WenoextError, match="cells"):
        wx.CachedCase(d, mesh=other.mesh_desc())
    with pytest.raises(wx.WenoextError):
        wx.CachedCase(tmp_path / "missing", mesh=case.mesh_desc())
    # truncated file
    blob = (d / "B").read_bytes()
    (d / "B").write_bytes(blob[:len(blob) // 2])
    with pytest.raises(wx.WenoextError, match="B"):
        wx.CachedCase(d, mesh=case.mesh_desc())


def test_stream_framing_known_answers(tmp_path):
    """Byte-level layout of OpenFOAM's binary OSstream as the reference drives it:
    numbers as text, raw blocks in parentheses (WENOBase.C:1273-1311, List3D.H:308-315,
    BlazeIO.C:112-131)."""
    case = wx.BoxCase((3, 3, 3), 2)
    d = tmp_path / "WENOBase2"
    case.write_cache(d, labelBytes=4, simdDoubles=4)
    a = _arrays(case)
    nF, nC, K = case.nFaces, case.nCells, case.nDvt
    # List<scalar>: "\nN\n(" raw ")"
    blob = (d / "refFacAr").read_bytes()
    head = b"\n%d\n(" % nF
    assert blob.startswith(head) and blob.endswith(b")") and len(blob) == len(head) + 8 * nF + 1
    assert np.array_equal(np.frombuffer(blob[len(head):-1], dtype="<f8"), a["rfa"])
    # labelList
    assert (d / "sendProcList").read_bytes() == b"\n1\n(" + struct.pack("<i", -1) + b")"
    # intBasTrans: "nFaces\n" then per face and side "3\n3\n3\n(" 27 doubles ")"
    blob = (d / "intBasTrans").read_bytes()
    head = b"%d\n" % nF
    rec = b"3\n3\n3\n("
    assert blob.startswith(head + rec)
    assert len(blob) == len(head) + nF * 2 * (len(rec) + 27 * 8 + 1)
    first = np.frombuffer(blob[len(head) + len(rec):len(head) + len(rec) + 27 * 8], dtype="<f8")
    assert np.array_equal(first, a["ibt"][0, 0])
    # B: list of row-major 9x9 matrices padded to 12 columns: "9\n9\n12\n(" 108 doubles ")"
    blob = (d / "B").read_bytes()
    head = b"\n%d\n(" % nC + b"\n9\n9\n12\n("
    assert blob.startswith(head)
    m = np.frombuffer(blob[len(head):len(head) + 9 * 12 * 8], dtype="<f8").reshape(9, 12)
    assert np.array_equal(m[:, :9], a["B"][0]) and not m[:, 9:].any()
    # Pseudoinverses: database size first, LSmatrix table after the blobs
    blob = (d / "Pseudoinverses").read_bytes()
    assert blob.startswith(b"%d\n" % case.desc.nUniqueMatrices)
    # DimLists: nCells entries "\n3\n(" + 3 labels + ")"
    blob = (d / "DimLists").read_bytes()
    one = b"\n3\n(" + struct.pack("<3i", 2, 2, 2) + b")"
    assert blob == b"\n%d\n(" % nC + one * nC + b"\n)\n"


def test_reader_accepts_compact_list_forms(tmp_path):
    """OpenFOAM writes short non-contiguous lists inline (`1(...)`, `0()`); the token
    stream is whitespace-insensitive and the reader must be too."""
    case = wx.BoxCase((3, 3, 3), 2)
    d = tmp_path / "WENOBase2"
    case.write_cache(d)
    blob = (d / "sendHaloCellIDList").read_bytes()
    assert blob == b"\n1\n(\n0\n\n)\n"
    (d / "sendHaloCellIDList").write_bytes(b"1(0())")
    (d / "receiveHaloSize").write_bytes(b"1(" + struct.pack("<i", 0) + b")")
    back = wx.CachedCase(d, mesh=case.mesh_desc())
    _same_plan(case, back)


def test_matrix_blob_is_blazes_memory_image(tmp_path):
    """The writer's padded blob must be what the real blaze::DynamicMatrix holds in
    memory (BlazeIO.C:126-129 dumps M.data()); oracle/_ref is compiled with AVX2, so
    Blaze's SIMD width is 4 doubles there."""
    orc = Oracle("blaze")
    fn = orc.BL.weno_oracle_blaze_image
    fn.restype = C.c_int
    fn.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    case = wx.BoxCase((4, 4, 4), 2)
    d = tmp_path / "WENOBase2"
    case.write_cache(d, simdDoubles=4)
    a = _arrays(case)
    K = case.nDvt
    blob = (d / "Pseudoinverses").read_bytes()
    # first record of the database: "nU\nkey\ncount\nrows\ncols\nspacing\n(" raw ")"
    pos, nums = 0, []
    for _ in range(6):
        end = blob.index(b"\n", pos)
        nums.append(int(blob[pos:end]))
        pos = end + 1
    nU, key, count, rows, cols, spacing = nums
    assert rows == K and blob[pos:pos + 1] == b"("
    raw = np.frombuffer(blob[pos + 1:pos + 1 + rows * spacing * 8], dtype="<f8")
    # which pool matrix is it? the one with the smallest key, first inserted
    cand = [u for u in range(nU) if a["mc"][u] == cols and
            np.array_equal(a["md"][a["ms"][u]:a["ms"][u + 1]].reshape(K, cols), raw.reshape(rows, spacing)[:, :cols])]
    assert cand
    A = np.ascontiguousarray(a["md"][a["ms"][cand[0]]:a["ms"][cand[0] + 1]])
    img = np.empty(rows * spacing + 64)
    sp = fn(rows, cols, A.ctypes.data, img.ctypes.data, img.size)
    assert sp == spacing
    assert np.array_equal(img[:rows * spacing], raw)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["hex3d", "tets_bestConditioned", "cavity2d"])
def test_gpu_correction_from_cache_is_bitwise_equal(tmp_path, name):
    """a plan built from the cache gives the same bits as the plan built from the setup"""
    case = wx.BoxCase(**CASES[name])
    d = tmp_path / "WENOBase"
    case.write_cache(d, labelBytes=8, simdDoubles=8)
    back = wx.CachedCase(d, mesh=case.mesh_desc(), geometry=(case.C, case.Sf))
    Cn = case.C / case.C.max(axis=0)
    vf = H.vector_field(Cn)
    phi = H.rotating_flux(case)
    for scheme, spec in ((wx.WENOUpwindFit, "2 1"), (wx.WENOUpwindFit01, "2")):
        want = scheme(case, phi, spec).correction(vf)
        got = scheme(back, phi, spec).correction(vf)
        assert np.array_equal(got, want)
        ref = Oracle("c").correction(back, scheme._scheme, True, vf, phi)
        assert H.normwise_err(got, ref) < 1e-12
    wx.WENOBase.clear()


# ---- case directories: constant/polyMesh/{owner,neighbour,boundary} + constant/WENOBase<order> ----
FOAM_BANNER = ("/*--------------------------------*- C++ -*----------------------------------*\\\n"
               "| =========                 |                                                 |\n"
               "| \\\\      /  F ield         | OpenFOAM: The Open Source CFD Toolbox           |\n"
               "\\*---------------------------------------------------------------------------*/\n")
SEP = "// * * * * * * * * * * * * * * * * * * * * * * * * * * * * * * * * * * * * * //\n\n"


def _foam_header(cls, obj, fmt, note=None, arch=None):
    h = FOAM_BANNER + "FoamFile\n{\n    version     2.0;\n    format      %s;\n" % fmt
    if arch:
        h += '    arch        "%s";\n' % arch
    h += "    class       %s;\n" % cls
    if note:
        h += '    note        "%s";\n' % note
    h += '    location    "constant/polyMesh";\n    object      %s;\n}\n' % obj
    return h + SEP


def write_polymesh(case, d, binary=False, labelBytes=4, with_note=True):
    """what blockMesh / decomposePar leave in constant/polyMesh for the addressing"""
    os.makedirs(d, exist_ok=True)
    note = (f"nPoints:0  nCells:{case.nCells}  nFaces:{case.nFaces}  nInternalFaces:{case.nInternalFaces}"
            if with_note else None)
    for name, arr in (("owner", case.owner), ("neighbour", case.neighbour)):
        if binary:
            arch = f"LSB;label={8 * labelBytes};scalar=64"
            raw = arr.astype("<i4" if labelBytes == 4 else "<i8").tobytes()
            body = _foam_header("labelList", name, "binary", note, arch).encode() + b"\n%d\n(" % arr.size + raw + b")\n"
        else:
            body = (_foam_header("labelList", name, "ascii", note) + "\n%d\n(\n" % arr.size +
                    "\n".join(str(int(v)) for v in arr) + "\n)\n").encode()
        with open(os.path.join(d, name), "wb") as f:
            f.write(body + b"\n\n// ************************************************************************* //\n")
    kinds = {wx.PATCH_GENERIC: "wall", wx.PATCH_EMPTY: "empty", wx.PATCH_CYCLIC: "cyclic",
             wx.PATCH_PROCESSOR: "processor"}
    txt = _foam_header("polyBoundaryMesh", "boundary", "ascii") + "%d\n(\n" % len(case.patches)
    for i, (typ, start, size, nbr) in enumerate(case.patches):
        txt += "    patch%d\n    {\n        type            %s;\n" % (i, kinds[typ])
        if typ == wx.PATCH_GENERIC:
            txt += "        inGroups        1(wall);\n"
        if typ == wx.PATCH_CYCLIC:
            txt += "        inGroups        1(cyclic);\n        matchTolerance  0.0001;\n"
            txt += "        transform       unknown; // comment inside\n        neighbourPatch  patch%d;\n" % nbr
        if typ == wx.PATCH_PROCESSOR:
            txt += "        inGroups        1(processor);\n        myProcNo        %d;\n" % case.myProc
            txt += "        neighbProcNo    %d;\n" % nbr
        txt += "        nFaces          %d;\n        startFace       %d;\n    }\n" % (size, start)
    txt += ")\n\n// ************************************************************************* //\n"
    with open(os.path.join(d, "boundary"), "w") as f:
        f.write(txt)


@pytest.mark.parametrize("name", ["hex3d", "cavity2d", "cyclic", "tets_bestConditioned"])
@pytest.mark.parametrize("binary,labelBytes,note", [(False, 4, True), (True, 4, False), (True, 8, True)])
def test_case_directory(tmp_path, name, binary, labelBytes, note):
    case = wx.BoxCase(**CASES[name])
    root = tmp_path / "case"
    write_polymesh(case, root / "constant" / "polyMesh", binary, labelBytes, note)
    case.write_cache(root / "constant" / f"WENOBase{case.order}", labelBytes=labelBytes)
    back = wx.CachedCase.from_case(root, case.order)
    a, b = _arrays(case), _arrays(back)
    assert back.patches == case.patches
    for k in ("owner", "neighbour", "cfs", "css", "ses", "sid", "B", "ibt", "rfa"):
        assert np.array_equal(a[k], b[k]), k
    with pytest.raises(wx.WenoextError, match=f"WENOBase{case.order + 1}"):  # no cache for that order
        wx.CachedCase.from_case(root, case.order + 1)


def test_case_directory_two_ranks(tmp_path):
    cases = [wx.BoxCase((8, 6, 6), 2, nProc=(2, 1, 1), myProc=r) for r in range(2)]
    wx.exchange_halo_lists(cases)
    for r, case in enumerate(cases):
        root = tmp_path / f"processor{r}"
        write_polymesh(case, root / "constant" / "polyMesh")
        case.write_cache(root / "constant" / "WENOBase2")
        back = wx.CachedCase.from_case(root, 2, myProc=r)
        assert back.patches == case.patches
        assert any(p[0] == wx.PATCH_PROCESSOR and p[3] == 1 - r for p in back.patches)
        assert back.desc.nProcs == 2


def test_case_directory_mismatch(tmp_path):
    case = wx.BoxCase((4, 4, 4), 2)
    other = wx.BoxCase((5, 4, 4), 2)
    root = tmp_path / "case"
    write_polymesh(other, root / "constant" / "polyMesh")
    case.write_cache(root / "constant" / "WENOBase2")
    with pytest.raises(wx.WenoextError, match="do not belong together"):
        wx.CachedCase.from_case(root, 2)


def test_polymesh_compact_and_commented_forms(tmp_path):
    """ASCII lists on one line, a uniform list `N{v}`, comments between tokens and a
    sub-dictionary inside a patch entry: all legal OpenFOAM, all must parse."""
    case = wx.BoxCase((2, 1, 1), 2, useClassCache=False)  # 2 cells, 1 internal face, 10 boundary faces
    root = tmp_path / "case"
    pm = root / "constant" / "polyMesh"
    write_polymesh(case, pm)
    hdr = _foam_header("labelList", "owner", "ascii")
    own = " ".join(str(int(v)) for v in case.owner)
    (pm / "owner").write_text(hdr + f"{case.nFaces} // count\n( {own} ) /* done */\n")
    (pm / "neighbour").write_text(_foam_header("labelList", "neighbour", "ascii") + "1{1}\n")
    b = (pm / "boundary").read_text()
    b = b.replace("        type            wall;", "        type            wall;\n        extra { a 1; b (1 2 3); }", 1)
    (pm / "boundary").write_text(b)
    case.write_cache(root / "constant" / "WENOBase2")
    back = wx.CachedCase.from_case(root, 2)
    assert np.array_equal(back.owner, case.owner) and np.array_equal(back.neighbour, case.neighbour)
    assert back.patches == case.patches
