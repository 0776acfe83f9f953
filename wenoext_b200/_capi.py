"""ctypes binding of include/wenoext_b200.h (the C-ABI drop-in boundary, libwenoext_b200.so) and of
include/wenoext_setup.h (host-side descriptor producers, libwenoext_setup.so).

No torch types cross this boundary: plain pointers and sizes only. The libraries must exist in-tree
(built by `python -m wenoext_b200.build`); there is no fallback of any kind when one is missing.
Each library is loaded on first use of one of its symbols, so CPU-only users of a descriptor (the
oracle, the reference build) never map the GPU library.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# WENOEXT_LIB: developer aid, an alternative build of the same library (kernel variants)
LIB_PATH = os.environ.get("WENOEXT_LIB") or os.path.join(_HERE, "libwenoext_b200.so")
SETUP_LIB_PATH = os.path.join(_HERE, "libwenoext_setup.so")

# constants of include/wenoext_b200.h
CELL_LOCAL, CELL_DELETED, CELL_EMPTY = -1, -4, -10
PATCH_GENERIC, PATCH_PROCESSOR, PATCH_CYCLIC, PATCH_EMPTY = 0, 1, 2, 3
SCHEME_UPWINDFIT, SCHEME_UPWINDFIT01 = 0, 1
MEM_HOST, MEM_DEVICE = 0, 1

i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
f64p = C.POINTER(C.c_double)


class Patch(C.Structure):
    _fields_ = [("type", C.c_int32), ("start", C.c_int32), ("size", C.c_int32), ("nbr", C.c_int32)]


class PlanDesc(C.Structure):
    _fields_ = [
        ("polOrder", C.c_int32), ("nDvt", C.c_int32), ("dim", C.c_int32 * 3),
        ("nCells", C.c_int32), ("nFaces", C.c_int32), ("nInternalFaces", C.c_int32),
        ("owner", i32p), ("neighbour", i32p), ("cellFaceStart", i32p), ("cellFaces", i32p),
        ("nPatches", C.c_int32), ("patches", C.POINTER(Patch)),
        ("nStencils", C.c_int64), ("cellStencilStart", i64p), ("stencilEntryStart", i64p),
        ("stencilsID", i32p), ("cellToProcMap", i32p),
        ("stencilMatrix", i32p), ("nUniqueMatrices", C.c_int32), ("matrixStart", i64p),
        ("matrixCols", i32p), ("matrixData", f64p),
        ("B", f64p), ("intBasTrans", f64p), ("refFacAr", f64p),
        ("nProcs", C.c_int32), ("myProc", C.c_int32),
        ("sendProcList", i32p), ("receiveProcList", i32p), ("receiveHaloSize", i32p),
        ("sendHaloStart", i64p), ("sendHaloCellID", i32p),
        ("patchNbrCell", i32p),
    ]


class Patch64(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("type", "start", "size", "nbr")]


class PlanDesc64(C.Structure):
    """wenoext_plan_desc64: the descriptor of a WM_LABEL_SIZE=64 build (labels are int64)"""
    _fields_ = [
        ("polOrder", C.c_int64), ("nDvt", C.c_int64), ("dim", C.c_int64 * 3),
        ("nCells", C.c_int64), ("nFaces", C.c_int64), ("nInternalFaces", C.c_int64),
        ("owner", i64p), ("neighbour", i64p), ("cellFaceStart", i64p), ("cellFaces", i64p),
        ("nPatches", C.c_int64), ("patches", C.POINTER(Patch64)),
        ("nStencils", C.c_int64), ("cellStencilStart", i64p), ("stencilEntryStart", i64p),
        ("stencilsID", i64p), ("cellToProcMap", i64p),
        ("stencilMatrix", i64p), ("nUniqueMatrices", C.c_int64), ("matrixStart", i64p),
        ("matrixCols", i64p), ("matrixData", f64p),
        ("B", f64p), ("intBasTrans", f64p), ("refFacAr", f64p),
        ("nProcs", C.c_int64), ("myProc", C.c_int64),
        ("sendProcList", i64p), ("receiveProcList", i64p), ("receiveHaloSize", i64p),
        ("sendHaloStart", i64p), ("sendHaloCellID", i64p),
        ("patchNbrCell", i64p),
    ]


class LsBatch(C.Structure):
    """wenoext_ls_batch: least-squares rows, matrices and pseudo-inverses in one device batch"""
    _fields_ = [
        ("polOrder", C.c_int32), ("nDvt", C.c_int32), ("dim", C.c_int32 * 3),
        ("nPoints", C.c_int32), ("nSrcCells", C.c_int32),
        ("points", f64p), ("triStart", i64p), ("triPoints", i32p), ("cellCentres", f64p), ("cellVolumes", f64p),
        ("nFrames", C.c_int32), ("JInv", f64p), ("refPoint", f64p), ("refDet", f64p), ("frameCell", i32p),
        ("volI", f64p),
        ("nPairs", C.c_int64), ("pairFrame", i32p), ("pairCell", i32p),
        ("nMat", C.c_int32), ("rows", i32p), ("start", i64p), ("rowPair", i64p),
    ]


class PlanInfo(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "nCells", "nFaces", "nInternalFaces", "nCoupledFaces", "nLiveStencils", "nHaloCells",
        "nInteriorCells", "recordBytes", "maxRecordBytes", "algBytesFixed", "algBytesPerComp",
        "algFlopsPerComp")]


class BoxSpec(C.Structure):
    _fields_ = [("n", C.c_int32 * 3), ("lo", C.c_double * 3), ("hi", C.c_double * 3),
                ("sideType", C.c_int32 * 6), ("tetSplit", C.c_int32), ("nProc", C.c_int32 * 3)]


class SetupParams(C.Structure):
    _fields_ = [("polOrder", C.c_int32), ("extendRatio", C.c_double),
                ("bestConditioned", C.c_int32), ("maxCondition", C.c_double),
                ("checkCondition", C.c_int32), ("nThreads", C.c_int32),
                ("useClassCache", C.c_int32), ("gpuPinv", C.c_int32), ("device", C.c_int32)]


class MeshDesc(C.Structure):
    _fields_ = [("nCells", C.c_int32), ("nFaces", C.c_int32), ("nInternalFaces", C.c_int32),
                ("owner", i32p), ("neighbour", i32p), ("cellFaceStart", i32p), ("cellFaces", i32p),
                ("nPatches", C.c_int32), ("patches", C.POINTER(Patch)), ("patchNbrCell", i32p)]


# every symbol include/wenoext_b200.h declares: name -> (restype, argtypes)
vp = C.c_void_p
SYMBOLS = {
    "wenoext_last_error": (C.c_char_p, []),
    "wenoext_device_count": (C.c_int, []),
    "wenoext_plan_create": (C.c_int, [C.POINTER(PlanDesc), C.c_int, C.POINTER(vp)]),
    "wenoext_plan_create_labels64": (C.c_int, [C.POINTER(PlanDesc64), C.c_int, C.POINTER(vp)]),
    "wenoext_plan_destroy": (C.c_int, [vp]),
    "wenoext_plan_get_info": (C.c_int, [vp, C.POINTER(PlanInfo)]),
    "wenoext_set_params": (C.c_int, [vp, C.c_double, C.c_double, C.c_double]),
    "wenoext_set_fit01_bug_compat": (C.c_int, [vp, C.c_int]),
    "wenoext_nccl_unique_id": (C.c_int, [C.c_char_p]),
    "wenoext_attach_nccl": (C.c_int, [vp, C.c_char_p, C.c_int, C.c_int]),
    "wenoext_correction": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, C.c_int, vp]),
    "wenoext_interpolate": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_int, vp]),
    "wenoext_div": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, C.c_int, vp]),
    "wenoext_host_alloc": (C.c_int, [C.c_uint64, C.POINTER(vp)]),
    "wenoext_host_free": (C.c_int, [vp]),
    "wenoext_set_face_flux": (C.c_int, [vp, vp, C.c_int, vp]),
    "wenoext_correction_noalloc": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp]),
    "wenoext_attach_local": (C.c_int, [C.POINTER(vp), C.c_int]),
    "wenoext_correction_begin": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp]),
    "wenoext_correction_exchange": (C.c_int, [vp]),
    "wenoext_correction_end": (C.c_int, [vp]),
    "wenoext_last_launch_count": (C.c_int, [vp]),
    "wenoext_enable_kernel_timing": (C.c_int, [vp, C.c_int]),
    "wenoext_kernel_time": (C.c_int, [vp, C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_int]),
    "wenoext_ls_pinv_batched": (C.c_int, [C.POINTER(LsBatch), C.c_double, f64p, i32p, f64p, i32p, f64p, f64p, C.c_int]),
    "wenoext_pinv_batched": (C.c_int, [C.c_int32, C.c_int32, i32p, i64p, f64p, C.c_double, f64p, i32p, f64p,
                                        i32p, C.c_int]),
}

# every symbol include/wenoext_setup.h declares
SETUP_SYMBOLS = {
    "wenoext_setup_last_error": (C.c_char_p, []),
    "wenoext_setup_box": (C.c_int, [C.POINTER(BoxSpec), C.POINTER(SetupParams), C.c_int32, C.POINTER(vp)]),
    "wenoext_setup_destroy": (C.c_int, [vp]),
    "wenoext_setup_desc": (C.c_int, [vp, C.POINTER(PlanDesc)]),
    "wenoext_setup_halo_request": (C.c_int, [vp, C.c_int32, C.POINTER(i32p), C.POINTER(C.c_int64)]),
    "wenoext_setup_set_send_halo": (C.c_int, [vp, C.c_int32, i32p, C.c_int64]),
    "wenoext_setup_geometry": (C.c_int, [vp, C.POINTER(f64p), C.POINTER(f64p)]),
    "wenoext_cache_write": (C.c_int, [vp, C.c_char_p, C.c_int, C.c_int]),
    "wenoext_cache_read": (C.c_int, [C.c_char_p, C.POINTER(MeshDesc), C.c_int32, C.c_int, C.POINTER(vp)]),
    "wenoext_case_read": (C.c_int, [C.c_char_p, C.c_int32, C.c_int32, C.c_int, C.POINTER(vp)]),
    "wenoext_setup_mesh": (C.c_int, [vp, C.POINTER(MeshDesc)]),
}

class WenoextError(RuntimeError):
    pass


class _Libs:
    """Attribute access resolves a symbol in the library that exports it (loaded on first use)."""

    def __init__(self):
        self._handles = {}

    def _load(self, path, table):
        if path not in self._handles:
            if not os.path.exists(path):
                raise WenoextError(
                    f"{path} is missing: build it with `python -m wenoext_b200.build` "
                    "(wenoext_b200 has no CPU or eager fallback)")
            L = C.CDLL(path)
            for name, (res, args) in table.items():
                fn = getattr(L, name)
                fn.restype = res
                fn.argtypes = args
            self._handles[path] = L
        return self._handles[path]

    def product(self):
        return self._load(LIB_PATH, SYMBOLS)

    def setup(self):
        return self._load(SETUP_LIB_PATH, SETUP_SYMBOLS)

    def __getattr__(self, name):
        if name in SETUP_SYMBOLS:
            return getattr(self.setup(), name)
        if name in SYMBOLS:
            return getattr(self.product(), name)
        raise AttributeError(name)


_libs = _Libs()


def lib():
    """The C ABI; raises when a library has not been built (no fallback)."""
    return _libs


def product_loaded():
    return LIB_PATH in _libs._handles


def check(rc):
    if rc != 0:
        raise WenoextError(f"wenoext error {rc}: {_libs.wenoext_last_error().decode()}")


def check_setup(rc):
    if rc != 0:
        raise WenoextError(f"wenoext setup error {rc}: {_libs.wenoext_setup_last_error().decode()}")
