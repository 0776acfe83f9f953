"""Build the in-tree native libraries.

  libwenoext_b200.so   product: device plan + kernels (CUDA, sm_100a), C ABI (include/wenoext_b200.h)
  libwenoext_setup.so  host-side producers of plan descriptors (plain C++, no CUDA): box stand-in of the
                       WENOBase constructor, reader/writer of constant/WENOBase<order>/, polyMesh reader
                       (include/wenoext_setup.h)
  oracle/libweno_oracle.so, oracle/_ref/libweno_oracle_blaze.so   checker (tests only)

nvcc cross-compiles sm_100a without a GPU. Run as `python -m wenoext_b200.build`.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "wenoext_b200", "csrc")
LIB = os.path.join(ROOT, "wenoext_b200", "libwenoext_b200.so")
SETUP_LIB = os.path.join(ROOT, "wenoext_b200", "libwenoext_setup.so")

CU_SOURCES = ["kernels.cu", "plan.cu", "pinv.cu", "noalloc.cu", "ls_rows.cu"]
# ls_rows.cu repeats the host setup arithmetic operation for operation: no FMA contraction, so that its
# rows are bitwise the host's (the host library is built for baseline x86-64, which has no FMA)
CU_EXTRA_FLAGS = {"ls_rows.cu": ["--fmad=false"]}
CPP_SOURCES = [
    "capi_setup.cpp",
    "setup/polymesh.cpp",
    "setup/geometry_weno.cpp",
    "setup/svd.cpp",
    "setup/weno_setup.cpp",
    "setup/box_fast.cpp",
    "setup/wenobase_cache.cpp",
]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _all_deps():
    deps = []
    for base, _, files in os.walk(CSRC):
        deps += [os.path.join(base, f) for f in files if f.endswith((".cu", ".cpp", ".h"))]
    deps.append(os.path.join(ROOT, "include", "wenoext_b200.h"))
    return deps


def _headers():
    hs = [os.path.join(ROOT, "include", "wenoext_b200.h")]
    for base, _, files in os.walk(CSRC):
        hs += [os.path.join(base, f) for f in files if f.endswith(".h")]
    return hs


def build_product(force=False, verbose=False):
    """Per-object incremental: an object is rebuilt when its source or a header it can see
    changed (setup/*.cpp see only setup/*.h and the public header)."""
    objdir = os.path.join(ROOT, "build", "obj")
    os.makedirs(objdir, exist_ok=True)
    objs, rebuilt = [], False
    common = ["-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-I", os.path.join(ROOT, "include")]
    kdebug = bool(os.environ.get("WENOEXT_KDEBUG"))
    if kdebug:  # developer aid: device-side bounds checks
        common.append("-DWENOEXT_KDEBUG")
    common += os.environ.get("WENOEXT_CFLAGS", "").split()  # developer aid: kernel variants
    stamp = os.path.join(objdir, "kdebug.stamp")
    if kdebug != os.path.exists(stamp):
        force = True
        (open(stamp, "w").close() if kdebug else os.remove(stamp))
    for src in CU_SOURCES:
        path = os.path.join(CSRC, src)
        obj = os.path.join(objdir, src.replace("/", "_") + ".o")
        hs = _headers()
        if src.startswith("setup/"):
            hs = [h for h in hs if os.sep + "setup" + os.sep in h or h.endswith("wenoext_b200.h")]
        if force or _newer(obj, [path] + hs):
            cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo"] + common
            if src.endswith(".cu"):
                cmd += ["-Xptxas", "-v"] if verbose else []
            cmd += CU_EXTRA_FLAGS.get(src, [])
            cmd += ["-c", path, "-o", obj]
            subprocess.run(cmd, check=True)
            rebuilt = True
        objs.append(obj)
    if rebuilt or not os.path.exists(LIB):
        cmd = ["nvcc", "-shared", "-o", LIB] + objs + ["-lpthread", "-ldl", "-cudart", "static"]
        subprocess.run(cmd, check=True)
    return LIB


def build_setup(force=False):
    """libwenoext_setup.so with the host compiler (no CUDA headers, no CUDA runtime)."""
    objdir = os.path.join(ROOT, "build", "obj_setup")
    os.makedirs(objdir, exist_ok=True)
    hs = [h for h in _headers() if os.sep + "setup" + os.sep in h or h.endswith("wenoext_b200.h")]
    hs.append(os.path.join(ROOT, "include", "wenoext_setup.h"))
    objs, rebuilt = [], False
    for src in CPP_SOURCES:
        path = os.path.join(CSRC, src)
        obj = os.path.join(objdir, src.replace("/", "_") + ".o")
        if force or _newer(obj, [path] + hs):
            subprocess.run(["g++", "-O3", "-std=c++17", "-fPIC", "-I", os.path.join(ROOT, "include"),
                            "-I", CSRC, "-c", path, "-o", obj], check=True)
            rebuilt = True
        objs.append(obj)
    if rebuilt or not os.path.exists(SETUP_LIB):
        subprocess.run(["g++", "-shared", "-o", SETUP_LIB] + objs + ["-lpthread", "-ldl"], check=True)
    return SETUP_LIB


def build_oracle(force=False):
    odir = os.path.join(ROOT, "oracle")
    subprocess.run(["make", "-C", odir] + (["-B"] if force else []), check=True,
                   stdout=subprocess.DEVNULL)
    return os.path.join(odir, "libweno_oracle.so")


def build_shim_check():
    """shim/_build/libweno_shim_check.so: the OpenFOAM-side shim compiled against the OpenFOAM stand-in
    (needs /root/reference for List3D.H / Blaze; elsewhere the prebuilt file is kept)."""
    subprocess.run(["make", "-C", os.path.join(ROOT, "shim"), "check"], check=True, stdout=subprocess.DEVNULL)
    return os.path.join(ROOT, "shim", "_build", "libweno_shim_check.so")


def build_test_hooks(force=False):
    """tests/native/libsetup_hooks.so: test-only C hooks into the host setup helpers."""
    tdir = os.path.join(ROOT, "tests", "native")
    lib = os.path.join(tdir, "libsetup_hooks.so")
    srcs = [os.path.join(tdir, "setup_hooks.cpp")] + [
        os.path.join(CSRC, "setup", f) for f in ("geometry_weno.cpp", "polymesh.cpp")]
    if not force and not _newer(lib, srcs + _all_deps()):
        return lib
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I", CSRC, "-o", lib] + srcs,
                   check=True)
    return lib


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_product(force=force, verbose="-v" in sys.argv))
    print(build_setup(force=force))
    print(build_oracle(force=force))
    print(build_test_hooks(force=force))
    print(build_shim_check())
