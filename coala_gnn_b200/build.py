"""In-tree build of libcoala_b200.so (sm_100a only). `python -m coala_gnn_b200.build [--force]`."""
import glob
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(OUT_DIR, "libcoala_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
HOST_CXX = "/usr/bin/g++"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O2,-Wall,-Wno-unused-function", "-ccbin", HOST_CXX,
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cpp")))


def _deps():
    return sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
        glob.glob(os.path.join(HERE, "..", "include", "*.h"))


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(f) > t for f in _deps())


def _compile(src, verbose):
    obj = os.path.join(OUT_DIR, "obj", os.path.basename(src) + ".o")
    cmd = [NVCC] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-x", "cu", "-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed on %s:\n%s%s" % (src, r.stdout, r.stderr))
    return obj, r.stderr


def build(force=False, verbose=False):
    """Compile + link in-tree.  Several ranks may import the package at once (torchrun): the whole check-and-build runs
    under an exclusive file lock, and the library is linked under a temporary name and renamed into place, so no
    process ever maps a half-written file."""
    import fcntl
    os.makedirs(os.path.join(OUT_DIR, "obj"), exist_ok=True)
    with open(os.path.join(OUT_DIR, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not needs_build():
                return LIB
            if not os.path.exists(NVCC):
                raise RuntimeError("nvcc not found at %s and no prebuilt %s" % (NVCC, LIB))
            with ThreadPoolExecutor(max_workers=8) as ex:
                results = list(ex.map(lambda s: _compile(s, verbose), sources()))
            if verbose:
                for _, log in results:
                    sys.stderr.write(log)
            objs = [o for o, _ in results]
            tmp = LIB + ".tmp.%d" % os.getpid()
            cmd = [NVCC, "-shared", "-ccbin", HOST_CXX, "-o", tmp] + objs + ["-lcudart", "-lrt", "-lpthread", "-ldl"]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
            os.replace(tmp, LIB)
            return LIB
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


PYBIND_SRC = os.path.join(HERE, "csrc_pybind", "COALA_GNN_Pybind.cpp")


def pybind_path():
    import sysconfig
    return os.path.join(HERE, "..", "COALA_GNN_Pybind", "_native" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_pybind(force=False):
    """The compiled pybind11 binding of the C ABI (COALA_GNN_Pybind/_native*.so): plain g++, linked against libcoala_b200.so."""
    import fcntl
    import sysconfig
    out = os.path.abspath(pybind_path())
    build()
    with open(os.path.join(OUT_DIR, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            deps = [PYBIND_SRC, LIB] + glob.glob(os.path.join(HERE, "..", "include", "*.h"))
            if not force and os.path.exists(out) and all(os.path.getmtime(f) <= os.path.getmtime(out) for f in deps):
                return out
            if not os.path.exists(HOST_CXX):
                if os.path.exists(out):
                    return out
                raise RuntimeError("g++ not found and no prebuilt %s" % out)
            import pybind11
            cuda_home = os.path.dirname(os.path.dirname(NVCC))
            tmp = out + ".tmp.%d" % os.getpid()
            cmd = [HOST_CXX, "-O2", "-shared", "-fPIC", "-std=c++17", "-I" + sysconfig.get_paths()["include"], "-I" + pybind11.get_include(),
                   "-I" + os.path.join(HERE, "..", "include"), "-I" + os.path.join(cuda_home, "include"), PYBIND_SRC, "-o", tmp,
                   "-L" + OUT_DIR, "-lcoala_b200", "-L" + os.path.join(cuda_home, "lib64"), "-lcudart", "-Wl,-rpath,$ORIGIN/../coala_gnn_b200/lib"]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("pybind11 module build failed:\n" + r.stdout + r.stderr)
            os.replace(tmp, out)
            return out
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_pybind(force="--force" in sys.argv))
