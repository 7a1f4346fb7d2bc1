"""In-tree build of libmcrt_b200.so (CUDA kernels + C ABI) for sm_100a with nvcc.

The shared library is git-ignored but travels with the repo snapshot to the GPU box.  There is no
JIT and no fallback: if the library is missing or stale and nvcc is unavailable, loading fails.
"""
import hashlib
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(_HERE), "include")
LIB_PATH = os.path.join(_HERE, "libmcrt_b200.so")
STAMP_PATH = os.path.join(_HERE, ".libmcrt_b200.stamp")

CUDA_SOURCES = ["rrb_api.cu", "rrb_kernels.cu", "ptb_api.cu", "pt_kernels.cu", "pt_stage_kernels.cu", "bdpt_kernels.cu",
                "postfx_kernels.cu", "refit_kernels.cu", "lbvh_build.cu", "rrb_noalloc.cu"]
CXX_SOURCES = ["bvh8_build.cpp"]

# fp32 modes of the SHADING arithmetic.  The traversal, the triangle test and the hit
# barycentrics are written with explicit __fmul_rn/__fadd_rn/__fdiv_rn and are identical in every mode.
#   strict  no FMA contraction, IEEE division and sqrt: operation-for-operation the oracle's arithmetic
#   fma     contraction allowed, IEEE division and sqrt
#   fast    contraction allowed, approximate division / sqrt (2 ulp)
FP_MODES = {
    "strict": ["--fmad=false", "--prec-div=true", "--prec-sqrt=true", "--ftz=false"],
    "fma": ["--fmad=true", "--prec-div=true", "--prec-sqrt=true", "--ftz=false"],
    "fast": ["--fmad=true", "--prec-div=false", "--prec-sqrt=false", "--ftz=false"],
}
FP_MODE = os.environ.get("MCRT_FP_MODE", "strict")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
] + FP_MODES[FP_MODE] + [
    "-Xcompiler", "-fPIC,-O3,-ffp-contract=off,-fno-fast-math", "-Xptxas", "-v",
] + os.environ.get("MCRT_NVCC_DEFS", "").replace(",", " ").split()


def _nccl():
    """(include dir, lib dir) of NCCL: the copy PyTorch ships (the one a torch.distributed process has already
    loaded), else the system's.  The library is linked by soname with an rpath, so whichever process loads
    libmcrt_b200.so first -- torch or a plain ctypes host -- resolves to the same libnccl.so.2."""
    import importlib.util
    spec = importlib.util.find_spec("nvidia.nccl")
    if spec and spec.submodule_search_locations:
        base = list(spec.submodule_search_locations)[0]
        if os.path.exists(os.path.join(base, "lib", "libnccl.so.2")):
            return os.path.join(base, "include"), os.path.join(base, "lib")
    return "/usr/include", "/usr/lib/x86_64-linux-gnu"


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _sources():
    srcs = [os.path.join(CSRC, s) for s in CUDA_SOURCES + CXX_SOURCES]
    return [s for s in srcs if os.path.exists(s)]


def _fingerprint():
    h = hashlib.sha256()
    files = sorted(os.listdir(CSRC)) + [os.path.join(INCLUDE, f) for f in sorted(os.listdir(INCLUDE))]
    for f in files:
        p = f if os.path.isabs(f) else os.path.join(CSRC, f)
        if os.path.isfile(p):
            h.update(os.path.basename(p).encode())      # names, not absolute paths: the same on every box
            with open(p, "rb") as fh:
                h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_current() -> bool:
    if not (os.path.exists(LIB_PATH) and os.path.exists(STAMP_PATH)):
        return False
    with open(STAMP_PATH) as f:
        return f.read().strip() == _fingerprint()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source of the package for sm_100a into libmcrt_b200.so."""
    if not force and is_current():
        return LIB_PATH
    build_dir = os.path.join(_HERE, "build")
    os.makedirs(build_dir, exist_ok=True)
    objs, procs = [], []
    for src in _sources():
        obj = os.path.join(build_dir, os.path.basename(src) + ".o")
        cmd = [_nvcc()] + NVCC_FLAGS + ["-I", INCLUDE, "-I", CSRC, "-I", _nccl()[0], "-c", src, "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    log = []
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        log.append("== %s\n%s" % (os.path.basename(src), out))
        failed = failed or p.returncode != 0
    with open(os.path.join(build_dir, "build.log"), "w") as f:
        f.write("\n".join(log))
    if verbose or failed:
        print("\n".join(log))
    if failed:
        raise RuntimeError("nvcc failed; see %s" % os.path.join(build_dir, "build.log"))
    nccl_lib = _nccl()[1]
    link = [_nvcc(), "-shared", "-o", LIB_PATH] + objs + ["-lcudart", "-lpthread", "-L" + nccl_lib, "-l:libnccl.so.2",
                                                           "-Xlinker", "-rpath=" + nccl_lib]
    subprocess.check_call(link)
    with open(STAMP_PATH, "w") as f:
        f.write(_fingerprint())
    return LIB_PATH


if __name__ == "__main__":
    print(build(force=True, verbose=True))
