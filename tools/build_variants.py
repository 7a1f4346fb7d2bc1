"""Build experiment variants of libmcrt_b200.so HERE (nvcc cross-compiles without a GPU) so that the GPU box only
has to run them:   python tools/build_variants.py name1="-DA=1 -DB=2" name2="..."
Each lands in monte-carlo-raytracer_b200/build/variants/libmcrt_<name>.so and is selected with
MCRT_LIB_VARIANT=<name> (rr.lib_path).  Only the sources listed in --only are recompiled with the knobs when given."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200  # noqa: E402
from mcrt_b200 import build as b  # noqa: E402


def main():
    out_dir = os.path.join(os.path.dirname(b.LIB_PATH), "build", "variants")
    os.makedirs(out_dir, exist_ok=True)
    jobs = []
    for arg in sys.argv[1:]:
        name, defs = arg.split("=", 1)
        odir = os.path.join(out_dir, name)
        os.makedirs(odir, exist_ok=True)
        objs, procs = [], []
        for src in b._sources():
            obj = os.path.join(odir, os.path.basename(src) + ".o")
            cmd = [b._nvcc()] + b.NVCC_FLAGS + defs.split() + ["-I", b.INCLUDE, "-I", b.CSRC, "-I", b._nccl()[0], "-c", src, "-o", obj]
            procs.append(subprocess.Popen(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True))
            objs.append(obj)
        jobs.append((name, objs, procs))
    for name, objs, procs in jobs:
        for p in procs:
            _, err = p.communicate()
            if p.returncode != 0:
                print(err)
                raise SystemExit("variant %s failed" % name)
        lib = os.path.join(out_dir, "libmcrt_%s.so" % name)
        nccl_lib = b._nccl()[1]
        subprocess.check_call([b._nvcc(), "-shared", "-o", lib] + objs + ["-lcudart", "-lpthread", "-L" + nccl_lib, "-l:libnccl.so.2",
                                                                          "-Xlinker", "-rpath=" + nccl_lib])
        print("built", lib)


if __name__ == "__main__":
    main()
