#!/usr/bin/env python3
"""cl2cpp.py -- TEST INFRASTRUCTURE (oracle/_ref build step). Not part of the product path.

Reads the reference's OpenCL-C sources where they lie under /root/reference and writes
C++-compilable copies to oracle/_ref/gen/ (git-ignored; never committed).  The rewrite is
purely mechanical -- no arithmetic, control flow or constant is touched:

  1. OpenCL vector literals   `(float3)(a,b,c)` -> `float3{a,b,c}`   (not valid C++)
     for float2/3/4, int2/3, uint2.  Braced initialisation is used because it evaluates the
     components left to right, as OpenCL compilers (clang) do; a function call would let g++
     evaluate them right to left, which swaps the two draws of `(float2)(randFloat(..),
     randFloat(..))` in samplers.cl:getSample2D.
  2. `#ifdef __cplusplus` in kernel_data.h -> `#ifdef MCRT_REF_HOST_BRANCH` so that the
     OpenCL branch of that shared header (the device-side struct layouts) is the one compiled.
  3. `#include <../RadeonRays/src/kernels/CL/common.cl>` -> `#include "common.cl"`.
  4. samplers.cl's compile-time sampler choice `#define RT_SAMPLER RT_SAMPLER_RANDOM` is wrapped in
     `#ifndef RT_SAMPLER` so the Sobol variant of the same source can be built with -DRT_SAMPLER=0
     (the reference selects it by editing that line).

Everything else (swizzles, address-space qualifiers, builtins) is supplied by clshim.hpp.
"""
import os
import re
import sys

REF = os.environ.get("MCRT_REFERENCE", "/root/reference")
APP_KERNELS = ["PathTracing.cl", "BDPT.cl", "bxdfs.cl", "materials.cl", "lights.cl", "geometry.cl",
               "textures.cl", "samplers.cl", "rng.cl", "cameras.cl", "math.cl", "matrix.cl",
               "image_samplers.cl", "kernel_data.h", "reconstruction.cl", "filters.cl", "ToneMapping.cl", "Denoise.cl",
               "colors.cl"]
RR_KERNELS = ["common.cl", "intersect_bvh2_lds.cl"]

VEC_LITERAL = re.compile(r"\((float[234]|int[23]|uint2)\)\s*\(")


def rewrite_literals(text: str) -> str:
    while True:
        m = VEC_LITERAL.search(text)
        if not m:
            return text
        depth, j = 1, m.end()
        while depth:
            c = text[j]
            depth += (c == "(") - (c == ")")
            j += 1
        text = text[:m.start()] + m.group(1) + "{" + text[m.end():j - 1] + "}" + text[j:]


def rewrite(text: str) -> str:
    text = rewrite_literals(text)
    text = text.replace("#ifdef __cplusplus", "#ifdef MCRT_REF_HOST_BRANCH")
    text = text.replace("#include <../RadeonRays/src/kernels/CL/common.cl>", '#include "common.cl"')
    text = re.sub(r"^#define RT_SAMPLER RT_SAMPLER_RANDOM\s*$",
                  "#ifndef RT_SAMPLER\n#define RT_SAMPLER RT_SAMPLER_RANDOM\n#endif", text, flags=re.M)
    return text


def main(out_dir: str) -> int:
    app_dir = os.path.join(REF, "assets", "kernels")
    rr_dir = os.path.join(REF, "third_party", "RadeonRays", "RadeonRays", "src", "kernels", "CL")
    os.makedirs(os.path.join(out_dir, "app"), exist_ok=True)
    os.makedirs(os.path.join(out_dir, "rr"), exist_ok=True)
    for name in APP_KERNELS:
        with open(os.path.join(app_dir, name), "r", encoding="latin-1") as f:
            src = f.read()
        with open(os.path.join(out_dir, "app", name), "w", encoding="latin-1") as f:
            f.write(rewrite(src))
    for name in RR_KERNELS:
        with open(os.path.join(rr_dir, name), "r", encoding="latin-1") as f:
            src = f.read()
        with open(os.path.join(out_dir, "rr", name), "w", encoding="latin-1") as f:
            f.write(rewrite(src))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1]))
