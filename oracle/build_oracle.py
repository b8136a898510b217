"""Build recipes for the parity oracle.  TEST INFRASTRUCTURE ONLY (never imported by the product).

build_c_oracle()      gcc  oracle/dh_oracle.c                      -> oracle/libdh_oracle.so
build_reference_gpu() nvcc/g++ on the UNMODIFIED reference sources, compiled from where they lie under
                      /root/reference (nothing is copied into this repo)     -> oracle/_ref/dh_ref_{fast,ieee}*.so
        sources: packages/draw_heatmap/accvlab/draw_heatmap/csrc/draw_heatmap.cpp
                 packages/draw_heatmap/accvlab/draw_heatmap/csrc/draw_heatmap_cuda.cu
                 (+ include/draw_heatmap_cuda_kernel.cuh)
        flags:   those of build_config/accvlab_build_config/helpers/build_utils.py:316-382 with the default
                 config (O3, c++17, USE_FAST_MATH=1 -> --use_fast_math) for `dh_ref_fast`, and USE_FAST_MATH=0
                 (--ftz=false) for `dh_ref_ieee`; arch fixed to sm_100a (the reference auto-detects the GPU).
        The reference's own setup.py / build system is NOT run.

build_reference_dali() g++ on the UNMODIFIED reference DALI operator source
                      /root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.cc (+ .h), compiled from
                      where it lies against oracle/dali_stub (a minimal stand-in for the DALI headers it includes:
                      DALI itself is not installed) plus oracle/dg_ref_driver.cc       -> oracle/_ref/libdg_ref.so
                      The reference's CMake build is NOT run.

oracle/_ref/ is git-ignored but travels to the GPU box; /root/reference does not exist there, so
build_reference_gpu() is a no-op when the sources are absent.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DH = "/root/reference/packages/draw_heatmap/accvlab/draw_heatmap"
REF_OUT = os.path.join(HERE, "_ref")


def _run(cmd, **kw):
    print("[oracle] " + " ".join(cmd), flush=True)
    subprocess.run(cmd, check=True, **kw)


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def build_c_oracle(force: bool = False) -> str:
    src = os.path.join(HERE, "dh_oracle.c")
    out = os.path.join(HERE, "libdh_oracle.so")
    if force or not _newer(out, [src]):
        _run(["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fvisibility=hidden", "-shared", "-fPIC",
              "-o", out, src, "-lm", "-lpthread"])
    return out


def reference_gpu_paths():
    return {name: os.path.join(REF_OUT, name + ".so") for name in ("dh_ref_fast", "dh_ref_ieee")}


def build_reference_gpu(force: bool = False) -> dict:
    """Compile the reference CUDA extension twice (fast-math default, and IEEE). Returns {name: path or None}."""
    paths = reference_gpu_paths()
    cpp = os.path.join(REF_DH, "csrc", "draw_heatmap.cpp")
    cu = os.path.join(REF_DH, "csrc", "draw_heatmap_cuda.cu")
    inc = os.path.join(REF_DH, "include")
    if not (os.path.exists(cpp) and os.path.exists(cu)):
        return {k: (v if os.path.exists(v) else None) for k, v in paths.items()}
    import torch
    from torch.utils import cpp_extension as ce

    os.makedirs(REF_OUT, exist_ok=True)
    tinc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", f"-I{inc}"]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"
    for name, math_flag in (("dh_ref_fast", "--use_fast_math"), ("dh_ref_ieee", "--ftz=false")):
        out = paths[name]
        if not force and _newer(out, [cpp, cu]):
            continue
        defs = [f"-DTORCH_EXTENSION_NAME={name}", "-DTORCH_API_INCLUDE_EXTENSION_H", abi]
        obj_cu = os.path.join(REF_OUT, name + "_cuda.o")
        obj_cpp = os.path.join(REF_OUT, name + "_cpp.o")
        _run(["nvcc", "-c", cu, "-o", obj_cu, "-gencode=arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", math_flag,
              "--generate-line-info", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"] + defs + tinc)
        _run(["g++", "-c", cpp, "-o", obj_cpp, "-std=c++17", "-O3", "-fPIC", "-pthread"] + defs + tinc
             + ["-I/usr/local/cuda/include"])
        _run(["g++", "-shared", "-o", out, obj_cu, obj_cpp, f"-L{tlib}", "-L/usr/local/cuda/lib64", "-lc10", "-lc10_cuda",
              "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart", f"-Wl,-rpath,{tlib}"])
        for o in (obj_cu, obj_cpp):
            os.remove(o)
    return {k: (v if os.path.exists(v) else None) for k, v in paths.items()}


REF_DPF = "/root/reference/packages/dali_pipeline_framework/ext_impl"


def build_reference_dali(force: bool = False):
    """Compile the reference's draw_gaussians operator source against the DALI stub. Returns the path or None."""
    out = os.path.join(REF_OUT, "libdg_ref.so")
    cc, hh = os.path.join(REF_DPF, "DrawGaussians.cc"), os.path.join(REF_DPF, "DrawGaussians.h")
    drv = os.path.join(HERE, "dg_ref_driver.cc")
    if not (os.path.exists(cc) and os.path.exists(hh)):
        return out if os.path.exists(out) else None
    if force or not _newer(out, [cc, hh, drv]):
        os.makedirs(REF_OUT, exist_ok=True)
        _run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-fvisibility=hidden",
              f"-I{os.path.join(HERE, 'dali_stub')}", f"-I{REF_DPF}", drv, cc, "-o", out])
    return out


if __name__ == "__main__":
    build_c_oracle(force="--force" in sys.argv)
    if "--ref" in sys.argv:
        print(build_reference_gpu(force="--force" in sys.argv))
        print(build_reference_dali(force="--force" in sys.argv))
