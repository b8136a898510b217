"""In-tree build of the product: libdh_b200.so (CUDA kernels + C ABI) and draw_heatmap_ext (torch binding).

    python accv-lab_b200/build.py [--force]

Both artefacts land in accv-lab_b200/accvlab/draw_heatmap/ (git-ignored, shipped to the GPU box by gpurun).
nvcc cross-compiles for sm_100a without a GPU; the binding is host-only C++ (g++), linked against libdh_b200.so
with an $ORIGIN rpath. No setuptools / JIT cache is involved so the build is reproducible and seconds-to-a-minute.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
PKG = os.path.join(HERE, "accvlab", "draw_heatmap")
INCLUDE = os.path.join(ROOT, "include")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]

LIB_PATH = os.path.join(PKG, "libdh_b200.so")
EXT_PATH = os.path.join(PKG, "draw_heatmap_ext" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))
SHIM_PATH = os.path.join(PKG, "libdh_ref_shim.so")


def _run(cmd):
    print("[build] " + " ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)


def _fresh(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def cuda_sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _includes(src, seen=None):
    """Quoted includes of a source file, transitively (csrc headers + the public header): its rebuild dependencies."""
    import re

    seen = [] if seen is None else seen
    if src in seen:
        return seen
    seen.append(src)
    for name in re.findall(r'^\s*#\s*include\s+"([^"]+)"', open(src).read(), flags=re.M):
        for d in (CSRC, INCLUDE):
            q = os.path.join(d, name)
            if os.path.exists(q):
                _includes(q, seen)
    return seen


def build_cuda_lib(force=False, verbose_ptxas=False):
    """One object per .cu (compiled in parallel, rebuilt only when the file or a header it includes changed), then one
    link. Objects live in accv-lab_b200/_obj/ (git- and gpurun-ignored)."""
    from concurrent.futures import ThreadPoolExecutor

    srcs = cuda_sources()
    objdir = os.path.join(HERE, "_obj")
    os.makedirs(objdir, exist_ok=True)
    objs, todo = [], []
    for s in srcs:
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        objs.append(o)
        if force or not _fresh(o, _includes(s)):
            todo.append(["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose_ptxas else []) + [f"-I{INCLUDE}", "-c", s, "-o", o])
    if todo:
        with ThreadPoolExecutor(max_workers=min(len(todo), os.cpu_count() or 1)) as ex:
            list(ex.map(_run, todo))
    if todo or not _fresh(LIB_PATH, objs):
        _run(["nvcc"] + NVCC_FLAGS + ["-shared"] + objs + ["-o", LIB_PATH, "-lcudart"])
    return LIB_PATH


def build_torch_ext(force=False):
    src = os.path.join(CSRC, "dh_torch_binding.cpp")
    deps = [src, os.path.join(INCLUDE, "dh_b200.h")]
    if force or not _fresh(EXT_PATH, deps):
        import torch
        from torch.utils import cpp_extension as ce

        tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
        inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", f"-I{INCLUDE}",
                                                        "-I/usr/local/cuda/include"]
        cmd = ["g++", "-shared", "-fPIC", "-O2", "-std=c++17", "-fvisibility=hidden", src, "-o", EXT_PATH,
               "-DTORCH_EXTENSION_NAME=draw_heatmap_ext", "-DTORCH_API_INCLUDE_EXTENSION_H",
               f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"] + inc + [
            f"-L{tlib}", f"-L{PKG}", "-L/usr/local/cuda/lib64", "-ldh_b200", "-lc10", "-lc10_cuda", "-ltorch_cpu",
            "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}"]
        _run(cmd)
    return EXT_PATH


def build_ref_shim(force=False):
    """libdh_ref_shim.so: the reference's two C++-linkage launchers (draw_heatmap_cuda.cu:29-31,43-46) over the C ABI."""
    src = os.path.join(CSRC, "dh_ref_shim.cpp")
    deps = [src, os.path.join(INCLUDE, "dh_b200.h"), LIB_PATH]
    if force or not _fresh(SHIM_PATH, deps):
        _run(["g++", "-shared", "-fPIC", "-O2", "-std=c++17", "-fvisibility=hidden", src, "-o", SHIM_PATH, f"-I{INCLUDE}",
              "-I/usr/local/cuda/include", f"-L{PKG}", "-L/usr/local/cuda/lib64", "-ldh_b200", "-lcudart",
              "-Wl,-rpath,$ORIGIN"])
    return SHIM_PATH


def build_all(force=False):
    build_cuda_lib(force)
    build_torch_ext(force)
    build_ref_shim(force)
    return LIB_PATH, EXT_PATH, SHIM_PATH


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))
