"""Stage the reference's OWN callers of the hot path so they can be run, unmodified, against this package on the GPU box.
TEST INFRASTRUCTURE ONLY (never imported by the product).

    python oracle/stage_reference_tests.py [--force]

/root/reference does not exist on the GPU box, and nothing of it may be committed here, so this recipe (run in the build
container, where /root/reference is present; __graft_entry__.build() calls it) writes everything into the git-ignored,
gpurun-shipped directory oracle/_ref/:

  oracle/_ref/site/accvlab/batching_helpers/   the reference's batching_helpers package: its Python files verbatim
        (packages/batching_helpers/accvlab/batching_helpers/**.py) + its two torch extensions compiled from where
        they lie (cpp_impl/batched_indexing_access_{cpu,cuda}*.{cpp,cu}; flags of packages/batching_helpers/setup.py:65-80
        with build_config's defaults: -O3 -std=c++17 --use_fast_math, arch fixed to sm_100a) -> the REAL RaggedBatch
        (data_format/ragged_batch.py:31-171) is importable next to this repo's accvlab.draw_heatmap (`accvlab` is a
        PEP 420 namespace package in both trees).
  oracle/_ref/dh/tests/      packages/draw_heatmap/tests/*.py            (the three test files + their two helpers)
  oracle/_ref/dh/examples/   packages/draw_heatmap/examples/*.py
  oracle/_ref/dh/benchmark/  packages/draw_heatmap/benchmark/*.py
  oracle/_ref/dh/bench_cpp/  packages/draw_heatmap/benchmark_cpp/bench_*.cpp compiled from where they lie and linked
        against libdh_ref_shim.so (this repo's C++-linkage launchers) instead of the reference's draw_heatmap_cuda.cu.

tests/test_reference_suite.py (-m gpu) then runs those files in subprocesses with accv-lab_b200/ and
oracle/_ref/site/ on PYTHONPATH. The reference's setup.py / CMake builds are NOT run.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference/packages"
REF_BH = os.path.join(REF, "batching_helpers", "accvlab", "batching_helpers")
REF_DH = os.path.join(REF, "draw_heatmap")
OUT = os.path.join(HERE, "_ref")
SITE = os.path.join(OUT, "site")
BH_OUT = os.path.join(SITE, "accvlab", "batching_helpers")
DH_OUT = os.path.join(OUT, "dh")
PKG = os.path.join(ROOT, "accv-lab_b200", "accvlab", "draw_heatmap")


def _run(cmd, **kw):
    print("[stage] " + " ".join(cmd), flush=True)
    subprocess.run(cmd, check=True, **kw)


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def available() -> bool:
    return os.path.isdir(REF_BH) and os.path.isdir(REF_DH)


def staged() -> bool:
    """True when a previous run left a complete staging area (what the GPU box sees)."""
    ext = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    need = [os.path.join(BH_OUT, "__init__.py"), os.path.join(BH_OUT, "batched_indexing_access_cuda" + ext),
            os.path.join(BH_OUT, "batched_indexing_access_cpu" + ext),
            os.path.join(DH_OUT, "tests", "test_draw_heatmap_batched_classwise.py"),
            os.path.join(DH_OUT, "examples", "draw_heatmap_batched.py")]
    return all(os.path.exists(p) for p in need)


def _copy_py_tree(src, dst):
    for base, _dirs, files in os.walk(src):
        rel = os.path.relpath(base, src)
        if "__pycache__" in rel:
            continue
        for f in files:
            if f.endswith(".py"):
                os.makedirs(os.path.join(dst, rel), exist_ok=True)
                shutil.copyfile(os.path.join(base, f), os.path.join(dst, rel, f))


def build_batching_helpers(force: bool = False):
    import torch
    from torch.utils import cpp_extension as ce

    ext = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    _copy_py_tree(REF_BH, BH_OUT)
    cpp_impl = os.path.join(REF_BH, "cpp_impl")
    tinc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", f"-I{cpp_impl}",
                                                      "-I/usr/local/cuda/include"]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"
    link = [f"-L{tlib}", "-L/usr/local/cuda/lib64", "-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python",
            f"-Wl,-rpath,{tlib}"]
    tmp = os.path.join(OUT, "_obj")
    os.makedirs(tmp, exist_ok=True)
    for name, srcs in (("batched_indexing_access_cpu", ["batched_indexing_access_cpu.cpp",
                                                        "batched_indexing_access_cpu_impl.cpp"]),
                       ("batched_indexing_access_cuda", ["batched_indexing_access_cuda.cpp",
                                                         "batched_indexing_access_cuda_impl.cu"])):
        out = os.path.join(BH_OUT, name + ext)
        paths = [os.path.join(cpp_impl, s) for s in srcs]
        if not force and _newer(out, paths):
            continue
        defs = [f"-DTORCH_EXTENSION_NAME={name}", "-DTORCH_API_INCLUDE_EXTENSION_H", abi]
        objs = []
        for p in paths:
            o = os.path.join(tmp, os.path.basename(p) + ".o")
            if p.endswith(".cu"):
                _run(["nvcc", "-c", p, "-o", o, "-gencode=arch=compute_100a,code=sm_100a", "-std=c++17", "-O3",
                      "--use_fast_math", "--generate-line-info", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]
                     + defs + tinc)
            else:
                _run(["g++", "-c", p, "-o", o, "-std=c++17", "-O3", "-fPIC", "-pthread"] + defs + tinc)
            objs.append(o)
        cuda_link = ["-lc10_cuda", "-ltorch_cuda", "-lcudart"] if name.endswith("cuda") else []
        _run(["g++", "-shared", "-o", out] + objs + link + cuda_link)
    shutil.rmtree(tmp, ignore_errors=True)
    return BH_OUT


def stage_draw_heatmap_callers():
    for sub in ("tests", "examples", "benchmark"):
        src, dst = os.path.join(REF_DH, sub), os.path.join(DH_OUT, sub)
        os.makedirs(dst, exist_ok=True)
        for f in sorted(os.listdir(src)):
            if f.endswith(".py"):
                shutil.copyfile(os.path.join(src, f), os.path.join(dst, f))
    return DH_OUT


def build_cpp_benchmarks(force: bool = False):
    """The reference's C++ benchmarks (benchmark_cpp/bench_*.cpp:24-27 declare the C++-linkage launchers of
    draw_heatmap_cuda.cu:29-31,43-46) linked, unmodified, against libdh_ref_shim.so."""
    shim = os.path.join(PKG, "libdh_ref_shim.so")
    if not os.path.exists(shim):
        return None
    dst = os.path.join(DH_OUT, "bench_cpp")
    os.makedirs(dst, exist_ok=True)
    for name in ("bench_flattened", "bench_batched", "bench_batched_classwise"):
        src = os.path.join(REF_DH, "benchmark_cpp", name + ".cpp")
        out = os.path.join(dst, name)
        if force or not _newer(out, [src, shim]):
            _run(["g++", "-O3", "-std=c++17", src, "-o", out, "-I/usr/local/cuda/include", f"-L{PKG}",
                  "-L/usr/local/cuda/lib64", "-ldh_ref_shim", "-ldh_b200", "-lcudart", f"-Wl,-rpath,{PKG}",
                  "-Wl,-rpath,$ORIGIN/../../../../accv-lab_b200/accvlab/draw_heatmap"])
    return dst


def stage_all(force: bool = False):
    if not available():
        return staged()
    build_batching_helpers(force)
    stage_draw_heatmap_callers()
    build_cpp_benchmarks(force)
    return staged()


if __name__ == "__main__":
    print("staged:", stage_all(force="--force" in sys.argv))
