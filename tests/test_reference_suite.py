"""The reference's OWN callers of the hot path, run UNMODIFIED against this package — the drop-in claim, proven.

oracle/stage_reference_tests.py (run by __graft_entry__.build() in the build container, where /root/reference exists)
stages, under the git-ignored / gpurun-shipped oracle/_ref/:
  * the reference's three test files + helpers       (packages/draw_heatmap/tests/*.py),
  * its three examples and two Python benchmarks     (packages/draw_heatmap/examples/*.py, benchmark/*.py),
  * its three C++ benchmarks, compiled from the reference's own .cpp files and linked against libdh_ref_shim.so
    (benchmark_cpp/bench_*.cpp:24-27 declare the launchers of draw_heatmap_cuda.cu:29-31,43-46),
  * the REAL accvlab.batching_helpers package (RaggedBatch: data_format/ragged_batch.py:31-171) with its two torch
    extensions compiled from the reference's sources.
Here every one of them runs in a subprocess whose PYTHONPATH puts accv-lab_b200/ (this repo's accvlab.draw_heatmap)
next to oracle/_ref/site/ (the reference's accvlab.batching_helpers); nothing is patched.
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
SITE = os.path.join(REF, "site")
DH = os.path.join(REF, "dh")
PKG_ROOT = os.path.join(ROOT, "accv-lab_b200")

needs_staging = pytest.mark.skipif(
    not os.path.exists(os.path.join(SITE, "accvlab", "batching_helpers", "__init__.py")),
    reason="oracle/_ref/ not staged (run `python oracle/stage_reference_tests.py` where /root/reference exists)")


def _env():
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([PKG_ROOT, SITE] + ([env["PYTHONPATH"]] if env.get("PYTHONPATH") else []))
    return env


def _run(cmd, cwd, timeout=600):
    r = subprocess.run(cmd, cwd=cwd, env=_env(), capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, f"{' '.join(cmd)} (cwd {cwd}) exited {r.returncode}\n--- stdout\n{r.stdout[-3000:]}\n--- stderr\n{r.stderr[-3000:]}"
    return r.stdout


# ---- CPU: the real RaggedBatch class is what the wrapper's attribute reads see ----------------------------------------
@needs_staging
def test_real_ragged_batch_feeds_the_wrapper_cpu():
    """No GPU needed: the REAL RaggedBatch (CPU tensors) goes through this repo's draw_heatmap_batched wrapper — its
    attribute reads and assertions run on the real class — and the extension then rejects the CPU tensors with the
    reference's own message (draw_heatmap_cuda.cu:23)."""
    code = r"""
import torch
import accvlab.draw_heatmap as dh, accvlab.batching_helpers as bh
assert 'accv-lab_b200' in dh.__file__ and '_ref' in bh.__file__, (dh.__file__, bh.__file__)
from accvlab.batching_helpers import RaggedBatch
B, T = 3, 5
sizes = torch.tensor([5, 0, 2], dtype=torch.int64)
c = RaggedBatch(torch.zeros(B, T, 2, dtype=torch.int32), sample_sizes=sizes)
r = RaggedBatch(torch.ones(B, T, dtype=torch.int32), sample_sizes=sizes)
l = RaggedBatch(torch.zeros(B, T, dtype=torch.int32), sample_sizes=sizes)
assert c.tensor.shape == (B, T, 2) and c.sample_sizes.dtype == torch.int64 and c.non_uniform_dim == 1
try:
    dh.draw_heatmap_batched(torch.zeros(B, 4, 8, 8), c, r, 6.0, 1.0, l)
except RuntimeError as e:
    assert 'must be a CUDA tensor' in str(e), e
else:
    raise SystemExit('CPU tensors were accepted')
bad = RaggedBatch(torch.ones(B, T + 1, dtype=torch.int32), sample_sizes=sizes)
try:
    dh.draw_heatmap_batched(torch.zeros(B, 8, 8), c, bad)
except AssertionError as e:
    assert 'maximum number of objects' in str(e)
else:
    raise SystemExit('shape mismatch was accepted')
print('ok')
"""
    assert "ok" in _run([sys.executable, "-c", code], cwd=ROOT)


# ---- GPU: the reference's tests / examples / benchmarks, unmodified ---------------------------------------------------
@pytest.mark.gpu
@needs_staging
@pytest.mark.parametrize("name", ["test_draw_heatmap.py", "test_draw_heatmap_batched.py",
                                  "test_draw_heatmap_batched_classwise.py"])
def test_reference_test_file_passes_unmodified(cuda_dev, name):
    """DH/tests/test_draw_heatmap.py:34-85, test_draw_heatmap_batched.py:36-97, test_draw_heatmap_batched_classwise.py:59-130."""
    out = _run([sys.executable, "-m", "pytest", "-x", "-q", "-p", "no:cacheprovider", name], cwd=os.path.join(DH, "tests"))
    assert "1 passed" in out, out


@pytest.mark.gpu
@needs_staging
@pytest.mark.parametrize("name,shape", [("draw_heatmap.py", None), ("draw_heatmap_batched.py", "torch.Size([4, 100, 100])"),
                                        ("draw_heatmap_batched_classwise.py", None)])
def test_reference_example_runs_unmodified(cuda_dev, name, shape):
    """DH/examples/*.py (draw_heatmap_batched.py:42-61 builds the real RaggedBatch) print the heatmap shape."""
    out = _run([sys.executable, name], cwd=os.path.join(DH, "examples"))
    assert "torch.Size" in out, out
    if shape:
        assert shape in out, out


@pytest.mark.gpu
@needs_staging
def test_reference_examples_match_the_reference_cpu_goldens(cuda_dev, golden):
    """The example inputs (DH/examples/input_data.py:35-59) through the staged example functions == the outputs of the
    reference's CPU drawer on the same inputs (tests/golden/ref_examples.npz, made by tests/golden/make_golden.py)."""
    code = r"""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import draw_heatmap_batched as ex_b, draw_heatmap as ex_f
hb = ex_b.draw_heatmap_example_batch().cpu().numpy()
np.save(sys.argv[1], hb)
"""
    import numpy as np
    import tempfile

    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "hb.npy")
        _run([sys.executable, "-c", code, path], cwd=os.path.join(DH, "examples"))
        got = np.load(path)
    want = golden("ref_examples")["heat"]  # [4,100,100]: the reference CPU drawer on the same example inputs
    assert want.shape == got.shape
    assert np.array_equal(got > 0, want > 0)
    assert float(np.abs(got - want).max()) <= 1e-6


@pytest.mark.gpu
@needs_staging
@pytest.mark.parametrize("name", ["benchmark.py", "benchmark_batched_classwise.py"])
def test_reference_python_benchmark_runs_unmodified(cuda_dev, name):
    """DH/benchmark/benchmark.py:141-159, benchmark_batched_classwise.py:139-158 (their per-object 'PyTorch' loop included)."""
    out = _run([sys.executable, name], cwd=os.path.join(DH, "benchmark"), timeout=900)
    assert "ms" in out.lower() or "speed" in out.lower(), out


@pytest.mark.gpu
@needs_staging
@pytest.mark.parametrize("name", ["bench_flattened", "bench_batched", "bench_batched_classwise"])
def test_reference_cpp_benchmark_links_and_runs_against_the_shim(cuda_dev, name):
    """DH/benchmark_cpp/bench_*.cpp compiled from the reference's files, linked against libdh_ref_shim.so."""
    exe = os.path.join(DH, "bench_cpp", name)
    if not os.path.exists(exe):
        pytest.skip("C++ benchmarks not staged")
    out = _run([exe], cwd=os.path.join(DH, "bench_cpp"))
    assert "Average time per iteration" in out and "Error in draw_heatmap_cuda" not in out, out
