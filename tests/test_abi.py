"""CPU: the C-ABI library loads, exports every symbol include/dh_b200.h declares, validates arguments before it
touches a device, and the product never routes through the oracle. No compute calls (no GPU here)."""
import ctypes
import os
import re
import subprocess

import _cabi

ROOT = _cabi.ROOT


def test_library_exports_every_declared_symbol(cabi):
    names = _cabi.declared_functions()
    assert len(names) >= 12
    for n in names:
        assert hasattr(cabi, n), f"{n} is declared in include/dh_b200.h but not exported by libdh_b200.so"
    # and the ctypes table used by the tests covers the whole header
    assert set(names) == set(_cabi.SIGNATURES)


def test_exports_are_c_linkage():
    out = subprocess.run(["nm", "-D", "--defined-only", _cabi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT (dh_[a-z0-9_]+)\b", out))
    assert set(_cabi.declared_functions()) <= exported


def test_version_and_strings(cabi):
    assert cabi.dh_version() == 100
    info = cabi.dh_build_info().decode()
    assert "sm_100a" in info and "no CPU fallback" in info
    assert cabi.dh_status_string(0) == b"ok"
    assert b"workspace" in cabi.dh_status_string(_cabi.DH_ERR_WORKSPACE_TOO_SMALL)


def test_argument_validation_needs_no_device(cabi):
    one = ctypes.c_void_p(16)  # never dereferenced: validation fails first
    f = cabi.dh_draw_heatmap_batched
    assert f(None, one, one, one, one, 0, None, -1, 1, 1, 8, 8, 6.0, 1.0, 0, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert f(None, one, one, one, one, 0, None, 1, 1, 1, 8, 8, 6.0, 1.0, 7, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert f(None, one, one, one, one, 9, None, 1, 1, 1, 8, 8, 6.0, 1.0, 0, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert f(None, one, one, one, one, 0, None, 1, 1, 1, 0, 8, 6.0, 1.0, 0, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert f(None, one, one, one, one, 0, None, 1, 1, 1, 1 << 20, 1 << 20, 6.0, 1.0, 0, None) == _cabi.DH_ERR_UNSUPPORTED_SHAPE
    assert f(None, None, one, one, one, 0, None, 1, 1, 1, 8, 8, 6.0, 1.0, 0, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    # empty problems are OK without a device
    assert f(None, one, one, one, one, 0, None, 0, 5, 1, 8, 8, 6.0, 1.0, 0, None) == _cabi.DH_OK
    assert cabi.dh_radius_center(None, None, None, 1.0, 0, None, None) == _cabi.DH_OK
    assert cabi.dh_radius_center(None, None, None, 1.0, 5, None, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert cabi.dh_draw_heatmap(None, one, one, one, one, 3, 0, 8, 8, 6.0, 1.0, 0, None, 0, None) == _cabi.DH_OK
    assert cabi.dh_draw_heatmap(None, one, one, one, one, -3, 1, 8, 8, 6.0, 1.0, 0, None, 0, None) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert cabi.dh_fill_probe(None, one, 8) == _cabi.DH_ERR_INVALID_ARGUMENT
    assert cabi.dh_checksum(None, None, 0, 0, one) == _cabi.DH_OK
    assert cabi.dh_draw_heatmap_workspace_bytes(1000, 10) >= 4 * 1000 + 3 * 4 * 10
    n_plain = cabi.dh_host_scratch_bytes(16, 200, 10, 180, 180, 1, 4, 0)
    n_stage = cabi.dh_host_scratch_bytes(16, 200, 10, 180, 180, 1, 4, 1)
    assert n_stage - n_plain >= 2 * 4 * 10 * 180 * 180 * 4
    assert n_plain >= 16 * 200 * (8 + 16 + 4) + 16 * 8


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure; the product path must not reference it (or any CPU fallback)."""
    pkg = os.path.join(ROOT, "accv-lab_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f"{f} imports the oracle"
                assert "libdh_oracle" not in src and "dh_oracle" not in src, f"{f} references the oracle"


def test_python_surface_matches_reference():
    """Same names / argument names / defaults as the reference Python API (draw_heatmap_batched.py:27-34,
    draw_heatmap.cpp:131-145). Importing needs the compiled extension (loads libcudart, no device needed)."""
    import inspect

    import accvlab.draw_heatmap as dh

    assert dh.__all__ == ["__version__", "draw_heatmap", "draw_heatmap_batched"]
    sig = inspect.signature(dh.draw_heatmap_batched)
    positional = [n for n, q in sig.parameters.items() if q.kind is not inspect.Parameter.KEYWORD_ONLY]
    assert positional == ["heatmap", "centers", "radii", "diameter_to_sigma_factor", "k_scale", "labels"]
    assert [n for n in sig.parameters if n not in positional] == ["assume_zero_init"]  # keyword-only addition, default False
    assert sig.parameters["diameter_to_sigma_factor"].default == 6.0
    assert sig.parameters["k_scale"].default == 1.0 and sig.parameters["labels"].default is None
    from accvlab.draw_heatmap import draw_heatmap_ext as ext
    from accvlab.draw_heatmap.funtions import draw_heatmap, draw_heatmap_batched  # misspelt upstream; kept

    assert draw_heatmap is ext.draw_heatmap and draw_heatmap_batched is dh.draw_heatmap_batched
    doc = ext.draw_heatmap.__doc__
    assert "heatmaps" in doc and "heatmap_idxes" in doc and "diameter_to_sigma_factor" in doc and "k_scale" in doc
    for name in ("draw_heatmap_batched_impl", "draw_heatmap_batched_classwise_impl"):
        assert "nums_targets" in getattr(ext, name).__doc__


def test_python_errors_match_reference_conventions():
    """Host-side validation: RuntimeError from the extension (TORCH_CHECK), AssertionError from the wrapper."""
    import pytest
    import torch

    import accvlab.draw_heatmap as dh
    from accvlab.draw_heatmap.b200 import RaggedBatch

    heat = torch.zeros(2, 8, 8)
    c = torch.zeros(2, 3, 2, dtype=torch.int32)
    r = torch.zeros(2, 3, dtype=torch.int32)
    n = torch.tensor([1, 2])
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):  # draw_heatmap_cuda.cu:23
        dh.draw_heatmap_batched(heat, RaggedBatch(c, sample_sizes=n), RaggedBatch(r, sample_sizes=n))
    with pytest.raises(AssertionError, match="same maximum number of objects"):  # draw_heatmap_batched.py:58-61
        dh.draw_heatmap_batched(heat, RaggedBatch(c, sample_sizes=n), RaggedBatch(r[:, :2], sample_sizes=n))
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        dh.draw_heatmap(heat, c[0], r[0], r[0])
