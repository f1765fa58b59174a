"""The C-ABI library loads and exports every symbol include/madpose_b200.h declares (no compute)."""
import os
import re

from madpose_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "madpose_b200.h")).read()
    declared = set(re.findall(r"\b(mp_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"mp_ctx"}
    assert {"mp_estimate_batch", "mp_score_batch", "mp_inliers", "mp_minimal_solve_batch", "mp_lm_batch"} <= declared
    lib = capi.load_library()
    for s in sorted(declared):
        assert hasattr(lib, s), f"libmadpose_b200.so does not export {s}"
    assert set(capi._SYMBOLS) == declared


def test_struct_sizes_match_header(tmp_path):
    """ctypes mirrors vs the sizes gcc gives the structs of the header."""
    import ctypes as C
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "madpose_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu\\n",'
                   'sizeof(mp_options),sizeof(mp_config),sizeof(mp_pair),sizeof(mp_stats),sizeof(mp_model));return 0;}')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(capi.Options), C.sizeof(capi.Config), C.sizeof(capi.Pair), C.sizeof(capi.Stats),
                     8 * capi.MODEL_DOUBLES]


def test_no_device_fails_loudly():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CUDA device|CPU"):
        capi.Context(0)
