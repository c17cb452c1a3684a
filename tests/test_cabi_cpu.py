"""CPU-side checks of the boundary: the CUDA library loads, exports every symbol include/dagcon_b200.h declares,
and refuses to run without a device (there is no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    import __graft_entry__ as ge
    ge.build_product()
    p = os.path.join(ROOT, "pbdagcon_b200", "csrc", "libdagcon_b200.so")
    assert os.path.exists(p)
    return p


def test_exports_match_header(lib_path):
    hdr = open(os.path.join(ROOT, "include", "dagcon_b200.h")).read()
    declared = sorted(set(re.findall(r"^(?:int|void|const char\*)\s+(dgc_\w+)\s*\(", hdr, re.M)))
    assert len(declared) >= 17
    lib = C.CDLL(lib_path)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    import pbdagcon_b200.api as api
    assert sorted(api.EXPORTS) == declared
    assert api.version() == int(re.search(r"#define DGC_ABI_VERSION (\d+)", hdr).group(1))


def test_struct_layouts_match_header():
    import pbdagcon_b200.api as api
    assert C.sizeof(api.Opts) == 32
    assert C.sizeof(api.CBatch) == 8 + 8 * 8 + 8 + 8
    assert C.sizeof(api.CResults) == 8 + 8 * 7 + 8
    assert C.sizeof(api.Timings) == 10 * 4 + 2 * 4 + 4 * 8 + 2 * 4


def test_no_cpu_fallback(lib_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import pbdagcon_b200.api as api
    with pytest.raises(api.DagconError, match="no CUDA device|CPU"):
        api.Consensus()


def test_product_does_not_reference_oracle():
    """The product path must never import, link or call anything under oracle/."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "pbdagcon_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", "Makefile")) and f != "smoke.py":
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "libdagcon_oracle" not in txt and "oracle/" not in txt and "orc_" not in txt, os.path.join(dirpath, f)
