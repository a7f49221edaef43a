"""CPU-side checks of the drop-in boundary: the shared library loads without a GPU and exports every
symbol include/gcr_b200.h declares (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    if not os.path.exists(g.LIB):
        g.build()
    import gcr_capi
    return gcr_capi.load_library()


def test_header_symbols_are_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "gcr_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(gcr_[a-z_]+)\s*\(", hdr))
    assert {"gcr_create", "gcr_destroy", "gcr_set_features", "gcr_run", "gcr_best", "gcr_last_error"} <= declared
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in gcr_b200.h but not exported"
    import gcr_capi
    assert declared == set(gcr_capi.EXPORTS)


def test_version_and_struct_layout(lib):
    import gcr_capi
    assert b"sm_100a" in lib.gcr_version()
    assert ctypes.sizeof(gcr_capi.GcrResult) == 8 + 8 + 32 + 8 + 8 + 6 * 8 + 8 + 8 + 24 + 8 + 8 + 8
    assert gcr_capi.RECORD_DTYPE.itemsize == 64


def test_ctypes_mirror_matches_the_header_as_compiled(tmp_path):
    """gcr_result as gcc lays it out from include/gcr_b200.h == the ctypes mirror in gcr_capi (size and every field offset)."""
    import subprocess
    import gcr_capi
    fields = [name for name, _ in gcr_capi.GcrResult._fields_]
    src = tmp_path / "layout.c"
    src.write_text("#include <stdio.h>\n#include <stddef.h>\n#include \"gcr_b200.h\"\nint main(void) {\n"
                   "  printf(\"%zu\\n\", sizeof(gcr_result));\n"
                   + "".join(f"  printf(\"%zu\\n\", offsetof(gcr_result, {f}));\n" for f in fields)
                   + "  printf(\"%zu\\n\", sizeof(gcr_record));\n"
                   + "  printf(\"%zu\\n\", sizeof(gcr_lo_round_io));\n"
                   + "".join(f"  printf(\"%zu\\n\", offsetof(gcr_lo_round_io, {f}));\n" for f, _ in gcr_capi.GcrLoRoundIo._fields_)
                   + "  return 0;\n}\n")
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    nf = len(fields)
    assert out[0] == ctypes.sizeof(gcr_capi.GcrResult)
    assert out[1:1 + nf] == [getattr(gcr_capi.GcrResult, f).offset for f in fields]
    assert out[1 + nf] == gcr_capi.RECORD_DTYPE.itemsize
    assert out[2 + nf] == ctypes.sizeof(gcr_capi.GcrLoRoundIo)  # gcr_lo_round_io and its ctypes mirror
    assert out[3 + nf:] == [getattr(gcr_capi.GcrLoRoundIo, f).offset for f, _ in gcr_capi.GcrLoRoundIo._fields_]


def test_null_handle_is_an_error_not_a_crash(lib):
    assert lib.gcr_destroy(None) == -1
    assert lib.gcr_set_features(None, None, 0, None, 0, 0) == -1
    assert lib.gcr_last_error(None) == b"null handle"


def test_no_cpu_fallback_in_product_sources():
    """The product path must not reach into oracle/ (BASELINE tier rule)."""
    pkg = os.path.join(ROOT, "graph-cut-ransac_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in text and "libgcr_oracle" not in text and "libgcr_ref" not in text, f
