"""The C-ABI library loads on a machine without a GPU, exports every symbol that
include/rti_b200.h declares, and validates its arguments before touching CUDA."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


def header_functions():
    text = open(os.path.join(ROOT, "include", "rti_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rti_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(lib):
    names = header_functions()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), "librti_b200.so does not export %s" % name


def test_python_signature_table_matches_header(lib):
    from rti_b200 import _capi
    assert sorted(_capi.SIGNATURES) == header_functions()


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "rti_b200.h"\nint main(void){return RTI_OK;}\n')
    import subprocess
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    "-c", str(src), "-o", str(tmp_path / "t.o")], check=True)


def test_version_and_error_string(lib):
    assert lib.rti_version() >= 100
    assert isinstance(lib.rti_last_error_string(), bytes)


def test_argument_errors_without_gpu(lib):
    from rti_b200 import _capi
    # NULL pointers
    assert lib.rti_gather_soa(None, 0, None, 4, 0, 1, 1, None, None, None, None) == _capi.RTI_ERR_ARG
    assert b"NULL" in lib.rti_last_error_string()
    buf = (C.c_double * 64)()
    p = C.cast(buf, C.c_void_p)
    # PTM with fewer than 6 samples: the reference raises ValueError (analysis.py:296-298)
    assert lib.rti_ptm_fit_frames(p, 4, p, 5, 0, 2, 2, p, 4, None, None) == _capi.RTI_ERR_TOO_FEW
    assert lib.rti_ptm_fit_soa(p, p, p, 4, 3, p, 4, None, None) == _capi.RTI_ERR_TOO_FEW
    with pytest.raises(ValueError):
        _capi.check(_capi.RTI_ERR_TOO_FEW)
    # inconsistent strides / sizes
    assert lib.rti_ptm_fit_frames(p, 3, p, 8, 0, 2, 2, p, 4, None, None) == _capi.RTI_ERR_ARG
    assert lib.rti_ptm_table(p, 2, 4, p, 100, p, 4, None, None) == _capi.RTI_ERR_ARG
    assert lib.rti_ptm_table(p, 4, 4, p, 0, p, 4, None, None) == _capi.RTI_ERR_ARG
    assert lib.rti_relight_ptm(p, 4, 0, 0, 0.0, 0.0, None, p, None) == _capi.RTI_ERR_ARG
    assert lib.rti_ptm_compute_host(p, p, 8, 4, 4, 2, 4, p, 100, p, None, 0) == _capi.RTI_ERR_ARG
    with pytest.raises(_capi.RtiError):
        _capi.check(_capi.RTI_ERR_ARG)
    with pytest.raises(_capi.RtiSingularError):
        _capi.check(_capi.RTI_ERR_SINGULAR)


def test_missing_library_fails_loudly(monkeypatch):
    from rti_b200 import _capi
    monkeypatch.setattr(_capi, "_lib", None)
    monkeypatch.setattr(_capi, "LIB_PATH", "/nonexistent/librti_b200.so")
    with pytest.raises(_capi.RtiLibraryMissing):
        _capi.load()
