"""CPU: the C-ABI library loads and exports every symbol include/cdx.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "cdx.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(cdx_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from compliancedex_b200 import _lib
    if not os.path.isfile(_lib.LIB_PATH):
        import __graft_entry__ as ge
        ge.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"libcdx.so does not export {n}"
    assert set(names) == set(_lib.SIGNATURES), "python binding table out of sync with include/cdx.h"


def test_no_cpu_fallback_without_gpu():
    import torch
    from compliancedex_b200 import _lib, GPIS
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    h = ctypes.c_void_p()
    rc = lib.cdx_create(0, ctypes.byref(h))
    assert rc == _lib.CDX_ERR_CUDA and b"no CPU fallback" in lib.cdx_last_error()
    g = GPIS(0.08, 1.0)
    with pytest.raises(RuntimeError):
        g.fit(torch.zeros(4, 3, dtype=torch.float64), torch.zeros(4, 1, dtype=torch.float64))


def test_params_struct_layout_matches_header():
    from compliancedex_b200 import _lib
    # field order / count must follow struct cdx_pregrasp_params in include/cdx.h
    txt = open(os.path.join(ROOT, "include", "cdx.h")).read()
    body = txt[txt.index("typedef struct cdx_pregrasp_params {"):txt.index("} cdx_pregrasp_params;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split("{", 1)[1].split(";"):
        decl = decl.strip()
        if not decl:
            continue
        _, rest = decl.split(None, 1)
        for part in rest.split(","):
            fields.append(re.match(r"\s*([A-Za-z_0-9]+)", part).group(1))
    assert fields == [f[0] for f in _lib.PregraspParams._fields_]
