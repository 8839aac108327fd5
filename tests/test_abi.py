"""CPU-side checks of the drop-in boundary: libcpc_b200.so loads and exports every symbol that
include/cpc_b200.h declares, the ctypes descriptor structs match the header's field lists, and the host-side
modules keep the reference's constructor / state_dict surface.  No kernel is launched here."""
import ctypes
import os
import re

import pytest
import torch

from cpc2_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "cpc_b200.h")


def header_text():
    with open(HEADER) as f:
        return f.read()


def declared_symbols():
    return sorted(set(re.findall(r"^(?:int|const char\*)\s+(cpc_[a-z0-9_]+)\s*\(", header_text(), flags=re.M)))


@pytest.fixture(scope="module")
def lib():
    if _lib.needs_build():
        _lib.build()
    return _lib.lib()


def test_header_and_binding_list_agree():
    assert declared_symbols() == sorted(_lib.EXPORTED)


def test_library_exports_every_declared_symbol(lib):
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_version_and_strerror_do_not_need_a_gpu(lib):
    assert lib.cpc_version() >= 1
    assert lib.cpc_strerror(0).decode().lower().startswith("ok")
    assert lib.cpc_strerror(-1).decode() != lib.cpc_strerror(0).decode()


STRUCTS = {"CpcGemmDesc": _lib.GemmDesc, "CpcRowNormFwd": _lib.RowNormFwd, "CpcRowNormBwd": _lib.RowNormBwd,
           "CpcConv0": _lib.Conv0, "CpcNce": _lib.Nce, "CpcLstm": _lib.Lstm, "CpcAttn": _lib.Attn,
           "CpcRowMap": _lib.RowMap, "CpcAdamSpan": _lib.AdamSpan, "CpcColNorm": _lib.ColNorm}


def c_struct_fields(name):
    text = re.sub(r"/\*.*?\*/", "", header_text(), flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    m = re.search(r"typedef struct\s*\{([^}]*)\}\s*" + name + r"\s*;", text)
    assert m, f"{name} not found in the header"
    body = m.group(1)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        names = decl.split(",")
        first = names[0].split()[-1]
        fields.append(first.lstrip("*"))
        fields += [n.strip().lstrip("*") for n in names[1:]]
    return fields


@pytest.mark.parametrize("cname", sorted(STRUCTS))
def test_ctypes_struct_matches_header(cname):
    alias = {"inp": "in"}  # `in` is a Python keyword
    assert [alias.get(f[0], f[0]) for f in STRUCTS[cname]._fields_] == c_struct_fields(cname)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU / eager fallback"):
        _lib.lib()


def test_cpu_tensors_are_rejected():
    from cpc2_b200 import model
    enc = model.CPCEncoder(32)
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        enc(torch.zeros(1, 1, 320))
