"""The adaptor classes against the REAL libMesh headers: every function the reference declares pure virtual in
include/numerics/sparse_matrix.h and include/numerics/numeric_vector.h must have an `override` of the same name, parameter
count and constness in libmesh_b200/shim/gpu_sparse_matrix.h / gpu_numeric_vector.h.  (The adaptor is compiled against the mock
of tests/shim/libmesh_mock.h because libMesh cannot be built here; this check reads the reference's own headers, so a virtual
missing from the mock cannot hide a missing override.)  Runs where /root/reference exists."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/include/numerics"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="/root/reference is not present on this machine")


def _strip_comments(s):
    s = re.sub(r"/\*.*?\*/", "", s, flags=re.S)
    return re.sub(r"//[^\n]*", "", s)


def _split_params(p):
    p = p.strip()
    if not p or p == "void":
        return []
    out, depth, cur = [], 0, ""
    for ch in p:
        if ch in "<(":
            depth += 1
        elif ch in ">)":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    out.append(cur)
    return out


def _signature(decl):
    """'ret name (params) const' -> (name, n_params, is_const)"""
    m = re.search(r"(operator\s*(?:\(\s*\)|[^\s(]+)|[A-Za-z_]\w*)\s*\((.*)\)\s*(const)?\s*$", " ".join(decl.split()), flags=re.S)
    assert m, decl
    name = re.sub(r"\s+", "", m.group(1))
    return name, len(_split_params(m.group(2))), bool(m.group(3))


def pure_virtuals(header):
    s = _strip_comments(open(os.path.join(REF, header)).read())
    return [_signature(m.group(1)) for m in re.finditer(r"virtual\s+([^;{}]*?)=\s*0\s*;", s, flags=re.S)]


def overrides(header):
    s = _strip_comments(open(os.path.join(ROOT, "libmesh_b200", "shim", header)).read())
    found = set()
    for m in re.finditer(r"([^;{}]*?\))\s*(const)?\s*override", s, flags=re.S):
        decl = m.group(1) + (" const" if m.group(2) else "")
        decl = decl.split("\n\n")[-1]
        try:
            found.add(_signature(decl))
        except AssertionError:
            pass
    return found


@pytest.mark.parametrize("ref_header,shim_header,expected", [("sparse_matrix.h", "gpu_sparse_matrix.h", 27), ("numeric_vector.h", "gpu_numeric_vector.h", 47)])
def test_every_pure_virtual_of_the_reference_has_an_override(ref_header, shim_header, expected):
    want = pure_virtuals(ref_header)
    assert len(want) == expected, f"the reference's {ref_header} declares {len(want)} pure virtuals"
    have = overrides(shim_header)
    missing = sorted({w for w in want if w not in have})
    assert not missing, f"{shim_header} lacks overrides of {missing}"
