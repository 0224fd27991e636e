"""CPU suite, part 2: the C-ABI library loads and exports every symbol include/fdsm.h declares,
and the host-side adapters mirror the reference's argument/error behaviour.  No compute calls."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            txt = open(os.path.join(ROOT, "include", fn)).read()
            txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
            names |= set(re.findall(r"\b(fdsm_[a-z0-9_]+)\s*\(", txt))
    return names


def test_library_exports_every_declared_symbol(fdsm):
    lib = fdsm._lib.load()
    syms = _header_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(lib, s), f"libfdsm.so does not export {s}"
    assert syms == set(fdsm._lib.PROTOTYPES), "ctypes prototype table out of sync with include/*.h"
    hdr = open(os.path.join(ROOT, "include", "fdsm.h")).read()
    assert lib.fdsm_version() == int(re.search(r"#define FDSM_ABI_VERSION (\d+)", hdr).group(1)) == fdsm._lib.ABI_VERSION


def test_stack_args_struct_matches_header(fdsm):
    """Field order of the ctypes mirror == field order of the C struct."""
    txt = open(os.path.join(ROOT, "include", "fdsm.h")).read()
    body = txt[txt.index("typedef struct fdsm_stack_args {"):txt.index("} fdsm_stack_args;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split("{", 1)[1].split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            fields.append(re.findall(r"([A-Za-z_0-9]+)\s*$", part.strip())[0])
    assert fields == [f[0] for f in fdsm._lib.StackArgs._fields_]


def test_integration_doc_struct_matches_header(fdsm):
    """The ctypes mirror a maintainer would paste from INTEGRATION.md has the fields (names, order, ctypes) of _lib.py's."""
    txt = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    body = txt[txt.index("class StackArgs(C.Structure):"):]
    body = body[body.index("_fields_ = ["):body.index("]\n") + 1]
    doc = re.findall(r'\("([A-Za-z_0-9]+)",\s*C\.(c_[a-z0-9_]+)\)', body)
    # same names in the same order, each field of the same size and kind (pointer / integer / float)
    kind = lambda t: ("p" if t in (ctypes.c_void_p, fdsm._lib._f32p) else "f" if t is ctypes.c_float else "i", ctypes.sizeof(t))
    assert [(n, kind(getattr(ctypes, t))) for n, t in doc] == [(n, kind(t)) for n, t in fdsm._lib.StackArgs._fields_]


def test_no_cpu_fallback(fdsm):
    """CPU tensors are refused loudly: the product has no CPU path."""
    x = torch.randn(4, 8)
    with pytest.raises(fdsm._lib.FdsmError):
        fdsm.efficient_score_matching_conjugate(lambda z: z.sum(-1), x)


def test_reference_error_behaviour(fdsm):
    x = torch.randn(4, 3, 8, 8)
    with pytest.raises(RuntimeError):  # reference is only shape-correct for n_particles=1
        fdsm.anneal_ESM_scorenet_VR(lambda z, y: z, x, torch.ones(3), n_particles=2)
    with pytest.raises(fdsm._lib.FdsmError):  # sign projections are implemented (efficient_sm.py:10-11) but only on the GPU
        fdsm.single_efficient_score_matching(lambda z: z.sum(-1), torch.randn(4, 8), noise_type="radermacher")
    with pytest.raises(ValueError):
        fdsm.fd_loss(lambda z: z, x, mode="nope")


def test_product_never_imports_oracle():
    for fn in os.listdir(os.path.join(ROOT, "scripts")):      # measurement scripts time the product: no checker in them
        if fn.endswith(".py"):
            txt = open(os.path.join(ROOT, "scripts", fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f"scripts/{fn} imports oracle/"
    pkg = os.path.join(ROOT, "fd-scorematching_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f"{fn} imports oracle/"
                assert "/root/reference" not in txt


def test_integration_md_names_every_entry_point():
    """INTEGRATION.md is the maintainer's map from C entry point to the reference code it replaces: every function
    include/fdsm.h declares appears there (the `a/b` and `_fwd/bwd` shorthands of its table are expanded)."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "fdsm.h")).read()
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    names = sorted(set(re.findall(r"\b(fdsm_[a-z0-9_]+)\s*\(", hdr)))
    mentioned = set(re.findall(r"fdsm_[a-z0-9_]+", doc))
    for m in re.findall(r"`(fdsm_[a-z0-9_]+(?:/[a-z0-9_]+)+)`", doc):      # `fdsm_mlp_sizes/prepare/forward/backward`, `..._fwd/bwd`
        head, *rest = m.split("/")
        stem = head[:head.rfind("_") + 1]
        mentioned.update(stem + r for r in rest)
    missing = [n for n in names if n not in mentioned]
    assert not missing, missing
