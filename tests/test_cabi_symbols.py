"""CPU checks of the drop-in boundary: libb200sparse.so loads without a GPU, exports every
function include/b200_sparse.h declares, and the ctypes table mirrors the header one to one.
No compute entry point is called here (no GPU in this container)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200_sparse.h")


def header_decls():
    """{name: n_args} of every function prototype in the header."""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    out = {}
    for m in re.finditer(r"\b(b200_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out


def test_header_declares_the_path():
    d = header_decls()
    for name in ("b200_tbe_forward", "b200_tbe_backward", "b200_tbe_sort_indices", "b200_cumsum",
                 "b200_permute_data", "b200_block_bucketize", "b200_bounds_check", "b200_tbe_int8_forward",
                 "b200_interaction_forward", "b200_interaction_backward", "b200_pooled_a2a_recat"):
        assert name in d, name


def test_library_exports_every_declared_symbol():
    from b200_sparse import cabi
    assert os.path.exists(cabi.LIB_PATH), "run `python __graft_entry__.py` (build()) first"
    lib = ctypes.CDLL(cabi.LIB_PATH)
    for name in header_decls():
        assert hasattr(lib, name), f"{name} declared in b200_sparse.h but not exported"


def test_ctypes_table_mirrors_header():
    from b200_sparse import cabi
    decl = header_decls()
    assert set(cabi.SIGNATURES) == set(decl)
    for name, (_res, argtypes) in cabi.SIGNATURES.items():
        assert len(argtypes) == decl[name], f"{name}: ctypes has {len(argtypes)} args, header {decl[name]}"


def test_host_only_entry_points():
    """The three entry points that touch no device are callable on CPU."""
    from b200_sparse import cabi
    lib = cabi.load()
    assert lib.b200_abi_version() == 5
    assert lib.b200_scan_workspace_bytes() > 0
    assert lib.b200_error_string(-1) and lib.b200_error_string(-1) != lib.b200_error_string(0)
    assert lib.b200_tbe_sort_workspace_bytes(1 << 20, 32, 26) > 0
    assert lib.b200_tbe_backward_scratch_bytes(1 << 20, 128) > 0


def test_no_signature_takes_torch_types():
    src = open(HEADER).read()
    code = re.sub(r"//[^\n]*", " ", re.sub(r"/\*.*?\*/", " ", src, flags=re.S))
    assert "torch" not in code.lower() and "at::" not in code and "Tensor" not in code and 'extern "C"' in code


def test_product_has_no_cpu_fallback():
    """ops refuses CPU tensors instead of silently computing elsewhere."""
    import torch
    from b200_sparse import ops
    with pytest.raises((ValueError, RuntimeError, AssertionError, TypeError)):
        ops.cumsum(torch.arange(4, dtype=torch.int64), 0)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "torchrec-oldfork_b200")
    for dp, _dn, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), os.path.join(dp, f)
