"""The C-ABI library loads on a GPU-less host and exports every entry point include/ccn_b200.h declares
(no compute calls here).  Also: the product package never imports the oracle."""
import os
import re
import subprocess

import pytest

from curvecloudnet_b200 import native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(native.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return native.load()


def test_every_declared_symbol_is_exported(lib):
    assert len(native.SIGNATURES) >= 30
    for name in native.SIGNATURES:
        assert hasattr(lib, name), name
    # entry points whose return type is not int / int64_t (the signature parser of native.py skips them): every `ccn_*(` the
    # header declares must resolve in the library
    hdr = open(os.path.join(ROOT, "include", "ccn_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ccn_\w+)\s*\(", hdr)) - {"ccn_draw_fn"}
    assert {"ccn_forward_run", "ccn_exec_create", "ccn_exec_destroy", "ccn_exec_topology_flags", "ccn_sgcnn_fused",
            "ccn_pointconv_fused", "ccn_exec_profile", "ccn_exec_host_stats"} <= declared
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ccn_version() >= 1
    assert native.grid_num_cells(1000) == 2048
    assert native.kernel_launches() == 0 or native.kernel_launches() > 0


def test_fast_binding_wraps_every_entry_point(lib):
    """The generated CPython extension (curvecloudnet_b200/_fastcall_build.py): built from THIS header, one wrapper per
    entry point, and an argument error comes back through it exactly as through ctypes (no GPU call is reached)."""
    from curvecloudnet_b200 import _fastcall_build
    so = [f for f in os.listdir(os.path.dirname(native.LIB_PATH)) if f.startswith("_ccn_cffi.") and f.endswith(".so")]
    if not so:
        _fastcall_build.build()
    from curvecloudnet_b200 import _ccn_cffi
    assert _ccn_cffi.ffi.string(_ccn_cffi.lib.w_header_hash()).decode() == _fastcall_build.header_hash()
    for name in native.SIGNATURES:
        assert hasattr(_ccn_cffi.lib, "w_" + name), name
    args = (0, 64, 0, 100, 9, 64, 0, 64, 0)                  # feat = NULL: rejected by the argument check
    assert _ccn_cffi.lib.w_ccn_group_max_dense(*args) == lib.ccn_group_max_dense(*args) == -1
    assert _ccn_cffi.lib.w_ccn_grid_num_cells(1000) == 2048
    if os.environ.get("CCN_BINDING", "") != "ctypes":
        assert native.BINDING == "cffi"


def test_library_has_sm100a_code_only(lib):
    out = subprocess.run(["cuobjdump", "--list-elf", native.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs


def test_cpu_tensors_are_rejected_not_emulated(lib):
    import torch
    with pytest.raises(native.CcnError):
        native.call("ccn_gather_i64", torch.zeros(4, dtype=torch.int64), torch.zeros(4, dtype=torch.int64), 4,
                    torch.zeros(4, dtype=torch.int64))


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "curvecloudnet_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), os.path.join(dirpath, f)
