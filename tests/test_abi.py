"""The C-ABI library loads on a CPU-only box and exports every entry point the header declares.
No compute calls are made here (they need a GPU and fail loudly without one)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "dartray_engine.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dre_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    ge.build()
    from dartray_b200 import engine
    return engine.lib()


def test_header_symbols_exported(lib):
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dartray_engine.h but not exported"


def test_python_binding_lists_every_symbol():
    from dartray_b200 import engine
    assert sorted(engine.SYMBOLS) == _declared()


def test_struct_layouts_match_header():
    from dartray_b200 import abi
    # dre_params: 4 int32 + 4 double + 4 int32 + 6 double ; dre_counters: 10 int64 + 2 double + uint64
    assert C.sizeof(abi.Params) == 4 * 4 + 4 * 8 + 4 * 4 + 6 * 8
    assert C.sizeof(abi.Counters) == 17 * 8
    assert abi.ISECT_DTYPE.itemsize == 32


def test_no_cpu_fallback(lib):
    """Without a CUDA device dre_create must fail; with one it must succeed."""
    import torch
    h = C.c_void_p()
    rc = lib.dre_create(C.byref(h), 0, 0, 1, None)
    if torch.cuda.is_available():
        assert rc == 0
        lib.dre_destroy(h)
    else:
        assert rc != 0 and not h.value


def build_abi_host(out_dir):
    """gcc-compiles tests/abi_host.c (a C host that calls the ABI the way the Fortran wrappers do) against the library."""
    import subprocess
    exe = os.path.join(str(out_dir), "abi_host")
    libdir = os.path.join(ROOT, "dart-ray_b200")
    subprocess.check_call(["gcc", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "abi_host.c"),
                           "-L" + libdir, "-ldartray_engine", "-Wl,-rpath," + libdir, "-o", exe])
    return exe


def test_c_host_compiles_and_links(lib, tmp_path):
    """The header is valid C (no C++-isms, no torch types) and a plain C program links against the library; without a GPU
    the program stops at dre_create with a message, it never computes on the CPU."""
    import subprocess
    import torch
    exe = build_abi_host(tmp_path)
    if not torch.cuda.is_available():
        r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
        assert r.returncode != 0          # no scene files and no device: a loud failure, not a CPU result
