"""CPU: the C-ABI library loads and exports every symbol include/chs.h declares; struct layouts seen by the
ctypes binding equal the C compiler's.  No compute calls (no GPU here)."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
HEADER = os.path.join(ROOT, "include", "chs.h")


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    ge.build()
    import rl_cattle_herding_b200 as pkg
    return pkg._lib


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(chs_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    l = lib.load()
    names = declared_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(l, n), f"{n} declared in include/chs.h but not exported by libchs.so"
    assert set(names) == set(lib.SYMBOLS), (set(names) ^ set(lib.SYMBOLS))
    assert l.chs_abi_version() == lib.ABI_VERSION


def test_struct_layouts_match_c(lib, tmp_path):
    prog = tmp_path / "sizes.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "chs.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu\\n",'
                    "sizeof(ChsParams), sizeof(ChsDims), sizeof(ChsStateView), offsetof(ChsParams, seed),"
                    "offsetof(ChsStateView, step_counter), offsetof(ChsStateView, episode));return 0;}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(lib.ChsParams), C.sizeof(lib.ChsDims), C.sizeof(lib.ChsStateView), lib.ChsParams.seed.offset,
            lib.ChsStateView.step_counter.offset, lib.ChsStateView.episode.offset]
    assert got == want


def test_header_compiles_as_plain_c():
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", HEADER], check=True)


def test_argument_errors_without_gpu(lib):
    """Validation happens before any CUDA call, so it can be exercised on the CPU box."""
    l = lib.load()
    h = C.c_void_p(0)
    p = lib.ChsParams()
    p.struct_size = 4   # wrong on purpose
    assert l.chs_create(C.byref(p), 8, 0, C.byref(h)) == -1
    assert b"struct_size" in l.chs_last_error()
    p.struct_size = C.sizeof(lib.ChsParams)
    p.num_cattle, p.max_drones, p.pyb_freq, p.ctrl_freq = 20, 3, 240, 7
    assert l.chs_create(C.byref(p), 8, 0, C.byref(h)) == -1
    assert b"not divisible" in l.chs_last_error()   # same message as BaseAviary.py:102-103
    p.ctrl_freq, p.max_drones = 30, 13
    assert l.chs_create(C.byref(p), 8, 0, C.byref(h)) == -1
    assert l.chs_step(None, None, None, None, None, None, None, None) == -1
    assert l.chs_destroy(None) == 0


def test_product_path_has_no_oracle_or_cpu_fallback():
    """The package never imports the oracle and raises without a GPU instead of falling back."""
    pkg_dir = os.path.join(ROOT, "rl-cattle-herding_b200")
    for fn in os.listdir(pkg_dir):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg_dir, fn)).read()
            assert "cattle_oracle" not in src and "oracle" not in src.replace("oracle/ref_harness", ""), fn
    import torch
    if not torch.cuda.is_available():
        import rl_cattle_herding_b200 as pkg
        with pytest.raises(RuntimeError):
            pkg.CattleHerdingBatchedEnv(4)
