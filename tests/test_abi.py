"""The C-ABI library builds for sm_100a, loads without a GPU, exports every symbol that
include/ihm2b.h declares, and its config struct has the layout the ctypes mirror assumes."""
import ctypes as C
import os
import re
import subprocess

import pytest

from ihm2_b200 import _abi, engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(built):
    lib = engine.load_library()
    hdr = open(os.path.join(ROOT, "include", "ihm2b.h")).read()
    declared = set(re.findall(r"^(?:int|int64_t|const char\*)\s+(ihm2b_[a-z0-9_]+)\s*\(", hdr, re.M))
    assert declared == set(engine.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ihm2b_version() >= 100


def test_config_struct_layout_matches_header(built, tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "ihm2b.h"\nint main(){printf("%zu %zu %zu %zu %zu\\n",'
                   "sizeof(ihm2b_config), offsetof(ihm2b_config, dt), offsetof(ihm2b_config, W),"
                   "offsetof(ihm2b_config, irk_A), offsetof(ihm2b_config, slack_Z));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    vals = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert vals == [C.sizeof(_abi.Config), _abi.Config.dt.offset, _abi.Config.W.offset, _abi.Config.irk_A.offset,
                    _abi.Config.slack_Z.offset]


def test_binary_is_sm100a_only(built):
    so = os.path.join(ROOT, "ihm2_b200", "lib", "libihm2b.so")
    out = subprocess.run(["cuobjdump", "--list-elf", so], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, out


def test_calls_fail_loudly_without_a_gpu(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ihm2_b200.track import load_packaged_track
    mp = load_packaged_track("skidpad")
    cfg = _abi.make_config(batch=4, n_grid=len(mp.s_ref))
    with pytest.raises(engine.Ihm2bError):
        engine.Engine(cfg, mp.s_ref, mp.kappa_ref)
    # the C entry point itself reports the missing device instead of computing anything
    lib = engine.load_library()
    h = C.c_void_p(0)
    assert lib.ihm2b_create(C.byref(cfg), C.byref(h)) != 0
    assert b"CUDA" in lib.ihm2b_last_error() or b"cuda" in lib.ihm2b_last_error()


def test_no_oracle_in_product_code():
    """The product package must never import / link the oracle."""
    pkg = os.path.join(ROOT, "ihm2_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", "Makefile")):
                txt = open(os.path.join(d, f), errors="ignore").read()
                assert "import oracle" not in txt and "from oracle" not in txt and "ihm2o_" not in txt, f
