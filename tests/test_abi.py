"""The C-ABI shared library: loads on a CPU-only machine, exports every symbol that include/b200svs.h
declares with the declared arity, and fails loudly (never falls back) when asked to compute without a GPU."""
import ctypes as C
import os
import re

import pytest
import torch

from music_generation_b200 import _lib, ops

_INC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include")
HEADER = os.path.join(_INC, "b200svs.h")
DEBUG_HEADER = os.path.join(_INC, "b200svs_debug.h")       # bring-up probes: exported, not part of the product ABI


def declared_functions():
    src = open(HEADER).read() + open(DEBUG_HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(?:int|const char\*)\s+(b200svs_\w+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        n = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
        out[m.group(1)] = n
    return out


def test_header_and_library_agree(built_lib):
    decl = declared_functions()
    assert len(decl) >= 18
    for name, nargs in decl.items():
        assert hasattr(built_lib, name), f"{name} declared in b200svs.h but not exported"
        if name in _lib.PROTOTYPES:
            assert len(_lib.PROTOTYPES[name]) == nargs, f"{name}: binding has {len(_lib.PROTOTYPES[name])} args, header {nargs}"
    for name in _lib.PROTOTYPES:
        assert name in decl, f"{name} bound in _lib.py but not declared in the header"


def test_struct_layout_matches_header():
    # field order of the ctypes mirrors follows the C declarations
    src = open(HEADER).read()
    body = src[src.index("typedef struct b200svs_epilogue {"):src.index("} b200svs_epilogue;")]
    names = re.findall(r"\b(\w+);", re.sub(r"/\*.*?\*/", "", body, flags=re.S))
    assert names == [f[0] for f in _lib.Epilogue._fields_]
    body = src[src.index("typedef struct b200svs_conv_args {"):src.index("} b200svs_conv_args;")]
    names = []
    for line in re.sub(r"/\*.*?\*/", "", body, flags=re.S).split(";"):
        line = line.strip()
        decl = line.split("{")[-1].strip()
        if not decl:
            continue
        parts = decl.replace("*", " ").split()
        names += [p.strip() for p in " ".join(parts[1:] if parts[0] != "const" else parts[2:]).split(",")]
    assert names == [f[0] for f in _lib.ConvArgs._fields_]


def test_abi_version(built_lib):
    assert built_lib.b200svs_abi_version() == 2


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only check")
def test_no_device_is_an_error_not_a_fallback(built_lib):
    assert built_lib.b200svs_check_device() != 0
    assert b"CUDA" in built_lib.b200svs_last_error() or b"device" in built_lib.b200svs_last_error()
    a = _lib.ConvArgs()
    assert built_lib.b200svs_conv1d_tc(C.byref(a), None) == -1          # EINVAL before touching a device
    with pytest.raises(ValueError, match="no CPU path"):
        x = torch.zeros(1, 4, 8)
        ops._btc(x, "x")


def test_sass_contains_blackwell_instructions():
    """UTCHMMA = tcgen05.mma, UTMALDG = TMA load, LDTM = tcgen05.ld (B200_PROFILING.md)."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump") or not os.path.exists(_lib.LIB_PATH):
        pytest.skip("cuobjdump or library missing")
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):
        assert mnemonic in sass, mnemonic
    assert "sm_100a" in subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
