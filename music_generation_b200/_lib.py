"""ctypes binding of libb200svs.so (C ABI declared in include/b200svs.h).

There is deliberately NO fallback: if the shared library is missing or a call fails the caller
gets an exception.  `build()` compiles it in-tree with nvcc for sm_100a.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200svs.so")
CSRC_DIR = os.path.join(_HERE, "csrc")

ACT_NONE, ACT_RELU, ACT_GELU, ACT_SIGMOID, ACT_TANH, ACT_GATE = range(6)
DT_F16, DT_BF16, DT_F16_SPLIT = 0, 1, 2


class Epilogue(C.Structure):
    _fields_ = [
        ("bias", C.c_void_p), ("gbias", C.c_void_p), ("gbias_ld", C.c_int32), ("act", C.c_int32),
        ("alpha", C.c_float), ("scale", C.c_float),
        ("res1", C.c_void_p), ("res1_ld", C.c_int32),
        ("res2", C.c_void_p), ("res2_ld", C.c_int32),
        ("lens", C.c_void_p), ("lens_div", C.c_int32),
        ("out32", C.c_void_p), ("out32_ld", C.c_int32),
        ("out16", C.c_void_p), ("out16_ld", C.c_int32), ("out16_dtype", C.c_int32),
        ("post_lrelu", C.c_int32), ("post_slope", C.c_float), ("n_valid", C.c_int32),
        ("res16", C.c_void_p), ("res16_ld", C.c_int32),
        ("out16r", C.c_void_p), ("out16r_ld", C.c_int32),
        ("res2_is16", C.c_int32),
        ("res16_unlrelu", C.c_float),
    ]


class ConvArgs(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("x_dtype", C.c_int32),
        ("B", C.c_int32), ("T", C.c_int32), ("Cin", C.c_int32), ("x_ld", C.c_int32),
        ("w", C.c_void_p),
        ("ntaps", C.c_int32), ("Cout_pad", C.c_int32), ("Cin_pad", C.c_int32), ("dil", C.c_int32),
        ("pad_left", C.c_int32), ("block_n", C.c_int32),
        ("ep", Epilogue),
    ]


_i32, _f32, _vp = C.c_int32, C.c_float, C.c_void_p

# name -> argtypes (all return int); mirrors include/b200svs.h one to one
PROTOTYPES = {
    "b200svs_abi_version": [],
    "b200svs_check_device": [],
    "b200svs_saturation_events": [C.POINTER(C.c_uint32), _i32, _vp],
    "b200svs_conv1d_tc": [C.POINTER(ConvArgs), _vp],
    "b200svs_conv1d_tc_set_trace": [_vp, _i32],
    "b200svs_song_assemble": [_vp, C.c_int64, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp],
    "b200svs_song_assemble_noalloc": [_vp, C.c_int64, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp],
    "b200svs_conv1d_f32": [C.POINTER(ConvArgs), _vp],
    "b200svs_attention_f32": [_vp, _i32, _vp, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp],
    "b200svs_attn_prep": [_vp, _i32, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp],
    "b200svs_attention_tc": [_vp, _vp, _vp, _vp, _i32, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _i32, _i32,
                             _i32, _i32, _i32, _vp],
    "b200svs_layernorm": [_vp, _i32, _vp, _i32, _vp, _vp, _f32, _vp, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _vp],
    "b200svs_text_embed": [_vp] * 12 + [_i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp],
    "b200svs_duration_quantize": [_vp, _vp, _vp, _i32, _vp],
    "b200svs_lr_scan": [_vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp],
    "b200svs_lr_gather": [_vp, _i32, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _i32, _vp, _vp],
    "b200svs_add_pos_mask": [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _vp],
    "b200svs_f0_embed": [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp],
    "b200svs_sample_prior": [_vp, _i32, _vp, _f32, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp],
    "b200svs_conv_post_tanh": [_vp, _i32, _i32, _vp, _i32, _i32, _i32, _i32, _vp, _vp],
    "b200svs_bct_to_btc": [_vp, _i32, _i32, _i32, _vp, _i32, _vp, _i32, _i32, _vp, _vp],
    "b200svs_btc_to_bct": [_vp, _i32, _i32, _i32, _i32, _vp, _vp],
    "b200svs_probe_rowshift": [_vp, _vp, _vp, _i32, _i32, _i32, _vp],
}

_lib = None


def build(verbose=False):
    """Compile every CUDA source for sm_100a into music_generation_b200/lib/libb200svs.so."""
    r = subprocess.run(["make", "-C", CSRC_DIR, "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libb200svs.so failed (see output above)")
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU / PyTorch fallback for the synthesis kernels)")
        h = C.CDLL(LIB_PATH)
        h.b200svs_last_error.restype = C.c_char_p
        h.b200svs_last_error.argtypes = []
        for name, args in PROTOTYPES.items():
            fn = getattr(h, name)          # AttributeError if the .so does not export it
            fn.restype = C.c_int
            fn.argtypes = args
        _lib = h
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().b200svs_last_error().decode(errors="replace")
        raise RuntimeError(f"b200svs {what} failed (code {rc}): {msg}")


# number of C-ABI compute calls issued by this process; each one launches exactly one kernel
LAUNCHES = 0


# bench.py's instrumented pass: when PROFILE is a list every call appends
# (entry point, tag, algorithmic flops, start event, end event, algorithmic HBM bytes); NEXT_META is set by ops.conv1d
PROFILE = None
NEXT_META = None


def call(name, *args):
    global LAUNCHES, NEXT_META
    LAUNCHES += 1
    if PROFILE is None:
        check(getattr(lib(), name)(*args), name)
        return
    import torch
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    check(getattr(lib(), name)(*args), name)
    ev1.record()
    tag, flops, nbytes = NEXT_META or ("", 0.0, 0.0)
    NEXT_META = None
    PROFILE.append((name.replace("b200svs_", ""), tag, flops, ev0, ev1, nbytes))
