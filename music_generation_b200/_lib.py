"""ctypes binding of libb200svs.so (C ABI declared in include/b200svs.h).

There is deliberately NO fallback: if the shared library is missing or a call fails the caller
gets an exception.  `build()` compiles it in-tree with nvcc for sm_100a.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200SVS_LIB") or os.path.join(_HERE, "lib", "libb200svs.so")      # override: A/B builds
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

# ---- stage-level API (include/b200svs.h "Stage-level entry points")
MAX_STAGES, MAX_KERNELS, MAX_PAIRS, MAX_WN_LAYERS = 6, 4, 4, 16


class ConvW(C.Structure):
    _fields_ = [("w", _vp), ("bias", _vp),
                ("ntaps", _i32), ("dil", _i32), ("pad_left", _i32), ("cin", _i32), ("cin_pad", _i32), ("cout", _i32),
                ("cout_pad", _i32), ("block_n", _i32), ("gate", _i32), ("f32", _i32), ("fold", _i32), ("n_out", _i32),
                ("out_scale", _f32)]


class GeneratorW(C.Structure):
    _fields_ = [("n_stages", _i32), ("n_kernels", _i32), ("n_pairs", _i32), ("dtype", _i32),
                ("up_rate", _i32 * MAX_STAGES), ("channels", _i32 * MAX_STAGES), ("pre_channels", _i32),
                ("chunk_frames", _i32), ("wide_chunk_frames", _i32), ("wide_stages", _i32), ("fuse_pairs", _i32), ("lrelu_slope", _f32),
                ("pre", ConvW), ("ups", ConvW * MAX_STAGES),
                ("rb", ((ConvW * 2) * MAX_PAIRS) * MAX_KERNELS * MAX_STAGES),
                ("post_w", _vp), ("post_taps", _i32)]


class WnW(C.Structure):
    _fields_ = [("n_layers", _i32), ("hidden", _i32), ("dtype", _i32),
                ("in_", ConvW * MAX_WN_LAYERS), ("rs", ConvW * MAX_WN_LAYERS), ("skip_cat", ConvW)]


class EncoderLayerW(C.Structure):
    _fields_ = [("hidden", _i32), ("filter", _i32), ("heads", _i32), ("window", _i32), ("mode", _i32), ("dtype", _i32),
                ("qkv", ConvW), ("o", ConvW), ("ffn1", ConvW), ("ffn2", ConvW),
                ("rel_k", _vp), ("rel_v", _vp), ("ln1_g", _vp), ("ln1_b", _vp), ("ln2_g", _vp), ("ln2_b", _vp)]


# name -> argtypes (all return int); mirrors include/b200svs.h one to one
PROTOTYPES = {
    "b200svs_abi_version": [],
    "b200svs_check_device": [],
    "b200svs_abi_struct_size": [_i32],
    "b200svs_saturation_events": [C.POINTER(C.c_uint32), _i32, _vp],
    "b200svs_conv1d_tc": [C.POINTER(ConvArgs), _vp],
    "b200svs_conv1d_tc_set_trace": [_vp, _i32],
    "b200svs_song_assemble": [_vp, C.c_int64, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp],
    "b200svs_song_assemble_noalloc": [_vp, C.c_int64, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp],
    "b200svs_generator_fwd": [C.POINTER(GeneratorW), _vp, _i32, _i32, _i32, _vp, _i32, _vp, _vp, C.c_size_t, _vp],
    "b200svs_wn_stack_fwd": [C.POINTER(WnW), _vp, _vp, _vp, _i32, _i32, _vp, _i32, _vp, _vp, C.c_size_t, _vp],
    "b200svs_encoder_fwd": [C.POINTER(EncoderLayerW), _i32, _vp, _vp, _vp, _i32, _i32, _vp, C.c_size_t, _vp],
    "b200svs_resblock_pair": [C.POINTER(ConvW), C.POINTER(ConvW), _vp, _i32, _i32, _vp, _vp, _i32, _f32, _f32, _f32, _vp],
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
    "b200svs_conv1d_ws": [_vp, _vp, _i32, _vp, _i32, _i32, _vp, _f32, _vp, _vp, _i32, _f32, _f32, _vp],
    "b200svs_resblock_pair_set_trace": [_vp, _i32],
    "b200svs_attention_set_trace": [_vp, _i32],
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
        h.b200svs_launch_count.restype = C.c_int64
        h.b200svs_launch_count.argtypes = []
        for name, wt in (("b200svs_generator_workspace_bytes", GeneratorW), ("b200svs_wn_workspace_bytes", WnW),
                         ("b200svs_encoder_workspace_bytes", EncoderLayerW)):
            fn = getattr(h, name)
            fn.restype = C.c_size_t
            fn.argtypes = [C.POINTER(wt), _i32, _i32]
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


def launches():
    """Kernels launched through the library by this process (b200svs_launch_count)."""
    return int(lib().b200svs_launch_count())



# bench.py's instrumented pass: when PROFILE is a list every call appends
# (entry point, tag, algorithmic flops, start event, end event, algorithmic HBM bytes); NEXT_META is set by ops.conv1d
PROFILE = None
NEXT_META = None


def call(name, *args):
    global NEXT_META
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
