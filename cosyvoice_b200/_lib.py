"""ctypes binding of libcvb200.so (include/cvb200.h).

The shared library is the product: if it is missing or fails to load this module raises --
there is no PyTorch/CPU fallback anywhere in the package.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CVB200_LIB: another build of the same library (A/B measurements of kernel changes on one GPU box)
LIB_PATH = os.environ.get("CVB200_LIB") or os.path.join(_HERE, "lib", "libcvb200.so")

F32, BF16 = 0, 1
ENGINE_SIMT, ENGINE_TCGEN05 = 0, 1
MAX_SEGS = 32
ACT_NONE, ACT_GELU, ACT_ELU, ACT_SNAKE, ACT_LRELU, ACT_MISH, ACT_TANH, ACT_SILU = range(8)
RES_NONE, RES_F32, RES_ACT = range(3)
OUT1_NONE, OUT1_COPY, OUT1_SNAKE, OUT1_LRELU = range(4)
ATTN_TCGEN05, ATTN_MMA_SYNC, ATTN_TCGEN05_V2 = 0, 1, 2


class Epilogue(C.Structure):
    _fields_ = [
        ("bias", C.c_void_p), ("seq_bias", C.c_void_p), ("act_alpha", C.c_void_p), ("res", C.c_void_p),
        ("out0", C.c_void_p), ("out1", C.c_void_p), ("out1_alpha", C.c_void_p), ("gn_stats", C.c_void_p),
        ("seq_bias_ld", C.c_int32), ("act", C.c_int32), ("act_slope", C.c_float), ("res_kind", C.c_int32),
        ("ldr", C.c_int32), ("res_dt", C.c_int32), ("out_scale", C.c_float), ("accumulate", C.c_int32),
        ("ldo0", C.c_int32), ("out0_dt", C.c_int32), ("phase_col", C.c_int32), ("out1_mode", C.c_int32),
        ("ldo1", C.c_int32), ("out1_dt", C.c_int32), ("out1_slope", C.c_float), ("gn_groups", C.c_int32),
        ("row_stats", C.c_void_p), ("row_u", C.c_void_p),
        ("row_parts", C.c_void_p), ("row_parts_out", C.c_void_p), ("row_eps", C.c_float), ("reserved0", C.c_int32),
    ]


class Gemm(C.Structure):
    _fields_ = [
        ("A", C.c_void_p), ("W", C.c_void_p), ("tile_seq", C.c_void_p), ("tile_row0", C.c_void_p),
        ("seq_len", C.c_void_p), ("out_limit", C.c_void_p),
        ("lda", C.c_int32), ("a_rows", C.c_int32), ("a_cols", C.c_int32),
        ("n_pad", C.c_int32), ("N", C.c_int32), ("n_phase", C.c_int32), ("n_seg", C.c_int32), ("seg_len", C.c_int32),
        ("dt", C.c_int32), ("m_tiles", C.c_int32), ("bn", C.c_int32),
        ("seg_row", C.c_int16 * MAX_SEGS), ("seg_col", C.c_int16 * MAX_SEGS),
        ("epi", Epilogue),
    ]


class Layout(C.Structure):
    _fields_ = [
        ("tile_seq", C.c_void_p), ("tile_row0", C.c_void_p), ("seq_len", C.c_void_p), ("seq_row0", C.c_void_p),
        ("n_seq", C.c_int32), ("rows", C.c_int32), ("max_len", C.c_int32),
    ]


# name -> argtypes; every function returns int status except the three listed in _SPECIAL
_P, _I, _F, _L = C.c_void_p, C.c_int, C.c_float, C.c_int64
_LAY = C.POINTER(Layout)
SIGNATURES = {
    "cvb200_tapgemm": [C.POINTER(Gemm), _I, _P],
    "cvb200_ffn_fused": [_P, _I, _P, _P, _P, _P, _P, _I, _P, _I, _LAY, _P],
    "cvb200_block_ffn": [_P, _I, _F, _P, _P, _P, _P, _P, _P, _I, _P, _LAY, _P],
    "cvb200_block_out_ffn": [_P, _I, _P, _I, _P, _P, _F, _P, _P, _P, _P, _P, _P, _I, _P, _LAY, _P],
    "cvb200_split3_bf16": [_P, _I, _I, _I, _P, _I, _L, _P],
    "cvb200_fade_in_out": [_P, _L, _I, _P, _L, _I, _P, _I, _I, _P],
    "cvb200_interp_linear": [_P, _L, _I, _P, _L, _I, _I, _P],
    "cvb200_pcm16": [_P, _P, _F, _L, _P],
    "cvb200_layernorm": [_P, _I, _P, _I, _I, _P, _P, _I, _I, _F, _P],
    "cvb200_layernorm_act": [_P, _I, _I, _P, _I, _I, _P, _P, _I, _I, _F, _P],
    "cvb200_norm_mish": [_P, _I, _I, _P, _I, _I, _P, _P, _P, _P, _I, _LAY, _I, _F, _P],
    "cvb200_attention": [_P, _I, _P, _I, _I, _LAY, _I, _F, _P],
    "cvb200_attention_ex": [_P, _I, _P, _I, _I, _LAY, _I, _F, _I, _P],
    "cvb200_time_sinusoid": [_P, _I, _I, _F, _P, _P],
    "cvb200_small_linear": [_P, _I, _I, _P, _P, _I, _I, _I, _P, _P],
    "cvb200_pack_cfg_input": [_P, _P, _P, _P, _P, _I, _LAY, _P],
    "cvb200_euler_cfg": [_P, _P, _L, _F, _F, _P],
    "cvb200_rowdot_abs": [_P, _I, _I, _P, _P, _I, _P, _I, _P],
    "cvb200_sine_source": [_P, _LAY, _I, _F, _P, _P, _P, _P, _P, _P, _L, _P],
    "cvb200_sine_source_noalloc": [_P, _LAY, _I, _F, _P, _P, _P, _P, _P, _L, _P],
    "cvb200_stft16": [_P, _L, _P, _I, _I, _LAY, _P],
    "cvb200_istft16": [_P, _I, _LAY, _P, _L, _F, _P],
    "cvb200_snake": [_P, _I, _I, _P, _I, _I, _P, _I, _LAY, _P],
    "cvb200_reflect_row0": [_P, _I, _I, _I, _LAY, _P],
    "cvb200_anti_alias_act_nct": [_P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P],
    "cvb200_anti_alias_act_rows": [_P, _I, _I, _P, _I, _I, _P, _P, _P, _I, _LAY, _P],
    "cvb200_nct_to_rows": [_P, _L, _L, _I, _P, _I, _I, _LAY, _P],
    "cvb200_rows_to_nct": [_P, _I, _I, _P, _L, _L, _LAY, _P],
    "cvb200_embed_tokens": [_P, _I, _P, _I, _I, _P, _I, _I, _LAY, _P],
    "cvb200_layernorm_seq": [_P, _I, _P, _I, _I, _P, _P, _LAY, _I, _F, _P],
    "cvb200_relpos_attention": [_P, _I, _P, _I, _I, _I, _P, _P, _P, _I, _I, _LAY, _I, _F, _I, _P],
    "cvb200_groupnorm1_mish": [_P, _I, _P, _I, _I, _P, _P, _I, _LAY, _F, _P],
    "cvb200_interp_rows": [_P, _I, _LAY, _P, _I, _I, _LAY, _P, _I, _P],
    "cvb200_l2_normalize": [_P, _P, _I, _I, _P],
}
_SPECIAL = {
    "cvb200_last_error": ([], C.c_char_p),
    "cvb200_version": ([], C.c_int),
    "cvb200_launch_count": ([], C.c_int64),
    "cvb200_set_pdl": ([C.c_int], C.c_int),
}

_lib = None


def load():
    """Load libcvb200.so (once).  Raises RuntimeError when it has not been built:
    run `python -c "import __graft_entry__ as g; g.build()"` (nvcc, sm_100a)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: the CUDA library must be built first (__graft_entry__.build()); "
            "cosyvoice_b200 has no fallback path")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = C.c_int
    for name, (argtypes, restype) in _SPECIAL.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype
    _lib = lib
    return lib


def exported_symbols():
    return list(SIGNATURES) + list(_SPECIAL)


def check(status: int, what: str = ""):
    if status != 0:
        msg = load().cvb200_last_error()
        raise RuntimeError(f"cvb200 {what} failed (status {status}): {msg.decode() if msg else '?'}")


def launch_count() -> int:
    return int(load().cvb200_launch_count())
