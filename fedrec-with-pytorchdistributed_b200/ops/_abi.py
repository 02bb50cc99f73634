"""ctypes signatures of the C ABI in ``include/fedrec_b200.h`` (kept in the same order)."""
from ctypes import (c_char_p, c_float, c_int, c_longlong, c_uint, c_ulonglong, c_void_p, POINTER)

P = c_void_p
I = c_int
F = c_float
LL = c_longlong

SIGNATURES = {
    # runtime.cu
    "fr_last_error": (c_char_p, []),
    "fr_version": (I, []),
    "fr_device_info": (I, [POINTER(c_int)]),
    # small_ops.cu
    "fr_score_ce": (I, [P, P, P, P, P, P, P, P, F, I, I, I, P]),
    "fr_segment_sum": (I, [P, P, P, I, I, I, P]),
    "fr_adam_advance": (I, [P, P, F, F, P]),
    "fr_adam_flat": (I, [P, P, P, P, LL, F, F, F, F, F, F, P, F, P]),
    "fr_dedup_workspace_ints": (LL, [I]),
    "fr_dedup": (I, [P, I, I, P, P, P, P, P]),
    "fr_sq_norm": (I, [P, LL, P, P]),
    "fr_clip_noise": (I, [P, LL, P, F, F, c_ulonglong, c_uint, P]),
    # bert_ops.cu
    "fr_bert_embed_ln": (I, [P, P, P, I, I, P, P, P, P, P, F, I, P]),
    "fr_layernorm_bf16": (I, [P, P, P, P, I, P, I, F, I, P]),
    "fr_bert_attention": (I, [P, P, P, P, I, I, I, I, P, P]),
    "fr_text_pool": (I, [P, P, P, P, P, P, P, I, I, I, I, P, P, P]),
    "fr_cast_f32_to_bf16": (I, [P, P, LL, P]),
    # gemm_sm100.cu
    "fr_gemm_bf16": (I, [P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, P, I, P]),
}
