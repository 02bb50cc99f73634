"""ctypes signatures of the C ABI in ``include/fedrec_b200.h`` (kept in the same order)."""
from ctypes import (c_char_p, c_double, c_float, c_int, c_longlong, c_uint, c_ulonglong, c_void_p, POINTER)

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
    "fr_dedup_static_max_rows": (I, []),
    "fr_dedup_static": (I, [P, I, I, P, P, P, P]),
    "fr_sq_norm": (I, [P, LL, P, P]),
    "fr_clip_noise": (I, [P, LL, P, F, F, c_ulonglong, c_uint, P]),
    "fr_timestamp": (I, [P, I, P, P]),
    "fr_host_mapped_alloc": (I, [LL, P, P]),
    "fr_host_mapped_free": (I, [P]),
    "fr_hostio_fetch": (I, [P, P, I, I, P, P]),
    "fr_hostio_publish": (I, [P, P, I, P, P]),
    # bert_ops.cu
    "fr_bert_embed_ln": (I, [P, P, P, I, I, P, P, P, P, P, F, I, P]),
    "fr_layernorm_bf16": (I, [P, P, P, P, I, P, I, F, I, P]),
    "fr_bert_attention": (I, [P, P, P, P, I, I, I, I, P, P]),
    "fr_bert_attention_tc": (I, [P, P, P, P, I, I, I, I, P, P]),
    "fr_text_pool": (I, [P, P, P, P, P, P, P, I, I, I, I, P, P, P]),
    "fr_cast_f32_to_bf16": (I, [P, P, LL, P]),
    "fr_cache_select": (I, [P, P, I, P, P, P, P]),
    "fr_rows_copy": (I, [P, P, P, P, P, I, I, I, P, P]),
    "fr_text_pool_bwd": (I, [P, P, P, P, P, P, I, I, I, I, P, P, P, P, P, P]),
    # user_encoder.cu
    "fr_ue_gather_dropout": (I, [P, P, P, I, I, F, c_ulonglong, P, P]),
    "fr_counter_advance": (I, [P, c_ulonglong, P]),
    "fr_ue_scatter_dx": (I, [P, P, P, I, I, F, c_ulonglong, F, c_ulonglong, P, P]),
    "fr_ue_mha_fwd": (I, [P, P, P, P, I, I, I, I, I, P]),
    "fr_ue_attention_tc": (I, [P, P, P, P, I, I, I, I, I, P]),
    "fr_ue_attention_bwd_tc": (I, [P, P, P, P, P, I, I, I, I, I, P]),
    "fr_ue_mha_bwd": (I, [P, P, P, P, P, I, I, I, I, I, P]),
    "fr_ue_pool_score": (I, [P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, I, F, c_ulonglong, P, P]),
    "fr_colsum": (I, [P, I, P, I, I, I, P, I, P]),
    "fr_colsum_weighted": (I, [P, I, P, P, I, I, I, P, I, P]),
    "fr_transpose_cast": (I, [P, P, I, I, I, P]),
    # gemm_grad_sm100.cu
    "fr_gemm_grad_bf16": (I, [P, P, P, I, I, I, I, I, I, P, I, I, P]),
    # aggregate.cu
    "fr_agg_chain_launches": (I, [I, I, I, F, F, I, I]),
    "fr_grad_allreduce": (I, [P, P, P, I, I, I, LL, F, F, F, c_ulonglong, I, I, LL, P, P]),
    "fr_grad_allreduce_adam": (I, [P, P, P, I, I, I, LL, F, F, F, c_ulonglong, P, P, P, F, c_double, c_double, F, I, I, I,
                                   LL, P, P]),
    "fr_param_average": (I, [P, P, P, P, I, I, I, LL, I, LL, P, P]),
    "fr_agg_mark_pending": (I, [P, I, I, P]),
    "fr_agg_pad_read": (I, [P, P, I, I]),
    "fr_agg_pad_set_step": (I, [P, c_uint]),
    "fr_agg_pad_reset": (I, [P, I]),
    "fr_signal_pad_bytes": (I, []),
    "fr_agg_errbuf_alloc": (I, [P, P]),
    "fr_agg_errbuf_free": (I, [P]),
    "fr_symm_alloc": (I, [P, LL]),
    "fr_symm_free": (I, [P]),
    "fr_symm_export": (I, [P, P]),
    "fr_symm_import": (I, [P, P]),
    "fr_symm_close": (I, [P]),
    # host_loader.cu (host code only)
    "fr_loader_create": (P, [P, P, P, P, LL, I, I, I, c_ulonglong, I]),
    "fr_loader_next": (I, [P, P, P]),
    "fr_loader_batch_ints": (LL, [P]),
    "fr_loader_destroy": (None, [P]),
    # gemm_sm100.cu
    "fr_gemm_bf16": (I, [P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, P, I, P]),
    "fr_gemm_bf16_ex": (I, [P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, P, I, I, I, P]),
    "fr_gemm_bf16_fold": (I, [P, P, P, P, P, I, I, I, I, I, I, I, I, P, I, P, P, P, P, P, P, F, P]),
}
