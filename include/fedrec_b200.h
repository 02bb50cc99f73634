/* fedrec_b200 — C ABI of libfedrec_b200.so (hand-written sm_100a kernels).
 *
 * Conventions: every pointer is a device pointer unless its name ends in _host; `stream` is a
 * cudaStream_t passed as void*; functions return 0 on success, a cudaError_t value or -1 otherwise,
 * and fr_last_error() then describes the failure. Nothing here synchronises the stream.
 * Each declaration cites the reference code whose behaviour it replaces (paths relative to the
 * upstream repository).
 */
#ifndef FEDREC_B200_H
#define FEDREC_B200_H
#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ---------------------------------------------------------------------------------- */
const char* fr_last_error(void);
int fr_version(void);
int fr_device_info(int* out4_host); /* {sm_count, cc_major, cc_minor, max_optin_smem} */

/* ---- small fused ops (csrc/small_ops.cu) -------------------------------------------------------- */

/* score = sigmoid(cand·user); loss = mean_b CE(score_b, label_b); if d_cand != NULL also writes
 * d_cand, d_user (gradients of grad_scale*loss). labels NULL = all zero. fp32, C <= 32.
 * Replaces model.py:121-126 (bmm + sigmoid + CrossEntropyLoss) and its autograd. */
int fr_score_ce(const float* cand, const float* user, const int* labels, float* score, float* loss_rows,
                float* loss_mean, float* d_cand, float* d_user, float grad_scale, int B, int C, int D, void* stream);

/* out[S,D] = 0; out[seg[k],:] += vals[k,:]. Replaces the Python dict loop main.py:20-42. */
int fr_segment_sum(const float* vals, const int* seg, float* out, int K, int S, int D, void* stream);

/* Adam over one flat fp32 buffer (n % 4 == 0). bc_dev (device float[2]) overrides bc1/bc2 when non-NULL;
 * fr_adam_advance increments the device step counter and refreshes bc_dev. Replaces the two
 * torch.optim.Adam steps model.py:22-23,89,93. */
int fr_adam_advance(int* step_dev, float* bc_dev, float b1, float b2, void* stream);
int fr_adam_flat(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2, float eps,
                 float bc1, float bc2, const float* bc_dev, float grad_scale, void* stream);

/* Sorted unique of K ids in [0,N): uniq[K] (tail padded with 0), *count, inv[K] with uniq[inv[k]] == ids[k].
 * workspace: fr_dedup_workspace_ints(N) ints of scratch. No host synchronisation.
 * Replaces nothing upstream (model.py:41-61 encodes every duplicate). */
long long fr_dedup_workspace_ints(int N);
int fr_dedup(const int* ids, int K, int N, int* workspace, int* uniq, int* count, int* inv, void* stream);

/* *out = sum(x^2);  x = x*min(1, clip/sqrt(*sq)) + N(0, noise_std^2) with a Philox stream (seed, stream_id).
 * Single-GPU halves of the fused gradient aggregation; client.py:87-89 is the upstream noise. */
int fr_sq_norm(const float* x, long long n, float* out, void* stream);
int fr_clip_noise(float* x, long long n, const float* sq, float clip, float noise_std, unsigned long long seed,
                  unsigned stream_id, void* stream);

/* ---- DistilBERT trunk / text head, memory-bound parts (csrc/bert_ops.cu) ----------------------------- */

/* out[(n*T+t),:] = LayerNorm(word[tokens[items[n],t]] + pos[t]) in bf16, n < min(n_items, *count_dev).
 * Replaces the HF Embeddings module reached from encoder.py:27. hidden must be 768. */
int fr_bert_embed_ln(const int* tokens, const int* items, const int* count_dev, int n_items, int T, const void* word,
                     const void* pos, const float* gamma, const float* beta, void* out, float eps, int hidden,
                     void* stream);

/* Row LayerNorm, bf16 in/out (the residual was already added by the GEMM epilogue). rows limited to
 * *count_dev * rows_per_item when count_dev != NULL. Replaces sa_layer_norm / output_layer_norm. */
int fr_layernorm_bf16(const void* in, void* out, const float* gamma, const float* beta, int rows, const int* count_dev,
                      int rows_per_item, float eps, int hidden, void* stream);

/* Multi-head self-attention over one title: qkv bf16 [n*T, 3*heads*64] -> out bf16 [n*T, heads*64], q scaled by
 * 1/8, key-padding mask from mask_table[items[n], :]. Replaces HF MultiHeadSelfAttention (encoder.py:27). */
int fr_bert_attention(const void* qkv, const int* mask_table, const int* items, const int* count_dev, int n_items,
                      int T, int heads, int head_dim, void* out, void* stream);

/* Text-head additive pooling: alpha = exp(e.w2+b2)/(sum+1e-8) over T, pooled = sum alpha*h. e comes from the GEMM
 * with the tanh epilogue. Replaces attention.py:18-25 as called from encoder.py:28. */
int fr_text_pool(const void* hidden, const void* e, const float* w2, const float* b2, const int* mask_table,
                 const int* items, const int* count_dev, int n_items, int T, int Hq, int hidden_dim, void* pooled,
                 float* alpha_out, void* stream);

int fr_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream);

/* ---- tcgen05 GEMM (csrc/gemm_sm100.cu) ------------------------------------------------------------ */

/* out = epi(A[M,K]·W[N,K]^T + bias); A, W bf16 K-major; out bf16 or fp32. epi: 0 bias, 1 bias+GELU(erf),
 * 2 bias+residual(bf16), 3 bias+tanh. block_n 0/64/128/256. m_dev: optional device int, effective
 * M = min(M, *m_dev * m_scale). Replaces every nn.Linear reached from encoder.py:27-29 and attention.py:16,72-74. */
int fr_gemm_bf16(const void* A, const void* W, const float* bias, const void* resid, void* out, int out_fp32, int M,
                 int N, int K, int lda, int ldw, int ldo, int ldr, int epi, int block_n, const int* m_dev, int m_scale,
                 void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FEDREC_B200_H */
