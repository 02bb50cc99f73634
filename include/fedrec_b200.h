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
 * workspace: fr_dedup_workspace_ints(N) ints of scratch. No host synchronisation. An id outside [0,N) is not an
 * error: it is left out of uniq/count and its inv entry is 0 (the slot of the smallest id present), so downstream
 * gathers stay in range; validate ids on the host if such input must be rejected.
 * Replaces nothing upstream (model.py:41-61 encodes every duplicate). */
long long fr_dedup_workspace_ints(int N);
int fr_dedup(const int* ids, int K, int N, int* workspace, int* uniq, int* count, int* inv, void* stream);

/* fr_dedup with the scratch built into the library: for hosts that cannot hand over device memory (a plain C/C++
 * caller with no allocator of its own, a process whose memory budget another pool already holds, a request loop
 * where cudaMalloc/cudaFree are banned because they stall every stream). No counterpart in the reference. Same
 * outputs as fr_dedup. Limit: N <= fr_dedup_static_max_rows() (1,048,576 rows; 4,198,404 bytes of device memory
 * per GPU, held for the life of the process); larger N is refused with an error. Allocates and frees nothing.
 * The first call on each device resolves the scratch address and must come before any stream capture on that
 * device. All calls on a device share the scratch: calls that can overlap on the GPU race; order them. Ids outside
 * [0,N) are treated as in fr_dedup (skipped, inv = 0). */
int fr_dedup_static_max_rows(void);
int fr_dedup_static(const int* ids, int K, int N, int* uniq, int* count, int* inv, void* stream);

/* *out = sum(x^2);  x = x*min(1, clip/sqrt(*sq)) + N(0, noise_std^2) with a Philox stream (seed, stream_id).
 * Single-GPU halves of the fused gradient aggregation; client.py:87-89 is the upstream noise. */
int fr_sq_norm(const float* x, long long n, float* out, void* stream);
int fr_clip_noise(float* x, long long n, const float* sq, float clip, float noise_std, unsigned long long seed,
                  unsigned stream_id, void* stream);

/* Host I/O of a step through host-mapped pinned rings, usable from inside a CUDA graph: fr_hostio_fetch copies slot
 * (*step_ctr % slots) of ring_mapped (slots x n ints, n % 4 == 0, device-visible address of pinned host memory) into dst and
 * increments *step_ctr; fr_hostio_publish writes {value bits, *step_ctr} into entry ((*step_ctr - 1) % slots) of
 * out_ring_mapped (slots x 4 uints), sequence number last, for the host to poll. fr_host_mapped_alloc returns both
 * addresses of a zeroed pinned allocation. Replaces the per-batch host/device copies of main.py:59-64,72-79,90. */
/* ring[(*ctr)++ % slots] = %globaltimer in ns (one thread): device-side timestamps around a captured step, for
 * measuring idle time between graph replays without a profiler. No upstream counterpart. */
int fr_timestamp(unsigned long long* ring, int slots, unsigned long long* ctr, void* stream);
int fr_host_mapped_alloc(long long bytes, void** host_out_host, void** dev_out_host);
int fr_host_mapped_free(void* host_ptr_host);
int fr_hostio_fetch(int* dst, const int* ring_mapped, int n, int slots, unsigned long long* step_ctr, void* stream);
int fr_hostio_publish(void* out_ring_mapped, const float* value, int slots, const unsigned long long* step_ctr,
                      void* stream);

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

/* The same attention on the tensor cores (csrc/bert_attention_sm100.cu): floor(128/T) titles share one 128-row
 * tile, S = QK^T and O = PV are tcgen05.mma with TMEM accumulators, P goes through shared memory in the swizzled
 * K-major layout. Same arguments and results as fr_bert_attention; this is the default path. */
int fr_bert_attention_tc(const void* qkv, const int* mask_table, const int* items, const int* count_dev, int n_items,
                         int T, int heads, int head_dim, void* out, void* stream);

/* Text-head additive pooling: alpha = exp(e.w2+b2)/(sum+1e-8) over T, pooled = sum alpha*h. e comes from the GEMM
 * with the tanh epilogue. Replaces attention.py:18-25 as called from encoder.py:28. */
int fr_text_pool(const void* hidden, const void* e, const float* w2, const float* b2, const int* mask_table,
                 const int* items, const int* count_dev, int n_items, int T, int Hq, int hidden_dim, void* pooled,
                 float* alpha_out, void* stream);

int fr_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream);

/* Trunk-output cache helpers (opt-in feature; the frozen trunk's hidden states of a news item never change, so they can
 * stay resident in HBM). fr_cache_select: miss[0..*miss_count) = ids in uniq[0..count) with valid[id] == 0.
 * fr_rows_copy: dst[dst_idx?[i]] = src[src_idx?[i]] for item blocks of rows_per_item rows, i < count; raises valid[] of
 * the destination item when given. No counterpart upstream (model.py:41-61 re-encodes every id every batch). */
int fr_cache_select(const int* uniq, const int* count_dev, int n_items, const int* valid, int* miss, int* miss_count,
                    void* stream);
int fr_rows_copy(const void* src, const int* src_idx, void* dst, const int* dst_idx, const int* count_dev, int n_items,
                 int rows_per_item, int row_bytes, int* valid, void* stream);

/* Backward of fr_text_pool given d_pooled (bf16 [n,768]): writes d_pre = d(tanh input) bf16 [n*T,Hq] (tail rows
 * zeroed to a 64-row boundary), accumulates d_w2 [Hq], d_b2 [1] and the fc1 bias gradient d_b1 [Hq]. da_scratch:
 * n*T floats of scratch (per-token d(logit)). The trunk is frozen, so no d_hidden.
 * Replaces the autograd of attention.py:16-25 inside news_vecs.backward(news_grad), model.py:86-87. */
int fr_text_pool_bwd(const void* hidden, const void* e, const float* alpha, const void* d_pooled, const float* w2,
                     const int* count_dev, int n_items, int T, int Hq, int hidden_dim, void* d_pre, float* da_scratch,
                     float* d_w2, float* d_b2, float* d_b1, void* stream);

/* ---- user encoder + loss, forward and backward (csrc/user_encoder.cu) ---------------------------------- */

/* X[r,:] = bf16(dropout(news_vecs[inv_his[r],:], p)) with a Philox mask (seed); encoder.py:50 + the gather that
 * model.py:112-117 does through numpy. fr_ue_scatter_dx is its backward fused with the news-gradient
 * segment-sum (main.py:20-42) and the optional LDP noise on news-vector gradients (client.py:87-89):
 * d_news[inv_his[r],:] += dX[r,:]*mask + N(0, noise_std^2). */
int fr_ue_gather_dropout(const float* news_vecs, const int* inv_his, void* X, int rows, int D, float p,
                         unsigned long long seed, const unsigned long long* seed_dev, void* stream);
int fr_ue_scatter_dx(const float* dX, const int* inv_his, float* d_news, int rows, int D, float p,
                     unsigned long long seed, float noise_std, unsigned long long noise_seed,
                     const unsigned long long* seed_dev, void* stream);
/* seed_dev (optional, device): added to the seeds at run time, so a step replayed from a CUDA graph still draws a
 * fresh dropout mask / noise every replay; fr_counter_advance bumps it on the stream. */
int fr_counter_advance(unsigned long long* ctr, unsigned long long inc, void* stream);

/* Per-(user, head) attention exp(QK^T/sqrt(dk))*mask/(rowsum+1e-8) · V; qkv bf16 [B*L, 3*H*dk] -> ctx bf16
 * [B*L, H*dk], stats fp32 [B*H*L, 2]. attention.py:37-45,69-81. The backward recomputes the probabilities. */
int fr_ue_mha_fwd(const void* qkv, const int* his_ids, void* ctx, float* stats, int B, int L, int H, int dk,
                  int use_mask, void* stream);
/* The same attention on the tensor cores (csrc/bert_attention_sm100.cu): two users share a 128-row tile, S = QK^T and
 * O = PV are tcgen05.mma with TMEM accumulators (head dim zero-padded to 32), the soft-max keeps the reference's
 * exp(s)/(sum + 1e-8) form. Same arguments, outputs and stats as fr_ue_mha_fwd (L <= 64, dk % 4 == 0, dk <= 32); this is
 * the default path of the training step. attention.py:32-45,69-81. */
int fr_ue_attention_tc(const void* qkv, const int* his_ids, void* ctx, float* stats, int B, int L, int H, int dk,
                       int use_mask, void* stream);
int fr_ue_mha_bwd(const void* qkv, const void* dctx, const float* stats, const int* his_ids, void* dqkv, int B, int L,
                  int H, int dk, int use_mask, void* stream);
/* The same backward on the tensor cores (csrc/bert_attention_sm100.cu): S = QK^T and dP = dO V^T, then dQ = dS K,
 * dK = dS^T Q, dV = P^T dO as tcgen05.mma with TMEM accumulators; dS^T and P^T are written transposed into shared
 * memory so every product uses the operand forms of the forward kernel. Same arguments and results as fr_ue_mha_bwd;
 * the default path of the training step. Autograd of attention.py:32-45,69-81. */
int fr_ue_attention_bwd_tc(const void* qkv, const void* dctx, const float* stats, const int* his_ids, void* dqkv, int B,
                           int L, int H, int dk, int use_mask, void* stream);

/* Additive pooling over the history + candidate scoring + sigmoid + soft-max CE (label 0), and with train != 0 the
 * whole backward down to the GEMM operands: d_ctx_pool, d_pre (bf16), d_w2/d_b2 (+=), candidate gradients
 * scattered into d_news (+=). attention.py:14-26, model.py:121-126. */
int fr_ue_pool_score(const void* ctx, const void* e, const float* w2, const float* b2, const float* news_vecs,
                     const int* inv_cand, const int* his_ids, float* score, float* loss_mean, float* user_vec,
                     void* d_ctx_pool, void* d_pre, float* d_w2, float* d_b2, float* d_news, int B, int L, int C, int D,
                     int Q, int use_mask, int train, float noise_std, unsigned long long noise_seed,
                     const unsigned long long* seed_dev, void* stream);

/* out[n] += sum_m X[m,n] (bias gradients); X bf16 or fp32 with leading dimension ld. */
int fr_colsum(const void* X, int is_bf16, float* out, int M, int N, int ld, const int* m_dev, int m_scale, void* stream);
/* out[n] += sum_m row_w[m] * X[m, n] (row_w fp32 [M]): the transposed GEMV behind d_w2 = e^T · da of the text head
 * (autograd of attention.py:18 upstream). */
int fr_colsum_weighted(const void* X, int is_bf16, const float* row_w, float* out, int M, int N, int ld, const int* m_dev,
                       int m_scale, void* stream);
/* out[c,r] = bf16(in[r,c]): K-major weight copies for the data-gradient GEMMs. */
int fr_transpose_cast(const float* in, void* out, int R, int Ccols, int ld_out, void* stream);

/* ---- federated aggregation over NVLink (csrc/aggregate.cu) -------------------------------------------- */

/* Every rank owns one fp32 buffer of n floats and one signal pad (fr_signal_pad_bytes(), zero-initialised) in
 * peer-mapped memory; bufs_host[i]/pads_host[i] are THIS process's mappings of rank i's. mc_buf: NVLS multicast
 * mapping of the buffers or NULL (plain peer loads/stores). ranks_here = 1 normally (rank = rank_base);
 * ranks_here > 1 runs that many logical ranks as blockIdx.y slices of every launch on the local GPU (single-GPU
 * testing). blocks <= 0 = default grid. Every rank must issue the same sequence of aggregations: the epoch
 * counter lives in the pad, not on the host, so every function below is capturable in a CUDA graph.
 *
 * One call enqueues a short chain of kernels on `stream` (sum of squares, publish, wait, reduce, wait, Adam,
 * commit; see csrc/aggregate.cu). The only place a rank waits for its peers is a one-warp wait kernel with a
 * %globaltimer deadline of timeout_ms (0 = for ever). On expiry it fills the host-mapped record err_mapped
 * (fr_agg_errbuf_alloc: int[16] = {status, rank, peer, phase, have, want, waited_ms}; status 1 = my wait ran
 * out, 2 = a peer gave up first), raises the abort word in every pad so that all ranks stop together, and the
 * rest of the chain does nothing; nothing traps. After such a failure call fr_agg_pad_reset on every rank
 * (no kernel in flight) before the pads are used again. gated != 0: the chain runs only if fr_agg_mark_pending
 * ran on the stream since the previous gated chain (software-pipelined training steps).
 *
 * fr_grad_allreduce: buf_r <- scale * sum_i [ buf_i * min(1, clip/||buf_i||) + N(0, noise_std^2) ] on every rank.
 *   Replaces DDP's gradient all-reduce and Gradient_Averaging_main.py:146; noise: client.py:87-89.
 * fr_param_average: buf_r <- sum_i weights_host[i] * buf_i on every rank (weighted FedAvg + redistribution).
 *   Replaces Parameter_Averaging_main.py:144-148, server.py:37-55,76-77 and client.py:261-264. */
int fr_grad_allreduce(void* const* bufs_host, void* const* pads_host, void* mc_buf, int W, int rank_base,
                      int ranks_here, long long n, float clip, float noise_std, float scale, unsigned long long seed,
                      int gated, int blocks, long long timeout_ms, void* err_mapped, void* stream);
/* fr_grad_allreduce followed on the same stream by one Adam step (torch.optim.Adam defaults) of each local rank's flat
 * parameters on the averaged gradient. params/m/v_host: W pointers (only the local ranks' entries are read). step > 0:
 * 1-based step count for the bias corrections; step == 0: the pad's device-side counter (fr_agg_pad_set_step), which
 * the chain advances itself. W == 1 is allowed (clip + noise + Adam locally, pads_host[0] = any zeroed pad).
 * Replaces DDP all-reduce + model.py:89,93 (two Adam steps). */
int fr_grad_allreduce_adam(void* const* bufs_host, void* const* pads_host, void* mc_buf, int W, int rank_base,
                           int ranks_here, long long n, float clip, float noise_std, float scale,
                           unsigned long long seed, void* const* params_host, void* const* m_host, void* const* v_host,
                           float lr, double b1, double b2, float eps, int step, int gated, int blocks,
                           long long timeout_ms, void* err_mapped, void* stream);
int fr_param_average(void* const* bufs_host, void* const* pads_host, void* mc_buf, const float* weights_host, int W,
                     int rank_base, int ranks_here, long long n, int blocks, long long timeout_ms, void* err_mapped,
                     void* stream);
/* Kernels one chain launches for this configuration (launch accounting). */
int fr_agg_chain_launches(int W, int mode, int nvls, float clip, float noise_std, int uniform_weights, int adam);
/* Arm the next gated chain of up to 4 local pads (one 1-warp kernel; arm == 0 only launches it, for warm-up passes
 * before a graph capture). No upstream counterpart: upstream applies its optimiser inline (model.py:66-70). */
int fr_agg_mark_pending(void* const* pads_host, int count, int arm, void* stream);
/* Synchronous host-side pad access for set-up, recovery and tests: read words [first, first+count), set the Adam
 * step counter, reset a pad to its initial state (keep_step != 0 preserves the step counter). */
int fr_agg_pad_read(const void* pad, unsigned* out_words_host, int first, int count);
int fr_agg_pad_set_step(void* pad, unsigned step);
int fr_agg_pad_reset(void* pad, int keep_step);
int fr_signal_pad_bytes(void);
/* Host-mapped error record (int[16]) the wait kernels report into: *host_out for the CPU, *dev_out for err_mapped. */
int fr_agg_errbuf_alloc(void** host_out_host, void** dev_out_host);
int fr_agg_errbuf_free(void* host_ptr_host);

/* Peer-mappable allocations through CUDA IPC: alloc (zeroed), export a 64-byte handle, import a peer's handle
 * (enables peer access lazily), close an imported mapping, free. Replaces init_process_group's role as the
 * only transport (main.py:144 etc.): here the process group is used for the handle exchange only. */
int fr_symm_alloc(void** ptr_out_host, long long bytes);
int fr_symm_free(void* ptr);
int fr_symm_export(void* ptr, void* handle64_host);
int fr_symm_import(const void* handle64_host, void** ptr_out_host);
int fr_symm_close(void* ptr);

/* ---- tcgen05 GEMM (csrc/gemm_sm100.cu) ------------------------------------------------------------ */

/* out = epi(A[M,K]·W[N,K]^T + bias); A, W bf16 K-major; out bf16 or fp32. epi: 0 bias, 1 bias+GELU(erf),
 * 2 bias+residual(bf16), 3 bias+tanh. block_n 0/64/128/256. m_dev: optional device int, effective
 * M = min(M, *m_dev * m_scale). Replaces every nn.Linear reached from encoder.py:27-29 and attention.py:16,72-74. */
int fr_gemm_bf16(const void* A, const void* W, const float* bias, const void* resid, void* out, int out_fp32, int M,
                 int N, int K, int lda, int ldw, int ldo, int ldr, int epi, int block_n, const int* m_dev, int m_scale,
                 void* stream);

/* fr_gemm_bf16 with its two scheduling choices explicit. cta_group: 0 choose, 1 one CTA per 128 x block_n tile,
 * 2 a CTA pair per 256 x 256 tile (tcgen05 cta_group::2, each CTA stages half of the B tile; bf16 out, block_n 256).
 * tma_epilogue: 0 choose, -1 per-row global stores, 1 swizzled shared-memory boxes + TMA store / TMA residual load. */
int fr_gemm_bf16_ex(const void* A, const void* W, const float* bias, const void* resid, void* out, int out_fp32, int M,
                    int N, int K, int lda, int ldw, int ldo, int ldr, int epi, int block_n, const int* m_dev,
                    int m_scale, int cta_group, int tma_epilogue, void* stream);

/* fr_gemm_bf16 with LayerNorm folded around it, for trunks whose weights are frozen (HF DistilBERT's sa_layer_norm /
 * output_layer_norm reached from encoder.py:27): the normalisation never runs as a kernel of its own. 256-wide tile,
 * bf16 output, epi in {bias, gelu, resid}; every extra pointer is optional.
 *   Row statistics are (sum, sum of squares) pairs, one per 64-column slice of the row: a_stats [M, K/64, 2],
 *   r_stats and stats_out [M, N/64, 2]. The producer stores its slices (no atomics, nothing to zero beforehand), the
 *   consumer adds them in a fixed order, so results are bit-reproducible.
 *   a_stats + col_c [N]:
 *       out = rstd*(A·W^T) - rstd*mu*col_c + bias  ==  LayerNorm(A)·W0^T + b0   for  W = W0∘gamma, col_c = rowsum(W),
 *       bias = b0 + W0·beta.
 *   r_stats, r_gamma [N], r_beta [N] (resid epilogue): adds LayerNorm(resid) instead of resid.
 *   stats_out (resid epilogue): statistics of each output row. */
int fr_gemm_bf16_fold(const void* A, const void* W, const float* bias, const void* resid, void* out, int M, int N, int K,
                      int lda, int ldw, int ldo, int ldr, int epi, const int* m_dev, int m_scale, const float* a_stats,
                      const float* col_c, const float* r_stats, const float* r_gamma, const float* r_beta,
                      float* stats_out, float ln_eps, void* stream);

/* out[M1,N1] (fp32) += A[K,M1]^T · B[K,N1]: weight gradients dW = dY^T X on tcgen05 with MN-major operands and
 * split-K (csrc/gemm_grad_sm100.cu). K_eff = min(K, *k_dev * k_scale); rows of A in [K_eff, round_up(K_eff,64))
 * must be zero. Replaces the autograd weight-gradient GEMMs of every trainable nn.Linear. */
int fr_gemm_grad_bf16(const void* A, const void* B, float* out, int M1, int N1, int K, int lda, int ldb, int ldo,
                      const int* k_dev, int k_scale, int splits, void* stream);

/* ---- host-side batch producer (csrc/host_loader.cu; no CUDA, callable on a CPU-only box) ------------------------- */

/* A client's impression shard (pos [S], his [S, L] 0-padded, negatives as CSR neg_ptr [S+1] / neg_idx) -> packed id
 * batches on a background thread: B impressions per batch in a fresh permutation per pass, K distinct sampled negatives
 * each (all of them + <unk> = 0 when an impression has fewer), `depth` batches prepared ahead. The arrays are copied.
 * fr_loader_next blocks until a batch is ready and copies its B*L + B*(1+K) int32 news rows into out (histories first,
 * then candidates with the clicked news in column 0: the layout fr_hostio_fetch feeds to the step); *epoch_out = pass
 * over the shard (1-based) the batch ended in. Replaces the DataLoader over TrainDataset (dataset.py:69-86, :10-14;
 * main.py:55-64). Returns NULL / non-zero with fr_last_error() set on bad arguments. */
void* fr_loader_create(const int* pos, const int* his, const long long* neg_ptr, const int* neg_idx, long long S, int L,
                       int B, int K, unsigned long long seed, int depth);
int fr_loader_next(void* loader, int* out, long long* epoch_out);
long long fr_loader_batch_ints(void* loader);
void fr_loader_destroy(void* loader);

#ifdef __cplusplus
}
#endif
#endif /* FEDREC_B200_H */
