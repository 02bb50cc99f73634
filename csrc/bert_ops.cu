// Memory-bound pieces of the DistilBERT trunk and of the text head:
//   N9  embedding gather (word + position) + LayerNorm                     (HF DistilBERT Embeddings)
//   -   LayerNorm over rows that already hold x + residual (the GEMM epilogue added it)
//   N8  12-head self-attention over one title (T <= 64 tokens), key-padding mask
//   N10 additive-attention pooling of the text head: alpha = exp(e·w2 + b2) / (sum + 1e-8), sum alpha*h
// Reference call sites: encoder.py:27 (DistilBertModel), encoder.py:28 → attention.py:14-26.
// All activations are bf16, statistics and accumulation fp32. Row counts may come from device
// memory (count_dev * rows_per_item) so a captured graph follows the per-step de-dup count.
#include "common.cuh"

namespace fr {

constexpr int kHid = 768;            // DistilBERT width; one warp owns a row: 24 elements per lane
constexpr int kPerLane = kHid / 32;  // 24 = 3 x 8 bf16 (16-byte vectors)

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float* v) {
  const uint4 q = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float* v) {
  uint4 q;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = q;
}

// gamma/beta slices of a lane (its 24 columns), loaded once per warp by the persistent kernels
__device__ __forceinline__ void load_affine(const float* __restrict__ gamma, const float* __restrict__ beta, int lane,
                                            float* g, float* b) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int col = c * 256 + lane * 8;
#pragma unroll
    for (int i = 0; i < 8; i += 4) {
      const float4 gv = *reinterpret_cast<const float4*>(gamma + col + i);
      const float4 bv = *reinterpret_cast<const float4*>(beta + col + i);
      g[c * 8 + i] = gv.x; g[c * 8 + i + 1] = gv.y; g[c * 8 + i + 2] = gv.z; g[c * 8 + i + 3] = gv.w;
      b[c * 8 + i] = bv.x; b[c * 8 + i + 1] = bv.y; b[c * 8 + i + 2] = bv.z; b[c * 8 + i + 3] = bv.w;
    }
  }
}
__device__ __forceinline__ void warp_layernorm_reg(float* x, const float* g, const float* b, float eps) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kPerLane; ++i) s += x[i];
  const float mean = warp_sum(s) * (1.f / kHid);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kPerLane; ++i) {
    const float d = x[i] - mean;
    q += d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / kHid) + eps);
#pragma unroll
  for (int i = 0; i < kPerLane; ++i) x[i] = (x[i] - mean) * rstd * g[i] + b[i];
}

// x[24] per lane → normalised in place. Two-pass (mean, then centred variance) in registers.
__device__ __forceinline__ void warp_layernorm(float* x, const float* __restrict__ gamma, const float* __restrict__ beta,
                                               int lane, float eps) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kPerLane; ++i) s += x[i];
  const float mean = warp_sum(s) * (1.f / kHid);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kPerLane; ++i) {
    const float d = x[i] - mean;
    q += d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / kHid) + eps);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int col = c * 256 + lane * 8;
#pragma unroll
    for (int i = 0; i < 8; ++i) x[c * 8 + i] = (x[c * 8 + i] - mean) * rstd * gamma[col + i] + beta[col + i];
  }
}

// out[(n*T + t), :] = LN(word[tokens[items[n], t]] + pos[t])
__global__ void __launch_bounds__(256) embed_ln_kernel(const int* __restrict__ tokens, const int* __restrict__ items,
                                                       const int* __restrict__ count_dev, int n_items, int T,
                                                       const __nv_bfloat16* __restrict__ word,
                                                       const __nv_bfloat16* __restrict__ pos,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       __nv_bfloat16* __restrict__ out, float eps) {
  const int lane = threadIdx.x & 31;
  const int n_eff = count_dev ? min(n_items, *count_dev) : n_items;
  const int rows = n_eff * T;
  float g[kPerLane], b[kPerLane];
  load_affine(gamma, beta, lane, g, b);
  for (int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += gridDim.x * (blockDim.x >> 5)) {
    const int n = row / T, t = row - n * T;
    const int item = items ? items[n] : n;
    const int tok = tokens[(size_t)item * T + t];
    float x[kPerLane], p[8];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const int col = c * 256 + lane * 8;
      load8(word + (size_t)tok * kHid + col, x + c * 8);
      load8(pos + (size_t)t * kHid + col, p);
#pragma unroll
      for (int i = 0; i < 8; ++i) x[c * 8 + i] += p[i];
    }
    warp_layernorm_reg(x, g, b, eps);
#pragma unroll
    for (int c = 0; c < 3; ++c) store8(out + (size_t)row * kHid + c * 256 + lane * 8, x + c * 8);
  }
}

// Persistent: a few fat blocks, every warp walks rows with two independent rows in flight so the
// load latency of one row hides behind the reductions of the other.
__global__ void __launch_bounds__(512) layernorm_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                        const float* __restrict__ gamma, const float* __restrict__ beta,
                                                        int rows_max, const int* __restrict__ count_dev, int rows_per_item,
                                                        float eps) {
  const int lane = threadIdx.x & 31;
  pdl_launch();                                      // a following GEMM may start its prologue
  pdl_wait();                                        // launched with the PDL attribute: the producer has finished from here on
  int rows = rows_max;
  if (count_dev) rows = min(rows_max, *count_dev * rows_per_item);
  const int nwarps = gridDim.x * (blockDim.x >> 5);
  const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  float g[kPerLane], b[kPerLane];
  load_affine(gamma, beta, lane, g, b);
  for (int row = w; row < rows; row += 2 * nwarps) {
    const int row2 = row + nwarps;
    const bool has2 = row2 < rows;
    float x[kPerLane], y[kPerLane];
#pragma unroll
    for (int c = 0; c < 3; ++c) load8(in + (size_t)row * kHid + c * 256 + lane * 8, x + c * 8);
    if (has2) {
#pragma unroll
      for (int c = 0; c < 3; ++c) load8(in + (size_t)row2 * kHid + c * 256 + lane * 8, y + c * 8);
    }
    warp_layernorm_reg(x, g, b, eps);
#pragma unroll
    for (int c = 0; c < 3; ++c) store8(out + (size_t)row * kHid + c * 256 + lane * 8, x + c * 8);
    if (has2) {
      warp_layernorm_reg(y, g, b, eps);
#pragma unroll
      for (int c = 0; c < 3; ++c) store8(out + (size_t)row2 * kHid + c * 256 + lane * 8, y + c * 8);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Self-attention over one title. qkv [rows, 3*768] = (q | k | v), heads of 64. One warp per
// (item, head): K and V of that head staged in shared memory, each lane owns query rows
// lane and lane+32 and runs an online soft-max over the keys. q is scaled by 1/8 (HF scales q).
// Masked keys get -inf (HF: masked_fill(finfo.min)); an item whose mask is all zero attends
// uniformly, exactly as soft-max over a constant row does.
// ---------------------------------------------------------------------------------------------
constexpr int kHeadDim = 64;
constexpr int kMaxT = 64;
constexpr int kAttnWarps = 2;

__global__ void __launch_bounds__(kAttnWarps * 32) bert_attention_kernel(
    const __nv_bfloat16* __restrict__ qkv, const int* __restrict__ mask_table, const int* __restrict__ items,
    const int* __restrict__ count_dev, int n_items, int T, int heads, __nv_bfloat16* __restrict__ out) {
  __shared__ __align__(16) __nv_bfloat16 sK[kAttnWarps][kMaxT][kHeadDim];
  __shared__ __align__(16) __nv_bfloat16 sV[kAttnWarps][kMaxT][kHeadDim];
  __shared__ float sBias[kAttnWarps][kMaxT];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_eff = count_dev ? min(n_items, *count_dev) : n_items;
  const int total = n_eff * heads;
  const int ld = 3 * heads * kHeadDim;
  for (int job = blockIdx.x * kAttnWarps + w; job < total; job += gridDim.x * kAttnWarps) {
    const int n = job / heads, h = job - n * heads;
    const int item = items ? items[n] : n;
    const size_t row0 = (size_t)n * T;
    // stage K, V: T rows x 64 bf16 = T x 8 uint4
    for (int i = lane; i < T * 8; i += 32) {
      const int r = i >> 3, c = i & 7;
      const __nv_bfloat16* base = qkv + (row0 + r) * ld + h * kHeadDim + c * 8;
      *reinterpret_cast<uint4*>(&sK[w][r][c * 8]) = *reinterpret_cast<const uint4*>(base + heads * kHeadDim);
      *reinterpret_cast<uint4*>(&sV[w][r][c * 8]) = *reinterpret_cast<const uint4*>(base + 2 * heads * kHeadDim);
    }
    bool any = false;
    for (int j = lane; j < T; j += 32) {
      const int mk = mask_table ? mask_table[(size_t)item * T + j] : 1;
      sBias[w][j] = mk ? 0.f : -INFINITY;
      any |= (mk != 0);
    }
    any = __any_sync(0xffffffffu, any);
    __syncwarp();
    for (int qr = lane; qr < T; qr += 32) {
      float q[kHeadDim];
      const __nv_bfloat16* qp = qkv + (row0 + qr) * ld + h * kHeadDim;
#pragma unroll
      for (int c = 0; c < 8; ++c) load8(qp + c * 8, q + c * 8);
#pragma unroll
      for (int d = 0; d < kHeadDim; ++d) q[d] *= 0.125f;
      float m = -INFINITY, l = 0.f;
      float acc[kHeadDim];
#pragma unroll
      for (int d = 0; d < kHeadDim; ++d) acc[d] = 0.f;
      for (int j = 0; j < T; ++j) {
        const float bias = any ? sBias[w][j] : 0.f;
        if (bias == -INFINITY) continue;            // uniform across the warp (same item)
        float s = 0.f;
        const __nv_bfloat162* kp = reinterpret_cast<const __nv_bfloat162*>(&sK[w][j][0]);
#pragma unroll
        for (int d = 0; d < kHeadDim / 2; ++d) {
          const float2 kf = __bfloat1622float2(kp[d]);
          s = fmaf(q[2 * d], kf.x, s);
          s = fmaf(q[2 * d + 1], kf.y, s);
        }
        if (!any) s = 0.f;                          // fully masked title: every score is the same constant in HF
        const float m_new = fmaxf(m, s);
        const float corr = __expf(m - m_new);
        const float p = __expf(s - m_new);
        l = l * corr + p;
        const __nv_bfloat162* vp = reinterpret_cast<const __nv_bfloat162*>(&sV[w][j][0]);
#pragma unroll
        for (int d = 0; d < kHeadDim / 2; ++d) {
          const float2 vf = __bfloat1622float2(vp[d]);
          acc[2 * d] = fmaf(p, vf.x, acc[2 * d] * corr);
          acc[2 * d + 1] = fmaf(p, vf.y, acc[2 * d + 1] * corr);
        }
        m = m_new;
      }
      const float inv = 1.f / l;
#pragma unroll
      for (int d = 0; d < kHeadDim; ++d) acc[d] *= inv;
      __nv_bfloat16* op = out + (row0 + qr) * (size_t)(heads * kHeadDim) + h * kHeadDim;
#pragma unroll
      for (int c = 0; c < 8; ++c) store8(op + c * 8, acc + c * 8);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------
// Text-head pooling (forward). e = tanh(h·W1^T + b1) comes from the GEMM (tanh epilogue), bf16 [rows, Hq].
// One block (8 warps) per item, tokens dealt round-robin to the warps so a whole title's rows are in flight at once
// (one warp per item left ~6 warps per SM and the kernel latency-bound):
//   a_t = e_t·w2 + b2; alpha = exp(a - m) / (sum exp(a - m) + 1e-8 exp(-m));
//   pooled = sum_t alpha_t h_t  (bf16 [n, 768]); alpha is kept (fp32 [n, T]) for the backward.
// ---------------------------------------------------------------------------------------------
constexpr int kPoolWarps = 8;
__global__ void __launch_bounds__(kPoolWarps * 32) text_pool_kernel(const __nv_bfloat16* __restrict__ hidden,
                                                        const __nv_bfloat16* __restrict__ e, const float* __restrict__ w2,
                                                        const float* __restrict__ b2, const int* __restrict__ mask_table,
                                                        const int* __restrict__ items, const int* __restrict__ count_dev,
                                                        int n_items, int T, int Hq, __nv_bfloat16* __restrict__ pooled,
                                                        float* __restrict__ alpha_out) {
  __shared__ float sA[kMaxT];
  __shared__ __align__(16) float sPart[kPoolWarps][kHid];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_wait();
  pdl_launch();
  const int n_eff = count_dev ? min(n_items, *count_dev) : n_items;
  for (int n = blockIdx.x; n < n_eff; n += gridDim.x) {
    const size_t row0 = (size_t)n * T;
    const int item = items ? items[n] : n;
    // logits a_t = e_t · w2 + b2
    for (int t = w; t < T; t += kPoolWarps) {
      float s = 0.f;
      for (int c = lane * 8; c < Hq; c += 256) {
        float ev[8];
        load8(e + (row0 + t) * Hq + c, ev);
#pragma unroll
        for (int i = 0; i < 8; ++i) s = fmaf(ev[i], w2[c + i], s);
      }
      s = warp_sum(s);
      if (lane == 0) sA[t] = s + b2[0];
    }
    __syncthreads();
    // every warp normalises redundantly: lane holds tokens lane and lane + 32 (T <= 64)
    const int t0 = lane, t1 = lane + 32;
    const float v0 = t0 < T ? sA[t0] : -INFINITY, v1 = t1 < T ? sA[t1] : -INFINITY;
    const float mx = warp_max(fmaxf(v0, v1));
    const float mk0 = (t0 < T && !(mask_table && mask_table[(size_t)item * T + t0] == 0)) ? 1.f : 0.f;
    const float mk1 = (t1 < T && !(mask_table && mask_table[(size_t)item * T + t1] == 0)) ? 1.f : 0.f;
    const float ex0 = t0 < T ? __expf(v0 - mx) * mk0 : 0.f, ex1 = t1 < T ? __expf(v1 - mx) * mk1 : 0.f;
    const float invz = 1.f / (warp_sum(ex0 + ex1) + 1e-8f * __expf(-mx));
    const float a0 = ex0 * invz, a1 = ex1 * invz;
    if (w == 0 && alpha_out) {
      if (t0 < T) alpha_out[row0 + t0] = a0;
      if (t1 < T) alpha_out[row0 + t1] = a1;
    }
    float acc[kPerLane];
#pragma unroll
    for (int i = 0; i < kPerLane; ++i) acc[i] = 0.f;
    for (int t = w; t < T; t += kPoolWarps) {
      const float a = __shfl_sync(0xffffffffu, t < 32 ? a0 : a1, t & 31);
      float hv[8];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        load8(hidden + (row0 + t) * kHid + c * 256 + lane * 8, hv);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[c * 8 + i] = fmaf(a, hv[i], acc[c * 8 + i]);
      }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float4* dst = reinterpret_cast<float4*>(&sPart[w][c * 256 + lane * 8]);
      dst[0] = make_float4(acc[c * 8], acc[c * 8 + 1], acc[c * 8 + 2], acc[c * 8 + 3]);
      dst[1] = make_float4(acc[c * 8 + 4], acc[c * 8 + 5], acc[c * 8 + 6], acc[c * 8 + 7]);
    }
    __syncthreads();
    if (threadIdx.x < kHid / 8) {
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = 0.f;
#pragma unroll
      for (int k = 0; k < kPoolWarps; ++k) {
        const float4* src = reinterpret_cast<const float4*>(&sPart[k][threadIdx.x * 8]);
        const float4 x = src[0], y = src[1];
        o[0] += x.x; o[1] += x.y; o[2] += x.z; o[3] += x.w; o[4] += y.x; o[5] += y.y; o[6] += y.z; o[7] += y.w;
      }
      store8(pooled + (size_t)n * kHid + threadIdx.x * 8, o);
    }
    __syncthreads();                       // sA / sPart are reused by the next item
  }
}

// ---------------------------------------------------------------------------------------------
// Text-head pooling backward (the "injected gradient" pass of model.py:72-90, head only: the trunk is frozen).
//   d_alpha_t = d_pooled · h_t ;  d_a_t = alpha_t (d_alpha_t - sum_s alpha_s d_alpha_s)
//   d_b2 += sum d_a ; d_w2 += sum_t d_a_t e_t ; d_pre[t] = d_a_t * w2 * (1 - e_t^2)      (bf16, GEMM operand)
// One warp per item, grid-stride; per-block partial sums for d_w2 so atomics stay rare. Rows of d_pre in
// [count*T, round_up(count*T, 64)) are zeroed: the weight-gradient GEMM reduces over whole 64-row blocks.
// ---------------------------------------------------------------------------------------------
constexpr int kMaxHq = 512;
// One block (8 warps) per item, tokens dealt round-robin to the warps:
//   da[n*T + t] (fp32), d_b2 += sum da, then d_pre[r, c] = da[r] * w2[c] * (1 - e[r,c]^2) (bf16, the weight-gradient GEMM
//   operand) for the item's T rows. Block 0 also writes zeros to the rows in [rows_eff, round_up(rows_eff, 64)).
// The two column reductions that need every row (d_w2 = e^T·da, d_b1 = colsum(d_pre)) are separate launches
// (colsum kernels in user_encoder.cu), so the training step can take them off its critical path.
__global__ void __launch_bounds__(kPoolWarps * 32) text_pool_bwd_kernel(const __nv_bfloat16* __restrict__ hidden,
                                                            const __nv_bfloat16* __restrict__ e,
                                                            const float* __restrict__ alpha,
                                                            const __nv_bfloat16* __restrict__ d_pooled,
                                                            const float* __restrict__ w2,
                                                            const int* __restrict__ count_dev, int n_items, int T, int Hq,
                                                            float* __restrict__ da_out, __nv_bfloat16* __restrict__ d_pre,
                                                            float* __restrict__ d_b2) {
  __shared__ float sS[kMaxT], sDa[kMaxT];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_wait();
  pdl_launch();
  const int n_eff = count_dev ? min(n_items, *count_dev) : n_items;
  const int hq8 = Hq >> 3;
  float db2 = 0.f;
  for (int n = blockIdx.x; n < n_eff; n += gridDim.x) {
    const size_t row0 = (size_t)n * T;
    float dp[kPerLane];
#pragma unroll
    for (int c = 0; c < 3; ++c) load8(d_pooled + (size_t)n * kHid + c * 256 + lane * 8, dp + c * 8);
    for (int t = w; t < T; t += kPoolWarps) {
      float hv[8], s = 0.f;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        load8(hidden + (row0 + t) * kHid + c * 256 + lane * 8, hv);
#pragma unroll
        for (int i = 0; i < 8; ++i) s = fmaf(dp[c * 8 + i], hv[i], s);
      }
      s = warp_sum(s);
      if (lane == 0) sS[t] = s;
    }
    __syncthreads();
    if (w == 0) {
      const int t0 = lane, t1 = lane + 32;
      const float a0 = t0 < T ? alpha[row0 + t0] : 0.f, a1 = t1 < T ? alpha[row0 + t1] : 0.f;
      const float s0 = t0 < T ? sS[t0] : 0.f, s1 = t1 < T ? sS[t1] : 0.f;
      const float dot = warp_sum(fmaf(a0, s0, a1 * s1));
      const float da0 = a0 * (s0 - dot), da1 = a1 * (s1 - dot);
      if (t0 < T) { da_out[row0 + t0] = da0; sDa[t0] = da0; }
      if (t1 < T) { da_out[row0 + t1] = da1; sDa[t1] = da1; }
      db2 += da0 + da1;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < T * hq8; i += kPoolWarps * 32) {
      const int t = i / hq8, c = (i - t * hq8) * 8;
      float ev[8], out[8];
      load8(e + (row0 + t) * Hq + c, ev);
      const float d = sDa[t];
      const float4 wa = *reinterpret_cast<const float4*>(w2 + c), wb = *reinterpret_cast<const float4*>(w2 + c + 4);
      const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
      for (int k = 0; k < 8; ++k) out[k] = d * wv[k] * (1.f - ev[k] * ev[k]);
      store8(d_pre + (row0 + t) * Hq + c, out);
    }
    __syncthreads();                       // sS / sDa are reused by the next item
  }
  if (w == 0) {
    db2 = warp_sum(db2);
    if (lane == 0 && db2 != 0.f) atomicAdd(d_b2, db2);
  }
  if (blockIdx.x == 0) {
    const long long rows_max = (long long)n_items * T, rows_eff = (long long)n_eff * T;
    long long rows_pad = (rows_eff + 63) / 64 * 64;
    if (rows_pad > rows_max) rows_pad = rows_max;
    const float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (long long i = threadIdx.x; i < (rows_pad - rows_eff) * hq8; i += kPoolWarps * 32)
      store8(d_pre + rows_eff * Hq + i * 8, z);
  }
}

// ---------------------------------------------------------------------------------------------
// Trunk-output cache (opt-in). DistilBERT is frozen, so the hidden states of a news item never change: with 180 GB of
// HBM the [N, T, 768] bf16 table of a whole news corpus (3 GB for 65k items) simply stays resident and the trunk only
// runs for items seen for the first time. No counterpart upstream (it re-encodes every id every batch, model.py:41-61).
//   cache_select: ids among uniq[0..count) whose `valid` flag is 0 → miss[0..*miss_count)   (order unspecified)
//   rows_copy:    dst[(di)*R + r, :] = src[(si)*R + r, :] for i < count, si/di = optional index arrays (gather / scatter),
//                 optionally raising valid[dst_idx[i]]
// ---------------------------------------------------------------------------------------------
__global__ void cache_select_kernel(const int* __restrict__ uniq, const int* __restrict__ count_dev, int n_items,
                                    const int* __restrict__ valid, int* __restrict__ miss, int* __restrict__ miss_count) {
  const int n_eff = count_dev ? min(n_items, *count_dev) : n_items;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n_eff) {
    const int id = uniq[k];
    if (valid[id] == 0) miss[atomicAdd(miss_count, 1)] = id;
  }
}

__global__ void __launch_bounds__(256) rows_copy_kernel(const uint4* __restrict__ src, const int* __restrict__ src_idx,
                                                        uint4* __restrict__ dst, const int* __restrict__ dst_idx,
                                                        const int* __restrict__ count_dev, int n_items, int rows_per_item,
                                                        int vec_per_row, int* __restrict__ valid) {
  const int n_eff = count_dev ? min(n_items, *count_dev) : n_items;
  const size_t per_item = (size_t)rows_per_item * vec_per_row;
  const size_t total = (size_t)n_eff * per_item;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int it = (int)(i / per_item);
    const size_t off = i - (size_t)it * per_item;
    const size_t s = (size_t)(src_idx ? src_idx[it] : it) * per_item + off;
    const int di = dst_idx ? dst_idx[it] : it;
    dst[(size_t)di * per_item + off] = src[s];
    if (valid && off == 0) valid[di] = 1;
  }
}

// fp32 → bf16 cast of a contiguous buffer (weight packing, activations entering a GEMM)
__global__ void __launch_bounds__(256) cast_bf16_kernel(const float4* __restrict__ in, uint2* __restrict__ out, size_t n4) {
  pdl_wait();
  pdl_launch();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = in[i];
    uint2 q;
    *reinterpret_cast<__nv_bfloat162*>(&q.x) = __floats2bfloat162_rn(v.x, v.y);
    *reinterpret_cast<__nv_bfloat162*>(&q.y) = __floats2bfloat162_rn(v.z, v.w);
    out[i] = q;
  }
}

}  // namespace fr

using namespace fr;

static inline int grid_for(long long warps_needed, int warps_per_block, int cap_blocks) {
  long long b = (warps_needed + warps_per_block - 1) / warps_per_block;
  if (b > cap_blocks) b = cap_blocks;
  if (b < 1) b = 1;
  return (int)b;
}

FR_API int fr_bert_embed_ln(const int* tokens, const int* items, const int* count_dev, int n_items, int T,
                            const void* word, const void* pos, const float* gamma, const float* beta, void* out,
                            float eps, int hidden, void* stream) {
  FR_REQUIRE(hidden == kHid, "fr_bert_embed_ln: hidden=%d (only %d is built)", hidden, kHid);
  if (n_items == 0) return 0;
  embed_ln_kernel<<<grid_for((long long)n_items * T, 8, num_sms() * 4), 256, 0, (cudaStream_t)stream>>>(
      tokens, items, count_dev, n_items, T, (const __nv_bfloat16*)word, (const __nv_bfloat16*)pos, gamma, beta,
      (__nv_bfloat16*)out, eps);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_layernorm_bf16(const void* in, void* out, const float* gamma, const float* beta, int rows,
                             const int* count_dev, int rows_per_item, float eps, int hidden, void* stream) {
  FR_REQUIRE(hidden == kHid, "fr_layernorm_bf16: hidden=%d (only %d is built)", hidden, kHid);
  if (rows == 0) return 0;
  // 117 registers x 512 threads = one block per SM: exactly one wave of blocks, so gamma/beta are loaded once per SM and
  // no block waits for another to retire (measured: +0.8% step throughput over a 3x oversubscribed grid).
  FR_CHECK(launch_pdl(layernorm_kernel, dim3(grid_for(rows, 32, num_sms())), dim3(512), 0, (cudaStream_t)stream,
                      (const __nv_bfloat16*)in, (__nv_bfloat16*)out, gamma, beta, rows, count_dev, rows_per_item, eps));
  return 0;
}

FR_API int fr_bert_attention(const void* qkv, const int* mask_table, const int* items, const int* count_dev,
                             int n_items, int T, int heads, int head_dim, void* out, void* stream) {
  FR_REQUIRE(head_dim == kHeadDim, "fr_bert_attention: head_dim=%d (only %d is built)", head_dim, kHeadDim);
  FR_REQUIRE(T >= 1 && T <= kMaxT, "fr_bert_attention: T=%d must be in [1,%d]", T, kMaxT);
  if (n_items == 0) return 0;
  bert_attention_kernel<<<grid_for((long long)n_items * heads, kAttnWarps, num_sms() * 16), kAttnWarps * 32, 0,
                          (cudaStream_t)stream>>>((const __nv_bfloat16*)qkv, mask_table, items, count_dev, n_items, T,
                                                  heads, (__nv_bfloat16*)out);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_text_pool(const void* hidden, const void* e, const float* w2, const float* b2, const int* mask_table,
                        const int* items, const int* count_dev, int n_items, int T, int Hq, int hidden_dim,
                        void* pooled, float* alpha_out, void* stream) {
  FR_REQUIRE(hidden_dim == kHid, "fr_text_pool: hidden=%d (only %d is built)", hidden_dim, kHid);
  FR_REQUIRE(T >= 1 && T <= kMaxT && Hq % 8 == 0, "fr_text_pool: T=%d Hq=%d (Hq must be a multiple of 8)", T, Hq);
  if (n_items == 0) return 0;
  FR_CHECK(launch_pdl(text_pool_kernel, dim3(n_items < num_sms() * 8 ? n_items : num_sms() * 8), dim3(kPoolWarps * 32), 0,
                      (cudaStream_t)stream, (const __nv_bfloat16*)hidden, (const __nv_bfloat16*)e, w2, b2, mask_table, items,
                      count_dev, n_items, T, Hq, (__nv_bfloat16*)pooled, alpha_out));
  return 0;
}

namespace fr {
int colsum_launch(const void* X, int is_bf16, const float* row_w, float* out, int M, int N, int ld, const int* m_dev,
                  int m_scale, cudaStream_t s);   // user_encoder.cu
}

// d_w2 / d_b1 may both be null: the caller then runs the two column reductions itself (fr_colsum on d_pre for d_b1,
// fr_colsum_weighted(e, da) for d_w2), e.g. on another stream.
FR_API int fr_text_pool_bwd(const void* hidden, const void* e, const float* alpha, const void* d_pooled, const float* w2,
                            const int* count_dev, int n_items, int T, int Hq, int hidden_dim, void* d_pre, float* da_scratch,
                            float* d_w2, float* d_b2, float* d_b1, void* stream) {
  FR_REQUIRE(hidden_dim == kHid, "fr_text_pool_bwd: hidden=%d (only %d is built)", hidden_dim, kHid);
  FR_REQUIRE(T >= 1 && T <= kMaxT && Hq >= 8 && Hq <= kMaxHq && Hq % 8 == 0, "fr_text_pool_bwd: T=%d Hq=%d", T, Hq);
  FR_REQUIRE(((uintptr_t)w2 & 15) == 0, "fr_text_pool_bwd: w2 must be 16-byte aligned");
  if (n_items == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  FR_CHECK(launch_pdl(text_pool_bwd_kernel, dim3(n_items < num_sms() * 8 ? n_items : num_sms() * 8), dim3(kPoolWarps * 32), 0, s,
                      (const __nv_bfloat16*)hidden, (const __nv_bfloat16*)e, alpha, (const __nv_bfloat16*)d_pooled, w2, count_dev,
                      n_items, T, Hq, da_scratch, (__nv_bfloat16*)d_pre, d_b2));
  const int rows = n_items * T;
  if (d_w2)
    if (int err = colsum_launch(e, 1, da_scratch, d_w2, rows, Hq, Hq, count_dev, T, s)) return err;
  if (d_b1)
    if (int err = colsum_launch(d_pre, 1, nullptr, d_b1, rows, Hq, Hq, count_dev, T, s)) return err;
  return 0;
}

FR_API int fr_cache_select(const int* uniq, const int* count_dev, int n_items, const int* valid, int* miss, int* miss_count,
                           void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  FR_CHECK(cudaMemsetAsync(miss_count, 0, sizeof(int), s));
  if (n_items == 0) return 0;
  cache_select_kernel<<<(n_items + 255) / 256, 256, 0, s>>>(uniq, count_dev, n_items, valid, miss, miss_count);
  FR_LAUNCH_CHECK();
  return 0;
}

// Row-block copy with optional gather (src_idx) / scatter (dst_idx) by item; row_bytes must be a multiple of 16.
FR_API int fr_rows_copy(const void* src, const int* src_idx, void* dst, const int* dst_idx, const int* count_dev,
                        int n_items, int rows_per_item, int row_bytes, int* valid, void* stream) {
  FR_REQUIRE(row_bytes % 16 == 0, "fr_rows_copy: row_bytes=%d must be a multiple of 16", row_bytes);
  if (n_items == 0) return 0;
  const size_t total = (size_t)n_items * rows_per_item * (row_bytes / 16);
  int blocks = (int)((total + 255) / 256);
  if (blocks > num_sms() * 16) blocks = num_sms() * 16;
  rows_copy_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((const uint4*)src, src_idx, (uint4*)dst, dst_idx, count_dev,
                                                            n_items, rows_per_item, row_bytes / 16, valid);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream) {
  FR_REQUIRE(n % 4 == 0, "fr_cast_f32_to_bf16: n=%lld must be a multiple of 4", n);
  if (n == 0) return 0;
  const size_t n4 = (size_t)n / 4;
  int blocks = (int)((n4 + 255) / 256);
  if (blocks > num_sms() * 16) blocks = num_sms() * 16;
  FR_CHECK(launch_pdl(cast_bf16_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const float4*)in, (uint2*)out, n4));
  return 0;
}
