// N4 / N5 / N6 / N11: the user encoder and the loss, forward and backward, around the tcgen05 GEMMs.
//
// Reference math (encoder.py:49-56, attention.py:14-26,37-45,69-81, model.py:121-126):
//   x   = dropout(his_vec, 0.2)                       his_vec = news_vecs[his ids]
//   QKV = x Wqkv^T + b                                (GEMM, gemm_sm100.cu)
//   A   = exp(QK^T/sqrt(dk)) * mask / (rowsum + 1e-8) ; ctx = merge_heads(A V)          → ue_mha_fwd
//   e   = tanh(ctx W1^T + b1)                         (GEMM with tanh epilogue)
//   a   = exp(e w2 + b2) * mask / (sum_L + 1e-8) ; u = sum_L a ctx
//   s   = sigmoid(cand · u) ; loss = mean_B CE(s, 0)                                    → ue_pool_score
// The backward kernels emit the operands of the weight-gradient GEMMs (gemm_grad_sm100.cu) and scatter
// the gradients of the news vectors straight into d_news[unique row] (the reference's Python dict
// segment-sum, main.py:20-42), optionally with the reference's LDP noise (client.py:87-89).
//
// Kernels here are latency-bound glue between GEMMs (≈3 GFLOP per 50-user round in total): one warp per
// (user, head) or per user, fp32 math, bf16 tensors where a GEMM consumes them.
#include "common.cuh"

namespace fr {

constexpr int kMaxL = 64;       // history length limit (reference: 50)
constexpr int kJobThreads = 64; // one (user, head) job per block; thread t owns query row t (and key row t in the backward)

__device__ __forceinline__ void red_add_f32x4(float* addr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Inverted-dropout multiplier for element (row, 4*q4..4*q4+3): four keep decisions from one Philox call.
__device__ __forceinline__ float4 dropout_mult4(const Philox& rng, uint64_t idx4, float p, float inv_keep) {
  uint32_t r[4];
  rng(idx4, 0xD0u, r);
  const uint32_t thr = (uint32_t)(p * 4294967296.0);
  return make_float4(r[0] >= thr ? inv_keep : 0.f, r[1] >= thr ? inv_keep : 0.f, r[2] >= thr ? inv_keep : 0.f,
                     r[3] >= thr ? inv_keep : 0.f);
}

// X[b*L + l, :] = bf16( dropout( news_vecs[inv_his[b*L + l], :] ) )
__global__ void __launch_bounds__(256) ue_gather_dropout_kernel(const float* __restrict__ news_vecs, const int* __restrict__ inv_his,
                                                                __nv_bfloat16* __restrict__ X, int rows, int D4, float p,
                                                                unsigned long long seed,
                                                                const unsigned long long* __restrict__ seed_dev) {
  pdl_wait();      // may have been launched early (PDL): the producer has finished from here on
  pdl_launch();
  if (seed_dev) seed += *seed_dev;            // per-replay salt when the step lives in a CUDA graph
  const Philox rng(seed);
  const float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  const size_t total = (size_t)rows * D4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / D4), q = (int)(i - (size_t)row * D4);
    float4 v = reinterpret_cast<const float4*>(news_vecs)[(size_t)inv_his[row] * D4 + q];
    if (p > 0.f) {
      const float4 m = dropout_mult4(rng, i, p, inv_keep);
      v.x *= m.x; v.y *= m.y; v.z *= m.z; v.w *= m.w;
    }
    uint2 o;
    *reinterpret_cast<__nv_bfloat162*>(&o.x) = __floats2bfloat162_rn(v.x, v.y);
    *reinterpret_cast<__nv_bfloat162*>(&o.y) = __floats2bfloat162_rn(v.z, v.w);
    reinterpret_cast<uint2*>(X)[i] = o;
  }
}

// d_news[inv_his[row], :] += dX[row, :] * dropout_mult (+ N(0, noise_std^2))
__global__ void __launch_bounds__(256) ue_scatter_dx_kernel(const float* __restrict__ dX, const int* __restrict__ inv_his,
                                                            float* __restrict__ d_news, int rows, int D4, float p,
                                                            unsigned long long seed, float noise_std,
                                                            unsigned long long noise_seed,
                                                            const unsigned long long* __restrict__ seed_dev) {
  pdl_wait();      // may have been launched early (PDL): the producer has finished from here on
  pdl_launch();
  if (seed_dev) { seed += *seed_dev; noise_seed += *seed_dev; }
  const Philox rng(seed), nrng(noise_seed);
  const float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  const size_t total = (size_t)rows * D4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / D4), q = (int)(i - (size_t)row * D4);
    float4 v = reinterpret_cast<const float4*>(dX)[i];
    if (p > 0.f) {
      const float4 m = dropout_mult4(rng, i, p, inv_keep);
      v.x *= m.x; v.y *= m.y; v.z *= m.z; v.w *= m.w;
    }
    if (noise_std > 0.f) {
      uint32_t r[4]; nrng(i, 0x415u, r);
      float z0, z1, z2, z3; box_muller(r[0], r[1], z0, z1); box_muller(r[2], r[3], z2, z3);
      v.x += noise_std * z0; v.y += noise_std * z1; v.z += noise_std * z2; v.w += noise_std * z3;
    }
    red_add_f32x4(d_news + ((size_t)inv_his[row] * D4 + q) * 4, v);
  }
}

// ---------------------------------------------------------------------------------------------------
// Multi-head self-attention over the history, one 64-thread block per (user, head).
// qkv bf16 [B*L, 3*H*dk] (q | k | v), ctx bf16 [B*L, H*dk], stats fp32 [B*H*L, 2] = (row max, denominator).
// Thread i owns query row i.
// ---------------------------------------------------------------------------------------------------
template <int DK>
__global__ void __launch_bounds__(kJobThreads) ue_mha_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, const int* __restrict__ his_ids,
                                                                 __nv_bfloat16* __restrict__ ctx, float* __restrict__ stats,
                                                                 int B, int L, int H, float scale, int use_mask) {
  pdl_wait();      // may have been launched early (PDL): the producer has finished from here on
  pdl_launch();
  __shared__ float sK[kMaxL][DK + 1];
  __shared__ float sV[kMaxL][DK + 1];
  __shared__ float sM[kMaxL];
  const int tid = threadIdx.x;
  const int job = blockIdx.x;
  const int b = job / H, h = job - b * H;
  const int ld = 3 * H * DK;
  const size_t row0 = (size_t)b * L;
  for (int i = tid; i < L * DK; i += kJobThreads) {
    const int r = i / DK, d = i - r * DK;
    const __nv_bfloat16* base = qkv + (row0 + r) * ld + h * DK + d;
    sK[r][d] = __bfloat162float(base[H * DK]);
    sV[r][d] = __bfloat162float(base[2 * H * DK]);
  }
  for (int j = tid; j < L; j += kJobThreads) sM[j] = (use_mask && his_ids[row0 + j] == 0) ? 0.f : 1.f;
  __syncthreads();
  for (int i = tid; i < L; i += kJobThreads) {
    float q[DK];
#pragma unroll
    for (int d = 0; d < DK; ++d) q[d] = __bfloat162float(qkv[(row0 + i) * ld + h * DK + d]) * scale;
    float m = -INFINITY;
    for (int j = 0; j < L; ++j) {
      float s = 0.f;
#pragma unroll
      for (int d = 0; d < DK; ++d) s = fmaf(q[d], sK[j][d], s);
      m = fmaxf(m, s);
    }
    float z = 0.f, acc[DK];
#pragma unroll
    for (int d = 0; d < DK; ++d) acc[d] = 0.f;
    for (int j = 0; j < L; ++j) {
      float s = 0.f;
#pragma unroll
      for (int d = 0; d < DK; ++d) s = fmaf(q[d], sK[j][d], s);
      const float e = __expf(s - m) * sM[j];
      z += e;
#pragma unroll
      for (int d = 0; d < DK; ++d) acc[d] = fmaf(e, sV[j][d], acc[d]);
    }
    z += 1e-8f * __expf(-m);
    const float inv = 1.f / z;
    __nv_bfloat16* op = ctx + (row0 + i) * (size_t)(H * DK) + h * DK;
#pragma unroll
    for (int d = 0; d < DK; d += 2)
      *reinterpret_cast<__nv_bfloat162*>(op + d) = __floats2bfloat162_rn(acc[d] * inv, acc[d + 1] * inv);
    stats[((size_t)job * L + i) * 2] = m;
    stats[((size_t)job * L + i) * 2 + 1] = z;
  }
}

// Backward of the attention: dctx bf16 [B*L, H*dk] → dqkv bf16 [B*L, 3*H*dk]. Probabilities are
// recomputed from q, k and the saved (max, denominator).
// dot of a register row with a shared-memory row read as float4 (all threads read the same row: broadcast)
template <int DKP>
__device__ __forceinline__ float dot_row(const float* reg, const float* srow) {
  float s = 0.f;
#pragma unroll
  for (int d = 0; d < DKP; d += 4) {
    const float4 k = *reinterpret_cast<const float4*>(srow + d);
    s = fmaf(reg[d], k.x, s); s = fmaf(reg[d + 1], k.y, s); s = fmaf(reg[d + 2], k.z, s); s = fmaf(reg[d + 3], k.w, s);
  }
  return s;
}
template <int DKP>
__device__ __forceinline__ void axpy_row(float* reg, float a, const float* srow) {
#pragma unroll
  for (int d = 0; d < DKP; d += 4) {
    const float4 k = *reinterpret_cast<const float4*>(srow + d);
    reg[d] = fmaf(a, k.x, reg[d]); reg[d + 1] = fmaf(a, k.y, reg[d + 1]);
    reg[d + 2] = fmaf(a, k.z, reg[d + 2]); reg[d + 3] = fmaf(a, k.w, reg[d + 3]);
  }
}

// One (user, head) job per block of 128 threads: thread (r, hf) = (tid & 63, tid >> 6) owns row r and one half of the
// work of that row, so four warps share a job and nothing is computed twice:
//   A  (r = query i, hf = half of the keys j): P_ij = exp(s_ij - m_i) mask_j / z_i and dP_ij = dctx_i · v_j → shared
//      memory; partial Dd_i = sum_j P_ij dP_ij
//   B1 (same split): dS_ij = P_ij (dP_ij - Dd_i), overwriting dP
//   B2 (r = query i, hf = half of the head dims): dQ_i = scale * sum_j dS_ij K_j
//   C  (r = key j,   hf = half of the head dims): dV_j = sum_i P_ij dctx_i ; dK_j = sum_i dS_ij (q_i * scale)
// P and dS rows have an odd stride, so row-wise (thread = i) and column-wise (thread = j) accesses are conflict-free.
constexpr int kBwdThreads = 128;
template <int DK>
__global__ void __launch_bounds__(kBwdThreads) ue_mha_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dctx,
                                                                 const float* __restrict__ stats, const int* __restrict__ his_ids,
                                                                 __nv_bfloat16* __restrict__ dqkv, int B, int L, int H, float scale,
                                                                 int use_mask) {
  pdl_wait();      // may have been launched early (PDL): the producer has finished from here on
  pdl_launch();
  constexpr int DKP = (DK + 3) / 4 * 4;            // rows padded to 16 bytes: the hot loops read them as float4 broadcasts
  constexpr int DH = DK / 2;                       // head dims per thread in B2 / C
  static_assert(DK % 4 == 0 && DH % 2 == 0, "head dim must be a multiple of 4");
  extern __shared__ __align__(16) float smem_f[];
  const int LP = L | 1;
  float* sQ = smem_f;                              // [L][DKP] q * scale
  float* sK = sQ + L * DKP;
  float* sV = sK + L * DKP;
  float* sG = sV + L * DKP;                        // dctx
  float* sP = sG + L * DKP;                        // [L][LP]
  float* sDS = sP + L * LP;                        // [L][LP] dP, then dS
  float* sMx = sDS + L * LP;
  float* sZ = sMx + L;
  float* sMask = sZ + L;
  float* sDd = sMask + L;                          // [2][L] partial Dd of the two key halves
  const int tid = threadIdx.x;
  const int r = tid & 63, hf = tid >> 6;
  const int job = blockIdx.x;
  const int b = job / H, h = job - b * H;
  const int ld = 3 * H * DK;
  const size_t row0 = (size_t)b * L;
  for (int i = tid; i < L * DKP; i += kBwdThreads) {
    const int rr = i / DKP, d = i - rr * DKP;
    float q = 0.f, k = 0.f, v = 0.f, g = 0.f;
    if (d < DK) {
      const __nv_bfloat16* base = qkv + (row0 + rr) * ld + h * DK + d;
      q = __bfloat162float(base[0]) * scale;
      k = __bfloat162float(base[H * DK]);
      v = __bfloat162float(base[2 * H * DK]);
      g = __bfloat162float(dctx[(row0 + rr) * (size_t)(H * DK) + h * DK + d]);
    }
    sQ[i] = q; sK[i] = k; sV[i] = v; sG[i] = g;
  }
  for (int j = tid; j < L; j += kBwdThreads) {
    sMx[j] = stats[((size_t)job * L + j) * 2];
    sZ[j] = stats[((size_t)job * L + j) * 2 + 1];
    sMask[j] = (use_mask && his_ids[row0 + j] == 0) ? 0.f : 1.f;
  }
  __syncthreads();
  const int jh = (L + 1) / 2;
  const int j0 = hf * jh, j1 = min(L, j0 + jh);
  // ---- A ----
  if (r < L) {
    float q[DKP], g[DKP];
#pragma unroll
    for (int d = 0; d < DKP; ++d) { q[d] = sQ[r * DKP + d]; g[d] = sG[r * DKP + d]; }
    const float m = sMx[r], invz = 1.f / sZ[r];
    float dd = 0.f;
    for (int j = j0; j < j1; ++j) {
      const float sc = dot_row<DKP>(q, sK + j * DKP), dp = dot_row<DKP>(g, sV + j * DKP);
      const float pij = __expf(sc - m) * sMask[j] * invz;
      sP[r * LP + j] = pij;
      sDS[r * LP + j] = dp;
      dd = fmaf(pij, dp, dd);
    }
    sDd[hf * L + r] = dd;
  }
  __syncthreads();
  // ---- B1 ----
  if (r < L) {
    const float dd = sDd[r] + sDd[L + r];
    for (int j = j0; j < j1; ++j) sDS[r * LP + j] = sP[r * LP + j] * (sDS[r * LP + j] - dd);
  }
  __syncthreads();
  // ---- B2 ----
  if (r < L) {
    float dq[DH];
#pragma unroll
    for (int d = 0; d < DH; ++d) dq[d] = 0.f;
    for (int j = 0; j < L; ++j) {
      const float ds = sDS[r * LP + j];
      const float* kr = sK + j * DKP + hf * DH;
#pragma unroll
      for (int d = 0; d < DH; d += 2) {
        const float2 k2 = *reinterpret_cast<const float2*>(kr + d);
        dq[d] = fmaf(ds, k2.x, dq[d]);
        dq[d + 1] = fmaf(ds, k2.y, dq[d + 1]);
      }
    }
    __nv_bfloat16* op = dqkv + (row0 + r) * ld + h * DK + hf * DH;
#pragma unroll
    for (int d = 0; d < DH; d += 2)
      *reinterpret_cast<__nv_bfloat162*>(op + d) = __floats2bfloat162_rn(dq[d] * scale, dq[d + 1] * scale);
  }
  // ---- C ---- (reads only what A / B1 wrote: no barrier needed after B2)
  if (r < L) {
    float dk[DH], dv[DH];
#pragma unroll
    for (int d = 0; d < DH; ++d) { dk[d] = 0.f; dv[d] = 0.f; }
    for (int i = 0; i < L; ++i) {
      const float pij = sP[i * LP + r], ds = sDS[i * LP + r];
      const float* gr = sG + i * DKP + hf * DH;
      const float* qr = sQ + i * DKP + hf * DH;
#pragma unroll
      for (int d = 0; d < DH; d += 2) {
        const float2 g2 = *reinterpret_cast<const float2*>(gr + d), q2 = *reinterpret_cast<const float2*>(qr + d);
        dv[d] = fmaf(pij, g2.x, dv[d]);
        dv[d + 1] = fmaf(pij, g2.y, dv[d + 1]);
        dk[d] = fmaf(ds, q2.x, dk[d]);                  // sQ already carries the 1/sqrt(dk) factor
        dk[d + 1] = fmaf(ds, q2.y, dk[d + 1]);
      }
    }
    __nv_bfloat16* op = dqkv + (row0 + r) * ld + h * DK + hf * DH;
#pragma unroll
    for (int d = 0; d < DH; d += 2) {
      *reinterpret_cast<__nv_bfloat162*>(op + H * DK + d) = __floats2bfloat162_rn(dk[d], dk[d + 1]);
      *reinterpret_cast<__nv_bfloat162*>(op + 2 * H * DK + d) = __floats2bfloat162_rn(dv[d], dv[d + 1]);
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Additive pooling + candidate scoring + sigmoid + soft-max CE, forward and (optionally) backward.
// One block (512 threads) per user. D <= 512, Q <= 256, L <= 64, C <= 32.
// ---------------------------------------------------------------------------------------------------
constexpr int kPoolThreads = 512;   // one block per user; B is small (50), so latency per user is what matters
struct PoolScoreArgs {
  const __nv_bfloat16* ctx;     // [B*L, D]
  const __nv_bfloat16* e;       // [B*L, Q]  tanh(fc1)
  const float* w2;              // [Q]
  const float* b2;              // [1]
  const float* news_vecs;       // [U, D]
  const int* inv_cand;          // [B*C]
  const int* his_ids;           // [B*L] (mask) or null
  float* score;                 // [B, C]
  float* loss_mean;             // [1], zeroed by the entry point
  float* user_vec;              // [B, D] or null
  // backward outputs (null in inference)
  __nv_bfloat16* d_ctx_pool;    // [B*L, D]   alpha_l * d_u
  __nv_bfloat16* d_pre;         // [B*L, Q]   d a_l * w2 * (1 - e^2)
  float* d_w2;                  // [Q]  (+=)
  float* d_b2;                  // [1]  (+=)
  float* d_news;                // [U, D] (+=)  candidate gradients scattered by unique row
  int B, L, C, D, Q, use_mask, train;
  float noise_std;
  unsigned long long noise_seed;
  const unsigned long long* seed_dev;
};

__global__ void __launch_bounds__(kPoolThreads) ue_pool_score_kernel(const PoolScoreArgs a) {
  pdl_wait();      // may have been launched early (PDL): the producer has finished from here on
  pdl_launch();
  __shared__ float sA[kMaxL];          // logits → exp → alpha
  __shared__ float sU[512];            // user vector, later d_u
  __shared__ float sS[32];             // candidate logits / ds
  __shared__ float sDal[kMaxL];        // d_alpha, later d_a
  __shared__ float red[32];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int L = a.L, D = a.D, Q = a.Q, C = a.C;
  const size_t row0 = (size_t)b * L;
  // a_l = e_l · w2 + b2  (one warp per row, 4 rows in flight)
  for (int l = w; l < L; l += kPoolThreads / 32) {
    float s = 0.f;
    for (int c = lane; c < Q; c += 32) s = fmaf(__bfloat162float(a.e[(row0 + l) * Q + c]), a.w2[c], s);
    s = warp_sum(s);
    if (lane == 0) sA[l] = s + a.b2[0];
  }
  __syncthreads();
  float mx = -INFINITY;
  for (int l = tid; l < L; l += kPoolThreads) mx = fmaxf(mx, sA[l]);
  mx = block_max(mx, red);
  float z = 0.f;
  for (int l = tid; l < L; l += kPoolThreads) {
    const float mk = (a.use_mask && a.his_ids[row0 + l] == 0) ? 0.f : 1.f;
    const float ex = __expf(sA[l] - mx) * mk;
    sA[l] = ex;
    z += ex;
  }
  z = block_sum(z, red) + 1e-8f * __expf(-mx);
  __syncthreads();
  for (int l = tid; l < L; l += kPoolThreads) sA[l] *= 1.f / z;
  __syncthreads();
  // u = sum_l alpha_l ctx_l
  for (int d = tid; d < D; d += kPoolThreads) {
    float acc = 0.f;
    for (int l = 0; l < L; ++l) acc = fmaf(sA[l], __bfloat162float(a.ctx[(row0 + l) * D + d]), acc);
    sU[d] = acc;
    if (a.user_vec) a.user_vec[(size_t)b * D + d] = acc;
  }
  __syncthreads();
  // candidate logits
  for (int c = w; c < C; c += kPoolThreads / 32) {
    const float* cv = a.news_vecs + (size_t)a.inv_cand[b * C + c] * D;
    float s = 0.f;
    for (int d = lane; d < D; d += 32) s = fmaf(cv[d], sU[d], s);
    s = warp_sum(s);
    if (lane == 0) sS[c] = s;
  }
  __syncthreads();
  if (w == 0) {
    const bool act = lane < C;
    const float p = act ? 1.f / (1.f + __expf(-sS[lane])) : -INFINITY;
    const float pm = warp_max(p);
    const float ex = act ? __expf(p - pm) : 0.f;
    const float zz = warp_sum(ex);
    const float p0 = __shfl_sync(0xffffffffu, p, 0);
    if (act && a.score) a.score[b * C + lane] = p;
    if (lane == 0 && a.loss_mean) atomicAdd(a.loss_mean, -(p0 - pm - __logf(zz)) / (float)a.B);
    // dL/ds_c = (softmax_c - [c == 0]) / B * p (1 - p)
    const float ds = act ? (ex / zz - (lane == 0 ? 1.f : 0.f)) / (float)a.B * p * (1.f - p) : 0.f;
    __syncwarp();
    sS[lane] = ds;
  }
  if (!a.train) return;
  __syncthreads();
  // d_cand_c = ds_c * u → scattered into d_news ; d_u = sum_c ds_c cand_c
  {
    const Philox nrng(a.noise_seed + (a.seed_dev ? *a.seed_dev : 0ull));
    const int D4 = D / 4;
    for (int i = tid; i < C * D4; i += kPoolThreads) {
      const int c = i / D4, q = i - c * D4;
      const float ds = sS[c];
      float4 v = make_float4(ds * sU[4 * q], ds * sU[4 * q + 1], ds * sU[4 * q + 2], ds * sU[4 * q + 3]);
      if (a.noise_std > 0.f) {
        uint32_t r[4]; nrng((uint64_t)(b * C + c) * D4 + q, 0xCA4Du, r);
        float z0, z1, z2, z3; box_muller(r[0], r[1], z0, z1); box_muller(r[2], r[3], z2, z3);
        v.x += a.noise_std * z0; v.y += a.noise_std * z1; v.z += a.noise_std * z2; v.w += a.noise_std * z3;
      }
      red_add_f32x4(a.d_news + ((size_t)a.inv_cand[b * C + c] * D4 + q) * 4, v);
    }
  }
  __syncthreads();
  for (int d = tid; d < D; d += kPoolThreads) {
    float du = 0.f;
    for (int c = 0; c < C; ++c) du = fmaf(sS[c], a.news_vecs[(size_t)a.inv_cand[b * C + c] * D + d], du);
    sU[d] = du;                       // from here on sU holds d_u
  }
  __syncthreads();
  // d_alpha_l = d_u · ctx_l ; d_ctx_pool[l] = alpha_l * d_u
  for (int l = w; l < L; l += kPoolThreads / 32) {
    float s = 0.f;
    const float al = sA[l];
    for (int d = lane * 2; d < D; d += 64) {
      const float2 cf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(a.ctx + (row0 + l) * D + d));
      s = fmaf(sU[d], cf.x, s);
      s = fmaf(sU[d + 1], cf.y, s);
      *reinterpret_cast<__nv_bfloat162*>(a.d_ctx_pool + (row0 + l) * D + d) = __floats2bfloat162_rn(al * sU[d], al * sU[d + 1]);
    }
    s = warp_sum(s);
    if (lane == 0) sDal[l] = s;
  }
  __syncthreads();
  float dot = 0.f;
  for (int l = tid; l < L; l += kPoolThreads) dot = fmaf(sA[l], sDal[l], dot);
  dot = block_sum(dot, red);
  __syncthreads();
  float db2 = 0.f;
  for (int l = tid; l < L; l += kPoolThreads) {
    const float da = sA[l] * (sDal[l] - dot);
    sDal[l] = da;
    db2 += da;
  }
  db2 = block_sum(db2, red);
  if (tid == 0) atomicAdd(a.d_b2, db2);
  __syncthreads();
  // d_w2 += sum_l da_l e_l ; d_pre[l] = da_l * w2 * (1 - e^2)
  for (int c = tid; c < Q; c += kPoolThreads) {
    float acc = 0.f;
    const float wc = a.w2[c];
    for (int l = 0; l < L; ++l) {
      const float ev = __bfloat162float(a.e[(row0 + l) * Q + c]);
      acc = fmaf(sDal[l], ev, acc);
      a.d_pre[(row0 + l) * Q + c] = __float2bfloat16(sDal[l] * wc * (1.f - ev * ev));
    }
    atomicAdd(a.d_w2 + c, acc);
  }
}

// out[n] += sum_m w[m] X[m, n]   (w = 1 if row_w is null: bias gradients; with weights: a transposed GEMV).
// X [M, N] with leading dimension ld; rows limited by *m_dev * m_scale.
// A thread owns 8 (bf16) or 4 (fp32) consecutive columns = one 16-byte load per row; 8 row lanes per block.
template <typename T, int VEC>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ X, const float* __restrict__ row_w,
                                                     float* __restrict__ out, int M, int N, int ld,
                                                     const int* __restrict__ m_dev, int m_scale) {
  if (m_dev) M = min(M, *m_dev * m_scale);
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cl) * VEC;
  __shared__ float sm[8][32][VEC + 1];
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) acc[i] = 0.f;
  if (col < N) {
    // four rows in flight per thread: the loop is a chain of dependent-free 16-byte loads, so latency is what costs
    const int rstep = gridDim.y * 8;
    for (int r = blockIdx.y * 8 + rl; r < M; r += 4 * rstep) {
      uint4 q[4];
      float wr[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int ru = r + u * rstep;
        const bool ok = ru < M;
        q[u] = ok ? *reinterpret_cast<const uint4*>(X + (size_t)ru * ld + col) : make_uint4(0, 0, 0, 0);
        wr[u] = ok ? (row_w ? row_w[ru] : 1.f) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (VEC == 8) {
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q[u]);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 f = __bfloat1622float2(h[i]);
            acc[2 * i] = fmaf(wr[u], f.x, acc[2 * i]);
            acc[2 * i + 1] = fmaf(wr[u], f.y, acc[2 * i + 1]);
          }
        } else {
          const float* f = reinterpret_cast<const float*>(&q[u]);
#pragma unroll
          for (int i = 0; i < VEC; ++i) acc[i] = fmaf(wr[u], f[i], acc[i]);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < VEC; ++i) sm[rl][cl][i] = acc[i];
  __syncthreads();
  if (rl == 0 && col < N) {
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      float t = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) t += sm[k][cl][i];
      if (col + i < N) atomicAdd(out + col + i, t);
    }
  }
}

// out[c, r] = bf16(in[r, c])  — K-major copies of weights for the data-gradient GEMMs (tiny matrices)
__global__ void __launch_bounds__(256) transpose_cast_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                             int R, int Ccols, int ld_out) {
  __shared__ float t[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    t[i][threadIdx.x] = (r < R && c < Ccols) ? in[(size_t)r * Ccols + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (c < Ccols && r < R) out[(size_t)c * ld_out + r] = __float2bfloat16(t[threadIdx.x][i]);
  }
}

}  // namespace fr

using namespace fr;

__global__ void counter_advance_kernel(unsigned long long* ctr, unsigned long long inc) { *ctr += inc; }

// *ctr += inc on the stream: the per-step salt of the RNG streams when a step is replayed from a CUDA graph.
FR_API int fr_counter_advance(unsigned long long* ctr, unsigned long long inc, void* stream) {
  counter_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(ctr, inc);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_ue_gather_dropout(const float* news_vecs, const int* inv_his, void* X, int rows, int D, float p,
                                unsigned long long seed, const unsigned long long* seed_dev, void* stream) {
  FR_REQUIRE(D % 4 == 0 && p >= 0.f && p < 1.f, "fr_ue_gather_dropout: D=%d p=%f", D, p);
  if (rows == 0) return 0;
  const size_t total = (size_t)rows * (D / 4);
  int blocks = (int)((total + 255) / 256);
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  FR_CHECK(launch_pdl(ue_gather_dropout_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, news_vecs, inv_his,
                      (__nv_bfloat16*)X, rows, D / 4, p, seed, seed_dev));
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_ue_scatter_dx(const float* dX, const int* inv_his, float* d_news, int rows, int D, float p,
                            unsigned long long seed, float noise_std, unsigned long long noise_seed,
                            const unsigned long long* seed_dev, void* stream) {
  FR_REQUIRE(D % 4 == 0, "fr_ue_scatter_dx: D=%d", D);
  if (rows == 0) return 0;
  const size_t total = (size_t)rows * (D / 4);
  int blocks = (int)((total + 255) / 256);
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  FR_CHECK(launch_pdl(ue_scatter_dx_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, dX, inv_his, d_news, rows, D / 4, p,
                      seed, noise_std, noise_seed, seed_dev));
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_ue_mha_fwd(const void* qkv, const int* his_ids, void* ctx, float* stats, int B, int L, int H, int dk,
                         int use_mask, void* stream) {
  FR_REQUIRE(L >= 1 && L <= kMaxL, "fr_ue_mha_fwd: L=%d must be in [1,%d]", L, kMaxL);
  FR_REQUIRE(dk == 20 || dk == 32 || dk == 16, "fr_ue_mha_fwd: head dim %d not built (16, 20, 32)", dk);
  FR_REQUIRE(!use_mask || his_ids, "fr_ue_mha_fwd: mask needs his_ids");
  if (B == 0) return 0;
  const int blocks = B * H;
  const float scale = 1.f / sqrtf((float)dk);
  cudaStream_t s = (cudaStream_t)stream;
#define FR_GO(DK) FR_CHECK(launch_pdl(ue_mha_fwd_kernel<DK>, dim3(blocks), dim3(kJobThreads), 0, s, (const __nv_bfloat16*)qkv, his_ids, (__nv_bfloat16*)ctx, stats, B, L, H, scale, use_mask))
  if (dk == 20) FR_GO(20); else if (dk == 32) FR_GO(32); else FR_GO(16);
#undef FR_GO
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_ue_mha_bwd(const void* qkv, const void* dctx, const float* stats, const int* his_ids, void* dqkv, int B,
                         int L, int H, int dk, int use_mask, void* stream) {
  FR_REQUIRE(L >= 1 && L <= kMaxL, "fr_ue_mha_bwd: L=%d must be in [1,%d]", L, kMaxL);
  FR_REQUIRE(dk == 20 || dk == 32 || dk == 16, "fr_ue_mha_bwd: head dim %d not built (16, 20, 32)", dk);
  if (B == 0) return 0;
  const int blocks = B * H;
  const float scale = 1.f / sqrtf((float)dk);
  cudaStream_t s = (cudaStream_t)stream;
  const int dkp = (dk + 3) / 4 * 4, lp = L | 1;
  const size_t smem = sizeof(float) * ((size_t)4 * L * dkp + (size_t)2 * L * lp + (size_t)5 * L);
#define FR_GO(DK)                                                                                                        \
  do {                                                                                                                   \
    static bool attr_set = false;                                                                                        \
    if (!attr_set) {                                                                                                     \
      FR_CHECK(cudaFuncSetAttribute(ue_mha_bwd_kernel<DK>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024));    \
      attr_set = true;                                                                                                   \
    }                                                                                                                    \
    FR_CHECK(launch_pdl(ue_mha_bwd_kernel<DK>, dim3(blocks), dim3(kBwdThreads), smem, s, (const __nv_bfloat16*)qkv,   \
                        (const __nv_bfloat16*)dctx, stats, his_ids, (__nv_bfloat16*)dqkv, B, L, H, scale, use_mask));      \
  } while (0)
  if (dk == 20) FR_GO(20); else if (dk == 32) FR_GO(32); else FR_GO(16);
#undef FR_GO
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_ue_pool_score(const void* ctx, const void* e, const float* w2, const float* b2, const float* news_vecs,
                            const int* inv_cand, const int* his_ids, float* score, float* loss_mean, float* user_vec,
                            void* d_ctx_pool, void* d_pre, float* d_w2, float* d_b2, float* d_news, int B, int L, int C,
                            int D, int Q, int use_mask, int train, float noise_std, unsigned long long noise_seed,
                            const unsigned long long* seed_dev, void* stream) {
  FR_REQUIRE(L >= 1 && L <= kMaxL && C >= 1 && C <= 32 && D <= 512 && D % 4 == 0 && Q <= 256, "fr_ue_pool_score: bad shape L=%d C=%d D=%d Q=%d", L, C, D, Q);
  FR_REQUIRE(!train || (d_ctx_pool && d_pre && d_w2 && d_b2 && d_news), "fr_ue_pool_score: backward outputs missing");
  FR_REQUIRE(!use_mask || his_ids, "fr_ue_pool_score: mask needs his_ids");
  cudaStream_t s = (cudaStream_t)stream;
  if (loss_mean) FR_CHECK(cudaMemsetAsync(loss_mean, 0, sizeof(float), s));
  if (B == 0) return 0;
  PoolScoreArgs a;
  a.ctx = (const __nv_bfloat16*)ctx; a.e = (const __nv_bfloat16*)e; a.w2 = w2; a.b2 = b2; a.news_vecs = news_vecs;
  a.inv_cand = inv_cand; a.his_ids = his_ids; a.score = score; a.loss_mean = loss_mean; a.user_vec = user_vec;
  a.d_ctx_pool = (__nv_bfloat16*)d_ctx_pool; a.d_pre = (__nv_bfloat16*)d_pre; a.d_w2 = d_w2; a.d_b2 = d_b2; a.d_news = d_news;
  a.B = B; a.L = L; a.C = C; a.D = D; a.Q = Q; a.use_mask = use_mask; a.train = train; a.noise_std = noise_std;
  a.noise_seed = noise_seed; a.seed_dev = seed_dev;
  FR_CHECK(launch_pdl(ue_pool_score_kernel, dim3(B), dim3(kPoolThreads), 0, s, a));
  FR_LAUNCH_CHECK();
  return 0;
}

namespace fr {
int colsum_launch(const void* X, int is_bf16, const float* row_w, float* out, int M, int N, int ld, const int* m_dev,
                  int m_scale, cudaStream_t s) {
  if (M == 0 || N == 0) return 0;
  const int vec = is_bf16 ? 8 : 4;
  FR_REQUIRE(N % vec == 0 && ld % vec == 0 && ((uintptr_t)X & 15) == 0, "fr_colsum: N=%d, ld=%d must be multiples of %d and X 16-byte aligned", N, ld, vec);
  int rows_blocks = (M + 63) / 64;
  if (rows_blocks > 128) rows_blocks = 128;
  if (rows_blocks < 1) rows_blocks = 1;
  dim3 grid((N / vec + 31) / 32, rows_blocks);
  if (is_bf16) colsum_kernel<__nv_bfloat16, 8><<<grid, 256, 0, s>>>((const __nv_bfloat16*)X, row_w, out, M, N, ld, m_dev, m_scale);
  else colsum_kernel<float, 4><<<grid, 256, 0, s>>>((const float*)X, row_w, out, M, N, ld, m_dev, m_scale);
  FR_LAUNCH_CHECK();
  return 0;
}
}  // namespace fr

FR_API int fr_colsum(const void* X, int is_bf16, float* out, int M, int N, int ld, const int* m_dev, int m_scale,
                     void* stream) {
  return colsum_launch(X, is_bf16, nullptr, out, M, N, ld, m_dev, m_scale, (cudaStream_t)stream);
}

// out[n] += sum_m row_w[m] * X[m, n]  (row_w fp32 [M]); same row limit as fr_colsum.
FR_API int fr_colsum_weighted(const void* X, int is_bf16, const float* row_w, float* out, int M, int N, int ld,
                              const int* m_dev, int m_scale, void* stream) {
  FR_REQUIRE(row_w != nullptr, "fr_colsum_weighted: row_w is null");
  return colsum_launch(X, is_bf16, row_w, out, M, N, ld, m_dev, m_scale, (cudaStream_t)stream);
}

FR_API int fr_transpose_cast(const float* in, void* out, int R, int Ccols, int ld_out, void* stream) {
  FR_REQUIRE(ld_out >= R, "fr_transpose_cast: ld_out=%d < R=%d", ld_out, R);
  if (R == 0 || Ccols == 0) return 0;
  dim3 grid((Ccols + 31) / 32, (R + 31) / 32), block(32, 8);
  transpose_cast_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(in, (__nv_bfloat16*)out, R, Ccols, ld_out);
  FR_LAUNCH_CHECK();
  return 0;
}
