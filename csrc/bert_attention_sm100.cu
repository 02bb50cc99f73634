// N8 on the tensor cores: DistilBERT self-attention for short titles (T <= 64 tokens, 12 heads x 64).
//
// A title is far smaller than a tcgen05 tile, so several titles share one 128-row tile: every title gets a slot of
// SLOT = 32 (T <= 32) or 64 rows, TP = 128 / SLOT titles per tile, and the attention is block-diagonal inside it:
//   S[128 x 128] = Q_h K_h^T        tcgen05.mma, both operands K-major, straight from the QKV GEMM output via TMA
//                                   (one {64 x SLOT} box per title and operand, so title t sits at rows SLOT*t)
//   P = exp(S/8 - rowmax), masked to valid keys (HF: masked_fill(finfo.min) + softmax); a row only ever needs the
//       SLOT columns of its own title = exactly the TMEM chunk(s) its warp owns, kept in registers
//   O[128 x 64]  = P V_h            tcgen05.mma, A = P written to shared memory in the K-major 128B-swizzled layout
//                                   (only the diagonal blocks are rewritten; the rest was zeroed once),
//                                   B = V_h as loaded (tokens x 64 = MN-major)
//   out = O / rowsum                (deferred normalisation), bf16
// The block-diagonal waste (x TP in MMA work) is irrelevant: the tile is latency-bound, not throughput-bound.
// One CTA = 4 warps; thread r owns row r of the tile for the soft-max and the epilogue; lane 0 of warp 0 issues
// the TMA loads (one job ahead of the math) and both MMAs. 80 KB of shared memory and 256 TMEM columns per CTA,
// so two CTAs share an SM and hide each other's latencies. (Measured: a third CTA per SM, with O aliased onto S and
// 24 KB of P, did not shorten the kernel, so the SM's issue/TMEM path rather than latency is the limit.) Reference: HF MultiHeadSelfAttention reached from
// encoder.py:27.
#include "sm100.cuh"
#include <cuda.h>

namespace fr {
using namespace sm100;

int make_tmap_bf16_2d_box(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows,
                          int box_cols);

constexpr int kAttTile = 128;
constexpr int kAttHd = 64;
constexpr int kAttPBlk = 16384;                // bytes of one 64-column block of P
constexpr int kAttBar = 49152 + 32768;         // barriers follow Q, K, V (16 KB each) and P (32 KB)

__device__ __forceinline__ uint64_t att_desc_k_sw128(uint32_t addr) { return make_smem_desc_sw128(addr); }
__device__ __forceinline__ uint64_t att_desc_mn_sw128(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(1024 >> 4) << 16;      // LBO: single 64-wide atom along N, value unused
  d |= (uint64_t)(1024 >> 4) << 32;      // SBO: 8-token groups are 1024 bytes apart
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t att_idesc(int M, int N, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

struct AttParams {
  const int* mask_table;     // [table rows, T] or null
  const int* items;          // [n_items] table row of every title, or null (identity)
  const int* count_dev;      // optional device count of valid titles
  __nv_bfloat16* out;        // [n_items*T, heads*64]
  int n_items, T, heads;
};

template <int SLOT>
__global__ void __launch_bounds__(128, 2)
bert_attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const AttParams p) {
  constexpr int TP = kAttTile / SLOT;          // titles per tile
  constexpr int NCH = SLOT / 32;               // 32-column chunks of S a row needs
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sQ = smem;                     // 128 x 64 bf16, K-major SW128
  uint8_t* sK = smem + 16384;
  uint8_t* sV = smem + 32768;             // tokens x 64: MN-major SW128 for the second MMA
  uint8_t* sP = smem + 49152;             // 128 x 128 bf16 as two K-major SW128 blocks of 64 columns
  uint64_t* bar_qk = reinterpret_cast<uint64_t*>(smem + kAttBar);    // Q and K of the next (tile, head) have landed
  uint64_t* bar_v = bar_qk + 1;                                     // V of the next (tile, head) has landed
  uint64_t* bar_mma = bar_v + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
  float* sKeyBias = reinterpret_cast<float*>(smem + kAttBar + 64);  // [128] per key column: 0 or -inf, pre-scaled domain

  const int tid = threadIdx.x, warp = tid >> 5;
  const int T = p.T, heads = p.heads;
  // Launched with the PDL attribute: the CTA may already be resident while the QKV GEMM drains (it cannot share an SM
  // with a GEMM CTA: shared memory), and from here on that GEMM has completed.
  griddep_wait();
  const int n_eff = p.count_dev ? min(p.n_items, *p.count_dev) : p.n_items;
  const int num_tiles = (n_eff + TP - 1) / TP;
  const int hd_all = heads * kAttHd;

  if (tid == 0) {
    prefetch_tmap(&tmap_qkv);
    mbar_init(bar_qk, 1);
    mbar_init(bar_v, 1);
    mbar_init(bar_mma, 1);
    fence_barrier_init();
  }
  // off-diagonal blocks of P stay zero for the whole kernel
  {
    const uint32_t pz = smem_u32(sP) + (uint32_t)tid * 256;
#pragma unroll
    for (int c = 0; c < 16; ++c)
      asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(pz + (uint32_t)c * 16), "r"(0u) : "memory");
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 256);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // The out-projection GEMM may start its prologue from here on. Not earlier: its CTAs allocate all 512 TMEM columns
  // and then wait for this grid, so one that lands on this SM before our own allocation would deadlock both.
  griddep_launch();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + 128;
  uint32_t ph_qk = 0, ph_v = 0, ph_mma = 0;

  // Loads run one (tile, head) ahead of the math: Q/K of the next job are requested as soon as S = QK^T has been
  // produced (their buffers are free), V as soon as O = PV has been produced. Title t of a tile lands at rows SLOT*t.
  auto load_qk = [&](int tile_, int h_) {
    mbar_arrive_expect_tx(bar_qk, 2 * 16384);
#pragma unroll
    for (int t = 0; t < TP; ++t) {
      const int r0 = (tile_ * TP + t) * T;
      tma_load_2d(sQ + t * SLOT * 128, &tmap_qkv, bar_qk, h_ * kAttHd, r0);
      tma_load_2d(sK + t * SLOT * 128, &tmap_qkv, bar_qk, hd_all + h_ * kAttHd, r0);
    }
  };
  auto load_v = [&](int tile_, int h_) {
    mbar_arrive_expect_tx(bar_v, 16384);
#pragma unroll
    for (int t = 0; t < TP; ++t)
      tma_load_2d(sV + t * SLOT * 128, &tmap_qkv, bar_v, 2 * hd_all + h_ * kAttHd, (tile_ * TP + t) * T);
  };
  // Work unit = one (tile, head) job; every CTA takes a contiguous, balanced range of them so that the grid stays
  // full when there are fewer tiles than CTA slots (tiles * heads / gridDim jobs each instead of whole tiles).
  const long long total_jobs = (long long)num_tiles * heads;
  const int job_begin = (int)(total_jobs * blockIdx.x / gridDim.x);
  const int job_end = (int)(total_jobs * (blockIdx.x + 1) / gridDim.x);
  if (tid == 0 && job_begin < job_end) {
    load_qk(job_begin / heads, job_begin % heads);
    load_v(job_begin / heads, job_begin % heads);
  }

  const int my_slot = tid / SLOT;                     // title slot of my row
  const int my_tok = tid - my_slot * SLOT;            // token index of my row inside its title
  const int col0 = my_slot * SLOT;                    // first key column of my title = first column I read from S
  const uint32_t lane_base = ((uint32_t)(warp * 32) << 16);
  constexpr float kScaleLog2 = 0.125f * 1.4426950408889634f;     // 1/sqrt(64) * log2(e): soft-max in the exp2 domain

  int cur_tile = -1;
  bool my_uniform = false, row_ok = false;
  size_t out_row = 0;
  for (int job = job_begin; job < job_end; ++job) {
    const int tile = job / heads, h = job - tile * heads;
    if (tile != cur_tile) {
      cur_tile = tile;
      const int title0 = tile * TP;
      // key bias for the 128 key columns of this tile (key column tid and row tid belong to the same title slot)
      my_uniform = false;
      float bias = -INFINITY;
      if (my_tok < T && title0 + my_slot < n_eff) {
        const int item = p.items ? p.items[title0 + my_slot] : (title0 + my_slot);
        const int* mrow = p.mask_table ? p.mask_table + (size_t)item * T : nullptr;
        bool any = false;
        if (mrow) for (int j = 0; j < T; ++j) any |= (mrow[j] != 0);
        my_uniform = mrow && !any;                    // fully masked title: HF's soft-max of a constant row = uniform
        bias = (!mrow || mrow[my_tok] != 0 || my_uniform) ? 0.f : -INFINITY;
      }
      sKeyBias[tid] = bias;                           // the previous job ended with __syncthreads: no reader is left
      __syncthreads();                                // sKeyBias is complete (every valid row of a title computed the same flag)
      row_ok = my_tok < T && (title0 + my_slot) < n_eff;
      out_row = (size_t)(title0 + my_slot) * T + my_tok;
    }
    {
      // the job after this one
      const int njob = job + 1;
      const bool has_next = njob < job_end;
      const int ntile = njob / heads, nh = njob - ntile * heads;
      // Only the MMA-issuing thread consumes Q/K, so only it waits for them. Round 1 had every thread wait here: thread 0
      // could then pass, run S = QK^T, and re-arm this barrier for the next job (load_qk below) before a slower warp had
      // looked at it; once that second phase completed, the parity the slow warp was polling for read "not yet" again,
      // the warp never left, and thread 0 stopped at the __syncthreads below: a hang that needed an unlucky few
      // microseconds of warp scheduling (measured: 1 run in 3 with a high-priority branch sharing the SMs; consistent with
      // the round-1 4-GPU scaling failure, where one rank sat in a step kernel while its peers timed out waiting for it).
      // bar_mma is waited on by everyone, but its two completions per job are separated by a __syncthreads each.
      if (tid == 0) {
        mbar_wait(bar_qk, ph_qk, 1);
        tc_fence_after();
        constexpr uint32_t id1 = att_idesc(128, 128, 0);
        const uint32_t q = smem_u32(sQ), k = smem_u32(sK);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          umma_bf16_ss(tS, att_desc_k_sw128(q + ks * 32), att_desc_k_sw128(k + ks * 32), id1, ks > 0);
        umma_commit(bar_mma);
      }
      ph_qk ^= 1;
      mbar_wait(bar_mma, ph_mma, 2);
      ph_mma ^= 1;
      tc_fence_after();
      if (tid == 0 && has_next) load_qk(ntile, nh);          // sQ / sK are free again

      // ---- soft-max over my row: only my title's SLOT key columns, held in registers ----
      float t[NCH * 32];
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tS + lane_base + (uint32_t)(col0 + c * 32), r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 kb = *reinterpret_cast<const float4*>(&sKeyBias[col0 + c * 32 + j]);
          t[c * 32 + j + 0] = my_uniform ? kb.x : fmaf(__uint_as_float(r[j + 0]), kScaleLog2, kb.x);
          t[c * 32 + j + 1] = my_uniform ? kb.y : fmaf(__uint_as_float(r[j + 1]), kScaleLog2, kb.y);
          t[c * 32 + j + 2] = my_uniform ? kb.z : fmaf(__uint_as_float(r[j + 2]), kScaleLog2, kb.z);
          t[c * 32 + j + 3] = my_uniform ? kb.w : fmaf(__uint_as_float(r[j + 3]), kScaleLog2, kb.w);
        }
      }
#pragma unroll
      for (int j = 0; j < NCH * 32; ++j) mx = fmaxf(mx, t[j]);
      if (mx == -INFINITY) mx = 0.f;                         // rows outside any valid title: keep the math finite
      float l = 0.f;
      const uint32_t prow = smem_u32(sP) + (uint32_t)tid * 128;
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        uint32_t packed[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float e0 = exp2f(t[c * 32 + 2 * j] - mx), e1 = exp2f(t[c * 32 + 2 * j + 1] - mx);
          l += e0 + e1;
          const __nv_bfloat162 b2 = __floats2bfloat162_rn(e0, e1);
          packed[j] = *reinterpret_cast<const uint32_t*>(&b2);
        }
        // 32 columns = 4 chunks of 16 bytes inside 64-column block; chunk index XOR (row % 8) = 128B swizzle
        const int cc = col0 + c * 32;
        const uint32_t blk = prow + (uint32_t)(cc >> 6) * kAttPBlk;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const uint32_t chunk = (uint32_t)(((cc & 63) >> 3) + q4) ^ (uint32_t)(tid & 7);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(blk + chunk * 16), "r"(packed[4 * q4]),
                       "r"(packed[4 * q4 + 1]), "r"(packed[4 * q4 + 2]), "r"(packed[4 * q4 + 3])
                       : "memory");
        }
      }
      fence_proxy_async();                 // generic-proxy writes of P → visible to the tensor core (async proxy)
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        mbar_wait(bar_v, ph_v, 3);
        tc_fence_after();
        constexpr uint32_t id2 = att_idesc(128, 64, 1);
        const uint32_t pa = smem_u32(sP), v = smem_u32(sV);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          umma_bf16_ss(tO, att_desc_k_sw128(pa + (ks >> 2) * kAttPBlk + (ks & 3) * 32), att_desc_mn_sw128(v + ks * 2048), id2, ks > 0);
        umma_commit(bar_mma);
      }
      ph_v ^= 1;
      mbar_wait(bar_mma, ph_mma, 4);
      ph_mma ^= 1;
      tc_fence_after();
      if (tid == 0 && has_next) load_v(ntile, nh);           // sV is free again
      // ---- epilogue: O / l → bf16 → global ----
      {
        const float inv = 1.f / l;
        __nv_bfloat16* op = p.out + out_row * hd_all + h * kAttHd;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(tO + lane_base + c * 32, r);
          tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint4 qv;
              __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&qv);
#pragma unroll
              for (int u = 0; u < 4; ++u)
                hh[u] = __floats2bfloat162_rn(__uint_as_float(r[j + 2 * u]) * inv, __uint_as_float(r[j + 2 * u + 1]) * inv);
              *reinterpret_cast<uint4*>(op + c * 32 + j) = qv;
            }
          }
        }
      }
      tc_fence_before();
      __syncthreads();                     // TMEM and shared tiles are free for the next head
    }
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// N4 on the tensor cores: the user encoder's multi-head self-attention over the click history (reference:
// attention.py:32-45 ScaledDotProductAttention, :69-81 MultiHeadAttention; 20 heads x 20 dims over L <= 64 clicks).
//
// Same block-diagonal tiling as the title kernel above with 64-row slots: two users share a 128-row tile,
//   S[128 x 128] = Q_h K_h^T (head dim zero-padded to 32: two UMMA K steps),  O[128 x 64] = P V_h (dims >= dk are zero).
// What differs from the title attention:
//   * the reference soft-max has no max subtraction and a +1e-8 in the denominator (attention.py:39-42). Subtracting the
//     row maximum m for range, the same value is  exp(s - m) / (sum_j exp(s_j - m) mask_j + 1e-8 exp(-m)) ;
//   * a masked history slot multiplies its probability by zero (attention.py:40-41) instead of adding -inf to its score,
//     so it still takes part in the maximum; key columns outside the history (j >= L) are excluded outright;
//   * a head is 20 bf16 = 40 bytes wide, which no TMA box can address (boxes are multiples of 16 bytes), so each thread
//     copies its own row of Q, K and V into the 128B-swizzled K-major layout with plain 8-byte loads (15 KB per job);
//   * (row maximum, denominator) are stored per query for the backward kernel, exactly like the CUDA-core forward.
// One CTA = 4 warps, thread r owns row r; 80 KB shared memory, 256 TMEM columns, two CTAs per SM.
// ---------------------------------------------------------------------------------------------------------------------
struct UeAttParams {
  const __nv_bfloat16* qkv;   // [B*L, 3*H*dk]  (q | k | v)
  const int* his_ids;         // [B*L] or null: id 0 = padding slot (only read with use_mask)
  __nv_bfloat16* ctx;         // [B*L, H*dk]
  float* stats;               // [B*H*L, 2] = (row maximum in natural-log units, denominator)
  int B, L, H, dk, use_mask;
  float scale;                // 1 / sqrt(dk)
};

__global__ void __launch_bounds__(128, 2) ue_attention_tc_kernel(const UeAttParams p) {
  constexpr int SLOT = 64, TP = 2, NCH = 2;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sQ = smem;
  uint8_t* sK = smem + 16384;
  uint8_t* sV = smem + 32768;
  uint8_t* sP = smem + 49152;
  uint64_t* bar_mma = reinterpret_cast<uint64_t*>(smem + kAttBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
  float* sKeyBias = reinterpret_cast<float*>(smem + kAttBar + 64);   // [128] 0 or -inf (key column outside the history)
  float* sKeyMul = sKeyBias + 128;                                   // [128] 0 (masked padding slot) or 1

  const int tid = threadIdx.x, warp = tid >> 5;
  griddep_wait();                                   // launched with the PDL attribute: the QKV GEMM is complete from here on
  const int B = p.B, L = p.L, H = p.H, dk = p.dk;
  const int ld = 3 * H * dk, hd = H * dk;
  if (tid == 0) {
    mbar_init(bar_mma, 1);
    fence_barrier_init();
  }
  {   // Q, K, V and P start as zeros: padding dims, rows outside the history and P's off-diagonal blocks stay zero
    const uint32_t z0 = smem_u32(smem) + (uint32_t)tid * 16;
    for (int c = 0; c < (49152 + 32768) / (128 * 16); ++c)
      asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(z0 + (uint32_t)c * 2048), "r"(0u) : "memory");
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 256);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  griddep_launch();                                 // after our TMEM allocation (see the title kernel)
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + 128;
  uint32_t ph_mma = 0;

  const int num_tiles = (B + TP - 1) / TP;
  const long long total_jobs = (long long)num_tiles * H;
  const int job_begin = (int)(total_jobs * blockIdx.x / gridDim.x);
  const int job_end = (int)(total_jobs * (blockIdx.x + 1) / gridDim.x);
  const int my_slot = tid / SLOT, my_tok = tid - my_slot * SLOT, col0 = my_slot * SLOT;
  const uint32_t lane_base = ((uint32_t)(warp * 32) << 16);
  const float scale_log2 = p.scale * 1.4426950408889634f;
  const uint32_t row_off = (uint32_t)tid * 128;     // my row inside a K-major 128B-swizzled tile
  const uint32_t swz = (uint32_t)(tid & 7);
  const int nquads = dk >> 2;                       // 8-byte groups of a head row (dk % 4 == 0)

  int cur_tile = -1;
  bool row_ok = false;
  size_t grow = 0;
  for (int job = job_begin; job < job_end; ++job) {
    const int tile = job / H, h = job - tile * H;
    if (tile != cur_tile) {
      cur_tile = tile;
      const int user = tile * TP + my_slot;
      row_ok = my_tok < L && user < B;
      grow = (size_t)user * L + my_tok;
      sKeyBias[tid] = row_ok ? 0.f : -INFINITY;     // key column tid and row tid belong to the same (user, token)
      sKeyMul[tid] = (row_ok && !(p.use_mask && p.his_ids[grow] == 0)) ? 1.f : 0.f;
    }
    // ---- my row of Q, K, V: 8-byte loads, 16-byte swizzled stores (chunks 0..3 = 32 dims; dims >= dk are zero) ----
    {
      uint32_t w[3][16];
#pragma unroll
      for (int o = 0; o < 3; ++o)
#pragma unroll
        for (int j = 0; j < 16; ++j) w[o][j] = 0u;
      if (row_ok) {
        const uint2* src = reinterpret_cast<const uint2*>(p.qkv + grow * ld + h * dk);
#pragma unroll
        for (int o = 0; o < 3; ++o) {
          const uint2* so = src + (size_t)o * (hd >> 2);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            if (q < nquads) {
              const uint2 v = __ldg(so + q);
              w[o][2 * q] = v.x;
              w[o][2 * q + 1] = v.y;
            }
          }
        }
      }
      const uint32_t base[3] = {smem_u32(sQ) + row_off, smem_u32(sK) + row_off, smem_u32(sV) + row_off};
#pragma unroll
      for (int o = 0; o < 3; ++o)
#pragma unroll
        for (int c = 0; c < 4; ++c)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(base[o] + (((uint32_t)c ^ swz) << 4)),
                       "r"(w[o][4 * c]), "r"(w[o][4 * c + 1]), "r"(w[o][4 * c + 2]), "r"(w[o][4 * c + 3]) : "memory");
    }
    fence_proxy_async();                            // generic-proxy writes → visible to the tensor core
    tc_fence_before();
    __syncthreads();                                // also publishes sKeyBias / sKeyMul
    if (tid == 0) {
      tc_fence_after();
      constexpr uint32_t id1 = att_idesc(128, 128, 0);
      const uint32_t q = smem_u32(sQ), k = smem_u32(sK);
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)                // head dim padded to 32 = two K steps of 16
        umma_bf16_ss(tS, att_desc_k_sw128(q + ks * 32), att_desc_k_sw128(k + ks * 32), id1, ks > 0);
      umma_commit(bar_mma);
    }
    mbar_wait(bar_mma, ph_mma, 5);
    ph_mma ^= 1;
    tc_fence_after();
    // ---- soft-max over my row: my user's 64 key columns ----
    float t[NCH * 32];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(tS + lane_base + (uint32_t)(col0 + c * 32), r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) t[c * 32 + j] = fmaf(__uint_as_float(r[j]), scale_log2, sKeyBias[col0 + c * 32 + j]);
    }
#pragma unroll
    for (int j = 0; j < NCH * 32; ++j) mx = fmaxf(mx, t[j]);
    if (mx == -INFINITY) mx = 0.f;                  // rows outside the history: keep the math finite
    float l = 0.f;
    const uint32_t prow = smem_u32(sP) + (uint32_t)tid * 128;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      uint32_t packed[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float e0 = exp2f(t[c * 32 + 2 * j] - mx) * sKeyMul[col0 + c * 32 + 2 * j];
        const float e1 = exp2f(t[c * 32 + 2 * j + 1] - mx) * sKeyMul[col0 + c * 32 + 2 * j + 1];
        l += e0 + e1;
        const __nv_bfloat162 b2 = __floats2bfloat162_rn(e0, e1);
        packed[j] = *reinterpret_cast<const uint32_t*>(&b2);
      }
      const int cc = col0 + c * 32;
      const uint32_t blk = prow + (uint32_t)(cc >> 6) * kAttPBlk;
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {
        const uint32_t chunk = (uint32_t)(((cc & 63) >> 3) + q4) ^ swz;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(blk + chunk * 16), "r"(packed[4 * q4]),
                     "r"(packed[4 * q4 + 1]), "r"(packed[4 * q4 + 2]), "r"(packed[4 * q4 + 3]) : "memory");
      }
    }
    const float z = l + 1e-8f * exp2f(-mx);         // the reference's "+ 1e-8", carried into the max-subtracted form
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      constexpr uint32_t id2 = att_idesc(128, 64, 1);
      const uint32_t pa = smem_u32(sP), v = smem_u32(sV);
#pragma unroll
      for (int ks = 0; ks < 8; ++ks)
        umma_bf16_ss(tO, att_desc_k_sw128(pa + (ks >> 2) * kAttPBlk + (ks & 3) * 32), att_desc_mn_sw128(v + ks * 2048), id2, ks > 0);
      umma_commit(bar_mma);
    }
    mbar_wait(bar_mma, ph_mma, 6);
    ph_mma ^= 1;
    tc_fence_after();
    {
      uint32_t r[32];
      tmem_ld_32x32(tO + lane_base, r);             // dims 0..31 of my row; only the first dk are real
      tmem_ld_wait();
      if (row_ok) {
        const float inv = 1.f / z;
        uint2* op = reinterpret_cast<uint2*>(p.ctx + grow * hd + h * dk);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (q < nquads) {
            const __nv_bfloat162 a = __floats2bfloat162_rn(__uint_as_float(r[4 * q]) * inv, __uint_as_float(r[4 * q + 1]) * inv);
            const __nv_bfloat162 b = __floats2bfloat162_rn(__uint_as_float(r[4 * q + 2]) * inv, __uint_as_float(r[4 * q + 3]) * inv);
            op[q] = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
          }
        }
        const int user = tile * TP + my_slot;
        float2* sp = reinterpret_cast<float2*>(p.stats) + ((size_t)(user * H + h) * L + my_tok);
        *sp = make_float2(mx * 0.6931471805599453f, z);       // maximum back in natural-log units, as the backward expects
      }
    }
    tc_fence_before();
    __syncthreads();                                // TMEM and the shared tiles are free for the next job
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// N5 on the tensor cores: backward of the user encoder's self-attention. dctx bf16 [B*L, H*dk] -> dqkv bf16
// [B*L, 3*H*dk], probabilities rebuilt from q, k and the forward's (row maximum, denominator). Autograd of
// attention.py:32-45 for the click history.
//
// Per (two-user tile, head), every product with the two operand forms the forward already uses (A K-major x B K-major,
// A K-major x B MN-major):
//   S  = Q K^T                     -> P_ij = exp(s_ij - m_i) mask_j / z_i         (thread i owns row i: registers)
//   dP = dO V^T                       dS_ij = P_ij (dP_ij - sum_j P_ij dP_ij)
//   dQ = scale * dS K                 A = dS as written row by row (K-major),              B = K as loaded (MN-major)
//   dK = scale * dS^T Q               A = dS^T: thread i writes its row as column i,       B = Q as loaded (MN-major)
//   dV = P^T dO                       A = P^T, same transposed write,                      B = dO as loaded (MN-major)
// The transposed writes are 2-byte shared-memory stores whose 32 lanes cover 64 contiguous bytes of one row: no bank
// conflicts. Off-diagonal (other user's) blocks of dS, dS^T and P^T are zeroed once and never touched.
// 160 KB of shared memory (Q, K, V, dO 16 KB each; dS, dS^T, P^T 32 KB each) and 448 TMEM columns: one CTA per SM.
// ---------------------------------------------------------------------------------------------------------------------
struct UeAttBwdParams {
  const __nv_bfloat16* qkv;   // [B*L, 3*H*dk]
  const __nv_bfloat16* dctx;  // [B*L, H*dk]
  const float* stats;         // [B*H*L, 2] from the forward
  const int* his_ids;         // [B*L] or null
  __nv_bfloat16* dqkv;        // [B*L, 3*H*dk]
  int B, L, H, dk, use_mask;
  float scale;
};
constexpr int kUeBwdTile = 16384;                    // one 128 x 64 bf16 operand tile
constexpr int kUeBwdMat = 32768;                     // one 128 x 128 bf16 matrix (two 64-column blocks)
constexpr int kUeBwdBar = 4 * kUeBwdTile + 3 * kUeBwdMat;

__global__ void __launch_bounds__(128, 1) ue_attention_bwd_tc_kernel(const UeAttBwdParams p) {
  constexpr int SLOT = 64, TP = 2;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sQ = smem;
  uint8_t* sK = smem + kUeBwdTile;
  uint8_t* sV = smem + 2 * kUeBwdTile;
  uint8_t* sG = smem + 3 * kUeBwdTile;               // dO
  uint8_t* sDS = smem + 4 * kUeBwdTile;              // dS   [query][key]
  uint8_t* sDST = sDS + kUeBwdMat;                   // dS^T [key][query]
  uint8_t* sPT = sDST + kUeBwdMat;                   // P^T  [key][query]
  uint64_t* bar_mma = reinterpret_cast<uint64_t*>(smem + kUeBwdBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
  float* sKeyBias = reinterpret_cast<float*>(smem + kUeBwdBar + 64);
  float* sKeyMul = sKeyBias + 128;

  const int tid = threadIdx.x, warp = tid >> 5;
  griddep_wait();
  const int B = p.B, L = p.L, H = p.H, dk = p.dk;
  const int ld = 3 * H * dk, hd = H * dk;
  if (tid == 0) {
    mbar_init(bar_mma, 1);
    fence_barrier_init();
  }
  {
    const uint32_t z0 = smem_u32(smem) + (uint32_t)tid * 16;
    for (int c = 0; c < kUeBwdBar / (128 * 16); ++c)
      asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(z0 + (uint32_t)c * 2048), "r"(0u) : "memory");
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  griddep_launch();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tDP = tmem + 128, tDQ = tmem + 256, tDK = tmem + 320, tDV = tmem + 384;
  uint32_t ph_mma = 0;

  const int num_tiles = (B + TP - 1) / TP;
  const long long total_jobs = (long long)num_tiles * H;
  const int job_begin = (int)(total_jobs * blockIdx.x / gridDim.x);
  const int job_end = (int)(total_jobs * (blockIdx.x + 1) / gridDim.x);
  const int my_slot = tid / SLOT, my_tok = tid - my_slot * SLOT, col0 = my_slot * SLOT;
  const uint32_t lane_base = ((uint32_t)(warp * 32) << 16);
  const float scale_log2 = p.scale * 1.4426950408889634f;
  const uint32_t row_off = (uint32_t)tid * 128;
  const uint32_t swz = (uint32_t)(tid & 7);
  const int nquads = dk >> 2;
  // my element (row j, column tid) of a transposed 128 x 128 matrix: 64-column block tid / 64, 16-byte chunk
  // ((tid % 64) / 8) ^ (j % 8) of row j, bf16 (tid % 8) inside it
  const uint32_t t_blk = (uint32_t)(tid >> 6) * 16384u, t_chunk = (uint32_t)((tid & 63) >> 3), t_in = (uint32_t)(tid & 7) * 2u;

  int cur_tile = -1;
  bool row_ok = false;
  size_t grow = 0;
  for (int job = job_begin; job < job_end; ++job) {
    const int tile = job / H, h = job - tile * H;
    if (tile != cur_tile) {
      cur_tile = tile;
      const int user = tile * TP + my_slot;
      row_ok = my_tok < L && user < B;
      grow = (size_t)user * L + my_tok;
      sKeyBias[tid] = row_ok ? 0.f : -INFINITY;
      sKeyMul[tid] = (row_ok && !(p.use_mask && p.his_ids[grow] == 0)) ? 1.f : 0.f;
    }
    // ---- my row of Q, K, V, dO ----
    {
      uint32_t w[4][16];
#pragma unroll
      for (int o = 0; o < 4; ++o)
#pragma unroll
        for (int j = 0; j < 16; ++j) w[o][j] = 0u;
      if (row_ok) {
        const uint2* src = reinterpret_cast<const uint2*>(p.qkv + grow * ld + h * dk);
        const uint2* gsrc = reinterpret_cast<const uint2*>(p.dctx + grow * hd + h * dk);
#pragma unroll
        for (int o = 0; o < 4; ++o) {
          const uint2* so = o < 3 ? src + (size_t)o * (hd >> 2) : gsrc;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            if (q < nquads) {
              const uint2 v = __ldg(so + q);
              w[o][2 * q] = v.x;
              w[o][2 * q + 1] = v.y;
            }
          }
        }
      }
      const uint32_t base[4] = {smem_u32(sQ) + row_off, smem_u32(sK) + row_off, smem_u32(sV) + row_off, smem_u32(sG) + row_off};
#pragma unroll
      for (int o = 0; o < 4; ++o)
#pragma unroll
        for (int c = 0; c < 4; ++c)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(base[o] + (((uint32_t)c ^ swz) << 4)),
                       "r"(w[o][4 * c]), "r"(w[o][4 * c + 1]), "r"(w[o][4 * c + 2]), "r"(w[o][4 * c + 3]) : "memory");
    }
    float m2 = 0.f, invz = 0.f;
    if (row_ok) {
      const float2 st = *(reinterpret_cast<const float2*>(p.stats) + ((size_t)((tile * TP + my_slot) * H + h) * L + my_tok));
      m2 = st.x * 1.4426950408889634f;
      invz = 1.f / st.y;
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      constexpr uint32_t id1 = att_idesc(128, 128, 0);
      const uint32_t q = smem_u32(sQ), k = smem_u32(sK), v = smem_u32(sV), g = smem_u32(sG);
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        umma_bf16_ss(tS, att_desc_k_sw128(q + ks * 32), att_desc_k_sw128(k + ks * 32), id1, ks > 0);
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        umma_bf16_ss(tDP, att_desc_k_sw128(g + ks * 32), att_desc_k_sw128(v + ks * 32), id1, ks > 0);
      umma_commit(bar_mma);
    }
    mbar_wait(bar_mma, ph_mma, 7);
    ph_mma ^= 1;
    tc_fence_after();
    // ---- P, dP -> dS for my row (my user's 64 key columns); dS row-wise, dS^T and P^T column-wise ----
    {
      float pr[64];
      float dd = 0.f;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tS + lane_base + (uint32_t)(col0 + c * 32), r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float t = fmaf(__uint_as_float(r[j]), scale_log2, sKeyBias[col0 + c * 32 + j]);
          pr[c * 32 + j] = row_ok ? exp2f(t - m2) * sKeyMul[col0 + c * 32 + j] * invz : 0.f;
        }
      }
      float dp[64];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tDP + lane_base + (uint32_t)(col0 + c * 32), r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          dp[c * 32 + j] = __uint_as_float(r[j]);
          dd = fmaf(pr[c * 32 + j], dp[c * 32 + j], dd);
        }
      }
      const uint32_t ds_row = smem_u32(sDS) + (uint32_t)my_slot * 16384u + row_off;
      const uint32_t dst_base = smem_u32(sDST) + t_blk + t_in, pt_base = smem_u32(sPT) + t_blk + t_in;
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {               // 8 key columns = one 16-byte chunk of my dS row
        uint32_t packed[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = c8 * 8 + 2 * u;
          const float d0 = pr[j] * (dp[j] - dd), d1 = pr[j + 1] * (dp[j + 1] - dd);
          const __nv_bfloat162 b2 = __floats2bfloat162_rn(d0, d1);
          packed[u] = *reinterpret_cast<const uint32_t*>(&b2);
          const __nv_bfloat162 p2 = __floats2bfloat162_rn(pr[j], pr[j + 1]);
          // transposed copies: element (key col0 + j, query tid)
          const uint32_t kr0 = (uint32_t)(col0 + j), kr1 = kr0 + 1u;
          const uint32_t o0 = kr0 * 128u + ((t_chunk ^ (kr0 & 7u)) << 4), o1 = kr1 * 128u + ((t_chunk ^ (kr1 & 7u)) << 4);
          asm volatile("st.shared.b16 [%0], %1;" ::"r"(dst_base + o0), "h"(__bfloat16_as_ushort(b2.x)) : "memory");
          asm volatile("st.shared.b16 [%0], %1;" ::"r"(dst_base + o1), "h"(__bfloat16_as_ushort(b2.y)) : "memory");
          asm volatile("st.shared.b16 [%0], %1;" ::"r"(pt_base + o0), "h"(__bfloat16_as_ushort(p2.x)) : "memory");
          asm volatile("st.shared.b16 [%0], %1;" ::"r"(pt_base + o1), "h"(__bfloat16_as_ushort(p2.y)) : "memory");
        }
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(ds_row + (((uint32_t)c8 ^ swz) << 4)), "r"(packed[0]),
                     "r"(packed[1]), "r"(packed[2]), "r"(packed[3]) : "memory");
      }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      constexpr uint32_t id2 = att_idesc(128, 64, 1);
      const uint32_t ds = smem_u32(sDS), dst = smem_u32(sDST), pt = smem_u32(sPT);
      const uint32_t q = smem_u32(sQ), k = smem_u32(sK), g = smem_u32(sG);
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint32_t a_off = (uint32_t)(ks >> 2) * 16384u + (uint32_t)(ks & 3) * 32u;
        umma_bf16_ss(tDQ, att_desc_k_sw128(ds + a_off), att_desc_mn_sw128(k + ks * 2048), id2, ks > 0);
        umma_bf16_ss(tDK, att_desc_k_sw128(dst + a_off), att_desc_mn_sw128(q + ks * 2048), id2, ks > 0);
        umma_bf16_ss(tDV, att_desc_k_sw128(pt + a_off), att_desc_mn_sw128(g + ks * 2048), id2, ks > 0);
      }
      umma_commit(bar_mma);
    }
    mbar_wait(bar_mma, ph_mma, 8);
    ph_mma ^= 1;
    tc_fence_after();
    {
      uint2* op = reinterpret_cast<uint2*>(p.dqkv + grow * ld + h * dk);
#pragma unroll
      for (int o = 0; o < 3; ++o) {
        uint32_t r[32];
        tmem_ld_32x32((o == 0 ? tDQ : (o == 1 ? tDK : tDV)) + lane_base, r);
        tmem_ld_wait();
        const float f = o < 2 ? p.scale : 1.f;
        if (row_ok) {
          uint2* oo = op + (size_t)o * (hd >> 2);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            if (q < nquads) {
              const __nv_bfloat162 a = __floats2bfloat162_rn(__uint_as_float(r[4 * q]) * f, __uint_as_float(r[4 * q + 1]) * f);
              const __nv_bfloat162 b = __floats2bfloat162_rn(__uint_as_float(r[4 * q + 2]) * f, __uint_as_float(r[4 * q + 3]) * f);
              oo[q] = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
            }
          }
        }
      }
    }
    tc_fence_before();
    __syncthreads();
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

}  // namespace fr

using namespace fr;

// Tensor-core version of fr_bert_attention (same contract). T <= 64, head_dim 64.
FR_API int fr_bert_attention_tc(const void* qkv, const int* mask_table, const int* items, const int* count_dev,
                                int n_items, int T, int heads, int head_dim, void* out, void* stream) {
  FR_REQUIRE(head_dim == kAttHd, "fr_bert_attention_tc: head_dim=%d (only 64 is built)", head_dim);
  FR_REQUIRE(T >= 1 && T <= 64, "fr_bert_attention_tc: T=%d must be in [1,64]", T);
  if (n_items == 0) return 0;
  const int slot = T <= 32 ? 32 : 64;
  const int TP = kAttTile / slot;
  CUtensorMap tm;
  if (int e = make_tmap_bf16_2d_box(&tm, qkv, (long long)n_items * T, 3LL * heads * kAttHd, 3LL * heads * kAttHd, slot, 64)) return e;
  AttParams p;
  p.mask_table = mask_table; p.items = items; p.count_dev = count_dev; p.out = (__nv_bfloat16*)out;
  p.n_items = n_items; p.T = T; p.heads = heads;
  constexpr int smem = kAttBar + 64 + 512 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    FR_CHECK(cudaFuncSetAttribute(bert_attention_tc_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    FR_CHECK(cudaFuncSetAttribute(bert_attention_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  long long jobs = (long long)((n_items + TP - 1) / TP) * heads;      // (tile, head) jobs, split evenly over the CTAs
  int grid = jobs < 2 * num_sms() ? (int)jobs : 2 * num_sms();
  if (slot == 32) FR_CHECK(launch_pdl(bert_attention_tc_kernel<32>, dim3(grid), dim3(128), smem, (cudaStream_t)stream, tm, p));
  else FR_CHECK(launch_pdl(bert_attention_tc_kernel<64>, dim3(grid), dim3(128), smem, (cudaStream_t)stream, tm, p));
  return 0;
}

// The user encoder's attention on the tensor cores (same contract as fr_ue_mha_fwd in user_encoder.cu, which stays as
// the cross-check): L <= 64, dk in {4, 8, .., 32}. Replaces attention.py:32-45 / :69-81 for the click history.
FR_API int fr_ue_attention_tc(const void* qkv, const int* his_ids, void* ctx, float* stats, int B, int L, int H, int dk,
                              int use_mask, void* stream) {
  FR_REQUIRE(L >= 1 && L <= 64, "fr_ue_attention_tc: L=%d must be in [1,64]", L);
  FR_REQUIRE(dk >= 4 && dk <= 32 && dk % 4 == 0, "fr_ue_attention_tc: dk=%d must be a multiple of 4 in [4,32]", dk);
  FR_REQUIRE(!use_mask || his_ids != nullptr, "fr_ue_attention_tc: use_mask needs his_ids");
  if (B == 0) return 0;
  UeAttParams p;
  p.qkv = (const __nv_bfloat16*)qkv; p.his_ids = his_ids; p.ctx = (__nv_bfloat16*)ctx; p.stats = stats;
  p.B = B; p.L = L; p.H = H; p.dk = dk; p.use_mask = use_mask; p.scale = 1.f / sqrtf((float)dk);
  constexpr int smem = kAttBar + 64 + 1024 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    FR_CHECK(cudaFuncSetAttribute(ue_attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  const long long jobs = (long long)((B + 1) / 2) * H;
  const int grid = jobs < 2LL * num_sms() ? (int)jobs : 2 * num_sms();
  FR_CHECK(launch_pdl(ue_attention_tc_kernel, dim3(grid), dim3(128), smem, (cudaStream_t)stream, p));
  return 0;
}

// Backward of the user-encoder attention on the tensor cores (same contract as fr_ue_mha_bwd in user_encoder.cu, which
// stays as the cross-check). Autograd of attention.py:32-45,69-81.
FR_API int fr_ue_attention_bwd_tc(const void* qkv, const void* dctx, const float* stats, const int* his_ids, void* dqkv,
                                  int B, int L, int H, int dk, int use_mask, void* stream) {
  FR_REQUIRE(L >= 1 && L <= 64, "fr_ue_attention_bwd_tc: L=%d must be in [1,64]", L);
  FR_REQUIRE(dk >= 4 && dk <= 32 && dk % 4 == 0, "fr_ue_attention_bwd_tc: dk=%d must be a multiple of 4 in [4,32]", dk);
  FR_REQUIRE(!use_mask || his_ids != nullptr, "fr_ue_attention_bwd_tc: use_mask needs his_ids");
  if (B == 0) return 0;
  UeAttBwdParams p;
  p.qkv = (const __nv_bfloat16*)qkv; p.dctx = (const __nv_bfloat16*)dctx; p.stats = stats; p.his_ids = his_ids;
  p.dqkv = (__nv_bfloat16*)dqkv; p.B = B; p.L = L; p.H = H; p.dk = dk; p.use_mask = use_mask;
  p.scale = 1.f / sqrtf((float)dk);
  constexpr int smem = kUeBwdBar + 64 + 1024 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    FR_CHECK(cudaFuncSetAttribute(ue_attention_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  const long long jobs = (long long)((B + 1) / 2) * H;
  const int grid = jobs < (long long)num_sms() ? (int)jobs : num_sms();
  FR_CHECK(launch_pdl(ue_attention_bwd_tc_kernel, dim3(grid), dim3(128), smem, (cudaStream_t)stream, p));
  return 0;
}
