// N7: bf16 "TN" GEMM on the 5th-generation tensor cores.
//
//   out[M, N] = epilogue( A[M, K] · W[N, K]^T + bias[N] )        A, W bf16 row-major (K-major)
//
// This is every Linear of the model (reference: HF DistilBERT q/k/v/out/ffn Linears reached from
// encoder.py:27; text head encoder.py:20-22; user-encoder projections attention.py:72-74, :16).
// nn.Linear stores W as [N, K], i.e. already K-major, so both operands are loaded by TMA with the
// 128-byte swizzle straight into the layout tcgen05.mma wants; no transposes anywhere.
//
// Structure: persistent, warp-specialised, one CTA per SM.
//   warp 0      TMA producer   (one elected lane)  global → smem ring, kStages deep
//   warp 1      MMA issuer     (one elected lane)  tcgen05.mma 128 x BN x 16, fp32 accumulators in TMEM
//   warps 2..17 epilogue       tcgen05.ld → bias / GELU / tanh / residual → bf16|fp32 → global.
//                              bf16 outputs with BN >= 128 go through a per-warp swizzled staging box in shared
//                              memory and leave with a TMA store (and the residual arrives by a TMA load into the
//                              same box), so the epilogue issues no per-row global loads or stores at all.
// TMEM holds two accumulator stages (2 x BN columns) so the epilogue of tile i overlaps the MMAs of
// tile i+1. M may be supplied from device memory (m_dev) so a CUDA-graph-captured step can run on the
// de-duplicated row count of the current batch without a host round trip.
#include "sm100.cuh"
#include <cuda.h>
#include <mutex>
#include <stdlib.h>
#include <string.h>

namespace fr {
using namespace sm100;

constexpr int BM = 128;
constexpr int BK = 64;        // 64 bf16 = 128 bytes = one swizzle row
constexpr int UMMA_K = 16;
constexpr int kEpiWarps = 16;     // four per TMEM lane quarter: the epilogue is latency-bound, so more warps = more overlap
constexpr int kThreads = 32 * (2 + kEpiWarps);

enum Epilogue { EPI_BIAS = 0, EPI_BIAS_GELU = 1, EPI_BIAS_RESID = 2, EPI_BIAS_TANH = 3 };

struct GemmParams {
  const float* bias;          // [N] or null
  const __nv_bfloat16* resid; // [M, ldr] or null
  void* out;                  // bf16 or fp32 [M, ldo]
  const int* m_dev;           // optional: effective M = min(M, *m_dev * m_scale)
  int m_scale;
  int M, N, K;
  int ldo, ldr;
  // LayerNorm folding (FOLD kernels only; every pointer optional). Row statistics are (sum, sum of squares) pairs, one
  // per 64-column slice of the row they describe.
  const float* a_stats;       // [M, K/64, 2] of the rows of A: out = rstd*(A·W'^T) - rstd*mu*col_c + bias  (W' = W∘gamma)
  const float* col_c;         // [N]   c_n = sum_k W'_nk
  const float* r_stats;       // [M, N/64, 2] of the residual rows: the residual is LayerNorm(resid) computed on the fly
  const float* r_gamma;       // [N]
  const float* r_beta;        // [N]
  float* stats_out;           // [M, N/64, 2] = (sum, M2 about the slice mean) of every 64-column slice of the output rows
  float ln_eps;
  int a_dim, r_dim;           // widths the two statistics were taken over
  int tail_split;             // CTA-pair 256-wide kernels: cut the tiles of a short last round in two (see the kernel)
};

// CG = 1: one CTA per 128 x BN tile. CG = 2: a CTA pair (cta_group::2) per 256 x BN tile; each CTA stages its own
// 128 rows of A and half of the B tile, so the L2 → shared-memory traffic per FLOP drops by a third.
template <int BN, bool TMA_EPI, int CG, bool FOLD = false>
struct Cfg {
  static constexpr int kStages = CG == 2 ? (BN == 192 ? (TMA_EPI ? 5 : 7) : (TMA_EPI ? 4 : 6))
                                         : (TMA_EPI ? (BN >= 256 ? 3 : 4) : (BN >= 256 ? 4 : (BN >= 128 ? 6 : 8)));
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = (BN / CG) * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = BN == 192 ? 512 : (2 * BN < 32 ? 32 : 2 * BN);   // two accumulators, a power of two
  static constexpr int kColsPerWarp = BN == 192 ? 64 : (BN / 4 < 32 ? 32 : BN / 4);  // columns one epilogue warp owns
  static constexpr int kColGroups = BN / kColsPerWarp;                 // 4 (BN = 128, 256), 3 (BN = 192) or 2 (BN = 64)
  static constexpr int kVecs = FOLD ? 4 : 1;                           // per-column vectors staged per warp: bias (+ c, gamma, beta)
  static constexpr int kBiasBytes = kVecs * kEpiWarps * kColsPerWarp * 4;
  static constexpr int kBoxBytes = 4096;                               // per-warp output box: 32 rows x (64|32) bf16
  static constexpr int kStagingBytes = TMA_EPI ? kEpiWarps * kBoxBytes : 0;
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024 /*align slack*/ + 512 /*barriers*/ + kBiasBytes;
};

// GELU in the epilogue. The epilogue shares the SM's issue slots with nothing else but is on the critical path for the
// K = 768 GEMMs (12 k-blocks per tile), so the form is chosen for instruction count:
//   default        0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))   5 FP32 ops + 1 MUFU (tanh.approx) per element.
//                  Distance to the exact erf form: < 5e-4 absolute everywhere (the approximation's 3e-4 plus
//                  tanh.approx's 2^-11), a tenth of the bf16 rounding of the stored result. Measured on the FFN1 shape
//                  (26880 x 3072 x 768): 101 us against 120 us for the erf form and 103 us for a bias-only epilogue.
//   -DFEDREC_GELU_ERF  0.5 x (1 + erf(x / sqrt 2)) with erf from Abramowitz–Stegun 7.1.25 (|err| < 2.5e-5):
//                  9 FP32 ops + 2 MUFU (rcp.approx, ex2.approx) per element; HF DistilBERT's "gelu" to 2.5e-5.
__device__ __forceinline__ float gelu_erf(float x) {
#ifdef FEDREC_GELU_ERF
  float t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.33267435f /* 0.47047 / sqrt 2 */, fabsf(x), 1.0f)));
  const float u = x * 0.84932180f;                                  // sqrt(log2(e) / 2): exp(-x^2/2) = 2^-(u^2)
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-u * u));
  float poly = fmaf(0.7478556f, t, -0.0958798f);
  poly = fmaf(poly, t, 0.3480242f);
  poly *= t;
  const float erf_abs = fmaf(-poly, e, 1.0f);
  const float hx = 0.5f * x;
  return fmaf(hx, copysignf(erf_abs, x), hx);
#else
  const float inner = x * fmaf(0.0356774081f, x * x, 0.7978845608f);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(inner));
  const float hx = 0.5f * x;
  return fmaf(hx, t, hx);
#endif
}
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// byte offset of 16-byte unit j of row r inside a swizzled staging box (rows of kCols bf16)
template <int kCols>
__device__ __forceinline__ uint32_t box_off(int r, int j) {
  if (kCols == 64) return (uint32_t)(r * 128 + ((j ^ (r & 7)) << 4));          // SWIZZLE_128B
  return (uint32_t)(r * 64 + ((j ^ ((r >> 1) & 3)) << 4));                       // SWIZZLE_64B
}

// Mean and (biased) variance of a row from its per-slice statistics. Every 64-column slice carries (sum, M2) with
// M2 = sum((v - slice mean)^2); the slices are merged one after the other with the parallel-variance update
//   d = mean_s - mean;  mean += d / (k + 1);  M2 += M2_s + d^2 * 64 k / (k + 1)        (k slices merged so far),
// one pass, fixed order (bit-reproducible). The term between slices is formed from differences of means, so a large
// common offset or a few outlier channels (a pretrained DistilBERT's residual stream) cost no precision across slices;
// inside a slice the producer's M2 = sum v^2 - (sum v)^2 / 64 loses about mean^2 / var * 1e-7 of relative accuracy, far
// below bf16 for any stream LayerNorm is usable on (mean 50, std 1: 3e-4).
__device__ __forceinline__ void merge_row_stats(const float2* sp, int ns, float& mean, float& var) {
  if (ns == 12) {
    // 768-wide rows (every LayerNorm of this trunk): all twelve loads are issued before the first is consumed. With a
    // run-time trip count the compiler keeps one load in flight per iteration, a chain of twelve L2 round trips per row
    // (the hottest stall of the short-K GEMMs in profiles/ncu_gemm_fold_out.md before this).
    float2 t[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) t[k] = __ldg(sp + k);
    mean = t[0].x * (1.f / 64.f);
    float m2 = t[0].y;
#pragma unroll
    for (int k = 1; k < 12; ++k) {
      const float d = fmaf(t[k].x, 1.f / 64.f, -mean);
      const float w = 1.f / (float)(k + 1);
      mean = fmaf(d, w, mean);
      m2 += fmaf(d * d, 64.f * (float)k * w, t[k].y);
    }
    var = fmaxf(m2 * (1.f / 768.f), 0.f);
    return;
  }
  float2 t = __ldg(sp);
  mean = t.x * (1.f / 64.f);
  float m2 = t.y;
  for (int k = 1; k < ns; ++k) {
    t = __ldg(sp + k);
    const float d = fmaf(t.x, 1.f / 64.f, -mean);
    const float w = __fdividef(1.f, (float)(k + 1));
    mean = fmaf(d, w, mean);
    m2 += fmaf(d * d, 64.f * (float)k * w, t.y);
  }
  var = fmaxf(__fdividef(m2, 64.f * (float)ns), 0.f);
}

template <int BN, int EPI, bool OUT_F32, bool TMA_EPI, int CG, bool FOLD = false>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_o, const __grid_constant__ CUtensorMap tmap_r,
                    const __grid_constant__ CUtensorMap tmap_bh, const GemmParams p) {
  using C = Cfg<BN, TMA_EPI, CG, FOLD>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* staging = smem + C::kStages * C::kStageBytes;                     // 1024-aligned (stage sizes are)
  uint8_t* ctrl = staging + C::kStagingBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ctrl);
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tmem_full = empty_bar + C::kStages;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;           // [2]
  uint64_t* resid_bar = tmem_empty + 2;           // [kEpiWarps]
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(resid_bar + kEpiWarps);
  float* bias_stage = reinterpret_cast<float*>(ctrl + 512);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  int M = p.M;
  if (p.m_dev) {
    const long long md = (long long)(*p.m_dev) * p.m_scale;
    if (md < M) M = (int)md;
  }
  const uint32_t cta_rank = CG == 2 ? cluster_ctarank() : 0u;     // 0 = leader of the pair
  const int m_blocks = (M + BM * CG - 1) / (BM * CG);
  const int n_blocks = (p.N + BN - 1) / BN;
  const int k_blocks = (p.K + BK - 1) / BK;
  const int num_tiles = m_blocks * n_blocks;
  const int tile_first = (int)blockIdx.x / CG, tile_step = (int)gridDim.x / CG;
  // Tail split (CTA pairs, 256-wide tiles). The tiles are dealt round-robin to the pairs, so when their number is not a
  // multiple of the pair count the last round keeps a few pairs busy for a whole tile while the rest idle (N = 768:
  // 315 tiles on 74 pairs = 4 full rounds + 19 tiles). If that last round fills at most half of the pairs, each of its
  // tiles is cut into two 256 x 128 halves (cta_group::2 MMAs with N = 128, B boxes of 64 rows per CTA through a second
  // tensor map), which halves the length of the last round. Work items >= split_from are such halves.
  constexpr bool kTailSplit = (CG == 2 && BN == 256);
  int split_from = num_tiles, num_items = num_tiles;
  if (kTailSplit && p.tail_split) {
    const int full = (num_tiles / tile_step) * tile_step, tail = num_tiles - full;
    if (tail > 0 && 2 * tail <= tile_step) { split_from = full; num_items = full + 2 * tail; }
  }
  // item -> (tile, half): half = -1 for a whole tile, 0 / 1 for the left / right 128 columns of a split tile
  auto decode = [&](int item, int& tile, int& half) {
    if (item < split_from) { tile = item; half = -1; }
    else { tile = split_from + ((item - split_from) >> 1); half = (item - split_from) & 1; }
  };

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
    if (TMA_EPI) {
      prefetch_tmap(&tmap_o);
      if (EPI == EPI_BIAS_RESID) prefetch_tmap(&tmap_r);
    }
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < C::kStages; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tmem_full[i], 1);
        mbar_init(&tmem_empty[i], kEpiWarps * CG);
      }
      for (int i = 0; i < kEpiWarps; ++i) mbar_init(&resid_bar[i], 1);
      fence_barrier_init();
    }
    __syncwarp();
    if (CG == 2) tmem_alloc_pair(tmem_base_slot, C::kTmemCols);
    else tmem_alloc(tmem_base_slot, C::kTmemCols);
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all();       // the peer's barriers must exist before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;
  // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) may overlap the tail of the
  // previous kernel in the stream; from here on we read its output. Our own dependents may begin their prologue now.
  griddep_launch();
  griddep_wait();

  if (warp == 0) {
    // ------------------------------- TMA producer -------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = tile_first; item < num_items; item += tile_step) {
        int tile, half;
        decode(item, tile, half);
        const int m0 = (tile / n_blocks) * (BM * CG) + (int)cta_rank * BM;
        const int n0 = half < 0 ? (tile % n_blocks) * BN + (int)cta_rank * (BN / CG) * (CG - 1)
                                : (tile % n_blocks) * BN + half * (BN / 2) + (int)cta_rank * (BN / 4);
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1, 11);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + C::kABytes;
          if (CG == 2) {
            // both CTAs' bytes are credited to the leader's barrier; only the leader arrives on it
            const uint32_t leader_bar = mapa_u32(smem_u32(&full_bar[stage]), 0);
            if (cta_rank == 0)
              mbar_arrive_expect_tx(&full_bar[stage], half < 0 ? 2 * C::kStageBytes : 2 * (C::kABytes + C::kBBytes / 2));
            tma_load_2d_pair(sa, &tmap_a, leader_bar, kb * BK, m0);
            tma_load_2d_pair(sb, half < 0 ? &tmap_b : &tmap_bh, leader_bar, kb * BK, n0);
          } else {
            mbar_arrive_expect_tx(&full_bar[stage], C::kStageBytes);
            tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * BK, m0);
            tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * BK, n0);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------- MMA issuer ---------------------------------
    if (lane == 0 && cta_rank == 0) {
      constexpr uint32_t idesc_full = make_idesc_bf16(BM * CG, BN);
      constexpr uint32_t idesc_half = make_idesc_bf16(BM * CG, BN / 2);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int item = tile_first; item < num_items; item += tile_step) {
        const uint32_t idesc = (kTailSplit && item >= split_from) ? idesc_half : idesc_full;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1, 12);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase, 13);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint64_t da = make_smem_desc_sw128(sa);
          const uint64_t db = make_smem_desc_sw128(sa + C::kABytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advance 32 bytes (16 bf16) along K inside the swizzle row: +2 in 16-byte units
            if (CG == 2) umma_bf16_ss_pair(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
            else umma_bf16_ss(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
          }
          if (CG == 2) umma_commit_pair(&empty_bar[stage], 3);   // frees the slot in both CTAs
          else umma_commit(&empty_bar[stage]);                    // smem slot free once these MMAs retire
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        if (CG == 2) umma_commit_pair(&tmem_full[acc], 3);       // accumulator complete → both epilogues
        else umma_commit(&tmem_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ------------------------------- epilogue -----------------------------------
    const int ew = warp - 2;                 // 0..15
    const int quarter = warp & 3;            // TMEM lane quarter this warp may touch
    const int cgroup = ew >> 2;              // which slice of the BN columns
    constexpr int kCols = C::kColsPerWarp;
    const bool active = cgroup < C::kColGroups;
    float* my_bias = bias_stage + ew * kCols;
    float* my_c = bias_stage + (kEpiWarps + ew) * kCols;          // FOLD only: the three extra per-column vectors
    float* my_g = bias_stage + (2 * kEpiWarps + ew) * kCols;
    float* my_be = bias_stage + (3 * kEpiWarps + ew) * kCols;
    uint8_t* my_box = staging + (TMA_EPI ? ew * C::kBoxBytes : 0);
    const uint32_t box_u32 = smem_u32(my_box);
    uint32_t rphase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    // the accumulator stage is handed back to the leader's MMA warp: locally, or across the pair
    const uint32_t empty_remote[2] = {CG == 2 ? mapa_u32(smem_u32(&tmem_empty[0]), 0) : 0u,
                                      CG == 2 ? mapa_u32(smem_u32(&tmem_empty[1]), 0) : 0u};
    auto release_acc = [&](int a) {
      if (CG == 2 && cta_rank != 0) mbar_arrive_cluster(empty_remote[a]);
      else mbar_arrive(&tmem_empty[a]);
    };
    for (int item = tile_first; item < num_items; item += tile_step) {
      int tile, half;
      decode(item, tile, half);
      const int m0 = (tile / n_blocks) * (BM * CG) + (int)cta_rank * BM;
      const int n0 = (tile % n_blocks) * BN + (half > 0 ? BN / 2 : 0);
      const int brow0 = m0 + quarter * 32;                   // first row of this warp's 32-row box
      const int row = brow0 + lane;
      const bool row_ok = row < M;
      const int wcol0 = n0 + cgroup * kCols;                 // first column of this warp
      const bool work = active && wcol0 < p.N && (half < 0 || cgroup * kCols < BN / 2);   // warp-uniform
      const bool box_full = brow0 + 32 <= M;                 // TMA store may cover the whole box
      // While the MMAs of this tile are still running: stage the bias slice, make sure the previous TMA store has
      // drained this warp's box, and start the residual load — nothing after "accumulator ready" waits on memory.
      if (work) {
        __syncwarp();
#pragma unroll
        for (int j = lane; j < kCols; j += 32) my_bias[j] = (p.bias && wcol0 + j < p.N) ? __ldg(p.bias + wcol0 + j) : 0.f;
        if constexpr (FOLD) {
#pragma unroll
          for (int j = lane; j < kCols; j += 32) {
            const bool ok = wcol0 + j < p.N;
            my_c[j] = (p.col_c && ok) ? __ldg(p.col_c + wcol0 + j) : 0.f;
            my_g[j] = (p.r_gamma && ok) ? __ldg(p.r_gamma + wcol0 + j) : 1.f;
            my_be[j] = (p.r_beta && ok) ? __ldg(p.r_beta + wcol0 + j) : 0.f;
          }
        }
        if (TMA_EPI && lane == 0) {
          tma_store_wait_read<0>();
          if (EPI == EPI_BIAS_RESID) {
            mbar_arrive_expect_tx(&resid_bar[ew], 32 * kCols * 2);
            tma_load_2d(my_box, &tmap_r, &resid_bar[ew], wcol0, brow0);
          }
        }
        __syncwarp();
      }
      uint4 rres[kCols / 8];
      if (!TMA_EPI && EPI == EPI_BIAS_RESID && work) {
        const uint4* rp = reinterpret_cast<const uint4*>(p.resid + (size_t)row * p.ldr + wcol0);
#pragma unroll
        for (int q = 0; q < kCols / 8; ++q)
          rres[q] = (row_ok && wcol0 + q * 8 < p.N) ? __ldg(rp + q) : make_uint4(0, 0, 0, 0);
      }
      // FOLD: per-row LayerNorm scalars (this thread owns row `row`), fetched while the MMAs run
      float ra = 1.f, rb = 0.f;            // out = ra * acc + rb * c[col] + bias[col]
      float rr = 1.f, rm = 0.f;            // residual = (resid - rm) * rr * gamma[col] + beta[col]
      float st_s = 0.f, st_q = 0.f;        // running (sum, sum of squares) of this thread's 64-column slice of its output row
      if constexpr (FOLD) {
        if (work && row_ok) {
          // A row's statistics arrive as one (sum, M2) pair per 64-column slice of the producer's output (written without
          // atomics, see below) and are merged here in a fixed order: bit-reproducible.
          if (p.a_stats) {
            float mu, var;
            merge_row_stats(reinterpret_cast<const float2*>(p.a_stats) + (size_t)row * (p.a_dim >> 6), p.a_dim >> 6, mu, var);
            ra = rsqrtf(var + p.ln_eps);
            rb = -ra * mu;
          }
          if (p.r_stats) {
            float var;
            merge_row_stats(reinterpret_cast<const float2*>(p.r_stats) + (size_t)row * (p.r_dim >> 6), p.r_dim >> 6, rm, var);
            rr = rsqrtf(var + p.ln_eps);
          }
        }
      }
      mbar_wait(&tmem_full[acc], acc_phase, 14);
      tc_fence_after();
      if (work) {
        if (TMA_EPI && EPI == EPI_BIAS_RESID) {
          mbar_wait(&resid_bar[ew], rphase, 15);
          rphase ^= 1;
        }
        uint32_t packed[kCols / 2];                          // bf16x2 results of the whole slice (TMA path)
#pragma unroll
        for (int j = 0; j < kCols / 2; ++j) packed[j] = 0u;
#pragma unroll
        for (int cc = 0; cc < kCols; cc += 32) {
          const int col0 = wcol0 + cc;
          if (col0 >= p.N) break;              // warp-uniform
          uint32_t r[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + cgroup * kCols + cc);
          tmem_ld_32x32(taddr, r);
          if (TMA_EPI && EPI == EPI_BIAS_RESID) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint4 t4;
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(t4.x), "=r"(t4.y), "=r"(t4.z), "=r"(t4.w)
                           : "r"(box_u32 + box_off<kCols>(lane, cc / 8 + q)));
              rres[cc / 8 + q] = t4;
            }
          }
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            if (FOLD) v[j] = fmaf(ra, __uint_as_float(r[j]), fmaf(rb, my_c[cc + j], my_bias[cc + j]));
            else v[j] = __uint_as_float(r[j]) + my_bias[cc + j];
          }
          const int ncols = min(32, p.N - col0);       // multiple of 8 (N % 8 == 0)
          if (EPI == EPI_BIAS_RESID) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&rres[cc / 8 + q]);
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                const float2 f = __bfloat1622float2(h[t]);
                if (FOLD) {
                  const int c = cc + q * 8 + 2 * t;
                  v[q * 8 + 2 * t] += fmaf((f.x - rm) * rr, my_g[c], my_be[c]);
                  v[q * 8 + 2 * t + 1] += fmaf((f.y - rm) * rr, my_g[c + 1], my_be[c + 1]);
                } else {
                  v[q * 8 + 2 * t] += f.x;
                  v[q * 8 + 2 * t + 1] += f.y;
                }
              }
            }
          }
          if constexpr (FOLD) {
            if (p.stats_out) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                if (j < ncols) {
                  st_s += v[j];
                  st_q = fmaf(v[j], v[j], st_q);
                }
              }
            }
          }
          if (EPI == EPI_BIAS_GELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
          }
          if (EPI == EPI_BIAS_TANH) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = tanh_fast(v[j]);
          }
          if (TMA_EPI) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const __nv_bfloat162 b2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
              packed[cc / 2 + j] = *reinterpret_cast<const uint32_t*>(&b2);
            }
          } else if (row_ok) {
            if (OUT_F32) {
              float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + (size_t)row * p.ldo + col0);
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                if (j < ncols) op[j / 4] = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
              uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)row * p.ldo + col0);
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                if (j < ncols) {
                  uint4 q;
                  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
                  for (int t = 0; t < 4; ++t) h[t] = __floats2bfloat162_rn(v[j + 2 * t], v[j + 2 * t + 1]);
                  op[j / 8] = q;
                }
              }
            }
          }
        }
        // the accumulator has been read out: hand the TMEM stage back to the MMA warp before storing
        tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(acc);
        if constexpr (FOLD) {
          static_assert(!FOLD || kCols == 64, "row statistics are kept per 64-column slice");
          if (p.stats_out && row_ok)       // this warp's 64 columns are slice wcol0 / 64 of the row: a plain store
            reinterpret_cast<float2*>(p.stats_out)[(size_t)row * (p.N >> 6) + (wcol0 >> 6)] =
                make_float2(st_s, fmaxf(fmaf(-st_s * (1.f / 64.f), st_s, st_q), 0.f));   // (sum, M2 about the slice mean)
        }
        if (TMA_EPI) {
          if (box_full) {
#pragma unroll
            for (int q = 0; q < kCols / 8; ++q) {
              {
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(box_u32 + box_off<kCols>(lane, q)),
                             "r"(packed[4 * q]), "r"(packed[4 * q + 1]), "r"(packed[4 * q + 2]), "r"(packed[4 * q + 3])
                             : "memory");
              }
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmap_o, my_box, wcol0, brow0);     // columns past N are clipped by the tensor map
              tma_store_commit();
            }
          } else if (row_ok) {                                   // last, partial row block under a device-side M
            uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)row * p.ldo + wcol0);
#pragma unroll
            for (int q = 0; q < kCols / 8; ++q)
              if (wcol0 + q * 8 < p.N) op[q] = make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
          }
        }
      } else {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(acc);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (TMA_EPI && lane == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  if (CG == 2) cluster_sync_all();       // neither CTA may leave while the pair still uses its memory
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (CG == 2) tmem_dealloc_pair(tmem_base, C::kTmemCols);
    else tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------
// Host side: tensor-map construction (driver entry point fetched through the runtime, so the
// library does not link libcuda) and launch.
// ------------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)f;
  });
  return fn;
}

// [rows, cols] bf16 row-major with leading dimension ld (elements); box = box_rows x box_cols, 128B swizzle.
int make_tmap_bf16_2d_box(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows,
                          int box_cols);
int make_tmap_bf16_2d(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows) {
  return make_tmap_bf16_2d_box(out, base, rows, cols, ld, box_rows, BK);
}
int make_tmap_bf16_2d_box(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows,
                          int box_cols) {
  PFN_encodeTiled enc = get_encode();
  FR_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  FR_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * 2) % 16 == 0, "TMA operand must be 16-byte aligned with ld %% 8 == 0 (ld=%lld)", ld);
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  FR_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld ld=%lld box_rows=%d)",
             (int)r, rows, cols, ld, box_rows);
  return 0;
}

struct Maps { CUtensorMap a, b, o, r, bh; };

template <int BN, int EPI, bool OUT_F32, bool TMA_EPI, int CG, bool FOLD = false>
static int launch(const Maps& m, const GemmParams& p, int max_ctas, cudaStream_t s) {
  auto kern = gemm_bf16_tn_kernel<BN, EPI, OUT_F32, TMA_EPI, CG, FOLD>;
  using C = Cfg<BN, TMA_EPI, CG, FOLD>;
  static bool attr_set = false;   // per instantiation
  if (!attr_set) {
    FR_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    attr_set = true;
  }
  const int tiles = ((p.M + BM * CG - 1) / (BM * CG)) * ((p.N + BN - 1) / BN);
  int grid = tiles * CG < max_ctas ? tiles * CG : max_ctas / CG * CG;
  if (grid < CG) grid = CG;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  int na = CG == 2 ? 1 : 0;
  static const bool use_pdl = [] { const char* e = getenv("FEDREC_PDL"); return !(e && atoi(e) == 0); }();
  if (use_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  FR_CHECK(cudaLaunchKernelEx(&cfg, kern, m.a, m.b, m.o, m.r, m.bh, p));
  return 0;
}

template <int BN, bool OUT_F32, bool TMA_EPI, int CG = 1>
static int dispatch_epi(int epi, const Maps& m, const GemmParams& p, int max_ctas, cudaStream_t s) {
  switch (epi) {
    case EPI_BIAS: return launch<BN, EPI_BIAS, OUT_F32, TMA_EPI, CG>(m, p, max_ctas, s);
    case EPI_BIAS_GELU: return launch<BN, EPI_BIAS_GELU, OUT_F32, TMA_EPI, CG>(m, p, max_ctas, s);
    case EPI_BIAS_RESID: return launch<BN, EPI_BIAS_RESID, OUT_F32, TMA_EPI, CG>(m, p, max_ctas, s);
    case EPI_BIAS_TANH: return launch<BN, EPI_BIAS_TANH, OUT_F32, TMA_EPI, CG>(m, p, max_ctas, s);
  }
  fr::set_error("fr_gemm_bf16: unknown epilogue %d", epi);
  return -1;
}

// LayerNorm-folding variants: the 256-wide tile, bf16 output, three of the epilogues.
template <bool TMA_EPI, int CG>
static int dispatch_fold(int epi, const Maps& m, const GemmParams& p, int max_ctas, cudaStream_t s) {
  switch (epi) {
    case EPI_BIAS: return launch<256, EPI_BIAS, false, TMA_EPI, CG, true>(m, p, max_ctas, s);
    case EPI_BIAS_GELU: return launch<256, EPI_BIAS_GELU, false, TMA_EPI, CG, true>(m, p, max_ctas, s);
    case EPI_BIAS_RESID: return launch<256, EPI_BIAS_RESID, false, TMA_EPI, CG, true>(m, p, max_ctas, s);
  }
  fr::set_error("fr_gemm_bf16_fold: epilogue %d has no LayerNorm-folding variant (bias, gelu, resid only)", epi);
  return -1;
}

// generic 2D bf16 map with an explicit swizzle mode (epilogue boxes)
static int make_tmap_box_swz(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows,
                             int box_cols, CUtensorMapSwizzle swz) {
  PFN_encodeTiled enc = get_encode();
  FR_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  FR_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (epilogue box) failed with CUresult %d", (int)r);
  return 0;
}

static int sm_count() { return num_sms(); }

}  // namespace fr

using namespace fr;

// out = epi(A · W^T + bias).  A [M,K] ld=lda, W [N,K] ld=ldw (both bf16, K-major), out bf16 or fp32 ld=ldo.
// block_n: 0 = choose, else 64/128/192/256 (192: CTA pairs only).  m_dev (optional, device int): effective M = min(M, *m_dev*m_scale).
struct FoldArgs {
  const float* a_stats; const float* col_c; const float* r_stats; const float* r_gamma; const float* r_beta;
  float* stats_out; float ln_eps;
};

static int gemm_bf16_impl(const void* A, const void* W, const float* bias, const void* resid, void* out, int out_fp32,
                          int M, int N, int K, int lda, int ldw, int ldo, int ldr, int epi, int block_n,
                          const int* m_dev, int m_scale, int cta_group, int tma_epilogue, void* stream,
                          const FoldArgs* fold = nullptr) {
  FR_REQUIRE(M >= 1 && N >= 8 && K >= 8, "fr_gemm_bf16: bad shape M=%d N=%d K=%d", M, N, K);
  FR_REQUIRE(N % 8 == 0 && K % 8 == 0, "fr_gemm_bf16: N=%d and K=%d must be multiples of 8", N, K);
  FR_REQUIRE(ldo % 8 == 0 && ((uintptr_t)out & 15) == 0, "fr_gemm_bf16: out must be 16-byte aligned, ldo %% 8 == 0");
  FR_REQUIRE(epi != EPI_BIAS_RESID || (resid && ldr % 8 == 0 && ((uintptr_t)resid & 15) == 0),
             "fr_gemm_bf16: residual epilogue needs an aligned resid pointer");
  FR_REQUIRE(!bias || ((uintptr_t)bias & 15) == 0, "fr_gemm_bf16: bias must be 16-byte aligned");
  if (block_n == 0) {
    // fewest wasted columns first, then the widest tile (fewer A re-reads)
    int best = 256, best_waste = ((N + 255) / 256) * 256 - N;
    for (int bn : {128, 64}) {
      const int waste = ((N + bn - 1) / bn) * bn - N;
      if (waste < best_waste) { best = bn; best_waste = waste; }
    }
    block_n = best;
    if (fold) block_n = 256;
    // N = 768 (attention out-projection, FFN2): 4 x 192 columns quantise into waves better than 3 x 256
    static const int env_bn768 = [] { const char* e = getenv("FEDREC_GEMM_BN768"); return e ? atoi(e) : 256; }();
    if (N == 768 && !out_fp32 && M >= 4096 && env_bn768 == 192) block_n = 192;
  }
  FR_REQUIRE(block_n == 64 || block_n == 128 || block_n == 192 || block_n == 256, "fr_gemm_bf16: block_n=%d", block_n);
  // cta_group: 0 = choose. Pairs need the wide tile, a bf16 output and enough rows to fill the GPU with pairs.
  static const int env_cg = [] { const char* e = getenv("FEDREC_GEMM_CTA_GROUP"); return e ? atoi(e) : 0; }();
  if (cta_group == 0) cta_group = env_cg;
  // 192-wide tiles exist for CTA pairs only (N = 768 = 4 x 192: six waves of narrow tiles instead of five of wide ones)
  if (block_n == 192 && (out_fp32 || cta_group == 1)) block_n = 256;
  const bool pair_ok = !out_fp32 && (block_n == 256 || block_n == 192);
  if (cta_group == 0) cta_group = (pair_ok && (M >= 4096 || block_n == 192)) ? 2 : 1;
  if (cta_group == 2 && !pair_ok) cta_group = 1;
  FR_REQUIRE(cta_group == 1 || cta_group == 2, "fr_gemm_bf16: cta_group=%d", cta_group);
  Maps maps;
  if (int e = make_tmap_bf16_2d(&maps.a, A, M, K, lda, BM)) return e;
  if (int e = make_tmap_bf16_2d(&maps.b, W, N, K, ldw, block_n / cta_group)) return e;
  maps.bh = maps.b;
  static const bool env_tail = [] { const char* e = getenv("FEDREC_GEMM_TAIL_SPLIT"); return !(e && atoi(e) == 0); }();
  const bool tail_split = env_tail && cta_group == 2 && block_n == 256;
  if (tail_split)                                   // B boxes of the half-width tiles of a short last round: 64 rows per CTA
    if (int e = make_tmap_bf16_2d(&maps.bh, W, N, K, ldw, block_n / 4)) return e;
  // bf16 outputs of the wide tiles leave through shared memory + TMA (FEDREC_GEMM_EPI=direct disables it).
  // For single CTAs the staging boxes cost a pipeline stage, which only pays off when the epilogue dominates (short K).
  static const int env_epi = [] { const char* e = getenv("FEDREC_GEMM_EPI"); return !e ? 0 : (strcmp(e, "direct") == 0 ? -1 : 1); }();
  if (tma_epilogue == 0) tma_epilogue = env_epi;
  bool tma_epi = !out_fp32 && block_n >= 192 && K <= 1536;     // long K: the deeper pipeline wins (measured)
  if (tma_epilogue < 0) tma_epi = false;
  if (tma_epilogue > 0) tma_epi = !out_fp32 && block_n >= 192;
  maps.o = maps.a; maps.r = maps.a;
  if (tma_epi) {
    const int kc = block_n == 192 ? 64 : block_n / 4;
    const CUtensorMapSwizzle swz = kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    if (int e = make_tmap_box_swz(&maps.o, out, M, N, ldo, 32, kc, swz)) return e;
    if (epi == EPI_BIAS_RESID)
      if (int e = make_tmap_box_swz(&maps.r, resid, M, N, ldr, 32, kc, swz)) return e;
  }
  GemmParams p;
  p.bias = bias; p.resid = (const __nv_bfloat16*)resid; p.out = out; p.m_dev = m_dev; p.m_scale = m_scale;
  p.M = M; p.N = N; p.K = K; p.ldo = ldo; p.ldr = ldr;
  p.a_stats = nullptr; p.col_c = nullptr; p.r_stats = nullptr; p.r_gamma = nullptr; p.r_beta = nullptr; p.stats_out = nullptr;
  p.ln_eps = 0.f; p.a_dim = K; p.r_dim = N; p.tail_split = tail_split ? 1 : 0;
  cudaStream_t s = (cudaStream_t)stream;
  const int ctas = sm_count();
  if (fold) {
    FR_REQUIRE(!out_fp32 && block_n == 256, "fr_gemm_bf16_fold: bf16 output and the 256-wide tile only (block_n=%d)", block_n);
    FR_REQUIRE(!fold->a_stats || fold->col_c, "fr_gemm_bf16_fold: a_stats needs col_c");
    FR_REQUIRE(!fold->r_stats || (fold->r_gamma && fold->r_beta && epi == EPI_BIAS_RESID),
               "fr_gemm_bf16_fold: r_stats needs r_gamma, r_beta and the residual epilogue");
    FR_REQUIRE(!fold->stats_out || epi == EPI_BIAS_RESID, "fr_gemm_bf16_fold: stats_out is built for the residual epilogue");
    FR_REQUIRE((!fold->a_stats || K % 64 == 0) && ((!fold->r_stats && !fold->stats_out) || N % 64 == 0),
               "fr_gemm_bf16_fold: row statistics are kept per 64 columns (K=%d, N=%d)", K, N);
    p.a_stats = fold->a_stats; p.col_c = fold->col_c; p.r_stats = fold->r_stats; p.r_gamma = fold->r_gamma;
    p.r_beta = fold->r_beta; p.stats_out = fold->stats_out; p.ln_eps = fold->ln_eps;
    if (cta_group == 2) return tma_epi ? dispatch_fold<true, 2>(epi, maps, p, ctas, s) : dispatch_fold<false, 2>(epi, maps, p, ctas, s);
    return tma_epi ? dispatch_fold<true, 1>(epi, maps, p, ctas, s) : dispatch_fold<false, 1>(epi, maps, p, ctas, s);
  }
  if (out_fp32) {
    switch (block_n) {
      case 64: return dispatch_epi<64, true, false>(epi, maps, p, ctas, s);
      case 128: return dispatch_epi<128, true, false>(epi, maps, p, ctas, s);
      default: return dispatch_epi<256, true, false>(epi, maps, p, ctas, s);
    }
  }
  if (cta_group == 2) {
    if (block_n == 192) {
      if (tma_epi) return dispatch_epi<192, false, true, 2>(epi, maps, p, ctas, s);
      return dispatch_epi<192, false, false, 2>(epi, maps, p, ctas, s);
    }
    if (tma_epi) return dispatch_epi<256, false, true, 2>(epi, maps, p, ctas, s);
    return dispatch_epi<256, false, false, 2>(epi, maps, p, ctas, s);
  }
  if (tma_epi) return dispatch_epi<256, false, true>(epi, maps, p, ctas, s);
  switch (block_n) {
    case 64: return dispatch_epi<64, false, false>(epi, maps, p, ctas, s);
    case 128: return dispatch_epi<128, false, false>(epi, maps, p, ctas, s);
    default: return dispatch_epi<256, false, false>(epi, maps, p, ctas, s);
  }
}

FR_API int fr_gemm_bf16(const void* A, const void* W, const float* bias, const void* resid, void* out, int out_fp32,
                        int M, int N, int K, int lda, int ldw, int ldo, int ldr, int epi, int block_n,
                        const int* m_dev, int m_scale, void* stream) {
  return gemm_bf16_impl(A, W, bias, resid, out, out_fp32, M, N, K, lda, ldw, ldo, ldr, epi, block_n, m_dev, m_scale, 0, 0,
                        stream);
}

// Same with the two scheduling choices made explicit: cta_group 0 = choose, 1 = single CTAs, 2 = CTA pairs
// (cta_group::2, 256-row MMAs); tma_epilogue 0 = choose, -1 = per-row global stores, 1 = staged TMA stores.
FR_API int fr_gemm_bf16_ex(const void* A, const void* W, const float* bias, const void* resid, void* out, int out_fp32,
                           int M, int N, int K, int lda, int ldw, int ldo, int ldr, int epi, int block_n,
                           const int* m_dev, int m_scale, int cta_group, int tma_epilogue, void* stream) {
  return gemm_bf16_impl(A, W, bias, resid, out, out_fp32, M, N, K, lda, ldw, ldo, ldr, epi, block_n, m_dev, m_scale,
                        cta_group, tma_epilogue, stream);
}

// fr_gemm_bf16 with LayerNorm folded around it (frozen-weight trunks; all extra pointers optional, 256-wide tile):
//   Row statistics are (sum, M2 = sum of squared deviations from the slice mean) per 64-column slice: a_stats [M, K/64, 2], r_stats / stats_out [M, N/64, 2];
//   the producer stores its slices (no atomics, nothing to zero) and the consumer adds them in a fixed order.
//   a_stats + col_c [N]:
//       out = rstd_A * (A · W^T) - rstd_A * mu_A * col_c + bias      = LayerNorm_gamma,beta(A) · W0^T + b0
//       when W = W0∘gamma, col_c = rowsum(W), bias = b0 + W0·beta.
//   r_stats + r_gamma, r_beta [N] (residual epilogue): the residual added is LayerNorm(resid) computed on the fly.
//   stats_out (residual epilogue): statistics of every output row, for the next consumer.
FR_API int fr_gemm_bf16_fold(const void* A, const void* W, const float* bias, const void* resid, void* out, int M, int N,
                             int K, int lda, int ldw, int ldo, int ldr, int epi, const int* m_dev, int m_scale,
                             const float* a_stats, const float* col_c, const float* r_stats, const float* r_gamma,
                             const float* r_beta, float* stats_out, float ln_eps, void* stream) {
  FoldArgs f{a_stats, col_c, r_stats, r_gamma, r_beta, stats_out, ln_eps};
  return gemm_bf16_impl(A, W, bias, resid, out, 0, M, N, K, lda, ldw, ldo, ldr, epi, 0, m_dev, m_scale, 0, 0, stream, &f);
}
