// Weight-gradient GEMM on tcgen05 with MN-major operands and split-K:
//
//   out[M1, N1] (fp32) += A[K, M1]^T · B[K, N1]          A, B bf16 row-major, reduction over rows
//
// This is dW = dY^T · X for every trainable Linear (text head fc / att_fc1: model.py:86-87 backward through
// encoder.py:20-22; user encoder W_Q/W_K/W_V/att_fc1: autograd of attention.py:72-74,16). Both operands
// are exactly the activations/gradients as the forward/backward kernels left them ([rows, features]), so
// the reduction dimension is the *outer* one: operands are "MN-major" in tcgen05 terms. TMA loads
// 64-feature x 64-row boxes with the 128-byte swizzle; the shared-memory descriptors describe them with
// LBO = distance between 64-feature atoms and SBO = distance between 8-row groups, and the instruction
// descriptor sets the A/B major bits to MN. No transposed copies of activations are ever made.
//
// K (row count) can be taken from device memory so the de-dup count of the step drives the reduction
// length; rows in [K_eff, round_up(K_eff, 64)) of A must be zero (the producers guarantee it).
// Parallelism comes from split-K: grid = tiles x splits, partial tiles are accumulated into `out` with
// vector red.global.add (out is the flat gradient buffer, zeroed at the start of the step).
#include "sm100.cuh"
#include <cuda.h>

namespace fr {
using namespace sm100;

int make_tmap_bf16_2d_box(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows,
                          int box_cols);

constexpr int GBM = 128, GBK = 64, GStages = 4;
constexpr int kGradThreads = 32 * 6;

struct GradParams {
  float* out;
  const int* k_dev;
  int k_scale;
  int M1, N1, K, ldo, splits;
};

// MN-major, 128B-swizzled operand: atoms of 64 features x 8 rows (1024 B); `lbo` bytes between feature atoms.
__device__ __forceinline__ uint64_t make_smem_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t make_idesc_bf16_mn(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

template <int BN>
__global__ void __launch_bounds__(kGradThreads, 1)
gemm_grad_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const GradParams p) {
  constexpr int kABytes = GBK * GBM * 2;           // two 64-feature atoms of [64 rows x 128 B]
  constexpr int kBBytes = GBK * BN * 2;
  constexpr int kStageBytes = kABytes + kBBytes;
  constexpr int kAtomBytes = GBK * 128;            // one [64 rows x 64 features] block
  constexpr int kTmemCols = BN < 32 ? 32 : BN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + GStages * kStageBytes);
  uint64_t* empty_bar = full_bar + GStages;
  uint64_t* done_bar = empty_bar + GStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int K = p.K;
  if (p.k_dev) {
    const long long kd = (long long)(*p.k_dev) * p.k_scale;
    if (kd < K) K = (int)kd;
  }
  const int kb_total = (K + GBK - 1) / GBK;
  const int kb_per = (kb_total + p.splits - 1) / p.splits;
  const int kb0 = blockIdx.y * kb_per;
  const int kb1 = min(kb_total, kb0 + kb_per);
  if (kb0 >= kb1) return;                           // nothing to reduce in this split (uniform over the CTA)
  const int n_tiles = (p.N1 + BN - 1) / BN;
  const int m0 = (blockIdx.x / n_tiles) * GBM;
  const int n0 = (blockIdx.x % n_tiles) * BN;

  if (warp == 0 && lane == 0) { prefetch_tmap(&tmap_a); prefetch_tmap(&tmap_b); }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < GStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
      mbar_init(done_bar, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, kTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1, 21);
        uint8_t* sa = smem + stage * kStageBytes;
        uint8_t* sb = sa + kABytes;
        mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
#pragma unroll
        for (int j = 0; j < GBM / 64; ++j) tma_load_2d(sa + j * kAtomBytes, &tmap_a, &full_bar[stage], m0 + 64 * j, kb * GBK);
#pragma unroll
        for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * kAtomBytes, &tmap_b, &full_bar[stage], n0 + 64 * j, kb * GBK);
        if (++stage == GStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16_mn(GBM, BN);
      int stage = 0; uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase, 22);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * kStageBytes);
        const uint32_t sb = sa + kABytes;
#pragma unroll
        for (int k = 0; k < GBK / 16; ++k) {
          // 16 reduction rows = two 8-row groups = 2048 bytes further into every feature atom
          const uint64_t da = make_smem_desc_mn_sw128(sa + k * 2048, kAtomBytes);
          const uint64_t db = make_smem_desc_mn_sw128(sb + k * 2048, kAtomBytes);
          umma_bf16_ss(tmem_base, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);
        if (++stage == GStages) { stage = 0; phase ^= 1; }
      }
      umma_commit(done_bar);
    }
  } else {
    const int quarter = warp & 3;
    mbar_wait(done_bar, 0, 23);
    tc_fence_after();
    const int row = m0 + quarter * 32 + lane;
#pragma unroll 1
    for (int cc = 0; cc < BN; cc += 32) {
      const int col0 = n0 + cc;
      if (col0 >= p.N1) break;
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)cc, r);
      tmem_ld_wait();
      if (row < p.M1) {
        float* op = p.out + (size_t)row * p.ldo + col0;
        const int ncols = min(32, p.N1 - col0);
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          if (j < ncols)
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(op + j), "f"(__uint_as_float(r[j])),
                         "f"(__uint_as_float(r[j + 1])), "f"(__uint_as_float(r[j + 2])), "f"(__uint_as_float(r[j + 3]))
                         : "memory");
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <int BN>
static int launch_grad(const CUtensorMap& ta, const CUtensorMap& tb, const GradParams& p, cudaStream_t s) {
  auto kern = gemm_grad_kernel<BN>;
  constexpr int smem = GStages * (GBK * GBM * 2 + GBK * BN * 2) + 1024 + 256;
  static bool attr_set = false;
  if (!attr_set) {
    FR_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  const int tiles = ((p.M1 + GBM - 1) / GBM) * ((p.N1 + BN - 1) / BN);
  dim3 grid(tiles, p.splits);
  kern<<<grid, kGradThreads, smem, s>>>(ta, tb, p);
  FR_LAUNCH_CHECK();
  return 0;
}

}  // namespace fr

using namespace fr;

// out[M1,N1] += A[K,M1]^T B[K,N1]. A, B bf16 row-major (lda, ldb in elements, multiples of 8), out fp32 (ldo % 4 == 0).
// k_dev optional device int: K_eff = min(K, *k_dev * k_scale). splits <= 0 chooses enough to fill the GPU.
FR_API int fr_gemm_grad_bf16(const void* A, const void* B, float* out, int M1, int N1, int K, int lda, int ldb, int ldo,
                             const int* k_dev, int k_scale, int splits, void* stream) {
  FR_REQUIRE(M1 >= 1 && N1 >= 8 && K >= 1, "fr_gemm_grad_bf16: bad shape M1=%d N1=%d K=%d", M1, N1, K);
  FR_REQUIRE(M1 % 8 == 0 && N1 % 8 == 0 && ldo % 4 == 0 && ((uintptr_t)out & 15) == 0,
             "fr_gemm_grad_bf16: M1=%d, N1=%d must be multiples of 8 and out 16-byte aligned with ldo %% 4 == 0", M1, N1);
  const int BN = N1 > 128 ? 256 : (N1 > 64 ? 128 : 64);
  CUtensorMap ta, tb;
  if (int e = make_tmap_bf16_2d_box(&ta, A, K, M1, lda, GBK, 64)) return e;
  if (int e = make_tmap_bf16_2d_box(&tb, B, K, N1, ldb, GBK, 64)) return e;
  const int tiles = ((M1 + GBM - 1) / GBM) * ((N1 + BN - 1) / BN);
  const int kb_total = (K + GBK - 1) / GBK;
  if (splits <= 0) {
    // One wave of CTAs (197 KB of shared memory = one CTA per SM): measured on the step's shapes
    // (scripts/bench_gemm_grad.py), 9 tiles x 16 splits takes 30.7 us where two waves (33 splits) take 39.9 us, since
    // every CTA pays the pipeline fill and a 128 KB red.global epilogue. At least four k-blocks per split.
    splits = num_sms() / tiles;
    if (splits > kb_total / 4) splits = kb_total / 4;
    if (splits < 1) splits = 1;
  }
  GradParams p;
  p.out = out; p.k_dev = k_dev; p.k_scale = k_scale; p.M1 = M1; p.N1 = N1; p.K = K; p.ldo = ldo; p.splits = splits;
  cudaStream_t s = (cudaStream_t)stream;
  switch (BN) {
    case 64: return launch_grad<64>(ta, tb, p, s);
    case 128: return launch_grad<128>(ta, tb, p, s);
    default: return launch_grad<256>(ta, tb, p, s);
  }
}
