// N1 / N2 / N3 device side: the two federated aggregation paths as single kernels that move data
// over NVLink with plain peer loads/stores (or NVLS multimem instructions), no NCCL.
//
//   fr_grad_allreduce   per-client L2 clip + LDP Gaussian noise + sum over clients + 1/N, result on every rank.
//                       Replaces DDP's bucketed all-reduce, the loss all-reduce (X5) and the
//                       upstream noise add (client.py:87-89).
//   fr_param_average    buf_r <- sum_i w_i * buf_i on every rank r: weighted FedAvg + redistribution of the
//                       global model. Replaces Parameter_Averaging_main.py:144-148 (116 all-reduces), the
//                       server mean server.py:37-55 and the per-tensor broadcasts client.py:261-264.
//
// Memory model. Every rank owns one buffer of identical layout in peer-mapped ("symmetric") memory and a
// small signal pad. `bufs[r]` / `pads[r]` are this process's mappings of rank r's buffer and pad.
// Algorithm (two-shot, in place): rank r owns slice r of the vector. It waits until every peer has
// entered the kernel (their data is final), reduces slice r over all peers' buffers, and stores the
// result into slice r of *every* peer's buffer; a closing barrier makes the full vector visible.
// Slice r is written only by rank r and only after rank r has read it everywhere, so no staging copy.
//
// Clipping needs each client's global norm before any scaled value is used. Instead of a separate
// "scale in place" pass, each rank computes its own sum of squares (grid-wide reduction), publishes
// scale_r = min(1, C/||g_r||) into every peer's pad together with its arrival flag, and the *reader*
// applies scale_r and regenerates client r's noise from the counter-based RNG (seed, r, index). The raw
// gradient is never rewritten and nothing but the final mean crosses NVLink twice.
//
// Cross-GPU ordering: st.release.sys / ld.acquire.sys on monotonically increasing epoch words, so pads
// never need resetting. Single-GPU testing: `ranks_here` > 1 runs several logical ranks as blockIdx.y
// slices of ONE cooperative launch over buffers that all live on the local GPU (ranks never run as
// separate spinning kernels on one device).
#include "common.cuh"
#include <cooperative_groups.h>
#include <string.h>

namespace fr {

constexpr int kMaxRanks = 16;
constexpr int kPadWords = 256;          // uint32 words per signal pad
// pad layout (uint32 index): [0,16) arrive flags, [16,32) done flags, [32,48) scale bits, 64 grid counter, 65 sumsq
constexpr int kPadArrive = 0, kPadDone = 16, kPadScale = 32, kPadGridCtr = 64, kPadSumSq = 65, kPadGridCtr2 = 66,
              kPadEndCtr = 68 /* and 69: alternate by epoch parity */;

struct AggParams {
  float* bufs[kMaxRanks];
  uint32_t* pads[kMaxRanks];
  float weights[kMaxRanks];   // param average: w_i (already normalised); grad: unused
  float* mc_buf;              // NVLS multicast mapping of the buffers (or null)
  int W, rank_base, ranks_here;
  long long n;                // floats, multiple of 4
  uint32_t epoch;             // > 0, strictly increasing per launch on this pad set
  float clip, noise_std, scale;
  unsigned long long seed;
  int uniform_weights;        // param average: all participating weights equal → no in-place pre-scaling needed for NVLS
  // optional fused Adam tail (gradient mode): parameters / moments of each logical rank handled by this launch
  float* adam_p[kMaxRanks];
  float* adam_m[kMaxRanks];
  float* adam_v[kMaxRanks];
  float lr, b1, b2, eps, bc1, bc2;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void spin_until_ge(const uint32_t* p, uint32_t target) {
  uint32_t spins = 0;
  while ((int32_t)(ld_acquire_sys(p) - target) < 0) {
    if (++spins > (1u << 26)) {
      printf("fedrec_b200: cross-rank wait timed out (block %d,%d)\n", (int)blockIdx.x, (int)blockIdx.y);
      __trap();
    }
    __nanosleep(64);
  }
}

// Barrier among the blocks of one logical rank (blockIdx.y fixed). The counter lives in that rank's own pad,
// is zero when the launch starts and is reset by block 0 once no block can still be waiting on it.
__device__ __forceinline__ void rank_grid_sync(uint32_t* ctr, uint32_t target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(ctr) - target) < 0) {
      if (++spins > (1u << 26)) { printf("fedrec_b200: grid sync timed out\n"); __trap(); }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_peer_v4(const float* p) {
  float4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_v4(float* p, float4 v) {
  asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 multimem_ld_reduce_v4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_v4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// kMode 0: gradient (clip + noise + mean).  kMode 1: weighted parameter average.
// kNvls: reduction and redistribution through the switch (values must already be final in memory,
//        so clip/noise/weights are applied in place first).
template <int kMode, bool kNvls>
__global__ void __launch_bounds__(512) aggregate_kernel(const AggParams p) {
  const int r = p.rank_base + blockIdx.y;                  // my logical rank
  const int W = p.W;
  uint32_t* my_pad = p.pads[r];
  float* my_buf = p.bufs[r];
  const uint32_t epoch = p.epoch;
  const size_t n4 = (size_t)p.n / 4;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nth = gridDim.x * blockDim.x;
  __shared__ float red[32];
  __shared__ float s_scale[kMaxRanks];

  float my_scale = 1.f;
  if (kMode == 0 && p.clip > 0.f) {
    // ---- phase A: this client's L2 norm (grid-wide) -------------------------------------------
    float acc = 0.f;
    for (size_t i = tid; i < n4; i += nth) {
      const float4 v = reinterpret_cast<const float4*>(my_buf)[i];
      acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) atomicAdd(reinterpret_cast<float*>(my_pad + kPadSumSq), acc);
    rank_grid_sync(my_pad + kPadGridCtr, gridDim.x);
    const float sq = __uint_as_float(ld_relaxed_sys(my_pad + kPadSumSq));
    my_scale = fminf(1.f, p.clip / fmaxf(sqrtf(sq), 1e-30f));
  }
  // The switch reduces what is in memory, so per-client factors have to be applied in place first — unless there are
  // none (plain mean, or FedAvg with equal weights: one common factor after the reduction).
  const bool prepass = kNvls && (kMode == 1 ? !p.uniform_weights : (p.clip > 0.f || p.noise_std > 0.f));
  if (prepass) {
    const float wgt = (kMode == 1) ? p.weights[r] : my_scale;
    Philox rng(p.seed);
    for (size_t i = tid; i < n4; i += nth) {
      float4 v = reinterpret_cast<float4*>(my_buf)[i];
      v.x *= wgt; v.y *= wgt; v.z *= wgt; v.w *= wgt;
      if (kMode == 0 && p.noise_std > 0.f) {
        uint32_t q[4]; rng(i, (uint32_t)r, q);
        float z0, z1, z2, z3; box_muller(q[0], q[1], z0, z1); box_muller(q[2], q[3], z2, z3);
        v.x += p.noise_std * z0; v.y += p.noise_std * z1; v.z += p.noise_std * z2; v.w += p.noise_std * z3;
      }
      reinterpret_cast<float4*>(my_buf)[i] = v;
    }
    rank_grid_sync(my_pad + kPadGridCtr2, gridDim.x);
  }

  // ---- arrive: tell every peer my buffer is final (and my clip scale) ------------------------------
  if (blockIdx.x == 0 && threadIdx.x < W) {
    const int peer = threadIdx.x;
    if (kMode == 0) p.pads[peer][kPadScale + r] = __float_as_uint(my_scale);
    __threadfence_system();
    st_release_sys(p.pads[peer] + kPadArrive + r, epoch);
  }
  if (threadIdx.x < W) {
    spin_until_ge(my_pad + kPadArrive + threadIdx.x, epoch);
    s_scale[threadIdx.x] = (kMode == 0) ? __uint_as_float(ld_relaxed_sys(my_pad + kPadScale + threadIdx.x))
                                        : p.weights[threadIdx.x];
  }
  __syncthreads();

  // ---- reduce my slice over all clients, store the result into every client's slice ----------------------
  const size_t per = (n4 + W - 1) / W;
  const size_t lo = (size_t)r * per;
  const size_t hi = lo + per < n4 ? lo + per : n4;
  Philox rng(p.seed);
  if (kNvls) {
    // in-switch reduction + multicast store, four independent 16-byte requests in flight per thread
    const float post = (kMode == 0) ? p.scale : (p.uniform_weights ? p.weights[r] : 1.f);
    for (size_t i0 = lo + tid; i0 < hi; i0 += 4 * (size_t)nth) {
      float4 a[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const size_t i = i0 + (size_t)u * nth;
        if (i < hi) a[u] = multimem_ld_reduce_v4(p.mc_buf + 4 * i);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const size_t i = i0 + (size_t)u * nth;
        if (i < hi) {
          a[u].x *= post; a[u].y *= post; a[u].z *= post; a[u].w *= post;
          multimem_st_v4(p.mc_buf + 4 * i, a[u]);
        }
      }
    }
  }
  for (size_t i = lo + tid; i < hi && !kNvls; i += nth) {
    float4 acc;
    {
      float4 v[kMaxRanks];
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q)
        if (q < W) v[q] = ld_peer_v4(p.bufs[(r + q) % W] + 4 * i);      // start with my own copy, stagger peers
      acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q) {
        if (q < W) {
          const int src = (r + q) % W;
          const float s = s_scale[src];
          float4 t = v[q];
          t.x *= s; t.y *= s; t.z *= s; t.w *= s;
          if (kMode == 0 && p.noise_std > 0.f) {
            uint32_t u[4]; rng(i, (uint32_t)src, u);
            float z0, z1, z2, z3; box_muller(u[0], u[1], z0, z1); box_muller(u[2], u[3], z2, z3);
            t.x += p.noise_std * z0; t.y += p.noise_std * z1; t.z += p.noise_std * z2; t.w += p.noise_std * z3;
          }
          acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
        }
      }
      if (kMode == 0) { acc.x *= p.scale; acc.y *= p.scale; acc.z *= p.scale; acc.w *= p.scale; }
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q)
        if (q < W) st_peer_v4(p.bufs[(r + q) % W] + 4 * i, acc);
    }
  }

  // ---- done: all my blocks finished writing → tell every peer; wait for everyone ---------------------------
  __threadfence_system();
  rank_grid_sync(my_pad + kPadEndCtr + (epoch & 1u), gridDim.x);
  if (blockIdx.x == 0 && threadIdx.x < W) st_release_sys(p.pads[threadIdx.x] + kPadDone + r, epoch);
  if (kMode == 0 && p.adam_p[r] != nullptr) {
    // ---- fused optimiser tail: once every peer has delivered its slice, my buffer holds the averaged gradient;
    //      every block waits on the (local) done flags and applies Adam to its share of the flat parameter vector.
    if (threadIdx.x < W) spin_until_ge(my_pad + kPadDone + threadIdx.x, epoch);
    __syncthreads();
    float4* pp = reinterpret_cast<float4*>(p.adam_p[r]);
    float4* mm = reinterpret_cast<float4*>(p.adam_m[r]);
    float4* vv = reinterpret_cast<float4*>(p.adam_v[r]);
    const float step_size = p.lr / p.bc1, inv_sqrt_bc2 = rsqrtf(p.bc2);
    for (size_t i = tid; i < n4; i += nth) {
      const float4 g = reinterpret_cast<const float4*>(my_buf)[i];
      float4 w = pp[i], m1 = mm[i], v2 = vv[i];
#define FR_ADAM1(c)                                                        \
      m1.c = p.b1 * m1.c + (1.f - p.b1) * g.c;                             \
      v2.c = p.b2 * v2.c + (1.f - p.b2) * g.c * g.c;                       \
      w.c -= step_size * (m1.c / (sqrtf(v2.c) * inv_sqrt_bc2 + p.eps));
      FR_ADAM1(x) FR_ADAM1(y) FR_ADAM1(z) FR_ADAM1(w)
#undef FR_ADAM1
      pp[i] = w; mm[i] = m1; vv[i] = v2;
    }
  }
  if (blockIdx.x == 0) {
    if (threadIdx.x < W) spin_until_ge(my_pad + kPadDone + threadIdx.x, epoch);
    __syncthreads();
    if (threadIdx.x == 0) {
      // every block of this rank is past the earlier barriers: make the pad ready for the next launch.
      // The end counter of the *other* parity was last used by the previous launch, which has retired.
      my_pad[kPadGridCtr] = 0u;
      my_pad[kPadGridCtr2] = 0u;
      my_pad[kPadSumSq] = 0u;
      my_pad[kPadEndCtr + ((epoch + 1u) & 1u)] = 0u;
    }
  }
}

}  // namespace fr

using namespace fr;

static int launch_agg(int mode, int nvls, const AggParams& p, int blocks, cudaStream_t s) {
  void* kern = nullptr;
  if (mode == 0) kern = nvls ? (void*)aggregate_kernel<0, true> : (void*)aggregate_kernel<0, false>;
  else kern = nvls ? (void*)aggregate_kernel<1, true> : (void*)aggregate_kernel<1, false>;
  AggParams pc = p;
  void* args[] = {&pc};
  // all blocks of all logical ranks must be co-resident: clamp to what the device can hold
  int dev = 0, sms = 0, per_sm = 0;
  FR_CHECK(cudaGetDevice(&dev));
  FR_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  FR_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 512, 0));
  const int cap = (sms * per_sm) / p.ranks_here;
  FR_REQUIRE(cap >= 1, "aggregate: %d logical ranks do not fit on one GPU", p.ranks_here);
  if (blocks > cap) blocks = cap;
  dim3 grid(blocks, p.ranks_here), block(512);
  // cooperative launch: every block of every logical rank must be co-resident (they wait on each other)
  FR_CHECK(cudaLaunchCooperativeKernel(kern, grid, block, args, 0, s));
  return 0;
}

// bufs_host/pads_host: W device pointers each (this process's mappings). ranks_here = 1 for a real multi-GPU
// run (rank = rank_base); > 1 emulates that many consecutive ranks on the local GPU in one launch.
// epoch must increase by one per call on a given pad set (both entry points share the counter).
FR_API int fr_grad_allreduce(void* const* bufs_host, void* const* pads_host, void* mc_buf, int W, int rank_base,
                             int ranks_here, long long n, unsigned epoch, float clip, float noise_std, float scale,
                             unsigned long long seed, int blocks, void* stream) {
  FR_REQUIRE(W >= 1 && W <= kMaxRanks, "fr_grad_allreduce: W=%d must be in [1,%d]", W, kMaxRanks);
  FR_REQUIRE(n % 4 == 0 && epoch > 0, "fr_grad_allreduce: n=%lld must be a multiple of 4 and epoch > 0", n);
  FR_REQUIRE(ranks_here >= 1 && rank_base >= 0 && rank_base + ranks_here <= W, "fr_grad_allreduce: bad rank range");
  AggParams p{};
  for (int i = 0; i < W; ++i) { p.bufs[i] = (float*)bufs_host[i]; p.pads[i] = (uint32_t*)pads_host[i]; }
  p.mc_buf = (float*)mc_buf; p.W = W; p.rank_base = rank_base; p.ranks_here = ranks_here; p.n = n; p.epoch = epoch;
  p.clip = clip; p.noise_std = noise_std; p.scale = scale; p.seed = seed;
  if (blocks <= 0) blocks = 128;
  return launch_agg(0, mc_buf != nullptr, p, blocks, (cudaStream_t)stream);
}

// fr_grad_allreduce followed, inside the same kernel, by one Adam step of every rank on the averaged gradient
// (params/m/v: per-rank fp32 buffers of n floats, local memory; only entries rank_base .. rank_base+ranks_here-1 are used).
// bc1 = 1 - b1^t, bc2 = 1 - b2^t for the 1-based step count t. One launch replaces DDP all-reduce + clip/noise + Adam.
FR_API int fr_grad_allreduce_adam(void* const* bufs_host, void* const* pads_host, void* mc_buf, int W, int rank_base,
                                  int ranks_here, long long n, unsigned epoch, float clip, float noise_std, float scale,
                                  unsigned long long seed, void* const* params_host, void* const* m_host,
                                  void* const* v_host, float lr, float b1, float b2, float eps, float bc1, float bc2,
                                  int blocks, void* stream) {
  FR_REQUIRE(W >= 1 && W <= kMaxRanks, "fr_grad_allreduce_adam: W=%d must be in [1,%d]", W, kMaxRanks);
  FR_REQUIRE(n % 4 == 0 && epoch > 0, "fr_grad_allreduce_adam: n=%lld must be a multiple of 4 and epoch > 0", n);
  FR_REQUIRE(ranks_here >= 1 && rank_base >= 0 && rank_base + ranks_here <= W, "fr_grad_allreduce_adam: bad rank range");
  AggParams p{};
  for (int i = 0; i < W; ++i) { p.bufs[i] = (float*)bufs_host[i]; p.pads[i] = (uint32_t*)pads_host[i]; }
  for (int i = rank_base; i < rank_base + ranks_here; ++i) {
    p.adam_p[i] = (float*)params_host[i]; p.adam_m[i] = (float*)m_host[i]; p.adam_v[i] = (float*)v_host[i];
    FR_REQUIRE(p.adam_p[i] && p.adam_m[i] && p.adam_v[i], "fr_grad_allreduce_adam: missing optimiser buffers of rank %d", i);
  }
  p.mc_buf = (float*)mc_buf; p.W = W; p.rank_base = rank_base; p.ranks_here = ranks_here; p.n = n; p.epoch = epoch;
  p.clip = clip; p.noise_std = noise_std; p.scale = scale; p.seed = seed;
  p.lr = lr; p.b1 = b1; p.b2 = b2; p.eps = eps; p.bc1 = bc1; p.bc2 = bc2;
  if (blocks <= 0) blocks = 128;
  return launch_agg(0, mc_buf != nullptr, p, blocks, (cudaStream_t)stream);
}

FR_API int fr_param_average(void* const* bufs_host, void* const* pads_host, void* mc_buf, const float* weights_host,
                            int W, int rank_base, int ranks_here, long long n, unsigned epoch, int blocks,
                            void* stream) {
  FR_REQUIRE(W >= 1 && W <= kMaxRanks, "fr_param_average: W=%d must be in [1,%d]", W, kMaxRanks);
  FR_REQUIRE(n % 4 == 0 && epoch > 0, "fr_param_average: n=%lld must be a multiple of 4 and epoch > 0", n);
  FR_REQUIRE(ranks_here >= 1 && rank_base >= 0 && rank_base + ranks_here <= W, "fr_param_average: bad rank range");
  AggParams p{};
  for (int i = 0; i < W; ++i) {
    p.bufs[i] = (float*)bufs_host[i]; p.pads[i] = (uint32_t*)pads_host[i]; p.weights[i] = weights_host[i];
  }
  p.mc_buf = (float*)mc_buf; p.W = W; p.rank_base = rank_base; p.ranks_here = ranks_here; p.n = n; p.epoch = epoch;
  p.scale = 1.f;
  p.uniform_weights = 1;                     // every weight equal (a zero-weight server rank breaks uniformity)
  for (int i = 1; i < W; ++i)
    if (weights_host[i] != weights_host[0]) p.uniform_weights = 0;
  if (blocks <= 0) blocks = 128;
  return launch_agg(1, mc_buf != nullptr, p, blocks, (cudaStream_t)stream);
}

FR_API int fr_signal_pad_bytes() { return kPadWords * 4; }

// ---- N3 host side: peer-mapped allocations through CUDA IPC -------------------------------------------------
FR_API int fr_symm_alloc(void** ptr_out, long long bytes) {
  FR_CHECK(cudaMalloc(ptr_out, (size_t)bytes));
  FR_CHECK(cudaMemset(*ptr_out, 0, (size_t)bytes));
  return 0;
}
FR_API int fr_symm_free(void* ptr) { FR_CHECK(cudaFree(ptr)); return 0; }
FR_API int fr_symm_export(void* ptr, void* handle64_host) {
  cudaIpcMemHandle_t h;
  FR_CHECK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle64_host, &h, sizeof(h));
  return 0;
}
FR_API int fr_symm_import(const void* handle64_host, void** ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64_host, sizeof(h));
  FR_CHECK(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
FR_API int fr_symm_close(void* ptr) { FR_CHECK(cudaIpcCloseMemHandle(ptr)); return 0; }
