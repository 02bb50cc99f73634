// N1 / N2 / N3 device side: the two federated aggregation paths over NVLink with plain peer loads/stores or NVLS
// multimem instructions, no NCCL.
//
//   gradient mode   per-client L2 clip + LDP Gaussian noise + sum over clients + 1/N (+ Adam on the result).
//                   Replaces DDP's bucketed all-reduce, the loss all-reduce (X5), the upstream noise
//                   add (client.py:87-89) and the two Adam steps (model.py:89,93).
//   parameter mode  buf_r <- sum_i w_i * buf_i on every rank r: weighted FedAvg + redistribution of the global model.
//                   Replaces Parameter_Averaging_main.py:144-148 (116 all-reduces), the server mean
//                   server.py:37-55 and the per-tensor broadcasts client.py:261-264.
//
// Memory model. Every rank owns one buffer of identical layout in peer-mapped ("symmetric") memory and a small
// signal pad. `bufs[r]` / `pads[r]` are this process's mappings of rank r's buffer and pad.
//
// One aggregation is a short chain of kernels on one stream (round 2 redesign; round 1 was a single cooperative
// kernel whose blocks spun on each other and on the peers, which (a) held SMs for as long as the slowest rank was
// late, (b) needed every block co-resident and (c) could only trap when a peer never showed up):
//
//   sumsq     (gradient mode with clipping) this client's sum of squares -> pad
//   publish   clip scale from the sum of squares; in NVLS mode the per-client factors (clip scale, noise, FedAvg
//             weight) are applied in place because the switch reduces what is in memory; the LAST block to finish
//             releases scale + arrival flag into every peer's pad (st.release.sys). No block waits on anything.
//   wait      ONE warp: lane q polls "rank q has arrived" with a %globaltimer deadline. This is the only place a
//             rank waits for its peers, it occupies one warp slot (it fits beside a resident GEMM CTA), and on expiry
//             it writes {rank, peer, phase, have, want} to a host-mapped record, raises the abort word in every pad
//             and returns: the kernels behind it see the abort word and do nothing. The host turns the record into a
//             Python exception naming the silent rank (no printf storm, no __trap, the context stays usable).
//   reduce    THE fused compute+collective kernel: rank r owns slice r. It reduces slice r over all peers (plain peer
//             loads applying each client's clip scale and regenerating each client's Philox noise, or one
//             multimem.ld_reduce through the switch), scales by 1/N and stores the result into slice r of EVERY
//             peer's buffer (peer stores or multimem.st). The last block to finish releases the "done" flag into
//             every pad. No block waits on anything.
//   wait      lane q polls "rank q is done": my buffer now holds the full result.
//   adam      (gradient mode, optional) Adam over the flat parameters from the averaged gradient; bias corrections
//             from a device-side step counter so the chain can sit in a CUDA graph.
//   commit    one thread: step counter, "pending" word.
//
// Epochs live in the pad (device memory), not on the host: the chain is replayable from a CUDA graph, and the host
// cannot get out of step with the device. Every rank must issue the same sequence of aggregations.
// `gated` chains run only while the pad's "pending" word is set (fr_agg_mark_pending): the training step sets it
// after its backward pass and the chain that applies the update sits at the START of the next step's graph, beside
// the frozen trunk, which needs nothing the aggregation writes (software pipelining).
//
// Single-GPU testing: `ranks_here` > 1 runs several logical ranks as blockIdx.y slices of every launch over buffers
// that all live on the local GPU. No kernel waits on another block, so no cooperative launch is needed and the
// emulated ranks run exactly the code of the real ones.
#include "common.cuh"
#include <string.h>

namespace fr {

constexpr int kMaxRanks = 16;
constexpr int kPadWords = 256;          // uint32 words per signal pad
// pad layout (uint32 index)
constexpr int kPadArrive = 0;           // [0,16)  arrival epoch of rank q
constexpr int kPadDone = 16;            // [16,32) done epoch of rank q
constexpr int kPadScale = 32;           // [32,48) clip scale of rank q (float bits)
constexpr int kPadEpoch = 64;           // aggregations published by this rank
constexpr int kPadSumSq = 65;           // float: this rank's sum of squares (zero between aggregations)
constexpr int kPadCtrPublish = 66;      // blocks of the publish kernel that have finished (zero between launches)
constexpr int kPadCtrReduce = 67;       // same for the reduce kernel
constexpr int kPadAbort = 70;           // != 0: an aggregation failed; every kernel of the chain returns at once
constexpr int kPadErrLock = 71;         // first failing thread takes it and fills the error record
constexpr int kPadPending = 72;         // != 0: gradients of a finished step wait to be aggregated (gated chains)
constexpr int kPadStep = 73;            // Adam steps applied (device-side copy of flat.step)
constexpr int kErrWords = 16;           // int32 words of one error record (host-mapped)

struct AggParams {
  float* bufs[kMaxRanks];
  uint32_t* pads[kMaxRanks];
  float weights[kMaxRanks];   // parameter mode: w_i (already normalised)
  float* mc_buf;              // NVLS multicast mapping of the buffers (or null)
  int W, rank_base, ranks_here;
  long long n;                // floats, multiple of 4
  float clip, noise_std, scale;
  unsigned long long seed;
  int uniform_weights;        // parameter mode: all weights equal -> no in-place pre-scaling needed for NVLS
  int gated;                  // run only while pad[kPadPending] != 0
  unsigned long long timeout_ns;
  int* err;                   // host-mapped error record (kErrWords ints) or null
  // optional Adam (gradient mode): parameters / moments of each logical rank handled by this launch
  float* adam_p[kMaxRanks];
  float* adam_m[kMaxRanks];
  float* adam_v[kMaxRanks];
  float lr, eps;
  double b1, b2;
  int step_host;              // > 0: 1-based step count from the host; 0: pad[kPadStep] + 1
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
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
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
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

// true when this launch has nothing to do: a failed aggregation is pending on the host, or a gated chain without
// pending gradients. The pending word only changes in kernels that are alone on the stream, so it is uniform over the
// grid. The abort word can be raised by a peer's waiter at any moment: a launch that some blocks skipped leaves a
// half-written buffer and an unreleased flag behind, which is fine because nothing after an abort is used - the waiters
// report, the host resets the pads (fr_agg_pad_reset) and re-synchronises the state (FederatedRunner._degrade).
__device__ __forceinline__ bool chain_idle(const AggParams& p, const uint32_t* my_pad) {
  if (ld_relaxed_sys(my_pad + kPadAbort) != 0u) return true;
  return p.gated && ld_relaxed_sys(my_pad + kPadPending) == 0u;
}

// Every thread of the block has finished its global writes. Returns true in every thread of the LAST block of this
// logical rank to get here (its view includes all the other blocks' writes); resets the counter for the next launch.
__device__ __forceinline__ bool last_block_done(uint32_t* ctr, int* s_flag) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t old = atomicAdd(ctr, 1u);
    const int last = (old == gridDim.x - 1u);
    if (last) { *ctr = 0u; __threadfence_system(); }
    *s_flag = last;
  }
  __syncthreads();
  return *s_flag != 0;
}

__device__ __forceinline__ uint64_t noise_key(const AggParams& p, uint32_t epoch) {
  return p.seed + 7919ull * (uint64_t)epoch;
}

// ---- sumsq: pad[kPadSumSq] += sum(buf^2) ---------------------------------------------------------------------------
__global__ void __launch_bounds__(512) agg_sumsq_kernel(const AggParams p) {
  const int r = p.rank_base + blockIdx.y;
  uint32_t* my_pad = p.pads[r];
  if (chain_idle(p, my_pad)) return;
  __shared__ float red[32];
  const float4* x = reinterpret_cast<const float4*>(p.bufs[r]);
  const size_t n4 = (size_t)p.n / 4;
  float acc = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = x[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(reinterpret_cast<float*>(my_pad + kPadSumSq), acc);
}

// ---- publish: per-client factors in place when the switch will reduce; arrival flag + clip scale to every peer -------
// kMode 0: gradient.  kMode 1: weighted parameter average.  kInPlace: apply the per-client factor (and noise) to my buffer.
template <int kMode, bool kInPlace>
__global__ void __launch_bounds__(512) agg_publish_kernel(const AggParams p) {
  const int r = p.rank_base + blockIdx.y;
  uint32_t* my_pad = p.pads[r];
  if (chain_idle(p, my_pad)) return;
  __shared__ int s_last;
  const uint32_t epoch = ld_relaxed_sys(my_pad + kPadEpoch) + 1u;
  float my_scale = 1.f;
  if (kMode == 0 && p.clip > 0.f) {
    const float sq = __uint_as_float(ld_relaxed_sys(my_pad + kPadSumSq));
    my_scale = fminf(1.f, p.clip / fmaxf(sqrtf(sq), 1e-30f));
  }
  if (kInPlace) {
    const float wgt = (kMode == 1) ? p.weights[r] : my_scale;
    const bool noisy = (kMode == 0 && p.noise_std > 0.f);
    if (wgt != 1.f || noisy) {
      float4* x = reinterpret_cast<float4*>(p.bufs[r]);
      const size_t n4 = (size_t)p.n / 4;
      Philox rng(noise_key(p, epoch));
      for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 v = x[i];
        v.x *= wgt; v.y *= wgt; v.z *= wgt; v.w *= wgt;
        if (noisy) {
          uint32_t q[4]; rng(i, (uint32_t)r, q);
          float z0, z1, z2, z3; box_muller(q[0], q[1], z0, z1); box_muller(q[2], q[3], z2, z3);
          v.x += p.noise_std * z0; v.y += p.noise_std * z1; v.z += p.noise_std * z2; v.w += p.noise_std * z3;
        }
        x[i] = v;
      }
    }
  }
  if (last_block_done(my_pad + kPadCtrPublish, &s_last)) {
    // my buffer is final: tell every peer (and myself), then advance my own epoch word
    if (threadIdx.x < p.W) {
      uint32_t* peer = p.pads[threadIdx.x];
      if (kMode == 0) st_relaxed_sys(peer + kPadScale + r, __float_as_uint(kInPlace ? 1.f : my_scale));
      __threadfence_system();
      st_release_sys(peer + kPadArrive + r, epoch);
    }
    __syncthreads();
    if (threadIdx.x == 0) st_relaxed_sys(my_pad + kPadEpoch, epoch);
  }
}

// ---- wait: lane q waits until rank q has published `which` (arrive / done) for my current epoch --------------------
__device__ __noinline__ void report_failure(const AggParams& p, uint32_t* my_pad, int status, int r, int peer, int which,
                                            uint32_t have, uint32_t epoch, unsigned long long waited_ns) {
  // status 1: this rank's wait ran out (peer `peer` never published); 2: a peer raised the abort word first
  if (atomicCAS(my_pad + kPadErrLock, 0u, 1u) == 0u && p.err != nullptr) {
    volatile int* e = p.err;
    e[1] = r; e[2] = peer; e[3] = which == kPadArrive ? 1 : 2; e[4] = (int)have; e[5] = (int)epoch;
    e[6] = (int)(waited_ns / 1000000ull);
    __threadfence_system();
    e[0] = status;
  }
  // everyone stops: the peers' waiters see the word and return without waiting for their own deadline
  for (int k = 0; k < p.W; ++k) st_relaxed_sys(p.pads[k] + kPadAbort, epoch ? epoch : 1u);
  __threadfence_system();
}

__global__ void __launch_bounds__(32) agg_wait_kernel(const AggParams p, const int which /* kPadArrive | kPadDone */) {
  const int r = p.rank_base + blockIdx.y;
  uint32_t* my_pad = p.pads[r];
  const int q = threadIdx.x;
  const uint32_t epoch = ld_relaxed_sys(my_pad + kPadEpoch);
  if (ld_relaxed_sys(my_pad + kPadAbort) != 0u) {      // a peer gave up before this rank even got here: tell the host
    if (q == 0) report_failure(p, my_pad, 2, r, -1, which, 0u, epoch, 0ull);
    return;
  }
  if (p.gated && ld_relaxed_sys(my_pad + kPadPending) == 0u) return;
  if (q >= p.W) return;
  const uint32_t* word = my_pad + which + q;
  const unsigned long long t0 = global_timer_ns();
  uint32_t have = ld_acquire_sys(word), spins = 0;
  while ((int32_t)(have - epoch) < 0) {
    __nanosleep(40);
    if ((++spins & 31u) == 0u) {
      const bool aborted = ld_relaxed_sys(my_pad + kPadAbort) != 0u;
      const unsigned long long waited = global_timer_ns() - t0;
      if (aborted || (p.timeout_ns != 0ull && waited > p.timeout_ns)) {
        report_failure(p, my_pad, aborted ? 2 : 1, r, q, which, have, epoch, waited);
        return;
      }
    }
    have = ld_acquire_sys(word);
  }
}

// ---- reduce: my slice over all clients, result into every client's slice -------------------------------------------
template <int kMode, bool kNvls>
__global__ void __launch_bounds__(512) agg_reduce_kernel(const AggParams p) {
  const int r = p.rank_base + blockIdx.y;
  const int W = p.W;
  uint32_t* my_pad = p.pads[r];
  if (chain_idle(p, my_pad)) return;
  __shared__ int s_last;
  __shared__ float s_scale[kMaxRanks];
  const uint32_t epoch = ld_relaxed_sys(my_pad + kPadEpoch);
  const size_t n4 = (size_t)p.n / 4;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nth = (size_t)gridDim.x * blockDim.x;
  if (threadIdx.x < W)
    s_scale[threadIdx.x] = (kMode == 0) ? __uint_as_float(ld_relaxed_sys(my_pad + kPadScale + threadIdx.x))
                                        : p.weights[threadIdx.x];
  __syncthreads();
  const size_t per = (n4 + W - 1) / W;
  const size_t lo = (size_t)r * per;
  const size_t hi = lo + per < n4 ? lo + per : n4;
  if (kNvls) {
    // in-switch reduction + multicast store, four independent 16-byte requests in flight per thread
    const float post = (kMode == 0) ? p.scale : (p.uniform_weights ? p.weights[r] : 1.f);
    for (size_t i0 = lo + tid; i0 < hi; i0 += 4 * nth) {
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
  } else {
    Philox rng(noise_key(p, epoch));
    for (size_t i = lo + tid; i < hi; i += nth) {
      float4 v[kMaxRanks];
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q)
        if (q < W) v[q] = ld_peer_v4(p.bufs[(r + q) % W] + 4 * i);      // start with my own copy, stagger peers
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
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
  if (last_block_done(my_pad + kPadCtrReduce, &s_last)) {
    if (threadIdx.x < W) st_release_sys(p.pads[threadIdx.x] + kPadDone + r, epoch);
    if (threadIdx.x == 0) st_relaxed_sys(my_pad + kPadSumSq, 0u);
  }
}

// ---- adam over the flat parameters from the averaged gradient (my buffer) ---------------------------------------------
__global__ void __launch_bounds__(512) agg_adam_kernel(const AggParams p) {
  const int r = p.rank_base + blockIdx.y;
  uint32_t* my_pad = p.pads[r];
  if (chain_idle(p, my_pad)) return;
  __shared__ float s_bc[2];
  if (threadIdx.x == 0) {
    const double t = p.step_host > 0 ? (double)p.step_host : (double)(ld_relaxed_sys(my_pad + kPadStep) + 1u);
    s_bc[0] = (float)(1.0 - pow(p.b1, t));
    s_bc[1] = (float)(1.0 - pow(p.b2, t));
  }
  __syncthreads();
  const float b1 = (float)p.b1, b2 = (float)p.b2;
  const float step_size = p.lr / s_bc[0], inv_sqrt_bc2 = rsqrtf(s_bc[1]);
  const size_t n4 = (size_t)p.n / 4;
  const float4* gg = reinterpret_cast<const float4*>(p.bufs[r]);
  float4* pp = reinterpret_cast<float4*>(p.adam_p[r]);
  float4* mm = reinterpret_cast<float4*>(p.adam_m[r]);
  float4* vv = reinterpret_cast<float4*>(p.adam_v[r]);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 g = gg[i];
    float4 w = pp[i], m1 = mm[i], v2 = vv[i];
#define FR_ADAM1(c)                                                        \
    m1.c = b1 * m1.c + (1.f - b1) * g.c;                                   \
    v2.c = b2 * v2.c + (1.f - b2) * g.c * g.c;                             \
    w.c -= step_size * (m1.c / (sqrtf(v2.c) * inv_sqrt_bc2 + p.eps));
    FR_ADAM1(x) FR_ADAM1(y) FR_ADAM1(z) FR_ADAM1(w)
#undef FR_ADAM1
    pp[i] = w; mm[i] = m1; vv[i] = v2;
  }
}

// ---- commit: the chain has run; one thread per logical rank ----------------------------------------------------------
__global__ void agg_commit_kernel(const AggParams p, const int with_adam) {
  const int r = p.rank_base + blockIdx.x;
  uint32_t* my_pad = p.pads[r];
  if (threadIdx.x != 0 || chain_idle(p, my_pad)) return;
  if (with_adam) my_pad[kPadStep] = p.step_host > 0 ? (uint32_t)p.step_host : my_pad[kPadStep] + 1u;
  my_pad[kPadPending] = 0u;
  my_pad[kPadSumSq] = 0u;               // (the reduce kernel also clears it; with one rank there is no reduce kernel)
}

__global__ void agg_mark_pending_kernel(uint32_t* pad0, uint32_t* pad1, uint32_t* pad2, uint32_t* pad3, uint32_t arm) {
  uint32_t* pads[4] = {pad0, pad1, pad2, pad3};
  if (arm != 0u && threadIdx.x < 4 && pads[threadIdx.x] != nullptr) pads[threadIdx.x][kPadPending] = 1u;
}

}  // namespace fr

using namespace fr;

namespace {

struct ChainOpts {
  int mode = 0;            // 0 gradient, 1 parameter average
  bool adam = false;
  int blocks = 0;
};

int fill_common(AggParams& p, void* const* bufs_host, void* const* pads_host, void* mc_buf, int W, int rank_base,
                int ranks_here, long long n, const char* who) {
  FR_REQUIRE(W >= 1 && W <= kMaxRanks, "%s: W=%d must be in [1,%d]", who, W, kMaxRanks);
  FR_REQUIRE(n % 4 == 0 && n > 0, "%s: n=%lld must be a positive multiple of 4", who, n);
  FR_REQUIRE(ranks_here >= 1 && rank_base >= 0 && rank_base + ranks_here <= W, "%s: bad rank range", who);
  for (int i = 0; i < W; ++i) {
    p.bufs[i] = (float*)bufs_host[i];
    p.pads[i] = (uint32_t*)pads_host[i];
    FR_REQUIRE(p.bufs[i] && p.pads[i], "%s: missing buffer or pad of rank %d", who, i);
  }
  p.mc_buf = (float*)mc_buf; p.W = W; p.rank_base = rank_base; p.ranks_here = ranks_here; p.n = n;
  return 0;
}

// Launches the whole chain on `s`. Nothing here synchronises or allocates; every launch is graph-capturable.
int launch_chain(const AggParams& p, const ChainOpts& o, cudaStream_t s) {
  int sms = num_sms(), dev = 0;
  FR_CHECK(cudaGetDevice(&dev));
  FR_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // default grid: 64 blocks for the 4.66 MB gradient (it shares the GPU with the trunk), more for buffers whose in-place pass
  // and Adam would otherwise be bandwidth-starved (one block per 256 KB, at most two per SM)
  int blocks = o.blocks;
  if (blocks <= 0) {
    const long long by_size = p.n / (64 * 1024);
    blocks = by_size < 64 ? 64 : (by_size > 2LL * sms ? 2 * sms : (int)by_size);
  }
  if (blocks > 2 * sms) blocks = 2 * sms;
  const bool nvls = p.mc_buf != nullptr && p.W > 1;
  const dim3 grid(blocks, p.ranks_here), block(512), one(1, p.ranks_here);
  // the per-client factors have to be in memory when the switch (or nobody, W == 1) reduces
  bool inplace;
  if (o.mode == 0) inplace = (nvls || p.W == 1) && (p.clip > 0.f || p.noise_std > 0.f);
  else inplace = nvls && !p.uniform_weights;
  if (o.mode == 0 && p.clip > 0.f) agg_sumsq_kernel<<<grid, block, 0, s>>>(p);
  if (p.W > 1 || inplace) {
    if (o.mode == 0) {
      if (inplace) agg_publish_kernel<0, true><<<grid, block, 0, s>>>(p);
      else agg_publish_kernel<0, false><<<dim3(1, p.ranks_here), block, 0, s>>>(p);
    } else {
      if (inplace) agg_publish_kernel<1, true><<<grid, block, 0, s>>>(p);
      else agg_publish_kernel<1, false><<<dim3(1, p.ranks_here), block, 0, s>>>(p);
    }
  }
  if (p.W > 1) {
    agg_wait_kernel<<<one, 32, 0, s>>>(p, kPadArrive);
    if (o.mode == 0) {
      if (nvls) agg_reduce_kernel<0, true><<<grid, block, 0, s>>>(p);
      else agg_reduce_kernel<0, false><<<grid, block, 0, s>>>(p);
    } else {
      if (nvls) agg_reduce_kernel<1, true><<<grid, block, 0, s>>>(p);
      else agg_reduce_kernel<1, false><<<grid, block, 0, s>>>(p);
    }
    agg_wait_kernel<<<one, 32, 0, s>>>(p, kPadDone);
  }
  if (o.adam) agg_adam_kernel<<<grid, block, 0, s>>>(p);
  agg_commit_kernel<<<p.ranks_here, 32, 0, s>>>(p, o.adam ? 1 : 0);
  FR_LAUNCH_CHECK();
  return 0;
}

unsigned long long ms_to_ns(long long ms) { return ms > 0 ? (unsigned long long)ms * 1000000ull : 0ull; }

}  // namespace

// Number of kernels one chain launches (bench.py counts launches).
FR_API int fr_agg_chain_launches(int W, int mode, int nvls, float clip, float noise_std, int uniform_weights, int adam) {
  const bool sw = nvls && W > 1;
  bool inplace = mode == 0 ? ((sw || W == 1) && (clip > 0.f || noise_std > 0.f)) : (sw && !uniform_weights);
  int k = 1;                                             // commit
  if (mode == 0 && clip > 0.f) ++k;
  if (W > 1 || inplace) ++k;
  if (W > 1) k += 3;
  if (adam) ++k;
  return k;
}

// bufs_host/pads_host: W device pointers each (this process's mappings). ranks_here = 1 for a real multi-GPU
// run (rank = rank_base); > 1 emulates that many consecutive ranks on the local GPU.
// timeout_ms bounds each cross-rank wait (0 = wait for ever); err_mapped: device-visible address of a host-mapped
// error record (fr_agg_errbuf_alloc) or NULL. gated != 0: do nothing unless fr_agg_mark_pending ran since the last chain.
FR_API int fr_grad_allreduce(void* const* bufs_host, void* const* pads_host, void* mc_buf, int W, int rank_base,
                             int ranks_here, long long n, float clip, float noise_std, float scale,
                             unsigned long long seed, int gated, int blocks, long long timeout_ms, void* err_mapped,
                             void* stream) {
  AggParams p{};
  if (int rc = fill_common(p, bufs_host, pads_host, mc_buf, W, rank_base, ranks_here, n, "fr_grad_allreduce")) return rc;
  p.clip = clip; p.noise_std = noise_std; p.scale = scale; p.seed = seed; p.gated = gated;
  p.timeout_ns = ms_to_ns(timeout_ms); p.err = (int*)err_mapped;
  ChainOpts o; o.mode = 0; o.blocks = blocks;
  return launch_chain(p, o, (cudaStream_t)stream);
}

// fr_grad_allreduce followed on the same stream by one Adam step of every rank on the averaged gradient
// (params/m/v: per-rank fp32 buffers of n floats, local memory; only entries rank_base .. rank_base+ranks_here-1 are
// used). step > 0: the 1-based step count; step == 0: the device-side counter in the pad (CUDA-graph replays).
FR_API int fr_grad_allreduce_adam(void* const* bufs_host, void* const* pads_host, void* mc_buf, int W, int rank_base,
                                  int ranks_here, long long n, float clip, float noise_std, float scale,
                                  unsigned long long seed, void* const* params_host, void* const* m_host,
                                  void* const* v_host, float lr, double b1, double b2, float eps, int step, int gated,
                                  int blocks, long long timeout_ms, void* err_mapped, void* stream) {
  AggParams p{};
  if (int rc = fill_common(p, bufs_host, pads_host, mc_buf, W, rank_base, ranks_here, n, "fr_grad_allreduce_adam")) return rc;
  for (int i = rank_base; i < rank_base + ranks_here; ++i) {
    p.adam_p[i] = (float*)params_host[i]; p.adam_m[i] = (float*)m_host[i]; p.adam_v[i] = (float*)v_host[i];
    FR_REQUIRE(p.adam_p[i] && p.adam_m[i] && p.adam_v[i], "fr_grad_allreduce_adam: missing optimiser buffers of rank %d", i);
  }
  p.clip = clip; p.noise_std = noise_std; p.scale = scale; p.seed = seed; p.gated = gated;
  p.timeout_ns = ms_to_ns(timeout_ms); p.err = (int*)err_mapped;
  p.lr = lr; p.b1 = b1; p.b2 = b2; p.eps = eps; p.step_host = step;
  ChainOpts o; o.mode = 0; o.adam = true; o.blocks = blocks;
  return launch_chain(p, o, (cudaStream_t)stream);
}

FR_API int fr_param_average(void* const* bufs_host, void* const* pads_host, void* mc_buf, const float* weights_host,
                            int W, int rank_base, int ranks_here, long long n, int blocks, long long timeout_ms,
                            void* err_mapped, void* stream) {
  AggParams p{};
  if (int rc = fill_common(p, bufs_host, pads_host, mc_buf, W, rank_base, ranks_here, n, "fr_param_average")) return rc;
  p.scale = 1.f;
  p.uniform_weights = 1;                     // every weight equal (a zero-weight server rank breaks uniformity)
  for (int i = 0; i < W; ++i) {
    p.weights[i] = weights_host[i];
    if (weights_host[i] != weights_host[0]) p.uniform_weights = 0;
  }
  p.timeout_ns = ms_to_ns(timeout_ms); p.err = (int*)err_mapped;
  ChainOpts o; o.mode = 1; o.blocks = blocks > 0 ? blocks : (n >= (32LL << 20) ? 0 : 128);
  return launch_chain(p, o, (cudaStream_t)stream);
}

// The step's backward pass has left fresh gradients in the flat buffer: arm the next gated chain (up to 4 local pads).
// arm == 0 launches the kernel without arming anything (warm-up passes before a graph capture).
FR_API int fr_agg_mark_pending(void* const* pads_host, int count, int arm, void* stream) {
  FR_REQUIRE(count >= 1 && count <= 4, "fr_agg_mark_pending: count=%d must be in [1,4]", count);
  uint32_t* q[4] = {nullptr, nullptr, nullptr, nullptr};
  for (int i = 0; i < count; ++i) q[i] = (uint32_t*)pads_host[i];
  agg_mark_pending_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(q[0], q[1], q[2], q[3], arm ? 1u : 0u);
  FR_LAUNCH_CHECK();
  return 0;
}

// Host-side view of / writes to one pad (synchronous; for set-up, recovery and tests, never on the step path).
FR_API int fr_agg_pad_read(const void* pad, unsigned* out_words_host, int first, int count) {
  FR_REQUIRE(first >= 0 && count >= 1 && first + count <= kPadWords, "fr_agg_pad_read: bad range");
  FR_CHECK(cudaMemcpy(out_words_host, (const uint32_t*)pad + first, (size_t)count * 4, cudaMemcpyDeviceToHost));
  return 0;
}
FR_API int fr_agg_pad_set_step(void* pad, unsigned step) {
  FR_CHECK(cudaMemcpy((uint32_t*)pad + kPadStep, &step, 4, cudaMemcpyHostToDevice));
  return 0;
}
// After a failed aggregation, with no kernel of any rank in flight: back to the initial state (epoch 0, no flags).
FR_API int fr_agg_pad_reset(void* pad, int keep_step) {
  uint32_t step = 0;
  if (keep_step) FR_CHECK(cudaMemcpy(&step, (uint32_t*)pad + kPadStep, 4, cudaMemcpyDeviceToHost));
  FR_CHECK(cudaMemset(pad, 0, kPadWords * 4));
  if (keep_step) FR_CHECK(cudaMemcpy((uint32_t*)pad + kPadStep, &step, 4, cudaMemcpyHostToDevice));
  return 0;
}

FR_API int fr_signal_pad_bytes() { return kPadWords * 4; }

// Host-mapped error record: the wait kernel writes it, the host reads it without a device synchronisation.
FR_API int fr_agg_errbuf_alloc(void** host_out, void** dev_out) {
  void* h = nullptr;
  FR_CHECK(cudaHostAlloc(&h, kErrWords * sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable));
  memset(h, 0, kErrWords * sizeof(int));
  void* d = nullptr;
  FR_CHECK(cudaHostGetDevicePointer(&d, h, 0));
  *host_out = h; *dev_out = d;
  return 0;
}
FR_API int fr_agg_errbuf_free(void* host) { FR_CHECK(cudaFreeHost(host)); return 0; }

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
