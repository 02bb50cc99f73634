// Small fused ops: candidate scoring + sigmoid + softmax-CE (N6), news-gradient
// segment-sum (N11), fused Adam over the flat trainable buffer (N12), and the
// sync-free per-step news-id de-duplication that feeds the text encoder.
//
// Reference behaviour being replaced: model.py:121-126 (bmm, sigmoid, CrossEntropyLoss),
// main.py:20-42 (Python dict segment-sum), model.py:22-23,89,93 (two torch Adam steps).
#include "common.cuh"
#include <string.h>

namespace fr {

// ---------------------------------------------------------------------------------------
// score = sigmoid(cand · user), loss = mean_b CE(score_b, label_b); optional backward.
// One warp per impression. C <= 32 candidates, any D (fp32, 16-byte aligned rows if D%4==0).
// ---------------------------------------------------------------------------------------
template <bool kBackward>
__global__ void __launch_bounds__(128) score_ce_kernel(
    const float* __restrict__ cand, const float* __restrict__ user, const int* __restrict__ labels,
    float* __restrict__ score, float* __restrict__ loss_rows, float* __restrict__ loss_mean,
    float* __restrict__ d_cand, float* __restrict__ d_user, float grad_scale, int B, int C, int D) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= B) return;
  const float* u = user + (size_t)warp * D;
  const float* cb = cand + (size_t)warp * C * D;
  float my_s = 0.f;  // lane c holds the logit of candidate c
  for (int c = 0; c < C; ++c) {
    float acc = 0.f;
    for (int d = lane; d < D; d += 32) acc = fmaf(cb[(size_t)c * D + d], u[d], acc);
    acc = warp_sum(acc);
    if (lane == c) my_s = acc;
  }
  const bool act = lane < C;
  const float p = act ? 1.f / (1.f + __expf(-my_s)) : -INFINITY;   // sigmoid, then softmax over p
  const float mx = warp_max(p);
  const float e = act ? __expf(p - mx) : 0.f;
  const float z = warp_sum(e);
  const int lab = labels ? labels[warp] : 0;
  const float p_lab = __shfl_sync(0xffffffffu, p, lab);
  const float loss = -(p_lab - mx - __logf(z));
  if (act && score) score[(size_t)warp * C + lane] = p;
  if (lane == 0) {
    if (loss_rows) loss_rows[warp] = loss;
    if (loss_mean) atomicAdd(loss_mean, loss / (float)B);
  }
  if (kBackward) {
    // dL/dp_c = (softmax_c - [c==lab]) * grad_scale/B ; dL/ds_c = dL/dp_c * p(1-p)
    const float sm = e / z;
    float ds = act ? (sm - (lane == lab ? 1.f : 0.f)) * (grad_scale / (float)B) * p * (1.f - p) : 0.f;
    for (int d = lane; d < D; d += 32) {
      const float ud = u[d];
      float du = 0.f;
      for (int c = 0; c < C; ++c) {
        const float dsc = __shfl_sync(0xffffffffu, ds, c);
        d_cand[((size_t)warp * C + c) * D + d] = dsc * ud;
        du = fmaf(dsc, cb[(size_t)c * D + d], du);
      }
      d_user[(size_t)warp * D + d] = du;
    }
  }
}

// ---------------------------------------------------------------------------------------
// out[seg[k], :] += vals[k, :]   (out zeroed by the caller entry point). D % 4 == 0.
// One warp per source row; vector reductions (red.global.add.v4.f32, sm_90+).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void red_add_v4(float* addr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

__global__ void __launch_bounds__(256) segment_sum_kernel(const float* __restrict__ vals, const int* __restrict__ seg,
                                                          float* __restrict__ out, int K, int S, int D4) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= K) return;
  const int s = seg[warp];
  if (s < 0 || s >= S) return;
  const float4* src = reinterpret_cast<const float4*>(vals) + (size_t)warp * D4;
  float* dst = out + (size_t)s * D4 * 4;
  for (int i = lane; i < D4; i += 32) red_add_v4(dst + 4 * i, src[i]);
}

// ---------------------------------------------------------------------------------------
// Adam (torch.optim.Adam defaults: no weight decay, no amsgrad), whole flat buffer, float4.
// bias corrections come from the host (bc == nullptr) or from a device pair written by
// adam_advance_kernel so the step can sit inside a CUDA graph.
// ---------------------------------------------------------------------------------------
__global__ void adam_advance_kernel(int* step, float* bc, float b1, float b2) {
  const int t = *step + 1;
  *step = t;
  bc[0] = 1.f - powf(b1, (float)t);
  bc[1] = 1.f - powf(b2, (float)t);
}

__global__ void __launch_bounds__(256) adam_kernel(float4* __restrict__ p, const float4* __restrict__ g,
                                                   float4* __restrict__ m, float4* __restrict__ v, size_t n4,
                                                   float lr, float b1, float b2, float eps, float bc1, float bc2,
                                                   const float* __restrict__ bc_dev, float grad_scale) {
  if (bc_dev) { bc1 = bc_dev[0]; bc2 = bc_dev[1]; }
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 pp = p[i], gg = g[i], mm = m[i], vv = v[i];
#define FR_ADAM1(c)                                                   \
    {                                                                 \
      const float gr = gg.c * grad_scale;                             \
      mm.c = b1 * mm.c + (1.f - b1) * gr;                             \
      vv.c = b2 * vv.c + (1.f - b2) * gr * gr;                        \
      const float denom = sqrtf(vv.c) * inv_sqrt_bc2 + eps;           \
      pp.c -= step_size * (mm.c / denom);                             \
    }
    FR_ADAM1(x) FR_ADAM1(y) FR_ADAM1(z) FR_ADAM1(w)
#undef FR_ADAM1
    p[i] = pp; m[i] = mm; v[i] = vv;
  }
}

// ---------------------------------------------------------------------------------------
// De-duplication of K news ids in [0, N): sorted unique list, count, inverse map.
// Scratch: ws[0..N) flag→slot table, ws[N..N+nb] per-block counts / offsets.
// ---------------------------------------------------------------------------------------
constexpr int kDedupBlock = 1024;

__global__ void dedup_mark_kernel(const int* __restrict__ ids, int K, int N, int* __restrict__ flag) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < K) {
    const int id = ids[k];
    if (id >= 0 && id < N) flag[id] = 1;
  }
}

__global__ void __launch_bounds__(kDedupBlock) dedup_count_kernel(const int* __restrict__ flag, int N,
                                                                  int* __restrict__ block_counts) {
  __shared__ int wsum[32];
  const int i = blockIdx.x * kDedupBlock + threadIdx.x;
  const int f = (i < N) ? flag[i] : 0;
  const unsigned b = __ballot_sync(0xffffffffu, f != 0);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = __popc(b);
  __syncthreads();
  if (threadIdx.x < 32) {
    int v = wsum[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = v;
  }
}

// single block: exclusive scan of block_counts[0..nb) in place, total → *count
__global__ void __launch_bounds__(1024) dedup_scan_kernel(int* __restrict__ block_counts, int nb, int* __restrict__ count) {
  __shared__ int wsum[32];
  __shared__ int carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = (i < nb) ? block_counts[i] : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
      int w = wsum[threadIdx.x], wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, wi, o);
        if (threadIdx.x >= o) wi += t;
      }
      wsum[threadIdx.x] = wi - w;  // exclusive warp offsets
    }
    __syncthreads();
    const int carry = carry_s;
    const int excl = carry + wsum[threadIdx.x >> 5] + incl - v;
    if (i < nb) block_counts[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = carry_s;
}

__global__ void __launch_bounds__(kDedupBlock) dedup_compact_kernel(int* __restrict__ flag_slot, int N,
                                                                    const int* __restrict__ block_off,
                                                                    int* __restrict__ uniq) {
  __shared__ int wsum[32];
  const int i = blockIdx.x * kDedupBlock + threadIdx.x;
  const int f = (i < N) ? flag_slot[i] : 0;
  const unsigned b = __ballot_sync(0xffffffffu, f != 0);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) wsum[w] = __popc(b);
  __syncthreads();
  if (threadIdx.x < 32) {
    int v = wsum[threadIdx.x], vi = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, vi, o);
      if (threadIdx.x >= o) vi += t;
    }
    wsum[threadIdx.x] = vi - v;
  }
  __syncthreads();
  if (f) {
    const int pos = block_off[blockIdx.x] + wsum[w] + __popc(b & ((1u << lane) - 1u));
    uniq[pos] = i;
    flag_slot[i] = pos;
  }
}

__global__ void dedup_inverse_kernel(const int* __restrict__ ids, int K, int N, const int* __restrict__ slot,
                                     int* __restrict__ inv, const int* __restrict__ count, int* __restrict__ uniq) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < K) {
    const int id = ids[k];
    inv[k] = (id >= 0 && id < N) ? slot[id] : 0;
    if (k >= *count) uniq[k] = 0;  // pad the tail with row 0 (<unk>) so downstream gathers stay in range
  }
}

// sum of squares of a flat fp32 buffer → *out (accumulated; caller zeroes)
__global__ void __launch_bounds__(256) sq_norm_kernel(const float4* __restrict__ x, size_t n4, float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = x[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(out, acc);
}

// x = x * min(1, clip/||x||) + N(0, noise_std^2); sq = sum of squares computed beforehand.
__global__ void __launch_bounds__(256) clip_noise_kernel(float4* __restrict__ x, size_t n4, const float* __restrict__ sq,
                                                         float clip, float noise_std, uint64_t seed, uint32_t stream_id) {
  float scale = 1.f;
  if (clip > 0.f) {
    const float nrm = sqrtf(*sq);
    scale = fminf(1.f, clip / fmaxf(nrm, 1e-30f));
  }
  Philox rng(seed);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 v = x[i];
    v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
    if (noise_std > 0.f) {
      uint32_t r[4];
      rng(i, stream_id, r);
      float z0, z1, z2, z3;
      box_muller(r[0], r[1], z0, z1);
      box_muller(r[2], r[3], z2, z3);
      v.x += noise_std * z0; v.y += noise_std * z1; v.z += noise_std * z2; v.w += noise_std * z3;
    }
    x[i] = v;
  }
}

}  // namespace fr

using namespace fr;

FR_API int fr_score_ce(const float* cand, const float* user, const int* labels, float* score, float* loss_rows,
                       float* loss_mean, float* d_cand, float* d_user, float grad_scale, int B, int C, int D,
                       void* stream) {
  FR_REQUIRE(C >= 1 && C <= 32, "fr_score_ce: C=%d must be in [1,32]", C);
  FR_REQUIRE(B >= 0 && D >= 1, "fr_score_ce: bad shape B=%d D=%d", B, D);
  if (B == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (loss_mean) FR_CHECK(cudaMemsetAsync(loss_mean, 0, sizeof(float), s));
  const int threads = 128, blocks = (B * 32 + threads - 1) / threads;
  if (d_cand != nullptr) {
    FR_REQUIRE(d_user != nullptr, "fr_score_ce: d_user is required with d_cand");
    score_ce_kernel<true><<<blocks, threads, 0, s>>>(cand, user, labels, score, loss_rows, loss_mean, d_cand, d_user,
                                                      grad_scale, B, C, D);
  } else {
    score_ce_kernel<false><<<blocks, threads, 0, s>>>(cand, user, labels, score, loss_rows, loss_mean, nullptr,
                                                       nullptr, grad_scale, B, C, D);
  }
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_segment_sum(const float* vals, const int* seg, float* out, int K, int S, int D, void* stream) {
  FR_REQUIRE(D % 4 == 0, "fr_segment_sum: D=%d must be a multiple of 4", D);
  cudaStream_t s = (cudaStream_t)stream;
  FR_CHECK(cudaMemsetAsync(out, 0, (size_t)S * D * sizeof(float), s));
  if (K == 0) return 0;
  const int threads = 256, blocks = (int)(((size_t)K * 32 + threads - 1) / threads);
  segment_sum_kernel<<<blocks, threads, 0, s>>>(vals, seg, out, K, S, D / 4);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_adam_advance(int* step_dev, float* bc_dev, float b1, float b2, void* stream) {
  adam_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev, bc_dev, b1, b2);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_adam_flat(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                        float eps, float bc1, float bc2, const float* bc_dev, float grad_scale, void* stream) {
  FR_REQUIRE(n % 4 == 0, "fr_adam_flat: n=%lld must be a multiple of 4", n);
  if (n == 0) return 0;
  const size_t n4 = (size_t)n / 4;
  int blocks = (int)((n4 + 255) / 256);
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  adam_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((float4*)p, (const float4*)g, (float4*)m, (float4*)v, n4, lr, b1,
                                                        b2, eps, bc1, bc2, bc_dev, grad_scale);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API long long fr_dedup_workspace_ints(int N) { return (long long)N + (N + kDedupBlock - 1) / kDedupBlock + 1; }

FR_API int fr_dedup(const int* ids, int K, int N, int* workspace, int* uniq, int* count, int* inv, void* stream) {
  FR_REQUIRE(K >= 1 && N >= 1, "fr_dedup: bad sizes K=%d N=%d", K, N);
  cudaStream_t s = (cudaStream_t)stream;
  const int nb = (N + kDedupBlock - 1) / kDedupBlock;
  int* flag = workspace;
  int* block_counts = workspace + N;
  FR_CHECK(cudaMemsetAsync(flag, 0, (size_t)N * sizeof(int), s));
  dedup_mark_kernel<<<(K + 255) / 256, 256, 0, s>>>(ids, K, N, flag);
  dedup_count_kernel<<<nb, kDedupBlock, 0, s>>>(flag, N, block_counts);
  dedup_scan_kernel<<<1, 1024, 0, s>>>(block_counts, nb, count);
  dedup_compact_kernel<<<nb, kDedupBlock, 0, s>>>(flag, N, block_counts, uniq);
  dedup_inverse_kernel<<<(K + 255) / 256, 256, 0, s>>>(ids, K, N, flag, inv, count, uniq);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_sq_norm(const float* x, long long n, float* out, void* stream) {
  FR_REQUIRE(n % 4 == 0, "fr_sq_norm: n=%lld must be a multiple of 4", n);
  cudaStream_t s = (cudaStream_t)stream;
  FR_CHECK(cudaMemsetAsync(out, 0, sizeof(float), s));
  if (n == 0) return 0;
  const size_t n4 = (size_t)n / 4;
  int blocks = (int)((n4 + 255) / 256);
  if (blocks > num_sms() * 4) blocks = num_sms() * 4;
  sq_norm_kernel<<<blocks, 256, 0, s>>>((const float4*)x, n4, out);
  FR_LAUNCH_CHECK();
  return 0;
}

FR_API int fr_clip_noise(float* x, long long n, const float* sq, float clip, float noise_std, unsigned long long seed,
                         unsigned stream_id, void* stream) {
  FR_REQUIRE(n % 4 == 0, "fr_clip_noise: n=%lld must be a multiple of 4", n);
  FR_REQUIRE(clip <= 0.f || sq != nullptr, "fr_clip_noise: clipping needs the squared norm");
  if (n == 0) return 0;
  const size_t n4 = (size_t)n / 4;
  int blocks = (int)((n4 + 255) / 256);
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  clip_noise_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((float4*)x, n4, sq, clip, noise_std, seed, stream_id);
  FR_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// fr_dedup_static: the same de-duplication as fr_dedup for callers that have no memory to hand over.
//
// fr_dedup needs fr_dedup_workspace_ints(N) ints of scratch from its caller. Some callers cannot supply it:
// a C or C++ host that drives this library directly and owns no device allocator, a serving process in which
// another framework's pool has already claimed the whole device-memory budget (a fresh cudaMalloc there can
// fail or fragment the pool), and request loops that forbid cudaMalloc/cudaFree outright because both
// serialise against every stream on the device. For them the scratch lives here, in a file-scope device array
// that every GPU loading this library carries until the process exits:
//   kStaticDedupMaxRows = 1,048,576 table rows  ->  (1,048,576 + 1024 + 1) ints = 4,198,404 bytes per GPU.
// Inputs and outputs are still the caller's. The function allocates nothing and frees nothing.
//
// Set-up: the array's name is only a host-side handle, so the device address is looked up with
// cudaGetSymbolAddress once per device, on the first call made while that device is current; the result is
// cached. Make that first call before starting any stream capture on the device.
// Sharing: every call on a device uses the same array and finds the previous call's contents in it; the flag
// region is cleared at the start of each call and the rest is fully rewritten before it is read. Two calls that
// may overlap on the GPU (different streams, no dependency) race on the array and must be ordered by the caller.
// ---------------------------------------------------------------------------------------------------------
namespace fr {
constexpr int kStaticDedupMaxRows = 1 << 20;
constexpr int kStaticDedupInts = kStaticDedupMaxRows + (kStaticDedupMaxRows + kDedupBlock - 1) / kDedupBlock + 1;
static __device__ __align__(256) int g_static_dedup_scratch[kStaticDedupInts];
constexpr int kMaxDevices = 64;
static int* g_static_dedup_ptr[kMaxDevices] = {nullptr};
}  // namespace fr

FR_API int fr_dedup_static_max_rows() { return kStaticDedupMaxRows; }

FR_API int fr_dedup_static(const int* ids, int K, int N, int* uniq, int* count, int* inv, void* stream) {
  FR_REQUIRE(K >= 1 && N >= 1, "fr_dedup_static: bad sizes K=%d N=%d", K, N);
  FR_REQUIRE(N <= kStaticDedupMaxRows,
             "fr_dedup_static: N=%d table rows exceed the built-in scratch (%d rows); use fr_dedup with caller scratch",
             N, kStaticDedupMaxRows);
  int dev = 0;
  FR_CHECK(cudaGetDevice(&dev));
  FR_REQUIRE(dev >= 0 && dev < kMaxDevices, "fr_dedup_static: device ordinal %d out of range", dev);
  int* ws = __atomic_load_n(&g_static_dedup_ptr[dev], __ATOMIC_ACQUIRE);
  if (ws == nullptr) {
    // first call on this device: resolve the device address of the file-scope array (its host-side name is not one)
    void* p = nullptr;
    FR_CHECK(cudaGetSymbolAddress(&p, g_static_dedup_scratch));
    FR_REQUIRE(p != nullptr, "fr_dedup_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    ws = static_cast<int*>(p);
    __atomic_store_n(&g_static_dedup_ptr[dev], ws, __ATOMIC_RELEASE);
  }
  cudaStream_t s = (cudaStream_t)stream;
  const int nb = (N + kDedupBlock - 1) / kDedupBlock;
  int* flag = ws;
  int* block_counts = ws + N;
  FR_CHECK(cudaMemsetAsync(flag, 0, (size_t)N * sizeof(int), s));     // must start at zero; the array keeps old contents
  dedup_mark_kernel<<<(K + 255) / 256, 256, 0, s>>>(ids, K, N, flag);
  dedup_count_kernel<<<nb, kDedupBlock, 0, s>>>(flag, N, block_counts);
  dedup_scan_kernel<<<1, 1024, 0, s>>>(block_counts, nb, count);
  dedup_compact_kernel<<<nb, kDedupBlock, 0, s>>>(flag, N, block_counts, uniq);
  dedup_inverse_kernel<<<(K + 255) / 256, 256, 0, s>>>(ids, K, N, flag, inv, count, uniq);
  FR_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Host I/O of a training step without copy-engine hops or CUDA API calls per step.
//
// The reference moves every batch and every result through blocking copies inside its batch loop (main.py:59-64,
// 72-79, 90). Here the step's ids wait in a small ring of slots in host-mapped pinned memory; the first kernel of the
// (captured) step pulls this step's slot over PCIe into the step's static id buffer, and the last kernel pushes
// {loss, sequence number} back into a second host-mapped ring that the host polls. Which slot is a function of a
// device-side step counter, so the very same launches replay from a CUDA graph. The host must stay fewer than
// `slots` steps ahead of the device (engine/fused.py enforces that by waiting on the result ring).
// ---------------------------------------------------------------------------------------------------------
namespace fr {
__global__ void __launch_bounds__(1024) hostio_fetch_kernel(int4* __restrict__ dst, const int4* __restrict__ ring, int n4,
                                                             int slots, unsigned long long* step_ctr) {
  const unsigned long long s = *reinterpret_cast<volatile unsigned long long*>(step_ctr);
  const int4* src = ring + (size_t)(s % (unsigned long long)slots) * (size_t)n4;
  // 16-byte reads over PCIe, all in flight at once (4-byte reads of host memory were measured to cost ~25 ns each, 70 us
  // for the 11 KB of a step); host memory is not cached across kernels, so plain loads see what the host wrote
  for (int i = threadIdx.x; i < n4; i += blockDim.x) {
    int4 v;
    asm volatile("ld.global.relaxed.sys.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src + i) : "memory");
    dst[i] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) *step_ctr = s + 1ull;
}
__global__ void hostio_publish_kernel(volatile unsigned* out_ring, const float* __restrict__ value, int slots,
                                      const unsigned long long* step_ctr) {
  const unsigned long long seq = *step_ctr;                    // 1-based number of the step that just finished
  volatile unsigned* slot = out_ring + 4 * (size_t)((seq - 1ull) % (unsigned long long)slots);
  slot[0] = __float_as_uint(*value);
  __threadfence_system();
  slot[1] = (unsigned)seq;
}
}  // namespace fr

FR_API int fr_host_mapped_alloc(long long bytes, void** host_out, void** dev_out) {
  FR_REQUIRE(bytes > 0, "fr_host_mapped_alloc: bytes=%lld", bytes);
  void* h = nullptr;
  FR_CHECK(cudaHostAlloc(&h, (size_t)bytes, cudaHostAllocMapped | cudaHostAllocPortable));
  memset(h, 0, (size_t)bytes);
  void* d = nullptr;
  FR_CHECK(cudaHostGetDevicePointer(&d, h, 0));
  *host_out = h; *dev_out = d;
  return 0;
}
FR_API int fr_host_mapped_free(void* host) { FR_CHECK(cudaFreeHost(host)); return 0; }

FR_API int fr_hostio_fetch(int* dst, const int* ring_mapped, int n, int slots, unsigned long long* step_ctr, void* stream) {
  FR_REQUIRE(n >= 4 && n % 4 == 0 && slots >= 1, "fr_hostio_fetch: n=%d must be a positive multiple of 4 (pad the slots), slots=%d", n, slots);
  FR_REQUIRE(((uintptr_t)dst & 15) == 0 && ((uintptr_t)ring_mapped & 15) == 0, "fr_hostio_fetch: 16-byte alignment");
  hostio_fetch_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>((int4*)dst, (const int4*)ring_mapped, n / 4, slots, step_ctr);
  FR_LAUNCH_CHECK();
  return 0;
}
FR_API int fr_hostio_publish(void* out_ring_mapped, const float* value, int slots, const unsigned long long* step_ctr,
                             void* stream) {
  FR_REQUIRE(slots >= 1, "fr_hostio_publish: slots=%d", slots);
  hostio_publish_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((volatile unsigned*)out_ring_mapped, value, slots, step_ctr);
  FR_LAUNCH_CHECK();
  return 0;
}

// Device-side timestamps (debugging / profiling without a profiler): ring[(*ctr)++ % slots] = %globaltimer (ns).
// engine/fused.py puts one at the start and one at the end of a captured step when FEDREC_STEP_TIMESTAMPS=1, so
// scripts/step_gaps.py can report how long the device sits idle between two graph replays in the real loop.
namespace fr {
__global__ void timestamp_kernel(unsigned long long* ring, int slots, unsigned long long* ctr) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  const unsigned long long i = *ctr;
  ring[i % (unsigned long long)slots] = t;
  *ctr = i + 1ull;
}
}  // namespace fr
FR_API int fr_timestamp(unsigned long long* ring, int slots, unsigned long long* ctr, void* stream) {
  FR_REQUIRE(slots >= 1, "fr_timestamp: slots=%d", slots);
  timestamp_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(ring, slots, ctr);
  FR_LAUNCH_CHECK();
  return 0;
}
