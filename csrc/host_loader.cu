// Host-side batch producer: a client's impression shard → packed id batches, on a background thread.
//
// What it replaces: upstream's DataLoader over TrainDataset (dataset.py:69-86, newsample :10-14; main.py:55-64): per
// impression the clicked news, npratio sampled negatives (padded with <unk> when there are fewer) and the 0-padded click
// history. Python produces a batch of 50 in 130 us with vectorised numpy (data/dataset.py:BatchSampler); here a C++
// thread keeps a small queue of ready batches in the packed layout the training step reads (histories [B, L] first, then
// candidates [B, 1 + K], int32 news rows), so the launch loop only memcpy's 11 KB into the step's host-mapped ring.
//
// No CUDA in this file: it is host code living in the same library, callable (and tested) on a CPU-only box.
#include "common.cuh"

#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

namespace {

struct Rng {                                         // xoshiro256** seeded through splitmix64
  uint64_t s[4];
  explicit Rng(uint64_t seed) {
    for (int i = 0; i < 4; ++i) {
      seed += 0x9E3779B97F4A7C15ull;
      uint64_t z = seed;
      z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
      z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
      s[i] = z ^ (z >> 31);
    }
  }
  static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
  uint64_t next() {
    const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return r;
  }
  uint64_t below(uint64_t n) {                       // unbiased integer in [0, n)
    if (n <= 1) return 0;
    const uint64_t limit = UINT64_MAX - UINT64_MAX % n;
    uint64_t r;
    do { r = next(); } while (r >= limit);
    return r % n;
  }
};

struct Batch {
  std::vector<int32_t> ids;
  long long epoch;
};

struct Loader {
  std::vector<int32_t> pos, his, neg_idx;
  std::vector<int64_t> neg_ptr;
  long long S = 0;
  int L = 0, B = 0, K = 0, depth = 0;
  Rng rng;
  std::vector<int64_t> order;
  size_t cursor = 0;
  long long epoch = 0;
  std::mutex mu;
  std::condition_variable not_empty, not_full;
  std::deque<Batch> queue;
  std::vector<std::vector<int32_t>> spare;           // buffers handed back by the consumer
  bool stop = false;
  std::thread worker;

  explicit Loader(uint64_t seed) : rng(seed) {}

  int64_t next_index() {
    if (cursor >= order.size()) {                    // a fresh permutation every pass over the shard (Fisher-Yates)
      order.resize((size_t)S);
      for (long long i = 0; i < S; ++i) order[(size_t)i] = i;
      for (long long i = S - 1; i > 0; --i) std::swap(order[(size_t)i], order[(size_t)rng.below((uint64_t)i + 1)]);
      cursor = 0;
      ++epoch;
    }
    return order[cursor++];
  }

  void fill(Batch& b) {
    const int C = 1 + K;
    b.ids.resize((size_t)B * (L + C));                // histories are copied whole; candidate rows are zeroed below
    int32_t* h_out = b.ids.data();
    int32_t* c_out = b.ids.data() + (size_t)B * L;
    int64_t pick[64];
    for (int j = 0; j < B; ++j) {
      const int64_t s = next_index();
      std::memcpy(h_out + (size_t)j * L, his.data() + (size_t)s * L, sizeof(int32_t) * (size_t)L);
      int32_t* c = c_out + (size_t)j * C;
      c[0] = pos[(size_t)s];
      for (int k = 1; k < C; ++k) c[k] = 0;
      const int64_t lo = neg_ptr[(size_t)s], n = neg_ptr[(size_t)s + 1] - lo;
      if (n >= K) {
        // Floyd: K distinct values of [0, n); then a shuffle, because Floyd's output order is not uniform
        for (int k = 0; k < K; ++k) {
          const int64_t top = n - K + k;
          int64_t t = (int64_t)rng.below((uint64_t)top + 1);
          for (int q = 0; q < k; ++q) if (pick[q] == t) { t = top; break; }
          pick[k] = t;
        }
        for (int k = K - 1; k > 0; --k) std::swap(pick[k], pick[rng.below((uint64_t)k + 1)]);
        for (int k = 0; k < K; ++k) c[1 + k] = neg_idx[(size_t)(lo + pick[k])];
      } else {
        for (int64_t k = 0; k < n; ++k) c[1 + k] = neg_idx[(size_t)(lo + k)];   // all of them, then <unk> = 0
      }
    }
    b.epoch = epoch;
  }

  void run() {
    for (;;) {
      Batch b;
      {
        std::lock_guard<std::mutex> lk(mu);
        if (!spare.empty()) { b.ids = std::move(spare.back()); spare.pop_back(); }
      }
      fill(b);
      std::unique_lock<std::mutex> lk(mu);
      not_full.wait(lk, [&] { return stop || (int)queue.size() < depth; });
      if (stop) return;
      queue.push_back(std::move(b));
      not_empty.notify_one();
    }
  }
};

}  // namespace

// pos [S], his [S, L] (0-padded), neg_ptr [S + 1], neg_idx [neg_ptr[S]]: the arrays of data/dataset.py:ImpressionTable
// (copied; the caller may free them). B impressions per batch, K sampled negatives each, `depth` batches prepared ahead.
FR_API void* fr_loader_create(const int32_t* pos, const int32_t* his, const int64_t* neg_ptr, const int32_t* neg_idx,
                              long long S, int L, int B, int K, unsigned long long seed, int depth) {
  if (S < 1 || L < 1 || B < 1 || K < 0 || K > 64 || depth < 1 || !pos || !his || !neg_ptr) {
    fr::set_error("fr_loader_create: need S, L, B >= 1, 0 <= K <= 64, depth >= 1 and the table arrays");
    return nullptr;
  }
  Loader* ld = new Loader(seed);
  ld->S = S; ld->L = L; ld->B = B; ld->K = K; ld->depth = depth;
  ld->pos.assign(pos, pos + S);
  ld->his.assign(his, his + (size_t)S * L);
  ld->neg_ptr.assign(neg_ptr, neg_ptr + S + 1);
  const int64_t nn = neg_ptr[S];
  if (nn > 0 && neg_idx) ld->neg_idx.assign(neg_idx, neg_idx + nn);
  for (long long i = 0; i < S; ++i)
    if (neg_ptr[i + 1] < neg_ptr[i] || neg_ptr[i + 1] > nn) {
      fr::set_error("fr_loader_create: neg_ptr is not a monotone CSR pointer array");
      delete ld;
      return nullptr;
    }
  ld->worker = std::thread([ld] { ld->run(); });
  return ld;
}

// Blocks until a batch is ready, copies its B*L + B*(1+K) ids into `out` (histories first) and reports the number of the
// pass over the shard (1-based) its LAST impression came from.
FR_API int fr_loader_next(void* handle, int32_t* out, long long* epoch_out) {
  FR_REQUIRE(handle != nullptr && out != nullptr, "fr_loader_next: null handle or output");
  Loader* ld = static_cast<Loader*>(handle);
  Batch b;
  {
    std::unique_lock<std::mutex> lk(ld->mu);
    ld->not_empty.wait(lk, [&] { return !ld->queue.empty(); });
    b = std::move(ld->queue.front());
    ld->queue.pop_front();
    ld->not_full.notify_one();
  }
  std::memcpy(out, b.ids.data(), sizeof(int32_t) * b.ids.size());
  if (epoch_out) *epoch_out = b.epoch;
  {
    std::lock_guard<std::mutex> lk(ld->mu);
    if ((int)ld->spare.size() < ld->depth + 2) ld->spare.push_back(std::move(b.ids));
  }
  return 0;
}

FR_API long long fr_loader_batch_ints(void* handle) {
  Loader* ld = static_cast<Loader*>(handle);
  return ld ? (long long)ld->B * (ld->L + 1 + ld->K) : 0;
}

FR_API void fr_loader_destroy(void* handle) {
  Loader* ld = static_cast<Loader*>(handle);
  if (!ld) return;
  {
    std::lock_guard<std::mutex> lk(ld->mu);
    ld->stop = true;
  }
  ld->not_full.notify_all();
  if (ld->worker.joinable()) ld->worker.join();
  delete ld;
}
