// Item-sharded evaluation, last step: every rank has scored all users of a tile against ITS block
// of item rows and kept a local top-k; the all-gathered lists [shard][user][k] are merged into the
// global top-k in the canonical order (score descending, item id ascending) and AP@k is read off
// the merged ids. One warp per user; entries are disjoint across shards (different item blocks).
//
// Reference: collie/metrics.py mapk :144-198 (one torch.topk over the whole score row there; the
// reference has no sharded form, so the merge itself has no counterpart).
#include "common.cuh"

namespace cmf {

// 64-bit ranking key: larger = ranked first (score descending, item id ascending); 0 = padding
__device__ __forceinline__ uint64_t merge_key(float s, int id) {
  if (id < 0) return 0ull;
  uint32_t u = __float_as_uint(s + 0.0f);            // (-0 -> +0: equal scores, equal keys)
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);   // monotone: larger float -> larger uint
  return ((uint64_t)u << 32) | (uint32_t)(0x7fffffff - id);
}

__global__ void __launch_bounds__(256)
k_merge_topk(const int32_t* __restrict__ ids, const float* __restrict__ scores, int n_shards,
             int64_t n_users, int k, const int64_t* __restrict__ tgt_indptr,
             const int32_t* __restrict__ tgt_items, int32_t* __restrict__ out_ids,
             float* __restrict__ out_scores, double* __restrict__ ap) {
  const int lane = threadIdx.x & 31;
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_users) return;
  const int n_ent = n_shards * k;
  // entry e = (shard e / k, rank e % k)
  auto key_of = [&](int e) {
    const int sh = e / k, j = e - sh * k;
    const int64_t at = ((int64_t)sh * n_users + r) * k + j;
    return merge_key(scores[at], ids[at]);
  };
  constexpr int EPL = 4;          // entries per lane kept in registers (128 per user), else re-read
  const bool in_regs = n_ent <= 32 * EPL;
  uint64_t mine[EPL];
#pragma unroll
  for (int c = 0; c < EPL; ++c) {
    const int e = lane + 32 * c;
    mine[c] = (in_regs && e < n_ent) ? key_of(e) : 0ull;
  }
  uint64_t prev = ~0ull, my_pick = 0;
  for (int round = 0; round < k; ++round) {
    uint64_t best = 0;
    if (in_regs) {
#pragma unroll
      for (int c = 0; c < EPL; ++c)
        if (mine[c] < prev && mine[c] > best) best = mine[c];
    } else {
      for (int e = lane; e < n_ent; e += 32) {
        const uint64_t key = key_of(e);
        if (key < prev && key > best) best = key;
      }
    }
    const uint32_t hi = (uint32_t)(best >> 32);
    const uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t ml = __reduce_max_sync(0xffffffffu, hi == mh ? (uint32_t)best : 0u);
    best = ((uint64_t)mh << 32) | ml;
    if (round == lane) my_pick = best;
    if (best == 0) break;
    prev = best;
  }
  bool hit = false;
  int64_t t0 = 0;
  int n_tgt = 0;
  if (tgt_indptr) {
    t0 = tgt_indptr[r];
    n_tgt = (int)(tgt_indptr[r + 1] - t0);
  }
  if (lane < k) {
    const int id = my_pick ? 0x7fffffff - (int)(uint32_t)my_pick : -1;
    uint32_t u = (uint32_t)(my_pick >> 32);
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    if (out_ids) out_ids[r * k + lane] = id;
    if (out_scores) out_scores[r * k + lane] = my_pick ? __uint_as_float(u) : -INFINITY;
    hit = id >= 0 && n_tgt > 0 && csr_contains(tgt_items + t0, 0, n_tgt, id);
  }
  const uint32_t hits = __ballot_sync(0xffffffffu, hit);
  if (lane == 0 && ap) {
    // sum_j hit_j * precision@j / min(k, #targets), fp64 like the reference's weights
    double acc = 0.0;
    for (int j = 0; j < k; ++j)
      if ((hits >> j) & 1u) acc += (double)__popc(hits & ((2u << j) - 1u)) / (double)(j + 1);
    const int denom = k < n_tgt ? k : n_tgt;
    ap[r] = denom > 0 ? acc / (double)denom : 0.0;
  }
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_merge_topk(const int32_t* ids, const float* scores, int32_t n_shards,
                              int64_t n_users, int32_t k, const int64_t* tgt_indptr,
                              const int32_t* tgt_items, int32_t* out_ids, float* out_scores,
                              double* ap, void* stream) {
  CMF_REQUIRE(ids && scores, "cmf_merge_topk: null candidate lists");
  CMF_REQUIRE(n_shards >= 1 && n_users >= 0 && k >= 1 && k <= 32,
              "cmf_merge_topk: n_shards %d, n_users %lld, k %d (k <= 32)", n_shards,
              (long long)n_users, k);
  CMF_REQUIRE(!ap || (tgt_indptr && tgt_items), "cmf_merge_topk: AP@k needs the targets");
  if (n_users == 0) return CMF_OK;
  k_merge_topk<<<(unsigned)((n_users * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      ids, scores, n_shards, n_users, k, tgt_indptr, tgt_items, out_ids, out_scores, ap);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}
