// Stand-alone negative sampler (the fused step draws the same stream in-kernel).
// Reference: Interactions._negative_sample, collie/interactions/datasets.py:390-445.
#include "common.cuh"

namespace cmf {

__global__ void __launch_bounds__(256)
k_sample_negatives(uint64_t seed, uint32_t epoch, const int64_t* __restrict__ gidx, int64_t gidx_base,
                   const int32_t* __restrict__ users, int64_t B, int N, uint32_t num_items,
                   const int64_t* __restrict__ indptr, const int32_t* __restrict__ csr_items,
                   int max_tries, int64_t* __restrict__ out) {
  if (indptr == nullptr) {
    // approximate: one thread per (interaction, slot)   (datasets.py:433-443)
    const int64_t total = B * N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
      const int64_t b = i / N;
      const int n = (int)(i - b * N);
      const int64_t gi = gidx ? gidx[b] : gidx_base + b;
      out[i] = draw_item(seed, epoch, gi, n, num_items);
    }
    return;
  }
  // exact: one thread per interaction walks the rejection loop (datasets.py:399-426)
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < B;
       b += (int64_t)gridDim.x * blockDim.x) {
    const int64_t gi = gidx ? gidx[b] : gidx_base + b;
    const int u = users[b];
    const int64_t lo = indptr[u], hi = indptr[u + 1];
    int64_t* row = out + b * N;
    uint32_t d = 0;
    int taken = 0, checked = 0;
    while (taken < N) {
      int cand = draw_item(seed, epoch, gi, d++, num_items);
      for (;;) {
        bool clash = csr_contains(csr_items, lo, hi, cand);
        for (int j = 0; j < taken && !clash; ++j) clash = (row[j] == cand);
        if (!clash) break;
        if (checked >= max_tries) {
          const int left = N - taken - 1;
          for (int j = 0; j < left; ++j) row[taken++] = draw_item(seed, epoch, gi, d++, num_items);
          break;
        }
        cand = draw_item(seed, epoch, gi, d++, num_items);
        ++checked;
      }
      row[taken++] = cand;
    }
  }
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_sample_negatives(uint64_t seed, uint32_t epoch, const int64_t* gidx,
                                    int64_t gidx_base, const int32_t* users, int64_t B, int32_t N,
                                    int64_t num_items, const int64_t* csr_indptr,
                                    const int32_t* csr_items, int32_t max_tries, int64_t* negs_out,
                                    void* stream) {
  CMF_REQUIRE(B >= 0 && N >= 1, "cmf_sample_negatives: bad B/N");
  CMF_REQUIRE(num_items >= 1 && num_items < (1ll << 31), "cmf_sample_negatives: bad num_items");
  if (B == 0) return CMF_OK;
  CMF_REQUIRE(negs_out, "cmf_sample_negatives: null output");
  CMF_REQUIRE(csr_indptr == nullptr || (csr_items && users),
              "cmf_sample_negatives: exact sampling needs users and csr_items");
  const int64_t work = csr_indptr ? B : B * N;
  int64_t blocks = (work + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  k_sample_negatives<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      seed, epoch, gidx, gidx_base, users, B, N, (uint32_t)num_items, csr_indptr, csr_items,
      max_tries < 0 ? 0 : max_tries, negs_out);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}
