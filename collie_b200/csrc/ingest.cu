// Dataset ingest on the device: de-duplication of (user, item) pairs and the positives CSR in one
// pass over a radix-sorted key list.
//
// Reference: collie/interactions/datasets.py:81-155 pushes the COO entries through a dict
// (dok_matrix._update + tocoo): a duplicated pair stays at the position of its FIRST occurrence
// and takes the value of its LAST one; :359-366 builds set((row, col)) of the positives for the
// exact negative sampler (here: CSR, sorted item ids per user, csr_contains in common.cuh).
//
//   keys  user * num_items + item (64 bit), values = dataset position
//   stable radix sort by key   -> equal pairs adjacent, in dataset order
//   run heads / tails          -> first and last position of every distinct pair; the distinct
//                                 keys in sorted order ARE the CSR (user-major, items ascending)
//   radix sort of the runs by first position -> the kept entries in dataset order
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "common.cuh"

namespace cmf {

__global__ void k_ingest_keys(const int32_t* __restrict__ users, const int32_t* __restrict__ items,
                              int64_t n, int64_t num_items, uint64_t* __restrict__ keys,
                              uint32_t* __restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = (uint64_t)users[i] * (uint64_t)num_items + (uint64_t)items[i];
  vals[i] = (uint32_t)i;
}

__global__ void k_ingest_heads(const uint64_t* __restrict__ keys, int64_t n,
                               int32_t* __restrict__ head) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

// run r of entry i = (heads up to and including i) - 1; at a head: first position and key, at a
// tail: last position. The last thread writes the number of runs.
__global__ void k_ingest_runs(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                              const int32_t* __restrict__ head, const int32_t* __restrict__ run_of,
                              int64_t n, int64_t num_items, uint32_t* __restrict__ first,
                              uint32_t* __restrict__ last, uint64_t* __restrict__ run_key,
                              int32_t* __restrict__ csr_items, int64_t* __restrict__ n_out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int r = run_of[i] + head[i] - 1;   // (exclusive scan: heads BEFORE i)
  if (head[i]) {
    first[r] = vals[i];
    run_key[r] = keys[i];
    if (csr_items) csr_items[r] = (int32_t)(keys[i] % (uint64_t)num_items);
  }
  if (i == n - 1 || keys[i + 1] != keys[i]) last[r] = vals[i];
  if (i == n - 1) *n_out = (int64_t)r + 1;
}

// indptr[u] = number of distinct pairs whose user is < u
__global__ void k_ingest_indptr(const uint64_t* __restrict__ run_key,
                                const int64_t* __restrict__ n_runs, int64_t num_users,
                                int64_t num_items, int64_t* __restrict__ indptr) {
  const int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u > num_users) return;
  const uint64_t want = (uint64_t)u * (uint64_t)num_items;
  int64_t lo = 0, hi = *n_runs;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (run_key[mid] < want) lo = mid + 1; else hi = mid;
  }
  indptr[u] = lo;
}

struct IngestLayout {
  size_t keys, keys2, vals, vals2, head, run_of, first, last, first2, run_key, temp, temp_bytes, total;
};

static IngestLayout ingest_layout(int64_t n) {
  IngestLayout L;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
  L.keys = take(sizeof(uint64_t) * n);
  L.keys2 = take(sizeof(uint64_t) * n);
  L.vals = take(sizeof(uint32_t) * n);
  L.vals2 = take(sizeof(uint32_t) * n);
  L.head = take(sizeof(int32_t) * n);
  L.run_of = take(sizeof(int32_t) * n);
  L.first = take(sizeof(uint32_t) * n);
  L.last = take(sizeof(uint32_t) * n);
  L.first2 = take(sizeof(uint32_t) * n);
  L.run_key = take(sizeof(uint64_t) * n);
  size_t t64 = 0, t32 = 0, tscan = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, t64, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n);
  cub::DeviceRadixSort::SortPairs((void*)nullptr, t32, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n);
  cub::DeviceScan::ExclusiveSum((void*)nullptr, tscan, (const int32_t*)nullptr, (int32_t*)nullptr,
                                (int)n);
  L.temp_bytes = t64 > t32 ? t64 : t32;
  if (tscan > L.temp_bytes) L.temp_bytes = tscan;
  L.temp = take(L.temp_bytes);
  L.total = o;
  return L;
}

}  // namespace cmf

using namespace cmf;

extern "C" size_t cmf_ingest_workspace_bytes(int64_t n) {
  if (n < 1 || n >= ((int64_t)1 << 31)) return 0;
  return ingest_layout(n).total;
}

extern "C" int cmf_ingest_pairs(const int32_t* users, const int32_t* items, int64_t n,
                                int64_t num_users, int64_t num_items, int32_t* keep_pos,
                                int32_t* keep_val, int64_t* n_out, int64_t* indptr,
                                int32_t* csr_items, void* workspace, size_t workspace_bytes,
                                void* stream) {
  CMF_REQUIRE(users && items && keep_pos && keep_val && n_out, "cmf_ingest_pairs: null pointer");
  CMF_REQUIRE(n >= 1 && n < ((int64_t)1 << 31), "cmf_ingest_pairs: n = %lld (1 .. 2^31 - 1)",
              (long long)n);
  CMF_REQUIRE(num_users >= 1 && num_items >= 1, "cmf_ingest_pairs: bad table shape");
  CMF_REQUIRE((indptr == nullptr) == (csr_items == nullptr),
              "cmf_ingest_pairs: indptr and csr_items go together");
  const IngestLayout L = ingest_layout(n);
  if (!workspace || workspace_bytes < L.total)
    return fail(CMF_ERR_WORKSPACE, "cmf_ingest_pairs: workspace of %zu bytes needed, %zu given",
                L.total, workspace ? workspace_bytes : (size_t)0);
  cudaStream_t s = (cudaStream_t)stream;
  char* base = (char*)workspace;
  uint64_t* keys = (uint64_t*)(base + L.keys);
  uint64_t* keys2 = (uint64_t*)(base + L.keys2);
  uint32_t* vals = (uint32_t*)(base + L.vals);
  uint32_t* vals2 = (uint32_t*)(base + L.vals2);
  int32_t* head = (int32_t*)(base + L.head);
  int32_t* run_of = (int32_t*)(base + L.run_of);
  uint32_t* first = (uint32_t*)(base + L.first);
  uint32_t* last = (uint32_t*)(base + L.last);
  uint32_t* first2 = (uint32_t*)(base + L.first2);
  uint64_t* run_key = (uint64_t*)(base + L.run_key);
  const unsigned grid = (unsigned)((n + 255) / 256);

  k_ingest_keys<<<grid, 256, 0, s>>>(users, items, n, num_items, keys, vals);
  CMF_LAUNCH_CHECK();
  int bits = 1;
  const unsigned __int128 span = (unsigned __int128)num_users * (unsigned __int128)num_items;
  while (bits < 64 && ((unsigned __int128)1 << bits) < span) ++bits;
  size_t temp = L.temp_bytes;
  CMF_CUDA(cub::DeviceRadixSort::SortPairs(base + L.temp, temp, keys, keys2, vals, vals2, (int)n, 0,
                                           bits, s));
  k_ingest_heads<<<grid, 256, 0, s>>>(keys2, n, head);
  CMF_LAUNCH_CHECK();
  temp = L.temp_bytes;
  CMF_CUDA(cub::DeviceScan::ExclusiveSum(base + L.temp, temp, head, run_of, (int)n, s));
  // (runs that do not exist sort behind every real one)
  CMF_CUDA(cudaMemsetAsync(first, 0xFF, sizeof(uint32_t) * n, s));
  k_ingest_runs<<<grid, 256, 0, s>>>(keys2, vals2, head, run_of, n, num_items, first, last, run_key,
                                     csr_items, n_out);
  CMF_LAUNCH_CHECK();
  if (indptr) {
    k_ingest_indptr<<<(unsigned)((num_users + 1 + 255) / 256), 256, 0, s>>>(run_key, n_out, num_users,
                                                                         num_items, indptr);
    CMF_LAUNCH_CHECK();
  }
  // the kept entries in dataset order: runs sorted by their first position
  temp = L.temp_bytes;
  CMF_CUDA(cub::DeviceRadixSort::SortPairs(base + L.temp, temp, first, first2, last,
                                           (uint32_t*)keep_val, (int)n, 0, 32, s));
  CMF_CUDA(cudaMemcpyAsync(keep_pos, first2, sizeof(uint32_t) * n, cudaMemcpyDeviceToDevice, s));
  return CMF_OK;
}
