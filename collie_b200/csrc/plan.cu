// Epoch plan of the data loader, on the device: one pass over the dataset as a random permutation
// cut into batches, every batch ordered by user (so the training step skips its own sort).
//
// Reference: the loaders shuffle an index array on the host and slice it into batches
// (collie/interactions/samplers.py:51-60 via dataloaders.py:70-173); the order inside a batch is
// free there. Here: 32-bit Philox keys per dataset position -> radix sort = the permutation
// (positions with equal keys, ~n^2 / 2^33 pairs, keep their dataset order); then one radix sort of
// (batch number, user) keys over the permuted positions; then one gather.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace cmf {

struct PlanLayout {
  size_t kA, kB, vA, vB, temp, temp_bytes, total;
};

static PlanLayout plan_layout(int64_t n) {
  PlanLayout L;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
  L.kA = take(sizeof(uint64_t) * (size_t)n);
  L.kB = take(sizeof(uint64_t) * (size_t)n);
  L.vA = take(sizeof(uint32_t) * (size_t)n);
  L.vB = take(sizeof(uint32_t) * (size_t)n);
  size_t t32 = 0, t64 = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, t32, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n);
  cub::DeviceRadixSort::SortPairs((void*)nullptr, t64, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n);
  L.temp_bytes = t32 > t64 ? t32 : t64;
  L.temp = take(L.temp_bytes);
  L.total = o;
  return L;
}

// key of dataset position i for this (seed, epoch); a stream of its own (the tag keeps it apart
// from the negative sampler's counters)
__global__ void __launch_bounds__(256)
k_plan_keys(uint64_t seed, uint32_t epoch, int64_t n, uint32_t* __restrict__ keys,
            uint32_t* __restrict__ pos) {
  // four positions per Philox call
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t i0 = q * 4;
  if (i0 >= n) return;
  const uint4 w = philox4x32_10(make_uint4((uint32_t)q, (uint32_t)(q >> 32), epoch, 0x706c616eu),
                                make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (i0 + j < n) { keys[i0 + j] = ws[j]; pos[i0 + j] = (uint32_t)(i0 + j); }
}

__global__ void __launch_bounds__(256)
k_plan_iota(int64_t n, uint32_t* __restrict__ pos) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pos[i] = (uint32_t)i;
}

// (batch number of rank r, user of the position ranked r)
template <typename K>
__global__ void __launch_bounds__(256)
k_plan_batch_user(const uint32_t* __restrict__ perm, const int32_t* __restrict__ users, int64_t n,
                  int64_t batch, int ubits, K* __restrict__ keys) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) keys[r] = ((K)(r / batch) << ubits) | (K)(uint32_t)users[perm[r]];
}

__global__ void __launch_bounds__(256)
k_plan_gather(const uint32_t* __restrict__ perm, const int32_t* __restrict__ users,
              const int32_t* __restrict__ items, const float* __restrict__ ratings,
              const int64_t* __restrict__ gpos, int64_t n, int32_t* __restrict__ out_users,
              int32_t* __restrict__ out_items, float* __restrict__ out_ratings,
              int64_t* __restrict__ out_gidx) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const uint32_t p = perm[r];
  out_users[r] = users[p];
  out_items[r] = items[p];
  if (out_ratings) out_ratings[r] = ratings[p];
  out_gidx[r] = gpos ? gpos[p] : (int64_t)p;
}

// ---- fast path: the dataset arrives ordered by user ------------------------------------------
// A keyed bijection sigma of [0, n) (Feistel network over the bits of n - 1, cycle-walked back into
// range) is the epoch's permutation: entry p goes to batch sigma(p) / batch.
// A STABLE partition of the user-ordered entries by batch number (one radix pass per 8 bits of
// batch number) then leaves every batch ordered by user with no further sort.
struct FeistelKey { uint32_t k[6]; int lbits, rbits; };   // lbits + rbits = bits of n - 1

__device__ __forceinline__ uint32_t fmix32(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
// Unbalanced Feistel network over exactly the bits of n - 1 (the halves swap sizes every round;
// an even number of rounds brings them back), cycle-walked into [0, n): the domain is < 2 n, so
// fewer than two turns on average.
__device__ __forceinline__ uint32_t feistel_perm(uint32_t x, const FeistelKey& fk, uint32_t n) {
  do {
    int lb = fk.lbits, rb = fk.rbits;
    uint32_t l = x >> rb, r = x & ((1u << rb) - 1u);
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const uint32_t t = l ^ (fmix32(r * 0x9e3779b1u + fk.k[i]) & ((1u << lb) - 1u));
      l = r;
      r = t;
      const int tmp = lb; lb = rb; rb = tmp;
    }
    x = (l << rb) | r;
  } while (x >= n);
  return x;
}

__global__ void __launch_bounds__(256)
k_plan_batch_of(FeistelKey fk, int64_t n, uint32_t batch, uint32_t* __restrict__ keys,
                uint32_t* __restrict__ pos) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  keys[p] = feistel_perm((uint32_t)p, fk, (uint32_t)n) / (uint32_t)batch;
  pos[p] = (uint32_t)p;
}

// rec[p] = {user, item, gidx low, gidx high} of the p-th entry in user order
__global__ void __launch_bounds__(256)
k_plan_gather_rec(const uint32_t* __restrict__ perm, const uint4* __restrict__ rec,
                  const float* __restrict__ ratings, int64_t n, int32_t* __restrict__ out_users,
                  int32_t* __restrict__ out_items, float* __restrict__ out_ratings,
                  int64_t* __restrict__ out_gidx) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const uint32_t p = perm[r];
  const uint4 e = __ldg(rec + p);
  out_users[r] = (int32_t)e.x;
  out_items[r] = (int32_t)e.y;
  out_gidx[r] = (int64_t)(((uint64_t)e.w << 32) | e.z);
  if (out_ratings) out_ratings[r] = ratings[p];
}

static int bits_for(int64_t count) {   // bits that hold 0 .. count - 1
  int b = 1;
  while (b < 63 && ((int64_t)1 << b) < count) ++b;
  return b;
}

}  // namespace cmf

using namespace cmf;

extern "C" size_t cmf_epoch_plan_workspace_bytes(int64_t n) {
  if (n <= 0 || n >= ((int64_t)1 << 31)) return 0;
  return plan_layout(n).total;
}

extern "C" int cmf_epoch_plan(const int32_t* users, const int32_t* items, const float* ratings,
                              const int64_t* gpos, int64_t n, int64_t batch, int64_t num_users,
                              int32_t shuffle, int32_t sort_within_batch, uint64_t seed,
                              uint32_t epoch, int32_t* out_users, int32_t* out_items,
                              float* out_ratings, int64_t* out_gidx, void* workspace,
                              size_t workspace_bytes, void* stream) {
  if (n == 0) return CMF_OK;
  if (n < 0 || n >= ((int64_t)1 << 31))
    return fail(CMF_ERR_INVALID, "cmf_epoch_plan: n = %lld is outside [0, 2^31)", (long long)n);
  if (batch <= 0) return fail(CMF_ERR_INVALID, "cmf_epoch_plan: batch must be positive");
  if (!users || !items || !out_users || !out_items || !out_gidx)
    return fail(CMF_ERR_INVALID, "cmf_epoch_plan: null users / items / output pointer");
  if ((out_ratings != nullptr) != (ratings != nullptr))
    return fail(CMF_ERR_INVALID, "cmf_epoch_plan: ratings and out_ratings go together");
  const PlanLayout L = plan_layout(n);
  if (!workspace || workspace_bytes < L.total)
    return fail(CMF_ERR_WORKSPACE, "cmf_epoch_plan: workspace of %zu bytes needed, %zu given",
                L.total, workspace ? workspace_bytes : (size_t)0);
  cudaStream_t s = (cudaStream_t)stream;
  char* base = (char*)workspace;
  uint32_t* vA = (uint32_t*)(base + L.vA);
  uint32_t* vB = (uint32_t*)(base + L.vB);
  const unsigned blocks = (unsigned)((n + 255) / 256);
  size_t temp = L.temp_bytes;
  uint32_t* perm = vA;
  if (shuffle) {
    uint32_t* kA = (uint32_t*)(base + L.kA);
    uint32_t* kB = (uint32_t*)(base + L.kB);
    k_plan_keys<<<(unsigned)(((n + 3) / 4 + 255) / 256), 256, 0, s>>>(seed, epoch, n, kA, vA);
    CMF_LAUNCH_CHECK();
    CMF_CUDA(cub::DeviceRadixSort::SortPairs(base + L.temp, temp, kA, kB, vA, vB, (int)n, 0, 32, s));
    perm = vB;
  } else {
    k_plan_iota<<<blocks, 256, 0, s>>>(n, vA);
    CMF_LAUNCH_CHECK();
  }
  if (sort_within_batch) {
    uint32_t* other = perm == vA ? vB : vA;
    const int ubits = bits_for(num_users > 1 ? num_users : 1);
    const int bbits = bits_for((n + batch - 1) / batch);
    temp = L.temp_bytes;
    if (ubits + bbits <= 32) {
      uint32_t* kA = (uint32_t*)(base + L.kA);
      uint32_t* kB = (uint32_t*)(base + L.kB);
      k_plan_batch_user<uint32_t><<<blocks, 256, 0, s>>>(perm, users, n, batch, ubits, kA);
      CMF_LAUNCH_CHECK();
      CMF_CUDA(cub::DeviceRadixSort::SortPairs(base + L.temp, temp, kA, kB, perm, other, (int)n, 0,
                                               ubits + bbits, s));
    } else {
      uint64_t* kA = (uint64_t*)(base + L.kA);
      uint64_t* kB = (uint64_t*)(base + L.kB);
      k_plan_batch_user<uint64_t><<<blocks, 256, 0, s>>>(perm, users, n, batch, ubits, kA);
      CMF_LAUNCH_CHECK();
      CMF_CUDA(cub::DeviceRadixSort::SortPairs(base + L.temp, temp, kA, kB, perm, other, (int)n, 0,
                                               ubits + bbits, s));
    }
    perm = other;
  }
  k_plan_gather<<<blocks, 256, 0, s>>>(perm, users, items, ratings, gpos, n, out_users, out_items,
                                       out_ratings, out_gidx);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

extern "C" int cmf_epoch_plan_presorted(const void* rec, const float* ratings, int64_t n,
                                        int64_t batch, uint64_t seed, uint32_t epoch,
                                        int32_t* out_users, int32_t* out_items, float* out_ratings,
                                        int64_t* out_gidx, void* workspace, size_t workspace_bytes,
                                        void* stream) {
  if (n == 0) return CMF_OK;
  if (n < 0 || n >= ((int64_t)1 << 31))
    return fail(CMF_ERR_INVALID, "cmf_epoch_plan_presorted: n = %lld is outside [0, 2^31)",
                (long long)n);
  if (batch <= 0) return fail(CMF_ERR_INVALID, "cmf_epoch_plan_presorted: batch must be positive");
  if (!rec || !out_users || !out_items || !out_gidx)
    return fail(CMF_ERR_INVALID, "cmf_epoch_plan_presorted: null record / output pointer");
  if ((reinterpret_cast<uintptr_t>(rec) & 15) != 0)
    return fail(CMF_ERR_INVALID, "cmf_epoch_plan_presorted: records must be 16-byte aligned");
  if ((out_ratings != nullptr) != (ratings != nullptr))
    return fail(CMF_ERR_INVALID, "cmf_epoch_plan_presorted: ratings and out_ratings go together");
  const PlanLayout L = plan_layout(n);
  if (!workspace || workspace_bytes < L.total)
    return fail(CMF_ERR_WORKSPACE,
                "cmf_epoch_plan_presorted: workspace of %zu bytes needed, %zu given", L.total,
                workspace ? workspace_bytes : (size_t)0);
  cudaStream_t s = (cudaStream_t)stream;
  char* base = (char*)workspace;
  uint32_t* kA = (uint32_t*)(base + L.kA);
  uint32_t* kB = (uint32_t*)(base + L.kB);
  uint32_t* vA = (uint32_t*)(base + L.vA);
  uint32_t* vB = (uint32_t*)(base + L.vB);
  // round keys of this (seed, epoch): splitmix64 on the host
  FeistelKey fk;
  uint64_t z = seed * 0x9e3779b97f4a7c15ull + ((uint64_t)epoch << 32 | 0x706c616eu);
  for (int i = 0; i < 6; ++i) {
    z += 0x9e3779b97f4a7c15ull;
    uint64_t x = z;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    fk.k[i] = (uint32_t)((x ^ (x >> 31)) >> 16);
  }
  int bits = bits_for(n);
  if (bits < 2) bits = 2;
  fk.lbits = bits / 2;
  fk.rbits = bits - fk.lbits;
  const unsigned blocks = (unsigned)((n + 255) / 256);
  // (a batch larger than the dataset is one batch)
  k_plan_batch_of<<<blocks, 256, 0, s>>>(fk, n, (uint32_t)(batch < n ? batch : n), kA, vA);
  CMF_LAUNCH_CHECK();
  size_t temp = L.temp_bytes;
  CMF_CUDA(cub::DeviceRadixSort::SortPairs(base + L.temp, temp, kA, kB, vA, vB, (int)n, 0,
                                           bits_for((n + batch - 1) / batch), s));
  k_plan_gather_rec<<<blocks, 256, 0, s>>>(vB, (const uint4*)rec, ratings, n, out_users, out_items,
                                           out_ratings, out_gidx);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}
