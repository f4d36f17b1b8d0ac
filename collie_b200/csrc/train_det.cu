// Deterministic (bit-repeatable) form of the training step's backward pass, on request
// (cmf_step_t.deterministic). The fast step adds item-row gradient contributions with vector RED
// in arrival order and lets users that straddle two chunks accumulate through RED as well; fp32
// sums then depend on the schedule. Here:
//   item side  (item, contribution) pairs of the batch -> stable radix sort by item -> one group
//              of lanes per item row sums its contributions IN BATCH ORDER and stores the
//              gradient row (no atomics); k_update_rows applies it as usual;
//   user side  k_bwd_update runs over chunks cut at user boundaries (k_aligned_chunks), so every
//              user is summed by one warp in batch order and finished in place.
// The reference's own CPU path is deterministic (index_add on one thread); torch's CUDA path is not.
// Reference semantics: torch.optim.SparseAdam / Adam on the coalesced gradient,
// collie/model/base/base_pipeline.py:418-433.
#include <cub/device/device_radix_sort.cuh>

#include "train_common.cuh"

namespace cmf {

// chunk c starts at the first user boundary at or after c * L
__global__ void k_aligned_chunks(const int32_t* __restrict__ su, int64_t B, int L,
                                 int64_t num_chunks, int32_t* __restrict__ starts) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c > num_chunks) return;
  int64_t e = c * L;
  if (e > B) e = B;
  while (e > 0 && e < B && su[e] == su[e - 1]) ++e;
  starts[c] = (int32_t)e;
}

// two pairs per interaction: (positive item, 2b) and (selected negative, 2b + 1); explicit
// losses have the first only. Inactive contributions (g == 0) get the key num_items (sorted last).
__global__ void k_det_pairs(const int32_t* __restrict__ items, const int32_t* __restrict__ sel,
                            const float* __restrict__ g, int64_t B, int implicit,
                            uint32_t num_items, uint32_t* __restrict__ keys,
                            uint32_t* __restrict__ vals) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const bool active = g[b] != 0.f;
  if (implicit) {
    keys[2 * b] = active ? (uint32_t)items[b] : num_items;
    vals[2 * b] = (uint32_t)(2 * b);
    keys[2 * b + 1] = (active && sel[b] >= 0) ? (uint32_t)sel[b] : num_items;
    vals[2 * b + 1] = (uint32_t)(2 * b + 1);
  } else {
    keys[b] = active ? (uint32_t)items[b] : num_items;
    vals[b] = (uint32_t)(2 * b);
  }
}

// starts[i] = first position of key i in the sorted keys (i = 0 .. num_items)
__global__ void k_det_item_starts(const uint32_t* __restrict__ keys, int64_t n,
                                  int64_t num_items, int32_t* __restrict__ starts) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > num_items) return;
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (keys[mid] < (uint32_t)i) lo = mid + 1; else hi = mid;
  }
  starts[i] = (int32_t)lo;
}

// One group of G lanes per item row: gradient row = sum over its contributions, in the order of
// the sorted pairs (= batch order), of round(+-g * user row); bias gradient likewise. Four
// contributions' rows are requested at once, added one after the other.
template <int V, int G, int R>
__global__ void __launch_bounds__(kThreads)
k_det_item_reduce(const int32_t* __restrict__ starts, const uint32_t* __restrict__ vals,
                  const int32_t* __restrict__ users, const float* __restrict__ g_arr,
                  const float* __restrict__ U, int D, int64_t num_items, float* __restrict__ GI,
                  float* __restrict__ gbi) {
  constexpr int kGroups = kThreads / G;
  const int l = threadIdx.x & (G - 1);
  const int64_t group = (int64_t)blockIdx.x * kGroups + threadIdx.x / G;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  for (int64_t i = group; i < num_items; i += n_groups) {
    const int j0 = starts[i], j1 = starts[i + 1];
    if (j0 >= j1) continue;
    Row<V, G, R> acc;
    acc.zero();
    float gb = 0.f;
    for (int j = j0; j < j1; j += 4) {
      Row<V, G, R> x[4];
      float gg[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        gg[q] = 0.f;
        x[q].zero();
        if (j + q < j1) {
          const uint32_t v = vals[j + q];
          const int64_t b = (int64_t)(v >> 1);
          const float g = g_arr[b];
          gg[q] = (v & 1u) ? -g : g;
          x[q].load(U + (int64_t)users[b] * D, D, l);
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (j + q < j1) {
          // products rounded before the add, as in k_bwd_update (+g*x and -g*x cancel exactly)
#pragma unroll
          for (int e = 0; e < R * V; ++e) acc.f[e] = __fadd_rn(acc.f[e], __fmul_rn(gg[q], x[q].f[e]));
          gb += gg[q];
        }
      }
    }
    acc.store(GI + i * D, D, l);
    if (l == 0) gbi[i] = gb;
  }
}

size_t det_sort_bytes(int64_t n_pairs) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, bytes, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (int)n_pairs, 0, 32, (cudaStream_t)0);
  return bytes;
}

int launch_aligned_chunks(const StepDev& p, int32_t* starts, cudaStream_t s) {
  const int64_t n = p.num_chunks + 1;
  k_aligned_chunks<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(p.su, p.B, p.L, p.num_chunks, starts);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

// item-row and item-bias gradients of the batch into GI / gbi (zero before the call), in a fixed order
int launch_det_item_side(const StepDev& p, int64_t num_item_rows, uint32_t* keys, uint32_t* keys2,
                         uint32_t* vals, uint32_t* vals2, int32_t* starts, void* temp,
                         size_t temp_bytes, cudaStream_t s) {
  const bool implicit = p.loss <= CMF_LOSS_WARP;
  const int64_t n = implicit ? 2 * p.B : p.B;
  k_det_pairs<<<(unsigned)((p.B + 255) / 256), 256, 0, s>>>(
      p.items, p.sel_arr, p.g_arr, p.B, implicit ? 1 : 0, (uint32_t)num_item_rows, keys, vals);
  CMF_LAUNCH_CHECK();
  int bits = 1;
  while (bits < 32 && ((int64_t)1 << bits) <= num_item_rows) ++bits;
  CMF_CUDA(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys, keys2, vals, vals2, (int)n, 0,
                                           bits, s));
  k_det_item_starts<<<(unsigned)((num_item_rows + 1 + 255) / 256), 256, 0, s>>>(
      keys2, n, num_item_rows, starts);
  CMF_LAUNCH_CHECK();
  const RowGeom geom = row_geom(p.D);
  const int groups_per_block = kThreads / geom.G;
  int64_t blocks = (num_item_rows + groups_per_block - 1) / groups_per_block;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
#define CALL(V_, G_, R_)                                                                  \
  k_det_item_reduce<V_, G_, R_><<<(unsigned)blocks, kThreads, 0, s>>>(                   \
      starts, vals2, p.users, p.g_arr, p.U, p.D, num_item_rows, p.GI, p.gbi)
  CMF_DISPATCH_GEOM(geom, CALL);
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

}  // namespace cmf
