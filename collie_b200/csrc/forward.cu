// Fused embedding-gather + dot + bias forward and its backward.
// Reference: collie/model/matrix_factorization.py:120-159 (forward) and autograd through it.
#include "common.cuh"

namespace cmf {

template <typename IdxT, int V, int G, int R>
__global__ void __launch_bounds__(256)
k_forward_scores(const float* __restrict__ U, const float* __restrict__ I,
                 const float* __restrict__ bu, const float* __restrict__ bi, int D,
                 const IdxT* __restrict__ users, const IdxT* __restrict__ items, int64_t M,
                 int has_range, float y_lo, float y_hi, float* __restrict__ out) {
  const int lane_w = threadIdx.x & 31;
  const int l = threadIdx.x & (G - 1);
  const unsigned mask = group_mask<G>(lane_w);
  const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t n_groups = ((int64_t)gridDim.x * blockDim.x) / G;
  // two pairs per iteration keep 4 row loads in flight per lane
  for (int64_t m = group * 2; m < M; m += n_groups * 2) {
    const bool two = (m + 1) < M;
    const int64_t u0 = users[m], i0 = items[m];
    const int64_t u1 = two ? (int64_t)users[m + 1] : u0, i1 = two ? (int64_t)items[m + 1] : i0;
    Row<V, G, R> a0, b0, a1, b1;
    a0.load(U + u0 * D, D, l);
    b0.load(I + i0 * D, D, l);
    a1.load(U + u1 * D, D, l);
    b1.load(I + i1 * D, D, l);
    float s0 = group_sum<G>(a0.dot(b0), mask);
    float s1 = group_sum<G>(a1.dot(b1), mask);
    if (l == 0) {
      s0 = s0 + bu[u0] + bi[i0];
      if (has_range) s0 = sigmoidf_(s0) * (y_hi - y_lo) + y_lo;
      out[m] = s0;
      if (two) {
        s1 = s1 + bu[u1] + bi[i1];
        if (has_range) s1 = sigmoidf_(s1) * (y_hi - y_lo) + y_lo;
        out[m + 1] = s1;
      }
    }
  }
}

template <typename IdxT, int V, int G, int R>
__global__ void __launch_bounds__(256)
k_backward_scores(const float* __restrict__ U, const float* __restrict__ I, int D,
                  const IdxT* __restrict__ users, const IdxT* __restrict__ items, int64_t M,
                  const float* __restrict__ dpreds, const float* __restrict__ preds, int has_range,
                  float y_lo, float y_hi, float* __restrict__ gU, float* __restrict__ gI,
                  float* __restrict__ gbu, float* __restrict__ gbi) {
  const int l = threadIdx.x & (G - 1);
  const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t n_groups = ((int64_t)gridDim.x * blockDim.x) / G;
  for (int64_t m = group; m < M; m += n_groups) {
    const int64_t u = users[m], i = items[m];
    float c = dpreds[m];
    if (has_range) {
      // preds = sigmoid(s) * (hi - lo) + lo  ->  d preds / d s = sig * (1 - sig) * (hi - lo)
      const float sg = (preds[m] - y_lo) / (y_hi - y_lo);
      c *= sg * (1.f - sg) * (y_hi - y_lo);
    }
    Row<V, G, R> a, b, ga, gb;
    a.load(U + u * D, D, l);
    b.load(I + i * D, D, l);
    ga.zero();
    gb.zero();
    ga.axpy(c, b);
    gb.axpy(c, a);
    ga.red_add(gU + u * D, D, l);
    gb.red_add(gI + i * D, D, l);
    if (l == 0) {
      atomicAdd(gbu + u, c);
      atomicAdd(gbi + i, c);
    }
  }
}

static int grid_for(int64_t groups_needed, int G) {
  const int groups_per_block = 256 / G;
  int64_t blocks = (groups_needed + groups_per_block - 1) / groups_per_block;
  const int64_t cap = (int64_t)num_sms() * 8;  // 8 x 256-thread CTAs = full occupancy per SM
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_forward_scores(const float* U, const float* I, const float* bu, const float* bi,
                                  int32_t D, const void* users, const void* items, int32_t idx_bytes,
                                  int64_t M, const float* host_y_range, float* out, void* stream) {
  CMF_REQUIRE(idx_bytes == 4 || idx_bytes == 8, "cmf_forward_scores: idx_bytes must be 4 or 8");
  CMF_REQUIRE(geom_supported(D), "cmf_forward_scores: embedding_dim %d unsupported (1..1024)", D);
  CMF_REQUIRE(M >= 0, "cmf_forward_scores: negative M");
  CMF_REQUIRE(U && I && bu && bi, "cmf_forward_scores: null table pointer");
  if (M == 0) return CMF_OK;
  CMF_REQUIRE(users && items && out, "cmf_forward_scores: null id or output pointer");
  const RowGeom g = row_geom(D);
  const int has_range = host_y_range != nullptr;
  const float lo = has_range ? host_y_range[0] : 0.f, hi = has_range ? host_y_range[1] : 1.f;
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = grid_for((M + 1) / 2, g.G);
#define CALL(V_, G_, R_)                                                                         \
  if (idx_bytes == 8)                                                                            \
    k_forward_scores<int64_t, V_, G_, R_><<<grid, 256, 0, s>>>(                                  \
        U, I, bu, bi, D, (const int64_t*)users, (const int64_t*)items, M, has_range, lo, hi, out); \
  else                                                                                           \
    k_forward_scores<int32_t, V_, G_, R_><<<grid, 256, 0, s>>>(                                  \
        U, I, bu, bi, D, (const int32_t*)users, (const int32_t*)items, M, has_range, lo, hi, out)
  CMF_DISPATCH_GEOM(g, CALL);
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

// ---- cosine similarity of one row against every row of a table
// (BasePipeline._calculate_embedding_similarity, collie/model/base/base_pipeline.py:849-859:
//  rows are divided by their L2 norm, then dotted with the normalised seed row)
template <int V, int G, int R>
__global__ void __launch_bounds__(256)
k_row_similarity(const float* __restrict__ E, int64_t rows, int D, int64_t id,
                 float* __restrict__ out) {
  constexpr int kGroups = 256 / G;
  const int lane_w = threadIdx.x & 31;
  const int l = threadIdx.x & (G - 1);
  const unsigned mask = group_mask<G>(lane_w);
  const int64_t group = (int64_t)blockIdx.x * kGroups + threadIdx.x / G;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  Row<V, G, R> seed;
  seed.load(E + id * D, D, l);
  float sn = seed.dot(seed);
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) sn += __shfl_xor_sync(mask, sn, o);
  sn = sqrtf(sn);
#pragma unroll
  for (int i = 0; i < R * V; ++i) seed.f[i] = seed.f[i] / sn;
  for (int64_t r = group; r < rows; r += n_groups) {
    Row<V, G, R> x;
    x.load(E + r * D, D, l);
    float n = x.dot(x);
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) n += __shfl_xor_sync(mask, n, o);
    n = sqrtf(n);
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < R * V; ++i) acc = fmaf(x.f[i] / n, seed.f[i], acc);
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(mask, acc, o);
    if (l == 0) out[r] = acc;
  }
}

extern "C" int cmf_row_similarity(const float* E, int64_t rows, int32_t D, int64_t id, float* out,
                                  void* stream) {
  CMF_REQUIRE(E && out, "cmf_row_similarity: null pointer");
  CMF_REQUIRE(geom_supported(D), "cmf_row_similarity: embedding_dim %d unsupported (1..1024)", D);
  CMF_REQUIRE(id >= 0 && id < rows, "cmf_row_similarity: id %lld is not a row of the table",
              (long long)id);
  const RowGeom g = row_geom(D);
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = grid_for(rows, g.G);
#define CALL(V_, G_, R_) k_row_similarity<V_, G_, R_><<<grid, 256, 0, s>>>(E, rows, D, id, out)
  CMF_DISPATCH_GEOM(g, CALL);
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

extern "C" int cmf_backward_scores(const float* U, const float* I, int32_t D, const void* users,
                                   const void* items, int32_t idx_bytes, int64_t M,
                                   const float* dpreds, const float* preds,
                                   const float* host_y_range, float* gU, float* gI, float* gbu,
                                   float* gbi, void* stream) {
  CMF_REQUIRE(U && I && gU && gI && gbu && gbi, "cmf_backward_scores: null pointer");
  CMF_REQUIRE(idx_bytes == 4 || idx_bytes == 8, "cmf_backward_scores: idx_bytes must be 4 or 8");
  CMF_REQUIRE(geom_supported(D), "cmf_backward_scores: embedding_dim %d unsupported", D);
  if (M == 0) return CMF_OK;
  CMF_REQUIRE(users && items && dpreds, "cmf_backward_scores: null id/grad pointer");
  const int has_range = host_y_range != nullptr;
  CMF_REQUIRE(!has_range || preds, "cmf_backward_scores: preds required with y_range");
  const float lo = has_range ? host_y_range[0] : 0.f, hi = has_range ? host_y_range[1] : 1.f;
  const RowGeom g = row_geom(D);
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = grid_for(M, g.G);
#define CALL(V_, G_, R_)                                                                          \
  if (idx_bytes == 8)                                                                             \
    k_backward_scores<int64_t, V_, G_, R_><<<grid, 256, 0, s>>>(                                  \
        U, I, D, (const int64_t*)users, (const int64_t*)items, M, dpreds, preds, has_range, lo,   \
        hi, gU, gI, gbu, gbi);                                                                    \
  else                                                                                            \
    k_backward_scores<int32_t, V_, G_, R_><<<grid, 256, 0, s>>>(                                  \
        U, I, D, (const int32_t*)users, (const int32_t*)items, M, dpreds, preds, has_range, lo,   \
        hi, gU, gI, gbu, gbi)
  CMF_DISPATCH_GEOM(g, CALL);
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}
