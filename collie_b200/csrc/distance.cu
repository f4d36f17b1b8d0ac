// Pairwise-distance forward / backward of the metric-learning model
// (collie/model/collaborative_metric_learning.py:100-124): the score of a (user, item) pair is
//   || U[u] - I[i] + eps ||_2,   eps = 1e-6 added to every element of the difference
// (torch.nn.functional.pairwise_distance), one gather of each row and a reduction per pair in
// one kernel; the backward adds  g * (U[u] - I[i] + eps) / distance  into a dense gradient table
// of the users and its negative into the items' (what autograd's norm + embedding backward do,
// 0 where the distance is 0).
#include "common.cuh"

namespace cmf {

constexpr float kPairwiseEps = 1e-6f;

// one warp per pair
template <typename IdT>
__global__ void __launch_bounds__(256)
k_pairwise_distance(const float* __restrict__ U, const float* __restrict__ I, int D,
                    const IdT* __restrict__ users, const IdT* __restrict__ items, int64_t n,
                    float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp0; r < n; r += n_warps) {
    const float* u = U + (int64_t)users[r] * D;
    const float* v = I + (int64_t)items[r] * D;
    float ss = 0.f;
    for (int d = lane; d < D; d += 32) {
      const float x = (u[d] - v[d]) + kPairwiseEps;
      ss = fmaf(x, x, ss);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (lane == 0) out[r] = sqrtf(ss);
  }
}

template <typename IdT>
__global__ void __launch_bounds__(256)
k_pairwise_distance_bwd(const float* __restrict__ U, const float* __restrict__ I, int D,
                        const IdT* __restrict__ users, const IdT* __restrict__ items, int64_t n,
                        const float* __restrict__ grad_out, const float* __restrict__ dist,
                        float* __restrict__ gU, float* __restrict__ gI) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp0; r < n; r += n_warps) {
    const float dd = dist[r];
    const float coef = dd > 0.f ? grad_out[r] / dd : 0.f;
    if (coef == 0.f) continue;
    const int64_t uo = (int64_t)users[r] * D, io = (int64_t)items[r] * D;
    for (int d = lane; d < D; d += 32) {
      const float x = (U[uo + d] - I[io + d]) + kPairwiseEps;
      const float g = coef * x;
      if (gU) atomicAdd(gU + uo + d, g);
      if (gI) atomicAdd(gI + io + d, -g);
    }
  }
}

static int pair_grid(int64_t n) {
  int64_t blocks = (n * 32 + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  return (int)(blocks < 1 ? 1 : blocks);
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_pairwise_distance(const float* U, const float* I, int32_t D, const void* users,
                                     const void* items, int32_t idx_bytes, int64_t n, float* out,
                                     void* stream) {
  CMF_REQUIRE(U && I && out, "cmf_pairwise_distance: null pointer");
  CMF_REQUIRE(D >= 1, "cmf_pairwise_distance: bad embedding_dim %d", D);
  CMF_REQUIRE(idx_bytes == 4 || idx_bytes == 8, "cmf_pairwise_distance: ids must be int32 or int64");
  if (n == 0) return CMF_OK;
  CMF_REQUIRE(users && items, "cmf_pairwise_distance: null ids");
  cudaStream_t s = (cudaStream_t)stream;
  if (idx_bytes == 4)
    k_pairwise_distance<int32_t><<<pair_grid(n), 256, 0, s>>>(U, I, D, (const int32_t*)users,
                                                              (const int32_t*)items, n, out);
  else
    k_pairwise_distance<int64_t><<<pair_grid(n), 256, 0, s>>>(U, I, D, (const int64_t*)users,
                                                              (const int64_t*)items, n, out);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

extern "C" int cmf_pairwise_distance_backward(const float* U, const float* I, int32_t D,
                                              const void* users, const void* items,
                                              int32_t idx_bytes, int64_t n, const float* grad_out,
                                              const float* dist, float* gU, float* gI,
                                              void* stream) {
  CMF_REQUIRE(U && I && grad_out && dist, "cmf_pairwise_distance_backward: null pointer");
  CMF_REQUIRE(D >= 1, "cmf_pairwise_distance_backward: bad embedding_dim %d", D);
  CMF_REQUIRE(idx_bytes == 4 || idx_bytes == 8,
              "cmf_pairwise_distance_backward: ids must be int32 or int64");
  if (n == 0 || (!gU && !gI)) return CMF_OK;
  CMF_REQUIRE(users && items, "cmf_pairwise_distance_backward: null ids");
  cudaStream_t s = (cudaStream_t)stream;
  if (idx_bytes == 4)
    k_pairwise_distance_bwd<int32_t><<<pair_grid(n), 256, 0, s>>>(
        U, I, D, (const int32_t*)users, (const int32_t*)items, n, grad_out, dist, gU, gI);
  else
    k_pairwise_distance_bwd<int64_t><<<pair_grid(n), 256, 0, s>>>(
        U, I, D, (const int64_t*)users, (const int64_t*)items, n, grad_out, dist, gU, gI);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}
