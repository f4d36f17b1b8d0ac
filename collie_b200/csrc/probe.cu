// Read-bandwidth probe for the roofline denominators of bench.py: streams a buffer `repeats` times
// with 16-byte loads from a persistent grid. A buffer that fits the 126 MB L2 gives the
// L2-resident read bandwidth, a larger one the HBM read bandwidth. No reference counterpart.
#include "common.cuh"

namespace cmf {

__global__ void __launch_bounds__(256)
k_read_probe(const float4* __restrict__ buf, int64_t n_vec, int repeats, float* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (int r = 0; r < repeats; ++r) {
    // four independent loads per iteration
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n_vec; i += 4 * stride) {
      // ld.global.cg: served by L2, never by this SM's L1
      const float4 a = __ldcg(buf + i), b = __ldcg(buf + i + stride), c = __ldcg(buf + i + 2 * stride),
                   d = __ldcg(buf + i + 3 * stride);
      acc += (a.x + b.y) + (c.z + d.w);
    }
    for (; i < n_vec; i += stride) acc += __ldcg(buf + i).x;
  }
  if (acc == 123.456f) out[0] = acc;   // keeps the loads alive without a store per thread
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_read_probe(const void* buf, size_t bytes, int32_t repeats, float* out,
                              void* stream) {
  CMF_REQUIRE(buf && out && bytes >= 16 && repeats >= 1, "cmf_read_probe: bad argument");
  CMF_REQUIRE((reinterpret_cast<uintptr_t>(buf) & 15) == 0, "cmf_read_probe: buffer not 16-byte aligned");
  k_read_probe<<<num_sms() * 8, 256, 0, (cudaStream_t)stream>>>((const float4*)buf,
                                                                 (int64_t)(bytes / 16), repeats, out);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

// Gather probe: reads 512-byte rows of `table` in the order given by `idx` (the forward kernel's
// access pattern without its arithmetic), U rows in flight per warp. layout 0: one warp instruction
// = one row (32 lanes x 16 bytes); layout 1: the g8 layout (8 lanes per row, 4 instructions per row,
// 4 rows per instruction).
namespace cmf {

template <int U, int LAYOUT>
__global__ void __launch_bounds__(256)
k_gather_probe(const float4* __restrict__ table, const int32_t* __restrict__ idx, int64_t n_idx,
               float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  float acc = 0.f;
  if (LAYOUT == 0) {
    for (int64_t i = warp * U; i + U <= n_idx; i += n_warps * U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = __ldcg(table + (int64_t)idx[i + u] * 32 + lane);
#pragma unroll
      for (int u = 0; u < U; ++u) acc += v[u].x + v[u].w;
    }
  } else {
    const int grp = lane >> 3, l = lane & 7;
    for (int64_t i = warp * 4 * U; i + 4 * U <= n_idx; i += n_warps * 4 * U) {
      float4 v[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const float4* row = table + (int64_t)idx[i + u * 4 + grp] * 32 + l;
#pragma unroll
        for (int k = 0; k < 4; ++k) v[u][k] = __ldcg(row + k * 8);
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc += v[u][k].x + v[u][k].w;
    }
  }
  if (acc == 123.456f) out[0] = acc;
}

}  // namespace cmf

namespace cmf {
// RED probe: vector atomic adds of 512-byte rows into `table` at idx[...] (k_bwd_update's item
// gradient pattern): one row per warp instruction, fire and forget.
__global__ void __launch_bounds__(256)
k_red_probe(float4* __restrict__ table, const int32_t* __restrict__ idx, int64_t n_idx) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const float4 one = make_float4(1.f, 1.f, 1.f, 1.f);
  for (int64_t i = warp; i < n_idx; i += n_warps)
    atomicAdd(table + (int64_t)idx[i] * 32 + lane, one);
}
}  // namespace cmf

extern "C" int cmf_red_probe(void* table, const int32_t* idx, int64_t n_idx, int32_t ctas_per_sm,
                             void* stream) {
  CMF_REQUIRE(table && idx && n_idx > 0 && ctas_per_sm >= 1 && ctas_per_sm <= 8,
              "cmf_red_probe: bad argument");
  k_red_probe<<<num_sms() * ctas_per_sm, 256, 0, (cudaStream_t)stream>>>((float4*)table, idx, n_idx);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

extern "C" int cmf_gather_probe(const void* table, const int32_t* idx, int64_t n_idx,
                                int32_t unroll, int32_t layout, int32_t ctas_per_sm, float* out,
                                void* stream) {
  CMF_REQUIRE(table && idx && out && n_idx > 0, "cmf_gather_probe: bad argument");
  CMF_REQUIRE(ctas_per_sm >= 1 && ctas_per_sm <= 8, "cmf_gather_probe: 1..8 CTAs per SM");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = num_sms() * ctas_per_sm;
#define GP(U_, L_) k_gather_probe<U_, L_><<<grid, 256, 0, s>>>((const float4*)table, idx, n_idx, out)
  if (layout == 0) {
    if (unroll <= 2) GP(2, 0); else if (unroll <= 4) GP(4, 0); else if (unroll <= 8) GP(8, 0); else GP(16, 0);
  } else {
    if (unroll <= 1) GP(1, 1); else if (unroll <= 2) GP(2, 1); else GP(4, 1);
  }
#undef GP
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}
