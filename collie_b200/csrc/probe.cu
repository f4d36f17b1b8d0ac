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
