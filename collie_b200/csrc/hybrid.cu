// Row gather / scatter-add for models that feed embedding rows into dense layers
// (HybridPretrainedModel): the concatenated input of the first dense layer,
//   X[r] = [ user_meta[u_r] | U[u_r] | I[i_r] | item_meta[i_r] ],
// is laid out by ONE pass per source table straight into its column block of X (no torch.cat of
// four gathered tensors), and the gradient of X flows back into the embedding tables by vector RED.
//
// Reference: collie/model/hybrid_pretrained_matrix_factorization.py forward :401-464 (four
// nn.Embedding / fancy-index gathers + torch.cat), autograd's embedding backward.
#include "common.cuh"

namespace cmf {

// out[r, col0 : col0 + D] = table[ids[r], :]   (out has row pitch `ld` floats)
template <typename IdT, int V>
__global__ void __launch_bounds__(256)
k_gather_rows(const float* __restrict__ table, const IdT* __restrict__ ids, int64_t n, int D,
              float* __restrict__ out, int64_t ld, int col0) {
  const int vecs = D / V;
  const int64_t total = n * vecs;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = t / vecs;
    const int c = (int)(t - r * vecs) * V;
    float f[V];
    ld_vec<V>(table + (int64_t)ids[r] * D + c, f);
    st_vec<V>(out + r * ld + col0 + c, f);
  }
}

// grad_table[ids[r], :] += grad_out[r, col0 : col0 + D]
template <typename IdT, int V>
__global__ void __launch_bounds__(256)
k_scatter_add_rows(float* __restrict__ grad_table, const IdT* __restrict__ ids, int64_t n, int D,
                   const float* __restrict__ grad_out, int64_t ld, int col0) {
  const int vecs = D / V;
  const int64_t total = n * vecs;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = t / vecs;
    const int c = (int)(t - r * vecs) * V;
    float f[V];
    ld_vec<V>(grad_out + r * ld + col0 + c, f);
    red_vec<V>(grad_table + (int64_t)ids[r] * D + c, f);
  }
}

static int vec_width(const void* a, const void* b, int D, int64_t ld, int col0) {
  const uintptr_t bits = (uintptr_t)a | (uintptr_t)b;
  if (D % 4 == 0 && ld % 4 == 0 && col0 % 4 == 0 && (bits & 15) == 0) return 4;
  if (D % 2 == 0 && ld % 2 == 0 && col0 % 2 == 0 && (bits & 7) == 0) return 2;
  return 1;
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_gather_rows(const float* table, int64_t num_rows, int32_t D, const void* ids,
                               int32_t idx_bytes, int64_t n, float* out, int64_t ld, int32_t col0,
                               void* stream) {
  CMF_REQUIRE(table && out, "cmf_gather_rows: null pointer");
  CMF_REQUIRE(idx_bytes == 4 || idx_bytes == 8, "cmf_gather_rows: idx_bytes must be 4 or 8");
  CMF_REQUIRE(D >= 1 && num_rows >= 1 && col0 >= 0 && ld >= (int64_t)col0 + D,
              "cmf_gather_rows: bad shape (D = %d, ld = %lld, col0 = %d)", D, (long long)ld, col0);
  if (n == 0) return CMF_OK;
  CMF_REQUIRE(ids, "cmf_gather_rows: null ids");
  const int V = vec_width(table, out, D, ld, col0);
  const int64_t work = n * (D / V);
  int64_t blocks = (work + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = (cudaStream_t)stream;
#define CALL(T, V_) \
  k_gather_rows<T, V_><<<(unsigned)blocks, 256, 0, s>>>(table, (const T*)ids, n, D, out, ld, col0)
  if (idx_bytes == 8) { if (V == 4) CALL(int64_t, 4); else if (V == 2) CALL(int64_t, 2); else CALL(int64_t, 1); }
  else { if (V == 4) CALL(int32_t, 4); else if (V == 2) CALL(int32_t, 2); else CALL(int32_t, 1); }
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

extern "C" int cmf_scatter_add_rows(float* grad_table, int64_t num_rows, int32_t D, const void* ids,
                                    int32_t idx_bytes, int64_t n, const float* grad_out, int64_t ld,
                                    int32_t col0, void* stream) {
  CMF_REQUIRE(grad_table && grad_out, "cmf_scatter_add_rows: null pointer");
  CMF_REQUIRE(idx_bytes == 4 || idx_bytes == 8, "cmf_scatter_add_rows: idx_bytes must be 4 or 8");
  CMF_REQUIRE(D >= 1 && num_rows >= 1 && col0 >= 0 && ld >= (int64_t)col0 + D,
              "cmf_scatter_add_rows: bad shape (D = %d, ld = %lld, col0 = %d)", D, (long long)ld,
              col0);
  if (n == 0) return CMF_OK;
  CMF_REQUIRE(ids, "cmf_scatter_add_rows: null ids");
  const int V = vec_width(grad_table, grad_out, D, ld, col0);
  const int64_t work = n * (D / V);
  int64_t blocks = (work + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = (cudaStream_t)stream;
#define CALL(T, V_)                                                                              \
  k_scatter_add_rows<T, V_><<<(unsigned)blocks, 256, 0, s>>>(grad_table, (const T*)ids, n, D,    \
                                                             grad_out, ld, col0)
  if (idx_bytes == 8) { if (V == 4) CALL(int64_t, 4); else if (V == 2) CALL(int64_t, 2); else CALL(int64_t, 1); }
  else { if (V == 4) CALL(int32_t, 4); else if (V == 2) CALL(int32_t, 2); else CALL(int32_t, 1); }
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}
