// Shared device/host helpers for the collie_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <string>

#include "../../include/collie_b200.h"

namespace cmf {

// ---------------------------------------------------------------------------------- host side
extern thread_local std::string g_last_error;
extern std::atomic<long long> g_launches;

inline int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}

#define CMF_CUDA(expr)                                                                        \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess)                                                                    \
      return ::cmf::fail(CMF_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                         __FILE__, __LINE__);                                                 \
  } while (0)

#define CMF_LAUNCH_CHECK()                                                                   \
  do {                                                                                       \
    ::cmf::g_launches.fetch_add(1, std::memory_order_relaxed);                               \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess)                                                                   \
      return ::cmf::fail(CMF_ERR_CUDA, "kernel launch failed: %s (%s:%d)",                   \
                         cudaGetErrorString(_e), __FILE__, __LINE__);                        \
  } while (0)

#define CMF_REQUIRE(cond, ...)                                  \
  do {                                                          \
    if (!(cond)) return ::cmf::fail(CMF_ERR_INVALID, __VA_ARGS__); \
  } while (0)

int num_sms();

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Row geometry: V floats per vector access, G lanes cooperate on one row, R vectors per lane.
// A warp works on 32 / G rows (interactions) at once, so a smaller G amortises the per-interaction
// scalar work (ids, RNG, loss) over more lanes at the price of R * V registers per row fragment.
struct RowGeom {
  int V, G, R;
};
inline RowGeom row_geom(int D, int max_g = 32) {
  RowGeom g;
  g.V = (D % 4 == 0) ? 4 : (D % 2 == 0) ? 2 : 1;
  const int vecs = D / g.V;
  if (max_g < 8) max_g = 8;
  g.G = 4;
  while (g.G < max_g && g.G < vecs) g.G <<= 1;
  if (g.G == 4 && vecs > 4) g.G = 8;
  g.R = (vecs + g.G - 1) / g.G;
  while (g.G < 32 && (g.R > 4 || (g.G == 4 && g.R > 1))) {  // instantiated shapes only
    g.G <<= 1;
    g.R = (vecs + g.G - 1) / g.G;
  }
  return g;
}

// Dispatch on <V, G, R>; R is rounded up to the instantiated {1, 2, 4, 8}. D <= 1024.
#define CMF_DISPATCH_GR(V_, geom, CALL)                                              \
  do {                                                                               \
    const int _G = (geom).G, _R = (geom).R;                                          \
    if (_G == 4) { CALL(V_, 4, 1); }                                                 \
    else if (_G == 8) {                                                              \
      if (_R == 1) { CALL(V_, 8, 1); } else if (_R == 2) { CALL(V_, 8, 2); }         \
      else { CALL(V_, 8, 4); }                                                       \
    } else if (_G == 16) {                                                           \
      if (_R == 1) { CALL(V_, 16, 1); } else if (_R == 2) { CALL(V_, 16, 2); }       \
      else { CALL(V_, 16, 4); }                                                      \
    } else {                                                                         \
      if (_R == 1) { CALL(V_, 32, 1); } else if (_R == 2) { CALL(V_, 32, 2); }       \
      else if (_R <= 4) { CALL(V_, 32, 4); } else { CALL(V_, 32, 8); }               \
    }                                                                                \
  } while (0)

#define CMF_DISPATCH_GEOM(geom, CALL)                                   \
  do {                                                                  \
    if ((geom).V == 4) { CMF_DISPATCH_GR(4, geom, CALL); }              \
    else if ((geom).V == 2) { CMF_DISPATCH_GR(2, geom, CALL); }         \
    else { CMF_DISPATCH_GR(1, geom, CALL); }                            \
  } while (0)

inline bool geom_supported(int D) {
  if (D < 1) return false;
  RowGeom g = row_geom(D);
  return g.R <= 8;
}

// ---------------------------------------------------------------------------------- device side
#ifdef __CUDACC__

template <int V>
struct VecT;
template <>
struct VecT<4> {
  using type = float4;
};
template <>
struct VecT<2> {
  using type = float2;
};
template <>
struct VecT<1> {
  using type = float;
};

template <int V>
__device__ __forceinline__ void ld_vec(const float* p, float* f) {
  if constexpr (V == 4) {
    float4 t = *reinterpret_cast<const float4*>(p);
    f[0] = t.x; f[1] = t.y; f[2] = t.z; f[3] = t.w;
  } else if constexpr (V == 2) {
    float2 t = *reinterpret_cast<const float2*>(p);
    f[0] = t.x; f[1] = t.y;
  } else {
    f[0] = *p;
  }
}

template <int V>
__device__ __forceinline__ void st_vec(float* p, const float* f) {
  if constexpr (V == 4) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  } else if constexpr (V == 2) {
    *reinterpret_cast<float2*>(p) = make_float2(f[0], f[1]);
  } else {
    *p = f[0];
  }
}

// vectorised fp32 reduction into global memory (RED.ADD.F32x4 / x2 on sm_90+)
template <int V>
__device__ __forceinline__ void red_vec(float* p, const float* f) {
  if constexpr (V == 4) {
    atomicAdd(reinterpret_cast<float4*>(p), make_float4(f[0], f[1], f[2], f[3]));
  } else if constexpr (V == 2) {
    atomicAdd(reinterpret_cast<float2*>(p), make_float2(f[0], f[1]));
  } else {
    atomicAdd(p, f[0]);
  }
}

// A row fragment: lane `l` of a G-lane group holds elements (r*G + l)*V + j, r < R, j < V.
template <int V, int G, int R>
struct Row {
  float f[R * V];

  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < R * V; ++i) f[i] = 0.f;
  }
  __device__ __forceinline__ void load(const float* __restrict__ row, int D, int l) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int d = (r * G + l) * V;
      if (d < D) {
        ld_vec<V>(row + d, f + r * V);
      } else {
#pragma unroll
        for (int j = 0; j < V; ++j) f[r * V + j] = 0.f;
      }
    }
  }
  __device__ __forceinline__ void store(float* __restrict__ row, int D, int l) const {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int d = (r * G + l) * V;
      if (d < D) st_vec<V>(row + d, f + r * V);
    }
  }
  __device__ __forceinline__ void red_add(float* __restrict__ row, int D, int l) const {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int d = (r * G + l) * V;
      if (d < D) red_vec<V>(row + d, f + r * V);
    }
  }
  // partial dot product of this lane's elements (reduce across the group afterwards)
  __device__ __forceinline__ float dot(const Row& o) const {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < R * V; ++i) s = fmaf(f[i], o.f[i], s);
    return s;
  }
  // f += round(a * x): the product is rounded BEFORE the add (no FMA), as autograd does when it
  // scales an embedding row by a gradient coefficient and then sums contributions. This makes
  // +g*x and -g*x cancel exactly, which sign-sensitive rules (Adam's first steps) depend on.
  __device__ __forceinline__ void axpy(float a, const Row& x) {
#pragma unroll
    for (int i = 0; i < R * V; ++i) f[i] = __fadd_rn(f[i], __fmul_rn(a, x.f[i]));
  }
};

template <int G>
__device__ __forceinline__ unsigned group_mask(int lane_in_warp) {
  if constexpr (G == 32) {
    return 0xffffffffu;
  } else {
    return ((1u << G) - 1u) << (lane_in_warp & ~(G - 1));
  }
}

template <int G>
__device__ __forceinline__ float group_sum(float v, unsigned mask) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o, G);
  return v;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// Philox4x32-10 (Salmon et al. 2011). counter = (c0..c3), key = (k0, k1).
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// d-th uniform item id of the interaction at dataset position gidx (see oracle `_draw`).
__device__ __forceinline__ int32_t draw_item(uint64_t seed, uint32_t epoch, int64_t gidx, uint32_t d,
                                             uint32_t num_items) {
  const uint4 w = philox4x32_10(
      make_uint4((uint32_t)gidx, (uint32_t)((uint64_t)gidx >> 32), epoch, d >> 2),
      make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  const uint32_t word = (d & 3) == 0 ? w.x : (d & 3) == 1 ? w.y : (d & 3) == 2 ? w.z : w.w;
  return (int32_t)__umulhi(word, num_items);
}

// is `item` in the sorted id list [lo, hi)?
__device__ __forceinline__ bool csr_contains(const int32_t* __restrict__ items, int64_t lo, int64_t hi,
                                             int32_t item) {
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    const int32_t v = items[mid];
    if (v == item) return true;
    if (v < item) lo = mid + 1; else hi = mid;
  }
  return false;
}

#endif  // __CUDACC__
}  // namespace cmf
