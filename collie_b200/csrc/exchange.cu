// Item-side exchange of a data-parallel step as ONE kernel over peer memory (NVLink / NVSwitch):
// the owner of an item row sums the ranks' gradient rows straight out of their workspaces
// (reduce-scatter), applies the optimizer rule to its shard, stores the new row into every rank's
// working copy of the table (all-gather) and clears the accumulators it consumed.
//
// Reference: PyTorch Lightning DDP all-reduces dense gradients of the replicated model between
// backward and optimizer.step (collie/model/base/trainer.py:463-470 is the single-process form);
// here users are partitioned over the ranks and only the item side is exchanged.
//
// Ordering is the caller's: every rank's step kernels must have finished before any rank runs
// this kernel, and every rank's kernel must have finished before any rank's next step reads its
// table (a stream-ordered collective on both sides; see collie_b200/distributed.py).
#include "train_common.cuh"

namespace cmf {

struct PeersDev {
  int rank, world;
  int64_t rows_per_rank;
  float* I[CMF_MAX_PEERS];
  float* bi[CMF_MAX_PEERS];
  float* GI[CMF_MAX_PEERS];
  float* gbi[CMF_MAX_PEERS];
  int32_t* flag[CMF_MAX_PEERS];
  float *mc_I, *mc_bi, *mc_GI, *mc_gbi;
  int32_t* mc_flag;
};

// ---- NVSwitch multicast (NVLS): one address stands for the same offset in every rank's buffer
template <int V>
__device__ __forceinline__ void mc_ld_add(const float* mc, float* out);
template <>
__device__ __forceinline__ void mc_ld_add<4>(const float* mc, float* out) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(out[0]), "=f"(out[1]), "=f"(out[2]), "=f"(out[3]) : "l"(mc) : "memory");
}
template <>
__device__ __forceinline__ void mc_ld_add<2>(const float* mc, float* out) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v2.f32 {%0, %1}, [%2];"
               : "=f"(out[0]), "=f"(out[1]) : "l"(mc) : "memory");
}
template <>
__device__ __forceinline__ void mc_ld_add<1>(const float* mc, float* out) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];"
               : "=f"(out[0]) : "l"(mc) : "memory");
}
template <int V>
__device__ __forceinline__ void mc_st(float* mc, const float* v);
template <>
__device__ __forceinline__ void mc_st<4>(float* mc, const float* v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               :: "l"(mc), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
}
template <>
__device__ __forceinline__ void mc_st<2>(float* mc, const float* v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};"
               :: "l"(mc), "f"(v[0]), "f"(v[1]) : "memory");
}
template <>
__device__ __forceinline__ void mc_st<1>(float* mc, const float* v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" :: "l"(mc), "f"(v[0]) : "memory");
}

// the same exchange with the reduction done by the switch and one store reaching every rank
template <int V, int G, int R>
__global__ void __launch_bounds__(kThreads)
k_item_exchange_update_mc(const PeersDev P, float* __restrict__ M, float* __restrict__ Vv,
                          float* __restrict__ bM, float* __restrict__ bV, int D, const OptDev o,
                          const OptDev ob) {
  constexpr int kGroups = kThreads / G;
  const int l = threadIdx.x & (G - 1);
  const int lane_w = threadIdx.x & 31;
  const unsigned gmask = group_mask<G>(lane_w);
  const int64_t group = (int64_t)blockIdx.x * kGroups + threadIdx.x / G;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  const float zeros[4] = {0.f, 0.f, 0.f, 0.f};
  for (int64_t lr = group; lr < P.rows_per_rank; lr += n_groups) {
    const int64_t row = (int64_t)P.rank * P.rows_per_rank + lr;
    uint32_t touched = 0;
    if (l == 0)
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];"
                   : "=r"(touched) : "l"(P.mc_flag + row) : "memory");
    touched = __shfl_sync(gmask, touched, lane_w & ~(G - 1));
    if (!touched && !o.dense && !ob.dense) continue;
    Row<V, G, R> g;
    g.zero();
    float gb = 0.f;
    if (touched) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int d = (r * G + l) * V;
        if (d < D) {
          mc_ld_add<V>(P.mc_GI + row * D + d, g.f + r * V);
          mc_st<V>(P.mc_GI + row * D + d, zeros);
        }
      }
      if (l == 0) {
        mc_ld_add<1>(P.mc_gbi + row, &gb);
        mc_st<1>(P.mc_gbi + row, zeros);
        asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;"
                     :: "l"(P.mc_flag + row), "r"(0u) : "memory");
      }
    }
    if (o.dense || touched) {
      Row<V, G, R> p, m, v;
      p.load(P.I[P.rank] + row * D, D, l);
      if (rule_has_m(o.rule)) m.load(M + lr * D, D, l); else m.zero();
      if (rule_has_v(o.rule)) v.load(Vv + lr * D, D, l); else v.zero();
#pragma unroll
      for (int i = 0; i < R * V; ++i) apply_rule(o, p.f[i], m.f[i], v.f[i], g.f[i]);
      if (rule_has_m(o.rule)) m.store(M + lr * D, D, l);
      if (rule_has_v(o.rule)) v.store(Vv + lr * D, D, l);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int d = (r * G + l) * V;
        if (d < D) mc_st<V>(P.mc_I + row * D + d, p.f + r * V);
      }
    }
    if (l == 0 && (ob.dense || touched)) {
      float p = P.bi[P.rank][row];
      float m = rule_has_m(ob.rule) ? bM[lr] : 0.f;
      float v = rule_has_v(ob.rule) ? bV[lr] : 0.f;
      apply_rule(ob, p, m, v, gb);
      if (rule_has_m(ob.rule)) bM[lr] = m;
      if (rule_has_v(ob.rule)) bV[lr] = v;
      mc_st<1>(P.mc_bi + row, &p);
    }
  }
}

template <int V, int G, int R>
__global__ void __launch_bounds__(kThreads)
k_item_exchange_update(const PeersDev P, float* __restrict__ M, float* __restrict__ Vv,
                       float* __restrict__ bM, float* __restrict__ bV, int D, const OptDev o,
                       const OptDev ob) {
  constexpr int kGroups = kThreads / G;
  const int l = threadIdx.x & (G - 1);
  const int64_t group = (int64_t)blockIdx.x * kGroups + threadIdx.x / G;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  for (int64_t lr = group; lr < P.rows_per_rank; lr += n_groups) {
    const int64_t row = (int64_t)P.rank * P.rows_per_rank + lr;
    // which ranks gathered this row (their accumulators hold something and need clearing)
    uint32_t tmask = 0;
    for (int q = 0; q < P.world; ++q)
      if (P.flag[q][row] != 0) tmask |= 1u << q;
    if (!tmask && !o.dense && !ob.dense) continue;
    // gradient row = sum over ranks in rank order (a fixed order: runs repeat bit for bit)
    Row<V, G, R> g;
    g.zero();
    float gb = 0.f;
    for (int q = 0; q < P.world; ++q) {
      if (!((tmask >> q) & 1u)) continue;
      Row<V, G, R> t;
      float* gq = P.GI[q] + row * D;
      t.load(gq, D, l);
#pragma unroll
      for (int i = 0; i < R * V; ++i) g.f[i] += t.f[i];
      t.zero();
      t.store(gq, D, l);
      if (l == 0) {
        gb += P.gbi[q][row];
        P.gbi[q][row] = 0.f;
        P.flag[q][row] = 0;
      }
    }
    if (o.dense || tmask) {
      Row<V, G, R> p, m, v;
      p.load(P.I[P.rank] + row * D, D, l);
      if (rule_has_m(o.rule)) m.load(M + lr * D, D, l); else m.zero();
      if (rule_has_v(o.rule)) v.load(Vv + lr * D, D, l); else v.zero();
#pragma unroll
      for (int i = 0; i < R * V; ++i) apply_rule(o, p.f[i], m.f[i], v.f[i], g.f[i]);
      if (rule_has_m(o.rule)) m.store(M + lr * D, D, l);
      if (rule_has_v(o.rule)) v.store(Vv + lr * D, D, l);
      for (int q = 0; q < P.world; ++q) p.store(P.I[q] + row * D, D, l);
    }
    if (l == 0 && (ob.dense || tmask)) {
      float p = P.bi[P.rank][row];
      float m = rule_has_m(ob.rule) ? bM[lr] : 0.f;
      float v = rule_has_v(ob.rule) ? bV[lr] : 0.f;
      apply_rule(ob, p, m, v, gb);
      if (rule_has_m(ob.rule)) bM[lr] = m;
      if (rule_has_v(ob.rule)) bV[lr] = v;
      for (int q = 0; q < P.world; ++q) P.bi[q][row] = p;
    }
  }
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_item_exchange_update(const cmf_peers_t* peers, float* M, float* V, float* bM,
                                        float* bV, const cmf_opt_t* opt, const cmf_opt_t* bias_opt,
                                        void* stream) {
  CMF_REQUIRE(peers && opt && bias_opt, "cmf_item_exchange_update: null argument");
  CMF_REQUIRE(peers->world >= 1 && peers->world <= CMF_MAX_PEERS && peers->rank >= 0 &&
                  peers->rank < peers->world,
              "cmf_item_exchange_update: rank %d of %d is out of range (at most %d peers)",
              peers->rank, peers->world, CMF_MAX_PEERS);
  CMF_REQUIRE(geom_supported(peers->D), "cmf_item_exchange_update: embedding_dim %d unsupported",
              peers->D);
  int rc = validate_opt(*opt, "embedding", M, V);
  if (rc) return rc;
  rc = validate_opt(*bias_opt, "bias", bM, bV);
  if (rc) return rc;
  PeersDev P{};
  P.rank = peers->rank;
  P.world = peers->world;
  P.rows_per_rank = peers->rows_per_rank;
  for (int q = 0; q < peers->world; ++q) {
    CMF_REQUIRE(peers->I[q] && peers->bi[q] && peers->GI[q] && peers->gbi[q] && peers->flag[q],
                "cmf_item_exchange_update: null buffer of rank %d", q);
    P.I[q] = peers->I[q]; P.bi[q] = peers->bi[q]; P.GI[q] = peers->GI[q]; P.gbi[q] = peers->gbi[q];
    P.flag[q] = peers->flag[q];
  }
  P.mc_I = peers->mc_I; P.mc_bi = peers->mc_bi; P.mc_GI = peers->mc_GI; P.mc_gbi = peers->mc_gbi;
  P.mc_flag = peers->mc_flag;
  const bool mc = P.mc_I && P.mc_bi && P.mc_GI && P.mc_gbi && P.mc_flag;
  CMF_REQUIRE(mc || !(P.mc_I || P.mc_bi || P.mc_GI || P.mc_gbi || P.mc_flag),
              "cmf_item_exchange_update: give all five multicast addresses or none");
  if (P.rows_per_rank <= 0) return CMF_OK;
  const OptDev o = make_opt(*opt), ob = make_opt(*bias_opt);
  const RowGeom geom = row_geom(peers->D);
  const int groups_per_block = kThreads / geom.G;
  int64_t blocks = (P.rows_per_rank + groups_per_block - 1) / groups_per_block;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = (cudaStream_t)stream;
  const int D = peers->D;
#define CALL(V_, G_, R_)                                                                        \
  do {                                                                                          \
    if (mc)                                                                                     \
      k_item_exchange_update_mc<V_, G_, R_><<<(unsigned)blocks, kThreads, 0, s>>>(P, M, V, bM,  \
                                                                                  bV, D, o, ob); \
    else                                                                                        \
      k_item_exchange_update<V_, G_, R_><<<(unsigned)blocks, kThreads, 0, s>>>(P, M, V, bM, bV, \
                                                                               D, o, ob);      \
  } while (0)
  CMF_DISPATCH_GEOM(geom, CALL);
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}
