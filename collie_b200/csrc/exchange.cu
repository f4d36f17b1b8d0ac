// Item-side exchange of a data-parallel step as ONE kernel over peer memory (NVLink / NVSwitch):
// the owner of an item row sums the ranks' gradient rows straight out of their workspaces
// (reduce-scatter), applies the optimizer rule to its shard and stores the new row into every
// rank's working copy of the table (all-gather). The rendezvous of the ranks is part of the same
// kernel: an entry barrier (every rank's step kernels have finished: its accumulators are
// complete) and an exit barrier (every owner's stores have landed in every copy) over a small
// signal pad in symmetric memory, so one step costs one launch instead of fence + kernel + fence.
// After the kernel (i.e. after the exit barrier) every rank clears its OWN accumulators with a
// local memset (nobody reads them any more), which keeps the zeroing off the links.
//
// Reference: PyTorch Lightning DDP all-reduces dense gradients of the replicated model between
// backward and optimizer.step (collie/model/base/trainer.py:463-470 is the single-process form);
// here users are partitioned over the ranks and only the item side is exchanged.
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
  uint32_t* pad[CMF_MAX_PEERS];   // signal pads (kPad* words each), NULL: the caller fences
  uint32_t epoch;
  uint4* clear_ptr;               // this rank's accumulators, cleared after the exit barrier
  int64_t clear_vec;              // ... in 16-byte units
};

// signal pad words
constexpr int kPadEntry = 0;     // [world] entry arrivals, written by rank q into word q
constexpr int kPadExit = 8;      // [world] exit arrivals
constexpr int kPadGoEntry = 16;  // local: block 0 passed the entry barrier
constexpr int kPadGoExit = 17;   // local: the last block passed the exit barrier
constexpr int kPadDone = 18;     // local: blocks that finished their rows
constexpr int kPadWords = 32;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t global_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// wait until *p has reached `epoch` (counters only grow; wrap-safe). A peer that never arrives
// (its process died) must not hang this GPU: after 20 s the kernel traps and the error surfaces
// at the caller's next synchronisation.
template <bool SYS>
__device__ __forceinline__ void spin_until(const uint32_t* p, uint32_t epoch) {
  const uint64_t t0 = global_ns();
  for (;;) {
    const uint32_t v = SYS ? ld_acquire_sys(p) : ld_acquire_gpu(p);
    if ((int32_t)(v - epoch) >= 0) return;
    __nanosleep(64);
    if (global_ns() - t0 > 20000000000ull) __trap();
  }
}

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

// MC: sums through multimem.ld_reduce (reduced inside the switch; its summation order is the
// switch's) and stores through multimem.st; otherwise plain peer loads summed in rank order and
// one store per copy.
template <int V, int G, int R, bool MC>
__global__ void __launch_bounds__(kThreads)
k_item_exchange_update(const PeersDev P, float* __restrict__ M, float* __restrict__ Vv,
                       float* __restrict__ bM, float* __restrict__ bV, int D, const OptDev o,
                       const OptDev ob) {
  constexpr int kGroups = kThreads / G;
  const int l = threadIdx.x & (G - 1);
  const int lane_w = threadIdx.x & 31;
  const unsigned gmask = group_mask<G>(lane_w);
  const int64_t group = (int64_t)blockIdx.x * kGroups + threadIdx.x / G;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  const bool fenced = P.pad[0] != nullptr;
  uint32_t* my = fenced ? P.pad[P.rank] : nullptr;

  // ---- entry barrier: every rank's accumulators are complete (their kernel is here too)
  if (fenced) {
    if (blockIdx.x == 0) {
      if ((int)threadIdx.x < P.world) {
        st_release_sys(P.pad[threadIdx.x] + kPadEntry + P.rank, P.epoch);
        spin_until<true>(my + kPadEntry + threadIdx.x, P.epoch);
      }
      __syncthreads();
      if (threadIdx.x == 0) st_release_gpu(my + kPadGoEntry, P.epoch);
    } else {
      if (threadIdx.x == 0) spin_until<false>(my + kPadGoEntry, P.epoch);
      __syncthreads();
    }
  }

  for (int64_t lr = group; lr < P.rows_per_rank; lr += n_groups) {
    const int64_t row = (int64_t)P.rank * P.rows_per_rank + lr;
    // everything this row needs from the peers is requested at once (one round trip): the
    // accumulators of a rank that did not gather the row are zero
    uint32_t touched = 0;
    Row<V, G, R> g;
    g.zero();
    float gb = 0.f;
    if (MC) {
      if (l == 0) {
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];"
                     : "=r"(touched) : "l"(P.mc_flag + row) : "memory");
        mc_ld_add<1>(P.mc_gbi + row, &gb);
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int d = (r * G + l) * V;
        if (d < D) mc_ld_add<V>(P.mc_GI + row * D + d, g.f + r * V);
      }
    } else if (R * V <= 8) {
      Row<V, G, R> t[CMF_MAX_PEERS];
      float tb[CMF_MAX_PEERS];
      uint32_t tf[CMF_MAX_PEERS];
#pragma unroll
      for (int q = 0; q < CMF_MAX_PEERS; ++q) {
        if (q < P.world) {
          t[q].load(P.GI[q] + row * D, D, l);
          tb[q] = 0.f;
          tf[q] = 0;
          if (l == 0) { tb[q] = P.gbi[q][row]; tf[q] = (uint32_t)P.flag[q][row]; }
        }
      }
      // gradient row = sum over ranks in rank order (a fixed order: runs repeat bit for bit)
#pragma unroll
      for (int q = 0; q < CMF_MAX_PEERS; ++q) {
        if (q < P.world) {
#pragma unroll
          for (int i = 0; i < R * V; ++i) g.f[i] += t[q].f[i];
          gb += tb[q];
          touched |= tf[q];
        }
      }
    } else {
      // long rows: one rank at a time (the registers of eight rows in flight would spill)
      for (int q = 0; q < P.world; ++q) {
        Row<V, G, R> t;
        t.load(P.GI[q] + row * D, D, l);
#pragma unroll
        for (int i = 0; i < R * V; ++i) g.f[i] += t.f[i];
        if (l == 0) { gb += P.gbi[q][row]; touched |= (uint32_t)P.flag[q][row]; }
      }
    }
    touched = __shfl_sync(gmask, touched, lane_w & ~(G - 1));
    if (!touched && !o.dense && !ob.dense) continue;
    if (!fenced && touched) {
      // caller-fenced mode: the owner clears what it consumed on every rank
      Row<V, G, R> z;
      z.zero();
      if (MC) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int d = (r * G + l) * V;
          if (d < D) mc_st<V>(P.mc_GI + row * D + d, z.f);
        }
        if (l == 0) {
          mc_st<1>(P.mc_gbi + row, z.f);
          asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;"
                       :: "l"(P.mc_flag + row), "r"(0u) : "memory");
        }
      } else {
        for (int q = 0; q < P.world; ++q) {
          z.store(P.GI[q] + row * D, D, l);
          if (l == 0) { P.gbi[q][row] = 0.f; P.flag[q][row] = 0; }
        }
      }
    }
    if (o.dense || touched) {
      Row<V, G, R> p, m, v;
      p.load(P.I[P.rank] + row * D, D, l);
      if (rule_has_m(o.rule)) m.load(M + lr * D, D, l); else m.zero();
      if (rule_has_v(o.rule)) v.load(Vv + lr * D, D, l); else v.zero();
      apply_rule_n<R * V>(o, p.f, m.f, v.f, g.f);
      if (rule_has_m(o.rule)) m.store(M + lr * D, D, l);
      if (rule_has_v(o.rule)) v.store(Vv + lr * D, D, l);
      if (MC) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int d = (r * G + l) * V;
          if (d < D) mc_st<V>(P.mc_I + row * D + d, p.f + r * V);
        }
      } else {
        for (int q = 0; q < P.world; ++q) p.store(P.I[q] + row * D, D, l);
      }
    }
    if (l == 0 && (ob.dense || touched)) {
      float p = P.bi[P.rank][row];
      float m = rule_has_m(ob.rule) ? bM[lr] : 0.f;
      float v = rule_has_v(ob.rule) ? bV[lr] : 0.f;
      apply_rule(ob, p, m, v, gb);
      if (rule_has_m(ob.rule)) bM[lr] = m;
      if (rule_has_v(ob.rule)) bV[lr] = v;
      if (MC) mc_st<1>(P.mc_bi + row, &p);
      else for (int q = 0; q < P.world; ++q) P.bi[q][row] = p;
    }
  }

  if (!fenced) return;
  // ---- exit barrier: this rank's stores are ordered before its exit signal; it leaves when every
  // rank has signalled, i.e. when every copy of the table is complete
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned ticket = atomicAdd(my + kPadDone, 1u);
    if (ticket == gridDim.x - 1) {
      my[kPadDone] = 0;
      __threadfence_system();
      for (int q = 0; q < P.world; ++q) st_release_sys(P.pad[q] + kPadExit + P.rank, P.epoch);
      for (int q = 0; q < P.world; ++q) spin_until<true>(my + kPadExit + q, P.epoch);
      st_release_gpu(my + kPadGoExit, P.epoch);
    }
  }
  // (nobody reads this rank's accumulators once the kernel has completed: the host clears them
  // with a memset behind it. Doing it here needed every CTA to wait for the last one, i.e. the
  // whole grid resident at once next to whatever else runs on the GPU: with a host-fed pipeline's
  // high-priority copy stream alongside, that hung.)
}

}  // namespace cmf

using namespace cmf;

extern "C" size_t cmf_exchange_pad_bytes(void) { return sizeof(uint32_t) * kPadWords; }

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
  const bool fenced = peers->pad[0] != nullptr;
  for (int q = 0; q < peers->world; ++q) {
    CMF_REQUIRE(peers->I[q] && peers->bi[q] && peers->GI[q] && peers->gbi[q] && peers->flag[q],
                "cmf_item_exchange_update: null buffer of rank %d", q);
    CMF_REQUIRE((peers->pad[q] != nullptr) == fenced,
                "cmf_item_exchange_update: give every rank's signal pad or none");
    P.I[q] = peers->I[q]; P.bi[q] = peers->bi[q]; P.GI[q] = peers->GI[q]; P.gbi[q] = peers->gbi[q];
    P.flag[q] = peers->flag[q];
    P.pad[q] = peers->pad[q];
  }
  P.mc_I = peers->mc_I; P.mc_bi = peers->mc_bi; P.mc_GI = peers->mc_GI; P.mc_gbi = peers->mc_gbi;
  P.mc_flag = peers->mc_flag;
  const bool mc = P.mc_I && P.mc_bi && P.mc_GI && P.mc_gbi && P.mc_flag;
  CMF_REQUIRE(mc || !(P.mc_I || P.mc_bi || P.mc_GI || P.mc_gbi || P.mc_flag),
              "cmf_item_exchange_update: give all five multicast addresses or none");
  if (fenced) {
    CMF_REQUIRE(peers->epoch != 0, "cmf_item_exchange_update: the barrier epoch starts at 1");
    CMF_REQUIRE(peers->clear_bytes % 16 == 0 && ((uintptr_t)peers->clear_ptr & 15) == 0,
                "cmf_item_exchange_update: the region to clear must be 16-byte aligned");
    P.epoch = peers->epoch;
    P.clear_ptr = (uint4*)peers->clear_ptr;
    P.clear_vec = peers->clear_ptr ? (int64_t)(peers->clear_bytes / 16) : 0;
  }
  if (P.rows_per_rank <= 0 && !fenced) return CMF_OK;
  const OptDev o = make_opt(*opt), ob = make_opt(*bias_opt);
  const RowGeom geom = row_geom(peers->D);
  const int groups_per_block = kThreads / geom.G;
  int64_t blocks = (P.rows_per_rank + groups_per_block - 1) / groups_per_block;
  const int64_t sms = num_sms();
  if (blocks < 1) blocks = 1;
  cudaStream_t s = (cudaStream_t)stream;
  const int D = peers->D;
  // CTAs spin at the entry barrier (and one at the exit barrier): the grid is capped so that they
  // leave room on every SM for the step's own kernels
#define LAUNCH(KERNEL)                                                                           \
  do {                                                                                           \
    int occ = 1;                                                                                 \
    if (fenced) {                                                                                \
      CMF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, KERNEL, kThreads, 0));        \
      CMF_REQUIRE(occ >= 1, "cmf_item_exchange_update: the kernel does not fit on an SM");       \
      if (occ > 2) occ = 2; /* (its CTAs spin at the barriers: leave the SMs to the step) */     \
    } else {                                                                                     \
      occ = 8;                                                                                   \
    }                                                                                            \
    const int64_t grid = blocks > sms * occ ? sms * occ : blocks;                                \
    KERNEL<<<(unsigned)grid, kThreads, 0, s>>>(P, M, V, bM, bV, D, o, ob);                       \
  } while (0)
#define CALL(V_, G_, R_)                                                                         \
  do {                                                                                           \
    if (mc) LAUNCH((k_item_exchange_update<V_, G_, R_, true>));                                  \
    else LAUNCH((k_item_exchange_update<V_, G_, R_, false>));                                    \
  } while (0)
  CMF_DISPATCH_GEOM(geom, CALL);
#undef CALL
#undef LAUNCH
  CMF_LAUNCH_CHECK();
  // the kernel completes after the exit barrier: every rank has finished reading this rank's
  // accumulators, which are cleared locally for the next step
  if (fenced && peers->clear_ptr && peers->clear_bytes > 0)
    CMF_CUDA(cudaMemsetAsync(peers->clear_ptr, 0, peers->clear_bytes, s));
  return CMF_OK;
}
