// Evaluation, tensor-core path (precision = 1): S = U_tile . I^T as tcgen05 GEMM tiles fed by TMA,
// fp32 accumulators in TMEM, and the per-user top-k (+ the MRR counter) in the epilogue, so the
// score matrix never exists in memory.
//
// Reference: collie/metrics.py get_preds :77-107 (Bu * n_items independent gather-dots there),
// mapk :144-198, mrr :201-240.
//
// fp32 parity on bf16 tensor cores: every fp32 value x is split as hi = bf16(x), lo = bf16(x - hi)
// and the GEMM runs over the concatenated K axis  A' = [Ah | Ah | Al],  B' = [Bh | Bl | Bh]
// (K' = 3 D + 4 bias slots), i.e. S = Ah.Bh + Ah.Bl + Al.Bh with fp32 accumulation; the dropped Al.Bl term is
// < 2^-16 relative. Values that are exactly representable in bf16 (the exact-arithmetic parity
// fixture) give bit-exact scores.
//
// The (128-user tile, 256-item tile) pairs are cut into equal contiguous chunks, one per CTA, one
// CTA per SM: a CTA walks its chunk as segments that each lie inside one user tile, leaves every
// segment's top-k candidates and counters in scratch, and k_tc_merge combines the segments of a
// user in canonical order (score desc, id asc).
//   warp 0      TMA producer: A' once per segment (resident), then B' K-blocks through a 3-stage mbarrier ring
//   warp 1      MMA issuer: tcgen05.mma.cta_group::1.kind::f16, M = 128, N = 256, K = 16 per
//               instruction, two accumulators in TMEM (all 512 columns) so tile t+1 is being multiplied
//               while tile t is ranked
//   warp 2      TMEM allocation
//   warps 4..11 epilogue, two warps per scheduler (each takes half of a tile's columns):
//               tcgen05.ld 32x32b (thread = user row); the biases are folded into the GEMM (four
//               extra K slots), so the accumulator IS the score: a branch-free pass builds 32-bit
//               predicate masks per 32 columns (beats the row's k-th best / beats the best
//               target), then lane-parallel inserts into an unsorted per-row list in shared memory
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace cmf {

int g_tc_debug = 0;   // experiments only (cmf_set_tuning key 3): 1 = epilogue skips the ranking

constexpr int TC_BM = 128;        // users per CTA
constexpr int TC_BN = 256;        // items per MMA tile (one instruction: M = 128, N = 256, K = 16)
constexpr int TC_BK = 64;         // bf16 per K block = one 128-byte swizzle row
constexpr int TC_STAGES = 3;
constexpr int TC_MAXK = 16;       // top-k list length supported by the epilogue
constexpr int TC_A_BYTES = TC_BM * TC_BK * 2;   // 16 KB: one K-block of the user tile
constexpr int TC_B_BYTES = TC_BN * TC_BK * 2;   // 32 KB: one K-block of an item tile

// ---------------------------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t s_addr(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
// one lane of the (converged) warp, the same one every time
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, bf16 inputs, fp32 accumulate; issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier when all MMAs issued so far by this thread have completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
// 32 columns of this thread's TMEM lane; the caller issues tcgen05.wait::ld before reading v
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, float (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]),
        "=f"(v[7]), "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]),
        "=f"(v[14]), "=f"(v[15]), "=f"(v[16]), "=f"(v[17]), "=f"(v[18]), "=f"(v[19]), "=f"(v[20]),
        "=f"(v[21]), "=f"(v[22]), "=f"(v[23]), "=f"(v[24]), "=f"(v[25]), "=f"(v[26]), "=f"(v[27]),
        "=f"(v[28]), "=f"(v[29]), "=f"(v[30]), "=f"(v[31])
      : "r"(taddr)
      : "memory");
}
// tcgen05.wait::ld with the loaded registers as in/out operands, so that no use of them can be
// scheduled above the wait (an empty `ins` makes it a pure dependency carrier)
#define CMF_V32_INOUT "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]), "+f"(v[8]), "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15]), \
  "+f"(v[16]), "+f"(v[17]), "+f"(v[18]), "+f"(v[19]), "+f"(v[20]), "+f"(v[21]), "+f"(v[22]), "+f"(v[23]), "+f"(v[24]), "+f"(v[25]), "+f"(v[26]), "+f"(v[27]), "+f"(v[28]), "+f"(v[29]), "+f"(v[30]), "+f"(v[31])
__device__ __forceinline__ void tmem_wait_ld(float (&v)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" : CMF_V32_INOUT : : "memory");
}
__device__ __forceinline__ void tmem_ld_dep(float (&v)[32]) {
  asm volatile("" : CMF_V32_INOUT : : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// K-major operand tile with 128-byte swizzle: rows of 64 bf16 at a 128-byte pitch, 8-row groups
// 1024 bytes apart (SBO = 64 x 16 B), descriptor version 1 (sm_100), layout type 2 = SWIZZLE_128B.
// Field layout: cute/arch/mma_sm100_desc.hpp (SmemDescriptor).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) |
         (2ull << 61);
}
// the same descriptor from its low word (address field + LBO); the high word is constant
__device__ __forceinline__ uint32_t smem_desc_lo(uint32_t addr) {
  return ((addr >> 4) & 0x3FFF) | (1u << 16);
}
__device__ __forceinline__ uint64_t smem_desc_from_lo(uint32_t lo) {
  return ((uint64_t)0x40004040u << 32) | lo;
}
// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major, N = TC_BN, M = 128
// (InstrDescriptor in cute/arch/mma_sm100_desc.hpp: c_format [4,6), a_format [7,10),
// b_format [10,13), n_dim = N >> 3 at [17,23), m_dim = M >> 4 at [24,29)).
constexpr uint32_t kIdescBf16 =
    (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

// v[j] for a run-time j without spilling v to local memory (whose traffic would compete with the
// MMA's operand reads for the shared-memory pipe): a 5-level select tree, 31 selects
__device__ __forceinline__ float pick32(const float (&v)[32], int j) {
  float a[16], b[8], c[4];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = (j & 1) ? v[2 * i + 1] : v[2 * i];
#pragma unroll
  for (int i = 0; i < 8; ++i) b[i] = (j & 2) ? a[2 * i + 1] : a[2 * i];
#pragma unroll
  for (int i = 0; i < 4; ++i) c[i] = (j & 4) ? b[2 * i + 1] : b[2 * i];
  const float d0 = (j & 8) ? c[1] : c[0], d1 = (j & 8) ? c[3] : c[2];
  return (j & 16) ? d1 : d0;
}

// ---------------------------------------------------------------------------------- prep kernels
__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(x);
  lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}

// rows of `src` (optionally gathered through `ids`) -> bf16 rows of length Kp:
//   items: [hi | lo | hi | 1, 1, bi_hi, bi_lo | 0...]   users: [hi | hi | lo | bu_hi, bu_lo, 1, 1 | 0...]
// so the GEMM yields  <u, i> + bu + bi  directly (the biases are split like the rows). Rows
// >= n_rows are zero.
// One warp per row; a lane produces two adjacent bf16 at a time (one 4-byte store, 128 bytes per
// warp), the row id is read once per row.
__device__ __forceinline__ __nv_bfloat16 tc_split_elem(const float* __restrict__ srow, float b,
                                                       int c, int D, int users) {
  if (c < 3 * D) {
    const int seg = c / D, d = c - seg * D;
    __nv_bfloat16 hi, lo;
    split_bf16(srow[d], hi, lo);
    const bool want_lo = users ? (seg == 2) : (seg == 1);
    return want_lo ? lo : hi;
  }
  if (c < 3 * D + 4) {
    const int q = c - 3 * D;
    __nv_bfloat16 hi, lo;
    split_bf16(b, hi, lo);
    const bool own = users ? (q < 2) : (q >= 2);     // own bias in two slots, 1.0 in the others
    return own ? ((q & 1) ? lo : hi) : __float2bfloat16_rn(1.f);
  }
  return __float2bfloat16_rn(0.f);
}

__global__ void __launch_bounds__(256)
k_tc_split(const float* __restrict__ src, const float* __restrict__ bias,
           const int32_t* __restrict__ ids, int64_t row_offset, int64_t n_rows, int64_t n_rows_pad,
           int D, int Kp, int users, __nv_bfloat16* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp0; r < n_rows_pad; r += n_warps) {
    __nv_bfloat162* orow = reinterpret_cast<__nv_bfloat162*>(out + r * Kp);   // Kp is even
    if (r >= n_rows) {
      for (int c2 = lane; c2 < Kp / 2; c2 += 32) orow[c2] = __floats2bfloat162_rn(0.f, 0.f);
      continue;
    }
    const int64_t row = ids ? (int64_t)ids[r] : row_offset + r;
    const float* srow = src + row * D;
    const float b = bias[row];
    if ((D & 1) == 0 && (reinterpret_cast<uintptr_t>(src) & 7) == 0) {
      // each value is split once and stored to its three segments
      for (int d2 = lane; d2 < D / 2; d2 += 32) {
        const float2 x = *reinterpret_cast<const float2*>(srow + 2 * d2);
        __nv_bfloat162 hi, lo;
        split_bf16(x.x, hi.x, lo.x);
        split_bf16(x.y, hi.y, lo.y);
        orow[d2] = hi;
        orow[D / 2 + d2] = users ? hi : lo;
        orow[D + d2] = users ? lo : hi;
      }
      for (int c2 = 3 * D / 2 + lane; c2 < Kp / 2; c2 += 32) {   // bias slots + zero tail
        __nv_bfloat162 v;
        v.x = tc_split_elem(srow, b, 2 * c2, D, users);
        v.y = tc_split_elem(srow, b, 2 * c2 + 1, D, users);
        orow[c2] = v;
      }
      continue;
    }
    for (int c2 = lane; c2 < Kp / 2; c2 += 32) {
      __nv_bfloat162 v;
      v.x = tc_split_elem(srow, b, 2 * c2, D, users);
      v.y = tc_split_elem(srow, b, 2 * c2 + 1, D, users);
      orow[c2] = v;
    }
  }
}

// exact fp32 score of every target of every user of the tile; keeps the best (score desc, id asc)
__global__ void __launch_bounds__(256)
k_tc_target_best(const float* __restrict__ U, const float* __restrict__ I,
                 const float* __restrict__ bu, const float* __restrict__ bi, int D,
                 const int32_t* __restrict__ user_ids, int64_t n_users, int64_t item_lo,
                 int64_t item_hi, const int64_t* __restrict__ tgt_indptr,
                 const int32_t* __restrict__ tgt_items, float* __restrict__ best_s,
                 int32_t* __restrict__ best_id) {
  const int lane = threadIdx.x & 31;
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_users) return;
  const int64_t u = user_ids[r];
  float bs = -INFINITY;
  int bid = -1;
  for (int64_t t = tgt_indptr[r]; t < tgt_indptr[r + 1]; ++t) {
    const int id = tgt_items[t];
    if (id < item_lo || id >= item_hi) continue;
    float acc = 0.f;
    for (int d = lane; d < D; d += 32) acc = fmaf(U[u * D + d], I[(int64_t)id * D + d], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    const float s = acc + bu[u] + bi[id];
    if (bid < 0 || s > bs || (s == bs && id < bid)) { bs = s; bid = id; }
  }
  if (lane == 0) { best_s[r] = bs; best_id[r] = bid; }
}

// ---------------------------------------------------------------------------------- GEMM + top-k
struct TcParams {
  int n_users, k, KB, n_tiles, m_tiles, n_slots, want_mrr, debug;
  int last_steps;        // K = 16 steps of the last K block that hold data (1..4)
  int64_t chunk, total_work;   // tiles per CTA / tiles in all (m_tiles * n_tiles)
  int64_t n_items, item_offset;
  const float* best_s;   // MRR threshold per user (may be NULL)
  const int32_t* best_id;
  float* cand_s;         // [n_users_pad][n_slots][k]
  int32_t* cand_i;
  int32_t* cnt;          // [n_users_pad][n_slots] items ranked before the best target
};

// The (user tile, item tile) pairs are numbered m * n_tiles + n and cut into equal contiguous
// chunks, one per CTA (about one CTA per SM, whatever the tile counts): a CTA walks its chunk as
// one or more segments, each inside a single user tile. Slot of a segment inside its user tile:
__device__ __forceinline__ int tc_slot(int64_t lin_start, int m, int n_tiles, int64_t chunk) {
  return (int)(lin_start / chunk - ((int64_t)m * n_tiles) / chunk);
}

constexpr int TC_EPI_WARPS = 8;   // two per scheduler; each pair splits a tile's columns in halves
constexpr int TC_THREADS = 128 + 32 * TC_EPI_WARPS;

template <int KL>   // per-row list length kept in registers, >= k
__global__ void __launch_bounds__(TC_THREADS, 1)
k_tc_gemm_topk(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const TcParams p) {
  extern __shared__ unsigned char smem_raw[];
  // 128-byte-swizzled operand tiles need 1024-byte alignment
  unsigned char* smem = smem_raw + ((1024u - (s_addr(smem_raw) & 1023u)) & 1023u);
  // carve-up: A' K-blocks | B' stages | barriers
  unsigned char* sA = smem;
  unsigned char* sB = sA + (size_t)p.KB * TC_A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + (size_t)TC_STAGES * TC_B_BYTES);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 2 * TC_STAGES + 6);

  const uint32_t bar_full = s_addr(&bars[0]);                 // [TC_STAGES]
  const uint32_t bar_empty = s_addr(&bars[TC_STAGES]);        // [TC_STAGES]
  const uint32_t bar_a = s_addr(&bars[2 * TC_STAGES]);           // A' of the segment has landed
  const uint32_t bar_afree = s_addr(&bars[2 * TC_STAGES + 1]);   // A' is no longer read
  const uint32_t bar_tfull = s_addr(&bars[2 * TC_STAGES + 2]);   // [2]
  const uint32_t bar_tempty = s_addr(&bars[2 * TC_STAGES + 4]);  // [2]

  // warp index through a shuffle: the compiler then treats it (and the role branches) as uniform
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int64_t w_lo = (int64_t)blockIdx.x * p.chunk;
  const int64_t w_hi = min(p.total_work, w_lo + p.chunk);

  if (threadIdx.x == 0) {
    for (int i = 0; i < TC_STAGES; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    mbar_init(bar_a, 1);
    mbar_init(bar_afree, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull + 8 * i, 1); mbar_init(bar_tempty + 8 * i, TC_EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    // all 512 TMEM columns: two 128 x 256 fp32 accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     s_addr(s_tmem)), "r"(2 * TC_BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ===================================================================== TMA producer
    // (whole warp in the loop, one elected lane issues: see the MMA warp)
    if (w_lo < w_hi) {
      if (elect_one()) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
      }
      int stage = 0, seg = 0;
      uint32_t phase = 0;
      const uint32_t sA0 = s_addr(sA), sB0 = s_addr(sB);
      for (int64_t w = w_lo; w < w_hi; ++seg) {
        const int m = (int)(w / p.n_tiles);
        const int64_t seg_hi = min(w_hi, (int64_t)(m + 1) * p.n_tiles);
        if (seg > 0) mbar_wait(bar_afree, (seg - 1) & 1);   // the previous segment's MMAs are done
        if (elect_one()) {
          mbar_expect_tx(bar_a, (uint32_t)p.KB * TC_A_BYTES);
          for (int kb = 0; kb < p.KB; ++kb)
            tma_load_2d(sA0 + (uint32_t)kb * TC_A_BYTES, &tmA, kb * TC_BK, m * TC_BM, bar_a);
        }
        __syncwarp();
        for (; w < seg_hi; ++w) {
          const int n0 = (int)(w - (int64_t)m * p.n_tiles) * TC_BN;
          for (int kb = 0; kb < p.KB; ++kb) {
            mbar_wait(bar_empty + 8 * stage, phase ^ 1);
            if (elect_one()) {
              mbar_expect_tx(bar_full + 8 * stage, TC_B_BYTES);
              tma_load_2d(sB0 + (uint32_t)stage * TC_B_BYTES, &tmB, kb * TC_BK, n0,
                          bar_full + 8 * stage);
            }
            __syncwarp();
            if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    // The whole warp walks the loop (warp-uniform control flow and operands, so the descriptors
    // live in uniform registers) and one elected lane issues: with a `lane == 0` branch around
    // the loop the compiler wrapped every tcgen05 instruction in an elect/R2UR/branch sequence,
    // ~100 issue slots per K block, and the tensor pipe idled 60% of the time behind it.
    if (w_lo < w_hi) {
      int stage = 0, seg = 0, t = 0;
      uint32_t phase = 0;
      const uint32_t a_lo0 = smem_desc_lo(s_addr(sA)), b_lo0 = smem_desc_lo(s_addr(sB));
      for (int64_t w = w_lo; w < w_hi; ++seg) {
        const int m = (int)(w / p.n_tiles);
        const int64_t seg_hi = min(w_hi, (int64_t)(m + 1) * p.n_tiles);
        mbar_wait(bar_a, seg & 1);
        tc_fence_after();
        for (; w < seg_hi; ++w, ++t) {
          const int acc = t & 1;
          mbar_wait(bar_tempty + 8 * acc, ((t >> 1) & 1) ^ 1);   // the epilogue drained this buffer
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TC_BN);
          for (int kb = 0; kb < p.KB; ++kb) {
            mbar_wait(bar_full + 8 * stage, phase);
            tc_fence_after();
            const uint32_t a_lo = a_lo0 + (uint32_t)kb * (TC_A_BYTES >> 4);
            const uint32_t b_lo = b_lo0 + (uint32_t)stage * (TC_B_BYTES >> 4);
            // (the zero-padded tail of the last K block is not multiplied)
            const int steps = kb == p.KB - 1 ? p.last_steps : TC_BK / 16;
            if (elect_one()) {
#pragma unroll
              for (int j = 0; j < TC_BK / 16; ++j)   // +32 bytes (2 x 16 B) per K = 16 step
                if (j < steps)
                  umma_bf16(d_tmem, smem_desc_from_lo(a_lo + 2 * j),
                            smem_desc_from_lo(b_lo + 2 * j), kIdescBf16, (kb | j) != 0);
              umma_commit(bar_empty + 8 * stage);   // frees the stage when these MMAs are done
            }
            __syncwarp();
            if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
          }
          if (elect_one()) umma_commit(bar_tfull + 8 * acc);   // accumulator complete -> epilogue
          __syncwarp();
        }
        if (elect_one()) umma_commit(bar_afree);   // A' may be overwritten by the next segment
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===================================================================== epilogue
    const int q = warp & 3;               // TMEM lane quarter this warp may read
    const int half = (warp - 4) >> 2;     // which half of every tile's columns
    const int row = q * 32 + lane;        // user row inside the tile
    int t = 0;
    for (int64_t w = w_lo; w < w_hi;) {
      const int m = (int)(w / p.n_tiles);
      const int64_t seg_hi = min(w_hi, (int64_t)(m + 1) * p.n_tiles);
      const int slot = tc_slot(w, m, p.n_tiles, p.chunk) * 2 + half;
      const int urow = m * TC_BM + row;
      const bool valid = urow < p.n_users;
      // the row's best KL so far, sorted (score desc, id asc), in registers; thr = the last one
      float ls[KL];
      int li[KL];
#pragma unroll
      for (int e = 0; e < KL; ++e) { ls[e] = -INFINITY; li[e] = -1; }
      float thr = -INFINITY;
      const bool mrr = p.want_mrr && valid;
      const float best_s = mrr ? p.best_s[urow] : INFINITY;   // (rows past the end: m_rng is 0)
      const int best_id = mrr ? p.best_id[urow] : -1;
      int before = 0;

      for (; w < seg_hi; ++w, ++t) {
        const int acc = t & 1;
        const int64_t n0 = (w - (int64_t)m * p.n_tiles) * TC_BN;
        mbar_wait(bar_tfull + 8 * acc, (t >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) +
                               (uint32_t)(acc * TC_BN + half * (TC_BN / 2));
        const int ncols = (int)min((int64_t)TC_BN, p.n_items - n0);   // < TC_BN only in the last tile
        const int gid0 = (int)(p.item_offset + n0);
        // 64 columns at a time go to registers; the accumulator is handed back as soon as the
        // last of them has been read, before they are ranked
#pragma unroll 1
        for (int pp = 0; pp < TC_BN / 128; ++pp) {
        float v[2][32];
        tmem_ld32_nowait(taddr + pp * 64, v[0]);
        tmem_ld32_nowait(taddr + pp * 64 + 32, v[1]);
        tmem_wait_ld(v[0]);
        tmem_ld_dep(v[1]);
        if (pp == TC_BN / 128 - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
        }
        if (p.debug == 1) continue;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int c0 = half * (TC_BN / 2) + pp * 64 + h * 32;
          // branch-free pass: bit j of each mask = predicate of column c0 + j. x > y <=> the sign
          // bit of (y - x) (no flush-to-zero here, so a nonzero difference keeps its sign; equal
          // values give +0), and one funnel shift moves that bit into the mask: two instructions
          // per column and mask, in four independent chains
          uint32_t mi[4] = {0, 0, 0, 0};
#pragma unroll
          for (int j = 7; j >= 0; --j) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
              mi[c] = __funnelshift_l(__float_as_uint(thr - v[h][8 * c + j]), mi[c], 1);
          }
          uint32_t m_ins = mi[0] | (mi[1] << 8) | (mi[2] << 16) | (mi[3] << 24);
          const int left = ncols - c0;
          const uint32_t m_rng = (!valid || left <= 0) ? 0u
                                 : (left >= 32 ? 0xffffffffu : ((1u << left) - 1u));
          if (p.want_mrr) {   // (uniform branch: no predicated-off copies on the top-k-only path)
            uint32_t mg[4] = {0, 0, 0, 0}, ml[4] = {0, 0, 0, 0};
#pragma unroll
            for (int j = 7; j >= 0; --j) {
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                const float x = v[h][8 * c + j];
                mg[c] = __funnelshift_l(__float_as_uint(best_s - x), mg[c], 1);   // x > best
                ml[c] = __funnelshift_l(__float_as_uint(x - best_s), ml[c], 1);   // x < best
              }
            }
            const uint32_t m_gt = mg[0] | (mg[1] << 8) | (mg[2] << 16) | (mg[3] << 24);
            const uint32_t m_eq = ~(m_gt | ml[0] | (ml[1] << 8) | (ml[2] << 16) | (ml[3] << 24));
            // ties with the best target count when their id is smaller; the target never does
            const int d = best_id - (gid0 + c0);      // columns with gid < best_id: j < d
            const uint32_t m_idlt = d >= 32 ? 0xffffffffu : (d > 0 ? ((1u << d) - 1u) : 0u);
            const uint32_t m_self = (d >= 0 && d < 32) ? (1u << d) : 0u;
            before += __popc(m_gt & m_rng & ~m_self) + __popc(m_eq & m_rng & m_idlt);
          }
          m_ins &= m_rng;
          if (p.debug == 2) m_ins &= 1u;   // experiments: (almost) no inserts
          // lanes insert side by side, each its own next candidate (column j differs per lane)
          while (__any_sync(0xffffffffu, m_ins != 0)) {
            if (m_ins) {
              const int j = __ffs(m_ins) - 1;
              m_ins &= m_ins - 1;
              const float sc = pick32(v[h], j);
              if (sc > thr) {     // thr may have risen within the chunk
                // sorted insert; an equal score stays behind the entries already there (ids
                // only grow along a segment), which is the canonical order
                float cs = sc;
                int ci = gid0 + c0 + j;
                bool up = false;   // sticky: below the insertion point everything shifts down
#pragma unroll
                for (int e = 0; e < KL; ++e) {
                  up = up || cs > ls[e];
                  const float os = ls[e];
                  const int oi = li[e];
                  ls[e] = up ? cs : os;
                  li[e] = up ? ci : oi;
                  cs = up ? os : cs;
                  ci = up ? oi : ci;
                }
                thr = ls[KL - 1];
              }
            }
          }
        }
        }
      }
      if (valid) {
        const int K = p.k;
        float* cs = p.cand_s + ((int64_t)urow * p.n_slots + slot) * K;
        int32_t* ci = p.cand_i + ((int64_t)urow * p.n_slots + slot) * K;
#pragma unroll
        for (int j = 0; j < KL; ++j)
          if (j < K) { cs[j] = ls[j]; ci[j] = li[j]; }
        if (p.cnt) p.cnt[(int64_t)urow * p.n_slots + slot] = before;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * TC_BN)
                 : "memory");
  }
}

// Merge the candidates of a user's segments. Every slot list arrives sorted (score desc, id asc)
// and item ids are unique across a user's slots, so this is a k-round merge of sorted lists: G
// lanes per user (G = n_slots rounded up to a power of two, 32 / G users per warp), one slot per
// lane with its head as a 64-bit key (ordered score bits | inverted id; larger = ranked first), a
// log2(G)-level max per round, the winning lane advances. The picks go through shared memory so
// the target lookups of all k picks run side by side; AP@k from the hit mask in rank order.
__device__ __forceinline__ uint64_t tc_key(float s, int id) {
  if (id < 0) return 0ull;
  uint32_t u = __float_as_uint(s + 0.0f);            // (-0 -> +0: equal scores, equal keys)
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);   // monotone: larger float -> larger uint
  return ((uint64_t)u << 32) | (uint32_t)(0x7fffffff - id);
}
__device__ __forceinline__ float tc_key_score(uint64_t key) {
  uint32_t u = (uint32_t)(key >> 32);
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
  return __uint_as_float(u);
}

template <int G>
__global__ void __launch_bounds__(256)
k_tc_merge(const float* __restrict__ cand_s, const int32_t* __restrict__ cand_i,
           const int32_t* __restrict__ cnt, int n_slots, int k, int64_t n_users,
           const int64_t* __restrict__ tgt_indptr, const int32_t* __restrict__ tgt_items,
           const int32_t* __restrict__ best_id, int32_t* __restrict__ topk_ids,
           float* __restrict__ topk_scores, double* __restrict__ ap,
           int32_t* __restrict__ n_before_best) {
  constexpr int UPW = 32 / G;                       // users per warp
  __shared__ uint64_t s_pick[8][UPW][TC_MAXK];      // [warp][user][rank]
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int gl = lane & (G - 1), ug = lane / G;
  const int64_t r = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * UPW + ug;
  const bool live = r < n_users;
  // slots gl, gl + G, ...: (only n_slots > 32 gives a lane more than one; then the best head of
  // its slots is found by a scan every round)
  const int total = n_slots * k;
  const float* cs = cand_s + (live ? r : 0) * total;
  const int32_t* ci = cand_i + (live ? r : 0) * total;
  int head = 0;                                     // (G >= n_slots: position in slot gl)
  uint64_t key = 0;
  uint64_t prev = ~0ull;
  auto load_head = [&]() {
    if (G >= 32 && n_slots > 32) {
      // several slots per lane: best candidate below the previous pick
      key = 0;
      for (int sl = gl; sl < n_slots; sl += G)
        for (int e = 0; e < k; ++e) {
          const uint64_t kk = tc_key(cs[sl * k + e], ci[sl * k + e]);
          if (kk < prev && kk > key) key = kk;
        }
    } else {
      key = (live && gl < n_slots && head < k) ? tc_key(cs[gl * k + head], ci[gl * k + head]) : 0ull;
    }
  };
  load_head();
  for (int j = 0; j < k; ++j) {
    uint64_t best = key;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
      const uint64_t other = __shfl_xor_sync(0xffffffffu, best, o);
      best = other > best ? other : best;
    }
    if (gl == 0) s_pick[wib][ug][j] = best;
    if (best != 0 && key == best) {   // the winner moves on
      ++head;
      prev = best;
      load_head();
    } else if (G >= 32) {
      prev = best ? best : prev;
    }
  }
  __syncwarp();
  int64_t t0 = 0;
  int n_tgt = 0;
  if (live) { t0 = tgt_indptr[r]; n_tgt = (int)(tgt_indptr[r + 1] - t0); }
  uint32_t hits = 0;
  for (int j = gl; j < k; j += G) {
    const uint64_t pk = s_pick[wib][ug][j];
    const int id = pk ? 0x7fffffff - (int)(uint32_t)pk : -1;
    if (live) {
      if (topk_ids) topk_ids[r * k + j] = id;
      if (topk_scores) topk_scores[r * k + j] = pk ? tc_key_score(pk) : -INFINITY;
      if (id >= 0 && csr_contains(tgt_items + t0, 0, n_tgt, id)) hits |= 1u << j;
    }
  }
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) hits |= __shfl_xor_sync(0xffffffffu, hits, o);
  if (live && gl == 0) {
    if (ap) {
      double ap_acc = 0.0;
      for (int j = 0; j < k; ++j)
        if ((hits >> j) & 1u) ap_acc += (double)__popc(hits & ((2u << j) - 1u)) / (double)(j + 1);
      const int denom = k < n_tgt ? k : n_tgt;
      ap[r] = denom > 0 ? ap_acc / (double)denom : 0.0;
    }
    if (n_before_best) {
      int tot = 0;
      if (cnt) for (int sl = 0; sl < n_slots; ++sl) tot += cnt[r * n_slots + sl];
      n_before_best[r] = (best_id && best_id[r] >= 0) ? tot : -1;
    }
  }
}

// ---------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// [rows, Kp] bf16 row-major, box = 64 columns x box_rows rows, 128-byte swizzle, OOB rows read as zero
static int make_map(CUtensorMap* map, void* base, int64_t rows, int Kp, int box_rows) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return fail(CMF_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {(cuuint64_t)Kp, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)Kp * 2};
  cuuint32_t box[2] = {TC_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(CMF_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return CMF_OK;
}

struct TcLayout {
  int Kp, KB, m_tiles, n_tiles, n_ctas, n_slots;
  int64_t m_pad, chunk, total_work;
  size_t Bp, Ap, bu, best_s, best_id, cand_s, cand_i, cnt, total;
};

static TcLayout tc_layout(int64_t n_users, int64_t n_items, int D, int k) {
  TcLayout L;
  L.Kp = (int)align_up((size_t)3 * D + 4, TC_BK);   // + the bias slots
  L.KB = L.Kp / TC_BK;
  L.m_tiles = (int)((n_users + TC_BM - 1) / TC_BM);
  L.m_pad = (int64_t)L.m_tiles * TC_BM;
  L.n_tiles = (int)((n_items + TC_BN - 1) / TC_BN);
  // equal contiguous chunks of the (user tile, item tile) pairs, one per CTA, one CTA per SM
  L.total_work = (int64_t)L.m_tiles * L.n_tiles;
  int64_t ctas = num_sms();
  if (ctas > L.total_work) ctas = L.total_work;
  L.chunk = (L.total_work + ctas - 1) / ctas;
  L.n_ctas = (int)((L.total_work + L.chunk - 1) / L.chunk);
  // segments a user tile can be cut into (x 2: the epilogue keeps one list per column half)
  L.n_slots = 2 * (int)((L.n_tiles + L.chunk - 1) / L.chunk + 1);
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 1024); return at; };
  L.Bp = take((size_t)n_items * L.Kp * 2);
  L.Ap = take((size_t)L.m_pad * L.Kp * 2);
  L.bu = take(sizeof(float) * L.m_pad);
  L.best_s = take(sizeof(float) * L.m_pad);
  L.best_id = take(sizeof(int32_t) * L.m_pad);
  L.cand_s = take(sizeof(float) * L.m_pad * L.n_slots * k);
  L.cand_i = take(sizeof(int32_t) * L.m_pad * L.n_slots * k);
  L.cnt = take(sizeof(int32_t) * L.m_pad * L.n_slots);
  L.total = o;
  return L;
}

bool eval_tc_supported(int D, int k) { return D >= 1 && D <= 128 && k >= 1 && k <= TC_MAXK; }

size_t eval_tc_workspace_bytes(int64_t n_users, int64_t n_items, int D, int k) {
  if (!eval_tc_supported(D, k)) return 0;
  return tc_layout(n_users, n_items, D, k).total;
}

int eval_tile_tc(const cmf_eval_t* a, cudaStream_t s) {
  CMF_REQUIRE(!a->want_auc, "cmf_eval_tile: AUC is computed by the exact path (precision = 0)");
  if (!eval_tc_supported(a->D, a->k))
    return fail(CMF_ERR_UNSUPPORTED,
                "cmf_eval_tile: the tensor-core path takes embedding_dim <= 128 and k <= %d "
                "(got D = %d, k = %d); use precision = 0", TC_MAXK, a->D, a->k);
  const TcLayout L = tc_layout(a->n_users_tile, a->num_items_shard, a->D, a->k);
  if (!a->workspace || a->workspace_bytes < L.total)
    return fail(CMF_ERR_WORKSPACE, "cmf_eval_tile: workspace of %zu bytes needed, %zu given",
                L.total, a->workspace ? a->workspace_bytes : (size_t)0);
  char* base = (char*)a->workspace;
  __nv_bfloat16* Bp = (__nv_bfloat16*)(base + L.Bp);
  __nv_bfloat16* Ap = (__nv_bfloat16*)(base + L.Ap);
  float* best_s = (float*)(base + L.best_s);
  int32_t* best_id = (int32_t*)(base + L.best_id);
  const int grid = num_sms() * 8;

  if (!a->reuse_item_prep) {
    k_tc_split<<<grid, 256, 0, s>>>(a->I, a->bi, nullptr, a->item_offset, a->num_items_shard,
                                    a->num_items_shard, a->D, L.Kp, 0, Bp);
    CMF_LAUNCH_CHECK();
  }
  k_tc_split<<<grid, 256, 0, s>>>(a->U, a->bu, a->user_ids, 0, a->n_users_tile, L.m_pad, a->D,
                                  L.Kp, 1, Ap);
  CMF_LAUNCH_CHECK();
  const bool mrr = a->want_mrr != 0;
  if (mrr) {
    if (a->ext_best_id) {
      CMF_CUDA(cudaMemcpyAsync(best_s, a->ext_best_score, sizeof(float) * a->n_users_tile,
                               cudaMemcpyDeviceToDevice, s));
      CMF_CUDA(cudaMemcpyAsync(best_id, a->ext_best_id, sizeof(int32_t) * a->n_users_tile,
                               cudaMemcpyDeviceToDevice, s));
    } else {
      k_tc_target_best<<<(unsigned)((a->n_users_tile * 32 + 255) / 256), 256, 0, s>>>(
          a->U, a->I, a->bu, a->bi, a->D, a->user_ids, a->n_users_tile, a->item_offset,
          a->item_offset + a->num_items_shard, a->tgt_indptr, a->tgt_items, best_s, best_id);
      CMF_LAUNCH_CHECK();
    }
    if (a->best_score)
      CMF_CUDA(cudaMemcpyAsync(a->best_score, best_s, sizeof(float) * a->n_users_tile,
                               cudaMemcpyDeviceToDevice, s));
    if (a->best_id)
      CMF_CUDA(cudaMemcpyAsync(a->best_id, best_id, sizeof(int32_t) * a->n_users_tile,
                               cudaMemcpyDeviceToDevice, s));
  }

  CUtensorMap tmA, tmB;
  int rc = make_map(&tmA, Ap, L.m_pad, L.Kp, TC_BM);
  if (rc) return rc;
  rc = make_map(&tmB, Bp, a->num_items_shard, L.Kp, TC_BN);
  if (rc) return rc;

  TcParams p{};
  p.n_users = (int)a->n_users_tile; p.k = a->k; p.KB = L.KB; p.n_tiles = L.n_tiles;
  p.last_steps = (3 * a->D + 4 - (L.KB - 1) * TC_BK + 15) / 16;
  p.m_tiles = L.m_tiles; p.n_slots = L.n_slots; p.chunk = L.chunk; p.total_work = L.total_work;
  p.want_mrr = mrr ? 1 : 0;
  p.n_items = a->num_items_shard; p.item_offset = a->item_offset; p.debug = g_tc_debug;
  p.best_s = best_s; p.best_id = best_id;
  p.cand_s = (float*)(base + L.cand_s); p.cand_i = (int32_t*)(base + L.cand_i);
  p.cnt = (int32_t*)(base + L.cnt);

  const size_t smem = (size_t)L.KB * TC_A_BYTES + (size_t)TC_STAGES * TC_B_BYTES +
                      (2 * TC_STAGES + 6) * sizeof(uint64_t) + 64 + 1024;
  // slots a user tile does not use stay empty
  CMF_CUDA(cudaMemsetAsync(p.cand_i, 0xFF, sizeof(int32_t) * L.m_pad * L.n_slots * a->k, s));
  CMF_CUDA(cudaMemsetAsync(p.cnt, 0, sizeof(int32_t) * L.m_pad * L.n_slots, s));
  // list length: the smallest instantiation that holds k
#define CMF_TC_LAUNCH(KL)                                                                        \
  do {                                                                                           \
    CMF_CUDA(cudaFuncSetAttribute(k_tc_gemm_topk<KL>,                                            \
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));      \
    k_tc_gemm_topk<KL><<<(unsigned)L.n_ctas, TC_THREADS, smem, s>>>(tmA, tmB, p);                \
  } while (0)
  if (a->k <= 4) CMF_TC_LAUNCH(4);
  else if (a->k <= 8) CMF_TC_LAUNCH(8);
  else if (a->k <= 10) CMF_TC_LAUNCH(10);
  else CMF_TC_LAUNCH(16);
#undef CMF_TC_LAUNCH
  CMF_LAUNCH_CHECK();

#define CMF_TC_MERGE(G)                                                                         \
  k_tc_merge<G><<<(unsigned)((a->n_users_tile * (G) + 255) / 256), 256, 0, s>>>(p.cand_s, p.cand_i, mrr ? p.cnt : nullptr, L.n_slots, a->k, a->n_users_tile, a->tgt_indptr, a->tgt_items, mrr ? best_id : nullptr, a->topk_ids, a->topk_scores, a->ap, a->n_before_best)
  if (L.n_slots <= 4) CMF_TC_MERGE(4);
  else if (L.n_slots <= 8) CMF_TC_MERGE(8);
  else if (L.n_slots <= 16) CMF_TC_MERGE(16);
  else CMF_TC_MERGE(32);
#undef CMF_TC_MERGE
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

}  // namespace cmf
