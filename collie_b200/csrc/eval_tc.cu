// Evaluation, tensor-core path (precision = 1): S = U_tile . I^T as tcgen05 GEMM tiles fed by TMA,
// fp32 accumulators in TMEM, and the per-user ranking in the epilogue, so the score matrix never
// exists in memory. The tensor cores SELECT, fp32 decides: the GEMM's scores carry a stated error
// bound, the epilogue keeps every item that could be in the top-k (and every item whose order
// against the best target is in doubt), and a small kernel re-scores exactly those in fp32 in the
// exact path's own summation order (eval.cu). Ids, scores, AP@k and the MRR counter therefore
// equal precision = 0 bit for bit on arbitrary fp32 weights; users for which the bound cannot
// separate the candidates (more near-ties than the lists hold) are re-ranked by the exact path.
//
// Reference: collie/metrics.py get_preds :77-107 (Bu * n_items independent gather-dots there),
// mapk :144-198, mrr :201-240.
//
// Operands: x = hi + lo with hi = bf16(x), lo = bf16(x - hi).
//   terms = 1 (top-k / MAP@k only):  A' = [Uh | bu as 3 bf16 pieces, 1, 1, 1 | a_u, a2_u]
//                                    B' = [Ih | 1, 1, 1, bi as 3 pieces     | b_i, 1   ]
//     K' = D + 8: one bf16 product per element. |S~ - S| <= c1 |u| |i| + 2^-20 (|bu| + |bi|),
//     c1 = 4.1e-3 (2^-8 for the two roundings, Cauchy-Schwarz over d, plus both accumulations).
//   terms = 3 (with the MRR counter): A' = [Uh | Uh | Ul | ...], B' = [Ih | Il | Ih | ...],
//     K' = 3 D + 8, S~ = Uh.Ih + Uh.Il + Ul.Ih: c3 = 8e-5 (3 * 2^-18 for the dropped terms, the
//     rest for fp32 accumulation of 3 D + 4 products).
//   The last two slots carry the bound ITSELF into the accumulator: a_u = c |u|, b_i = |i| (both
//   rounded UP to bf16), a2_u = 2^-20 (|bu| + max |bi|), so the GEMM yields
//       B = S~ + a_u b_i + a2_u >= S,
//   an upper bound of the exact score with the margin e = a_u b_i + a2_u of that very pair (no
//   shard-wide largest norm: a few large item rows do not widen everybody's margin), and
//   B - 2 e (1 + 2^-10) <= S is the matching lower bound, recomputed by k_tc_finish for the few
//   entries it looks at. Lists are kept by B. Sufficiency: if the largest B any list dropped lies
//   below the k-th largest LOWER bound of the entries, no dropped item can be in the exact top-k;
//   candidates = entries whose B reaches that k-th lower bound.
//   MRR: B < best target's exact score: surely not before it; B > that score + 2 e_tile (e_tile =
//   the largest margin among the tile's 256 columns): surely before it; between: re-scored. For
//   this form the split copy holds the item rows ORDERED BY NORM (radix sort of the squared norms,
//   once per evaluation), which makes e_tile tight; lists and band then hold positions in that
//   copy and k_tc_finish maps them back to ids.
//
// The (128-user tile, 256-item tile) pairs are cut into equal contiguous chunks, one per CTA, one
// CTA per SM: a CTA walks its chunk as segments that each lie inside one user tile and leaves every
// segment's best KL > k candidates (per column half) in a scratch slot.
//   warp 0      TMA producer: A' once per segment (resident), then B' K-blocks through a 3-stage mbarrier ring
//   warp 1      MMA issuer: tcgen05.mma.cta_group::1.kind::f16, M = 128, N = 256, K = 16 per
//               instruction, two accumulators in TMEM (all 512 columns) so tile t+1 is being multiplied
//               while tile t is ranked
//   warp 2      TMEM allocation
//   warps 4..11 epilogue, two warps per scheduler (each takes half of a tile's columns):
//               tcgen05.ld 32x32b (thread = user row); the biases are folded into the GEMM (four
//               extra K slots), so the accumulator IS the (upper bound of the) score: a branch-free pass builds 32-bit
//               predicate masks per 32 columns (beats the row's KL-th best / surely beats the best
//               target / cannot be told from it), then lane-parallel inserts into a sorted list
//               held in registers
// k_tc_finish (one warp per user): k-th best of the union of the slots' lists, the proof that
// nothing outside the lists can reach the top-k, exact re-scoring, exact top-k, AP@k, MRR counter.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace cmf {

int g_tc_debug = 0;   // experiments only (cmf_set_tuning key 3): 1 = epilogue skips the ranking

constexpr int TC_BM = 128;        // users per CTA
constexpr int TC_BN = 256;        // items per MMA tile (one instruction: M = 128, N = 256, K = 16)
constexpr int TC_BK = 64;         // bf16 per K block = one 128-byte swizzle row
constexpr int TC_STAGES = 3;
constexpr int TC_MAXK = 16;       // largest k
constexpr int TC_BIAS_SLOTS = 8;   // extra K slots: the two biases (three bf16 pieces each) + two error-bound slots
constexpr int TC_BCAP = 512;      // band entries kept per USER for the MRR counter (all its slots share them)
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

// a - b as one FFMA (b * -1 + a: the same bits as the subtraction). On sm_100 FADD issues on the
// alu pipe, where the funnel shifts of the mask pass already are; FFMA with an immediate
// multiplier issues on the fma pipe at twice the rate, so the pass is no longer alu-bound.
__device__ __forceinline__ float sub_on_fma_pipe(float a, float b) {
  float d;
  asm("fma.rn.f32 %0, %1, 0fBF800000, %2;" : "=f"(d) : "f"(b), "f"(a));
  return d;
}

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

// Error-bound coefficients (see the file header). Kept generous: a bound that is too large only
// sends a few more items to the exact re-scoring, one that is too small breaks parity.
constexpr float kTcC1 = 4.1e-3f;
constexpr float kTcC3 = 8.0e-5f;
constexpr float kTcCBias = 9.5367431640625e-7f;   // 2^-20: the biases are exact, their sums are not

// rows of `src` (optionally gathered through `ids`) -> bf16 rows of length Kp, T = 1 or 3 terms:
//   T = 3  items: [hi | lo | hi | 1, 1, 1, bi x3 | 0...]   users: [hi | hi | lo | bu x3, 1, 1, 1 | 0...]
//   T = 1  items: [hi | 1, 1, 1, bi x3 | 0...]             users: [hi | bu x3, 1, 1, 1 | 0...]
// so the GEMM yields  <u, i> + bu + bi  directly (each bias as three bf16 pieces: exact). Rows
// >= n_rows are zero. Also the row's L2 norm and |bias| (users: per row; items: the maximum over
// the shard, by atomicMax on the bit pattern of a non-negative float).
// One warp per row; a lane produces two adjacent bf16 at a time (one 4-byte store, 128 bytes per
// warp), the row id is read once per row.
// smallest bf16 >= x (x >= 0)
__device__ __forceinline__ __nv_bfloat16 bf16_up(float x) {
  __nv_bfloat16 h = __float2bfloat16_rn(x);
  if (__bfloat162float(h) < x) h = __ushort_as_bfloat16((unsigned short)(__bfloat16_as_ushort(h) + 1));
  return h;
}

// ea / eb: this row's two error-bound slots (users: a_u, a2_u; items: b_i, 1), see the file header
__device__ __forceinline__ __nv_bfloat16 tc_split_elem(const float* __restrict__ srow, float b,
                                                       int c, int D, int users, int T,
                                                       __nv_bfloat16 ea, __nv_bfloat16 eb) {
  if (c < T * D) {
    const int seg = c / D, d = c - seg * D;
    __nv_bfloat16 hi, lo;
    split_bf16(srow[d], hi, lo);
    const bool want_lo = T == 3 && (users ? (seg == 2) : (seg == 1));
    return want_lo ? lo : hi;
  }
  if (c < T * D + 6) {
    // the bias as THREE bf16 pieces (24 mantissa bits: the whole fp32 value) against 1.0, so the
    // biases enter the accumulator exactly; users own the first three slots, items the last three
    const int q = c - T * D;
    const bool own = users ? (q < 3) : (q >= 3);
    if (!own) return __float2bfloat16_rn(1.f);
    const __nv_bfloat16 p0 = __float2bfloat16_rn(b);
    const float r1 = b - __bfloat162float(p0);
    const __nv_bfloat16 p1 = __float2bfloat16_rn(r1);
    const __nv_bfloat16 p2 = __float2bfloat16_rn(r1 - __bfloat162float(p1));
    const int piece = q % 3;
    return piece == 0 ? p0 : (piece == 1 ? p1 : p2);
  }
  if (c == T * D + 6) return ea;
  if (c == T * D + 7) return eb;
  return __float2bfloat16_rn(0.f);
}

// Error-bound slots. Users carry a_u = bf16_up(c |u|) and a2_u = bf16_up(2^-20 (|bu| + max|bi|)),
// items b_i = bf16_up(|i|) and 1: the accumulator becomes B = S~ + a_u b_i + a2_u >= S, an UPPER
// bound of the exact score with the margin of that very pair (no shard-wide largest norm). The
// values as stored (as floats) go to row_a / row_a2 so that k_tc_finish can take the margin off
// again: users row_a = a_u, row_a2 = a2_u; items row_a = b_i. stats[1] = max |bi| of the shard
// (items are split first).
__global__ void __launch_bounds__(256)
k_tc_split(const float* __restrict__ src, const float* __restrict__ bias,
           const int32_t* __restrict__ ids, int64_t row_offset, int64_t n_rows, int64_t n_rows_pad,
           int D, int Kp, int users, int T, float c, __nv_bfloat16* __restrict__ out,
           float* __restrict__ row_a, float* __restrict__ row_a2, uint32_t* __restrict__ stats) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const float bmax = users ? __uint_as_float(stats[1]) : 0.f;
  for (int64_t r = warp0; r < n_rows_pad; r += n_warps) {
    __nv_bfloat162* orow = reinterpret_cast<__nv_bfloat162*>(out + r * Kp);   // Kp is even
    if (r >= n_rows) {
      for (int c2 = lane; c2 < Kp / 2; c2 += 32) orow[c2] = __floats2bfloat162_rn(0.f, 0.f);
      if (lane == 0) {
        if (row_a) row_a[r] = 0.f;
        if (row_a2) row_a2[r] = 0.f;
      }
      continue;
    }
    const int64_t row = row_offset + (ids ? (int64_t)ids[r] : r);
    const float* srow = src + row * D;
    const float b = bias[row];
    float ss = 0.f;
    const bool fast = (D & 1) == 0 && (reinterpret_cast<uintptr_t>(src) & 7) == 0;
    if (fast) {
      // each value is split once and stored to its segments
      for (int d2 = lane; d2 < D / 2; d2 += 32) {
        const float2 x = *reinterpret_cast<const float2*>(srow + 2 * d2);
        ss = fmaf(x.x, x.x, ss);
        ss = fmaf(x.y, x.y, ss);
        __nv_bfloat162 hi, lo;
        split_bf16(x.x, hi.x, lo.x);
        split_bf16(x.y, hi.y, lo.y);
        orow[d2] = hi;
        if (T == 3) {
          orow[D / 2 + d2] = users ? hi : lo;
          orow[D + d2] = users ? lo : hi;
        }
      }
    } else {
      for (int d = lane; d < D; d += 32) ss = fmaf(srow[d], srow[d], ss);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float nrm = sqrtf(ss) * 1.00002f;     // (fp32 sum of squares: D * 2^-24 relative)
    const __nv_bfloat16 ea = bf16_up(users ? c * nrm : nrm);
    const __nv_bfloat16 eb = users ? bf16_up(kTcCBias * (fabsf(b) + bmax) + 1e-37f)
                                   : __float2bfloat16_rn(1.f);
    for (int c2 = (fast ? T * D / 2 : 0) + lane; c2 < Kp / 2; c2 += 32) {   // (bias, bound slots, zero tail)
      __nv_bfloat162 v;
      v.x = tc_split_elem(srow, b, 2 * c2, D, users, T, ea, eb);
      v.y = tc_split_elem(srow, b, 2 * c2 + 1, D, users, T, ea, eb);
      orow[c2] = v;
    }
    if (lane == 0) {
      if (row_a) row_a[r] = __bfloat162float(ea);
      if (row_a2) row_a2[r] = __bfloat162float(eb);
      if (!users && stats) {   // (read first: all rows hammering one word was most of the kernel's time)
        if (__float_as_uint(fabsf(b)) > *(volatile uint32_t*)(stats + 1))
          atomicMax(stats + 1, __float_as_uint(fabsf(b)));
      }
    }
  }
}

// sort keys for the item order of the MRR form: squared norm of every row (non-negative floats
// order like their bit patterns) and the row's number
__global__ void __launch_bounds__(256)
k_tc_norm_keys(const float* __restrict__ src, int64_t row_offset, int64_t n_rows, int D,
               uint32_t* __restrict__ keys, int32_t* __restrict__ vals) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp0; r < n_rows; r += n_warps) {
    const float* srow = src + (row_offset + r) * D;
    float ss = 0.f;
    for (int d = lane; d < D; d += 32) ss = fmaf(srow[d], srow[d], ss);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (lane == 0) { keys[r] = __float_as_uint(ss); vals[r] = (int32_t)r; }
  }
}

// largest b_i of every item tile (TC_BN rows): what the epilogue can know about a column's margin
__global__ void __launch_bounds__(TC_BN)
k_tc_tile_max(const float* __restrict__ item_b, int64_t n_items, float* __restrict__ tile_bmax) {
  const int64_t i = (int64_t)blockIdx.x * TC_BN + threadIdx.x;
  float v = i < n_items ? item_b[i] : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  __shared__ float sm[TC_BN / 32];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = 0.f;
    for (int w = 0; w < TC_BN / 32; ++w) m = fmaxf(m, sm[w]);
    tile_bmax[blockIdx.x] = m;
  }
}

// the exact path's score of one (user, item) pair: fmaf over d ascending from 0, then + bu, + bi
// (k_score_tile / k_score_tile_128 in eval.cu): same bits
__device__ __forceinline__ float tc_exact_score(const float* __restrict__ urow,
                                                const float* __restrict__ irow, int D, float ub,
                                                float ib) {
  float acc = 0.f;
  if ((D & 3) == 0 && ((reinterpret_cast<uintptr_t>(urow) | reinterpret_cast<uintptr_t>(irow)) & 15) == 0) {
    for (int d = 0; d < D; d += 4) {
      const float4 a = *reinterpret_cast<const float4*>(urow + d);
      const float4 b = *reinterpret_cast<const float4*>(irow + d);
      acc = fmaf(a.x, b.x, acc);
      acc = fmaf(a.y, b.y, acc);
      acc = fmaf(a.z, b.z, acc);
      acc = fmaf(a.w, b.w, acc);
    }
  } else {
    for (int d = 0; d < D; ++d) acc = fmaf(urow[d], irow[d], acc);
  }
  return acc + ub + ib;
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
  const float ub = bu[u];
  float bs = -INFINITY;
  int bid = -1;
  // lanes take targets side by side; every score is one lane's sequential sum (the exact order)
  for (int64_t t = tgt_indptr[r] + lane; t < tgt_indptr[r + 1]; t += 32) {
    const int id = tgt_items[t];
    if (id < item_lo || id >= item_hi) continue;
    const float s = tc_exact_score(U + u * D, I + (int64_t)id * D, D, ub, bi[id]);
    if (bid < 0 || s > bs || (s == bs && id < bid)) { bs = s; bid = id; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float os = __shfl_xor_sync(0xffffffffu, bs, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bid, o);
    if (oi >= 0 && (bid < 0 || os > bs || (os == bs && oi < bid))) { bs = os; bid = oi; }
  }
  if (lane == 0) { best_s[r] = bs; best_id[r] = bid; }
}

// ---------------------------------------------------------------------------------- GEMM + top-k
struct TcParams {
  int n_users, k, KB, n_tiles, m_tiles, n_slots, want_mrr, debug;
  int last_steps;        // K = 16 steps of the last K block that hold data (1..4)
  int64_t chunk, total_work;   // tiles per CTA / tiles in all (m_tiles * n_tiles)
  int64_t n_items, item_offset;
  const float* best_s;   // exact score of the user's best target (MRR), may be NULL
  const int32_t* best_id;
  const float* ua;       // a_u, a2_u per user row and the largest b_i per item tile: the margin
  const float* ub2;      // of a pair is a_u b_i + a2_u (k_tc_split)
  const float* tile_bmax;
  float* cand_s;         // [n_users_pad][n_slots][KL]
  int32_t* cand_i;
  int32_t* cnt;          // [n_users_pad][n_slots] items surely ranked before the best target
  int32_t* bcnt;         // [n_users_pad] items in the band around it (may exceed TC_BCAP)
  int32_t* band_i;       // [n_users_pad][TC_BCAP] their positions, appended through bcnt
  unsigned long long* dbg;   // experiments (debug == 4): [0] list insertions, [1] insert-loop rounds
};

// The (user tile, item tile) pairs are numbered m * n_tiles + n and cut into equal contiguous
// chunks, one per CTA (about one CTA per SM, whatever the tile counts): a CTA walks its chunk as
// one or more segments, each inside a single user tile. Slot of a segment inside its user tile:
__device__ __forceinline__ int tc_slot(int64_t lin_start, int m, int n_tiles, int64_t chunk) {
  return (int)(lin_start / chunk - ((int64_t)m * n_tiles) / chunk);
}

// EW epilogue warps (8 or 16): EW / 4 per scheduler, each takes 1 / (EW / 4) of a tile's columns.
// The epilogue is bound by issue latency, not by instruction count (two instructions per score on
// two different pipes): four warps per scheduler keep both pipes fed where two leave them idle
// more than half of the time. Sixteen warps leave 96 registers per thread: lists up to KL = 16.
// DBG: the experiment switches of cmf_set_tuning key 3 and the insert counters (kept out of the
// shipped instantiations: the counter updates alone were 5 % of the kernel's instructions).
template <int KL, int EW, bool DBG>   // KL = per-row list length kept in registers, > k
__global__ void __launch_bounds__(128 + 32 * EW, 1)
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
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull + 8 * i, 1); mbar_init(bar_tempty + 8 * i, EW); }
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
    constexpr int NSUB = EW / 4;          // column parts of a tile, one per warp of a quarter
    constexpr int SUBW = TC_BN / NSUB;    // columns per part (128 or 64)
    const int sub = (warp - 4) >> 2;      // which part of every tile's columns
    const int row = q * 32 + lane;        // user row inside the tile
    int t = 0;
    for (int64_t w = w_lo; w < w_hi;) {
      const int m = (int)(w / p.n_tiles);
      const int64_t seg_hi = min(w_hi, (int64_t)(m + 1) * p.n_tiles);
      const int slot = tc_slot(w, m, p.n_tiles, p.chunk) * NSUB + sub;
      const int urow = m * TC_BM + row;
      const bool valid = urow < p.n_users;
      // the row's best KL so far, sorted (score desc, id asc), in registers; thr = the last one
      float ls[KL];
      int li[KL];
#pragma unroll
      for (int e = 0; e < KL; ++e) { ls[e] = -INFINITY; li[e] = -1; }
      float thr = -INFINITY;
      // MRR: x > t_hi is surely ranked before the best target, x < t_lo surely after it; what
      // lies between cannot be told from it with these scores and is re-scored by k_tc_finish
      const bool mrr = p.want_mrr && valid && p.best_id[urow] >= 0;
      // The accumulator is an upper bound B >= S. B < bs: surely not before the target. B beyond
      // bs + twice the largest margin of the tile's columns: surely before it.
      float t_hi = INFINITY, t_lo = INFINITY;   // (no target: nothing before, nothing in the band)
      float m_ua = 0.f, m_ub2 = 0.f, m_bs = 0.f;
      if (mrr) {
        m_bs = p.best_s[urow];
        m_ua = p.ua[urow];
        m_ub2 = p.ub2[urow];
        t_lo = m_bs;
      }
      int before = 0;
      int32_t* band = p.band_i + (int64_t)urow * TC_BCAP;
      unsigned dbg_ins = 0, dbg_rounds = 0;

      for (; w < seg_hi; ++w, ++t) {
        const int acc = t & 1;
        const int64_t n0 = (w - (int64_t)m * p.n_tiles) * TC_BN;
        mbar_wait(bar_tfull + 8 * acc, (t >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) +
                               (uint32_t)(acc * TC_BN + sub * SUBW);
        const int ncols = (int)min((int64_t)TC_BN, p.n_items - n0);   // < TC_BN only in the last tile
        const int gid0 = (int)n0;   // lists and band hold POSITIONS in the split copy; k_tc_finish maps them to ids
        if (mrr) t_hi = m_bs + 2.002f * fmaf(m_ua, p.tile_bmax[n0 / TC_BN], m_ub2);
        // 32 columns at a time go to registers, the next 32 already on their way from TMEM while
        // these are ranked; the accumulator is handed back as soon as the last of them has been
        // read, before they are ranked
        auto rank32 = [&](const float (&v)[32], int g) {
          if (DBG && p.debug == 1) return;
          const int c0 = sub * SUBW + g * 32;
          // branch-free pass: bit j of each mask = predicate of column c0 + j. x > y <=> the sign
          // bit of (y - x) (no flush-to-zero here, so a nonzero difference keeps its sign; equal
          // values give +0), and one funnel shift moves that bit into the mask: two instructions
          // per column and mask (one on the fma pipe, one on the alu pipe), four independent chains
          uint32_t mi[4] = {0, 0, 0, 0};
#pragma unroll
          for (int j = 7; j >= 0; --j) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
              mi[c] = __funnelshift_l(__float_as_uint(sub_on_fma_pipe(thr, v[8 * c + j])), mi[c], 1);
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
                const float x = v[8 * c + j];
                mg[c] = __funnelshift_l(__float_as_uint(sub_on_fma_pipe(t_hi, x)), mg[c], 1);   // x > t_hi
                ml[c] = __funnelshift_l(__float_as_uint(sub_on_fma_pipe(x, t_lo)), ml[c], 1);   // x < t_lo
              }
            }
            const uint32_t m_gt = mg[0] | (mg[1] << 8) | (mg[2] << 16) | (mg[3] << 24);
            const uint32_t m_lt = ml[0] | (ml[1] << 8) | (ml[2] << 16) | (ml[3] << 24);
            before += __popc(m_gt & m_rng);
            uint32_t m_band = mrr ? (~(m_gt | m_lt) & m_rng) : 0u;
            while (m_band) {          // rare: a handful of items per user
              const int j = __ffs(m_band) - 1;
              m_band &= m_band - 1;
              const int at = atomicAdd(p.bcnt + urow, 1);   // (the user's segments share the buffer)
              if (at < TC_BCAP) band[at] = gid0 + c0 + j;
            }
          }
          m_ins &= m_rng;
          if (DBG && p.debug == 2) m_ins &= 1u;   // experiments: (almost) no inserts
          // lanes insert side by side, each its own next candidate (column j differs per lane)
          while (__any_sync(0xffffffffu, m_ins != 0)) {
            if (DBG) ++dbg_rounds;
            if (m_ins) {
              const int j = __ffs(m_ins) - 1;
              m_ins &= m_ins - 1;
              const float sc = pick32(v, j);
              if (sc > thr) {     // thr may have risen within the chunk
                // sorted insert; an equal score stays behind the entries already there. The list
                // is non-increasing, so "sc beats entry e" is monotone in e and the new entry e is
                // old e - 1 where sc beats e - 1, sc where it beats e only, old e otherwise: every
                // entry on its own, no chain through the list (a carried insert runs at the
                // latency of 3 * KL dependent instructions with two warps per scheduler)
                const int ci = gid0 + c0 + j;
                bool beats[KL];
#pragma unroll
                for (int e = 0; e < KL; ++e) beats[e] = sc > ls[e];
#pragma unroll
                for (int e = KL - 1; e >= 1; --e) {
                  ls[e] = beats[e - 1] ? ls[e - 1] : (beats[e] ? sc : ls[e]);
                  li[e] = beats[e - 1] ? li[e - 1] : (beats[e] ? ci : li[e]);
                }
                ls[0] = beats[0] ? sc : ls[0];
                li[0] = beats[0] ? ci : li[0];
                thr = ls[KL - 1];
                if (DBG) ++dbg_ins;
              }
            }
          }
        };
        constexpr int NG = SUBW / 32;
        static_assert(NG % 2 == 0, "chunks are processed in pairs");
        auto release_acc = [&]() {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
        };
        float va[32], vb[32];
        tmem_ld32_nowait(taddr, va);
#pragma unroll 1
        for (int g = 0; g < NG; g += 2) {
          tmem_wait_ld(va);
          tmem_ld32_nowait(taddr + (g + 1) * 32, vb);
          rank32(va, g);
          tmem_wait_ld(vb);
          if (g + 2 < NG) tmem_ld32_nowait(taddr + (g + 2) * 32, va);
          else release_acc();
          rank32(vb, g + 1);
        }
      }
      if (DBG && p.debug == 4) {
        atomicAdd(p.dbg + 0, (unsigned long long)dbg_ins);
        if (lane == 0) atomicAdd(p.dbg + 1, (unsigned long long)dbg_rounds);
      }
      if (valid) {
        float* cs = p.cand_s + ((int64_t)urow * p.n_slots + slot) * KL;
        int32_t* ci = p.cand_i + ((int64_t)urow * p.n_slots + slot) * KL;
#pragma unroll
        for (int j = 0; j < KL; ++j) { cs[j] = ls[j]; ci[j] = li[j]; }
        if (p.want_mrr) {
          p.cnt[(int64_t)urow * p.n_slots + slot] = before;
        }
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

// ---------------------------------------------------------------------------------- finish
// 64-bit ranking key: larger = ranked first (score descending, item id ascending)
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
__device__ __forceinline__ int tc_key_id(uint64_t key) { return 0x7fffffff - (int)(uint32_t)key; }
// warp-wide maximum of 64-bit keys: two REDUX instructions (the high words, then the low words of
// the lanes that hold that high word) instead of five shuffle levels of 64-bit compares
__device__ __forceinline__ uint64_t warp_max_u64(uint64_t v) {
  const uint32_t hi = (uint32_t)(v >> 32);
  const uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
  const uint32_t ml = __reduce_max_sync(0xffffffffu, hi == mh ? (uint32_t)v : 0u);
  return ((uint64_t)mh << 32) | ml;
}
// six keys in descending order (12 compare-exchanges, the optimal network for six inputs)
__device__ __forceinline__ void sort6_desc(uint64_t (&a)[6]) {
#define CMF_CE(i, j)                                   \
  do {                                                 \
    const uint64_t x = a[i], y = a[j];                 \
    const bool sw = x < y;                             \
    a[i] = sw ? y : x;                                 \
    a[j] = sw ? x : y;                                 \
  } while (0)
  CMF_CE(0, 5); CMF_CE(1, 3); CMF_CE(2, 4);
  CMF_CE(1, 2); CMF_CE(3, 4);
  CMF_CE(0, 3); CMF_CE(2, 5);
  CMF_CE(0, 1); CMF_CE(2, 3); CMF_CE(4, 5);
  CMF_CE(1, 2); CMF_CE(3, 4);
#undef CMF_CE
}
// the lane whose head is `best` drops it (distinct keys: item ids differ)
__device__ __forceinline__ void pop6_if(uint64_t (&a)[6], uint64_t best) {
  if (a[0] == best) {
#pragma unroll
    for (int c = 0; c < 5; ++c) a[c] = a[c + 1];
    a[5] = 0;
  }
}

struct TcFinish {
  const float* cand_s;
  const int32_t* cand_i;
  const int32_t *cnt, *bcnt, *band_i;
  int n_slots, KL, k, D, want_mrr;
  int64_t n_users;
  const float *ua, *ub2;     // a_u, a2_u per user row; b_i per item of the shard (k_tc_split)
  const float* item_b;       // ... by position in the split copy
  const int32_t* perm;       // position -> row of the shard (NULL: identity)
  int64_t item_offset;
  const float *U, *I, *bu, *bi;
  const int32_t* user_ids;
  const int64_t* tgt_indptr;
  const int32_t* tgt_items;
  const float* best_s;
  const int32_t* best_id;
  int32_t* topk_ids;
  float* topk_scores;
  double* ap;
  int32_t* n_before_best;
  int32_t* fb_list;     // users handed to the exact path (rows of the tile) ...
  int32_t* fb_uid;      // ... and their user ids
  int32_t* status;      // [0] their number, [1] how many did not fit into fb_list
  int fb_cap;
  // optional: the rows that did not fit, for the caller to re-rank (row_base + row of the tile,
  // appended at the cursor caller_status[2])
  int32_t* overflow_rows;
  int32_t* caller_status;
  int64_t row_base;
  int overflow_cap;
};

constexpr int TC_LANE_CAND = 6;   // exact re-scorings one lane can hold; more -> the exact path
static_assert(TC_LANE_CAND == 6, "sort6_desc");
constexpr int TC_EPL = 6;         // list entries per lane kept in registers (192 per user)

// One warp per user. (1) A_k = the k-th best key of the union of the slots' lists; T = the best
// score any item OUTSIDE the lists can have (the largest last entry of a full list). (2) If
// T < A_k - 2 E, every item of the true top-k is inside the lists and has S~ >= A_k - 2 E: those
// entries are re-scored exactly and ranked; otherwise the user goes to the exact path. (3) MRR:
// the slots' sure counts plus the exact verdict on every band item.
__global__ void __launch_bounds__(256)
k_tc_finish(const TcFinish p) {
  const int lane = threadIdx.x & 31;
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= p.n_users) return;
  const int n_ent = p.n_slots * p.KL;
  const float* cs = p.cand_s + r * n_ent;
  const int32_t* ci = p.cand_i + r * n_ent;
  const float ua = p.ua[r], ub2 = p.ub2[r];
  // entry e of the walk = rank j of slot s with e = j * n_slots + s: the heads of the lists,
  // where the candidates are, spread over the lanes. s = the pair's UPPER bound B (what the
  // accumulator held); lo = B minus twice the pair's own margin, a LOWER bound of the exact score.
  auto entry = [&](int e, float& s, float& lo, int& id) {
    const int j = e / p.n_slots, sl = e - j * p.n_slots;
    s = cs[sl * p.KL + j];
    const int pos = ci[sl * p.KL + j];
    id = pos >= 0 ? (int)(p.item_offset + (p.perm ? p.perm[pos] : pos)) : -1;
    lo = pos >= 0 ? s - 2.002f * fmaf(ua, p.item_b[pos], ub2) : -INFINITY;
  };
  bool fallback = false;
  // the lists are read ONCE: up to TC_EPL entries per lane stay in registers (the k selection
  // rounds below re-read them from memory before: ten round trips per user)
  const bool in_regs = n_ent <= 32 * TC_EPL;
  uint64_t lk[TC_EPL];    // ranking keys by lower bound
  float ub_s[TC_EPL];     // upper bounds
#pragma unroll
  for (int c = 0; c < TC_EPL; ++c) {
    lk[c] = 0;
    ub_s[c] = -INFINITY;
    const int e = lane + 32 * c;
    if (in_regs && e < n_ent) {
      float s, lo; int id;
      entry(e, s, lo, id);
      lk[c] = tc_key(lo, id);
      ub_s[c] = s;
    }
  }
  // ---- (1) the k-th largest lower bound: at least k items have an exact score >= it
  uint64_t prev = ~0ull;
  int found = 0;
  uint64_t sk[TC_EPL];    // this lane's keys in descending order; the head is its candidate
#pragma unroll
  for (int c = 0; c < TC_EPL; ++c) sk[c] = lk[c];
  if (in_regs) sort6_desc(sk);
  for (int round = 0; round < p.k; ++round) {
    uint64_t best = 0;
    if (in_regs) {
      best = sk[0];
    } else {
      for (int e = lane; e < n_ent; e += 32) {
        float s, lo; int id;
        entry(e, s, lo, id);
        const uint64_t key = tc_key(lo, id);
        if (key < prev && key > best) best = key;
      }
    }
    best = warp_max_u64(best);
    if (best == 0) break;
    if (in_regs) pop6_if(sk, best);
    prev = best;
    ++found;
  }
  // everything outside the lists was pushed out of a full list: its upper bound lies at or below
  // that list's last entry (lists that never filled have lost nothing)
  float T = -INFINITY;
  for (int sl = lane; sl < p.n_slots; sl += 32)
    if (ci[sl * p.KL + p.KL - 1] >= 0) T = fmaxf(T, cs[sl * p.KL + p.KL - 1]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) T = fmaxf(T, __shfl_xor_sync(0xffffffffu, T, o));
  // an item can be in the exact top-k only if its upper bound reaches the k-th lower bound
  const float thresh = found == p.k ? tc_key_score(prev) : -INFINITY;
  if (!(T < thresh) && T != -INFINITY) fallback = true;
  // ---- MRR ingredients (before anything is written: a band that overflowed also falls back)
  const bool mrr = p.want_mrr != 0;
  const int bid = mrr ? p.best_id[r] : -1;
  const float bs = mrr && bid >= 0 ? p.best_s[r] : 0.f;
  int sure = 0;
  if (mrr && bid >= 0) {
    for (int sl = lane; sl < p.n_slots; sl += 32) sure += p.cnt[r * p.n_slots + sl];
    if (p.bcnt[r] > TC_BCAP) fallback = true;
  }
  // ---- (2) this lane's candidates
  const int64_t u = p.user_ids[r];
  const float* urow = p.U + u * p.D;
  const float ub = p.bu[u];
  uint64_t ek[TC_LANE_CAND];
  int n_mine = 0;
#pragma unroll
  for (int c = 0; c < TC_LANE_CAND; ++c) ek[c] = 0;
  auto rescore = [&](float s, int id) {
    if (id < 0 || !(s >= thresh) || fallback) return;
    if (n_mine == TC_LANE_CAND) { fallback = true; return; }
    const float ex = tc_exact_score(urow, p.I + (int64_t)id * p.D, p.D, ub, p.bi[id]);
    const uint64_t key = tc_key(ex, id);
#pragma unroll
    for (int c = 0; c < TC_LANE_CAND; ++c)
      if (c == n_mine) ek[c] = key;
    ++n_mine;
  };
  if (in_regs) {
#pragma unroll
    for (int c = 0; c < TC_EPL; ++c)
      if (lk[c]) rescore(ub_s[c], tc_key_id(lk[c]));
  } else {
    for (int e = lane; e < n_ent; e += 32) {
      float s, lo; int id;
      entry(e, s, lo, id);
      rescore(s, id);
    }
  }
  fallback = __any_sync(0xffffffffu, fallback);
  if (fallback) {
    if (lane == 0) {
      const int at = atomicAdd(p.status + 0, 1);
      if (at < p.fb_cap) {
        p.fb_list[at] = (int32_t)r;
        p.fb_uid[at] = (int32_t)p.user_ids[r];
      } else {
        atomicAdd(p.status + 1, 1);
        if (p.overflow_rows) {
          const int idx = atomicAdd(p.caller_status + 2, 1);
          if (idx < p.overflow_cap) p.overflow_rows[idx] = (int32_t)(p.row_base + r);
        }
      }
    }
    return;
  }
  // ---- exact top-k of the candidates: k rounds of "best key below the previous pick"
  const int64_t t0 = p.tgt_indptr[r];
  const int n_tgt = (int)(p.tgt_indptr[r + 1] - t0);
  uint64_t my_pick = 0;
  sort6_desc(ek);
  for (int round = 0; round < p.k; ++round) {
    const uint64_t best = warp_max_u64(ek[0]);
    if (round == lane) my_pick = best;
    if (best == 0) break;
    pop6_if(ek, best);
  }
  bool hit = false;
  if (lane < p.k) {
    const int id = my_pick ? tc_key_id(my_pick) : -1;
    if (p.topk_ids) p.topk_ids[r * p.k + lane] = id;
    if (p.topk_scores) p.topk_scores[r * p.k + lane] = my_pick ? tc_key_score(my_pick) : -INFINITY;
    hit = id >= 0 && csr_contains(p.tgt_items + t0, 0, n_tgt, id);
  }
  const uint32_t hits = __ballot_sync(0xffffffffu, hit);
  if (lane == 0 && p.ap) {
    double ap_acc = 0.0;
    for (int j = 0; j < p.k; ++j)
      if ((hits >> j) & 1u) ap_acc += (double)__popc(hits & ((2u << j) - 1u)) / (double)(j + 1);
    const int denom = p.k < n_tgt ? p.k : n_tgt;
    p.ap[r] = denom > 0 ? ap_acc / (double)denom : 0.0;
  }
  // ---- (3)
  if (mrr && p.n_before_best) {
    int more = 0;
    if (bid >= 0) {
      const int n_band = p.bcnt[r];
      for (int e = lane; e < n_band; e += 32) {
        const int pos = p.band_i[r * TC_BCAP + e];
        const int id = (int)(p.item_offset + (p.perm ? p.perm[pos] : pos));
        if (id == bid) continue;     // the target never precedes itself
        const float ex = tc_exact_score(urow, p.I + (int64_t)id * p.D, p.D, ub, p.bi[id]);
        more += (ex > bs || (ex == bs && id < bid)) ? 1 : 0;
      }
    }
    int tot = sure + more;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if (lane == 0) p.n_before_best[r] = bid >= 0 ? tot : -1;
  }
}

__global__ void k_tc_status_add(const int32_t* __restrict__ status, int32_t* __restrict__ out,
                                int with_dbg) {
  if (threadIdx.x < 2 && status[threadIdx.x]) atomicAdd(out + threadIdx.x, status[threadIdx.x]);
  // experiments: the caller's array is int32[8] then; [4..7] = two 64-bit counters of the last call
  if (with_dbg && threadIdx.x >= 4 && threadIdx.x < 8) out[threadIdx.x] = status[threadIdx.x];
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

int g_tc_terms = 0;   // cmf_set_tuning key 4: 0 = automatic (1 without the MRR counter, else 3)

int g_tc_epi_warps = 0;  // cmf_set_tuning key 7: 0 = automatic, 8 or 16 epilogue warps
int g_tc_list_len = 0;   // cmf_set_tuning key 5: 0 = automatic, else 8 / 16 / 24 (must exceed k)

static int tc_list_len(int k) {
  if (g_tc_list_len > k) return g_tc_list_len;
  // (a list only has to hold what ITS segment contributes to the top-k plus a margin for the
  // error bound: the union of a row's eight and more lists is what k_tc_finish works on)
  return k <= 4 ? 8 : (k <= 10 ? 16 : 24);
}

struct TcLayout {
  int Kp, KB, KL, epi_warps, m_tiles, n_tiles, n_ctas, n_slots, fb_cap;
  int64_t m_pad, chunk, total_work;
  size_t Bp, Ap, ua, ub2, item_b, tile_bmax, perm, sort_keys, sort_keys2, sort_vals, sort_temp, sort_temp_bytes, stats, best_s, best_id, cand_s, cand_i, cnt, bcnt, band_i,
      fb_list, fb_uid, status, fb_scores, total;
};

// (the layout is sized for three terms, so one workspace serves both forms)
static TcLayout tc_layout(int64_t n_users, int64_t n_items, int D, int k, int terms) {
  TcLayout L;
  const int Kp_max = (int)align_up((size_t)3 * D + TC_BIAS_SLOTS, TC_BK);
  L.Kp = (int)align_up((size_t)terms * D + TC_BIAS_SLOTS, TC_BK);
  L.KB = L.Kp / TC_BK;
  L.KL = tc_list_len(k);
  L.m_tiles = (int)((n_users + TC_BM - 1) / TC_BM);
  L.m_pad = (int64_t)L.m_tiles * TC_BM;
  L.n_tiles = (int)((n_items + TC_BN - 1) / TC_BN);
  // equal contiguous chunks of the (user tile, item tile) pairs, one per CTA, one CTA per SM
  L.total_work = (int64_t)L.m_tiles * L.n_tiles;
  int64_t ctas = num_sms();
  if (ctas > L.total_work) ctas = L.total_work;
  L.chunk = (L.total_work + ctas - 1) / ctas;
  L.n_ctas = (int)((L.total_work + L.chunk - 1) / L.chunk);
  // segments a user tile can be cut into (x the column parts the epilogue keeps a list for)
  // (measured: sixteen epilogue warps of 32-column groups are ~15 % slower than eight)
  L.epi_warps = (g_tc_epi_warps == 16 && L.KL <= 16) ? 16 : 8;
  L.n_slots = (L.epi_warps / 4) * (int)((L.n_tiles + L.chunk - 1) / L.chunk + 1);
  // users the exact path can take over per call: up to 128 MB of score rows
  int64_t cap = ((int64_t)128 << 20) / (int64_t)(sizeof(float) * (size_t)n_items);
  L.fb_cap = (int)(cap < 1 ? 1 : (cap > 1024 ? 1024 : cap));
  if (L.fb_cap > n_users) L.fb_cap = (int)n_users;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 1024); return at; };
  // item side first: what a call with reuse_item_prep finds again must not move with the number
  // of users of the tile (the last tile of an evaluation is usually smaller)
  L.Bp = take((size_t)n_items * Kp_max * 2);
  L.item_b = take(sizeof(float) * (size_t)n_items);
  L.tile_bmax = take(sizeof(float) * (size_t)L.n_tiles);
  // item order by norm (MRR form): permutation + radix-sort scratch
  L.perm = take(sizeof(int32_t) * (size_t)n_items);
  L.sort_keys = take(sizeof(uint32_t) * (size_t)n_items);
  L.sort_keys2 = take(sizeof(uint32_t) * (size_t)n_items);
  L.sort_vals = take(sizeof(int32_t) * (size_t)n_items);
  L.sort_temp_bytes = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, L.sort_temp_bytes, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (const int32_t*)nullptr, (int32_t*)nullptr,
                                  (int)n_items);
  L.sort_temp = take(L.sort_temp_bytes);
  L.stats = take(sizeof(uint32_t) * 8);     // [1] largest |item bias| of the shard
  L.Ap = take((size_t)L.m_pad * Kp_max * 2);
  L.ua = take(sizeof(float) * L.m_pad);
  L.ub2 = take(sizeof(float) * L.m_pad);
  L.best_s = take(sizeof(float) * L.m_pad);
  L.best_id = take(sizeof(int32_t) * L.m_pad);
  L.cand_s = take(sizeof(float) * L.m_pad * L.n_slots * L.KL);
  L.cand_i = take(sizeof(int32_t) * L.m_pad * L.n_slots * L.KL);
  L.cnt = take(sizeof(int32_t) * L.m_pad * L.n_slots);
  L.bcnt = take(sizeof(int32_t) * L.m_pad);
  L.band_i = take(sizeof(int32_t) * L.m_pad * TC_BCAP);
  L.fb_list = take(sizeof(int32_t) * L.fb_cap);
  L.fb_uid = take(sizeof(int32_t) * L.fb_cap);
  L.status = take(sizeof(int32_t) * 8);
  L.fb_scores = take(sizeof(float) * (size_t)L.fb_cap * n_items);
  L.total = o;
  return L;
}

bool eval_tc_supported(int D, int k) { return D >= 1 && D <= 128 && k >= 1 && k <= TC_MAXK; }

size_t eval_tc_workspace_bytes(int64_t n_users, int64_t n_items, int D, int k) {
  if (!eval_tc_supported(D, k)) return 0;
  return tc_layout(n_users, n_items, D, k, 3).total;
}

// eval.cu: k_rank_rows over score rows [n_rows][n_items] whose user is row_map[row]
int launch_rank_rows_mapped(const cmf_eval_t* a, const float* scores, float* sig_scratch,
                            const int32_t* row_map, const int32_t* row_count, int row_cap,
                            cudaStream_t s);
// eval.cu: the exact path's score kernel over `cap` user ids of which only *count exist
int launch_score_rows_counted(const cmf_eval_t* a, const int32_t* user_ids, int cap,
                              const int32_t* count, float* out, cudaStream_t s);

int eval_tile_tc(const cmf_eval_t* a, cudaStream_t s) {
  CMF_REQUIRE(!a->want_auc, "cmf_eval_tile: AUC is computed by the exact path (precision = 0)");
  if (!eval_tc_supported(a->D, a->k))
    return fail(CMF_ERR_UNSUPPORTED,
                "cmf_eval_tile: the tensor-core path takes embedding_dim <= 128 and k <= %d "
                "(got D = %d, k = %d); use precision = 0", TC_MAXK, a->D, a->k);
  const bool mrr = a->want_mrr != 0;
  CMF_REQUIRE(a->tc_terms == 0 || a->tc_terms == 1 || a->tc_terms == 3,
              "cmf_eval_tile: tc_terms must be 0, 1 or 3 (got %d)", a->tc_terms);
  const int terms = a->tc_terms ? a->tc_terms
                                : (g_tc_terms == 1 || g_tc_terms == 3 ? g_tc_terms : (mrr ? 3 : 1));
  const TcLayout L = tc_layout(a->n_users_tile, a->num_items_shard, a->D, a->k, terms);
  const size_t need = eval_tc_workspace_bytes(a->n_users_tile, a->num_items_shard, a->D, a->k);
  if (!a->workspace || a->workspace_bytes < need)
    return fail(CMF_ERR_WORKSPACE, "cmf_eval_tile: workspace of %zu bytes needed, %zu given",
                need, a->workspace ? a->workspace_bytes : (size_t)0);
  char* base = (char*)a->workspace;
  __nv_bfloat16* Bp = (__nv_bfloat16*)(base + L.Bp);
  __nv_bfloat16* Ap = (__nv_bfloat16*)(base + L.Ap);
  float* ua = (float*)(base + L.ua);
  float* ub2 = (float*)(base + L.ub2);
  float* item_b = (float*)(base + L.item_b);
  float* tile_bmax = (float*)(base + L.tile_bmax);
  uint32_t* stats = (uint32_t*)(base + L.stats);
  float* best_s = (float*)(base + L.best_s);
  int32_t* best_id = (int32_t*)(base + L.best_id);
  int32_t* status = (int32_t*)(base + L.status);
  const int grid = num_sms() * 8;
  const float c = terms == 1 ? kTcC1 : kTcC3;
  int32_t* perm = (int32_t*)(base + L.perm);
  const bool sorted = mrr;   // (a caller that reuses the item side keeps asking for the same outputs)

  // the item side of the workspace (split copy, per-item and per-tile bounds, largest |bias|) is
  // reused across calls on request
  const bool reuse = a->reuse_item_prep != 0;
  if (!reuse) {
    CMF_CUDA(cudaMemsetAsync(stats, 0, sizeof(uint32_t) * 8, s));
    if (sorted) {
      // MRR form: the split copy holds the item rows in order of their norm, so that the largest
      // margin of a tile (all the epilogue knows about a column) is close to every column's own
      uint32_t* k1 = (uint32_t*)(base + L.sort_keys);
      uint32_t* k2 = (uint32_t*)(base + L.sort_keys2);
      int32_t* v1 = (int32_t*)(base + L.sort_vals);
      k_tc_norm_keys<<<grid, 256, 0, s>>>(a->I, a->item_offset, a->num_items_shard, a->D, k1, v1);
      CMF_LAUNCH_CHECK();
      size_t temp = L.sort_temp_bytes;
      CMF_CUDA(cub::DeviceRadixSort::SortPairs(base + L.sort_temp, temp, k1, k2, v1, perm,
                                               (int)a->num_items_shard, 0, 32, s));
    }
    k_tc_split<<<grid, 256, 0, s>>>(a->I, a->bi, sorted ? perm : nullptr, a->item_offset,
                                    a->num_items_shard, a->num_items_shard, a->D, L.Kp, 0, terms,
                                    c, Bp, item_b, nullptr, stats);
    CMF_LAUNCH_CHECK();
    k_tc_tile_max<<<(unsigned)L.n_tiles, TC_BN, 0, s>>>(item_b, a->num_items_shard, tile_bmax);
    CMF_LAUNCH_CHECK();
  }
  k_tc_split<<<grid, 256, 0, s>>>(a->U, a->bu, a->user_ids, 0, a->n_users_tile, L.m_pad, a->D,
                                  L.Kp, 1, terms, c, Ap, ua, ub2, stats);
  CMF_LAUNCH_CHECK();
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
  p.last_steps = (terms * a->D + TC_BIAS_SLOTS - (L.KB - 1) * TC_BK + 15) / 16;
  p.m_tiles = L.m_tiles; p.n_slots = L.n_slots; p.chunk = L.chunk; p.total_work = L.total_work;
  p.want_mrr = mrr ? 1 : 0;
  p.n_items = a->num_items_shard; p.item_offset = a->item_offset; p.debug = g_tc_debug;
  p.best_s = best_s; p.best_id = best_id; p.ua = ua; p.ub2 = ub2; p.tile_bmax = tile_bmax;
  p.dbg = (unsigned long long*)(status + 4);     // (status[4..7]; cleared with it)
  p.cand_s = (float*)(base + L.cand_s); p.cand_i = (int32_t*)(base + L.cand_i);
  p.cnt = (int32_t*)(base + L.cnt); p.bcnt = (int32_t*)(base + L.bcnt);
  p.band_i = (int32_t*)(base + L.band_i);

  const size_t smem = (size_t)L.KB * TC_A_BYTES + (size_t)TC_STAGES * TC_B_BYTES +
                      (2 * TC_STAGES + 6) * sizeof(uint64_t) + 64 + 1024;
  // slots a user tile does not use stay empty
  CMF_CUDA(cudaMemsetAsync(p.cand_i, 0xFF, sizeof(int32_t) * L.m_pad * L.n_slots * L.KL, s));
  if (mrr) {
    CMF_CUDA(cudaMemsetAsync(p.cnt, 0, sizeof(int32_t) * L.m_pad * L.n_slots, s));
    CMF_CUDA(cudaMemsetAsync(p.bcnt, 0, sizeof(int32_t) * L.m_pad, s));
  }
  CMF_CUDA(cudaMemsetAsync(status, 0, sizeof(int32_t) * 8, s));
  // list length: the smallest instantiation that leaves a margin above k
#define CMF_TC_LAUNCH(KL, EW)                                                                    \
  do {                                                                                           \
    CMF_CUDA(cudaFuncSetAttribute(k_tc_gemm_topk<KL, EW, false>,                                 \
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));      \
    k_tc_gemm_topk<KL, EW, false><<<(unsigned)L.n_ctas, 128 + 32 * EW, smem, s>>>(tmA, tmB, p);  \
  } while (0)
  if (g_tc_debug != 0) {   // experiments: one instantiation carries the switches and counters
    CMF_REQUIRE(L.KL == 16 && L.epi_warps == 8,
                "cmf_set_tuning(3, ...): the debug kernel exists for list length 16, 8 epilogue warps");
    CMF_CUDA(cudaFuncSetAttribute(k_tc_gemm_topk<16, 8, true>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_tc_gemm_topk<16, 8, true><<<(unsigned)L.n_ctas, 128 + 32 * 8, smem, s>>>(tmA, tmB, p);
  } else if (L.KL == 8 && L.epi_warps == 16) CMF_TC_LAUNCH(8, 16);
  else if (L.KL == 8) CMF_TC_LAUNCH(8, 8);
  else if (L.KL == 12) CMF_TC_LAUNCH(12, 8);
  else if (L.KL == 16 && L.epi_warps == 16) CMF_TC_LAUNCH(16, 16);
  else if (L.KL == 16) CMF_TC_LAUNCH(16, 8);
  else CMF_TC_LAUNCH(24, 8);
#undef CMF_TC_LAUNCH
  CMF_LAUNCH_CHECK();

  TcFinish f{};
  f.cand_s = p.cand_s; f.cand_i = p.cand_i; f.cnt = p.cnt; f.bcnt = p.bcnt; f.band_i = p.band_i;
  f.n_slots = L.n_slots; f.KL = L.KL; f.k = a->k; f.D = a->D; f.want_mrr = mrr ? 1 : 0;
  f.n_users = a->n_users_tile; f.ua = ua; f.ub2 = ub2; f.item_b = item_b;
  f.item_offset = a->item_offset; f.perm = sorted ? perm : nullptr;
  f.U = a->U; f.I = a->I; f.bu = a->bu; f.bi = a->bi; f.user_ids = a->user_ids;
  f.tgt_indptr = a->tgt_indptr; f.tgt_items = a->tgt_items;
  f.best_s = best_s; f.best_id = best_id;
  f.topk_ids = a->topk_ids; f.topk_scores = a->topk_scores; f.ap = a->ap;
  f.n_before_best = a->n_before_best;
  f.fb_list = (int32_t*)(base + L.fb_list); f.fb_uid = (int32_t*)(base + L.fb_uid);
  f.status = status; f.fb_cap = L.fb_cap;
  if (a->tc_status && a->tc_overflow_rows && a->tc_overflow_cap > 0) {
    f.overflow_rows = a->tc_overflow_rows; f.caller_status = a->tc_status;
    f.row_base = a->tc_row_base; f.overflow_cap = a->tc_overflow_cap;
  }
  k_tc_finish<<<(unsigned)((a->n_users_tile * 32 + 255) / 256), 256, 0, s>>>(f);
  CMF_LAUNCH_CHECK();

  // the users the bound could not settle: exact scores of their rows, ranked by the exact path's
  // own kernel (both grids are sized for fb_cap users and return at once when there are none)
  float* fb_scores = (float*)(base + L.fb_scores);
  rc = launch_score_rows_counted(a, f.fb_uid, L.fb_cap, status, fb_scores, s);
  if (rc) return rc;
  cmf_eval_t b = *a;
  if (mrr) { b.ext_best_score = best_s; b.ext_best_id = best_id; b.best_score = nullptr; b.best_id = nullptr; }
  rc = launch_rank_rows_mapped(&b, fb_scores, nullptr /* (AUC scratch: not on this path) */, f.fb_list, status,
                               L.fb_cap, s);
  if (rc) return rc;
  if (a->tc_status) {
    k_tc_status_add<<<1, 32, 0, s>>>(status, a->tc_status, g_tc_debug == 4);
    CMF_LAUNCH_CHECK();
  }
  return CMF_OK;
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_eval_tc_info(int32_t D, double* out /* [4] */) {
  CMF_REQUIRE(out && D >= 1, "cmf_eval_tc_info: bad arguments");
  out[0] = (double)(((D + TC_BIAS_SLOTS) + 15) / 16 * 16);        // K extent issued per output, 1 term
  out[1] = (double)(((3 * D + TC_BIAS_SLOTS) + 15) / 16 * 16);    // ... 3 terms
  out[2] = (double)kTcC1;
  out[3] = (double)kTcC3;
  return CMF_OK;
}
