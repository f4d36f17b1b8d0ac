// Definitions shared by the training-step kernels (train_step.cu, train_step_v2.cu).
#pragma once

#include <cmath>

#include "common.cuh"

namespace cmf {

constexpr int kMaxNeg = 64;          // negatives per interaction supported by the fused step
constexpr int kMaxPartials = 8192;   // upper bound on the fused kernel's grid size
constexpr int kThreads = 256;

struct WsHeader {
  int item_count;
  int user_count;
  unsigned blocks_done;
  int pad[61];
};
static_assert(sizeof(WsHeader) == 256, "header is one 256-byte line group");

// ------------------------------------------------------------------------------- optimizer rules
struct OptDev {
  int rule;
  int dense;        // every row of the table moves each step
  float lr, wd, eps;
  float b2;         // Adam: beta2
  float omb1, omb2; // 1 - beta
  float step_size;  // SparseAdam: lr*sqrt(1-b2^t)/(1-b1^t); Adam: lr/(1-b1^t)
  float bc2_sqrt;   // Adam: sqrt(1-b2^t)
};

inline OptDev make_opt(const cmf_opt_t& o) {
  OptDev d{};
  d.rule = o.rule;
  d.lr = (float)o.lr;
  d.wd = (float)o.weight_decay;
  d.eps = (float)(o.rule == CMF_OPT_ADAGRAD ? 1e-10 : o.eps);
  d.b2 = (float)o.beta2;
  d.omb1 = (float)(1.0 - o.beta1);
  d.omb2 = (float)(1.0 - o.beta2);
  const double t = (double)(o.step < 1 ? 1 : o.step);
  const double bc1 = 1.0 - std::pow(o.beta1, t), bc2 = 1.0 - std::pow(o.beta2, t);
  if (o.rule == CMF_OPT_SPARSE_ADAM) d.step_size = (float)(o.lr * std::sqrt(bc2) / bc1);
  if (o.rule == CMF_OPT_ADAM) { d.step_size = (float)(o.lr / bc1); d.bc2_sqrt = (float)std::sqrt(bc2); }
  d.dense = (o.rule == CMF_OPT_ADAM) || (o.weight_decay != 0.0);
  return d;
}

inline int validate_opt(const cmf_opt_t& o, const char* which, const float* m, const float* v) {
  CMF_REQUIRE(o.rule >= CMF_OPT_SPARSE_ADAM && o.rule <= CMF_OPT_ADAGRAD,
              "cmf_train_step: unknown %s optimizer rule %d", which, o.rule);
  const bool need_m = o.rule != CMF_OPT_SGD;
  const bool need_v = o.rule == CMF_OPT_SPARSE_ADAM || o.rule == CMF_OPT_ADAM;
  CMF_REQUIRE(!need_m || m, "cmf_train_step: %s optimizer needs its first state tensor", which);
  CMF_REQUIRE(!need_v || v, "cmf_train_step: %s optimizer needs its second state tensor", which);
  CMF_REQUIRE(o.step >= 1, "cmf_train_step: %s optimizer step must be >= 1", which);
  return CMF_OK;
}

// sqrt / division of the update term with the SFU approximations (<= 2 ulp each): the term is
// multiplied by lr before it touches the parameter, so the parameter moves by < 1e-9 relative.
// The bare MUFU forms (no denormal rescaling around them): arguments are second moments or sums
// of squares; below 1e-30 their root (< 1e-15) vanishes next to eps and is taken as zero, and the
// divisors are >= eps. For normal arguments these are the bits rsqrtf / __fdividef return.
__device__ __forceinline__ float rsqrt_ftz(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_ftz(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_sqrt(float x) { return x > 1e-30f ? x * rsqrt_ftz(x) : 0.f; }
__device__ __forceinline__ float fast_div(float a, float b) { return a * rcp_ftz(b); }

// N parameter elements. m/v are the rule's state (Adagrad keeps its sum in m). The rule is chosen
// ONCE per call, outside the element loop (per element, the four-way branch and the library
// forms of rsqrt / division made the optimizer arithmetic 70 % of k_bwd_update's instructions).
template <int N>
__device__ __forceinline__ void apply_rule_n(const OptDev& o, float* p, float* m, float* v,
                                             const float* g_in) {
  switch (o.rule) {
    case CMF_OPT_SPARSE_ADAM: {  // torch/optim/sparse_adam.py (functional sparse_adam)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const float g = g_in[i];
        const float mn = m[i] + (g - m[i]) * o.omb1;
        const float vn = v[i] + (g * g - v[i]) * o.omb2;
        m[i] = mn;
        v[i] = vn;
        p[i] = p[i] - o.step_size * fast_div(mn, fast_sqrt(vn) + o.eps);
      }
      break;
    }
    case CMF_OPT_ADAM: {  // torch/optim/adam.py _single_tensor_adam
#pragma unroll
      for (int i = 0; i < N; ++i) {
        float g = g_in[i];
        if (o.wd != 0.f) g = g + o.wd * p[i];
        m[i] = m[i] + o.omb1 * (g - m[i]);
        v[i] = v[i] * o.b2 + (o.omb2 * g) * g;
        const float denom = fast_div(fast_sqrt(v[i]), o.bc2_sqrt) + o.eps;
        p[i] = p[i] - o.step_size * fast_div(m[i], denom);
      }
      break;
    }
    case CMF_OPT_ADAGRAD: {  // torch/optim/adagrad.py _single_tensor_adagrad
#pragma unroll
      for (int i = 0; i < N; ++i) {
        float g = g_in[i];
        if (o.wd != 0.f) g = g + o.wd * p[i];
        m[i] = m[i] + g * g;
        p[i] = p[i] - o.lr * fast_div(g, fast_sqrt(m[i]) + o.eps);
      }
      break;
    }
    default: {  // SGD
#pragma unroll
      for (int i = 0; i < N; ++i) {
        float g = g_in[i];
        if (o.wd != 0.f) g = g + o.wd * p[i];
        p[i] = p[i] - o.lr * g;
      }
      break;
    }
  }
}
__device__ __forceinline__ void apply_rule(const OptDev& o, float& p, float& m, float& v, float g) {
  apply_rule_n<1>(o, &p, &m, &v, &g);
}

__device__ __forceinline__ bool rule_has_m(int rule) { return rule != CMF_OPT_SGD; }
__device__ __forceinline__ bool rule_has_v(int rule) {
  return rule == CMF_OPT_SPARSE_ADAM || rule == CMF_OPT_ADAM;
}

template <int V, int G, int R>
__device__ __forceinline__ void update_row(const OptDev& o, float* __restrict__ P,
                                           float* __restrict__ M, float* __restrict__ Vv, int D, int l,
                                           int64_t row, const Row<V, G, R>& g) {
  Row<V, G, R> p, m, v;
  p.load(P + row * D, D, l);
  if (rule_has_m(o.rule)) m.load(M + row * D, D, l); else m.zero();
  if (rule_has_v(o.rule)) v.load(Vv + row * D, D, l); else v.zero();
  apply_rule_n<R * V>(o, p.f, m.f, v.f, g.f);
  p.store(P + row * D, D, l);
  if (rule_has_m(o.rule)) m.store(M + row * D, D, l);
  if (rule_has_v(o.rule)) v.store(Vv + row * D, D, l);
}

__device__ __forceinline__ void update_scalar(const OptDev& o, float* P, float* M, float* Vv,
                                              int64_t row, float g) {
  float p = P[row];
  float m = rule_has_m(o.rule) ? M[row] : 0.f;
  float v = rule_has_v(o.rule) ? Vv[row] : 0.f;
  apply_rule(o, p, m, v, g);
  P[row] = p;
  if (rule_has_m(o.rule)) M[row] = m;
  if (rule_has_v(o.rule)) Vv[row] = v;
}

// ------------------------------------------------------------------------------- fused step
struct StepDev {
  float *U, *I, *bu, *bi, *U_m, *U_v, *bu_m, *bu_v;
  int D;
  uint32_t num_items;
  const int32_t* users; // users in batch order
  const int32_t* su;    // users in sorted order
  const uint32_t* perm; // sorted position -> batch position (NULL: identity)
  const int32_t* items;
  const float* ratings;
  const int64_t* negs_in;
  const int64_t* gidx;
  int64_t gidx_base;
  int64_t B;
  int N, L;
  int64_t num_chunks;
  uint64_t seed;
  uint32_t epoch;
  const int64_t* indptr;
  const int32_t* csr_items;
  int max_tries;
  int loss, has_range;
  float y_lo, y_hi;
  int n_meta;
  const int32_t* meta[CMF_MAX_METADATA];
  float meta_w[CMF_MAX_METADATA];
  OptDev emb, bias;
  int train, user_inplace, use_lists;
  float inv_B;
  double loss_div;  // the loss is a mean over the (global) batch
  float *GU, *gbu, *GI, *gbi;
  int *user_flag, *item_flag, *user_list, *item_list;
  WsHeader* hdr;
  double* partials;
  float* loss_out;
  int64_t* negs_out;
  int32_t* sel_out;
  float* pos_out;
  float* neg_out;
  float* g_arr;     // [B] d loss / d positive score (split path: k_fwd_loss -> k_bwd_update)
  int32_t* sel_arr; // [B] item id of the selected negative, -1 when none
  // dropout on both embeddings of every scored pair (matrix_factorization.py:118, :147): keep an
  // element when its Philox word >= drop_thr, scaled by drop_scale = 1 / (1 - p); 0 = off
  uint32_t drop_thr;
  float drop_scale;
  uint32_t drop_key_hi;   // optimizer step: a batch that is repeated draws fresh masks
  int red_in_fwd;         // the forward kernel sends the item-row gradient contributions itself
  int bwd_skip_item;      // k_bwd_update leaves the item side alone (red_in_fwd, or the deterministic reduce)
  // deterministic step: chunk c of k_bwd_update = [chunk_starts[c], chunk_starts[c + 1]), cut at
  // user boundaries, so every user is summed by one warp in batch order (NULL: chunks of L)
  const int32_t* chunk_starts;
};

// The dropout mask of one row of one scored pair: interaction at dataset position gidx, slot 0 =
// the positive / 1 + n = negative n, side 0 = user row / 1 = item row. Element d takes word d % 4
// of the Philox block (gidx, epoch, slot, side, d / 4); see oracle.dropout_mask.
__device__ __forceinline__ float drop_factor(const StepDev& p, int64_t gidx, int slot, int side,
                                             int d) {
  const uint4 w = philox4x32_10(
      make_uint4((uint32_t)gidx, (uint32_t)((uint64_t)gidx >> 32), p.epoch,
                 ((uint32_t)slot << 16) | ((uint32_t)side << 15) | (uint32_t)(d >> 2)),
      make_uint2((uint32_t)p.seed ^ 0x85EBCA6Bu, (uint32_t)(p.seed >> 32) ^ p.drop_key_hi));
  const uint32_t word = (d & 3) == 0 ? w.x : (d & 3) == 1 ? w.y : (d & 3) == 2 ? w.z : w.w;
  return word >= p.drop_thr ? p.drop_scale : 0.f;
}
template <int V, int G, int R>
__device__ __forceinline__ void drop_row(const StepDev& p, int64_t gidx, int slot, int side, int D,
                                         int l, Row<V, G, R>& x) {
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int d0 = (r * G + l) * V;
    if (d0 >= D) continue;
    if (V == 4) {
      const uint4 w = philox4x32_10(
          make_uint4((uint32_t)gidx, (uint32_t)((uint64_t)gidx >> 32), p.epoch,
                     ((uint32_t)slot << 16) | ((uint32_t)side << 15) | (uint32_t)(d0 >> 2)),
          make_uint2((uint32_t)p.seed ^ 0x85EBCA6Bu, (uint32_t)(p.seed >> 32) ^ p.drop_key_hi));
      x.f[r * V + 0] *= w.x >= p.drop_thr ? p.drop_scale : 0.f;
      x.f[r * V + 1] *= w.y >= p.drop_thr ? p.drop_scale : 0.f;
      x.f[r * V + 2] *= w.z >= p.drop_thr ? p.drop_scale : 0.f;
      x.f[r * V + 3] *= w.w >= p.drop_thr ? p.drop_scale : 0.f;
    } else {
#pragma unroll
      for (int j = 0; j < V; ++j) x.f[r * V + j] *= drop_factor(p, gidx, slot, side, d0 + j);
    }
  }
}
// the same two masks applied to a gradient direction: out = mask_user * mask_item * x
template <int V, int G, int R>
__device__ __forceinline__ Row<V, G, R> drop_both(const StepDev& p, int64_t gidx, int slot, int D,
                                                  int l, const Row<V, G, R>& x) {
  Row<V, G, R> y = x;
  drop_row<V, G, R>(p, gidx, slot, 0, D, l, y);
  drop_row<V, G, R>(p, gidx, slot, 1, D, l, y);
  return y;
}

__device__ __forceinline__ void mark_row(int* flag, int* list, int* count, int id, int use_list) {
  if (*(volatile int*)(flag + id) == 0) {
    if (atomicExch(flag + id, 1) == 0 && use_list) list[atomicAdd(count, 1)] = id;
  }
}

__device__ __forceinline__ float ideal_of(const StepDev& p, int pos, int neg) {
  // 1 - sum_k w_k [meta_k[pos] == meta_k[neg]]   (collie/loss/metadata_utils.py:60-71)
  float frac = 0.f;
  for (int k = 0; k < p.n_meta; ++k)
    frac += (p.meta[k][pos] == p.meta[k][neg]) ? p.meta_w[k] : 0.f;
  return 1.0f - frac;
}


// train_step_v3.cu: the split fast path (k_fwd_loss + k_bwd_update)
bool v3_supported(int D, int N);
int launch_fwd_loss_v3(const StepDev& p, bool sort_runs_alongside, cudaStream_t s);
int launch_bwd_update_v3(const StepDev& p, int max_chunk_len, cudaStream_t s,
                         int32_t* aligned_chunk_starts = nullptr);

}  // namespace cmf
