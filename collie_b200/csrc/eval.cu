// Evaluation, exact-fp32 path: tiled SIMT scoring of a user tile against the item table, then one
// CTA per user ranks its score row (top-k in canonical order, MAP@k, best-target rank for MRR,
// Mann-Whitney pair count for AUC).
//
// Reference: collie/metrics.py  get_preds :77-107, mapk :144-198, mrr :201-240, auc :243-282,
// one iteration of the evaluate_in_batches loop :379-384.
// Canonical order everywhere: score descending, item id ascending (torch.topk's tie order is
// unpinned in the reference).
#include "common.cuh"

namespace cmf {

int eval_tile_tc(const cmf_eval_t* a, cudaStream_t s);  // eval_tc.cu (tcgen05 path)
size_t eval_tc_workspace_bytes(int64_t n_users, int64_t n_items, int D, int k);

// ------------------------------------------------------------------ scores: 64x64 tile, K step 16
constexpr int TM = 64, TN = 64, TK = 16;

__global__ void __launch_bounds__(256)
k_score_tile(const float* __restrict__ U, const float* __restrict__ I, const float* __restrict__ bu,
             const float* __restrict__ bi, int D, const int32_t* __restrict__ user_ids, int64_t n_users,
             int64_t item_offset, int64_t n_items, float* __restrict__ out, int64_t ld_out,
             const int32_t* __restrict__ n_users_dev) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // 16 x 16 threads, 4x4 outputs each
  const int64_t row0 = (int64_t)blockIdx.y * TM, col0 = (int64_t)blockIdx.x * TN;
  if (n_users_dev) {   // the row count lives on the device (the tensor-core path's hand-over)
    if ((int64_t)*n_users_dev < n_users) n_users = *n_users_dev;
    if (row0 >= n_users) return;
  }
  float acc[4][4] = {};
  // loader mapping: 256 threads fetch 64 rows x 16 k-values; thread -> (row = t / 4, 4 k's)
  const int lr = threadIdx.x >> 2, lk = (threadIdx.x & 3) * 4;
  const int64_t arow = row0 + lr, brow = col0 + lr;
  const float* aptr = (arow < n_users) ? U + (int64_t)user_ids[arow] * D : nullptr;
  const float* bptr = (brow < n_items) ? I + (item_offset + brow) * D : nullptr;
  for (int k0 = 0; k0 < D; k0 += TK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int d = k0 + lk + j;
      As[lk + j][lr] = (aptr && d < D) ? aptr[d] : 0.f;
      Bs[lk + j][lr] = (bptr && d < D) ? bptr[d] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[k][ty * 4 + i]; b[i] = Bs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t r = row0 + ty * 4 + i;
    if (r >= n_users) continue;
    const float ub = bu[user_ids[r]];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t c = col0 + tx * 4 + j;
      if (c < n_items) out[r * ld_out + c] = acc[i][j] + ub + bi[item_offset + c];
    }
  }
}

// ---- the same contraction for full-size tiles: 128 x 128 outputs per CTA, 8 x 8 per thread
// (two 4-wide halves 64 apart on each axis, so shared-memory reads are conflict-free float4s), K step
// 8, global loads of the next K step in flight while the current one is multiplied. Every output is
// accumulated over d in ascending order with fmaf, exactly like k_score_tile: same bits.
constexpr int BM = 128, BN = 128, BK = 8;

__global__ void __launch_bounds__(256)
k_score_tile_128(const float* __restrict__ U, const float* __restrict__ I,
                 const float* __restrict__ bu, const float* __restrict__ bi, int D,
                 const int32_t* __restrict__ user_ids, int64_t n_users, int64_t item_offset,
                 int64_t n_items, float* __restrict__ out, int64_t ld_out,
                 const int32_t* __restrict__ n_users_dev) {
  __shared__ __align__(16) float As[2][BK][BM];
  __shared__ __align__(16) float Bs[2][BK][BN];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t row0 = (int64_t)blockIdx.y * BM, col0 = (int64_t)blockIdx.x * BN;
  if (n_users_dev) {
    if ((int64_t)*n_users_dev < n_users) n_users = *n_users_dev;
    if (row0 >= n_users) return;
  }
  // loader: thread -> (row = tid / 2, four consecutive k's)
  const int lr = tid >> 1, lk = (tid & 1) * 4;
  const int64_t arow = row0 + lr, brow = col0 + lr;
  const float* aptr = (arow < n_users) ? U + (int64_t)user_ids[arow] * D : nullptr;
  const float* bptr = (brow < n_items) ? I + (item_offset + brow) * D : nullptr;
  const bool vec = (D & 3) == 0;
  auto fetch = [&](const float* src, int k0, float (&v)[4]) {
    const int d = k0 + lk;
    if (src && vec && d + 3 < D) {
      const float4 t = *reinterpret_cast<const float4*>(src + d);
      v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = (src && d + j < D) ? src[d + j] : 0.f;
    }
  };
  float acc[8][8] = {};
  float ra[4], rb[4];
  fetch(aptr, 0, ra);
  fetch(bptr, 0, rb);
#pragma unroll
  for (int j = 0; j < 4; ++j) { As[0][lk + j][lr] = ra[j]; Bs[0][lk + j][lr] = rb[j]; }
  __syncthreads();
  int buf = 0;
  for (int k0 = 0; k0 < D; k0 += BK) {
    const bool more = k0 + BK < D;
    if (more) { fetch(aptr, k0 + BK, ra); fetch(bptr, k0 + BK, rb); }
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][k][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (more) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { As[buf ^ 1][lk + j][lr] = ra[j]; Bs[buf ^ 1][lk + j][lr] = rb[j]; }
    }
    __syncthreads();
    buf ^= 1;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t r = row0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (r >= n_users) continue;
    const float ub = bu[user_ids[r]];
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int64_t c = col0 + jh * 64 + tx * 4;
      if (c + 3 < n_items && ((ld_out & 3) == 0) &&
          ((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
        float4 o;
        o.x = acc[i][jh * 4 + 0] + ub + bi[item_offset + c + 0];
        o.y = acc[i][jh * 4 + 1] + ub + bi[item_offset + c + 1];
        o.z = acc[i][jh * 4 + 2] + ub + bi[item_offset + c + 2];
        o.w = acc[i][jh * 4 + 3] + ub + bi[item_offset + c + 3];
        *reinterpret_cast<float4*>(out + r * ld_out + c) = o;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (c + j < n_items) out[r * ld_out + c + j] = acc[i][jh * 4 + j] + ub + bi[item_offset + c + j];
      }
    }
  }
}

// ------------------------------------------------------------------ ranking: one CTA per user row
struct Key {
  float s;
  int id;
};
// canonical order: a before b  <=>  a.s > b.s or (a.s == b.s and a.id < b.id)
__device__ __forceinline__ bool before(const Key& a, const Key& b) {
  return a.s > b.s || (a.s == b.s && a.id < b.id);
}

__device__ __forceinline__ Key block_best(Key v, Key* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    Key w;
    w.s = __shfl_xor_sync(0xffffffffu, v.s, o);
    w.id = __shfl_xor_sync(0xffffffffu, v.id, o);
    if (w.id >= 0 && (v.id < 0 || before(w, v))) v = w;
  }
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
  __syncthreads();
  Key r = s_red[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) {
    const Key c = s_red[w];
    if (c.id >= 0 && (r.id < 0 || before(c, r))) r = c;
  }
  __syncthreads();
  return r;
}

template <typename T>
__device__ __forceinline__ T block_sum(T v, T* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
  __syncthreads();
  T r = 0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) r += s_red[w];
  __syncthreads();
  return r;
}

constexpr int kSigSmem = 2048;

struct RankDev {
  const float* S;  // [n, ld]
  int64_t ld, n_items, item_offset;
  int k;
  const int64_t* tgt_indptr;
  const int32_t* tgt_items;
  int want_mrr, want_auc;
  const float* ext_best_score;
  const int32_t* ext_best_id;
  int32_t* topk_ids;
  float* topk_scores;
  double* ap;
  int32_t* n_before_best;
  float* best_score;
  int32_t* best_id;
  double* auc_pairs;
  float* sig_scratch;  // [total targets of the tile], same offsets as tgt_items (minus indptr[0])
  // optional indirection (the tensor-core path's hand-over): score row b belongs to user
  // row_map[b]; only the first min(*row_count, row_cap) rows exist
  const int32_t* row_map;
  const int32_t* row_count;
  int row_cap;
};

constexpr int kListL = 4;    // entries of a thread's list in the one-pass ranking

// One CTA per user row. Everything that needs the whole row is gathered in ONE pass over it:
// per-thread top-k lists (shared memory, sorted), the count of items ranked before the best
// target (MRR), the Mann-Whitney pair count against the sorted target sigmoids (AUC); the lists
// are then merged by k block-wide selections. A thread keeps only its best kListL items and the
// best score it let go: if that could belong to the top-k after all (more than kListL of the top-k
// in one thread's 1/256 of the row, or ties at the k-th score), the row is ranked again by k
// selection passes over the scores.
__global__ void __launch_bounds__(256)
k_rank_rows(const RankDev p) {
  __shared__ Key s_key[8];
  __shared__ long long s_ll[8];
  __shared__ double s_dbl[8];
  __shared__ float s_sig[kSigSmem];
  __shared__ float s_raw[kSigSmem];
  __shared__ float s_ls[kListL][256];   // [rank in the thread's list][thread]: conflict-free
  __shared__ int s_li[kListL][256];
  const int tid = threadIdx.x;
  int r = blockIdx.x;
  const float* row = p.S + (int64_t)blockIdx.x * p.ld;
  if (p.row_map) {
    const int n_rows = *p.row_count < p.row_cap ? *p.row_count : p.row_cap;
    if ((int)blockIdx.x >= n_rows) return;
    r = p.row_map[blockIdx.x];
  }
  const int64_t t0 = p.tgt_indptr ? p.tgt_indptr[r] : 0, t1 = p.tgt_indptr ? p.tgt_indptr[r + 1] : 0;
  const int n_tgt = (int)(t1 - t0);
  const int32_t* tgt = p.tgt_items + t0;
  const int lo_id = (int)p.item_offset, hi_id = (int)(p.item_offset + p.n_items);
  const int K = p.k;
  const bool want_topk = p.topk_ids || p.topk_scores || p.ap;
  const bool lists = want_topk;
  const int L = K < kListL ? K : kListL;   // list length
  float dropped = -INFINITY;               // best score this thread did not keep

  // ---- MRR: the best-ranked target (its row entry is read directly)
  Key best;
  best.s = -INFINITY;
  best.id = -1;
  if (p.want_mrr) {
    for (int t = tid; t < n_tgt; t += blockDim.x) {
      const int id = tgt[t];
      if (id < lo_id || id >= hi_id) continue;
      Key cur;
      cur.s = row[id - lo_id];
      cur.id = id;
      if (best.id < 0 || before(cur, best)) best = cur;
    }
    best = block_best(best, s_key);
    if (tid == 0) {
      if (p.best_score) p.best_score[r] = best.id >= 0 ? best.s : -INFINITY;
      if (p.best_id) p.best_id[r] = best.id;
    }
    if (p.ext_best_id) {
      best.id = p.ext_best_id[r];
      best.s = p.ext_best_score[r];
    }
  }

  // ---- AUC: the targets' sigmoids, rank-sorted ascending by counting
  const bool auc = p.want_auc && n_tgt > 0;
  float* sig = s_sig;
  int sig_pow2 = 0;   // > 0: sig is padded to this power of two (shared-memory case)
  if (auc) {
    const bool in_smem = n_tgt <= kSigSmem;
    sig = in_smem ? s_sig : p.sig_scratch + 2 * (t0 - p.tgt_indptr[0]);
    float* raw = in_smem ? s_raw : sig + n_tgt;
    for (int t = tid; t < n_tgt; t += blockDim.x) raw[t] = sigmoidf_(row[tgt[t] - lo_id]);
    __syncthreads();
    for (int t = tid; t < n_tgt; t += blockDim.x) {
      const float v = raw[t];
      int rank = 0;
      for (int q = 0; q < n_tgt; ++q) {
        const float w = raw[q];
        rank += (w < v || (w == v && q < t)) ? 1 : 0;
      }
      sig[rank] = v;
    }
    if (in_smem) {
      sig_pow2 = 1;
      while (sig_pow2 < n_tgt) sig_pow2 <<= 1;
      for (int t = n_tgt + tid; t < sig_pow2; t += blockDim.x) sig[t] = INFINITY;
    }
    __syncthreads();
  }

  // ---- the pass over the row
  int have = 0;             // entries in this thread's list
  float worst = -INFINITY;  // score of the list's last entry once it is full
  long long cnt = 0;
  double total = 0.0;
  const bool one_pass = lists || (p.want_mrr && best.id >= 0) || auc;
  if (one_pass) {
    auto visit = [&](float score, int64_t c) {
      Key cur;
      cur.s = score;
      cur.id = (int)(p.item_offset + c);
      if (lists) {
        if (have < L || cur.s > worst) {
          // ids grow along a thread's walk: an equal score never displaces an entry; sorted insert
          if (have == L) dropped = fmaxf(dropped, worst);
          int pos = have < L ? have : L - 1;
          while (pos > 0 && s_ls[pos - 1][tid] < cur.s) {
            s_ls[pos][tid] = s_ls[pos - 1][tid];
            s_li[pos][tid] = s_li[pos - 1][tid];
            --pos;
          }
          s_ls[pos][tid] = cur.s;
          s_li[pos][tid] = cur.id;
          if (have < L) ++have;
          if (have == L) worst = s_ls[L - 1][tid];
        } else {
          dropped = fmaxf(dropped, cur.s);
        }
      }
      // (the target never precedes itself, whatever arithmetic produced an external threshold)
      if (p.want_mrr && best.id >= 0) cnt += (cur.id != best.id && before(cur, best)) ? 1 : 0;
      if (auc) {
        // every item adds (#targets above it) + (#targets equal to it) / 2
        const float v = sigmoidf_(cur.s);
        int lb, lo;
        if (sig_pow2 > 0) {
          // the sorted sigmoids padded with +inf to a power of two: a lower bound in log2 steps
          // of select arithmetic, the same for all lanes (the two branching searches were 30 %
          // of this kernel's instructions); equal targets are then counted forward (rare)
          int pos = 0;
          for (int step = sig_pow2 >> 1; step > 0; step >>= 1) pos += (sig[pos + step - 1] < v) ? step : 0;
          pos += (sig[pos] < v) ? 1 : 0;
          lb = pos;
          while (pos < n_tgt && sig[pos] == v) ++pos;
          lo = pos;
        } else {
          int hi = n_tgt;  // lower_bound
          lo = 0;
          while (lo < hi) { const int mid = (lo + hi) >> 1; if (sig[mid] < v) lo = mid + 1; else hi = mid; }
          lb = lo;
          hi = n_tgt;              // upper_bound
          while (lo < hi) { const int mid = (lo + hi) >> 1; if (sig[mid] <= v) lo = mid + 1; else hi = mid; }
        }
        total += (double)(n_tgt - lo) + 0.5 * (double)(lo - lb);
      }
    };
    // 16-byte loads, four scores per thread and iteration (a CTA alone on its SM, as in the
    // tensor-core path's hand-over, runs at the latency of its loads); the elements before the
    // first aligned one and after the last whole vector go one by one. A thread's ids still grow
    // along its walk.
    const int64_t mis = (int64_t)((reinterpret_cast<uintptr_t>(row) >> 2) & 3);
    int64_t lead = (4 - mis) & 3;
    if (lead > p.n_items) lead = p.n_items;
    const int64_t n_vec = (p.n_items - lead) >> 2;
    if (tid < lead) visit(row[tid], tid);
    const float4* vrow = reinterpret_cast<const float4*>(row + lead);
    for (int64_t v4 = tid; v4 < n_vec; v4 += blockDim.x) {
      const float4 x = vrow[v4];
      const int64_t c = lead + 4 * v4;
      visit(x.x, c);
      visit(x.y, c + 1);
      visit(x.z, c + 2);
      visit(x.w, c + 3);
    }
    for (int64_t c = lead + 4 * n_vec + tid; c < p.n_items; c += blockDim.x) visit(row[c], c);
  }

  // ---- top-k: k block-wide selections, over the list heads (or over the row when k is large)
  for (int attempt = 0; want_topk && attempt < 2; ++attempt) {
    const bool use_lists = lists && attempt == 0;
    int hits = 0, head = 0, picks = 0;
    double ap_acc = 0.0;
    Key prev, pick;
    prev.s = INFINITY;
    prev.id = -1;
    pick = prev;
    for (int j = 0; j < K; ++j) {
      Key mine;
      mine.s = -INFINITY;
      mine.id = -1;
      if (use_lists) {
        if (head < have) { mine.s = s_ls[head][tid]; mine.id = s_li[head][tid]; }
      } else {
        for (int64_t c = tid; c < p.n_items; c += blockDim.x) {
          Key cur;
          cur.s = row[c];
          cur.id = (int)(p.item_offset + c);
          const bool after_prev = (prev.id < 0) || before(prev, cur);
          if (after_prev && (mine.id < 0 || before(cur, mine))) mine = cur;
        }
      }
      pick = block_best(mine, s_key);
      if (use_lists && pick.id >= 0 && pick.id == mine.id) ++head;
      if (tid == 0) {
        if (p.topk_ids) p.topk_ids[(int64_t)r * K + j] = pick.id;
        if (p.topk_scores) p.topk_scores[(int64_t)r * K + j] = pick.id >= 0 ? pick.s : -INFINITY;
        if (pick.id >= 0 && csr_contains(tgt, 0, n_tgt, pick.id)) {
          ++hits;
          ap_acc += (double)hits / (double)(j + 1);
        }
      }
      prev = pick;
      if (pick.id < 0) {  // fewer than k items in this shard: pad
        if (tid == 0)
          for (int jj = j + 1; jj < K; ++jj) {
            if (p.topk_ids) p.topk_ids[(int64_t)r * K + jj] = -1;
            if (p.topk_scores) p.topk_scores[(int64_t)r * K + jj] = -INFINITY;
          }
        break;
      }
      ++picks;
    }
    if (tid == 0 && p.ap) {
      // AP@k = sum_j acc_j * cumsum_j / j / min(k, nnz_u); 0 when the user has no targets
      const int denom = K < n_tgt ? K : n_tgt;
      p.ap[r] = denom > 0 ? ap_acc / (double)denom : 0.0;
    }
    if (!use_lists) break;
    // could something a thread let go belong to the top-k? (then rank again without the lists)
    const bool doubt = dropped > -INFINITY && (picks < K || dropped >= pick.s);
    if (!__syncthreads_or(doubt ? 1 : 0)) break;
  }

  if (p.want_mrr) {
    cnt = block_sum<long long>(cnt, s_ll);
    if (tid == 0 && p.n_before_best) p.n_before_best[r] = best.id >= 0 ? (int)cnt : -1;
  }
  if (p.want_auc) {
    total = block_sum<double>(total, s_dbl);
    // target-target pairs (self pairs included) contribute exactly n_tgt^2 / 2
    if (tid == 0 && p.auc_pairs)
      p.auc_pairs[r] = n_tgt > 0 ? total - 0.5 * (double)n_tgt * (double)n_tgt : 0.0;
  }
}

static size_t eval_ws_scores(int64_t n, int64_t items) { return align_up(sizeof(float) * n * items, 256); }

// 128 x 128 tiles when the user tile fills them, else the 64 x 64 kernel (identical results)
static void launch_score_tile(const float* U, const float* I, const float* bu, const float* bi,
                              int D, const int32_t* user_ids, int64_t n_users, int64_t item_offset,
                              int64_t n_items, float* out, int64_t ld_out, cudaStream_t s,
                              const int32_t* n_users_dev = nullptr) {
  if (n_users >= 96 && n_items >= 96) {
    dim3 grid((unsigned)((n_items + BN - 1) / BN), (unsigned)((n_users + BM - 1) / BM));
    k_score_tile_128<<<grid, 256, 0, s>>>(U, I, bu, bi, D, user_ids, n_users, item_offset, n_items,
                                          out, ld_out, n_users_dev);
  } else {
    dim3 grid((unsigned)((n_items + TN - 1) / TN), (unsigned)((n_users + TM - 1) / TM));
    k_score_tile<<<grid, 256, 0, s>>>(U, I, bu, bi, D, user_ids, n_users, item_offset, n_items, out,
                                      ld_out, n_users_dev);
  }
}

// exact scores of up to `cap` users whose number is only known on the device (both score kernels
// give the same bits, so the tile shape does not matter)
int launch_score_rows_counted(const cmf_eval_t* a, const int32_t* user_ids, int cap,
                              const int32_t* count, float* out, cudaStream_t s) {
  launch_score_tile(a->U, a->I, a->bu, a->bi, a->D, user_ids, cap, a->item_offset,
                    a->num_items_shard, out, a->num_items_shard, s, count);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

static RankDev rank_args(const cmf_eval_t* a, const float* S, float* sig) {
  RankDev p{};
  p.S = S; p.ld = a->num_items_shard; p.n_items = a->num_items_shard; p.item_offset = a->item_offset;
  p.k = a->k; p.tgt_indptr = a->tgt_indptr; p.tgt_items = a->tgt_items;
  p.want_mrr = a->want_mrr; p.want_auc = a->want_auc;
  p.ext_best_score = a->ext_best_score; p.ext_best_id = a->ext_best_id;
  p.topk_ids = a->topk_ids; p.topk_scores = a->topk_scores; p.ap = a->ap;
  p.n_before_best = a->n_before_best; p.best_score = a->best_score; p.best_id = a->best_id;
  p.auc_pairs = a->auc_pairs; p.sig_scratch = sig;
  return p;
}

int launch_rank_rows_mapped(const cmf_eval_t* a, const float* scores, float* sig_scratch,
                            const int32_t* row_map, const int32_t* row_count, int row_cap,
                            cudaStream_t s) {
  RankDev p = rank_args(a, scores, sig_scratch);
  p.row_map = row_map; p.row_count = row_count; p.row_cap = row_cap;
  k_rank_rows<<<(unsigned)row_cap, 256, 0, s>>>(p);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

}  // namespace cmf

using namespace cmf;

extern "C" size_t cmf_eval_workspace_bytes(int64_t n_users_tile, int64_t num_items_shard, int32_t D,
                                           int32_t k, int32_t precision) {
  if (n_users_tile < 1 || num_items_shard < 1) return 0;
  if (precision == 1) return eval_tc_workspace_bytes(n_users_tile, num_items_shard, D, k);
  // score tile + sigmoid scratch for users with more than kSigSmem targets (bounded by the shard)
  return eval_ws_scores(n_users_tile, num_items_shard) +
         align_up(2 * sizeof(float) * n_users_tile * num_items_shard, 256);
}

extern "C" int cmf_score_tile(const float* U, const float* I, const float* bu, const float* bi,
                              int32_t D, const int32_t* user_ids, int64_t n_users_tile,
                              int64_t num_items, float* out, void* stream) {
  CMF_REQUIRE(U && I && bu && bi && out, "cmf_score_tile: null pointer");
  CMF_REQUIRE(D >= 1, "cmf_score_tile: bad embedding_dim");
  if (n_users_tile == 0 || num_items == 0) return CMF_OK;
  CMF_REQUIRE(user_ids, "cmf_score_tile: null user ids");
  launch_score_tile(U, I, bu, bi, D, user_ids, n_users_tile, 0, num_items, out, num_items,
                    (cudaStream_t)stream);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

extern "C" int cmf_eval_tile(const cmf_eval_t* a, void* stream) {
  CMF_REQUIRE(a, "cmf_eval_tile: null args");
  CMF_REQUIRE(a->scores_in || (a->U && a->I && a->bu && a->bi),
              "cmf_eval_tile: null parameter table");
  CMF_REQUIRE(a->scores_in || a->D >= 1, "cmf_eval_tile: bad embedding_dim");
  CMF_REQUIRE(a->n_users_tile >= 0 && a->num_items_shard >= 1 && a->item_offset >= 0,
              "cmf_eval_tile: bad tile shape");
  CMF_REQUIRE(a->k >= 1, "cmf_eval_tile: k must be >= 1");
  if (a->n_users_tile == 0) return CMF_OK;
  CMF_REQUIRE(a->scores_in || a->user_ids, "cmf_eval_tile: null user ids");
  CMF_REQUIRE(a->tgt_indptr && a->tgt_items, "cmf_eval_tile: target CSR required");
  CMF_REQUIRE(!a->want_auc || a->item_offset == 0,
              "cmf_eval_tile: AUC needs the whole item table in one shard (item_offset must be 0)");
  CMF_REQUIRE((a->ext_best_id == nullptr) == (a->ext_best_score == nullptr),
              "cmf_eval_tile: ext_best_score and ext_best_id go together");
  cudaStream_t s = (cudaStream_t)stream;
  if (a->precision == 1 && !a->scores_in) return eval_tile_tc(a, s);
  CMF_REQUIRE(a->precision == 0 || a->scores_in, "cmf_eval_tile: unknown precision %d",
              a->precision);
  const size_t need = cmf_eval_workspace_bytes(a->n_users_tile, a->num_items_shard, a->D, a->k, 0);
  if (!a->workspace || a->workspace_bytes < need)
    return fail(CMF_ERR_WORKSPACE, "cmf_eval_tile: workspace of %zu bytes needed, %zu given", need,
                a->workspace ? a->workspace_bytes : (size_t)0);
  const float* S = a->scores_in;
  float* sig = (float*)((char*)a->workspace + eval_ws_scores(a->n_users_tile, a->num_items_shard));
  if (!S) {
    float* Sw = (float*)a->workspace;
    launch_score_tile(a->U, a->I, a->bu, a->bi, a->D, a->user_ids, a->n_users_tile, a->item_offset,
                      a->num_items_shard, Sw, a->num_items_shard, s);
    CMF_LAUNCH_CHECK();
    S = Sw;
  }
  const RankDev p = rank_args(a, S, sig);
  k_rank_rows<<<(unsigned)a->n_users_tile, 256, 0, s>>>(p);
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

// ------------------------------------------------------------------ allocation-free evaluation
// cmf_eval_tile_noalloc: the same evaluation as cmf_eval_tile, for hosts that must not (or cannot)
// obtain device memory themselves: the scratch tile lives in the library image.
namespace cmf {
constexpr size_t kNoAllocScratchBytes = 16u << 20;   // 16 MiB, resident on every GPU that loads us
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];
}  // namespace cmf

extern "C" int cmf_eval_tile_noalloc(const cmf_eval_t* a, void* stream) {
  CMF_REQUIRE(a, "cmf_eval_tile_noalloc: null args");
  // In host code `g_noalloc_scratch` is only a handle for the symbol; the device address has to be
  // asked for, once per device, with that device current.
  static std::atomic<void*> dev_ptr[64];   // (zero-initialised; threads may race to fill a slot
                                           // with the same address)
  int dev = -1;
  CMF_CUDA(cudaGetDevice(&dev));
  CMF_REQUIRE(dev >= 0 && dev < 64, "cmf_eval_tile_noalloc: device ordinal %d not supported", dev);
  // the scratch belongs to the CURRENT device: the tables must live there too
  if (a->I) {
    cudaPointerAttributes attr{};
    CMF_CUDA(cudaPointerGetAttributes(&attr, a->I));
    CMF_REQUIRE(attr.type != cudaMemoryTypeDevice || attr.device == dev,
                "cmf_eval_tile_noalloc: the tables live on device %d but device %d is current; "
                "make the owning device current before the call", attr.device, dev);
  }
  if (dev_ptr[dev].load(std::memory_order_acquire) == nullptr) {
    void* ptr = nullptr;
    CMF_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
    dev_ptr[dev].store(ptr, std::memory_order_release);
  }
  const size_t need = cmf_eval_workspace_bytes(a->n_users_tile, a->num_items_shard, a->D, a->k,
                                               a->precision);
  if (need > kNoAllocScratchBytes)
    return fail(CMF_ERR_WORKSPACE,
                "cmf_eval_tile_noalloc: a tile of %lld users x %lld items needs %zu bytes of scratch, "
                "the built-in scratch holds %zu; use fewer users per tile or cmf_eval_tile",
                (long long)a->n_users_tile, (long long)a->num_items_shard, need,
                (size_t)kNoAllocScratchBytes);
  cmf_eval_t b = *a;
  b.workspace = dev_ptr[dev].load(std::memory_order_acquire);
  b.workspace_bytes = kNoAllocScratchBytes;
  return cmf_eval_tile(&b, stream);
}
