// Fused MF training step: on-device negative sampling + gather + scores + loss + backward +
// in-place row optimizer for whole user segments, vector-RED item gradients, and one row-update
// kernel for everything that could not be finished in place.
//
// Reference call sites replaced (one iteration of collie/model/base/trainer.py:458-470):
//   Interactions._negative_sample        collie/interactions/datasets.py:390-445
//   BasePipeline.calculate_loss          collie/model/base/base_pipeline.py:582-654
//   MatrixFactorizationModel.forward x2  collie/model/matrix_factorization.py:120-159
//   losses                               collie/loss/{bpr,hinge,warp,metadata_utils}.py
//   loss.backward() + optimizer.step()   torch autograd / torch.optim.{SparseAdam,Adam,SGD,Adagrad}
//
// Work decomposition: the batch is ordered by user (radix sort of <=B 32-bit keys);
// the ordered batch is cut into chunks of L consecutive interactions; one G-lane group walks one
// chunk. A user whose interactions all lie inside one chunk is finished in place (its row, Adam
// moments and bias are read once and written once, no atomics, no gradient scratch). Users that
// straddle a chunk boundary, and all item rows, accumulate into dense fp32 gradient tables with
// vector RED operations and are finished by k_update_rows, which walks the de-duplicated lists of
// touched rows. Every chunk has the same amount of work, so the grid is balanced whatever the skew.
#include <cub/device/device_radix_sort.cuh>

#include <cmath>
#include <cstdlib>
#include <vector>

#include "train_common.cuh"

namespace cmf {

extern int g_tc_debug;
extern int g_tc_terms;
extern int g_tc_list_len;
extern int g_tc_epi_warps;
extern int g_pdl;
extern int g_red_in_fwd;
size_t det_sort_bytes(int64_t n_pairs);
int launch_aligned_chunks(const StepDev& p, int32_t* starts, cudaStream_t s);
int launch_det_item_side(const StepDev& p, int64_t num_item_rows, uint32_t* keys, uint32_t* keys2,
                         uint32_t* vals, uint32_t* vals2, int32_t* starts, void* temp,
                         size_t temp_bytes, cudaStream_t s);

struct WsLayout {
  // persistent (zero between steps; the kernels restore the invariant)
  size_t item_flag, user_flag, GI, gbi, GU, gbu, persistent_end;
  // per step
  size_t hdr, partials, item_list, user_list, keys_out, vals_in, vals_out, iota, g_arr, sel_arr, cub_temp;
  size_t cub_temp_bytes, total;
  // large item tables: the touched-row list sorted by row, so the update streams through memory
  size_t item_list_sorted, cub_items_temp, cub_items_temp_bytes;
  // deterministic step (train_det.cu): sorted (item, contribution) pairs, item segment starts,
  // user-aligned chunk starts
  size_t det_keys, det_keys2, det_vals, det_vals2, det_starts, det_chunks, det_temp, det_temp_bytes;
};

static size_t cub_sort_bytes(int64_t B) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, bytes, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (int)B, 0, 32, (cudaStream_t)0);
  return bytes;
}

static WsLayout ws_layout(int64_t B, int64_t num_users, int64_t num_items, int D) {
  WsLayout w;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
  // item side first: its offsets depend only on (num_items, D), so they are the same on every
  // rank of a user-partitioned job (the multicast view of the workspaces relies on that)
  w.item_flag = take(sizeof(int) * num_items);
  w.GI = take(sizeof(float) * num_items * D);
  w.gbi = take(sizeof(float) * num_items);
  w.user_flag = take(sizeof(int) * num_users);
  w.GU = take(sizeof(float) * num_users * D);
  w.gbu = take(sizeof(float) * num_users);
  w.persistent_end = o;
  w.hdr = take(sizeof(WsHeader));
  w.partials = take(sizeof(double) * kMaxPartials);
  w.item_list = take(sizeof(int) * num_items);
  w.user_list = take(sizeof(int) * num_users);
  w.keys_out = take(sizeof(uint32_t) * B);
  w.vals_out = take(sizeof(uint32_t) * B);
  w.iota = take(sizeof(uint32_t) * B);
  w.g_arr = take(sizeof(float) * B);
  w.sel_arr = take(sizeof(int32_t) * B);
  w.cub_temp_bytes = cub_sort_bytes(B);
  w.cub_temp = take(w.cub_temp_bytes);
  w.item_list_sorted = take(sizeof(int) * num_items);
  w.cub_items_temp_bytes = 0;
  cub::DeviceRadixSort::SortKeys((void*)nullptr, w.cub_items_temp_bytes, (const uint32_t*)nullptr,
                                 (uint32_t*)nullptr, (int)num_items, 0, 32, (cudaStream_t)0);
  w.cub_items_temp = take(w.cub_items_temp_bytes);
  w.det_keys = take(sizeof(uint32_t) * 2 * B);
  w.det_keys2 = take(sizeof(uint32_t) * 2 * B);
  w.det_vals = take(sizeof(uint32_t) * 2 * B);
  w.det_vals2 = take(sizeof(uint32_t) * 2 * B);
  w.det_starts = take(sizeof(int32_t) * (num_items + 1));
  w.det_chunks = take(sizeof(int32_t) * (B + 2));
  w.det_temp_bytes = det_sort_bytes(2 * B);
  w.det_temp = take(w.det_temp_bytes);
  w.total = o;
  return w;
}

template <int V, int G, int R>
__global__ void __launch_bounds__(kThreads)
k_fused_step(const StepDev p) {
  constexpr int NC = (R * V <= 4) ? 4 : (R * V <= 8 ? 2 : 1);
  constexpr int kGroups = kThreads / G;
  __shared__ int s_negs[kGroups][kMaxNeg];
  __shared__ float s_warp_loss[kThreads / 32];

  const int lane_w = threadIdx.x & 31;
  const int l = threadIdx.x & (G - 1);
  const int gib = threadIdx.x / G;  // group in block
  const unsigned mask = group_mask<G>(lane_w);
  const int64_t group = (int64_t)blockIdx.x * kGroups + gib;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  const int D = p.D, N = p.N;
  const bool implicit = p.loss <= CMF_LOSS_WARP;
  int* negs = s_negs[gib];

  float loss_acc = 0.f;

  for (int64_t c = group; c < p.num_chunks; c += n_groups) {
    const int64_t e0 = c * p.L;
    const int64_t e1 = (e0 + p.L < p.B) ? e0 + p.L : p.B;
    int cur_user = -1;
    bool left_whole = false;
    Row<V, G, R> urow, gu;
    float ubias = 0.f, gbu_acc = 0.f;

    for (int64_t e = e0; e <= e1; ++e) {
      const int u = (e < e1) ? p.su[e] : -2;
      if (u != cur_user) {
        // ---- finish the previous user's segment
        if (cur_user >= 0 && p.train) {
          // e < e1: the next entry (inside this chunk) belongs to another user; e == e1: look
          // across the chunk boundary
          const bool right_whole = (e < e1) || (e1 == p.B) || (p.su[e1] != cur_user);
          if (left_whole && right_whole && p.user_inplace) {
            update_row<V, G, R>(p.emb, p.U, p.U_m, p.U_v, D, l, cur_user, gu);
            if (l == 0) update_scalar(p.bias, p.bu, p.bu_m, p.bu_v, cur_user, gbu_acc);
          } else {
            gu.red_add(p.GU + (int64_t)cur_user * D, D, l);
            if (l == 0) {
              atomicAdd(p.gbu + cur_user, gbu_acc);
              mark_row(p.user_flag, p.user_list, &p.hdr->user_count, cur_user, p.use_lists);
            }
          }
        }
        if (e == e1) break;
        cur_user = u;
        left_whole = (e == 0) || (p.su[e - 1] != u);
        urow.load(p.U + (int64_t)u * D, D, l);
        ubias = p.bu[u];
        gu.zero();
        gbu_acc = 0.f;
      }

      const int64_t b = p.perm ? (int64_t)p.perm[e] : e;
      const int pos = p.items[b];
      Row<V, G, R> irow;
      irow.load(p.I + (int64_t)pos * D, D, l);

      const bool drop = p.drop_thr != 0;
      const int64_t gi_b = p.gidx ? p.gidx[b] : p.gidx_base + b;
      if (!implicit) {
        // ---------------- explicit: mse / mae (base_pipeline.py:635-650)
        Row<V, G, R> um = urow, im = irow;
        if (drop) {
          drop_row<V, G, R>(p, gi_b, 0, 0, D, l, um);
          drop_row<V, G, R>(p, gi_b, 0, 1, D, l, im);
        }
        const float s = group_sum<G>(um.dot(im), mask) + ubias + p.bi[pos];
        float sg = 0.f, pred = s;
        if (p.has_range) { sg = sigmoidf_(s); pred = sg * (p.y_hi - p.y_lo) + p.y_lo; }
        const float diff = pred - p.ratings[b];
        float dp;
        if (p.loss == CMF_LOSS_MSE) { loss_acc += diff * diff; dp = 2.f * diff * p.inv_B; }
        else { loss_acc += fabsf(diff); dp = (diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f)) * p.inv_B; }
        const float ds = p.has_range ? dp * (sg * (1.f - sg) * (p.y_hi - p.y_lo)) : dp;
        if (p.pos_out && l == 0) p.pos_out[b] = pred;
        if (p.train) {
          // (with dropout the score is <mu * u, mi * i>: both masks reach both gradients)
          gu.axpy(ds, drop ? drop_both<V, G, R>(p, gi_b, 0, D, l, irow) : irow);
          gbu_acc += ds;
          Row<V, G, R> t;
          t.zero();
          t.axpy(ds, drop ? drop_both<V, G, R>(p, gi_b, 0, D, l, urow) : urow);
          t.red_add(p.GI + (int64_t)pos * D, D, l);
          if (l == 0) {
            atomicAdd(p.gbi + pos, ds);
            mark_row(p.item_flag, p.item_list, &p.hdr->item_count, pos, p.use_lists);
          }
        }
        continue;
      }

      // ---------------- implicit
      float sp;
      if (drop) {
        Row<V, G, R> um = urow, im = irow;
        drop_row<V, G, R>(p, gi_b, 0, 0, D, l, um);
        drop_row<V, G, R>(p, gi_b, 0, 1, D, l, im);
        sp = group_sum<G>(um.dot(im), mask) + ubias + p.bi[pos];
      } else {
        sp = group_sum<G>(urow.dot(irow), mask) + ubias + p.bi[pos];
      }
      if (p.pos_out && l == 0) p.pos_out[b] = sp;

      // negatives of this interaction -> shared memory
      __syncwarp(mask);
      if (p.negs_in) {
        for (int n = l; n < N; n += G) negs[n] = (int)p.negs_in[b * N + n];
      } else {
        const int64_t gi = p.gidx ? p.gidx[b] : p.gidx_base + b;
        if (p.indptr == nullptr) {
          for (int n = l; n < N; n += G) negs[n] = draw_item(p.seed, p.epoch, gi, n, p.num_items);
        } else if (l == 0) {
          // exact sampling: the reference's rejection loop (datasets.py:399-426) over our stream
          const int64_t lo = p.indptr[cur_user], hi = p.indptr[cur_user + 1];
          uint32_t d = 0;
          int taken = 0, checked = 0;
          while (taken < N) {
            int cand = draw_item(p.seed, p.epoch, gi, d++, p.num_items);
            for (;;) {
              bool clash = csr_contains(p.csr_items, lo, hi, cand);
              for (int j = 0; j < taken && !clash; ++j) clash = (negs[j] == cand);
              if (!clash) break;
              if (checked >= p.max_tries) {
                const int left = N - taken - 1;
                for (int j = 0; j < left; ++j)
                  negs[taken++] = draw_item(p.seed, p.epoch, gi, d++, p.num_items);
                break;
              }
              cand = draw_item(p.seed, p.epoch, gi, d++, p.num_items);
              ++checked;
            }
            negs[taken++] = cand;
          }
        }
      }
      __syncwarp(mask);
      if (p.negs_out) for (int n = l; n < N; n += G) p.negs_out[b * N + n] = negs[n];

      // scan the negatives; keep the row of the selected one for the backward
      Row<V, G, R> sel_row;
      sel_row.zero();
      int sel = (p.loss == CMF_LOSS_WARP) ? N : 0;
      int sel_id = -1;
      float sel_s = -INFINITY, sel_h = 0.f;
      const bool adaptive = p.loss == CMF_LOSS_ADAPTIVE_BPR || p.loss == CMF_LOSS_ADAPTIVE_HINGE;
      for (int n0 = 0; n0 < N; n0 += NC) {
        Row<V, G, R> rows[NC];
        int ids[NC];
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          ids[j] = (n0 + j < N) ? negs[n0 + j] : -1;
          if (ids[j] >= 0) rows[j].load(p.I + (int64_t)ids[j] * D, D, l);
        }
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          if (ids[j] < 0) continue;
          const int n = n0 + j;
          float s;
          if (drop) {
            Row<V, G, R> um = urow, im = rows[j];
            drop_row<V, G, R>(p, gi_b, 1 + n, 0, D, l, um);
            drop_row<V, G, R>(p, gi_b, 1 + n, 1, D, l, im);
            s = group_sum<G>(um.dot(im), mask) + ubias + p.bi[ids[j]];
          } else {
            s = group_sum<G>(urow.dot(rows[j]), mask) + ubias + p.bi[ids[j]];
          }
          if (p.neg_out && l == 0) p.neg_out[b * N + n] = s;
          if (p.loss == CMF_LOSS_WARP) {
            // h = ideal - pos + neg, first strict violation wins (warp.py:95-105,136-154)
            const float h = (ideal_of(p, pos, ids[j]) - sp) + s;
            if (sel == N && h > 0.f) { sel = n; sel_id = ids[j]; sel_h = h; sel_row = rows[j]; }
          } else if (adaptive) {
            // torch.max over the N axis: first index of the maximum (bpr.py:153, hinge.py:157)
            if (n == 0 || s > sel_s) { sel = n; sel_id = ids[j]; sel_s = s; sel_row = rows[j]; }
          } else {
            if (n == 0) { sel = 0; sel_id = ids[j]; sel_s = s; sel_row = rows[j]; }
          }
        }
      }
      if (p.sel_out && l == 0) p.sel_out[b] = sel;

      // per-interaction loss value and d loss / d pos score (d / d selected neg = -g)
      float g = 0.f;
      if (p.loss == CMF_LOSS_WARP) {
        if (sel < N) {
          const float w = logf((float)p.num_items / (float)(sel + 1));
          const float Lw = w * sel_h;
          loss_acc += Lw + Lw * Lw;
          g = -((1.f + 2.f * Lw) * w) * p.inv_B;
        }
      } else {
        const float ideal = ideal_of(p, pos, sel_id);
        const float x = sp - sel_s;
        if (p.loss == CMF_LOSS_BPR || p.loss == CMF_LOSS_ADAPTIVE_BPR) {
          const float sg = sigmoidf_(x);
          const float lv = ideal - sg;
          loss_acc += lv + lv * lv;
          g = -((1.f + 2.f * lv) * (sg * (1.f - sg))) * p.inv_B;
        } else {
          const float raw = ideal - x;
          const float lv = raw > 0.f ? raw : 0.f;
          loss_acc += lv + lv * lv;
          g = (raw >= 0.f) ? -(1.f + 2.f * lv) * p.inv_B : 0.f;
        }
      }

      if (p.train) {
        if (g != 0.f && drop) {
          // <mu * u, mi * i>: the pair's two masks reach both gradients; the positive and the
          // selected negative were scored under different masks
          gu.axpy(g, drop_both<V, G, R>(p, gi_b, 0, D, l, irow));
          gu.axpy(-g, drop_both<V, G, R>(p, gi_b, 1 + sel, D, l, sel_row));
          gbu_acc += g;
          gbu_acc += -g;
          Row<V, G, R> t;
          t.zero();
          t.axpy(g, drop_both<V, G, R>(p, gi_b, 0, D, l, urow));
          t.red_add(p.GI + (int64_t)pos * D, D, l);
          t.zero();
          t.axpy(-g, drop_both<V, G, R>(p, gi_b, 1 + sel, D, l, urow));
          t.red_add(p.GI + (int64_t)sel_id * D, D, l);
          if (l == 0) {
            atomicAdd(p.gbi + pos, g);
            atomicAdd(p.gbi + sel_id, -g);
          }
        } else if (g != 0.f) {
          gu.axpy(g, irow);
          gu.axpy(-g, sel_row);
          gbu_acc += g;
          gbu_acc += -g;
          Row<V, G, R> t;
          t.zero();
          t.axpy(g, urow);
          t.red_add(p.GI + (int64_t)pos * D, D, l);
#pragma unroll
          for (int i = 0; i < R * V; ++i) t.f[i] = -t.f[i];
          t.red_add(p.GI + (int64_t)sel_id * D, D, l);
          if (l == 0) {
            atomicAdd(p.gbi + pos, g);
            atomicAdd(p.gbi + sel_id, -g);
          }
        }
        // every gathered row is "touched" (SparseAdam moves rows whose gradient is exactly zero)
        for (int n = l; n <= N; n += G)
          mark_row(p.item_flag, p.item_list, &p.hdr->item_count, n == N ? pos : negs[n],
                   p.use_lists);
      }
    }
  }

  // ---- loss: lane 0 of every group holds its share; deterministic two-level reduction
  float v = (l == 0) ? loss_acc : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane_w == 0) s_warp_loss[threadIdx.x >> 5] = v;
  __syncthreads();
  __shared__ bool s_last;
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int w = 0; w < kThreads / 32; ++w) tot += (double)s_warp_loss[w];
    p.partials[blockIdx.x] = tot;
    __threadfence();
    const unsigned ticket = atomicAdd(&p.hdr->blocks_done, 1u);
    s_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    double tot = 0.0;
    for (unsigned i = 0; i < gridDim.x; ++i) tot += ((volatile double*)p.partials)[i];
    p.loss_out[0] = (float)(tot / p.loss_div);
  }
}

// ------------------------------------------------------------------------------- row update
struct UpdDev {
  // table 0 = items, table 1 = users
  float *P[2], *M[2], *Vv[2], *G[2];
  float *bP[2], *bM[2], *bV[2], *bG[2];
  int* flag[2];
  const int* list[2];
  int64_t rows[2];
  int D;
  int dense;  // walk every row instead of the touched lists
  int skip_items;  // leave the item table (and its gradients) to the caller
  OptDev emb, bias;
  WsHeader* hdr;
};

// One row of one table: everything it needs is requested up front (gradient row, parameter row,
// moments, bias scalars), so a group has 4 * (number of rows in flight) independent loads
// outstanding instead of a chain of dependent ones.
template <int V, int G, int R>
struct RowJob {
  Row<V, G, R> g, p, m, v;
  float gb, pb, mb, vb;
  int64_t row;
  int t;
  bool touched;

  __device__ __forceinline__ void fetch(const UpdDev& q, int64_t w, int64_t n0, int l) {
    t = w < n0 ? 0 : 1;
    const int64_t k = t == 0 ? w : w - n0;
    row = q.dense ? k : (int64_t)q.list[t][k];
    touched = q.dense ? (q.flag[t][row] != 0) : true;
    const int64_t off = row * q.D;
    g.load(q.G[t] + off, q.D, l);
    p.load(q.P[t] + off, q.D, l);
    if (rule_has_m(q.emb.rule)) m.load(q.M[t] + off, q.D, l); else m.zero();
    if (rule_has_v(q.emb.rule)) v.load(q.Vv[t] + off, q.D, l); else v.zero();
    gb = pb = mb = vb = 0.f;
    if (l == 0) {
      gb = q.bG[t][row];
      pb = q.bP[t][row];
      if (rule_has_m(q.bias.rule)) mb = q.bM[t][row];
      if (rule_has_v(q.bias.rule)) vb = q.bV[t][row];
    }
  }

  __device__ __forceinline__ void finish(const UpdDev& q, int l) {
    const int64_t off = row * q.D;
    if (q.emb.dense || touched) {
      if (touched) {
        Row<V, G, R> z;
        z.zero();
        z.store(q.G[t] + off, q.D, l);
      }
      apply_rule_n<R * V>(q.emb, p.f, m.f, v.f, g.f);
      p.store(q.P[t] + off, q.D, l);
      if (rule_has_m(q.emb.rule)) m.store(q.M[t] + off, q.D, l);
      if (rule_has_v(q.emb.rule)) v.store(q.Vv[t] + off, q.D, l);
    }
    if (l == 0) {
      if (q.bias.dense || touched) {
        if (touched) q.bG[t][row] = 0.f;
        apply_rule(q.bias, pb, mb, vb, gb);
        q.bP[t][row] = pb;
        if (rule_has_m(q.bias.rule)) q.bM[t][row] = mb;
        if (rule_has_v(q.bias.rule)) q.bV[t][row] = vb;
      }
      if (touched) q.flag[t][row] = 0;
    }
  }
};

template <int V, int G, int R>
__global__ void __launch_bounds__(kThreads)
k_update_rows(const UpdDev p) {
  asm volatile("griddepcontrol.wait;" ::: "memory");   // (programmatic dependent of k_bwd_update)
  constexpr int kGroups = kThreads / G;
  const int l = threadIdx.x & (G - 1);
  const int64_t group = (int64_t)blockIdx.x * kGroups + threadIdx.x / G;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  const int64_t n0 = p.skip_items ? 0 : (p.dense ? p.rows[0] : (int64_t)p.hdr->item_count);
  const int64_t n1 = p.dense ? p.rows[1] : (int64_t)p.hdr->user_count;
  const int64_t total = n0 + n1;
  // NJ rows per round: all their loads are requested before the first one is finished
  constexpr int NJ = (R * V <= 4) ? 4 : 2;
  for (int64_t w = group; w < total; w += NJ * n_groups) {
    RowJob<V, G, R> job[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j)
      if (w + j * n_groups < total) job[j].fetch(p, w + j * n_groups, n0, l);
#pragma unroll
    for (int j = 0; j < NJ; ++j)
      if (w + j * n_groups < total) job[j].finish(p, l);
  }
}

__global__ void k_iota(uint32_t* out, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint32_t)i;
}

// standalone optimizer over a dense gradient table (autograd-compatible path)
template <int V, int G, int R>
__global__ void __launch_bounds__(kThreads)
k_optimizer_rows(float* P, float* M, float* Vv, float* Gr, int64_t num_rows, int D,
                 const int32_t* rows, int64_t n_rows, const OptDev o, int zero_grad) {
  constexpr int kGroups = kThreads / G;
  const int l = threadIdx.x & (G - 1);
  const int64_t group = (int64_t)blockIdx.x * kGroups + threadIdx.x / G;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  const int64_t total = rows ? n_rows : num_rows;
  for (int64_t w = group; w < total; w += n_groups) {
    const int64_t row = rows ? (int64_t)rows[w] : w;
    Row<V, G, R> g;
    g.load(Gr + row * D, D, l);
    if (zero_grad) {
      Row<V, G, R> z;
      z.zero();
      z.store(Gr + row * D, D, l);
    }
    update_row<V, G, R>(o, P, M, Vv, D, l, row, g);
  }
}

// owner-side update of a sharded table: rows with touched != 0 (or every row for a dense rule)
template <int V, int G, int R>
__global__ void __launch_bounds__(kThreads)
k_optimizer_rows_masked(float* P, float* M, float* Vv, const float* Gr, float* bP, float* bM,
                        float* bV, const float* bG, int64_t num_rows, int D,
                        const int32_t* touched, const OptDev o, const OptDev ob) {
  constexpr int kGroups = kThreads / G;
  const int l = threadIdx.x & (G - 1);
  const int64_t group = (int64_t)blockIdx.x * kGroups + threadIdx.x / G;
  const int64_t n_groups = (int64_t)gridDim.x * kGroups;
  for (int64_t row = group; row < num_rows; row += n_groups) {
    const bool t = touched[row] != 0;
    if (o.dense || t) {
      Row<V, G, R> g;
      g.load(Gr + row * D, D, l);
      update_row<V, G, R>(o, P, M, Vv, D, l, row, g);
    }
    if (bG && l == 0 && (ob.dense || t)) update_scalar(ob, bP, bM, bV, row, bG[row]);
  }
}

static std::atomic<int> g_forced_chunk_len{-1};
static std::atomic<int> g_max_group{32};
static std::atomic<int> g_kernel_version{0};  // 0 auto (v2 when the shape allows), 1 generic, 2 v2

// ---- optional phase timing (bench.py roofline): events on the launching stream
struct ProfRec { cudaEvent_t e[5]; int split; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static WsHeader* g_prof_last_hdr = nullptr;
static cudaStream_t g_prof_stream = nullptr;

// side stream + events that let the batch sort run next to k_fwd_loss (one set per device)
struct SideStream {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
static SideStream* side_stream() {
  static SideStream per_dev[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  SideStream& ss = per_dev[dev];
  if (!ss.stream) {
    if (cudaStreamCreateWithFlags(&ss.stream, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming);
  }
  return &ss;
}

static int chunk_len(int64_t B, int G) {
  int forced = g_forced_chunk_len.load();
  if (forced < 0) {
    const char* e = getenv("CMF_CHUNK_LEN");
    forced = e ? atoi(e) : 0;
    g_forced_chunk_len.store(forced);
  }
  if (forced > 0) return forced;
  // aim for >= 2 chunks per resident group so the static round-robin stays balanced
  const int64_t resident = (int64_t)num_sms() * (2048 / G);
  int L = 16;
  while (L > 1 && B / L < 2 * resident) L >>= 1;
  return L;
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_set_tuning(int32_t key, int32_t value) {
  if (key == 1) {
    CMF_REQUIRE(value == 8 || value == 16 || value == 32, "cmf_set_tuning: group width %d", value);
    return g_max_group.exchange(value);
  }
  if (key == 3) {
    const int prev = g_tc_debug;
    g_tc_debug = value;
    return prev;
  }
  if (key == 4) {
    CMF_REQUIRE(value == 0 || value == 1 || value == 3, "cmf_set_tuning: terms %d", value);
    const int prev = g_tc_terms;
    g_tc_terms = value;
    return prev;
  }
  if (key == 5) {
    CMF_REQUIRE(value == 0 || value == 8 || value == 12 || value == 16 || value == 24,
                "cmf_set_tuning: list length %d", value);
    const int prev = g_tc_list_len;
    g_tc_list_len = value;
    return prev;
  }
  if (key == 8) {
    const int prev = g_red_in_fwd;
    g_red_in_fwd = value ? 1 : 0;
    return prev;
  }
  if (key == 7) {
    CMF_REQUIRE(value == 0 || value == 8 || value == 16, "cmf_set_tuning: epilogue warps %d", value);
    const int prev = g_tc_epi_warps;
    g_tc_epi_warps = value;
    return prev;
  }
  if (key == 6) {
    const int prev = g_pdl;
    g_pdl = value;
    return prev;
  }
  if (key == 2) {
    CMF_REQUIRE(value >= 0 && value <= 2, "cmf_set_tuning: kernel version %d", value);
    return g_kernel_version.exchange(value);
  }
  CMF_REQUIRE(key == 0, "cmf_set_tuning: unknown key %d", key);
  CMF_REQUIRE(value >= 0 && value <= 4096, "cmf_set_tuning: chunk length %d out of range", value);
  const int prev = g_forced_chunk_len.exchange(value);
  return prev < 0 ? 0 : prev;
}

extern "C" size_t cmf_train_workspace_bytes(int64_t max_B, int32_t max_N, int64_t num_users,
                                            int64_t num_items, int32_t D) {
  (void)max_N;
  if (max_B < 1 || num_users < 1 || num_items < 1 || D < 1) return 0;
  return ws_layout(max_B, num_users, num_items, D).total;
}

extern "C" int cmf_train_workspace_items(void* workspace, int64_t num_users, int64_t num_items,
                                         int32_t D, float** GI, float** gbi, int32_t** item_flag) {
  CMF_REQUIRE(workspace && num_users > 0 && num_items > 0 && D > 0,
              "cmf_train_workspace_items: bad arguments");
  const WsLayout w = ws_layout(1, num_users, num_items, D);   // the persistent part ignores B
  char* base = (char*)workspace;
  if (GI) *GI = (float*)(base + w.GI);
  if (gbi) *gbi = (float*)(base + w.gbi);
  if (item_flag) *item_flag = (int32_t*)(base + w.item_flag);
  return CMF_OK;
}

extern "C" int cmf_optimizer_rows_masked(float* P, float* M, float* V, const float* G, float* bP,
                                         float* bM, float* bV, const float* bG, int64_t num_rows,
                                         int32_t D, const int32_t* touched, const cmf_opt_t* opt,
                                         const cmf_opt_t* bias_opt, void* stream) {
  CMF_REQUIRE(P && G && touched && opt, "cmf_optimizer_rows_masked: null pointer");
  CMF_REQUIRE(geom_supported(D), "cmf_optimizer_rows_masked: embedding_dim %d unsupported", D);
  int rc = validate_opt(*opt, "embedding", M, V);
  if (rc) return rc;
  OptDev ob{};
  if (bG) {
    CMF_REQUIRE(bP && bias_opt, "cmf_optimizer_rows_masked: bias table / rule missing");
    rc = validate_opt(*bias_opt, "bias", bM, bV);
    if (rc) return rc;
    ob = make_opt(*bias_opt);
  }
  if (num_rows <= 0) return CMF_OK;
  const OptDev o = make_opt(*opt);
  const RowGeom geom = row_geom(D);
  const int groups_per_block = kThreads / geom.G;
  int64_t blocks = (num_rows + groups_per_block - 1) / groups_per_block;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = (cudaStream_t)stream;
#define CALL(V_, G_, R_)                                                                     \
  k_optimizer_rows_masked<V_, G_, R_><<<(unsigned)blocks, kThreads, 0, s>>>(                 \
      P, M, V, G, bP, bM, bV, bG, num_rows, D, touched, o, ob)
  CMF_DISPATCH_GEOM(geom, CALL);
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

extern "C" int cmf_workspace_init(void* workspace, size_t workspace_bytes, void* stream) {
  CMF_REQUIRE(workspace && workspace_bytes > 0, "cmf_workspace_init: null workspace");
  CMF_CUDA(cudaMemsetAsync(workspace, 0, workspace_bytes, (cudaStream_t)stream));
  return CMF_OK;
}

extern "C" int cmf_train_step(const cmf_step_t* a, void* stream) {
  CMF_REQUIRE(a, "cmf_train_step: null args");
  CMF_REQUIRE(a->U && a->I && a->bu && a->bi, "cmf_train_step: null parameter table");
  CMF_REQUIRE(a->num_users > 0 && a->num_items > 0 && a->num_items < (1ll << 31) &&
                  a->num_users < (1ll << 31), "cmf_train_step: bad table shape");
  CMF_REQUIRE(geom_supported(a->D), "cmf_train_step: embedding_dim %d unsupported (1..1024)", a->D);
  CMF_REQUIRE(a->B >= 1 && a->B < (1ll << 31), "cmf_train_step: batch size %lld out of range",
              (long long)a->B);
  CMF_REQUIRE(a->users && a->items && a->loss_out, "cmf_train_step: null batch/output pointer");
  CMF_REQUIRE(a->loss >= CMF_LOSS_BPR && a->loss <= CMF_LOSS_MAE, "cmf_train_step: unknown loss %d",
              a->loss);
  const bool implicit = a->loss <= CMF_LOSS_WARP;
  if (implicit) {
    CMF_REQUIRE(a->N >= 1 && a->N <= kMaxNeg, "cmf_train_step: num_negative_samples %d not in 1..%d",
                a->N, kMaxNeg);
    CMF_REQUIRE(a->loss != CMF_LOSS_WARP || a->N > 1,
                "Cannot use WARP loss with a single negative sample!");
    CMF_REQUIRE((a->loss != CMF_LOSS_BPR && a->loss != CMF_LOSS_HINGE) || a->N == 1,
                "cmf_train_step: bpr/hinge take exactly one negative (use the adaptive form)");
    CMF_REQUIRE(!a->has_y_range, "cmf_train_step: y_range applies to explicit losses only");
    CMF_REQUIRE(a->n_meta >= 0 && a->n_meta <= CMF_MAX_METADATA, "cmf_train_step: bad n_meta");
    double wsum = 0;
    for (int k = 0; k < a->n_meta; ++k) {
      CMF_REQUIRE(a->meta[k], "cmf_train_step: null metadata array %d", k);
      wsum += a->meta_w[k];
    }
    CMF_REQUIRE(wsum <= 1.0 + 1e-6, "sum of metadata weights was %g, must be <=1", wsum);
    CMF_REQUIRE(a->negs_in || a->csr_indptr == nullptr || a->csr_items,
                "cmf_train_step: csr_items missing for exact sampling");
  } else {
    CMF_REQUIRE(a->ratings, "cmf_train_step: explicit loss needs ratings");
    CMF_REQUIRE(a->N == 0, "cmf_train_step: explicit loss takes no negatives");
  }
  if (a->train) {
    int rc = validate_opt(a->emb_opt, "embedding", a->U_m, a->U_v);
    if (rc) return rc;
    rc = validate_opt(a->emb_opt, "embedding", a->I_m, a->I_v);
    if (rc) return rc;
    rc = validate_opt(a->bias_opt, "bias", a->bu_m, a->bu_v);
    if (rc) return rc;
    rc = validate_opt(a->bias_opt, "bias", a->bi_m, a->bi_v);
    if (rc) return rc;
  }
  const WsLayout w = ws_layout(a->B, a->num_users, a->num_items, a->D);
  if (!a->workspace || a->workspace_bytes < w.total)
    return fail(CMF_ERR_WORKSPACE, "cmf_train_step: workspace of %zu bytes needed, %zu given",
                w.total, a->workspace ? a->workspace_bytes : (size_t)0);

  cudaStream_t s = (cudaStream_t)stream;
  char* base = (char*)a->workspace;
  const RowGeom geom = row_geom(a->D, g_max_group.load());

  StepDev p{};
  p.U = a->U; p.I = a->I; p.bu = a->bu; p.bi = a->bi;
  p.U_m = a->U_m; p.U_v = a->U_v; p.bu_m = a->bu_m; p.bu_v = a->bu_v;
  p.D = a->D;
  CMF_REQUIRE(a->num_items_logical >= 0 && a->num_items_logical <= a->num_items,
              "cmf_train_step: num_items_logical out of range");
  p.num_items = (uint32_t)(a->num_items_logical > 0 ? a->num_items_logical : a->num_items);
  p.items = a->items; p.ratings = a->ratings; p.negs_in = a->negs_in;
  p.gidx = a->gidx; p.gidx_base = a->gidx_base;
  p.B = a->B; p.N = a->N;
  p.L = chunk_len(a->B, geom.G);
  p.num_chunks = (a->B + p.L - 1) / p.L;
  p.seed = a->seed; p.epoch = a->epoch;
  p.indptr = a->csr_indptr; p.csr_items = a->csr_items;
  p.max_tries = a->max_tries < 0 ? 0 : a->max_tries;
  p.loss = a->loss; p.has_range = a->has_y_range; p.y_lo = a->y_lo; p.y_hi = a->y_hi;
  p.n_meta = implicit ? a->n_meta : 0;
  for (int k = 0; k < p.n_meta; ++k) { p.meta[k] = a->meta[k]; p.meta_w[k] = a->meta_w[k]; }
  p.train = a->train ? 1 : 0;
  if (p.train) { p.emb = make_opt(a->emb_opt); p.bias = make_opt(a->bias_opt); }
  CMF_REQUIRE(a->B_global == 0 || a->B_global >= a->B, "cmf_train_step: B_global smaller than B");
  const int dense = p.train && (p.emb.dense || p.bias.dense);
  p.user_inplace = p.train && !dense;
  p.use_lists = p.train && !dense;
  p.inv_B = 1.0f / (float)(a->B_global > 0 ? a->B_global : a->B);
  p.loss_div = (double)(a->B_global > 0 ? a->B_global : a->B);
  p.GU = (float*)(base + w.GU); p.gbu = (float*)(base + w.gbu);
  p.GI = (float*)(base + w.GI); p.gbi = (float*)(base + w.gbi);
  p.user_flag = (int*)(base + w.user_flag); p.item_flag = (int*)(base + w.item_flag);
  p.user_list = (int*)(base + w.user_list); p.item_list = (int*)(base + w.item_list);
  p.hdr = (WsHeader*)(base + w.hdr);
  p.partials = (double*)(base + w.partials);
  p.loss_out = a->loss_out;
  p.negs_out = a->negs_out; p.sel_out = a->sel_out;
  p.pos_out = a->pos_scores_out; p.neg_out = a->neg_scores_out;

  p.users = a->users;
  p.g_arr = (float*)(base + w.g_arr);
  p.sel_arr = (int32_t*)(base + w.sel_arr);

  CMF_REQUIRE(a->dropout_p >= 0.f && a->dropout_p < 1.f, "cmf_train_step: dropout_p %g not in [0, 1)",
              (double)a->dropout_p);
  if (a->dropout_p > 0.f) {
    p.drop_thr = (uint32_t)((double)a->dropout_p * 4294967296.0);
    if (p.drop_thr == 0) p.drop_thr = 1;
    p.drop_scale = 1.0f / (1.0f - a->dropout_p);
    p.drop_key_hi = (uint32_t)a->emb_opt.step;
  }
  const int version = g_kernel_version.load();
  // (dropout lives in the generic kernel only: it keeps the scored rows and their masks together)
  const bool split = version != 1 && v3_supported(a->D, a->N) && a->dropout_p == 0.f;
  CMF_REQUIRE(split || !(a->train && a->deterministic),
              "cmf_train_step: the deterministic step needs embedding_dim %% 4 == 0, at most %d "
              "negatives and no dropout (got D = %d, N = %d, dropout_p = %g)", kMaxNeg, a->D, a->N,
              (double)a->dropout_p);
  CMF_REQUIRE(split || version != 2,
              "cmf_train_step: the split kernels do not support this shape (D=%d, N=%d)", a->D, a->N);
  const int forced_len = g_forced_chunk_len.load() > 0 ? g_forced_chunk_len.load() : 0;

  CMF_CUDA(cudaMemsetAsync(p.hdr, 0, sizeof(WsHeader), s));
  // Item tables that do not stay in L2: k_update_rows walks the touched rows in row order (the
  // list is sorted before it runs), so its reads and writes stream instead of landing in
  // first-touch order. The unused tail of the list is pre-set above every valid id.
  const int64_t list_cap = std::min<int64_t>(a->num_items, a->B * (int64_t)(a->N + 1));
  const bool sort_item_list = p.train && p.use_lists && !a->skip_item_update &&
                              (int64_t)a->num_items * a->D * 4 * 3 > ((int64_t)64 << 20);
  if (sort_item_list)
    CMF_CUDA(cudaMemsetAsync(p.item_list, 0x7F, sizeof(int) * (size_t)list_cap, s));
  ProfRec prof{};
  prof.split = split ? 1 : 0;
  if (g_prof_on) {
    for (auto& e : prof.e) CMF_CUDA(cudaEventCreate(&e));
    CMF_CUDA(cudaEventRecord(prof.e[0], s));
  }

  // ---- order the batch by user. On the split path only k_bwd_update needs the order, so the sort
  // runs on a side stream next to k_fwd_loss.
  const bool need_sort = !a->users_sorted && (p.train || !split);
  SideStream* side = (split && need_sort) ? side_stream() : nullptr;
  cudaStream_t sort_stream = side ? side->stream : s;
  if (side) {
    CMF_CUDA(cudaEventRecord(side->fork, s));
    CMF_CUDA(cudaStreamWaitEvent(side->stream, side->fork, 0));
  }
  if (!need_sort) {
    p.su = a->users;
    p.perm = nullptr;
  } else {
    uint32_t* iota = (uint32_t*)(base + w.iota);
    uint32_t* keys_out = (uint32_t*)(base + w.keys_out);
    uint32_t* vals_out = (uint32_t*)(base + w.vals_out);
    k_iota<<<(unsigned)((a->B + 255) / 256), 256, 0, sort_stream>>>(iota, a->B);
    CMF_LAUNCH_CHECK();
    int bits = 1;
    while (bits < 32 && (1ll << bits) < a->num_users) ++bits;
    size_t temp = w.cub_temp_bytes;
    CMF_CUDA(cub::DeviceRadixSort::SortPairs(base + w.cub_temp, temp, (const uint32_t*)a->users,
                                             keys_out, (const uint32_t*)iota, vals_out, (int)a->B,
                                             0, bits, sort_stream));
    p.su = (const int32_t*)keys_out;
    p.perm = vals_out;
  }
  if (side) CMF_CUDA(cudaEventRecord(side->join, side->stream));

  if (split) {
    // e0 -> e1: k_fwd_loss;  e1 -> e2: what is left of the sort;  e2 -> e3: k_bwd_update
    const bool det = p.train && a->deterministic != 0;
    p.red_in_fwd = (!det && g_red_in_fwd && p.train && a->D / 4 <= 32) ? 1 : 0;
    p.bwd_skip_item = (p.red_in_fwd || det) ? 1 : 0;
    int rc = launch_fwd_loss_v3(p, side != nullptr, s);
    if (rc != CMF_OK) return rc;
    if (g_prof_on) CMF_CUDA(cudaEventRecord(prof.e[1], s));
    if (side) CMF_CUDA(cudaStreamWaitEvent(s, side->join, 0));
    if (g_prof_on) CMF_CUDA(cudaEventRecord(prof.e[2], s));
    if (det) {
      // item side first: it reads the user rows the backward kernel is about to update in place
      rc = launch_det_item_side(p, a->num_items, (uint32_t*)(base + w.det_keys),
                                (uint32_t*)(base + w.det_keys2), (uint32_t*)(base + w.det_vals),
                                (uint32_t*)(base + w.det_vals2), (int32_t*)(base + w.det_starts),
                                base + w.det_temp, w.det_temp_bytes, s);
      if (rc != CMF_OK) return rc;
    }
    if (p.train) {
      rc = launch_bwd_update_v3(p, forced_len, s, det ? (int32_t*)(base + w.det_chunks) : nullptr);
      if (rc != CMF_OK) return rc;
    }
  } else {
    // e0 -> e1: sort;  e1 -> e2: generic fused kernel;  e2 -> e3: nothing
    if (g_prof_on) CMF_CUDA(cudaEventRecord(prof.e[1], s));
    const int groups_per_block = kThreads / geom.G;
    int64_t blocks = (p.num_chunks + groups_per_block - 1) / groups_per_block;
    const int64_t cap = (int64_t)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks > kMaxPartials) blocks = kMaxPartials;
#define CALL(V_, G_, R_) k_fused_step<V_, G_, R_><<<(unsigned)blocks, kThreads, 0, s>>>(p)
    CMF_DISPATCH_GEOM(geom, CALL);
#undef CALL
    CMF_LAUNCH_CHECK();
    if (g_prof_on) CMF_CUDA(cudaEventRecord(prof.e[2], s));
  }
  if (g_prof_on) CMF_CUDA(cudaEventRecord(prof.e[3], s));
  // the item-side accumulators are complete from here on
  if (a->item_grads_ready_event)
    CMF_CUDA(cudaEventRecord((cudaEvent_t)a->item_grads_ready_event, s));

  // ---- finish items (always) and the users that were not finished in place
  if (p.train) {
    UpdDev q{};
    q.P[0] = a->I; q.M[0] = a->I_m; q.Vv[0] = a->I_v; q.G[0] = p.GI;
    q.bP[0] = a->bi; q.bM[0] = a->bi_m; q.bV[0] = a->bi_v; q.bG[0] = p.gbi;
    q.flag[0] = p.item_flag; q.list[0] = p.item_list; q.rows[0] = a->num_items;
    if (sort_item_list) {
      int bits = 1;
      while (bits < 31 && ((int64_t)1 << bits) < a->num_items) ++bits;
      size_t temp = w.cub_items_temp_bytes;
      uint32_t* sorted = (uint32_t*)(base + w.item_list_sorted);
      CMF_CUDA(cub::DeviceRadixSort::SortKeys(base + w.cub_items_temp, temp,
                                              (const uint32_t*)p.item_list, sorted, (int)list_cap,
                                              0, 31, s));
      (void)bits;
      q.list[0] = (const int*)sorted;
    }
    q.P[1] = a->U; q.M[1] = a->U_m; q.Vv[1] = a->U_v; q.G[1] = p.GU;
    q.bP[1] = a->bu; q.bM[1] = a->bu_m; q.bV[1] = a->bu_v; q.bG[1] = p.gbu;
    q.flag[1] = p.user_flag; q.list[1] = p.user_list; q.rows[1] = a->num_users;
    q.D = a->D; q.dense = dense; q.emb = p.emb; q.bias = p.bias; q.hdr = p.hdr;
    q.skip_items = a->skip_item_update ? 1 : 0;
    const int groups_per_block = kThreads / geom.G;
    // sparse mode: the list lengths live on the device; size the grid for the worst case but cap it
    int64_t work = dense ? a->num_items + a->num_users
                         : std::min<int64_t>(a->num_items, a->B * (int64_t)(a->N + 1)) +
                               std::min<int64_t>(a->num_users, a->B);
    int64_t blocks = (work + groups_per_block - 1) / groups_per_block;
    const int64_t cap = (int64_t)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3(kThreads);
    cfg.stream = s;
    cudaLaunchAttribute pdl[1];
    pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pdl[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = pdl;
    cfg.numAttrs = (g_pdl && split) ? 1 : 0;
#define CALL(V_, G_, R_) cudaLaunchKernelEx(&cfg, k_update_rows<V_, G_, R_>, q)
    CMF_DISPATCH_GEOM(geom, CALL);
#undef CALL
    CMF_LAUNCH_CHECK();
  }
  if (g_prof_on) {
    CMF_CUDA(cudaEventRecord(prof.e[4], s));
    g_prof.push_back(prof);
    g_prof_last_hdr = p.hdr;
    g_prof_stream = s;
  }
  return CMF_OK;
}

extern "C" int cmf_profile_enable(int32_t on) {
  g_prof_on = on != 0;
  return CMF_OK;
}

extern "C" int cmf_profile_read(double* host_ms /* [4] */, int32_t* host_split, int64_t* steps,
                                int64_t* last_item_rows, int64_t* last_user_rows) {
  double t[4] = {0, 0, 0, 0};
  int split = 0;
  for (auto& r : g_prof) {
    CMF_CUDA(cudaEventSynchronize(r.e[4]));
    for (int i = 0; i < 4; ++i) {
      float ms = 0.f;
      CMF_CUDA(cudaEventElapsedTime(&ms, r.e[i], r.e[i + 1]));
      t[i] += ms;
    }
    split = r.split;
    for (auto& e : r.e) cudaEventDestroy(e);
  }
  if (host_ms) for (int i = 0; i < 4; ++i) host_ms[i] = t[i];
  if (host_split) *host_split = split;
  if (steps) *steps = (int64_t)g_prof.size();
  if (g_prof_last_hdr && (last_item_rows || last_user_rows)) {
    WsHeader h{};
    CMF_CUDA(cudaMemcpyAsync(&h, g_prof_last_hdr, sizeof(h), cudaMemcpyDeviceToHost, g_prof_stream));
    CMF_CUDA(cudaStreamSynchronize(g_prof_stream));
    if (last_item_rows) *last_item_rows = h.item_count;
    if (last_user_rows) *last_user_rows = h.user_count;
  }
  g_prof.clear();
  return CMF_OK;
}

extern "C" int cmf_train_step_implicit(const cmf_step_t* a, void* stream) {
  CMF_REQUIRE(a, "cmf_train_step_implicit: null args");
  CMF_REQUIRE(a->loss <= CMF_LOSS_WARP, "Explicit loss with implicit data is invalid!");
  return cmf_train_step(a, stream);
}

extern "C" int cmf_train_step_explicit(const cmf_step_t* a, void* stream) {
  CMF_REQUIRE(a, "cmf_train_step_explicit: null args");
  CMF_REQUIRE(a->loss == CMF_LOSS_MSE || a->loss == CMF_LOSS_MAE,
              "Implicit loss with explicit data is invalid!");
  return cmf_train_step(a, stream);
}

extern "C" int cmf_optimizer_rows(float* P, float* M, float* V, float* G, int64_t num_rows,
                                  int32_t D, const int32_t* rows, int64_t n_rows,
                                  const cmf_opt_t* opt, int32_t zero_grad, void* stream) {
  CMF_REQUIRE(P && G && opt, "cmf_optimizer_rows: null pointer");
  CMF_REQUIRE(geom_supported(D), "cmf_optimizer_rows: embedding_dim %d unsupported", D);
  CMF_REQUIRE(opt->rule >= CMF_OPT_SPARSE_ADAM && opt->rule <= CMF_OPT_ADAGRAD,
              "cmf_optimizer_rows: unknown rule %d", opt->rule);
  CMF_REQUIRE(opt->rule == CMF_OPT_SGD || M, "cmf_optimizer_rows: state tensor M required");
  CMF_REQUIRE((opt->rule != CMF_OPT_SPARSE_ADAM && opt->rule != CMF_OPT_ADAM) || V,
              "cmf_optimizer_rows: state tensor V required");
  CMF_REQUIRE(opt->step >= 1, "cmf_optimizer_rows: step must be >= 1");
  const int64_t total = rows ? n_rows : num_rows;
  if (total <= 0) return CMF_OK;
  const OptDev o = make_opt(*opt);
  const RowGeom geom = row_geom(D);
  const int groups_per_block = kThreads / geom.G;
  int64_t blocks = (total + groups_per_block - 1) / groups_per_block;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = (cudaStream_t)stream;
#define CALL(V_, G_, R_)                                                                     \
  k_optimizer_rows<V_, G_, R_><<<(unsigned)blocks, kThreads, 0, s>>>(P, M, V, G, num_rows, D, \
                                                                     rows, n_rows, o, zero_grad)
  CMF_DISPATCH_GEOM(geom, CALL);
#undef CALL
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

// One pass over a planned epoch (collie_b200 loaders lay an epoch out as contiguous arrays, every
// batch already ordered by user): the steps are issued back to back from here, so a small-batch
// job is not paced by its host language's per-step overhead.
extern "C" int cmf_train_epoch(const cmf_step_t* first, int64_t n_total, float* losses,
                               void* stream) {
  CMF_REQUIRE(first && losses, "cmf_train_epoch: null argument");
  CMF_REQUIRE(first->B >= 1, "cmf_train_epoch: batch size must be positive");
  CMF_REQUIRE(n_total >= 0, "cmf_train_epoch: negative length");
  CMF_REQUIRE(!first->negs_out && !first->sel_out && !first->pos_scores_out &&
                  !first->neg_scores_out,
              "cmf_train_epoch: per-batch inspection outputs are not available here");
  const int64_t B = first->B;
  int64_t s = 0;
  for (int64_t off = 0; off < n_total; off += B, ++s) {
    cmf_step_t a = *first;
    a.B = n_total - off < B ? n_total - off : B;
    a.users = first->users + off;
    a.items = first->items + off;
    if (first->ratings) a.ratings = first->ratings + off;
    if (first->negs_in) a.negs_in = first->negs_in + off * first->N;
    if (first->gidx) a.gidx = first->gidx + off;
    a.gidx_base = first->gidx_base + off;
    a.emb_opt.step = first->emb_opt.step + s;
    a.bias_opt.step = first->bias_opt.step + s;
    a.loss_out = losses + s;
    const int rc = cmf_train_step(&a, stream);
    if (rc != CMF_OK) return rc;
  }
  return CMF_OK;
}
