// Training step as two kernels (the fast path; D % 4 == 0), along the reference's own seam between
// "calculate_loss" and "backward + optimizer.step()":
//
//   k_fwd_loss    negative sampling + gather + scores + loss + d loss / d score for every
//                 interaction of the batch, in batch order. Nothing here depends on the user order,
//                 so it runs concurrently with the radix sort of the batch, at high occupancy
//                 (no shared-memory staging: a warp keeps the 8 rows of one score group in flight
//                 in registers, 32+ warps per SM hide the latency).
//                 Replaces Interactions._negative_sample (datasets.py:390-445), forward x2
//                 (matrix_factorization.py:120-159) and collie/loss/*.py.
//   k_bwd_update  walks the user-ordered batch in chunks of L interactions: re-reads the three rows
//                 an interaction's gradient needs (user, positive, selected negative; L2 hits),
//                 accumulates the user's gradient row in registers, sends the two item-row
//                 contributions with vector RED, and finishes every user whose segment lies inside
//                 the chunk in place (row, Adam moments and bias read once, written once).
//                 Replaces loss.backward() + optimizer.step() (trainer.py:463-470).
//
// k_update_rows (train_step.cu) then finishes the item rows and the users that straddle chunks.
#include <cstdlib>

#include "train_common.cuh"

namespace cmf {

constexpr int kChunkA = 8;    // interactions per id table in k_fwd_loss
constexpr int kMaxChunkB = 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ float lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ int lds_i32(uint32_t a) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_i32(uint32_t a, int v) {
  asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_f32(uint32_t a, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
}
__device__ __forceinline__ float4 ldg_f4(const char* p) {
  return *reinterpret_cast<const float4*>(p);
}
// float -> uint32 whose unsigned order is the float order (NaN aside)
__device__ __forceinline__ uint32_t float_key(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// Sum 8 per-lane values v[0..7] over the warp at once: three halving exchanges then two plain
// ones (9 shuffles instead of 40). On return lane l holds the total of row
// ((l >> 4) & 1) * 4 + ((l >> 3) & 1) * 2 + ((l >> 2) & 1).
__device__ __forceinline__ float reduce8(float (&v)[8], int lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
  float w[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float keep = b4 ? v[k + 4] : v[k];
    const float send = b4 ? v[k] : v[k + 4];
    w[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  float x[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const float keep = b3 ? w[k + 2] : w[k];
    const float send = b3 ? w[k] : w[k + 2];
    x[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  const float keep = b2 ? x[1] : x[0];
  const float send = b2 ? x[0] : x[1];
  float y = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  y += __shfl_xor_sync(0xffffffffu, y, 2);
  y += __shfl_xor_sync(0xffffffffu, y, 1);
  return y;
}
__device__ __forceinline__ int reduce8_row(int lane) {
  return ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
}
// first lane that holds row j after reduce8
__device__ __forceinline__ int reduce8_lane(int j) {
  return ((j >> 2) & 1) * 16 + ((j >> 1) & 1) * 8 + (j & 1) * 4;
}

// ===================================================================================== k_fwd_loss
// R float4 per lane hold a 128 * R-float row. FULL: D == 128 * R.
template <int R, bool FULL>
__global__ void __launch_bounds__(256)
k_fwd_loss(const StepDev p) {
  __shared__ int s_ids[8][kChunkA][kMaxNeg + 2];   // per warp: user, positive, negatives
  __shared__ float s_warp_loss[8];
  __shared__ bool s_last;

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int D = p.D, N = p.N;
  const int D4 = FULL ? 32 * R : (D >> 2);
  const int Rw = N + 1;
  const uint32_t a_ids = smem_u32(&s_ids[warp][0][0]);
  constexpr uint32_t kRowStride = (kMaxNeg + 2) * 4;
  const char* baseU = reinterpret_cast<const char*>(p.U) + lane * 16;
  const char* baseI = reinterpret_cast<const char*>(p.I) + lane * 16;
  const uint64_t row_bytes = (uint64_t)D * 4;
  const uint32_t rb32 = FULL ? 512u * R : (uint32_t)D * 4u;

  const bool implicit = p.loss <= CMF_LOSS_WARP;
  const bool adaptive = p.loss == CMF_LOSS_ADAPTIVE_BPR || p.loss == CMF_LOSS_ADAPTIVE_HINGE;
  const int64_t gwarp = (int64_t)blockIdx.x * 8 + warp;
  const int64_t n_warps = (int64_t)gridDim.x * 8;
  const int64_t n_chunks = (p.B + kChunkA - 1) / kChunkA;
  const int jrow = reduce8_row(lane);
  float loss_acc = 0.f;

  for (int64_t ch = gwarp; ch < n_chunks; ch += n_warps) {
    const int64_t b0 = ch * kChunkA;
    const int n_e = (int)((b0 + kChunkA < p.B ? b0 + kChunkA : p.B) - b0);

    // ---- ids of the chunk, lanes in parallel
    if (lane < n_e) {
      sts_i32(a_ids + lane * kRowStride, p.users[b0 + lane]);
      sts_i32(a_ids + lane * kRowStride + 4, p.items[b0 + lane]);
    }
    if (implicit) {
      if (p.negs_in) {
        for (int t = lane; t < n_e * N; t += 32) {
          const int le = t / N, n = t - le * N;
          sts_i32(a_ids + le * kRowStride + 8 + n * 4, (int)p.negs_in[(b0 + le) * N + n]);
        }
      } else if (p.indptr == nullptr) {
        // approximate sampling: one Philox block yields four consecutive draws
        const int nblk = (N + 3) >> 2;
        for (int t = lane; t < n_e * nblk; t += 32) {
          const int le = t / nblk, blk = t - le * nblk;
          const int64_t b = b0 + le;
          const int64_t gi = p.gidx ? p.gidx[b] : p.gidx_base + b;
          const uint4 w = philox4x32_10(
              make_uint4((uint32_t)gi, (uint32_t)((uint64_t)gi >> 32), p.epoch, (uint32_t)blk),
              make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32)));
          const uint32_t dst = a_ids + le * kRowStride + 8 + blk * 16;
          const int left = N - blk * 4;
          sts_i32(dst, (int)__umulhi(w.x, p.num_items));
          if (left > 1) sts_i32(dst + 4, (int)__umulhi(w.y, p.num_items));
          if (left > 2) sts_i32(dst + 8, (int)__umulhi(w.z, p.num_items));
          if (left > 3) sts_i32(dst + 12, (int)__umulhi(w.w, p.num_items));
        }
      } else if (lane < n_e) {
        // exact sampling: the reference's rejection loop (datasets.py:399-426) over our stream
        const uint32_t negs = a_ids + lane * kRowStride + 8;
        const int64_t b = b0 + lane;
        const int u = p.users[b];
        const int64_t gi = p.gidx ? p.gidx[b] : p.gidx_base + b;
        const int64_t lo = p.indptr[u], hi = p.indptr[u + 1];
        uint32_t d = 0;
        int taken = 0, checked = 0;
        while (taken < N) {
          int cand = draw_item(p.seed, p.epoch, gi, d++, p.num_items);
          for (;;) {
            bool clash = csr_contains(p.csr_items, lo, hi, cand);
            for (int j = 0; j < taken && !clash; ++j) clash = (lds_i32(negs + j * 4) == cand);
            if (!clash) break;
            if (checked >= p.max_tries) {
              const int left = N - taken - 1;
              for (int j = 0; j < left; ++j)
                sts_i32(negs + (taken++) * 4, draw_item(p.seed, p.epoch, gi, d++, p.num_items));
              break;
            }
            cand = draw_item(p.seed, p.epoch, gi, d++, p.num_items);
            ++checked;
          }
          sts_i32(negs + (taken++) * 4, cand);
        }
      }
    }
    __syncwarp();
    if (p.negs_out)
      for (int t = lane; t < n_e * N; t += 32) {
        const int le = t / N, n = t - le * N;
        p.negs_out[(b0 + le) * N + n] = lds_i32(a_ids + le * kRowStride + 8 + n * 4);
      }
    if (p.train) {
      // every gathered item row is "touched" (SparseAdam moves rows whose gradient is zero)
      for (int t = lane; t < n_e * Rw; t += 32) {
        const int le = t / Rw, n = t - le * Rw;
        mark_row(p.item_flag, p.item_list, &p.hdr->item_count,
                 lds_i32(a_ids + le * kRowStride + 4 + n * 4), p.use_lists);
      }
    }

    // ---- one interaction at a time; its rows are scored in groups of 8 held in registers
    for (int le = 0; le < n_e; ++le) {
      const int64_t b = b0 + le;
      const uint32_t t_ids = a_ids + le * kRowStride;
      const int u = lds_i32(t_ids);
      const int pos = lds_i32(t_ids + 4);
      float4 ur[R];
      {
        const char* src = baseU + (uint64_t)(uint32_t)u * row_bytes;
#pragma unroll
        for (int k = 0; k < R; ++k)
          ur[k] = (FULL || k * 32 + lane < D4) ? ldg_f4(src + k * 512) : make_float4(0, 0, 0, 0);
      }
      const float b_user = p.bu[u];

      float sp = 0.f, sel_s = 0.f, sel_h = 0.f;
      int sel = (p.loss == CMF_LOSS_WARP) ? N : 0;
      uint32_t best_key = 0;
      for (int j0 = 0; j0 < Rw; j0 += 8) {
        // the lane that will hold row j0 + jrow after the reduction fetches its id and bias
        const int myj = j0 + jrow;
        int my_id = 0;
        float my_bias = 0.f;
        if (myj < Rw) {
          my_id = lds_i32(t_ids + 4 + myj * 4);
          my_bias = p.bi[my_id];
        }
        float v[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          v[r] = 0.f;
          if (j0 + r < Rw) {   // warp-uniform
            const int id = lds_i32(t_ids + 4 + (j0 + r) * 4);
            const char* src = baseI + (uint64_t)(uint32_t)id * rb32;
#pragma unroll
            for (int k = 0; k < R; ++k) {
              if (FULL || k * 32 + lane < D4) {
                const float4 x = ldg_f4(src + k * 512);
                v[r] = fmaf(ur[k].x, x.x, v[r]);
                v[r] = fmaf(ur[k].y, x.y, v[r]);
                v[r] = fmaf(ur[k].z, x.z, v[r]);
                v[r] = fmaf(ur[k].w, x.w, v[r]);
              }
            }
          }
        }
        const float tot = reduce8(v, lane);
        const float my_score = (myj < Rw) ? tot + b_user + my_bias : -INFINITY;
        if (j0 == 0) {
          sp = __shfl_sync(0xffffffffu, my_score, 0);   // row 0 lives in lane 0
          if (p.pos_out && lane == 0) p.pos_out[b] = sp;
        }
        if (!implicit) break;
        const bool is_neg = (myj >= 1) && (myj < Rw);
        if (p.neg_out && is_neg && (lane & 3) == 0) p.neg_out[b * N + (myj - 1)] = my_score;
        // lanes are not ordered by row after reduce8, so "first" is found through a row-ordered key
        if (p.loss == CMF_LOSS_WARP) {
          // h = ideal - pos + neg; the first strict violation wins (warp.py:95-105, 136-154)
          float h = 0.f;
          if (is_neg) h = (ideal_of(p, pos, my_id) - sp) + my_score;
          const uint32_t key = (is_neg && h > 0.f) ? (uint32_t)(64 - jrow) : 0u;   // lowest row wins
          const uint32_t mx = __reduce_max_sync(0xffffffffu, key);
          if (sel == N && mx) {
            const int jr = 64 - (int)mx;
            sel = j0 + jr - 1;
            sel_h = __shfl_sync(0xffffffffu, h, reduce8_lane(jr));
          }
        } else if (adaptive) {
          // torch.max over the N axis: the first index of the maximum (bpr.py:153, hinge.py:157)
          const uint32_t key = is_neg ? float_key(my_score) : 0u;
          const uint32_t mx = __reduce_max_sync(0xffffffffu, key);
          if (mx > best_key) {
            const uint32_t rk = (is_neg && key == mx) ? (uint32_t)(64 - jrow) : 0u;
            const int jr = 64 - (int)__reduce_max_sync(0xffffffffu, rk);
            sel = j0 + jr - 1;
            sel_s = __shfl_sync(0xffffffffu, my_score, reduce8_lane(jr));
            best_key = mx;
          }
        } else if (j0 == 0) {
          sel = 0;
          sel_s = __shfl_sync(0xffffffffu, my_score, reduce8_lane(1));
        }
      }

      // ---- loss value and d loss / d score
      float g = 0.f;       // implicit: d/d pos (d/d selected neg is -g); explicit: d/d score
      int sel_id = -1;
      if (implicit) {
        if (p.sel_out && lane == 0) p.sel_out[b] = sel;
        if (p.loss == CMF_LOSS_WARP) {
          if (sel < N) {
            sel_id = lds_i32(t_ids + 8 + sel * 4);
            const float w = logf((float)p.num_items / (float)(sel + 1));
            const float Lw = w * sel_h;
            loss_acc += Lw + Lw * Lw;
            g = -((1.f + 2.f * Lw) * w) * p.inv_B;
          }
        } else {
          sel_id = lds_i32(t_ids + 8 + sel * 4);
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
      } else {
        float sg = 0.f, pred = sp;
        if (p.has_range) { sg = sigmoidf_(sp); pred = sg * (p.y_hi - p.y_lo) + p.y_lo; }
        if (p.pos_out && lane == 0) p.pos_out[b] = pred;
        const float diff = pred - p.ratings[b];
        float dp;
        if (p.loss == CMF_LOSS_MSE) { loss_acc += diff * diff; dp = 2.f * diff * p.inv_B; }
        else { loss_acc += fabsf(diff); dp = (diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f)) * p.inv_B; }
        g = p.has_range ? dp * (sg * (1.f - sg) * (p.y_hi - p.y_lo)) : dp;
      }
      if (p.train && lane == 0) {
        p.g_arr[b] = g;
        p.sel_arr[b] = sel_id;
      }
    }
    __syncwarp();   // the id table is rewritten by the next chunk
  }

  // ---- loss: deterministic two-level reduction (every lane carries the same loss_acc)
  if (lane == 0) s_warp_loss[warp] = loss_acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int w = 0; w < 8; ++w) tot += (double)s_warp_loss[w];
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

// ================================================================================ k_fwd_loss_g8
// Same work as k_fwd_loss for D <= 128: a group of 8 lanes scores one interaction (lane l owns
// float4 l, l + 8, ... of every row), so one warp instruction serves 4 interactions and the scalar
// work per interaction (ids, selection, loss) is shared four ways. KQ float4 per lane per row.
template <int KQ, bool FULL>
__global__ void __launch_bounds__(256, 2)
k_fwd_loss_g8(const StepDev p, const int NP) {
  // the backward kernel may become resident as this grid's CTAs retire (it waits for the whole
  // grid before it reads anything)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  extern __shared__ int s_dyn[];   // [8 warps][32 interactions][NP]: user, positive, negatives
  __shared__ float s_warp_loss[8];
  __shared__ bool s_last;

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int grp = lane >> 3, l = lane & 7;
  const unsigned gmask = 0xFFu << (grp * 8);
  const int D = p.D, N = p.N;
  const int D4 = FULL ? 8 * KQ : (D >> 2);
  const int Rw = N + 1;
  const uint32_t a_ids = smem_u32(s_dyn) + (uint32_t)(warp * 32 * NP) * 4;
  const uint32_t row_stride = (uint32_t)NP * 4;
  const uint32_t rb = FULL ? 128u * KQ : (uint32_t)D * 4u;   // row bytes
  const char* baseU = reinterpret_cast<const char*>(p.U) + l * 16;
  const char* baseI = reinterpret_cast<const char*>(p.I) + l * 16;

  const bool implicit = p.loss <= CMF_LOSS_WARP;
  const bool adaptive = p.loss == CMF_LOSS_ADAPTIVE_BPR || p.loss == CMF_LOSS_ADAPTIVE_HINGE;
  const bool is_warp = p.loss == CMF_LOSS_WARP;
  const bool is_bpr = p.loss == CMF_LOSS_BPR || p.loss == CMF_LOSS_ADAPTIVE_BPR;
  const int64_t gwarp = (int64_t)blockIdx.x * 8 + warp;
  const int64_t n_warps = (int64_t)gridDim.x * 8;
  const int64_t n_chunks = (p.B + 31) / 32;
  float loss_acc = 0.f;

  for (int64_t ch = gwarp; ch < n_chunks; ch += n_warps) {
    const int64_t b0 = ch * 32;
    const int n_e = (int)((b0 + 32 < p.B ? b0 + 32 : p.B) - b0);

    // ---- ids of 32 interactions, lanes in parallel
    if (lane < n_e) {
      sts_i32(a_ids + lane * row_stride, p.users[b0 + lane]);
      sts_i32(a_ids + lane * row_stride + 4, p.items[b0 + lane]);
    }
    if (implicit) {
      if (p.negs_in) {
        for (int t = lane; t < n_e * N; t += 32) {
          const int le = t / N, n = t - le * N;
          sts_i32(a_ids + le * row_stride + 8 + n * 4, (int)p.negs_in[(b0 + le) * N + n]);
        }
      } else if (p.indptr == nullptr) {
        // approximate sampling: one Philox block yields four consecutive draws
        const int nblk = (N + 3) >> 2;
        for (int t = lane; t < n_e * nblk; t += 32) {
          const int le = t / nblk, blk = t - le * nblk;
          const int64_t b = b0 + le;
          const int64_t gi = p.gidx ? p.gidx[b] : p.gidx_base + b;
          const uint4 w = philox4x32_10(
              make_uint4((uint32_t)gi, (uint32_t)((uint64_t)gi >> 32), p.epoch, (uint32_t)blk),
              make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32)));
          const uint32_t dst = a_ids + le * row_stride + 8 + blk * 16;
          const int left = N - blk * 4;
          sts_i32(dst, (int)__umulhi(w.x, p.num_items));
          if (left > 1) sts_i32(dst + 4, (int)__umulhi(w.y, p.num_items));
          if (left > 2) sts_i32(dst + 8, (int)__umulhi(w.z, p.num_items));
          if (left > 3) sts_i32(dst + 12, (int)__umulhi(w.w, p.num_items));
        }
      } else if (lane < n_e) {
        // exact sampling: the reference's rejection loop (datasets.py:399-426) over our stream
        const uint32_t negs = a_ids + lane * row_stride + 8;
        const int64_t b = b0 + lane;
        const int u = p.users[b];
        const int64_t gi = p.gidx ? p.gidx[b] : p.gidx_base + b;
        const int64_t lo = p.indptr[u], hi = p.indptr[u + 1];
        uint32_t d = 0;
        int taken = 0, checked = 0;
        while (taken < N) {
          int cand = draw_item(p.seed, p.epoch, gi, d++, p.num_items);
          for (;;) {
            bool clash = csr_contains(p.csr_items, lo, hi, cand);
            for (int j = 0; j < taken && !clash; ++j) clash = (lds_i32(negs + j * 4) == cand);
            if (!clash) break;
            if (checked >= p.max_tries) {
              const int left = N - taken - 1;
              for (int j = 0; j < left; ++j)
                sts_i32(negs + (taken++) * 4, draw_item(p.seed, p.epoch, gi, d++, p.num_items));
              break;
            }
            cand = draw_item(p.seed, p.epoch, gi, d++, p.num_items);
            ++checked;
          }
          sts_i32(negs + (taken++) * 4, cand);
        }
      }
    }
    __syncwarp();
    if (p.negs_out)
      for (int t = lane; t < n_e * N; t += 32) {
        const int le = t / N, n = t - le * N;
        p.negs_out[(b0 + le) * N + n] = lds_i32(a_ids + le * row_stride + 8 + n * 4);
      }
    if (p.train) {
      // every gathered item row is "touched" (SparseAdam moves rows whose gradient is zero)
      for (int t = lane; t < n_e * Rw; t += 32) {
        const int le = t / Rw, n = t - le * Rw;
        mark_row(p.item_flag, p.item_list, &p.hdr->item_count,
                 lds_i32(a_ids + le * row_stride + 4 + n * 4), p.use_lists);
      }
    }

    // ---- 8 rounds; in each, the 4 lane groups score 4 interactions side by side
    for (int it = 0; it < 8; ++it) {
      const int le = it * 4 + grp;
      if (le >= n_e) continue;   // only in the last chunk; shuffles below use the group mask
      const int64_t b = b0 + le;
      const uint32_t t_ids = a_ids + le * row_stride;
      const int u = lds_i32(t_ids);
      const int pos = lds_i32(t_ids + 4);
      float4 ur[KQ];
      {
        const char* src = baseU + (uint64_t)(uint32_t)u * rb;
#pragma unroll
        for (int k = 0; k < KQ; ++k)
          ur[k] = (FULL || k * 8 + l < D4) ? ldg_f4(src + k * 128) : make_float4(0, 0, 0, 0);
      }
      const float b_user = p.bu[u];

      float sp = 0.f, sel_s = -INFINITY, sel_h = 0.f;
      int sel = is_warp ? N : 0;
      // rows 0 (positive), 1..N (negatives), three at a time so 3 * KQ loads are in flight
      for (int j0 = 0; j0 < Rw; j0 += 3) {
        int id[3];
        float4 x[3][KQ];
        float bia[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const bool on = j0 + r < Rw;
          id[r] = on ? lds_i32(t_ids + 4 + (j0 + r) * 4) : pos;
          const char* src = baseI + (uint64_t)(uint32_t)id[r] * rb;
#pragma unroll
          for (int k = 0; k < KQ; ++k)
            x[r][k] = (FULL || k * 8 + l < D4) ? ldg_f4(src + k * 128) : make_float4(0, 0, 0, 0);
          bia[r] = p.bi[id[r]];
        }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          float a0 = 0.f, a1 = 0.f;
#pragma unroll
          for (int k = 0; k < KQ; ++k) {
            a0 = fmaf(ur[k].x, x[r][k].x, a0);
            a1 = fmaf(ur[k].y, x[r][k].y, a1);
            a0 = fmaf(ur[k].z, x[r][k].z, a0);
            a1 = fmaf(ur[k].w, x[r][k].w, a1);
          }
          float acc = a0 + a1;
          acc += __shfl_xor_sync(gmask, acc, 4);
          acc += __shfl_xor_sync(gmask, acc, 2);
          acc += __shfl_xor_sync(gmask, acc, 1);
          const int j = j0 + r;
          if (j >= Rw) continue;
          const float sc = acc + b_user + bia[r];
          if (j == 0) {
            sp = sc;
            if (p.pos_out && l == 0) p.pos_out[b] = sp;
            continue;
          }
          if (p.neg_out && l == 0) p.neg_out[b * N + (j - 1)] = sc;
          if (is_warp) {
            // h = ideal - pos + neg; the first strict violation wins (warp.py:95-105, 136-154)
            const float h = (ideal_of(p, pos, id[r]) - sp) + sc;
            if (sel == N && h > 0.f) { sel = j - 1; sel_h = h; }
          } else if (adaptive) {
            // torch.max over the N axis: the first index of the maximum (bpr.py:153, hinge.py:157)
            if (j == 1 || sc > sel_s) { sel = j - 1; sel_s = sc; }
          } else if (j == 1) {
            sel = 0;
            sel_s = sc;
          }
        }
      }

      // ---- loss value and d loss / d score
      float g = 0.f, lval = 0.f;   // implicit: d/d pos (d/d selected neg is -g)
      int sel_id = -1;
      if (implicit) {
        if (p.sel_out && l == 0) p.sel_out[b] = sel;
        if (is_warp) {
          if (sel < N) {
            sel_id = lds_i32(t_ids + 8 + sel * 4);
            const float w = logf((float)p.num_items / (float)(sel + 1));
            const float Lw = w * sel_h;
            lval = Lw + Lw * Lw;
            g = -((1.f + 2.f * Lw) * w) * p.inv_B;
          }
        } else {
          sel_id = lds_i32(t_ids + 8 + sel * 4);
          const float ideal = ideal_of(p, pos, sel_id);
          const float xd = sp - sel_s;
          if (is_bpr) {
            const float sg = sigmoidf_(xd);
            const float lv = ideal - sg;
            lval = lv + lv * lv;
            g = -((1.f + 2.f * lv) * (sg * (1.f - sg))) * p.inv_B;
          } else {
            const float raw = ideal - xd;
            const float lv = raw > 0.f ? raw : 0.f;
            lval = lv + lv * lv;
            g = (raw >= 0.f) ? -(1.f + 2.f * lv) * p.inv_B : 0.f;
          }
        }
      } else {
        float sg = 0.f, pred = sp;
        if (p.has_range) { sg = sigmoidf_(sp); pred = sg * (p.y_hi - p.y_lo) + p.y_lo; }
        if (p.pos_out && l == 0) p.pos_out[b] = pred;
        const float diff = pred - p.ratings[b];
        float dp;
        if (p.loss == CMF_LOSS_MSE) { lval = diff * diff; dp = 2.f * diff * p.inv_B; }
        else { lval = fabsf(diff); dp = (diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f)) * p.inv_B; }
        g = p.has_range ? dp * (sg * (1.f - sg) * (p.y_hi - p.y_lo)) : dp;
      }
      if (l == 0) {
        loss_acc += lval;
        if (p.train) {
          p.g_arr[b] = g;
          p.sel_arr[b] = sel_id;
        }
      }
      if (p.red_in_fwd && g != 0.f) {
        // item side of the backward pass, here, where the user row is in registers anyway:
        // d loss / d item row = +-g * user row (products rounded before they are added, as in
        // k_bwd_update), sent with vector RED; the backward kernel keeps the user side
        const int tgt = implicit ? sel_id : pos;
        float* gp = p.GI + (int64_t)pos * D + l * 4;
        float* gs = p.GI + (int64_t)tgt * D + l * 4;
#pragma unroll
        for (int k = 0; k < KQ; ++k) {
          if (FULL || k * 8 + l < D4) {
            float tt[4] = {__fmul_rn(g, ur[k].x), __fmul_rn(g, ur[k].y), __fmul_rn(g, ur[k].z),
                           __fmul_rn(g, ur[k].w)};
            red_vec<4>(gp + k * 32, tt);
            if (implicit) {
              tt[0] = -tt[0]; tt[1] = -tt[1]; tt[2] = -tt[2]; tt[3] = -tt[3];
              red_vec<4>(gs + k * 32, tt);
            }
          }
        }
        if (l == 0) {
          atomicAdd(p.gbi + pos, g);
          if (implicit) atomicAdd(p.gbi + sel_id, -g);
        }
      }
    }
    __syncwarp();   // the id table is rewritten by the next chunk
  }

  // ---- loss: lane 0 of every group carries its share; deterministic two-level reduction
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) loss_acc += __shfl_xor_sync(0xffffffffu, loss_acc, o);
  if (lane == 0) s_warp_loss[warp] = loss_acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int w = 0; w < 8; ++w) tot += (double)s_warp_loss[w];
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

// =================================================================================== k_bwd_update
template <int R, bool FULL>
__global__ void __launch_bounds__(256)
k_bwd_update(const StepDev p) {
  // (launched as a programmatic dependent of the forward kernel: resident early, starts here)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // k_update_rows may line up
  // per warp and interaction of the chunk: user, positive, selected negative, coefficient
  __shared__ int s_tab[8][kMaxChunkB][4];
  __shared__ int s_meta[8][2];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int D = p.D;
  const int D4 = FULL ? 32 * R : (D >> 2);
  const uint32_t a_tab = smem_u32(&s_tab[warp][0][0]);
  const uint32_t a_meta = smem_u32(&s_meta[warp][0]);
  const char* baseU = reinterpret_cast<const char*>(p.U) + lane * 16;
  const char* baseI = reinterpret_cast<const char*>(p.I) + lane * 16;
  const uint64_t row_bytes = (uint64_t)D * 4;
  const bool implicit = p.loss <= CMF_LOSS_WARP;
  const bool emb_m = rule_has_m(p.emb.rule), emb_v = rule_has_v(p.emb.rule);
  const bool bias_m = rule_has_m(p.bias.rule), bias_v = rule_has_v(p.bias.rule);
  const int64_t gwarp = (int64_t)blockIdx.x * 8 + warp;
  const int64_t n_warps = (int64_t)gridDim.x * 8;

  const bool aligned = p.chunk_starts != nullptr;   // chunks cut at user boundaries: no straddlers
  for (int64_t ch = gwarp; ch < p.num_chunks; ch += n_warps) {
    const int64_t c0 = aligned ? (int64_t)p.chunk_starts[ch] : ch * p.L;
    const int64_t c1 = aligned ? (int64_t)p.chunk_starts[ch + 1]
                               : (c0 + p.L < p.B ? c0 + p.L : p.B);
    if (c0 >= c1) continue;
    // (an aligned chunk can be longer than the table: it is walked in table-sized pieces)
    int64_t e0 = c0;
    int n_e = (int)(c1 - c0 < kMaxChunkB ? c1 - c0 : kMaxChunkB);
    auto load_table = [&]() {
      if (lane < n_e) {
        const int64_t e = e0 + lane;
        const int64_t b = p.perm ? (int64_t)p.perm[e] : e;
        const uint32_t t = a_tab + lane * 16;
        sts_i32(t, p.su[e]);
        sts_i32(t + 4, p.items[b]);
        sts_i32(t + 8, p.sel_arr[b]);
        sts_f32(t + 12, p.g_arr[b]);
      }
    };
    load_table();
    // the users just outside the chunk decide whether a boundary segment is whole
    if (!aligned) {
      if (lane == 30) sts_i32(a_meta, (e0 > 0) ? p.su[e0 - 1] : -1);
      if (lane == 31) sts_i32(a_meta + 4, (e0 + n_e < p.B) ? p.su[e0 + n_e] : -1);
    }
    __syncwarp();

    int cur_user = -1;
    bool left_whole = false;
    Row<4, 32, R> gu, pr, m_pre, v_pre;  // gradient, parameter row, Adam moments of the user
    float gbu_acc = 0.f, bu_cur = 0.f, bm_pre = 0.f, bv_pre = 0.f;

    auto finish_user = [&](bool right_whole) {
      if (cur_user < 0) return;
      if (left_whole && right_whole && p.user_inplace) {
        const int64_t off = (int64_t)cur_user * D;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int i = r * 32 + lane;
          if (FULL || i < D4) {
            apply_rule_n<4>(p.emb, pr.f + r * 4, m_pre.f + r * 4, v_pre.f + r * 4, gu.f + r * 4);
            st_vec<4>(p.U + off + i * 4, pr.f + r * 4);
            if (emb_m) st_vec<4>(p.U_m + off + i * 4, m_pre.f + r * 4);
            if (emb_v) st_vec<4>(p.U_v + off + i * 4, v_pre.f + r * 4);
          }
        }
        if (lane == 0) {
          float pb = bu_cur;
          apply_rule(p.bias, pb, bm_pre, bv_pre, gbu_acc);
          p.bu[cur_user] = pb;
          if (bias_m) p.bu_m[cur_user] = bm_pre;
          if (bias_v) p.bu_v[cur_user] = bv_pre;
        }
      } else {
        gu.red_add(p.GU + (int64_t)cur_user * D, D, lane);
        if (lane == 0) {
          atomicAdd(p.gbu + cur_user, gbu_acc);
          mark_row(p.user_flag, p.user_list, &p.hdr->user_count, cur_user, p.use_lists);
        }
      }
    };

    for (;;) {
    for (int le = 0; le < n_e; ++le) {
      const uint32_t t = a_tab + le * 16;
      const int u = lds_i32(t);
      const int pos = lds_i32(t + 4);
      const int sel_id = lds_i32(t + 8);
      const float g = lds_f32(t + 12);
      // issue this interaction's item-row loads first so they overlap the user bookkeeping
      const bool active = g != 0.f;
      float4 xp[R], xs[R];
      if (active) {
        const char* sp_ = baseI + (uint64_t)(uint32_t)pos * row_bytes;
        const char* ss_ = baseI + (uint64_t)(uint32_t)(implicit ? sel_id : pos) * row_bytes;
#pragma unroll
        for (int k = 0; k < R; ++k) {
          const bool ok = FULL || k * 32 + lane < D4;
          xp[k] = ok ? ldg_f4(sp_ + k * 512) : make_float4(0, 0, 0, 0);
          xs[k] = (ok && implicit) ? ldg_f4(ss_ + k * 512) : make_float4(0, 0, 0, 0);
        }
      }
      if (u != cur_user) {
        finish_user(true);   // another user follows inside the chunk
        cur_user = u;
        left_whole = aligned || (le > 0) || (lds_i32(a_meta) != u);
        gu.zero();
        gbu_acc = 0.f;
        m_pre.zero();
        v_pre.zero();
        bm_pre = 0.f;
        bv_pre = 0.f;
        const int64_t off = (int64_t)u * D;
        pr.load(p.U + off, D, lane);
        bu_cur = p.bu[u];
        if (p.user_inplace) {
          if (emb_m) m_pre.load(p.U_m + off, D, lane);
          if (emb_v) v_pre.load(p.U_v + off, D, lane);
          if (lane == 0) {
            if (bias_m) bm_pre = p.bu_m[u];
            if (bias_v) bv_pre = p.bu_v[u];
          }
        }
      }
      if (active) {
        float* gi_pos = p.GI + (int64_t)pos * D + lane * 4;
        float* gi_sel = p.GI + (int64_t)(implicit ? sel_id : pos) * D + lane * 4;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          if (FULL || r * 32 + lane < D4) {
            float* f = gu.f + r * 4;
            const float* pu = pr.f + r * 4;
            // products rounded before the add, so +g*x and -g*x cancel exactly (see Row::axpy)
            f[0] = __fadd_rn(f[0], __fmul_rn(g, xp[r].x));
            f[1] = __fadd_rn(f[1], __fmul_rn(g, xp[r].y));
            f[2] = __fadd_rn(f[2], __fmul_rn(g, xp[r].z));
            f[3] = __fadd_rn(f[3], __fmul_rn(g, xp[r].w));
            float tt[4] = {__fmul_rn(g, pu[0]), __fmul_rn(g, pu[1]), __fmul_rn(g, pu[2]),
                           __fmul_rn(g, pu[3])};
            if (!p.bwd_skip_item) red_vec<4>(gi_pos + r * 128, tt);
            if (implicit) {
              f[0] = __fadd_rn(f[0], __fmul_rn(-g, xs[r].x));
              f[1] = __fadd_rn(f[1], __fmul_rn(-g, xs[r].y));
              f[2] = __fadd_rn(f[2], __fmul_rn(-g, xs[r].z));
              f[3] = __fadd_rn(f[3], __fmul_rn(-g, xs[r].w));
              tt[0] = -tt[0]; tt[1] = -tt[1]; tt[2] = -tt[2]; tt[3] = -tt[3];
              if (!p.bwd_skip_item) red_vec<4>(gi_sel + r * 128, tt);
            }
          }
        }
        if (implicit) {
          gbu_acc += g;
          gbu_acc += -g;
          if (lane == 0 && !p.bwd_skip_item) {
            atomicAdd(p.gbi + pos, g);
            atomicAdd(p.gbi + sel_id, -g);
          }
        } else {
          gbu_acc += g;
          if (lane == 0 && !p.bwd_skip_item) atomicAdd(p.gbi + pos, g);
        }
      }
    }
      e0 += n_e;
      if (e0 >= c1) break;
      __syncwarp();   // next piece of a long aligned chunk: the user in progress carries over
      n_e = (int)(c1 - e0 < kMaxChunkB ? c1 - e0 : kMaxChunkB);
      load_table();
      __syncwarp();
    }
    finish_user(aligned || lds_i32(a_meta + 4) != cur_user);
    __syncwarp();   // the table is rewritten by the next chunk
  }
}

static int blocks_for(int64_t warps_of_work, int per_sm) {
  int64_t blocks = (warps_of_work + 7) / 8;
  const int64_t cap = (int64_t)num_sms() * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (blocks > kMaxPartials) blocks = kMaxPartials;
  return (int)blocks;
}

bool v3_supported(int D, int N) { return D % 4 == 0 && D <= 1024 && N <= kMaxNeg; }

template <int KQ>
static cudaError_t launch_g8(const StepDev& p, bool full, int grid, cudaStream_t s) {
  const int NP = p.N + 2;
  const size_t smem = (size_t)8 * 32 * NP * sizeof(int);
  cudaError_t e = cudaSuccess;
  if (full) {
    if (smem > 48 * 1024)
      e = cudaFuncSetAttribute(k_fwd_loss_g8<KQ, true>,
                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) k_fwd_loss_g8<KQ, true><<<grid, 256, smem, s>>>(p, NP);
  } else {
    if (smem > 48 * 1024)
      e = cudaFuncSetAttribute(k_fwd_loss_g8<KQ, false>,
                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) k_fwd_loss_g8<KQ, false><<<grid, 256, smem, s>>>(p, NP);
  }
  return e == cudaSuccess ? cudaGetLastError() : e;
}

int launch_fwd_loss_v3(const StepDev& p, bool sort_runs_alongside, cudaStream_t s) {
  const int D4 = p.D / 4;
  if (D4 <= 32) {
    // 8 lanes per interaction, 4 interactions per warp instruction
    const int KQ = (D4 + 7) / 8;
    const bool full = (D4 == 8 * KQ);
    // one resident wave (2 CTAs of 124 registers per SM): a second wave costs 8 % of the kernel.
    // With the batch sort running next to it on the side stream, smaller CTAs-in-waves leave the
    // sort's kernels a way in.
    const int grid = blocks_for((p.B + 31) / 32, sort_runs_alongside ? 4 : 2);
    cudaError_t e;
    if (KQ == 1) e = launch_g8<1>(p, full, grid, s);
    else if (KQ == 2) e = launch_g8<2>(p, full, grid, s);
    else if (KQ == 3) e = launch_g8<3>(p, full, grid, s);
    else e = launch_g8<4>(p, full, grid, s);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (e != cudaSuccess)
      return fail(CMF_ERR_CUDA, "k_fwd_loss_g8 launch failed: %s", cudaGetErrorString(e));
    return CMF_OK;
  }
  const int R = (D4 + 31) / 32;
  const bool full = (p.D == 128 * R);
  const int grid = blocks_for((p.B + kChunkA - 1) / kChunkA, 8);
#define K1(R_)                                                         \
  do {                                                                 \
    if (full) k_fwd_loss<R_, true><<<grid, 256, 0, s>>>(p);            \
    else k_fwd_loss<R_, false><<<grid, 256, 0, s>>>(p);                \
  } while (0)
  if (R == 2) K1(2); else if (R <= 4) K1(4); else K1(8);
#undef K1
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

static int pdl_default() {
  const char* e = getenv("CMF_PDL");
  return e ? atoi(e) : 0;
}
// cmf_set_tuning key 6 (or CMF_PDL=1): programmatic dependent launch of k_bwd_update / k_update_rows.
// Off by default: it takes 1.3 % off a device-resident step, but the dependents fill every SM the
// moment the previous grid drains, so kernels of OTHER streams (a host-fed pipeline's batch sort on
// its copy stream) find no way in until the step ends: end-to-end throughput fell by 12 %.
int g_pdl = pdl_default();
// cmf_set_tuning key 8 (or CMF_RED_IN_FWD=1): item-row gradient REDs issued by k_fwd_loss_g8
static int red_in_fwd_default() {
  const char* e = getenv("CMF_RED_IN_FWD");
  return e ? atoi(e) : 0;
}
int g_red_in_fwd = red_in_fwd_default();

int launch_aligned_chunks(const StepDev& p, int32_t* starts, cudaStream_t s);

int launch_bwd_update_v3(const StepDev& p_in, int max_chunk_len, cudaStream_t s,
                         int32_t* aligned_chunk_starts /* = nullptr */) {
  StepDev p = p_in;
  const int R = (p.D / 4 + 31) / 32;
  const bool full = (p.D == 128 * R);
  // long chunks leave fewer straddling users; every warp should still get a few chunks
  int L = kMaxChunkB;
  const int64_t warps = (int64_t)num_sms() * 8 * 4;
  while (L > 1 && p.B / L < 2 * warps) L >>= 1;
  if (max_chunk_len > 0) L = max_chunk_len < kMaxChunkB ? max_chunk_len : kMaxChunkB;
  p.L = L;
  p.num_chunks = (p.B + L - 1) / L;
  if (aligned_chunk_starts) {
    // deterministic step: chunks cut at user boundaries
    const int rc = launch_aligned_chunks(p, aligned_chunk_starts, s);
    if (rc != CMF_OK) return rc;
    p.chunk_starts = aligned_chunk_starts;
  }
  const int grid = blocks_for(p.num_chunks, 8);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(256);
  cfg.stream = s;
  cudaLaunchAttribute pdl[1];
  pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  pdl[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = pdl;
  cfg.numAttrs = g_pdl ? 1 : 0;
#define K2(R_)                                                                       \
  do {                                                                               \
    if (full) cudaLaunchKernelEx(&cfg, k_bwd_update<R_, true>, p);                   \
    else cudaLaunchKernelEx(&cfg, k_bwd_update<R_, false>, p);                       \
  } while (0)
  if (R <= 1) K2(1); else if (R == 2) K2(2); else if (R <= 4) K2(4); else K2(8);
#undef K2
  CMF_LAUNCH_CHECK();
  return CMF_OK;
}

}  // namespace cmf
