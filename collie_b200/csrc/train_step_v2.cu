// Fused MF training step, TMA-staged variant (the fast path; D % 4 == 0).
//
// Same semantics and work decomposition as k_fused_step in train_step.cu (batch ordered by user,
// cut into chunks of L interactions, whole user segments finished in place, straddlers and items
// through vector RED into the gradient tables) -- what changes is how a warp spends instructions:
//
//   phase A  per chunk, lanes in parallel: batch positions, item ids, Philox negatives, biases and
//            "touched" marks of all L * (N + 1) item rows land in a small shared-memory table.
//   phase B  per interaction: the N + 2 rows it needs (user, positive, negatives) are fetched by
//            1-D TMA bulk copies (cp.async.bulk, one instruction per 512-byte row) into a
//            kStages-deep shared-memory ring guarded by mbarriers, so the loads of interaction
//            e + kStages - 1 are in flight while interaction e is being computed; the N + 1 dot
//            products are done at once with a (row, segment) lane mapping that needs log2(St)
//            shuffle steps instead of 5 per dot.
//
// Persistent launch: one CTA per SM, W warps, W * warp_bytes of dynamic shared memory.
#include "train_common.cuh"

namespace cmf {

constexpr int kStages = 3;
constexpr int kMaxChunk = 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
// 1-D TMA bulk copy global -> shared, completion counted on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ float4 lds128(const unsigned char* p) {
  return *reinterpret_cast<const float4*>(p);
}
// float -> uint32 whose unsigned order is the float order (NaN aside)
__device__ __forceinline__ uint32_t float_key(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

struct V2Cfg {
  int W;            // warps per CTA
  int L;            // interactions per chunk (<= kMaxChunk)
  int NP;           // table columns per interaction: user, positive, N negatives
  int St, Jt;       // lanes per row segment group / rows handled at once (St * Jt == 32)
  int D4;           // float4 per row
  int row_bytes;    // D * 4
  int stride;       // shared-memory row pitch (bank-conflict-free for the (row, segment) mapping)
  int off_ids, off_bias, off_aux, off_rows, stage_bytes, warp_bytes;
};

template <int R>  // R float4 per lane hold a 128 * R-float gradient row
__global__ void __launch_bounds__(512, 1)
k_fused_step_v2(const StepDev p, const V2Cfg c) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ float s_warp_loss[16];
  __shared__ bool s_last;

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  unsigned char* wbase = smem + (size_t)warp * c.warp_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(wbase);
  int* ids = reinterpret_cast<int*>(wbase + c.off_ids);       // [L][NP]
  float* bias = reinterpret_cast<float*>(wbase + c.off_bias); // [L][NP]
  int* bpos = reinterpret_cast<int*>(wbase + c.off_aux);      // [L] batch position
  float* aux = reinterpret_cast<float*>(wbase + c.off_aux) + kMaxChunk;  // [L] rating
  unsigned char* rows = wbase + c.off_rows;

  if (lane == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();

  const int D = p.D, N = p.N, NP = c.NP, St = c.St;
  const int Rw = N + 1;  // item rows per interaction
  const bool implicit = p.loss <= CMF_LOSS_WARP;
  const bool adaptive = p.loss == CMF_LOSS_ADAPTIVE_BPR || p.loss == CMF_LOSS_ADAPTIVE_HINGE;
  const int64_t gwarp = (int64_t)blockIdx.x * c.W + warp;
  const int64_t n_warps = (int64_t)gridDim.x * c.W;
  const int seg = lane % St;   // (row, segment) mapping of the dot products
  const int jrow = lane / St;
  const uint32_t tx_bytes = (uint32_t)(NP * c.row_bytes);

  uint32_t issued = 0, consumed = 0;
  float loss_acc = 0.f;

  for (int64_t ch = gwarp; ch < p.num_chunks; ch += n_warps) {
    const int64_t e0 = ch * c.L;
    const int n_e = (int)((e0 + c.L < p.B ? e0 + c.L : p.B) - e0);

    // ------------------------------------------------------------------ phase A: the id table
    if (lane < n_e) {
      const int64_t e = e0 + lane;
      const int u = p.su[e];
      const int64_t b = p.perm ? (int64_t)p.perm[e] : e;
      const int pos = p.items[b];
      ids[lane * NP + 0] = u;
      ids[lane * NP + 1] = pos;
      bpos[lane] = (int)b;
      bias[lane * NP + 0] = p.bu[u];
      bias[lane * NP + 1] = p.bi[pos];
      if (!implicit) aux[lane] = p.ratings[b];
    }
    __syncwarp();
    if (implicit) {
      if (p.negs_in) {
        for (int t = lane; t < n_e * N; t += 32) {
          const int le = t / N, n = t - le * N;
          ids[le * NP + 2 + n] = (int)p.negs_in[(int64_t)bpos[le] * N + n];
        }
      } else if (p.indptr == nullptr) {
        // approximate sampling: one Philox block yields four consecutive draws
        const int nblk = (N + 3) >> 2;
        for (int t = lane; t < n_e * nblk; t += 32) {
          const int le = t / nblk, blk = t - le * nblk;
          const int64_t b = bpos[le];
          const int64_t gi = p.gidx ? p.gidx[b] : p.gidx_base + b;
          const uint4 w = philox4x32_10(
              make_uint4((uint32_t)gi, (uint32_t)((uint64_t)gi >> 32), p.epoch, (uint32_t)blk),
              make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32)));
          int* dst = ids + le * NP + 2 + blk * 4;
          const int left = N - blk * 4;
          dst[0] = (int)__umulhi(w.x, p.num_items);
          if (left > 1) dst[1] = (int)__umulhi(w.y, p.num_items);
          if (left > 2) dst[2] = (int)__umulhi(w.z, p.num_items);
          if (left > 3) dst[3] = (int)__umulhi(w.w, p.num_items);
        }
      } else if (lane < n_e) {
        // exact sampling: the reference's rejection loop (datasets.py:399-426) over our stream
        int* negs = ids + lane * NP + 2;
        const int u = ids[lane * NP];
        const int64_t b = bpos[lane];
        const int64_t gi = p.gidx ? p.gidx[b] : p.gidx_base + b;
        const int64_t lo = p.indptr[u], hi = p.indptr[u + 1];
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
      __syncwarp();
      for (int t = lane; t < n_e * N; t += 32) {
        const int le = t / N, n = t - le * N;
        const int id = ids[le * NP + 2 + n];
        bias[le * NP + 2 + n] = p.bi[id];
        if (p.negs_out) p.negs_out[(int64_t)bpos[le] * N + n] = id;
      }
    }
    if (p.train) {
      // every gathered item row is "touched" (SparseAdam moves rows whose gradient is zero)
      for (int t = lane; t < n_e * Rw; t += 32) {
        const int le = t / Rw, n = t - le * Rw;
        mark_row(p.item_flag, p.item_list, &p.hdr->item_count, ids[le * NP + 1 + n], p.use_lists);
      }
    }
    __syncwarp();

    // ------------------------------------------------------------------ phase B: the row pipeline
    auto issue = [&](int le) {
      unsigned char* dst = rows + (size_t)(issued % kStages) * c.stage_bytes;
      uint64_t* bar = &bars[issued % kStages];
      if (lane == 0) mbar_expect_tx(bar, tx_bytes);
      __syncwarp();
      if (lane < NP) {
        const int id = ids[le * NP + lane];
        const float* src = (lane == 0 ? p.U : p.I) + (int64_t)id * D;
        bulk_g2s(dst + (size_t)lane * c.stride, src, (uint32_t)c.row_bytes, bar);
      }
      for (int r = 32; r < NP; r += 32) {  // more than 30 negatives
        if (r + lane < NP) {
          const int id = ids[le * NP + r + lane];
          bulk_g2s(dst + (size_t)(r + lane) * c.stride, p.I + (int64_t)id * D,
                   (uint32_t)c.row_bytes, bar);
        }
      }
      ++issued;
    };
    for (int k = 0; k < kStages - 1 && k < n_e; ++k) issue(k);

    int cur_user = -1;
    bool left_whole = false;
    Row<4, 32, R> gu;
    float gbu_acc = 0.f;

    auto finish_user = [&](int next_le) {
      // next_le < n_e: the following entry of this chunk belongs to another user
      if (cur_user < 0 || !p.train) return;
      const int64_t e1 = e0 + n_e;
      const bool right_whole = (next_le < n_e) || (e1 == p.B) || (p.su[e1] != cur_user);
      if (left_whole && right_whole && p.user_inplace) {
        update_row<4, 32, R>(p.emb, p.U, p.U_m, p.U_v, D, lane, cur_user, gu);
        if (lane == 0) update_scalar(p.bias, p.bu, p.bu_m, p.bu_v, cur_user, gbu_acc);
      } else {
        gu.red_add(p.GU + (int64_t)cur_user * D, D, lane);
        if (lane == 0) {
          atomicAdd(p.gbu + cur_user, gbu_acc);
          mark_row(p.user_flag, p.user_list, &p.hdr->user_count, cur_user, p.use_lists);
        }
      }
    };

    for (int le = 0; le < n_e; ++le) {
      if (le + kStages - 1 < n_e) issue(le + kStages - 1);
      const unsigned char* st = rows + (size_t)(consumed % kStages) * c.stage_bytes;
      mbar_wait(&bars[consumed % kStages], (consumed / kStages) & 1u);
      ++consumed;

      const int* idr = ids + le * NP;
      const float* br = bias + le * NP;
      const int u = idr[0];
      if (u != cur_user) {
        finish_user(le);
        cur_user = u;
        left_whole = (e0 + le == 0) || (p.su[e0 + le - 1] != u);
        gu.zero();
        gbu_acc = 0.f;
      }
      const int64_t b = bpos[le];
      const int pos = idr[1];
      const unsigned char* urow = st;

      // ---- all N + 1 dot products at once: lane (jrow, seg) owns float4 seg, seg + St, ... of row
      float my_score = -INFINITY;  // score of row j0 + jrow (valid when j < Rw), same in all St lanes
      float sp = 0.f;
      int sel = (p.loss == CMF_LOSS_WARP) ? N : 0;
      float sel_s = 0.f, sel_h = 0.f;
      uint32_t best_key = 0;
      for (int j0 = 0; j0 < Rw; j0 += c.Jt) {
        const int j = j0 + jrow;
        float acc = 0.f;
        if (j < Rw) {
          const unsigned char* rj = st + (size_t)(1 + j) * c.stride;
          for (int i = seg; i < c.D4; i += St) {
            const float4 a = lds128(urow + i * 16);
            const float4 x = lds128(rj + i * 16);
            acc = fmaf(a.x, x.x, acc);
            acc = fmaf(a.y, x.y, acc);
            acc = fmaf(a.z, x.z, acc);
            acc = fmaf(a.w, x.w, acc);
          }
        }
        for (int o = St >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        my_score = (j < Rw) ? acc + br[0] + br[1 + j] : -INFINITY;
        if (j0 == 0) {
          sp = __shfl_sync(0xffffffffu, my_score, 0);
          if (p.pos_out && lane == 0) p.pos_out[b] = sp;
        }
        if (!implicit) break;
        const bool is_neg = (j >= 1) && (j < Rw);
        if (p.neg_out && is_neg && seg == 0) p.neg_out[b * N + (j - 1)] = my_score;
        if (p.loss == CMF_LOSS_WARP) {
          // h = ideal - pos + neg; the first strict violation wins (warp.py:95-105, 136-154)
          float h = 0.f;
          if (is_neg) h = (ideal_of(p, pos, idr[1 + j]) - sp) + my_score;
          const unsigned viol = __ballot_sync(0xffffffffu, is_neg && h > 0.f);
          if (sel == N && viol) {
            const int src = __ffs(viol) - 1;
            sel = j0 + src / St - 1;
            sel_h = __shfl_sync(0xffffffffu, h, src);
          }
        } else if (adaptive) {
          // torch.max over the N axis: the first index of the maximum (bpr.py:153, hinge.py:157)
          const uint32_t key = is_neg ? float_key(my_score) : 0u;
          const uint32_t mx = __reduce_max_sync(0xffffffffu, key);
          if (mx > best_key) {
            const unsigned who = __ballot_sync(0xffffffffu, is_neg && key == mx);
            const int src = __ffs(who) - 1;
            sel = j0 + src / St - 1;
            sel_s = __shfl_sync(0xffffffffu, my_score, src);
            best_key = mx;
          }
        } else if (j0 == 0) {
          sel = 0;
          sel_s = __shfl_sync(0xffffffffu, my_score, St);
        }
      }

      // ---- loss value and d loss / d score
      float g = 0.f;       // implicit: d/d pos (d/d selected neg is -g); explicit: d/d score
      int sel_id = -1;
      if (implicit) {
        if (p.sel_out && lane == 0) p.sel_out[b] = sel;
        if (p.loss == CMF_LOSS_WARP) {
          if (sel < N) {
            sel_id = idr[2 + sel];
            const float w = logf((float)p.num_items / (float)(sel + 1));
            const float Lw = w * sel_h;
            loss_acc += Lw + Lw * Lw;
            g = -((1.f + 2.f * Lw) * w) * p.inv_B;
          }
        } else {
          sel_id = idr[2 + sel];
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
        const float diff = pred - aux[le];
        float dp;
        if (p.loss == CMF_LOSS_MSE) { loss_acc += diff * diff; dp = 2.f * diff * p.inv_B; }
        else { loss_acc += fabsf(diff); dp = (diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f)) * p.inv_B; }
        g = p.has_range ? dp * (sg * (1.f - sg) * (p.y_hi - p.y_lo)) : dp;
      }

      // ---- backward: lane owns float4 lane, lane + 32, ... of the gradient rows
      if (p.train && g != 0.f) {
        const unsigned char* prow = st + c.stride;
        const unsigned char* srow = implicit ? st + (size_t)(2 + sel) * c.stride : nullptr;
        float* gi_pos = p.GI + (int64_t)pos * D;
        float* gi_sel = implicit ? p.GI + (int64_t)sel_id * D : nullptr;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int i = r * 32 + lane;
          if (i < c.D4) {
            const float4 xp = lds128(prow + i * 16);
            const float4 xu = lds128(urow + i * 16);
            float* f = gu.f + r * 4;
            // products rounded before the add, so +g*x and -g*x cancel exactly (see Row::axpy)
            f[0] = __fadd_rn(f[0], __fmul_rn(g, xp.x));
            f[1] = __fadd_rn(f[1], __fmul_rn(g, xp.y));
            f[2] = __fadd_rn(f[2], __fmul_rn(g, xp.z));
            f[3] = __fadd_rn(f[3], __fmul_rn(g, xp.w));
            float t[4] = {__fmul_rn(g, xu.x), __fmul_rn(g, xu.y), __fmul_rn(g, xu.z),
                          __fmul_rn(g, xu.w)};
            red_vec<4>(gi_pos + i * 4, t);
            if (implicit) {
              const float4 xs = lds128(srow + i * 16);
              f[0] = __fadd_rn(f[0], __fmul_rn(-g, xs.x));
              f[1] = __fadd_rn(f[1], __fmul_rn(-g, xs.y));
              f[2] = __fadd_rn(f[2], __fmul_rn(-g, xs.z));
              f[3] = __fadd_rn(f[3], __fmul_rn(-g, xs.w));
              t[0] = -t[0]; t[1] = -t[1]; t[2] = -t[2]; t[3] = -t[3];
              red_vec<4>(gi_sel + i * 4, t);
            }
          }
        }
        if (implicit) {
          gbu_acc += g;
          gbu_acc += -g;
          if (lane == 0) {
            atomicAdd(p.gbi + pos, g);
            atomicAdd(p.gbi + sel_id, -g);
          }
        } else {
          gbu_acc += g;
          if (lane == 0) atomicAdd(p.gbi + pos, g);
        }
      }
      // the stage may be refilled (by the async proxy) only after every lane has read it
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }
    finish_user(n_e);
    __syncwarp();
  }

  // ---- loss: deterministic two-level reduction (every lane carries the same loss_acc)
  if (lane == 0) s_warp_loss[warp] = loss_acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int w = 0; w < c.W; ++w) tot += (double)s_warp_loss[w];
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
    p.loss_out[0] = (float)(tot / (double)p.B);
  }
}

bool launch_fused_step_v2(const StepDev& p_in, int max_chunk_len, cudaStream_t s, int* rc) {
  *rc = CMF_OK;
  const int D = p_in.D, N = p_in.N;
  if (D % 4 != 0 || D > 1024) return false;
  V2Cfg c{};
  c.NP = N + 2;
  const int Rw = N + 1;
  c.Jt = Rw <= 8 ? 8 : (Rw <= 16 ? 16 : 32);
  c.St = 32 / c.Jt;
  c.D4 = D / 4;
  c.row_bytes = D * 4;
  c.stride = (int)align_up((size_t)c.row_bytes, 128) + 16 * c.St;
  c.off_ids = 64;
  c.off_bias = c.off_ids + (int)align_up(sizeof(int) * kMaxChunk * c.NP, 16);
  c.off_aux = c.off_bias + (int)align_up(sizeof(float) * kMaxChunk * c.NP, 16);
  c.off_rows = (int)align_up((size_t)c.off_aux + 2 * sizeof(int) * kMaxChunk, 128);
  c.stage_bytes = c.NP * c.stride;
  c.warp_bytes = (int)align_up((size_t)c.off_rows + (size_t)kStages * c.stage_bytes, 128);
  const int budget = 200 * 1024;
  c.W = budget / c.warp_bytes;
  if (c.W > 16) c.W = 16;
  if (c.W < 2) return false;  // rows too large for the staged pipeline: generic kernel
  const int sms = num_sms();
  // chunk length: long chunks amortise phase A and leave fewer straddling users, but every warp
  // should still get several chunks
  int L = kMaxChunk;
  const int64_t warps = (int64_t)sms * c.W;
  while (L > 1 && p_in.B / L < 4 * warps) L >>= 1;
  if (max_chunk_len > 0) L = max_chunk_len < kMaxChunk ? max_chunk_len : kMaxChunk;
  c.L = L;
  StepDev p = p_in;
  p.L = L;
  p.num_chunks = (p.B + L - 1) / L;
  int64_t blocks = (p.num_chunks + c.W - 1) / c.W;
  if (blocks > sms) blocks = sms;
  const size_t smem = (size_t)c.W * c.warp_bytes;
  const int R = (c.D4 + 31) / 32;
  cudaError_t e = cudaSuccess;
#define V2_LAUNCH(R_)                                                                            \
  do {                                                                                           \
    e = cudaFuncSetAttribute(k_fused_step_v2<R_>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                             (int)smem);                                                         \
    if (e == cudaSuccess)                                                                        \
      k_fused_step_v2<R_><<<(unsigned)blocks, 32 * c.W, smem, s>>>(p, c);                        \
  } while (0)
  if (R <= 1) V2_LAUNCH(1);
  else if (R == 2) V2_LAUNCH(2);
  else if (R <= 4) V2_LAUNCH(4);
  else V2_LAUNCH(8);
#undef V2_LAUNCH
  if (e == cudaSuccess) e = cudaGetLastError();
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (e != cudaSuccess)
    *rc = fail(CMF_ERR_CUDA, "k_fused_step_v2 launch failed: %s", cudaGetErrorString(e));
  return true;
}

}  // namespace cmf
