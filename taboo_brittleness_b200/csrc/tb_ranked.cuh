// K1b for nested ablation budgets (Execution Plan:126, budgets as prefixes of ONE ranking per prompt / per seed):
//
//   tb_rank_lists_kernel     once per ranking: every row's active latents that appear in the ranking
//                            (rank < 1024), re-ordered by rank -> (key = rank << 20 | latent, value) lists
//   tb_ablate_ranked_kernel  per budget b: the latents of a row to ablate are exactly the leading entries of its
//                            ranked list with rank < b -- at most b of them. No rank-table lookups per budget, no
//                            pass over the whole active list (about 150 entries per row, of which a handful are
//                            ablated at small budgets), no compaction: read the prefix, gather those W_dec rows.
//
// Same arithmetic as tb_ablate_decode_kernel in mode 0 (x' = x - sum a_i W_dec[i], bf16 rounding, fused final
// RMSNorm), with the sum taken in rank order instead of latent order.
#pragma once
#include "tb_rows.cuh"

namespace tb {

constexpr int kRankedMax = 1024;           // ranks 0 .. 1023 (TB_TOPM_MAX)
constexpr int kRankListThreads = 256;

// One CTA per row (loop over rows). Ranks are unique within a ranking, so the row's selected entries are placed
// by rank into 1024 slots of shared memory (a counting sort without collisions) and then compacted in slot order.
__global__ void __launch_bounds__(kRankListThreads)
tb_rank_lists_kernel(const int64_t* __restrict__ row_start, int cap, const int32_t* __restrict__ count,
                     const int32_t* __restrict__ idx, const float* __restrict__ val,
                     const uint16_t* __restrict__ rank, int rank_rows, int n_rows, int S,
                     int32_t* __restrict__ out_key, float* __restrict__ out_val, int32_t* __restrict__ out_count) {
  __shared__ int32_t s_lat[kRankedMax];
  __shared__ float s_val[kRankedMax];
  __shared__ int s_warp[kRankListThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int row = blockIdx.x; row < n_rows; row += gridDim.x) {
    for (int i = tid; i < kRankedMax; i += kRankListThreads) s_lat[i] = -1;
    __syncthreads();
    const int64_t start = row_start ? row_start[row] : static_cast<int64_t>(row) * cap;
    int cnt = count[row];
    if (row_start == nullptr) cnt = min(cnt, cap);
    const uint16_t* rk = rank + (rank_rows > 0 ? static_cast<size_t>(row / rank_rows) * S : 0);
    for (int i = tid; i < cnt; i += kRankListThreads) {
      const int32_t lat = idx[start + i];
      const int r = static_cast<int>(rk[lat]);
      if (r < kRankedMax) {
        s_lat[r] = lat;
        s_val[r] = val[start + i];
      }
    }
    __syncthreads();
    // compaction in rank order: each thread owns 4 consecutive slots
    int mine = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) mine += (s_lat[tid * 4 + q] >= 0) ? 1 : 0;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kRankListThreads / 32; ++w) {
      if (w < warp) base += s_warp[w];
      total += s_warp[w];
    }
    int pos = base + incl - mine;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int slot = tid * 4 + q;
      const int32_t lat = s_lat[slot];
      if (lat >= 0) {
        out_key[start + pos] = (slot << 20) | lat;
        out_val[start + pos] = s_val[slot];
        ++pos;
      }
    }
    if (tid == 0) out_count[row] = total;
    __syncthreads();   // s_lat / s_warp are reused by the next row
  }
}

struct AblateRankedArgs {
  const __nv_bfloat16* x;      // [n_rows, d]
  const int64_t* row_start;    // [n_rows] or nullptr -> row * cap (the layout of the active lists)
  int cap;
  const int32_t* sel_count;    // [n_rows] entries of the ranked list
  const int32_t* sel_key;      // rank << 20 | latent, ascending
  const float* sel_val;
  const __nv_bfloat16* w_dec;  // [S, d]
  int budget;
  int heavy_min;               // rows with at least this many entries to ablate are left to tb_ablate_heavy_kernel
  const float* w_norm;         // [d] or nullptr
  float eps;
  __nv_bfloat16* out_x;        // nullable
  __nv_bfloat16* out_h;        // nullable
  float* out_rstd;             // nullable, [n_rows]: the RMSNorm's 1 / sqrt(mean(x'^2) + eps) of the rounded x' (rank-update lens)
  int n_rows, d;
};

// Row-group layout of tb_ablate_decode_kernel: four rows per CTA, 128 threads per row on their own named barrier,
// a CTA owns a contiguous run of rows.
template <int NV>
__global__ void __launch_bounds__(kAblThreads, (NV <= 4 ? 2 : 1)) tb_ablate_ranked_kernel(const AblateRankedArgs a) {
  __shared__ float s_red[kRowGroups][4];
  __shared__ int32_t s_id[kRowGroups][kRowGroupThreads];
  __shared__ float s_a[kRowGroups][kRowGroupThreads];
  __shared__ int s_wcnt[kRowGroups][4];
  const int group = threadIdx.x / kRowGroupThreads;
  const int gt = threadIdx.x - group * kRowGroupThreads;
  const int gwarp = gt >> 5, lane = gt & 31;
  const int nvec = a.d >> 3;
  const int rows_per_cta = (a.n_rows + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int cta_row0 = static_cast<int>(blockIdx.x) * rows_per_cta;
  const int cta_row1 = min(a.n_rows, cta_row0 + rows_per_cta);
  for (int row = cta_row0 + group; row < cta_row1; row += kRowGroups) {
    // the list prefix first (its latency heads the dependent chain), then x
    const int64_t start = a.row_start ? a.row_start[row] : static_cast<int64_t>(row) * a.cap;
    const int n_list = min(a.sel_count[row], a.budget);     // ranks are unique: at most `budget` entries qualify
    if (n_list >= a.heavy_min) continue;                    // (uniform over the row group) a whole CTA takes this row
    float acc[NV][8];
    const int4* xr = reinterpret_cast<const int4*>(a.x + static_cast<size_t>(row) * a.d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      if (vec < nvec) unpack8(__ldg(xr + vec), acc[i]);
    }
    for (int b0 = 0; b0 < n_list; b0 += kRowGroupThreads) {
      const int i = b0 + gt;
      int32_t key = 0x7fffffff;
      float av = 0.0f;
      if (i < n_list) {
        key = a.sel_key[start + i];
        av = a.sel_val[start + i];
      }
      const bool sel = (key >> 20) < a.budget;               // sorted by rank: the selected entries are a prefix
      const uint32_t bal = __ballot_sync(0xffffffffu, sel);
      group_bar_sync(group);                                 // the previous batch's gathers are done with s_id / s_a
      if (lane == 0) s_wcnt[group][gwarp] = __popc(bal);
      if (sel) {
        s_id[group][gt] = key & 0xfffff;
        s_a[group][gt] = -av;
      }
      group_bar_sync(group);
      const int n = (s_wcnt[group][0] + s_wcnt[group][1]) + (s_wcnt[group][2] + s_wcnt[group][3]);
#pragma unroll 2
      for (int j = 0; j < n; ++j) {
        const int4* wr = reinterpret_cast<const int4*>(a.w_dec + static_cast<size_t>(s_id[group][j]) * a.d);
        const float aj = s_a[group][j];
#pragma unroll
        for (int q = 0; q < NV; ++q) {
          const int vec = gt + q * kRowGroupThreads;
          if (vec < nvec) {
            float wv[8];
            unpack8(__ldg(wr + vec), wv);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[q][e] = fmaf(aj, wv[e], acc[q][e]);
          }
        }
      }
      if (n < kRowGroupThreads) break;                       // the prefix ended inside this batch
    }
    float ss = 0.0f;
#pragma unroll
    for (int q = 0; q < NV; ++q) {
      const int vec = gt + q * kRowGroupThreads;
      if (vec < nvec) {
        const Bf16x8 p = pack8(acc[q]);
        if (a.out_x) reinterpret_cast<Bf16x8*>(a.out_x + static_cast<size_t>(row) * a.d)[vec] = p;
        unpack8(p, acc[q]);
#pragma unroll
        for (int e = 0; e < 8; ++e) ss = fmaf(acc[q][e], acc[q][e], ss);
      }
    }
    if (a.out_h || a.out_rstd) {
      ss = group_sum(ss, s_red[group], group, gwarp, lane);
      const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, static_cast<float>(a.d)), a.eps)));
      if (a.out_rstd && gt == 0) a.out_rstd[row] = rstd;
#pragma unroll
      for (int q = 0; q < NV; ++q) {
        const int vec = gt + q * kRowGroupThreads;
        if (a.out_h && vec < nvec) {
          const float4 w0 = __ldg(reinterpret_cast<const float4*>(a.w_norm) + 2 * vec);
          const float4 w1 = __ldg(reinterpret_cast<const float4*>(a.w_norm) + 2 * vec + 1);
          const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
          float h[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) h[e] = __fmul_rn(__fmul_rn(acc[q][e], rstd), __fadd_rn(1.0f, ww[e]));
          reinterpret_cast<Bf16x8*>(a.out_h + static_cast<size_t>(row) * a.d)[vec] = pack8(h);
        }
      }
    }
  }
}

// The rows with long prefixes -- in practice the <bos>-scale row of every prompt, in which nearly all of the
// prompt's ranked latents are active, so a budget of 1024 means 1024 W_dec rows (7 MB from L2) for that one row
// while an ordinary row gathers about ten. In the row-group kernel such a row keeps one group of 128 threads
// busy with a serial chain of gathers for hundreds of microseconds. Here a whole CTA takes it: 512 threads, one
// 16-byte column slice each (two above d = 4096), eight W_dec rows in flight per thread. The sum runs over the
// entries in rank order with the same fmaf chain per element as the row-group kernel: same bits.
constexpr int kHeavyThreads = 512;
constexpr int kHeavyBatch = 512;
constexpr int kHeavyUnroll = 8;

template <int NV>
__global__ void __launch_bounds__(kHeavyThreads) tb_ablate_heavy_kernel(const AblateRankedArgs a) {
  __shared__ float s_red[kRowMaxWarps];
  __shared__ int32_t s_id[kHeavyBatch];
  __shared__ float s_a[kHeavyBatch];
  __shared__ int s_rows[kHeavyThreads];
  __shared__ int s_nheavy;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nvec = a.d >> 3;
  // a CTA scans a contiguous run of rows, kHeavyThreads at a time, and collects the heavy ones
  const int rows_per_cta = (a.n_rows + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int row0 = static_cast<int>(blockIdx.x) * rows_per_cta;
  const int row1 = min(a.n_rows, row0 + rows_per_cta);
  for (int base = row0; base < row1; base += kHeavyThreads) {
    __syncthreads();
    if (tid == 0) s_nheavy = 0;
    __syncthreads();
    const int r = base + tid;
    if (r < row1 && min(a.sel_count[r], a.budget) >= a.heavy_min) s_rows[atomicAdd(&s_nheavy, 1)] = r;
    __syncthreads();
    const int nh = s_nheavy;
    for (int hi = 0; hi < nh; ++hi) {
      // (the order in which heavy rows are taken does not matter: every row is computed independently)
      const int row = s_rows[hi];
      const int64_t start = a.row_start ? a.row_start[row] : static_cast<int64_t>(row) * a.cap;
      const int n_list = min(a.sel_count[row], a.budget);
      float acc[NV][8];
      const int4* xr = reinterpret_cast<const int4*>(a.x + static_cast<size_t>(row) * a.d);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = tid + i * kHeavyThreads;
        if (vec < nvec) unpack8(__ldg(xr + vec), acc[i]);
      }
      for (int b0 = 0; b0 < n_list; b0 += kHeavyBatch) {
        __syncthreads();                                     // the previous batch's gathers are done with s_id / s_a
        const int i = b0 + tid;
        int32_t key = 0x7fffffff;
        float av = 0.0f;
        if (i < n_list) {
          key = a.sel_key[start + i];
          av = a.sel_val[start + i];
        }
        const bool sel = (key >> 20) < a.budget;
        s_id[tid] = key & 0xfffff;
        s_a[tid] = -av;
        const int n = __syncthreads_count(sel);              // sorted by rank: the selected entries are a prefix
#pragma unroll
        for (int q = 0; q < NV; ++q) {
          const int vec = tid + q * kHeavyThreads;
          if (vec < nvec) {
            int j = 0;
            for (; j + kHeavyUnroll <= n; j += kHeavyUnroll) {
              int4 w[kHeavyUnroll];
#pragma unroll
              for (int u = 0; u < kHeavyUnroll; ++u)
                w[u] = __ldg(reinterpret_cast<const int4*>(a.w_dec + static_cast<size_t>(s_id[j + u]) * a.d) + vec);
#pragma unroll
              for (int u = 0; u < kHeavyUnroll; ++u) {
                float wv[8];
                unpack8(w[u], wv);
                const float aj = s_a[j + u];
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[q][e] = fmaf(aj, wv[e], acc[q][e]);
              }
            }
            for (; j < n; ++j) {
              float wv[8];
              unpack8(__ldg(reinterpret_cast<const int4*>(a.w_dec + static_cast<size_t>(s_id[j]) * a.d) + vec), wv);
              const float aj = s_a[j];
#pragma unroll
              for (int e = 0; e < 8; ++e) acc[q][e] = fmaf(aj, wv[e], acc[q][e]);
            }
          }
        }
        if (n < kHeavyBatch) break;
      }
      float xb[NV][8];
#pragma unroll
      for (int q = 0; q < NV; ++q) {
        const int vec = tid + q * kHeavyThreads;
        if (vec < nvec) {
          const Bf16x8 p = pack8(acc[q]);
          if (a.out_x) reinterpret_cast<Bf16x8*>(a.out_x + static_cast<size_t>(row) * a.d)[vec] = p;
          unpack8(p, xb[q]);
        }
      }
      if (a.out_h || a.out_rstd)
        rmsnorm_store<NV>(xb, a.w_norm, a.d, a.eps, a.out_h ? a.out_h + static_cast<size_t>(row) * a.d : nullptr, s_red,
                          a.out_rstd ? a.out_rstd + row : nullptr);
    }
  }
  (void)lane; (void)warp;
}

// -----------------------------------------------------------------------------------------------------------------
// All nested budgets of a chain in ONE pass over the rows. The budgets are prefixes of one rank-ordered list, so
// x'_{b(k)} = x'_{b(k-1)} - (the entries with rank in [b(k-1), b(k))): the row is read once, every W_dec row it needs is
// gathered once (not once per budget that contains it), and after each budget's entries the row is rounded, normalised
// and written to that budget's output. Per row: d * 2 bytes in, n_budgets * d * 2 out, against n_budgets * d * 4 for one
// launch per budget. The fmaf chain per element is the per-budget kernel's: the same bits.
// Outputs of budget k live at out + k * out_stride elements (out_stride >= n_rows * d), rstd at out_rstd + k * n_rows.
// -----------------------------------------------------------------------------------------------------------------
constexpr int kMaxBudgets = 16;

struct AblateRankedMultiArgs {
  AblateRankedArgs base;       // base.budget = the largest budget; base.out_x / out_h / out_rstd = budget 0's outputs
  int n_budgets;
  int budgets[kMaxBudgets];    // ascending
  int64_t out_stride;          // elements between consecutive budgets' out_x / out_h
};

// round, (store x'), sum of squares, rstd, (store h) for the outputs k0 .. k1 - 1, which all see the same x'
template <int NV>
__device__ __forceinline__ void emit_row_group(const AblateRankedMultiArgs& m, const float (&acc)[NV][8], int row, int k0, int k1,
                                               float* red4, int group, int gt, int gwarp, int lane) {
  const AblateRankedArgs& a = m.base;
  const int nvec = a.d >> 3;
  float ss = 0.0f;
#pragma unroll
  for (int q = 0; q < NV; ++q) {
    const int vec = gt + q * kRowGroupThreads;
    if (vec < nvec) {
      const Bf16x8 p = pack8(acc[q]);
      if (a.out_x)
        for (int k = k0; k < k1; ++k) reinterpret_cast<Bf16x8*>(a.out_x + k * m.out_stride + static_cast<size_t>(row) * a.d)[vec] = p;
      float t[8];
      unpack8(p, t);
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(t[e], t[e], ss);
    }
  }
  if (!a.out_h && !a.out_rstd) return;
  ss = group_sum(ss, red4, group, gwarp, lane);
  const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, static_cast<float>(a.d)), a.eps)));
  if (a.out_rstd && gt == 0)
    for (int k = k0; k < k1; ++k) a.out_rstd[static_cast<size_t>(k) * a.n_rows + row] = rstd;
  if (!a.out_h) return;
#pragma unroll
  for (int q = 0; q < NV; ++q) {
    const int vec = gt + q * kRowGroupThreads;
    if (vec < nvec) {
      float t[8];
      unpack8(pack8(acc[q]), t);
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(a.w_norm) + 2 * vec);
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(a.w_norm) + 2 * vec + 1);
      const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
      float h[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) h[e] = __fmul_rn(__fmul_rn(t[e], rstd), __fadd_rn(1.0f, ww[e]));
      const Bf16x8 ph = pack8(h);
      for (int k = k0; k < k1; ++k) reinterpret_cast<Bf16x8*>(a.out_h + k * m.out_stride + static_cast<size_t>(row) * a.d)[vec] = ph;
    }
  }
}

template <int NV>
__global__ void __launch_bounds__(kAblThreads, (NV <= 4 ? 2 : 1)) tb_ablate_ranked_multi_kernel(const AblateRankedMultiArgs m) {
  const AblateRankedArgs& a = m.base;
  __shared__ float s_red[kRowGroups][4];
  __shared__ int32_t s_id[kRowGroups][kRowGroupThreads];
  __shared__ float s_a[kRowGroups][kRowGroupThreads];
  __shared__ int s_wcnt[kRowGroups][4];
  const int group = threadIdx.x / kRowGroupThreads;
  const int gt = threadIdx.x - group * kRowGroupThreads;
  const int gwarp = gt >> 5, lane = gt & 31;
  const int nvec = a.d >> 3;
  const int rows_per_cta = (a.n_rows + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int cta_row0 = static_cast<int>(blockIdx.x) * rows_per_cta;
  const int cta_row1 = min(a.n_rows, cta_row0 + rows_per_cta);
  for (int row = cta_row0 + group; row < cta_row1; row += kRowGroups) {
    const int64_t start = a.row_start ? a.row_start[row] : static_cast<int64_t>(row) * a.cap;
    const int n_list = min(a.sel_count[row], a.budget);
    if (n_list >= a.heavy_min) continue;                    // (uniform over the row group) a whole CTA takes this row
    float acc[NV][8];
    const int4* xr = reinterpret_cast<const int4*>(a.x + static_cast<size_t>(row) * a.d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      if (vec < nvec) unpack8(__ldg(xr + vec), acc[i]);
    }
    int pos = 0;                                            // entries applied so far
    int k = 0;
    while (k < m.n_budgets) {
      // a run of budgets no entry separates sees the same x': find the run [k, k1) that ends at the next entry's rank
      int k1 = k + 1;
      if (pos >= n_list) k1 = m.n_budgets;                  // the list is used up: all remaining budgets are equal
      const int bk = m.budgets[k1 - 1];
      // apply the entries with rank < bk from pos on
      while (pos < n_list) {
        const int i = pos + gt;
        int32_t key = 0x7fffffff;
        float av = 0.0f;
        if (i < n_list) {
          key = a.sel_key[start + i];
          av = a.sel_val[start + i];
        }
        const bool sel = (key >> 20) < bk;                   // sorted by rank: the selected entries are a prefix
        const uint32_t bal = __ballot_sync(0xffffffffu, sel);
        group_bar_sync(group);                               // the previous batch's gathers are done with s_id / s_a
        if (lane == 0) s_wcnt[group][gwarp] = __popc(bal);
        if (sel) {
          s_id[group][gt] = key & 0xfffff;
          s_a[group][gt] = -av;
        }
        group_bar_sync(group);
        const int n = (s_wcnt[group][0] + s_wcnt[group][1]) + (s_wcnt[group][2] + s_wcnt[group][3]);
#pragma unroll 2
        for (int j = 0; j < n; ++j) {
          const int4* wr = reinterpret_cast<const int4*>(a.w_dec + static_cast<size_t>(s_id[group][j]) * a.d);
          const float aj = s_a[group][j];
#pragma unroll
          for (int q = 0; q < NV; ++q) {
            const int vec = gt + q * kRowGroupThreads;
            if (vec < nvec) {
              float wv[8];
              unpack8(__ldg(wr + vec), wv);
#pragma unroll
              for (int e = 0; e < 8; ++e) acc[q][e] = fmaf(aj, wv[e], acc[q][e]);
            }
          }
        }
        pos += n;
        if (n < kRowGroupThreads) break;                     // the budget's entries ended inside this batch
      }
      emit_row_group<NV>(m, acc, row, k, k1, s_red[group], group, gt, gwarp, lane);
      k = k1;
    }
  }
}

template <int NV>
__global__ void __launch_bounds__(kHeavyThreads) tb_ablate_heavy_multi_kernel(const AblateRankedMultiArgs m) {
  const AblateRankedArgs& a = m.base;
  __shared__ float s_red[kRowMaxWarps];
  __shared__ int32_t s_id[kHeavyBatch];
  __shared__ float s_a[kHeavyBatch];
  __shared__ int s_rows[kHeavyThreads];
  __shared__ int s_nheavy;
  const int tid = threadIdx.x;
  const int nvec = a.d >> 3;
  const int rows_per_cta = (a.n_rows + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int row0 = static_cast<int>(blockIdx.x) * rows_per_cta;
  const int row1 = min(a.n_rows, row0 + rows_per_cta);
  for (int base = row0; base < row1; base += kHeavyThreads) {
    __syncthreads();
    if (tid == 0) s_nheavy = 0;
    __syncthreads();
    const int r = base + tid;
    if (r < row1 && min(a.sel_count[r], a.budget) >= a.heavy_min) s_rows[atomicAdd(&s_nheavy, 1)] = r;
    __syncthreads();
    const int nh = s_nheavy;
    for (int hi = 0; hi < nh; ++hi) {
      const int row = s_rows[hi];
      const int64_t start = a.row_start ? a.row_start[row] : static_cast<int64_t>(row) * a.cap;
      const int n_list = min(a.sel_count[row], a.budget);
      float acc[NV][8];
      const int4* xr = reinterpret_cast<const int4*>(a.x + static_cast<size_t>(row) * a.d);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = tid + i * kHeavyThreads;
        if (vec < nvec) unpack8(__ldg(xr + vec), acc[i]);
      }
      int pos = 0;
      for (int k = 0; k < m.n_budgets; ++k) {
        const int bk = m.budgets[k];
        while (pos < n_list) {
          __syncthreads();                                   // the previous batch's gathers are done with s_id / s_a
          const int i = pos + tid;
          int32_t key = 0x7fffffff;
          float av = 0.0f;
          if (i < n_list) {
            key = a.sel_key[start + i];
            av = a.sel_val[start + i];
          }
          const bool sel = (key >> 20) < bk;
          s_id[tid] = key & 0xfffff;
          s_a[tid] = -av;
          const int n = __syncthreads_count(sel);            // sorted by rank: the selected entries are a prefix
#pragma unroll
          for (int q = 0; q < NV; ++q) {
            const int vec = tid + q * kHeavyThreads;
            if (vec < nvec) {
              int j = 0;
              for (; j + kHeavyUnroll <= n; j += kHeavyUnroll) {
                int4 w[kHeavyUnroll];
#pragma unroll
                for (int u = 0; u < kHeavyUnroll; ++u)
                  w[u] = __ldg(reinterpret_cast<const int4*>(a.w_dec + static_cast<size_t>(s_id[j + u]) * a.d) + vec);
#pragma unroll
                for (int u = 0; u < kHeavyUnroll; ++u) {
                  float wv[8];
                  unpack8(w[u], wv);
                  const float aj = s_a[j + u];
#pragma unroll
                  for (int e = 0; e < 8; ++e) acc[q][e] = fmaf(aj, wv[e], acc[q][e]);
                }
              }
              for (; j < n; ++j) {
                float wv[8];
                unpack8(__ldg(reinterpret_cast<const int4*>(a.w_dec + static_cast<size_t>(s_id[j]) * a.d) + vec), wv);
                const float aj = s_a[j];
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[q][e] = fmaf(aj, wv[e], acc[q][e]);
              }
            }
          }
          pos += n;
          if (n < kHeavyBatch) break;
        }
        float xb[NV][8];
#pragma unroll
        for (int q = 0; q < NV; ++q) {
          const int vec = tid + q * kHeavyThreads;
          if (vec < nvec) {
            const Bf16x8 p = pack8(acc[q]);
            if (a.out_x) reinterpret_cast<Bf16x8*>(a.out_x + k * m.out_stride + static_cast<size_t>(row) * a.d)[vec] = p;
            unpack8(p, xb[q]);
          }
        }
        if (a.out_h || a.out_rstd)
          rmsnorm_store<NV>(xb, a.w_norm, a.d, a.eps,
                            a.out_h ? a.out_h + k * m.out_stride + static_cast<size_t>(row) * a.d : nullptr, s_red,
                            a.out_rstd ? a.out_rstd + static_cast<size_t>(k) * a.n_rows + row : nullptr);
      }
    }
  }
}

}  // namespace tb
