// Row-wise, HBM/L2-bound kernels of the intervention-and-readout path:
//   tb_rmsnorm_kernel          a6  final Gemma2RMSNorm -> bf16 GEMM operand h
//   tb_ablate_decode_kernel    a4  SAE latent ablation + sparse decode (+ fused a6)
//   tb_latent_mean_kernel      a2  mean over a prompt's tokens of the sparse activations
//   tb_topm_select_kernel      a2  top-m latents of the mean (sorted) + rank table
//   tb_lens_merge_kernel       a8-a10 merge of (max, sumexp, secret, top-10) partials
// A residual row (d bf16 values, d % 8 == 0) is held by one CTA as 16-byte vectors, NV per
// thread, so every global access is a coalesced 128-bit load/store.
#pragma once
#include <cuda_bf16.h>
#include <math_constants.h>
#include <stdint.h>

namespace tb {

constexpr int kRowMaxWarps = 32;

struct alignas(16) Bf16x8 {
  __nv_bfloat162 v[4];
};

__device__ __forceinline__ void unpack8(const Bf16x8& p, float (&f)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(p.v[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
// unpack8 overload for a raw 16-byte load
__device__ __forceinline__ void unpack8(const int4& raw, float (&f)[8]) {
  Bf16x8 p;
  *reinterpret_cast<int4*>(&p) = raw;
  unpack8(p, f);
}

__device__ __forceinline__ Bf16x8 pack8(const float (&f)[8]) {
  Bf16x8 p;
#pragma unroll
  for (int i = 0; i < 4; ++i) p.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return p;
}

// sum over the CTA; every thread gets the result. red must hold kRowMaxWarps floats.
__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = (blockDim.x + 31) >> 5;
  __syncthreads();  // red may still be read from a previous call
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = 0.0f;
  for (int w = 0; w < nwarps; ++w) t += red[w];  // fixed order: deterministic
  return t;
}

// Final RMSNorm exactly as Gemma2RMSNorm.forward computes it on a bf16 residual:
//   h = bf16( (x * rsqrt(mean(x^2) + eps)) * (1 + w) ), fp32 intermediates, no FMA contraction.
// xb holds the row already rounded to bf16 (what the reference's bf16 residual stream holds).
template <int NV>
__device__ __forceinline__ void rmsnorm_store(const float (&xb)[NV][8], const float* __restrict__ w,
                                              int d, float eps, __nv_bfloat16* __restrict__ out_h,
                                              float* red, float* out_rstd = nullptr) {
  const int nvec = d >> 3;
  float ss = 0.0f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vec = threadIdx.x + i * blockDim.x;
    if (vec < nvec) {
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(xb[i][e], xb[i][e], ss);
    }
  }
  ss = block_sum(ss, red);
  const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, static_cast<float>(d)), eps)));
  if (out_rstd && threadIdx.x == 0) *out_rstd = rstd;
  if (out_h == nullptr) return;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vec = threadIdx.x + i * blockDim.x;
    if (vec < nvec) {
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(w) + 2 * vec);
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(w) + 2 * vec + 1);
      const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
      float h[8];
#pragma unroll
      for (int e = 0; e < 8; ++e)
        h[e] = __fmul_rn(__fmul_rn(xb[i][e], rstd), __fadd_rn(1.0f, ww[e]));
      reinterpret_cast<Bf16x8*>(out_h)[vec] = pack8(h);
    }
  }
}

// -----------------------------------------------------------------------------------------
// a4: ablate + decode. For every row, walk that row's active-latent list; a latent is
// ablated when its rank in the row's ablation table is below the budget.
//   mode 0 (error-term form, north_star): x' = x - sum_{ablated active i} a_i W_dec[i, :]
//        == decode(acts with idx zeroed) + (x - decode(acts))
//   mode 1 (plain decode, Plan:126):      x' = b_dec + sum_{surviving active i} a_i W_dec[i, :]
// x' is rounded to bf16 (the residual stream dtype); optionally the final RMSNorm of that
// rounded row is produced in the same pass (out_h), so the lens GEMM reads its operand
// without a separate normalisation pass.
// -----------------------------------------------------------------------------------------

struct AblateArgs {
  const __nv_bfloat16* x;      // [n_rows, d]
  const int64_t* row_start;    // [n_rows] or nullptr -> row * cap
  int cap;
  const int32_t* count;        // [n_rows]
  const int32_t* idx;
  const float* val;
  const __nv_bfloat16* w_dec;  // [S, d]
  const float* b_dec;          // [d] (mode 1)
  const uint16_t* rank;        // [n_tables, S]
  const int32_t* row_table;    // [n_rows] table index per row, or nullptr
  int rank_rows;               // rows per table when row_table == nullptr (0: one table)
  int budget;
  int mode;
  int S;
  const float* w_norm;         // [d] or nullptr
  float eps;
  __nv_bfloat16* out_x;        // [n_rows, d] or nullptr
  __nv_bfloat16* out_h;        // [n_rows, d] or nullptr
  int n_rows, d;
};

// Four rows per CTA, 128 threads (4 warps) per row, each row group on its own named barrier:
// up to 16 rows in flight per SM keeps enough 16-byte loads outstanding to stream x / h at HBM
// speed, and a row's latency chain (count -> list -> rank lookups -> W_dec rows) overlaps with
// the other groups' instead of idling a whole CTA.
constexpr int kRowGroupThreads = 128;
constexpr int kRowGroups = 4;
constexpr int kListPerThread = 1;                                   // list entries a thread examines per batch
constexpr int kListBatch = kRowGroupThreads * kListPerThread;
constexpr int kAblThreads = kRowGroupThreads * kRowGroups;

__device__ __forceinline__ void group_bar_sync(int group) {
  asm volatile("bar.sync %0, 128;" ::"r"(group + 1) : "memory");
}
// sum over the 128 threads of a row group (4 warps); fixed order, every thread gets the result
__device__ __forceinline__ float group_sum(float v, float* red4, int group, int gwarp, int lane) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  group_bar_sync(group);   // red4 may still be read from a previous use
  if (lane == 0) red4[gwarp] = v;
  group_bar_sync(group);
  return (red4[0] + red4[1]) + (red4[2] + red4[3]);
}

// a6 alone: four rows per CTA, 128 threads per row (the layout of tb_ablate_decode_kernel), so that
// 8-16 rows per SM are in flight.
template <int NV>
__global__ void __launch_bounds__(kAblThreads) tb_rmsnorm_kernel(const __nv_bfloat16* __restrict__ x,
                                                                 const float* __restrict__ w, int n_rows, int d,
                                                                 float eps, __nv_bfloat16* __restrict__ out_h) {
  __shared__ float s_red[kRowGroups][4];
  const int group = threadIdx.x / kRowGroupThreads;
  const int gt = threadIdx.x - group * kRowGroupThreads;
  const int gwarp = gt >> 5, lane = gt & 31;
  const int nvec = d >> 3;
  // the next row's x is requested (raw, 16 registers) before this row's reduction: twice the loads in flight
  const int row_step = gridDim.x * kRowGroups;
  int4 nxt[NV];
  {
    const int row = blockIdx.x * kRowGroups + group;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      nxt[i] = make_int4(0, 0, 0, 0);
      if (row < n_rows && vec < nvec) nxt[i] = __ldg(reinterpret_cast<const int4*>(x + static_cast<size_t>(row) * d) + vec);
    }
  }
  for (int row = blockIdx.x * kRowGroups + group; row < n_rows; row += row_step) {
    float xb[NV][8];
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      unpack8(nxt[i], xb[i]);
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(xb[i][e], xb[i][e], ss);   // (vectors past the row hold zeros)
    }
    if (row + row_step < n_rows) {
      const int4* xn = reinterpret_cast<const int4*>(x + static_cast<size_t>(row + row_step) * d);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = gt + i * kRowGroupThreads;
        if (vec < nvec) nxt[i] = __ldg(xn + vec);
      }
    }
    ss = group_sum(ss, s_red[group], group, gwarp, lane);
    const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, static_cast<float>(d)), eps)));
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      if (vec < nvec) {
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(w) + 2 * vec);
        const float4 w1 = __ldg(reinterpret_cast<const float4*>(w) + 2 * vec + 1);
        const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
        float h[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) h[e] = __fmul_rn(__fmul_rn(xb[i][e], rstd), __fadd_rn(1.0f, ww[e]));
        reinterpret_cast<Bf16x8*>(out_h + static_cast<size_t>(row) * d)[vec] = pack8(h);
      }
    }
  }
}

// K2 on the tensor cores, step 3: h = bf16((x' * rstd) * (1 + w)) with rstd from the per-(part, row) sums of
// squares that step 2's epilogue left behind (tb_epilogues.cuh: ProjApplyEpi). Same row-group layout and
// arithmetic as tb_rmsnorm_kernel, minus the reduction. In place (xin == out_h) is fine.
template <int NV>
__global__ void __launch_bounds__(kAblThreads) tb_scale_rows_kernel(const __nv_bfloat16* xin,   // (may alias out_h)
                                                                    const float* __restrict__ ss, int n_parts,
                                                                    const float* __restrict__ w, int n_rows, int d,
                                                                    float eps, __nv_bfloat16* out_h) {
  const int group = threadIdx.x / kRowGroupThreads;
  const int gt = threadIdx.x - group * kRowGroupThreads;
  const int nvec = d >> 3;
  const int row_step = gridDim.x * kRowGroups;
  int4 nxt[NV];   // the next row, requested one row ahead
  {
    const int row = blockIdx.x * kRowGroups + group;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      nxt[i] = make_int4(0, 0, 0, 0);
      if (row < n_rows && vec < nvec) nxt[i] = reinterpret_cast<const int4*>(xin + static_cast<size_t>(row) * d)[vec];
    }
  }
  for (int row = blockIdx.x * kRowGroups + group; row < n_rows; row += row_step) {
    int4 raw[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) raw[i] = nxt[i];
    if (row + row_step < n_rows) {
      // (in place: another group's row, not yet rewritten -- rows are only ever touched by their own group)
      const int4* xn = reinterpret_cast<const int4*>(xin + static_cast<size_t>(row + row_step) * d);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = gt + i * kRowGroupThreads;
        if (vec < nvec) nxt[i] = xn[vec];
      }
    }
    float sum = 0.0f;
    for (int p = 0; p < n_parts; ++p) sum += __ldg(ss + static_cast<size_t>(p) * n_rows + row);   // fixed order
    const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(sum, static_cast<float>(d)), eps)));
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      if (vec < nvec) {
        float xb[8], h[8];
        unpack8(raw[i], xb);
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(w) + 2 * vec);
        const float4 w1 = __ldg(reinterpret_cast<const float4*>(w) + 2 * vec + 1);
        const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) h[e] = __fmul_rn(__fmul_rn(xb[e], rstd), __fadd_rn(1.0f, ww[e]));
        reinterpret_cast<Bf16x8*>(out_h + static_cast<size_t>(row) * d)[vec] = pack8(h);
      }
    }
  }
}

// K2 on the tensor cores: the B operand of step 2, ub [d, 2 * rp] = [U | U] with U = ut^T (ut is [r, d]),
// zero in the padding columns r <= k < rp.
__global__ void tb_lowrank_ub_kernel(const __nv_bfloat16* __restrict__ ut, int r, int rp, int d,
                                     __nv_bfloat16* __restrict__ ub) {
  const int kc = 2 * rp;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < static_cast<int64_t>(d) * kc;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int j = static_cast<int>(i / kc), k = static_cast<int>(i % kc) % rp;
    ub[i] = (k < r) ? ut[static_cast<size_t>(k) * d + j] : __float2bfloat16_rn(0.0f);
  }
}

template <int NV>
__global__ void __launch_bounds__(kAblThreads, (NV <= 4 ? 2 : 1)) tb_ablate_decode_kernel(const AblateArgs a) {
  __shared__ float s_red[kRowGroups][4];
  __shared__ int32_t s_id[kRowGroups][kListBatch];
  __shared__ float s_a[kRowGroups][kListBatch];
  __shared__ int s_wcnt[kRowGroups][4 * kListPerThread];
  const int group = threadIdx.x / kRowGroupThreads;
  const int gt = threadIdx.x - group * kRowGroupThreads;
  const int gwarp = gt >> 5, lane = gt & 31;
  const int nvec = a.d >> 3;
  const float sign = (a.mode == 0) ? -1.0f : 1.0f;
  // A CTA owns a contiguous run of rows (its groups interleave inside it). Rows cost very different
  // amounts -- a <bos>-scale row has ~100x the active latents -- and they recur with the prompt length;
  // a grid-stride assignment puts all of them on the few CTAs whose index resonates with that period.
  const int rows_per_cta = (a.n_rows + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int cta_row0 = static_cast<int>(blockIdx.x) * rows_per_cta;
  const int cta_row1 = min(a.n_rows, cta_row0 + rows_per_cta);
  for (int row = cta_row0 + group; row < cta_row1; row += kRowGroups) {
    float acc[NV][8];
    if (a.mode == 0) {
      const int4* xr = reinterpret_cast<const int4*>(a.x + static_cast<size_t>(row) * a.d);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = gt + i * kRowGroupThreads;
        if (vec < nvec) unpack8(__ldg(xr + vec), acc[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = gt + i * kRowGroupThreads;
        if (vec < nvec) {
          const float4 b0 = __ldg(reinterpret_cast<const float4*>(a.b_dec) + 2 * vec);
          const float4 b1 = __ldg(reinterpret_cast<const float4*>(a.b_dec) + 2 * vec + 1);
          acc[i][0] = b0.x; acc[i][1] = b0.y; acc[i][2] = b0.z; acc[i][3] = b0.w;
          acc[i][4] = b1.x; acc[i][5] = b1.y; acc[i][6] = b1.z; acc[i][7] = b1.w;
        }
      }
    }
    const int64_t start = a.row_start ? a.row_start[row] : static_cast<int64_t>(row) * a.cap;
    int cnt = a.count[row];
    if (a.row_start == nullptr) cnt = min(cnt, a.cap);  // an overflowed row only holds cap entries
    const int table = a.row_table ? a.row_table[row] : (a.rank_rows > 0 ? row / a.rank_rows : 0);
    const uint16_t* rk = a.rank ? a.rank + static_cast<size_t>(table) * a.S : nullptr;

    for (int b0 = 0; b0 < cnt; b0 += kListBatch) {
      // stage 1: every thread looks at kListPerThread list entries (entry = b0 + q * 128 + gt, all
      // loads in flight together); ordered compaction (list order is kept, so the fp32 accumulation
      // order is deterministic): ballots per (q, warp) + prefix over the 16 of them
      bool sel[kListPerThread];
      int32_t id[kListPerThread];
      float av[kListPerThread];
#pragma unroll
      for (int q = 0; q < kListPerThread; ++q) {
        const int i = b0 + q * kRowGroupThreads + gt;
        sel[q] = false;
        id[q] = 0;
        av[q] = 0.0f;
        if (i < cnt) {
          id[q] = a.idx[start + i];
          av[q] = a.val[start + i];
        }
      }
#pragma unroll
      for (int q = 0; q < kListPerThread; ++q) {
        const int i = b0 + q * kRowGroupThreads + gt;
        if (i < cnt) {
          const bool abl = rk ? (static_cast<int>(rk[id[q]]) < a.budget) : false;
          sel[q] = (a.mode == 0) ? abl : !abl;
        }
      }
      uint32_t bal[kListPerThread];
#pragma unroll
      for (int q = 0; q < kListPerThread; ++q) bal[q] = __ballot_sync(0xffffffffu, sel[q]);
      group_bar_sync(group);          // previous batch's stage 2 is done with s_id / s_a / s_wcnt
      if (lane == 0) {
#pragma unroll
        for (int q = 0; q < kListPerThread; ++q) s_wcnt[group][q * 4 + gwarp] = __popc(bal[q]);
      }
      group_bar_sync(group);
      int n = 0;
      int before[kListPerThread];
#pragma unroll
      for (int q = 0; q < kListPerThread; ++q) {
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          if (w == gwarp) before[q] = n;
          n += s_wcnt[group][q * 4 + w];
        }
      }
#pragma unroll
      for (int q = 0; q < kListPerThread; ++q) {
        if (sel[q]) {
          const int pos = before[q] + __popc(bal[q] & ((1u << lane) - 1u));
          s_id[group][pos] = id[q];
          s_a[group][pos] = sign * av[q];
        }
      }
      group_bar_sync(group);
      // stage 2: acc += a_i * W_dec[i, :]; independent 16-byte gathers, L2-resident rows
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
    }
    // round to the residual dtype; everything downstream sees the bf16 row
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
    if (a.out_h) {   // fused final RMSNorm (same arithmetic as rmsnorm_store)
      ss = group_sum(ss, s_red[group], group, gwarp, lane);
      const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, static_cast<float>(a.d)), a.eps)));
#pragma unroll
      for (int q = 0; q < NV; ++q) {
        const int vec = gt + q * kRowGroupThreads;
        if (vec < nvec) {
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

// -----------------------------------------------------------------------------------------
// a2 (first half): mean over a prompt's rows of the sparse activations -> dense [P, S] fp32.
// One CTA per prompt; rows are added in token order (a latent occurs at most once per row,
// so there are no intra-row conflicts and the fp32 sum order is fixed).
// Reference: torch.mean(sae.encode(residual[start:]), dim=0), src/02_run_sae_baseline.py:67-70.
// -----------------------------------------------------------------------------------------
__global__ void tb_latent_mean_kernel(const int64_t* __restrict__ row_start, int cap,
                                      const int32_t* __restrict__ count,
                                      const int32_t* __restrict__ idx, const float* __restrict__ val,
                                      const int64_t* __restrict__ seg, int rows_per_prompt, int S,
                                      float* __restrict__ out_mean) {
  const int p = blockIdx.x;
  const int64_t r0 = seg ? seg[p] : static_cast<int64_t>(p) * rows_per_prompt;
  const int64_t r1 = seg ? seg[p + 1] : r0 + rows_per_prompt;
  float* out = out_mean + static_cast<size_t>(p) * S;
  for (int j = threadIdx.x; j < S; j += blockDim.x) out[j] = 0.0f;
  __syncthreads();
  for (int64_t r = r0; r < r1; ++r) {
    const int64_t start = row_start ? row_start[r] : r * cap;
    int cnt = count[r];
    if (row_start == nullptr) cnt = min(cnt, cap);
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) out[idx[start + i]] += val[start + i];
    __syncthreads();
  }
  const float n = static_cast<float>(r1 - r0);
  if (r1 > r0)
    for (int j = threadIdx.x; j < S; j += blockDim.x) out[j] = __fdiv_rn(out[j], n);
}

// -----------------------------------------------------------------------------------------
// Weighted per-prompt latent sums for the Execution Plan's "secret score" (Plan:116-124):
//   out[p, s] = sum over the prompt's rows t of w[t] * a[t, s]^power,  power in {1, 2}
// (w = 1 / spike indicator / secret probability give the sums the score's mean-at-spikes and
// Pearson correlation are made of). Same structure and determinism as tb_latent_mean_kernel.
// -----------------------------------------------------------------------------------------
__global__ void tb_latent_wsum_kernel(const int64_t* __restrict__ row_start, int cap,
                                      const int32_t* __restrict__ count, const int32_t* __restrict__ idx,
                                      const float* __restrict__ val, const int64_t* __restrict__ seg,
                                      int rows_per_prompt, int S, const float* __restrict__ row_weight,
                                      int power, float* __restrict__ out_sum) {
  const int p = blockIdx.x;
  const int64_t r0 = seg ? seg[p] : static_cast<int64_t>(p) * rows_per_prompt;
  const int64_t r1 = seg ? seg[p + 1] : r0 + rows_per_prompt;
  float* out = out_sum + static_cast<size_t>(p) * S;
  for (int j = threadIdx.x; j < S; j += blockDim.x) out[j] = 0.0f;
  __syncthreads();
  for (int64_t r = r0; r < r1; ++r) {
    const float w = row_weight ? row_weight[r] : 1.0f;
    if (w != 0.0f) {   // uniform over the CTA
      const int64_t start = row_start ? row_start[r] : r * cap;
      int cnt = count[r];
      if (row_start == nullptr) cnt = min(cnt, cap);
      for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
        const float a = val[start + i];
        out[idx[start + i]] += w * (power == 2 ? a * a : a);
      }
    }
    __syncthreads();
  }
}

// -----------------------------------------------------------------------------------------
// a2 (second half): per prompt, the m largest means, sorted descending (ties: lower latent id
// first), plus a rank table rank[latent] = position (0xFFFF if not selected). Nested budgets
// m' <= m are prefixes of the sorted list, i.e. "rank < m'".
// Radix select (11 + 11 + 10 bits) for the threshold key, ordered compaction, bitonic sort.
// Reference: torch.topk(avg_sae_acts, k).indices, src/02_run_sae_baseline.py:73.
// -----------------------------------------------------------------------------------------
constexpr int kTopmThreads = 1024;
constexpr int kTopmMax = 1024;

__device__ __forceinline__ uint32_t float_key(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// find the bin (scanning from the top) in which the cumulative count reaches `need`
__device__ __forceinline__ void pick_bin(const uint32_t* hist, int nbins, uint32_t need,
                                         uint32_t* s_bin, uint32_t* s_above) {
  if (threadIdx.x == 0) {
    uint32_t cum = 0;
    int b = nbins - 1;
    for (; b > 0; --b) {
      if (cum + hist[b] >= need) break;
      cum += hist[b];
    }
    *s_bin = static_cast<uint32_t>(b);
    *s_above = cum;
  }
}

__global__ void __launch_bounds__(kTopmThreads, 1)
tb_topm_select_kernel(const float* __restrict__ mean, int S, int m, int32_t* __restrict__ out_idx,
                      float* __restrict__ out_val, uint16_t* __restrict__ rank) {
  __shared__ uint32_t hist[2048];
  __shared__ unsigned long long sel[kTopmMax];
  __shared__ uint32_t s_bin, s_above, s_warp_gt[32], s_warp_eq[32];
  __shared__ uint32_t s_run_sel, s_run_eq;
  const int p = blockIdx.x;
  const float* v = mean + static_cast<size_t>(p) * S;
  const int tid = threadIdx.x;
  const uint32_t mm = static_cast<uint32_t>(min(m, S));

  // ---- pass 1: top 11 bits
  for (int i = tid; i < 2048; i += kTopmThreads) hist[i] = 0;
  __syncthreads();
  for (int base = 0; base < S; base += 4 * kTopmThreads) {
    // four loads in flight per thread (one block reads a whole row of up to V values: latency-bound otherwise); lanes of a
    // warp that hit the same bin add once (column sums of probabilities crowd into a handful of bins: one shared-memory
    // atomic per element serialised the whole pass at V = 256000)
    uint32_t bin[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = base + u * kTopmThreads + tid;
      bin[u] = j < S ? (float_key(v[j]) >> 21) : 0xffffffffu;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t peers = __match_any_sync(0xffffffffu, bin[u]);
      if (bin[u] != 0xffffffffu && (tid & 31) == __ffs(peers) - 1) atomicAdd(&hist[bin[u]], static_cast<uint32_t>(__popc(peers)));
    }
  }
  __syncthreads();
  pick_bin(hist, 2048, mm, &s_bin, &s_above);
  __syncthreads();
  const uint32_t b1 = s_bin;
  uint32_t need = mm - s_above;
  __syncthreads();
  // ---- pass 2: middle 11 bits inside bin b1
  for (int i = tid; i < 2048; i += kTopmThreads) hist[i] = 0;
  __syncthreads();
  for (int base = 0; base < S; base += 4 * kTopmThreads) {
    uint32_t k[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = base + u * kTopmThreads + tid;
      k[u] = j < S ? float_key(v[j]) : 0u;       // (key 0 = -NaN pattern, never a value's key inside a chosen bin with need > 0)
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (base + u * kTopmThreads + tid < S && (k[u] >> 21) == b1) atomicAdd(&hist[(k[u] >> 10) & 0x7ffu], 1u);
  }
  __syncthreads();
  pick_bin(hist, 2048, need, &s_bin, &s_above);
  __syncthreads();
  const uint32_t b2 = s_bin;
  need -= s_above;
  __syncthreads();
  // ---- pass 3: low 10 bits inside (b1, b2)
  for (int i = tid; i < 1024; i += kTopmThreads) hist[i] = 0;
  __syncthreads();
  const uint32_t hi22 = (b1 << 11) | b2;
  for (int base = 0; base < S; base += 4 * kTopmThreads) {
    uint32_t k[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = base + u * kTopmThreads + tid;
      k[u] = j < S ? float_key(v[j]) : 0u;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (base + u * kTopmThreads + tid < S && (k[u] >> 10) == hi22) atomicAdd(&hist[k[u] & 0x3ffu], 1u);
  }
  __syncthreads();
  pick_bin(hist, 1024, need, &s_bin, &s_above);
  __syncthreads();
  const uint32_t kth = (hi22 << 10) | s_bin;  // key of the m-th largest value
  const uint32_t need_eq = need - s_above;    // how many elements equal to it are taken

  // ---- ordered compaction: key > kth, plus the first need_eq (by latent id) with key == kth
  if (tid == 0) { s_run_sel = 0; s_run_eq = 0; }
  for (int i = tid; i < kTopmMax; i += kTopmThreads) sel[i] = 0ull;
  __syncthreads();
  const int warp = tid >> 5, lane = tid & 31;
  uint32_t kpre[4] = {0u, 0u, 0u, 0u};
  bool live = true;                               // does the current group of four rounds hold anything at or above the threshold?
  for (int base = 0; base < S; base += kTopmThreads) {
    const int u4 = (base / kTopmThreads) & 3;
    if (u4 == 0) {                                // four rounds' values at once, one barrier to rule all four out
      bool any = false;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int jj = base + u * kTopmThreads + tid;
        kpre[u] = jj < S ? float_key(v[jj]) : 0u;
        any = any || (jj < S && kpre[u] >= kth);
      }
      live = __syncthreads_or(any) != 0;
    }
    if (!live) continue;
    const int j = base + tid;
    uint32_t k = 0;
    bool gt = false, eq = false;
    if (j < S) {
      k = u4 == 0 ? kpre[0] : u4 == 1 ? kpre[1] : u4 == 2 ? kpre[2] : kpre[3];
      gt = k > kth;
      eq = k == kth;
    }
    if (!__syncthreads_or(gt || eq)) continue;      // nothing at or above the threshold in these 1024 values (the usual case for large S)
    const uint32_t bal_eq = __ballot_sync(0xffffffffu, eq);
    const uint32_t bal_gt = __ballot_sync(0xffffffffu, gt);
    if (lane == 0) { s_warp_eq[warp] = __popc(bal_eq); s_warp_gt[warp] = __popc(bal_gt); }
    __syncthreads();
    uint32_t eq_before = s_run_eq, gt_before = 0, eq_total = 0, gt_total = 0;
    for (int w = 0; w < kTopmThreads / 32; ++w) {
      if (w < warp) { eq_before += s_warp_eq[w]; gt_before += s_warp_gt[w]; }
      eq_total += s_warp_eq[w];
      gt_total += s_warp_gt[w];
    }
    const uint32_t lane_mask = (1u << lane) - 1u;
    const uint32_t my_eq_rank = eq_before + __popc(bal_eq & lane_mask);  // global tie order
    const bool take = gt || (eq && my_eq_rank < need_eq);
    // position: selected-so-far + gt before me + eq taken before me (this round)
    const uint32_t eq_taken_before_round = min(s_run_eq, need_eq);
    const uint32_t eq_taken_before_me = min(my_eq_rank, need_eq) - eq_taken_before_round;
    if (take) {
      const uint32_t pos = s_run_sel + gt_before + __popc(bal_gt & lane_mask) + eq_taken_before_me;
      if (pos < kTopmMax)
        sel[pos] = (static_cast<unsigned long long>(k) << 32) |
                   static_cast<unsigned long long>(0xffffffffu - static_cast<uint32_t>(j));
    }
    __syncthreads();
    if (tid == 0) {
      const uint32_t eq_taken_round = min(s_run_eq + eq_total, need_eq) - eq_taken_before_round;
      s_run_sel += gt_total + eq_taken_round;
      s_run_eq += eq_total;
    }
    __syncthreads();
  }

  // ---- bitonic sort, descending on (key, ~id): larger value first, lower id first on ties
  for (int size = 2; size <= kTopmMax; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const int partner = tid ^ stride;
      if (partner > tid) {
        const bool desc = ((tid & size) == 0);
        const unsigned long long a = sel[tid], b = sel[partner];
        if ((a < b) == desc) { sel[tid] = b; sel[partner] = a; }
      }
      __syncthreads();
    }
  }

  // ---- outputs
  uint16_t* rk = rank ? rank + static_cast<size_t>(p) * S : nullptr;     // (nullable: only the sorted list is wanted)
  if (rk)
    for (int j = tid; j < S; j += kTopmThreads) rk[j] = 0xffffu;
  __syncthreads();
  if (tid < static_cast<int>(mm)) {
    const unsigned long long e = sel[tid];
    const int32_t id = static_cast<int32_t>(0xffffffffu - static_cast<uint32_t>(e & 0xffffffffull));
    out_idx[static_cast<size_t>(p) * m + tid] = id;
    out_val[static_cast<size_t>(p) * m + tid] = key_float(static_cast<uint32_t>(e >> 32));
    if (rk) rk[id] = static_cast<uint16_t>(tid);
  } else if (tid < m) {
    out_idx[static_cast<size_t>(p) * m + tid] = -1;
    out_val[static_cast<size_t>(p) * m + tid] = 0.0f;
  }
}

// -----------------------------------------------------------------------------------------
// K4: merge per-(row, part) lens partials: parts are vocab chunks of one GPU and/or the
// vocab shards of several GPUs after an all_gather. finalize == 0 writes one merged partial
// (same layout, n_parts = 1); finalize == 1 writes lse, secret probability, top-10 probs/ids.
// -----------------------------------------------------------------------------------------
struct MergeArgs {
  const float* pm; const float* ps; const float* psec; const float* ptv; const int32_t* pti;
  int n_parts, n_rows;
  int finalize;
  float* om; float* os; float* osec; float* otv; int32_t* oti;   // finalize == 0
  float* lse; float* secret_prob; float* top_prob; int32_t* top_id;  // finalize == 1
};

__global__ void tb_lens_merge_kernel(const MergeArgs a) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= a.n_rows) return;
  float m = -CUDART_INF_F, s = 0.0f, sec = -CUDART_INF_F;
  float tv[10];
  int32_t ti[10];
#pragma unroll
  for (int k = 0; k < 10; ++k) { tv[k] = -CUDART_INF_F; ti[k] = -1; }
  for (int part = 0; part < a.n_parts; ++part) {
    const size_t o = static_cast<size_t>(part) * a.n_rows + row;
    const float pm = a.pm[o], ps = a.ps[o];
    if (pm > m) { s *= expf(m - pm); m = pm; }
    if (pm > -CUDART_INF_F) s += ps * expf(pm - m);
    sec = fmaxf(sec, a.psec[o]);
    for (int c = 0; c < 10; ++c) {
      const float v = a.ptv[o * 10 + c];
      if (!(v > tv[9])) break;  // a part's list is sorted: nothing further can enter
      tv[9] = v;
      ti[9] = a.pti[o * 10 + c];
#pragma unroll
      for (int k = 9; k > 0; --k) {
        const bool sw = tv[k] > tv[k - 1];
        const float fa = tv[k - 1], fb = tv[k];
        const int32_t ia = ti[k - 1], ib = ti[k];
        tv[k - 1] = sw ? fb : fa; tv[k] = sw ? fa : fb;
        ti[k - 1] = sw ? ib : ia; ti[k] = sw ? ia : ib;
      }
    }
  }
  if (!a.finalize) {
    a.om[row] = m; a.os[row] = s; a.osec[row] = sec;
#pragma unroll
    for (int k = 0; k < 10; ++k) { a.otv[row * 10 + k] = tv[k]; a.oti[row * 10 + k] = ti[k]; }
  } else {
    const float lse = m + logf(s);
    a.lse[row] = lse;
    if (a.secret_prob) a.secret_prob[row] = (sec > -CUDART_INF_F) ? expf(sec - lse) : 0.0f;
#pragma unroll
    for (int k = 0; k < 10; ++k) {
      a.top_prob[row * 10 + k] = (tv[k] > -CUDART_INF_F) ? expf(tv[k] - lse) : 0.0f;
      a.top_id[row * 10 + k] = ti[k];
    }
  }
}

// -----------------------------------------------------------------------------------------
// K5 merge: per prompt, the k best of n_parts sorted (sum, id) lists of 10 (vocab chunks, CTA
// halves, and -- after an all_gather -- vocab shards). Output sorted descending; ids -1 / sums
// -inf where fewer than k entries exist.
// -----------------------------------------------------------------------------------------
__global__ void tb_ll_topk_merge_kernel(const float* __restrict__ pv, const int32_t* __restrict__ pi,
                                        int n_parts, int n_prompts, int k, float* __restrict__ ov,
                                        int32_t* __restrict__ oi) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_prompts) return;
  float tv[10];
  int32_t ti[10];
#pragma unroll
  for (int j = 0; j < 10; ++j) { tv[j] = -CUDART_INF_F; ti[j] = -1; }
  for (int part = 0; part < n_parts; ++part) {
    const size_t o = (static_cast<size_t>(part) * n_prompts + p) * 10;
    for (int c = 0; c < 10; ++c) {
      const float v = pv[o + c];
      const int32_t id = pi[o + c];
      // a part's list is sorted: once an entry cannot enter, none after it can
      if (id < 0 || !(v > tv[9] || (v == tv[9] && (ti[9] < 0 || id < ti[9])))) break;
      tv[9] = v;
      ti[9] = id;
#pragma unroll
      for (int j = 9; j > 0; --j) {
        const bool sw = tv[j] > tv[j - 1] || (tv[j] == tv[j - 1] && ti[j] >= 0 && (ti[j - 1] < 0 || ti[j] < ti[j - 1]));
        const float fa = tv[j - 1], fb = tv[j];
        const int32_t ia = ti[j - 1], ib = ti[j];
        tv[j - 1] = sw ? fb : fa; tv[j] = sw ? fa : fb;
        ti[j - 1] = sw ? ib : ia; ti[j] = sw ? ia : ib;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 10; ++j)
    if (j < k) { ov[static_cast<size_t>(p) * k + j] = tv[j]; oi[static_cast<size_t>(p) * k + j] = ti[j]; }
}

// -----------------------------------------------------------------------------------------
// a11 on CACHED dense probabilities (the cached-analysis branch of the reference's driver,
// src/01_reproduce_logit_lens.py:35-71 called from :120-150): acc[v] = sum over response
// positions i of probs[i, v], with probs[i, cur_id[i]] and probs[i, prev_id[i]] taken as 0.
// One thread per vocabulary column, positions added in order with plain fp32 adds -- the same
// order and rounding as the reference's `acc += p` loop, so acc is bit-identical. HBM-bound:
// T * V * 4 bytes read once, fully coalesced.
// -----------------------------------------------------------------------------------------
constexpr int kColsumMaxT = 1024;

__global__ void tb_probs_colsum_kernel(const float* __restrict__ probs, int T, int V,
                                       const int32_t* __restrict__ cur_id,
                                       const int32_t* __restrict__ prev_id, float* __restrict__ acc) {
  __shared__ int32_t s_cur[kColsumMaxT], s_prev[kColsumMaxT];
  for (int t = threadIdx.x; t < T; t += blockDim.x) {
    s_cur[t] = cur_id ? cur_id[t] : -1;
    s_prev[t] = prev_id ? prev_id[t] : -1;
  }
  __syncthreads();
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= V) return;
  float a = 0.0f;
  for (int t = 0; t < T; ++t) {
    float p = __ldg(probs + static_cast<size_t>(t) * V + col);
    if (col == s_cur[t] || col == s_prev[t]) p = 0.0f;
    a = __fadd_rn(a, p);
  }
  acc[col] = a;
}

// -----------------------------------------------------------------------------------------
// K2 / a5: low-rank direction removal  x' = x - (x U) U^T,  U [d, r] orthonormal, r <= 64
// (Execution Plan:148, r_edited = r - P_U(r)). The caller passes U^T [r, d] (bf16) so that
// both thin products read 16-byte contiguous slices: phase 1 forms c = x U for a tile of
// kProjRows rows staged in shared memory (one direction per warp, fp32 accumulate, shuffle
// reduce); phase 2 subtracts sum_j c_j U[:, j] with each thread owning fixed 8-element slices
// of the rows. c stays in fp32. Output is rounded to bf16 and optionally RMS-normed (a6).
// -----------------------------------------------------------------------------------------
constexpr int kProjRows = 8;
constexpr int kProjGroup = 4;  // rows accumulated at once in phase 2
constexpr int kProjMaxRank = 64;

struct ProjArgs {
  const __nv_bfloat16* x;   // [n_rows, d]
  const __nv_bfloat16* ut;  // [r, d]
  int r;
  const float* w_norm;      // [d] or nullptr
  float eps;
  __nv_bfloat16* out_x;     // nullable
  __nv_bfloat16* out_h;     // nullable
  int n_rows, d;
};

template <int NV>
__global__ void __launch_bounds__(512) tb_lowrank_remove_kernel(const ProjArgs a) {
  extern __shared__ __align__(16) uint8_t proj_smem[];
  __shared__ float red[kRowMaxWarps];
  __shared__ float s_c[kProjRows][kProjMaxRank];
  Bf16x8* s_x = reinterpret_cast<Bf16x8*>(proj_smem);  // [kProjRows][nvec]
  const int nvec = a.d >> 3;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int n_tiles = (a.n_rows + kProjRows - 1) / kProjRows;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int row0 = tile * kProjRows;
    __syncthreads();
    // stage the row tile (zero rows past the end)
    for (int rr = 0; rr < kProjRows; ++rr) {
      const int row = row0 + rr;
      for (int vec = threadIdx.x; vec < nvec; vec += blockDim.x) {
        Bf16x8 z;
        if (row < a.n_rows) z = reinterpret_cast<const Bf16x8*>(a.x + static_cast<size_t>(row) * a.d)[vec];
        else *reinterpret_cast<int4*>(&z) = make_int4(0, 0, 0, 0);
        s_x[rr * nvec + vec] = z;
      }
    }
    __syncthreads();
    // phase 1: c[rr][j] = <x_rr, U[:, j]>
    for (int j = warp; j < a.r; j += nwarps) {
      float c[kProjRows];
#pragma unroll
      for (int rr = 0; rr < kProjRows; ++rr) c[rr] = 0.0f;
      const int4* ur = reinterpret_cast<const int4*>(a.ut + static_cast<size_t>(j) * a.d);
      for (int vec = lane; vec < nvec; vec += 32) {
        float u[8];
        unpack8(__ldg(ur + vec), u);
#pragma unroll
        for (int rr = 0; rr < kProjRows; ++rr) {
          float xv[8];
          unpack8(s_x[rr * nvec + vec], xv);
#pragma unroll
          for (int e = 0; e < 8; ++e) c[rr] = fmaf(xv[e], u[e], c[rr]);
        }
      }
#pragma unroll
      for (int rr = 0; rr < kProjRows; ++rr) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c[rr] += __shfl_xor_sync(0xffffffffu, c[rr], o);
        if (lane == 0) s_c[rr][j] = c[rr];
      }
    }
    __syncthreads();
    // phase 2: x' = x - sum_j c_j U[:, j], kProjGroup rows at a time (register budget), then
    // round, store, and the optional fused final RMSNorm
#pragma unroll 1
    for (int g0 = 0; g0 < kProjRows; g0 += kProjGroup) {
      float acc[kProjGroup][NV][8];
#pragma unroll
      for (int rr = 0; rr < kProjGroup; ++rr)
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int vec = threadIdx.x + i * blockDim.x;
          if (vec < nvec) unpack8(s_x[(g0 + rr) * nvec + vec], acc[rr][i]);
        }
      for (int j = 0; j < a.r; ++j) {
        const int4* ur = reinterpret_cast<const int4*>(a.ut + static_cast<size_t>(j) * a.d);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int vec = threadIdx.x + i * blockDim.x;
          if (vec < nvec) {
            float u[8];
            unpack8(__ldg(ur + vec), u);
#pragma unroll
            for (int rr = 0; rr < kProjGroup; ++rr) {
              const float cj = -s_c[g0 + rr][j];
#pragma unroll
              for (int e = 0; e < 8; ++e) acc[rr][i][e] = fmaf(cj, u[e], acc[rr][i][e]);
            }
          }
        }
      }
#pragma unroll
      for (int rr = 0; rr < kProjGroup; ++rr) {
        const int row = row0 + g0 + rr;
        const bool valid = row < a.n_rows;  // uniform across the CTA
        float xb[NV][8];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int vec = threadIdx.x + i * blockDim.x;
          if (vec < nvec) {
            const Bf16x8 p = pack8(acc[rr][i]);
            if (valid && a.out_x) reinterpret_cast<Bf16x8*>(a.out_x + static_cast<size_t>(row) * a.d)[vec] = p;
            unpack8(p, xb[i]);
          }
        }
        if (valid && a.out_h)
          rmsnorm_store<NV>(xb, a.w_norm, a.d, a.eps, a.out_h + static_cast<size_t>(row) * a.d, red);
      }
    }
  }
}


// -----------------------------------------------------------------------------------------
// K2 for the ranks the Execution Plan actually sweeps (r <= 8, Plan:144-150): U^T (r x d bf16,
// <= 57 KB) is loaded into shared memory once per CTA and stays there; a CTA streams tiles of
// four rows, each thread holding its 16-byte column slices of all four rows in registers, so
// every U element read from shared memory is used for four rows in both thin products
// (c = x U, x' = x - c U^T). Per row the kernel moves the algorithmic 2 * d * 2 bytes of HBM
// traffic and nothing else. Larger ranks use tb_lowrank_remove_kernel (U^T slices from L2).
// -----------------------------------------------------------------------------------------
constexpr int kProjSmallRank = 8;
constexpr int kProjSmallRows = 4;
constexpr int kProjSmallThreads = 256;

template <int NV>
__global__ void __launch_bounds__(kProjSmallThreads, 2) tb_lowrank_small_kernel(const ProjArgs a) {
  extern __shared__ __align__(16) uint8_t proj_small_smem[];
  int4* s_ut = reinterpret_cast<int4*>(proj_small_smem);            // [r][nvec] 16-byte vectors of U^T
  __shared__ float s_part[kProjSmallThreads / 32][kProjSmallRows * kProjSmallRank];
  __shared__ float s_c[kProjSmallRows * kProjSmallRank];
  __shared__ float s_ss[kProjSmallThreads / 32][kProjSmallRows];
  const int nvec = a.d >> 3;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int kWarps = kProjSmallThreads / 32;
  for (int i = tid; i < a.r * nvec; i += kProjSmallThreads)
    s_ut[i] = __ldg(reinterpret_cast<const int4*>(a.ut) + i);
  __syncthreads();
  const int n_tiles = (a.n_rows + kProjSmallRows - 1) / kProjSmallRows;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int row0 = tile * kProjSmallRows;
    float x[kProjSmallRows][NV][8];
#pragma unroll
    for (int rr = 0; rr < kProjSmallRows; ++rr) {
      const int row = row0 + rr;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = tid + i * kProjSmallThreads;
        int4 raw = make_int4(0, 0, 0, 0);
        if (row < a.n_rows && vec < nvec) raw = __ldg(reinterpret_cast<const int4*>(a.x + static_cast<size_t>(row) * a.d) + vec);
        unpack8(raw, x[rr][i]);
      }
    }
    // phase 1: partial <x_rr, U_j> over this thread's columns, then over the CTA
    float part[kProjSmallRows][kProjSmallRank];
#pragma unroll
    for (int rr = 0; rr < kProjSmallRows; ++rr)
#pragma unroll
      for (int j = 0; j < kProjSmallRank; ++j) part[rr][j] = 0.0f;
#pragma unroll
    for (int j = 0; j < kProjSmallRank; ++j) {
      if (j < a.r) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int vec = tid + i * kProjSmallThreads;
          if (vec < nvec) {
            float u[8];
            unpack8(s_ut[j * nvec + vec], u);
#pragma unroll
            for (int rr = 0; rr < kProjSmallRows; ++rr)
#pragma unroll
              for (int e = 0; e < 8; ++e) part[rr][j] = fmaf(x[rr][i][e], u[e], part[rr][j]);
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < kProjSmallRank; ++j) {
      if (j < a.r) {   // uniform over the CTA
#pragma unroll
        for (int rr = 0; rr < kProjSmallRows; ++rr) {
          float v = part[rr][j];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          if (lane == 0) s_part[warp][rr * kProjSmallRank + j] = v;
        }
      }
    }
    __syncthreads();
    if (tid < kProjSmallRows * kProjSmallRank) {
      float c = 0.0f;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) c += s_part[w][tid];   // fixed order
      s_c[tid] = c;
    }
    __syncthreads();
    // phase 2: x' = x - sum_j c_j U[:, j]
#pragma unroll
    for (int j = 0; j < kProjSmallRank; ++j) {
      if (j < a.r) {
        float cj[kProjSmallRows];
#pragma unroll
        for (int rr = 0; rr < kProjSmallRows; ++rr) cj[rr] = -s_c[rr * kProjSmallRank + j];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int vec = tid + i * kProjSmallThreads;
          if (vec < nvec) {
            float u[8];
            unpack8(s_ut[j * nvec + vec], u);
#pragma unroll
            for (int rr = 0; rr < kProjSmallRows; ++rr)
#pragma unroll
              for (int e = 0; e < 8; ++e) x[rr][i][e] = fmaf(cj[rr], u[e], x[rr][i][e]);
          }
        }
      }
    }
    // round to the residual dtype, store, and the fused final RMSNorm
    float ss[kProjSmallRows];
#pragma unroll
    for (int rr = 0; rr < kProjSmallRows; ++rr) {
      const int row = row0 + rr;
      ss[rr] = 0.0f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = tid + i * kProjSmallThreads;
        if (vec < nvec) {
          const Bf16x8 pk = pack8(x[rr][i]);
          if (row < a.n_rows && a.out_x) reinterpret_cast<Bf16x8*>(a.out_x + static_cast<size_t>(row) * a.d)[vec] = pk;
          unpack8(pk, x[rr][i]);
#pragma unroll
          for (int e = 0; e < 8; ++e) ss[rr] = fmaf(x[rr][i][e], x[rr][i][e], ss[rr]);
        }
      }
    }
    if (a.out_h) {
#pragma unroll
      for (int rr = 0; rr < kProjSmallRows; ++rr) {
        float v = ss[rr];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_ss[warp][rr] = v;
      }
      __syncthreads();
#pragma unroll
      for (int rr = 0; rr < kProjSmallRows; ++rr) {
        const int row = row0 + rr;
        float tot = 0.0f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) tot += s_ss[w][rr];
        const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(tot, static_cast<float>(a.d)), a.eps)));
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int vec = tid + i * kProjSmallThreads;
          if (row < a.n_rows && vec < nvec) {
            const float4 w0 = __ldg(reinterpret_cast<const float4*>(a.w_norm) + 2 * vec);
            const float4 w1 = __ldg(reinterpret_cast<const float4*>(a.w_norm) + 2 * vec + 1);
            const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
            float h[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) h[e] = __fmul_rn(__fmul_rn(x[rr][i][e], rstd), __fadd_rn(1.0f, ww[e]));
            reinterpret_cast<Bf16x8*>(a.out_h + static_cast<size_t>(row) * a.d)[vec] = pack8(h);
          }
        }
      }
    }
    __syncthreads();   // s_part / s_c / s_ss are reused by the next tile
  }
}


// -----------------------------------------------------------------------------------------
// K2 for ranks 1..2 in the row-group layout of tb_rmsnorm_kernel (four rows per CTA, 128 threads
// per row, U^T in shared memory): at these ranks the two thin products are a handful of FMAs
// per element, and what matters is the number of rows in flight -- this is the norm kernel with
// a projection spliced in (one extra pair of group barriers for the r dot products).
// -----------------------------------------------------------------------------------------
constexpr int kProjTinyRank = 2;

template <int NV>
__global__ void __launch_bounds__(kAblThreads) tb_lowrank_tiny_kernel(const ProjArgs a) {
  extern __shared__ __align__(16) uint8_t proj_tiny_smem[];
  int4* s_ut = reinterpret_cast<int4*>(proj_tiny_smem);            // [r][nvec]
  __shared__ float s_red[kRowGroups][4];
  __shared__ float s_dot[kRowGroups][4][kProjTinyRank];
  const int group = threadIdx.x / kRowGroupThreads;
  const int gt = threadIdx.x - group * kRowGroupThreads;
  const int gwarp = gt >> 5, lane = gt & 31;
  const int nvec = a.d >> 3;
  for (int i = threadIdx.x; i < a.r * nvec; i += kAblThreads) s_ut[i] = __ldg(reinterpret_cast<const int4*>(a.ut) + i);
  __syncthreads();
  // the next row's x is requested (raw, 16 registers) before this row's two reductions: twice the loads in flight
  const int row_step = gridDim.x * kRowGroups;
  int4 nxt[NV];
  {
    const int row = blockIdx.x * kRowGroups + group;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      nxt[i] = make_int4(0, 0, 0, 0);
      if (row < a.n_rows && vec < nvec) nxt[i] = __ldg(reinterpret_cast<const int4*>(a.x + static_cast<size_t>(row) * a.d) + vec);
    }
  }
  for (int row = blockIdx.x * kRowGroups + group; row < a.n_rows; row += row_step) {
    float xb[NV][8];
    float dot[kProjTinyRank] = {0.0f, 0.0f};
#pragma unroll
    for (int i = 0; i < NV; ++i) unpack8(nxt[i], xb[i]);
    if (row + row_step < a.n_rows) {
      const int4* xn = reinterpret_cast<const int4*>(a.x + static_cast<size_t>(row + row_step) * a.d);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = gt + i * kRowGroupThreads;
        if (vec < nvec) nxt[i] = __ldg(xn + vec);
      }
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      if (vec < nvec) {
#pragma unroll
        for (int j = 0; j < kProjTinyRank; ++j) {
          if (j < a.r) {
            float u[8];
            unpack8(s_ut[j * nvec + vec], u);
#pragma unroll
            for (int e = 0; e < 8; ++e) dot[j] = fmaf(xb[i][e], u[e], dot[j]);
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < kProjTinyRank; ++j) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dot[j] += __shfl_xor_sync(0xffffffffu, dot[j], o);
    }
    group_bar_sync(group);          // s_dot of the previous row is no longer read
    if (lane == 0) {
#pragma unroll
      for (int j = 0; j < kProjTinyRank; ++j) s_dot[group][gwarp][j] = dot[j];
    }
    group_bar_sync(group);
    float c[kProjTinyRank];
#pragma unroll
    for (int j = 0; j < kProjTinyRank; ++j)
      c[j] = -((s_dot[group][0][j] + s_dot[group][1][j]) + (s_dot[group][2][j] + s_dot[group][3][j]));
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = gt + i * kRowGroupThreads;
      if (vec < nvec) {
#pragma unroll
        for (int j = 0; j < kProjTinyRank; ++j) {
          if (j < a.r) {
            float u[8];
            unpack8(s_ut[j * nvec + vec], u);
#pragma unroll
            for (int e = 0; e < 8; ++e) xb[i][e] = fmaf(c[j], u[e], xb[i][e]);
          }
        }
        const Bf16x8 pk = pack8(xb[i]);
        if (a.out_x) reinterpret_cast<Bf16x8*>(a.out_x + static_cast<size_t>(row) * a.d)[vec] = pk;
        unpack8(pk, xb[i]);
#pragma unroll
        for (int e = 0; e < 8; ++e) ss = fmaf(xb[i][e], xb[i][e], ss);
      }
    }
    if (a.out_h) {
      ss = group_sum(ss, s_red[group], group, gwarp, lane);
      const float rstd = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, static_cast<float>(a.d)), a.eps)));
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vec = gt + i * kRowGroupThreads;
        if (vec < nvec) {
          const float4 w0 = __ldg(reinterpret_cast<const float4*>(a.w_norm) + 2 * vec);
          const float4 w1 = __ldg(reinterpret_cast<const float4*>(a.w_norm) + 2 * vec + 1);
          const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
          float h[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) h[e] = __fmul_rn(__fmul_rn(xb[i][e], rstd), __fadd_rn(1.0f, ww[e]));
          reinterpret_cast<Bf16x8*>(a.out_h + static_cast<size_t>(row) * a.d)[vec] = pack8(h);
        }
      }
    }
  }
}


// -----------------------------------------------------------------------------------------
// K6 on the tensor cores, the two small kernels around the block-diagonal GEMM (CandSumEpi):
//   gather   : copy the W_U rows of each prompt's listed candidates into 128 slots per prompt
//              (one warp per slot, 16-byte copies; empty slots are left untouched -- nobody reads
//              their sums)
//   finalize : per prompt, the k best exact sums (value desc, id asc on ties), theta, and the
//              certificate flag.
// -----------------------------------------------------------------------------------------
constexpr int kCandSlots = 128;

__global__ void tb_ll_gather_candidates_kernel(const __nv_bfloat16* __restrict__ w_u, int V, int d, int vocab_offset,
                                               const int32_t* __restrict__ cand_count,
                                               const int32_t* __restrict__ cand_id, int cap, int n_prompts,
                                               __nv_bfloat16* __restrict__ wc) {
  const int warps_per_block = blockDim.x >> 5;
  const int64_t slot = static_cast<int64_t>(blockIdx.x) * warps_per_block + (threadIdx.x >> 5);
  const int p = static_cast<int>(slot / kCandSlots), s = static_cast<int>(slot % kCandSlots);
  if (p >= n_prompts || s >= min(cand_count[p], min(cap, kCandSlots))) return;
  const int v = cand_id[static_cast<size_t>(p) * cap + s] - vocab_offset;
  int4* dst = reinterpret_cast<int4*>(wc + slot * d);
  const int nvec = d >> 3;
  if (v < 0 || v >= V) {
    for (int i = threadIdx.x & 31; i < nvec; i += 32) dst[i] = make_int4(0, 0, 0, 0);
    return;
  }
  const int4* src = reinterpret_cast<const int4*>(w_u + static_cast<size_t>(v) * d);
  for (int i = threadIdx.x & 31; i < nvec; i += 32) dst[i] = __ldg(src + i);
}

struct CandFinalArgs {
  const float* sums;           // [n_prompts, 128]
  const float* lse;            // [n_rows]
  const float* ref;            // [n_rows]
  const float* tau;            // [n_prompts]
  const int32_t* cand_count;   // [n_prompts]
  const int32_t* cand_id;      // [n_prompts, cap]
  int cap, k, n_rows;
  float* out_val;              // [n_prompts, k]
  int32_t* out_id;             // [n_prompts, k]
  float* out_theta;            // [n_prompts]
  int32_t* out_flag;           // [n_prompts]
};

__global__ void __launch_bounds__(kCandSlots) tb_ll_candidates_finalize_kernel(const CandFinalArgs a) {
  __shared__ float s_v[4];
  __shared__ int s_i[4], s_pos[4];
  __shared__ float s_g[4];
  const int p = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_all = a.cand_count[p];
  const int n = min(n_all, min(a.cap, kCandSlots));
  // g_max = max_t exp(ref_t - lse_t) over the slots that hold a token
  float g = 0.0f;
  const int64_t r = static_cast<int64_t>(p) * 128 + tid;
  if (r < a.n_rows && a.ref[r] < CUDART_INF_F) g = expf(a.ref[r] - a.lse[r]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) g = fmaxf(g, __shfl_xor_sync(0xffffffffu, g, o));
  if (lane == 0) s_g[warp] = g;
  float val = -CUDART_INF_F;
  int id = -1;
  if (tid < n) {
    val = a.sums[static_cast<size_t>(p) * kCandSlots + tid];
    id = a.cand_id[static_cast<size_t>(p) * a.cap + tid];
  }
  __syncthreads();
  const float g_max = fmaxf(fmaxf(s_g[0], s_g[1]), fmaxf(s_g[2], s_g[3]));
  float theta = -CUDART_INF_F;
  for (int round = 0; round < a.k; ++round) {
    float bv = val;
    int bi = id, bp = tid;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      const int op = __shfl_xor_sync(0xffffffffu, bp, o);
      if (ov > bv || (ov == bv && oi >= 0 && (bi < 0 || oi < bi))) { bv = ov; bi = oi; bp = op; }
    }
    __syncthreads();
    if (lane == 0) { s_v[warp] = bv; s_i[warp] = bi; s_pos[warp] = bp; }
    __syncthreads();
    float wv = s_v[0];
    int wi = s_i[0], wp = s_pos[0];
#pragma unroll
    for (int w = 1; w < 4; ++w)
      if (s_v[w] > wv || (s_v[w] == wv && s_i[w] >= 0 && (wi < 0 || s_i[w] < wi))) { wv = s_v[w]; wi = s_i[w]; wp = s_pos[w]; }
    if (wi >= 0 && tid == wp) { val = -CUDART_INF_F; id = -1; }   // taken
    if (tid == 0) {
      a.out_val[static_cast<size_t>(p) * a.k + round] = wi >= 0 ? wv : -CUDART_INF_F;
      a.out_id[static_cast<size_t>(p) * a.k + round] = wi;
    }
    theta = wi >= 0 ? wv : -CUDART_INF_F;
  }
  if (tid == 0) {
    a.out_theta[p] = theta;
    const bool ok = n_all <= min(a.cap, kCandSlots) && theta > 0.0f &&
                    theta >= g_max * a.tau[p] * (1.0f + 0.0009765625f);
    a.out_flag[p] = ok ? 1 : 0;
  }
}

// p[r, v] = exp(logit[r, v] - lse[r]) in place over a dense [n_rows, V] fp32 array (the compatibility surface's
// all_probs, src/models.py:138). HBM-bound: one read and one write of the array.
__global__ void tb_probs_from_logits_kernel(float* __restrict__ logits, const float* __restrict__ lse, int64_t n_rows, int V) {
  const int64_t total = n_rows * static_cast<int64_t>(V);
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    logits[i] = expf(logits[i] - __ldg(lse + i / V));
}

// K1 with latent chunks (EncodeEpi::Params::sub_lists): a row's list arrives as `n_sub` sub-lists, each at the start
// of its own cap / n_sub slots. One warp per row moves sub-lists 1.. down to follow their predecessor (in place:
// every round reads 32 entries ahead of anything it writes), writes the total count and flags the row if any
// sub-list did not fit its slots (such rows are encoded once more on their own, like rows whose whole list overflows).
__global__ void tb_merge_sublists_kernel(const int32_t* __restrict__ sub_count, int n_sub, int cap, int n_rows,
                                         int32_t* __restrict__ idx, float* __restrict__ val,
                                         int32_t* __restrict__ out_count, int32_t* __restrict__ out_truncated) {
  const int lane = threadIdx.x & 31;
  const int row = static_cast<int>((blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5);
  if (row >= n_rows) return;
  const int sub = cap / n_sub;
  const int64_t base = static_cast<int64_t>(row) * cap;
  int total = 0, filled = 0, trunc = 0;
  for (int c = 0; c < n_sub; ++c) {
    const int cnt = sub_count[static_cast<size_t>(c) * n_rows + row];
    const int have = min(cnt, sub);
    trunc |= cnt > sub;
    if (c > 0 && filled < c * sub) {
      for (int j = 0; j < have; j += 32) {
        const int i = j + lane;
        int32_t id = 0;
        float v = 0.0f;
        if (i < have) {
          id = idx[base + c * sub + i];
          v = val[base + c * sub + i];
        }
        __syncwarp();
        if (i < have) {
          idx[base + filled + i] = id;
          val[base + filled + i] = v;
        }
      }
      __syncwarp();
    }
    filled += have;
    total += cnt;
  }
  if (lane == 0) {
    out_count[row] = total;
    out_truncated[row] = trunc;
  }
}

// a11 from dense LOGITS, all prompts of a batch in one launch: acc[p, v] = sum over the prompt's 128 row slots of
// exp(logit[t, v] - lse[t]) with the row's current / previous token id zeroed (01_reproduce_logit_lens.py:35-71).
// The first condition of a sweep has no previous logsumexp to lean on (K3F) -- instead of a second sweep of the GEMM
// (K5) the lens kernel writes its soft-capped logits in the one sweep it does (LensDenseEpi, row sub-batches), and
// this kernel reduces them: HBM-bound, rows * V * 4 bytes read once, coalesced. grid = (ceil(V / 256), prompts).
constexpr int kLlSlots = 128;

__global__ void __launch_bounds__(256) tb_logits_ll_colsum_kernel(const float* __restrict__ logits, const float* __restrict__ lse,
                                                                  const int32_t* __restrict__ cur_id,
                                                                  const int32_t* __restrict__ prev_id, int64_t n_rows, int V,
                                                                  float* __restrict__ acc) {
  // a thread owns four consecutive columns (one 16-byte load per row), eight rows in flight: 128 bytes per thread
  __shared__ float s_lse[kLlSlots];
  __shared__ int32_t s_cur[kLlSlots], s_prev[kLlSlots];
  const int p = blockIdx.y;
  const int64_t row0 = static_cast<int64_t>(p) * kLlSlots;
  if (threadIdx.x < kLlSlots) {
    const int64_t r = row0 + threadIdx.x;
    const bool valid = r < n_rows;
    s_lse[threadIdx.x] = valid ? lse[r] : CUDART_INF_F;            // +inf (also the caller's mark of an empty slot): weight 0
    s_cur[threadIdx.x] = (valid && cur_id) ? cur_id[r] : -1;
    s_prev[threadIdx.x] = (valid && prev_id) ? prev_id[r] : -1;
  }
  __syncthreads();
  const int col = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (col >= V) return;
  const bool vec = (V & 3) == 0;                                   // rows are 16-byte aligned and col + 4 <= V
  float a[4] = {0.0f, 0.0f, 0.0f, 0.0f};
  const float* lp = logits + row0 * V + col;
  for (int t0 = 0; t0 < kLlSlots; t0 += 8) {
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      v[u] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      if (s_lse[t0 + u] < CUDART_INF_F) {                          // uniform over the block
        const float* q = lp + static_cast<int64_t>(t0 + u) * V;
        if (vec) v[u] = __ldcs(reinterpret_cast<const float4*>(q));     // streamed: read once
        else {
          v[u].x = __ldg(q);
          if (col + 1 < V) v[u].y = __ldg(q + 1);
          if (col + 2 < V) v[u].z = __ldg(q + 2);
          if (col + 3 < V) v[u].w = __ldg(q + 3);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const float l = s_lse[t0 + u];
      if (l < CUDART_INF_F) {
        const int c = s_cur[t0 + u], q = s_prev[t0 + u];
        const float e[4] = {expf(v[u].x - l), expf(v[u].y - l), expf(v[u].z - l), expf(v[u].w - l)};
#pragma unroll
        for (int j = 0; j < 4; ++j) a[j] = __fadd_rn(a[j], (col + j == c || col + j == q) ? 0.0f : e[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (col + j < V) acc[static_cast<size_t>(p) * V + col + j] = a[j];
}

// The same from logits stored in PROMPT TILES [prompt][V / 32][128 rows][32 columns], in log2 units (LensDenseEpi with tma == 2): a
// prompt's 128 x 32 piece of the array is 16 KB of consecutive memory, so a block streams 8 of them (256 columns,
// 128 KB) front to back -- whole DRAM pages, few TLB entries -- where the row-major array gives 512 bytes per row, 1 MB
// apart. Thread t takes the 16-byte piece t + 256 i of a tile (row t / 8 + 32 i, columns 4 (t % 8) ...), sums its four
// rows, the 32 threads of a column piece are combined by shuffles (4 lanes) and through shared memory (8 warps), in a
// fixed order. grid = (ceil(V / 256), prompts). V % 32 == 0, whole prompts.
constexpr int kLlTilesPerBlock = 8;
__device__ __forceinline__ float ex2_approx_rows(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(256) tb_logits_ll_colsum_tiled_kernel(const float4* __restrict__ lt, const float* __restrict__ lse,
                                                                        const int32_t* __restrict__ cur_id,
                                                                        const int32_t* __restrict__ prev_id, int n_tiles, int V,
                                                                        float* __restrict__ acc) {
  __shared__ float s_lse[kLlSlots];
  __shared__ int32_t s_cur[kLlSlots], s_prev[kLlSlots];
  __shared__ float4 s_red[kLlTilesPerBlock][8][8];                 // [tile][warp][column piece]
  const int p = blockIdx.y, t = threadIdx.x;
  const int tile0 = blockIdx.x * kLlTilesPerBlock;
  if (t < kLlSlots) {
    const int64_t r = static_cast<int64_t>(p) * kLlSlots + t;
    s_lse[t] = lse[r];                                             // +inf (the caller's mark of an empty slot): weight 0
    s_cur[t] = cur_id ? cur_id[r] : -1;
    s_prev[t] = prev_id ? prev_id[r] : -1;
  }
  __syncthreads();
  const int c4 = t & 7, r0 = t >> 3, warp = t >> 5, lane = t & 31;
  float l[4];                                                      // the row's logsumexp in log2 units, like the logits
  int32_t cu[4], pv[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    l[i] = s_lse[r0 + 32 * i] * 1.4426950408889634f;
    cu[i] = s_cur[r0 + 32 * i];
    pv[i] = s_prev[r0 + 32 * i];
  }
  const float4* base = lt + (static_cast<size_t>(p) * n_tiles + tile0) * 1024 + t;
#pragma unroll 1
  for (int tj = 0; tj < kLlTilesPerBlock; tj += 2) {                // two tiles = eight 16-byte loads in flight per thread
    float4 v[2][4];
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int i = 0; i < 4; ++i)
        v[u][i] = (tile0 + tj + u < n_tiles) ? __ldcs(base + (tj + u) * 1024 + 256 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int col = (tile0 + tj + u) * 32 + c4 * 4;
      float a[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float e[4] = {ex2_approx_rows(v[u][i].x - l[i]), ex2_approx_rows(v[u][i].y - l[i]), ex2_approx_rows(v[u][i].z - l[i]),
                      ex2_approx_rows(v[u][i].w - l[i])};
        const uint32_t dc = static_cast<uint32_t>(cu[i] - col), dp = static_cast<uint32_t>(pv[i] - col);
        if (dc < 4u || dp < 4u) {                                 // a masked id among my four columns: rare
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (dc == static_cast<uint32_t>(j) || dp == static_cast<uint32_t>(j)) e[j] = 0.0f;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) a[j] = __fadd_rn(a[j], e[j]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        a[j] = __fadd_rn(a[j], __shfl_xor_sync(0xffffffffu, a[j], 8));
        a[j] = __fadd_rn(a[j], __shfl_xor_sync(0xffffffffu, a[j], 16));
      }
      if (lane < 8) s_red[tj + u][warp][lane] = make_float4(a[0], a[1], a[2], a[3]);
    }
  }
  __syncthreads();
  if (t < kLlTilesPerBlock * 8) {
    const int tile = t >> 3, c = t & 7;
    if (tile0 + tile < n_tiles) {
      float4 r = s_red[tile][0][c];
#pragma unroll
      for (int w = 1; w < 8; ++w) {
        const float4 o = s_red[tile][w][c];
        r.x = __fadd_rn(r.x, o.x); r.y = __fadd_rn(r.y, o.y); r.z = __fadd_rn(r.z, o.z); r.w = __fadd_rn(r.w, o.w);
      }
      *reinterpret_cast<float4*>(acc + static_cast<size_t>(p) * V + (tile0 + tile) * 32 + c * 4) = r;
    }
  }
}

}  // namespace tb
