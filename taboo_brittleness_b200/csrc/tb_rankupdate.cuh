// Nested ablation budgets as rank updates of stored base logits (Execution Plan:126 budgets).
//
// For the error-term ablation x' = x - sum_{j in S_m} a_j W_dec[j] the lens logits of budget m are
//   l_m[t, v] = rstd_m[t] * ( L0[t, v] - sum_{j in S_m} a_j[t] D[j, v] ),
//   L0 = (x (.) (1 + w)) W_U^T,   D = (W_dec (.) (1 + w)) W_U^T   [S, V] (computed once per weight set, kept tile-major),
// with rstd_m the RMSNorm factor of the rounded x' (K1b writes it, tb_sae_ablate_ranked out_rstd). L0 is what
// costs 2 d V FLOP per row; it is swept ONCE per row sub-batch (LensDenseEpi, raw mode, prompt tiles) and every
// budget then needs a product of depth m <= 1024 instead of d = 3584, plus the same readout epilogue.
//
// One CTA per (prompt, vocabulary chunk): M = the prompt's 128 rows, N = 256 ids per tile, K = the prompt's
// ranked latents in slabs of 64.
//   A [rows, 1024 (+1024)] bf16, K-major: a_j[t] in column rank(j) (tb_rank_acts_dense_kernel; optionally the bf16
//     remainders in columns 1024 + rank), loaded by TMA;
//   B = the rows D[id_j, tile] of the prompt's own latents: an MN-major operand GATHERED by two or four producer warps with
//     16-byte cp.async straight into the 128-byte-swizzled layout the UMMA descriptor expects (one warp instruction
//     = one latent's 512 bytes); rows at or beyond the budget are zero-filled, so one A matrix serves all budgets;
//   accumulator U in TMEM (two tiles in flight); the epilogue warps fetch their [32 rows x 32 ids] piece of L0 with a
//     TMA load of their own (4 KB, double-buffered), form rstd * (L0 - U) and hand it to the lens readout epilogue
//     (LensEpi's state and arithmetic; 2 or 4 partials per (chunk, row): tb_lens_rank_update_parts). With `store` the
//     soft-capped logits (log2 units) go back into the buffer the L0 piece came in and one lane TMA-stores it to a second
//     prompt-tile array, from which tb_logits_ll_colsum gets the budget's LL-Top-k.
// All CTAs work on the same vocabulary chunk at the same time (unit = chunk * prompts + prompt), so a tile's slab of
// D (8 MB for 16384 latents, contiguous in the tile-major layout) can be served from L2.
//
// This changes the arithmetic of the reference path (the normalised residual no longer passes through bf16 per
// budget; D is a rounded product): it is a separate workload with its drift reported, never the dense line.
#pragma once
#include "tb_epilogues.cuh"
#include "tb_cluster_rows.cuh"   // umma_desc_mn_sw128, kIdescBMajorMn

namespace tb {

constexpr int kRuSlabK = 64;
constexpr int kRuABytes = 128 * kRuSlabK * 2;            // 16 KB: [128 rows x 64 k], K-major, 128-byte swizzle
constexpr int kRuBGroupBytes = kRuSlabK * 128;           // 8 KB: 64 k-rows x 64 ids of one column group
constexpr int kRuBBytes = 4 * kRuBGroupBytes;            // 32 KB: four column groups = 256 ids
constexpr int kRuStageBytes = kRuABytes + kRuBBytes;
constexpr int kRuL0Bytes = 32 * 128;                     // one epilogue warp's [32 rows x 32 ids] fp32 piece
// What bounds the kernel (profiles/r02ru_update_kernel_what_bounds_it.log, 65,536 rows, budgets <= 128): not the L0
// stream (deeper prefetch changes nothing), not the gather, but the readout arithmetic of eight epilogue warps -- two
// per scheduler, each a chain of TMEM load -> soft cap -> max / exp-sum / top-10 with little to overlap: 23 of 37 ms.
// Hence kEW = 16 epilogue warps: four per TMEM lane quadrant, each owning 64 of a tile's 256 ids (two L0 pieces, worked
// through in 16-id steps to stay inside 96 registers): 37.5 -> 20 ms per budget. That leaves room for two operand
// stages; budget 1024 (16 slabs per tile, 768 KB of A and D rows from L2 per tile: bound by the gather) keeps eight
// epilogue warps, three stages and four gather warps (65 vs 84 ms; 60 with the tile-major D).
template <int kStages, int kNB, int kEW>
struct RuCfg {
  static constexpr int kGatherWarps = kEW == 16 ? 2 : 4;
  static constexpr int kGatherWarp0 = kEW == 16 ? 2 : 4;
  static constexpr int kEpiWarp0 = kEW == 16 ? 4 : 8;     // a multiple of 4: warp % 4 is the TMEM lane quadrant
  static constexpr int kAllocWarp = kEW == 16 ? 0 : 2;
  static constexpr int kThreads = 32 * (kEpiWarp0 + kEW);
  static constexpr int kColGroups = kEW / 4;              // column groups of a tile, one partial each
  static constexpr int kColsPerWarp = 256 / kColGroups;   // 128 or 64 ids of every tile
  static constexpr int kPiecesPerTile = kColsPerWarp / 32;
  static constexpr int kOffL0 = kStages * kRuStageBytes;
  static constexpr int kOffBars = kOffL0 + kEW * kNB * kRuL0Bytes;
  static constexpr int kSmemBytes = kOffBars + 512 + 1024;    // + barriers + alignment slack
  static_assert(kSmemBytes <= 232448, "shared memory");
  static_assert(2 * kStages + 4 + kEW * kNB + 1 <= 64, "barrier block");
};

struct RankUpdateArgs {
  int n_prompts;               // rows = 128 * n_prompts
  int V, n_tiles, n_chunks, tiles_per_chunk;
  int m;                       // budget: ranks < m are ablated (1 .. 1024)
  int n_slabs;                 // ceil(m / 64)
  int parts;                   // 1: A holds bf16(a); 2: also the remainders a - bf16(a), a_lo_off columns further
  int a_lo_off;
  const int32_t* rank_ids;     // [n_prompts, rank_stride]: latent id per rank
  int rank_stride;
  const __nv_bfloat16* D;      // [n_tiles][S][256]: tile-major, so that the rows one tile step gathers lie within S * 512 bytes
                               // (a handful of pages) instead of one 2 MB page per four rows of the [S, V] matrix
  int S;
  const float* rstd;           // [rows]
  int store;                   // 1: also write the soft-capped logits (log2 units, prompt tiles: tmOut) for tb_logits_ll_colsum
  int debug;                   // experiments only (TB_RU_DEBUG; results are garbage): 1 = no L0 loads, 2 = no gather traffic,
                               // 4 = no readout arithmetic
};

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int32_t c0, int32_t c1,
                                            int32_t c2, int32_t c3) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

// A[row, rank] = bf16(a) (and A[row, lo_off + rank] = bf16(a - bf16(a))) from the rank-ordered lists of
// tb_rank_lists_kernel; everything else zero. One warp per row.
__global__ void __launch_bounds__(256) tb_rank_acts_dense_kernel(const int64_t* __restrict__ row_start, int cap,
                                                                 const int32_t* __restrict__ sel_count,
                                                                 const int32_t* __restrict__ sel_key,
                                                                 const float* __restrict__ sel_val, int64_t n_rows, int a_cols,
                                                                 int lo_off, __nv_bfloat16* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= n_rows) return;
  __nv_bfloat16* o = out + row * a_cols;
  for (int i = lane; i < a_cols / 8; i += 32) reinterpret_cast<uint4*>(o)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncwarp();
  const int64_t base = row_start ? row_start[row] : row * cap;
  const int n = sel_count[row];
  for (int e = lane; e < n; e += 32) {
    const int rank = sel_key[base + e] >> 20;
    if (rank >= kRankedMax) continue;
    const float a = sel_val[base + e];
    const __nv_bfloat16 hi = __float2bfloat16_rn(a);
    o[rank] = hi;
    if (lo_off > 0) o[lo_off + rank] = __float2bfloat16_rn(a - __bfloat162float(hi));
  }
}

// out[r, c] = bf16(x[r, c] * (1 + w[c])): the operand of the base sweep L0 = (x (.) (1 + w)) W_U^T
__global__ void __launch_bounds__(256) tb_scale_cols_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                                                            int64_t n_vec, int d_vec, __nv_bfloat16* __restrict__ out) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_vec;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % d_vec);
    float v[8];
    unpack8(__ldcs(reinterpret_cast<const int4*>(x) + i), v);
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(w) + 2 * c), w1 = __ldg(reinterpret_cast<const float4*>(w) + 2 * c + 1);
    const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = __fmul_rn(v[e], __fadd_rn(1.0f, ww[e]));
    const Bf16x8 p = pack8(v);
    reinterpret_cast<int4*>(out)[i] = *reinterpret_cast<const int4*>(&p);
  }
}

template <int kRuStages, int kNB, int kEW>
__global__ void __launch_bounds__((RuCfg<kRuStages, kNB, kEW>::kThreads), 1)
tb_rank_update_lens_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmL0,
                           const __grid_constant__ CUtensorMap tmOut, const RankUpdateArgs a, const LensEpi::Params lens) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const uint32_t smem0 = smem_u32(smem);
  using Cfg = RuCfg<kRuStages, kNB, kEW>;
  constexpr int kRuOffL0 = Cfg::kOffL0;
  constexpr int kRuGatherWarps = Cfg::kGatherWarps;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kOffBars);
  uint64_t* full_bar = bars;                       // [stages] A bytes (TMA) + one arrive per gather warp
  uint64_t* empty_bar = bars + kRuStages;          // [stages] MMA commit
  uint64_t* tfull_bar = bars + 2 * kRuStages;      // [2]
  uint64_t* tempty_bar = bars + 2 * kRuStages + 2; // [2] one arrive per epilogue warp
  uint64_t* l0_bar = bars + 2 * kRuStages + 4;     // [kEW warps][kNB]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kRuStages + 4 + kEW * kNB);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(threadIdx.x & 31);
  const int total_units = a.n_prompts * a.n_chunks;
  const int slabs_per_tile = a.n_slabs * a.parts;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmL0);
    if (a.store) tma_prefetch_desc(&tmOut);
    for (int i = 0; i < kRuStages; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1 + kRuGatherWarps);
      mbar_init(smem_u32(&empty_bar[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&tfull_bar[i]), 1);
      mbar_init(smem_u32(&tempty_bar[i]), kEW);
    }
    for (int i = 0; i < kEW * kNB; ++i) mbar_init(smem_u32(&l0_bar[i]), 1);
    fence_mbar_init();
  }
  __syncwarp();
  if (warp == Cfg::kAllocWarp) tmem_alloc<1>(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  // unit -> (chunk, prompt, tile range): chunk-major, so that all CTAs sweep the same ids at the same time
  auto decode = [&](int unit, int& prompt, int& chunk, int& t0, int& t1) {
    chunk = unit / a.n_prompts;
    prompt = unit - chunk * a.n_prompts;
    t0 = chunk * a.tiles_per_chunk;
    t1 = min(t0 + a.tiles_per_chunk, a.n_tiles);
  };

  if (warp == 0) {
    // ================================ A slabs (TMA) ================================
    if (lane == 0) {
      uint32_t it = 0;
      for (int unit = blockIdx.x; unit < total_units; unit += gridDim.x) {
        int prompt, chunk, t0, t1;
        decode(unit, prompt, chunk, t0, t1);
        for (int nt = t0; nt < t1; ++nt)
          for (int s = 0; s < slabs_per_tile; ++s, ++it) {
            const uint32_t stage = it % kRuStages, ph = (it / kRuStages) & 1u;
            mbar_wait(smem_u32(&empty_bar[stage]), ph ^ 1u);
            mbar_arrive_expect_tx(smem_u32(&full_bar[stage]), kRuABytes);
            const int part = s / a.n_slabs, sl = s - part * a.n_slabs;
            tma_load_2d(smem0 + stage * kRuStageBytes, &tmA, smem_u32(&full_bar[stage]), part * a.a_lo_off + sl * kRuSlabK,
                        prompt * 128);
          }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    constexpr uint32_t idesc = umma_idesc_bf16_f32(128, 256) | kIdescBMajorMn;
    uint32_t it = 0, acc_it = 0;
    for (int unit = blockIdx.x; unit < total_units; unit += gridDim.x) {
      int prompt, chunk, t0, t1;
      decode(unit, prompt, chunk, t0, t1);
      for (int nt = t0; nt < t1; ++nt, ++acc_it) {
        const uint32_t acc = acc_it & 1u;
        mbar_wait(smem_u32(&tempty_bar[acc]), ((acc_it >> 1) & 1u) ^ 1u);
        tc_fence_after();
        for (int s = 0; s < slabs_per_tile; ++s, ++it) {
          const uint32_t stage = it % kRuStages, ph = (it / kRuStages) & 1u;
          mbar_wait(smem_u32(&full_bar[stage]), ph);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = smem0 + stage * kRuStageBytes, sb = sa + kRuABytes;
            const uint64_t da = umma_desc_k_sw128(sa);
#pragma unroll
            for (int k = 0; k < kRuSlabK / 16; ++k) {
              const uint64_t db = umma_desc_mn_sw128(sb + k * 16 * 128, kRuBGroupBytes);
              umma_bf16<1>(tmem_base + acc * 256, da + static_cast<uint64_t>(2 * k), db, idesc, (s | k) != 0 ? 1u : 0u);
            }
            umma_commit(smem_u32(&empty_bar[stage]));
            if (s == slabs_per_tile - 1) umma_commit(smem_u32(&tfull_bar[acc]));
          }
          __syncwarp();
        }
      }
    }
  } else if (warp >= Cfg::kGatherWarp0 && warp < Cfg::kGatherWarp0 + kRuGatherWarps) {
    // ================================ B gather (cp.async) ================================
    const int pw = warp - Cfg::kGatherWarp0;
    constexpr int kRowsPerWarp = kRuSlabK / kRuGatherWarps;   // 16 or 32 k-rows of every slab
    uint32_t it = 0;
    int prev_stage = -1;
    for (int unit = blockIdx.x; unit < total_units; unit += gridDim.x) {
      int prompt, chunk, t0, t1;
      decode(unit, prompt, chunk, t0, t1);
      const int32_t* ids = a.rank_ids + static_cast<size_t>(prompt) * a.rank_stride;
      for (int nt = t0; nt < t1; ++nt) {
        const __nv_bfloat16* d_tile = a.D + static_cast<size_t>(nt) * a.S * 256 + lane * 8;   // this lane's 8 ids of the tile
        for (int s = 0; s < slabs_per_tile; ++s, ++it) {
          const uint32_t stage = it % kRuStages, ph = (it / kRuStages) & 1u;
          const int sl = s % a.n_slabs;
          // ids of my rows of the slab (rows pw, pw + kRuGatherWarps, ...): lane jj fetches row jj's
          int my_id = 0;
          {
            const int r = sl * kRuSlabK + (lane % kRowsPerWarp) * kRuGatherWarps + pw;
            if (lane < kRowsPerWarp && r < a.m) my_id = __ldg(ids + r);
          }
          mbar_wait(smem_u32(&empty_bar[stage]), ph ^ 1u);
          const uint32_t sb = smem0 + stage * kRuStageBytes + kRuABytes + (lane >> 3) * kRuBGroupBytes;
#pragma unroll 4
          for (int jj = 0; jj < kRowsPerWarp; ++jj) {
            const int j = jj * kRuGatherWarps + pw;           // k-row of the slab
            const int id = __shfl_sync(0xffffffffu, my_id, jj);
            const bool ok = (sl * kRuSlabK + j < a.m) && !(a.debug & 2);      // (ids past V are zero padding in D)
            const __nv_bfloat16* src = d_tile + (ok ? static_cast<size_t>(id) * 256 : 0);
            cp_async16(sb + j * 128 + (((lane & 7) ^ (j & 7)) << 4), src, ok ? 16u : 0u);
          }
          cp_async_commit();
          if (prev_stage >= 0) {                              // the previous slab's rows have landed: publish it
            cp_async_wait<1>();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&full_bar[prev_stage]));
          }
          prev_stage = static_cast<int>(stage);
        }
      }
    }
    if (prev_stage >= 0) {
      cp_async_wait<0>();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&full_bar[prev_stage]));
    }
  } else if (warp >= Cfg::kEpiWarp0) {
    // ================================ epilogue ================================
    const int ew = warp - Cfg::kEpiWarp0, quad = warp & 3, cg = ew >> 2;   // cg: which column group of every tile
    constexpr int kPPT = Cfg::kPiecesPerTile;
    const int col_base = cg * Cfg::kColsPerWarp;
    const int row_in_blk = quad * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(quad * 32) << 16;
    const uint32_t l0_buf = smem0 + kRuOffL0 + ew * kNB * kRuL0Bytes;
    const uint32_t l0_b0 = smem_u32(&l0_bar[ew * kNB]);
    GemmArgs g;
    g.M = a.n_prompts * 128;
    g.N = a.V;
    const CapLog2 cap2(lens.softcap);
    const uint64_t evict_first = l2_policy_evict_first();
    LensEpi::State st;
    uint32_t acc_it = 0, pc = 0;                              // tiles / L0 pieces consumed so far
    for (int unit = blockIdx.x; unit < total_units; unit += gridDim.x) {
      int prompt, chunk, t0, t1;
      decode(unit, prompt, chunk, t0, t1);
      const int row = prompt * 128 + row_in_blk;
      const float rk1 = __ldg(a.rstd + row) * cap2.k1;      // the row's RMSNorm factor and the soft cap's input scale in one
      LensEpi::begin_unit(st, lens, g, row, chunk * Cfg::kColGroups + cg);
      const int n_pieces = (t1 - t0) * kPPT;
      // piece q of the unit: tile t0 + q / kPPT, columns col_base + (q % kPPT) * 32 of it
      auto request = [&](int q, uint32_t slot) {
        if (lane == 0 && !(a.debug & 1)) {
          if (a.store) bulk_wait_read0();                     // the logits stored from this buffer have been read out of it
          const int c32 = ((t0 + q / kPPT) * 256 + col_base + (q % kPPT) * 32) >> 5;
          mbar_arrive_expect_tx(l0_b0 + slot * 8, kRuL0Bytes);
          tma_load_4d(l0_buf + slot * kRuL0Bytes, &tmL0, l0_b0 + slot * 8, 0, quad * 32, c32, prompt);
        }
      };
#pragma unroll
      for (int q = 0; q < kNB - 1; ++q)                       // kNB - 1 pieces ahead
        if (q < n_pieces) request(q, (pc + q) % kNB);
      for (int nt = t0; nt < t1; ++nt, ++acc_it) {
        const uint32_t acc = acc_it & 1u;
        mbar_wait(smem_u32(&tfull_bar[acc]), (acc_it >> 1) & 1u);
        tc_fence_after();
#pragma unroll 1
        for (int cq = 0; cq < kPPT; ++cq, ++pc) {
          const int q = (nt - t0) * kPPT + cq;
          const uint32_t slot = pc % kNB;
          __syncwarp();                                       // everybody is done with the buffer of the piece before this one
          if (q + kNB - 1 < n_pieces) request(q + kNB - 1, (pc + kNB - 1) % kNB);
          const int col0 = nt * 256 + col_base + cq * 32;
          const uint32_t taddr = tmem_base + lane_base + acc * 256 + col_base + cq * 32;
          const uint32_t l0_row = l0_buf + slot * kRuL0Bytes + lane * 128;
          if constexpr (kEW == 16) {
#pragma unroll 1
            for (int hf = 0; hf < 2; ++hf) {                  // 16 ids at a time: 96 registers per thread
              uint32_t v[16];
              tmem_ld_32x16(taddr + hf * 16, v);
              if (hf == 0 && !(a.debug & 1)) mbar_wait(l0_b0 + slot * 8, (pc / kNB) & 1u);
              tmem_ld_wait();
              float c[16];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                float4 l0;
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(l0.x), "=f"(l0.y), "=f"(l0.z), "=f"(l0.w)
                             : "r"(l0_row + (((hf * 4 + k) ^ (lane & 7)) << 4)));
                c[4 * k] = cap2.from_scaled(rk1 * (l0.x - __uint_as_float(v[4 * k])));
                c[4 * k + 1] = cap2.from_scaled(rk1 * (l0.y - __uint_as_float(v[4 * k + 1])));
                c[4 * k + 2] = cap2.from_scaled(rk1 * (l0.z - __uint_as_float(v[4 * k + 2])));
                c[4 * k + 3] = cap2.from_scaled(rk1 * (l0.w - __uint_as_float(v[4 * k + 3])));
              }
              if (a.store) {                                 // the capped logits replace the base logits in the buffer
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(l0_row + (((hf * 4 + k) ^ (lane & 7)) << 4)),
                               "f"(c[4 * k]), "f"(c[4 * k + 1]), "f"(c[4 * k + 2]), "f"(c[4 * k + 3]) : "memory");
              }
              if (!(a.debug & 4)) LensEpi::process_capped_w<16>(st, lens, g, c, col0 + hf * 16);
              else {
#pragma unroll
                for (int k = 0; k < 16; ++k) st.s += c[k];
              }
            }
          } else {
            uint32_t v[32];
            tmem_ld_32x32(taddr, v);
            if (!(a.debug & 1)) mbar_wait(l0_b0 + slot * 8, (pc / kNB) & 1u);
            tmem_ld_wait();
            float c[32];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              float4 l0;
              asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                           : "=f"(l0.x), "=f"(l0.y), "=f"(l0.z), "=f"(l0.w)
                           : "r"(l0_row + ((k ^ (lane & 7)) << 4)));
              c[4 * k] = cap2.from_scaled(rk1 * (l0.x - __uint_as_float(v[4 * k])));
              c[4 * k + 1] = cap2.from_scaled(rk1 * (l0.y - __uint_as_float(v[4 * k + 1])));
              c[4 * k + 2] = cap2.from_scaled(rk1 * (l0.z - __uint_as_float(v[4 * k + 2])));
              c[4 * k + 3] = cap2.from_scaled(rk1 * (l0.w - __uint_as_float(v[4 * k + 3])));
            }
            if (a.store) {
#pragma unroll
              for (int k = 0; k < 8; ++k)
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(l0_row + ((k ^ (lane & 7)) << 4)),
                             "f"(c[4 * k]), "f"(c[4 * k + 1]), "f"(c[4 * k + 2]), "f"(c[4 * k + 3]) : "memory");
            }
            if (!(a.debug & 4)) LensEpi::process_capped(st, lens, g, c, col0);
            else {
#pragma unroll
              for (int k = 0; k < 32; ++k) st.s += c[k];
            }
          }
          if (a.store) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && col0 < a.V) {
              tma_store_4d_hint(&tmOut, l0_buf + slot * kRuL0Bytes, 0, quad * 32, col0 >> 5, prompt, evict_first);
              bulk_commit();
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
      }
      if (a.store && lane == 0) bulk_wait0();                 // this unit's logits are written
      LensEpi::end_unit(st, lens, g, row, chunk * Cfg::kColGroups + cg);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == Cfg::kAllocWarp) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

}  // namespace tb
