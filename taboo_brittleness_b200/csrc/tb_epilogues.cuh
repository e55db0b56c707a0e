// Per-row epilogues plugged into tb_umma_gemm_kernel. Each epilogue thread owns one row of
// the CTA's 128-row block and sees that row's accumulator 32 columns at a time.
#pragma once
#include <math_constants.h>
#include "tb_gemm.cuh"
#include "tb_rows.cuh"

namespace tb {

constexpr int kTopK = 10;
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// (tanh.approx.f32 is only good to 2^-11 relative -- 0.015 in a logit capped at 30, too coarse for 1e-3 in p --
// so the softcap below goes through ex2 + rcp: two MUFU ops, absolute error ~1e-7.)
constexpr float kLn2 = 0.6931471805599453f;
// Soft-capped logit in log2 units, 2 MUFU + 3 FP ops: with k1 = 2 log2(e) / cap and k2 = cap log2(e),
//   cap * tanh(x / cap) * log2(e) = k2 - 2 k2 / (1 + 2^(x k1)).
// Everything downstream (running max, exponentials, top-10 comparisons) works in log2 units; values
// are converted back (x ln 2) only where they leave the kernel. cap <= 0: plain x log2(e).
struct CapLog2 {
  float k1, k2, m2k2;
  bool on;
  __device__ __forceinline__ explicit CapLog2(float cap)
      : k1(cap > 0.0f ? 2.0f * kLog2e / cap : kLog2e), k2(cap * kLog2e), m2k2(-2.0f * cap * kLog2e), on(cap > 0.0f) {}
  __device__ __forceinline__ float operator()(float x) const {
    if (!on) return x * k1;
    const float r = rcp_approx(1.0f + ex2_approx(x * k1));
    return fmaf(r, m2k2, k2);
  }
  // the same for an argument that already carries the factor k1 (a caller with a per-row scale folds the two)
  __device__ __forceinline__ float from_scaled(float xk1) const {
    if (!on) return xk1;
    const float r = rcp_approx(1.0f + ex2_approx(xk1));
    return fmaf(r, m2k2, k2);
  }
};

// =========================================================================================
// K3: logit-lens readout. A = h (final-RMSNorm'ed residual, bf16), B = W_U shard.
// Per (row, vocab chunk) partial: running max m, sumexp s (relative to m), the secret
// token's (capped) logit, and the ten largest (capped) logits with their global vocab ids.
// Reference: src/models.py:135-138 (norm -> lm_head -> softmax), plots.py:10 / 01:142 (secret
// probability gather), models.py:147 (per-row max); softcap per installed Gemma2ForCausalLM.
// =========================================================================================
struct LensEpi {
  static constexpr int kEpiWarps = 8;   // two threads per row, each sweeping half of every tile's columns
  static constexpr bool kSweepA = false;
  static constexpr bool kTileHooks = false;
  static constexpr int kExtraSmem = 0;
  struct Params {
    const int32_t* secret_id;  // [M] global vocab id per row, or nullptr
    int vocab_offset;          // global id of this shard's column 0
    float softcap;             // <= 0: none (reference parity mode); Gemma-2: 30.0
    float* pm;                 // [n_chunks, M]
    float* ps;                 // [n_chunks, M]
    float* psec;               // [n_chunks, M]  (-inf where the secret id is not in the chunk)
    float* ptv;                // [n_chunks, M, 10]
    int32_t* pti;              // [n_chunks, M, 10]
  };
  struct State {
    float m, s, sec;
    float tv[kTopK];
    int32_t ti[kTopK];
    int32_t sid;  // secret column relative to this shard (or a value that never matches)
  };

  __device__ __forceinline__ static void begin_unit(State& st, const Params& p, const GemmArgs& g,
                                                    int row, int /*chunk*/) {
    st.m = -CUDART_INF_F;
    st.s = 0.0f;
    st.sec = -CUDART_INF_F;
#pragma unroll
    for (int k = 0; k < kTopK; ++k) { st.tv[k] = -CUDART_INF_F; st.ti[k] = -1; }
    st.sid = -0x40000000;
    if (p.secret_id != nullptr && row < g.M) st.sid = p.secret_id[row] - p.vocab_offset;
  }

  __device__ __forceinline__ static void insert(State& st, float v, int32_t id) {
    st.tv[kTopK - 1] = v;
    st.ti[kTopK - 1] = id;
#pragma unroll
    for (int k = kTopK - 1; k > 0; --k) {
      const bool sw = st.tv[k] > st.tv[k - 1];
      const float a = st.tv[k - 1], b = st.tv[k];
      const int32_t ia = st.ti[k - 1], ib = st.ti[k];
      st.tv[k - 1] = sw ? b : a;
      st.tv[k] = sw ? a : b;
      st.ti[k - 1] = sw ? ib : ia;
      st.ti[k] = sw ? ia : ib;
    }
  }

  __device__ __forceinline__ static void process(State& st, const Params& p, const GemmArgs& g,
                                                 const uint32_t (&v)[32], int /*row*/, int col0) {
    // c: soft-capped logits in log2 units
    const CapLog2 cap2(p.softcap);
    float c[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) c[j] = cap2(__uint_as_float(v[j]));
    process_capped(st, p, g, c, col0);
  }
  // The same with the soft-capped logits (log2 units) already in hand.
  __device__ __forceinline__ static void process_capped(State& st, const Params& p, const GemmArgs& g,
                                                        float (&c)[32], int col0) {
    process_capped_w<32>(st, p, g, c, col0);
  }
  template <int W>
  __device__ __forceinline__ static void process_capped_w(State& st, const Params& p, const GemmArgs& g,
                                                          float (&c)[W], int col0) {
    if (col0 + W > g.N) {  // ragged last tile: columns past N hold zero-filled garbage
#pragma unroll
      for (int j = 0; j < W; ++j)
        if (col0 + j >= g.N) c[j] = -CUDART_INF_F;
    }
    float cmax = c[0];
#pragma unroll
    for (int j = 1; j < W; ++j) cmax = fmaxf(cmax, c[j]);

    // online logsumexp (base 2), rescaled at most once per 32 columns
    if (cmax > st.m) {
      st.s *= ex2_approx(st.m - cmax);  // m = -inf -> factor 0 (s is 0 then anyway)
      st.m = cmax;
    }
    if (st.m > -CUDART_INF_F) {
      float acc0 = 0.0f, acc1 = 0.0f;
#pragma unroll
      for (int j = 0; j < W; j += 2) {
        acc0 += ex2_approx(c[j] - st.m);
        acc1 += ex2_approx(c[j + 1] - st.m);
      }
      st.s += acc0 + acc1;
    }

    // running top-10 (strict '>' keeps the lowest column among equal values)
    if (cmax > st.tv[kTopK - 1]) {
#pragma unroll
      for (int j = 0; j < W; ++j)
        if (c[j] > st.tv[kTopK - 1]) insert(st, c[j], p.vocab_offset + col0 + j);
    }

    // secret-token logit
    const uint32_t rel = static_cast<uint32_t>(st.sid - col0);
    if (rel < static_cast<uint32_t>(W)) {
#pragma unroll
      for (int j = 0; j < W; ++j)
        if (rel == static_cast<uint32_t>(j)) st.sec = c[j];
    }
  }

  __device__ __forceinline__ static void end_unit(State& st, const Params& p, const GemmArgs& g,
                                                  int row, int chunk) {
    if (row >= g.M) return;
    const size_t o = static_cast<size_t>(chunk) * g.M + row;
    p.pm[o] = st.m * kLn2;      // state is in log2 units; partials are in natural-log units
    p.ps[o] = st.s;
    p.psec[o] = st.sec * kLn2;
#pragma unroll
    for (int k = 0; k < kTopK; ++k) {
      p.ptv[o * kTopK + k] = st.tv[k] * kLn2;
      p.pti[o * kTopK + k] = st.ti[k];
    }
  }
};

// =========================================================================================
// K1: Gemma Scope JumpReLU SAE encode. A = x (bf16 residual rows), B = W_enc^T [S, d].
//   pre = x . W_enc + b_enc ; act = relu(pre) * (pre > theta)        (sae_lens JumpReLU,
//   call site src/02_run_sae_baseline.py:67)
// The dense [rows, S] activation matrix is never formed: each row's non-zero latents are
// appended (in latent order) to that row's list. An optional ablation table drops latents
// whose rank is below the budget right here (random-control / second-pass masking).
// =========================================================================================
struct EncodeEpi {
  static constexpr int kEpiWarps = 4;   // one thread per row: a row's list is appended in latent order
  static constexpr bool kSweepA = false;
  static constexpr bool kTileHooks = false;
  static constexpr int kExtraSmem = 0;
  struct Params {
    const float* b_enc;        // [S]
    const float* theta;        // [S]
    const int64_t* row_start;  // [M] start of each row's list, or nullptr -> row * cap
    int cap;                   // list capacity per row when row_start == nullptr
    const int32_t* row_cap;    // [M] capacity per row when row_start != nullptr (nullable = unchecked)
    int32_t* out_idx;          // latent ids   (nullable: count only)
    float* out_val;            // activations
    int32_t* out_count;        // [M] number of active latents (may exceed the capacity)
    const uint16_t* rank;      // ablation rank table(s) [n_tables, S] or nullptr
    int rank_rows;             // rows sharing one table (0: a single table for all rows)
    int budget;                // latents with rank < budget are zeroed
    int sub_lists;             // > 1 (fixed-stride lists only): the sweep is cut into that many latent chunks, chunk c
                               // of a row writes slots [c * cap / sub_lists, (c + 1) * cap / sub_lists) and its count to
                               // out_count[c * M + row]; tb_merge_sublists_kernel closes the gaps afterwards
  };
  struct State {
    int k;
    int limit;
    int64_t base;
    const uint16_t* rk;
  };

  __device__ __forceinline__ static void begin_unit(State& st, const Params& p, const GemmArgs& g,
                                                    int row, int chunk) {
    st.k = 0;
    st.limit = 0;
    st.base = 0;
    st.rk = nullptr;
    if (row >= g.M) return;
    if (p.out_idx != nullptr) {
      if (p.row_start != nullptr) {
        st.base = p.row_start[row];
        st.limit = p.row_cap != nullptr ? p.row_cap[row] : 0x7fffffff;
      } else if (p.sub_lists > 1) {
        st.limit = p.cap / p.sub_lists;
        st.base = static_cast<int64_t>(row) * p.cap + static_cast<int64_t>(chunk) * st.limit;
      } else {
        st.base = static_cast<int64_t>(row) * p.cap;
        st.limit = p.cap;
      }
    }
    if (p.rank != nullptr)
      st.rk = p.rank + (p.rank_rows > 0 ? static_cast<size_t>(row / p.rank_rows) * g.N : 0);
  }

  __device__ __forceinline__ static void emit(State& st, const Params& p, int col, float pre) {
    if (st.rk != nullptr && static_cast<int>(st.rk[col]) < p.budget) return;  // ablated
    if (st.k < st.limit) {
      p.out_idx[st.base + st.k] = col;
      p.out_val[st.base + st.k] = pre;
    }
    ++st.k;
  }

  __device__ __forceinline__ static void process(State& st, const Params& p, const GemmArgs& g,
                                                 const uint32_t (&v)[32], int row, int col0) {
    if (row >= g.M) return;
    if (col0 + 32 <= g.N) {
      const float4* b4 = reinterpret_cast<const float4*>(p.b_enc + col0);
      const float4* t4 = reinterpret_cast<const float4*>(p.theta + col0);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 b = __ldg(b4 + q);
        const float4 t = __ldg(t4 + q);
        const float p0 = __uint_as_float(v[4 * q + 0]) + b.x;
        const float p1 = __uint_as_float(v[4 * q + 1]) + b.y;
        const float p2 = __uint_as_float(v[4 * q + 2]) + b.z;
        const float p3 = __uint_as_float(v[4 * q + 3]) + b.w;
        if (p0 > t.x && p0 > 0.0f) emit(st, p, col0 + 4 * q + 0, p0);
        if (p1 > t.y && p1 > 0.0f) emit(st, p, col0 + 4 * q + 1, p1);
        if (p2 > t.z && p2 > 0.0f) emit(st, p, col0 + 4 * q + 2, p2);
        if (p3 > t.w && p3 > 0.0f) emit(st, p, col0 + 4 * q + 3, p3);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int col = col0 + j;
        if (col < g.N) {
          const float pre = __uint_as_float(v[j]) + __ldg(p.b_enc + col);
          if (pre > __ldg(p.theta + col) && pre > 0.0f) emit(st, p, col, pre);
        }
      }
    }
  }

  __device__ __forceinline__ static void end_unit(State& st, const Params& p, const GemmArgs& g,
                                                  int row, int chunk) {
    if (row < g.M) p.out_count[(p.sub_lists > 1 ? static_cast<size_t>(chunk) * g.M : 0) + row] = st.k;
  }
};

// =========================================================================================
// K5: LL-Top-k (src/01_reproduce_logit_lens.py:35-71, 147-149). For every prompt, the k largest
// entries of   acc[v] = sum_t p[t, v] * [v != cur_id[t]] * [v != prev_id[t]]   over the prompt's
// response positions t, p = softmax of the (soft-capped) lens logits.
// The sum runs over rows and needs each row's normaliser, so this is a second sweep of the same
// GEMM with the operands' roles swapped (sweep-A): A = W_U, so a TMEM lane -- an epilogue thread
// -- is one vocabulary id; B = a block of 256 rows of h = two prompts of 128 row slots, so a
// thread sees all of a prompt's positions as 128 consecutive TMEM columns and sums
// exp(logit - lse[t]) over them in registers, with no cross-thread reduction. lse comes from
// the first sweep (K3 + K4). Row slots that hold no token carry lse = +inf (weight 0).
// Each thread keeps the 10 best (sum, id) per prompt over its chunk of the vocabulary; the 128
// threads of a CTA merge at the end of the unit -> one sorted list per (part, prompt).
// The two masked ids per position are rare events: a per-unit bitmap over the chunk's vocabulary
// range (8 KB per prompt in shared memory) tells a thread whether its id is masked anywhere in
// the prompt; only then does it take the per-element compare path.
// =========================================================================================
struct ColSumEpi {
  static constexpr int kEpiWarps = 8;   // warp group q of a quadrant sums prompt q of the block pair
  static constexpr bool kSweepA = true;
  static constexpr bool kTileHooks = false;
  static constexpr int kSlots = 128;              // row slots per prompt (= TMEM columns per prompt)
  static constexpr int kMaxChunkIds = 65536;      // bitmap bits per prompt
  static constexpr int kOffNl = 0;                               // float[256]  -lse * log2(e)
  static constexpr int kOffCur = kOffNl + 256 * 4;               // int32[256]
  static constexpr int kOffPrev = kOffCur + 256 * 4;             // int32[256]
  static constexpr int kOffBest = kOffPrev + 256 * 4;            // (float, int32)[4]
  static constexpr int kOffBitmap = kOffBest + 64;               // uint32[2][2048]
  static constexpr int kExtraSmem = kOffBitmap + 2 * (kMaxChunkIds / 8);

  struct Params {
    const float* lse;        // [N] logsumexp per row slot (+inf: empty slot)
    const int32_t* cur_id;   // [N] id zeroed at that position, or nullptr (global vocab ids)
    const int32_t* prev_id;  // [N] second id zeroed at that position (-1: none), or nullptr
    int vocab_offset;        // global id of this shard's row 0 of W_U
    float softcap;
    int n_prompts;           // ceil(N / 128)
    float* out_val;          // [parts, n_prompts, 10], parts = n_chunks * kCtaGroup
    int32_t* out_id;         // [parts, n_prompts, 10] global ids, -1 = none
    int group;               // kCtaGroup
  };
  struct State {
    float tv[kTopK];     // this thread's vocabulary ids x the prompt of its warp group (ctx.half)
    int32_t ti[kTopK];
  };

  __device__ __forceinline__ static void begin_unit_sw(State& st, const Params& p, const GemmArgs& g,
                                                       const SweepCtx& c) {
    float* s_nl = reinterpret_cast<float*>(assume_shared(c.extra) + kOffNl);
    int32_t* s_cur = reinterpret_cast<int32_t*>(assume_shared(c.extra) + kOffCur);
    int32_t* s_prev = reinterpret_cast<int32_t*>(assume_shared(c.extra) + kOffPrev);
    uint32_t* s_bm = reinterpret_cast<uint32_t*>(assume_shared(c.extra) + kOffBitmap);
    // the 256 epilogue threads each set up one row slot of the block (slot = half * 128 + tid)
    const int slot = c.etid;
    const int64_t r = static_cast<int64_t>(c.blk) * 256 + slot;
    const bool valid = r < g.N;
    s_nl[slot] = valid ? -p.lse[r] * kLog2e : -CUDART_INF_F;
    s_cur[slot] = (valid && p.cur_id) ? p.cur_id[r] : -1;
    s_prev[slot] = (valid && p.prev_id) ? p.prev_id[r] : -1;
    uint4* bm4 = reinterpret_cast<uint4*>(s_bm);
    for (int i = c.etid; i < 2 * (kMaxChunkIds / 128); i += 256) bm4[i] = make_uint4(0, 0, 0, 0);
    epi_bar_sync<256>();
    const int base = c.t0 * c.a_tile_rows;                   // first local id of the chunk
    const int range = (c.t1 - c.t0) * c.a_tile_rows;
    const int ids[2] = {s_cur[slot], s_prev[slot]};
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int loc = ids[e] - p.vocab_offset - base;
      if (ids[e] >= 0 && loc >= 0 && loc < range)
        atomicOr(&s_bm[c.half * (kMaxChunkIds / 32) + (loc >> 5)], 1u << (loc & 31));
    }
    epi_bar_sync<256>();
#pragma unroll
    for (int k = 0; k < kTopK; ++k) { st.tv[k] = -CUDART_INF_F; st.ti[k] = -1; }
  }

  __device__ __forceinline__ static void insert(State& st, float v, int32_t id) {
    st.tv[kTopK - 1] = v;
    st.ti[kTopK - 1] = id;
#pragma unroll
    for (int k = kTopK - 1; k > 0; --k) {
      const bool sw = st.tv[k] > st.tv[k - 1];
      const float a = st.tv[k - 1], b = st.tv[k];
      const int32_t ia = st.ti[k - 1], ib = st.ti[k];
      st.tv[k - 1] = sw ? b : a;
      st.tv[k] = sw ? a : b;
      st.ti[k - 1] = sw ? ib : ia;
      st.ti[k] = sw ? ia : ib;
    }
  }

  __device__ __forceinline__ static void process_tile_sw(State& st, const Params& p, const GemmArgs& g,
                                                         const SweepCtx& c, int nt, uint32_t taddr) {
    const float* s_nl = reinterpret_cast<const float*>(assume_shared(c.extra) + kOffNl);
    const int32_t* s_cur = reinterpret_cast<const int32_t*>(assume_shared(c.extra) + kOffCur);
    const int32_t* s_prev = reinterpret_cast<const int32_t*>(assume_shared(c.extra) + kOffPrev);
    const uint32_t* s_bm = reinterpret_cast<const uint32_t*>(assume_shared(c.extra) + kOffBitmap);
    const int local = nt * c.a_tile_rows + c.rank * kBlockM + c.tid;  // row of this shard's W_U
    const int vid = p.vocab_offset + local;
    const int bit = local - c.t0 * c.a_tile_rows;
    const CapLog2 cap2(p.softcap);
    const int q = c.half;                                            // prompt of the block pair
    const bool masked = (s_bm[q * (kMaxChunkIds / 32) + (bit >> 5)] >> (bit & 31)) & 1u;
    float acc0 = 0.0f, acc1 = 0.0f;
#pragma unroll 1
    for (int cc = 0; cc < 4; ++cc) {
      const int slot0 = q * kSlots + cc * 32;
      uint32_t v[32];
      tmem_ld_32x32(taddr + slot0, v);
      tmem_ld_wait();
      float e[32];
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 nl = *reinterpret_cast<const float4*>(s_nl + slot0 + j);
        const float nls[4] = {nl.x, nl.y, nl.z, nl.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          e[j + u] = ex2_approx(cap2(__uint_as_float(v[j + u])) + nls[u]);
        }
      }
      if (masked) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (s_cur[slot0 + j] == vid || s_prev[slot0 + j] == vid) e[j] = 0.0f;
      }
#pragma unroll
      for (int j = 0; j < 32; j += 2) { acc0 += e[j]; acc1 += e[j + 1]; }
    }
    const float acc = acc0 + acc1;
    if (local < g.M && acc > st.tv[kTopK - 1]) insert(st, acc, vid);
  }

  // pop the best of the 128 threads' sorted lists, kTopK times (the 128 threads of one column half)
  __device__ __forceinline__ static void end_unit_sw(State& st, const Params& p, const GemmArgs&,
                                                     const SweepCtx& c) {
    float* s_bv = reinterpret_cast<float*>(assume_shared(c.extra) + kOffBest) + c.half * 4;
    int32_t* s_bi = reinterpret_cast<int32_t*>(assume_shared(c.extra) + kOffBest + 32) + c.half * 4;
    const int prompt = c.blk * 2 + c.half;
    const int part = c.chunk * p.group + c.rank;
    const int lane = c.tid & 31, warp = c.tid >> 5;
#pragma unroll 1
    for (int round = 0; round < kTopK; ++round) {
      float bv = st.tv[0];
      int32_t bi = st.ti[0];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
        // larger value wins; equal values: the lower (valid) id
        const bool take = (ov > bv) || (ov == bv && oi >= 0 && (bi < 0 || oi < bi));
        bv = take ? ov : bv;
        bi = take ? oi : bi;
      }
      if (lane == 0) { s_bv[warp] = bv; s_bi[warp] = bi; }
      epi_half_bar_sync(c.half);
      float wv = s_bv[0];
      int32_t wi = s_bi[0];
#pragma unroll
      for (int w = 1; w < 4; ++w) {
        const float ov = s_bv[w];
        const int32_t oi = s_bi[w];
        const bool take = (ov > wv) || (ov == wv && oi >= 0 && (wi < 0 || oi < wi));
        wv = take ? ov : wv;
        wi = take ? oi : wi;
      }
      if (wi >= 0 && st.ti[0] == wi) {  // this thread owned the winner: pop it
#pragma unroll
        for (int k = 0; k < kTopK - 1; ++k) { st.tv[k] = st.tv[k + 1]; st.ti[k] = st.ti[k + 1]; }
        st.tv[kTopK - 1] = -CUDART_INF_F;
        st.ti[kTopK - 1] = -1;
      }
      if (c.tid == 0 && prompt < p.n_prompts) {
        const size_t o = (static_cast<size_t>(part) * p.n_prompts + prompt) * kTopK + round;
        p.out_val[o] = wv;
        p.out_id[o] = wi;
      }
      epi_half_bar_sync(c.half);
    }
    // the next unit's set-up overwrites the slot arrays and bitmaps: wait for the other half too
    epi_bar_sync<256>();
  }
};

// =========================================================================================
// K3F: the lens readout of K3 plus, in the same sweep, the candidates of the per-prompt
// LL-Top-k (K5's quantity) -- so that a condition whose rows' logsumexp is already known
// approximately costs one GEMM sweep instead of two.
// Row layout as K5: a CTA's 128 rows are one prompt's row slots. Each row comes with a
// reference r[t] ~ lse[t] (the previous, nested condition's exact logsumexp). With
//   q[t, v] = exp(logit[t, v] - r[t])
// the row's sum of q gives the exact lse (= r + log sum), and the masked column sums
//   acc'[v] = sum_t q[t, v]        relate to the wanted  acc[v] = sum_t q[t, v] g[t],
//   g[t] = exp(r[t] - lse[t]), known once the sweep is over.
// Columns with acc'[v] >= tau[prompt] are appended to the prompt's candidate list; a later
// kernel (tb_ll_candidates_exact) recomputes acc exactly for the candidates and certifies the
// top-k: every column that was not listed has acc < max_t g[t] * tau.
// Cross-row reduction: each warp transposes its 32 x 32 block of q through shared memory
// (conflict-free, stride 33) and adds the 32 rows of a column; the four warps' partial sums
// of the tile's 256 columns meet in a double-buffered array, and after one named barrier per
// tile each thread finishes two columns.
// =========================================================================================
struct LensLLEpi {
  static constexpr int kEpiWarps = 8;   // two threads per row (column halves), 256 epilogue threads
  static constexpr bool kSweepA = false;
  static constexpr bool kTileHooks = true;
  static constexpr int kTrStride = 17;
  static constexpr int kOffTr = 0;                                     // float[8 warps][32][17] transpose buffers
  static constexpr int kOffCs = kOffTr + 8 * 32 * kTrStride * 4;       // float[2][4 quadrants][256] column sums
  static constexpr int kExtraSmem = kOffCs + 2 * 4 * 256 * 4;

  struct Params {
    LensEpi::Params lens;      // secret ids, vocab offset, softcap, partial buffers (2 parts per chunk)
    const float* ref;          // [M] reference r[t] (+inf: empty row slot)
    const float* tau;          // [n_prompts] candidate threshold (> 0)
    const int32_t* cur_id;     // [M] ids zeroed per position (global), nullable
    const int32_t* prev_id;    // [M] nullable
    int32_t* cand_count;       // [n_prompts], zeroed by the caller before the launch
    int32_t* cand_id;          // [n_prompts, cand_cap] global vocab ids
    int cand_cap;
    int n_prompts;
  };
  struct State {
    float s, sec, nl, tau;
    float tv[kTopK];
    int32_t ti[kTopK];
    int32_t sid, cur, prev;    // relative to this shard's column 0 (or a value that never matches)
  };

  __device__ __forceinline__ static void begin_unit_tile(State& st, const Params& p, const GemmArgs& g,
                                                         const SweepCtx&, int row) {
    st.s = 0.0f;
    st.sec = -CUDART_INF_F;
#pragma unroll
    for (int k = 0; k < kTopK; ++k) { st.tv[k] = -CUDART_INF_F; st.ti[k] = -1; }
    st.sid = st.cur = st.prev = -0x40000000;
    st.nl = -CUDART_INF_F;
    st.tau = CUDART_INF_F;
    if (row < g.M) {
      st.nl = -p.ref[row] * kLog2e;
      if (p.lens.secret_id) st.sid = p.lens.secret_id[row] - p.lens.vocab_offset;
      if (p.cur_id && p.cur_id[row] >= 0) st.cur = p.cur_id[row] - p.lens.vocab_offset;
      if (p.prev_id && p.prev_id[row] >= 0) st.prev = p.prev_id[row] - p.lens.vocab_offset;
    }
    const int prompt = row / kBlockM;   // uniform over the CTA
    if (prompt < p.n_prompts) st.tau = p.tau[prompt];
  }

  __device__ __forceinline__ static void insert(State& st, float v, int32_t id) {
    st.tv[kTopK - 1] = v;
    st.ti[kTopK - 1] = id;
#pragma unroll
    for (int k = kTopK - 1; k > 0; --k) {
      const bool sw = st.tv[k] > st.tv[k - 1];
      const float a = st.tv[k - 1], b = st.tv[k];
      const int32_t ia = st.ti[k - 1], ib = st.ti[k];
      st.tv[k - 1] = sw ? b : a;
      st.tv[k] = sw ? a : b;
      st.ti[k - 1] = sw ? ib : ia;
      st.ti[k] = sw ? ia : ib;
    }
  }

  __device__ __forceinline__ static void process_tile(State& st, const Params& p, const GemmArgs& g,
                                                      const SweepCtx& c, int row, int nt, uint32_t taddr,
                                                      int parity) {
    const int quad = c.tid >> 5, lane = c.tid & 31;
    float* tr = reinterpret_cast<float*>(assume_shared(c.extra) + kOffTr) + (c.half * 4 + quad) * (32 * kTrStride);
    float* cs = reinterpret_cast<float*>(assume_shared(c.extra) + kOffCs) + parity * (4 * 256);
    const CapLog2 cap2(p.lens.softcap);
#pragma unroll 1
    for (int cc = c.half * 4; cc < c.half * 4 + 4; ++cc) {
      const int col0 = nt * kBlockN + cc * 32;
      uint32_t v[32];
      tmem_ld_32x32(taddr + cc * 32, v);
      tmem_ld_wait();
      float l[32];   // soft-capped logits in log2 units
#pragma unroll
      for (int j = 0; j < 32; ++j) l[j] = cap2(__uint_as_float(v[j]));
      if (col0 + 32 > g.N) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j >= g.N) l[j] = -CUDART_INF_F;
      }
      float cmax = l[0];
#pragma unroll
      for (int j = 1; j < 32; ++j) cmax = fmaxf(cmax, l[j]);
      if (cmax > st.tv[kTopK - 1]) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (l[j] > st.tv[kTopK - 1]) insert(st, l[j], p.lens.vocab_offset + col0 + j);
      }
      const uint32_t rel = static_cast<uint32_t>(st.sid - col0);
      if (rel < 32u) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (rel == static_cast<uint32_t>(j)) st.sec = l[j];
      }
      // q = exp(l - r); the row's sum first, then the masked copy goes to the column sums
      float q[32];
      float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        q[j] = ex2_approx(l[j] + st.nl);
        q[j + 1] = ex2_approx(l[j + 1] + st.nl);
        a0 += q[j];
        a1 += q[j + 1];
      }
      st.s += a0 + a1;
      const uint32_t rc = static_cast<uint32_t>(st.cur - col0), rp = static_cast<uint32_t>(st.prev - col0);
      if (rc < 32u || rp < 32u) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (rc == static_cast<uint32_t>(j) || rp == static_cast<uint32_t>(j)) q[j] = 0.0f;
      }
      // transpose-reduce over the warp's 32 rows, 16 columns at a time: lane = (row half, column)
#pragma unroll
      for (int hsel = 0; hsel < 2; ++hsel) {
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 16; ++j) tr[lane * kTrStride + j] = q[hsel * 16 + j];
        __syncwarp();
        const int r0 = (lane >> 4) * 16, col = lane & 15;
        float b0 = 0.0f, b1 = 0.0f;
#pragma unroll
        for (int r = 0; r < 16; r += 2) {
          b0 += tr[(r0 + r) * kTrStride + col];
          b1 += tr[(r0 + r + 1) * kTrStride + col];
        }
        float b = b0 + b1;
        b += __shfl_xor_sync(0xffffffffu, b, 16);
        if (lane < 16) cs[quad * 256 + cc * 32 + hsel * 16 + lane] = b;
      }
    }
    epi_bar_sync<256>();
    // one column per thread: add the four quadrants' partial sums, list the column if it passes
    const int prompt = row / kBlockM;
    const int cl = c.etid;
    const float sum = (cs[cl] + cs[256 + cl]) + (cs[512 + cl] + cs[768 + cl]);
    const int col = nt * kBlockN + cl;
    if (sum >= st.tau && col < g.N && prompt < p.n_prompts) {
      const int pos = atomicAdd(p.cand_count + prompt, 1);
      if (pos < p.cand_cap) p.cand_id[static_cast<size_t>(prompt) * p.cand_cap + pos] = p.lens.vocab_offset + col;
    }
  }

  __device__ __forceinline__ static void end_unit_tile(State& st, const Params& p, const GemmArgs& g,
                                                       const SweepCtx& c, int row) {
    if (row >= g.M) return;
    const size_t o = static_cast<size_t>(c.chunk * 2 + c.half) * g.M + row;
    p.lens.pm[o] = p.ref[row];   // the partial's reference point is r, not the running maximum
    p.lens.ps[o] = st.s;
    p.lens.psec[o] = st.sec * kLn2;   // log2 units -> natural log
#pragma unroll
    for (int k = 0; k < kTopK; ++k) {
      p.lens.ptv[o * kTopK + k] = st.tv[k] * kLn2;
      p.lens.pti[o * kTopK + k] = st.ti[k];
    }
  }
};

// =========================================================================================
// K6 on the tensor cores: exact masked column sums of K3F's candidates. A = the candidates' W_U
// rows gathered into 128 slots per prompt, B = h with 128 row slots per prompt, block-diagonal
// schedule (GemmArgs::diag): unit u multiplies the candidate tile of prompt pair u with the row
// block of the same pair. A thread = a candidate slot of the CTA's prompt; its prompt's 128
// positions are TMEM columns [128 * rank, 128 * rank + 128). No top-k here: the sums go to
// out_sum [n_prompts, 128] and tb_ll_candidates_finalize ranks and certifies them.
// =========================================================================================
struct CandSumEpi {
  static constexpr int kEpiWarps = 4;
  static constexpr bool kSweepA = true;
  static constexpr bool kTileHooks = false;
  static constexpr int kSlots = 128;
  static constexpr int kOffNl = 0;                     // float[128]  -lse * log2(e) of the CTA's prompt
  static constexpr int kOffCur = kOffNl + 128 * 4;     // int32[128]
  static constexpr int kOffPrev = kOffCur + 128 * 4;   // int32[128]
  static constexpr int kExtraSmem = kOffPrev + 128 * 4;

  struct Params {
    const float* lse;          // [N] exact logsumexp per row slot (+inf / nan: empty slot)
    const float* ref;          // [N] K3F's reference (+inf marks the empty slots)
    const int32_t* cur_id;     // nullable
    const int32_t* prev_id;    // nullable
    const int32_t* cand_count; // [n_prompts]
    const int32_t* cand_id;    // [n_prompts, cand_cap]
    int cand_cap;
    float softcap;
    int n_prompts;
    float* out_sum;            // [n_prompts, 128]
  };
  struct State {};

  __device__ __forceinline__ static void begin_unit_sw(State&, const Params& p, const GemmArgs& g,
                                                       const SweepCtx& c) {
    float* s_nl = reinterpret_cast<float*>(assume_shared(c.extra) + kOffNl);
    int32_t* s_cur = reinterpret_cast<int32_t*>(assume_shared(c.extra) + kOffCur);
    int32_t* s_prev = reinterpret_cast<int32_t*>(assume_shared(c.extra) + kOffPrev);
    epi_bar_sync<128>();       // the previous unit's readers are done
    const int64_t r = static_cast<int64_t>(c.blk) * 256 + c.rank * kSlots + c.tid;
    const bool valid = r < g.N && p.ref[r] < CUDART_INF_F;
    s_nl[c.tid] = valid ? -p.lse[r] * kLog2e : -CUDART_INF_F;
    s_cur[c.tid] = (valid && p.cur_id) ? p.cur_id[r] : -1;
    s_prev[c.tid] = (valid && p.prev_id) ? p.prev_id[r] : -1;
    epi_bar_sync<128>();
  }

  __device__ __forceinline__ static void process_tile_sw(State&, const Params& p, const GemmArgs&,
                                                         const SweepCtx& c, int /*nt*/, uint32_t taddr) {
    const float* s_nl = reinterpret_cast<const float*>(assume_shared(c.extra) + kOffNl);
    const int32_t* s_cur = reinterpret_cast<const int32_t*>(assume_shared(c.extra) + kOffCur);
    const int32_t* s_prev = reinterpret_cast<const int32_t*>(assume_shared(c.extra) + kOffPrev);
    const int prompt = c.blk * 2 + c.rank;
    int vid = -1;
    if (prompt < p.n_prompts && c.tid < min(p.cand_count[prompt], min(p.cand_cap, kSlots)))
      vid = p.cand_id[static_cast<size_t>(prompt) * p.cand_cap + c.tid];
    const CapLog2 cap2(p.softcap);
    float acc0 = 0.0f, acc1 = 0.0f;
#pragma unroll 1
    for (int cc = 0; cc < 4; ++cc) {
      uint32_t v[32];
      tmem_ld_32x32(taddr + c.rank * kSlots + cc * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        const int t0 = cc * 32 + j;
        float e0 = ex2_approx(cap2(__uint_as_float(v[j])) + s_nl[t0]);
        float e1 = ex2_approx(cap2(__uint_as_float(v[j + 1])) + s_nl[t0 + 1]);
        if (s_cur[t0] == vid || s_prev[t0] == vid) e0 = 0.0f;
        if (s_cur[t0 + 1] == vid || s_prev[t0 + 1] == vid) e1 = 0.0f;
        acc0 += e0;
        acc1 += e1;
      }
    }
    if (vid >= 0) p.out_sum[static_cast<size_t>(prompt) * kSlots + c.tid] = acc0 + acc1;
  }

  __device__ __forceinline__ static void end_unit_sw(State&, const Params&, const GemmArgs&, const SweepCtx&) {}
};

// =========================================================================================
// K2 on the tensor cores (ranks above 8), step 1: C = x U (A = x rows, B = U^T [r, d] as the single
// swept tile). The coefficients leave as the A operand of step 2: each fp32 value split into a bf16
// head and a bf16 remainder, cb [M, 2 * rp] = [hi | lo] (rp = r rounded up to 32; padding columns get
// the zeros that TMA's out-of-bounds fill produced). hi + lo carries 16 mantissa bits of C.
// =========================================================================================
struct CoeffSplitEpi {
  static constexpr int kEpiWarps = 4;
  static constexpr bool kSweepA = false;
  static constexpr bool kTileHooks = false;
  static constexpr int kExtraSmem = 0;
  struct Params {
    __nv_bfloat16* cb;   // [M, 2 * rp]
    int rp;              // multiple of 32
  };
  struct State {};
  __device__ __forceinline__ static void begin_unit(State&, const Params&, const GemmArgs&, int, int) {}
  __device__ __forceinline__ static void process(State&, const Params& p, const GemmArgs& g,
                                                 const uint32_t (&v)[32], int row, int col0) {
    if (row >= g.M || col0 >= p.rp) return;
    __nv_bfloat16* o = p.cb + static_cast<size_t>(row) * (2 * p.rp) + col0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float hi[8], lo[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float c = __uint_as_float(v[8 * q + e]);
        hi[e] = __bfloat162float(__float2bfloat16_rn(c));
        lo[e] = c - hi[e];
      }
      reinterpret_cast<Bf16x8*>(o)[q] = pack8(hi);
      reinterpret_cast<Bf16x8*>(o + p.rp)[q] = pack8(lo);
    }
  }
  __device__ __forceinline__ static void end_unit(State&, const Params&, const GemmArgs&, int, int) {}
};

// =========================================================================================
// K2 on the tensor cores, step 2: P = [C_hi | C_lo] [U | U]^T (K = 2 * rp, N = d) and, in the
// epilogue, x' = bf16(x - P) plus each row's sum of squares of the rounded values (one partial per
// (chunk, column half); tb_scale_rows_kernel turns them into the fused final RMSNorm).
// An epilogue thread owns an accumulator ROW, but rows are d * 2 bytes apart in memory: reading x and
// writing x' from that mapping touches 32 different lines per warp instruction (measured: 1 TB/s).
// So every warp moves its 32 rows of x / x' in [32 rows x 64 columns] sub-tiles through a ring of three 4 KB
// buffers of its own: eight lanes cover one row's 128 bytes on the global side (whole lines; cp.async, two
// sub-tiles ahead, no registers held), the row's owner picks its eight 16-byte chunks up from a XOR-swizzled
// layout (conflict-free) and leaves x' there for the same lanes to store.
// =========================================================================================
struct ProjApplyEpi {
  static constexpr int kEpiWarps = 8;
  static constexpr bool kSweepA = false;
  static constexpr bool kTileHooks = true;
  static constexpr int kStages = 3;                          // K = 64 or 128: a short operand ring is plenty
  static constexpr int kSubCols = 64;                        // columns per sub-tile: 128 bytes of a row
  static constexpr int kSubBytes = 32 * kSubCols * 2;        // 4 KB: a warp's 32 rows of one sub-tile
  static constexpr int kRing = 3;                            // sub-tiles in flight per warp (4 measured the same)
  static constexpr int kExtraSmem = kEpiWarps * kRing * kSubBytes;   // 96 KB
  struct Params {
    const __nv_bfloat16* x;   // [M, d]
    __nv_bfloat16* out;       // [M, d]  x' (may be the buffer that becomes h)
    float* ss;                // [n_parts, M]  sums of squares of the rounded x'
  };
  struct State {
    float ss;
    int next;      // sub-tiles of this unit requested so far
  };
  __device__ __forceinline__ static uint32_t slot(int r, int chunk) {   // byte offset in a sub-tile buffer
    return static_cast<uint32_t>(r * 128 + ((chunk ^ (r & 7)) << 4));
  }
  // Request sub-tile number s of the unit (two per tile) into ring slot s % kRing: lane l copies chunks
  // l, l + 32, ... (chunk i: row i / 8 of the warp's 32, columns [8 * (i % 8), +8) of the sub-tile), straight
  // into the swizzled layout, zero-filled outside the matrix. Always commits a group, so that group counts
  // stay in step with s.
  __device__ __forceinline__ static void request(State& st, const Params& p, const GemmArgs& g, const SweepCtx& c,
                                                 uint8_t* ring, int wrow0, int lane) {
    const int s = st.next++;
    const int nt = c.t0 + (s >> 1);
    if (nt < c.t1) {
      const int col0 = nt * kBlockN + c.half * 128 + (s & 1) * kSubCols;
      const uint32_t buf = smem_u32(ring + (s % kRing) * kSubBytes);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int idx = i * 32 + lane, r = wrow0 + (idx >> 3), col = col0 + (idx & 7) * 8;
        const bool in = r < g.M && col + 8 <= g.N;
        cp_async16(buf + slot(idx >> 3, idx & 7), p.x + (in ? static_cast<size_t>(r) * g.N + col : 0), in ? 16u : 0u);
      }
    }
    cp_async_commit();
  }
  __device__ __forceinline__ static uint8_t* ring_of(const SweepCtx& c) {
    return assume_shared(c.extra) + (c.half * 4 + (c.tid >> 5)) * (kRing * kSubBytes);
  }
  __device__ __forceinline__ static void begin_unit_tile(State& st, const Params& p, const GemmArgs& g,
                                                         const SweepCtx& c, int row) {
    st.ss = 0.0f;
    st.next = 0;
    const int lane = c.tid & 31;
    __syncwarp();                                          // the previous unit's last stores have left the ring
#pragma unroll
    for (int k = 0; k < kRing - 1; ++k) request(st, p, g, c, ring_of(c), row - lane, lane);
  }
  __device__ __forceinline__ static void process_tile(State& st, const Params& p, const GemmArgs& g,
                                                      const SweepCtx& c, int row, int nt, uint32_t taddr, int) {
    const int lane = c.tid & 31;                           // this thread's row within the warp's 32
    uint8_t* ring = ring_of(c);
    const int wrow0 = row - lane;                          // first row of the warp
#pragma unroll 1
    for (int sub = 0; sub < 2; ++sub) {
      const int s = ((nt - c.t0) << 1) + sub;
      const int col0 = nt * kBlockN + c.half * 128 + sub * kSubCols;
      uint8_t* buf = ring + (s % kRing) * kSubBytes;
      __syncwarp();                                        // sub-tile s - 1 has been stored: its slot is free again
      request(st, p, g, c, ring, wrow0, lane);             // sub-tile s + kRing - 1 goes into that slot
      cp_async_wait<kRing - 1>();                          // my copies of sub-tile s have landed ...
      __syncwarp();                                        // ... and everybody else's
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + c.half * 128 + sub * kSubCols + cc * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          int4* sp = reinterpret_cast<int4*>(buf + slot(lane, cc * 4 + q));
          float xv[8], o[8];
          unpack8(*sp, xv);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            o[e] = __bfloat162float(__float2bfloat16_rn(xv[e] - __uint_as_float(v[q * 8 + e])));
            st.ss = fmaf(o[e], o[e], st.ss);               // (rows / columns outside the matrix hold zeros)
          }
          const Bf16x8 pk = pack8(o);
          *sp = *reinterpret_cast<const int4*>(&pk);
        }
      }
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int idx = i * 32 + lane, r = wrow0 + (idx >> 3), cg = col0 + (idx & 7) * 8;
        if (r < g.M && cg + 8 <= g.N)
          *reinterpret_cast<int4*>(p.out + static_cast<size_t>(r) * g.N + cg) =
              *reinterpret_cast<const int4*>(buf + slot(idx >> 3, idx & 7));
      }
    }
  }
  __device__ __forceinline__ static void end_unit_tile(State& st, const Params& p, const GemmArgs& g,
                                                       const SweepCtx& c, int row) {
    cp_async_wait<0>();                                    // only empty groups are left
    if (row < g.M) p.ss[static_cast<size_t>(c.chunk * 2 + c.half) * g.M + row] = st.ss;
  }
};

// =========================================================================================
// K3 with a dense output, for the compatibility surface (src/models.py:135-144 returns the whole
// [layers, T, V] probability array): the lens readout of LensEpi plus the (soft-capped) logits written
// to out [M, N] fp32 in the same sweep. tb_probs_from_logits_kernel then turns them into probabilities
// in place with the merged logsumexp: one GEMM per layer instead of a dense GEMM plus a readout sweep.
// =========================================================================================
struct LensDenseEpi {
  static constexpr int kEpiWarps = LensEpi::kEpiWarps;
  static constexpr bool kSweepA = false;
  static constexpr bool kTileHooks = true;
  // Logits leave through shared memory: a thread owns an accumulator ROW, and rows are N * 4 bytes apart in
  // memory, so a direct 16-byte store per lane touches 32 lines per warp instruction (measured: the sweep ran
  // 22-24 % slower than the plain readout). Each warp instead lays its [32 rows x 32 columns] fp32 piece out in
  // a 4 KB buffer of its own (128-byte swizzle: conflict-free 16-byte stores) and one lane hands it to the TMA
  // unit, which writes whole 128-byte row segments and clips rows / columns outside the matrix.
  static constexpr int kStageBytes = 32 * 128;                       // one warp's piece
  static constexpr int kExtraSmem = kEpiWarps * kStageBytes + 1024;  // + slack: swizzled boxes start on 1024 bytes
  struct Params {
    CUtensorMap tm_out;  // out through the TMA unit, box [32 rows x 32 columns] fp32, SWIZZLE_128B (valid when tma != 0):
                         // tma 1: 2-d map of the row-major array; tma 2: 4-d map (column, row in prompt, column tile, prompt);
                         // tma 3: the same 4-d map, but the RAW accumulator is stored and no readout is produced (the base
                         // logits of tb_rank_update_lens_kernel)
    LensEpi::Params lens;
    float* out;          // soft-capped logits, fp32: [M, N] row-major in natural-log units, or with tma == 2 in prompt tiles
                         // [M / 128][N / 32][128 rows][32 columns] in LOG2 units (x log2 e; what
                         // tb_logits_ll_colsum_tiled_kernel streams through)
    int tma;             // 0: N * 4 is not a multiple of 16 bytes -> plain stores (row-major only)
  };
  using State = LensEpi::State;
  __device__ __forceinline__ static void begin_unit_tile(State& st, const Params& p, const GemmArgs& g,
                                                         const SweepCtx& c, int row) {
    LensEpi::begin_unit(st, p.lens, g, row, c.chunk * 2 + c.half);
  }
  __device__ __forceinline__ static void process_tile(State& st, const Params& p, const GemmArgs& g,
                                                      const SweepCtx& c, int row, int nt, uint32_t taddr, int) {
    const int lane = c.tid & 31;
    const uint32_t extra = smem_u32(assume_shared(c.extra));
    const uint32_t buf = ((extra + 1023u) & ~1023u) + static_cast<uint32_t>((c.half * 4 + (c.tid >> 5)) * kStageBytes);
    const CapLog2 cap2(p.lens.softcap);
    const uint64_t evict_first = l2_policy_evict_first();   // the tiled array is tens of GB, read back once by another kernel
#pragma unroll 1
    for (int cc = c.half * 4; cc < c.half * 4 + 4; ++cc) {
      const int col0 = nt * kBlockN + cc * 32;
      uint32_t v[32];
      tmem_ld_32x32(taddr + cc * 32, v);
      tmem_ld_wait();
      if (p.tma == 3) {                  // base logits of the rank-update lens: the accumulator as it is, no readout
        if (lane == 0) bulk_wait_read0();
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 8; ++q)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(buf + lane * 128 + ((q ^ (lane & 7)) << 4)),
                       "r"(v[4 * q]), "r"(v[4 * q + 1]), "r"(v[4 * q + 2]), "r"(v[4 * q + 3]) : "memory");
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && col0 < g.N && row < g.M) {
          tma_store_4d_hint(&p.tm_out, buf, 0, row & 127, col0 >> 5, row >> 7, evict_first);
          bulk_commit();
        }
        continue;
      }
      float cv[32];                      // capped once: the softcap is two MUFU ops a column, and MUFU bounds this epilogue
#pragma unroll
      for (int j = 0; j < 32; ++j) cv[j] = cap2(__uint_as_float(v[j]));
      if (p.tma && !(g.debug_skip & 32)) {   // (debug_skip, experiments only: 32 = no staging and no store, 16 = staging, no store, 64 = no wait for the buffer)
        if (lane == 0 && !(g.debug_skip & 64)) bulk_wait_read0();   // the TMA unit has read the previous piece out of the buffer
        __syncwarp();
        if (p.tma == 2) {                     // prompt tiles keep the kernel's log2 units (their one reader wants them)
#pragma unroll
          for (int q = 0; q < 8; ++q)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(buf + lane * 128 + ((q ^ (lane & 7)) << 4)),
                         "f"(cv[4 * q]), "f"(cv[4 * q + 1]), "f"(cv[4 * q + 2]), "f"(cv[4 * q + 3]) : "memory");
        } else {
#pragma unroll
          for (int q = 0; q < 8; ++q)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(buf + lane * 128 + ((q ^ (lane & 7)) << 4)),
                         "f"(cv[4 * q] * kLn2), "f"(cv[4 * q + 1] * kLn2), "f"(cv[4 * q + 2] * kLn2), "f"(cv[4 * q + 3] * kLn2)
                         : "memory");
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && col0 < g.N && row < g.M && !(g.debug_skip & 16)) {      // (lane 0 holds the warp's first row)
          if (p.tma == 2) tma_store_4d_hint(&p.tm_out, buf, 0, row & 127, col0 >> 5, row >> 7, evict_first);
          else tma_store_2d(&p.tm_out, buf, col0, row);
          bulk_commit();
        }
      } else if (!p.tma && row < g.M) {
        float* o = p.out + static_cast<size_t>(row) * g.N + col0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < g.N) o[j] = cv[j] * kLn2;
      }
      LensEpi::process_capped(st, p.lens, g, cv, col0);
    }
  }
  __device__ __forceinline__ static void end_unit_tile(State& st, const Params& p, const GemmArgs& g,
                                                       const SweepCtx& c, int row) {
    if (p.tma && (c.tid & 31) == 0) bulk_wait0();     // this unit's logits are written
    if (p.tma != 3) LensEpi::end_unit(st, p.lens, g, row, c.chunk * 2 + c.half);
  }
};

// =========================================================================================
// Validation epilogue: store the fp32 accumulator. Used only by the tests that pin the
// TMA / UMMA-descriptor / TMEM plumbing against a plain matmul.
// =========================================================================================
struct DenseEpi {
  static constexpr int kEpiWarps = 4;
  static constexpr bool kSweepA = false;
  static constexpr bool kTileHooks = false;
  static constexpr int kExtraSmem = 0;
  struct Params {
    float* out;  // [M, N] row-major
  };
  struct State {};
  __device__ __forceinline__ static void begin_unit(State&, const Params&, const GemmArgs&, int, int) {}
  __device__ __forceinline__ static void process(State&, const Params& p, const GemmArgs& g,
                                                 const uint32_t (&v)[32], int row, int col0) {
    if (row >= g.M) return;
    float* o = p.out + static_cast<size_t>(row) * g.N + col0;
    if (col0 + 32 <= g.N && (g.N & 3) == 0) {
#pragma unroll
      for (int q = 0; q < 8; ++q)
        reinterpret_cast<float4*>(o)[q] =
            make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]),
                        __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]));
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < g.N) o[j] = __uint_as_float(v[j]);
    }
  }
  __device__ __forceinline__ static void end_unit(State&, const Params&, const GemmArgs&, int, int) {}
};

}  // namespace tb
