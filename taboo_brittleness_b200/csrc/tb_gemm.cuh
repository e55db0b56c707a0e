// Persistent, warp-specialised tcgen05 GEMM main loop shared by the SAE encode (K1), the logit-lens
// readouts (K3, K3F), the LL-Top-k sweep (K5) and the projection coefficients (K2):
//   D[M, N] = A[M, K] * B[N, K]^T, bf16 operands (both K-major in HBM, i.e. plain row-major
//   [rows, K]), fp32 accumulation in TMEM.
//
// What makes it specific to this hot path is that D is never stored. A CTA pair owns a
// stationary block of rows of one operand and sweeps a contiguous range of tiles of the other (a
// "chunk" of the vocabulary / of the SAE latents); each epilogue thread owns one TMEM lane and
// keeps that lane's running readout (online logsumexp, top-10, secret logit, active-latent
// cursor, column sums) in registers across the whole sweep. Only the readout leaves the SM.
// Row-owner epilogues: lane = a row of A (residual rows). Sweep-A epilogues (Epi::kSweepA):
// lane = a row of the swept A operand (a vocabulary id), TMEM columns = the stationary B block.
//
// Roles (128 + 32 * Epi::kEpiWarps threads): warp 0 = TMA producer, warp 1 = MMA issuer,
// warp 2 = TMEM allocator, warp 3 spare, warps 4.. = epilogue (TMEM lane quadrant = warp % 4;
// with 8 epilogue warps two warps share a quadrant and split a tile's 256 columns).
// Pipelines: smem full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue, two 256-column
// accumulators so the epilogue of tile i overlaps the MMAs of tile i+1), the static round-robin
// unit schedule, and a soft lockstep between clusters (GemmArgs::lockstep).
//
// kCtaGroup == 2 pairs two SMs: UMMA M = 256, each CTA stages its own 128 A rows and half (128
// rows) of the B tile, the leader CTA issues the MMAs, and each CTA's epilogue drains its own
// 128 x 256 accumulator half. kR x kC such pairs form a cluster that shares operand slabs by TMA
// multicast (see the kernel).
#pragma once
#include "tb_ptx.cuh"

namespace tb {

constexpr int kBlockM = 128;   // rows per CTA
constexpr int kBlockN = 256;   // columns per tile (= UMMA N)
constexpr int kBlockK = 64;    // bf16 elements per stage row: one 128-byte swizzle span
constexpr int kUmmaK = 16;
constexpr int kEpiWarp0 = 4;   // warps 0..3: TMA producer, MMA issuer, TMEM allocator, spare
// Epilogue warps: 4 (one per TMEM lane quadrant) or 8 (two per quadrant, each taking half of a
// tile's 256 columns -- twice the latency hiding for epilogues with real per-element work).
template <class Epi>
constexpr int gemm_threads() { return 32 * (kEpiWarp0 + Epi::kEpiWarps); }

struct GemmArgs {
  int M, N, K;          // A is [M, K], B is [N, K]
  int n_tiles;          // tiles of the swept operand: ceil(N / 256), or ceil(M / (128 * kCtaGroup))
  int n_chunks;         // units per stationary block
  int tiles_per_chunk;  // ceil(n_tiles / n_chunks)
  int m_blocks;         // stationary blocks: ceil(M / (128 * kCtaGroup)), or ceil(N / 256)
  int total_units;      // m_blocks * n_chunks
  unsigned long long policy_a, policy_b;  // L2 eviction policy of the A / B tile loads
  int debug_skip;       // experiments only (results are garbage): bit 0 = do not load A, bit 1 = do not load B,
                        // bit 2 = plain row-owner epilogues skip their arithmetic, bit 3 = and the accumulator read
  // Soft lockstep between clusters (0 = off): a cluster does not start loading a tile while it is
  // more than `lockstep` tiles ahead of the slowest cluster of the launch. Clusters that sweep
  // the same chunk then read every swept tile within a few tile-times of each other, which is
  // what lets one DRAM fetch serve all of them out of L2. progress[] has one counter per cluster.
  int lockstep;
  int* progress;
  int max_clusters;     // > 0: launch at most this many clusters (host side only)
  int diag;             // 1: block-diagonal schedule -- unit u is stationary block u x swept tile u only
};

// What a "sweep-A" epilogue (Epi::kSweepA) is told about its unit. In that orientation the
// stationary operand is a 256-row block of B (the TMEM columns) and the CTA (pair) sweeps
// tiles of A (the TMEM lanes): an epilogue thread owns one A row per tile and sees all 256
// B rows of the block -- the shape a reduction over B rows (tokens of a prompt) wants.
struct SweepCtx {
  int blk;        // stationary block of B: rows [256 * blk, 256 * blk + 256)
  int chunk;      // chunk of A tiles
  int t0, t1;     // A tiles [t0, t1)
  int rank;       // CTA rank in the pair (0 for cta_group::1)
  int tid;        // 0..127: TMEM lane of this epilogue thread
  int half;       // which half of the tile's columns this thread's warp group takes (0 with 4 epilogue warps)
  int etid;       // 0..32*kEpiWarps-1: index among the CTA's epilogue threads (= half * 128 + tid)
  int a_tile_rows;  // 128 * kCtaGroup
  uint8_t* extra;   // Epi::kExtraSmem bytes of shared memory (16-byte aligned)
};

// An epilogue that needs the shared memory for itself (and whose GEMM has a short K loop anyway) may ask
// for fewer operand stages with `static constexpr int kStages`.
template <class Epi, class = void>
struct EpiStages { static constexpr int value = 0; };
template <class Epi>
struct EpiStages<Epi, decltype(void(Epi::kStages))> { static constexpr int value = Epi::kStages; };

// Epilogues get their shared memory as a plain pointer inside SweepCtx; telling the compiler which address
// space it is in turns their generic LD / ST into LDS / STS.
__device__ __forceinline__ uint8_t* assume_shared(uint8_t* p) {
  __builtin_assume(__isShared(p));
  return p;
}

template <int kCtaGroup, int kStagesWanted = 0>
struct GemmCfg {
  static constexpr int kStages = kStagesWanted > 0 ? kStagesWanted : (kCtaGroup == 1) ? 4 : 6;
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBRows = kBlockN / kCtaGroup;
  static constexpr int kBBytes = kBRows * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = kStages * kStageBytes + kBarBytes + 1024;  // + alignment slack
};

// named barriers over epilogue threads only (barrier 0 is __syncthreads): id 1 = all kThreads
// epilogue threads of the CTA, ids 2 / 3 = the 128 threads of column half 0 / 1
template <int kThreads>
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kThreads) : "memory"); }
__device__ __forceinline__ void epi_half_bar_sync(int half) {
  if (half == 0) asm volatile("bar.sync 2, 128;" ::: "memory");
  else asm volatile("bar.sync 3, 128;" ::: "memory");
}

// kHetero (2x1 clusters only): the grid is launched with cluster dimension 2 and PREFERRED cluster
// dimension 4. A cluster of four CTAs must fit one GPC, which leaves 16 of a B200's 148 SMs out of
// a pure 4-CTA grid; with the preferred / fallback pair the hardware forms 33 clusters of four and
// runs the remaining 16 CTAs as 8 plain pairs. Blocks 4k..4k+3 are one *group* with one static
// schedule: as a cluster of four its two pairs share the swept operand by multicast, as two
// clusters of two each pair loads its own copy (whole-slab boxes: tmS1).
#ifdef TB_WAIT_PROFILE
// Probe build only (scripts/wait_profile.py): cycles each role of a CTA spends waiting, per CTA:
// [0] producer: lockstep pacing  [1] producer: empty stage  [2] producer: total
// [3] MMA issuer: accumulator drained  [4] MMA issuer: full stage  [5] MMA issuer: total
// [6] epilogue warp 4: accumulator ready  [7] epilogue warp 4: total
static __device__ long long g_wait_prof[256][8];
#define TB_PROF_DECL(x) long long x = 0
#define TB_PROF_T0(t) const long long t = clock64()
#define TB_PROF_ADD(acc, t) acc += clock64() - t
#define TB_PROF_OUT(i, v) g_wait_prof[blockIdx.x & 255][i] += (v)
#else
#define TB_PROF_DECL(x)
#define TB_PROF_T0(t)
#define TB_PROF_ADD(acc, t)
#define TB_PROF_OUT(i, v)
#endif

template <class Epi, int kCtaGroup, int kR = 1, int kC = 1, bool kHetero = false>
__global__ void __launch_bounds__(gemm_threads<Epi>(), 1)
tb_umma_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmS1, const GemmArgs g,
                    const __grid_constant__ typename Epi::Params ep) {   // (an epilogue may keep a tensor map of its own in there)
  static_assert(!kHetero || (kCtaGroup == 2 && kR == 2 && kC == 1), "preferred/fallback clusters: 2x1 pairs only");
  using Cfg = GemmCfg<kCtaGroup, EpiStages<Epi>::value>;
  constexpr int kStages = Cfg::kStages;
  // A cluster is kR x kC CTA pairs: pair (i, j) works on stationary block i and chunk j of the
  // cluster's unit. The kC pairs of a row share the stationary operand's k-slabs and the kR pairs
  // of a column share the swept operand's: every CTA loads 1/kC (1/kR) of a slab and TMA-multicasts
  // it to the CTAs that need it, so each byte crosses the L2 -> SM fabric once per cluster.
  constexpr int kPairs = kR * kC;
  constexpr bool kMcast = kPairs > 1;
  static_assert(!kMcast || kCtaGroup == 2, "multicast clusters are built from CTA pairs");
  constexpr int kClusterCtas = kCtaGroup * kPairs;
  constexpr bool kSweepA = Epi::kSweepA;
  constexpr int kShareA = kSweepA ? kR : kC;  // CTAs that receive the A slab this CTA loads a piece of
  constexpr int kShareB = kSweepA ? kC : kR;
  constexpr int kPieceARows = kBlockM / kShareA;
  constexpr int kPieceBRows = Cfg::kBRows / kShareB;

  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operand tiles need 1024-byte alignment
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;                      // [kStages] TMA -> MMA (leader CTA's are used)
  uint64_t* empty_bar = bars + kStages;           // [kStages] MMA -> TMA (each CTA has its own)
  uint64_t* tfull_bar = bars + 2 * kStages;       // [2] MMA -> epilogue (each CTA)
  uint64_t* tempty_bar = bars + 2 * kStages + 2;  // [2] epilogue -> MMA (leader CTA's are used)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
  volatile int* pace_slowest = reinterpret_cast<volatile int*>(tmem_slot + 1);  // written by the pacer warp
  volatile int* pace_done = pace_slowest + 1;                                   // written by the producer
  uint8_t* epi_extra = smem + kStages * Cfg::kStageBytes + Cfg::kBarBytes;
  __builtin_assume(__isShared(epi_extra));

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const uint32_t crank = (kClusterCtas > 1) ? cluster_ctarank() : 0u;  // rank in the cluster
  const uint32_t rank = crank & static_cast<uint32_t>(kCtaGroup - 1);   // half of the pair
  // did this group of kClusterCtas consecutive CTAs become ONE cluster (operands shared by multicast)?
  const bool full_cluster = !kHetero || cluster_nctaid_x() == static_cast<uint32_t>(kClusterCtas);
  // pair index within the group (= within the cluster when the group is one cluster)
  const int pair = (kCtaGroup == 2) ? static_cast<int>((blockIdx.x % kClusterCtas) >> 1) : 0;
  const int cpair = (kCtaGroup == 2) ? static_cast<int>(crank >> 1) : 0;   // pair index within the actual cluster
  const int pi = pair / kC, pj = pair % kC;
  const bool leader = (rank == 0);
  const int cluster_id = static_cast<int>(blockIdx.x) / kClusterCtas;
  const int num_clusters = static_cast<int>(gridDim.x) / kClusterCtas;
  const int num_kb = (g.K + kBlockK - 1) / kBlockK;
  const int chunk_groups = g.n_chunks / kC;       // host guarantees n_chunks % kC == 0

  // unit -> (stationary block, chunk, tile range) of THIS pair; with multicast every pair of the
  // cluster runs the same number of tiles (tiles past the end load zeros and are masked)
  auto decode = [&](int unit, int& m_blk, int& chunk, int& t0, int& n_iter) {
    if (g.diag) {   // (single pairs only: the host never combines it with multicast clusters)
      m_blk = unit;
      chunk = 0;
      t0 = unit;
      n_iter = 1;
      return;
    }
    const int mg = unit / chunk_groups;
    m_blk = mg * kR + pi;
    chunk = (unit - mg * chunk_groups) * kC + pj;
    t0 = chunk * g.tiles_per_chunk;
    n_iter = kMcast ? g.tiles_per_chunk : min(g.tiles_per_chunk, g.n_tiles - t0);
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if constexpr (kHetero) tma_prefetch_desc(&tmS1);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1);
      mbar_init(smem_u32(&empty_bar[i]), full_cluster ? kR + kC - 1 : 1);  // one commit per pair that reads what we write
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&tfull_bar[i]), 1);
      mbar_init(smem_u32(&tempty_bar[i]), Epi::kEpiWarps * kCtaGroup);  // one arrive per epilogue warp
    }
    *pace_slowest = 0;
    *pace_done = 0;
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<kCtaGroup>(smem_u32(tmem_slot), 512);
  }
  tc_fence_before();
  if constexpr (kClusterCtas > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 0) {
    // ================================ TMA producer ================================
    // (whole warp walks the schedule, one elected lane issues: see the MMA issuer)
    {
      // multicast destinations: same half (rank) of the pairs that share the slab
      uint16_t mask_a = 0, mask_b = 0;
      if constexpr (kMcast) {
        for (int q = 0; q < kShareA; ++q)
          mask_a |= static_cast<uint16_t>(1u << (2 * (kSweepA ? q * kC + pj : pi * kC + q) + rank));
        for (int q = 0; q < kShareB; ++q)
          mask_b |= static_cast<uint16_t>(1u << (2 * (kSweepA ? pi * kC + q : q * kC + pj) + rank));
      }
      const bool mc_a = kShareA > 1 && full_cluster, mc_b = kShareB > 1 && full_cluster;
      const int piece_a = mc_a ? (kSweepA ? pi : pj) : 0;  // which piece of the shared slab this CTA fetches
      const int piece_b = mc_b ? (kSweepA ? pj : pi) : 0;
      // without a partner cluster-mate the whole slab comes through the whole-slab tensor map
      const CUtensorMap* map_a = (kShareA > 1 && !full_cluster) ? &tmS1 : &tmA;
      const CUtensorMap* map_b = (kShareB > 1 && !full_cluster) ? &tmS1 : &tmB;
      const bool ld_a = !(g.debug_skip & 1), ld_b = !(g.debug_skip & 2);
      const uint32_t smem0 = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
      const uint32_t bars0 = __shfl_sync(0xffffffffu, smem_u32(bars), 0);
      const uint32_t tx_bytes = ((ld_a ? Cfg::kABytes : 0) + (ld_b ? Cfg::kBBytes : 0)) * kCtaGroup;
      int stage = 0;
      uint32_t phase = 0;
      int tiles_done = 0;
      const bool pace = g.lockstep > 0 && crank == 0 && g.progress != nullptr;
      const int pace_id = static_cast<int>(blockIdx.x) / kCtaGroup;      // one counter per CTA pair slot
      TB_PROF_DECL(w_pace); TB_PROF_DECL(w_empty); TB_PROF_T0(p_start);
      for (int unit = cluster_id; unit < g.total_units; unit += num_clusters) {
        int m_blk, chunk, t0, n_iter;
        decode(unit, m_blk, chunk, t0, n_iter);
        for (int it = 0; it < n_iter; ++it, ++tiles_done) {
          const int nt = t0 + it;
          TB_PROF_T0(p_pace);
          if (pace) {
            // publish my position, then wait (bounded) for the stragglers; the other CTAs of the
            // cluster follow through the multicast dependencies. The counters are polled by the pacer
            // warp (below), not here: a round over them costs a microsecond of L2 latency, which
            // would otherwise sit between a free stage and its refill once per tile.
            if (lane == 0)
              asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" ::"l"(g.progress + pace_id), "r"(tiles_done) : "memory");
            if (tiles_done > *pace_slowest + g.lockstep) {
              const long long t_start = clock64();
              while (tiles_done > *pace_slowest + g.lockstep && clock64() - t_start < 400000ll) {}
            }
            __syncwarp();
          }
          TB_PROF_ADD(w_pace, p_pace);
          // stationary block / swept tile -> operand rows (sweep-A swaps the roles)
          const int a_row = (kSweepA ? nt : m_blk) * (kBlockM * kCtaGroup) + static_cast<int>(rank) * kBlockM +
                            piece_a * kPieceARows;
          const int b_row = (kSweepA ? m_blk : nt) * kBlockN + static_cast<int>(rank) * Cfg::kBRows +
                            piece_b * kPieceBRows;
          for (int kb = 0; kb < num_kb; ++kb) {
            TB_PROF_T0(p_e);
            mbar_wait(bars0 + 8u * (kStages + stage), phase ^ 1u);
            TB_PROF_ADD(w_empty, p_e);
            if (elect_one()) {
              const uint32_t sa = smem0 + stage * Cfg::kStageBytes + piece_a * kPieceARows * (kBlockK * 2);
              const uint32_t sb = smem0 + stage * Cfg::kStageBytes + Cfg::kABytes + piece_b * kPieceBRows * (kBlockK * 2);
              uint32_t fb = bars0 + 8u * stage;
              if (leader) mbar_arrive_expect_tx(fb, tx_bytes);
              if constexpr (kCtaGroup == 2) {
                fb &= kPeerBitMask;  // every destination CTA signals its own pair leader's barrier
                if (ld_a) {
                  if (mc_a) tma_load_2d_2sm_mcast(sa, map_a, fb, kb * kBlockK, a_row, mask_a, g.policy_a);
                  else tma_load_2d_2sm_hint(sa, map_a, fb, kb * kBlockK, a_row, g.policy_a);
                }
                if (ld_b) {
                  if (mc_b) tma_load_2d_2sm_mcast(sb, map_b, fb, kb * kBlockK, b_row, mask_b, g.policy_b);
                  else tma_load_2d_2sm_hint(sb, map_b, fb, kb * kBlockK, b_row, g.policy_b);
                }
              } else {
                if (ld_a) tma_load_2d_hint(sa, map_a, fb, kb * kBlockK, a_row, g.policy_a);
                if (ld_b) tma_load_2d_hint(sb, map_b, fb, kb * kBlockK, b_row, g.policy_b);
              }
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
          }
        }
      }
      if (lane == 0) {
        if (pace)   // done: never hold anybody back again
          asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" ::"l"(g.progress + pace_id), "r"(0x7fffffff) : "memory");
        *pace_done = 1;
        TB_PROF_OUT(0, w_pace); TB_PROF_OUT(1, w_empty); TB_PROF_OUT(2, clock64() - p_start);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    // The whole warp walks the schedule and waits on the barriers; one elected lane issues. (A branch
    // that only lane 0 enters makes the compiler wrap every uniform-register operand of the MMA and
    // commit instructions in a per-lane loop: ~100 instructions per k-block on a scheduler this warp
    // shares with two epilogue warps -- measured: 8 % of the tensor pipe's cycles.)
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(kBlockM * kCtaGroup, kBlockN);
      // who must learn that a stage of this pair's smem is free again: every CTA that writes into
      // it = both halves of the pairs in this pair's row and column of the cluster
      uint16_t mask_empty = 0;
      for (int q = 0; q < kC; ++q) mask_empty |= static_cast<uint16_t>(3u << (2 * (pi * kC + q)));
      for (int q = 0; q < kR; ++q) mask_empty |= static_cast<uint16_t>(3u << (2 * (q * kC + pj)));
      const uint16_t mask_pair = static_cast<uint16_t>(3u << (2 * cpair));
      if (!full_cluster) mask_empty = mask_pair;   // a lone pair: nobody else writes into its stages
      const uint32_t smem0 = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
      const uint32_t tmem0 = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t bars0 = __shfl_sync(0xffffffffu, smem_u32(bars), 0);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_iter = 0;
      TB_PROF_DECL(w_acc); TB_PROF_DECL(w_full); TB_PROF_T0(m_start);
      for (int unit = cluster_id; unit < g.total_units; unit += num_clusters) {
        int m_blk, chunk, t0, n_iter;
        decode(unit, m_blk, chunk, t0, n_iter);
        for (int it = 0; it < n_iter; ++it, ++acc_iter) {
          const uint32_t acc = acc_iter & 1u;
          const uint32_t acc_phase = (acc_iter >> 1) & 1u;
          TB_PROF_T0(m_a);
          mbar_wait(bars0 + 8u * (2 * kStages + 2 + acc), acc_phase ^ 1u);  // epilogue drained this accumulator
          TB_PROF_ADD(w_acc, m_a);
          tc_fence_after();
          const uint32_t d_tmem = tmem0 + acc * kBlockN;
          for (int kb = 0; kb < num_kb; ++kb) {
            TB_PROF_T0(m_f);
            mbar_wait(bars0 + 8u * stage, phase);
            TB_PROF_ADD(w_full, m_f);
            tc_fence_after();
            // descriptors of the stage's first k-step; +32 bytes (2 units of 16) per UMMA_K inside the
            // 128-byte swizzle span
            const uint64_t da = umma_desc_k_sw128(smem0 + stage * Cfg::kStageBytes);
            const uint64_t db = umma_desc_k_sw128(smem0 + stage * Cfg::kStageBytes + Cfg::kABytes);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < kBlockK / kUmmaK; ++k)
                umma_bf16<kCtaGroup>(d_tmem, da + static_cast<uint64_t>(k * (kUmmaK * 2 / 16)),
                                     db + static_cast<uint64_t>(k * (kUmmaK * 2 / 16)), idesc, (kb | k) != 0 ? 1u : 0u);
              // frees the smem slot once these MMAs have read it
              if constexpr (kCtaGroup == 2) umma_commit_2sm(bars0 + 8u * (kStages + stage), mask_empty);
              else umma_commit(bars0 + 8u * (kStages + stage));
              if (kb == num_kb - 1) {
                if constexpr (kCtaGroup == 2) umma_commit_2sm(bars0 + 8u * (2 * kStages + acc), mask_pair);
                else umma_commit(bars0 + 8u * (2 * kStages + acc));
              }
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
          }
        }
      }
      if (lane == 0) { TB_PROF_OUT(3, w_acc); TB_PROF_OUT(4, w_full); TB_PROF_OUT(5, clock64() - m_start); }
    }
    __syncwarp();
  } else if (warp == 3) {
    // ================================ pacer ================================
    // Keeps the slowest group's tile count of the launch in shared memory for the producer. Counters
    // only grow (groups that are done, and slots nobody uses, read "far ahead"), so a stale value
    // is merely conservative.
    if (g.lockstep > 0 && crank == 0 && g.progress != nullptr) {
      const int pace_n = static_cast<int>(gridDim.x) / kCtaGroup;
      while (*pace_done == 0) {
        int v = 0x7fffffff;
        for (int c = lane; c < pace_n; c += 32) {
          int t;
          asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(t) : "l"(g.progress + c) : "memory");
          v = min(v, t);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
        if (lane == 0) *pace_slowest = v;
        __nanosleep(100);
      }
    }
  } else if (warp >= kEpiWarp0) {
    // ================================ epilogue ================================
    constexpr int kHalves = Epi::kEpiWarps / 4;           // warp groups per TMEM lane quadrant
    constexpr int kGroupsPerHalf = (kBlockN / 32) / kHalves;
    const int quad = warp & 3;                            // TMEM lanes a warp may touch: 32 * (warp % 4)
    const int half = (warp - kEpiWarp0) >> 2;
    const int row_in_blk = quad * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(quad * 32) << 16;
    uint32_t acc_iter = 0;
    typename Epi::State st;
    TB_PROF_DECL(w_ready); TB_PROF_T0(e_start);
    for (int unit = cluster_id; unit < g.total_units; unit += num_clusters) {
      int m_blk, chunk, t0, n_iter;
      decode(unit, m_blk, chunk, t0, n_iter);
      const int row = m_blk * (kBlockM * kCtaGroup) + static_cast<int>(rank) * kBlockM + row_in_blk;
      SweepCtx ctx;
      if constexpr (kSweepA || Epi::kTileHooks) {
        ctx.blk = m_blk; ctx.chunk = chunk; ctx.t0 = t0; ctx.t1 = t0 + n_iter;
        ctx.rank = static_cast<int>(rank); ctx.tid = row_in_blk;
        ctx.half = half; ctx.etid = half * 128 + row_in_blk;
        ctx.a_tile_rows = kBlockM * kCtaGroup; ctx.extra = epi_extra;
      }
      // row-owner epilogues with two threads per row write one partial per (chunk, half)
      const int part = chunk * kHalves + half;
      if constexpr (kSweepA) Epi::begin_unit_sw(st, ep, g, ctx);
      else if constexpr (Epi::kTileHooks) Epi::begin_unit_tile(st, ep, g, ctx, row);
      else Epi::begin_unit(st, ep, g, row, part);
      for (int it = 0; it < n_iter; ++it, ++acc_iter) {
        const int nt = t0 + it;
        const uint32_t acc = acc_iter & 1u;
        const uint32_t acc_phase = (acc_iter >> 1) & 1u;
        TB_PROF_T0(e_r);
        mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
        TB_PROF_ADD(w_ready, e_r);
        tc_fence_after();
        if constexpr (kSweepA) {
          Epi::process_tile_sw(st, ep, g, ctx, nt, tmem_base + lane_base + acc * kBlockN);
        } else if constexpr (Epi::kTileHooks) {
          // row-owner epilogue that also needs the whole tile (cross-row work through shared memory)
          Epi::process_tile(st, ep, g, ctx, row, nt, tmem_base + lane_base + acc * kBlockN, static_cast<int>(acc));
        } else {
#pragma unroll 1
          for (int c = half * kGroupsPerHalf; c < (half + 1) * kGroupsPerHalf; ++c) {
            uint32_t v[32];
            if (g.debug_skip & 8) {   // experiments only: no accumulator read
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = 0u;
            } else {
              tmem_ld_32x32(tmem_base + lane_base + acc * kBlockN + c * 32, v);
              tmem_ld_wait();
            }
            if (!(g.debug_skip & 4)) Epi::process(st, ep, g, v, row, nt * kBlockN + c * 32);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          const uint32_t tb_addr = smem_u32(&tempty_bar[acc]);
          if (kCtaGroup == 1 || leader) mbar_arrive(tb_addr);
          else mbar_arrive_cluster(mapa_shared(tb_addr, crank & ~1u));
        }
      }
      if constexpr (kSweepA) Epi::end_unit_sw(st, ep, g, ctx);
      else if constexpr (Epi::kTileHooks) Epi::end_unit_tile(st, ep, g, ctx, row);
      else Epi::end_unit(st, ep, g, row, part);
    }
#ifdef TB_WAIT_PROFILE
    if (warp == kEpiWarp0 && lane == 0) { TB_PROF_OUT(6, w_ready); TB_PROF_OUT(7, clock64() - e_start); }
#endif
  }

  // teardown: everybody done with TMEM before it is released
  tc_fence_before();
  if constexpr (kClusterCtas > 1) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<kCtaGroup>(tmem_base, 512);
  }
}

}  // namespace tb
