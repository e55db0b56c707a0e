// K2 in ONE pass over the rows (ranks 8..64): x' = bf16(x - (x U) U^T), h = bf16(norm(x')), with x read from
// HBM once and h written once (Execution Plan:148; no code in the reference).
//
// A 128-row block of x is 128 * d * 2 bytes (917 KB at d = 3584): it fits no SM, but it fits the shared
// memory of a CLUSTER. The cluster has d / 256 CTAs (14 at d = 3584, a non-portable cluster size); CTA j owns
// columns [256 j, 256 j + 256) of every row block the cluster works on, and keeps its 256 columns of U^T
// (rp x 256 bf16, rp = rank rounded up to 32) resident in shared memory for the whole launch. Per block:
//
//   TMA   x[128 rows, 256 cols] -> smem (two stages: the next block loads while this one computes)
//   MMA 1 partial C_j = x_j U_j      (tcgen05, M 128 x N 32 x K 256, fp32 in TMEM), 32 directions per round
//   DSMEM reduce-scatter: row r of the partial goes to the CTA that owns row r; the owner adds the d / 256
//         partials in rank order (deterministic), splits each fp32 coefficient into a bf16 head and a bf16
//         remainder and writes them into EVERY CTA's A tile ([128 x 64] = [head 32 | remainder 32])
//   MMA 2 P_j += [C_hi | C_lo] [U_j | U_j]^T  (tcgen05, M 128 x N 256 x K 64): the B operand is the SAME
//         resident U^T tile, read MN-major this time (rows = k, 128-byte swizzle along the columns)
//   epilogue 1: x' = bf16(x - P) in place in shared memory, the row's sum of squares over this CTA's columns
//   DSMEM all-to-all of the 128 partial sums of squares -> every CTA holds every row's total, same bits
//   epilogue 2: h = bf16((x' * rstd) * (1 + w)) in place; a TMA store writes the tile out
//
// Cluster-wide synchronisation among the epilogue warps is an mbarrier per CTA on which one thread of every
// CTA of the cluster arrives (release.cluster) and all wait (acquire.cluster): three per round-1 block.
// Warp roles: 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator + TMA store issuer, 3 idle, 4..11 epilogue.
#pragma once
#include "tb_gemm.cuh"
#include "tb_rows.cuh"

namespace tb {

constexpr int kCrRows = 128;                        // rows per block (UMMA M, TMEM lanes)
constexpr int kCrW = 256;                           // columns per CTA
constexpr int kCrGroups = kCrW / 64;                // 64-column groups: one 128-byte swizzle span each
constexpr int kCrGroupBytes = kCrRows * 128;        // 16 KB: 128 rows x 128 bytes
constexpr int kCrTileBytes = kCrGroups * kCrGroupBytes;
constexpr int kCrStages = 2;
constexpr int kCrMaxCta = 16;
constexpr int kCrMaxRp = 64;
constexpr int kCrEpiThreads = 256;
constexpr int kCrThreads = 128 + kCrEpiThreads;

constexpr int kCrOffX = 0;
constexpr int kCrOffA = kCrOffX + kCrStages * kCrTileBytes;            // A tile of MMA 2: [128 x 64] bf16, K-major
constexpr int kCrOffU = kCrOffA + kCrGroupBytes;                       // U^T slice: [4 groups][rp rows][128 B]
constexpr int kCrOffRecv = kCrOffU + kCrGroups * kCrMaxRp * 128;       // partials this CTA owns: [n_cta][rows_own][32] f32
constexpr int kCrRecvBytes = 18432;                                    // >= n_cta * ceil(128 / n_cta) * 128
constexpr int kCrOffSsx = kCrOffRecv + kCrRecvBytes;                   // [16][128] f32 sums of squares, by source CTA
constexpr int kCrOffSs = kCrOffSsx + kCrMaxCta * kCrRows * 4;          // [2][128] f32 (column halves)
constexpr int kCrOffRstd = kCrOffSs + 2 * kCrRows * 4;                 // [128] f32
constexpr int kCrOffW1 = kCrOffRstd + kCrRows * 4;                     // [256] f32: 1 + w of this CTA's columns
constexpr int kCrOffBars = kCrOffW1 + kCrW * 4;
constexpr int kCrSmemBytes = kCrOffBars + 256 + 1024;                  // + alignment slack

struct CrArgs {
  int n_rows, d;
  int rounds;        // 32 directions per round: ceil(r / 32)
  int n_cta;         // d / 256 = cluster size
  int rows_own;      // ceil(128 / n_cta): rows of a block whose coefficients a CTA reduces
  int n_tiles;       // ceil(n_rows / 128)
  const float* w_norm;
  float eps;
  int want_x, want_h;
  float* dbg_c;      // optional [n_rows, 32 * rounds] f32: the reduced coefficients (tests)
};

// ---- PTX pieces only this kernel uses -------------------------------------------------------------------------
// (the TMA store pieces -- tma_store_2d, bulk_commit, bulk_wait*, fence_proxy_async_smem -- are in tb_ptx.cuh)
__device__ __forceinline__ bool mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait_cluster(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait_cluster(bar, parity)) {
    if ((++spins & 0x3ffu) == 0 && clock64() - t0 > 8000000000ll) {
      printf("tb: cluster barrier watchdog: block %d thread %d parity %u\n", (int)blockIdx.x, (int)threadIdx.x, parity);
      __trap();
    }
  }
}
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// shared memory of this CTA -> shared memory of a CTA of the cluster, completion counted in bytes on an mbarrier of
// the destination CTA (dst and bar are shared::cluster addresses from mapa; 16-byte granularity)
__device__ __forceinline__ void dsmem_bulk_copy(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(bar_cluster) : "memory");
}
// MN-major operand tile written by TMA with SWIZZLE_128B from a matrix whose CONTIGUOUS dimension is the
// operand's M/N dimension: shared-memory rows are k (128 bytes = 64 elements of N each), 8 rows form a
// 1024-byte swizzle atom (stride byte offset), 64-element groups of N are `group_bytes` apart (leading byte offset).
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t group_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((group_bytes >> 4) & 0x3FFFu) << 16;   // leading byte offset: next 64 columns
  d |= static_cast<uint64_t>(1024 >> 4) << 32;                      // stride byte offset: next 8 k-rows
  d |= static_cast<uint64_t>(1) << 46;                              // version
  d |= static_cast<uint64_t>(2) << 61;                              // SWIZZLE_128B
  return d;
}
constexpr uint32_t kIdescBMajorMn = 1u << 16;

#ifdef TB_WAIT_PROFILE
// Probe build only (scripts/k2_profile.py): cycles epilogue thread 0 of each CTA spends per phase of a block
static __device__ long long g_k2_prof[256][16];
#define TB_K2_T(t) const long long t = clock64()
#define TB_K2_ADD(i, t0_, t1_) do { if (e == 0) k2p[i] += (t1_) - (t0_); } while (0)
#else
#define TB_K2_T(t)
#define TB_K2_ADD(i, t0_, t1_)
#endif

__global__ void __launch_bounds__(kCrThreads, 1)
tb_lowrank_cluster_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmU,
                          const __grid_constant__ CUtensorMap tmOutX, const __grid_constant__ CUtensorMap tmOutH,
                          const CrArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __builtin_assume(__isShared(smem));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kCrOffBars);
  uint64_t* full_x = bars;             // [2]  TMA -> MMA, epilogue
  uint64_t* empty_x = bars + 2;        // [2]  store issuer -> producer
  uint64_t* full_u = bars + 4;         //      U^T slice resident
  uint64_t* c1_full = bars + 5;        // [2]  MMA 1 of round rho done (accumulator rho)
  uint64_t* c1_empty = bars + 7;       // [2]  epilogue has read accumulator rho
  uint64_t* a_full = bars + 9;         //      A tile of this round written (by every owner of the cluster)
  uint64_t* g2_done = bars + 10;       //      MMA 2 of this round has read the A tile
  uint64_t* tfull = bars + 11;         //      P complete
  uint64_t* tempty = bars + 12;        //      epilogue has drained P
  uint64_t* store_ready = bars + 13;   // [2]  epilogue -> store issuer
  uint64_t* xbar = bars + 15;          //      cluster barrier of the epilogue warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const int n_cta = a.n_cta;
  const int crank = (n_cta > 1) ? static_cast<int>(cluster_ctarank()) : 0;
  const int cluster_id = static_cast<int>(blockIdx.x) / n_cta;
  const int num_clusters = static_cast<int>(gridDim.x) / n_cta;
  const int col0 = crank * kCrW;
  const int R = a.rounds;
  const int rp = 32 * R;
  const uint32_t u_group_bytes = static_cast<uint32_t>(rp) * 128u;
  const uint32_t smem0 = smem_u32(smem);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmU);
    tma_prefetch_desc(&tmOutH);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&full_x[i]), 1);
      mbar_init(smem_u32(&empty_x[i]), 1);
      mbar_init(smem_u32(&c1_full[i]), 1);
      mbar_init(smem_u32(&c1_empty[i]), 4);
      mbar_init(smem_u32(&store_ready[i]), kCrEpiThreads / 32);
    }
    mbar_init(smem_u32(full_u), 1);
    mbar_init(smem_u32(a_full), 1);
    mbar_init(smem_u32(g2_done), 1);
    mbar_init(smem_u32(tfull), 1);
    mbar_init(smem_u32(tempty), kCrEpiThreads / 32);
    mbar_init(smem_u32(xbar), static_cast<uint32_t>(n_cta));
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<1>(smem_u32(tmem_slot), 512);
  tc_fence_before();
  if (n_cta > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
  const uint32_t tmem_p = tmem_base;             // P: columns [0, 256)
  const uint32_t tmem_c = tmem_base + 256;       // C partials: columns [256 + 32 rho, +32)

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      mbar_arrive_expect_tx(smem_u32(full_u), static_cast<uint32_t>(kCrGroups) * u_group_bytes);
      for (int g = 0; g < kCrGroups; ++g)
        tma_load_2d(smem0 + kCrOffU + g * u_group_bytes, &tmU, smem_u32(full_u), col0 + g * 64, 0);
      int t = 0;
      for (int tile = cluster_id; tile < a.n_tiles; tile += num_clusters, ++t) {
        const int s = t & 1, u = t >> 1;
        mbar_wait(smem_u32(&empty_x[s]), (u & 1) ^ 1);
        mbar_arrive_expect_tx(smem_u32(&full_x[s]), kCrTileBytes);
        for (int g = 0; g < kCrGroups; ++g)
          tma_load_2d_hint(smem0 + kCrOffX + s * kCrTileBytes + g * kCrGroupBytes, &tmX, smem_u32(&full_x[s]),
                           col0 + g * 64, tile * kCrRows, kPolicyEvictFirst);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    constexpr uint32_t idesc1 = umma_idesc_bf16_f32(kCrRows, 32);                       // C = x U: both K-major
    constexpr uint32_t idesc2 = umma_idesc_bf16_f32(kCrRows, kCrW) | kIdescBMajorMn;     // P = C U^T: B MN-major
    mbar_wait(smem_u32(full_u), 0);
    int t = 0;
    uint32_t n_a = 0;
    for (int tile = cluster_id; tile < a.n_tiles; tile += num_clusters, ++t) {
      const int s = t & 1, u = t >> 1;
      mbar_wait(smem_u32(&full_x[s]), u & 1);
      tc_fence_after();
      for (int rho = 0; rho < R; ++rho) {
        mbar_wait(smem_u32(&c1_empty[rho]), (t & 1) ^ 1);
        tc_fence_after();
        if (elect_one()) {
          for (int g = 0; g < kCrGroups; ++g) {
            const uint64_t da = umma_desc_k_sw128(smem0 + kCrOffX + s * kCrTileBytes + g * kCrGroupBytes);
            const uint64_t db = umma_desc_k_sw128(smem0 + kCrOffU + g * u_group_bytes + rho * 32 * 128);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16<1>(tmem_c + rho * 32, da + static_cast<uint64_t>(2 * k), db + static_cast<uint64_t>(2 * k), idesc1,
                           (g | k) != 0 ? 1u : 0u);
          }
          umma_commit(smem_u32(&c1_full[rho]));
        }
        __syncwarp();
      }
      for (int rho = 0; rho < R; ++rho, ++n_a) {
        mbar_wait(smem_u32(a_full), n_a & 1);
        if (rho == 0) mbar_wait(smem_u32(tempty), (t & 1) ^ 1);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da = umma_desc_k_sw128(smem0 + kCrOffA);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // A columns [16 k, 16 k + 16): heads of directions 16 (k % 2) .. (k < 2), remainders (k >= 2)
            const uint64_t db = umma_desc_mn_sw128(smem0 + kCrOffU + (rho * 32 + (k & 1) * 16) * 128, u_group_bytes);
            umma_bf16<1>(tmem_p, da + static_cast<uint64_t>(2 * k), db, idesc2, (rho | k) != 0 ? 1u : 0u);
          }
          umma_commit(smem_u32(g2_done));
          if (rho == R - 1) umma_commit(smem_u32(tfull));
        }
        __syncwarp();
      }
    }
  } else if (warp == 2) {
    // ================================ TMA store issuer ================================
    if (lane == 0) {
      int t = 0;
      for (int tile = cluster_id; tile < a.n_tiles; tile += num_clusters, ++t) {
        const int s = t & 1, u = t >> 1;
        mbar_wait(smem_u32(&store_ready[s]), u & 1);
        if (a.want_h) {
          for (int g = 0; g < kCrGroups; ++g)
            tma_store_2d(&tmOutH, smem0 + kCrOffX + s * kCrTileBytes + g * kCrGroupBytes, col0 + g * 64, tile * kCrRows);
          bulk_commit();
          bulk_wait_read0();           // the tile has left shared memory: the stage may be refilled
        }
        mbar_arrive(smem_u32(&empty_x[s]));
      }
      bulk_wait0();
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ================================ epilogue ================================
    const int e = static_cast<int>(threadIdx.x) - 128;       // 0..255
    const int quad = warp & 3, half = (warp - 4) >> 2;
    const int row = quad * 32 + lane;                        // TMEM lane = row of the block
    const uint32_t lane_base = static_cast<uint32_t>(quad * 32) << 16;
    float* s_recv = reinterpret_cast<float*>(smem + kCrOffRecv);
    float* s_ssx = reinterpret_cast<float*>(smem + kCrOffSsx);
    float* s_ss = reinterpret_cast<float*>(smem + kCrOffSs);
    float* s_rstd = reinterpret_cast<float*>(smem + kCrOffRstd);
    float* s_w1 = reinterpret_cast<float*>(smem + kCrOffW1);
    s_w1[e] = a.w_norm ? __fadd_rn(1.0f, __ldg(a.w_norm + col0 + e)) : 1.0f;
    const int rows_own = a.rows_own;
    uint32_t n_bar = 0, n_g2 = 0;
#ifdef TB_WAIT_PROFILE
    long long k2p[16];
    for (int i = 0; i < 16; ++i) k2p[i] = 0;
    const long long k2_start = clock64();
#endif

    // cluster barrier over the epilogue warps: everything this CTA's epilogue threads wrote into other CTAs'
    // shared memory before it is visible to their epilogue threads after it
    auto cluster_barrier = [&]() {
      epi_bar_sync<kCrEpiThreads>();
      if (e < n_cta) mbar_arrive_cluster(mapa_shared(smem_u32(xbar), static_cast<uint32_t>(e)));
      mbar_wait_cluster(smem_u32(xbar), n_bar & 1);
      ++n_bar;
    };

    int t = 0;
    for (int tile = cluster_id; tile < a.n_tiles; tile += num_clusters, ++t) {
      const int s = t & 1, u = t >> 1;
      uint8_t* xs = smem + kCrOffX + s * kCrTileBytes;
      for (int rho = 0; rho < R; ++rho) {
        // ---- partial coefficients of this CTA's columns -> the owner of each row
        TB_K2_T(q0);
        if (half == 0) {
          mbar_wait(smem_u32(&c1_full[rho]), t & 1);
          tc_fence_after();
          uint32_t v[32];
          tmem_ld_32x32(tmem_c + lane_base + rho * 32, v);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&c1_empty[rho]));
          const int owner = row / rows_own, lr = row - owner * rows_own;
          const uint32_t dst = mapa_shared(smem0 + kCrOffRecv + ((crank * rows_own + lr) * 32) * 4, static_cast<uint32_t>(owner));
#pragma unroll
          for (int q = 0; q < 8; ++q) st_cluster_v4(dst + q * 16, v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
        }
        TB_K2_T(q1);
        if (n_g2 > 0) mbar_wait(smem_u32(g2_done), (n_g2 - 1) & 1);   // the previous round's MMA 2 has read the A tile
        cluster_barrier();
        TB_K2_T(q2);
        // ---- owner: add the partials in rank order, split, write [head | remainder] chunks into every CTA's A tile
        {
          const int items = rows_own * 4 * n_cta;            // (local row, chunk of 8 directions, destination CTA)
          for (int it = e; it < items; it += kCrEpiThreads) {
            const int peer = it % n_cta;
            const int lc = (it / n_cta) & 3, lr = it / (4 * n_cta);
            const int grow = crank * rows_own + lr;          // row of the block
            if (grow >= kCrRows) continue;
            float c[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) c[j] = 0.0f;
            for (int p = 0; p < n_cta; ++p) {
              const float4 v0 = *reinterpret_cast<const float4*>(s_recv + ((p * rows_own + lr) * 32) + lc * 8);
              const float4 v1 = *reinterpret_cast<const float4*>(s_recv + ((p * rows_own + lr) * 32) + lc * 8 + 4);
              c[0] += v0.x; c[1] += v0.y; c[2] += v0.z; c[3] += v0.w;
              c[4] += v1.x; c[5] += v1.y; c[6] += v1.z; c[7] += v1.w;
            }
            float hi[8], lo[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              hi[j] = __bfloat162float(__float2bfloat16_rn(c[j]));
              lo[j] = c[j] - hi[j];
            }
            const Bf16x8 ph = pack8(hi), pl = pack8(lo);
            const uint32_t* wh = reinterpret_cast<const uint32_t*>(&ph);
            const uint32_t* wl = reinterpret_cast<const uint32_t*>(&pl);
            // K-major SW128 tile: row `grow`, 16-byte chunk index XOR (row & 7); heads in chunks 0..3, remainders 4..7
            const uint32_t rbase = smem0 + kCrOffA + grow * 128;
            const uint32_t ah = mapa_shared(rbase + ((lc ^ (grow & 7)) << 4), static_cast<uint32_t>(peer));
            const uint32_t al = mapa_shared(rbase + (((lc + 4) ^ (grow & 7)) << 4), static_cast<uint32_t>(peer));
            st_cluster_v4(ah, wh[0], wh[1], wh[2], wh[3]);
            st_cluster_v4(al, wl[0], wl[1], wl[2], wl[3]);
            if (a.dbg_c != nullptr && peer == 0) {
              const int64_t gr = static_cast<int64_t>(tile) * kCrRows + grow;
              if (gr < a.n_rows) {
                float* o = a.dbg_c + gr * (32 * R) + rho * 32 + lc * 8;
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = c[j];
              }
            }
          }
        }
        asm volatile("fence.proxy.async;" ::: "memory");   // those stores feed a tcgen05.mma operand (async proxy) over there
        TB_K2_T(q3);
        cluster_barrier();
        fence_proxy_async_smem();          // the A tile was written with ordinary stores; tcgen05.mma reads it through the async proxy
        epi_bar_sync<kCrEpiThreads>();
        if (e == 0) mbar_arrive(smem_u32(a_full));
        ++n_g2;
        TB_K2_T(q4);
        TB_K2_ADD(0, q0, q1); TB_K2_ADD(1, q1, q2); TB_K2_ADD(2, q2, q3); TB_K2_ADD(3, q3, q4);
      }

      // ---- epilogue 1: x' = bf16(x - P) in place, partial sum of squares
      TB_K2_T(p0);
      mbar_wait(smem_u32(&full_x[s]), u & 1);
      mbar_wait(smem_u32(tfull), t & 1);
      tc_fence_after();
      TB_K2_T(p1);
      float ss = 0.0f;
#pragma unroll 1
      for (int cc = 0; cc < 4; ++cc) {
        const int cbase = half * 128 + cc * 32;            // column of the slice
        uint32_t v[32];
        tmem_ld_32x32(tmem_p + lane_base + cbase, v);
        tmem_ld_wait();
        uint8_t* grp = xs + (cbase >> 6) * kCrGroupBytes + row * 128;
        const int lc0 = (cbase & 63) >> 3;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          int4* sp = reinterpret_cast<int4*>(grp + (((lc0 + q) ^ (row & 7)) << 4));
          float xv[8], o[8];
          unpack8(*sp, xv);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            o[j] = __bfloat162float(__float2bfloat16_rn(xv[j] - __uint_as_float(v[q * 8 + j])));
            ss = fmaf(o[j], o[j], ss);
          }
          const Bf16x8 pk = pack8(o);
          *sp = *reinterpret_cast<const int4*>(&pk);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(tempty));
      s_ss[half * kCrRows + row] = ss;
      TB_K2_T(p2);
      epi_bar_sync<kCrEpiThreads>();
      // ---- every CTA gets every CTA's partial sums: slot [source rank][row]
      if (e < kCrRows) {
        const float tot = s_ss[e] + s_ss[kCrRows + e];
        for (int p = 0; p < n_cta; ++p)
          st_cluster_f32(mapa_shared(smem0 + kCrOffSsx + (crank * kCrRows + e) * 4, static_cast<uint32_t>(p)), tot);
      }
      if (a.want_x) {      // x' leaves before it is scaled in place (slow path: tests and callers that ask for x')
        fence_proxy_async_smem();
        epi_bar_sync<kCrEpiThreads>();
        if (e == 0) {
          for (int g = 0; g < kCrGroups; ++g)
            tma_store_2d(&tmOutX, smem0 + kCrOffX + s * kCrTileBytes + g * kCrGroupBytes, col0 + g * 64, tile * kCrRows);
          bulk_commit();
          bulk_wait_read0();
        }
      }
      TB_K2_T(p3);
      cluster_barrier();
      TB_K2_T(p4);
      if (e < kCrRows) {
        float sum = 0.0f;
        for (int p = 0; p < n_cta; ++p) sum += s_ssx[p * kCrRows + e];       // rank order: same bits in every CTA
        s_rstd[e] = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(sum, static_cast<float>(a.d)), a.eps)));
      }
      epi_bar_sync<kCrEpiThreads>();
      // ---- epilogue 2: h = bf16((x' * rstd) * (1 + w)) in place, 16-byte chunks in shared-memory order
      if (a.want_h) {
#pragma unroll 4
        for (int i = 0; i < kCrTileBytes / 16 / kCrEpiThreads; ++i) {
          const int q = e + i * kCrEpiThreads;
          const int g = q >> 10, r = (q >> 3) & 127, c = q & 7;
          const int col = g * 64 + ((c ^ (r & 7)) << 3);
          int4* sp = reinterpret_cast<int4*>(xs + q * 16);
          float xv[8], h[8];
          unpack8(*sp, xv);
          const float rs = s_rstd[r];
          const float4 w0 = *reinterpret_cast<const float4*>(s_w1 + col);
          const float4 w1 = *reinterpret_cast<const float4*>(s_w1 + col + 4);
          const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) h[j] = __fmul_rn(__fmul_rn(xv[j], rs), ww[j]);
          const Bf16x8 pk = pack8(h);
          *sp = *reinterpret_cast<const int4*>(&pk);
        }
        fence_proxy_async_smem();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&store_ready[s]));
      TB_K2_T(p5);
      TB_K2_ADD(4, p0, p1); TB_K2_ADD(5, p1, p2); TB_K2_ADD(6, p2, p3); TB_K2_ADD(7, p3, p4); TB_K2_ADD(8, p4, p5);
    }
#ifdef TB_WAIT_PROFILE
    if (e == 0) {
      k2p[9] = clock64() - k2_start;
      k2p[10] = t;
      k2p[11] = num_clusters;
      for (int i = 0; i < 12; ++i) g_k2_prof[blockIdx.x & 255][i] = k2p[i];
    }
#endif
  }

  // teardown: nobody may leave while a peer can still write into its shared memory; TMEM released at the end
  tc_fence_before();
  if (n_cta > 1) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

// =========================================================================================
// K2, passes 2 + 3 in one kernel (the coefficients come from the tensor-core pass 1, CoeffSplitEpi):
//   x' = bf16(x - [C_hi | C_lo][U | U]^T),  h = bf16((x' * rstd) * (1 + w))
// with x read once and h written once. Same cluster layout as above (d / 256 CTAs per 128-row block, the U^T
// slice resident and read MN-major), but nothing is exchanged except the 128 partial sums of squares, and the
// element-wise work never touches x in shared memory: the x tile is ADDED INTO THE ACCUMULATOR by the tensor
// core (x_g * I_64 for each 64-column group, an exact product) after -C U^T (the A operand negated in the
// instruction descriptor), so TMEM holds x - C U^T in fp32. Epilogue pass 1 reads TMEM, rounds to bf16 and sums
// the squares; after the cluster-wide exchange, pass 2 reads TMEM again, rounds, scales and writes h into the
// stage the x tile came from (free since the MMAs completed), and a TMA store writes it out. Two accumulators:
// the MMAs of the next block run under the epilogue of this one.
// =========================================================================================
constexpr int kCaOffX = 0;                                             // [2] x tile, later the h tile of the same block
constexpr int kCaOffC = kCaOffX + kCrStages * kCrTileBytes;            // coefficients [128 x 2 rp] bf16, K-major (1 or 2 blocks)
constexpr int kCaOffU = kCaOffC + 2 * kCrGroupBytes;                   // U^T slice [4 groups][rp][128 B]
constexpr int kCaOffI = kCaOffU + kCrGroups * kCrMaxRp * 128;          // identity [64 x 64] bf16, K-major, 128-byte swizzle
constexpr int kCaOffSsx = kCaOffI + 64 * 128;                          // [2 (block parity)][16 ranks][128] f32
constexpr int kCaOffSs = kCaOffSsx + 2 * kCrMaxCta * kCrRows * 4;      // [2 (block parity)][128] f32
constexpr int kCaOffRstd = kCaOffSs + 4 * kCrRows * 4;
constexpr int kCaOffW1 = kCaOffRstd + kCrRows * 4;
constexpr int kCaOffBars = kCaOffW1 + kCrW * 4;
constexpr int kCaSmemBytes = kCaOffBars + 256 + 1024;
constexpr uint32_t kIdescANegate = 1u << 13;

struct CaArgs {
  int n_rows, d;
  int rp;            // 32 or 64: coefficients used (columns of C_hi and of C_lo that enter the product)
  int c_rp;          // 32 or 64 >= rp: columns of C_hi (= of C_lo) in cb's layout [C_hi(c_rp) | C_lo(c_rp)] -- larger than rp
                     // when one coefficient pass for the largest rank serves a whole sweep of nested ranks
  int n_cta, n_tiles;
  const float* w_norm;
  float eps;
  int want_x, want_h;
};

__global__ void __launch_bounds__(kCrThreads, 1)
tb_lowrank_apply_cluster_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmC,
                                const __grid_constant__ CUtensorMap tmU, const __grid_constant__ CUtensorMap tmOutX,
                                const __grid_constant__ CUtensorMap tmOutH, const CaArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __builtin_assume(__isShared(smem));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kCaOffBars);
  uint64_t* full_x = bars;             // [2]
  uint64_t* empty_x = bars + 2;        // [2]
  uint64_t* full_u = bars + 4;
  uint64_t* full_c = bars + 5;
  uint64_t* empty_c = bars + 6;
  uint64_t* tfull = bars + 7;          // [2]
  uint64_t* tempty = bars + 9;         // [2]
  uint64_t* store_ready = bars + 11;   // [2]
  uint64_t* xbar = bars + 13;          // [2] by block parity: this CTA's arrival + the bytes of every CTA's partial sums
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const int n_cta = a.n_cta;
  const int crank = (n_cta > 1) ? static_cast<int>(cluster_ctarank()) : 0;
  const int cluster_id = static_cast<int>(blockIdx.x) / n_cta;
  const int num_clusters = static_cast<int>(gridDim.x) / n_cta;
  const int col0 = crank * kCrW;
  const int rp = a.rp;
  const int kc = 2 * a.c_rp / 64;                             // 64-column blocks of [C_hi | C_lo] in cb: 1 or 2
  const uint32_t u_group_bytes = static_cast<uint32_t>(rp) * 128u;
  const uint32_t smem0 = smem_u32(smem);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmC);
    tma_prefetch_desc(&tmU);
    tma_prefetch_desc(&tmOutH);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&full_x[i]), 1);
      mbar_init(smem_u32(&empty_x[i]), 1);
      mbar_init(smem_u32(&tfull[i]), 1);
      mbar_init(smem_u32(&tempty[i]), kCrEpiThreads / 32);
      mbar_init(smem_u32(&store_ready[i]), kCrEpiThreads / 32);
    }
    mbar_init(smem_u32(full_u), 1);
    mbar_init(smem_u32(full_c), 1);
    mbar_init(smem_u32(empty_c), 1);
    mbar_init(smem_u32(&xbar[0]), 1);
    mbar_init(smem_u32(&xbar[1]), 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<1>(smem_u32(tmem_slot), 512);
  if (warp == 3) {
    // identity tile: row n holds a one in column n; chunk c of row n sits at chunk position c ^ (n & 7)
    for (int i = lane; i < 64 * 8; i += 32) {
      const int n = i >> 3, c = i & 7;
      uint32_t w[4] = {0u, 0u, 0u, 0u};
      if ((n >> 3) == c) w[(n & 7) >> 1] = (n & 1) ? 0x3F800000u : 0x00003F80u;     // bf16 1.0 in the high / low half
      *reinterpret_cast<uint4*>(smem + kCaOffI + n * 128 + ((c ^ (n & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
    }
    fence_proxy_async_smem();
  }
  tc_fence_before();
  if (n_cta > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      mbar_arrive_expect_tx(smem_u32(full_u), static_cast<uint32_t>(kCrGroups) * u_group_bytes);
      for (int g = 0; g < kCrGroups; ++g)
        tma_load_2d(smem0 + kCaOffU + g * u_group_bytes, &tmU, smem_u32(full_u), col0 + g * 64, 0);
      int t = 0;
      for (int tile = cluster_id; tile < a.n_tiles; tile += num_clusters, ++t) {
        const int s = t & 1, u = t >> 1;
        mbar_wait(smem_u32(&empty_x[s]), (u & 1) ^ 1);
        mbar_arrive_expect_tx(smem_u32(&full_x[s]), kCrTileBytes);
        for (int g = 0; g < kCrGroups; ++g)
          tma_load_2d_hint(smem0 + kCaOffX + s * kCrTileBytes + g * kCrGroupBytes, &tmX, smem_u32(&full_x[s]),
                           col0 + g * 64, tile * kCrRows, kPolicyEvictFirst);
        mbar_wait(smem_u32(empty_c), (t & 1) ^ 1);
        mbar_arrive_expect_tx(smem_u32(full_c), static_cast<uint32_t>(kc) * kCrGroupBytes);
        for (int kb = 0; kb < kc; ++kb)
          tma_load_2d(smem0 + kCaOffC + kb * kCrGroupBytes, &tmC, smem_u32(full_c), kb * 64, tile * kCrRows);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    constexpr uint32_t idesc_p = umma_idesc_bf16_f32(kCrRows, kCrW) | kIdescBMajorMn | kIdescANegate;   // -C U^T
    constexpr uint32_t idesc_i = umma_idesc_bf16_f32(kCrRows, 64);                                       // + x I
    mbar_wait(smem_u32(full_u), 0);
    int t = 0;
    for (int tile = cluster_id; tile < a.n_tiles; tile += num_clusters, ++t) {
      const int s = t & 1, u = t >> 1;
      mbar_wait(smem_u32(full_c), t & 1);
      mbar_wait(smem_u32(&full_x[s]), u & 1);
      mbar_wait(smem_u32(&tempty[s]), (u & 1) ^ 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t d_tmem = tmem_base + s * kCrW;
        // heads, then remainders, 16 coefficients per step; both multiply the same rows of U^T. Column of cb =
        // part * c_rp + 16 k: the same sequence of products whatever c_rp is (same bits)
        for (int part = 0; part < 2; ++part) {
          for (int k = 0; k < rp / 16; ++k) {
            const int ccol = part * a.c_rp + k * 16;
            const uint64_t da = umma_desc_k_sw128(smem0 + kCaOffC + (ccol >> 6) * kCrGroupBytes);
            const uint64_t db = umma_desc_mn_sw128(smem0 + kCaOffU + k * 16 * 128, u_group_bytes);
            umma_bf16<1>(d_tmem, da + static_cast<uint64_t>(2 * ((ccol & 63) >> 4)), db, idesc_p, (part | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(smem_u32(empty_c));
        const uint64_t di = umma_desc_k_sw128(smem0 + kCaOffI);
        for (int g = 0; g < kCrGroups; ++g) {
          const uint64_t da = umma_desc_k_sw128(smem0 + kCaOffX + s * kCrTileBytes + g * kCrGroupBytes);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16<1>(d_tmem + g * 64, da + static_cast<uint64_t>(2 * k), di + static_cast<uint64_t>(2 * k), idesc_i, 1u);
        }
        umma_commit(smem_u32(&tfull[s]));
      }
      __syncwarp();
    }
  } else if (warp == 2) {
    // ================================ TMA store issuer ================================
    if (lane == 0) {
      int t = 0;
      for (int tile = cluster_id; tile < a.n_tiles; tile += num_clusters, ++t) {
        const int s = t & 1, u = t >> 1;
        mbar_wait(smem_u32(&store_ready[s]), u & 1);
        if (a.want_h) {
          for (int g = 0; g < kCrGroups; ++g)
            tma_store_2d(&tmOutH, smem0 + kCaOffX + s * kCrTileBytes + g * kCrGroupBytes, col0 + g * 64, tile * kCrRows);
          bulk_commit();
          bulk_wait_read0();
        }
        mbar_arrive(smem_u32(&empty_x[s]));
      }
      bulk_wait0();
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ================================ epilogue ================================
    const int e = static_cast<int>(threadIdx.x) - 128;
    const int quad = warp & 3, half = (warp - 4) >> 2;
    const int row = quad * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(quad * 32) << 16;
    float* s_ssx = reinterpret_cast<float*>(smem + kCaOffSsx);
    float* s_ss = reinterpret_cast<float*>(smem + kCaOffSs);
    float* s_rstd = reinterpret_cast<float*>(smem + kCaOffRstd);
    float* s_w1 = reinterpret_cast<float*>(smem + kCaOffW1);
    s_w1[e] = a.w_norm ? __fadd_rn(1.0f, __ldg(a.w_norm + col0 + e)) : 1.0f;
    s_ss[e] = 0.0f;                                          // [2 parities][128]
    epi_bar_sync<kCrEpiThreads>();
#ifdef TB_WAIT_PROFILE
    long long k2p[16];
    for (int i = 0; i < 16; ++i) k2p[i] = 0;
    const long long k2_start = clock64();
#endif
    int t = 0;
    for (int tile = cluster_id; tile < a.n_tiles; tile += num_clusters, ++t) {
      const int s = t & 1, u = t >> 1;
      uint8_t* xs = smem + kCaOffX + s * kCrTileBytes;
      const uint32_t tacc = tmem_base + lane_base + s * kCrW + half * 128;
      TB_K2_T(p0);
      mbar_wait(smem_u32(&tfull[s]), u & 1);
      tc_fence_after();
      TB_K2_T(p1);
      // ---- pass 1: sum of squares of the rounded x' (and x' itself into the stage when it is wanted)
      float ss = 0.0f;
      {
        uint32_t va[32], vb[32];
        tmem_ld_32x32(tacc, va);
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t (&v)[32] = (cc & 1) ? vb : va;
          tmem_ld_wait();
          if (cc + 1 < 4) tmem_ld_32x32(tacc + (cc + 1) * 32, (cc & 1) ? va : vb);   // in flight under this slice's arithmetic
          const int cbase = half * 128 + cc * 32;
          uint8_t* grp = xs + (cbase >> 6) * kCrGroupBytes + row * 128;
          const int lc0 = (cbase & 63) >> 3;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = __uint_as_float(v[q * 8 + j]);
            const Bf16x8 pk = pack8(o);
            unpack8(pk, o);
#pragma unroll
            for (int j = 0; j < 8; ++j) ss = fmaf(o[j], o[j], ss);
            if (a.want_x) *reinterpret_cast<int4*>(grp + (((lc0 + q) ^ (row & 7)) << 4)) = *reinterpret_cast<const int4*>(&pk);
          }
        }
      }
      // the row's two column halves meet in one zero-initialised slot (two addends: the order cannot change the sum)
      float* s_half = s_ss + (t & 1) * kCrRows;              // [parity][128]: the source of the copies below
      atomicAdd(&s_half[row], ss);
      fence_proxy_async_smem();
      TB_K2_T(p2);
      if (a.want_x) {
        epi_bar_sync<kCrEpiThreads>();
        if (e == 0) {
          for (int g = 0; g < kCrGroups; ++g)
            tma_store_2d(&tmOutX, smem0 + kCaOffX + s * kCrTileBytes + g * kCrGroupBytes, col0 + g * 64, tile * kCrRows);
          bulk_commit();
          bulk_wait_read0();
        }
      }
      epi_bar_sync<kCrEpiThreads>();
      // ---- every CTA gets every CTA's partial sums (128 rows = 512 bytes): one bulk copy per peer into
      // slot [parity][source rank], counted on the peer's mbarrier of this parity. (Scattered st.shared::cluster
      // stores cost a packet per lane: 1792 of them plus an arrive / wait round measured 4.6 k cycles per block.)
      const uint32_t xb = smem_u32(&xbar[t & 1]);
      if (e == 0) mbar_arrive_expect_tx(xb, static_cast<uint32_t>(n_cta) * kCrRows * 4);
      if (e < n_cta)
        dsmem_bulk_copy(mapa_shared(smem0 + kCaOffSsx + (((t & 1) * kCrMaxCta + crank) * kCrRows) * 4, static_cast<uint32_t>(e)),
                        smem_u32(s_half), kCrRows * 4, mapa_shared(xb, static_cast<uint32_t>(e)));
      TB_K2_T(p3);
      mbar_wait_cluster(xb, (t >> 1) & 1);
      TB_K2_T(p4);
      if (e < kCrRows) {
        float sum = 0.0f;
        for (int p = 0; p < n_cta; ++p) sum += s_ssx[((t & 1) * kCrMaxCta + p) * kCrRows + e];   // rank order: same bits everywhere
        // every peer has passed its wait of the previous block, so the copies that read the other parity's slot are
        // done: clear it for the next block
        s_ss[((t & 1) ^ 1) * kCrRows + e] = 0.0f;
        s_rstd[e] = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(sum, static_cast<float>(a.d)), a.eps)));
      }
      epi_bar_sync<kCrEpiThreads>();
      // ---- pass 2: h = bf16((x' * rstd) * (1 + w)) from the accumulator into the stage, in the swizzled tile layout
      if (a.want_h) {
        const float rs = s_rstd[row];
        uint32_t va[32], vb[32];
        tmem_ld_32x32(tacc, va);
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t (&v)[32] = (cc & 1) ? vb : va;
          tmem_ld_wait();
          if (cc + 1 < 4) tmem_ld_32x32(tacc + (cc + 1) * 32, (cc & 1) ? va : vb);
          const int cbase = half * 128 + cc * 32;
          uint8_t* grp = xs + (cbase >> 6) * kCrGroupBytes + row * 128;
          const int lc0 = (cbase & 63) >> 3;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = __uint_as_float(v[q * 8 + j]);
            const Bf16x8 pk = pack8(o);
            unpack8(pk, o);
            const float4 w0 = *reinterpret_cast<const float4*>(s_w1 + cbase + q * 8);
            const float4 w1 = *reinterpret_cast<const float4*>(s_w1 + cbase + q * 8 + 4);
            const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
            float h[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) h[j] = __fmul_rn(__fmul_rn(o[j], rs), ww[j]);
            const Bf16x8 ph = pack8(h);
            *reinterpret_cast<int4*>(grp + (((lc0 + q) ^ (row & 7)) << 4)) = *reinterpret_cast<const int4*>(&ph);
          }
        }
        fence_proxy_async_smem();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(smem_u32(&tempty[s]));
        mbar_arrive(smem_u32(&store_ready[s]));
      }
      TB_K2_T(p5);
      TB_K2_ADD(4, p0, p1); TB_K2_ADD(5, p1, p2); TB_K2_ADD(6, p2, p3); TB_K2_ADD(7, p3, p4); TB_K2_ADD(8, p4, p5);
    }
#ifdef TB_WAIT_PROFILE
    if (e == 0) {
      k2p[9] = clock64() - k2_start;
      k2p[10] = t;
      k2p[11] = num_clusters;
      for (int i = 0; i < 12; ++i) g_k2_prof[blockIdx.x & 255][i] = k2p[i];
    }
#endif
  }

  tc_fence_before();
  if (n_cta > 1) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

}  // namespace tb
