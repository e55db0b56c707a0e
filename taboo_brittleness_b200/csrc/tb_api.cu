// C-ABI entry points of libtb_b200.so (declared in include/tb_b200.h): argument checks,
// TMA tensor-map construction and kernel launches. No allocation, no synchronisation.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "../../include/tb_b200.h"
#include "tb_epilogues.cuh"
#include "tb_gemm.cuh"
#include "tb_rows.cuh"
#include "tb_cluster_rows.cuh"
#include "tb_ranked.cuh"
#include "tb_rankupdate.cuh"

// Progress counters of the soft lockstep between clusters (tb_gemm.cuh): a few rotating slots of
// 64 counters, so that launches in flight on different streams do not share one.
static __device__ int g_cluster_progress[8][128];

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define TB_CUDA(call)                                                                      \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess)                                                                 \
      return fail(TB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),     \
                  __FILE__, __LINE__);                                                     \
  } while (0)

struct DevInfo {
  int ok = 0;
  int sms = 0;
};

// per-device capability cache (plain ints; filled on first use of each device)
int device_info(DevInfo* out) {
  static int cached_major[64];
  static int cached_sms[64];
  static bool have[64];
  int dev = 0;
  TB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(TB_ERR_INVALID, "device ordinal %d out of range", dev);
  if (!have[dev]) {
    int major = 0, sms = 0;
    TB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    TB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    cached_major[dev] = major;
    cached_sms[dev] = sms;
    have[dev] = true;
  }
  if (cached_major[dev] != 10)
    return fail(TB_ERR_ARCH, "device %d is compute capability %d.x; these kernels are sm_100a only",
                dev, cached_major[dev]);
  out->ok = 1;
  out->sms = cached_sms[dev];
  return TB_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int get_encode_fn(EncodeTiledFn* fn) {
  static EncodeTiledFn cached = nullptr;
  if (!cached) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    TB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    if (q != cudaDriverEntryPointSuccess || p == nullptr)
      return fail(TB_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    cached = reinterpret_cast<EncodeTiledFn>(p);
  }
  *fn = cached;
  return TB_OK;
}

// bf16 row-major [rows, K] matrix, box = [box_rows, 64 elements] with the 128-byte swizzle the
// UMMA descriptors expect. Out-of-bounds box elements are zero-filled.
int make_tmap(CUtensorMap* map, const void* base, int64_t rows, int K, int box_rows) {
  EncodeTiledFn enc;
  int rc = get_encode_fn(&enc);
  if (rc != TB_OK) return rc;
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(K) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(tb::kBlockK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(TB_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld K=%d box_rows=%d)",
                static_cast<int>(r), static_cast<long long>(rows), K, box_rows);
  return TB_OK;
}

// fp32 logits in prompt tiles [n_rows / 128][V / 32][128 rows][32 ids] as a 4-d map (id in tile, row in prompt, tile,
// prompt), box = one warp's [32 rows x 32 ids], 128-byte swizzle
int make_tile_map(CUtensorMap* map, const float* base, int64_t n_rows, int V) {
  EncodeTiledFn enc;
  int rc = get_encode_fn(&enc);
  if (rc != TB_OK) return rc;
  const cuuint64_t n_ct = static_cast<cuuint64_t>(V / 32);
  cuuint64_t dims[4] = {32, 128, n_ct, static_cast<cuuint64_t>(n_rows / 128)};
  cuuint64_t strides[3] = {128, 128ull * 128, 128ull * 128 * n_ct};
  cuuint32_t box[4] = {32, 32, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(TB_ERR_CUDA, "cuTensorMapEncodeTiled (prompt tiles) failed with CUresult %d (rows=%lld V=%d)", static_cast<int>(r),
                static_cast<long long>(n_rows), V);
  return TB_OK;
}

int check_gemm_operands(const void* a, const void* b, int64_t M, int N, int K) {
  if (!a || !b) return fail(TB_ERR_INVALID, "null operand pointer");
  if (M <= 0 || N <= 0 || K <= 0) return fail(TB_ERR_INVALID, "empty GEMM shape M=%lld N=%d K=%d", (long long)M, N, K);
  if (M > 0x7fffff00ll) return fail(TB_ERR_INVALID, "too many rows (%lld)", (long long)M);
  if (K % 8 != 0) return fail(TB_ERR_INVALID, "K=%d must be a multiple of 8 (16-byte TMA row pitch)", K);
  if ((reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15))
    return fail(TB_ERR_INVALID, "operands must be 16-byte aligned");
  return TB_OK;
}

// L2 policy of the two operand streams. The stationary block is re-read once per swept tile
// and should stay resident; swept tiles are read by the few CTAs that share a chunk and are
// dead afterwards. TB_L2_POLICY=<stationary><swept> with n / f / l (normal, evict-first,
// evict-last) overrides the default for experiments.
void set_policies(tb::GemmArgs* g, bool stationary_is_a) {
  unsigned long long pol[2];
  const char* e = getenv("TB_L2_POLICY");
  const char* d = (e && strlen(e) == 2) ? e : "nn";
  for (int i = 0; i < 2; ++i)
    pol[i] = d[i] == 'l' ? tb::kPolicyEvictLast : d[i] == 'f' ? tb::kPolicyEvictFirst : tb::kPolicyEvictNormal;
  const char* dbg = getenv("TB_DEBUG_SKIP");
  g->debug_skip = dbg ? atoi(dbg) : 0;
  // soft lockstep between clusters (tb_gemm.cuh): on by default, slack of one tile; it only engages
  // for multicast clusters (launch_gemm), where the L2 -> SM fabric has the headroom to profit
  const char* mc = getenv("TB_MAX_CLUSTERS");   // experiments: cap the grid (side launches on leftover SMs)
  g->max_clusters = mc ? atoi(mc) : 0;
  const char* ls = getenv("TB_LOCKSTEP");
  g->lockstep = ls ? atoi(ls) : 1;
  g->progress = nullptr;
  g->diag = 0;
  g->policy_a = stationary_is_a ? pol[0] : pol[1];
  g->policy_b = stationary_is_a ? pol[1] : pol[0];
}

// group = CTAs per pair (1 or 2); R x C = pairs per cluster (stationary blocks x chunks)
tb::GemmArgs make_args(int64_t M, int N, int K, int n_chunks, int group, int R = 1, int C = 1) {
  tb::GemmArgs g;
  g.M = static_cast<int>(M);
  g.N = N;
  g.K = K;
  g.n_tiles = (N + tb::kBlockN - 1) / tb::kBlockN;
  if (n_chunks < 1) n_chunks = 1;
  if (n_chunks > g.n_tiles) n_chunks = g.n_tiles;
  g.tiles_per_chunk = (g.n_tiles + n_chunks - 1) / n_chunks;
  g.n_chunks = (g.n_tiles + g.tiles_per_chunk - 1) / g.tiles_per_chunk;  // no empty chunks
  g.m_blocks = static_cast<int>((M + tb::kBlockM * group - 1) / (tb::kBlockM * group));
  g.total_units = ((g.m_blocks + R - 1) / R) * (g.n_chunks / C);  // callers check n_chunks % C == 0
  set_policies(&g, /*stationary_is_a=*/true);
  return g;
}

// sweep-A orientation (tb_gemm.cuh): A [M, K] is swept in tiles of 128 * group rows, B [N, K] is
// cut into stationary blocks of 256 rows
tb::GemmArgs make_args_sweep_a(int M, int64_t N, int K, int n_chunks, int group, int R = 1, int C = 1) {
  tb::GemmArgs g;
  g.M = M;
  g.N = static_cast<int>(N);
  g.K = K;
  g.n_tiles = (M + tb::kBlockM * group - 1) / (tb::kBlockM * group);
  if (n_chunks < 1) n_chunks = 1;
  if (n_chunks > g.n_tiles) n_chunks = g.n_tiles;
  g.tiles_per_chunk = (g.n_tiles + n_chunks - 1) / n_chunks;
  g.n_chunks = (g.n_tiles + g.tiles_per_chunk - 1) / g.tiles_per_chunk;
  g.m_blocks = static_cast<int>((N + tb::kBlockN - 1) / tb::kBlockN);
  g.total_units = ((g.m_blocks + R - 1) / R) * (g.n_chunks / C);
  set_policies(&g, /*stationary_is_a=*/false);
  return g;
}

// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another
// GPU's memory hung the GPU: an operand of a multicast cluster mode must be plain device memory of the current
// device (peer, IPC-mapped, managed or host memory is refused; cluster modes 0 and 1 do not multicast).
// The whole extent [p, p + bytes) is checked, not only its first byte: the range is walked mapping by mapping
// (a cudaMalloc allocation is one mapping; a range stitched together with cuMemMap has one per mapped chunk), and
// every piece must be device memory of the current GPU. A range with more than kMaxMappings pieces is refused.
typedef CUresult (*PointerGetAttributesFn)(unsigned int, CUpointer_attribute*, void**, CUdeviceptr);

int get_pointer_attrs_fn(PointerGetAttributesFn* fn) {
  static PointerGetAttributesFn cached = nullptr;
  if (!cached) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    TB_CUDA(cudaGetDriverEntryPoint("cuPointerGetAttributes", &p, cudaEnableDefault, &q));
    if (q != cudaDriverEntryPointSuccess || p == nullptr)
      return fail(TB_ERR_CUDA, "cuPointerGetAttributes not available from the driver");
    cached = reinterpret_cast<PointerGetAttributesFn>(p);
  }
  *fn = cached;
  return TB_OK;
}

int operand_is_local(const void* p, size_t bytes, const char* what) {
  constexpr int kMaxMappings = 8192;
  PointerGetAttributesFn attrs_fn;
  int rc = get_pointer_attrs_fn(&attrs_fn);
  if (rc != TB_OK) return rc;
  int dev = -1;
  TB_CUDA(cudaGetDevice(&dev));
  const uintptr_t begin = reinterpret_cast<uintptr_t>(p);
  const uintptr_t end = begin + (bytes ? bytes : 1);
  uintptr_t q = begin;
  for (int n = 0; q < end; ++n) {
    if (n >= kMaxMappings)
      return fail(TB_ERR_INVALID, "%s spans more than %d separate mappings: the multicast cluster modes refuse it, "
                  "use cta_pair 0 or 1", what, kMaxMappings);
    unsigned int mem_type = 0, managed = 0;
    int ordinal = -1;
    CUdeviceptr map_base = 0, range_base = 0;
    size_t map_size = 0, range_size = 0;
    CUpointer_attribute which[7] = {CU_POINTER_ATTRIBUTE_MEMORY_TYPE, CU_POINTER_ATTRIBUTE_IS_MANAGED,
                                    CU_POINTER_ATTRIBUTE_DEVICE_ORDINAL, CU_POINTER_ATTRIBUTE_MAPPING_BASE_ADDR,
                                    CU_POINTER_ATTRIBUTE_MAPPING_SIZE, CU_POINTER_ATTRIBUTE_RANGE_START_ADDR,
                                    CU_POINTER_ATTRIBUTE_RANGE_SIZE};
    void* data[7] = {&mem_type, &managed, &ordinal, &map_base, &map_size, &range_base, &range_size};
    const CUresult r = attrs_fn(7, which, data, static_cast<CUdeviceptr>(q));
    if (r != CUDA_SUCCESS || mem_type != CU_MEMORYTYPE_DEVICE || managed || ordinal != dev)
      return fail(TB_ERR_INVALID, "%s is not memory of the current GPU (device %d) at byte %llu of %llu: the multicast "
                  "cluster modes take local operands only, use cta_pair 0 or 1", what, dev,
                  static_cast<unsigned long long>(q - begin), static_cast<unsigned long long>(bytes));
    // the piece that holds q: the mapped chunk where the driver reports one, else the allocation
    uintptr_t piece_end = 0;
    if (map_size > 0 && map_base <= q && q < map_base + map_size) piece_end = static_cast<uintptr_t>(map_base) + map_size;
    else if (range_size > 0 && range_base <= q && q < range_base + range_size) piece_end = static_cast<uintptr_t>(range_base) + range_size;
    if (piece_end <= q)
      return fail(TB_ERR_INVALID, "%s: the driver reports no mapping around byte %llu: the multicast cluster modes "
                  "refuse it, use cta_pair 0 or 1", what, static_cast<unsigned long long>(q - begin));
    q = piece_end;
  }
  return TB_OK;
}

template <class Epi, int G, int R = 1, int C = 1, bool H = false>
int launch_gemm(const void* a, const void* b, const tb::GemmArgs& g, const typename Epi::Params& ep,
                cudaStream_t stream, int sms) {
  using Cfg = tb::GemmCfg<G, tb::EpiStages<Epi>::value>;
  constexpr int kClusterCtas = G * R * C;
  if constexpr (R * C > 1) {
    // what the two tensor maps cover: [rows, K] bf16 each
    int lrc = operand_is_local(a, static_cast<size_t>(g.M) * g.K * 2, "operand A");
    if (lrc != TB_OK) return lrc;
    lrc = operand_is_local(b, static_cast<size_t>(g.N) * g.K * 2, "operand B");
    if (lrc != TB_OK) return lrc;
  }
  // every CTA loads 1/share of a slab (and multicasts it): the TMA box shrinks accordingly
  constexpr int kShareA = Epi::kSweepA ? R : C;
  constexpr int kShareB = Epi::kSweepA ? C : R;
  CUtensorMap tmA, tmB, tmS1;
  int rc = make_tmap(&tmA, a, g.M, g.K, tb::kBlockM / kShareA);
  if (rc != TB_OK) return rc;
  rc = make_tmap(&tmB, b, g.N, g.K, Cfg::kBRows / kShareB);
  if (rc != TB_OK) return rc;
  // whole-slab boxes of the shared operand, for the groups that run as plain pairs (H only)
  if (Epi::kSweepA) rc = make_tmap(&tmS1, a, g.M, g.K, tb::kBlockM);
  else rc = make_tmap(&tmS1, b, g.N, g.K, Cfg::kBRows);
  if (rc != TB_OK) return rc;
  auto kern = tb::tb_umma_gemm_kernel<Epi, G, R, C, H>;
  constexpr int kSmem = Cfg::kSmemBytes + Epi::kExtraSmem;
  TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.blockDim = dim3(tb::gemm_threads<Epi>(), 1, 1);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = H ? G : kClusterCtas;   // H: pairs are the fallback ...
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributePreferredClusterDimension;   // ... clusters of pairs the preference
  attr[1].val.clusterDim.x = kClusterCtas;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = H ? 2 : 1;
  // co-resident groups of kClusterCtas CTAs: sms / size for pairs and for preferred/fallback launches
  // (what does not fit a GPC as one cluster runs as pairs); fewer for plain larger clusters
  int clusters = sms / kClusterCtas;
  if (kClusterCtas > 2 && !H) {
    static int cached[64];
    int dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    if (cached[dev] == 0) {
      cfg.gridDim = dim3(static_cast<unsigned>(clusters * kClusterCtas), 1, 1);
      int n = 0;
      TB_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
      if (n < 1) return fail(TB_ERR_CUDA, "no cluster of %d CTAs fits on this device", kClusterCtas);
      cached[dev] = n;
    }
    clusters = cached[dev];
  }
  if (clusters > g.total_units) clusters = g.total_units;
  if (g.max_clusters > 0 && clusters > g.max_clusters) clusters = g.max_clusters;
  cfg.gridDim = dim3(static_cast<unsigned>(clusters * kClusterCtas), 1, 1);
  tb::GemmArgs ga = g;
  const int pace_slots = clusters * kClusterCtas / G;
  if (ga.lockstep > 0 && kClusterCtas > 2 && pace_slots <= 128 && clusters > 1) {
    // progress counters: one of 8 rotating 128-counter slots of a static device array (no allocation;
    // up to 8 launches may be in flight on different streams before two of them share a slot, and
    // sharing only costs pacing accuracy: every wait is bounded). They start "far ahead" (0x7f7f7f7f)
    // so that a counter nobody publishes to never holds anybody back.
    static std::atomic<unsigned> slot{0};
    int* base = nullptr;
    TB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&base), g_cluster_progress));
    ga.progress = base + 128 * (slot.fetch_add(1) & 7u);
    TB_CUDA(cudaMemsetAsync(ga.progress, 0x7f, 128 * sizeof(int), stream));
  } else {
    ga.lockstep = 0;
  }
  TB_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmS1, ga, ep));
  return TB_OK;
}

// cluster shape of the lens kernels, selected by the cta_pair argument of the C ABI:
//   0: single CTAs   1: CTA pairs   2: 2x1 pairs (two row blocks share W_U tiles)
//   3: 1x2 pairs (two chunks share the row block)   4: 2x2 pairs (both)
//   5: 2x1 pairs as the PREFERRED cluster shape with plain pairs as fallback (all SMs busy)
struct ClusterMode { int group, R, C; int hetero = 0; };
int cluster_mode(int cta_pair, ClusterMode* m) {
  switch (cta_pair) {
    case 0: *m = {1, 1, 1}; return TB_OK;
    case 1: *m = {2, 1, 1}; return TB_OK;
    case 2: *m = {2, 2, 1}; return TB_OK;
    case 3: *m = {2, 1, 2}; return TB_OK;
    case 4: *m = {2, 2, 2}; return TB_OK;
    case 5: *m = {2, 2, 1, 1}; return TB_OK;
  }
  return fail(TB_ERR_INVALID, "cta_pair=%d: cluster mode must be 0..5", cta_pair);
}

template <class Epi>
int launch_gemm_mode(const ClusterMode& m, const void* a, const void* b, const tb::GemmArgs& g,
                     const typename Epi::Params& ep, cudaStream_t s, int sms) {
  if (m.group == 1) return launch_gemm<Epi, 1>(a, b, g, ep, s, sms);
  if (m.R == 1 && m.C == 1) return launch_gemm<Epi, 2, 1, 1>(a, b, g, ep, s, sms);
  if (m.R == 2 && m.C == 1 && m.hetero) return launch_gemm<Epi, 2, 2, 1, true>(a, b, g, ep, s, sms);
  if (m.R == 2 && m.C == 1) return launch_gemm<Epi, 2, 2, 1>(a, b, g, ep, s, sms);
  if (m.R == 1 && m.C == 2) return launch_gemm<Epi, 2, 1, 2>(a, b, g, ep, s, sms);
  return launch_gemm<Epi, 2, 2, 2>(a, b, g, ep, s, sms);
}

// choose vectors-per-thread and block size for the row kernels: d/8 16-byte vectors per row
int row_launch_shape(int d, int* nv, int* threads) {
  if (d <= 0 || d % 8 != 0) return fail(TB_ERR_INVALID, "d=%d must be a positive multiple of 8", d);
  const int nvec = d / 8;
  int v = 1;
  while (v <= 4 && (nvec + v - 1) / v > 512) v *= 2;
  if (v > 4) return fail(TB_ERR_INVALID, "d=%d too large for the row kernels (max 16384)", d);
  *nv = v;
  int t = (nvec + v - 1) / v;
  t = (t + 31) / 32 * 32;
  if (t < 64) t = 64;
  *threads = t;
  return TB_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Built-in scratch for tb_logit_lens_static: TB_STATIC_LENS_MAX_SLOTS (row, chunk) partials of
// 23 words each (max, sumexp, secret logit, 10 values, 10 ids) = 6,029,312 bytes that are part
// of the library image on every GPU that loads it, for the life of the process.
// In host code `g_lens_static_scratch` is only a handle for the symbol -- not a device address.
// The address a kernel may use is fetched once per device with cudaGetSymbolAddress (below),
// with that device current.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_lens_static_scratch[static_cast<size_t>(TB_STATIC_LENS_MAX_SLOTS) * 23];

namespace {

std::mutex g_static_mu;
float* g_static_ptr[64];
bool g_static_ready[64];

// device address of the built-in scratch on the CURRENT device (resolved on first use)
int static_scratch_address(float** out) {
  int dev = 0;
  TB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(TB_ERR_INVALID, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ready[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_lens_static_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail(TB_ERR_CUDA, "cudaGetSymbolAddress(g_lens_static_scratch) failed on device %d: %s", dev,
                  cudaGetErrorString(e));
    g_static_ptr[dev] = static_cast<float*>(p);
    g_static_ready[dev] = true;
  }
  *out = g_static_ptr[dev];
  return TB_OK;
}

}  // namespace

extern "C" {

int tb_version(void) { return 101; }

int tb_lens_parts_per_chunk(void) { return tb::LensEpi::kEpiWarps / 4; }

const char* tb_last_error(void) { return g_err; }

int tb_device_check(void) {
  DevInfo di;
  return device_info(&di);
}

int tb_gemm_bf16_nt(const void* a, const void* b, int64_t M, int N, int K, int n_chunks, int cta_pair,
                    float* out, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  rc = check_gemm_operands(a, b, M, N, K);
  if (rc != TB_OK) return rc;
  if (!out) return fail(TB_ERR_INVALID, "null output");
  const int G = cta_pair ? 2 : 1;
  tb::GemmArgs g = make_args(M, N, K, n_chunks, G);
  tb::DenseEpi::Params ep{out};
  if (G == 2) return launch_gemm<tb::DenseEpi, 2>(a, b, g, ep, static_cast<cudaStream_t>(stream), di.sms);
  return launch_gemm<tb::DenseEpi, 1>(a, b, g, ep, static_cast<cudaStream_t>(stream), di.sms);
}

int tb_sae_encode(const void* x, const void* w_enc_t, const float* b_enc, const float* theta,
                  int64_t n_rows, int d, int S, const int64_t* row_start, const int32_t* row_cap,
                  int cap, int32_t* out_idx, float* out_val, int32_t* out_count,
                  const uint16_t* rank, int rank_rows, int budget, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  rc = check_gemm_operands(x, w_enc_t, n_rows, S, d);
  if (rc != TB_OK) return rc;
  if (!b_enc || !theta || !out_count) return fail(TB_ERR_INVALID, "null b_enc / theta / out_count");
  if ((reinterpret_cast<uintptr_t>(b_enc) & 15) || (reinterpret_cast<uintptr_t>(theta) & 15))
    return fail(TB_ERR_INVALID, "b_enc / theta must be 16-byte aligned");
  if (out_idx && !out_val) return fail(TB_ERR_INVALID, "out_idx given without out_val");
  if (out_idx && !row_start && cap <= 0) return fail(TB_ERR_INVALID, "fixed-stride lists need cap > 0");
  if (rank && rank_rows < 0) return fail(TB_ERR_INVALID, "rank_rows < 0");
  tb::EncodeEpi::Params ep;
  ep.b_enc = b_enc;
  ep.theta = theta;
  ep.row_start = row_start;
  ep.cap = cap;
  ep.row_cap = row_cap;
  ep.out_idx = out_idx;
  ep.out_val = out_val;
  ep.out_count = out_count;
  ep.rank = rank;
  ep.rank_rows = rank_rows;
  ep.budget = budget;
  ep.sub_lists = 1;
  // one unit per row block: a row's list is written by a single thread, in latent order.
  // CTA pairs (UMMA 256 x 256) once there are enough 256-row blocks to fill them; single CTAs for small inputs
  if (n_rows >= static_cast<int64_t>(tb::kBlockM) * di.sms) {
    // Large batches: 2x1 clusters of pairs as the preferred shape (two row blocks share every W_enc^T tile through
    // TMA multicast, the clusters are paced against each other), plain pairs on the SMs that form no cluster -- the
    // lens kernels' default. Plain pairs alone re-read W_enc^T from DRAM once per pair: the ncu capture of round 2a
    // shows 35.7 GB of DRAM reads per 75,776-row launch for 0.66 GB of operands (profiles/r02b_encode_*).
    // Operands that are not this GPU's own memory keep the plain pairs (multicast loads read local memory only).
    if (operand_is_local(x, static_cast<size_t>(n_rows) * d * 2, "x") == TB_OK &&
        operand_is_local(w_enc_t, static_cast<size_t>(S) * d * 2, "w_enc_t") == TB_OK) {
      tb::GemmArgs g = make_args(n_rows, S, d, 1, 2, 2, 1);
      return launch_gemm<tb::EncodeEpi, 2, 2, 1, true>(x, w_enc_t, g, ep, static_cast<cudaStream_t>(stream), di.sms);
    }
    tb::GemmArgs g = make_args(n_rows, S, d, 1, 2);
    return launch_gemm<tb::EncodeEpi, 2>(x, w_enc_t, g, ep, static_cast<cudaStream_t>(stream), di.sms);
  }
  tb::GemmArgs g = make_args(n_rows, S, d, 1, 1);
  return launch_gemm<tb::EncodeEpi, 1>(x, w_enc_t, g, ep, static_cast<cudaStream_t>(stream), di.sms);
}

int tb_sae_encode_chunked(const void* x, const void* w_enc_t, const float* b_enc, const float* theta, int64_t n_rows,
                          int d, int S, int cap, int n_sub, int32_t* out_idx, float* out_val, int32_t* sub_count,
                          int32_t* out_count, int32_t* out_truncated, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  rc = check_gemm_operands(x, w_enc_t, n_rows, S, d);
  if (rc != TB_OK) return rc;
  if (!b_enc || !theta || !out_idx || !out_val || !sub_count || !out_count || !out_truncated)
    return fail(TB_ERR_INVALID, "null pointer");
  if ((reinterpret_cast<uintptr_t>(b_enc) & 15) || (reinterpret_cast<uintptr_t>(theta) & 15))
    return fail(TB_ERR_INVALID, "b_enc / theta must be 16-byte aligned");
  if (n_sub < 2 || n_sub > 8 || cap <= 0 || cap % n_sub) return fail(TB_ERR_INVALID, "n_sub=%d must be 2..8 and divide cap=%d", n_sub, cap);
  tb::GemmArgs g = make_args(n_rows, S, d, n_sub, 2, 2, 1);
  if (g.n_chunks != n_sub)
    return fail(TB_ERR_INVALID, "S=%d: %d tiles do not split into %d non-empty chunks", S, g.n_tiles, n_sub);
  // operands of another GPU would be read through TMA multicast: refuse (the caller keeps tb_sae_encode)
  rc = operand_is_local(x, static_cast<size_t>(n_rows) * d * 2, "x");
  if (rc != TB_OK) return rc;
  rc = operand_is_local(w_enc_t, static_cast<size_t>(S) * d * 2, "w_enc_t");
  if (rc != TB_OK) return rc;
  tb::EncodeEpi::Params ep;
  ep.b_enc = b_enc;
  ep.theta = theta;
  ep.row_start = nullptr;
  ep.cap = cap;
  ep.row_cap = nullptr;
  ep.out_idx = out_idx;
  ep.out_val = out_val;
  ep.out_count = sub_count;
  ep.rank = nullptr;
  ep.rank_rows = 0;
  ep.budget = 0;
  ep.sub_lists = n_sub;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  rc = launch_gemm<tb::EncodeEpi, 2, 2, 1, true>(x, w_enc_t, g, ep, s, di.sms);
  if (rc != TB_OK) return rc;
  const int64_t threads = n_rows * 32;
  tb::tb_merge_sublists_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, s>>>(
      sub_count, n_sub, cap, static_cast<int>(n_rows), out_idx, out_val, out_count, out_truncated);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_sae_latent_mean(const int64_t* row_start, int cap, const int32_t* count, const int32_t* idx,
                       const float* val, const int64_t* seg, int rows_per_prompt, int64_t n_prompts,
                       int S, float* out_mean, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!count || !idx || !val || !out_mean) return fail(TB_ERR_INVALID, "null pointer");
  if (n_prompts < 0 || S <= 0) return fail(TB_ERR_INVALID, "bad shape");
  if (!seg && rows_per_prompt <= 0) return fail(TB_ERR_INVALID, "need seg or rows_per_prompt > 0");
  if (!row_start && cap <= 0) return fail(TB_ERR_INVALID, "fixed-stride lists need cap > 0");
  if (n_prompts == 0) return TB_OK;
  tb::tb_latent_mean_kernel<<<static_cast<unsigned>(n_prompts), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      row_start, cap, count, idx, val, seg, rows_per_prompt, S, out_mean);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_sae_latent_wsum(const int64_t* row_start, int cap, const int32_t* count, const int32_t* idx,
                       const float* val, const int64_t* seg, int rows_per_prompt, int64_t n_prompts, int S,
                       const float* row_weight, int power, float* out_sum, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!count || !idx || !val || !out_sum) return fail(TB_ERR_INVALID, "null pointer");
  if (n_prompts < 0 || S <= 0) return fail(TB_ERR_INVALID, "bad shape");
  if (power != 1 && power != 2) return fail(TB_ERR_INVALID, "power must be 1 or 2");
  if (!seg && rows_per_prompt <= 0) return fail(TB_ERR_INVALID, "need seg or rows_per_prompt > 0");
  if (!row_start && cap <= 0) return fail(TB_ERR_INVALID, "fixed-stride lists need cap > 0");
  if (n_prompts == 0) return TB_OK;
  tb::tb_latent_wsum_kernel<<<static_cast<unsigned>(n_prompts), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      row_start, cap, count, idx, val, seg, rows_per_prompt, S, row_weight, power, out_sum);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_topm_select(const float* mean, int64_t n_prompts, int S, int m, int32_t* out_idx, float* out_val,
                   uint16_t* rank, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!mean || !out_idx || !out_val) return fail(TB_ERR_INVALID, "null pointer");
  if (m < 1 || m > TB_TOPM_MAX) return fail(TB_ERR_INVALID, "m=%d outside [1, %d]", m, TB_TOPM_MAX);
  if (S < 1) return fail(TB_ERR_INVALID, "S=%d", S);
  if (n_prompts == 0) return TB_OK;
  tb::tb_topm_select_kernel<<<static_cast<unsigned>(n_prompts), tb::kTopmThreads, 0,
                              static_cast<cudaStream_t>(stream)>>>(mean, S, m, out_idx, out_val, rank);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_sae_ablate_decode(const void* x, const int64_t* row_start, int cap, const int32_t* count,
                         const int32_t* idx, const float* val, const void* w_dec, const float* b_dec,
                         const uint16_t* rank, const int32_t* row_table, int rank_rows, int budget,
                         int mode, int64_t n_rows, int d, int S, const float* w_norm, float eps,
                         void* out_x, void* out_h, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!count || !idx || !val || !w_dec) return fail(TB_ERR_INVALID, "null pointer");
  if (mode != 0 && mode != 1) return fail(TB_ERR_INVALID, "mode must be 0 (error term) or 1 (plain decode)");
  if (mode == 0 && !x) return fail(TB_ERR_INVALID, "mode 0 needs x");
  if (mode == 1 && !b_dec) return fail(TB_ERR_INVALID, "mode 1 needs b_dec");
  if (!out_x && !out_h) return fail(TB_ERR_INVALID, "no output requested");
  if (out_h && !w_norm) return fail(TB_ERR_INVALID, "out_h needs w_norm");
  if (!row_start && cap <= 0) return fail(TB_ERR_INVALID, "fixed-stride lists need cap > 0");
  if (n_rows == 0) return TB_OK;
  if (d <= 0 || d % 8 != 0) return fail(TB_ERR_INVALID, "d=%d must be a positive multiple of 8", d);
  const int nv = (d / 8 + tb::kRowGroupThreads - 1) / tb::kRowGroupThreads;   // 16-byte vectors per thread
  if (nv > 8) return fail(TB_ERR_INVALID, "d=%d too large for the ablate/decode kernel (max 8192)", d);
  tb::AblateArgs a;
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.row_start = row_start;
  a.cap = cap;
  a.count = count;
  a.idx = idx;
  a.val = val;
  a.w_dec = static_cast<const __nv_bfloat16*>(w_dec);
  a.b_dec = b_dec;
  a.rank = rank;
  a.row_table = row_table;
  a.rank_rows = rank_rows;
  a.budget = budget;
  a.mode = mode;
  a.S = S;
  a.w_norm = w_norm;
  a.eps = eps;
  a.out_x = static_cast<__nv_bfloat16*>(out_x);
  a.out_h = static_cast<__nv_bfloat16*>(out_h);
  a.n_rows = static_cast<int>(n_rows);
  a.d = d;
  const int64_t blocks = (n_rows + tb::kRowGroups - 1) / tb::kRowGroups;
  const int64_t max_grid = static_cast<int64_t>(di.sms) * 4;
  const unsigned grid = static_cast<unsigned>(blocks < max_grid ? blocks : max_grid);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (nv == 1) tb::tb_ablate_decode_kernel<1><<<grid, tb::kAblThreads, 0, s>>>(a);
  else if (nv == 2) tb::tb_ablate_decode_kernel<2><<<grid, tb::kAblThreads, 0, s>>>(a);
  else if (nv <= 4) tb::tb_ablate_decode_kernel<4><<<grid, tb::kAblThreads, 0, s>>>(a);
  else tb::tb_ablate_decode_kernel<8><<<grid, tb::kAblThreads, 0, s>>>(a);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_sae_rank_lists(const int64_t* row_start, int cap, const int32_t* count, const int32_t* idx, const float* val,
                      const uint16_t* rank, int rank_rows, int64_t n_rows, int S, int32_t* out_key, float* out_val,
                      int32_t* out_count, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!count || !idx || !val || !rank || !out_key || !out_val || !out_count) return fail(TB_ERR_INVALID, "null pointer");
  if (!row_start && cap <= 0) return fail(TB_ERR_INVALID, "fixed-stride lists need cap > 0");
  if (rank_rows < 0) return fail(TB_ERR_INVALID, "rank_rows < 0");
  if (S < 1 || S > (1 << 20)) return fail(TB_ERR_INVALID, "S=%d outside [1, 2^20] (latent ids are packed into 20 bits)", S);
  if (n_rows == 0) return TB_OK;
  const int64_t max_grid = static_cast<int64_t>(di.sms) * 8;
  const unsigned grid = static_cast<unsigned>(n_rows < max_grid ? n_rows : max_grid);
  tb::tb_rank_lists_kernel<<<grid, tb::kRankListThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      row_start, cap, count, idx, val, rank, rank_rows, static_cast<int>(n_rows), S, out_key, out_val, out_count);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_sae_ablate_ranked(const void* x, const int64_t* row_start, int cap, const int32_t* sel_count,
                         const int32_t* sel_key, const float* sel_val, const void* w_dec, int budget, int64_t n_rows,
                         int d, const float* w_norm, float eps, void* out_x, void* out_h, float* out_rstd,
                         void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!x || !sel_count || !sel_key || !sel_val || !w_dec) return fail(TB_ERR_INVALID, "null pointer");
  if (budget < 0 || budget > tb::kRankedMax) return fail(TB_ERR_INVALID, "budget %d outside [0, %d]", budget, tb::kRankedMax);
  if (!out_x && !out_h && !out_rstd) return fail(TB_ERR_INVALID, "no output requested");
  if (out_h && !w_norm) return fail(TB_ERR_INVALID, "out_h needs w_norm");
  if (!row_start && cap <= 0) return fail(TB_ERR_INVALID, "fixed-stride lists need cap > 0");
  if (n_rows == 0) return TB_OK;
  if (d <= 0 || d % 8 != 0) return fail(TB_ERR_INVALID, "d=%d must be a positive multiple of 8", d);
  const int nv = (d / 8 + tb::kRowGroupThreads - 1) / tb::kRowGroupThreads;
  if (nv > 8) return fail(TB_ERR_INVALID, "d=%d too large for the ablate/decode kernel (max 8192)", d);
  tb::AblateRankedArgs a;
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.row_start = row_start;
  a.cap = cap;
  a.sel_count = sel_count;
  a.sel_key = sel_key;
  a.sel_val = sel_val;
  a.w_dec = static_cast<const __nv_bfloat16*>(w_dec);
  a.budget = budget;
  {
    // prefixes this long go to the whole-CTA kernel (TB_HEAVY_MIN overrides; larger than any budget = never)
    const char* hm = getenv("TB_HEAVY_MIN");
    a.heavy_min = hm ? atoi(hm) : 48;
    if (a.heavy_min < 1 || d > 8 * 2 * tb::kHeavyThreads) a.heavy_min = tb::kRankedMax + 1;
  }
  a.w_norm = w_norm;
  a.eps = eps;
  a.out_x = static_cast<__nv_bfloat16*>(out_x);
  a.out_h = static_cast<__nv_bfloat16*>(out_h);
  a.out_rstd = out_rstd;
  a.n_rows = static_cast<int>(n_rows);
  a.d = d;
  const int64_t blocks = (n_rows + tb::kRowGroups - 1) / tb::kRowGroups;
  const int64_t max_grid = static_cast<int64_t>(di.sms) * 4;
  const unsigned grid = static_cast<unsigned>(blocks < max_grid ? blocks : max_grid);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (nv == 1) tb::tb_ablate_ranked_kernel<1><<<grid, tb::kAblThreads, 0, s>>>(a);
  else if (nv == 2) tb::tb_ablate_ranked_kernel<2><<<grid, tb::kAblThreads, 0, s>>>(a);
  else if (nv <= 4) tb::tb_ablate_ranked_kernel<4><<<grid, tb::kAblThreads, 0, s>>>(a);
  else tb::tb_ablate_ranked_kernel<8><<<grid, tb::kAblThreads, 0, s>>>(a);
  TB_CUDA(cudaGetLastError());
  if (a.heavy_min <= budget && d <= 8 * 2 * tb::kHeavyThreads) {
    // the rows the row-group kernel skipped (a prefix of heavy_min entries or more): a whole CTA per row
    const int64_t hgrid = static_cast<int64_t>(di.sms) * 2;      // one resident wave (two 512-thread CTAs per SM)
    const unsigned hg = static_cast<unsigned>(n_rows < hgrid ? n_rows : hgrid);
    if (d <= 8 * tb::kHeavyThreads) tb::tb_ablate_heavy_kernel<1><<<hg, tb::kHeavyThreads, 0, s>>>(a);
    else tb::tb_ablate_heavy_kernel<2><<<hg, tb::kHeavyThreads, 0, s>>>(a);
    TB_CUDA(cudaGetLastError());
  }
  return TB_OK;
}

int tb_sae_ablate_ranked_multi(const void* x, const int64_t* row_start, int cap, const int32_t* sel_count,
                               const int32_t* sel_key, const float* sel_val, const void* w_dec, const int32_t* budgets_host,
                               int n_budgets, int64_t n_rows, int d, const float* w_norm, float eps, void* out_x, void* out_h,
                               float* out_rstd, int64_t out_stride, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!x || !sel_count || !sel_key || !sel_val || !w_dec || !budgets_host) return fail(TB_ERR_INVALID, "null pointer");
  if (n_budgets < 1 || n_budgets > tb::kMaxBudgets) return fail(TB_ERR_INVALID, "n_budgets=%d outside [1, %d]", n_budgets, tb::kMaxBudgets);
  for (int k = 0; k < n_budgets; ++k)
    if (budgets_host[k] < 0 || budgets_host[k] > tb::kRankedMax || (k > 0 && budgets_host[k] <= budgets_host[k - 1]))
      return fail(TB_ERR_INVALID, "budgets must be strictly ascending within [0, %d]", tb::kRankedMax);
  if (!out_x && !out_h && !out_rstd) return fail(TB_ERR_INVALID, "no output requested");
  if (out_h && !w_norm) return fail(TB_ERR_INVALID, "out_h needs w_norm");
  if (!row_start && cap <= 0) return fail(TB_ERR_INVALID, "fixed-stride lists need cap > 0");
  if (n_rows == 0) return TB_OK;
  if (d <= 0 || d % 8 != 0) return fail(TB_ERR_INVALID, "d=%d must be a positive multiple of 8", d);
  if ((out_x || out_h) && n_budgets > 1 && out_stride < n_rows * static_cast<int64_t>(d))
    return fail(TB_ERR_INVALID, "out_stride=%lld smaller than one output (%lld elements)", static_cast<long long>(out_stride),
                static_cast<long long>(n_rows * static_cast<int64_t>(d)));
  const int nv = (d / 8 + tb::kRowGroupThreads - 1) / tb::kRowGroupThreads;
  if (nv > 8) return fail(TB_ERR_INVALID, "d=%d too large for the ablate/decode kernel (max 8192)", d);
  tb::AblateRankedMultiArgs m;
  tb::AblateRankedArgs& a = m.base;
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.row_start = row_start;
  a.cap = cap;
  a.sel_count = sel_count;
  a.sel_key = sel_key;
  a.sel_val = sel_val;
  a.w_dec = static_cast<const __nv_bfloat16*>(w_dec);
  a.budget = budgets_host[n_budgets - 1];
  {
    const char* hm = getenv("TB_HEAVY_MIN");
    a.heavy_min = hm ? atoi(hm) : 48;
    if (a.heavy_min < 1 || d > 8 * 2 * tb::kHeavyThreads) a.heavy_min = tb::kRankedMax + 1;
  }
  a.w_norm = w_norm;
  a.eps = eps;
  a.out_x = static_cast<__nv_bfloat16*>(out_x);
  a.out_h = static_cast<__nv_bfloat16*>(out_h);
  a.out_rstd = out_rstd;
  a.n_rows = static_cast<int>(n_rows);
  a.d = d;
  m.n_budgets = n_budgets;
  for (int k = 0; k < tb::kMaxBudgets; ++k) m.budgets[k] = k < n_budgets ? budgets_host[k] : 0;
  m.out_stride = out_stride;
  const int64_t blocks = (n_rows + tb::kRowGroups - 1) / tb::kRowGroups;
  const int64_t max_grid = static_cast<int64_t>(di.sms) * 4;
  const unsigned grid = static_cast<unsigned>(blocks < max_grid ? blocks : max_grid);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (nv == 1) tb::tb_ablate_ranked_multi_kernel<1><<<grid, tb::kAblThreads, 0, st>>>(m);
  else if (nv == 2) tb::tb_ablate_ranked_multi_kernel<2><<<grid, tb::kAblThreads, 0, st>>>(m);
  else if (nv <= 4) tb::tb_ablate_ranked_multi_kernel<4><<<grid, tb::kAblThreads, 0, st>>>(m);
  else tb::tb_ablate_ranked_multi_kernel<8><<<grid, tb::kAblThreads, 0, st>>>(m);
  TB_CUDA(cudaGetLastError());
  if (a.heavy_min <= a.budget && d <= 8 * 2 * tb::kHeavyThreads) {
    const int64_t hgrid = static_cast<int64_t>(di.sms) * 2;
    const unsigned hg = static_cast<unsigned>(n_rows < hgrid ? n_rows : hgrid);
    if (d <= 8 * tb::kHeavyThreads) tb::tb_ablate_heavy_multi_kernel<1><<<hg, tb::kHeavyThreads, 0, st>>>(m);
    else tb::tb_ablate_heavy_multi_kernel<2><<<hg, tb::kHeavyThreads, 0, st>>>(m);
    TB_CUDA(cudaGetLastError());
  }
  return TB_OK;
}

int tb_lowrank_remove(const void* x, const void* ut, int r, int64_t n_rows, int d, const float* w_norm,
                      float eps, void* out_x, void* out_h, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!x || (!ut && r > 0)) return fail(TB_ERR_INVALID, "null pointer");
  if (r < 0 || r > tb::kProjMaxRank) return fail(TB_ERR_INVALID, "rank %d outside [0, %d]", r, tb::kProjMaxRank);
  if (!out_x && !out_h) return fail(TB_ERR_INVALID, "no output requested");
  if (out_h && !w_norm) return fail(TB_ERR_INVALID, "out_h needs w_norm");
  if (n_rows == 0) return TB_OK;
  int nv, threads;
  rc = row_launch_shape(d, &nv, &threads);
  if (rc != TB_OK) return rc;
  // kProjRows rows of fp32 accumulators per thread: keep NV small by using more threads
  if (nv > 2) return fail(TB_ERR_INVALID, "d=%d too large for the projection kernel (max 8192)", d);
  tb::ProjArgs a;
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.ut = static_cast<const __nv_bfloat16*>(ut);
  a.r = r;
  a.w_norm = w_norm;
  a.eps = eps;
  a.out_x = static_cast<__nv_bfloat16*>(out_x);
  a.out_h = static_cast<__nv_bfloat16*>(out_h);
  a.n_rows = static_cast<int>(n_rows);
  a.d = d;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (r >= 1 && r <= tb::kProjTinyRank && d <= 8 * 4 * tb::kRowGroupThreads) {
    // ranks 1-2: the norm kernel's row-group layout with the projection spliced in
    const int nvt = (d / 8 + tb::kRowGroupThreads - 1) / tb::kRowGroupThreads;
    const size_t sm = static_cast<size_t>(r) * d * 2;
    const int64_t blocks = (n_rows + tb::kRowGroups - 1) / tb::kRowGroups;
    const int64_t mg = static_cast<int64_t>(di.sms) * 8;
    const unsigned gr = static_cast<unsigned>(blocks < mg ? blocks : mg);
    if (nvt == 1) tb::tb_lowrank_tiny_kernel<1><<<gr, tb::kAblThreads, sm, s>>>(a);
    else if (nvt == 2) tb::tb_lowrank_tiny_kernel<2><<<gr, tb::kAblThreads, sm, s>>>(a);
    else tb::tb_lowrank_tiny_kernel<4><<<gr, tb::kAblThreads, sm, s>>>(a);
    TB_CUDA(cudaGetLastError());
    return TB_OK;
  }
  if (r >= 1 && r <= tb::kProjSmallRank && d <= 8 * 2 * tb::kProjSmallThreads) {
    // small ranks: U^T resident in shared memory, four rows per tile in registers
    const int nvs = (d / 8 + tb::kProjSmallThreads - 1) / tb::kProjSmallThreads;
    const size_t sm = static_cast<size_t>(r) * d * 2;   // U^T stays resident
    const int64_t tl = (n_rows + tb::kProjSmallRows - 1) / tb::kProjSmallRows;
    const int64_t mg = static_cast<int64_t>(di.sms) * 2;   // two CTAs per SM (register-limited)
    const unsigned gr = static_cast<unsigned>(tl < mg ? tl : mg);
    if (nvs == 1) {
      TB_CUDA(cudaFuncSetAttribute(tb::tb_lowrank_small_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      tb::tb_lowrank_small_kernel<1><<<gr, tb::kProjSmallThreads, sm, s>>>(a);
    } else if (nvs == 2) {
      TB_CUDA(cudaFuncSetAttribute(tb::tb_lowrank_small_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      tb::tb_lowrank_small_kernel<2><<<gr, tb::kProjSmallThreads, sm, s>>>(a);
    } else {
      return fail(TB_ERR_INVALID, "d=%d too large for the small-rank projection kernel", d);
    }
    TB_CUDA(cudaGetLastError());
    return TB_OK;
  }
  const size_t smem = static_cast<size_t>(tb::kProjRows) * d * 2;
  const int64_t tiles = (n_rows + tb::kProjRows - 1) / tb::kProjRows;
  const int64_t max_grid = static_cast<int64_t>(di.sms) * 4;
  const unsigned grid = static_cast<unsigned>(tiles < max_grid ? tiles : max_grid);
  if (nv == 1) {
    TB_CUDA(cudaFuncSetAttribute(tb::tb_lowrank_remove_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tb::tb_lowrank_remove_kernel<1><<<grid, threads, smem, s>>>(a);
  } else {
    TB_CUDA(cudaFuncSetAttribute(tb::tb_lowrank_remove_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tb::tb_lowrank_remove_kernel<2><<<grid, threads, smem, s>>>(a);
  }
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

namespace {
// K2 on the tensor cores: scratch layout (256-byte aligned pieces) and the chunking of step 2
struct ProjTcPlan {
  int rp, kc, n_chunks, n_parts;
  size_t off_cb, off_ub, off_ss, bytes;
};
ProjTcPlan proj_tc_plan(int64_t n_rows, int d, int r, int sms) {
  ProjTcPlan p;
  p.rp = (r + 31) / 32 * 32;
  p.kc = 2 * p.rp;
  const int n_tiles = (d + tb::kBlockN - 1) / tb::kBlockN;
  const int64_t m_blocks = (n_rows + 2 * tb::kBlockM - 1) / (2 * tb::kBlockM);
  int c = 1;
  while (m_blocks * c < sms && c < n_tiles) c *= 2;         // enough units to fill the pairs
  if (c > n_tiles) c = n_tiles;
  const int tpc = (n_tiles + c - 1) / c;
  p.n_chunks = (n_tiles + tpc - 1) / tpc;
  p.n_parts = 2 * p.n_chunks;
  auto up = [](size_t v) { return (v + 255) / 256 * 256; };
  p.off_cb = 0;
  p.off_ub = up(static_cast<size_t>(n_rows) * p.kc * 2);
  p.off_ss = p.off_ub + up(static_cast<size_t>(d) * p.kc * 2);
  p.bytes = p.off_ss + up(static_cast<size_t>(p.n_parts) * n_rows * 4);
  return p;
}
}  // namespace

// steps 2 + 3 of K2 on the tensor cores in one cluster kernel (tb_cluster_rows.cuh): rank r (padded to rp) removed with
// the coefficients in cb, whose layout is [C_hi(c_rp) | C_lo(c_rp)], c_rp >= rp
static bool apply_cluster_supported(int d) { return d % tb::kCrW == 0 && d / tb::kCrW <= tb::kCrMaxCta; }
static int launch_apply_cluster(const void* x, const __nv_bfloat16* cb, int c_rp, const void* ut, int r, int rp, int64_t n_rows,
                                int d, const float* w_norm, float eps, void* out_x, void* out_h, cudaStream_t s, int sms) {
  tb::CaArgs a;
  a.n_rows = static_cast<int>(n_rows);
  a.d = d;
  a.rp = rp;
  a.c_rp = c_rp;
  a.n_cta = d / tb::kCrW;
  a.n_tiles = static_cast<int>((n_rows + tb::kCrRows - 1) / tb::kCrRows);
  a.w_norm = w_norm;
  a.eps = eps;
  a.want_x = out_x ? 1 : 0;
  a.want_h = out_h ? 1 : 0;
  CUtensorMap tmX, tmC, tmU, tmOX, tmOH;
  int rc = make_tmap(&tmX, x, n_rows, d, tb::kCrRows);
  if (rc != TB_OK) return rc;
  rc = make_tmap(&tmC, cb, n_rows, 2 * c_rp, tb::kCrRows);
  if (rc != TB_OK) return rc;
  rc = make_tmap(&tmU, ut, r, d, rp);            // rows r .. rp - 1 of the box: zero-filled
  if (rc != TB_OK) return rc;
  rc = make_tmap(&tmOX, out_x ? out_x : out_h, n_rows, d, tb::kCrRows);
  if (rc != TB_OK) return rc;
  rc = make_tmap(&tmOH, out_h ? out_h : out_x, n_rows, d, tb::kCrRows);
  if (rc != TB_OK) return rc;
  auto kern = tb::tb_lowrank_apply_cluster_kernel;
  TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tb::kCaSmemBytes));
  if (a.n_cta > 8) TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.blockDim = dim3(tb::kCrThreads, 1, 1);
  cfg.dynamicSmemBytes = tb::kCaSmemBytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = static_cast<unsigned>(a.n_cta);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  static int cached[64][tb::kCrMaxCta + 1];
  int dev = 0;
  TB_CUDA(cudaGetDevice(&dev));
  if (cached[dev][a.n_cta] == 0) {
    cfg.gridDim = dim3(static_cast<unsigned>((sms / a.n_cta) * a.n_cta), 1, 1);
    int n = 0;
    TB_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
    if (n < 1) return fail(TB_ERR_CUDA, "no cluster of %d CTAs fits on this device", a.n_cta);
    cached[dev][a.n_cta] = n;
  }
  int clusters = cached[dev][a.n_cta];
  if (clusters > a.n_tiles) clusters = a.n_tiles;
  cfg.gridDim = dim3(static_cast<unsigned>(clusters * a.n_cta), 1, 1);
  TB_CUDA(cudaLaunchKernelEx(&cfg, kern, tmX, tmC, tmU, tmOX, tmOH, a));
  return TB_OK;
}

int64_t tb_lowrank_tc_scratch_bytes(int64_t n_rows, int d, int r) {
  DevInfo di;
  if (device_info(&di) != TB_OK || n_rows < 0 || d <= 0 || r < 1) return -1;
  return static_cast<int64_t>(proj_tc_plan(n_rows, d, r, di.sms).bytes);
}

int tb_lowrank_remove_tc(const void* x, const void* ut, int r, int64_t n_rows, int d, const float* w_norm,
                         float eps, void* scratch, int64_t scratch_bytes, void* out_x, void* out_h,
                         void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (r < 1 || r > tb::kProjMaxRank) return fail(TB_ERR_INVALID, "rank %d outside [1, %d]", r, tb::kProjMaxRank);
  rc = check_gemm_operands(x, ut, n_rows, r, d);
  if (rc != TB_OK) return rc;
  if (!out_x && !out_h) return fail(TB_ERR_INVALID, "no output requested");
  if (out_h && !w_norm) return fail(TB_ERR_INVALID, "out_h needs w_norm");
  const int nv = (d / 8 + tb::kRowGroupThreads - 1) / tb::kRowGroupThreads;
  if (nv > 8) return fail(TB_ERR_INVALID, "d=%d too large for the row kernels (max 8192)", d);
  if (n_rows == 0) return TB_OK;
  const ProjTcPlan pl = proj_tc_plan(n_rows, d, r, di.sms);
  if (!scratch || scratch_bytes < static_cast<int64_t>(pl.bytes))
    return fail(TB_ERR_INVALID, "scratch of %lld bytes needed (tb_lowrank_tc_scratch_bytes), got %lld",
                static_cast<long long>(pl.bytes), static_cast<long long>(scratch_bytes));
  if (reinterpret_cast<uintptr_t>(scratch) % 256) return fail(TB_ERR_INVALID, "scratch must be 256-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* sc = static_cast<uint8_t*>(scratch);
  __nv_bfloat16* cb = reinterpret_cast<__nv_bfloat16*>(sc + pl.off_cb);
  __nv_bfloat16* ub = reinterpret_cast<__nv_bfloat16*>(sc + pl.off_ub);
  float* ss = reinterpret_cast<float*>(sc + pl.off_ss);
  // step 1: C = x U, split into bf16 head + remainder
  {
    tb::GemmArgs g = make_args(n_rows, r, d, 1, 2);
    tb::CoeffSplitEpi::Params ep{cb, pl.rp};
    rc = launch_gemm<tb::CoeffSplitEpi, 2>(x, ut, g, ep, s, di.sms);
    if (rc != TB_OK) return rc;
  }
  // steps 2 + 3 in one kernel when the rows split into 256-column slices for a cluster (d = 3584: 14 CTAs): x is
  // read once more and h written once, the sums of squares cross the cluster through distributed shared memory
  // (tb_cluster_rows.cuh). TB_K2_THREE_PASS=1 keeps the two separate passes below.
  {
    const char* tp = getenv("TB_K2_THREE_PASS");
    if (!(tp && atoi(tp) != 0) && apply_cluster_supported(d))
      return launch_apply_cluster(x, cb, pl.rp, ut, r, pl.rp, n_rows, d, w_norm, eps, out_x, out_h, s, di.sms);
  }
  // B operand of step 2
  {
    const int64_t n = static_cast<int64_t>(d) * pl.kc;
    tb::tb_lowrank_ub_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, s>>>(
        static_cast<const __nv_bfloat16*>(ut), r, pl.rp, d, ub);
    TB_CUDA(cudaGetLastError());
  }
  // step 2: x' = bf16(x - C U^T) and the rows' sums of squares
  __nv_bfloat16* xp = static_cast<__nv_bfloat16*>(out_x ? out_x : out_h);
  {
    tb::GemmArgs g = make_args(n_rows, d, pl.kc, pl.n_chunks, 2);
    if (g.n_chunks != pl.n_chunks) return fail(TB_ERR_INVALID, "internal: chunk plan mismatch");
    tb::ProjApplyEpi::Params ep{static_cast<const __nv_bfloat16*>(x), xp, ss};
    rc = launch_gemm<tb::ProjApplyEpi, 2>(cb, ub, g, ep, s, di.sms);
    if (rc != TB_OK) return rc;
  }
  // step 3: the fused final RMSNorm
  if (out_h) {
    const int64_t blocks = (n_rows + tb::kRowGroups - 1) / tb::kRowGroups;
    const int64_t max_grid = static_cast<int64_t>(di.sms) * 8;
    const unsigned grid = static_cast<unsigned>(blocks < max_grid ? blocks : max_grid);
    __nv_bfloat16* hp = static_cast<__nv_bfloat16*>(out_h);
    const int n = static_cast<int>(n_rows);
    if (nv == 1) tb::tb_scale_rows_kernel<1><<<grid, tb::kAblThreads, 0, s>>>(xp, ss, pl.n_parts, w_norm, n, d, eps, hp);
    else if (nv == 2) tb::tb_scale_rows_kernel<2><<<grid, tb::kAblThreads, 0, s>>>(xp, ss, pl.n_parts, w_norm, n, d, eps, hp);
    else if (nv <= 4) tb::tb_scale_rows_kernel<4><<<grid, tb::kAblThreads, 0, s>>>(xp, ss, pl.n_parts, w_norm, n, d, eps, hp);
    else tb::tb_scale_rows_kernel<8><<<grid, tb::kAblThreads, 0, s>>>(xp, ss, pl.n_parts, w_norm, n, d, eps, hp);
    TB_CUDA(cudaGetLastError());
  }
  return TB_OK;
}

namespace {
// generic 2-D bf16 tensor map: box = [box_rows, 64 elements], 128-byte swizzle
int make_tmap_box(CUtensorMap* map, const void* base, int64_t rows, int cols, int box_rows) {
  return make_tmap(map, base, rows, cols, box_rows);
}
}  // namespace

int tb_lowrank_remove_tc_multi(const void* x, const void* ut, const int32_t* ranks_host, int n_ranks, int64_t n_rows, int d,
                               const float* w_norm, float eps, void* scratch, int64_t scratch_bytes, void* out_x, void* out_h,
                               int64_t out_stride, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!ranks_host || n_ranks < 1 || n_ranks > 16) return fail(TB_ERR_INVALID, "1 .. 16 ranks");
  for (int k = 0; k < n_ranks; ++k)
    if (ranks_host[k] < 1 || ranks_host[k] > tb::kProjMaxRank || (k > 0 && ranks_host[k] <= ranks_host[k - 1]))
      return fail(TB_ERR_INVALID, "ranks must be strictly ascending within [1, %d]", tb::kProjMaxRank);
  const int r_max = ranks_host[n_ranks - 1];
  rc = check_gemm_operands(x, ut, n_rows, r_max, d);
  if (rc != TB_OK) return rc;
  if (!out_x && !out_h) return fail(TB_ERR_INVALID, "no output requested");
  if (out_h && !w_norm) return fail(TB_ERR_INVALID, "out_h needs w_norm");
  if (!apply_cluster_supported(d)) return fail(TB_ERR_INVALID, "d=%d: needs d %% 256 == 0 and d <= %d", d, tb::kCrW * tb::kCrMaxCta);
  if (n_ranks > 1 && out_stride < n_rows * static_cast<int64_t>(d)) return fail(TB_ERR_INVALID, "out_stride smaller than one output");
  if (n_rows == 0) return TB_OK;
  const ProjTcPlan pl = proj_tc_plan(n_rows, d, r_max, di.sms);
  if (!scratch || scratch_bytes < static_cast<int64_t>(pl.bytes))
    return fail(TB_ERR_INVALID, "scratch of %lld bytes needed (tb_lowrank_tc_scratch_bytes of the largest rank), got %lld",
                static_cast<long long>(pl.bytes), static_cast<long long>(scratch_bytes));
  if (reinterpret_cast<uintptr_t>(scratch) % 256) return fail(TB_ERR_INVALID, "scratch must be 256-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  __nv_bfloat16* cb = reinterpret_cast<__nv_bfloat16*>(static_cast<uint8_t*>(scratch) + pl.off_cb);
  {   // the coefficients of all r_max directions, once: rank r uses the first r of them
    tb::GemmArgs g = make_args(n_rows, r_max, d, 1, 2);
    tb::CoeffSplitEpi::Params ep{cb, pl.rp};
    rc = launch_gemm<tb::CoeffSplitEpi, 2>(x, ut, g, ep, s, di.sms);
    if (rc != TB_OK) return rc;
  }
  for (int k = 0; k < n_ranks; ++k) {
    const int r = ranks_host[k], rp = (r + 31) / 32 * 32;
    __nv_bfloat16* ox = out_x ? static_cast<__nv_bfloat16*>(out_x) + k * out_stride : nullptr;
    __nv_bfloat16* oh = out_h ? static_cast<__nv_bfloat16*>(out_h) + k * out_stride : nullptr;
    rc = launch_apply_cluster(x, cb, pl.rp, ut, r, rp, n_rows, d, w_norm, eps, ox, oh, s, di.sms);
    if (rc != TB_OK) return rc;
  }
  return TB_OK;
}

int tb_lowrank_fused_supported(int d, int r) {
  DevInfo di;
  if (device_info(&di) != TB_OK) return 0;
  return (r >= 1 && r <= tb::kCrMaxRp && d > 0 && d % tb::kCrW == 0 && d / tb::kCrW <= tb::kCrMaxCta) ? 1 : 0;
}

int tb_lowrank_remove_fused(const void* x, const void* ut, int r, int64_t n_rows, int d, const float* w_norm,
                            float eps, void* out_x, void* out_h, float* dbg_c, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (r < 1 || r > tb::kCrMaxRp) return fail(TB_ERR_INVALID, "rank %d outside [1, %d]", r, tb::kCrMaxRp);
  rc = check_gemm_operands(x, ut, n_rows, r, d);
  if (rc != TB_OK) return rc;
  if (d % tb::kCrW != 0 || d / tb::kCrW > tb::kCrMaxCta)
    return fail(TB_ERR_INVALID, "d=%d: the one-pass kernel wants a multiple of %d, at most %d (use tb_lowrank_remove_tc)", d,
                tb::kCrW, tb::kCrW * tb::kCrMaxCta);
  if (!out_x && !out_h) return fail(TB_ERR_INVALID, "no output requested");
  if (out_h && !w_norm) return fail(TB_ERR_INVALID, "out_h needs w_norm");
  if (out_x && (reinterpret_cast<uintptr_t>(out_x) & 15)) return fail(TB_ERR_INVALID, "out_x must be 16-byte aligned");
  if (out_h && (reinterpret_cast<uintptr_t>(out_h) & 15)) return fail(TB_ERR_INVALID, "out_h must be 16-byte aligned");
  if (n_rows == 0) return TB_OK;
  tb::CrArgs a;
  a.n_rows = static_cast<int>(n_rows);
  a.d = d;
  a.rounds = (r + 31) / 32;
  a.n_cta = d / tb::kCrW;
  a.rows_own = (tb::kCrRows + a.n_cta - 1) / a.n_cta;
  a.n_tiles = static_cast<int>((n_rows + tb::kCrRows - 1) / tb::kCrRows);
  a.w_norm = w_norm;
  a.eps = eps;
  a.want_x = out_x ? 1 : 0;
  a.want_h = out_h ? 1 : 0;
  a.dbg_c = dbg_c;
  CUtensorMap tmX, tmU, tmOX, tmOH;
  rc = make_tmap_box(&tmX, x, n_rows, d, tb::kCrRows);
  if (rc != TB_OK) return rc;
  rc = make_tmap_box(&tmU, ut, r, d, 32 * a.rounds);      // rows past r are zero-filled
  if (rc != TB_OK) return rc;
  rc = make_tmap_box(&tmOX, out_x ? out_x : out_h, n_rows, d, tb::kCrRows);
  if (rc != TB_OK) return rc;
  rc = make_tmap_box(&tmOH, out_h ? out_h : out_x, n_rows, d, tb::kCrRows);
  if (rc != TB_OK) return rc;
  auto kern = tb::tb_lowrank_cluster_kernel;
  TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tb::kCrSmemBytes));
  if (a.n_cta > 8) TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.blockDim = dim3(tb::kCrThreads, 1, 1);
  cfg.dynamicSmemBytes = tb::kCrSmemBytes;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = static_cast<unsigned>(a.n_cta);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  static int cached[64][tb::kCrMaxCta + 1];
  int dev = 0;
  TB_CUDA(cudaGetDevice(&dev));
  if (cached[dev][a.n_cta] == 0) {
    cfg.gridDim = dim3(static_cast<unsigned>((di.sms / a.n_cta) * a.n_cta), 1, 1);
    int n = 0;
    TB_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
    if (n < 1) return fail(TB_ERR_CUDA, "no cluster of %d CTAs fits on this device", a.n_cta);
    cached[dev][a.n_cta] = n;
  }
  int clusters = cached[dev][a.n_cta];
  if (clusters > a.n_tiles) clusters = a.n_tiles;
  const char* mc = getenv("TB_MAX_CLUSTERS");
  if (mc && atoi(mc) > 0 && clusters > atoi(mc)) clusters = atoi(mc);
  cfg.gridDim = dim3(static_cast<unsigned>(clusters * a.n_cta), 1, 1);
  TB_CUDA(cudaLaunchKernelEx(&cfg, kern, tmX, tmU, tmOX, tmOH, a));
  return TB_OK;
}

int tb_rmsnorm(const void* x, const float* w_norm, float eps, int64_t n_rows, int d, void* out_h,
               void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!x || !w_norm || !out_h) return fail(TB_ERR_INVALID, "null pointer");
  if (d <= 0 || d % 8 != 0) return fail(TB_ERR_INVALID, "d=%d must be a positive multiple of 8", d);
  const int nv = (d / 8 + tb::kRowGroupThreads - 1) / tb::kRowGroupThreads;
  if (nv > 8) return fail(TB_ERR_INVALID, "d=%d too large for the row kernels (max 8192)", d);
  if (n_rows == 0) return TB_OK;
  const int64_t blocks = (n_rows + tb::kRowGroups - 1) / tb::kRowGroups;
  const int64_t max_grid = static_cast<int64_t>(di.sms) * 8;
  const unsigned grid = static_cast<unsigned>(blocks < max_grid ? blocks : max_grid);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* hp = static_cast<__nv_bfloat16*>(out_h);
  if (nv == 1) tb::tb_rmsnorm_kernel<1><<<grid, tb::kAblThreads, 0, s>>>(xp, w_norm, (int)n_rows, d, eps, hp);
  else if (nv == 2) tb::tb_rmsnorm_kernel<2><<<grid, tb::kAblThreads, 0, s>>>(xp, w_norm, (int)n_rows, d, eps, hp);
  else if (nv <= 4) tb::tb_rmsnorm_kernel<4><<<grid, tb::kAblThreads, 0, s>>>(xp, w_norm, (int)n_rows, d, eps, hp);
  else tb::tb_rmsnorm_kernel<8><<<grid, tb::kAblThreads, 0, s>>>(xp, w_norm, (int)n_rows, d, eps, hp);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_lens_partials(const void* h, const void* w_u, int64_t n_rows, int d, int V, int vocab_offset,
                     float softcap, const int32_t* secret_id, int n_chunks, int cta_pair, float* pm,
                     float* ps, float* psec, float* ptv, int32_t* pti, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  rc = check_gemm_operands(h, w_u, n_rows, V, d);
  if (rc != TB_OK) return rc;
  if (!pm || !ps || !psec || !ptv || !pti) return fail(TB_ERR_INVALID, "null partial buffer");
  ClusterMode cm;
  rc = cluster_mode(cta_pair, &cm);
  if (rc != TB_OK) return rc;
  tb::GemmArgs g = make_args(n_rows, V, d, n_chunks, cm.group, cm.R, cm.C);
  if (g.n_chunks != n_chunks)
    return fail(TB_ERR_INVALID, "n_chunks=%d does not divide %d tiles into non-empty chunks (use %d)",
                n_chunks, g.n_tiles, g.n_chunks);
  if (n_chunks % cm.C)
    return fail(TB_ERR_INVALID, "n_chunks=%d must be a multiple of %d in cluster mode %d", n_chunks, cm.C, cta_pair);
  tb::LensEpi::Params ep;
  ep.secret_id = secret_id;
  ep.vocab_offset = vocab_offset;
  ep.softcap = softcap;
  ep.pm = pm;
  ep.ps = ps;
  ep.psec = psec;
  ep.ptv = ptv;
  ep.pti = pti;
  return launch_gemm_mode<tb::LensEpi>(cm, h, w_u, g, ep, static_cast<cudaStream_t>(stream), di.sms);
}

int tb_lens_dense_partials(const void* h, const void* w_u, int64_t n_rows, int d, int V, float softcap,
                           const int32_t* secret_id, int n_chunks, int cta_pair, float* pm, float* ps, float* psec,
                           float* ptv, int32_t* pti, float* logits_out, int layout, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  rc = check_gemm_operands(h, w_u, n_rows, V, d);
  if (rc != TB_OK) return rc;
  const bool raw = layout == TB_LOGITS_PROMPT_TILES_RAW;
  if (!logits_out || (!raw && (!pm || !ps || !psec || !ptv || !pti))) return fail(TB_ERR_INVALID, "null buffer");
  if (layout != TB_LOGITS_ROW_MAJOR && layout != TB_LOGITS_PROMPT_TILES && !raw) return fail(TB_ERR_INVALID, "layout=%d", layout);
  if (layout != TB_LOGITS_ROW_MAJOR && (V % 32 != 0 || n_rows % tb::kLlSlots != 0))
    return fail(TB_ERR_INVALID, "prompt tiles need V %% 32 == 0 and whole prompts of %d rows (V=%d, n_rows=%lld)", tb::kLlSlots, V,
                static_cast<long long>(n_rows));
  if (reinterpret_cast<uintptr_t>(logits_out) & 15) return fail(TB_ERR_INVALID, "logits_out must be 16-byte aligned");
  ClusterMode cm;
  rc = cluster_mode(cta_pair, &cm);
  if (rc != TB_OK) return rc;
  tb::GemmArgs g = make_args(n_rows, V, d, n_chunks, cm.group, cm.R, cm.C);
  if (g.n_chunks != n_chunks)
    return fail(TB_ERR_INVALID, "n_chunks=%d does not divide %d tiles into non-empty chunks (use %d)", n_chunks, g.n_tiles,
                g.n_chunks);
  if (n_chunks % cm.C)
    return fail(TB_ERR_INVALID, "n_chunks=%d must be a multiple of %d in cluster mode %d", n_chunks, cm.C, cta_pair);
  tb::LensDenseEpi::Params ep;
  memset(&ep.tm_out, 0, sizeof(ep.tm_out));
  static const bool direct = getenv("TB_DENSE_DIRECT_STORE") != nullptr;   // experiments: plain stores from the row owners
  ep.tma = raw ? 3 : layout == TB_LOGITS_PROMPT_TILES ? 2 : (V % 4 == 0 && !direct) ? 1 : 0;   // a tensor map wants a 16-byte row pitch
  if (ep.tma) {                            // box = one warp's [32 rows x 32 columns] fp32, 128-byte swizzle
    EncodeTiledFn enc;
    rc = get_encode_fn(&enc);
    if (rc != TB_OK) return rc;
    const cuuint64_t n_ct = static_cast<cuuint64_t>(V / 32);
    // row-major [n_rows, V]; or (column in tile, row in prompt, column tile, prompt) over [P][V / 32][128][32]
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(V), static_cast<cuuint64_t>(n_rows), 1, 1};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(V) * 4, 0, 0};
    cuuint32_t box[4] = {32, 32, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (ep.tma >= 2) {
      dims[0] = 32; dims[1] = tb::kLlSlots; dims[2] = n_ct; dims[3] = static_cast<cuuint64_t>(n_rows / tb::kLlSlots);
      strides[0] = 128; strides[1] = 128ull * tb::kLlSlots; strides[2] = strides[1] * n_ct;
    }
    CUresult r = enc(&ep.tm_out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, ep.tma >= 2 ? 4 : 2, logits_out, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
      return fail(TB_ERR_CUDA, "cuTensorMapEncodeTiled (logits) failed with CUresult %d (rows=%lld V=%d)", static_cast<int>(r),
                  static_cast<long long>(n_rows), V);
  }
  ep.lens.secret_id = secret_id;
  ep.lens.vocab_offset = 0;
  ep.lens.softcap = softcap;
  ep.lens.pm = pm; ep.lens.ps = ps; ep.lens.psec = psec; ep.lens.ptv = ptv; ep.lens.pti = pti;
  ep.out = logits_out;
  return launch_gemm_mode<tb::LensDenseEpi>(cm, h, w_u, g, ep, static_cast<cudaStream_t>(stream), di.sms);
}

int tb_logits_ll_colsum(const float* logits, const float* lse, const int32_t* cur_id, const int32_t* prev_id,
                        int64_t n_rows, int V, float* acc, int layout, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!logits || !lse || !acc) return fail(TB_ERR_INVALID, "null pointer");
  if (prev_id && !cur_id) return fail(TB_ERR_INVALID, "prev_id given without cur_id");
  if (n_rows < 0 || V < 1) return fail(TB_ERR_INVALID, "bad shape");
  if (layout != TB_LOGITS_ROW_MAJOR && layout != TB_LOGITS_PROMPT_TILES) return fail(TB_ERR_INVALID, "layout=%d", layout);
  if (n_rows == 0) return TB_OK;
  const int64_t n_prompts = (n_rows + tb::kLlSlots - 1) / tb::kLlSlots;
  if (n_prompts > 65535) return fail(TB_ERR_INVALID, "at most 65535 prompts per call (%lld)", static_cast<long long>(n_prompts));
  if (layout == TB_LOGITS_PROMPT_TILES) {
    if (V % 32 != 0 || n_rows % tb::kLlSlots != 0)
      return fail(TB_ERR_INVALID, "prompt tiles need V %% 32 == 0 and whole prompts of %d rows (V=%d, n_rows=%lld)", tb::kLlSlots, V,
                  static_cast<long long>(n_rows));
    if ((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(acc)) & 15)
      return fail(TB_ERR_INVALID, "logits and acc must be 16-byte aligned");
    const int n_ct = V / 32;
    dim3 tgrid(static_cast<unsigned>((n_ct + tb::kLlTilesPerBlock - 1) / tb::kLlTilesPerBlock), static_cast<unsigned>(n_prompts), 1);
    tb::tb_logits_ll_colsum_tiled_kernel<<<tgrid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const float4*>(logits), lse, cur_id, prev_id, n_ct, V, acc);
    TB_CUDA(cudaGetLastError());
    return TB_OK;
  }
  dim3 grid(static_cast<unsigned>((V + 1023) / 1024), static_cast<unsigned>(n_prompts), 1);     // 4 columns per thread
  tb::tb_logits_ll_colsum_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(logits, lse, cur_id, prev_id, n_rows, V, acc);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_scale_cols(const void* x, const float* w, int64_t n_rows, int d, void* out, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!x || !w || !out) return fail(TB_ERR_INVALID, "null pointer");
  if (n_rows < 0 || d <= 0 || d % 8 != 0) return fail(TB_ERR_INVALID, "bad shape (d must be a multiple of 8)");
  if (n_rows == 0) return TB_OK;
  const int64_t n_vec = n_rows * (d / 8);
  const int64_t want = (n_vec + 255) / 256, cap = static_cast<int64_t>(di.sms) * 16;
  tb::tb_scale_cols_kernel<<<static_cast<unsigned>(want < cap ? want : cap), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), w, n_vec, d / 8, static_cast<__nv_bfloat16*>(out));
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_rank_acts_dense(const int64_t* row_start, int cap, const int32_t* sel_count, const int32_t* sel_key,
                       const float* sel_val, int64_t n_rows, int a_cols, int lo_off, void* out_bf16, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!sel_count || !sel_key || !sel_val || !out_bf16) return fail(TB_ERR_INVALID, "null pointer");
  if (!row_start && cap <= 0) return fail(TB_ERR_INVALID, "fixed-stride lists need cap > 0");
  if (a_cols % 64 != 0 || a_cols < tb::kRankedMax || (lo_off != 0 && (lo_off < tb::kRankedMax || lo_off + tb::kRankedMax > a_cols)))
    return fail(TB_ERR_INVALID, "a_cols=%d lo_off=%d: need a multiple of 64 >= %d, remainders inside", a_cols, lo_off, tb::kRankedMax);
  if (n_rows < 0) return fail(TB_ERR_INVALID, "bad shape");
  if (n_rows == 0) return TB_OK;
  tb::tb_rank_acts_dense_kernel<<<static_cast<unsigned>((n_rows + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      row_start, cap, sel_count, sel_key, sel_val, n_rows, a_cols, lo_off, static_cast<__nv_bfloat16*>(out_bf16));
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

// 16 epilogue warps (4 partials per chunk and row) for budgets of up to eight slabs per tile, 8 (2 partials) beyond;
// TB_RU_CONFIG=8 / 16 forces one of them (experiments)
static int ru_variant(int budget, int lo_off) {
  const char* force = getenv("TB_RU_CONFIG");
  if (force && atoi(force) == 8) return 8;
  if (force && atoi(force) == 16) return 16;
  const int slabs = (budget + tb::kRuSlabK - 1) / tb::kRuSlabK * (lo_off > 0 ? 2 : 1);
  return slabs <= 8 ? 16 : 8;      // measured at 65,536 rows: 256: 27 vs 43 ms, 512: 47 vs 48, 1024: 84 vs 65
}

int tb_lens_rank_update_parts(int budget, int lo_off) { return ru_variant(budget, lo_off) / 4; }

int tb_lens_rank_update_partials(const void* acts_bf16, int a_cols, int lo_off, const int32_t* rank_ids, int rank_stride,
                                 const void* d_bf16, int S, const float* l0_tiles, const float* rstd, int64_t n_rows, int V, int budget,
                                 float softcap, const int32_t* secret_id, int n_chunks, float* pm, float* ps, float* psec,
                                 float* ptv, int32_t* pti, float* logits_out_tiles, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!acts_bf16 || !rank_ids || !d_bf16 || !l0_tiles || !rstd || !pm || !ps || !psec || !ptv || !pti)
    return fail(TB_ERR_INVALID, "null pointer");
  if (n_rows <= 0 || n_rows % tb::kLlSlots != 0 || V % 32 != 0 || V <= 0)
    return fail(TB_ERR_INVALID, "whole prompts of %d rows and V %% 32 == 0 (n_rows=%lld, V=%d)", tb::kLlSlots,
                static_cast<long long>(n_rows), V);
  if (budget < 1 || budget > tb::kRankedMax || rank_stride < budget) return fail(TB_ERR_INVALID, "budget %d / rank_stride %d", budget, rank_stride);
  if (a_cols % 64 != 0 || a_cols < tb::kRankedMax || (lo_off != 0 && lo_off + tb::kRankedMax > a_cols))
    return fail(TB_ERR_INVALID, "a_cols=%d lo_off=%d", a_cols, lo_off);
  if ((reinterpret_cast<uintptr_t>(acts_bf16) | reinterpret_cast<uintptr_t>(d_bf16) | reinterpret_cast<uintptr_t>(l0_tiles)) & 15)
    return fail(TB_ERR_INVALID, "operands must be 16-byte aligned");
  tb::RankUpdateArgs a;
  a.n_prompts = static_cast<int>(n_rows / tb::kLlSlots);
  a.V = V;
  a.n_tiles = (V + 255) / 256;
  if (n_chunks < 1 || n_chunks > a.n_tiles) return fail(TB_ERR_INVALID, "n_chunks=%d", n_chunks);
  a.tiles_per_chunk = (a.n_tiles + n_chunks - 1) / n_chunks;
  if ((a.n_tiles + a.tiles_per_chunk - 1) / a.tiles_per_chunk != n_chunks)
    return fail(TB_ERR_INVALID, "n_chunks=%d does not divide %d tiles into non-empty chunks", n_chunks, a.n_tiles);
  a.n_chunks = n_chunks;
  a.m = budget;
  a.n_slabs = (budget + tb::kRuSlabK - 1) / tb::kRuSlabK;
  a.parts = lo_off > 0 ? 2 : 1;
  a.a_lo_off = lo_off;
  a.rank_ids = rank_ids;
  a.rank_stride = rank_stride;
  a.D = static_cast<const __nv_bfloat16*>(d_bf16);
  a.S = S;
  if (S <= 0) return fail(TB_ERR_INVALID, "S=%d", S);
  a.rstd = rstd;
  {
    const char* dbg = getenv("TB_RU_DEBUG");
    a.debug = dbg ? atoi(dbg) : 0;
  }
  tb::LensEpi::Params lens;
  lens.secret_id = secret_id;
  lens.vocab_offset = 0;
  lens.softcap = softcap;
  lens.pm = pm; lens.ps = ps; lens.psec = psec; lens.ptv = ptv; lens.pti = pti;
  CUtensorMap tmA, tmL0;
  rc = make_tmap(&tmA, acts_bf16, n_rows, a_cols, 128);
  if (rc != TB_OK) return rc;
  rc = make_tile_map(&tmL0, l0_tiles, n_rows, V);
  if (rc != TB_OK) return rc;
  CUtensorMap tmOut = tmL0;
  a.store = logits_out_tiles ? 1 : 0;
  if (logits_out_tiles) {
    if (reinterpret_cast<uintptr_t>(logits_out_tiles) & 15) return fail(TB_ERR_INVALID, "logits_out_tiles must be 16-byte aligned");
    rc = make_tile_map(&tmOut, logits_out_tiles, n_rows, V);
    if (rc != TB_OK) return rc;
  }
  const int units = a.n_prompts * a.n_chunks;
  const unsigned grid = static_cast<unsigned>(units < di.sms ? units : di.sms);
  const int variant = ru_variant(budget, lo_off);
  if (variant == 16) {
    auto k = tb::tb_rank_update_lens_kernel<2, 2, 16>;
    using Cfg = tb::RuCfg<2, 2, 16>;
    TB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    k<<<grid, Cfg::kThreads, Cfg::kSmemBytes, static_cast<cudaStream_t>(stream)>>>(tmA, tmL0, tmOut, a, lens);
  } else {
    auto k = tb::tb_rank_update_lens_kernel<3, 2, 8>;
    using Cfg = tb::RuCfg<3, 2, 8>;
    TB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    k<<<grid, Cfg::kThreads, Cfg::kSmemBytes, static_cast<cudaStream_t>(stream)>>>(tmA, tmL0, tmOut, a, lens);
  }
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_probs_from_logits(float* logits, const float* lse, int64_t n_rows, int V, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!logits || !lse) return fail(TB_ERR_INVALID, "null pointer");
  if (n_rows < 0 || V < 1) return fail(TB_ERR_INVALID, "bad shape");
  if (n_rows == 0) return TB_OK;
  const int64_t total = n_rows * static_cast<int64_t>(V);
  const int64_t want = (total + 255) / 256, cap = static_cast<int64_t>(di.sms) * 16;
  tb::tb_probs_from_logits_kernel<<<static_cast<unsigned>(want < cap ? want : cap), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      logits, lse, n_rows, V);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_lens_merge(const float* pm, const float* ps, const float* psec, const float* ptv, const int32_t* pti,
                  int n_parts, int64_t n_rows, int finalize, float* om, float* os, float* osec,
                  float* otv, int32_t* oti, float* lse, float* secret_prob, float* top_prob,
                  int32_t* top_id, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!pm || !ps || !psec || !ptv || !pti) return fail(TB_ERR_INVALID, "null partial buffer");
  if (n_parts < 1) return fail(TB_ERR_INVALID, "n_parts=%d", n_parts);
  if (finalize) {
    if (!lse || !top_prob || !top_id) return fail(TB_ERR_INVALID, "finalize needs lse, top_prob, top_id");
  } else if (!om || !os || !osec || !otv || !oti) {
    return fail(TB_ERR_INVALID, "merge needs om, os, osec, otv, oti");
  }
  if (n_rows == 0) return TB_OK;
  tb::MergeArgs a;
  a.pm = pm; a.ps = ps; a.psec = psec; a.ptv = ptv; a.pti = pti;
  a.n_parts = n_parts;
  a.n_rows = static_cast<int>(n_rows);
  a.finalize = finalize;
  a.om = om; a.os = os; a.osec = osec; a.otv = otv; a.oti = oti;
  a.lse = lse; a.secret_prob = secret_prob; a.top_prob = top_prob; a.top_id = top_id;
  const unsigned grid = static_cast<unsigned>((n_rows + 127) / 128);
  tb::tb_lens_merge_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(a);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_ll_topk_partials(const void* h, const void* w_u, int64_t n_rows, int d, int V, int vocab_offset,
                        float softcap, const float* lse, const int32_t* cur_id, const int32_t* prev_id,
                        int n_chunks, int cta_pair, float* part_val, int32_t* part_id, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  rc = check_gemm_operands(w_u, h, V, static_cast<int>(n_rows), d);
  if (rc != TB_OK) return rc;
  if (n_rows > 0x7fffff00ll) return fail(TB_ERR_INVALID, "too many rows (%lld)", (long long)n_rows);
  if (!lse || !part_val || !part_id) return fail(TB_ERR_INVALID, "null lse / part_val / part_id");
  if (prev_id && !cur_id) return fail(TB_ERR_INVALID, "prev_id given without cur_id");
  ClusterMode cm;
  rc = cluster_mode(cta_pair, &cm);
  if (rc != TB_OK) return rc;
  const int G = cm.group;
  tb::GemmArgs g = make_args_sweep_a(V, n_rows, d, n_chunks, G, cm.R, cm.C);
  if (g.n_chunks != n_chunks)
    return fail(TB_ERR_INVALID, "n_chunks=%d does not divide %d tiles into non-empty chunks (use %d)",
                n_chunks, g.n_tiles, g.n_chunks);
  if (n_chunks % cm.C)
    return fail(TB_ERR_INVALID, "n_chunks=%d must be a multiple of %d in cluster mode %d", n_chunks, cm.C, cta_pair);
  if (g.tiles_per_chunk * tb::kBlockM * G > tb::ColSumEpi::kMaxChunkIds)
    return fail(TB_ERR_INVALID, "a chunk spans %d vocabulary ids, more than the %d the mask bitmap holds: use "
                "n_chunks >= %d", g.tiles_per_chunk * tb::kBlockM * G, tb::ColSumEpi::kMaxChunkIds,
                (V + tb::ColSumEpi::kMaxChunkIds - 1) / tb::ColSumEpi::kMaxChunkIds);
  tb::ColSumEpi::Params ep;
  ep.lse = lse;
  ep.cur_id = cur_id;
  ep.prev_id = prev_id;
  ep.vocab_offset = vocab_offset;
  ep.softcap = softcap;
  ep.n_prompts = static_cast<int>((n_rows + tb::ColSumEpi::kSlots - 1) / tb::ColSumEpi::kSlots);
  ep.out_val = part_val;
  ep.out_id = part_id;
  ep.group = G;
  return launch_gemm_mode<tb::ColSumEpi>(cm, w_u, h, g, ep, static_cast<cudaStream_t>(stream), di.sms);
}

int tb_ll_topk_merge(const float* part_val, const int32_t* part_id, int n_parts, int64_t n_prompts, int k,
                     float* out_val, int32_t* out_id, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!part_val || !part_id || !out_val || !out_id) return fail(TB_ERR_INVALID, "null pointer");
  if (n_parts < 1) return fail(TB_ERR_INVALID, "n_parts=%d", n_parts);
  if (k < 1 || k > TB_TOPK) return fail(TB_ERR_INVALID, "k=%d outside [1, %d]", k, TB_TOPK);
  if (n_prompts == 0) return TB_OK;
  const unsigned grid = static_cast<unsigned>((n_prompts + 127) / 128);
  tb::tb_ll_topk_merge_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      part_val, part_id, n_parts, static_cast<int>(n_prompts), k, out_val, out_id);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_probs_ll_colsum(const float* probs, int T, int V, const int32_t* cur_id, const int32_t* prev_id,
                       float* acc, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!probs || !acc) return fail(TB_ERR_INVALID, "null pointer");
  if (T < 0 || T > tb::kColsumMaxT) return fail(TB_ERR_INVALID, "T=%d outside [0, %d]", T, tb::kColsumMaxT);
  if (V < 1) return fail(TB_ERR_INVALID, "V=%d", V);
  if (prev_id && !cur_id) return fail(TB_ERR_INVALID, "prev_id given without cur_id");
  const unsigned grid = static_cast<unsigned>((V + 255) / 256);
  tb::tb_probs_colsum_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(probs, T, V, cur_id, prev_id, acc);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_lens_ll_partials(const void* h, const void* w_u, int64_t n_rows, int d, int V, int vocab_offset,
                        float softcap, const int32_t* secret_id, const float* ref_lse, const float* tau,
                        const int32_t* cur_id, const int32_t* prev_id, int n_chunks, int cta_pair, float* pm,
                        float* ps, float* psec, float* ptv, int32_t* pti, int32_t* cand_count, int32_t* cand_id,
                        int cand_cap, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  rc = check_gemm_operands(h, w_u, n_rows, V, d);
  if (rc != TB_OK) return rc;
  if (!pm || !ps || !psec || !ptv || !pti) return fail(TB_ERR_INVALID, "null partial buffer");
  if (!ref_lse || !tau || !cand_count || !cand_id) return fail(TB_ERR_INVALID, "null ref_lse / tau / candidate buffer");
  if (cand_cap < 1) return fail(TB_ERR_INVALID, "cand_cap=%d", cand_cap);
  if (prev_id && !cur_id) return fail(TB_ERR_INVALID, "prev_id given without cur_id");
  ClusterMode cm;
  rc = cluster_mode(cta_pair, &cm);
  if (rc != TB_OK) return rc;
  tb::GemmArgs g = make_args(n_rows, V, d, n_chunks, cm.group, cm.R, cm.C);
  if (g.n_chunks != n_chunks)
    return fail(TB_ERR_INVALID, "n_chunks=%d does not divide %d tiles into non-empty chunks (use %d)",
                n_chunks, g.n_tiles, g.n_chunks);
  if (n_chunks % cm.C)
    return fail(TB_ERR_INVALID, "n_chunks=%d must be a multiple of %d in cluster mode %d", n_chunks, cm.C, cta_pair);
  tb::LensLLEpi::Params ep;
  ep.lens.secret_id = secret_id;
  ep.lens.vocab_offset = vocab_offset;
  ep.lens.softcap = softcap;
  ep.lens.pm = pm;
  ep.lens.ps = ps;
  ep.lens.psec = psec;
  ep.lens.ptv = ptv;
  ep.lens.pti = pti;
  ep.ref = ref_lse;
  ep.tau = tau;
  ep.cur_id = cur_id;
  ep.prev_id = prev_id;
  ep.cand_count = cand_count;
  ep.cand_id = cand_id;
  ep.cand_cap = cand_cap;
  ep.n_prompts = static_cast<int>((n_rows + 127) / 128);
  return launch_gemm_mode<tb::LensLLEpi>(cm, h, w_u, g, ep, static_cast<cudaStream_t>(stream), di.sms);
}

int tb_ll_candidates_exact_tc(const void* h, const void* w_u, int64_t n_rows, int d, int V, int vocab_offset,
                              float softcap, const float* lse, const float* ref_lse, const float* tau,
                              const int32_t* cur_id, const int32_t* prev_id, const int32_t* cand_count,
                              const int32_t* cand_id, int cand_cap, int k, void* wc_scratch, float* sum_scratch,
                              float* out_val, int32_t* out_id, float* out_theta, int32_t* out_flag, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  if (!h || !w_u || !lse || !ref_lse || !tau || !cand_count || !cand_id || !wc_scratch || !sum_scratch || !out_val ||
      !out_id || !out_theta || !out_flag)
    return fail(TB_ERR_INVALID, "null pointer");
  if (k < 1 || k > TB_TOPK) return fail(TB_ERR_INVALID, "k=%d outside [1, %d]", k, TB_TOPK);
  if (cand_cap < 1) return fail(TB_ERR_INVALID, "cand_cap=%d", cand_cap);
  if (n_rows <= 0) return TB_OK;
  if (n_rows % 128) return fail(TB_ERR_INVALID, "n_rows=%lld: 128 row slots per prompt", (long long)n_rows);
  rc = check_gemm_operands(wc_scratch, h, n_rows, static_cast<int>(n_rows), d);
  if (rc != TB_OK) return rc;
  const int n_prompts = static_cast<int>(n_rows / 128);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // 1. candidates' W_U rows -> 128 slots per prompt
  {
    const int64_t slots = static_cast<int64_t>(n_prompts) * tb::kCandSlots;
    const unsigned grid = static_cast<unsigned>((slots + 7) / 8);
    tb::tb_ll_gather_candidates_kernel<<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(w_u), V, d, vocab_offset,
                                                          cand_count, cand_id, cand_cap, n_prompts,
                                                          static_cast<__nv_bfloat16*>(wc_scratch));
    TB_CUDA(cudaGetLastError());
  }
  // 2. block-diagonal GEMM: prompt pair u's candidate tile x the same pair's row block
  {
    tb::GemmArgs g = make_args_sweep_a(static_cast<int>(n_rows), n_rows, d, 1, 2);
    g.diag = 1;
    g.total_units = g.m_blocks;
    tb::CandSumEpi::Params ep;
    ep.lse = lse; ep.ref = ref_lse; ep.cur_id = cur_id; ep.prev_id = prev_id;
    ep.cand_count = cand_count; ep.cand_id = cand_id; ep.cand_cap = cand_cap; ep.softcap = softcap;
    ep.n_prompts = n_prompts; ep.out_sum = sum_scratch;
    rc = launch_gemm<tb::CandSumEpi, 2>(wc_scratch, h, g, ep, s, di.sms);
    if (rc != TB_OK) return rc;
  }
  // 3. top-k, theta, certificate
  tb::CandFinalArgs a;
  a.sums = sum_scratch; a.lse = lse; a.ref = ref_lse; a.tau = tau; a.cand_count = cand_count; a.cand_id = cand_id;
  a.cap = cand_cap; a.k = k; a.n_rows = static_cast<int>(n_rows);
  a.out_val = out_val; a.out_id = out_id; a.out_theta = out_theta; a.out_flag = out_flag;
  tb::tb_ll_candidates_finalize_kernel<<<static_cast<unsigned>(n_prompts), tb::kCandSlots, 0, s>>>(a);
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

int tb_logit_lens_static(const void* h, const void* w_u, int64_t n_rows, int d, int V, int vocab_offset,
                         float softcap, const int32_t* secret_id, int n_chunks, int cta_pair, float* lse,
                         float* secret_prob, float* top_prob, int32_t* top_id, void* stream) {
  DevInfo di;
  int rc = device_info(&di);
  if (rc != TB_OK) return rc;
  rc = check_gemm_operands(h, w_u, n_rows, V, d);
  if (rc != TB_OK) return rc;
  if (!lse || !top_prob || !top_id) return fail(TB_ERR_INVALID, "lse, top_prob and top_id are required");
  const tb::GemmArgs g = make_args(n_rows, V, d, n_chunks, cta_pair ? 2 : 1);  // (shape check only)
  if (g.n_chunks != n_chunks)
    return fail(TB_ERR_INVALID, "n_chunks=%d does not divide %d tiles into non-empty chunks (use %d)", n_chunks,
                g.n_tiles, g.n_chunks);
  const int n_parts = n_chunks * tb_lens_parts_per_chunk();
  const int64_t slots = n_rows * static_cast<int64_t>(n_parts);
  if (slots > TB_STATIC_LENS_MAX_SLOTS)
    return fail(TB_ERR_INVALID,
                "n_rows * n_chunks * %d = %lld exceeds the built-in scratch of %d (row, part) slots; split the rows "
                "over several calls or use tb_lens_partials + tb_lens_merge with your own scratch",
                tb_lens_parts_per_chunk(), static_cast<long long>(slots), TB_STATIC_LENS_MAX_SLOTS);
  float* base = nullptr;
  rc = static_scratch_address(&base);
  if (rc != TB_OK) return rc;
  // carve the five partial arrays out of the built-in block; every slot that K4 reads is written
  // by K3 first, so nothing left over from an earlier call is ever observed
  float* pm = base;
  float* ps = pm + slots;
  float* psec = ps + slots;
  float* ptv = psec + slots;
  int32_t* pti = reinterpret_cast<int32_t*>(ptv + slots * TB_TOPK);
  rc = tb_lens_partials(h, w_u, n_rows, d, V, vocab_offset, softcap, secret_id, n_chunks, cta_pair, pm, ps, psec,
                        ptv, pti, stream);
  if (rc != TB_OK) return rc;
  return tb_lens_merge(pm, ps, psec, ptv, pti, n_parts, n_rows, 1, nullptr, nullptr, nullptr, nullptr, nullptr,
                       lse, secret_prob, top_prob, top_id, stream);
}

}  // extern "C"

#ifdef TB_WAIT_PROFILE
// Probe build only: per-CTA phase cycles of the one-pass K2 kernel (256 x 16 counters; scripts/k2_profile.py)
extern "C" int tb_debug_k2_profile(long long* out_host) {
  TB_CUDA(cudaDeviceSynchronize());
  TB_CUDA(cudaMemcpyFromSymbol(out_host, tb::g_k2_prof, sizeof(long long) * 256 * 16));
  return TB_OK;
}
// Probe build only: read and clear the per-CTA wait counters of tb_gemm.cuh (256 x 8 cycle counts).
extern "C" int tb_debug_wait_profile(long long* out_host) {
  TB_CUDA(cudaDeviceSynchronize());
  TB_CUDA(cudaMemcpyFromSymbol(out_host, tb::g_wait_prof, sizeof(long long) * 256 * 8));
  static long long zeros[256 * 8];
  TB_CUDA(cudaMemcpyToSymbol(tb::g_wait_prof, zeros, sizeof(zeros)));
  return TB_OK;
}
#endif
