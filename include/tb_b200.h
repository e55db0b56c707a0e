/* tb_b200.h -- C ABI of libtb_b200.so: the B200 (sm_100a) kernels behind the
 * intervention-and-readout hot path of lmmontoya-ai/taboo-brittleness.
 *
 * The reference has no FFI of its own: its boundary is the Python surface of src/models.py,
 * src/feature_map.py, src/metrics.py and the two drivers. The host
 * side (taboo_brittleness_b200/*.py) keeps that surface and binds these entry points with
 * ctypes; each entry point below names the reference line(s) whose arithmetic it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into memory the caller allocated (inputs, outputs and
 *     scratch alike); the library allocates nothing. Two blocks of device memory are part of the
 *     library image on every GPU that loads it and live until the process exits: the pacing counters
 *     of the multicast cluster modes (8 rotating slots of 128 ints; with more than 8 such launches in
 *     flight on different streams two launches share a slot, which only costs pacing accuracy -- every
 *     wait on them is bounded) and the 6 MB scratch of tb_logit_lens_static. Nothing else persists
 *     between calls;
 *   - matrices are dense row-major; "bf16" = __nv_bfloat16, stored as uint16_t bits here;
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it and never
 *     synchronise;
 *   - return value: TB_OK (0) or a negative TB_ERR_* code; tb_last_error() describes the last
 *     failure of the calling thread;
 *   - there is no CPU fallback: on a device that is not compute capability 10.x every compute
 *     entry point returns TB_ERR_ARCH.
 */
#ifndef TB_B200_H_
#define TB_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TB_OK 0
#define TB_ERR_INVALID (-1)  /* bad argument (null pointer, shape, alignment, range)        */
#define TB_ERR_ARCH (-2)     /* current device is not sm_100                                */
#define TB_ERR_CUDA (-3)     /* a CUDA runtime / driver call failed                         */

#define TB_TOPK 10           /* per-row top-k kept by the lens readout                      */
#define TB_TOPM_MAX 1024     /* largest ablation budget (rank tables are uint16)            */
#define TB_RANK_NONE 0xFFFFu /* rank-table entry of a latent that is not in the ablation set */

int tb_version(void);
/* Partials written per (chunk, row) by tb_lens_partials / tb_lens_ll_partials: every row is swept by
 * this many epilogue threads (column halves of each tile), each leaving its own partial (2). */
int tb_lens_parts_per_chunk(void);
const char* tb_last_error(void);
/* TB_OK iff the CURRENT device can run the kernels (compute capability 10.x). */
int tb_device_check(void);

/* ---- K1: JumpReLU SAE encode ---------------------------------------------------------
 * Replaces sae.encode(response_residuals), src/02_run_sae_baseline.py:67 (sae_lens JumpReLU:
 *   pre = x @ W_enc + b_enc;  acts = relu(pre) * (pre > threshold)).
 * tcgen05 GEMM [n_rows, d] x [S, d]^T with the bias, threshold and (optional) ablation mask
 * fused in the epilogue. The dense acts matrix is not formed: row r's non-zero latents are
 * written in increasing latent order to idx/val[start(r) ...], and count[r] receives their
 * number. start(r) = row_start[r] (CSR; row_cap[r] entries may be written, NULL = unchecked)
 * or, when row_start is NULL, r * cap (fixed stride; count[r] > cap flags an overflowed row,
 * only its first cap latents are stored). out_idx == NULL: count only.
 * rank (optional): uint16 table(s) [n_tables, S]; a latent whose rank < budget is treated as
 * zero (ablated). rank_rows = rows sharing a table (0: one table for all rows).
 *   w_enc_t : bf16 [S, d]  (W_enc transposed, K-major)     b_enc, theta : f32 [S]
 */
int tb_sae_encode(const void* x_bf16, const void* w_enc_t_bf16, const float* b_enc,
                  const float* theta, int64_t n_rows, int d, int S, const int64_t* row_start,
                  const int32_t* row_cap, int cap, int32_t* out_idx, float* out_val,
                  int32_t* out_count, const uint16_t* rank, int rank_rows, int budget,
                  void* stream);

/* K1 for large batches with the sweep over the latents cut into n_sub chunks (2..8, dividing cap): same result as
 * tb_sae_encode with fixed-stride lists, but several CTA pairs share a row block, so the resident row blocks fit L2
 * (with one chunk, 74 blocks of 1.8 MB are re-read from DRAM for every W_enc^T tile: 54x the algorithmic traffic in
 * the ncu capture). Chunk c of a row first writes its own cap / n_sub slots and sub_count[c * n_rows + row]; a merge
 * kernel then closes the gaps in place: out_count[row] = the row's total, out_truncated[row] = 1 if a sub-list did
 * not fit its slots (re-encode such rows alone with tb_sae_encode, as for rows whose whole list overflows).
 * Operands must be the current GPU's own memory (TMA multicast clusters). Replaces sae.encode, 02:67. */
int tb_sae_encode_chunked(const void* x_bf16, const void* w_enc_t_bf16, const float* b_enc, const float* theta,
                          int64_t n_rows, int d, int S, int cap, int n_sub, int32_t* out_idx, float* out_val,
                          int32_t* sub_count, int32_t* out_count, int32_t* out_truncated, void* stream);

/* ---- a2: targeted latent selection ------------------------------------------------------
 * Replaces torch.mean(acts, dim=0) and torch.topk(avg, k), src/02_run_sae_baseline.py:70-73,
 * on the sparse lists. seg (optional): int64 [n_prompts + 1] row offsets of each prompt's
 * response window; NULL = uniform rows_per_prompt. out_mean: f32 [n_prompts, S].
 */
int tb_sae_latent_mean(const int64_t* row_start, int cap, const int32_t* count, const int32_t* idx,
                       const float* val, const int64_t* seg, int rows_per_prompt,
                       int64_t n_prompts, int S, float* out_mean, void* stream);
/* Weighted per-prompt latent sums out[p, s] = sum_t w[t] * a[t, s]^power (power 1 or 2; w NULL = 1),
 * the building block of the Execution Plan's secret score (Plan:116-124: mean activation at the spike
 * tokens x max(0, correlation with the secret probability)); no code in the reference. */
int tb_sae_latent_wsum(const int64_t* row_start, int cap, const int32_t* count, const int32_t* idx,
                       const float* val, const int64_t* seg, int rows_per_prompt, int64_t n_prompts,
                       int S, const float* row_weight, int power, float* out_sum, void* stream);
/* Top-m (m <= TB_TOPM_MAX) of each prompt's mean, sorted descending (ties: lower latent id
 * first): out_idx/out_val [n_prompts, m]; rank u16 [n_prompts, S], rank[l] = position of
 * latent l in the list or TB_RANK_NONE. Nested budgets m' <= m are "rank < m'". */
int tb_topm_select(const float* mean, int64_t n_prompts, int S, int m, int32_t* out_idx,
                   float* out_val, uint16_t* rank /* nullable: sorted lists only */, void* stream);

/* ---- K1b: latent ablation + sparse decode (+ fused final RMSNorm) -----------------------
 * No code in the reference (Execution Plan:126 "zero m latents, decode back to residual";
 * random control Plan:128); sae.decode semantics x_hat = acts @ W_dec + b_dec.
 *   mode 0: x' = x - sum_{i in ablated & active(row)} a_i W_dec[i,:]   (decode of the ablated
 *           code plus the SAE reconstruction error x - x_hat, as north_star asks)
 *   mode 1: x' = b_dec + sum_{i active(row), not ablated} a_i W_dec[i,:]   (Plan:126 literally)
 * "ablated" = rank[table(row)][i] < budget; table(row) = row_table[row], or row / rank_rows,
 * or 0. x' is rounded to bf16. out_x and/or out_h may be NULL; out_h = Gemma2RMSNorm(x')
 * in bf16 (see tb_rmsnorm), produced in the same pass.
 */
int tb_sae_ablate_decode(const void* x_bf16, const int64_t* row_start, int cap,
                         const int32_t* count, const int32_t* idx, const float* val,
                         const void* w_dec_bf16, const float* b_dec, const uint16_t* rank,
                         const int32_t* row_table, int rank_rows, int budget, int mode,
                         int64_t n_rows, int d, int S, const float* w_norm, float eps,
                         void* out_x_bf16, void* out_h_bf16, void* stream);

/* K1b for NESTED budgets (Plan:126: budgets are prefixes of one ranking per prompt, or per seed for the random
 * control, Plan:128). tb_sae_rank_lists runs once per ranking: for every row, its active latents that appear in
 * the ranking (rank < 1024), ordered by rank, written as key = rank << 20 | latent and value into out_key /
 * out_val at the row's list position (same row_start / cap layout as idx / val, so the buffers have the same
 * size), their number into out_count. tb_sae_ablate_ranked then serves any budget: it reads the leading entries
 * with rank < budget (at most `budget`) and computes what tb_sae_ablate_decode computes in mode 0, the sum taken
 * in rank order. S <= 2^20. out_rstd (nullable, f32 [n_rows]): the final RMSNorm's 1 / sqrt(mean(x'^2) + eps) of the
 * rounded x' -- all the rank-update lens (tb_lens_rank_update_partials) needs of a budget. No code in the reference.
 */
int tb_sae_rank_lists(const int64_t* row_start, int cap, const int32_t* count, const int32_t* idx,
                      const float* val, const uint16_t* rank, int rank_rows, int64_t n_rows, int S,
                      int32_t* out_key, float* out_val, int32_t* out_count, void* stream);
int tb_sae_ablate_ranked(const void* x_bf16, const int64_t* row_start, int cap, const int32_t* sel_count,
                         const int32_t* sel_key, const float* sel_val, const void* w_dec_bf16, int budget,
                         int64_t n_rows, int d, const float* w_norm, float eps, void* out_x_bf16,
                         void* out_h_bf16, float* out_rstd, void* stream);
/* All budgets of a chain in one pass over the rows: x is read once and every W_dec row a row needs is gathered once;
 * after each budget's entries the row is rounded, normalised and written to that budget's output (the same bits as
 * one tb_sae_ablate_ranked call per budget). budgets_host: n_budgets (<= 16) strictly ascending values, a HOST array
 * read during the call. Output k at out_x / out_h + k * out_stride elements (out_stride >= n_rows * d), out_rstd +
 * k * n_rows. */
int tb_sae_ablate_ranked_multi(const void* x_bf16, const int64_t* row_start, int cap, const int32_t* sel_count,
                               const int32_t* sel_key, const float* sel_val, const void* w_dec_bf16,
                               const int32_t* budgets_host, int n_budgets, int64_t n_rows, int d, const float* w_norm,
                               float eps, void* out_x_bf16, void* out_h_bf16, float* out_rstd, int64_t out_stride,
                               void* stream);

/* ---- K2: low-rank direction removal x' = x - (x U) U^T ----------------------------------
 * No code in the reference (Execution Plan:148 r_edited = r - P_U(r); random control :150).
 * ut = U^T, bf16 [r, d], r <= 64, columns of U orthonormal. Outputs as for K1b.
 */
int tb_lowrank_remove(const void* x_bf16, const void* ut_bf16, int r, int64_t n_rows, int d,
                      const float* w_norm, float eps, void* out_x_bf16, void* out_h_bf16,
                      void* stream);

/* (When d is a multiple of 256, at most 4096, the second tcgen05 pass and the norm pass described here run as ONE
 * kernel on clusters of d / 256 CTAs -- x added into the accumulator by the tensor core, sums of squares exchanged
 * through distributed shared memory; TB_K2_THREE_PASS=1 in the environment keeps them separate. Same contract.)
 * K2 for ranks 9..64 on the tensor cores, one call: C = x U (tcgen05, fp32 accumulate), the
 * coefficients split into bf16 head + remainder, x' = bf16(x - [C_hi | C_lo] [U | U]^T) in the epilogue
 * of a second tcgen05 pass (K = 2 * ceil32(r), N = d), and the fused final RMSNorm from the rows' sums
 * of squares collected there. Same result contract as tb_lowrank_remove (Plan:148; no code in the
 * reference). scratch: device memory, 256-byte aligned, at least tb_lowrank_tc_scratch_bytes(n_rows,
 * d, r) bytes (-1: bad arguments / no device); nothing is allocated inside. d % 8 == 0, d <= 8192. */
int64_t tb_lowrank_tc_scratch_bytes(int64_t n_rows, int d, int r);
int tb_lowrank_remove_tc(const void* x_bf16, const void* ut_bf16, int r, int64_t n_rows, int d,
                         const float* w_norm, float eps, void* scratch, int64_t scratch_bytes,
                         void* out_x_bf16, void* out_h_bf16, void* stream);
/* Nested ranks (the first r columns of one U, for several r: Plan:148-150 sweeps r = 1 .. 64) in one call: the
 * coefficients x U of the largest rank are computed ONCE (one pass over x instead of one per rank) and every rank's
 * removal + norm pass uses its leading r of them -- the same products in the same order, so the same bits as one
 * tb_lowrank_remove_tc call per rank. ranks_host: n_ranks (<= 16) strictly ascending values, a HOST array; ut holds the
 * largest rank's rows; output k at out + k * out_stride elements; scratch as for the largest rank. d % 256 == 0,
 * d <= 4096. */
int tb_lowrank_remove_tc_multi(const void* x_bf16, const void* ut_bf16, const int32_t* ranks_host, int n_ranks,
                               int64_t n_rows, int d, const float* w_norm, float eps, void* scratch,
                               int64_t scratch_bytes, void* out_x_bf16, void* out_h_bf16, int64_t out_stride,
                               void* stream);

/* ---- a6: final RMSNorm -------------------------------------------------------------------
 * Replaces model.model.norm(...), src/models.py:135 (Gemma2RMSNorm.forward on a bf16
 * residual): h = bf16((x * rsqrt(mean(x^2) + eps)) * (1 + w)), fp32 intermediates.
 */
int tb_rmsnorm(const void* x_bf16, const float* w_norm, float eps, int64_t n_rows, int d,
               void* out_h_bf16, void* stream);

/* ---- K3: logit-lens readout ---------------------------------------------------------------
 * Replaces model.lm_head(h) -> softmax -> [secret-token gather, per-row max/top-k]:
 * src/models.py:135-138,147; src/plots.py:10; secret id src/01_reproduce_logit_lens.py:142.
 * tcgen05 GEMM h [n_rows, d] x W_U[vocab shard] [V, d]^T; the epilogue applies the optional
 * final-logit softcap (softcap <= 0: none, as the reference; Gemma-2: 30.0), keeps an online
 * (max, sum-exp), the secret token's logit and a top-10 per row. Logits never reach HBM.
 * The vocabulary shard is swept in n_chunks contiguous chunks (scheduling / L2 knob); the
 * per-(part, row) partials are written to caller scratch, n_parts = n_chunks *
 * tb_lens_parts_per_chunk() (pass n_parts to tb_lens_merge):
 *   pm, ps, psec : f32 [n_parts, n_rows]      ptv : f32 [n_parts, n_rows, 10]
 *   pti : i32 [n_parts, n_rows, 10]  (global vocab ids = vocab_offset + column)
 * secret_id: i32 [n_rows] global vocab id per row, or NULL. cta_pair selects the cluster shape:
 * 0 single CTAs, 1 CTA pairs (cta_group::2), 2 = 2x1 pairs (two row blocks share the W_U tiles
 * through TMA multicast), 3 = 1x2 pairs (two chunks share the row block), 4 = 2x2 pairs,
 * 5 = mode 2 as the PREFERRED cluster shape with plain pairs as the fallback (groups of four CTAs
 * that do not fit a GPC as one cluster run as two pairs: all 148 SMs work; the default of ops.py);
 * n_chunks must be even in modes 3 and 4. Modes 2-5 multicast operand tiles within a cluster and take
 * operands in the current GPU's own memory only (anything else is refused with TB_ERR_INVALID):
 * TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from
 * another GPU's memory hung the GPU. Modes 0 and 1 do not multicast.
 */
int tb_lens_partials(const void* h_bf16, const void* w_u_bf16, int64_t n_rows, int d, int V,
                     int vocab_offset, float softcap, const int32_t* secret_id, int n_chunks,
                     int cta_pair, float* pm, float* ps, float* psec, float* ptv, int32_t* pti,
                     void* stream);
/* K4: merge n_parts partials per row (chunks of one GPU, or all-gathered shards).
 * finalize == 0: one merged partial to (om, os, osec, otv, oti) (same layout, one part);
 * finalize != 0: lse, secret_prob (nullable), top_prob / top_id [n_rows, 10]. */
int tb_lens_merge(const float* pm, const float* ps, const float* psec, const float* ptv,
                  const int32_t* pti, int n_parts, int64_t n_rows, int finalize, float* om,
                  float* os, float* osec, float* otv, int32_t* oti, float* lse,
                  float* secret_prob, float* top_prob, int32_t* top_id, void* stream);

/* ---- K5: LL-Top-k ------------------------------------------------------------------------
 * Replaces aggregate_response_logits + torch.topk, src/01_reproduce_logit_lens.py:35-71 and
 * :147-149: per prompt, acc[v] = sum over response positions t of softmax(logits)[t, v] with
 * the current token's id and the previous token's id zeroed at each position; top-k of acc.
 * Second sweep of the lens GEMM with lse[t] from K3 + K4 (same h, W_U shard, softcap): a
 * thread owns a vocabulary id and sums over a prompt's positions, no [T, V] tensor exists.
 * Row layout: prompt p owns row slots [128 p, 128 p + 128) of h; slots without a token
 * carry lse = +inf (prompts longer than 128 positions are not supported by this kernel).
 *   lse     : f32 [n_rows]                      cur_id / prev_id : i32 [n_rows] global vocab ids
 *             zeroed at that position (-1 or out of range: nothing), both nullable
 *   part_val: f32 [n_parts, n_prompts, 10]      part_id : i32 [n_parts, n_prompts, 10]
 *             n_prompts = ceil(n_rows / 128), n_parts = n_chunks * (cta_pair ? 2 : 1); sorted
 * A chunk may span at most 65536 vocabulary ids (n_chunks >= 4 for V = 256000).
 */
int tb_ll_topk_partials(const void* h_bf16, const void* w_u_bf16, int64_t n_rows, int d, int V,
                        int vocab_offset, float softcap, const float* lse, const int32_t* cur_id,
                        const int32_t* prev_id, int n_chunks, int cta_pair, float* part_val,
                        int32_t* part_id, void* stream);
/* merge n_parts sorted lists per prompt -> the k <= 10 best (sum, id), descending; id -1 and
 * sum -inf where fewer exist. The reference returns no guesses when acc sums to zero
 * (01:148): that is out_val[p][0] <= 0. */
int tb_ll_topk_merge(const float* part_val, const int32_t* part_id, int n_parts, int64_t n_prompts,
                     int k, float* out_val, int32_t* out_id, void* stream);

/* ---- a11 on cached dense probabilities ------------------------------------------------------
 * Replaces aggregate_response_logits (src/01_reproduce_logit_lens.py:35-71) on the cached
 * all_probs[layer, start:] slice the reference's driver analyses (:120-150): acc[v] = sum over
 * positions of probs[i, v], with the entries at cur_id[i] and prev_id[i] (-1: none) taken as 0.
 * Positions are added in order in fp32, exactly like the reference's loop, so acc is
 * bit-identical; tb_topm_select(acc, 1, V, k, ...) then gives torch.topk(acc, k).
 *   probs : f32 [T, V] row-major, T <= 1024      acc : f32 [V]
 */
int tb_probs_ll_colsum(const float* probs, int T, int V, const int32_t* cur_id, const int32_t* prev_id,
                       float* acc, void* stream);

/* K2 for ranks 8..64 in ONE pass over the rows (x read once, h written once; same result contract as
 * tb_lowrank_remove, Plan:148; no code in the reference). A cluster of d / 256 CTAs holds a 128-row block in
 * its shared memory: C = x U as per-CTA partial tcgen05 products reduced across the cluster through
 * distributed shared memory (rank order: deterministic), x' = bf16(x - [C_hi | C_lo][U | U]^T) and the
 * fused final RMSNorm on the block while it is resident. Needs d % 256 == 0 and d <= 4096
 * (tb_lowrank_fused_supported returns 1); no scratch. dbg_c (nullable): f32 [n_rows, 32 * ceil(r / 32)],
 * receives the reduced coefficients x U (tests). */
int tb_lowrank_fused_supported(int d, int r);
int tb_lowrank_remove_fused(const void* x_bf16, const void* ut_bf16, int r, int64_t n_rows, int d,
                            const float* w_norm, float eps, void* out_x_bf16, void* out_h_bf16,
                            float* dbg_c, void* stream);

/* ---- K3 with dense output ------------------------------------------------------------------
 * tb_lens_dense_partials is tb_lens_partials (single shard, any cluster mode) that also writes the (soft-capped)
 * logits to logits_out f32 [n_rows, V]. Two users:
 *  - the compatibility surface of get_layer_logits (src/models.py:135-144 returns softmax(lm_head(norm(.))) for every
 *    position as a dense array): after tb_lens_merge has produced lse, tb_probs_from_logits turns the array into
 *    probabilities in place, p = exp(logit - lse[row]). One GEMM per call;
 *  - LL-Top-k of a condition that has no predecessor (01:35-71, 147-149): instead of a second sweep of the GEMM,
 *    tb_logits_ll_colsum reduces the stored logits of a row sub-batch to acc f32 [n_prompts, V] (128 row slots per
 *    prompt, lse = +inf marks an empty slot, cur_id / prev_id zeroed per position) and tb_topm_select (rank = NULL)
 *    ranks each prompt's sums. HBM-bound: n_rows * V * 4 bytes written once and read once.
 *
 * layout of the logits array (both calls): TB_LOGITS_ROW_MAJOR = [n_rows, V]; TB_LOGITS_PROMPT_TILES =
 * [n_rows / 128][V / 32][128 rows][32 columns], values in LOG2 units (logit x log2 e; V % 32 == 0, n_rows % 128 == 0;
 * an intermediate for tb_logits_ll_colsum, not an output format) -- a prompt's 128 x 32 piece is 16 KB of
 * consecutive memory, which is what lets tb_logits_ll_colsum stream; only the LL-Top-k user wants it. The lens kernel
 * writes either layout through the TMA unit (row-major with V % 4 != 0: plain stores). */
#define TB_LOGITS_ROW_MAJOR 0
#define TB_LOGITS_PROMPT_TILES 1
#define TB_LOGITS_PROMPT_TILES_RAW 2 /* prompt tiles holding the RAW products h W_U^T (no soft cap, natural units), no
                                        readout written (pm .. pti may be NULL): the base logits of the rank-update lens */
int tb_lens_dense_partials(const void* h_bf16, const void* w_u_bf16, int64_t n_rows, int d, int V, float softcap,
                           const int32_t* secret_id, int n_chunks, int cta_pair, float* pm, float* ps, float* psec,
                           float* ptv, int32_t* pti, float* logits_out, int layout, void* stream);
int tb_logits_ll_colsum(const float* logits, const float* lse, const int32_t* cur_id, const int32_t* prev_id,
                        int64_t n_rows, int V, float* acc, int layout, void* stream);
int tb_probs_from_logits(float* logits, const float* lse, int64_t n_rows, int V, void* stream);

/* ---- nested budgets as rank updates of stored base logits (budgets of Execution Plan:126) -----------
 * For the error-term ablation the lens logits of budget m are rstd_m[t] * (L0[t, v] - sum_{rank(j) < m} a_j[t] D[j, v]) with
 * L0 = (x (.) (1 + w)) W_U^T and D = (W_dec (.) (1 + w)) W_U^T [S, V] (bf16, computed once per weight set by the caller and
 * handed over TILE-MAJOR: d_tiles [ceil(V / 256)][S][256], ids past V zero).
 * Per row sub-batch of whole prompts:
 *   tb_scale_cols(x, w_norm) -> xw = bf16(x (.) (1 + w));  tb_lens_dense_partials(xw, W_U, ..., l0, TB_LOGITS_PROMPT_TILES_RAW);
 *   tb_rank_acts_dense: the rank-ordered lists of tb_sae_rank_lists as a dense matrix acts bf16 [n_rows, a_cols]:
 *     column r = the activation of the prompt's rank-r latent (a_cols a multiple of 64, >= 1024; lo_off != 0: the bf16
 *     remainders a - bf16(a) in column lo_off + r, which halves the rounding of the activations for twice the depth);
 *   per budget: tb_sae_ablate_ranked(..., out_rstd) and tb_lens_rank_update_partials: same partial layout as
 *     tb_lens_partials with tb_lens_rank_update_parts(budget, lo_off) partials per (chunk, row) -> tb_lens_merge. rank_ids i32 [n_rows / 128, rank_stride]: the
 *     latent id of each rank of each prompt (tb_topm_select's ids). logits_out_tiles (nullable, its own n_rows * V floats):
 *     the budget's soft-capped logits in the TB_LOGITS_PROMPT_TILES form, for LL-Top-k by tb_logits_ll_colsum.
 * Not the arithmetic of the reference path (no bf16 rounding of the normalised residual per budget, D is a rounded
 * product): drift against the dense kernels is reported in DESIGN.md; no code in the reference. */
int tb_scale_cols(const void* x_bf16, const float* w_norm, int64_t n_rows, int d, void* out_bf16, void* stream);
int tb_rank_acts_dense(const int64_t* row_start, int cap, const int32_t* sel_count, const int32_t* sel_key,
                       const float* sel_val, int64_t n_rows, int a_cols, int lo_off, void* out_bf16, void* stream);
int tb_lens_rank_update_parts(int budget, int lo_off);   /* partials per (chunk, row) the next call writes: 2 or 4 */
int tb_lens_rank_update_partials(const void* acts_bf16, int a_cols, int lo_off, const int32_t* rank_ids, int rank_stride,
                                 const void* d_tiles_bf16, int S, const float* l0_tiles, const float* rstd, int64_t n_rows, int V,
                                 int budget, float softcap, const int32_t* secret_id, int n_chunks, float* pm, float* ps,
                                 float* psec, float* ptv, int32_t* pti, float* logits_out_tiles, void* stream);

/* ---- K3F + K6: lens readout and LL-Top-k in one sweep -------------------------------------
 * Same readouts as K3 (tb_lens_partials) and K5, for a condition whose rows' logsumexp is
 * already known approximately -- in a sweep over nested ablation budgets, the previous budget's
 * exact lse. K3F sums exp(logit - ref_lse[t]) instead of tracking a running maximum (the
 * partials carry pm = ref_lse, tb_lens_merge needs no change) and, per prompt (128 row slots as
 * K5), lists every vocabulary id whose masked column sum of those terms reaches tau[prompt].
 *   cand_count : i32 [n_prompts], MUST be zero on entry      cand_id : i32 [n_prompts, cand_cap]
 * tb_ll_candidates_exact_tc then recomputes the true sums of the listed ids (with the exact lse
 * from tb_lens_merge) and writes the k best; out_flag[p] = 1 certifies that no unlisted id can
 * belong to them (list did not overflow and the k-th sum >= max_t exp(ref_lse - lse) * tau).
 * Prompts with flag 0 must be redone with tb_ll_topk_partials. out_theta = the k-th best sum
 * (callers derive the next condition's tau from it). Single vocabulary shard only.
 */
int tb_lens_ll_partials(const void* h_bf16, const void* w_u_bf16, int64_t n_rows, int d, int V,
                        int vocab_offset, float softcap, const int32_t* secret_id,
                        const float* ref_lse, const float* tau, const int32_t* cur_id,
                        const int32_t* prev_id, int n_chunks, int cta_pair, float* pm, float* ps,
                        float* psec, float* ptv, int32_t* pti, int32_t* cand_count,
                        int32_t* cand_id, int cand_cap, void* stream);
/* K6, with the sums on the tensor cores: the candidates' W_U rows are gathered into 128 slots per prompt (wc_scratch: bf16 [n_rows, d], n_rows = 128 *
 * n_prompts), a block-diagonal tcgen05 GEMM multiplies each prompt pair's candidate tile with the
 * same pair's rows (sum_scratch: f32 [n_prompts, 128]), a small kernel ranks and certifies. Prompts
 * with more than 128 listed candidates get flag 0. */
int tb_ll_candidates_exact_tc(const void* h_bf16, const void* w_u_bf16, int64_t n_rows, int d, int V,
                              int vocab_offset, float softcap, const float* lse, const float* ref_lse,
                              const float* tau, const int32_t* cur_id, const int32_t* prev_id,
                              const int32_t* cand_count, const int32_t* cand_id, int cand_cap, int k,
                              void* wc_scratch, float* sum_scratch, float* out_val, int32_t* out_id,
                              float* out_theta, int32_t* out_flag, void* stream);

/* ---- lens readout without caller scratch -------------------------------------------------
 * (No counterpart in the reference: this entry point exists for the way the library is
 * deployed, not for anything the reference computes.)
 * Same readout as tb_lens_partials followed by tb_lens_merge(finalize), but the per-(row,
 * chunk) partials live in a fixed block inside the library instead of in memory the caller
 * provides. It is meant for callers that hold nothing but their input and output buffers and
 * have no way to obtain more: a worker that was handed pre-sized tensors by a serving runtime
 * whose allocator already owns all of HBM, a plug-in written against this header alone (no
 * torch, no allocator of its own), or a latency-critical loop that must not enter
 * cudaMalloc/cudaFree because those serialise the device. The function itself performs no
 * allocation of any kind.
 *   - limit: n_rows * n_chunks * tb_lens_parts_per_chunk() <= TB_STATIC_LENS_MAX_SLOTS (65536),
 *     i.e. n_rows * n_chunks <= 32768; larger requests are rejected with TB_ERR_INVALID and
 *     nothing is launched. The block is 6,029,312 bytes
 *     (65536 x 23 words) and is resident on every GPU that loads the library until the
 *     process exits.
 *   - the first call on each device resolves the block's device address (one
 *     cudaGetSymbolAddress, with that device current). Make that first call during start-up,
 *     before any stream capture is begun.
 *   - all calls on a device share the one block. Calls whose kernels may overlap on the GPU
 *     (different streams, different host threads) race on it: serialise them, or give all
 *     but one of them caller scratch through tb_lens_partials. The block keeps whatever the
 *     last call left; that is harmless here because every slot read by the merge is first
 *     written by the lens kernel of the same call.
 * Arguments as tb_lens_partials / tb_lens_merge; secret_prob may be NULL.
 */
#define TB_STATIC_LENS_MAX_SLOTS 65536
int tb_logit_lens_static(const void* h_bf16, const void* w_u_bf16, int64_t n_rows, int d, int V,
                         int vocab_offset, float softcap, const int32_t* secret_id, int n_chunks,
                         int cta_pair, float* lse, float* secret_prob, float* top_prob,
                         int32_t* top_id, void* stream);

/* ---- validation only: D = A B^T through the same TMA/tcgen05 main loop, fp32 out --------- */
int tb_gemm_bf16_nt(const void* a_bf16, const void* b_bf16, int64_t M, int N, int K,
                    int n_chunks, int cta_pair, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TB_B200_H_ */
