/*
 * arborist_b200.h -- C ABI of libarborist_b200.so (hand-written sm_100a CUDA).
 *
 * The drop-in boundary for Arborist's per-tree variational fit.  The reference
 * (/root/reference/src/arborist/arborist.py, "A.py") has no FFI of its own; its
 * de-facto operator seam is the function-valued variable
 *     run = run_variational_inference                      (A.py:1076)
 * called once per candidate tree from the loop at A.py:1089-1126.  The entry
 * points below are what a binding for that path needs: build the device-resident
 * read-count matrix once (A.py:1034-1069), set the prior (A.py:1052), then fit a
 * whole SHARD of trees in one call (A.py:1089-1126 collapsed), and fetch results.
 * INTEGRATION.md shows the ctypes stub that replaces the loop.
 *
 * Conventions
 *   - every function returns 0 (ARB_OK) or a negative arb_status; the message for
 *     the calling thread's last failure is arb_last_error();
 *   - plain pointers and sizes only; pointers are HOST memory unless a name ends
 *     in _dev; outputs are caller-allocated; inputs are never modified;
 *   - one arb_ctx per GPU, used from one host thread at a time; every call is
 *     synchronous (work is enqueued on the context's stream and waited for);
 *   - no C++ exceptions cross the boundary; there is no CPU fallback: without a
 *     usable CUDA device arb_ctx_create fails with ARB_ERR_CUDA.
 *
 * Index conventions (the reference's): K = n_clones counts the normal/root clone,
 * C = K-1 clusters; clone and cluster indices are positions in the ascending label
 * order (A.py:1025-1028); cell/SNV indices are first-appearance order AFTER the
 * zero-alt filter (A.py:1041, A.py:851-855).  Posteriors are row-major fp64:
 * q_z [n_cells, K], q_y [n_snvs, C].
 */
#ifndef ARBORIST_B200_H
#define ARBORIST_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ARB_MAX_CLONES 32 /* K <= 32 in this release (one lane per clone) */
#define ARB_ABI_VERSION 2 /* 2: arb_fit_trees_multi, arb_get_fits, ARB_ERR_DECLINED, profile slots 20-23; "groups" option gone */

typedef enum arb_status {
    ARB_OK = 0,
    ARB_ERR_INVALID = -1,     /* bad argument / call order */
    ARB_ERR_CUDA = -2,        /* CUDA runtime failure (message has the detail) */
    ARB_ERR_UNSUPPORTED = -3, /* outside this release's limits (K > 32, > 65535 count pairs) */
    ARB_ERR_NOMEM = -4,       /* device or host allocation failed */
    ARB_ERR_DECLINED = -5     /* arb_csv_feed only: the file is not a plain integer table; use a host reader */
} arb_status;

typedef struct arb_ctx arb_ctx;

/* ---- library / context ------------------------------------------------------ */
int arb_abi_version(void);
const char *arb_last_error(void);
int arb_device_count(int *count);
int arb_ctx_create(int device, arb_ctx **out);
int arb_ctx_destroy(arb_ctx *ctx);

/* ---- ingest: replaces A.py:1034-1069 ----------------------------------------
 * arb_ingest_coo takes the long-format table as integer codes: cell_code/snv_code
 * are first-appearance factorisations of the UNFILTERED id columns (pandas
 * factorize / Series.unique order, A.py:851-852), a negative snv_code marks a row
 * to drop.  On the device: zero-alt SNV filter (A.py:1034-1041), re-indexing in
 * first-appearance order of the surviving rows, dictionary coding of the distinct
 * (alt,total) pairs, binomial log-pmf table for present/absent (A.py:838-846,
 * p_present = 0.5 - alpha + 0.5*alpha/3, p_absent = alpha/3), and cell-major (CSR)
 * + SNV-major (CSC) copies with rows sorted by index (A.py:423-456, 459-488). */
int arb_ingest_coo(arb_ctx *ctx, int64_t n_rows, const int32_t *cell_code,
                   const int32_t *snv_code, const int32_t *alt, const int32_t *total,
                   int32_t n_cell_codes, int32_t n_snv_codes, double alpha);
/* The same ingest for tables whose ids are 64-bit integers: the first-appearance
 * factorisation of the two id columns (pandas `unique` / `factorize` order, A.py:851-855) is done
 * on the device as well (stable radix sort by id, first row of every id, rank of the first rows),
 * then everything arb_ingest_coo does.  alt / total are int64 as pandas holds them.  The distinct
 * ids in code order come back through arb_get_id_uniques ([*n_cell_ids] / [*n_snv_ids] int64). */
int arb_ingest_ids(arb_ctx *ctx, int64_t n_rows, const int64_t *cell_id, const int64_t *snv_id,
                   const int64_t *alt, const int64_t *total, double alpha, int32_t *n_cell_ids,
                   int32_t *n_snv_ids);
int arb_get_id_uniques(arb_ctx *ctx, int64_t *cell_uniques, int64_t *snv_uniques);

/* Seam-level ingest from the reference's own arrays (the arguments of run(...),
 * A.py:1093-1107): CSR pointer, SNV index per entry, interleaved [log_present,
 * log_absent] per entry (A.py:478-486).  The CSC copy is derived on the device. */
int arb_ingest_csr(arb_ctx *ctx, int32_t n_cells, int32_t n_snvs, int64_t nnz,
                   const int32_t *cell_ptr, const int32_t *snv_idx, const double *log_likes);

/* sizes after ingest; any pointer may be NULL */
int arb_get_dims(arb_ctx *ctx, int32_t *n_cells, int32_t *n_snvs, int64_t *nnz,
                 int32_t *n_pairs);
/* cell_order[i] / snv_order[j] = input code of final row i / column j (coo ingest only) */
int arb_get_orders(arb_ctx *ctx, int32_t *cell_order, int32_t *snv_order);
/* the distinct (alt,total) pairs, in table order (coo ingest only) */
int arb_get_count_pairs(arb_ctx *ctx, int32_t *alt, int32_t *total);
/* read / overwrite the per-pair log-likelihood table (e.g. with scipy's doubles,
 * so the device uses bit-identical values to A.py:838-846) */
int arb_get_loglik_table(arb_ctx *ctx, double *log_present, double *log_absent);
int arb_set_loglik_table(arb_ctx *ctx, int32_t n_pairs, const double *log_present,
                         const double *log_absent);
/* canonical device matrix back on the host, reference layout (for tests):
 * CSR if snv_major == 0 else CSC; log_likes interleaved [lp, la] */
int arb_export_sparse(arb_ctx *ctx, int snv_major, int32_t *ptr, int32_t *idx,
                      double *log_likes);

/* ---- prior: replaces initialize_q_y, A.py:523-569 ---------------------------
 * label_idx[j] = cluster index of SNV j's input label, or -1 for the uniform row
 * (label missing / falsy / not a tree node, A.py:564-565). */
int arb_set_prior_labels(arb_ctx *ctx, const int32_t *label_idx, int32_t n_clusters,
                         double gamma);
int arb_set_prior(arb_ctx *ctx, const double *q_y_init, int32_t n_clusters);
int arb_get_prior(arb_ctx *ctx, double *q_y_init);

/* ---- the fit: replaces the tree loop A.py:1089-1126 --------------------------
 * parent [n_trees, K]: parent clone index of every clone, -1 for a root; the
 * ancestor-closure genotype matrix (enumerate_presence, A.py:491-520) is built on
 * the device.  normal_idx = clone index of the normal clone (in-degree-0 node of
 * tree 0, A.py:1012-1013).  Runs run_variational_inference (A.py:679-816) for
 * every tree: elbo[t] = best ELBO, iters[t] = loop bodies executed.  elbo_trace
 * (optional, [n_trees, max_iter+1], NaN-padded) receives ELBO_0 and every
 * iteration's ELBO.  The best q_z/q_y of every tree stay on the device when
 * keep_posteriors != 0 (else they are refit on demand by arb_get_fit). */
#define ARB_MODE_VI 0 /* run_variational_inference, A.py:679 */
#define ARB_MODE_ML 1 /* run_simple_max_likelihood, A.py:598 */
int arb_fit_trees(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, int32_t normal_idx,
                  const int32_t *parent, int32_t max_iter, double tolerance, int32_t mode,
                  int32_t keep_posteriors, double *elbo, int32_t *iters, double *elbo_trace);
/* same, from explicit genotype matrices presence [n_trees, K-1, K] (0/1 bytes) --
 * the literal first argument of run(...), for inputs that are not forests */
int arb_fit_presence(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint8_t *presence,
                     int32_t max_iter, double tolerance, int32_t mode, int32_t keep_posteriors,
                     double *elbo, int32_t *iters, double *elbo_trace);
/* The same fit spread over n_ctx contexts, normally one per GPU of the box, each holding the SAME ingested matrix
 * and prior: tree t is fitted by context t mod n_ctx as its local tree t / n_ctx (so arb_get_fit(ctxs[t % n_ctx],
 * t / n_ctx, ...) fetches it); one host thread per context, results written into the caller's arrays at the global
 * tree index.  Trees are independent (A.py:1089-1126), so the ELBOs need no collective inside one process; the
 * torch.distributed / NCCL gather in arborist_b200/distributed.py is only for one-process-per-GPU launches.
 * Bit-identical to arb_fit_trees on one context. */
int arb_fit_trees_multi(arb_ctx *const *ctxs, int32_t n_ctx, int32_t n_trees, int32_t n_clones,
                        int32_t normal_idx, const int32_t *parent, int32_t max_iter, double tolerance,
                        int32_t mode, int32_t keep_posteriors, double *elbo, int32_t *iters,
                        double *elbo_trace);
/* bytes the resident best posteriors of n_trees trees would need (for "auto") */
int arb_posterior_bytes(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, int64_t *bytes,
                        int64_t *free_bytes);

/* best posteriors / MAP labels (row argmax, first maximum wins -- treefit.py:105)
 * of tree t of the last fit */
int arb_get_fit(arb_ctx *ctx, int32_t t, double *q_z, double *q_y);
/* the same for trees t0 .. t0+count-1 at once: q_z [count, n_cells, K], q_y [count, n_snvs, K-1] (either may be
 * NULL).  When the posteriors were not kept on the device, the whole range is RE-FITTED AS ONE BATCH (the fit is
 * deterministic: same bits) and stays cached until another range is asked for, so a writer that walks the
 * candidate set (--pickle / --q_z of every fit, A.py:1161-1188) costs ceil(T / count) batched fits.
 * arb_last_fit_profile()[23] counts those re-fits. */
int arb_get_fits(arb_ctx *ctx, int32_t t0, int32_t count, double *q_z, double *q_y);
int arb_get_map(arb_ctx *ctx, int32_t t, int32_t *z, int32_t *y);
/* the genotype matrix the device built for tree t, [K-1, K] 0/1 bytes */
int arb_get_presence(arb_ctx *ctx, int32_t t, uint8_t *presence);

/* ---- kernel-level hooks (parity tests against A.py:200-420 called directly) ---
 * presence [K-1, K] 0/1 bytes; q_y [m, K-1]; q_z [n, K]. */
/* compute_q_z_sparse (A.py:200): q_z from q_y (NOT its log; zeros allowed) */
int arb_qz_update(arb_ctx *ctx, int32_t n_clones, const uint8_t *presence, const double *q_y,
                  double *q_z);
/* compute_q_y_sparse (A.py:325): q_y from q_z and the context's prior */
int arb_qy_update(arb_ctx *ctx, int32_t n_clones, const uint8_t *presence, const double *q_z,
                  double *q_y);
/* compute_likelihood_sparse (A.py:378), compute_entropy(q_z) (A.py:253),
 * compute_kl_divergence(q_y, prior) (A.py:276) */
int arb_elbo_terms(arb_ctx *ctx, int32_t n_clones, const uint8_t *presence, const double *q_z,
                   const double *q_y, double *likelihood, double *entropy, double *kl);

/* ---- counters for bench.py ------------------------------------------------------
 * kernels launched / device milliseconds spent inside the last arb_fit_* call, and
 * CAVI tree-iterations it executed (initial pass = half an iteration) */
int arb_last_fit_stats(arb_ctx *ctx, int64_t *kernel_launches, double *device_ms,
                       double *tree_iterations, double *sweep_ms);
/* per-kernel profile of the last arb_fit_* call, measured with CUDA events on the launching
 * stream: values[0..7] = q(z) sweep ms, q(y) sweep ms, finishing kernels ms, tree-sweeps done
 * by the q(z) / q(y) sweep launches, number of q(z) / q(y) sweep launches, rounds;
 * values[8..19] = q(z) / q(y) finishing ms, then the layout in use: KP, rows per panel, panels
 * (z, y), tiles per warp (z, y), range-proof flags (z, y), stream quads (z, y); values[20] = rounds the
 * host enqueued after the last tree had finished (they find nothing to do), values[21] = 1 when the sparse
 * stream uses 16-bit words, values[22] = times the host blocked on the device inside the round loop,
 * values[23] = batched re-fits done by arb_get_fits / arb_get_map since the last fit */
int arb_last_fit_profile(arb_ctx *ctx, double *values, int32_t n_values);
/* tuning knobs: "slots" (trees in flight), "split" (CTAs per (virtual slot, panel) of a sweep; 0 = chosen per
 * launch), "lookahead" (rounds enqueued ahead of the device, 0 = default 1), "panel_kb" (panel bytes cap),
 * "threads", "no_fold", "w32" (force 32-bit stream words); returns ARB_ERR_INVALID for unknown names */
int arb_set_option(arb_ctx *ctx, const char *name, int64_t value);
/* Device allocations released by a context (arb_ctx_destroy, workspace resizes) are parked in a
 * process-wide pool and handed to the next context that asks for a buffer of about that size, so
 * repeated arborist(...) calls do not pay cudaMalloc / cudaFree for the workspace every time.
 * arb_pool_trim returns everything parked to the driver; ARB_POOL=0 in the environment disables
 * the pool. */
int arb_pool_trim(void);

/* ---- read-count CSV parsed on the device: replaces pd.read_csv of the long table (A.py:1135) followed by
 * arb_ingest_ids, for files whose four columns are plain integers.  The host reads the header line itself
 * (column names -> field indices), then hands over the REST of the file as raw bytes in chunks that end on a
 * line boundary (at most 2^30 bytes each); the columns are parsed into device memory, factorised and ingested
 * there and never exist as host arrays.
 *   arb_csv_begin : field indices of the cell / snv / alt / total columns, fields per line, expected row count
 *                   (a sizing hint; the device columns grow when it is exceeded)
 *   arb_csv_feed  : one chunk; n_rows_total = rows parsed so far.  ARB_ERR_DECLINED (table discarded, device columns freed) when
 *                   a line is not n_fields comma-separated fields with plain integers in the selected columns
 *                   (quotes, empty lines, fractions, text ids): use the host reader + arb_ingest_ids/_coo then.
 *   arb_csv_finish: first-appearance factorisation of the ids + everything arb_ingest_ids does; ids in code
 *                   order through arb_get_id_uniques. */
int arb_csv_begin(arb_ctx *ctx, int32_t col_cell, int32_t col_snv, int32_t col_alt, int32_t col_total,
                  int32_t n_fields, int64_t expected_rows);
int arb_csv_feed(arb_ctx *ctx, const char *bytes, int64_t n_bytes, int64_t *n_rows_total);
int arb_csv_finish(arb_ctx *ctx, double alpha, int64_t *n_rows, int32_t *n_cell_ids, int32_t *n_snv_ids);

/* ---- legacy scorer: replaces the reference's arborist_base.py ---------------------------------------
 * find_cell_assignments (arborist_base.py:102-194) and the numeric part of rank_trees
 * (arborist_base.py:250-362).  The host mirror is arborist_b200/arborist_base.py.
 *
 * arb_base_ingest: the long read-count table restricted to rows with a cluster label.  cell_code[i] in
 * [0, n_cells), cluster_code[i] in [0, n_clusters) (n_clusters <= 32), total / alt the read counts.  Folds
 * the rows into per (cell, cluster) sums of binom.logpmf(alt, total, p) at p = 0.5 - alpha + 0.5 alpha / 3
 * and p = alpha / 3 (arborist_base.py:165-166, 186-188), kept on the device. */
int arb_base_ingest(arb_ctx *ctx, int64_t n_rows, const int32_t *cell_code, const int32_t *cluster_code,
                    const int32_t *total, const int32_t *alt, int32_t n_cells, int32_t n_clusters,
                    double alpha);
/* arb_base_score: likelihood[t] = sum over cells of max over clones of the cell's log-likelihood
 * (arborist_base.py:193).  anc_masks[t * n_clones + c]: bit k set when cluster code k is clone c or one of
 * its ancestors (build_genotypes, arborist_base.py:228-247); tree_masks[t]: bit k set when cluster code k is
 * a node of tree t -- rows of other clusters are dropped, cells left without rows do not count
 * (arborist_base.py:153-160).  n_clones <= 64. */
int arb_base_score(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint32_t *anc_masks,
                   const uint32_t *tree_masks, double *likelihood);
/* arb_base_score_static: arb_base_score for callers that must not allocate device memory once set-up is over.
 * (No counterpart in the reference, which has no device memory to manage; same arguments, same meaning and the same
 * scores as arb_base_score up to the rounding of a differently grouped sum -- about 1e-15 relative.)
 *
 * Who needs it: a host whose device memory is owned by another allocator (a tensor framework's caching allocator, an
 * RMM pool) so that a late cudaMalloc fails or fragments that arena; a scoring loop with a latency budget, because
 * cudaMalloc / cudaFree serialise on a driver lock and can wait for the whole device; a client under a fixed memory
 * quota (MIG slice, MPS limit).  arb_base_score takes three temporary buffers per call from the library's pool, which
 * ends in cudaMalloc when the pool has nothing of that size.  This variant calls no cudaMalloc / cudaFree and no
 * allocating library routine: the staged masks, the per-CTA partial sums of its kernel and the result vector live in
 * one file-scope device array of ARB_BASE_STATIC_BYTES that is part of the library image (every GPU this library's
 * module is loaded on holds those bytes until the process exits).  The tables of arb_base_ingest, which it reads, were
 * allocated by that earlier call.
 *
 * Limit: the call needs about n_trees * (8 * splits + 4 * n_clones + 16) bytes, each of its five sections rounded up to
 * 256 bytes, splits = CTAs per tree <= 64 (1 once there are about 300 trees); beyond ARB_BASE_STATIC_BYTES it returns
 * ARB_ERR_UNSUPPORTED and the message names both numbers -- score the trees in smaller groups then.
 * First call on a device: looks up the array's address on that device (and may load the module there); make it
 * during set-up, before any stream capture begins.
 * Sharing: there is ONE array per GPU, used from offset 0 by every call on that GPU whatever the context, and it keeps
 * the previous call's bytes; the call clears the part that has to start at zero itself.  Calls that overlap in time on
 * the same GPU (two host threads, two contexts) race on it: serialise them. */
#define ARB_BASE_STATIC_BYTES (1 << 20)
int arb_base_score_static(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint32_t *anc_masks,
                          const uint32_t *tree_masks, double *likelihood);
/* arb_base_cells: per-cell detail for the listed trees, all [n_trees][n_cells](x n_clones): log-likelihoods
 * (may be NULL), posteriors over clones exp(L - logsumexp(L)) (arborist_base.py:335-338), index of the first
 * best clone (arborist_base.py:296-298), entropy of the posterior in bits (arborist_base.py:341), and
 * present = 1 when the cell keeps at least one row under the tree's cluster filter. */
int arb_base_cells(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint32_t *anc_masks,
                   const uint32_t *tree_masks, double *loglik, double *posterior, int32_t *clone_idx,
                   double *entropy, uint8_t *present);

#ifdef __cplusplus
}
#endif
#endif /* ARBORIST_B200_H */
