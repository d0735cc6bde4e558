// arb_common.cuh -- shared declarations of libarborist_b200.so (sm_100a only).
//
// One arb_ctx per GPU.  Everything numeric lives on the device; the host side of
// this library only sequences kernels and moves the caller's buffers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/arborist_b200.h"

#define ARB_API extern "C" __attribute__((visibility("default")))

void arb_set_error(const char *fmt, ...);

#define ARB_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            arb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call,                  \
                          cudaGetErrorString(e__));                                      \
            return e__ == cudaErrorMemoryAllocation ? ARB_ERR_NOMEM : ARB_ERR_CUDA;      \
        }                                                                                \
    } while (0)

#define ARB_TRY(call)                 \
    do {                              \
        int rc__ = (call);            \
        if (rc__ != ARB_OK) return rc__; \
    } while (0)

#define ARB_REQUIRE(cond, code, ...)      \
    do {                                  \
        if (!(cond)) {                    \
            arb_set_error(__VA_ARGS__);   \
            return (code);                \
        }                                 \
    } while (0)

// ---- device buffer with size tracking -----------------------------------------------
// Released allocations go to a process-wide pool (arb_ingest.cu) instead of back to the driver: a
// service that calls arborist(...) repeatedly on same-shaped inputs would otherwise spend tens of
// milliseconds per call in cudaMalloc / cudaFree of its ~50 GB workspace.  ARB_POOL=0 disables it,
// arb_pool_trim() empties it.
void *arb_pool_alloc(size_t n, size_t *got);       // nullptr on failure (error already set)
void arb_pool_free(void *p, size_t n);
size_t arb_pool_bytes(int device);                 // bytes parked in the pool for this device
void arb_pool_flush_device(int device);            // return them to the driver

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
    int reserve(size_t n) {          // grows only; contents are NOT preserved
        if (n <= bytes && p) return ARB_OK;
        if (p) arb_pool_free(p, bytes);
        p = nullptr;
        bytes = 0;
        if (n == 0) return ARB_OK;
        size_t got = 0;
        p = arb_pool_alloc(n, &got);
        if (!p) return ARB_ERR_NOMEM;
        bytes = got;
        return ARB_OK;
    }
    void release() {
        if (p) arb_pool_free(p, bytes);
        p = nullptr;
        bytes = 0;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

// The fit enqueues rounds ahead of the device: the Control block of every round is copied to a pinned ring the
// host reads ARB_LOOKAHEAD rounds later (never waiting on the round it is about to enqueue).
#define ARB_RING 8
#define ARB_LOOKAHEAD 1      // measured 1 / 2 / 3 at configs 2 and 3 and at 125 trees: 1 is never slower (profiles/r2i_*)
#ifndef ARB_SWEEP_WARPS
#define ARB_SWEEP_WARPS 16    // warps per sweep CTA; the layouts are dealt out to this many warps
#endif

// Panelled sparse layout for one padded clone width KP (8, 16 or 32 doubles per row of
// the gathered operand).  "z": cell-major rows, SNVs cut into panels (q(z) sweep);
// "y": SNV-major rows, cells cut into panels (q(y) sweep).  Tile slot
// x = (panel * ARB_SWEEP_WARPS + warp) * tiles_per_warp + u: desc[x] = {first quad, quads in
// the A region, quads in total, 0}; rows[x * TW + g] = row of the tile's g-th segment or -1;
// stream = uint4 [quad][TW].  See arb_ingest.cu "panel layouts".
struct PanelLayout {
    int kp = 0;
    int panel_rows = 0;      // max(prz, pry): sizes the shared-memory panel
    // Panels are INTERLEAVED: gathered index c belongs to panel c % np and is row c / np of it, so the
    // nonzeros of a row spread evenly over the panels whatever the id order of the input (first-appearance
    // indices put the cells of the first SNVs next to each other).  The gathered operand is stored in that
    // order: row arb_perm(c, np, pr) of [np * pr] rows per virtual slot, every panel one contiguous block.
    int prz = 0, pry = 0;    // rows per panel of the q(z) sweep (S rows of SNVs) / the q(y) sweep (R rows of cells)
    int np_z = 0, np_y = 0;  // number of panels
    int uz = 0, uy = 0;      // tiles per warp per panel
    int fold_z = 0, fold_y = 0;   // 1: a tile's running product provably stays inside the double range
    int w16 = 0;                  // 1: 16-bit stream words {row | code << 11} (<= 32 pair codes, < 2048 panel rows)
    int64_t quads_z = 0, quads_y = 0;   // stream length in quads (x TW x 16 bytes)
    DevBuf zdesc, zrows, zstream;
    DevBuf ydesc, yrows, ystream;
    void release()
    {
        for (DevBuf *b : {&zdesc, &zrows, &zstream, &ydesc, &yrows, &ystream}) b->release();
        kp = 0;
    }
};

struct FitStats {
    int64_t launches = 0;
    double device_ms = 0, tree_iterations = 0, sweep_ms = 0;
    // per-kernel profile of the last fit (CUDA events on the launching stream)
    double sweep_z_ms = 0, sweep_y_ms = 0, finish_ms = 0, finish_z_ms = 0, finish_y_ms = 0;
    double tree_sweeps_z = 0, tree_sweeps_y = 0, launches_z = 0, launches_y = 0, rounds = 0;
    double host_waits = 0, idle_rounds = 0, refit_batches = 0;   // blocking waits of the round pipeline, rounds enqueued
                                                                   // after the last tree finished, arb_get_fits batches
};

struct arb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    int sm_count = 0;
    size_t smem_optin = 0;

    // ---- canonical matrix, sorted by (cell, snv) --------------------------------
    bool ingested = false, from_coo = false;
    int32_t n_cells = 0, n_snvs = 0, n_pairs = 0;
    int64_t nnz = 0;
    DevBuf coo_cell, coo_snv, coo_code;   // int32, int32, uint16
    DevBuf cell_order, snv_order;         // int32 (coo ingest)
    DevBuf cell_uniques, snv_uniques;     // int64 ids in first-appearance order (arb_ingest_ids)
    int32_t n_cell_uniques = 0, n_snv_uniques = 0;
    std::vector<int32_t> pair_alt, pair_total;
    std::vector<int64_t> pair_count;
    std::vector<double> tab_lp, tab_la;   // host copies of the table
    DevBuf tab_dw;                        // double2 [n_pairs]: (expm1(lp-la), lp-la); (0, 0) for "deep" pairs
    // "Deep" pairs: log-likelihood ratio w = lp - la outside [ARB_W_LO, ARB_W_HI] (deep coverage: alt = 0 with
    // total >= 14, or >= 83 alt reads, at the default alpha).  The product form 1 + expm1(w) S loses the reference's
    // accuracy there (cancellation against S -> 1 for very negative w, overflow for large w), so their nonzeros
    // are neutral in the sweeps (factor exactly 1) and are added in the log domain by the finishing kernels from
    // two small side lists (arb_build_side_lists; finish_row in arb_fit.cu).
    std::vector<uint8_t> pair_deep;
    bool side_dirty = true;
    int64_t n_side = 0;
    DevBuf side_z_ptr, side_z_idx, side_z_w;   // CSR over cells: SNV index and w of every deep nonzero
    DevBuf side_y_ptr, side_y_idx, side_y_w;   // CSR over SNVs: cell index and w
    DevBuf tab_w_dev, tab_deep_dev;            // per pair code: true w, deep flag (k_first_pass_general)
    std::vector<double> h_dw;             // host copy of tab_dw (d, w interleaved)
    double sum_la = 0;                    // sum over nonzeros of log_absent
    int renorm_every = 16;
    double alpha = 0;

    // ---- prior ---------------------------------------------------------------------
    int32_t n_clusters = 0;
    DevBuf q0, logq0, logq0e, q0_hard;    // [m, C]: q, log q (-inf for 0), log(q + 1e-12), one-hot argmax of q
    bool prior_general = false;           // some prior row does not sum to one (K = 2: the row is [gamma]; a caller's
                                          // own q_y_init): the first pass runs k_first_pass_general (arb_fit.cu)
    DevBuf row_ptr, tab_la_dev;           // CSR pointer of the canonical matrix / log_absent per pair code, for it

    // ---- layouts (index 0: KP 8, 1: KP 16, 2: KP 32) ----------------------------------
    PanelLayout layout[3];
    int64_t opt_slots = 0, opt_panel_kb = 0, opt_threads = 0, opt_no_fold = 0, opt_w32 = 0, opt_split = 0,
            opt_lookahead = 0;

    // ---- fit workspace -----------------------------------------------------------------
    int ws_kp = 0, ws_slots = 0, ws_k = 0;
    cudaEvent_t rev[ARB_RING][7] = {};            // per round in flight: 6 profile events + "Control snapshot arrived"
    cudaEvent_t fit_ev[2] = {};                   // start / end of a fit
    void *h_ctl = nullptr;                        // pinned: ARB_RING Control snapshots + the block uploaded at the start
    DevBuf xs, xr, qz_cur, qy_cur, part_z, part_g, part_y, red, slot_state, lists, masks;
    // results of the last fit
    int32_t fit_trees = 0, fit_k = 0, fit_max_iter = 0, fit_mode = 0;
    double fit_tol = 0;
    bool fit_kept = false;
    DevBuf best_qz, best_qy, d_elbo, d_iters, d_trace, tree_masks;
    std::vector<uint32_t> h_masks;        // per tree: K column masks then C row masks
    // posteriors re-fitted on demand when they were not kept: trees batch_t0 .. batch_t0 + batch_n - 1
    // of the last fit sit in batch_qz / batch_qy (arb_get_fits re-fits a whole batch in one go)
    int batch_t0 = -1, batch_n = 0;
    DevBuf batch_qz, batch_qy, batch_out;
    FitStats stats;
    DevBuf scratch;                       // cub temp storage

    // ---- legacy scorer (arb_base.cu): per (cell, cluster) log-pmf sums and row counts ----------
    DevBuf base_ta, base_tb, base_cnt;
    int32_t base_cells = 0, base_clusters = 0;

    // ---- CSV rows parsed so far (arb_csv_begin / feed / finish): int64 cell, snv, alt, total ----------
    DevBuf csv_col[4];
    int64_t csv_rows = 0, csv_cap = 0;
    int32_t csv_field[4] = {0, 0, 0, 0}, csv_fields = 0;
    bool csv_open = false;
};

__host__ __device__ static inline int64_t arb_perm(int64_t c, int np, int pr) { return (c % np) * (int64_t)pr + c / np; }
static inline int kp_index(int kp) { return kp == 8 ? 0 : kp == 16 ? 1 : 2; }
// The sweeps always run on 16-column packed rows ("virtual slots", arb_fit.cu): the columns of the
// trees in flight are laid side by side in units of four, so no lane works on a padding column.
#define ARB_W_LO (-9.0)     // exp(9) * 2^-53 ~ 1e-12: relative accuracy of 1 + expm1(w) S at S = 1
#define ARB_W_HI 600.0
#define ARB_VW 16
static inline int kp_for(int) { return ARB_VW; }

// ---- internal entry points shared between translation units ---------------------------
int arb_build_layout(arb_ctx *ctx, int kp);                       // arb_ingest.cu
int arb_refresh_tables(arb_ctx *ctx);                             // arb_ingest.cu
int arb_build_side_lists(arb_ctx *ctx);                           // arb_ingest.cu (no-op unless the table changed)
