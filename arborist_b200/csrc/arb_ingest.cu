// arb_ingest.cu -- context management and the read-count ingest of libarborist_b200.so.
//
// Replaces the host-side Python of the reference's ingest
// (/root/reference/src/arborist/arborist.py, "A.py"):
//   A.py:1034-1041  zero-alt SNV filter            -> k_alt_sum / k_keep_flags / compaction
//   A.py:838-846    binomial log-pmf per row       -> dictionary of distinct (alt,total) pairs
//                                                     + a per-pair table (k_logpmf_table)
//   A.py:851-855    first-appearance index maps    -> k_first_row + sort (cells), scan (SNVs)
//   A.py:459-488    COO build (Python loop)        -> k_compact
//   A.py:423-456    CSR / CSC build (Python loops) -> radix sorts into the panelled layouts
//
// Data layout in HBM after ingest: a canonical COO sorted by (cell, snv) with a 16-bit
// pair code per nonzero (10 B/nnz instead of the reference's 20 B/nnz per orientation),
// plus, per padded clone width KP, two panelled streams of 4 B/nnz (see PanelLayout).
#include "arb_common.cuh"

#include <cub/cub.cuh>

#include <algorithm>
#include <limits.h>
#include <math.h>
#include <numeric>
#include <stdarg.h>
#include <string.h>

// ---------------------------------------------------------------------------------------
// errors / library
// ---------------------------------------------------------------------------------------
static thread_local std::string g_last_error;

void arb_set_error(const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
}

ARB_API const char *arb_last_error(void) { return g_last_error.c_str(); }

// ---------------------------------------------------------------------------------------
// device allocation pool (see DevBuf in arb_common.cuh)
// ---------------------------------------------------------------------------------------
#include <mutex>
struct PoolBlock { void *p; size_t bytes; int device; };
static std::mutex g_pool_mu;
static std::vector<PoolBlock> g_pool;
static int pool_enabled()
{
    static int on = -1;
    if (on < 0) {
        const char *e = getenv("ARB_POOL");
        on = !(e && e[0] == '0');
    }
    return on;
}

// ARB_POOL_POISON=1 (debugging aid): fill every block handed out with 0xff bytes (NaN doubles, -1 ints),
// so that a kernel relying on the contents of fresh memory fails every time instead of now and then.
static void pool_poison(void *p, size_t n)
{
    static int on = -1;
    if (on < 0) {
        const char *e = getenv("ARB_POOL_POISON");
        on = e && e[0] == '1';
    }
    if (on && p) {
        cudaDeviceSynchronize();
        cudaMemset(p, 0xff, n);
        cudaDeviceSynchronize();
    }
}

static int pool_trace()
{
    static int on = -1;
    if (on < 0) {
        const char *e = getenv("ARB_POOL_TRACE");
        on = e && e[0] == '1';
    }
    return on;
}

static void pool_flush_locked(int device)
{
    if (pool_trace()) fprintf(stderr, "[arb pool] flush device %d (%zu blocks parked)\n", device, g_pool.size());
    for (size_t i = 0; i < g_pool.size();) {
        if (g_pool[i].device == device) {
            cudaFree(g_pool[i].p);
            g_pool[i] = g_pool.back();
            g_pool.pop_back();
        } else {
            ++i;
        }
    }
}

void *arb_pool_alloc(size_t n, size_t *got)
{
    int device = 0;
    cudaGetDevice(&device);
    std::lock_guard<std::mutex> lk(g_pool_mu);
    // best fit among the parked blocks of this device, at most 12.5 % larger than asked
    int best = -1;
    for (size_t i = 0; i < g_pool.size(); ++i)
        if (g_pool[i].device == device && g_pool[i].bytes >= n && g_pool[i].bytes <= n + n / 8 + 4096 &&
            (best < 0 || g_pool[i].bytes < g_pool[best].bytes))
            best = (int)i;
    if (best >= 0) {
        void *p = g_pool[best].p;
        *got = g_pool[best].bytes;
        g_pool[best] = g_pool.back();
        g_pool.pop_back();
        pool_poison(p, *got);
        return p;
    }
    void *p = nullptr;
    if (pool_trace() && n >= (1u << 20)) fprintf(stderr, "[arb pool] miss: cudaMalloc(%zu)\n", n);
    cudaError_t e = cudaMalloc(&p, n);
    if (e != cudaSuccess) {              // give the parked memory back to the driver and try again
        cudaGetLastError();
        pool_flush_locked(device);
        e = cudaMalloc(&p, n);
    }
    if (e != cudaSuccess) {
        arb_set_error("cudaMalloc(%zu bytes): %s", n, cudaGetErrorString(e));
        cudaGetLastError();
        return nullptr;
    }
    *got = n;
    pool_poison(p, n);
    return p;
}

void arb_pool_free(void *p, size_t n)
{
    if (!p) return;
    if (!pool_enabled()) {
        cudaFree(p);
        return;
    }
    int device = 0;
    cudaGetDevice(&device);
    std::lock_guard<std::mutex> lk(g_pool_mu);
    g_pool.push_back(PoolBlock{p, n, device});
}

// give this device's parked blocks back to the driver (memory-tight fits: what stays parked is invisible to every
// other allocator of the process -- NCCL, torch -- although cudaMemGetInfo-based sizing here counts it as free)
void arb_pool_flush_device(int device)
{
    std::lock_guard<std::mutex> lk(g_pool_mu);
    pool_flush_locked(device);
}

size_t arb_pool_bytes(int device)
{
    std::lock_guard<std::mutex> lk(g_pool_mu);
    size_t t = 0;
    for (const PoolBlock &b : g_pool)
        if (b.device == device) t += b.bytes;
    return t;
}

ARB_API int arb_pool_trim(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return ARB_OK; }
    int cur = 0;
    cudaGetDevice(&cur);
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (int d = 0; d < n; ++d) {
        bool any = false;
        for (const PoolBlock &b : g_pool) any |= b.device == d;
        if (!any) continue;
        cudaSetDevice(d);
        pool_flush_locked(d);
    }
    cudaSetDevice(cur);
    return ARB_OK;
}

ARB_API int arb_abi_version(void) { return ARB_ABI_VERSION; }

ARB_API int arb_device_count(int *count)
{
    ARB_REQUIRE(count, ARB_ERR_INVALID, "arb_device_count: null pointer");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    *count = n;
    return ARB_OK;
}

ARB_API int arb_ctx_create(int device, arb_ctx **out)
{
    ARB_REQUIRE(out, ARB_ERR_INVALID, "arb_ctx_create: null output pointer");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        arb_set_error("arb_ctx_create: no usable CUDA device (%s); this library has no CPU path",
                      e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        return ARB_ERR_CUDA;
    }
    ARB_REQUIRE(device >= 0 && device < n, ARB_ERR_INVALID,
                "arb_ctx_create: device %d out of range (0..%d)", device, n - 1);
    ARB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    ARB_CUDA(cudaGetDeviceProperties(&prop, device));
    ARB_REQUIRE(prop.major >= 10, ARB_ERR_UNSUPPORTED,
                "arb_ctx_create: device %d is sm_%d%d; this library is built for sm_100a only",
                device, prop.major, prop.minor);
    arb_ctx *ctx = new arb_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        arb_set_error("cudaStreamCreate: %s", cudaGetErrorString(e));
        delete ctx;
        return ARB_ERR_CUDA;
    }
    *out = ctx;
    return ARB_OK;
}

ARB_API int arb_ctx_destroy(arb_ctx *ctx)
{
    if (!ctx) return ARB_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf *bufs[] = {&ctx->coo_cell, &ctx->coo_snv, &ctx->coo_code, &ctx->cell_order,
                      &ctx->snv_order, &ctx->tab_dw, &ctx->q0, &ctx->logq0, &ctx->logq0e, &ctx->q0_hard,
                      &ctx->xs, &ctx->xr, &ctx->qz_cur, &ctx->qy_cur, &ctx->part_z, &ctx->part_g,
                      &ctx->part_y, &ctx->red, &ctx->slot_state, &ctx->lists, &ctx->masks,
                      &ctx->best_qz, &ctx->best_qy, &ctx->d_elbo, &ctx->d_iters, &ctx->d_trace,
                      &ctx->tree_masks, &ctx->batch_qz, &ctx->batch_qy, &ctx->batch_out, &ctx->row_ptr, &ctx->tab_la_dev,
                      &ctx->side_z_ptr, &ctx->side_z_idx, &ctx->side_z_w, &ctx->side_y_ptr, &ctx->side_y_idx, &ctx->side_y_w,
                      &ctx->tab_w_dev, &ctx->tab_deep_dev, &ctx->scratch, &ctx->cell_uniques,
                      &ctx->snv_uniques, &ctx->base_ta, &ctx->base_tb, &ctx->base_cnt,
                      &ctx->csv_col[0], &ctx->csv_col[1], &ctx->csv_col[2], &ctx->csv_col[3]};
    for (DevBuf *b : bufs) b->release();
    for (auto &l : ctx->layout) l.release();
    for (auto &r : ctx->rev)
        for (auto &e : r) if (e) cudaEventDestroy(e);
    for (auto &e : ctx->fit_ev) if (e) cudaEventDestroy(e);
    if (ctx->h_ctl) cudaFreeHost(ctx->h_ctl);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return ARB_OK;
}

ARB_API int arb_set_option(arb_ctx *ctx, const char *name, int64_t value)
{
    ARB_REQUIRE(ctx && name, ARB_ERR_INVALID, "arb_set_option: null argument");
    if (!strcmp(name, "slots")) ctx->opt_slots = value;
    else if (!strcmp(name, "panel_kb")) {
        ctx->opt_panel_kb = value;
        for (auto &l : ctx->layout) l.release();
        ctx->ws_kp = 0;                      // partial buffers are sized by the panel count
    } else if (!strcmp(name, "threads")) ctx->opt_threads = value;
    else if (!strcmp(name, "split")) ctx->opt_split = value;           // 0 = chosen per launch, n = forced
    else if (!strcmp(name, "lookahead")) {                            // rounds enqueued ahead of the device
        ARB_REQUIRE(value >= 0 && value <= ARB_RING - 3, ARB_ERR_INVALID, "arb_set_option: lookahead=%lld outside 0..%d",
                    (long long)value, ARB_RING - 3);
        ctx->opt_lookahead = value;
    }
    else if (!strcmp(name, "w32")) {          // force 32-bit stream words
        ctx->opt_w32 = value;
        for (auto &l : ctx->layout) l.release();
        ctx->ws_kp = 0;
    }
    else if (!strcmp(name, "no_fold")) {
        ctx->opt_no_fold = value;
        for (auto &l : ctx->layout) l.release();
        ctx->ws_kp = 0;
    }
    else {
        arb_set_error("arb_set_option: unknown option '%s'", name);
        return ARB_ERR_INVALID;
    }
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------------------
static inline int grid_for(int64_t n, int block) { return (int)std::max<int64_t>(1, (n + block - 1) / block); }

template <class F> static int with_cub(arb_ctx *ctx, F f)
{
    size_t bytes = 0;
    ARB_CUDA(f(nullptr, bytes));
    ARB_TRY(ctx->scratch.reserve(std::max<size_t>(bytes, 16)));
    bytes = ctx->scratch.bytes;
    ARB_CUDA(f(ctx->scratch.p, bytes));
    return ARB_OK;
}

static int bits_for(uint64_t max_value)
{
    int b = 1;
    while (b < 64 && (max_value >> b)) ++b;
    return b;
}

__global__ void k_iota(int32_t *v, int64_t n)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int32_t)i;
}

// ---------------------------------------------------------------------------------------
// COO ingest kernels
// ---------------------------------------------------------------------------------------
// A.py:1034: read_counts.groupby("snv")["alt"].sum()
__global__ void k_alt_sum(const int32_t *snv, const int32_t *alt, int64_t n, unsigned long long *sum)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n && snv[i] >= 0) atomicAdd(&sum[snv[i]], (unsigned long long)(long long)alt[i]);
}

// A.py:1037: valid SNVs are those with a strictly positive alt sum
__global__ void k_keep_flags(const unsigned long long *sum, int32_t n, int32_t *keep)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keep[i] = ((long long)sum[i] > 0) ? 1 : 0;
}

__global__ void k_row_keep(const int32_t *snv, const int32_t *keep_snv, int64_t n, int32_t *row_keep)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) row_keep[i] = (snv[i] >= 0 && keep_snv[snv[i]]) ? 1 : 0;
}

// first surviving row of every cell code: Series.unique() order after the filter (A.py:851)
__global__ void k_first_row(const int32_t *cell, const int32_t *row_keep, int64_t n, int32_t *first)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n && row_keep[i]) atomicMin(&first[cell[i]], (int32_t)i);
}

__global__ void k_fill_i32(int32_t *v, int64_t n, int32_t x)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) v[i] = x;
}

__global__ void k_cell_rank(const int32_t *sorted_first, const int32_t *sorted_cell, int32_t n,
                            int32_t *cell_new, int32_t *cell_order, int32_t *n_cells)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool live = sorted_first[i] != INT_MAX;
    cell_new[sorted_cell[i]] = live ? i : -1;
    if (live) cell_order[i] = sorted_cell[i];
    // the sorted array is [live..., dead...]; the boundary is the count
    if (live && (i == n - 1 || sorted_first[i + 1] == INT_MAX)) *n_cells = i + 1;
}

__global__ void k_snv_order(const int32_t *keep, const int32_t *snv_new, int32_t n, int32_t *order)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && keep[i]) order[snv_new[i]] = i;
}

// A.py:459-488: the COO triple of every surviving row, keyed for the canonical sort
__global__ void k_compact(const int32_t *cell, const int32_t *snv, const int32_t *alt,
                          const int32_t *total, const int32_t *row_keep, const int32_t *row_pos,
                          const int32_t *cell_new, const int32_t *snv_new, int64_t n, int32_t n_snvs,
                          uint64_t *key, uint64_t *pairkey)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n || !row_keep[i]) return;
    const int32_t o = row_pos[i];
    key[o] = (uint64_t)cell_new[cell[i]] * (uint64_t)n_snvs + (uint64_t)snv_new[snv[i]];
    pairkey[o] = ((uint64_t)(uint32_t)total[i] << 32) | (uint64_t)(uint32_t)alt[i];
}

__global__ void k_encode_pairs(const uint64_t *pairkey, int64_t n, const uint64_t *uniq, int32_t n_uniq,
                               const int32_t *code_of_rank, uint16_t *code)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t k = pairkey[i];
    int lo = 0, hi = n_uniq - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (uniq[mid] < k) lo = mid + 1; else hi = mid;
    }
    code[i] = (uint16_t)code_of_rank[lo];
}

__global__ void k_gather_canonical(const uint64_t *sorted_key, const int32_t *perm, const uint16_t *code,
                                   int64_t n, int32_t n_snvs, int32_t *coo_cell, int32_t *coo_snv,
                                   uint16_t *coo_code)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t k = sorted_key[i];
    coo_cell[i] = (int32_t)(k / (uint64_t)n_snvs);
    coo_snv[i] = (int32_t)(k % (uint64_t)n_snvs);
    coo_code[i] = code[perm[i]];
}

// scipy.stats.binom.logpmf as the reference calls it (A.py:838-846); scipy's formula
// (_discrete_distns.py, binom_gen._logpmf):
//   gammaln(n+1) - (gammaln(k+1) + gammaln(n-k+1)) + xlogy(k, p) + xlog1py(n-k, -p)
__global__ void k_logpmf_table(const int32_t *alt, const int32_t *total, int32_t n, double p_present,
                               double p_absent, double *lp, double *la)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double k = alt[i], nn = total[i];
    const double comb = lgamma(nn + 1.0) - (lgamma(k + 1.0) + lgamma(nn - k + 1.0));
    auto xlogy = [](double x, double y) { return x == 0.0 ? 0.0 : x * log(y); };
    auto xlog1py = [](double x, double y) { return x == 0.0 ? 0.0 : x * log1p(y); };
    const bool bad = (k < 0.0) || (k > nn);
    lp[i] = bad ? -INFINITY : comb + xlogy(k, p_present) + xlog1py(nn - k, -p_present);
    la[i] = bad ? -INFINITY : comb + xlogy(k, p_absent) + xlog1py(nn - k, -p_absent);
}

// per pair: d = exp(lp - la) - 1 (likelihood ratio minus one) and w = lp - la
__global__ void k_table_derive(const double *lp, const double *la, int32_t n, double2 *dw)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double w = lp[i] - la[i];
    dw[i] = make_double2(expm1(w), w);
}

// ---------------------------------------------------------------------------------------
// CSR (seam-level) ingest kernels
// ---------------------------------------------------------------------------------------
__global__ void k_csr_keys(const int32_t *ptr, const int32_t *col, int32_t n_rows, int32_t n_cols,
                           int64_t nnz, uint64_t *key)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nnz) return;
    int lo = 0, hi = n_rows - 1;          // last row with ptr[row] <= i
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((int64_t)ptr[mid] <= i) lo = mid; else hi = mid - 1;
    }
    key[i] = (uint64_t)lo * (uint64_t)n_cols + (uint64_t)col[i];
}

__global__ void k_split_ll(const double *ll, int64_t nnz, uint64_t *lp_bits, uint64_t *la_bits)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nnz) return;
    lp_bits[i] = (uint64_t)__double_as_longlong(ll[2 * i]);
    la_bits[i] = (uint64_t)__double_as_longlong(ll[2 * i + 1]);
}

__global__ void k_gather_u64(const uint64_t *src, const int32_t *idx, int64_t n, uint64_t *dst)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}

__global__ void k_pair_heads(const uint64_t *lp_bits, const uint64_t *la_bits, const int32_t *order,
                             int64_t n, int32_t *head)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i == 0) { head[0] = 1; return; }
    const int32_t a = order[i], b = order[i - 1];
    head[i] = (lp_bits[a] != lp_bits[b] || la_bits[a] != la_bits[b]) ? 1 : 0;
}

__global__ void k_pair_assign(const uint64_t *lp_bits, const uint64_t *la_bits, const int32_t *order,
                              const int32_t *head, const int32_t *rank_incl, int64_t n, int32_t max_pairs,
                              int32_t *raw_code, double *tab_lp, double *tab_la, unsigned long long *count)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t r = rank_incl[i] - 1, o = order[i];
    raw_code[o] = r;
    if (r < max_pairs) {
        atomicAdd(&count[r], 1ull);
        if (head[i]) {
            tab_lp[r] = __longlong_as_double((long long)lp_bits[o]);
            tab_la[r] = __longlong_as_double((long long)la_bits[o]);
        }
    }
}

__global__ void k_remap_codes(const int32_t *raw, const int32_t *map, int64_t n, uint16_t *code)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) code[i] = (uint16_t)map[raw[i]];
}

// ---------------------------------------------------------------------------------------
// panel layouts ("warp-tile ELL")
// ---------------------------------------------------------------------------------------
// The stream of one orientation is cut into segments, one per (panel of the gathered index,
// row).  Inside a segment the nonzeros of the most frequent pair code (code 0) come first,
// the rest follow ordered by code.  Inside a panel the segments are ordered by (code-0 length,
// rest length) descending and dealt out in TILES of TW = 128/KP consecutive segments: a tile is
// what one warp of the sweep kernel works on at a time (GL = KP/4 lanes per segment).  Tile
// number tau of a panel belongs to warp tau % ARB_SWEEP_WARPS; all tiles of one (panel, warp)
// are contiguous in memory, in the order the warp visits them, so the warp reads one unbroken
// stream.  A tile is stored step-major in QUADS: quad q holds steps 4q..4q+3 of every segment
// of the tile as one uint4 per segment ([quad][segment] -> a warp reads TW*16 contiguous bytes).
//   A region: the first 4*floor(min code-0 length / 4) steps; every segment has a code-0
//             nonzero there; word = byte offset of the gathered row inside the staged panel.
//   G region: the rest, up to the longest segment of the tile; word = code << 18 | byte offset.
//             Short segments are padded with NULL words (code 0, offset of an all-zero row kept
//             after the panel): their factor is 1 + d*0 = 1 exactly, so padding needs no predicate.
__global__ void k_panel_keys(const int32_t *row, const int32_t *col, const uint16_t *code, int64_t nnz,
                             int32_t n_rows, int32_t np, uint64_t *key)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nnz) return;
    // interleaved panels (PanelLayout in arb_common.cuh)
    const uint64_t panel = (uint64_t)(col[i] % np), local = (uint64_t)(col[i] / np);
    const uint64_t seg = panel * (uint64_t)n_rows + (uint64_t)row[i];
    key[i] = (((seg << 16) | (uint64_t)code[i]) << 16) | local;
}

// segment starts in key order + end of every segment's code-0 run
__global__ void k_seg_bounds(const uint64_t *sorted_key, int64_t nnz, int64_t n_seg, int32_t *start, int32_t *end_a)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nnz) return;
    const uint64_t k = sorted_key[i];
    const int64_t seg = (int64_t)(k >> 32);
    const int64_t prev = i == 0 ? -1 : (int64_t)(sorted_key[i - 1] >> 32);
    for (int64_t s = prev + 1; s <= seg; ++s) start[s] = (int32_t)i;
    if (i == nnz - 1)
        for (int64_t s = seg + 1; s <= n_seg; ++s) start[s] = (int32_t)nnz;
    if (((k >> 16) & 0xffffu) == 0) {
        const bool last = (i == nnz - 1) || ((sorted_key[i + 1] >> 16) != (k >> 16));
        if (last) end_a[seg] = (int32_t)(i + 1);
    }
}

__global__ void k_seg_keys(const int32_t *start, const int32_t *end_a, int64_t n_seg, int32_t n_rows,
                           uint64_t *key, int32_t *val)
{
    int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (s >= n_seg) return;
    const int32_t len = start[s + 1] - start[s];
    const int32_t la = end_a[s] ? end_a[s] - start[s] : 0, lb = len - la;
    const uint64_t panel = (uint64_t)(s / n_rows);
    key[s] = (panel << 40) | ((uint64_t)(0xFFFFF - min(la, 0xFFFFF)) << 20) | (uint64_t)(0xFFFFF - min(lb, 0xFFFFF));
    val[s] = (int32_t)s;
}

__global__ void k_seg_pos(const int32_t *order, int64_t n_seg, int32_t *pos_of_seg)
{
    int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p < n_seg) pos_of_seg[order[p]] = (int32_t)p;
}

// largest |log2| a segment's running product can reach: sum over its nonzeros of the per-code bound
__global__ void k_seg_expbound(const uint64_t *sorted_key, const int32_t *start, int64_t n_seg,
                               const float *code_up, const float *code_dn, int32_t *out_bits)
{
    int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    float worst = 0.f;
    if (s < n_seg) {
        float up = 0.f, dn = 0.f;
        for (int32_t i = start[s]; i < start[s + 1]; ++i) {
            const uint32_t code = (uint32_t)((sorted_key[i] >> 16) & 0xffffu);
            up += code_up[code];
            dn += code_dn[code];
        }
        worst = fmaxf(up, dn);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) worst = fmaxf(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    if ((threadIdx.x & 31) == 0 && worst > 0.f) atomicMax(out_bits, __float_as_int(worst));
}

// slot x = (panel * WARPS + warp) * U + u  <->  tile tau = u * WARPS + warp of that panel
__global__ void k_tile_shape(const int32_t *order, const int32_t *start, const int32_t *end_a, int32_t n_rows,
                             int64_t n_slots, int32_t U, int32_t tw, int32_t *nq, int32_t *naq)
{
    int64_t x = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (x > n_slots) return;
    if (x == n_slots) { nq[x] = 0; return; }
    const int u = (int)(x % U), w = (int)((x / U) % ARB_SWEEP_WARPS);
    const int64_t panel = x / ((int64_t)U * ARB_SWEEP_WARPS);
    const int64_t tau = (int64_t)u * ARB_SWEEP_WARPS + w, n_tiles = (n_rows + tw - 1) / tw;
    int aq = 0, tq = 0;
    if (tau < n_tiles) {
        int min_la = INT_MAX, max_len = 0;
        for (int g = 0; g < tw; ++g) {
            const int64_t i = tau * tw + g;
            int la = 0, len = 0;
            if (i < n_rows) {
                const int32_t s = order[panel * n_rows + i];
                len = start[s + 1] - start[s];
                la = end_a[s] ? end_a[s] - start[s] : 0;
            }
            min_la = min(min_la, la);
            max_len = max(max_len, len);
        }
        aq = min_la / 4;
        tq = aq + (max_len - 4 * aq + 3) / 4;
    }
    naq[x] = aq;
    nq[x] = tq;
}

__global__ void k_tile_desc(const int32_t *qoff, const int32_t *naq, const int32_t *nq, const int32_t *order,
                            int32_t n_rows, int64_t n_slots, int32_t U, int32_t tw, int4 *desc, int32_t *rows)
{
    int64_t x = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (x >= n_slots) return;
    desc[x] = make_int4(qoff[x], naq[x], nq[x], 0);
    const int u = (int)(x % U), w = (int)((x / U) % ARB_SWEEP_WARPS);
    const int64_t panel = x / ((int64_t)U * ARB_SWEEP_WARPS);
    const int64_t tau = (int64_t)u * ARB_SWEEP_WARPS + w;
    for (int g = 0; g < tw; ++g) {
        const int64_t i = tau * tw + g;
        rows[x * tw + g] = i < n_rows ? order[panel * n_rows + i] % n_rows : -1;
    }
}

__global__ void k_fill_u32(uint32_t *v, int64_t n, uint32_t x)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) v[i] = x;
}

// W16: 16-bit words {row in panel | code << 11} (the sweep shifts them left by 7 = log2 of the 128-byte row);
// else 32-bit words {byte offset of the row | code << 18}
template <bool W16>
__global__ void k_tile_scatter(const uint64_t *sorted_key, int64_t nnz, const int32_t *start,
                               const int32_t *pos_of_seg, const int32_t *qoff, const int32_t *naq, int32_t n_rows,
                               int32_t U, int32_t tw, uint32_t row_bytes, void *stream)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nnz) return;
    const uint64_t k = sorted_key[i];
    const int64_t seg = (int64_t)(k >> 32);
    const uint32_t code = (uint32_t)((k >> 16) & 0xffffu), local = (uint32_t)(k & 0xffffu);
    const int64_t panel = seg / n_rows, idx = (int64_t)pos_of_seg[seg] - panel * n_rows;
    const int64_t tau = idx / tw;
    const int g = (int)(idx % tw), w = (int)(tau % ARB_SWEEP_WARPS), u = (int)(tau / ARB_SWEEP_WARPS);
    const int64_t x = (panel * ARB_SWEEP_WARPS + w) * U + u;
    const int32_t t = (int32_t)(i - start[seg]);
    const bool in_a = t < 4 * naq[x];
    const int64_t at = (((int64_t)qoff[x] + t / 4) * tw + g) * 4 + (t & 3);
    if (W16) {
        reinterpret_cast<uint16_t *>(stream)[at] = (uint16_t)(in_a ? local : ((code << 11) | local));
    } else {
        const uint32_t off = local * row_bytes;
        reinterpret_cast<uint32_t *>(stream)[at] = in_a ? off : ((code << 18) | off);
    }
}

__global__ void k_fill_u16(uint16_t *v, int64_t n, uint16_t x)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) v[i] = x;
}

int arb_sweep_table_entries(const arb_ctx *ctx);

struct StreamOut {
    DevBuf *desc, *rows, *stream;
    int *tiles_per_warp, *fold;
    int64_t *quads;
};

static int build_stream(arb_ctx *ctx, const int32_t *row, const int32_t *col, int32_t n_rows, int32_t panel_rows,
                        int np, int kp, int w16, const float *code_up, const float *code_dn, StreamOut out)
{
    const int64_t nnz = ctx->nnz, n_seg = (int64_t)np * n_rows;
    const int tw = 128 / kp;
    const int U = (int)(((n_rows + tw - 1) / tw + ARB_SWEEP_WARPS - 1) / ARB_SWEEP_WARPS);
    const int64_t n_slots = (int64_t)np * ARB_SWEEP_WARPS * U;
    ARB_REQUIRE(n_seg + 1 < (int64_t)(1 << 30), ARB_ERR_UNSUPPORTED, "layout: %lld segments", (long long)n_seg);
    DevBuf k_in, k_out, start, end_a, sk_in, sk_out, sv_in, order, pos_of_seg, nq, naq, qoff, bound;
    ARB_TRY(k_in.reserve(nnz * 8));
    ARB_TRY(k_out.reserve(nnz * 8));
    for (DevBuf *b : {&start, &end_a, &sv_in, &order, &pos_of_seg}) ARB_TRY(b->reserve((n_seg + 1) * 4));
    for (DevBuf *b : {&nq, &naq, &qoff}) ARB_TRY(b->reserve((n_slots + 1) * 4));
    ARB_TRY(sk_in.reserve(n_seg * 8));
    ARB_TRY(sk_out.reserve(n_seg * 8));
    ARB_TRY(bound.reserve(4));
    ARB_TRY(out.desc->reserve((n_slots + 1) * 16));          // one spare slot: the sweep prefetches past the end
    ARB_TRY(out.rows->reserve((n_slots + 1) * tw * 4));
    ARB_CUDA(cudaMemsetAsync(out.desc->p, 0, (n_slots + 1) * 16, ctx->stream));
    ARB_CUDA(cudaMemsetAsync(out.rows->p, 0xff, (n_slots + 1) * tw * 4, ctx->stream));
    cudaStream_t st = ctx->stream;
    const int B = 256;
    k_panel_keys<<<grid_for(nnz, B), B, 0, st>>>(row, col, ctx->coo_code.as<uint16_t>(), nnz, n_rows, np,
                                                 k_in.as<uint64_t>());
    const int end_bit = std::min(64, 32 + bits_for((uint64_t)n_seg));
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortKeys(tmp, bytes, k_in.as<uint64_t>(), k_out.as<uint64_t>(), nnz, 0, end_bit, st);
    }));
    ARB_CUDA(cudaMemsetAsync(end_a.p, 0, (n_seg + 1) * 4, st));
    ARB_CUDA(cudaMemsetAsync(bound.p, 0, 4, st));
    k_seg_bounds<<<grid_for(nnz, B), B, 0, st>>>(k_out.as<uint64_t>(), nnz, n_seg, start.as<int32_t>(),
                                                 end_a.as<int32_t>());
    k_seg_expbound<<<grid_for(n_seg, B), B, 0, st>>>(k_out.as<uint64_t>(), start.as<int32_t>(), n_seg, code_up, code_dn,
                                                    bound.as<int32_t>());
    k_seg_keys<<<grid_for(n_seg, B), B, 0, st>>>(start.as<int32_t>(), end_a.as<int32_t>(), n_seg, n_rows,
                                                 sk_in.as<uint64_t>(), sv_in.as<int32_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, sk_in.as<uint64_t>(), sk_out.as<uint64_t>(),
                                               sv_in.as<int32_t>(), order.as<int32_t>(), n_seg, 0, 64, st);
    }));
    k_seg_pos<<<grid_for(n_seg, B), B, 0, st>>>(order.as<int32_t>(), n_seg, pos_of_seg.as<int32_t>());
    k_tile_shape<<<grid_for(n_slots + 1, B), B, 0, st>>>(order.as<int32_t>(), start.as<int32_t>(), end_a.as<int32_t>(),
                                                        n_rows, n_slots, U, tw, nq.as<int32_t>(), naq.as<int32_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceScan::ExclusiveSum(tmp, bytes, nq.as<int32_t>(), qoff.as<int32_t>(), n_slots + 1, st);
    }));
    int32_t total_q = 0, bound_bits = 0;
    ARB_CUDA(cudaMemcpyAsync(&total_q, qoff.as<int32_t>() + n_slots, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(&bound_bits, bound.p, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    float worst;
    memcpy(&worst, &bound_bits, 4);
    *out.fold = worst < 1000.f && ctx->n_pairs <= arb_sweep_table_entries(ctx);           // every running product and every partial stays a normal double
    *out.tiles_per_warp = U;
    *out.quads = total_q;
    // the sweep reads three quads ahead of the one it works on: keep that many spare at the end
    const int64_t n_words = ((int64_t)total_q + 4 + 64) * tw * 4;     // + room for the L1 prefetch distance
    ARB_REQUIRE(n_words < ((int64_t)1 << 31), ARB_ERR_UNSUPPORTED, "layout: stream of %lld words", (long long)n_words);
    ARB_TRY(out.stream->reserve(n_words * (w16 ? 2 : 4)));
    if (w16) {
        k_fill_u16<<<grid_for(n_words, B), B, 0, st>>>(out.stream->as<uint16_t>(), n_words, (uint16_t)panel_rows);   // code 0, the zero row
        k_tile_scatter<true><<<grid_for(nnz, B), B, 0, st>>>(k_out.as<uint64_t>(), nnz, start.as<int32_t>(),
                                                             pos_of_seg.as<int32_t>(), qoff.as<int32_t>(),
                                                             naq.as<int32_t>(), n_rows, U, tw, (uint32_t)kp * 8u, out.stream->p);
    } else {
        const uint32_t null_word = (uint32_t)panel_rows * (uint32_t)kp * 8u;       // code 0, the zero row
        k_fill_u32<<<grid_for(n_words, B), B, 0, st>>>(out.stream->as<uint32_t>(), n_words, null_word);
        k_tile_scatter<false><<<grid_for(nnz, B), B, 0, st>>>(k_out.as<uint64_t>(), nnz, start.as<int32_t>(),
                                                              pos_of_seg.as<int32_t>(), qoff.as<int32_t>(),
                                                              naq.as<int32_t>(), n_rows, U, tw, (uint32_t)kp * 8u, out.stream->p);
    }
    k_tile_desc<<<grid_for(n_slots, B), B, 0, st>>>(qoff.as<int32_t>(), naq.as<int32_t>(), nq.as<int32_t>(),
                                                   order.as<int32_t>(), n_rows, n_slots, U, tw, out.desc->as<int4>(),
                                                   out.rows->as<int32_t>());
    ARB_CUDA(cudaStreamSynchronize(st));
    ARB_CUDA(cudaGetLastError());
    for (DevBuf *b : {&k_in, &k_out, &start, &end_a, &sk_in, &sk_out, &sv_in, &order, &pos_of_seg, &nq, &naq, &qoff, &bound})
        b->release();
    return ARB_OK;
}

// smem budget of the sweep kernel: [panel_rows + 1][KP] doubles + pair table + one mbarrier
int arb_sweep_table_entries(const arb_ctx *ctx) { return std::min<int>(ctx->n_pairs, 512); }

int arb_build_layout(arb_ctx *ctx, int kp)
{
    PanelLayout &L = ctx->layout[kp_index(kp)];
    if (L.kp == kp) return ARB_OK;
    ARB_REQUIRE(ctx->ingested, ARB_ERR_INVALID, "no matrix ingested");
    ARB_REQUIRE(ctx->n_pairs <= 16384, ARB_ERR_UNSUPPORTED, "%d distinct (alt,total) pairs: the stream word holds 14 bits",
                ctx->n_pairs);
    size_t budget = ctx->smem_optin - 1024 - (size_t)arb_sweep_table_entries(ctx) * 16;
    if (ctx->opt_panel_kb > 0) budget = std::min<size_t>(budget, (size_t)ctx->opt_panel_kb * 1024);
    int panel_rows = (int)(budget / ((size_t)kp * 8)) - 1;        // one more row of zeros for the padding words
    panel_rows = std::min(panel_rows, 65504) / 32 * 32;
    ARB_REQUIRE(panel_rows >= 32, ARB_ERR_INVALID, "panel budget too small");
    L.np_z = (ctx->n_snvs + panel_rows - 1) / panel_rows;
    L.np_y = (ctx->n_cells + panel_rows - 1) / panel_rows;
    // equal panels: ceil(rows / panels) rows each
    L.prz = (ctx->n_snvs + L.np_z - 1) / L.np_z;
    L.pry = (ctx->n_cells + L.np_y - 1) / L.np_y;
    L.panel_rows = std::max(L.prz, L.pry);
    // 16-bit stream words when the row (11 bits, the zero row included) and the pair code (5 bits) fit
    L.w16 = kp == ARB_VW && L.panel_rows < 2048 && ctx->n_pairs <= 32 && !ctx->opt_w32;
    // per-code bounds on |log2(1 + d x)|, x in [0,1], with a little slack for rounding
    std::vector<float> up(ctx->n_pairs), dn(ctx->n_pairs);
    for (int c = 0; c < ctx->n_pairs; ++c) {
        const double w = (c < (int)ctx->pair_deep.size() && ctx->pair_deep[c]) ? 0.0 : ctx->tab_lp[c] - ctx->tab_la[c];
        up[c] = (float)(std::max(0.0, w) / 0.6931471805599453 * 1.001 + 1e-3);
        dn[c] = (float)(std::max(0.0, -w) / 0.6931471805599453 * 1.001 + 1e-3);
    }
    DevBuf d_up, d_dn;
    ARB_TRY(d_up.reserve(up.size() * 4));
    ARB_TRY(d_dn.reserve(dn.size() * 4));
    ARB_CUDA(cudaMemcpyAsync(d_up.p, up.data(), up.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    ARB_CUDA(cudaMemcpyAsync(d_dn.p, dn.data(), dn.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    ARB_TRY(build_stream(ctx, ctx->coo_cell.as<int32_t>(), ctx->coo_snv.as<int32_t>(), ctx->n_cells, L.prz,
                         L.np_z, kp, L.w16, d_up.as<float>(), d_dn.as<float>(),
                         StreamOut{&L.zdesc, &L.zrows, &L.zstream, &L.uz, &L.fold_z, &L.quads_z}));
    ARB_TRY(build_stream(ctx, ctx->coo_snv.as<int32_t>(), ctx->coo_cell.as<int32_t>(), ctx->n_snvs, L.pry,
                         L.np_y, kp, L.w16, d_up.as<float>(), d_dn.as<float>(),
                         StreamOut{&L.ydesc, &L.yrows, &L.ystream, &L.uy, &L.fold_y, &L.quads_y}));
    d_up.release();
    d_dn.release();
    if (ctx->opt_no_fold) L.fold_z = L.fold_y = 0;
    L.kp = kp;
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// table refresh: derived table, sum of log_absent, renormalisation interval
// ---------------------------------------------------------------------------------------
int arb_refresh_tables(arb_ctx *ctx)
{
    const int n = ctx->n_pairs;
    DevBuf lp, la;
    ARB_TRY(lp.reserve((size_t)n * 8));
    ARB_TRY(la.reserve((size_t)n * 8));
    ARB_TRY(ctx->tab_dw.reserve((size_t)n * 16));
    ARB_CUDA(cudaMemcpyAsync(lp.p, ctx->tab_lp.data(), (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
    ARB_CUDA(cudaMemcpyAsync(la.p, ctx->tab_la.data(), (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
    k_table_derive<<<grid_for(n, 128), 128, 0, ctx->stream>>>(lp.as<double>(), la.as<double>(), n,
                                                             ctx->tab_dw.as<double2>());
    ctx->h_dw.resize((size_t)n * 2);             // the device's own doubles, for the kernels' constant-bank copy
    ARB_CUDA(cudaMemcpyAsync(ctx->h_dw.data(), ctx->tab_dw.p, (size_t)n * 16, cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    lp.release();
    la.release();
    // deep pairs leave the product form: (d, w) = (0, 0) makes their factor exactly 1 and their likelihood term 0
    // in the sweeps; the finishing kernels add them in the log domain (side lists)
    ctx->pair_deep.assign((size_t)n, 0);
    bool any_deep = false;
    for (int c = 0; c < n; ++c) {
        const double w = ctx->tab_lp[c] - ctx->tab_la[c];
        if (ctx->pair_count[c] > 0 && (w < ARB_W_LO || w > ARB_W_HI)) {
            ctx->pair_deep[c] = 1;
            ctx->h_dw[2 * c] = ctx->h_dw[2 * c + 1] = 0.0;
            any_deep = true;
        }
    }
    if (any_deep)
        ARB_CUDA(cudaMemcpy(ctx->tab_dw.p, ctx->h_dw.data(), (size_t)n * 16, cudaMemcpyHostToDevice));
    ctx->side_dirty = true;
    // bookkeeping on the host: sum_k log_absent_k = sum_c count_c * la_c in table order
    double s = 0.0, max_abs_w = 0.0;
    for (int c = 0; c < n; ++c) {
        s += (double)ctx->pair_count[c] * ctx->tab_la[c];
        const double w = ctx->tab_lp[c] - ctx->tab_la[c];
        if (ctx->pair_count[c] > 0 && !ctx->pair_deep[c] && fabs(w) > max_abs_w) max_abs_w = fabs(w);
    }
    ctx->sum_la = s;
    // factors 1 + d*x lie in [min(1,e^w), max(1,e^w)]: where the layout has no range proof the sweep
    // renormalises its running products every renorm_every factors (|log2| < 900 in between);
    // renorm_every / 4 quads, 0 = after every factor
    const double max_log2 = max_abs_w / 0.6931471805599453;
    int every = max_log2 > 0 ? (int)floor(900.0 / max_log2) : 256;
    ctx->renorm_every = std::max(1, std::min(every, 256));
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// side lists of the deep pairs (log-domain path of the finishing kernels)
// ---------------------------------------------------------------------------------------
__global__ void k_side_flags(const uint16_t *__restrict__ code, const uint8_t *__restrict__ deep, int64_t nnz, int32_t *flag)
{
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < nnz) flag[k] = deep[code[k]] ? 1 : 0;
}

__global__ void k_side_compact(const int32_t *__restrict__ cell, const int32_t *__restrict__ snv, const uint16_t *__restrict__ code,
                               const int32_t *__restrict__ flag, const int32_t *__restrict__ pos, const double *__restrict__ w_tab,
                               int64_t nnz, int32_t *o_cell, int32_t *o_snv, double *o_w)
{
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= nnz || !flag[k]) return;
    const int32_t o = pos[k];
    o_cell[o] = cell[k];
    o_snv[o] = snv[k];
    o_w[o] = w_tab[code[k]];
}

__global__ void k_side_gather(const int32_t *__restrict__ order, const int32_t *__restrict__ cell, const double *__restrict__ w,
                              int64_t n, int32_t *o_cell, double *o_w)
{
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    o_cell[k] = cell[order[k]];
    o_w[k] = w[order[k]];
}

// CSR pointer of a sorted row-index list
__global__ void k_sorted_ptr(const int32_t *__restrict__ row, int64_t n_items, int32_t n_rows, int32_t *__restrict__ ptr)
{
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k > n_items) return;
    const int prev = k == 0 ? -1 : row[k - 1], cur = k == n_items ? n_rows : row[k];
    for (int r = prev + 1; r <= cur; ++r) ptr[r] = (int32_t)k;
}

int arb_build_side_lists(arb_ctx *ctx)
{
    if (!ctx->side_dirty) return ARB_OK;
    ARB_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int n_pairs = ctx->n_pairs;
    const int64_t nnz = ctx->nnz;
    const int B = 256;
    // per-code tables the general first pass reads (true w, deep flag)
    std::vector<double> w_true((size_t)n_pairs);
    bool any = false;
    for (int c = 0; c < n_pairs; ++c) {
        w_true[c] = ctx->tab_lp[c] - ctx->tab_la[c];
        any = any || (c < (int)ctx->pair_deep.size() && ctx->pair_deep[c]);
    }
    ARB_TRY(ctx->tab_w_dev.reserve((size_t)n_pairs * 8));
    ARB_TRY(ctx->tab_deep_dev.reserve((size_t)n_pairs + 16));
    ARB_CUDA(cudaMemcpy(ctx->tab_w_dev.p, w_true.data(), (size_t)n_pairs * 8, cudaMemcpyHostToDevice));
    ARB_CUDA(cudaMemcpy(ctx->tab_deep_dev.p, ctx->pair_deep.data(), (size_t)n_pairs, cudaMemcpyHostToDevice));
    ctx->n_side = 0;
    if (!any) {
        for (DevBuf *b : {&ctx->side_z_ptr, &ctx->side_z_idx, &ctx->side_z_w, &ctx->side_y_ptr, &ctx->side_y_idx, &ctx->side_y_w})
            b->release();
        ctx->side_dirty = false;
        return ARB_OK;
    }
    DevBuf flag, pos, s_cell, s_snv, key_out, order_in, order_out;
    ARB_TRY(flag.reserve((size_t)nnz * 4));
    ARB_TRY(pos.reserve((size_t)(nnz + 1) * 4));
    k_side_flags<<<grid_for(nnz, B), B, 0, st>>>(ctx->coo_code.as<uint16_t>(), ctx->tab_deep_dev.as<uint8_t>(), nnz,
                                                 flag.as<int32_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceScan::ExclusiveSum(tmp, bytes, flag.as<int32_t>(), pos.as<int32_t>(), (int)nnz, st);
    }));
    int32_t last_pos = 0, last_flag = 0;
    ARB_CUDA(cudaMemcpyAsync(&last_pos, pos.as<int32_t>() + (nnz - 1), 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(&last_flag, flag.as<int32_t>() + (nnz - 1), 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    const int64_t ns = (int64_t)last_pos + last_flag;
    ctx->n_side = ns;
    ARB_TRY(s_cell.reserve((size_t)ns * 4 + 16));
    ARB_TRY(ctx->side_z_idx.reserve((size_t)ns * 4 + 16));
    ARB_TRY(ctx->side_z_w.reserve((size_t)ns * 8 + 16));
    ARB_TRY(ctx->side_z_ptr.reserve(((size_t)ctx->n_cells + 1) * 4));
    ARB_TRY(ctx->side_y_idx.reserve((size_t)ns * 4 + 16));
    ARB_TRY(ctx->side_y_w.reserve((size_t)ns * 8 + 16));
    ARB_TRY(ctx->side_y_ptr.reserve(((size_t)ctx->n_snvs + 1) * 4));
    // canonical order is (cell, snv): the compacted list is the cell-major side list as it stands
    k_side_compact<<<grid_for(nnz, B), B, 0, st>>>(ctx->coo_cell.as<int32_t>(), ctx->coo_snv.as<int32_t>(),
                                                   ctx->coo_code.as<uint16_t>(), flag.as<int32_t>(), pos.as<int32_t>(),
                                                   ctx->tab_w_dev.as<double>(), nnz, s_cell.as<int32_t>(),
                                                   ctx->side_z_idx.as<int32_t>(), ctx->side_z_w.as<double>());
    k_sorted_ptr<<<grid_for(ns + 1, B), B, 0, st>>>(s_cell.as<int32_t>(), ns, ctx->n_cells, ctx->side_z_ptr.as<int32_t>());
    // SNV-major copy: stable sort of the positions by SNV (cells stay ascending inside an SNV)
    ARB_TRY(s_snv.reserve((size_t)ns * 4 + 16));
    ARB_TRY(key_out.reserve((size_t)ns * 4 + 16));
    ARB_TRY(order_in.reserve((size_t)ns * 4 + 16));
    ARB_TRY(order_out.reserve((size_t)ns * 4 + 16));
    k_iota<<<grid_for(ns, B), B, 0, st>>>(order_in.as<int32_t>(), ns);
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, ctx->side_z_idx.as<int32_t>(), key_out.as<int32_t>(),
                                               order_in.as<int32_t>(), order_out.as<int32_t>(), (int)ns, 0, 32, st);
    }));
    k_side_gather<<<grid_for(ns, B), B, 0, st>>>(order_out.as<int32_t>(), s_cell.as<int32_t>(), ctx->side_z_w.as<double>(), ns,
                                                 ctx->side_y_idx.as<int32_t>(), ctx->side_y_w.as<double>());
    k_sorted_ptr<<<grid_for(ns + 1, B), B, 0, st>>>(key_out.as<int32_t>(), ns, ctx->n_snvs, ctx->side_y_ptr.as<int32_t>());
    ARB_CUDA(cudaStreamSynchronize(st));
    ARB_CUDA(cudaGetLastError());
    for (DevBuf *b : {&flag, &pos, &s_cell, &s_snv, &key_out, &order_in, &order_out}) b->release();
    ctx->side_dirty = false;
    return ARB_OK;
}

static int check_table_range(arb_ctx *ctx)
{
    for (int c = 0; c < ctx->n_pairs; ++c) {
        if (ctx->pair_count[c] == 0) continue;
        const double w = ctx->tab_lp[c] - ctx->tab_la[c];
        ARB_REQUIRE(w == w && fabs(w) < 1e300, ARB_ERR_INVALID,
                    "count pair %d (alt %d, total %d) has no finite log-likelihoods (alt > total?)", c,
                    c < (int)ctx->pair_alt.size() ? ctx->pair_alt[c] : -1,
                    c < (int)ctx->pair_total.size() ? ctx->pair_total[c] : -1);
    }
    return arb_build_side_lists(ctx);        // deep pairs (w outside [ARB_W_LO, ARB_W_HI]) go through the log-domain path
}
int arb_check_table_range(arb_ctx *ctx) { return check_table_range(ctx); }

// order the dictionary by descending frequency (ties: ascending raw rank); returns map raw->final
static std::vector<int32_t> frequency_order(const std::vector<int64_t> &count)
{
    const int n = (int)count.size();
    std::vector<int32_t> order(n), map(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return count[a] > count[b]; });
    for (int f = 0; f < n; ++f) map[order[f]] = f;
    return map;
}

static void reset_matrix(arb_ctx *ctx)
{
    ctx->ingested = false;
    ctx->row_ptr.release();
    for (auto &l : ctx->layout) l.release();
    ctx->ws_kp = 0;
    ctx->fit_trees = 0;
    ctx->batch_t0 = -1;
    ctx->n_clusters = 0;
}

// canonical sort by (cell, snv) and gather; key/code are per compacted row
static int finish_canonical(arb_ctx *ctx, DevBuf &key, const uint16_t *code)
{
    const int64_t nnz = ctx->nnz;
    cudaStream_t st = ctx->stream;
    DevBuf k_out, v_in, v_out;
    ARB_TRY(k_out.reserve(nnz * 8));
    ARB_TRY(v_in.reserve(nnz * 4));
    ARB_TRY(v_out.reserve(nnz * 4));
    ARB_TRY(ctx->coo_cell.reserve(nnz * 4));
    ARB_TRY(ctx->coo_snv.reserve(nnz * 4));
    ARB_TRY(ctx->coo_code.reserve(nnz * 2));
    k_iota<<<grid_for(nnz, 256), 256, 0, st>>>(v_in.as<int32_t>(), nnz);
    const int end_bit = bits_for((uint64_t)ctx->n_cells * (uint64_t)ctx->n_snvs);
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, key.as<uint64_t>(), k_out.as<uint64_t>(),
                                               v_in.as<int32_t>(), v_out.as<int32_t>(), nnz, 0, end_bit, st);
    }));
    k_gather_canonical<<<grid_for(nnz, 256), 256, 0, st>>>(k_out.as<uint64_t>(), v_out.as<int32_t>(), code, nnz,
                                                           ctx->n_snvs, ctx->coo_cell.as<int32_t>(),
                                                           ctx->coo_snv.as<int32_t>(), ctx->coo_code.as<uint16_t>());
    ARB_CUDA(cudaStreamSynchronize(st));
    k_out.release(); v_in.release(); v_out.release();
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// arb_ingest_coo
// ---------------------------------------------------------------------------------------
// the table as device-resident first-appearance codes (d_cell, d_snv, d_alt, d_tot: int32 [n_rows], consumed)
static int ingest_device_codes(arb_ctx *ctx, int64_t n_rows, DevBuf &d_cell, DevBuf &d_snv, DevBuf &d_alt, DevBuf &d_tot,
                               int32_t n_cell_codes, int32_t n_snv_codes, double alpha);

ARB_API int arb_ingest_coo(arb_ctx *ctx, int64_t n_rows, const int32_t *cell_code, const int32_t *snv_code,
                           const int32_t *alt, const int32_t *total, int32_t n_cell_codes,
                           int32_t n_snv_codes, double alpha)
{
    ARB_REQUIRE(ctx && cell_code && snv_code && alt && total, ARB_ERR_INVALID, "arb_ingest_coo: null argument");
    ARB_REQUIRE(n_rows > 0 && n_rows < (int64_t)INT_MAX, ARB_ERR_INVALID, "arb_ingest_coo: n_rows=%lld",
                (long long)n_rows);
    ARB_REQUIRE(n_cell_codes > 0 && n_snv_codes > 0, ARB_ERR_INVALID, "arb_ingest_coo: empty code space");
    ARB_CUDA(cudaSetDevice(ctx->device));
    ctx->n_cell_uniques = ctx->n_snv_uniques = 0;
    cudaStream_t st = ctx->stream;
    DevBuf d_cell, d_snv, d_alt, d_tot;
    for (DevBuf *b : {&d_cell, &d_snv, &d_alt, &d_tot}) ARB_TRY(b->reserve(n_rows * 4));
    ARB_CUDA(cudaMemcpyAsync(d_cell.p, cell_code, n_rows * 4, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_snv.p, snv_code, n_rows * 4, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_alt.p, alt, n_rows * 4, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_tot.p, total, n_rows * 4, cudaMemcpyHostToDevice, st));
    return ingest_device_codes(ctx, n_rows, d_cell, d_snv, d_alt, d_tot, n_cell_codes, n_snv_codes, alpha);
}

static int ingest_device_codes(arb_ctx *ctx, int64_t n_rows, DevBuf &d_cell, DevBuf &d_snv, DevBuf &d_alt, DevBuf &d_tot,
                               int32_t n_cell_codes, int32_t n_snv_codes, double alpha)
{
    reset_matrix(ctx);
    cudaStream_t st = ctx->stream;
    const int B = 256;
    DevBuf alt_sum, keep, snv_new, row_keep, row_pos, first, iota, sfirst, scell, cell_new, counters;
    for (DevBuf *b : {&row_keep, &row_pos}) ARB_TRY(b->reserve(n_rows * 4));
    ARB_TRY(alt_sum.reserve((size_t)n_snv_codes * 8));
    ARB_TRY(keep.reserve((size_t)n_snv_codes * 4));
    ARB_TRY(snv_new.reserve((size_t)n_snv_codes * 4));
    for (DevBuf *b : {&first, &iota, &sfirst, &scell, &cell_new}) ARB_TRY(b->reserve((size_t)n_cell_codes * 4));
    ARB_TRY(counters.reserve(64));
    ARB_CUDA(cudaMemsetAsync(alt_sum.p, 0, (size_t)n_snv_codes * 8, st));
    ARB_CUDA(cudaMemsetAsync(counters.p, 0, 64, st));

    // zero-alt filter (A.py:1034-1041)
    k_alt_sum<<<grid_for(n_rows, B), B, 0, st>>>(d_snv.as<int32_t>(), d_alt.as<int32_t>(), n_rows,
                                                 alt_sum.as<unsigned long long>());
    k_keep_flags<<<grid_for(n_snv_codes, B), B, 0, st>>>(alt_sum.as<unsigned long long>(), n_snv_codes,
                                                         keep.as<int32_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceScan::ExclusiveSum(tmp, bytes, keep.as<int32_t>(), snv_new.as<int32_t>(), n_snv_codes, st);
    }));
    k_row_keep<<<grid_for(n_rows, B), B, 0, st>>>(d_snv.as<int32_t>(), keep.as<int32_t>(), n_rows,
                                                  row_keep.as<int32_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceScan::ExclusiveSum(tmp, bytes, row_keep.as<int32_t>(), row_pos.as<int32_t>(), n_rows, st);
    }));
    int32_t h_tail[4];
    ARB_CUDA(cudaMemcpyAsync(&h_tail[0], keep.as<int32_t>() + n_snv_codes - 1, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(&h_tail[1], snv_new.as<int32_t>() + n_snv_codes - 1, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(&h_tail[2], row_keep.as<int32_t>() + n_rows - 1, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(&h_tail[3], row_pos.as<int32_t>() + n_rows - 1, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    const int32_t n_snvs = h_tail[0] + h_tail[1];
    const int64_t nnz = (int64_t)h_tail[2] + h_tail[3];
    ARB_REQUIRE(nnz > 0 && n_snvs > 0, ARB_ERR_INVALID, "arb_ingest_coo: no row survives the zero-alt filter");

    // first-appearance cell order among the surviving rows (A.py:851)
    k_fill_i32<<<grid_for(n_cell_codes, B), B, 0, st>>>(first.as<int32_t>(), n_cell_codes, INT_MAX);
    k_first_row<<<grid_for(n_rows, B), B, 0, st>>>(d_cell.as<int32_t>(), row_keep.as<int32_t>(), n_rows,
                                                   first.as<int32_t>());
    k_iota<<<grid_for(n_cell_codes, B), B, 0, st>>>(iota.as<int32_t>(), n_cell_codes);
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, first.as<int32_t>(), sfirst.as<int32_t>(),
                                               iota.as<int32_t>(), scell.as<int32_t>(), n_cell_codes, 0, 32, st);
    }));
    ARB_TRY(ctx->cell_order.reserve((size_t)n_cell_codes * 4));
    ARB_TRY(ctx->snv_order.reserve((size_t)n_snvs * 4));
    k_cell_rank<<<grid_for(n_cell_codes, B), B, 0, st>>>(sfirst.as<int32_t>(), scell.as<int32_t>(), n_cell_codes,
                                                         cell_new.as<int32_t>(), ctx->cell_order.as<int32_t>(),
                                                         counters.as<int32_t>());
    k_snv_order<<<grid_for(n_snv_codes, B), B, 0, st>>>(keep.as<int32_t>(), snv_new.as<int32_t>(), n_snv_codes,
                                                        ctx->snv_order.as<int32_t>());
    int32_t n_cells = 0;
    ARB_CUDA(cudaMemcpyAsync(&n_cells, counters.p, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    ARB_REQUIRE(n_cells > 0, ARB_ERR_INVALID, "arb_ingest_coo: no cell survives");
    ctx->n_cells = n_cells;
    ctx->n_snvs = n_snvs;
    ctx->nnz = nnz;
    ctx->alpha = alpha;

    // compaction + dictionary of distinct (alt, total) pairs
    DevBuf key, pairkey, pk_sorted, uniq, ucount, nruns, code, code_of_rank;
    ARB_TRY(key.reserve(nnz * 8));
    ARB_TRY(pairkey.reserve(nnz * 8));
    ARB_TRY(pk_sorted.reserve(nnz * 8));
    ARB_TRY(code.reserve(nnz * 2));
    k_compact<<<grid_for(n_rows, B), B, 0, st>>>(d_cell.as<int32_t>(), d_snv.as<int32_t>(), d_alt.as<int32_t>(),
                                                 d_tot.as<int32_t>(), row_keep.as<int32_t>(), row_pos.as<int32_t>(),
                                                 cell_new.as<int32_t>(), snv_new.as<int32_t>(), n_rows, n_snvs,
                                                 key.as<uint64_t>(), pairkey.as<uint64_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortKeys(tmp, bytes, pairkey.as<uint64_t>(), pk_sorted.as<uint64_t>(), nnz, 0,
                                              64, st);
    }));
    // run-length encode; the number of runs can be as large as nnz in principle
    const int64_t max_runs = std::min<int64_t>(nnz, 1 << 20);
    ARB_TRY(uniq.reserve(nnz * 8));
    ARB_TRY(ucount.reserve(nnz * 4));
    ARB_TRY(nruns.reserve(16));
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRunLengthEncode::Encode(tmp, bytes, pk_sorted.as<uint64_t>(), uniq.as<uint64_t>(),
                                                  ucount.as<int32_t>(), nruns.as<int32_t>(), nnz, st);
    }));
    int32_t n_pairs = 0;
    ARB_CUDA(cudaMemcpyAsync(&n_pairs, nruns.p, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    (void)max_runs;
    ARB_REQUIRE(n_pairs <= 65535, ARB_ERR_UNSUPPORTED,
                "arb_ingest_coo: %d distinct (alt,total) pairs; this release supports 65535", n_pairs);
    std::vector<uint64_t> h_uniq(n_pairs);
    std::vector<int32_t> h_cnt(n_pairs);
    ARB_CUDA(cudaMemcpy(h_uniq.data(), uniq.p, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost));
    ARB_CUDA(cudaMemcpy(h_cnt.data(), ucount.p, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost));
    std::vector<int64_t> cnt64(h_cnt.begin(), h_cnt.end());
    std::vector<int32_t> map = frequency_order(cnt64);
    ctx->n_pairs = n_pairs;
    ctx->pair_alt.assign(n_pairs, 0);
    ctx->pair_total.assign(n_pairs, 0);
    ctx->pair_count.assign(n_pairs, 0);
    for (int r = 0; r < n_pairs; ++r) {
        ctx->pair_alt[map[r]] = (int32_t)(uint32_t)(h_uniq[r] & 0xffffffffu);
        ctx->pair_total[map[r]] = (int32_t)(uint32_t)(h_uniq[r] >> 32);
        ctx->pair_count[map[r]] = cnt64[r];
    }
    ARB_TRY(code_of_rank.reserve((size_t)n_pairs * 4));
    ARB_CUDA(cudaMemcpyAsync(code_of_rank.p, map.data(), (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
    k_encode_pairs<<<grid_for(nnz, B), B, 0, st>>>(pairkey.as<uint64_t>(), nnz, uniq.as<uint64_t>(), n_pairs,
                                                   code_of_rank.as<int32_t>(), code.as<uint16_t>());
    ARB_TRY(finish_canonical(ctx, key, code.as<uint16_t>()));

    // per-pair log-likelihood table (A.py:838-846)
    DevBuf t_alt, t_tot, t_lp, t_la;
    ARB_TRY(t_alt.reserve((size_t)n_pairs * 4));
    ARB_TRY(t_tot.reserve((size_t)n_pairs * 4));
    ARB_TRY(t_lp.reserve((size_t)n_pairs * 8));
    ARB_TRY(t_la.reserve((size_t)n_pairs * 8));
    ARB_CUDA(cudaMemcpyAsync(t_alt.p, ctx->pair_alt.data(), (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(t_tot.p, ctx->pair_total.data(), (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
    k_logpmf_table<<<grid_for(n_pairs, 128), 128, 0, st>>>(t_alt.as<int32_t>(), t_tot.as<int32_t>(), n_pairs,
                                                           0.5 - alpha + 0.5 * alpha / 3, alpha / 3,
                                                           t_lp.as<double>(), t_la.as<double>());
    ctx->tab_lp.resize(n_pairs);
    ctx->tab_la.resize(n_pairs);
    ARB_CUDA(cudaMemcpyAsync(ctx->tab_lp.data(), t_lp.p, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(ctx->tab_la.data(), t_la.p, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    for (DevBuf *b : {&d_cell, &d_snv, &d_alt, &d_tot, &alt_sum, &keep, &snv_new, &row_keep, &row_pos, &first,
                      &iota, &sfirst, &scell, &cell_new, &counters, &key, &pairkey, &pk_sorted, &uniq, &ucount,
                      &nruns, &code, &code_of_rank, &t_alt, &t_tot, &t_lp, &t_la})
        b->release();
    ARB_TRY(arb_refresh_tables(ctx));
    ctx->ingested = true;
    ctx->from_coo = true;
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// arb_ingest_ids: the long table with raw 64-bit integer ids; the first-appearance factorisation
// of the two id columns (pandas unique / factorize order, A.py:851-855) is done on the device
// ---------------------------------------------------------------------------------------
__global__ void k_narrow_i64(const int64_t *src, int64_t n, int32_t *dst, int32_t *bad)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t v = src[i];
    if (v < 0 || v > INT_MAX) atomicExch(bad, 1);
    dst[i] = (int32_t)v;
}

__global__ void k_head_flags(const uint64_t *sorted, int64_t n, int32_t *flag)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) flag[i] = (i == 0 || sorted[i] != sorted[i - 1]) ? 1 : 0;
}

// group g (ids in ascending order) first appears at row first[g]
__global__ void k_group_first(const int32_t *flag, const int32_t *gidx, const int32_t *row, int64_t n, int32_t *first)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n && flag[i]) first[gidx[i] - 1] = row[i]; // the sort is stable: the head holds the smallest row
}

__global__ void k_group_rank(const int32_t *group_by_first, int32_t n_groups, int32_t *code_of_group)
{
    int32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n_groups) code_of_group[group_by_first[r]] = r;
}

__global__ void k_scatter_codes(const uint64_t *sorted, const int32_t *flag, const int32_t *gidx, const int32_t *row,
                                const int32_t *code_of_group, int64_t n, int32_t *code, int64_t *uniques)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t c = code_of_group[gidx[i] - 1];
    code[row[i]] = c;
    if (flag[i]) uniques[c] = (int64_t)sorted[i];
}

// codes[row] = index of ids[row] in first-appearance order; uniques[code] = id
static int device_factorize(arb_ctx *ctx, const DevBuf &ids, int64_t n, DevBuf &codes, DevBuf &uniques, int32_t *n_codes)
{
    cudaStream_t st = ctx->stream;
    const int B = 256;
    DevBuf sorted, row_in, row_out, flag, gidx, first, gi, sfirst, gsorted, cog;
    ARB_TRY(sorted.reserve(n * 8));
    for (DevBuf *b : {&row_in, &row_out, &flag, &gidx}) ARB_TRY(b->reserve(n * 4));
    k_iota<<<grid_for(n, B), B, 0, st>>>(row_in.as<int32_t>(), n);
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, ids.as<uint64_t>(), sorted.as<uint64_t>(), row_in.as<int32_t>(),
                                               row_out.as<int32_t>(), n, 0, 64, st);
    }));
    k_head_flags<<<grid_for(n, B), B, 0, st>>>(sorted.as<uint64_t>(), n, flag.as<int32_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceScan::InclusiveSum(tmp, bytes, flag.as<int32_t>(), gidx.as<int32_t>(), n, st);
    }));
    int32_t ng = 0;                      // gidx[i] = 1 + group of element i (groups in ascending id order)
    ARB_CUDA(cudaMemcpyAsync(&ng, gidx.as<int32_t>() + n - 1, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    for (DevBuf *b : {&first, &gi, &sfirst, &gsorted, &cog}) ARB_TRY(b->reserve((size_t)ng * 4));
    ARB_TRY(uniques.reserve((size_t)ng * 8));
    ARB_TRY(codes.reserve(n * 4));
    k_group_first<<<grid_for(n, B), B, 0, st>>>(flag.as<int32_t>(), gidx.as<int32_t>(), row_out.as<int32_t>(), n,
                                                first.as<int32_t>());
    k_iota<<<grid_for(ng, B), B, 0, st>>>(gi.as<int32_t>(), ng);
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, first.as<int32_t>(), sfirst.as<int32_t>(), gi.as<int32_t>(),
                                               gsorted.as<int32_t>(), ng, 0, 32, st);
    }));
    k_group_rank<<<grid_for(ng, B), B, 0, st>>>(gsorted.as<int32_t>(), ng, cog.as<int32_t>());
    k_scatter_codes<<<grid_for(n, B), B, 0, st>>>(sorted.as<uint64_t>(), flag.as<int32_t>(), gidx.as<int32_t>(),
                                                  row_out.as<int32_t>(), cog.as<int32_t>(), n, codes.as<int32_t>(),
                                                  uniques.as<int64_t>());
    ARB_CUDA(cudaStreamSynchronize(st));
    for (DevBuf *b : {&sorted, &row_in, &row_out, &flag, &gidx, &first, &gi, &sfirst, &gsorted, &cog}) b->release();
    *n_codes = ng;
    return ARB_OK;
}

ARB_API int arb_ingest_ids(arb_ctx *ctx, int64_t n_rows, const int64_t *cell_id, const int64_t *snv_id,
                           const int64_t *alt, const int64_t *total, double alpha, int32_t *n_cell_ids,
                           int32_t *n_snv_ids)
{
    ARB_REQUIRE(ctx && cell_id && snv_id && alt && total, ARB_ERR_INVALID, "arb_ingest_ids: null argument");
    ARB_REQUIRE(n_rows > 0 && n_rows < (int64_t)INT_MAX, ARB_ERR_INVALID, "arb_ingest_ids: n_rows=%lld",
                (long long)n_rows);
    ARB_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int B = 256;
    DevBuf raw, d_cell, d_snv, d_alt, d_tot, bad;
    ARB_TRY(raw.reserve(n_rows * 8));
    ARB_TRY(bad.reserve(4));
    ARB_CUDA(cudaMemsetAsync(bad.p, 0, 4, st));
    ARB_TRY(d_alt.reserve(n_rows * 4));
    ARB_TRY(d_tot.reserve(n_rows * 4));
    ARB_CUDA(cudaMemcpyAsync(raw.p, alt, n_rows * 8, cudaMemcpyHostToDevice, st));
    k_narrow_i64<<<grid_for(n_rows, B), B, 0, st>>>(raw.as<int64_t>(), n_rows, d_alt.as<int32_t>(), bad.as<int32_t>());
    ARB_CUDA(cudaMemcpyAsync(raw.p, total, n_rows * 8, cudaMemcpyHostToDevice, st));
    k_narrow_i64<<<grid_for(n_rows, B), B, 0, st>>>(raw.as<int64_t>(), n_rows, d_tot.as<int32_t>(), bad.as<int32_t>());
    int32_t h_bad = 0;
    ARB_CUDA(cudaMemcpyAsync(&h_bad, bad.p, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(raw.p, cell_id, n_rows * 8, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    ARB_REQUIRE(!h_bad, ARB_ERR_INVALID, "arb_ingest_ids: alt / total outside 0..2^31-1");
    int32_t nc = 0, ns = 0;
    ARB_TRY(device_factorize(ctx, raw, n_rows, d_cell, ctx->cell_uniques, &nc));
    ARB_CUDA(cudaMemcpyAsync(raw.p, snv_id, n_rows * 8, cudaMemcpyHostToDevice, st));
    ARB_TRY(device_factorize(ctx, raw, n_rows, d_snv, ctx->snv_uniques, &ns));
    raw.release();
    bad.release();
    ctx->n_cell_uniques = nc;
    ctx->n_snv_uniques = ns;
    if (n_cell_ids) *n_cell_ids = nc;
    if (n_snv_ids) *n_snv_ids = ns;
    return ingest_device_codes(ctx, n_rows, d_cell, d_snv, d_alt, d_tot, nc, ns, alpha);
}

ARB_API int arb_get_id_uniques(arb_ctx *ctx, int64_t *cell_uniques, int64_t *snv_uniques)
{
    ARB_REQUIRE(ctx && ctx->n_cell_uniques > 0, ARB_ERR_INVALID, "arb_get_id_uniques: no arb_ingest_ids table");
    ARB_CUDA(cudaSetDevice(ctx->device));
    if (cell_uniques)
        ARB_CUDA(cudaMemcpy(cell_uniques, ctx->cell_uniques.p, (size_t)ctx->n_cell_uniques * 8, cudaMemcpyDeviceToHost));
    if (snv_uniques)
        ARB_CUDA(cudaMemcpy(snv_uniques, ctx->snv_uniques.p, (size_t)ctx->n_snv_uniques * 8, cudaMemcpyDeviceToHost));
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// read-count CSV parsed on the device (replaces pd.read_csv of A.py:1135 for integer tables): the host
// hands over the raw bytes of the file in chunks that end on a line boundary; the four columns are
// parsed straight into device memory and never exist as host arrays.
// ---------------------------------------------------------------------------------------
#define ARB_CSV_SPAN 128      // bytes per thread

__global__ void k_csv_count(const char *__restrict__ buf, int64_t n, int32_t *__restrict__ counts, int64_t n_spans)
{
    int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (s >= n_spans) return;
    int64_t lo = s * ARB_CSV_SPAN, hi = min(n, lo + ARB_CSV_SPAN);
    int c = 0;
    if (hi - lo == ARB_CSV_SPAN) {
        const uint4 *v = reinterpret_cast<const uint4 *>(buf + lo);
#pragma unroll
        for (int i = 0; i < ARB_CSV_SPAN / 16; ++i) {
            uint4 q = v[i];
            uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                c += __popc(__vcmpeq4(w[j], 0x0a0a0a0au)) >> 3;       // 0xff in every byte that is '\n'
            }
        }
    } else {
        for (int64_t i = lo; i < hi; ++i) c += buf[i] == '\n';
    }
    counts[s] = c;
}

// One thread per span: every line that ENDS in the span is parsed by this thread (its start is found by walking
// back to the previous newline).  Selected fields must be plain integers; a quote anywhere, a wrong field count or
// an empty line records the byte offset in `bad` (the host then falls back to its own reader).
__global__ void k_csv_parse(const char *__restrict__ buf, int64_t n, const int32_t *__restrict__ offs, int64_t n_spans,
                            int64_t row0, int c_cell, int c_snv, int c_alt, int c_tot, int n_fields,
                            int64_t *__restrict__ o_cell, int64_t *__restrict__ o_snv, int64_t *__restrict__ o_alt,
                            int64_t *__restrict__ o_tot, unsigned long long *__restrict__ bad)
{
    int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (s >= n_spans) return;
    int64_t lo = s * ARB_CSV_SPAN, hi = min(n, lo + ARB_CSV_SPAN);
    int64_t row = row0 + offs[s];
    for (int64_t p = lo; p < hi; ++p) {
        if (buf[p] != '\n') continue;
        int64_t q = p;
        while (q > 0 && buf[q - 1] != '\n') --q;
        int64_t e = p;
        if (e > q && buf[e - 1] == '\r') --e;
        bool ok = e > q;
        int f = 0, digits = 0;
        int64_t v = 0, out[4] = {0, 0, 0, 0};
        bool neg = false;
        for (int64_t i = q; i <= e && ok; ++i) {
            char c = i < e ? buf[i] : ',';
            bool sel = f == c_cell || f == c_snv || f == c_alt || f == c_tot;
            if (c == ',') {
                if (sel) {
                    ok = digits > 0;
                    int64_t val = neg ? -v : v;
                    if (f == c_cell) out[0] = val;
                    if (f == c_snv) out[1] = val;
                    if (f == c_alt) out[2] = val;
                    if (f == c_tot) out[3] = val;
                }
                ++f;
                v = 0;
                digits = 0;
                neg = false;
            } else if (c == '"') {
                ok = false;
            } else if (sel) {
                if (c >= '0' && c <= '9') {
                    v = v * 10 + (c - '0');
                    ok = ++digits <= 18;
                } else if (c == '-' && digits == 0 && !neg) {
                    neg = true;
                } else {
                    ok = false;
                }
            }
        }
        if (!ok || f != n_fields) {
            atomicMin(bad, (unsigned long long)q);
        } else {
            o_cell[row] = out[0];
            o_snv[row] = out[1];
            o_alt[row] = out[2];
            o_tot[row] = out[3];
        }
        ++row;
    }
}

static int csv_grow(arb_ctx *ctx, int64_t need)
{
    if (need <= ctx->csv_cap) return ARB_OK;
    int64_t cap = std::max<int64_t>(need, ctx->csv_cap * 2);
    for (DevBuf &b : ctx->csv_col) {
        DevBuf nb;
        ARB_TRY(nb.reserve((size_t)cap * 8));
        if (ctx->csv_rows)
            ARB_CUDA(cudaMemcpyAsync(nb.p, b.p, (size_t)ctx->csv_rows * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        ARB_CUDA(cudaStreamSynchronize(ctx->stream));
        b.release();
        b = nb;
    }
    ctx->csv_cap = cap;
    return ARB_OK;
}

ARB_API int arb_csv_begin(arb_ctx *ctx, int32_t col_cell, int32_t col_snv, int32_t col_alt, int32_t col_total,
                          int32_t n_fields, int64_t expected_rows)
{
    ARB_REQUIRE(ctx, ARB_ERR_INVALID, "arb_csv_begin: null context");
    const int32_t c[4] = {col_cell, col_snv, col_alt, col_total};
    for (int i = 0; i < 4; ++i) {
        ARB_REQUIRE(c[i] >= 0 && c[i] < n_fields, ARB_ERR_INVALID, "arb_csv_begin: column index %d outside 0..%d", c[i],
                    n_fields - 1);
        for (int j = 0; j < i; ++j) ARB_REQUIRE(c[i] != c[j], ARB_ERR_INVALID, "arb_csv_begin: column %d given twice", c[i]);
    }
    ARB_CUDA(cudaSetDevice(ctx->device));
    for (int i = 0; i < 4; ++i) ctx->csv_field[i] = c[i];
    ctx->csv_fields = n_fields;
    ctx->csv_rows = 0;
    ctx->csv_open = true;
    return csv_grow(ctx, std::max<int64_t>(expected_rows, 1024));
}

ARB_API int arb_csv_feed(arb_ctx *ctx, const char *bytes, int64_t n_bytes, int64_t *n_rows_total)
{
    ARB_REQUIRE(ctx && ctx->csv_open, ARB_ERR_INVALID, "arb_csv_feed: call arb_csv_begin first");
    ARB_REQUIRE(bytes && n_bytes > 0 && n_bytes <= ((int64_t)1 << 30), ARB_ERR_INVALID,
                "arb_csv_feed: chunk of %lld bytes (1..2^30 allowed)", (long long)n_bytes);
    ARB_REQUIRE(bytes[n_bytes - 1] == '\n', ARB_ERR_INVALID, "arb_csv_feed: a chunk has to end with a newline");
    ARB_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int B = 256;
    int64_t n_spans = (n_bytes + ARB_CSV_SPAN - 1) / ARB_CSV_SPAN;
    DevBuf buf, counts, offs, bad;
    ARB_TRY(buf.reserve((size_t)n_bytes + 16));
    ARB_TRY(counts.reserve((size_t)n_spans * 4));
    ARB_TRY(offs.reserve((size_t)n_spans * 4));
    ARB_TRY(bad.reserve(8));
    ARB_CUDA(cudaMemcpyAsync(buf.p, bytes, (size_t)n_bytes, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemsetAsync(bad.p, 0xff, 8, st));
    k_csv_count<<<grid_for(n_spans, B), B, 0, st>>>(buf.as<char>(), n_bytes, counts.as<int32_t>(), n_spans);
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &tb) {
        return cub::DeviceScan::ExclusiveSum(tmp, tb, counts.as<int32_t>(), offs.as<int32_t>(), (int)n_spans, st);
    }));
    int32_t last_off = 0, last_cnt = 0;
    ARB_CUDA(cudaMemcpyAsync(&last_off, offs.as<int32_t>() + (n_spans - 1), 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(&last_cnt, counts.as<int32_t>() + (n_spans - 1), 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    int64_t lines = (int64_t)last_off + last_cnt;
    ARB_REQUIRE(ctx->csv_rows + lines < (int64_t)INT_MAX, ARB_ERR_UNSUPPORTED, "arb_csv_feed: more than 2^31-1 rows");
    ARB_TRY(csv_grow(ctx, ctx->csv_rows + lines));
    k_csv_parse<<<grid_for(n_spans, B), B, 0, st>>>(buf.as<char>(), n_bytes, offs.as<int32_t>(), n_spans, ctx->csv_rows,
                                                    ctx->csv_field[0], ctx->csv_field[1], ctx->csv_field[2],
                                                    ctx->csv_field[3], ctx->csv_fields, ctx->csv_col[0].as<int64_t>(),
                                                    ctx->csv_col[1].as<int64_t>(), ctx->csv_col[2].as<int64_t>(),
                                                    ctx->csv_col[3].as<int64_t>(), bad.as<unsigned long long>());
    ARB_CUDA(cudaGetLastError());
    unsigned long long h_bad = 0;
    ARB_CUDA(cudaMemcpyAsync(&h_bad, bad.p, 8, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    for (DevBuf *b : {&buf, &counts, &offs, &bad}) b->release();
    if (h_bad != ~0ull) {
        // declined: the caller falls back to a host reader, so the columns parsed so far are dead weight
        ctx->csv_open = false;
        for (DevBuf &b : ctx->csv_col) b.release();
        ctx->csv_cap = ctx->csv_rows = 0;
        arb_set_error("arb_csv_feed: the line at byte %llu of this chunk is not %d comma-separated fields with plain "
                      "integers in the four selected columns (quoted, empty, fractional or text fields need the host reader)",
                      h_bad, ctx->csv_fields);
        return ARB_ERR_DECLINED;
    }
    ctx->csv_rows += lines;
    if (n_rows_total) *n_rows_total = ctx->csv_rows;
    return ARB_OK;
}

ARB_API int arb_csv_finish(arb_ctx *ctx, double alpha, int64_t *n_rows, int32_t *n_cell_ids, int32_t *n_snv_ids)
{
    ARB_REQUIRE(ctx && ctx->csv_open, ARB_ERR_INVALID, "arb_csv_finish: no open CSV table");
    ARB_REQUIRE(ctx->csv_rows > 0, ARB_ERR_INVALID, "arb_csv_finish: no rows were fed");
    ARB_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int B = 256;
    const int64_t n = ctx->csv_rows;
    ctx->csv_open = false;
    DevBuf d_cell, d_snv, d_alt, d_tot, bad;
    ARB_TRY(bad.reserve(4));
    ARB_CUDA(cudaMemsetAsync(bad.p, 0, 4, st));
    ARB_TRY(d_alt.reserve(n * 4));
    ARB_TRY(d_tot.reserve(n * 4));
    k_narrow_i64<<<grid_for(n, B), B, 0, st>>>(ctx->csv_col[2].as<int64_t>(), n, d_alt.as<int32_t>(), bad.as<int32_t>());
    k_narrow_i64<<<grid_for(n, B), B, 0, st>>>(ctx->csv_col[3].as<int64_t>(), n, d_tot.as<int32_t>(), bad.as<int32_t>());
    int32_t h_bad = 0;
    ARB_CUDA(cudaMemcpyAsync(&h_bad, bad.p, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    bad.release();
    ARB_REQUIRE(!h_bad, ARB_ERR_INVALID, "arb_csv_finish: alt / total outside 0..2^31-1");
    int32_t nc = 0, ns = 0;
    ARB_TRY(device_factorize(ctx, ctx->csv_col[0], n, d_cell, ctx->cell_uniques, &nc));
    ARB_TRY(device_factorize(ctx, ctx->csv_col[1], n, d_snv, ctx->snv_uniques, &ns));
    ARB_CUDA(cudaStreamSynchronize(st));
    for (DevBuf &b : ctx->csv_col) b.release();
    ctx->csv_cap = ctx->csv_rows = 0;
    ctx->n_cell_uniques = nc;
    ctx->n_snv_uniques = ns;
    if (n_rows) *n_rows = n;
    if (n_cell_ids) *n_cell_ids = nc;
    if (n_snv_ids) *n_snv_ids = ns;
    return ingest_device_codes(ctx, n, d_cell, d_snv, d_alt, d_tot, nc, ns, alpha);
}

// ---------------------------------------------------------------------------------------
// arb_ingest_csr (the arrays of the reference's run(...) seam, A.py:1093-1107)
// ---------------------------------------------------------------------------------------
ARB_API int arb_ingest_csr(arb_ctx *ctx, int32_t n_cells, int32_t n_snvs, int64_t nnz, const int32_t *cell_ptr,
                           const int32_t *snv_idx, const double *log_likes)
{
    ARB_REQUIRE(ctx && cell_ptr && snv_idx && log_likes, ARB_ERR_INVALID, "arb_ingest_csr: null argument");
    ARB_REQUIRE(n_cells > 0 && n_snvs > 0 && nnz > 0 && nnz < (int64_t)INT_MAX, ARB_ERR_INVALID,
                "arb_ingest_csr: bad sizes (%d, %d, %lld)", n_cells, n_snvs, (long long)nnz);
    ARB_REQUIRE(cell_ptr[0] == 0 && (int64_t)cell_ptr[n_cells] == nnz, ARB_ERR_INVALID,
                "arb_ingest_csr: cell_ptr does not span nnz");
    ARB_CUDA(cudaSetDevice(ctx->device));
    reset_matrix(ctx);
    cudaStream_t st = ctx->stream;
    const int B = 256;
    ctx->n_cells = n_cells;
    ctx->n_snvs = n_snvs;
    ctx->nnz = nnz;
    DevBuf d_ptr, d_col, d_ll, key, lpb, lab, v0, v1, v2, k1, k2, head, rank, raw, t_lp, t_la, cnt, map_d, code;
    ARB_TRY(d_ptr.reserve((size_t)(n_cells + 1) * 4));
    ARB_TRY(d_col.reserve(nnz * 4));
    ARB_TRY(d_ll.reserve(nnz * 16));
    for (DevBuf *b : {&key, &lpb, &lab, &k1, &k2}) ARB_TRY(b->reserve(nnz * 8));
    for (DevBuf *b : {&v0, &v1, &v2, &head, &rank, &raw}) ARB_TRY(b->reserve(nnz * 4));
    ARB_TRY(code.reserve(nnz * 2));
    const int max_pairs = 65536;
    ARB_TRY(t_lp.reserve(max_pairs * 8));
    ARB_TRY(t_la.reserve(max_pairs * 8));
    ARB_TRY(cnt.reserve(max_pairs * 8));
    ARB_CUDA(cudaMemcpyAsync(d_ptr.p, cell_ptr, (size_t)(n_cells + 1) * 4, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_col.p, snv_idx, nnz * 4, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_ll.p, log_likes, nnz * 16, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemsetAsync(cnt.p, 0, max_pairs * 8, st));
    k_csr_keys<<<grid_for(nnz, B), B, 0, st>>>(d_ptr.as<int32_t>(), d_col.as<int32_t>(), n_cells, n_snvs, nnz,
                                               key.as<uint64_t>());
    k_split_ll<<<grid_for(nnz, B), B, 0, st>>>(d_ll.as<double>(), nnz, lpb.as<uint64_t>(), lab.as<uint64_t>());
    k_iota<<<grid_for(nnz, B), B, 0, st>>>(v0.as<int32_t>(), nnz);
    // lexicographic (lp, la) order by two stable passes
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, lab.as<uint64_t>(), k1.as<uint64_t>(), v0.as<int32_t>(),
                                               v1.as<int32_t>(), nnz, 0, 64, st);
    }));
    k_gather_u64<<<grid_for(nnz, B), B, 0, st>>>(lpb.as<uint64_t>(), v1.as<int32_t>(), nnz, k1.as<uint64_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, k1.as<uint64_t>(), k2.as<uint64_t>(), v1.as<int32_t>(),
                                               v2.as<int32_t>(), nnz, 0, 64, st);
    }));
    k_pair_heads<<<grid_for(nnz, B), B, 0, st>>>(lpb.as<uint64_t>(), lab.as<uint64_t>(), v2.as<int32_t>(), nnz,
                                                 head.as<int32_t>());
    ARB_TRY(with_cub(ctx, [&](void *tmp, size_t &bytes) {
        return cub::DeviceScan::InclusiveSum(tmp, bytes, head.as<int32_t>(), rank.as<int32_t>(), nnz, st);
    }));
    int32_t n_pairs = 0;
    ARB_CUDA(cudaMemcpyAsync(&n_pairs, rank.as<int32_t>() + nnz - 1, 4, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    ARB_REQUIRE(n_pairs <= 65535, ARB_ERR_UNSUPPORTED,
                "arb_ingest_csr: %d distinct (log_present, log_absent) pairs; this release supports 65535", n_pairs);
    k_pair_assign<<<grid_for(nnz, B), B, 0, st>>>(lpb.as<uint64_t>(), lab.as<uint64_t>(), v2.as<int32_t>(),
                                                  head.as<int32_t>(), rank.as<int32_t>(), nnz, max_pairs,
                                                  raw.as<int32_t>(), t_lp.as<double>(), t_la.as<double>(),
                                                  cnt.as<unsigned long long>());
    std::vector<double> r_lp(n_pairs), r_la(n_pairs);
    std::vector<unsigned long long> r_cnt(n_pairs);
    ARB_CUDA(cudaMemcpyAsync(r_lp.data(), t_lp.p, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(r_la.data(), t_la.p, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(r_cnt.data(), cnt.p, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    std::vector<int64_t> cnt64(r_cnt.begin(), r_cnt.end());
    std::vector<int32_t> map = frequency_order(cnt64);
    ctx->n_pairs = n_pairs;
    ctx->pair_alt.assign(n_pairs, -1);
    ctx->pair_total.assign(n_pairs, -1);
    ctx->pair_count.assign(n_pairs, 0);
    ctx->tab_lp.assign(n_pairs, 0.0);
    ctx->tab_la.assign(n_pairs, 0.0);
    for (int r = 0; r < n_pairs; ++r) {
        ctx->pair_count[map[r]] = cnt64[r];
        ctx->tab_lp[map[r]] = r_lp[r];
        ctx->tab_la[map[r]] = r_la[r];
    }
    ARB_TRY(map_d.reserve((size_t)n_pairs * 4));
    ARB_CUDA(cudaMemcpyAsync(map_d.p, map.data(), (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
    k_remap_codes<<<grid_for(nnz, B), B, 0, st>>>(raw.as<int32_t>(), map_d.as<int32_t>(), nnz, code.as<uint16_t>());
    ARB_TRY(finish_canonical(ctx, key, code.as<uint16_t>()));
    for (DevBuf *b : {&d_ptr, &d_col, &d_ll, &key, &lpb, &lab, &v0, &v1, &v2, &k1, &k2, &head, &rank, &raw, &t_lp,
                      &t_la, &cnt, &map_d, &code})
        b->release();
    ctx->cell_order.release();
    ctx->snv_order.release();
    ARB_TRY(arb_refresh_tables(ctx));
    ctx->ingested = true;
    ctx->from_coo = false;
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// accessors
// ---------------------------------------------------------------------------------------
ARB_API int arb_get_dims(arb_ctx *ctx, int32_t *n_cells, int32_t *n_snvs, int64_t *nnz, int32_t *n_pairs)
{
    ARB_REQUIRE(ctx && ctx->ingested, ARB_ERR_INVALID, "arb_get_dims: no matrix ingested");
    if (n_cells) *n_cells = ctx->n_cells;
    if (n_snvs) *n_snvs = ctx->n_snvs;
    if (nnz) *nnz = ctx->nnz;
    if (n_pairs) *n_pairs = ctx->n_pairs;
    return ARB_OK;
}

ARB_API int arb_get_orders(arb_ctx *ctx, int32_t *cell_order, int32_t *snv_order)
{
    ARB_REQUIRE(ctx && ctx->ingested && ctx->from_coo, ARB_ERR_INVALID, "arb_get_orders: needs arb_ingest_coo");
    ARB_CUDA(cudaSetDevice(ctx->device));
    if (cell_order)
        ARB_CUDA(cudaMemcpy(cell_order, ctx->cell_order.p, (size_t)ctx->n_cells * 4, cudaMemcpyDeviceToHost));
    if (snv_order)
        ARB_CUDA(cudaMemcpy(snv_order, ctx->snv_order.p, (size_t)ctx->n_snvs * 4, cudaMemcpyDeviceToHost));
    return ARB_OK;
}

ARB_API int arb_get_count_pairs(arb_ctx *ctx, int32_t *alt, int32_t *total)
{
    ARB_REQUIRE(ctx && ctx->ingested && ctx->from_coo, ARB_ERR_INVALID, "arb_get_count_pairs: needs arb_ingest_coo");
    if (alt) memcpy(alt, ctx->pair_alt.data(), (size_t)ctx->n_pairs * 4);
    if (total) memcpy(total, ctx->pair_total.data(), (size_t)ctx->n_pairs * 4);
    return ARB_OK;
}

ARB_API int arb_get_loglik_table(arb_ctx *ctx, double *log_present, double *log_absent)
{
    ARB_REQUIRE(ctx && ctx->ingested, ARB_ERR_INVALID, "arb_get_loglik_table: no matrix ingested");
    if (log_present) memcpy(log_present, ctx->tab_lp.data(), (size_t)ctx->n_pairs * 8);
    if (log_absent) memcpy(log_absent, ctx->tab_la.data(), (size_t)ctx->n_pairs * 8);
    return ARB_OK;
}

ARB_API int arb_set_loglik_table(arb_ctx *ctx, int32_t n_pairs, const double *log_present, const double *log_absent)
{
    ARB_REQUIRE(ctx && ctx->ingested && log_present && log_absent, ARB_ERR_INVALID,
                "arb_set_loglik_table: bad argument");
    ARB_REQUIRE(n_pairs == ctx->n_pairs, ARB_ERR_INVALID, "arb_set_loglik_table: %d entries, table has %d", n_pairs,
                ctx->n_pairs);
    ARB_CUDA(cudaSetDevice(ctx->device));
    ctx->tab_lp.assign(log_present, log_present + n_pairs);
    ctx->tab_la.assign(log_absent, log_absent + n_pairs);
    ctx->batch_t0 = -1;
    // the layouts carry a range proof made with the old table: rebuild them at the next fit
    for (auto &l : ctx->layout) l.release();
    ctx->ws_kp = 0;
    return arb_refresh_tables(ctx);
}

// canonical matrix in the reference's layout; a test / interchange hook, so the host does the
// counting sort for the SNV-major copy
ARB_API int arb_export_sparse(arb_ctx *ctx, int snv_major, int32_t *ptr, int32_t *idx, double *log_likes)
{
    ARB_REQUIRE(ctx && ctx->ingested && ptr && idx && log_likes, ARB_ERR_INVALID, "arb_export_sparse: bad argument");
    ARB_CUDA(cudaSetDevice(ctx->device));
    const int64_t nnz = ctx->nnz;
    std::vector<int32_t> cell(nnz), snv(nnz);
    std::vector<uint16_t> code(nnz);
    ARB_CUDA(cudaMemcpy(cell.data(), ctx->coo_cell.p, nnz * 4, cudaMemcpyDeviceToHost));
    ARB_CUDA(cudaMemcpy(snv.data(), ctx->coo_snv.p, nnz * 4, cudaMemcpyDeviceToHost));
    ARB_CUDA(cudaMemcpy(code.data(), ctx->coo_code.p, nnz * 2, cudaMemcpyDeviceToHost));
    const int32_t n_rows = snv_major ? ctx->n_snvs : ctx->n_cells;
    const std::vector<int32_t> &row = snv_major ? snv : cell, &col = snv_major ? cell : snv;
    std::vector<int32_t> fill(n_rows + 1, 0);
    for (int64_t k = 0; k < nnz; ++k) ++fill[row[k] + 1];
    for (int32_t r = 0; r < n_rows; ++r) fill[r + 1] += fill[r];
    memcpy(ptr, fill.data(), (size_t)(n_rows + 1) * 4);
    for (int64_t k = 0; k < nnz; ++k) {       // stable: canonical order is (cell, snv)
        const int32_t o = fill[row[k]]++;
        idx[o] = col[k];
        log_likes[2 * (int64_t)o] = ctx->tab_lp[code[k]];
        log_likes[2 * (int64_t)o + 1] = ctx->tab_la[code[k]];
    }
    return ARB_OK;
}
