// arb_base.cu -- the reference's legacy tree scorer on the device (sm_100a).
//
// Reference: /root/reference/src/arborist/arborist_base.py
//   find_cell_assignments :102-194  per (cell, clone): sum over the cell's rows of binom.logpmf(alt, total, p),
//                                   p = 0.5 - e + 0.5 e / 3 when the row's SNV cluster is the clone or one of its
//                                   ancestors, e / 3 otherwise; tree score = sum over cells of the best clone
//   rank_trees            :250-362  scores of all trees, then per-cell posteriors / MAP clone / entropy (bits) of
//                                   the top-n trees
//
// The per-row log-pmf does not depend on the tree, so the rows are folded ONCE into two dense tables
// TA / TB [cell][cluster] (sum of log-pmf at the "mutated" / "not mutated" probability) plus a row-count table;
// every tree is then a pass over those tables (n_cells x n_clusters doubles, L2-resident) with its ancestor bit
// masks.  All sums run in a fixed order: results do not change from run to run.
#include "arb_common.cuh"

#include <algorithm>
#include <cub/cub.cuh>
#include <mutex>

#define ARB_BASE_MAXC 32      // clusters (bit masks are 32 bits wide)
#define ARB_BASE_MAXK 64      // clones per tree

namespace {

template <class F> int base_cub(arb_ctx *ctx, F f)
{
    size_t bytes = 0;
    ARB_CUDA(f(nullptr, bytes));
    ARB_TRY(ctx->scratch.reserve(std::max<size_t>(bytes, 16)));
    bytes = ctx->scratch.bytes;
    ARB_CUDA(f(ctx->scratch.p, bytes));
    return ARB_OK;
}

__global__ void k_base_iota(int32_t *v, int64_t n)
{
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int32_t)i;
}

// scipy.stats.binom.logpmf(k, n, p): combiln + xlogy(k, p) + xlog1py(n - k, -p)
__device__ __forceinline__ double binom_logpmf(int k, int n, double logp, double log1mp)
{
    if (k < 0 || k > n) return -INFINITY;
    double comb = lgamma((double)n + 1.0) - (lgamma((double)k + 1.0) + lgamma((double)(n - k) + 1.0));
    return comb + (k ? k * logp : 0.0) + (n - k ? (n - k) * log1mp : 0.0);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One warp per cell: its rows (sorted by cell, original order inside a cell) in chunks of 32; lane c keeps the
// accumulators of cluster c.
__global__ void __launch_bounds__(256) k_base_tables(const int32_t *__restrict__ sorted_cell, const int32_t *__restrict__ perm,
                                                     const int32_t *__restrict__ cluster, const int32_t *__restrict__ total,
                                                     const int32_t *__restrict__ alt, int64_t n_rows, int n_cells, int C,
                                                     double lpm, double l1pm, double lpn, double l1pn,
                                                     double *__restrict__ TA, double *__restrict__ TB, int32_t *__restrict__ CNT)
{
    int cell = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (cell >= n_cells) return;
    // [lo, hi) = rows of this cell
    int64_t lo = 0, hi = n_rows;
    {
        int64_t a = 0, b = n_rows;
        while (a < b) { int64_t mid = (a + b) >> 1; if (sorted_cell[mid] < cell) a = mid + 1; else b = mid; }
        lo = a;
        b = n_rows;
        while (a < b) { int64_t mid = (a + b) >> 1; if (sorted_cell[mid] <= cell) a = mid + 1; else b = mid; }
        hi = a;
    }
    double accA = 0, accB = 0;
    int cnt = 0;
    for (int64_t base = lo; base < hi; base += 32) {
        int64_t r = base + lane;
        int k = -1;
        double vA = 0, vB = 0;
        if (r < hi) {
            int row = perm[r];
            k = cluster[row];
            int a = alt[row], d = total[row];
            vA = binom_logpmf(a, d, lpm, l1pm);
            vB = binom_logpmf(a, d, lpn, l1pn);
        }
        for (int c = 0; c < C; ++c) {
            bool mine = k == c;
            unsigned m = __ballot_sync(0xffffffffu, mine);
            if (!m) continue;
            double sA = warp_sum(mine ? vA : 0.0), sB = warp_sum(mine ? vB : 0.0);
            if (lane == c) { accA += sA; accB += sB; cnt += __popc(m); }
        }
    }
    if (lane < C) {
        TA[(int64_t)cell * C + lane] = accA;
        TB[(int64_t)cell * C + lane] = accB;
        CNT[(int64_t)cell * C + lane] = cnt;
    }
}

// log-likelihood of one cell under every clone of one tree; returns the number of rows that survive the tree's
// cluster filter (arborist_base.py:153-155).
__device__ __forceinline__ int base_cell_loglik(const double *__restrict__ TA, const double *__restrict__ TB,
                                                const int32_t *__restrict__ CNT, int64_t cell, int C, int K,
                                                const uint32_t *anc, uint32_t tmask, double *L)
{
    double ta[ARB_BASE_MAXC], tb[ARB_BASE_MAXC];
    int rows = 0;
    for (int k = 0; k < C; ++k) {
        ta[k] = TA[cell * C + k];
        tb[k] = TB[cell * C + k];
        if ((tmask >> k) & 1u) rows += CNT[cell * C + k];
    }
    for (int c = 0; c < K; ++c) {
        uint32_t a = anc[c];
        double s = 0;
        for (int k = 0; k < C; ++k)
            if ((tmask >> k) & 1u) s += ((a >> k) & 1u) ? ta[k] : tb[k];
        L[c] = s;
    }
    return rows;
}

// One CTA per tree: sum over cells of the best clone's log-likelihood.
__global__ void __launch_bounds__(256) k_base_score(const double *__restrict__ TA, const double *__restrict__ TB,
                                                    const int32_t *__restrict__ CNT, int n_cells, int C, int K,
                                                    const uint32_t *__restrict__ anc_all, const uint32_t *__restrict__ tmask_all,
                                                    double *__restrict__ lik)
{
    __shared__ uint32_t anc[ARB_BASE_MAXK];
    __shared__ double part[256];
    int t = blockIdx.x;
    for (int c = threadIdx.x; c < K; c += blockDim.x) anc[c] = anc_all[(int64_t)t * K + c];
    __syncthreads();
    uint32_t tmask = tmask_all[t];
    double L[ARB_BASE_MAXK];
    double acc = 0;
    for (int cell = threadIdx.x; cell < n_cells; cell += blockDim.x) {
        int rows = base_cell_loglik(TA, TB, CNT, cell, C, K, anc, tmask, L);
        if (!rows) continue;
        double best = L[0];
        for (int c = 1; c < K; ++c) best = L[c] > best ? L[c] : best;
        acc += best;
    }
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if (threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) lik[t] = part[0];
}

// Per-cell detail of the selected trees (arborist_base.py:296-300, 335-342).
__global__ void __launch_bounds__(128) k_base_cells(const double *__restrict__ TA, const double *__restrict__ TB,
                                                    const int32_t *__restrict__ CNT, int n_cells, int C, int K,
                                                    const uint32_t *__restrict__ anc_all, const uint32_t *__restrict__ tmask_all,
                                                    double *__restrict__ loglik, double *__restrict__ post,
                                                    int32_t *__restrict__ clone_idx, double *__restrict__ entropy,
                                                    uint8_t *__restrict__ present)
{
    __shared__ uint32_t anc[ARB_BASE_MAXK];
    int s = blockIdx.y;
    for (int c = threadIdx.x; c < K; c += blockDim.x) anc[c] = anc_all[(int64_t)s * K + c];
    __syncthreads();
    int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= n_cells) return;
    double L[ARB_BASE_MAXK];
    int rows = base_cell_loglik(TA, TB, CNT, cell, C, K, anc, tmask_all[s], L);
    int64_t o = (int64_t)s * n_cells + cell;
    present[o] = rows > 0;
    int arg = 0;
    double best = L[0];
    for (int c = 1; c < K; ++c)
        if (L[c] > best) { best = L[c]; arg = c; }          // first maximum, like DataFrame.idxmax
    clone_idx[o] = arg;
    double sum = 0;
    for (int c = 0; c < K; ++c) sum += exp(L[c] - best);
    double lse = log(sum) + best;                           // scipy.special.logsumexp
    double psum = 0;
    for (int c = 0; c < K; ++c) {
        double p = exp(L[c] - lse);
        if (loglik) loglik[o * K + c] = L[c];
        post[o * K + c] = p;
        L[c] = p;
        psum += p;
    }
    double h = 0;                                           // scipy.stats.entropy(row, base=2)
    for (int c = 0; c < K; ++c) {
        double p = L[c] / psum;
        if (p > 0) h -= p * log(p);
    }
    entropy[o] = h / log(2.0);
}

}  // namespace

ARB_API int arb_base_ingest(arb_ctx *ctx, int64_t n_rows, const int32_t *cell_code, const int32_t *cluster_code,
                            const int32_t *total, const int32_t *alt, int32_t n_cells, int32_t n_clusters,
                            double alpha)
{
    ARB_REQUIRE(ctx && cell_code && cluster_code && total && alt, ARB_ERR_INVALID, "arb_base_ingest: null argument");
    ARB_REQUIRE(n_rows > 0 && n_rows < (int64_t)INT32_MAX, ARB_ERR_INVALID, "arb_base_ingest: n_rows %lld out of range",
                (long long)n_rows);
    ARB_REQUIRE(n_cells > 0, ARB_ERR_INVALID, "arb_base_ingest: no cells");
    ARB_REQUIRE(n_clusters > 0 && n_clusters <= ARB_BASE_MAXC, ARB_ERR_UNSUPPORTED,
                "arb_base_ingest: %d SNV clusters; at most %d are supported", n_clusters, ARB_BASE_MAXC);
    ARB_REQUIRE(alpha > 0 && alpha < 1, ARB_ERR_INVALID, "arb_base_ingest: alpha %g not in (0, 1)", alpha);
    for (int64_t i = 0; i < n_rows; ++i) {
        ARB_REQUIRE(cell_code[i] >= 0 && cell_code[i] < n_cells && cluster_code[i] >= 0 && cluster_code[i] < n_clusters,
                    ARB_ERR_INVALID, "arb_base_ingest: row %lld has a code out of range", (long long)i);
    }
    ARB_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    ctx->base_cells = ctx->base_clusters = 0;
    DevBuf d_cell, d_clu, d_tot, d_alt, d_idx, d_cell_sorted, d_perm;
    size_t nb = (size_t)n_rows * sizeof(int32_t);
    for (DevBuf *b : {&d_cell, &d_clu, &d_tot, &d_alt, &d_idx, &d_cell_sorted, &d_perm}) ARB_TRY(b->reserve(nb));
    ARB_CUDA(cudaMemcpyAsync(d_cell.p, cell_code, nb, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_clu.p, cluster_code, nb, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_tot.p, total, nb, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_alt.p, alt, nb, cudaMemcpyHostToDevice, st));
    k_base_iota<<<(unsigned)((n_rows + 255) / 256), 256, 0, st>>>(d_idx.as<int32_t>(), n_rows);
    int bits = 1;
    while (bits < 31 && (n_cells >> bits)) ++bits;
    ARB_TRY(base_cub(ctx, [&](void *tmp, size_t &bytes) {          // stable: rows of a cell keep their input order
        return cub::DeviceRadixSort::SortPairs(tmp, bytes, d_cell.as<int32_t>(), d_cell_sorted.as<int32_t>(),
                                               d_idx.as<int32_t>(), d_perm.as<int32_t>(), (int)n_rows, 0, bits, st);
    }));
    size_t tb = (size_t)n_cells * n_clusters;
    ARB_TRY(ctx->base_ta.reserve(tb * sizeof(double)));
    ARB_TRY(ctx->base_tb.reserve(tb * sizeof(double)));
    ARB_TRY(ctx->base_cnt.reserve(tb * sizeof(int32_t)));
    double p_mut = 0.5 - alpha + 0.5 * alpha / 3, p_not = alpha / 3;          // arborist_base.py:165-166
    k_base_tables<<<(unsigned)((n_cells + 7) / 8), 256, 0, st>>>(
        d_cell_sorted.as<int32_t>(), d_perm.as<int32_t>(), d_clu.as<int32_t>(), d_tot.as<int32_t>(), d_alt.as<int32_t>(),
        n_rows, n_cells, n_clusters, log(p_mut), log1p(-p_mut), log(p_not), log1p(-p_not), ctx->base_ta.as<double>(),
        ctx->base_tb.as<double>(), ctx->base_cnt.as<int32_t>());
    ARB_CUDA(cudaGetLastError());
    ARB_CUDA(cudaStreamSynchronize(st));
    for (DevBuf *b : {&d_cell, &d_clu, &d_tot, &d_alt, &d_idx, &d_cell_sorted, &d_perm}) b->release();
    ctx->base_cells = n_cells;
    ctx->base_clusters = n_clusters;
    return ARB_OK;
}

static int base_upload_masks(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint32_t *anc_masks,
                             const uint32_t *tree_masks, DevBuf &d_anc, DevBuf &d_tm)
{
    ARB_REQUIRE(ctx->base_cells > 0, ARB_ERR_INVALID, "legacy scorer: call arb_base_ingest first");
    ARB_REQUIRE(n_trees > 0 && anc_masks && tree_masks, ARB_ERR_INVALID, "legacy scorer: no trees");
    ARB_REQUIRE(n_clones > 0 && n_clones <= ARB_BASE_MAXK, ARB_ERR_UNSUPPORTED,
                "legacy scorer: %d clones per tree; at most %d are supported", n_clones, ARB_BASE_MAXK);
    ARB_CUDA(cudaSetDevice(ctx->device));
    ARB_TRY(d_anc.reserve((size_t)n_trees * n_clones * sizeof(uint32_t)));
    ARB_TRY(d_tm.reserve((size_t)n_trees * sizeof(uint32_t)));
    ARB_CUDA(cudaMemcpyAsync(d_anc.p, anc_masks, (size_t)n_trees * n_clones * sizeof(uint32_t), cudaMemcpyHostToDevice,
                             ctx->stream));
    ARB_CUDA(cudaMemcpyAsync(d_tm.p, tree_masks, (size_t)n_trees * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    return ARB_OK;
}

ARB_API int arb_base_score(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint32_t *anc_masks,
                           const uint32_t *tree_masks, double *likelihood)
{
    ARB_REQUIRE(ctx && likelihood, ARB_ERR_INVALID, "arb_base_score: null argument");
    DevBuf d_anc, d_tm, d_lik;
    ARB_TRY(base_upload_masks(ctx, n_trees, n_clones, anc_masks, tree_masks, d_anc, d_tm));
    ARB_TRY(d_lik.reserve((size_t)n_trees * sizeof(double)));
    k_base_score<<<n_trees, 256, 0, ctx->stream>>>(ctx->base_ta.as<double>(), ctx->base_tb.as<double>(),
                                                   ctx->base_cnt.as<int32_t>(), ctx->base_cells, ctx->base_clusters,
                                                   n_clones, d_anc.as<uint32_t>(), d_tm.as<uint32_t>(), d_lik.as<double>());
    ARB_CUDA(cudaGetLastError());
    ARB_CUDA(cudaMemcpyAsync(likelihood, d_lik.p, (size_t)n_trees * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    for (DevBuf *b : {&d_anc, &d_tm, &d_lik}) b->release();
    return ARB_OK;
}

ARB_API int arb_base_cells(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint32_t *anc_masks,
                           const uint32_t *tree_masks, double *loglik, double *posterior, int32_t *clone_idx,
                           double *entropy, uint8_t *present)
{
    ARB_REQUIRE(ctx && posterior && clone_idx && entropy && present, ARB_ERR_INVALID, "arb_base_cells: null argument");
    DevBuf d_anc, d_tm, d_ll, d_post, d_idx, d_ent, d_pres;
    ARB_TRY(base_upload_masks(ctx, n_trees, n_clones, anc_masks, tree_masks, d_anc, d_tm));
    size_t cells = (size_t)n_trees * ctx->base_cells;
    if (loglik) ARB_TRY(d_ll.reserve(cells * n_clones * sizeof(double)));
    ARB_TRY(d_post.reserve(cells * n_clones * sizeof(double)));
    ARB_TRY(d_idx.reserve(cells * sizeof(int32_t)));
    ARB_TRY(d_ent.reserve(cells * sizeof(double)));
    ARB_TRY(d_pres.reserve(cells));
    dim3 grid((unsigned)((ctx->base_cells + 127) / 128), (unsigned)n_trees);
    k_base_cells<<<grid, 128, 0, ctx->stream>>>(ctx->base_ta.as<double>(), ctx->base_tb.as<double>(),
                                                ctx->base_cnt.as<int32_t>(), ctx->base_cells, ctx->base_clusters, n_clones,
                                                d_anc.as<uint32_t>(), d_tm.as<uint32_t>(), loglik ? d_ll.as<double>() : nullptr,
                                                d_post.as<double>(), d_idx.as<int32_t>(), d_ent.as<double>(),
                                                d_pres.as<uint8_t>());
    ARB_CUDA(cudaGetLastError());
    cudaStream_t st = ctx->stream;
    if (loglik) ARB_CUDA(cudaMemcpyAsync(loglik, d_ll.p, cells * n_clones * sizeof(double), cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(posterior, d_post.p, cells * n_clones * sizeof(double), cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(clone_idx, d_idx.p, cells * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(entropy, d_ent.p, cells * sizeof(double), cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(present, d_pres.p, cells, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    for (DevBuf *b : {&d_anc, &d_tm, &d_ll, &d_post, &d_idx, &d_ent, &d_pres}) b->release();
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------------------------
// arb_base_score_static: the tree scorer for callers that must not allocate device memory
// ---------------------------------------------------------------------------------------------------------
// arb_base_score above takes three temporary device buffers per call (the uploaded masks and the result vector)
// from the allocation pool, which falls through to cudaMalloc whenever the pool has no block of that size.  Some
// hosts cannot afford that once their set-up phase is over:
//   - a process whose device memory already belongs to another allocator (a tensor framework's caching allocator,
//     an RMM pool, a JVM GPU plug-in): a late cudaMalloc there fails or fragments the other allocator's arena;
//   - a scoring loop with a latency budget: cudaMalloc / cudaFree take a process-wide driver lock and may wait for
//     all work on the device, so a call that allocates stalls every other stream of the process;
//   - a client running under a fixed device-memory quota (a MIG slice or an MPS client with a memory limit that the
//     tables of arb_base_ingest have already been budgeted against).
// For them every byte this variant needs beyond the caller's own arrays and the tables arb_base_ingest left in the
// context sits in ONE fixed array in the library's device image.  The CUDA runtime instantiates the array on every
// GPU on which this library's module gets loaded and keeps it until the process exits; that is the price, hence
// the modest size.
//
// What lives in the array during a call (all offsets multiples of 256 bytes):
//   tickets  uint32 [n_trees]           CTAs of a tree that have delivered their partial sum -- must start at ZERO
//   partial  double [n_trees * splits]  per-CTA partial scores (the real scratch of the split kernel)
//   anc      uint32 [n_trees * K]       staged copy of the caller's ancestor masks
//   tmask    uint32 [n_trees]           staged copy of the caller's tree masks
//   lik      double [n_trees]           result, copied back to the caller
// There is one array per GPU and every call on that GPU uses it from offset 0, and it still holds whatever the last
// call left there: the tickets are cleared at the start of each call (the kernel also resets them, but a call that
// died half-way must not poison the next), everything else is overwritten before it is read.  Two calls that are in
// flight on the same GPU at the same time (two contexts, two host threads) would overwrite each other's masks and
// tickets: the caller has to serialise them.
static __device__ __align__(256) unsigned char g_base_static[ARB_BASE_STATIC_BYTES];

namespace {

constexpr int BASE_STATIC_MAX_SPLIT = 64;      // CTAs per tree at most
constexpr int BASE_STATIC_MAX_DEVICES = 64;

// In host code `g_base_static` is only a handle for the runtime, not a device address: the address on each GPU has
// to be asked for with that GPU current.  Looked up by the first call on each device (which may also make the
// runtime load this library's module onto that GPU), then kept: let that first call happen during set-up, before any
// stream capture begins.
std::mutex g_base_static_mu;
unsigned char *g_base_static_addr[BASE_STATIC_MAX_DEVICES];

int base_static_address(int device, unsigned char **out)
{
    ARB_REQUIRE(device >= 0 && device < BASE_STATIC_MAX_DEVICES, ARB_ERR_UNSUPPORTED,
                "arb_base_score_static: device ordinal %d; ordinals below %d are supported", device,
                BASE_STATIC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(g_base_static_mu);
    if (!g_base_static_addr[device]) {
        ARB_CUDA(cudaSetDevice(device));
        void *p = nullptr;
        ARB_CUDA(cudaGetSymbolAddress(&p, g_base_static));
        ARB_REQUIRE(p != nullptr, ARB_ERR_CUDA, "arb_base_score_static: no device address for the static scratch array");
        g_base_static_addr[device] = static_cast<unsigned char *>(p);
    }
    *out = g_base_static_addr[device];
    return ARB_OK;
}

inline size_t up256(size_t n) { return (n + 255) & ~(size_t)255; }

// k_base_score with the cells of a tree dealt out to gridDim.x CTAs (a handful of trees no longer leaves most of
// the SMs idle).  CTA (s, t) sums the cells s * 256 + thread, stepping by gridDim.x * 256, and leaves its partial sum
// in the scratch; the CTA that draws the last ticket of tree t adds the partial sums in the order s = 0, 1, ... --
// the result does not depend on which CTA finishes last.  With one CTA per tree the additions are those of
// k_base_score.
__global__ void __launch_bounds__(256) k_base_score_split(const double *__restrict__ TA, const double *__restrict__ TB,
                                                          const int32_t *__restrict__ CNT, int n_cells, int C, int K,
                                                          const uint32_t *__restrict__ anc_all,
                                                          const uint32_t *__restrict__ tmask_all, double *partial,
                                                          unsigned int *tickets, double *__restrict__ lik)
{
    __shared__ uint32_t anc[ARB_BASE_MAXK];
    __shared__ double part[256];
    const int t = blockIdx.y, s = blockIdx.x, S = gridDim.x;
    for (int c = threadIdx.x; c < K; c += blockDim.x) anc[c] = anc_all[(int64_t)t * K + c];
    __syncthreads();
    const uint32_t tmask = tmask_all[t];
    double L[ARB_BASE_MAXK];
    double acc = 0;
    for (int cell = s * 256 + threadIdx.x; cell < n_cells; cell += S * 256) {
        int rows = base_cell_loglik(TA, TB, CNT, cell, C, K, anc, tmask, L);
        if (!rows) continue;
        double best = L[0];
        for (int c = 1; c < K; ++c) best = L[c] > best ? L[c] : best;
        acc += best;
    }
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if (threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        partial[(int64_t)t * S + s] = part[0];
        __threadfence();                                        // the partial sum is visible before the ticket is drawn
        if (atomicAdd(&tickets[t], 1u) == (unsigned)(S - 1)) {      // every other CTA of this tree has delivered
            __threadfence();
            const volatile double *p = partial + (int64_t)t * S;
            double sum = 0;
            for (int i = 0; i < S; ++i) sum += p[i];
            lik[t] = sum;
            tickets[t] = 0;
        }
    }
}

}  // namespace

ARB_API int arb_base_score_static(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint32_t *anc_masks,
                                  const uint32_t *tree_masks, double *likelihood)
{
    ARB_REQUIRE(ctx && likelihood, ARB_ERR_INVALID, "arb_base_score_static: null argument");
    ARB_REQUIRE(ctx->base_cells > 0, ARB_ERR_INVALID, "arb_base_score_static: call arb_base_ingest first");
    ARB_REQUIRE(n_trees > 0 && anc_masks && tree_masks, ARB_ERR_INVALID, "arb_base_score_static: no trees");
    ARB_REQUIRE(n_clones > 0 && n_clones <= ARB_BASE_MAXK, ARB_ERR_UNSUPPORTED,
                "arb_base_score_static: %d clones per tree; at most %d are supported", n_clones, ARB_BASE_MAXK);
    // CTAs per tree: about two waves over the SMs, no more than there are 256-cell slices
    const int64_t T = n_trees, K = n_clones;
    const int64_t slices = (ctx->base_cells + 255) / 256;
    int64_t S = (2 * (int64_t)std::max(ctx->sm_count, 1) + T - 1) / T;
    S = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(S, slices), BASE_STATIC_MAX_SPLIT));
    const size_t o_tick = 0;
    const size_t o_part = o_tick + up256((size_t)T * sizeof(unsigned int));
    const size_t o_anc = o_part + up256((size_t)(T * S) * sizeof(double));
    const size_t o_tm = o_anc + up256((size_t)(T * K) * sizeof(uint32_t));
    const size_t o_lik = o_tm + up256((size_t)T * sizeof(uint32_t));
    const size_t need = o_lik + up256((size_t)T * sizeof(double));
    ARB_REQUIRE(need <= (size_t)ARB_BASE_STATIC_BYTES && T <= 65535, ARB_ERR_UNSUPPORTED,
                "arb_base_score_static: %d trees of %d clones need %lld bytes of scratch, the static array holds %d; "
                "score the trees in smaller groups or use arb_base_score",
                n_trees, n_clones, (long long)need, (int)ARB_BASE_STATIC_BYTES);
    ARB_CUDA(cudaSetDevice(ctx->device));
    unsigned char *base = nullptr;
    ARB_TRY(base_static_address(ctx->device, &base));
    cudaStream_t st = ctx->stream;
    unsigned int *d_tick = reinterpret_cast<unsigned int *>(base + o_tick);
    double *d_part = reinterpret_cast<double *>(base + o_part);
    uint32_t *d_anc = reinterpret_cast<uint32_t *>(base + o_anc);
    uint32_t *d_tm = reinterpret_cast<uint32_t *>(base + o_tm);
    double *d_lik = reinterpret_cast<double *>(base + o_lik);
    ARB_CUDA(cudaMemsetAsync(d_tick, 0, (size_t)T * sizeof(unsigned int), st));
    ARB_CUDA(cudaMemcpyAsync(d_anc, anc_masks, (size_t)(T * K) * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaMemcpyAsync(d_tm, tree_masks, (size_t)T * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    k_base_score_split<<<dim3((unsigned)S, (unsigned)T), 256, 0, st>>>(
        ctx->base_ta.as<double>(), ctx->base_tb.as<double>(), ctx->base_cnt.as<int32_t>(), ctx->base_cells,
        ctx->base_clusters, n_clones, d_anc, d_tm, d_part, d_tick, d_lik);
    ARB_CUDA(cudaGetLastError());
    ARB_CUDA(cudaMemcpyAsync(likelihood, d_lik, (size_t)T * sizeof(double), cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    return ARB_OK;
}
