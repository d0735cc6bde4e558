// arb_fit.cu -- the per-tree variational fit of libarborist_b200.so (sm_100a).
//
// Replaces the reference's tree loop and everything under it
// (/root/reference/src/arborist/arborist.py, "A.py"):
//   A.py:491-520   enumerate_presence          -> k_masks_from_parents / k_masks_from_presence
//   A.py:200-250   compute_q_z_sparse          -> k_seed_s + k_sweep<16,WITH_G=true,..>  + k_finish<GL,IS_Z=true,..>
//   A.py:325-375   compute_q_y_sparse          -> k_seed_r + k_sweep<16,WITH_G=false,..> + k_finish<GL,IS_Z=false,..>
//   A.py:378-420   compute_likelihood_sparse   -> fused into the q(z) sweep (G accumulators)
//   A.py:253-299   entropy / KL                -> fused into k_finish (finish_row)
//   A.py:679-816   run_variational_inference   -> k_control (per-tree state machine on device)
//   A.py:1089-1126 the sequential tree loop    -> many trees in flight ("slots"), one round of
//                                                 kernels advances every active tree by one iteration
//
// Algebra (P is the 0/1 presence matrix, lp/la the per-nonzero
// log-likelihoods, w = lp - la, d = expm1(w)):
//   S[j,r] = sum_p q_y[j,p] P[p,r]            R[i,p] = sum_r q_z[i,r] P[p,r]
//   q_z logit[i,r] = sum_{k in row i} la_k + log(1 + d_k S[j_k,r])       (first term cancels)
//   q_y logit[j,p] = log q_y0[j,p] + sum_{k in col j} log(1 + d_k R[i_k,p])
//   E[log lik]     = sum_k la_k + sum_i sum_r q_z[i,r] G[i,r],  G[i,r] = sum_k w_k S[j_k,r]
// The sums of logs are accumulated as running products with the exponent peeled off every
// `renorm_every` factors; one log per (row, clone) is taken in the finishing kernel.
//
// Sweep kernel data flow: the columns of the trees in flight are packed side by side, four at a time, into
// "virtual slots" of ARB_VW = 16 columns (128-byte rows), so no lane works on a padding column.  A CTA owns
// (virtual slot, panel[, share of a split tail item]).  The panel of the gathered operand (S or R rows) is brought
// into shared memory with bulk async copies (cp.async.bulk + mbarrier, the TMA engine's 1-D path); each warp then
// walks its run of tiles of the sparse stream.  Four lanes share a segment (one row of the output); each lane owns
// four of the 16 columns and reads them with two LDS.128 per nonzero (sweep_step).  The per-row products
// (q(z) sweep: product and likelihood sum interleaved) leave as full 128-byte lines of "partial rows", one per
// (panel, row); k_finish adds the panels of a row, takes the one log per (row, clone), the softmax, the ELBO terms
// and writes the next sweep's operand.  See the comment above split_plan / k_sweep for the details.
#include "arb_common.cuh"

#include <algorithm>
#include <math.h>
#include <string.h>
#include <thread>

int arb_sweep_table_entries(const arb_ctx *ctx);
int arb_check_table_range(arb_ctx *ctx);

// Partial rows are written once by a sweep and read once by a finishing kernel: streaming 256-bit stores
// (st.global.cs.v4.f64) and loads.  Every lane's chunk is contiguous in the row.  Variants that were measured and
// lost (16-byte "pieces", an L1 prefetch of the stream, 128-bit accesses, the (d, w) table in shared memory, a
// private exp, stream groups) are recorded in profiles/r1i_summary.md and scripts/README.md and no longer built.
#define SWEEP_THREADS (ARB_SWEEP_WARPS * 32)
#define FIN_THREADS 256
#define FIN_ROWS 256          // rows of q per finishing CTA (fixed: the ELBO reduction order depends on it)
#define MASK_STRIDE 64        // per tree: 32 column masks (over clusters) + 32 row masks (over clones)
#define LN2 0.693147180559945309417232121458

enum { PH_EMPTY = 0, PH_INIT = 1, PH_ITER = 2 };

// q(z) / q(y) of a slot live in two planes: `cur` holds the current iterate, `bestb` the best one so far
// (A.py:762-763, 808-811 copy the arrays; here the next iterate is simply written to the plane that is
// not the best one, and the best plane is copied out once, when the tree is done).
struct SlotState {
    int32_t tree, phase, iter, converged;
    double best;
    int32_t cur, bestb;
};

// The fit's control block, on the device.  k_control rewrites the counts after every round; the sweeps, seeding
// and finishing kernels read them from here, so their grids can be sized by an upper bound the host knows and
// no launch waits for a read-back.  A copy of the block goes to a pinned ring after every round (fit_core).
struct Control {
    int32_t next_tree, n_trees, n_slots, max_iter;
    int32_t n_done, keep, mode, round;
    double tol, sum_la;
    int32_t n_y, n_z, n_init, n_act;      // trees in this round's q(y) list / q(z) list / first-pass list; slots in use
    int32_t n_vy, n_vz, chunks_y, chunks_z;   // virtual slots of the two sweeps, 4-column chunks per tree
    int32_t nbz, nby, nbmax, general_init;    // general_init: first passes run k_first_pass_general, not the sweep
    int32_t pad2[8];               // 128 bytes
};
static_assert(sizeof(Control) == 128, "Control is copied as one 128-byte block");

static inline int grid_for(int64_t n, int block) { return (int)std::max<int64_t>(1, (n + block - 1) / block); }

// ---------------------------------------------------------------------------------------
// genotype (ancestor-closure) masks: colmask[r] bit p = P[p,r], rowmask[p] bit r = P[p,r]
// ---------------------------------------------------------------------------------------
__global__ void k_masks_from_parents(const int32_t *__restrict__ parent, int32_t n_trees, int32_t K,
                                     int32_t normal, uint32_t *__restrict__ masks)
{
    __shared__ uint32_t col[32];
    const int t = blockIdx.x, r = threadIdx.x;
    if (r < 32) col[r] = 0;
    __syncthreads();
    if (r < K) {
        // P[p,r] = 1 iff cluster p is r itself or an ancestor of r (A.py:513-518, transposed)
        uint32_t m = 0;
        int a = r;
        for (int step = 0; step < K && a >= 0; ++step) {
            if (a != normal) m |= 1u << (a < normal ? a : a - 1);
            a = parent[(int64_t)t * K + a];
        }
        col[r] = m;
    }
    __syncthreads();
    uint32_t *out = masks + (int64_t)t * MASK_STRIDE;
    if (r < 32) out[r] = col[r];
    if (r < 32) {
        uint32_t m = 0;
        for (int c = 0; c < K; ++c) m |= ((col[c] >> r) & 1u) << c;
        out[32 + r] = (r < K - 1) ? m : 0u;
    }
}

__global__ void k_masks_from_presence(const uint8_t *__restrict__ presence, int32_t n_trees, int32_t K,
                                      uint32_t *__restrict__ masks)
{
    const int t = blockIdx.x, x = threadIdx.x, C = K - 1;
    const uint8_t *P = presence + (int64_t)t * C * K;
    uint32_t *out = masks + (int64_t)t * MASK_STRIDE;
    if (x < 32) {
        uint32_t m = 0;
        if (x < K) for (int p = 0; p < C; ++p) m |= (P[p * K + x] ? 1u : 0u) << p;
        out[x] = m;
        m = 0;
        if (x < C) for (int r = 0; r < K; ++r) m |= (P[x * K + r] ? 1u : 0u) << r;
        out[32 + x] = m;
    }
}

// ---------------------------------------------------------------------------------------
// Column packing ("virtual slots").  The sweep kernel works on rows of ARB_VW = 16 doubles
// (one 128-byte shared-memory wavefront per nonzero).  A tree needs kc = 4*ceil(cols/4) columns
// (cols = K for S rows / q(z), K-1 for R rows / q(y)); instead of padding every tree to 16,
// the trees active in a sweep are laid side by side in units of four columns ("chunks"):
// the tree at position `pos` of the sweep's list owns chunks pos*kc/4 .. (pos+1)*kc/4 - 1 of
// the flat chunk axis, chunk q lives in virtual slot q/4 at columns 4*(q%4)...  The sweep
// never learns which tree a column belongs to; the seeding and finishing kernels address
// the packed operand / partial rows through packed_off().
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ int64_t packed_off(int pos, int chunks_per_tree, int h, int64_t rows_per_slot, int64_t row)
{
    const int q = pos * chunks_per_tree + h;                 // flat chunk index
    return ((int64_t)(q >> 2) * rows_per_slot + row) * ARB_VW + (q & 3) * 4;
}

// Rows of q live in GL = ceil(K/4) consecutive lanes (any GL in 1..8; a warp holds TW = 32/GL rows, its last
// 32 - TW*GL lanes idle), lane h of the group holding columns 4h..4h+3.  `base` = first lane of the group.
// masked sums of a row spread over GL lanes:
// acc[c] = sum over source columns b < n_src with bit b of omask[c] set, in ascending b
// (fixed order => deterministic and independent of the tree shape).  Whole warps only.
template <int GL>
__device__ __forceinline__ void masked_rows(const double (&q)[4], const uint32_t (&omask)[4], int n_src, int base,
                                            double (&acc)[4])
{
    acc[0] = acc[1] = acc[2] = acc[3] = 0.0;
#pragma unroll
    for (int b = 0; b < 4 * GL; ++b) {
        if (b < n_src) {
            const double xb = __shfl_sync(0xffffffffu, q[b & 3], base + (b >> 2));
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if ((omask[c] >> b) & 1u) acc[c] += xb;
        }
    }
}

// fixed-order block reduction of per-thread partial sums; result valid in thread 0
__device__ __forceinline__ double block_sum(double v, double *sh)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[w] = v;
    __syncthreads();
    double s = 0.0;
    if (threadIdx.x == 0) for (int i = 0; i < nw; ++i) s += sh[i];
    return s;
}

#define SEED_THREADS 256

// four columns (32 bytes, 32-byte aligned) of a packed operand row in one 256-bit store
__device__ __forceinline__ void store_chunk(double2 *dst, const double (&v)[4])
{
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(dst), "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3]) : "memory");
}

// lane -> (row group of the warp, lane within the group); the lanes past the last whole group idle but keep
// shuffling with the last group
template <int GL> struct RowLanes {
    static constexpr int TW = 32 / GL;
    int grp, h, base;
    bool active;
    __device__ __forceinline__ RowLanes()
    {
        const int lane = threadIdx.x & 31, g0 = lane / GL;
        active = g0 < TW;
        grp = active ? g0 : TW - 1;
        h = lane - g0 * GL;
        base = grp * GL;
    }
};

// S rows (packed, at the tree's position in this round's q(z) list) from a q(y) matrix: the prior
// for a tree's first pass (A.py:733-743), or an explicit q(y) for the kernel-level hooks.
// grid (row blocks, trees in `list`)
template <int GL>
__global__ void __launch_bounds__(SEED_THREADS)
k_seed_s(const double *__restrict__ q_src, int64_t src_slot_stride, const int32_t *__restrict__ list,
         const int32_t *__restrict__ pos_z, const SlotState *__restrict__ st, const uint32_t *__restrict__ masks,
         int32_t m, int32_t K, double *__restrict__ xs, double *__restrict__ qy_cur, int copy_q, int np, int pr,
         int64_t plane_slots, const Control *__restrict__ ctl)
{
    if (ctl && (int)blockIdx.y >= ctl->n_init) return;       // grid.y is an upper bound
    const RowLanes<GL> rl;
    const int C = K - 1, h = rl.h;
    const int slot = list[blockIdx.y], pos = pos_z[slot];
    const uint32_t *mk = masks + (int64_t)st[slot].tree * MASK_STRIDE;
    uint32_t omask[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) omask[c] = 4 * h + c < K ? mk[4 * h + c] : 0u;
    const int64_t j = ((int64_t)blockIdx.x * (SEED_THREADS / 32) + (threadIdx.x >> 5)) * rl.TW + rl.grp;
    const bool live = rl.active && j < m;
    const double *src = q_src + (int64_t)slot * src_slot_stride;
    double q[4], acc[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) q[c] = (live && 4 * h + c < C) ? src[j * C + 4 * h + c] : 0.0;
    masked_rows<GL>(q, omask, C, rl.base, acc);
    if (live) {
        if (4 * h < K) {
            double2 *xn = reinterpret_cast<double2 *>(xs + packed_off(pos, (K + 3) / 4, h, (int64_t)np * pr, arb_perm(j, np, pr)));
            store_chunk(xn, acc);
        }
        if (copy_q)
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (4 * h + c < C) qy_cur[(((1 - st[slot].bestb) * plane_slots + slot) * m + j) * C + 4 * h + c] = q[c];
    }
}

// R rows (packed, at the tree's position in this round's q(y) list) from the tree's current q(z)
template <int GL>
__global__ void __launch_bounds__(SEED_THREADS)
k_seed_r(const double *__restrict__ qz_cur, const int32_t *__restrict__ list, const SlotState *__restrict__ st,
         const uint32_t *__restrict__ masks, int32_t n, int32_t K, double *__restrict__ xr, int np, int pr,
         int64_t plane_slots, const Control *__restrict__ ctl)
{
    if (ctl && (int)blockIdx.y >= ctl->n_y) return;          // grid.y is an upper bound
    const RowLanes<GL> rl;
    const int C = K - 1, h = rl.h;
    const int slot = list[blockIdx.y], pos = blockIdx.y;
    const uint32_t *mk = masks + (int64_t)st[slot].tree * MASK_STRIDE + 32;
    uint32_t omask[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) omask[c] = 4 * h + c < C ? mk[4 * h + c] : 0u;
    const int64_t i = ((int64_t)blockIdx.x * (SEED_THREADS / 32) + (threadIdx.x >> 5)) * rl.TW + rl.grp;
    const bool live = rl.active && i < n;
    double q[4], acc[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) q[c] = (live && 4 * h + c < K) ? qz_cur[((st[slot].cur * plane_slots + slot) * n + i) * K + 4 * h + c] : 0.0;
    masked_rows<GL>(q, omask, K, rl.base, acc);
    if (live && 4 * h < C) {
        double2 *xn = reinterpret_cast<double2 *>(xr + packed_off(pos, (C + 3) / 4, h, (int64_t)np * pr, arb_perm(i, np, pr)));
        store_chunk(xn, acc);
    }
}

// ---------------------------------------------------------------------------------------
// the sweep
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void renorm(double &p, int &e)
{
    const int hi = __double2hiint(p);
    if (hi >= 0x00100000) {          // positive normal: peel the exponent off, keep p in [1,2)
        e += (hi >> 20) - 1023;
        p = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(p));
    }
}

struct SweepArgs {
    const double *x;           // gathered operand, packed: [virtual slot][n_gather][KP], rows in interleaved
                               // panel order (arb_perm): panel p = rows p * panel_rows .. of the slot
    int64_t x_slot_stride;     // doubles per virtual slot
    int32_t n_gather;          // rows of the gathered operand per slot = n_panels * panel_rows (padded)
    const int4 *desc;          // [n_panels * WARPS * tiles_per_warp] {first quad, A quads, all quads, 0}
    const int32_t *rows;       // [.. * TW] row of every segment of every tile, -1 = none
    const void *stream;        // [quad][TW] of uint4 (32-bit words) or uint2 (16-bit words, W16)
    int32_t n_rows, panel_rows, n_panels, tiles_per_warp;
    int32_t n_vslots;          // virtual slots in this launch when n_vslots_dev is null
    const int32_t *n_vslots_dev;   // else read here (written by k_control): the host never learns the count, the
                               // grid is an upper bound and the CTAs past n_vslots * n_panels * split leave at once
    int32_t split;             // a (virtual slot, panel) is cut into `split` CTAs, each taking a share of every
                               // warp's run of tiles: fills the machine when few trees are active.  0: chosen
                               // ON THE DEVICE from the real n_vslots (split_for), the host only bounds the grid
    int32_t sm_count;
    const double2 *dw;         // per pair code: (d, w)
    int32_t n_pairs, tab_n, renorm_quads;   // not FOLD: renormalise every this many quads, 0 = after every factor
    double *out_m;             // [vslot][panel][row][KP] the segment's product (FOLD) or its mantissa; the q(z)
                               // sweep interleaves {product, sum_k w_k x} per column
    int32_t *out_e;            // [vslot][panel][row][KP] exponent (not FOLD)
    double2 tabc[32];          // W16 (<= 32 pair codes): the (d, w) table in the kernel's constant bank, so the
                               // look-ups of the G region stay off the shared-memory pipe
};

// Four clones of one nonzero: two LDS.128 from the staged panel, 4 x (DFMA + DMUL) for the
// running products and 4 x DFMA for the likelihood accumulators.
template <bool WITH_G>
__device__ __forceinline__ void sweep_step(const char *xa_base, const char *xb_base, uint32_t off, double d, double w,
                                           double (&p)[4], double (&g)[4])
{
    const double2 xa = *reinterpret_cast<const double2 *>(xa_base + off);
    const double2 xb = *reinterpret_cast<const double2 *>(xb_base + off);
    p[0] *= fma(d, xa.x, 1.0);
    p[1] *= fma(d, xa.y, 1.0);
    p[2] *= fma(d, xb.x, 1.0);
    p[3] *= fma(d, xb.y, 1.0);
    if (WITH_G) {
        g[0] = fma(w, xa.x, g[0]);
        g[1] = fma(w, xa.y, g[1]);
        g[2] = fma(w, xb.x, g[2]);
        g[3] = fma(w, xb.y, g[3]);
    }
}

// One CTA = (tree slot, panel, share); one warp = a run of tiles of that panel (see "panel layouts" in
// arb_ingest.cu).  GL = KP/4 lanes share a segment and each lane owns two 16-byte pieces (two columns each) of
// the gathered 128-byte row.  Which two differs between the sweeps, because the partial rows differ:
//   q(y) sweep (8 bytes per column):  lane h owns pieces 2h, 2h+1 (columns 4h..4h+3): its four products are
//       32 contiguous bytes, one 256-bit store, and a lane group writes one full 128-byte line;
//   q(z) sweep ({product, likelihood} = 16 bytes per column, 256-byte rows): lane h owns pieces h and h+4
//       (columns 2h, 2h+1, 8+2h, 8+2h+1), so that each of its two 256-bit stores lands, together with the other
//       three lanes of the group, on ONE full line (with pieces 2h, 2h+1 every store instruction touched both
//       lines of every row: 16 tag requests per instruction instead of 8).
// The two loads of a lane are swizzled on bit 2 of the lane id so that the eight lanes of a quarter-warp cover
// all 32 banks.  Every lane of a group loads the same uint2/uint4 of four stream words (a warp reads TW*8/16
// contiguous bytes per quad), three quads ahead of the one in use; there are no shuffles, no predicates and no
// per-segment pointers in the loop.
// A sweep launch is n_vslots * n_panels equal work items, one CTA per SM at a time.  Whole waves of them run as
// full-size CTAs; the items of the LAST, partial wave (all of them when the launch is less than one wave: small
// candidate sets, the last rounds of a fit, a strong-scaled shard) are cut into `s_tail` shares that each take
// a part of every warp's run of tiles, so the block scheduler can pack the tail onto all SMs.  A share re-loads
// the panel (224 KB from L2) and pays the kernel prologue, about one tile's time, which is what bounds the
// split.  Integer arithmetic only: the host (grid bound) and the device (actual plan) must agree bit for bit.
struct SplitPlan { int full, s_tail, total; };
__host__ __device__ static inline SplitPlan split_plan(int64_t items, int sms, int tiles_per_warp)
{
    SplitPlan pl;
    if (sms <= 0) sms = 1;
    pl.full = (int)(items / sms) * sms;
    const int rem = (int)(items - pl.full);
    const int smax = tiles_per_warp < 16 ? (tiles_per_warp < 1 ? 1 : tiles_per_warp) : 16;
    pl.s_tail = 1;
    if (rem > 0) {
        int64_t best = 0;
        for (int s = 1; s <= smax; ++s) {
            const int64_t sub_waves = ((int64_t)rem * s + sms - 1) / sms;
            const int64_t cost = sub_waves * ((tiles_per_warp + s - 1) / s + 1);     // in tile times
            if (s == 1 || cost < best) { best = cost; pl.s_tail = s; }
        }
    }
    pl.total = pl.full + rem * pl.s_tail;
    return pl;
}

template <bool W16> struct StreamQuad { typedef uint4 type; };
template <> struct StreamQuad<true> { typedef uint2 type; };
// the four stream words of a quad as {byte offset of the gathered row | code << 18}
__device__ __forceinline__ void quad_words(const uint4 &q, uint32_t (&w)[4]) { w[0] = q.x; w[1] = q.y; w[2] = q.z; w[3] = q.w; }
// 16-bit words {row | code << 11}; rows are 128 bytes (ARB_VW doubles), so word << 7 is the 32-bit form
__device__ __forceinline__ void quad_words(const uint2 &q, uint32_t (&w)[4])
{
    w[0] = (q.x << 7) & 0x7fff80u;
    w[1] = (q.x >> 9) & 0x7fff80u;
    w[2] = (q.y << 7) & 0x7fff80u;
    w[3] = (q.y >> 9) & 0x7fff80u;
}

template <int KP, bool WITH_G, bool FOLD, bool W16>
__global__ void __launch_bounds__(ARB_SWEEP_WARPS * 32, 1) k_sweep(const __grid_constant__ SweepArgs a)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *xs = reinterpret_cast<double *>(smem_raw);
    double2 *tab = reinterpret_cast<double2 *>(smem_raw + (size_t)(a.panel_rows + 1) * KP * 8);
    uint64_t *mbar = reinterpret_cast<uint64_t *>(tab + a.tab_n);

    const int n_vslots = a.n_vslots_dev ? __ldg(a.n_vslots_dev) : a.n_vslots;
    const int items = n_vslots * a.n_panels;
    int split, share, rest;
    if (a.split > 0) {                                     // forced (option "split"): every item cut alike
        split = a.split;
        if ((int)blockIdx.x >= items * split) return;     // the grid is an upper bound
        share = blockIdx.x / items;
        rest = blockIdx.x - share * items;
    } else {
        const SplitPlan pl = split_plan(items, a.sm_count, a.tiles_per_warp);
        if ((int)blockIdx.x >= pl.total) return;
        if ((int)blockIdx.x < pl.full) {
            split = 1; share = 0; rest = blockIdx.x;
        } else {
            const int t = blockIdx.x - pl.full;
            split = pl.s_tail; share = t % pl.s_tail; rest = pl.full + t / pl.s_tail;
        }
    }
    const int panel = rest / n_vslots;
    const int slot = rest % n_vslots;                    // virtual slot: 16 packed columns of the active trees
    const int row0 = panel * a.panel_rows;
    const int cnt_rows = min(a.panel_rows, a.n_gather - row0);
    const uint32_t bytes = (uint32_t)cnt_rows * KP * 8u;
    const uint32_t bar = smem_u32(mbar);

    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // TMA bulk copies of the panel: global -> shared, completion counted on the mbarrier
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        const char *src = reinterpret_cast<const char *>(a.x + (int64_t)slot * a.x_slot_stride + (int64_t)row0 * KP);
        const uint32_t dst = smem_u32(xs);
        for (uint32_t off = 0; off < bytes; off += 32768u) {
            const uint32_t sz = min(32768u, bytes - off);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst + off), "l"(src + off), "r"(sz), "r"(bar) : "memory");
        }
    }
    for (int c = threadIdx.x; c < a.tab_n; c += ARB_SWEEP_WARPS * 32) tab[c] = a.dw[c];
    if (threadIdx.x < KP) xs[(size_t)a.panel_rows * KP + threadIdx.x] = 0.0;      // the row NULL words point at
    const double2 dw0 = __ldg(a.dw);

    constexpr int GL = KP / 4;                     // lanes per segment; each lane owns four clones
    constexpr int TW = 32 / GL;                    // segments per tile
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int h = lane % GL, grp = lane / GL;
    const int swz = (lane >> 2) & 1;
    // columns (doubles) of the lane's two pieces: the one it loads first and the one it loads second
    const int off_a = WITH_G ? 2 * (h + GL * swz) : 4 * h + 2 * swz;
    const int off_b = WITH_G ? 2 * (h + GL * (1 - swz)) : 4 * h + 2 * (1 - swz);
    const char *xa_base = reinterpret_cast<const char *>(xs + off_a);
    const char *xb_base = reinterpret_cast<const char *>(xs + off_b);
    // this CTA's share of every warp's run of tiles
    const int U = a.tiles_per_warp;
    const int u0 = (int)((int64_t)share * U / split), u1 = (int)((int64_t)(share + 1) * U / split);
    const int64_t x0 = ((int64_t)panel * ARB_SWEEP_WARPS + warp) * U + u0;
    const int4 *__restrict__ desc = a.desc + x0;
    const int32_t *__restrict__ rows = a.rows + x0 * TW + grp;
    constexpr int RW = WITH_G ? 2 : 1;             // doubles per column of a partial row
    const int64_t out0 = ((int64_t)slot * a.n_panels + panel) * a.n_rows;
    const uint32_t out_stride = (uint32_t)KP * 8u * RW;
    char *out_a = reinterpret_cast<char *>(a.out_m) + (out0 * KP + off_a) * 8 * RW;
    char *out_b = reinterpret_cast<char *>(a.out_m) + (out0 * KP + off_b) * 8 * RW;
    int32_t *oute_a = FOLD ? nullptr : a.out_e + out0 * KP + off_a;
    const bool each = a.renorm_quads == 0;

    // start the stream before waiting for the panel
    int4 dn = __ldg(desc);
    int rown = __ldg(rows);
    typedef typename StreamQuad<W16>::type Quad;
    const Quad *__restrict__ sp = reinterpret_cast<const Quad *>(a.stream) + (int64_t)dn.x * TW + grp;
    Quad w0 = __ldg(sp), w1 = __ldg(sp + TW), w2 = __ldg(sp + 2 * TW);
    sp += 3 * TW;

    __syncthreads();
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_PANEL:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t"
                 "@p bra DONE_PANEL;\n\tbra WAIT_PANEL;\n\tDONE_PANEL:\n\t}" ::"r"(bar) : "memory");

    for (int u = u0; u < u1; ++u) {
        const int4 dc = dn;
        const int row = rown;
        // next tile's descriptor (the arrays carry one spare slot at the end)
        ++desc;
        rows += TW;
        dn = __ldg(desc);
        rown = __ldg(rows);
        double p[4] = {1.0, 1.0, 1.0, 1.0}, g[4] = {0.0, 0.0, 0.0, 0.0};
        int e[4] = {0, 0, 0, 0};
        int since = 0;
        // one quad = four steps of every segment of the tile; WC is consumed, then refilled three quads ahead
#define ARB_QUAD_A(WC)                                                                          \
    {                                                                                           \
        uint32_t w[4];                                                                          \
        quad_words(WC, w);                                                                      \
        WC = __ldg(sp);                                                                         \
        sp += TW;                                                                               \
        sweep_step<WITH_G>(xa_base, xb_base, w[0], dw0.x, dw0.y, p, g);                         \
        if (!FOLD && each) for (int c = 0; c < 4; ++c) renorm(p[c], e[c]);                      \
        sweep_step<WITH_G>(xa_base, xb_base, w[1], dw0.x, dw0.y, p, g);                         \
        if (!FOLD && each) for (int c = 0; c < 4; ++c) renorm(p[c], e[c]);                      \
        sweep_step<WITH_G>(xa_base, xb_base, w[2], dw0.x, dw0.y, p, g);                         \
        if (!FOLD && each) for (int c = 0; c < 4; ++c) renorm(p[c], e[c]);                      \
        sweep_step<WITH_G>(xa_base, xb_base, w[3], dw0.x, dw0.y, p, g);                         \
        if (!FOLD && ++since >= a.renorm_quads) {                                               \
            since = 0;                                                                          \
            for (int c = 0; c < 4; ++c) renorm(p[c], e[c]);                                     \
        }                                                                                       \
    }
#define ARB_QUAD_G(WC)                                                                          \
    {                                                                                           \
        uint32_t ww[4];                                                                         \
        quad_words(WC, ww);                                                                     \
        WC = __ldg(sp);                                                                         \
        sp += TW;                                                                               \
        _Pragma("unroll") for (int s = 0; s < 4; ++s) {                                         \
            const uint32_t code = ww[s] >> 18, off = ww[s] & 0x3ffffu;                          \
            const double2 dw = W16 ? a.tabc[code]                                               \
                               : (FOLD || code < (uint32_t)a.tab_n) ? tab[code] : __ldg(a.dw + code); \
            sweep_step<WITH_G>(xa_base, xb_base, off, dw.x, dw.y, p, g);                        \
            if (!FOLD && each) for (int c = 0; c < 4; ++c) renorm(p[c], e[c]);                  \
        }                                                                                       \
        if (!FOLD && ++since >= a.renorm_quads) {                                               \
            since = 0;                                                                          \
            for (int c = 0; c < 4; ++c) renorm(p[c], e[c]);                                     \
        }                                                                                       \
    }
#define ARB_ROT1 { const Quad t = w0; w0 = w1; w1 = w2; w2 = t; }
#define ARB_ROT2 { const Quad t = w2; w2 = w1; w1 = w0; w0 = t; }
        // ---- A region: the most frequent (alt,total) pair; d and w live in registers ------------
        {
            int q = dc.y;
            for (; q >= 3; q -= 3) { ARB_QUAD_A(w0) ARB_QUAD_A(w1) ARB_QUAD_A(w2) }
            if (q == 1) { ARB_QUAD_A(w0) ARB_ROT1 }
            else if (q == 2) { ARB_QUAD_A(w0) ARB_QUAD_A(w1) ARB_ROT2 }
        }
        // ---- G region: any pair; (d, w) from the table ---------------------------------------------
        {
            int q = dc.z - dc.y;
            for (; q >= 3; q -= 3) { ARB_QUAD_G(w0) ARB_QUAD_G(w1) ARB_QUAD_G(w2) }
            if (q == 1) { ARB_QUAD_G(w0) ARB_ROT1 }
            else if (q == 2) { ARB_QUAD_G(w0) ARB_QUAD_G(w1) ARB_ROT2 }
        }
#undef ARB_QUAD_A
#undef ARB_QUAD_G
#undef ARB_ROT1
#undef ARB_ROT2
        if (!FOLD) for (int c = 0; c < 4; ++c) renorm(p[c], e[c]);
        if (row >= 0) {
            // partial row: q(z) {m, g} per column, q(y) m; streaming 256-bit stores (written once, read once)
            const int64_t ro = (int64_t)row * out_stride;
            if (WITH_G) {
                asm volatile("st.global.cs.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(out_a + ro), "d"(p[0]), "d"(g[0]), "d"(p[1]), "d"(g[1]) : "memory");
                asm volatile("st.global.cs.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(out_b + ro), "d"(p[2]), "d"(g[2]), "d"(p[3]), "d"(g[3]) : "memory");
            } else {
                // the lane's chunk (columns 4h..4h+3, 32 bytes) in one store
                const double c0 = swz ? p[2] : p[0], c1 = swz ? p[3] : p[1], c2 = swz ? p[0] : p[2], c3 = swz ? p[1] : p[3];
                asm volatile("st.global.cs.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(out_a + ro - 16 * swz), "d"(c0), "d"(c1), "d"(c2), "d"(c3) : "memory");
            }
            if (!FOLD) {
                int32_t *ea = oute_a + (int64_t)row * KP;
                *reinterpret_cast<int2 *>(ea) = make_int2(e[0], e[1]);
                *reinterpret_cast<int2 *>(ea + (off_b - off_a)) = make_int2(e[2], e[3]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// finishing kernel: combine the panel partials, softmax, ELBO pieces, next gathered operand
// ---------------------------------------------------------------------------------------
// log(Sbar + S e^w) for a deep pair, as a two-term log-sum-exp (A.py:173-197): exact when either mass is 0 and
// for any |w| (e^w itself may overflow or underflow)
__device__ __forceinline__ double deep_term(double S, double Sb, double w)
{
    const double a = log(Sb), b = log(S) + w;           // log(0) = -inf is fine here
    const double hi = fmax(a, b), lo = fmin(a, b);
    return hi == -INFINITY ? hi : hi + log1p(exp(lo - hi));
}

struct FinishArgs {
    const double *part_m;           // [vslot][panel][row][16] ({m, g} interleaved for q(z)), packed columns
    const int32_t *part_e;          // exponents, null for range-proven (folded) partials
    int32_t n_panels, n_rows, K;
    const int32_t *slot_list;       // position in the sweep's list -> tree slot
    const int32_t *pos_out;         // q(y) only: tree slot -> position in this round's q(z) list
    const SlotState *st;
    const uint32_t *masks;
    const double *logq0, *logq0e;   // q(y) only: log prior, log(prior + 1e-12)
    double *q_cur;                  // [plane][slot][row][K or C]; the iterate goes to plane 1 - st[slot].bestb
    int64_t plane_slots;            // slots per plane
    double *x_next;                 // q(y) only: packed S rows for the q(z) sweep of the same round
    int32_t xn_np, xn_pr;           // panels / rows per panel of that sweep (interleaved row order)
    double *red;                    // [slot][3][nbmax]: lik, entropy, kl partial sums per CTA
    int32_t nbmax;
    int32_t given_q;                // elbo_terms hook: take q from q_cur instead of the softmax
    const Control *ctl;             // trees in the list are counted here (grid.y is an upper bound); null: grid.y
    // deep pairs (arb_common.cuh): CSR over this kernel's rows of the deep nonzeros (other-side index, w); null when
    // the matrix has none.  q_other = the posterior matrix of the OTHER side ([plane][slot][other_rows][other_cols]):
    // q(y) for the q(z) kernel (plane 1 - bestb: this round's iterate), q(z) for the q(y) kernel (plane cur).
    const int32_t *side_ptr, *side_idx;
    const double *side_w, *q_other;
    int32_t other_rows;
};

// The end of a row: A.py:245-250 softmax + A.py:268-273 entropy + likelihood term (q(z)),
// A.py:373-375 (+ log prior) softmax + A.py:293-299 KL + S rows (q(y)).  A row lives in the GL
// lanes of a group, lane h holds columns 4h..4h+3.  Executed by whole warps (shuffles); only memory
// accesses are guarded.  Masked sums run in ascending source order (deterministic, tree-independent).
template <int GL, bool IS_Z, bool HARD, bool SIDE>
__device__ __forceinline__ void finish_row(const FinishArgs &a, int slot, int pos_out, int64_t row, bool rowok, int h,
                                           int base, const uint32_t (&omask)[4], const uint32_t (&smask)[4],
                                           const double (&p)[4], const int (&e)[4], const double (&g_in)[4], double &acc1,
                                           double &acc2)
{
    const int K = a.K, C = K - 1, ncol = IS_Z ? K : C;
    const int64_t qbase = (((1 - a.st[slot].bestb) * a.plane_slots + slot) * a.n_rows + row) * ncol;
    double logit[4], g[4];
    bool v[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        v[c] = 4 * h + c < ncol;
        g[c] = g_in[c];
        double l = log(p[c]) + (double)e[c] * LN2;
        if (!IS_Z && v[c] && rowok) l += a.logq0[row * C + 4 * h + c];
        logit[c] = v[c] ? l : -INFINITY;
    }
    if (SIDE && rowok) {
        // Deep pairs, in the log domain like the reference (A.py:229-243 / 354-369): the nonzero's term is
        // log(Sbar e^la + S e^lp) - la = log(Sbar + S e^w), with S the mass of the other side's posterior row on the
        // ancestors / descendants and Sbar the mass on the rest, both summed directly (no 1 - S cancellation).
        const int ocols = IS_Z ? C : K;
        const int plane = IS_Z ? 1 - a.st[slot].bestb : a.st[slot].cur;
        const double *qo = a.q_other + ((int64_t)plane * a.plane_slots + slot) * (int64_t)a.other_rows * ocols;
        for (int k = a.side_ptr[row]; k < a.side_ptr[row + 1]; ++k) {
            const double *qr = qo + (int64_t)a.side_idx[k] * ocols;
            const double w = a.side_w[k];
            double S[4] = {0.0, 0.0, 0.0, 0.0}, Sb[4] = {0.0, 0.0, 0.0, 0.0};
            for (int b = 0; b < ocols; ++b) {
                const double x = qr[b];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    if ((smask[c] >> b) & 1u) S[c] += x;
                    else Sb[c] += x;
                }
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (v[c]) {
                    logit[c] += deep_term(S[c], Sb[c], w);
                    if (IS_Z) g[c] = fma(w, S[c], g[c]);
                }
            }
        }
    }
    double mx = fmax(fmax(logit[0], logit[1]), fmax(logit[2], logit[3]));
    constexpr bool POW2 = (GL & (GL - 1)) == 0;    // butterflies inside aligned groups, else gather the group
    if (POW2) {
#pragma unroll
        for (int o = GL / 2; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o, GL));
    } else {
        const double mine = mx;
#pragma unroll
        for (int k = 0; k < GL; ++k) mx = fmax(mx, __shfl_sync(0xffffffffu, mine, base + k));
    }
    double q[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) q[c] = v[c] ? exp(logit[c] - mx) : 0.0;
    double s = (q[0] + q[1]) + (q[2] + q[3]);
    if (POW2) {
#pragma unroll
        for (int o = GL / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o, GL);
    } else {
        const double mine = s;                    // same order on every lane of the group: same bits
        s = __shfl_sync(0xffffffffu, mine, base);
#pragma unroll
        for (int k = 1; k < GL; ++k) s += __shfl_sync(0xffffffffu, mine, base + k);
    }
    const double inv_s = 1.0 / s;                 // one division per row (A.py:169 divides every entry)
#pragma unroll
    for (int c = 0; c < 4; ++c) q[c] = q[c] * inv_s;
    if (IS_Z && a.given_q) {
#pragma unroll
        for (int c = 0; c < 4; ++c) q[c] = (v[c] && rowok) ? a.q_cur[qbase + 4 * h + c] : 0.0;
    }
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    if (!IS_Z) masked_rows<GL>(q, omask, ncol, base, acc);    // S rows of the new q(y)
    if (IS_Z && HARD) {
        // hard-assignment mode (A.py:596, 662-674): the likelihood term of the row's MAP clone (first maximum
        // of q(z), like numpy's argmax); the soft q(z) is still what the caller gets back
        double bv = -1.0;
        int bi = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (v[c] && q[c] > bv) { bv = q[c]; bi = 4 * h + c; }
        const double bv0 = bv;
        const int bi0 = bi;
#pragma unroll
        for (int k = 0; k < GL; ++k) {                // every lane scans the group in column order
            const double ov = __shfl_sync(0xffffffffu, bv0, base + k);
            const int oi = __shfl_sync(0xffffffffu, bi0, base + k);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (rowok) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (v[c]) {
                    if (!a.given_q) a.q_cur[qbase + 4 * h + c] = q[c];
                    if (bi == 4 * h + c) acc1 += g[c];
                }
            }
        }
        return;
    }
    if (rowok) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (v[c]) {
                if (!(IS_Z && a.given_q)) a.q_cur[qbase + 4 * h + c] = q[c];
                if (IS_Z) {
                    acc1 += q[c] * g[c];
                    if (q[c] > 0.0) acc2 -= q[c] * log(q[c] + 1e-12);
                } else {
                    acc1 += q[c] * (log(q[c] + 1e-12) - a.logq0e[row * C + 4 * h + c]);
                }
            }
        }
        if (!IS_Z && 4 * h < K) {
            double2 *xn = reinterpret_cast<double2 *>(a.x_next + packed_off(pos_out, (K + 3) / 4, h, (int64_t)a.xn_np * a.xn_pr,
                                                                          arb_perm(row, a.xn_np, a.xn_pr)));
            store_chunk(xn, acc);
        }
    }
}

#define FIN_WARPS (FIN_THREADS / 32)

// grid (row blocks of FIN_ROWS, trees of the sweep's list); a warp takes TW = 32/GL rows at a time
#ifndef ARB_FINZ_OCC
#define ARB_FINZ_OCC 4       // CTAs per SM the q(z) finishing kernel is compiled for (register cap 64 at 4, 85 at 3)
#endif
template <int GL, bool IS_Z, bool HARD, bool SIDE>
__global__ void __launch_bounds__(FIN_THREADS, SIDE ? 3 : (IS_Z ? ARB_FINZ_OCC : 4)) k_finish(const FinishArgs a)
{
    __shared__ double sh[FIN_WARPS];
    if (a.ctl && (int)blockIdx.y >= (IS_Z ? a.ctl->n_z : a.ctl->n_y)) return;
    constexpr int TW = 32 / GL, RW = IS_Z ? 2 : 1;
    const RowLanes<GL> rl;
    const int warp = threadIdx.x >> 5;
    const int h = rl.h, grp = rl.grp;
    const int pos = blockIdx.y, slot = a.slot_list[pos];
    const int np = a.n_panels;
    const int ncol = IS_Z ? a.K : a.K - 1, chunks = (ncol + 3) / 4;
    const bool keep = h < chunks;
    const int pos_out = IS_Z ? 0 : a.pos_out[slot];
    uint32_t omask[4] = {0u, 0u, 0u, 0u}, smask[4] = {0u, 0u, 0u, 0u};
    if (!IS_Z) {
        const uint32_t *mk = a.masks + (int64_t)a.st[slot].tree * MASK_STRIDE;
#pragma unroll
        for (int c = 0; c < 4; ++c) omask[c] = 4 * h + c < a.K ? mk[4 * h + c] : 0u;
    }
    if (SIDE) {
        // q(z) kernel: column masks of this lane's clones over the clusters; q(y) kernel: row masks of its clusters
        // over the clones
        const uint32_t *mk = a.masks + (int64_t)a.st[slot].tree * MASK_STRIDE + (IS_Z ? 0 : 32);
#pragma unroll
        for (int c = 0; c < 4; ++c) smask[c] = 4 * h + c < ncol ? mk[4 * h + c] : 0u;
    }
    const int64_t panel_stride = (int64_t)a.n_rows * ARB_VW;        // columns between panels
    double acc1 = 0.0, acc2 = 0.0;
    for (int it = 0; it < (FIN_ROWS + FIN_WARPS * TW - 1) / (FIN_WARPS * TW); ++it) {
        const int lr = (it * FIN_WARPS + warp) * TW + grp;             // row of this CTA's block of FIN_ROWS
        const int64_t row = (int64_t)blockIdx.x * FIN_ROWS + lr;
        const bool rowok = rl.active && lr < FIN_ROWS && row < a.n_rows;
        const bool mine = rowok && keep;
        double p[4] = {1.0, 1.0, 1.0, 1.0}, g[4] = {0.0, 0.0, 0.0, 0.0};
        int e[4] = {0, 0, 0, 0};
        if (mine) {
            // this lane's chunk of the packed partial rows: [vslot][panel][row][16 columns]
            const int q = pos * chunks + h;
            const int64_t o0 = ((int64_t)(q >> 2) * np * a.n_rows + row) * ARB_VW + (q & 3) * 4;
            const double2 *src = reinterpret_cast<const double2 *>(a.part_m + o0 * RW);
            const int64_t pstep = panel_stride * RW / 2;            // in double2
            // split every partial (positive, normal: 2^+-1000 at most) into mantissa and exponent with integer
            // ops; the mantissa product stays below 2^np, far inside the double range
            auto fold_in = [&](const double (&m)[4], int pn) {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int hi = __double2hiint(m[c]);
                    e[c] += (hi >> 20) - 1023;
                    p[c] *= __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(m[c]));
                }
                if ((pn & 255) == 255)
                    for (int c = 0; c < 4; ++c) renorm(p[c], e[c]);
            };
            if (IS_Z) {
                // HBM-bound (23 MB of partial rows per tree): the loads of panel pn+1 are issued before the
                // partials of panel pn are consumed
                double2 cur[4], nxt[4];
                // a chunk is 64 contiguous bytes: two 256-bit streaming loads (sm_100: ld.global.v4.f64)
                auto load4 = [&](double2 (&d)[4], const double2 *from) {
                    asm volatile("ld.global.cs.v4.f64 {%0, %1, %2, %3}, [%4];"
                                 : "=d"(d[0].x), "=d"(d[0].y), "=d"(d[1].x), "=d"(d[1].y) : "l"(from));
                    asm volatile("ld.global.cs.v4.f64 {%0, %1, %2, %3}, [%4];"
                                 : "=d"(d[2].x), "=d"(d[2].y), "=d"(d[3].x), "=d"(d[3].y) : "l"(from + 2));
                };
                load4(cur, src);
#pragma unroll 2
                for (int pn = 0; pn < np; ++pn) {
                    src += pstep;
                    if (pn + 1 < np) load4(nxt, src);
                    const double m[4] = {cur[0].x, cur[1].x, cur[2].x, cur[3].x};
                    g[0] += cur[0].y; g[1] += cur[1].y; g[2] += cur[2].y; g[3] += cur[3].y;
                    fold_in(m, pn);
#pragma unroll
                    for (int k = 0; k < 4; ++k) cur[k] = nxt[k];
                }
            } else {
                // issue-bound (three fp64 transcendentals per entry): the plain loop has the fewest instructions
#pragma unroll 2
                for (int pn = 0; pn < np; ++pn) {
                    double2 v0, v1;
                    asm volatile("ld.global.cs.v4.f64 {%0, %1, %2, %3}, [%4];"
                                 : "=d"(v0.x), "=d"(v0.y), "=d"(v1.x), "=d"(v1.y) : "l"(src));
                    const double m[4] = {v0.x, v0.y, v1.x, v1.y};
                    src += pstep;
                    fold_in(m, pn);
                }
            }
            if (a.part_e) {
                const int4 *se = reinterpret_cast<const int4 *>(a.part_e + o0);
                for (int pn = 0; pn < np; ++pn) {
                    const int4 v = __ldcs(se + pn * (panel_stride / 4));
                    e[0] += v.x; e[1] += v.y; e[2] += v.z; e[3] += v.w;
                }
            }
        }
        finish_row<GL, IS_Z, HARD, SIDE>(a, slot, pos_out, rowok ? row : 0, rowok, h, rl.base, omask, smask, p, e, g, acc1, acc2);
    }
    const double t1 = block_sum(acc1, sh);
    const double t2 = block_sum(acc2, sh);
    if (threadIdx.x == 0) {
        double *red = a.red + (int64_t)slot * 3 * a.nbmax;
        if (IS_Z) {
            red[blockIdx.x] = t1;
            red[a.nbmax + blockIdx.x] = t2;
        } else {
            red[2 * a.nbmax + blockIdx.x] = t1;
        }
    }
}

// ---------------------------------------------------------------------------------------
// First pass for priors whose rows do not sum to one.
// The sweeps use sum_p q_y[j,p] = 1: a nonzero's factor is (1 - S) e^la + S e^lp = e^la (1 + d S).  The reference
// works in the log domain on whatever q_y it is given (A.py:229-243, 404-418), i.e. with the row's own mass
// M_j = sum_p q_y[j,p]:  factor = e^la (M_j + d S)  and  E[log lik] = sum_k la_k M_jk + sum q_z G.  Every q(y)
// the fit computes is a softmax row (mass 1), so this only concerns the FIRST pass, which runs on the prior:
// initialize_q_y gives the row [gamma] when there is a single cluster (K = 2, A.py:552-567), and a caller may
// hand any q_y_init to the seam.  In that case the first pass of every tree runs here, straight from the
// canonical CSR matrix and the prior (one warp per cell, lanes over its nonzeros, S and M formed per nonzero
// in ascending cluster order like k_seed_s), and writes what k_finish<.,true,.> would: q(z), the entropy and
// likelihood partial sums (the latter with sum_k la_k (M_jk - 1) folded in, k_control adds sum_k la_k).
// ---------------------------------------------------------------------------------------
#define GEN_CH 8        // clones per walk over a row's nonzeros

__global__ void __launch_bounds__(FIN_THREADS)
k_first_pass_general(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ snv, const uint16_t *__restrict__ code,
                     const double2 *__restrict__ dw, const double *__restrict__ la_tab, const double *__restrict__ w_tab,
                     const uint8_t *__restrict__ deep_tab, const double *__restrict__ q0, int32_t n, int32_t m, int32_t K, const int32_t *__restrict__ list_init, const SlotState *__restrict__ st,
                     const uint32_t *__restrict__ masks, double *__restrict__ qz_cur, double *__restrict__ qy_cur,
                     int64_t plane_slots, double *__restrict__ red, int32_t nbmax, const Control *__restrict__ ctl)
{
    __shared__ double sh[FIN_WARPS];
    __shared__ double sh_l[FIN_WARPS][32], sh_g[FIN_WARPS][32];
    if ((int)blockIdx.y >= ctl->n_init) return;
    const int C = K - 1, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot = list_init[blockIdx.y];
    const uint32_t *mk = masks + (int64_t)st[slot].tree * MASK_STRIDE;
    const int plane = 1 - st[slot].bestb;
    // the prior is this tree's q(y) iterate 0 (A.py:762-763 keeps it as best_q_y)
    {
        double *dst = qy_cur + ((int64_t)plane * plane_slots + slot) * (int64_t)m * C;
        const int64_t total = (int64_t)m * C, per = (total + gridDim.x - 1) / gridDim.x;
        const int64_t lo = per * blockIdx.x, hi = min(total, lo + per);
        for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) dst[i] = q0[i];
    }
    double acc1 = 0.0, acc2 = 0.0;
    for (int it = 0; it < FIN_ROWS / FIN_WARPS; ++it) {
        const int64_t row = (int64_t)blockIdx.x * FIN_ROWS + it * FIN_WARPS + warp;
        const bool rowok = row < n;
        const int k0 = rowok ? row_ptr[row] : 0, k1 = rowok ? row_ptr[row + 1] : 0;
        for (int r0 = 0; r0 < K; r0 += GEN_CH) {
            uint32_t cm[GEN_CH];
#pragma unroll
            for (int c = 0; c < GEN_CH; ++c) cm[c] = r0 + c < K ? mk[r0 + c] : 0u;
            double p[GEN_CH], g[GEN_CH], ls[GEN_CH];
            int e[GEN_CH];
#pragma unroll
            for (int c = 0; c < GEN_CH; ++c) { p[c] = 1.0; g[c] = 0.0; e[c] = 0; ls[c] = 0.0; }
            for (int k = k0 + lane; k < k1; k += 32) {
                const int j = snv[k], cd = code[k];
                const double2 t = dw[cd];
                const double *qr = q0 + (int64_t)j * C;
                double S[GEN_CH], Sb[GEN_CH], M = 0.0;
#pragma unroll
                for (int c = 0; c < GEN_CH; ++c) S[c] = Sb[c] = 0.0;
                for (int b = 0; b < C; ++b) {               // ascending cluster order, like masked_rows
                    const double x = qr[b];
                    M += x;
#pragma unroll
                    for (int c = 0; c < GEN_CH; ++c) {
                        if ((cm[c] >> b) & 1u) S[c] += x;
                        else Sb[c] += x;
                    }
                }
                if (r0 == 0) acc1 += la_tab[cd] * (M - 1.0);
                if (deep_tab[cd]) {                          // deep pair: log domain, like finish_row
                    const double w = w_tab[cd];
#pragma unroll
                    for (int c = 0; c < GEN_CH; ++c) {
                        ls[c] += deep_term(S[c], Sb[c], w);
                        g[c] = fma(w, S[c], g[c]);
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < GEN_CH; ++c) {
                        p[c] *= fma(t.x, S[c], M);
                        renorm(p[c], e[c]);
                        g[c] = fma(t.y, S[c], g[c]);
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < GEN_CH; ++c) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    p[c] *= __shfl_xor_sync(0xffffffffu, p[c], o);
                    e[c] += __shfl_xor_sync(0xffffffffu, e[c], o);
                    g[c] += __shfl_xor_sync(0xffffffffu, g[c], o);
                    ls[c] += __shfl_xor_sync(0xffffffffu, ls[c], o);
                    renorm(p[c], e[c]);
                }
                if (lane == 0 && r0 + c < K) {
                    sh_l[warp][r0 + c] = log(p[c]) + (double)e[c] * LN2 + ls[c];
                    sh_g[warp][r0 + c] = g[c];
                }
            }
        }
        __syncwarp();
        // lane r holds clone r: softmax (A.py:245-250), entropy (A.py:268-273), likelihood term
        const bool v = lane < K;
        const double l = v ? sh_l[warp][lane] : -INFINITY;
        double mx = l;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        double q = v ? exp(l - mx) : 0.0, s = q;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        q = q * (1.0 / s);
        if (rowok && v) {
            qz_cur[(((int64_t)plane * plane_slots + slot) * n + row) * K + lane] = q;
            acc1 += q * sh_g[warp][lane];
            if (q > 0.0) acc2 -= q * log(q + 1e-12);
        }
        __syncwarp();
    }
    const double t1 = block_sum(acc1, sh);
    const double t2 = block_sum(acc2, sh);
    if (threadIdx.x == 0) {
        double *r = red + (int64_t)slot * 3 * nbmax;
        r[blockIdx.x] = t1;
        r[nbmax + blockIdx.x] = t2;
    }
}

// CSR pointer of the canonical (cell, snv)-sorted matrix
__global__ void k_row_ptr(const int32_t *__restrict__ cell, int64_t nnz, int32_t n, int32_t *__restrict__ ptr)
{
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k > nnz) return;
    const int prev = k == 0 ? -1 : cell[k - 1], cur = k == nnz ? n : cell[k];
    for (int r = prev + 1; r <= cur; ++r) ptr[r] = (int32_t)k;
}

// largest |row sum - 1| of a [m, C] matrix
__global__ void k_row_mass_dev(const double *__restrict__ q, int32_t m, int32_t C, unsigned long long *__restrict__ out)
{
    const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (j >= m) return;
    double s = 0.0;
    for (int c = 0; c < C; ++c) s += q[j * C + c];
    const double dev = fabs(s - 1.0);
    atomicMax(out, (unsigned long long)__double_as_longlong(dev == dev ? dev : 1.0));      // non-negative doubles order like integers
}

// entropy / KL of given matrices (kernel-level hook)
__global__ void k_entropy_kl(const double *qz, int64_t nz, const double *qy, const double *q0, int64_t ny,
                             double *out)
{
    __shared__ double sh[32];
    double ent = 0.0, kl = 0.0;
    for (int64_t i = threadIdx.x; i < nz; i += blockDim.x)
        if (qz[i] > 0.0) ent -= qz[i] * log(qz[i] + 1e-12);
    for (int64_t i = threadIdx.x; i < ny; i += blockDim.x) kl += qy[i] * (log(qy[i] + 1e-12) - log(q0[i] + 1e-12));
    const double e = block_sum(ent, sh);
    const double k = block_sum(kl, sh);
    if (threadIdx.x == 0) { out[0] = e; out[1] = k; }
}

// ---------------------------------------------------------------------------------------
// the per-tree state machine (A.py:745-816) and slot scheduling
// ---------------------------------------------------------------------------------------
__global__ void k_control(Control *ctl, SlotState *st, const double *__restrict__ red, int32_t *list_y,
                          int32_t *list_z, int32_t *list_init, int32_t *pos_z, int32_t *snap_tree, double *elbo_out,
                          int32_t *iters_out, double *trace, int first_call)
{
    const int s1 = ctl->n_slots, nbmax = ctl->nbmax;
    if (!first_call) {
        for (int s = threadIdx.x; s < s1; s += blockDim.x) {
            SlotState z = st[s];
            int snap = -1;
            if (z.phase != PH_EMPTY) {
                const int wr = 1 - z.bestb;                    // the plane this round's q(y) / q(z) went to
                z.cur = wr;
                const double *r = red + (int64_t)s * 3 * nbmax;
                double lik = ctl->sum_la, ent = 0.0, kl = 0.0;
                for (int b = 0; b < ctl->nbz; ++b) lik += r[b];
                for (int b = 0; b < ctl->nbz; ++b) ent += r[nbmax + b];
                bool done;
                if (z.phase == PH_INIT) {
                    // KL(q_y0 || q_y0) = 0 (A.py:750-759); hard-assignment mode returns the likelihood (A.py:664-676)
                    const double elbo = ctl->mode == ARB_MODE_ML ? lik : lik + ent - 0.0;
                    z.best = elbo;
                    z.iter = 0;
                    z.converged = 0;
                    z.phase = PH_ITER;
                    z.bestb = wr;                             // A.py:762-763
                    if (trace) trace[(int64_t)z.tree * (ctl->max_iter + 1)] = elbo;
                    done = ctl->max_iter == 0 || ctl->mode == ARB_MODE_ML;
                } else {
                    for (int b = 0; b < ctl->nby; ++b) kl += r[2 * nbmax + b];
                    const double elbo = lik + ent - kl;       // A.py:801-803
                    if (trace) trace[(int64_t)z.tree * (ctl->max_iter + 1) + z.iter + 1] = elbo;
                    if (fabs(elbo - z.best) < ctl->tol) z.converged = 1;      // A.py:805 (vs BEST)
                    if (elbo > z.best) {                                       // A.py:808-811
                        z.bestb = wr;
                        z.best = elbo;
                    }
                    z.iter += 1;
                    done = z.converged || z.iter >= ctl->max_iter;           // A.py:813 / loop bound
                }
                if (done) {
                    elbo_out[z.tree] = z.best;
                    iters_out[z.tree] = z.iter;
                    snap = z.tree * 2 + z.bestb;              // k_snapshot copies the best plane out
                    z.phase = PH_EMPTY;
                    z.tree = -1;
                    atomicAdd(&ctl->n_done, 1);
                }
            }
            st[s] = z;
            snap_tree[s] = snap;
        }
    }
    __syncthreads();
    // ---- slot refill and the next round's lists, all threads: thread t owns the slots [t * per, (t + 1) * per) so
    // that positions follow slot order; two block scans give every slot its place in the lists
    __shared__ int sc_a[256], sc_b[256];
    const int T = blockDim.x, t = threadIdx.x, per = (s1 + T - 1) / T;
    const int lo = min(s1, t * per), hi = min(s1, lo + per);
    const bool gen = ctl->general_init != 0;
    const int next0 = ctl->next_tree, avail = max(0, ctl->n_trees - next0);
    int ce = 0, ci = 0;
    for (int s = lo; s < hi; ++s) {
        const int ph = st[s].phase;
        ce += ph == PH_EMPTY;
        ci += ph == PH_ITER;
    }
    auto block_scan = [&](int v, int *buf, int &total) {        // exclusive prefix of v over the threads
        buf[t] = v;
        __syncthreads();
        for (int o = 1; o < T; o <<= 1) {
            const int add = t >= o ? buf[t - o] : 0;
            __syncthreads();
            buf[t] += add;
            __syncthreads();
        }
        total = buf[T - 1];
        const int ex = buf[t] - v;
        __syncthreads();
        return ex;
    };
    int tot_e = 0, tot_i = 0, tot_z = 0;
    const int pe = block_scan(ce, sc_a, tot_e);                 // empties before this thread's slots
    const int pi = block_scan(ci, sc_b, tot_i);                 // iterating slots before them
    const int my_new = max(0, min(ce, avail - pe));             // the first `avail` empties (slot order) get a tree
    const int cz = ci + (gen ? 0 : my_new);
    const int pz = block_scan(cz, sc_a, tot_z);
    int ke = 0, ki = 0, kz = 0;
    for (int s = lo; s < hi; ++s) {
        SlotState z = st[s];
        bool in_z = false;
        if (z.phase == PH_EMPTY) {
            const int rank = pe + ke++;
            if (rank < avail) {
                z.tree = next0 + rank;
                z.phase = PH_INIT;
                z.iter = 0;
                z.converged = 0;
                z.best = 0.0;
                z.cur = 0;
                z.bestb = 1;                                  // nothing kept yet: the first pass writes plane 0
                st[s] = z;
                list_init[rank] = s;
                in_z = !gen;
            }
        } else if (z.phase == PH_ITER) {
            list_y[pi + ki++] = s;
            in_z = true;
        }
        if (in_z) {
            pos_z[s] = pz + kz;                               // packed position in this round's q(z) sweep
            list_z[pz + kz] = s;
            ++kz;
        }
    }
    if (t == 0) {
        const int ni = min(tot_e, avail);
        ctl->next_tree = next0 + ni;
        ctl->n_y = tot_i;
        ctl->n_z = tot_z;
        ctl->n_init = ni;
        ctl->n_act = tot_i + ni;
        ctl->n_vy = (tot_i * ctl->chunks_y + 3) >> 2;
        ctl->n_vz = (tot_z * ctl->chunks_z + 3) >> 2;
        ctl->round += 1;
    }
}

// A.py:816: the posteriors of a finished tree's best iterate leave the slot (snap = 2 * tree + plane)
__global__ void k_snapshot(const int32_t *__restrict__ snap_tree, const double *__restrict__ qz_cur,
                           const double *__restrict__ qy_cur, int64_t nz, int64_t ny, double *__restrict__ best_qz,
                           double *__restrict__ best_qy, int64_t plane_slots)
{
    const int slot = blockIdx.y, snap = snap_tree[slot];
    if (snap < 0) return;
    const int tree = snap >> 1, b = snap & 1;
    const double *sz = qz_cur + (b * plane_slots + slot) * nz, *sy = qy_cur + (b * plane_slots + slot) * ny;
    double *dz = best_qz + (int64_t)tree * nz, *dy = best_qy + (int64_t)tree * ny;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nz; i += stride) dz[i] = sz[i];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < ny; i += stride) dy[i] = sy[i];
}

// row argmax, first maximum wins (numpy argmax, treefit.py:105)
__global__ void k_argmax(const double *__restrict__ q, int64_t rows, int cols, int32_t *__restrict__ out)
{
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= rows) return;
    const double *r = q + i * cols;
    int best = 0;
    double bv = r[0];
    for (int c = 1; c < cols; ++c)
        if (r[c] > bv) { bv = r[c]; best = c; }
    // numpy treats NaN as the maximum; keep the first NaN like np.argmax does
    for (int c = 0; c < cols; ++c)
        if (r[c] != r[c]) { best = c; break; }
    out[i] = best;
}

// prior (A.py:523-569): gamma on the input cluster, (1-gamma)/(k-1) elsewhere, uniform 1/k
__global__ void k_prior_from_labels(const int32_t *__restrict__ label, int32_t m, int32_t C, double gamma,
                                    double other, double uniform, double *__restrict__ q0)
{
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= (int64_t)m * C) return;
    const int j = (int)(i / C), p = (int)(i % C), l = label[j];
    q0[i] = l < 0 ? uniform : (p == l ? gamma : other);
}

__global__ void k_prior_logs(const double *__restrict__ q0, int64_t n, double *__restrict__ lq, double *__restrict__ lqe)
{
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    lq[i] = q0[i] > 0.0 ? log(q0[i]) : -INFINITY;      // log_array, A.py:317-321
    lqe[i] = log(q0[i] + 1e-12);                      // A.py:296
}

// hard_assign_q of the prior (A.py:572-595): 1 at the row's first maximum, 0 elsewhere
__global__ void k_prior_hard(const double *__restrict__ q0, int32_t m, int32_t C, double *__restrict__ qh)
{
    const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (j >= m) return;
    const double *r = q0 + j * C;
    int best = 0;
    for (int c = 1; c < C; ++c)
        if (r[c] > r[best]) best = c;
    for (int c = 0; c < C; ++c) qh[j * C + c] = c == best ? 1.0 : 0.0;
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
struct FitTarget {                 // where a fit's per-tree results go (all device pointers)
    const uint32_t *masks;
    double *elbo;
    int32_t *iters;
    double *trace;                 // may be null
    double *best_qz, *best_qy;     // null => posteriors are not kept
};

static size_t sweep_smem_bytes(const arb_ctx *ctx, const PanelLayout &L)
{
    return (size_t)(L.panel_rows + 1) * L.kp * 8 + (size_t)arb_sweep_table_entries(ctx) * 16 + 16;
}

static inline int chunks_for(int cols) { return (cols + 3) / 4; }              // 4-column chunks per tree
static inline int64_t vslots_for(int64_t trees, int cols) { return (trees * chunks_for(cols) + 3) / 4; }
// lanes per row in the seeding / finishing kernels: power of two covering ceil(K/4) chunks
static inline int gl_for(int K) { return chunks_for(K); }          // 1..8, need not be a power of two

static int configure_kernels(const arb_ctx *ctx, const PanelLayout &L)
{
    const int smem = (int)sweep_smem_bytes(ctx, L);
#define ARB_SET_SMEM(G, F, W) \
    ARB_CUDA(cudaFuncSetAttribute(k_sweep<ARB_VW, G, F, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem))
    ARB_SET_SMEM(true, false, false); ARB_SET_SMEM(false, false, false); ARB_SET_SMEM(true, true, false);
    ARB_SET_SMEM(false, true, false); ARB_SET_SMEM(true, false, true); ARB_SET_SMEM(false, false, true);
    ARB_SET_SMEM(true, true, true); ARB_SET_SMEM(false, true, true);
#undef ARB_SET_SMEM
    return ARB_OK;
}

struct Workspace {
    int kp, K, n_slots, nbz, nby, nbmax;
    int64_t per_slot_bytes;
};

// bytes one more tree in flight costs (packed buffers grow by kc/16 of a virtual slot per tree)
static int64_t per_slot_bytes(const arb_ctx *ctx, const PanelLayout &L, int K)
{
    const int64_t n = ctx->n_cells, m = ctx->n_snvs, C = K - 1;
    const int64_t kcz = 4 * chunks_for(K), kcy = 4 * chunks_for(C);
    const int64_t nb = std::max((n + FIN_ROWS - 1) / FIN_ROWS, (m + FIN_ROWS - 1) / FIN_ROWS);
    const int64_t ez = L.fold_z ? 0 : 4, ey = L.fold_y ? 0 : 4;
    const int64_t mg = (int64_t)L.np_z * L.prz, ng = (int64_t)L.np_y * L.pry;      // padded gathered rows
    return (mg * kcz + ng * kcy + 2 * (n * K + m * C)) * 8 + (int64_t)L.np_z * n * kcz * (16 + ez) +
           (int64_t)L.np_y * m * kcy * (8 + ey) + 3 * nb * 8 + 64;
}

static int ensure_workspace(arb_ctx *ctx, int K, int want_slots, int64_t reserve_bytes, Workspace &ws)
{
    const int kp = kp_for(K);
    ARB_TRY(arb_build_layout(ctx, kp));
    const PanelLayout &L = ctx->layout[kp_index(kp)];
    const int64_t n = ctx->n_cells, m = ctx->n_snvs, C = K - 1;
    ws.kp = kp;
    ws.K = K;
    ws.nbz = (int)((n + FIN_ROWS - 1) / FIN_ROWS);
    ws.nby = (int)((m + FIN_ROWS - 1) / FIN_ROWS);
    ws.nbmax = std::max(ws.nbz, ws.nby);
    ws.per_slot_bytes = per_slot_bytes(ctx, L, K);
    int n_slots = want_slots;
    if (ctx->opt_slots > 0) n_slots = (int)std::min<int64_t>(n_slots, ctx->opt_slots);
    if (!(ctx->ws_kp == kp && ctx->ws_k == K && ctx->ws_slots >= n_slots)) {
        // release first so the free-memory query sees what we can really use
        for (DevBuf *b : {&ctx->xs, &ctx->xr, &ctx->qz_cur, &ctx->qy_cur, &ctx->part_z, &ctx->part_g, &ctx->part_y,
                          &ctx->red, &ctx->slot_state, &ctx->lists})
            b->release();
        ctx->ws_slots = 0;
        ctx->ws_kp = 0;
        size_t free_b = 0, total_b = 0;
        ARB_CUDA(cudaMemGetInfo(&free_b, &total_b));
        free_b += arb_pool_bytes(ctx->device);            // parked allocations are ours to reuse
        // one virtual slot of slack for the rounding of the packed buffers
        const int64_t mg = (int64_t)L.np_z * L.prz, ng = (int64_t)L.np_y * L.pry;
        const int64_t slack = (mg + ng) * ARB_VW * 8 + ((int64_t)L.np_z * n * 20 + (int64_t)L.np_y * m * 12) * ARB_VW;
        // leave a tenth of the free memory, at least 4 GB, to the other allocators of the process (NCCL buffers,
        // torch): a candidate set that does not fit in flight takes everything else
        // (at most a quarter of what is free: a small or shared GPU keeps most of its memory for the fit)
        const int64_t headroom = std::min<int64_t>(std::max<int64_t>((int64_t)(free_b * 0.10), (int64_t)4 << 30),
                                                   (int64_t)(free_b / 4));
        const int64_t raw = (int64_t)free_b - reserve_bytes - slack;
        const int64_t usable = raw - headroom;
        int64_t fit = usable > 0 ? usable / ws.per_slot_bytes : 0;
        if (fit < 1 && raw >= ws.per_slot_bytes) fit = 1;      // the headroom never turns a fit that has room into an error
        const bool tight = fit < n_slots;
        ARB_REQUIRE(fit >= 1, ARB_ERR_NOMEM, "not enough device memory for one tree slot (%lld bytes needed, %lld free)",
                    (long long)ws.per_slot_bytes, (long long)raw);
        if (fit < n_slots) n_slots = (int)fit;
        const int64_t S = n_slots, vz = vslots_for(S, K), vy = vslots_for(S, C);
        ARB_TRY(ctx->xs.reserve(vz * mg * ARB_VW * 8));
        ARB_TRY(ctx->xr.reserve(vy * ng * ARB_VW * 8));
        // columns past the last tree of a sweep are never read back, but the sweep multiplies through
        // them: keep them finite (stale S / R values or zero)
        ARB_CUDA(cudaMemsetAsync(ctx->xs.p, 0, vz * mg * ARB_VW * 8, ctx->stream));
        ARB_CUDA(cudaMemsetAsync(ctx->xr.p, 0, vy * ng * ARB_VW * 8, ctx->stream));
        ARB_TRY(ctx->qz_cur.reserve(2 * S * n * K * 8));      // two planes (SlotState)
        ARB_TRY(ctx->qy_cur.reserve(2 * S * m * C * 8));
        // product (8) [+ exponent (4) when the layout could not prove the range]
        ARB_TRY(ctx->part_z.reserve(vz * L.np_z * n * ARB_VW * (L.fold_z ? 16 : 20) + 64));
        ARB_TRY(ctx->part_y.reserve(vy * L.np_y * m * ARB_VW * (L.fold_y ? 8 : 12) + 64));
        ARB_TRY(ctx->red.reserve(S * 3 * ws.nbmax * 8));
        ARB_TRY(ctx->slot_state.reserve(S * sizeof(SlotState) + sizeof(Control) + 64));
        ARB_TRY(ctx->lists.reserve(S * 5 * 4 + 64));
        ctx->ws_kp = kp;
        ctx->ws_k = K;
        ctx->ws_slots = (int)S;
        // memory-tight: blocks still parked in the pool (ingest scratch) were counted as free above but no other
        // allocator can have them; hand them back so that the headroom is real
        if (tight) arb_pool_flush_device(ctx->device);
    }
    ws.n_slots = std::min(n_slots, ctx->ws_slots);          // a fit may use fewer slots than the workspace holds
    for (auto &r : ctx->rev)
        for (int i = 0; i < 7; ++i)
            if (!r[i])
                ARB_CUDA(i < 6 ? cudaEventCreate(&r[i]) : cudaEventCreateWithFlags(&r[i], cudaEventDisableTiming));
    for (auto &e : ctx->fit_ev)
        if (!e) ARB_CUDA(cudaEventCreate(&e));
    if (!ctx->h_ctl) ARB_CUDA(cudaMallocHost(&ctx->h_ctl, (ARB_RING + 1) * sizeof(Control)));
    return ARB_OK;
}

// device pointers of the workspace (sized for ctx->ws_slots slots)
struct WsPtrs {
    cudaStream_t stream;
    SlotState *st;
    Control *ctl;
    int32_t *list_y, *list_z, *list_init, *pos_z, *snap_tree;
    double *xs, *xr, *pz_m, *py_m;
    int32_t *pz_e, *py_e;
};

static WsPtrs ws_ptrs(arb_ctx *ctx, const PanelLayout &L, cudaStream_t stream)
{
    const int K = ctx->ws_k, C = K - 1;
    WsPtrs w;
    const int64_t S = ctx->ws_slots;
    const int64_t vz = vslots_for(S, K), vy = vslots_for(S, C);
    const int64_t n = ctx->n_cells, m = ctx->n_snvs;
    w.stream = stream;
    w.st = ctx->slot_state.as<SlotState>();
    w.ctl = reinterpret_cast<Control *>(reinterpret_cast<char *>(ctx->slot_state.p) + S * sizeof(SlotState));
    int32_t *lists = ctx->lists.as<int32_t>();
    w.list_y = lists;
    w.list_z = lists + S;
    w.list_init = lists + 2 * S;
    w.pos_z = lists + 3 * S;
    w.snap_tree = lists + 4 * S;
    w.xs = ctx->xs.as<double>();
    w.xr = ctx->xr.as<double>();
    double *pz = ctx->part_z.as<double>(), *py = ctx->part_y.as<double>();
    const int64_t pz_cols = vz * L.np_z * n * ARB_VW, py_cols = vy * L.np_y * m * ARB_VW;
    w.pz_m = pz;
    w.pz_e = L.fold_z ? nullptr : reinterpret_cast<int32_t *>(pz + 2 * pz_cols);
    w.py_m = py;
    w.py_e = L.fold_y ? nullptr : reinterpret_cast<int32_t *>(py + py_cols);
    return w;
}

static void fill_tabc(const arb_ctx *ctx, SweepArgs &a)
{
    memset(a.tabc, 0, sizeof a.tabc);
    const int n = std::min<int>(32, (int)ctx->h_dw.size() / 2);
    for (int c = 0; c < n; ++c) a.tabc[c] = make_double2(ctx->h_dw[2 * c], ctx->h_dw[2 * c + 1]);
}

template <bool WITH_G>
static void launch_sweep(const SweepArgs &a, int grid, int fold, int w16, size_t smem, cudaStream_t st)
{
    if (fold) {
        if (w16) k_sweep<ARB_VW, WITH_G, true, true><<<grid, SWEEP_THREADS, smem, st>>>(a);
        else k_sweep<ARB_VW, WITH_G, true, false><<<grid, SWEEP_THREADS, smem, st>>>(a);
    } else {
        if (w16) k_sweep<ARB_VW, WITH_G, false, true><<<grid, SWEEP_THREADS, smem, st>>>(a);
        else k_sweep<ARB_VW, WITH_G, false, false><<<grid, SWEEP_THREADS, smem, st>>>(a);
    }
}

// Grid of a sweep launch: the exact plan when the tree count is known on the host (kernel-level hooks), else the
// largest total any count up to the bound can ask for (the CTAs past the real total leave at once).
static int sweep_grid(const arb_ctx *ctx, SweepArgs &a, int ub_vslots, bool device_count)
{
    a.sm_count = ctx->sm_count;
    if (ctx->opt_split > 0) {
        a.split = (int)std::min<int64_t>(ctx->opt_split, std::max(1, a.tiles_per_warp));
        return ub_vslots * a.n_panels * a.split;
    }
    a.split = 0;
    if (!device_count) return split_plan((int64_t)ub_vslots * a.n_panels, ctx->sm_count, a.tiles_per_warp).total;
    int grid = 1;
    for (int v = 1; v <= ub_vslots; ++v)
        grid = std::max(grid, split_plan((int64_t)v * a.n_panels, ctx->sm_count, a.tiles_per_warp).total);
    return grid;
}

// q(z) sweep over the packed S rows of the trees in the q(z) list: `ub_trees` bounds the grid, the count itself
// is read on the device (ctl) unless ctl is null (kernel-level hooks: exactly ub_trees trees)
static void launch_sweep_z(arb_ctx *ctx, const PanelLayout &L, const WsPtrs &w, int ub_trees, const Control *ctl)
{
    SweepArgs a;
    a.x = w.xs;
    a.x_slot_stride = (int64_t)L.np_z * L.prz * ARB_VW;
    a.n_gather = L.np_z * L.prz;
    a.desc = L.zdesc.as<int4>();
    a.rows = L.zrows.as<int32_t>();
    a.stream = L.zstream.p;
    a.n_rows = ctx->n_cells;
    a.panel_rows = L.prz;
    a.n_panels = L.np_z;
    a.tiles_per_warp = L.uz;
    a.n_vslots = (int)vslots_for(ub_trees, ctx->ws_k);
    a.n_vslots_dev = ctl ? &ctl->n_vz : nullptr;
    a.dw = ctx->tab_dw.as<double2>();
    a.n_pairs = ctx->n_pairs;
    a.tab_n = arb_sweep_table_entries(ctx);
    a.renorm_quads = ctx->renorm_every / 4;
    a.out_m = w.pz_m;
    a.out_e = w.pz_e;
    fill_tabc(ctx, a);
    const int grid = sweep_grid(ctx, a, a.n_vslots, ctl != nullptr);
    launch_sweep<true>(a, grid, L.fold_z, L.w16, sweep_smem_bytes(ctx, L), w.stream);
}

static void launch_sweep_y(arb_ctx *ctx, const PanelLayout &L, const WsPtrs &w, int ub_trees, const Control *ctl)
{
    SweepArgs a;
    a.x = w.xr;
    a.x_slot_stride = (int64_t)L.np_y * L.pry * ARB_VW;
    a.n_gather = L.np_y * L.pry;
    a.desc = L.ydesc.as<int4>();
    a.rows = L.yrows.as<int32_t>();
    a.stream = L.ystream.p;
    a.n_rows = ctx->n_snvs;
    a.panel_rows = L.pry;
    a.n_panels = L.np_y;
    a.tiles_per_warp = L.uy;
    a.n_vslots = (int)vslots_for(ub_trees, ctx->ws_k - 1);
    a.n_vslots_dev = ctl ? &ctl->n_vy : nullptr;
    a.dw = ctx->tab_dw.as<double2>();
    a.n_pairs = ctx->n_pairs;
    a.tab_n = arb_sweep_table_entries(ctx);
    a.renorm_quads = ctx->renorm_every / 4;
    a.out_m = w.py_m;
    a.out_e = w.py_e;
    fill_tabc(ctx, a);
    const int grid = sweep_grid(ctx, a, a.n_vslots, ctl != nullptr);
    launch_sweep<false>(a, grid, L.fold_y, L.w16, sweep_smem_bytes(ctx, L), w.stream);
}

template <int GL>
static void launch_finish_z(arb_ctx *ctx, const PanelLayout &L, const WsPtrs &w, const Workspace &ws,
                            const uint32_t *masks, const int32_t *list, int ub_trees, int given_q, int hard,
                            const Control *ctl)
{
    FinishArgs a;
    a.part_m = w.pz_m;
    a.part_e = w.pz_e;
    a.n_panels = L.np_z;
    a.n_rows = ctx->n_cells;
    a.K = ws.K;
    a.slot_list = list;
    a.pos_out = nullptr;
    a.st = w.st;
    a.masks = masks;
    a.logq0 = nullptr;
    a.logq0e = nullptr;
    a.q_cur = ctx->qz_cur.as<double>();
    a.plane_slots = ctx->ws_slots;
    a.x_next = nullptr;
    a.xn_np = a.xn_pr = 1;
    a.red = ctx->red.as<double>();
    a.nbmax = ws.nbmax;
    a.given_q = given_q;
    a.ctl = ctl;
    a.side_ptr = ctx->n_side > 0 ? ctx->side_z_ptr.as<int32_t>() : nullptr;
    a.side_idx = ctx->side_z_idx.as<int32_t>();
    a.side_w = ctx->side_z_w.as<double>();
    a.q_other = ctx->qy_cur.as<double>();
    a.other_rows = ctx->n_snvs;
    const dim3 grid(ws.nbz, ub_trees);
    if (a.side_ptr) {                   // the matrix has deep pairs: the variant with the log-domain side loop
        if (hard) k_finish<GL, true, true, true><<<grid, FIN_THREADS, 0, w.stream>>>(a);
        else k_finish<GL, true, false, true><<<grid, FIN_THREADS, 0, w.stream>>>(a);
    } else {
        if (hard) k_finish<GL, true, true, false><<<grid, FIN_THREADS, 0, w.stream>>>(a);
        else k_finish<GL, true, false, false><<<grid, FIN_THREADS, 0, w.stream>>>(a);
    }
}

template <int GL>
static void launch_finish_y(arb_ctx *ctx, const PanelLayout &L, const WsPtrs &w, const Workspace &ws,
                            const uint32_t *masks, const int32_t *list, int ub_trees, const Control *ctl)
{
    FinishArgs a;
    a.part_m = w.py_m;
    a.part_e = w.py_e;
    a.n_panels = L.np_y;
    a.n_rows = ctx->n_snvs;
    a.K = ws.K;
    a.slot_list = list;
    a.pos_out = w.pos_z;
    a.st = w.st;
    a.masks = masks;
    a.logq0 = ctx->logq0.as<double>();
    a.logq0e = ctx->logq0e.as<double>();
    a.q_cur = ctx->qy_cur.as<double>();
    a.plane_slots = ctx->ws_slots;
    a.x_next = w.xs;
    a.xn_np = L.np_z;
    a.xn_pr = L.prz;
    a.red = ctx->red.as<double>();
    a.nbmax = ws.nbmax;
    a.given_q = 0;
    a.ctl = ctl;
    a.side_ptr = ctx->n_side > 0 ? ctx->side_y_ptr.as<int32_t>() : nullptr;
    a.side_idx = ctx->side_y_idx.as<int32_t>();
    a.side_w = ctx->side_y_w.as<double>();
    a.q_other = ctx->qz_cur.as<double>();
    a.other_rows = ctx->n_cells;
    if (a.side_ptr) k_finish<GL, false, false, true><<<dim3(ws.nby, ub_trees), FIN_THREADS, 0, w.stream>>>(a);
    else k_finish<GL, false, false, false><<<dim3(ws.nby, ub_trees), FIN_THREADS, 0, w.stream>>>(a);
}

template <int GL>
static void launch_seed_s(arb_ctx *ctx, const WsPtrs &w, const uint32_t *masks, const double *q_src,
                          int64_t src_stride, const int32_t *list, int ub_trees, int K, int copy_q, const Control *ctl)
{
    k_seed_s<GL><<<dim3(grid_for(ctx->n_snvs, (SEED_THREADS / 32) * (32 / GL)), ub_trees), SEED_THREADS, 0, w.stream>>>(
        q_src, src_stride, list, w.pos_z, w.st, masks, ctx->n_snvs, K, w.xs, ctx->qy_cur.as<double>(), copy_q,
        ctx->layout[kp_index(ctx->ws_kp)].np_z, ctx->layout[kp_index(ctx->ws_kp)].prz, (int64_t)ctx->ws_slots, ctl);
}

template <int GL>
static void launch_seed_r(arb_ctx *ctx, const WsPtrs &w, const uint32_t *masks, const int32_t *list, int ub_trees,
                          int K, const Control *ctl)
{
    k_seed_r<GL><<<dim3(grid_for(ctx->n_cells, (SEED_THREADS / 32) * (32 / GL)), ub_trees), SEED_THREADS, 0, w.stream>>>(
        ctx->qz_cur.as<double>(), list, w.st, masks, ctx->n_cells, K, w.xr, ctx->layout[kp_index(ctx->ws_kp)].np_y,
        ctx->layout[kp_index(ctx->ws_kp)].pry, (int64_t)ctx->ws_slots, ctl);
}

// One fit of n_trees trees whose masks are on the device.  A round = seeds -> q(y) sweep + finish -> q(z) sweep
// + finish -> k_control (per-tree state machine, slot refill, next round's lists and counts) -> k_snapshot.  The
// kernels of a round take their tree counts from the device-side Control block and their grids from an upper
// bound, so the host enqueues rounds WITHOUT waiting for the device: every round's Control block is copied to a
// pinned ring, and the host only blocks on the copy that is ARB_LOOKAHEAD rounds old (to tighten the grid bounds
// and to notice that every tree is done).  The few rounds enqueued past the end find n_z = 0 and leave at once.
template <int GL>
static int fit_core(arb_ctx *ctx, int n_trees, int K, int max_iter, double tol, int mode, const FitTarget &tg,
                    int64_t reserve)
{
    Workspace ws;
    ARB_TRY(ensure_workspace(ctx, K, std::min(n_trees, 1024), reserve, ws));
    const PanelLayout &L = ctx->layout[kp_index(ws.kp)];
    ARB_TRY(configure_kernels(ctx, L));
    const bool hard = mode == ARB_MODE_ML;
    const int64_t nz = (int64_t)ctx->n_cells * K, ny = (int64_t)ctx->n_snvs * (K - 1);
    cudaStream_t st = ctx->stream;
    const WsPtrs w = ws_ptrs(ctx, L, st);
    Control *ring = reinterpret_cast<Control *>(ctx->h_ctl);
    Control &c0 = ring[ARB_RING];                       // the block uploaded at the start (not part of the ring)
    memset(&c0, 0, sizeof c0);
    c0.n_trees = n_trees;
    c0.n_slots = ws.n_slots;
    c0.max_iter = max_iter;
    c0.tol = tol;
    c0.sum_la = ctx->sum_la;
    c0.nbz = ws.nbz;
    c0.nby = ws.nby;
    c0.nbmax = ws.nbmax;
    c0.keep = tg.best_qz != nullptr;
    c0.mode = mode;
    c0.chunks_z = chunks_for(K);
    c0.chunks_y = chunks_for(K - 1);
    // a prior with rows of mass != 1: first passes leave the packed sweep (k_first_pass_general)
    const bool general = ctx->prior_general && !hard;
    c0.general_init = general ? 1 : 0;
    if (general) {
        if (ctx->row_ptr.bytes < ((size_t)ctx->n_cells + 1) * 4 || !ctx->row_ptr.p) {
            ARB_TRY(ctx->row_ptr.reserve(((size_t)ctx->n_cells + 1) * 4));
            k_row_ptr<<<grid_for(ctx->nnz + 1, 256), 256, 0, ctx->stream>>>(ctx->coo_cell.as<int32_t>(), ctx->nnz,
                                                                           ctx->n_cells, ctx->row_ptr.as<int32_t>());
        }
        ARB_TRY(ctx->tab_la_dev.reserve((size_t)ctx->n_pairs * 8));
        ARB_CUDA(cudaMemcpyAsync(ctx->tab_la_dev.p, ctx->tab_la.data(), (size_t)ctx->n_pairs * 8, cudaMemcpyHostToDevice,
                                 ctx->stream));
    }
    ARB_CUDA(cudaMemsetAsync(ctx->slot_state.p, 0, (size_t)ctx->ws_slots * sizeof(SlotState), st));
    ARB_CUDA(cudaMemcpyAsync(w.ctl, &c0, sizeof c0, cudaMemcpyHostToDevice, st));
    ARB_CUDA(cudaEventRecord(ctx->fit_ev[0], st));
    FitStats stats;
    const int lookahead = ctx->opt_lookahead > 0 ? (int)ctx->opt_lookahead : ARB_LOOKAHEAD;

    // control #s writes the counts round s runs with and is followed by snapshot s
    auto control = [&](int s, int first) -> int {
        k_control<<<1, 256, 0, st>>>(w.ctl, w.st, ctx->red.as<double>(), w.list_y, w.list_z, w.list_init, w.pos_z,
                                     w.snap_tree, tg.elbo, tg.iters, tg.trace, first);
        stats.launches++;
        if (!first && tg.best_qz) {
            k_snapshot<<<dim3(16, ws.n_slots), 256, 0, st>>>(w.snap_tree, ctx->qz_cur.as<double>(),
                                                            ctx->qy_cur.as<double>(), nz, ny, tg.best_qz, tg.best_qy,
                                                            (int64_t)ctx->ws_slots);
            stats.launches++;
        }
        ARB_CUDA(cudaMemcpyAsync(&ring[s % ARB_RING], w.ctl, sizeof(Control), cudaMemcpyDeviceToHost, st));
        ARB_CUDA(cudaEventRecord(ctx->rev[s % ARB_RING][6], st));
        return ARB_OK;
    };

    struct Counts { int n_y, n_z, n_init, n_act; };
    Counts used[ARB_RING] = {};                            // what round r ran with (from snapshot r)
    int seen = -1, enq = 0, ub = std::min(ws.n_slots, n_trees);
    bool done = false, more_inits = true;
    // snapshot s has arrived: account for round s - 1 (its events precede the snapshot in the stream)
    auto process = [&](int s) {
        const Control c = ring[s % ARB_RING];
        if (s >= 1) {
            const Counts u = used[(s - 1) % ARB_RING];
            cudaEvent_t *ev = ctx->rev[(s - 1) % ARB_RING];
            float t = 0;
            if (u.n_act > 0) {
                stats.rounds += 1;
                if (u.n_y > 0) {
                    cudaEventElapsedTime(&t, ev[0], ev[1]); stats.sweep_y_ms += t;
                    cudaEventElapsedTime(&t, ev[1], ev[2]); stats.finish_ms += t; stats.finish_y_ms += t;
                    stats.launches_y += 1;
                    stats.tree_sweeps_y += u.n_y;
                }
                if (u.n_z > 0) {
                    cudaEventElapsedTime(&t, ev[3], ev[4]); stats.sweep_z_ms += t;
                    cudaEventElapsedTime(&t, ev[4], ev[5]); stats.finish_ms += t; stats.finish_z_ms += t;
                    stats.launches_z += 1;
                    stats.tree_sweeps_z += u.n_z;
                }
            } else {
                stats.idle_rounds += 1;
            }
        }
        used[s % ARB_RING] = Counts{c.n_y, c.n_z, c.n_init, c.n_act};
        ub = std::max(0, std::min(ws.n_slots, n_trees - c.n_done));
        more_inits = c.next_tree < n_trees;
        if (c.n_act == 0) done = true;
    };

    int rc = control(0, 1);
    while (rc == ARB_OK && !done) {
        // snapshots 0..enq exist; the one `lookahead` rounds back must have arrived before round enq is enqueued
        while (seen < enq && !done) {
            const int s = seen + 1;
            cudaEvent_t ev = ctx->rev[s % ARB_RING][6];
            cudaError_t e = cudaEventQuery(ev);
            if (e == cudaErrorNotReady) {
                if (s > enq - lookahead) break;
                stats.host_waits += 1;
                e = cudaEventSynchronize(ev);
            }
            if (e != cudaSuccess) {
                arb_set_error("fit round %d: %s", s, cudaGetErrorString(e));
                rc = ARB_ERR_CUDA;
                break;
            }
            process(s);
            seen = s;
        }
        if (rc != ARB_OK || done) break;
        // ---- enqueue round enq --------------------------------------------------------------------
        cudaEvent_t *ev = ctx->rev[enq % ARB_RING];
        cudaError_t rec_err = cudaSuccess;              // first failure of this round's event records
        auto rec = [&](cudaEvent_t e) {
            const cudaError_t r = cudaEventRecord(e, st);
            if (r != cudaSuccess && rec_err == cudaSuccess) rec_err = r;
        };
        const int ubt = std::max(1, ub);
        if ((more_inits || enq == 0) && general) {
            k_first_pass_general<<<dim3(ws.nbz, ubt), FIN_THREADS, 0, st>>>(
                ctx->row_ptr.as<int32_t>(), ctx->coo_snv.as<int32_t>(), ctx->coo_code.as<uint16_t>(),
                ctx->tab_dw.as<double2>(), ctx->tab_la_dev.as<double>(), ctx->tab_w_dev.as<double>(),
                ctx->tab_deep_dev.as<uint8_t>(), ctx->q0.as<double>(), ctx->n_cells, ctx->n_snvs, K,
                w.list_init, w.st, tg.masks, ctx->qz_cur.as<double>(), ctx->qy_cur.as<double>(), (int64_t)ctx->ws_slots,
                ctx->red.as<double>(), ws.nbmax, w.ctl);
            stats.launches++;
        } else if (more_inits || enq == 0) {
            launch_seed_s<GL>(ctx, w, tg.masks, hard ? ctx->q0_hard.as<double>() : ctx->q0.as<double>(), 0, w.list_init,
                              ubt, K, 1, w.ctl);
            stats.launches++;
        }
        if (enq > 0) {                                  // the first round has first passes only
            launch_seed_r<GL>(ctx, w, tg.masks, w.list_y, ubt, K, w.ctl);
            rec(ev[0]);
            launch_sweep_y(ctx, L, w, ubt, w.ctl);
            rec(ev[1]);
            launch_finish_y<GL>(ctx, L, w, ws, tg.masks, w.list_y, ubt, w.ctl);
            rec(ev[2]);
            stats.launches += 3;
        } else {
            rec(ev[0]);
            rec(ev[1]);
            rec(ev[2]);
        }
        rec(ev[3]);
        launch_sweep_z(ctx, L, w, ubt, w.ctl);
        rec(ev[4]);
        launch_finish_z<GL>(ctx, L, w, ws, tg.masks, w.list_z, ubt, 0, hard, w.ctl);
        rec(ev[5]);
        stats.launches += 2;
        {
            cudaError_t e = cudaGetLastError();
            if (e == cudaSuccess) e = rec_err;
            if (e != cudaSuccess) {
                arb_set_error("fit round %d: launch or event record failed: %s", enq, cudaGetErrorString(e));
                rc = ARB_ERR_CUDA;
                break;
            }
        }
        ++enq;
        rc = control(enq, 0);
    }
    // drain what is in flight (also after an error, so the workspace can be reused or released)
    cudaEventRecord(ctx->fit_ev[1], st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (rc == ARB_OK && e != cudaSuccess) {
        arb_set_error("fit: %s", cudaGetErrorString(e));
        rc = ARB_ERR_CUDA;
    }
    if (rc == ARB_OK) {
        for (int s = seen + 1; s <= enq; ++s) process(s);
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->fit_ev[0], ctx->fit_ev[1]);
        stats.device_ms = ms;
        e = cudaGetLastError();
        if (e != cudaSuccess) {
            arb_set_error("fit: %s", cudaGetErrorString(e));
            rc = ARB_ERR_CUDA;
        }
    }
    stats.sweep_ms = stats.sweep_z_ms + stats.sweep_y_ms;
    ctx->stats = stats;
    return rc;
}

static int fit_dispatch(arb_ctx *ctx, int n_trees, int K, int max_iter, double tol, int mode, const FitTarget &tg,
                        int64_t reserve)
{
    switch (gl_for(K)) {
    case 1: return fit_core<1>(ctx, n_trees, K, max_iter, tol, mode, tg, reserve);
    case 2: return fit_core<2>(ctx, n_trees, K, max_iter, tol, mode, tg, reserve);
    case 3: return fit_core<3>(ctx, n_trees, K, max_iter, tol, mode, tg, reserve);
    case 4: return fit_core<4>(ctx, n_trees, K, max_iter, tol, mode, tg, reserve);
    case 5: return fit_core<5>(ctx, n_trees, K, max_iter, tol, mode, tg, reserve);
    case 6: return fit_core<6>(ctx, n_trees, K, max_iter, tol, mode, tg, reserve);
    case 7: return fit_core<7>(ctx, n_trees, K, max_iter, tol, mode, tg, reserve);
    default: return fit_core<8>(ctx, n_trees, K, max_iter, tol, mode, tg, reserve);
    }
}

static int check_ready(arb_ctx *ctx, int K, const char *who)
{
    ARB_REQUIRE(ctx, ARB_ERR_INVALID, "%s: null context", who);
    ARB_REQUIRE(ctx->ingested, ARB_ERR_INVALID, "%s: no matrix ingested", who);
    ARB_REQUIRE(K >= 2 && K <= ARB_MAX_CLONES, K > ARB_MAX_CLONES ? ARB_ERR_UNSUPPORTED : ARB_ERR_INVALID,
                "%s: n_clones=%d outside 2..%d", who, K, ARB_MAX_CLONES);
    ARB_REQUIRE(ctx->n_clusters == K - 1, ARB_ERR_INVALID, "%s: prior has %d clusters, trees have %d", who,
                ctx->n_clusters, K - 1);
    ARB_TRY(arb_check_table_range(ctx));
    return ARB_OK;
}

static int run_fit(arb_ctx *ctx, int32_t n_trees, int32_t K, int32_t max_iter, double tol, int32_t mode,
                   int32_t keep, double *elbo, int32_t *iters, double *elbo_trace)
{
    ARB_REQUIRE(mode == ARB_MODE_VI || mode == ARB_MODE_ML, ARB_ERR_INVALID, "unknown fit mode %d", mode);
    ARB_REQUIRE(max_iter >= 0, ARB_ERR_INVALID, "max_iter=%d", max_iter);
    cudaStream_t st = ctx->stream;
    const int64_t nz = (int64_t)ctx->n_cells * K, ny = (int64_t)ctx->n_snvs * (K - 1);
    ARB_TRY(ctx->d_elbo.reserve((size_t)n_trees * 8));
    ARB_TRY(ctx->d_iters.reserve((size_t)n_trees * 4));
    FitTarget tg;
    memset(&tg, 0, sizeof tg);
    tg.masks = ctx->tree_masks.as<uint32_t>();
    tg.elbo = ctx->d_elbo.as<double>();
    tg.iters = ctx->d_iters.as<int32_t>();
    const size_t trace_n = (size_t)n_trees * (max_iter + 1);
    if (elbo_trace) {
        ARB_TRY(ctx->d_trace.reserve(trace_n * 8));
        ARB_CUDA(cudaMemsetAsync(ctx->d_trace.p, 0xff, trace_n * 8, st));     // all-ones = NaN
        tg.trace = ctx->d_trace.as<double>();
    }
    int64_t reserve = 0;
    ctx->fit_kept = false;
    if (keep) {
        // release the workspace first if the posteriors would not fit beside it
        ARB_TRY(ctx->best_qz.reserve((size_t)n_trees * nz * 8));
        ARB_TRY(ctx->best_qy.reserve((size_t)n_trees * ny * 8));
        tg.best_qz = ctx->best_qz.as<double>();
        tg.best_qy = ctx->best_qy.as<double>();
    } else {
        ctx->best_qz.release();
        ctx->best_qy.release();
    }
    ctx->fit_trees = 0;
    ctx->batch_t0 = -1;
    ARB_TRY(fit_dispatch(ctx, n_trees, K, max_iter, tol, mode, tg, reserve));
    ARB_CUDA(cudaMemcpyAsync(elbo, tg.elbo, (size_t)n_trees * 8, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaMemcpyAsync(iters, tg.iters, (size_t)n_trees * 4, cudaMemcpyDeviceToHost, st));
    if (elbo_trace) ARB_CUDA(cudaMemcpyAsync(elbo_trace, tg.trace, trace_n * 8, cudaMemcpyDeviceToHost, st));
    ARB_CUDA(cudaStreamSynchronize(st));
    double ti = 0;
    for (int t = 0; t < n_trees; ++t) ti += iters[t] + 0.5;
    ctx->stats.tree_iterations = ti;
    ctx->fit_trees = n_trees;
    ctx->fit_k = K;
    ctx->fit_max_iter = max_iter;
    ctx->fit_tol = tol;
    ctx->fit_mode = mode;
    ctx->fit_kept = keep != 0;
    return ARB_OK;
}

ARB_API int arb_fit_trees(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, int32_t normal_idx, const int32_t *parent,
                          int32_t max_iter, double tolerance, int32_t mode, int32_t keep_posteriors, double *elbo,
                          int32_t *iters, double *elbo_trace)
{
    ARB_TRY(check_ready(ctx, n_clones, "arb_fit_trees"));
    ARB_REQUIRE(n_trees > 0 && parent && elbo && iters, ARB_ERR_INVALID, "arb_fit_trees: bad argument");
    ARB_REQUIRE(normal_idx >= 0 && normal_idx < n_clones, ARB_ERR_INVALID, "arb_fit_trees: normal_idx=%d", normal_idx);
    for (int64_t i = 0; i < (int64_t)n_trees * n_clones; ++i)
        ARB_REQUIRE(parent[i] >= -1 && parent[i] < n_clones, ARB_ERR_INVALID, "arb_fit_trees: parent[%lld]=%d",
                    (long long)i, parent[i]);
    ARB_CUDA(cudaSetDevice(ctx->device));
    DevBuf d_parent;
    ARB_TRY(d_parent.reserve((size_t)n_trees * n_clones * 4));
    ARB_TRY(ctx->tree_masks.reserve((size_t)n_trees * MASK_STRIDE * 4));
    ARB_CUDA(cudaMemcpyAsync(d_parent.p, parent, (size_t)n_trees * n_clones * 4, cudaMemcpyHostToDevice, ctx->stream));
    k_masks_from_parents<<<n_trees, 32, 0, ctx->stream>>>(d_parent.as<int32_t>(), n_trees, n_clones, normal_idx,
                                                          ctx->tree_masks.as<uint32_t>());
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    d_parent.release();
    return run_fit(ctx, n_trees, n_clones, max_iter, tolerance, mode, keep_posteriors, elbo, iters, elbo_trace);
}

// The same fit over several contexts (one per GPU, every one holding the same matrix and prior): tree t goes to
// context t mod n_ctx as its local tree t / n_ctx; one host thread per context runs arb_fit_trees on its shard and
// writes the shard's ELBOs / iteration counts / traces straight into the caller's arrays -- candidate trees are
// independent (A.py:1089-1126), so the only "exchange" is those 12 bytes per tree and a single process needs no
// collective for it.  Keeps arborist(...) an ordinary function call on a multi-GPU box.
ARB_API int arb_fit_trees_multi(arb_ctx *const *ctxs, int32_t n_ctx, int32_t n_trees, int32_t n_clones,
                                int32_t normal_idx, const int32_t *parent, int32_t max_iter, double tolerance,
                                int32_t mode, int32_t keep_posteriors, double *elbo, int32_t *iters, double *elbo_trace)
{
    ARB_REQUIRE(ctxs && n_ctx >= 1 && n_ctx <= 64, ARB_ERR_INVALID, "arb_fit_trees_multi: %d contexts", n_ctx);
    ARB_REQUIRE(n_trees > 0 && parent && elbo && iters, ARB_ERR_INVALID, "arb_fit_trees_multi: bad argument");
    for (int c = 0; c < n_ctx; ++c) {
        ARB_REQUIRE(ctxs[c], ARB_ERR_INVALID, "arb_fit_trees_multi: context %d is null", c);
        for (int d = 0; d < c; ++d)
            ARB_REQUIRE(ctxs[c] != ctxs[d], ARB_ERR_INVALID, "arb_fit_trees_multi: context %d given twice", c);
        ARB_REQUIRE(ctxs[c]->ingested && ctxs[c]->n_cells == ctxs[0]->n_cells && ctxs[c]->n_snvs == ctxs[0]->n_snvs &&
                        ctxs[c]->nnz == ctxs[0]->nnz,
                    ARB_ERR_INVALID, "arb_fit_trees_multi: context %d holds a different matrix than context 0", c);
    }
    const int K = n_clones, W = max_iter + 1;
    std::vector<int> rcs(n_ctx, ARB_OK);
    std::vector<std::string> errs(n_ctx);
    std::vector<std::thread> workers;
    for (int c = 0; c < n_ctx; ++c) {
        workers.emplace_back([&, c]() {
            const int local_n = (n_trees - c + n_ctx - 1) / n_ctx;
            if (local_n <= 0) {
                ctxs[c]->fit_trees = 0;
                return;
            }
            std::vector<int32_t> par((size_t)local_n * K), it(local_n);
            std::vector<double> el(local_n), tr(elbo_trace ? (size_t)local_n * W : 0);
            for (int l = 0; l < local_n; ++l)
                memcpy(&par[(size_t)l * K], parent + ((size_t)l * n_ctx + c) * K, (size_t)K * 4);
            rcs[c] = arb_fit_trees(ctxs[c], local_n, K, normal_idx, par.data(), max_iter, tolerance, mode,
                                   keep_posteriors, el.data(), it.data(), elbo_trace ? tr.data() : nullptr);
            if (rcs[c] != ARB_OK) {
                errs[c] = arb_last_error();            // thread-local: carry it over to the calling thread
                return;
            }
            for (int l = 0; l < local_n; ++l) {
                const size_t t = (size_t)l * n_ctx + c;
                elbo[t] = el[l];
                iters[t] = it[l];
                if (elbo_trace) memcpy(elbo_trace + t * W, &tr[(size_t)l * W], (size_t)W * 8);
            }
        });
    }
    for (auto &w : workers) w.join();
    for (int c = 0; c < n_ctx; ++c)
        if (rcs[c] != ARB_OK) {
            arb_set_error("arb_fit_trees_multi: context %d (device %d): %s", c, ctxs[c]->device, errs[c].c_str());
            return rcs[c];
        }
    return ARB_OK;
}

ARB_API int arb_fit_presence(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, const uint8_t *presence,
                             int32_t max_iter, double tolerance, int32_t mode, int32_t keep_posteriors, double *elbo,
                             int32_t *iters, double *elbo_trace)
{
    ARB_TRY(check_ready(ctx, n_clones, "arb_fit_presence"));
    ARB_REQUIRE(n_trees > 0 && presence && elbo && iters, ARB_ERR_INVALID, "arb_fit_presence: bad argument");
    ARB_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = (size_t)n_trees * (n_clones - 1) * n_clones;
    DevBuf d_p;
    ARB_TRY(d_p.reserve(bytes));
    ARB_TRY(ctx->tree_masks.reserve((size_t)n_trees * MASK_STRIDE * 4));
    ARB_CUDA(cudaMemcpyAsync(d_p.p, presence, bytes, cudaMemcpyHostToDevice, ctx->stream));
    k_masks_from_presence<<<n_trees, 32, 0, ctx->stream>>>(d_p.as<uint8_t>(), n_trees, n_clones,
                                                           ctx->tree_masks.as<uint32_t>());
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    d_p.release();
    return run_fit(ctx, n_trees, n_clones, max_iter, tolerance, mode, keep_posteriors, elbo, iters, elbo_trace);
}

ARB_API int arb_posterior_bytes(arb_ctx *ctx, int32_t n_trees, int32_t n_clones, int64_t *bytes, int64_t *free_bytes)
{
    ARB_REQUIRE(ctx && ctx->ingested, ARB_ERR_INVALID, "arb_posterior_bytes: no matrix ingested");
    ARB_CUDA(cudaSetDevice(ctx->device));
    if (bytes)
        *bytes = (int64_t)n_trees * ((int64_t)ctx->n_cells * n_clones + (int64_t)ctx->n_snvs * (n_clones - 1)) * 8;
    if (free_bytes) {
        size_t f = 0, t = 0;
        ARB_CUDA(cudaMemGetInfo(&f, &t));
        f += arb_pool_bytes(ctx->device);
        *free_bytes = (int64_t)f + (int64_t)ctx->best_qz.bytes + (int64_t)ctx->best_qy.bytes;
    }
    return ARB_OK;
}

// Posteriors of trees t0 .. t0 + count - 1 of the last fit: resident, or re-fitted on demand as ONE batch (the
// fit is deterministic, so the re-fit reproduces every bit).  A writer that walks the whole candidate set
// (--pickle, A.py:1187-1188) therefore costs ceil(T / count) batched fits at full occupancy instead of T
// single-tree fits.
static int locate_fits(arb_ctx *ctx, int32_t t0, int32_t count, const double **qz, const double **qy)
{
    ARB_REQUIRE(ctx && ctx->fit_trees > 0, ARB_ERR_INVALID, "no fit to read from");
    ARB_REQUIRE(t0 >= 0 && count >= 1 && (int64_t)t0 + count <= ctx->fit_trees, ARB_ERR_INVALID,
                "trees %d..%d outside 0..%d", t0, t0 + count - 1, ctx->fit_trees - 1);
    ARB_CUDA(cudaSetDevice(ctx->device));
    const int K = ctx->fit_k;
    const int64_t nz = (int64_t)ctx->n_cells * K, ny = (int64_t)ctx->n_snvs * (K - 1);
    if (ctx->fit_kept) {
        *qz = ctx->best_qz.as<double>() + (int64_t)t0 * nz;
        *qy = ctx->best_qy.as<double>() + (int64_t)t0 * ny;
        return ARB_OK;
    }
    if (!(ctx->batch_t0 >= 0 && t0 >= ctx->batch_t0 && t0 + count <= ctx->batch_t0 + ctx->batch_n)) {
        ctx->batch_t0 = -1;
        ARB_TRY(ctx->batch_qz.reserve((size_t)count * nz * 8));
        ARB_TRY(ctx->batch_qy.reserve((size_t)count * ny * 8));
        ARB_TRY(ctx->batch_out.reserve((size_t)count * 12 + 16));
        FitTarget tg;
        memset(&tg, 0, sizeof tg);
        tg.masks = ctx->tree_masks.as<uint32_t>() + (int64_t)t0 * MASK_STRIDE;
        tg.elbo = ctx->batch_out.as<double>();
        tg.iters = reinterpret_cast<int32_t *>(ctx->batch_out.as<double>() + count);
        tg.best_qz = ctx->batch_qz.as<double>();
        tg.best_qy = ctx->batch_qy.as<double>();
        const FitStats keep_stats = ctx->stats;
        int rc = fit_dispatch(ctx, count, K, ctx->fit_max_iter, ctx->fit_tol, ctx->fit_mode, tg, 0);
        ctx->stats = keep_stats;
        ctx->stats.refit_batches += 1;
        ARB_TRY(rc);
        ctx->batch_t0 = t0;
        ctx->batch_n = count;
    }
    *qz = ctx->batch_qz.as<double>() + (int64_t)(t0 - ctx->batch_t0) * nz;
    *qy = ctx->batch_qy.as<double>() + (int64_t)(t0 - ctx->batch_t0) * ny;
    return ARB_OK;
}

ARB_API int arb_get_fits(arb_ctx *ctx, int32_t t0, int32_t count, double *q_z, double *q_y)
{
    const double *dz, *dy;
    ARB_TRY(locate_fits(ctx, t0, count, &dz, &dy));
    const int K = ctx->fit_k;
    if (q_z)
        ARB_CUDA(cudaMemcpyAsync(q_z, dz, (size_t)count * ctx->n_cells * K * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (q_y)
        ARB_CUDA(cudaMemcpyAsync(q_y, dy, (size_t)count * ctx->n_snvs * (K - 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    return ARB_OK;
}

ARB_API int arb_get_fit(arb_ctx *ctx, int32_t t, double *q_z, double *q_y)
{
    return arb_get_fits(ctx, t, 1, q_z, q_y);
}

ARB_API int arb_get_map(arb_ctx *ctx, int32_t t, int32_t *z, int32_t *y)
{
    const double *dz, *dy;
    ARB_TRY(locate_fits(ctx, t, 1, &dz, &dy));
    const int K = ctx->fit_k;
    DevBuf out;
    ARB_TRY(out.reserve(((size_t)ctx->n_cells + ctx->n_snvs) * 4));
    int32_t *oz = out.as<int32_t>(), *oy = oz + ctx->n_cells;
    k_argmax<<<grid_for(ctx->n_cells, 256), 256, 0, ctx->stream>>>(dz, ctx->n_cells, K, oz);
    k_argmax<<<grid_for(ctx->n_snvs, 256), 256, 0, ctx->stream>>>(dy, ctx->n_snvs, K - 1, oy);
    if (z) ARB_CUDA(cudaMemcpyAsync(z, oz, (size_t)ctx->n_cells * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (y) ARB_CUDA(cudaMemcpyAsync(y, oy, (size_t)ctx->n_snvs * 4, cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    out.release();
    return ARB_OK;
}

ARB_API int arb_get_presence(arb_ctx *ctx, int32_t t, uint8_t *presence)
{
    ARB_REQUIRE(ctx && ctx->fit_trees > 0 && presence, ARB_ERR_INVALID, "arb_get_presence: no fit");
    ARB_REQUIRE(t >= 0 && t < ctx->fit_trees, ARB_ERR_INVALID, "arb_get_presence: tree %d", t);
    ARB_CUDA(cudaSetDevice(ctx->device));
    uint32_t mk[MASK_STRIDE];
    ARB_CUDA(cudaMemcpy(mk, ctx->tree_masks.as<uint32_t>() + (int64_t)t * MASK_STRIDE, sizeof mk,
                        cudaMemcpyDeviceToHost));
    const int K = ctx->fit_k, C = K - 1;
    for (int p = 0; p < C; ++p)
        for (int r = 0; r < K; ++r) presence[p * K + r] = (uint8_t)((mk[32 + p] >> r) & 1u);
    return ARB_OK;
}

ARB_API int arb_last_fit_stats(arb_ctx *ctx, int64_t *kernel_launches, double *device_ms, double *tree_iterations,
                               double *sweep_ms)
{
    ARB_REQUIRE(ctx, ARB_ERR_INVALID, "arb_last_fit_stats: null context");
    if (kernel_launches) *kernel_launches = ctx->stats.launches;
    if (device_ms) *device_ms = ctx->stats.device_ms;
    if (tree_iterations) *tree_iterations = ctx->stats.tree_iterations;
    if (sweep_ms) *sweep_ms = ctx->stats.sweep_ms;
    return ARB_OK;
}

// [sweep_z_ms, sweep_y_ms, finish_ms, tree_sweeps_z, tree_sweeps_y, launches_z, launches_y, rounds,
//  finish_z_ms, finish_y_ms, KP, panel_rows, panels_z, panels_y, tiles_per_warp_z, tiles_per_warp_y,
//  fold_z, fold_y, quads_z, quads_y, rounds enqueued past the end, 16-bit stream words, blocking host waits,
//  batched re-fits done by arb_get_fits since the last fit]
ARB_API int arb_last_fit_profile(arb_ctx *ctx, double *values, int32_t n_values)
{
    ARB_REQUIRE(ctx && values && n_values >= 0, ARB_ERR_INVALID, "arb_last_fit_profile: bad argument");
    const FitStats &s = ctx->stats;
    PanelLayout none;
    const PanelLayout &L = ctx->ws_kp ? ctx->layout[kp_index(ctx->ws_kp)] : none;
    const double v[24] = {s.sweep_z_ms, s.sweep_y_ms, s.finish_ms, s.tree_sweeps_z, s.tree_sweeps_y,
                          s.launches_z, s.launches_y, s.rounds, s.finish_z_ms, s.finish_y_ms,
                          (double)L.kp, (double)L.panel_rows, (double)L.np_z, (double)L.np_y, (double)L.uz,
                          (double)L.uy, (double)L.fold_z, (double)L.fold_y, (double)L.quads_z, (double)L.quads_y,
                          s.idle_rounds, (double)L.w16, s.host_waits, s.refit_batches};
    for (int i = 0; i < n_values && i < 24; ++i) values[i] = v[i];
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// prior
// ---------------------------------------------------------------------------------------
static int finish_prior(arb_ctx *ctx, int32_t C)
{
    const int64_t n = (int64_t)ctx->n_snvs * C;
    ARB_TRY(ctx->logq0.reserve(n * 8));
    ARB_TRY(ctx->logq0e.reserve(n * 8));
    ARB_TRY(ctx->q0_hard.reserve(n * 8));
    k_prior_logs<<<grid_for(n, 256), 256, 0, ctx->stream>>>(ctx->q0.as<double>(), n, ctx->logq0.as<double>(),
                                                           ctx->logq0e.as<double>());
    k_prior_hard<<<grid_for(ctx->n_snvs, 256), 256, 0, ctx->stream>>>(ctx->q0.as<double>(), ctx->n_snvs, C,
                                                                     ctx->q0_hard.as<double>());
    // rows that do not sum to one (K = 2: [gamma]; a caller's own q_y_init) send the first pass of every tree
    // through k_first_pass_general; 1e-9 is far above the rounding of gamma + (k-1) * (1-gamma)/(k-1)
    DevBuf dev;
    ARB_TRY(dev.reserve(8));
    ARB_CUDA(cudaMemsetAsync(dev.p, 0, 8, ctx->stream));
    k_row_mass_dev<<<grid_for(ctx->n_snvs, 256), 256, 0, ctx->stream>>>(ctx->q0.as<double>(), ctx->n_snvs, C,
                                                                       dev.as<unsigned long long>());
    double worst = 0.0;
    ARB_CUDA(cudaMemcpyAsync(&worst, dev.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    dev.release();
    ctx->prior_general = worst > 1e-9;
    ctx->n_clusters = C;
    ctx->batch_t0 = -1;
    return ARB_OK;
}

ARB_API int arb_set_prior_labels(arb_ctx *ctx, const int32_t *label_idx, int32_t n_clusters, double gamma)
{
    ARB_REQUIRE(ctx && ctx->ingested && label_idx, ARB_ERR_INVALID, "arb_set_prior_labels: bad argument");
    ARB_REQUIRE(n_clusters >= 1 && n_clusters < ARB_MAX_CLONES, ARB_ERR_INVALID, "arb_set_prior_labels: %d clusters",
                n_clusters);
    for (int32_t j = 0; j < ctx->n_snvs; ++j)
        ARB_REQUIRE(label_idx[j] >= -1 && label_idx[j] < n_clusters, ARB_ERR_INVALID,
                    "arb_set_prior_labels: label_idx[%d]=%d", j, label_idx[j]);
    ARB_CUDA(cudaSetDevice(ctx->device));
    const int64_t n = (int64_t)ctx->n_snvs * n_clusters;
    DevBuf lab;
    ARB_TRY(lab.reserve((size_t)ctx->n_snvs * 4));
    ARB_TRY(ctx->q0.reserve(n * 8));
    ARB_CUDA(cudaMemcpyAsync(lab.p, label_idx, (size_t)ctx->n_snvs * 4, cudaMemcpyHostToDevice, ctx->stream));
    // the same double arithmetic as A.py:552-553
    const double other = n_clusters > 1 ? (1 - gamma) / (n_clusters - 1) : 0.0, uniform = 1.0 / n_clusters;
    k_prior_from_labels<<<grid_for(n, 256), 256, 0, ctx->stream>>>(lab.as<int32_t>(), ctx->n_snvs, n_clusters, gamma,
                                                                  other, uniform, ctx->q0.as<double>());
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    lab.release();
    return finish_prior(ctx, n_clusters);
}

ARB_API int arb_set_prior(arb_ctx *ctx, const double *q_y_init, int32_t n_clusters)
{
    ARB_REQUIRE(ctx && ctx->ingested && q_y_init, ARB_ERR_INVALID, "arb_set_prior: bad argument");
    ARB_REQUIRE(n_clusters >= 1 && n_clusters < ARB_MAX_CLONES, ARB_ERR_INVALID, "arb_set_prior: %d clusters",
                n_clusters);
    ARB_CUDA(cudaSetDevice(ctx->device));
    const int64_t n = (int64_t)ctx->n_snvs * n_clusters;
    ARB_TRY(ctx->q0.reserve(n * 8));
    ARB_CUDA(cudaMemcpyAsync(ctx->q0.p, q_y_init, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    return finish_prior(ctx, n_clusters);
}

ARB_API int arb_get_prior(arb_ctx *ctx, double *q_y_init)
{
    ARB_REQUIRE(ctx && ctx->n_clusters > 0 && q_y_init, ARB_ERR_INVALID, "arb_get_prior: no prior set");
    ARB_CUDA(cudaSetDevice(ctx->device));
    ARB_CUDA(cudaMemcpy(q_y_init, ctx->q0.p, (size_t)ctx->n_snvs * ctx->n_clusters * 8, cudaMemcpyDeviceToHost));
    return ARB_OK;
}

// ---------------------------------------------------------------------------------------
// kernel-level hooks: one tree in slot 0 with explicit posteriors
// ---------------------------------------------------------------------------------------
struct HookEnv {
    Workspace ws;
    WsPtrs w;
    const PanelLayout *L;
    int32_t *list0;
};

static int hook_setup(arb_ctx *ctx, int K, const uint8_t *presence, HookEnv &env, const char *who)
{
    ARB_REQUIRE(ctx && ctx->ingested && presence, ARB_ERR_INVALID, "%s: bad argument", who);
    ARB_REQUIRE(K >= 2 && K <= ARB_MAX_CLONES, ARB_ERR_INVALID, "%s: n_clones=%d", who, K);
    ARB_TRY(arb_check_table_range(ctx));
    ARB_CUDA(cudaSetDevice(ctx->device));
    ARB_TRY(ensure_workspace(ctx, K, 1, 0, env.ws));
    env.L = &ctx->layout[kp_index(env.ws.kp)];
    env.w = ws_ptrs(ctx, *env.L, ctx->stream);
    ARB_TRY(configure_kernels(ctx, *env.L));
    const int C = K - 1;
    DevBuf d_p;
    ARB_TRY(d_p.reserve((size_t)C * K));
    ARB_TRY(ctx->masks.reserve(MASK_STRIDE * 4));
    ARB_CUDA(cudaMemcpyAsync(d_p.p, presence, (size_t)C * K, cudaMemcpyHostToDevice, ctx->stream));
    k_masks_from_presence<<<1, 32, 0, ctx->stream>>>(d_p.as<uint8_t>(), 1, K, ctx->masks.as<uint32_t>());
    SlotState s0;
    memset(&s0, 0, sizeof s0);
    s0.tree = 0;
    s0.phase = PH_INIT;
    s0.cur = 0;                         // explicit posteriors are read from / written to plane 0
    s0.bestb = 1;
    ARB_CUDA(cudaMemcpyAsync(env.w.st, &s0, sizeof s0, cudaMemcpyHostToDevice, ctx->stream));
    int32_t zero = 0;                   // the one tree sits in slot 0 at packed position 0
    ARB_CUDA(cudaMemcpyAsync(env.w.list_z, &zero, 4, cudaMemcpyHostToDevice, ctx->stream));
    ARB_CUDA(cudaMemcpyAsync(env.w.pos_z, &zero, 4, cudaMemcpyHostToDevice, ctx->stream));
    env.list0 = env.w.list_z;
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    d_p.release();
    ctx->batch_t0 = -1;
    return ARB_OK;
}

#define GL_SWITCH(K, ...)                                  \
    switch (gl_for(K)) {                                   \
    case 1: { constexpr int GLV = 1; __VA_ARGS__; } break;  \
    case 2: { constexpr int GLV = 2; __VA_ARGS__; } break;  \
    case 3: { constexpr int GLV = 3; __VA_ARGS__; } break;  \
    case 4: { constexpr int GLV = 4; __VA_ARGS__; } break;  \
    case 5: { constexpr int GLV = 5; __VA_ARGS__; } break;  \
    case 6: { constexpr int GLV = 6; __VA_ARGS__; } break;  \
    case 7: { constexpr int GLV = 7; __VA_ARGS__; } break;  \
    default: { constexpr int GLV = 8; __VA_ARGS__; } break; \
    }

ARB_API int arb_qz_update(arb_ctx *ctx, int32_t n_clones, const uint8_t *presence, const double *q_y, double *q_z)
{
    HookEnv env;
    ARB_TRY(hook_setup(ctx, n_clones, presence, env, "arb_qz_update"));
    ARB_REQUIRE(q_y && q_z, ARB_ERR_INVALID, "arb_qz_update: null argument");
    const int K = n_clones, C = K - 1;
    const uint32_t *mk = ctx->masks.as<uint32_t>();
    ARB_CUDA(cudaMemcpyAsync(ctx->qy_cur.p, q_y, (size_t)ctx->n_snvs * C * 8, cudaMemcpyHostToDevice, ctx->stream));
    GL_SWITCH(K, {
        launch_seed_s<GLV>(ctx, env.w, mk, ctx->qy_cur.as<double>(), (int64_t)ctx->n_snvs * C, env.list0, 1, K, 0, nullptr);
        launch_sweep_z(ctx, *env.L, env.w, 1, nullptr);
        launch_finish_z<GLV>(ctx, *env.L, env.w, env.ws, mk, env.list0, 1, 0, 0, nullptr);
    });
    ARB_CUDA(cudaMemcpyAsync(q_z, ctx->qz_cur.p, (size_t)ctx->n_cells * K * 8, cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    ARB_CUDA(cudaGetLastError());
    return ARB_OK;
}

ARB_API int arb_qy_update(arb_ctx *ctx, int32_t n_clones, const uint8_t *presence, const double *q_z, double *q_y)
{
    HookEnv env;
    ARB_TRY(hook_setup(ctx, n_clones, presence, env, "arb_qy_update"));
    ARB_REQUIRE(q_z && q_y, ARB_ERR_INVALID, "arb_qy_update: null argument");
    ARB_REQUIRE(ctx->n_clusters == n_clones - 1, ARB_ERR_INVALID, "arb_qy_update: prior has %d clusters",
                ctx->n_clusters);
    const int K = n_clones, C = K - 1;
    const uint32_t *mk = ctx->masks.as<uint32_t>();
    ARB_CUDA(cudaMemcpyAsync(ctx->qz_cur.p, q_z, (size_t)ctx->n_cells * K * 8, cudaMemcpyHostToDevice, ctx->stream));
    GL_SWITCH(K, {
        launch_seed_r<GLV>(ctx, env.w, mk, env.list0, 1, K, nullptr);
        launch_sweep_y(ctx, *env.L, env.w, 1, nullptr);
        launch_finish_y<GLV>(ctx, *env.L, env.w, env.ws, mk, env.list0, 1, nullptr);
    });
    ARB_CUDA(cudaMemcpyAsync(q_y, ctx->qy_cur.p, (size_t)ctx->n_snvs * C * 8, cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    ARB_CUDA(cudaGetLastError());
    return ARB_OK;
}

ARB_API int arb_elbo_terms(arb_ctx *ctx, int32_t n_clones, const uint8_t *presence, const double *q_z,
                           const double *q_y, double *likelihood, double *entropy, double *kl)
{
    HookEnv env;
    ARB_TRY(hook_setup(ctx, n_clones, presence, env, "arb_elbo_terms"));
    ARB_REQUIRE(q_z && q_y, ARB_ERR_INVALID, "arb_elbo_terms: null argument");
    ARB_REQUIRE(ctx->n_clusters == n_clones - 1, ARB_ERR_INVALID, "arb_elbo_terms: prior has %d clusters",
                ctx->n_clusters);
    const int K = n_clones, C = K - 1;
    const int64_t nz = (int64_t)ctx->n_cells * K, ny = (int64_t)ctx->n_snvs * C;
    const uint32_t *mk = ctx->masks.as<uint32_t>();
    ARB_CUDA(cudaMemcpyAsync(ctx->qy_cur.p, q_y, ny * 8, cudaMemcpyHostToDevice, ctx->stream));
    ARB_CUDA(cudaMemcpyAsync(ctx->qz_cur.p, q_z, nz * 8, cudaMemcpyHostToDevice, ctx->stream));
    GL_SWITCH(K, {
        launch_seed_s<GLV>(ctx, env.w, mk, ctx->qy_cur.as<double>(), ny, env.list0, 1, K, 0, nullptr);
        launch_sweep_z(ctx, *env.L, env.w, 1, nullptr);
        launch_finish_z<GLV>(ctx, *env.L, env.w, env.ws, mk, env.list0, 1, 1, 0, nullptr);
    });
    DevBuf out;
    ARB_TRY(out.reserve(16));
    k_entropy_kl<<<1, 1024, 0, ctx->stream>>>(ctx->qz_cur.as<double>(), nz, ctx->qy_cur.as<double>(),
                                              ctx->q0.as<double>(), ny, out.as<double>());
    std::vector<double> red((size_t)env.ws.nbz);
    double ek[2];
    ARB_CUDA(cudaMemcpyAsync(red.data(), ctx->red.p, red.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaMemcpyAsync(ek, out.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    ARB_CUDA(cudaStreamSynchronize(ctx->stream));
    ARB_CUDA(cudaGetLastError());
    out.release();
    double lik = ctx->sum_la;
    for (double v : red) lik += v;
    if (likelihood) *likelihood = lik;
    if (entropy) *entropy = ek[0];
    if (kl) *kl = ek[1];
    return ARB_OK;
}
