/*
 * arborist_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the reference's per-tree variational fit
 * (/root/reference/src/arborist/arborist.py, "A.py" below).  It exists so the
 * CUDA path has something to be checked against on a machine where the
 * reference itself (Python + numba) is not present.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library; the product (arborist_b200/) never does.
 *
 * Parity status: PINNED.  oracle/gen_golden.py runs the unmodified reference
 * in the build container and stores its outputs under tests/golden/;
 * tests/test_oracle_golden.py checks every function below against them.
 *
 * The arithmetic deliberately follows the reference term by term (log-sum-exp
 * over the other latent variable, one exp per (nonzero, clone, cluster)), so
 * it is as slow as the reference.  OpenMP "parallel for" stands in for numba's
 * prange (A.py:231, A.py:358, A.py:410); like the reference, the likelihood's
 * scalar reduction rounds differently for different thread counts.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

ORC_API int orc_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

ORC_API void orc_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* A.py:173-197 logsumexp_inline */
static inline double lse(const double *t, int n)
{
    double mx = -INFINITY;
    for (int i = 0; i < n; ++i)
        if (t[i] > mx) mx = t[i];
    double s = 0.0;
    for (int i = 0; i < n; ++i) s += exp(t[i] - mx);
    return mx + log(s);
}

/* A.py:151-170 soft_max (row-wise, in place) */
static void soft_max_rows(double *a, long rows, int cols)
{
    for (long j = 0; j < rows; ++j) {
        double *r = a + j * cols;
        double mx = r[0];
        for (int c = 1; c < cols; ++c)
            if (r[c] > mx) mx = r[c];
        double s = 0.0;
        for (int c = 0; c < cols; ++c) {
            r[c] = exp(r[c] - mx);
            s += r[c];
        }
        for (int c = 0; c < cols; ++c) r[c] /= s;
    }
}

/* A.py:302-322 log_array */
ORC_API void orc_log_array(const double *a, long count, double *out)
{
    for (long i = 0; i < count; ++i) out[i] = a[i] > 0.0 ? log(a[i]) : -INFINITY;
}

/* A.py:200-250 compute_q_z_sparse.  presence is int64 [(K-1), K] row-major. */
ORC_API void orc_q_z(const int32_t *cell_ptr, const int32_t *snv_idx, const double *log_likes,
                     const double *log_q_y, const int64_t *presence, int n_cells, int n_clones,
                     double *q_z)
{
    const int K = n_clones, C = n_clones - 1;
#pragma omp parallel
    {
        double *terms = (double *)malloc(sizeof(double) * (size_t)(C > 0 ? C : 1));
#pragma omp for schedule(static)
        for (int i = 0; i < n_cells; ++i) {
            for (int r = 0; r < K; ++r) {
                double acc = 0.0;
                for (int k = cell_ptr[i]; k < cell_ptr[i + 1]; ++k) {
                    const double lp = log_likes[2 * (long)k], la = log_likes[2 * (long)k + 1];
                    const long j = snv_idx[k];
                    for (int p = 0; p < C; ++p) {
                        const double pr = (double)presence[p * K + r];
                        /* A.py:241-243: lp*P + la*(1-P) */
                        terms[p] = log_q_y[j * C + p] + (lp * pr + la * (1.0 - pr));
                    }
                    acc += lse(terms, C);
                }
                q_z[(long)i * K + r] = acc;
            }
        }
        free(terms);
    }
    soft_max_rows(q_z, n_cells, K);
}

/* A.py:325-375 compute_q_y_sparse */
ORC_API void orc_q_y(const int32_t *snv_ptr, const int32_t *cell_idx, const double *log_likes,
                     const double *log_q_z, const int64_t *presence, int n_snvs, int n_clones,
                     const double *log_q_y_init, double *q_y)
{
    const int K = n_clones, C = n_clones - 1;
#pragma omp parallel
    {
        double *terms = (double *)malloc(sizeof(double) * (size_t)K);
#pragma omp for schedule(static)
        for (int j = 0; j < n_snvs; ++j) {
            for (int p = 0; p < C; ++p) {
                double acc = 0.0;
                for (int k = snv_ptr[j]; k < snv_ptr[j + 1]; ++k) {
                    const double lp = log_likes[2 * (long)k], la = log_likes[2 * (long)k + 1];
                    const long i = cell_idx[k];
                    for (int r = 0; r < K; ++r)
                        terms[r] = log_q_z[i * K + r] + (presence[p * K + r] ? lp : la);
                    acc += lse(terms, K);
                }
                q_y[(long)j * C + p] = acc + log_q_y_init[(long)j * C + p];
            }
        }
        free(terms);
    }
    soft_max_rows(q_y, n_snvs, C);
}

/* A.py:378-420 compute_likelihood_sparse */
ORC_API double orc_likelihood(const int32_t *cell_ptr, const int32_t *snv_idx,
                              const double *log_likes, const double *log_q_y,
                              const double *log_q_z, const int64_t *presence, int n_clones,
                              int n_cells)
{
    const int K = n_clones, C = n_clones - 1;
    double total = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : total)
    for (int i = 0; i < n_cells; ++i) {
        for (int k = cell_ptr[i]; k < cell_ptr[i + 1]; ++k) {
            const double lp = log_likes[2 * (long)k], la = log_likes[2 * (long)k + 1];
            const long j = snv_idx[k];
            for (int r = 0; r < K; ++r)
                for (int p = 0; p < C; ++p) {
                    const double lw = log_q_z[(long)i * K + r] + log_q_y[j * C + p];
                    total += exp(lw) * (presence[p * K + r] ? lp : la);
                }
        }
    }
    return total;
}

/* A.py:253-273 compute_entropy */
ORC_API double orc_entropy(const double *q, long rows, int cols)
{
    double ent = 0.0;
    for (long i = 0; i < rows * cols; ++i)
        if (q[i] > 0) ent -= q[i] * log(q[i] + 1e-12);
    return ent;
}

/* A.py:276-299 compute_kl_divergence */
ORC_API double orc_kl(const double *q, const double *q0, long rows, int cols)
{
    double kl = 0.0;
    for (long i = 0; i < rows * cols; ++i) kl += q[i] * (log(q[i] + 1e-12) - log(q0[i] + 1e-12));
    return kl;
}

/*
 * A.py:679-816 run_variational_inference.
 * Returns the best ELBO; best_q_z [n,K], best_q_y [m,K-1] are caller-allocated.
 * Extra outputs the reference does not expose (used by the tests to pin the
 * control flow): n_iters = number of loop bodies executed, elbo_trace
 * [max_iter+1] = ELBO_0 followed by each iteration's ELBO (may be NULL).
 */
ORC_API double orc_run_vi(const int64_t *presence, const double *ll_cell_sort,
                          const double *ll_snv_sort, const int32_t *cell_idx,
                          const int32_t *snv_idx, int n_cells, int n_snvs, int n_clones,
                          const double *q_y_init, const int32_t *cell_ptr,
                          const int32_t *snv_ptr, int max_iter, double tolerance,
                          double *best_q_z, double *best_q_y, int *n_iters, double *elbo_trace)
{
    const int K = n_clones, C = n_clones - 1;
    const long nz = (long)n_cells * K, ny = (long)n_snvs * C;
    double *log_q_y_init = (double *)malloc(sizeof(double) * (size_t)ny);
    double *log_q_y = (double *)malloc(sizeof(double) * (size_t)ny);
    double *q_y = (double *)malloc(sizeof(double) * (size_t)ny);
    double *q_z = (double *)malloc(sizeof(double) * (size_t)nz);
    double *log_q_z = (double *)malloc(sizeof(double) * (size_t)nz);

    orc_log_array(q_y_init, ny, log_q_y_init);                       /* A.py:733 */
    memcpy(log_q_y, log_q_y_init, sizeof(double) * (size_t)ny);      /* A.py:734 */
    orc_q_z(cell_ptr, snv_idx, ll_cell_sort, log_q_y, presence, n_cells, K, q_z);
    orc_log_array(q_z, nz, log_q_z);
    double best = orc_likelihood(cell_ptr, snv_idx, ll_cell_sort, log_q_y, log_q_z, presence, K,
                                 n_cells) +
                  orc_entropy(q_z, n_cells, K) - orc_kl(q_y_init, q_y_init, n_snvs, C);
    memcpy(best_q_z, q_z, sizeof(double) * (size_t)nz);              /* A.py:762 */
    memcpy(best_q_y, q_y_init, sizeof(double) * (size_t)ny);         /* A.py:763 */
    if (elbo_trace) elbo_trace[0] = best;

    int converged = 0, it = 0;
    for (it = 0; it < max_iter; ++it) {                              /* A.py:767 */
        orc_q_y(snv_ptr, cell_idx, ll_snv_sort, log_q_z, presence, n_snvs, K, log_q_y_init, q_y);
        orc_log_array(q_y, ny, log_q_y);
        orc_q_z(cell_ptr, snv_idx, ll_cell_sort, log_q_y, presence, n_cells, K, q_z);
        orc_log_array(q_z, nz, log_q_z);
        const double lik = orc_likelihood(cell_ptr, snv_idx, ll_cell_sort, log_q_y, log_q_z,
                                          presence, K, n_cells);
        const double kl = orc_kl(q_y, q_y_init, n_snvs, C);
        const double ent = orc_entropy(q_z, n_cells, K);
        const double elbo = lik + ent - kl;                          /* A.py:803 */
        if (elbo_trace) elbo_trace[it + 1] = elbo;
        if (fabs(elbo - best) < tolerance) converged = 1;            /* A.py:805 (vs BEST) */
        if (elbo > best) {                                           /* A.py:808 */
            memcpy(best_q_z, q_z, sizeof(double) * (size_t)nz);
            memcpy(best_q_y, q_y, sizeof(double) * (size_t)ny);
            best = elbo;
        }
        if (converged) {                                             /* A.py:813 */
            ++it;
            break;
        }
    }
    if (n_iters) *n_iters = it;
    free(log_q_y_init);
    free(log_q_y);
    free(q_y);
    free(q_z);
    free(log_q_z);
    return best;
}

/*
 * A.py:572-676 run_simple_max_likelihood (the hard-assignment mode; unreachable
 * from the reference CLI, kept for the SURVEY section 8(f) "next" row).
 * q_z_out [n,K] is the SOFT q_z the reference returns, q_y_hard [m,K-1].
 */
static void hard_assign(const double *q, long rows, int cols, double *out)
{
    for (long i = 0; i < rows; ++i) {
        int best = 0;
        for (int c = 1; c < cols; ++c)
            if (q[i * cols + c] > q[i * cols + best]) best = c; /* first max wins */
        for (int c = 0; c < cols; ++c) out[i * cols + c] = (c == best) ? 1.0 : 0.0;
    }
}

ORC_API double orc_run_simple_ml(const int64_t *presence, const double *ll_cell_sort,
                                 const int32_t *snv_idx, int n_cells, int n_snvs, int n_clones,
                                 const double *q_y_init, const int32_t *cell_ptr, double *q_z_out,
                                 double *q_y_hard)
{
    const int K = n_clones, C = n_clones - 1;
    const long nz = (long)n_cells * K, ny = (long)n_snvs * C;
    double *lqy = (double *)malloc(sizeof(double) * (size_t)ny);
    double *qzh = (double *)malloc(sizeof(double) * (size_t)nz);
    double *lqz = (double *)malloc(sizeof(double) * (size_t)nz);
    hard_assign(q_y_init, n_snvs, C, q_y_hard);
    orc_log_array(q_y_hard, ny, lqy);
    orc_q_z(cell_ptr, snv_idx, ll_cell_sort, lqy, presence, n_cells, K, q_z_out);
    hard_assign(q_z_out, n_cells, K, qzh);
    orc_log_array(qzh, nz, lqz);
    const double lik =
        orc_likelihood(cell_ptr, snv_idx, ll_cell_sort, lqy, lqz, presence, K, n_cells);
    free(lqy);
    free(qzh);
    free(lqz);
    return lik;
}
