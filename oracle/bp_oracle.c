/*
 * TEST INFRASTRUCTURE ONLY.  CPU oracle for the binary soft BP the reference calls at
 * simulate/hqc.py:694-708 and simulate/decode.py:155-171.
 *
 * PARITY UNPINNED: the arithmetic lives in a third-party dependency that is absent from the reference
 * tree -- PyPI package `ldpc`, pinned `ldpc==0.1.3` (simulate-with-python/requirements.txt:8), class
 * `ldpc.bp_decoder` with bp_method="product_sum".  The package is not installed and cannot be installed
 * here.  This file restates its published algorithm (flooding schedule, fp64 probability *ratios*
 * p1/p0, forward/backward running products, syndrome-match early stop) as summarised in SURVEY.md
 * appendix B.  The reference's own tests at this boundary pin decoding OUTCOMES only
 * (decode.py:139-149 -> 100/100 on rep_code(13); hqc.py:1229-1236 toy -> True; hqc.py:1277-1284 full
 * N=17669 -> True); tests/test_oracle_golden.py checks those.  No message or posterior values of the
 * upstream implementation are available, so numerical parity of this path is "unpinned".
 *
 * The product (sca_ldpc_b200/) never imports, links or executes anything in this directory.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    int m, n, E;
    int *row_ptr;  /* m+1; edges of a row in ascending column */
    int *col;      /* E */
    int *col_ptr;  /* n+1 */
    int *csc_edge; /* E; edges of a column in ascending row */
} bp_graph;

void bp_graph_free(bp_graph *g)
{
    if (!g) return;
    free(g->row_ptr);
    free(g->col);
    free(g->col_ptr);
    free(g->csc_edge);
    free(g);
}

/* Build from CSR (row_ptr, col); entries of a row must be ascending in column. */
int bp_graph_build(int m, int n, const int *row_ptr, const int *col, bp_graph **out)
{
    bp_graph *g = (bp_graph *)calloc(1, sizeof(*g));
    int E = row_ptr[m];
    g->m = m;
    g->n = n;
    g->E = E;
    g->row_ptr = (int *)malloc(sizeof(int) * (m + 1));
    g->col = (int *)malloc(sizeof(int) * (E ? E : 1));
    g->col_ptr = (int *)calloc(n + 1, sizeof(int));
    g->csc_edge = (int *)malloc(sizeof(int) * (E ? E : 1));
    memcpy(g->row_ptr, row_ptr, sizeof(int) * (m + 1));
    memcpy(g->col, col, sizeof(int) * E);
    for (int e = 0; e < E; ++e) {
        if (col[e] < 0 || col[e] >= n) {
            bp_graph_free(g);
            return -1;
        }
        g->col_ptr[col[e] + 1]++;
    }
    for (int c = 0; c < n; ++c) g->col_ptr[c + 1] += g->col_ptr[c];
    int *fill = (int *)calloc(n, sizeof(int));
    for (int r = 0; r < m; ++r)
        for (int p = row_ptr[r]; p < row_ptr[r + 1]; ++p) g->csc_edge[g->col_ptr[col[p]] + fill[col[p]]++] = p;
    free(fill);
    *out = g;
    return 0;
}

/* One decode.  synd: m bytes (0/1).  channel_probs: n doubles.  max_iter == 0 => n (upstream default).
 * Outputs: decoding (n bytes, the estimated ERROR pattern), log_prob_ratios (n doubles, log(p0/p1)),
 * *iters (iterations run), return value 1 = converged (syndrome matched), 0 = not. */
int bp_product_sum_decode(const bp_graph *g, const uint8_t *synd, const double *channel_probs, int max_iter,
                          uint8_t *decoding, double *log_prob_ratios, int *iters)
{
    const int m = g->m, n = g->n, E = g->E;
    double *b2c = (double *)malloc(sizeof(double) * (E ? E : 1));
    double *c2b = (double *)malloc(sizeof(double) * (E ? E : 1));
    if (max_iter == 0) max_iter = n;
    for (int j = 0; j < n; ++j) {
        double r = channel_probs[j] / (1 - channel_probs[j]);
        for (int p = g->col_ptr[j]; p < g->col_ptr[j + 1]; ++p) b2c[g->csc_edge[p]] = r;
    }
    int converged = 0, it = 0;
    for (it = 1; it <= max_iter; ++it) {
        /* check-to-bit: forward/backward running products of tanh(L/2) = 2/(1+ratio) - 1 */
        for (int i = 0; i < m; ++i) {
            double temp = synd[i] ? -1.0 : 1.0;
            for (int e = g->row_ptr[i]; e < g->row_ptr[i + 1]; ++e) {
                c2b[e] = temp;
                temp *= 2 / (1 + b2c[e]) - 1;
            }
            temp = 1.0;
            for (int e = g->row_ptr[i + 1] - 1; e >= g->row_ptr[i]; --e) {
                c2b[e] *= temp;
                c2b[e] = (1 - c2b[e]) / (1 + c2b[e]);
                temp *= 2 / (1 + b2c[e]) - 1;
            }
        }
        /* bit-to-check */
        for (int j = 0; j < n; ++j) {
            double temp = channel_probs[j] / (1 - channel_probs[j]);
            for (int p = g->col_ptr[j]; p < g->col_ptr[j + 1]; ++p) {
                int e = g->csc_edge[p];
                b2c[e] = temp;
                temp *= c2b[e];
                if (temp != temp) temp = 1.0;
            }
            log_prob_ratios[j] = log(1 / temp);
            decoding[j] = temp >= 1;
            temp = 1.0;
            for (int p = g->col_ptr[j + 1] - 1; p >= g->col_ptr[j]; --p) {
                int e = g->csc_edge[p];
                b2c[e] *= temp;
                temp *= c2b[e];
                if (temp != temp) temp = 1.0;
            }
        }
        int equal = 1;
        for (int i = 0; i < m && equal; ++i) {
            int s = 0;
            for (int e = g->row_ptr[i]; e < g->row_ptr[i + 1]; ++e) s ^= decoding[g->col[e]];
            if (s != synd[i]) equal = 0;
        }
        if (equal) {
            converged = 1;
            break;
        }
    }
    if (it > max_iter) it = max_iter;
    *iters = it;
    free(b2c);
    free(c2b);
    return converged;
}

/* Batch: T independent syndromes on one graph; channel_probs either shared (probs_stride == 0) or per
 * trial (probs_stride == n).  One decode per thread. */
void bp_product_sum_decode_batch(const bp_graph *g, int T, const uint8_t *synd, const double *channel_probs,
                                 int probs_stride, int max_iter, uint8_t *decoding, double *log_prob_ratios,
                                 int *iters, int *converged, int threads)
{
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (int t = 0; t < T; ++t) {
        converged[t] = bp_product_sum_decode(g, synd + (size_t)t * g->m, channel_probs + (size_t)t * probs_stride,
                                             max_iter, decoding + (size_t)t * g->n,
                                             log_prob_ratios + (size_t)t * g->n, iters + t);
    }
}
