/*
 * TEST INFRASTRUCTURE ONLY.  CPU oracle for the q-ary min-sum decoder of atneit/SCA-LDPC.
 *
 * This file restates, in plain C, the algorithm of the reference's Rust decoder core
 * (simulate-with-python/simulate_rs/src/decoder.rs and decoder_special.rs; see qary_impl.inc for the
 * per-function file:line citations).  It exists so that tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py can check and time the reference's algorithm on the
 * host.  The product (sca_ldpc_b200/) never imports, links or executes anything in this directory.
 *
 * Parity pinning: the reference's Rust crate cannot be built here (no cargo/rustc), so this oracle is
 * pinned against the golden vectors of the reference's own tests instead:
 *   decoder.rs:743-768   into_llr values               (tests/test_oracle_golden.py)
 *   decoder.rs:770-799   6x3 decode -> zeros
 *   decoder.rs:818-854   450x150 decode -> zeros       (H committed as tests/golden/parity_check_150_450.npz)
 *   decode.py:192-209    seed-1 end-to-end doctest -> 1 success
 * DecoderSpecial has no live test in the reference (decoder_special.rs:687-747 are commented out); its
 * restatement is cross-checked against the generic decoder's semantics and the Kyber end-to-end property
 * (decoded == secret, sum variables == -ssum).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_MAX_DC 24
#define ORACLE_MAX_Q 128
#define ORACLE_MAX_RANGE 512

#define ORACLE_ERR_ARG (-1)
#define ORACLE_ERR_PROB_SUM (-2)  /* decoder.rs:683-684 assert on the row sum */
#define ORACLE_ERR_NO_MAX (-3)    /* decoder.rs:680 expect("No maximum probability found") */
#define ORACLE_ERR_NO_CONFIG (-4) /* decoder.rs:618 assert!(!configurations.is_empty()) */
#define ORACLE_ERR_DEGREE (-5)    /* decoder.rs:465-473 insert_first_none panic */
#define ORACLE_ERR_STRUCTURE (-6) /* decoder_special.rs:388-395, :513-522 (H must be [H' | I]-shaped) */
#define ORACLE_ERR_RANGE (-7)     /* decoder_special.rs:546 b2i index out of bounds */

typedef struct {
    int R, N, E;
    int bvars;     /* number of B-variables: N (generic) or N-R (special) */
    int b, bsum;   /* alphabets [-b, b] and [-bsum, bsum]; bsum == 0 selects the generic decoder */
    int *row_ptr;  /* R+1, CSR: edges of check r are row_ptr[r]..row_ptr[r+1], columns ascending */
    int *col;      /* E, variable of each edge */
    int8_t *sign;  /* E, +1 / -1 (only the sign of H matters, decoder.rs:164-172) */
    int *col_ptr;  /* N+1, CSC */
    int *csc_edge; /* E, CSR edge ids of each variable, rows ascending */
} oracle_graph;

void oracle_graph_free(oracle_graph *g)
{
    if (!g) return;
    free(g->row_ptr);
    free(g->col);
    free(g->sign);
    free(g->col_ptr);
    free(g->csc_edge);
    free(g);
}

/* decoder.rs:494-553 (bsum == 0) / decoder_special.rs:387-464 (bsum > 0).  dv / dc are the const-generic
 * capacity limits of the reference class; exceeding them panics there (insert_first_none). */
int oracle_graph_build(const int8_t *H, int R, int N, int dv, int dc, int b, int bsum, oracle_graph **out)
{
    *out = NULL;
    if (R <= 0 || N <= 0 || b <= 0 || bsum < 0 || dc > ORACLE_MAX_DC) return ORACLE_ERR_ARG;
    if (2 * b + 1 > ORACLE_MAX_Q || 2 * bsum + 1 > ORACLE_MAX_Q) return ORACLE_ERR_ARG;
    if (bsum > 0 && (bsum % b) != 0) return ORACLE_ERR_STRUCTURE; /* decoder_special.rs:388-392 */
    if (bsum > 0 && N <= R) return ORACLE_ERR_STRUCTURE;
    oracle_graph *g = (oracle_graph *)calloc(1, sizeof(*g));
    g->R = R;
    g->N = N;
    g->b = b;
    g->bsum = bsum;
    g->bvars = bsum > 0 ? N - R : N;
    int E = 0;
    for (size_t i = 0; i < (size_t)R * N; ++i) E += H[i] != 0;
    g->E = E;
    g->row_ptr = (int *)calloc(R + 1, sizeof(int));
    g->col = (int *)calloc(E ? E : 1, sizeof(int));
    g->sign = (int8_t *)calloc(E ? E : 1, 1);
    g->col_ptr = (int *)calloc(N + 1, sizeof(int));
    g->csc_edge = (int *)calloc(E ? E : 1, sizeof(int));
    int e = 0, rc = 0;
    for (int r = 0; r < R; ++r) {
        g->row_ptr[r] = e;
        int nsum = 0;
        for (int c = 0; c < N; ++c) {
            int8_t h = H[(size_t)r * N + c];
            if (!h) continue;
            g->col[e] = c;
            g->sign[e] = h < 0 ? -1 : 1;
            g->col_ptr[c + 1]++;
            if (c >= g->bvars) ++nsum;
            ++e;
        }
        int L = e - g->row_ptr[r];
        if (L > dc) rc = ORACLE_ERR_DEGREE;
        if (bsum > 0) {
            /* decoder_special.rs:510-522: the last entry of every check must be its only sum-variable */
            if (L < 1 || nsum != 1 || g->col[e - 1] < g->bvars) rc = ORACLE_ERR_STRUCTURE;
            if ((L - 1) * b > bsum) rc = ORACLE_ERR_RANGE;
        } else if (L < 1) {
            rc = ORACLE_ERR_NO_CONFIG; /* decoder.rs:327 `self.len - 1` underflows: panic */
        }
    }
    g->row_ptr[R] = e;
    for (int c = 0; c < N; ++c) {
        int deg = g->col_ptr[c + 1];
        if (c < g->bvars ? deg > dv : deg > 1) rc = rc ? rc : ORACLE_ERR_DEGREE;
        g->col_ptr[c + 1] += g->col_ptr[c];
    }
    int *fill = (int *)calloc(N, sizeof(int));
    for (int r = 0; r < R; ++r)
        for (int p = g->row_ptr[r]; p < g->row_ptr[r + 1]; ++p) {
            int c = g->col[p];
            g->csc_edge[g->col_ptr[c] + fill[c]++] = p;
        }
    free(fill);
    if (rc) {
        oracle_graph_free(g);
        return rc;
    }
    *out = g;
    return 0;
}

int oracle_graph_edges(const oracle_graph *g) { return g->E; }

#define REAL float
#define SUFFIX _f32
#define RLOG logf
#include "qary_impl.inc"
#undef REAL
#undef SUFFIX
#undef RLOG

#define REAL double
#define SUFFIX _f64
#define RLOG log
#include "qary_impl.inc"
#undef REAL
#undef SUFFIX
#undef RLOG

/* Batched driver: one decode per thread, the reference's own threading model
 * (simulate/decode.py:247-262 ThreadPoolExecutor + pydecoder.rs:55 allow_threads).
 * probs (T x bvars x Q) and probs_sum (T x R x QS) are probabilities when from_probs != 0 (then
 * into_llr is applied, pydecoder.rs:62-63), else LLRs.  status[t] receives each decode's return code. */
int oracle_min_sum_batch_f32(const oracle_graph *g, int max_iter, int algo, int T, int from_probs,
                             const float *in, const float *in_sum, int8_t *hard, float *post,
                             float *post_sum, int *status, int threads)
{
    const int Q = 2 * g->b + 1, QS = 2 * g->bsum + 1;
    const size_t nb = (size_t)g->bvars * Q, ns = g->bsum > 0 ? (size_t)g->R * QS : 0;
    int worst = 0;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (int t = 0; t < T; ++t) {
        float *llr = (float *)malloc(sizeof(float) * (nb ? nb : 1));
        float *llr_sum = (float *)malloc(sizeof(float) * (ns ? ns : 1));
        int rc = 0;
        if (from_probs) {
            rc = oracle_into_llr_f32(in + t * nb, g->bvars, Q, llr);
            if (!rc && ns) rc = oracle_into_llr_f32(in_sum + t * ns, g->R, QS, llr_sum);
        } else {
            memcpy(llr, in + t * nb, sizeof(float) * nb);
            if (ns) memcpy(llr_sum, in_sum + t * ns, sizeof(float) * ns);
        }
        if (!rc)
            rc = oracle_min_sum_f32(g, max_iter, algo, llr, llr_sum, hard + (size_t)t * g->N,
                                    post ? post + t * nb : NULL, (post_sum && ns) ? post_sum + t * ns : NULL);
        status[t] = rc;
        if (rc) {
#ifdef _OPENMP
#pragma omp critical
#endif
            worst = rc;
        }
        free(llr);
        free(llr_sum);
    }
    return worst;
}

/* 1 when the batched drivers really run one decode per thread (built with -fopenmp). */
int oracle_has_openmp(void)
{
#ifdef _OPENMP
    return 1;
#else
    return 0;
#endif
}

int oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
