/*
 * scaldpc.h -- C ABI of libscaldpc.so, the B200-native (sm_100a) belief-propagation decoder library.
 *
 * This is the drop-in boundary for the decoder hot path of atneit/SCA-LDPC.  Every entry point states the
 * reference interface it replaces (paths relative to the reference's simulate-with-python/ directory).
 * Plain pointers and sizes only; no torch / pyo3 types.  All functions return 0 on success or a negative
 * SCALDPC_ERR_* code; scaldpc_last_error() returns a human-readable description of the last failure on the
 * calling thread.  The library never aborts the process (the reference panics, SURVEY.md 8b "Errors").
 *
 * There is NO CPU fallback anywhere behind this interface: every decode runs on the CUDA device given at
 * create time, and creation fails if no such device exists.
 *
 * Threading: a handle may be used from several host threads (the reference releases the GIL and is called
 * from a ThreadPoolExecutor, simulate/decode.py:247-262, simulate_rs/src/pydecoder.rs:49-55); calls on one
 * handle are serialised internally.  The batched calls are the intended fast path.
 */
#ifndef SCALDPC_H
#define SCALDPC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes ------------------------------------------------------------------------------------- */
#define SCALDPC_OK 0
#define SCALDPC_ERR_ARG (-1)        /* bad argument (NULL, negative size, unsupported alphabet) */
#define SCALDPC_ERR_PROB_SUM (-2)   /* a prior row does not sum to 1 +- 0.001   (decoder.rs:683-684 assert) */
#define SCALDPC_ERR_NO_MAX (-3)     /* a prior row is all-NaN                     (decoder.rs:680 expect) */
#define SCALDPC_ERR_NO_CONFIG (-4)  /* a check has no valid configuration        (decoder.rs:618 assert) */
#define SCALDPC_ERR_DEGREE (-5)     /* node degree exceeds the class's DV / DC   (decoder.rs:465-473 panic) */
#define SCALDPC_ERR_STRUCTURE (-6)  /* H is not [H' | I]-shaped, BSUM % B != 0, BVARS != N-R
                                       (decoder_special.rs:388-395, :513-522) */
#define SCALDPC_ERR_RANGE (-7)      /* a check's sum leaves [-BSUM, BSUM]        (decoder_special.rs:546) */
#define SCALDPC_ERR_CUDA (-8)       /* CUDA runtime failure (no device, out of memory, launch error) */
#define SCALDPC_ERR_UNSUPPORTED (-9)

const char *scaldpc_last_error(void);
/* Library / build identification: "scaldpc <version> sm_100a". */
const char *scaldpc_version(void);
/* Number of CUDA devices visible, or a negative error. */
int scaldpc_device_count(void);

/* ======================================================================================================
 * q-ary min-sum decoder
 *   replaces  simulate_rs::Decoder<N,R,DV,DC,Q,B,i8>            (simulate_rs/src/decoder.rs:417-666)
 *             simulate_rs::DecoderSpecial<N,R,BVARS,SW,DC,DV,B,..>(simulate_rs/src/decoder_special.rs:294-617)
 *   as bound by the pyo3 classes of simulate_rs/src/pydecoder.rs:13-71 and :85-150.
 * ==================================================================================================== */
typedef struct scaldpc_qary scaldpc_qary;

/* decode flags */
#define SCALDPC_IN_LLR 0x1u       /* inputs are LLRs (skip into_llr); default: probabilities (pydecoder.rs:62) */
#define SCALDPC_DEVICE_PTRS 0x2u  /* all data pointers are device pointers on the handle's device */
#define SCALDPC_CN_EXACT 0x4u     /* every trial with the check node in the reference's rounding order (decisions AND
                                     posteriors bit-exact).  Default: fast forward/backward (min,+) convolution,
                                     VERIFIED -- see SCALDPC_NO_VERIFY */
#define SCALDPC_CN_GENERIC 0x8u   /* force the runtime-shape kernels (check node AND variable node) even where a
                                     specialised one exists */
#define SCALDPC_NO_VERIFY 0x100u  /* skip the verification of the fast path.  By default the last iteration flags
                                     every trial whose hard decisions could depend on the rounding order of the check
                                     node -- a posterior whose two smallest entries are within the risk margin
                                     (default 1e-2, scaldpc_qary_set_risk_margin), or decisions that violate a parity
                                     check sum_j h_ij s_j (+ s_sum,i) = 0, i.e. every trial that did not converge --
                                     and decodes the flagged trials AGAIN with the exact-order check node
                                     (decoder_special.rs:539-553 arithmetic), overwriting their outputs.  Unflagged
                                     trials keep the fast path's posteriors (measured <= 6.1e-5 from the reference's;
                                     the flagged ones, where up to 2.6e-4 was seen, become bit-exact).  Costs one
                                     stream synchronisation per batch and the re-decode of ~1-3 % of a decodable
                                     workload's trials (all of an undecodable one's). */

#define SCALDPC_QARY_EARLY_STOP 0x200u /* OPT-IN per-trial early termination with active-trial compaction -- an EXTENSION:
                                     the reference has no syndrome test and always runs max_iter iterations
                                     (decoder_special.rs:501-504 "noop, we use only num iterations", exit at :611-613), so
                                     parity runs leave this off.  Every iteration emits hard decisions; a trial whose
                                     decisions satisfy every row, sum_j h_ij s_j (+ s_sum,i) = 0, and whose posteriors
                                     are clear of ties (risk margin) is frozen with that iteration's outputs.  Once at
                                     most 1/8 of a batch (chunk) of >= 256 trials is still running the stragglers are re-staged
                                     compactly and decoded from scratch (the same results, fewer idle lanes).  Trials
                                     still running at max_iter keep their last decisions (SCALDPC_NO_VERIFY) or are
                                     decoded with the exact-order check node and the fixed iteration count, i.e. they get
                                     the reference's result.  The sum-variables' messages make the full round trip in
                                     this mode (plain forward/backward flooding).  Frozen trials can differ from the
                                     fixed-iteration decode only if their decisions would still have moved after the row
                                     tests first held; bench.py reports that count.  Synchronises the stream once per
                                     iteration.  Ignored with SCALDPC_CN_EXACT. */

/*
 * scaldpc_qary_create -- replaces `#[new] fn new(py_parity_check, iterations)` (pydecoder.rs:24-45, 96-117),
 * i.e. Decoder::new (decoder.rs:494-553) when bsum == 0 and DecoderSpecial::new (decoder_special.rs:387-464)
 * when bsum > 0.
 *   H         dense row-major R x N int8 parity-check matrix; only the sign of an entry matters
 *             (mult_in_gf, decoder.rs:164-172)
 *   b         alphabet of the first N-R (special) / all N (generic) variables is [-b, b]
 *   bsum      0: generic single-alphabet decoder.  >0: special decoder, H = [H' | I], the last R variables
 *             have alphabet [-bsum, bsum]
 *   dv, dc    capacity limits of the reference class (const generics DV, DC); pass 0 to accept any degree
 *   max_iter  iterations; 0 and 1 both run one iteration (decoder.rs:575-663)
 *   device    CUDA device ordinal
 */
int scaldpc_qary_create(const int8_t *H, int R, int N, int b, int bsum, int dv, int dc, int max_iter, int device,
                        scaldpc_qary **out);
/* The same constructor for a SPARSE parity-check matrix (make_code.py:72-94 builds H = [H' | I] from
 * negacyclic blocks of column weight 1-3 -- 7 entries per row of 768-1280): CSR with columns strictly ascending inside a row
 * (= the order the reference's row-major scan of the dense matrix visits them, decoder.rs:507-539) and one int8 per entry
 * whose sign is the entry (no explicit zeros).  Identical handle to scaldpc_qary_create on the dense matrix. */
int scaldpc_qary_create_csr(int R, int N, const int32_t *row_ptr, const int32_t *col, const int8_t *val, int b, int bsum, int dv,
                            int dc, int max_iter, int device, scaldpc_qary **out);
void scaldpc_qary_destroy(scaldpc_qary *h);

/* Shape queries (sizes a caller needs to allocate buffers). */
int scaldpc_qary_info(const scaldpc_qary *h, int *R, int *N, int *bvars, int *q, int *qs, int *edges,
                      int *max_iter);
int scaldpc_qary_set_max_iter(scaldpc_qary *h, int max_iter);

/*
 * scaldpc_qary_decode -- replaces `fn min_sum(&self, py_channel_output[, py_channel_output_sum])`
 * (pydecoder.rs:53-65, 125-145) = into_llr (decoder.rs:668-692) + min_sum (decoder.rs:560-666 /
 * decoder_special.rs:471-617), batched over T independent trials.
 *   in        T x BVARS x Q   float32, priors of the B-variables (probabilities, or LLRs with SCALDPC_IN_LLR)
 *   in_sum    T x R x QS      float32, priors of the sum-variables (special decoder only, else NULL)
 *   hard      T x N           int8 out: hard decisions, B-variables then sum-variables
 *                             (decoder_special.rs:588, 607)
 *   post      T x BVARS x Q   float32 out or NULL: final `sum` vectors (decoder.rs:638-642)
 *   post_sum  T x R x QS      float32 out or NULL
 *   status    T               int32 out or NULL: per-trial 0 / SCALDPC_ERR_PROB_SUM / _NO_MAX / _NO_CONFIG
 *   stream    cudaStream_t (as void*), NULL = default stream.  With host pointers the call returns after the
 *             results are in `hard`.  With SCALDPC_DEVICE_PTRS all work is enqueued on `stream`; the default
 *             (verified) mode synchronises the stream once per pass of <= 65536 trials to read the flagged-trial
 *             count, the exact-order re-decode that may follow is asynchronous again.  With SCALDPC_DEVICE_PTRS and
 *             SCALDPC_NO_VERIFY or SCALDPC_CN_EXACT the call never synchronises.
 * Returns 0, or the most severe per-trial error (outputs of failed trials are unspecified).  Exception: a
 * SCALDPC_DEVICE_PTRS call that does not synchronise (previous sentence) cannot know the per-trial errors when it
 * returns -- it returns 0 and reports them through `status` only.
 */
int scaldpc_qary_decode(scaldpc_qary *h, int T, const float *in, const float *in_sum, int8_t *hard, float *post,
                        float *post_sum, int32_t *status, unsigned flags, void *stream);

/*
 * scaldpc_qary_decode_f64 -- the fp64 VERIFICATION mode (north_star: "fp32 log-domain; fp64 option for verification"): the
 * decode of scaldpc_qary_decode with every LLR, message and posterior in double -- into_llr (decoder.rs:668-692), check node in
 * the reference's order of additions (decoder.rs:600-629, decoder_special.rs:539-553), variable node (decoder.rs:634-658,
 * decoder_special.rs:566-609) with f64 in place of f32.  in / in_sum are the f32 priors of scaldpc_qary_decode (pydecoder.rs:53
 * takes f32); post T x bvars x q and post_sum T x R x qs are DOUBLE (may be NULL).  flags: 0 or SCALDPC_IN_LLR.  Host pointers,
 * synchronous, run-time-shape kernels, temporary buffers per call: a checker for the f32 path, not a fast path.
 */
int scaldpc_qary_decode_f64(scaldpc_qary *h, int T, const float *in, const float *in_sum, int8_t *hard, double *post,
                            double *post_sum, int32_t *status, unsigned flags);

/*
 * scaldpc_qary_decode_one -- ONE decode in the reference's call shape: pydecoder.rs:53-65 / 125-145
 * (min_sum(channel_output[, channel_output_sum]) -> Vec<i8>, GIL released), called from a thread pool at
 * simulate/decode.py:247-262.  in = bvars x q probabilities, in_sum = R x qs (NULL for the generic decoder), hard = N int8.
 * Blocking, host pointers.  Concurrent calls on one handle are COMBINED inside the library: the thread that finds no batch in
 * progress decodes every request queued so far as one batch (up to 32 = one tile, a CUDA graph replay) and wakes the others.
 * Returns the status of THIS trial: 0, or SCALDPC_ERR_PROB_SUM / NO_MAX / NO_CONFIG where the reference panics
 * (decoder.rs:618, 680, 683-684); other calls of the same batch are unaffected.
 * scaldpc_qary_one_stats: requests served and batches launched so far (requests / batches = the combining factor).
 */
int scaldpc_qary_decode_one(scaldpc_qary *h, const float *in, const float *in_sum, int8_t *hard);
int scaldpc_qary_one_stats(scaldpc_qary *h, long long *requests, long long *batches);

/*
 * scaldpc_qary_decode_observed -- the same decode fed by ORACLE OBSERVATIONS instead of prior vectors.
 * Replaces the reference-side composition that produces min_sum's arguments for every trial:
 *   max_likelihood.s_distribution_from_hard_y (simulate/max_likelihood.py:87-105)  posterior of a value given the
 *                                                                                  observed oracle bits y
 *   kyber.get_channel_probabilities           (simulate/kyber.py:362-376)          packing into channel_output[_sum]
 * followed by pydecoder.rs:125-145 min_sum.  The posterior depends only on y, so a batch of priors is a table with one
 * row per bit pattern plus one row index per (trial, variable); into_llr (decoder.rs:668-692) runs once per table row
 * and 2 bytes per variable instead of 4*Q cross PCIe / are read from HBM.
 *   table      n_rows x Q        float32: table[y] = Pr[S = . | Y = y] (LLRs with SCALDPC_IN_LLR); 1 <= n_rows <= 65535
 *   table_sum  n_rows_sum x QS   float32, special decoder only (already REVERSED as kyber.py:374-375 does)
 *   obs        T x BVARS         uint16 row index of every B-variable
 *   obs_sum    T x R             uint16 row index of every sum-variable (special decoder only, else NULL)
 * Outputs, flags, stream and return value as scaldpc_qary_decode; with SCALDPC_DEVICE_PTRS the tables are device
 * pointers too.  A row index >= n_rows gives that trial the status SCALDPC_ERR_ARG; a table row that violates
 * into_llr's checks gives every trial that uses it SCALDPC_ERR_PROB_SUM / SCALDPC_ERR_NO_MAX.
 * Results are bit-identical to scaldpc_qary_decode on the expanded priors table[obs].
 */
int scaldpc_qary_decode_observed(scaldpc_qary *h, int T, const float *table, int n_rows, const float *table_sum,
                                 int n_rows_sum, const uint16_t *obs, const uint16_t *obs_sum, int8_t *hard, float *post,
                                 float *post_sum, int32_t *status, unsigned flags, void *stream);

/*
 * scaldpc_qary_simulate -- T Monte-Carlo trials of the Kyber attack simulation, synthesised, decoded and scored on the
 * device: no per-trial data crosses PCIe.  Replaces, per trial, the reference-side composition
 *   kyber.sample_secret_coefs / generate_secret_for_H  (simulate/kyber.py:35-47, 100-105)   s ~ centred binomial(eta)
 *   kyber.compute_ssum                                 (simulate/kyber.py:85-92)            ssum = H' s
 *   SimpleOracle                                       (simulate/max_likelihood.py:9-18)    every bit of the value's code
 *                                                       word is observed correctly with probability `accuracy`
 *   s_distribution_from_hard_y + get_channel_probabilities + min_sum                        (= scaldpc_qary_decode_observed)
 *   success <=> decoded[:BVARS] == s  (and decoded[BVARS:] == -ssum)                        
 * The reference draws from os.urandom; here trial g = first_trial + t draws from Philox4x32-10 keyed by `seed` with the
 * counter layout documented in sca_ldpc_b200/csrc/qary_sim.cuh, so a campaign's trials do not depend on how it is cut
 * into calls or spread over GPUs.  oracle/kyber_sim.py restates the generator.
 *   table, table_sum   as scaldpc_qary_decode_observed (host pointers); row y = posterior given observed pattern y
 *   code       Q   uint16: code word (bit k = oracle bit k) of secret value v at index v + B; nbits <= 16 bits each
 *   code_sum   QS  uint16: code word of check-sum value v at index v + BSUM; nbits_sum bits each
 *                  (kyber.coding_from_patterns, simulate/kyber.py:50-57)
 *   accuracy   probability that one oracle bit is right; eta: secret coefficients are centred binomial(eta), eta <= B
 *   counters   5 x int64 out: trials, successes (secret and sum variables right), trials whose secret is right,
 *              wrong secret coefficients, wrong sum variables
 *   secret (T x BVARS int8), ssum (T x R int16), obs (T x BVARS uint16), obs_sum (T x R uint16), hard (T x N int8):
 *              optional host outputs of the synthesised trials and their decisions (NULL = not copied)
 * Needs a special decoder (BSUM > 0) whose BSUM covers row weight * eta.  Synchronous; flags as scaldpc_qary_decode
 * (SCALDPC_DEVICE_PTRS is rejected).
 */
int scaldpc_qary_simulate(scaldpc_qary *h, int T, uint64_t seed, uint64_t first_trial, const float *table, int n_rows,
                          const float *table_sum, int n_rows_sum, const uint16_t *code, int nbits, const uint16_t *code_sum,
                          int nbits_sum, double accuracy, int eta, int64_t *counters, int8_t *secret, int16_t *ssum,
                          uint16_t *obs, uint16_t *obs_sum, int8_t *hard, unsigned flags, void *stream);

/* Kernels launched by this handle so far (for bench.py's "gpu_launches") and device time is NOT tracked
 * here -- callers time with CUDA events on `stream`. */
long long scaldpc_qary_launch_count(const scaldpc_qary *h);
/* Statistics of the last decode / decode_observed / simulate call on this handle: trials decoded, how many of them the
 * verification re-decoded with the exact-order check node (0 with SCALDPC_NO_VERIFY / SCALDPC_CN_EXACT), and the
 * trial-iterations actually launched (restarted and re-decoded trials counted again) -- any pointer may be NULL. */
int scaldpc_qary_last_run(const scaldpc_qary *h, long long *trials, long long *redecoded, long long *trial_iterations);
/* Per-trial iteration counts of the last decode / decode_observed call, which must have used SCALDPC_QARY_EARLY_STOP on
 * at least T trials: the iteration in which the trial was frozen, or max_iter.  flags: SCALDPC_DEVICE_PTRS or 0. */
int scaldpc_qary_last_iterations(scaldpc_qary *h, int32_t *iters, int T, unsigned flags, void *stream);
/* Risk margin of the verification (absolute LLR difference between a posterior's two smallest entries below which
 * the trial is re-decoded); default 1e-2, or the environment variable SCALDPC_RISK_MARGIN at create time. */
int scaldpc_qary_set_risk_margin(scaldpc_qary *h, float eps);

/*
 * Per-kernel timing probe used by bench.py's roofline line: runs `reps` back-to-back launches of ONE kernel
 * of the decode pipeline on the state left behind by the previous decode of `T` trials and returns the
 * average duration in milliseconds measured with CUDA events on `stream`.
 *   which  0 = check-node kernel (middle iteration), 1 = variable-node kernel (middle iteration)
 */
int scaldpc_qary_time_kernel(scaldpc_qary *h, int which, int T, unsigned flags, int reps, void *stream,
                             float *avg_ms);

/* ======================================================================================================
 * binary soft BP (product-sum, flooding, syndrome-match early stop)
 *   replaces  ldpc.bp_decoder(H, max_iter=, bp_method="product_sum", channel_probs=) and .decode(vec)
 *   third-party PyPI package ldpc==0.1.3 (requirements.txt:8); call sites simulate/hqc.py:694-708 and
 *   simulate/decode.py:155-171.
 * ==================================================================================================== */
typedef struct scaldpc_bp scaldpc_bp;

/* CSR parity-check matrix: row_ptr (m+1), col (row_ptr[m]) ascending within a row. */
int scaldpc_bp_create(int m, int n, const int32_t *row_ptr, const int32_t *col, int max_iter, int device,
                      scaldpc_bp **out);
void scaldpc_bp_destroy(scaldpc_bp *h);

#define SCALDPC_BP_PROBS_PER_TRIAL 0x10u /* channel_probs is T x n instead of n */
#define SCALDPC_BP_NO_EARLY_STOP 0x20u   /* run max_iter iterations regardless of convergence */
#define SCALDPC_BP_NO_COMPACT 0x40u      /* keep converged trials in place (freeze only); default: running trials are
                                            compacted into fewer tiles whenever they fill <= half of the slots */
/*
 *   synd           T x m   uint8 syndromes
 *   channel_probs  n (or T x n) float64 prior error probabilities           (hqc.py:684-690)
 *   decoding       T x n   uint8 out: estimated error pattern
 *   llr            T x n   float32 out or NULL: log(p0/p1) posteriors ("log_prob_ratios")
 *   iters          T       int32 out or NULL: iterations run per trial ("iter")
 *   converged      T       int32 out or NULL                              ("converge")
 */
int scaldpc_bp_decode(scaldpc_bp *h, int T, const uint8_t *synd, const double *channel_probs, uint8_t *decoding,
                      float *llr, int32_t *iters, int32_t *converged, unsigned flags, void *stream);
/*
 * scaldpc_bp_decode_scored -- decode, then the statistics hqc.decode computes from the decoded vector (hqc.py:709-758)
 * on the device, so a campaign moves 24 bytes per trial back instead of n.  Needs H = [Hin | I_m]
 * (hqc.py:680): columns < n_msg = n - m are message bits, column n_msg + i is the measurement error of check i.
 *   truth     T x weight int32: support of the true message of every trial (y_sparse, hqc.py:1043-1046), distinct
 *   stats     T x 6 int32 out: unsatisfied, good_flips, bad_flips, found_bad_satisfied_checks,
 *             found_bad_unsatisfied_checks, success   -- the arguments of tracking.add_decoder_stats (hqc.py:751-759).
 *             hqc.decode scores the CORRECTED word msg ^ error (bpd.decode of a length-n vector, hqc.py:703-708): with
 *             c_i the measured check bit and e_i the decoded measurement error, found_bad_unsatisfied counts c_i = 1 and
 *             e_i = 1, found_bad_satisfied counts c_i = 0 and e_i = 1.
 *   decoding  T x n uint8 out, or NULL with host pointers (not copied back): the ERROR ESTIMATE as scaldpc_bp_decode
 *             returns it, not the corrected word
 *   n_checks  NULL, or T int32 as scaldpc_bp_decode_nested (the statistics then cover the active checks only)
 * Other arguments, flags and return value as scaldpc_bp_decode.
 */
int scaldpc_bp_decode_scored(scaldpc_bp *h, int T, int n_msg, const int32_t *n_checks, const uint8_t *synd,
                             const double *channel_probs, const int32_t *truth, int weight, int32_t *stats, uint8_t *decoding,
                             int32_t *iters, int32_t *converged, unsigned flags, void *stream);
/*
 * scaldpc_bp_decode_nested -- one batch over NESTED graphs.  The reference's decode_every loop
 * (hqc.py:953-984) appends measured checks to H and re-decodes on the first R_k rows: H_k = [Hin[:R_k] | I_{R_k}]
 * (hqc.py:680).  With the handle built on the largest graph [Hin | I_m], trial t decodes on its own prefix:
 *   n_checks  T int32, 0 <= n_checks[t] <= m: only the checks < n_checks[t] take part (messages, syndrome test);
 *             the others send the neutral message, so iterations, convergence and the decisions of the message bits and
 *             of the active checks' error bits equal a decode on H_k with channel_probs / synd cut to match.  The
 *             decisions reported for the inactive checks' error bits are their priors' (0 when p < 1/2).
 * Everything else as scaldpc_bp_decode.
 */
int scaldpc_bp_decode_nested(scaldpc_bp *h, int T, const int32_t *n_checks, const uint8_t *synd, const double *channel_probs,
                             uint8_t *decoding, float *llr, int32_t *iters, int32_t *converged, unsigned flags, void *stream);
/*
 * scaldpc_bp_create_multi / scaldpc_bp_decode_multi -- SEGMENTED CSR: one batch over UNRELATED graphs.
 * The reference draws a NEW code for every simulated attack (hqc.py:1056 make_random_ldpc_parity_check_matrix(N, weight,
 * rng)) and decodes that attack's growing ladder of check sets on it (hqc.py:953-984); a campaign of attacks is therefore
 * a batch whose trials live on different graphs of the same shape.  Here the n_graphs graphs share one handle and one
 * workspace: the graph tables are concatenated, every tile of 32 trial slots belongs to ONE graph (an attack's ladder
 * levels / noise realisations fill the lanes), so graph indices stay warp-uniform and every kernel is the single-graph
 * kernel with a per-tile table offset.
 *   n_graphs   G >= 1 graphs, all m x n
 *   row_ptr    G x (m+1) int32, every segment starts at 0
 *   col        the G column arrays back to back (segment g has row_ptr[g][m] entries), ascending within a row
 * scaldpc_bp_decode_multi:
 *   graph_of   T int32, HOST pointer even with SCALDPC_DEVICE_PTRS: graph of every trial, any order (trials are grouped
 *              by graph internally; outputs come back in the caller's order)
 *   n_checks   NULL or T int32 as scaldpc_bp_decode_nested (every trial its own prefix of its graph's rows)
 *   truth / weight / stats   all NULL / 0, or as scaldpc_bp_decode_scored (needs H_g = [Hin_g | I])
 *   decoding, llr, iters, converged, synd, channel_probs, flags, stream   as scaldpc_bp_decode; channel_probs is shared
 *              by all graphs (n values) or per trial (SCALDPC_BP_PROBS_PER_TRIAL)
 * Running trials are frozen as they converge but NOT compacted (a tile's lanes must share a graph).
 * Each trial's results equal a scaldpc_bp_decode on its own graph.
 */
int scaldpc_bp_create_multi(int n_graphs, int m, int n, const int32_t *row_ptr, const int32_t *col, int max_iter, int device,
                            scaldpc_bp **out);
/*
 * scaldpc_bp_create_circulant -- the graphs of scaldpc_bp_create_multi built ON THE DEVICE from what the reference builds them
 * from.  The reference makes a dense N x N circulant of a sparse first column per attack (make_code.py:220-244
 * make_random_ldpc_parity_check_matrix, 2.5 GB for HQC-128), takes the rows of the measured positions (add_check,
 * hqc.py:885-908) and appends the identity (hqc.py:680).  Here:
 *   first_col  G x weight int32: support of every circulant's first column (= r1_sparse, hqc.py:1059), distinct, any order
 *   rows       G x m int32: the measured positions b_0 .. b_{m-1} of every graph, in the order the checks were added
 * Row i of graph g is { (rows[g][i] - r) mod n_msg : r in first_col[g] } and the identity column n_msg + i.  CSR, CSC, the
 * degree-sorted variable order and every other table the decoder uses are produced by kernels (bp_build.cuh); the handle is
 * indistinguishable from one made by scaldpc_bp_create_multi on the same matrices.  weight <= 128.
 */
int scaldpc_bp_create_circulant(int n_graphs, int n_msg, int weight, int m, const int32_t *first_col, const int32_t *rows,
                                int max_iter, int device, scaldpc_bp **out);
int scaldpc_bp_decode_multi(scaldpc_bp *h, int T, const int32_t *graph_of, const int32_t *n_checks, const uint8_t *synd,
                            const double *channel_probs, const int32_t *truth, int weight, int32_t *stats, uint8_t *decoding,
                            float *llr, int32_t *iters, int32_t *converged, unsigned flags, void *stream);
/* Shape queries: checks, variables, edges, iteration cap (max_iter = 0 at create time means n, as upstream). */
/*
 * scaldpc_bp_decode_static -- scaldpc_bp_decode for callers that must not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid
 * cudaMalloc.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine) and never synchronises: it
 * is a fixed sequence of launches on `stream`; all of max_iter iterations are launched, converged trials freeze on the device
 * exactly as in scaldpc_bp_decode, so decisions, iteration counts and convergence flags are identical to that call's.
 *   synd T x m uint8, channel_probs n doubles (shared by the trials), decoding T x n uint8, iters / converged T int32 (may be
 *   NULL): all DEVICE pointers.  Single-graph handles only.
 * Every per-decode buffer lives in one file-scope device array of SCALDPC_BP_STATIC_BYTES that every GPU loading the library
 * holds until exit; a decode needs about T/32 * (128 * edges + 4 * (m + n)) bytes of it, and a call that needs more is
 * refused with SCALDPC_ERR_ARG (the message states both sizes).  All calls -- any handle, any stream -- share the array, which
 * keeps the last call's contents (the call clears what must start at zero): calls that overlap on the GPU race, so serialise
 * them on one stream or with events.  Calls on ONE handle also share its [n] prior table with scaldpc_bp_decode.
 */
#define SCALDPC_BP_STATIC_BYTES (8u << 20)
int scaldpc_bp_decode_static(scaldpc_bp *h, int T, const uint8_t *synd, const double *channel_probs, uint8_t *decoding,
                             int32_t *iters, int32_t *converged, void *stream);
int scaldpc_bp_info(const scaldpc_bp *h, int *m, int *n, int *edges, int *max_iter);
long long scaldpc_bp_launch_count(const scaldpc_bp *h);
/* Statistics of the last decode on this handle: iterations launched (the slowest trial's count) and the number of
 * active-trial compactions performed. */
int scaldpc_bp_last_run(const scaldpc_bp *h, int *iterations_launched, int *compactions);
/* Host-buffer decodes into PAGE-LOCKED output arrays send the T x n decision bytes (and LLRs) home while the non-converging
 * stragglers still iterate: at the first compaction that leaves <= 256 trials running, every other row is final and the whole
 * array is copied on a side stream; the stragglers' rows are sent again at the end.  rows = stragglers at that moment in the
 * last decode on this handle, -1 = the copy ran at the end as usual (pageable output, small batch, no such compaction). */
int scaldpc_bp_last_early_copy(const scaldpc_bp *h, int *rows);
/* Per-kernel timing probe (bench.py roofline leg), as scaldpc_qary_time_kernel:
 *   which  0 = check-to-bit kernel, 1 = bit-to-check kernel; every tile of the previous decode is re-activated. */
int scaldpc_bp_time_kernel(scaldpc_bp *h, int which, int T, unsigned flags, int reps, void *stream, float *avg_ms);

#ifdef __cplusplus
}
#endif
#endif /* SCALDPC_H */
