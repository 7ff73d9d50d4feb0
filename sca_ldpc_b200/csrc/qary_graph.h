// Host-side Tanner-graph build for the q-ary decoder: dense int8 H -> CSR (check-major, column ascending) +
// CSC (row ascending) + per-edge message-slot offsets, in exactly the orders of the reference's graph build
// (decoder.rs:494-553, decoder_special.rs:387-464), with the reference's construction-time panics turned into
// status codes.  Pure C++ (no CUDA) so the same code feeds the device upload and the CPU test model.
#pragma once
#include <stdint.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/scaldpc.h"

namespace scaldpc {

struct EdgeRef {
    int off;   // first message slot of the edge
    int sign;  // +1 / -1
};

struct HostGraph {
    int R = 0, N = 0, bvars = 0, B = 0, BSUM = 0, Q = 0, QS = 0, E = 0;
    int S = 0;       // message slots per trial (sum of the alphabets of all edges)
    int SL = 0;      // bvars * Q
    int SLS = 0;     // R * QS (special decoder) or 0
    int lmax = 0;    // max row degree
    int wmax = 0;    // max DP table width 2 * hh + 1, hh = max_k min(sum of a check's first k alphabets, sum of the others)
    int regular_sw = 0;  // >0: every check is exactly `regular_sw` B-variables + its sum-variable, BSUM == sw*B
    std::vector<int> row_ptr, col, edge_off, col_ptr;
    std::vector<int8_t> sign;
    std::vector<EdgeRef> csc;
};

// Sparse input: CSR with columns strictly ascending inside a row (the order the reference's row-major scan of the dense
// matrix visits them, decoder.rs:507-539) and one int8 value per entry, of which only the sign matters.
// Returns SCALDPC_OK or an error code; `err` receives the description.
inline int build_graph_csr(int R, int N, const int32_t *in_row_ptr, const int32_t *in_col, const int8_t *in_val, int b, int bsum,
                           int dv, int dc, HostGraph &g, std::string &err)
{
    auto bad = [&](int code, const std::string &m) {
        err = m;
        return code;
    };
    if (!in_row_ptr || !in_col || !in_val || R <= 0 || N <= 0) return bad(SCALDPC_ERR_ARG, "bad parity-check matrix shape");
    if (in_row_ptr[0] != 0) return bad(SCALDPC_ERR_ARG, "row_ptr[0] must be 0");
    for (int r = 0; r < R; ++r) {
        if (in_row_ptr[r + 1] < in_row_ptr[r]) return bad(SCALDPC_ERR_ARG, "row_ptr must be non-decreasing");
        for (int e = in_row_ptr[r]; e < in_row_ptr[r + 1]; ++e) {
            if (in_col[e] < 0 || in_col[e] >= N) return bad(SCALDPC_ERR_ARG, "column index out of range in row " + std::to_string(r));
            if (e > in_row_ptr[r] && in_col[e] <= in_col[e - 1])
                return bad(SCALDPC_ERR_ARG, "columns must be strictly ascending within row " + std::to_string(r));
            if (in_val[e] == 0) return bad(SCALDPC_ERR_ARG, "explicit zero entry in row " + std::to_string(r));
        }
    }
    if (R > 65535 || N > 65535) return bad(SCALDPC_ERR_ARG, "R and N must fit the reference's u16 keys (decoder.rs:211)");
    if (b <= 0 || 2 * b + 1 > 31) return bad(SCALDPC_ERR_ARG, "B must be in [1, 15]");
    if (bsum < 0 || 2 * bsum + 1 > 127) return bad(SCALDPC_ERR_ARG, "BSUM must be in [0, 63]");
    const bool special = bsum > 0;
    if (special && (bsum % b) != 0)  // decoder_special.rs:388-392
        return bad(SCALDPC_ERR_STRUCTURE,
                   "BSUM (" + std::to_string(bsum) + ") must be multiple of B (" + std::to_string(b) + ")");
    if (special && N <= R) return bad(SCALDPC_ERR_STRUCTURE, "special decoder needs N > R (BVARS = N - R)");

    g = HostGraph();
    g.R = R;
    g.N = N;
    g.B = b;
    g.BSUM = bsum;
    g.Q = 2 * b + 1;
    g.QS = special ? 2 * bsum + 1 : 0;
    g.bvars = special ? N - R : N;
    g.row_ptr.assign(R + 1, 0);
    g.col_ptr.assign(N + 1, 0);

    int slots = 0, sw0 = -1;
    bool regular = special;
    for (int r = 0; r < R; ++r) {
        g.row_ptr[r] = (int)g.col.size();
        int nsum = 0, tot = 0;
        for (int e_in = in_row_ptr[r]; e_in < in_row_ptr[r + 1]; ++e_in) {
            const int c = in_col[e_in];
            const int8_t v = in_val[e_in];
            const bool is_sum = c >= g.bvars;
            g.col.push_back(c);
            g.sign.push_back(v < 0 ? -1 : 1);  // only the sign matters (mult_in_gf, decoder.rs:164-172)
            g.edge_off.push_back(slots);
            slots += is_sum ? g.QS : g.Q;
            tot += is_sum ? bsum : b;
            nsum += is_sum;
            g.col_ptr[c + 1]++;
        }
        const int L = (int)g.col.size() - g.row_ptr[r];
        const std::string where = "check " + std::to_string(r) + ": ";
        if (dc > 0 && L > dc)
            return bad(SCALDPC_ERR_DEGREE, where + std::to_string(L) + " variables, more than DC=" + std::to_string(dc) +
                                               " (decoder.rs:465-473)");
        if (special) {
            // decoder_special.rs:510-522: the last entry of every check is its one sum-variable
            if (L < 1 || nsum != 1 || g.col.back() < g.bvars)
                return bad(SCALDPC_ERR_STRUCTURE,
                           where + "parity-check matrix must be [H' | I]-shaped (one sum-variable, last in the row)");
            if ((L - 1) * b > bsum)
                return bad(SCALDPC_ERR_RANGE, where + "its B-variables can sum outside [-BSUM, BSUM] (decoder_special.rs:546)");
            if (sw0 < 0) sw0 = L - 1;
            if (L - 1 != sw0) regular = false;
        } else if (L < 1) {
            return bad(SCALDPC_ERR_NO_CONFIG, where + "no variables (decoder.rs:327 underflow panic)");
        }
        g.lmax = std::max(g.lmax, L);
        {  // half-width of the DP tables (k_cn_generic): max over k of min(first k alphabets, the others)
            int hh = 0, pre = 0;
            for (int k = 0; k < L; ++k) {
                pre += (g.col[g.row_ptr[r] + k] >= g.bvars) ? bsum : b;
                hh = std::max(hh, std::min(pre, tot - pre));
            }
            g.wmax = std::max(g.wmax, 2 * hh + 1);
        }
    }
    g.row_ptr[R] = (int)g.col.size();
    g.edge_off.push_back(slots);
    g.E = (int)g.col.size();
    g.S = slots;
    g.SL = g.bvars * g.Q;
    g.SLS = special ? R * g.QS : 0;
    for (int c = 0; c < N; ++c) {
        const int deg = g.col_ptr[c + 1];
        if (c >= g.bvars) {
            if (deg > 1)
                return bad(SCALDPC_ERR_DEGREE, "sum-variable " + std::to_string(c) + " appears in " + std::to_string(deg) +
                                                   " checks (capacity 1, decoder_special.rs:423)");
        } else if (dv > 0 && deg > dv) {
            return bad(SCALDPC_ERR_DEGREE, "variable " + std::to_string(c) + " is in " + std::to_string(deg) +
                                               " checks, more than DV=" + std::to_string(dv) + " (decoder.rs:465-473)");
        }
        g.col_ptr[c + 1] += g.col_ptr[c];
    }
    g.csc.resize(g.E);
    std::vector<int> fill(N, 0);
    for (int r = 0; r < R; ++r)
        for (int e = g.row_ptr[r]; e < g.row_ptr[r + 1]; ++e) {
            const int c = g.col[e];
            g.csc[g.col_ptr[c] + fill[c]++] = EdgeRef{g.edge_off[e], g.sign[e]};
        }
    if (regular && sw0 > 0 && bsum == sw0 * b) g.regular_sw = sw0;
    return SCALDPC_OK;
}

// Dense input (pydecoder.rs:24-45: the reference's constructor takes the dense R x N int8 matrix): row-major scan -> CSR.
inline int build_graph(const int8_t *H, int R, int N, int b, int bsum, int dv, int dc, HostGraph &g, std::string &err)
{
    if (!H || R <= 0 || N <= 0) {
        err = "bad parity-check matrix shape";
        return SCALDPC_ERR_ARG;
    }
    std::vector<int32_t> rp(R + 1, 0), col;
    std::vector<int8_t> val;
    for (int r = 0; r < R; ++r) {
        for (int c = 0; c < N; ++c)
            if (H[(size_t)r * N + c]) {
                col.push_back(c);
                val.push_back(H[(size_t)r * N + c]);
            }
        rp[r + 1] = (int32_t)col.size();
    }
    col.push_back(0);  // never read; keeps data() non-null for an all-zero matrix
    val.push_back(1);
    return build_graph_csr(R, N, rp.data(), col.data(), val.data(), b, bsum, dv, dc, g, err);
}

// Shared-memory scratch (in floats per lane) of the runtime-shape check-node kernel.
inline size_t generic_scratch_floats(const HostGraph &g)
{
    const size_t qm = (size_t)std::max(g.Q, g.QS);
    return (size_t)g.lmax * qm + (size_t)(g.lmax + 1) * g.wmax + 2 * (size_t)g.wmax;
}

}  // namespace scaldpc
