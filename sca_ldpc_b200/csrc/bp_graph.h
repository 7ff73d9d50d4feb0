// Host-side Tanner-graph build for the binary BP decoder: CSR (rows ascending, columns ascending within a row --
// the visiting order of ldpc.bp_decoder's row lists) -> CSC (rows ascending within a column).  Pure C++ so the
// same code feeds the device upload and the CPU test model.
#pragma once
#include <stdint.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/scaldpc.h"

namespace scaldpc {

struct BpHostGraph {
    int m = 0, n = 0, E = 0;
    int dc_max = 0, dv_max = 0;
    std::vector<int> row_ptr, col, col_ptr, csc_edge;
    std::vector<int> vn_desc;       // 4 ints per sorted position i: {var, col_ptr[var], degree, 0} of var = vn_order[i]
    std::vector<uint32_t> csc_off;  // csc_edge * 128: byte offset of the edge's 128-byte line inside a tile's message block
    // variables sorted by degree (ascending, stable) and the end index of each degree class in that order:
    // class c holds the variables with degree <= kBpVnClass[c] (the last class, "generic", holds the rest)
    std::vector<int> vn_order;
    int class_end[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

constexpr int kBpVnClasses = 7;
constexpr int kBpVnClass[kBpVnClasses] = {1, 4, 8, 12, 16, 24, 32};

inline int build_bp_graph(int m, int n, const int32_t *row_ptr, const int32_t *col, BpHostGraph &g, std::string &err)
{
    auto bad = [&](int code, const std::string &msg) {
        err = msg;
        return code;
    };
    if (m <= 0 || n <= 0 || !row_ptr || !col) return bad(SCALDPC_ERR_ARG, "bad parity-check matrix");
    if (row_ptr[0] != 0) return bad(SCALDPC_ERR_ARG, "row_ptr[0] must be 0");
    g = BpHostGraph();
    g.m = m;
    g.n = n;
    g.E = row_ptr[m];
    if (g.E < 0) return bad(SCALDPC_ERR_ARG, "negative edge count");
    g.row_ptr.assign(row_ptr, row_ptr + m + 1);
    g.col.assign(col, col + g.E);
    g.col_ptr.assign(n + 1, 0);
    for (int r = 0; r < m; ++r) {
        const int a = row_ptr[r], b = row_ptr[r + 1];
        if (b < a) return bad(SCALDPC_ERR_ARG, "row_ptr must be non-decreasing");
        g.dc_max = std::max(g.dc_max, b - a);
        for (int e = a; e < b; ++e) {
            if (col[e] < 0 || col[e] >= n) return bad(SCALDPC_ERR_ARG, "column index out of range in row " + std::to_string(r));
            if (e > a && col[e] <= col[e - 1])
                return bad(SCALDPC_ERR_ARG, "columns must be strictly ascending within row " + std::to_string(r));
            g.col_ptr[col[e] + 1]++;
        }
    }
    for (int c = 0; c < n; ++c) {
        g.dv_max = std::max(g.dv_max, g.col_ptr[c + 1]);
        g.col_ptr[c + 1] += g.col_ptr[c];
    }
    g.csc_edge.resize(std::max(1, g.E));
    std::vector<int> fill(n, 0);
    for (int r = 0; r < m; ++r)
        for (int e = row_ptr[r]; e < row_ptr[r + 1]; ++e) g.csc_edge[g.col_ptr[col[e]] + fill[col[e]]++] = e;
    if ((long long)g.E * 128 > 0xffffffffLL) return bad(SCALDPC_ERR_UNSUPPORTED, "too many edges");
    g.csc_off.resize(g.csc_edge.size());
    for (size_t i = 0; i < g.csc_edge.size(); ++i) g.csc_off[i] = (uint32_t)g.csc_edge[i] * 128u;
    g.vn_order.resize(n);
    for (int c = 0; c < n; ++c) g.vn_order[c] = c;
    std::stable_sort(g.vn_order.begin(), g.vn_order.end(), [&](int a, int b) {
        return g.col_ptr[a + 1] - g.col_ptr[a] < g.col_ptr[b + 1] - g.col_ptr[b];
    });
    g.vn_desc.resize((size_t)4 * std::max(1, n));
    for (int i = 0; i < n; ++i) {
        const int v = g.vn_order[i];
        g.vn_desc[4 * i + 0] = v;
        g.vn_desc[4 * i + 1] = g.col_ptr[v];
        g.vn_desc[4 * i + 2] = g.col_ptr[v + 1] - g.col_ptr[v];
        g.vn_desc[4 * i + 3] = 0;
    }
    {
        int i = 0;
        for (int c = 0; c < kBpVnClasses; ++c) {
            while (i < n && g.col_ptr[g.vn_order[i] + 1] - g.col_ptr[g.vn_order[i]] <= kBpVnClass[c]) ++i;
            g.class_end[c] = i;
        }
        g.class_end[kBpVnClasses] = n;
    }
    return SCALDPC_OK;
}

}  // namespace scaldpc
