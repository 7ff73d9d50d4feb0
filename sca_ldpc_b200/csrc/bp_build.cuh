// Parity-check graphs of the HQC attack simulation built ON THE DEVICE.
//
// The reference draws a code per simulated attack as a dense N x N circulant (make_code.py:220-244
// make_random_ldpc_parity_check_matrix: scipy.linalg.circulant of a sparse first column, 2.5 GB for HQC-128), picks the rows
// of the measured positions (add_check, hqc.py:885-908) and appends the identity (hqc.py:680).  Here a graph is described by
// what it is made of -- the support r_0 .. r_{W-1} of the circulant's first column and the measured positions b_0 .. b_{m-1} --
// and every table the decoder needs is produced by the kernels below, for G graphs at once:
//     row i of graph g = { (b_i - r) mod N : r in first_col[g] }  (ascending)  followed by  N + i
// k_circ_rows     CSR columns.  With the support sorted ascending and t = #{r <= b}, the ascending row is
//                 b - r_{t-1}, ..., b - r_0, b + N - r_{W-1}, ..., b + N - r_t : a rotation of the reversed support, no sort.
// k_circ_count    column degrees (atomic histogram), k_circ_scan  col_ptr (one CTA per graph, chunked block scan),
// k_circ_fill     CSC entries in arbitrary order (atomic cursor), k_circ_order  every column's entries sorted by edge id =
//                 rows ascending (ldpc's column lists; degrees are ~10), + the CSC-ordered companions of bp_kernels.cuh
// k_circ_degsort  variables sorted by degree, stable (one compaction pass per degree value, one CTA per graph), class bounds
// Everything is integer work; the results are the tables build_bp_graph (bp_graph.h) makes on the host, entry for entry.
#pragma once
#include <stdint.h>

namespace scaldpc {

struct CircDev {
    int G, N, W, m, n, E;  // n = N + m, E = m * (W + 1)
    int *first_col;         // [G][W]  (sorted in place by k_circ_sort_support)
    const int *rows;        // [G][m]
    int *row_ptr, *col, *col_ptr, *vn_order, *vn_desc, *edge_row, *edge_csc, *csc_row, *class_end;
    uint32_t *csc_off;
    int *cursor;            // [G][n] scratch: fill cursors, then degrees
    int *err;               // [0] != 0: bad description (duplicate support entry, position out of range); [1] = max column degree
};

// One thread per graph: insertion sort of the support (W <= 128), duplicates / range flagged.
__global__ void k_circ_sort_support(const CircDev d)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= d.G) return;
    int *s = d.first_col + (size_t)g * d.W;
    for (int i = 1; i < d.W; ++i) {
        const int v = s[i];
        int j = i - 1;
        while (j >= 0 && s[j] > v) {
            s[j + 1] = s[j];
            --j;
        }
        s[j + 1] = v;
    }
    for (int i = 0; i < d.W; ++i)
        if (s[i] < 0 || s[i] >= d.N || (i && s[i] == s[i - 1])) atomicExch(d.err, 1);
}

// One thread per (graph, row, k).
__global__ void __launch_bounds__(256) k_circ_rows(const CircDev d)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int per = d.W + 1;
    if (idx >= (long long)d.G * d.m * per || d.err[0]) return;  // a bad support (k_circ_sort_support) stops everything
    const int k = (int)(idx % per);
    const int i = (int)((idx / per) % d.m);
    const int g = (int)(idx / ((long long)per * d.m));
    const int b = __ldg(d.rows + (size_t)g * d.m + i);
    if (b < 0 || b >= d.N) {
        atomicExch(d.err, 1);
        return;
    }
    const int *s = d.first_col + (size_t)g * d.W;
    int c;
    if (k == d.W) {
        c = d.N + i;  // the identity column of check i (hqc.py:680)
    } else {
        int t = 0;  // #{r <= b}: binary search in the sorted support
        for (int lo = 0, hi = d.W; lo < hi;) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(s + mid) <= b) lo = mid + 1;
            else hi = mid;
            t = lo;
        }
        c = k < t ? b - __ldg(s + t - 1 - k) : b + d.N - __ldg(s + d.W - 1 - (k - t));
    }
    const size_t e = (size_t)g * d.E + (size_t)i * per + k;
    d.col[e] = c;
    d.edge_row[e] = i;
    if (k == 0) d.row_ptr[(size_t)g * (d.m + 1) + i] = i * per;
    if (k == 0 && i == d.m - 1) d.row_ptr[(size_t)g * (d.m + 1) + d.m] = d.m * per;
    atomicAdd(d.cursor + (size_t)g * d.n + c, 1);  // column degree
}

// One CTA per graph: col_ptr = exclusive scan of the degrees; the degrees stay in `cursor` (k_circ_fill counts them down).
__global__ void __launch_bounds__(1024) k_circ_scan(const CircDev d)
{
    __shared__ int wsum[32];
    __shared__ int s_base;
    const int g = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int *deg = d.cursor + (size_t)g * d.n;
    int *cp = d.col_ptr + (size_t)g * (d.n + 1);
    if (tid == 0) s_base = 0;
    __syncthreads();
    int dmax = 0;
    for (int v0 = 0; v0 < d.n; v0 += 1024) {
        const int v = v0 + tid;
        const int c = v < d.n ? deg[v] : 0;
        dmax = max(dmax, c);
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int x = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += x;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int x = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += x;
            }
            wsum[lane] = w;
        }
        __syncthreads();
        if (v < d.n) cp[v] = s_base + (warp ? wsum[warp - 1] : 0) + incl - c;
        __syncthreads();
        if (tid == 0) s_base += wsum[31];
        __syncthreads();
    }
    if (tid == 0) cp[d.n] = s_base;
#pragma unroll
    for (int o = 16; o; o >>= 1) dmax = max(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    if (lane == 0) atomicMax(d.err + 1, dmax);
}

// One thread per edge: drop the edge into its column's segment (any order; k_circ_order sorts the segment).
__global__ void __launch_bounds__(256) k_circ_fill(const CircDev d, int *csc_edge)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)d.G * d.E || d.err[0]) return;  // after a bad description `col` is not fully written
    const int g = (int)(idx / d.E), e = (int)(idx % d.E);
    const int c = d.col[idx];
    const int left = atomicSub(d.cursor + (size_t)g * d.n + c, 1);  // degree .. 1
    csc_edge[(size_t)g * d.E + d.col_ptr[(size_t)g * (d.n + 1) + c] + left - 1] = e;
}

// One thread per (graph, variable): its CSC entries ascending by edge id (= by row), and the CSC-ordered tables.
__global__ void __launch_bounds__(256) k_circ_order(const CircDev d, int *csc_edge)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)d.G * d.n || d.err[0]) return;
    const int g = (int)(idx / d.n), v = (int)(idx % d.n);
    const int *cp = d.col_ptr + (size_t)g * (d.n + 1) + v;
    const int p0 = cp[0], deg = cp[1] - cp[0];
    int *seg = csc_edge + (size_t)g * d.E + p0;
    for (int i = 1; i < deg; ++i) {
        const int x = seg[i];
        int j = i - 1;
        while (j >= 0 && seg[j] > x) {
            seg[j + 1] = seg[j];
            --j;
        }
        seg[j + 1] = x;
    }
    for (int i = 0; i < deg; ++i) {
        const int e = seg[i];
        const size_t q = (size_t)g * d.E + p0 + i;
        d.csc_off[q] = (uint32_t)e * 128u;
        d.csc_row[q] = d.edge_row[(size_t)g * d.E + e];
        d.edge_csc[(size_t)g * d.E + e] = p0 + i;
    }
    d.cursor[(size_t)g * d.n + v] = deg;  // the degree, for k_circ_degsort
}

// One CTA per graph: variables ordered by degree (ascending, ties by index: a stable counting sort as one block-wide
// compaction per degree value), the {var, col_ptr, degree} descriptors, and the end of every degree class.
__global__ void __launch_bounds__(1024) k_circ_degsort(const CircDev d, int dv_max, const int *class_bound, int n_classes)
{
    __shared__ int wsum[32];
    __shared__ int s_base;
    const int g = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int *deg = d.cursor + (size_t)g * d.n;
    const int *cp = d.col_ptr + (size_t)g * (d.n + 1);
    int *order = d.vn_order + (size_t)g * d.n;
    int *desc = d.vn_desc + (size_t)g * 4 * d.n;
    if (tid == 0) s_base = 0;
    __syncthreads();
    int cls = 0;
    for (int dd = 0; dd <= dv_max; ++dd) {
        // classes whose bound lies below this degree end here
        while (cls < n_classes && class_bound[cls] < dd) {
            if (tid == 0) d.class_end[(size_t)g * (n_classes + 1) + cls] = s_base;
            ++cls;
        }
        for (int v0 = 0; v0 < d.n; v0 += 1024) {
            const int v = v0 + tid;
            const int hit = (v < d.n && deg[v] == dd) ? 1 : 0;
            const unsigned bal = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) wsum[warp] = __popc(bal);
            __syncthreads();
            if (warp == 0) {
                int w = wsum[lane];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int x = __shfl_up_sync(0xffffffffu, w, o);
                    if (lane >= o) w += x;
                }
                wsum[lane] = w;
            }
            __syncthreads();
            if (hit) {
                const int pos = s_base + (warp ? wsum[warp - 1] : 0) + __popc(bal & ((1u << lane) - 1u));
                order[pos] = v;
                desc[4 * pos + 0] = v;
                desc[4 * pos + 1] = cp[v];
                desc[4 * pos + 2] = dd;
                desc[4 * pos + 3] = 0;
            }
            __syncthreads();
            if (tid == 0) s_base += wsum[31];
            __syncthreads();
        }
    }
    if (tid == 0)
        for (; cls <= n_classes; ++cls) d.class_end[(size_t)g * (n_classes + 1) + cls] = cls < n_classes ? s_base : d.n;
}

}  // namespace scaldpc
