// Natural logarithm of an f32, bit-identical to glibc's logf (glibc >= 2.27, sysdeps/ieee754/flt-32/e_logf.c +
// e_logf_data.c, the ARM Optimized Routines algorithm by Szabolcs Nagy): 16-entry table {1/c, log c}, z = x / 2^k in
// [c_i - eps, c_i + eps), r = z/c - 1, log x = k ln2 + log c + r + r^2 (A2 + A1 r + A0 r^2) evaluated in fp64 and
// rounded once to f32.  The reference's into_llr calls Rust's f32::ln (decoder.rs:687), which is libm's logf on
// Linux, so THIS is the function whose rounding the priors must follow (it is not correctly rounded: 0.818 ULP).
// Restated from the published algorithm; the constants were checked against the table inside the local
// libm.so.6 and the function against libm's logf on all 2 139 095 040 positive finite floats (0 mismatches, with and
// without FMA contraction; tests/test_logf.py repeats a sample of that check).
// Replaces (float)log((double)x) of v1: ~12 fp64-pipe instructions instead of ~60, and no "last-bit" exception.
// Included INSIDE namespace scaldpc.
#pragma once

struct ScLogfTab {
    double invc, logc;
};
#ifdef SCALDPC_HOST_MODEL
static const ScLogfTab sc_logf_tab[16] = {
#else
__device__ const ScLogfTab sc_logf_tab[16] = {
#endif
    {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
    {0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2},  {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
    {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3},
    {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
    {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
    {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4},
    {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3},
    {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}};

// `tab` = sc_logf_tab or a copy of it in shared memory (the index differs from lane to lane).
__device__ __forceinline__ float sc_logf(float x, const ScLogfTab *tab)
{
    const double Ln2 = 0x1.62e42fefa39efp-1;
    const double A0 = -0x1.00ea348b88334p-2, A1 = 0x1.5575b0be00b6ap-2, A2 = -0x1.ffffef20a4123p-2;
    uint32_t ix;
    memcpy(&ix, &x, 4);
    if (ix == 0x3f800000u) return 0.0f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {  // zero, subnormal, inf, NaN, negative
        if (ix * 2u == 0u) return -SC_INF;
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2u >= 0xff000000u) return SC_INF - SC_INF;  // NaN
        const float xs = x * 8388608.0f;  // subnormal: normalise
        memcpy(&ix, &xs, 4);
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const int k = (int32_t)tmp >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    float zf;
    memcpy(&zf, &iz, 4);
    const double invc = tab[i].invc, logc = tab[i].logc;
    const double z = (double)zf;
    const double r = z * invc - 1.0;
    const double y0 = logc + (double)k * Ln2;
    const double r2 = r * r;
    double y = A1 * r + A2;
    y = A0 * r2 + y;
    y = y * r2 + (y0 + r);
    return (float)y;
}
