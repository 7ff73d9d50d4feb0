// Microbenchmark 2: which instruction mix folds (add, min) candidates fastest on sm_100a?  Every variant performs, per
// loop step and chain, 2 adds and 2 min-folds (one FADD2 + one FMNMX3 in the baseline).  Prints cycles per (2 add + 2 min)
// per scheduler for 2 / 4 / 8 warps per scheduler.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_mix2 pipe_mix2.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 add2s(u64 x, float s)
{
    u64 r;
    asm volatile("{\n\t.reg .b64 t;\n\tmov.b64 t, {%2, %2};\n\tadd.rn.f32x2 %0, %1, t;\n\t}" : "=l"(r) : "l"(x), "f"(s));
    return r;
}
__device__ __forceinline__ u64 add2p(u64 x, u64 y)
{
    u64 r;
    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
    return r;
}
__device__ __forceinline__ float lo(u64 x) { return __uint_as_float((unsigned)x); }
__device__ __forceinline__ float hi(u64 x) { return __uint_as_float((unsigned)(x >> 32)); }
__device__ __forceinline__ float min3(float a, float b, float c)
{
    float r;
    asm volatile("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float min2(float a, float b)
{
    float r;
    asm volatile("min.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float fadd(float a, float b)
{
    float r;
    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ int imin2(int a, int b)
{
    int r;
    asm volatile("min.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(float *out, long long *cyc, int iters)
{
    constexpr int N = 8;  // independent chains per thread
    u64 p[N];
    float m[N], m2[N], s = threadIdx.x * 1e-3f;
    int mi[N], mi2[N];
    for (int i = 0; i < N; ++i) {
        p[i] = (u64)__float_as_uint(1.0f + i) << 32 | __float_as_uint(2.0f + i);
        m[i] = 3.0f + i;
        mi[i] = 1000 + i;
        m2[i] = 5.0f + i;
        mi2[i] = 2000 + i;
    }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if (MODE == 0) {  // baseline: FADD2 (scalar broadcast) + FMNMX3
                p[i] = add2s(p[i], s);
                m[i] = min3(m[i], lo(p[i]), hi(p[i]));
            } else if (MODE == 1) {  // 2 x FADD + FMNMX3
                float a = fadd(lo(p[i]), s), b = fadd(hi(p[i]), s);
                p[i] = (u64)__float_as_uint(b) << 32 | __float_as_uint(a);
                m[i] = min3(m[i], a, b);
            } else if (MODE == 2) {  // FADD2 + 2 x FMNMX
                p[i] = add2s(p[i], s);
                m[i] = min2(min2(m[i], lo(p[i])), hi(p[i]));
            } else if (MODE == 3) {  // FADD2 + 2 x IMNMX (non-negative floats compare like ints)
                p[i] = add2s(p[i], s);
                mi[i] = imin2(imin2(mi[i], (int)(unsigned)p[i]), (int)(unsigned)(p[i] >> 32));
            } else if (MODE == 4) {  // FADD2 + DPX 3-input integer min
                p[i] = add2s(p[i], s);
                mi[i] = __vimin3_s32(mi[i], (int)(unsigned)p[i], (int)(unsigned)(p[i] >> 32));
            } else if (MODE == 5) {  // FADD2 with a real pair operand + FMNMX3
                p[i] = add2p(p[i], p[(i + 3) % N]);
                m[i] = min3(m[i], lo(p[i]), hi(p[i]));
            } else if (MODE == 6) {  // FADD2 only (2 adds, no mins)
                p[i] = add2s(p[i], s);
            } else if (MODE == 7) {  // FMNMX3 only
                m[i] = min3(m[i], m[(i + 3) % N], s);
            } else if (MODE == 8) {  // DPX fused add+min: min(a + b, c), twice (integer arithmetic)
                mi[i] = __viaddmin_s32(mi[(i + 1) % N], (int)threadIdx.x, mi[i]);
                mi[i] = __viaddmin_s32(mi[(i + 2) % N], (int)threadIdx.x, mi[i]);
            } else if (MODE == 10) {  // FADD2 + two INDEPENDENT 2-input FMNMX chains (no FMNMX3 fusion possible)
                p[i] = add2s(p[i], s);
                m[i] = min2(m[i], lo(p[i]));
                m2[i] = min2(m2[i], hi(p[i]));
            } else if (MODE == 11) {  // FADD2 + two independent IMNMX chains
                p[i] = add2s(p[i], s);
                mi[i] = imin2(mi[i], (int)(unsigned)p[i]);
                mi2[i] = imin2(mi2[i], (int)(unsigned)(p[i] >> 32));
            } else if (MODE == 12) {  // two independent 2-input FMNMX only
                m[i] = min2(m[i], s);
                m2[i] = min2(m2[i], m[(i + 3) % N]);
            } else if (MODE == 13) {  // 2 FADD + two independent 2-input FMNMX
                float a = fadd(lo(p[i]), s), b = fadd(hi(p[i]), s);
                p[i] = (u64)__float_as_uint(b) << 32 | __float_as_uint(a);
                m[i] = min2(m[i], a);
                m2[i] = min2(m2[i], b);
            } else if (MODE == 14) {  // two independent IMNMX only
                mi[i] = imin2(mi[i], (int)threadIdx.x);
                mi2[i] = imin2(mi2[i], mi[(i + 3) % N]);
            } else if (MODE == 9) {  // 2 x FADD + 2 x FMNMX (all scalar, all 2-input)
                float a = fadd(lo(p[i]), s), b = fadd(hi(p[i]), s);
                p[i] = (u64)__float_as_uint(b) << 32 | __float_as_uint(a);
                m[i] = min2(min2(m[i], a), b);
            }
        }
    }
    long long t1 = clock64();
    float acc = 0;
    for (int i = 0; i < N; ++i) acc += m[i] + m2[i] + lo(p[i]) + hi(p[i]) + (float)mi[i] + (float)mi2[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MODE>
void run(const char *name, float *out, long long *cyc)
{
    const int iters = 2000;
    long long h[148];
    printf("%-44s", name);
    for (int threads : {256, 512, 1024}) {
        for (int rep = 0; rep < 2; ++rep) {
            k<MODE><<<148, threads>>>(out, cyc, iters);
            cudaDeviceSynchronize();
        }
        cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
        const double wps = threads / 32 / 4.0;                       // warps per scheduler
        const double per = (double)h[0] / (iters * 8.0 * wps);        // cycles per (2 add + 2 min) per scheduler
        printf("  %d w/s: %.2f cyc", (int)wps, per);
    }
    printf("\n");
}
int main()
{
    float *out;
    long long *cyc;
    cudaMalloc(&out, 148 * 1024 * 4);
    cudaMalloc(&cyc, 148 * 8);
    run<0>("FADD2(bcast) + FMNMX3   [kernel today]", out, cyc);
    run<1>("2 FADD + FMNMX3", out, cyc);
    run<2>("FADD2 + 2 FMNMX", out, cyc);
    run<3>("FADD2 + 2 IMNMX (int compare)", out, cyc);
    run<4>("FADD2 + VIMNMX3 (DPX int min3)", out, cyc);
    run<5>("FADD2(pair) + FMNMX3", out, cyc);
    run<9>("2 FADD + 2 FMNMX", out, cyc);
    run<10>("FADD2 + 2 independent FMNMX (2-input)", out, cyc);
    run<11>("FADD2 + 2 independent IMNMX", out, cyc);
    run<13>("2 FADD + 2 independent FMNMX", out, cyc);
    run<12>("2 independent FMNMX only", out, cyc);
    run<14>("2 independent IMNMX only", out, cyc);
    run<6>("FADD2 only", out, cyc);
    run<7>("FMNMX3 only", out, cyc);
    run<8>("2 x VIADDMNMX (DPX fused int add+min)", out, cyc);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
    return 0;
}
