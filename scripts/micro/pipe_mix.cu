// Microbenchmark: issue rate of the instruction mix of k_cn_fast on sm_100a (FADD2 = add.rn.f32x2 with a broadcast
// scalar, FMNMX3 = 3-input min) with 2 warps per scheduler, as the kernel runs.  Prints warp-instructions per cycle
// and scheduler for each mix.      nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_mix pipe_mix.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long add2s(unsigned long long x, float s)
{
    unsigned long long r;
    asm volatile("{\n\t.reg .b64 t;\n\tmov.b64 t, {%2, %2};\n\tadd.rn.f32x2 %0, %1, t;\n\t}" : "=l"(r) : "l"(x), "f"(s));
    return r;
}
__device__ __forceinline__ float min3(float a, float b, float c)
{
    float r;
    asm volatile("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float fadd(float a, float b)
{
    float r;
    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
template <int MODE>
__global__ void __launch_bounds__(256, 1) k(float *out, long long *cyc, int iters)
{
    constexpr int N = 16;  // independent chains
    unsigned long long p[N];
    float m[N], s = threadIdx.x * 1e-3f;
    for (int i = 0; i < N; ++i) {
        p[i] = (unsigned long long)__float_as_uint(1.0f + i) << 32 | __float_as_uint(2.0f + i);
        m[i] = 3.0f + i;
    }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if (MODE == 0 || MODE == 2) p[i] = add2s(p[i], s);
            if (MODE == 1 || MODE == 2) m[i] = min3(m[i], m[(i + 5) % N], s);
            if (MODE == 3) m[i] = fadd(m[i], s);
            if (MODE == 4) { m[i] = fadd(m[i], s); p[i] = add2s(p[i], s); }
        }
    }
    long long t1 = clock64();
    float acc = 0;
    for (int i = 0; i < N; ++i) acc += m[i] + __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32));
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main()
{
    float *out;
    long long *cyc, h[148];
    cudaMalloc(&out, 148 * 256 * 4);
    cudaMalloc(&cyc, 148 * 8);
    const int iters = 2000;
    const char *names[5] = {"FADD2 only", "FMNMX3 only", "FADD2 + FMNMX3 alternating", "FADD only", "FADD + FADD2 alternating"};
    const int per[5] = {16, 16, 32, 16, 32};
    for (int mode = 0; mode < 5; ++mode) {
        for (int rep = 0; rep < 2; ++rep) {
            switch (mode) {
            case 0: k<0><<<148, 256>>>(out, cyc, iters); break;
            case 1: k<1><<<148, 256>>>(out, cyc, iters); break;
            case 2: k<2><<<148, 256>>>(out, cyc, iters); break;
            case 3: k<3><<<148, 256>>>(out, cyc, iters); break;
            case 4: k<4><<<148, 256>>>(out, cyc, iters); break;
            }
            cudaDeviceSynchronize();
        }
        cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
        // 8 warps per SM = 2 per scheduler; each warp issues per[mode] * iters instructions
        double ipc = 2.0 * per[mode] * iters / (double)h[0];
        printf("%-32s %8lld cycles  %.3f warp-instr / cycle / scheduler (2 warps per scheduler)\n", names[mode], h[0], ipc);
    }
    return 0;
}
