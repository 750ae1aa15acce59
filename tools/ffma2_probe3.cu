// ffma2_probe3.cu — marginal cost (SMSP cycles) of one extra instruction X inserted per 4 FFMA2.
#include <cuda_runtime.h>
#include <stdio.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
// X: 0 none, 1 FMNMX3, 2 FMNMX, 3 FSETP(+pred accumulate), 4 LOP3, 5 IADD3, 6 FADD, 7 FFMA, 8 two FFMA, 9 FMUL2?, 10 LDS, 11 IMAD
template <int X>
__global__ void __launch_bounds__(256, 2) kx(const float* __restrict__ in, float* out, int iters) {
    __shared__ float sm[256];
    sm[threadIdx.x] = in[threadIdx.x];
    __syncthreads();
    u64 acc[8], x[8];
    for (int i = 0; i < 8; ++i) { acc[i] = pk(in[threadIdx.x + i], in[threadIdx.x + 8 + i]); x[i] = pk(in[threadIdx.x + 16 + i], in[threadIdx.x + 24 + i]); }
    float s0 = in[threadIdx.x + 50];
    float m = -3e38f, f0 = in[threadIdx.x + 60], f1 = in[threadIdx.x + 61];
    unsigned im = __float_as_uint(in[threadIdx.x + 62]); int cnt = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float2 a = *reinterpret_cast<float2*>(&acc[i]), xi = *reinterpret_cast<float2*>(&x[i]);
                float2 r = __ffma2_rn(xi, make_float2(s0, s0), a);
                acc[i] = *reinterpret_cast<u64*>(&r);
                if ((i & 3) == 3) {
                    if (X == 1) { float t; asm volatile("max.f32 %0, %1, %2, %3;" : "=f"(t) : "f"(m), "f"(r.x), "f"(r.y)); m = t; }
                    if (X == 2) { float t; asm volatile("max.f32 %0, %1, %2;" : "=f"(t) : "f"(m), "f"(r.x)); m = t; }
                    if (X == 3) { asm volatile("{ .reg .pred p; setp.ge.f32 p, %1, %2; @p add.s32 %0, %0, 1; }" : "+r"(cnt) : "f"(r.x), "f"(f0)); }
                    if (X == 4) { unsigned t; asm volatile("lop3.b32 %0, %1, %2, %3, 0x80;" : "=r"(t) : "r"(im), "r"(__float_as_uint(r.x)), "r"(__float_as_uint(r.y))); im = t; }
                    if (X == 5) { asm volatile("add.s32 %0, %0, %1;" : "+r"(cnt) : "r"(__float_as_int(r.x))); }
                    if (X == 6) { asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f0) : "f"(r.x)); }
                    if (X == 7) { asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f0) : "f"(r.x), "f"(f1)); }
                    if (X == 8) { asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f0) : "f"(r.x), "f"(f1)); asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f1) : "f"(r.y), "f"(s0)); }
                    if (X == 10) { float t; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(t) : "r"((unsigned)__cvta_generic_to_shared(&sm[(i + u) & 255]))); f0 += t * 0.f; }
                    if (X == 11) { asm volatile("mad.lo.s32 %0, %1, %2, %0;" : "+r"(cnt) : "r"(__float_as_int(r.x)), "r"((int)im)); }
                }
            }
        }
    }
    float s = m + f0 + f1 + cnt + im;
    for (int i = 0; i < 8; ++i) { float2 a = *reinterpret_cast<float2*>(&acc[i]); s += a.x + a.y; }
    if (s == 12345.678f) out[0] = s;
}
template <class K>
static double run(const char* name, K kern, const float* in, float* out, double base) {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 16384, grid = sms * 2;
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0); kern<<<grid, 256>>>(in, out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double groups = (double)iters * 8 * 4;           // groups of 4 FFMA2 per warp, x4 warps/SMSP
    const double cyc = best * 1e-3 * 1.965e9 / groups;
    printf("%-34s %.2f cyc per [4 FFMA2 + X]   X costs %.2f cyc\n", name, cyc, cyc - base);
    return cyc;
}
int main() {
    float h[512]; for (int i = 0; i < 512; ++i) h[i] = 1.0f + 1e-6f * (i % 13) - 2e-6f * (i % 7);
    float* in; cudaMalloc(&in, sizeof(h)); cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
    float* out; cudaMalloc(&out, 64);
    double b = run("X = none", kx<0>, in, out, 0); 
    run("X = FMNMX3", kx<1>, in, out, b); run("X = FMNMX", kx<2>, in, out, b); run("X = FSETP + @p IADD", kx<3>, in, out, b);
    run("X = LOP3", kx<4>, in, out, b); run("X = IADD3", kx<5>, in, out, b); run("X = FADD", kx<6>, in, out, b);
    run("X = FFMA", kx<7>, in, out, b); run("X = 2 x FFMA", kx<8>, in, out, b); run("X = LDS", kx<10>, in, out, b); run("X = IMAD", kx<11>, in, out, b);
    return 0;
}
