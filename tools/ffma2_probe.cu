// ffma2_probe.cu — micro-benchmark: FFMA2 issue rate as a function of where its operands come from
// (register pairs, 32-bit broadcast register `R.F32`, uniform register `UR.F32` fed by LDCU from constant memory).
// Decides how k_isect2 feeds the warp-uniform sphere operand.  Build & run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2_probe tools/ffma2_probe.cu && /tmp/ffma2_probe
#include <cuda_runtime.h>
#include <stdio.h>

__constant__ float4 c_sph[2048];

// Shape of the k_isect2 pivot scan: per sphere 4 ray pairs x (3 FFMA2 chain + 1 square) = 16 FFMA2.
//   V0  sphere operand = register PAIR (x,x) from shared memory          (LDS.128 x2 per sphere)
//   V1  sphere operand = 32-bit register, broadcast form R.F32            (LDS.128 x1 per sphere)
//   V2  sphere operand = uniform register UR.F32 via LDCU from c[3]       (no LDS, no RF read)
//   V3  V1 without any loads (operands loop-invariant registers): pure issue/RF limit
template <int V>
__global__ void __launch_bounds__(256, 2) kscan(const float4* __restrict__ sph, const float2* __restrict__ rays, float* out, int n, int reps) {
    extern __shared__ float4 s_sph[];
    for (int i = threadIdx.x; i < 2 * n; i += 256) s_sph[i] = sph[i];
    __syncthreads();
    float2 dx[4], dy[4], dz[4], nod[4];
    for (int p = 0; p < 4; ++p) {
        dx[p] = rays[threadIdx.x * 16 + p]; dy[p] = rays[threadIdx.x * 16 + 4 + p];
        dz[p] = rays[threadIdx.x * 16 + 8 + p]; nod[p] = rays[threadIdx.x * 16 + 12 + p];
    }
    float m = -3e38f;
    for (int rep = 0; rep < reps; ++rep) {
#pragma unroll 4
        for (int s = 0; s < n; ++s) {
            float2 ax, ay, az, aw;
            if (V == 0) {
                const float4 u = s_sph[2 * s], w = s_sph[2 * s + 1];
                ax = make_float2(u.x, u.y); ay = make_float2(u.z, u.w); az = make_float2(w.x, w.y); aw = make_float2(w.z, w.w);
            } else if (V == 1) {
                const float4 a = s_sph[s];
                ax = make_float2(a.x, a.x); ay = make_float2(a.y, a.y); az = make_float2(a.z, a.z); aw = make_float2(a.w, a.w);
            } else if (V == 2) {
                const float4 a = c_sph[s];
                ax = make_float2(a.x, a.x); ay = make_float2(a.y, a.y); az = make_float2(a.z, a.z); aw = make_float2(a.w, a.w);
            } else {
                const float4 a = s_sph[0];
                ax = make_float2(a.x, a.x); ay = make_float2(a.y, a.y); az = make_float2(a.z, a.z); aw = make_float2(a.w + m * 1e-30f, a.w);
            }
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const float2 b = __ffma2_rn(ax, dx[p], __ffma2_rn(ay, dy[p], __ffma2_rn(az, dz[p], nod[p])));
                const float2 q = __ffma2_rn(b, b, aw);
                float t; asm("max.f32 %0, %1, %2, %3;" : "=f"(t) : "f"(m), "f"(q.x), "f"(q.y)); m = t;
            }
        }
    }
    if (m == 12345.678f) out[0] = m;
}


// Pair-major order: for one ray pair, K=4 spheres back to back, so the RAY operands (dz,nod / dy / dx) sit in the
// operand-reuse cache and the sphere operands come from uniform registers (V=4, LDCU) or 32-bit registers (V=5, LDS).
template <int V>
__global__ void __launch_bounds__(256, 2) kpm(const float4* __restrict__ sph, const float2* __restrict__ rays, float* out, int n, int reps) {
    extern __shared__ float4 s_sph[];
    for (int i = threadIdx.x; i < 2 * n; i += 256) s_sph[i] = sph[i];
    __syncthreads();
    float2 dx[4], dy[4], dz[4], nod[4];
    for (int p = 0; p < 4; ++p) {
        dx[p] = rays[threadIdx.x * 16 + p]; dy[p] = rays[threadIdx.x * 16 + 4 + p];
        dz[p] = rays[threadIdx.x * 16 + 8 + p]; nod[p] = rays[threadIdx.x * 16 + 12 + p];
    }
    float m = -3e38f;
    for (int rep = 0; rep < reps; ++rep) {
#pragma unroll 1
        for (int s = 0; s < n; s += 4) {
            float4 a[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) a[k] = (V == 4) ? c_sph[s + k] : s_sph[s + k];
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                float2 t[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) t[k] = __ffma2_rn(dz[p], make_float2(a[k].z, a[k].z), nod[p]);
#pragma unroll
                for (int k = 0; k < 4; ++k) t[k] = __ffma2_rn(dy[p], make_float2(a[k].y, a[k].y), t[k]);
#pragma unroll
                for (int k = 0; k < 4; ++k) t[k] = __ffma2_rn(dx[p], make_float2(a[k].x, a[k].x), t[k]);
#pragma unroll
                for (int k = 0; k < 4; ++k) t[k] = __ffma2_rn(t[k], t[k], make_float2(a[k].w, a[k].w));
#pragma unroll
                for (int k = 0; k < 4; ++k) { float r; asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(m), "f"(t[k].x), "f"(t[k].y)); m = r; }
            }
        }
    }
    if (m == 12345.678f) out[0] = m;
}


// V6/V7: lanes of every packed value are two SPHERES (constant-memory pairs -> uniform register pairs, no RF read);
// the ray operands are 32-bit broadcast registers.  Ray-major order over KP sphere pairs.
__constant__ float4 c_pair[2048];     // [2*j] = (cx0,cx1,cy0,cy1), [2*j+1] = (cz0,cz1,k0,k1)
template <int KP>
__global__ void __launch_bounds__(256, 2) ksp(const float* __restrict__ rays, float* out, int n, int reps) {
    float dx[8], dy[8], dz[8], nod[8];
    for (int r = 0; r < 8; ++r) {
        dx[r] = rays[threadIdx.x * 32 + r]; dy[r] = rays[threadIdx.x * 32 + 8 + r];
        dz[r] = rays[threadIdx.x * 32 + 16 + r]; nod[r] = rays[threadIdx.x * 32 + 24 + r];
    }
    float m = -3e38f;
    for (int rep = 0; rep < reps; ++rep) {
#pragma unroll 1
        for (int j = 0; j < n / 2; j += KP) {
            float2 cx[KP], cy[KP], cz[KP], kw[KP];
#pragma unroll
            for (int k = 0; k < KP; ++k) {
                const float4 u = c_pair[2 * (j + k)], w = c_pair[2 * (j + k) + 1];
                cx[k] = make_float2(u.x, u.y); cy[k] = make_float2(u.z, u.w); cz[k] = make_float2(w.x, w.y); kw[k] = make_float2(w.z, w.w);
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                float2 t[KP];
#pragma unroll
                for (int k = 0; k < KP; ++k) t[k] = __ffma2_rn(cz[k], make_float2(dz[r], dz[r]), make_float2(nod[r], nod[r]));
#pragma unroll
                for (int k = 0; k < KP; ++k) t[k] = __ffma2_rn(cy[k], make_float2(dy[r], dy[r]), t[k]);
#pragma unroll
                for (int k = 0; k < KP; ++k) t[k] = __ffma2_rn(cx[k], make_float2(dx[r], dx[r]), t[k]);
#pragma unroll
                for (int k = 0; k < KP; ++k) t[k] = __ffma2_rn(t[k], t[k], kw[k]);
#pragma unroll
                for (int k = 0; k < KP; ++k) { float q; asm("max.f32 %0, %1, %2, %3;" : "=f"(q) : "f"(m), "f"(t[k].x), "f"(t[k].y)); m = q; }
            }
        }
    }
    if (m == 12345.678f) out[0] = m;
}
template <class K>
static void run2(const char* name, K kern, const float* rays, float* out) {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int n = 1024, reps = 64, grid = sms * 2;
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
        cudaEventRecord(e0); kern<<<grid, 256>>>(rays, out, n, reps); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    cudaError_t e = cudaGetLastError();
    const double ffma2_per_warp = (double)n * reps * 16;
    const double cyc = best * 1e-3 * 1.965e9;
    printf("%-44s %.3f ms  %.2f cyc/FFMA2/SMSP (4 warps/SMSP)  %.1f TFLOP/s  %s\n", name, best, cyc / (ffma2_per_warp * 4),
           4.0 * 32 * ffma2_per_warp * (grid * 8.0) / (best * 1e-3) / 1e12, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

template <class K>
static void run(const char* name, K kern, size_t smem, const float4* sph, const float2* rays, float* out) {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int n = 1024, reps = 64, grid = sms * 2;
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
        cudaEventRecord(e0); kern<<<grid, 256, smem>>>(sph, rays, out, n, reps); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    cudaError_t e = cudaGetLastError();
    const double ffma2_per_warp = (double)n * reps * 16;
    const double cyc = best * 1e-3 * 1.965e9;
    printf("%-44s %.3f ms  %.2f cyc/FFMA2/SMSP (4 warps/SMSP)  %.1f TFLOP/s  %s\n", name, best, cyc / (ffma2_per_warp * 4),
           4.0 * 32 * ffma2_per_warp * (grid * 8.0) / (best * 1e-3) / 1e12, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main() {
    const int n = 1024;
    float4* h = new float4[2 * n];
    for (int i = 0; i < 2 * n; ++i) h[i] = make_float4(0.001f * i, 0.5f, -0.25f + 1e-4f * i, 0.125f);
    float4* sph; cudaMalloc(&sph, sizeof(float4) * 2 * n); cudaMemcpy(sph, h, sizeof(float4) * 2 * n, cudaMemcpyHostToDevice);
    cudaMemcpyToSymbol(c_sph, h, sizeof(float4) * n);
    float2* rays; cudaMalloc(&rays, sizeof(float2) * 256 * 16);
    float2* hr = new float2[256 * 16];
    for (int i = 0; i < 256 * 16; ++i) hr[i] = make_float2(1e-3f * (i % 97), -1e-3f * (i % 89));
    cudaMemcpy(rays, hr, sizeof(float2) * 256 * 16, cudaMemcpyHostToDevice);
    float* out; cudaMalloc(&out, 64);
    const size_t smem = sizeof(float4) * 2 * n;
    run("V0 sphere = register pair (LDS.128 x2)", kscan<0>, smem, sph, rays, out);
    run("V1 sphere = R.F32 broadcast (LDS.128 x1)", kscan<1>, smem, sph, rays, out);
    run("V2 sphere = UR.F32 via LDCU (constant mem)", kscan<2>, smem, sph, rays, out);
    run("V4 pair-major, sphere = UR.F32 via LDCU", kpm<4>, smem, sph, rays, out);
    run("V5 pair-major, sphere = R.F32 via LDS", kpm<5>, smem, sph, rays, out);
    cudaMemcpyToSymbol(c_pair, h, sizeof(float4) * n);
    float* rf; cudaMalloc(&rf, sizeof(float) * 256 * 32); cudaMemcpy(rf, hr, sizeof(float) * 256 * 32, cudaMemcpyHostToDevice);
    run2("V6 lanes = 2 spheres (UR pairs), KP=2", ksp<2>, rf, out);
    run2("V7 lanes = 2 spheres (UR pairs), KP=4", ksp<4>, rf, out);
    return 0;
}
