// b200rt.cu — hand-written CUDA (sm_100a) implementation of the python-ray-tracer `render` hot path and its C ABI.
//
// Design (DESIGN.md has the long form): a device-side WAVEFRONT.  One frame tile is a set of independent
// paths (1 or 9 samples per pixel).  Each reflection depth runs
//     k_closest   brute-force closest hit of every live ray         (reference trace.py:7-41, called at :53)
//     k_setup     hit point, normal, bias; compacts the hit list    (trace.py:56-83)
//     k_shadow    any-hit query of (hit, light) pairs               (trace.py:86-96)
//     k_shade     Lambert terms, weighted accumulate, reflection    (trace.py:99-110, :123-131)
// over dense, compacted ray queues, so the two intersection kernels — >99 % of the work at S=1024 — run with
// every lane busy regardless of how many paths have died.  k_raygen (kernels.py:16-23,32-55) feeds the first
// queue, k_resolve (kernels.py:58-73) sums the samples in reference order, quantises and writes the framebuffer
// with vector stores.  Queue lengths live in device memory; nothing synchronises with the host.
//
// Numerics: FP32 for intersection and shading (IEEE div/sqrt, explicit FMAs — rt_math.cuh); primary-ray
// generation in FP64 from the reference's float64 pixel grid, rounded once to FP32.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <mutex>
#include <vector>
#include <string>
#include <algorithm>
#include "../../include/b200rt.h"
#include "rt_math.cuh"
#include "isect3.cuh"
#include "shadow_grid.cuh"
#include "closest_grid.cuh"
#include "fused_small.cuh"

namespace b200rt {

// ------------------------------------------------------------------------------------------------
// Resident scene.  Everything that depends only on the scene arrays (packed float4 tables, the lights' solid-angle orders,
// the Morton order and cluster bounds of the closest-hit tables, the pivot tables of the lights and the origin-free table, the
// lights' direction grids, the bounding box of the regrouping) is kept on the device between calls and rebuilt only when
// the scene changes.  "Changes" is decided ON THE DEVICE, so it holds for device-resident scene arrays too and the launch
// sequence of a frame is the same whether or not anything is rebuilt (it can be captured in a CUDA graph): k_scene_cmp
// compares the caller's arrays bit for bit with the copy of the arrays the tables were built from and bumps a version
// counter when they differ; every builder kernel returns at once when its table group already carries that version, and
// k_cache_commit stamps the groups built in this call.  Only the camera pivot table (k_pivot, j == L) is per frame.
//   cs[0] scene version   cs[1..3] version the groups {base tables, light grids, regrouping box} were built from
//   cs[4] key of the regrouping box (position bits, direction cells)   cs[5..7] S, L, Pl of the saved copy   cs[8] CTA ticket
//   cs[9] version the closest-hit grid (closest_grid.cuh) was built from   cs[10] version of the packed arrays (k_pack; the small-scene path builds nothing else)
// ------------------------------------------------------------------------------------------------
enum { kCsVersion = 0, kCsBase = 1, kCsSgrid = 2, kCsGrid = 3, kCsGridKey = 4, kCsS = 5, kCsL = 6, kCsPl = 7, kCsTicket = 8, kCsCgrid = 9, kCsPack = 10, kCsWords = 16 };

__global__ void __launch_bounds__(256) k_scene_cmp(const float* __restrict__ spheres, int S, const float* __restrict__ lights, int L,
                                                  const float* __restrict__ planes, int Pl, uint32_t* saved, uint32_t* cs, float* cmax) {
    __shared__ int s_diff, s_last;
    const int nS = 7 * S, nL = 3 * L, n = nS + nL + 9 * Pl;
    if (threadIdx.x == 0) s_diff = (cs[kCsS] != (uint32_t)S || cs[kCsL] != (uint32_t)L || cs[kCsPl] != (uint32_t)Pl) ? 1 : 0;
    __syncthreads();
    bool diff = false;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float v = i < nS ? spheres[i] : (i < nS + nL ? lights[i - nS] : planes[i - nS - nL]);
        const uint32_t b = __float_as_uint(v);
        if (saved[i] != b) { saved[i] = b; diff = true; }            // compare and update in one pass (bit patterns: NaNs and -0 count)
    }
    if (__syncthreads_or(diff ? 1 : 0) && threadIdx.x == 0) s_diff = 1;
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_diff) atomicOr(&cs[kCsTicket], 0x80000000u);
        __threadfence();
        const uint32_t t = atomicAdd(&cs[kCsTicket], 1u);
        s_last = ((t & 0x7fffffffu) == gridDim.x - 1) ? 1 : 0;
        if (s_last) {                                                 // last CTA: publish
            __threadfence();
            const uint32_t all = atomicExch(&cs[kCsTicket], 0u);
            if (all & 0x80000000u) { cs[kCsVersion] += 1u; cs[kCsS] = (uint32_t)S; cs[kCsL] = (uint32_t)L; cs[kCsPl] = (uint32_t)Pl; }
            // k_pack's atomicMax bounds start from zero when (and only when) the base tables are about to be rebuilt
            if (cs[kCsPack] != cs[kCsVersion]) { cmax[0] = 0.f; cmax[1] = 0.f; }
        }
    }
}
__global__ void k_cache_commit(uint32_t* cs, int base, int sgrid, int grid, uint32_t gridkey, int cgrid) {
    const uint32_t v = cs[kCsVersion];
    cs[kCsPack] = v;                                             // k_pack runs in every call
    if (cgrid) cs[kCsCgrid] = v;
    if (base) cs[kCsBase] = v;
    if (sgrid) cs[kCsSgrid] = v;
    if (grid) { cs[kCsGrid] = v; cs[kCsGridKey] = gridkey; }
}
__device__ __forceinline__ bool cache_fresh(const uint32_t* cs, int group) { return cs[group] == cs[kCsVersion]; }

// (7,S) / (9,Pl) / (3,L) float32 SoA  ->  packed float4 arrays.  One thread per object.
struct PackOut { float4 *sphA, *sphB, *sphCol, *plO, *plN, *plNu, *plCol, *lightsOut; float* cmax; };
__device__ __forceinline__ void pack_item(int i, const float* __restrict__ spheres, int S, int Spad, const float* __restrict__ planes, int Pl,
                                          const float* __restrict__ lights, int L, const PackOut& o) {
    float4 *sphA = o.sphA, *sphB = o.sphB, *sphCol = o.sphCol, *plO = o.plO, *plN = o.plN, *plNu = o.plNu, *plCol = o.plCol, *lightsOut = o.lightsOut; float* cmax = o.cmax;
    if (i < Spad) {
        if (i < S) {
            float cx = spheres[0 * S + i], cy = spheres[1 * S + i], cz = spheres[2 * S + i], r = spheres[3 * S + i];
            float r2 = __fmul_rn(r, r);                                   // float32 product, intersections.py:21
            double cc = (double)cx * cx + (double)cy * cy + (double)cz * cz;
            sphA[i] = make_float4(cx, cy, cz, (float)((double)r2 - cc));
            sphB[i] = make_float4(cx, cy, cz, r2);
            sphCol[i] = make_float4(spheres[4 * S + i], spheres[5 * S + i], spheres[6 * S + i], 0.f);
            // scene bound for the filter margin: max |C| (atomicMax on the bit pattern of a non-negative float)
            // (a sphere with a NaN / infinite coordinate or radius is never hit by any path, like in the reference — intersections.py:25-36 with NaN —
            // and must not poison the bounds every other sphere's margins are computed from)
            if (isfinite(cc) && isfinite(r)) {
                atomicMax(reinterpret_cast<int*>(cmax), __float_as_int((float)sqrt(cc) * 1.0000002f));
                atomicMax(reinterpret_cast<int*>(cmax) + 1, __float_as_int(fabsf(r)));      // bounds[1] = max r
            }
        } else {
            sphA[i] = make_float4(0.f, 0.f, 0.f, -3.0e38f);               // padding: q = -3e38 never reaches a threshold
        }
    }
    if (i < Pl) {
        plO[i] = make_float4(planes[0 * Pl + i], planes[1 * Pl + i], planes[2 * Pl + i], 0.f);
        float nx = planes[3 * Pl + i], ny = planes[4 * Pl + i], nz = planes[5 * Pl + i];
        plN[i] = make_float4(nx, ny, nz, 0.f);
        // common.py:104-110 + :28-32 evaluated in float32, left to right, no contraction
        float nn = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(nx, nx), __fmul_rn(ny, ny)), __fmul_rn(nz, nz)));
        plNu[i] = make_float4(__fdiv_rn(nx, nn), __fdiv_rn(ny, nn), __fdiv_rn(nz, nn), 0.f);
        plCol[i] = make_float4(planes[6 * Pl + i], planes[7 * Pl + i], planes[8 * Pl + i], 0.f);
    }
    if (i < L) lightsOut[i] = make_float4(lights[0 * L + i], lights[1 * L + i], lights[2 * L + i], 0.f);}
__global__ void k_pack(const float* __restrict__ spheres, int S, int Spad,
                       const float* __restrict__ planes, int Pl,
                       const float* __restrict__ lights, int L,
                       float4* sphA, float4* sphB, float4* sphCol,
                       float4* plO, float4* plN, float4* plNu, float4* plCol, float4* lightsOut, float* cmax, const uint32_t* cs) {
    if (cache_fresh(cs, kCsPack)) return;
    pack_item(blockIdx.x * blockDim.x + threadIdx.x, spheres, S, Spad, planes, Pl, lights, L, PackOut{ sphA, sphB, sphCol, plO, plN, plNu, plCol, lightsOut, cmax });
}

// Small scenes (the fused path needs nothing but the packed arrays): k_scene_cmp + k_pack + k_cache_commit as ONE single-CTA launch.
__global__ void __launch_bounds__(256) k_small_prep(const float* __restrict__ spheres, int S, int Spad, const float* __restrict__ lights, int L,
                                                   const float* __restrict__ planes, int Pl, uint32_t* saved, uint32_t* cs, PackOut o) {
    __shared__ int s_diff;
    const int nS = 7 * S, nL = 3 * L, n = nS + nL + 9 * Pl;
    if (threadIdx.x == 0) s_diff = (cs[kCsS] != (uint32_t)S || cs[kCsL] != (uint32_t)L || cs[kCsPl] != (uint32_t)Pl) ? 1 : 0;
    __syncthreads();
    bool diff = false;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float v = i < nS ? spheres[i] : (i < nS + nL ? lights[i - nS] : planes[i - nS - nL]);
        const uint32_t b = __float_as_uint(v);
        if (saved[i] != b) { saved[i] = b; diff = true; }
    }
    if (__syncthreads_or(diff ? 1 : 0) && threadIdx.x == 0) s_diff = 1;
    __syncthreads();
    if (threadIdx.x == 0 && s_diff) { cs[kCsVersion] += 1u; cs[kCsS] = (uint32_t)S; cs[kCsL] = (uint32_t)L; cs[kCsPl] = (uint32_t)Pl; }
    __syncthreads();
    if (cs[kCsPack] == cs[kCsVersion]) return;                                   // (uniform: every thread reads the same words after the barrier)
    if (threadIdx.x == 0) { o.cmax[0] = 0.f; o.cmax[1] = 0.f; }
    __syncthreads();
    const int nobj = max(max(Spad, Pl), L);
    for (int i = threadIdx.x; i < nobj; i += blockDim.x) pack_item(i, spheres, S, Spad, planes, Pl, lights, L, o);
    __syncthreads();
    if (threadIdx.x == 0) cs[kCsPack] = cs[kCsVersion];
}

// ------------------------------------------------------------------------------------------------
// Frame / tile description shared by k_raygen and k_resolve
// ------------------------------------------------------------------------------------------------
// Per-frame camera: read through a device pointer so that a captured launch sequence (CUDA graph) can be replayed for a new
// camera pose by updating this block alone.
struct CamDev {
    double gf, gar, gsy, gsz;      // affine grid (camera.py:18-26)
    double ox, oy, oz;             // camera origin
    double rot[9];                 // camera rotation, row-major
    float cox, coy, coz, pad;      // camera origin rounded to float32 (pivot of the primary rays)
};
struct FrameDev {
    const double* pixel_loc;       // (3,W,H) or nullptr
    const CamDev* cam;
    int W, H;
    int x0, x1;                    // chunk LOCAL columns [x0, x1) (local = index among the columns this call renders)
    int xb;                        // tile origin: global x of local column 0
    int sw, srank, sworld;         // interleaved strips (sw == 0: contiguous tile)
    int ns;                        // samples per pixel slot: 1 or 9
    uint32_t npix;                 // (x1-x0)*H
    int tiled;                     // pixel slots in 16x16 tiles (H and the chunk's column count are multiples of 16), else column-major
};

// local column -> global x (array axis 1)
__device__ __forceinline__ int col_x(const FrameDev& f, int lc) {
    if (f.sw <= 0) return f.xb + lc;
    const int s = lc / f.sw;
    return f.xb + (s * f.sworld + f.srank) * f.sw + (lc - s * f.sw);
}

// Pixel slot p of the chunk -> (local column, y).  Tiled order: a ray group of 256 consecutive slots is a 16x16 pixel
// tile, i.e. a narrow beam (the filter's candidates are shared by the whole group and k_isect3's beam culling is tight).
__device__ __forceinline__ void slot_pixel(const FrameDev& f, uint32_t p, int& lc, int& y) {
    if (f.tiled) {
        const uint32_t tile = p >> 8, in = p & 255u, tiles_y = (uint32_t)f.H >> 4;
        const uint32_t tx = tile / tiles_y, ty = tile - tx * tiles_y;
        lc = (int)(tx * 16u + (in >> 4)); y = (int)(ty * 16u + (in & 15u));
    } else { lc = (int)(p / (uint32_t)f.H); y = (int)(p % (uint32_t)f.H); }
}

// pixel_loc[c, x, y] with the GPU's unchecked C-order flat addressing: elements past the end
// of the buffer are defined as 0.0.
__device__ __forceinline__ double pl_at(const FrameDev& f, const CamDev& cm, int c, int x, int y) {
    if (!f.pixel_loc && x >= 0 && x < f.W && y >= 0 && y < f.H) {           // in-range element of the closed-form grid: no 64-bit divisions
        if (c == 0) return cm.gf;
        if (c == 1) return __dadd_rn(__dmul_rn((double)x, cm.gsy), cm.gar);
        return __dadd_rn(__dmul_rn((double)y, cm.gsz), 1.0);
    }
    const long long W = f.W, H = f.H;
    long long flat = (long long)c * W * H + (long long)x * H + y;
    if (flat < 0 || flat >= 3 * W * H) return 0.0;
    if (f.pixel_loc) return f.pixel_loc[flat];
    int cc = (int)(flat / (W * H)); long long rem = flat - (long long)cc * W * H;
    int xx = (int)(rem / H), yy = (int)(rem - (long long)xx * H);
    if (cc == 0) return cm.gf;
    if (cc == 1) return __dadd_rn(__dmul_rn((double)xx, cm.gsy), cm.gar);
    return __dadd_rn(__dmul_rn((double)yy, cm.gsz), 1.0);
}


// AA neighbour order of kernels.py:32-53: left, right, top(y+1), bot(y-1), topleft, topright, bottomleft, bottomright
__constant__ int c_dx[9] = { 0, -1, +1, 0, 0, -1, +1, -1, +1 };
__constant__ int c_dy[9] = { 0, 0, 0, +1, -1, +1, +1, -1, -1 };

// Primary ray of sample slot `slot` = (sample k, pixel p): direction in FP64 (kernels.py:19-23 / :43-55) rounded once to FP32, slot bits in .w.
// False for the slots the reference never samples (AA border: x == 0 or y == 0, kernels.py:29).
__device__ __forceinline__ bool primary_dir(const FrameDev& f, const CamDev& cm, uint32_t slot, float4& d) {
    const uint32_t k = slot / f.npix, p = slot - k * f.npix;
    int lc, y; slot_pixel(f, p, lc, y);
    const int x = col_x(f, f.x0 + lc);
    if (k > 0 && !(x >= 1 && y >= 1)) return false;                       // kernels.py:29
    double px = pl_at(f, cm, 0, x, y), py = pl_at(f, cm, 1, x, y), pz = pl_at(f, cm, 2, x, y);
    if (k > 0) {                                                          // kernels.py:43-50  0.5*P + 0.5*P'
        const int xn = x + c_dx[k], yn = y + c_dy[k];
        px = __dadd_rn(__dmul_rn(0.5, px), __dmul_rn(0.5, pl_at(f, cm, 0, xn, yn)));
        py = __dadd_rn(__dmul_rn(0.5, py), __dmul_rn(0.5, pl_at(f, cm, 1, xn, yn)));
        pz = __dadd_rn(__dmul_rn(0.5, pz), __dmul_rn(0.5, pl_at(f, cm, 2, xn, yn)));
    }
    // matmul (common.py:40-49) then normalize (common.py:28-32), float64
    const double bx = cm.rot[0] * px + cm.rot[1] * py + cm.rot[2] * pz;
    const double by = cm.rot[3] * px + cm.rot[4] * py + cm.rot[5] * pz;
    const double bz = cm.rot[6] * px + cm.rot[7] * py + cm.rot[8] * pz;
    const double n = sqrt(bx * bx + by * by + bz * bz);
    d = make_float4((float)(bx / n), (float)(by / n), (float)(bz / n), __uint_as_float(slot));     // origin = the camera (CamDev.cox..coz = (float)ox..oz): not stored
    return true;
}

// One thread per (sample k, pixel p) slot.  Primary ray in FP64 (kernels.py:19-23 / :43-55), rounded to FP32,
// appended to the ray queue.  Slots that the reference never samples (AA border: x == 0 or y == 0) are skipped.
// Queue position: without anti-aliasing every slot has a ray, so position = slot (the queue is in slot order by
// construction); with it, border slots are skipped and the queue is compacted IN ORDER (ordered_append8: chunks of
// 256 slots in ticket order) — the queue order decides which rays share a group in k_isect3 at every later stage.
__global__ void k_raygen(FrameDev f, float4* __restrict__ rayD, unsigned long long* __restrict__ hit,
                         uint32_t* n_live, unsigned long long* ostat, uint32_t* ticket) {
    const uint32_t nslots = f.npix * (uint32_t)f.ns;
    const uint32_t nchunks = (nslots + 255u) / 256u;
    __shared__ uint32_t s_chunk;
    __shared__ CamDev s_cam;
    if (threadIdx.x < sizeof(CamDev) / 8) reinterpret_cast<double*>(&s_cam)[threadIdx.x] = reinterpret_cast<const double*>(f.cam)[threadIdx.x];
    __syncthreads();
    const CamDev& cm = s_cam;
    for (uint32_t it = blockIdx.x;; it += gridDim.x) {
        uint32_t c = it;
        if (f.ns > 1) {
            if (threadIdx.x == 0) s_chunk = atomicAdd(ticket, 1u);
            __syncthreads();
            c = s_chunk;
        }
        if (c >= nchunks) break;
        uint32_t slot = c * 256u + threadIdx.x;
        float4 d = make_float4(0, 0, 0, 0);
        const bool valid = slot < nslots && primary_dir(f, cm, slot, d);
        uint32_t pos = slot;
        if (f.ns > 1) { uint32_t pp[8]; ordered_append8(valid ? 1u : 0u, 1, c, ostat, n_live, c == nchunks - 1, pp); pos = pp[0]; }
        else if (slot == 0) *n_live = nslots;
        if (valid) { rayD[pos] = d; if (hit) hit[pos] = kNoHitKey; }
    }
}

// Small scenes (S < 64): one kernel per depth does the whole `trace` of every live path (fused_small.cuh); depth 0 also generates the ray.
__global__ void __launch_bounds__(256) k_fused_primary(SceneDev sc, FusedArgs a, FrameDev f) {
    __shared__ CamDev s_cam;
    if (threadIdx.x < sizeof(CamDev) / 8) reinterpret_cast<double*>(&s_cam)[threadIdx.x] = reinterpret_cast<const double*>(f.cam)[threadIdx.x];
    __syncthreads();
    if (a.tl && threadIdx.x == 0) atomicMin(a.tl, globaltimer_ns());
    fused_depth_body<true>(sc, a, f.npix * (uint32_t)f.ns, [&](uint32_t slot, V3& D) { float4 d; const bool v = primary_dir(f, s_cam, slot, d); D = mk(d.x, d.y, d.z); return v; });
    if (a.tl && threadIdx.x == 0) atomicMax(a.tl + 1, globaltimer_ns());
}
__global__ void __launch_bounds__(256) k_fused_bounce(SceneDev sc, FusedArgs a) {
    if (a.tl && threadIdx.x == 0) atomicMin(a.tl, globaltimer_ns());
    fused_depth_body<false>(sc, a, *a.n_in, [](uint32_t, V3&) { return false; });
    if (a.tl && threadIdx.x == 0) atomicMax(a.tl + 1, globaltimer_ns());
}

// ------------------------------------------------------------------------------------------------
// The hot kernel: brute-force intersection of a dense ray queue against all spheres and planes.
// ------------------------------------------------------------------------------------------------
constexpr int kRays    = 4;              // rays per thread (ILP; one broadcast LDS.128 feeds 4 tests)
constexpr int kGroup   = kThreads * kRays;
constexpr int kChunkS  = 4096;           // spheres staged in shared memory per pass (64 KB + 16 KB)


// Debug / proof kernel (B200RT_F_NO_FILTER): sphere_exact() on EVERY sphere for every ray, no conservative filter, no culling,
// no grids — the statement every production path must reproduce byte for byte (tests/test_gpu_parity.py).  Scalar, spheres
// staged through shared memory chunk by chunk.
template <bool ANY>
__global__ void __launch_bounds__(kThreads, 2)
k_intersect(SceneDev sc, QueueDev q) {
    extern __shared__ float4 s_sph[];                        // [chunk] filter operands
    float* s_r2 = reinterpret_cast<float*>(s_sph + min(sc.Spad, kChunkS));   // [chunk] r^2 for the exact test
    __shared__ uint32_t s_group;

    const uint32_t n_in = *q.n_in;
    const uint32_t n = ANY ? n_in * (uint32_t)sc.L : n_in;
    const uint32_t ngroups = (n + kGroup - 1) / kGroup;
    const float cmax = *sc.cmax;
    int staged = -1;                                          // sphere chunk currently in shared memory
    uint32_t slow_cnt = 0;                                    // thread-level entries into the exact test

    for (;;) {
        __syncthreads();                                      // protects s_group and the staged chunk
        if (threadIdx.x == 0) s_group = atomicAdd(q.work, 1u);
        __syncthreads();
        const uint32_t g = s_group;
        if (g >= ngroups) break;

        // ---- load / build this thread's rays ----
        float dx[kRays], dy[kRays], dz[kRays], o2x[kRays], o2y[kRays], o2z[kRays], thr[kRays];
        float best[kRays]; int bid[kRays];
#pragma unroll
        for (int r = 0; r < kRays; ++r) {
            const uint32_t i = g * kGroup + r * kThreads + threadIdx.x;
            V3 O = mk(0.f, 0.f, 0.f), D = mk(0.f, 0.f, 0.f);
            const bool active = i < n;
            if (active) {
                if (ANY) {
                    const uint32_t j = i / n_in, h = i - j * n_in;
                    const float4 p = q.hP[h]; const float4 l = __ldg(&sc.lights[j]);
                    O = mk(p.x, p.y, p.z);
                    D = normalize3(sub3(mk(l.x, l.y, l.z), O));            // common.py:84-91
                } else {
                    const float4 d = q.rayD[i];
                    D = mk(d.x, d.y, d.z);
                    uint32_t slot_; O = ray_origin(q.rayO, q.camo, i, D, d.w, slot_);
                }
            }
            dx[r] = D.x; dy[r] = D.y; dz[r] = D.z;
            o2x[r] = 2.f * O.x; o2y[r] = 2.f * O.y; o2z[r] = 2.f * O.z;
            const float oo = dot3(O, O);
            const float A = cmax + __fsqrt_rn(oo);
            const float margin = fmaf(3.8146973e-6f /* 64 * 2^-24 */, A * A, 1e-12f);
            thr[r] = active ? (oo - margin) : __int_as_float(0x7f800000);
            best[r] = kTMax; bid[r] = -1;
        }

        // ---- spheres, staged chunk by chunk ----
        for (int c0 = 0; c0 < sc.Spad; c0 += kChunkS) {
            const int cnt = min(kChunkS, sc.Spad - c0);
            if (staged != c0) {
                __syncthreads();
                for (int s = threadIdx.x; s < cnt; s += kThreads) {
                    s_sph[s] = sc.sphA[c0 + s];
                    s_r2[s] = (c0 + s < sc.S) ? sc.sphB[c0 + s].w : 0.f;
                }
                staged = c0;
                __syncthreads();
            }
#pragma unroll 2
            for (int s = 0; s < cnt; s += 2) {
                const float4 a0 = s_sph[s], a1 = s_sph[s + 1];
                bool p0[kRays], p1[kRays]; bool any = false;
#pragma unroll
                for (int r = 0; r < kRays; ++r) {
                    p0[r] = thr[r] < 3e38f && (c0 + s < sc.S); p1[r] = thr[r] < 3e38f && (c0 + s + 1 < sc.S);       // every sphere of every active ray
                    any |= p0[r] | p1[r];
                }
                if (any) {                                                // rare: a few spheres per ray
                    slow_cnt++;
#pragma unroll
                    for (int r = 0; r < kRays; ++r) {
                        const V3 O = mk(0.5f * o2x[r], 0.5f * o2y[r], 0.5f * o2z[r]);   // exact
                        const V3 D = mk(dx[r], dy[r], dz[r]);
                        if (p0[r]) {
                            const float t = sphere_exact(O, D, a0.x, a0.y, a0.z, s_r2[s]);
                            if (ANY) { if (t > 0.f && t < kTMax) { bid[r] = 0; thr[r] = __int_as_float(0x7f800000); } }
                            else if (t > 0.f && t < best[r]) { best[r] = t; bid[r] = c0 + s; }    // trace.py:26-29
                        }
                        if (p1[r]) {
                            const float t = sphere_exact(O, D, a1.x, a1.y, a1.z, s_r2[s + 1]);
                            if (ANY) { if (t > 0.f && t < kTMax) { bid[r] = 0; thr[r] = __int_as_float(0x7f800000); } }
                            else if (t > 0.f && t < best[r]) { best[r] = t; bid[r] = c0 + s + 1; }
                        }
                    }
                }
            }
        }

        // ---- planes (trace.py:32-39), then write back ----
#pragma unroll
        for (int r = 0; r < kRays; ++r) {
            const uint32_t i = g * kGroup + r * kThreads + threadIdx.x;
            if (i >= n) continue;
            const V3 O = mk(0.5f * o2x[r], 0.5f * o2y[r], 0.5f * o2z[r]);
            const V3 D = mk(dx[r], dy[r], dz[r]);
            if (ANY) {
                bool occluded = bid[r] >= 0;
                for (int p = 0; p < sc.Pl && !occluded; ++p) {
                    const float t = plane_exact(O, D, __ldg(&sc.plO[p]), __ldg(&sc.plN[p]));
                    occluded = (t > 0.f && t < kTMax);
                }
                q.occl[(size_t)(i / n_in) * q.ocap + (i % n_in)] = occluded ? 1 : 0;
            } else {
                for (int p = 0; p < sc.Pl; ++p) {
                    const float t = plane_exact(O, D, __ldg(&sc.plO[p]), __ldg(&sc.plN[p]));
                    if (t > 0.f && t < best[r]) { best[r] = t; bid[r] = sc.S + p; }
                }
                q.hit[i] = ((unsigned long long)__float_as_uint(best[r]) << 32) | (uint32_t)bid[r];
            }
        }
    }
    if (q.slow) {
        const uint32_t tot = __reduce_add_sync(0xffffffffu, slow_cnt);
        if (lane_id() == 0 && tot) atomicAdd(q.slow, (unsigned long long)tot);
    }
}

// Order of light j's table: spheres by decreasing solid angle seen from the light, ~ r^2 / |C - V_j|^2 — the likelier
// occluders come first, so most occluded shadow rays are finished after the first pass over a fraction of the spheres
// (C4: half of all shadow rays are occluded, 90 % of those by one of the first 120-350 spheres of 1024).  Any order
// gives the same image: a shadow query only asks WHETHER something is hit (trace.py:92-96).  One thread per
// (light j, sphere s): rank by counting (S is a few thousand at most); padding entries go last.
__global__ void __launch_bounds__(256) k_rank(SceneDev sc, int* perm, const uint32_t* cs) {
    if (cache_fresh(cs, kCsBase)) return;
    // grid = (ceil(Spad / 256), L); the keys of 1024 spheres at a time go through shared memory (one division per key
    // per CTA instead of one per comparison)
    __shared__ float skey[1024];
    const int j = blockIdx.y, s = blockIdx.x * blockDim.x + threadIdx.x;
    const float4 v = sc.lights[j];
    auto key = [&](int t) -> float {
        if (t >= sc.S) return -1.f;
        const float4 sp = sc.sphB[t];
        const float ex = sp.x - v.x, ey = sp.y - v.y, ez = sp.z - v.z;
        const float k = sp.w / fmaxf(ex * ex + ey * ey + ez * ez, 1e-30f);
        return (k == k) ? k : -0.5f;                          // NaN sphere: after every valid one (keys >= 0), before the padding (-1); the order must stay total
    };
    const float ks = key(s);
    int rank = 0;
    for (int t0 = 0; t0 < sc.Spad; t0 += 1024) {
        __syncthreads();
        for (int t = threadIdx.x; t < 1024; t += blockDim.x) skey[t] = key(t0 + t);      // (t0 + t >= Spad: key -1 and index beyond s: never counted twice)
        __syncthreads();
        const int n = min(1024, sc.Spad - t0);
        for (int t = 0; t < n; ++t) { const float kt = skey[t]; rank += (kt > ks || (kt == ks && t0 + t < s)) ? 1 : 0; }
    }
    if (s < sc.Spad) perm[(size_t)j * sc.Spad + rank] = s;
}

// Order of the two closest-hit tables (row L of perm): spheres along a Morton curve through their centres, padding last.
// Any order gives the same image (the closest hit is the minimum over (t, sphere id), and ids are mapped back through
// perm); a spatial order makes 64 consecutive table positions a compact cluster, which k_isect3's beam culling tests as
// a whole before it looks at the cluster's pairs (k_clus, isect3.cuh).
__device__ __forceinline__ uint32_t morton10(uint32_t v) {          // spread the low 10 bits: ..9876543210 -> 9..8..7.. (every third bit)
    v &= 1023u; v = (v | (v << 16)) & 0x030000ffu; v = (v | (v << 8)) & 0x0300f00fu; v = (v | (v << 4)) & 0x030c30c3u; v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__global__ void __launch_bounds__(256) k_mrank(SceneDev sc, int* perm, const uint32_t* cs) {
    if (cache_fresh(cs, kCsBase)) return;
    __shared__ uint32_t skey[1024];
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    const float cm = fmaxf(sc.cmax[0], 1e-20f);
    auto key = [&](int t) -> uint32_t {
        if (t >= sc.S) return 0xffffffffu;
        const float4 sp = sc.sphB[t];
        const uint32_t qx = (uint32_t)min(1023, max(0, (int)((sp.x / cm * 0.5f + 0.5f) * 1023.f)));
        const uint32_t qy = (uint32_t)min(1023, max(0, (int)((sp.y / cm * 0.5f + 0.5f) * 1023.f)));
        const uint32_t qz = (uint32_t)min(1023, max(0, (int)((sp.z / cm * 0.5f + 0.5f) * 1023.f)));
        return morton10(qx) | (morton10(qy) << 1) | (morton10(qz) << 2);
    };
    const uint32_t ks = key(s);
    int rank = 0;
    for (int t0 = 0; t0 < sc.Spad; t0 += 1024) {
        __syncthreads();
        for (int t = threadIdx.x; t < 1024; t += blockDim.x) skey[t] = key(t0 + t);
        __syncthreads();
        const int n = min(1024, sc.Spad - t0);
        for (int t = 0; t < n; ++t) { const uint32_t kt = skey[t]; rank += (kt < ks || (kt == ks && t0 + t < s)) ? 1 : 0; }
    }
    if (s < sc.Spad) perm[(size_t)sc.L * sc.Spad + rank] = s;
}
// One thread per cluster of 64 consecutive positions of the closest-hit tables: bounding sphere (centre = mean of the
// centres, radius = max |C - centre| + r, rounded up); a cluster of padding only gets radius -1 (never kept).
__global__ void __launch_bounds__(64) k_clus(SceneDev sc, const int* perm, float4* clus, const uint32_t* cs) {
    if (cache_fresh(cs, kCsBase)) return;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c * 64 >= sc.Spad) return;
    const int* pr = perm + (size_t)sc.L * sc.Spad;
    float mx = 0.f, my = 0.f, mz = 0.f; int n = 0;
    auto finite4 = [](const float4& v) { return isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w); };      // (a non-finite sphere is never hit: it needs no bound)
    for (int k = 0; k < 64 && c * 64 + k < sc.Spad; ++k) { const int s = pr[c * 64 + k]; if (s < sc.S) { const float4 sp = sc.sphB[s]; if (!finite4(sp)) continue; mx += sp.x; my += sp.y; mz += sp.z; ++n; } }
    if (n == 0) { clus[c] = make_float4(0.f, 0.f, 0.f, -1.f); return; }
    mx /= (float)n; my /= (float)n; mz /= (float)n;
    float rad = 0.f;
    for (int k = 0; k < 64 && c * 64 + k < sc.Spad; ++k) {
        const int s = pr[c * 64 + k];
        if (s < sc.S) { const float4 sp = sc.sphB[s]; if (!finite4(sp)) continue; const float ex = sp.x - mx, ey = sp.y - my, ez = sp.z - mz;
                        rad = fmaxf(rad, __fsqrt_ru(ex * ex + ey * ey + ez * ez) * 1.00001f + __fsqrt_ru(fmaxf(sp.w, 0.f))); }
    }
    clus[c] = make_float4(mx, my, mz, rad * 1.00001f + 1e-6f);
}

// One thread per (table j, pair i of table positions): k = r^2 - |C-V_j|^2 (j <= L) or r^2 - |C|^2 (j == L+1), float64
// rounded once.  Light tables (j < L) are in k_rank's order, the two closest-hit tables in array order (ids matter there).
// The camera's table (j == L) is rebuilt every frame, the others only with the scene.
__global__ void k_pivot(SceneDev sc, const CamDev* cam, const int* perm, float4* tab, float4* piv, float* rad, const uint32_t* cs) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int npair = sc.Spad / 2;
    const int j = idx / npair, i = idx - j * npair;
    if (j > sc.L + 1) return;
    if (j != sc.L && cache_fresh(cs, kCsBase)) return;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (j < sc.L) v = sc.lights[j]; else if (j == sc.L) v = make_float4(cam->cox, cam->coy, cam->coz, 0.f);
    if (i == 0 && j <= sc.L) piv[j] = v;
    float c[2][4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int s = perm[(size_t)min(j, sc.L) * sc.Spad + 2 * i + h];         // light j's order, or the closest-hit tables' (row L)
        if (s < sc.S) {
            const float4 sp = sc.sphB[s];
            const double ex = (double)sp.x - v.x, ey = (double)sp.y - v.y, ez = (double)sp.z - v.z;
            c[h][0] = sp.x; c[h][1] = sp.y; c[h][2] = sp.z; c[h][3] = (float)((double)sp.w - (ex * ex + ey * ey + ez * ez));
            rad[(size_t)j * sc.Spad + 2 * i + h] = __fsqrt_ru(sp.w);              // radius, rounded up (beam culling)
        } else { c[h][0] = c[h][1] = c[h][2] = 0.f; c[h][3] = -3.0e38f; rad[(size_t)j * sc.Spad + 2 * i + h] = 0.f; }     // padding never reaches a threshold
    }
    float4* o = tab + ((size_t)j * npair + i) * 2;
    o[0] = make_float4(c[0][0], c[1][0], c[0][1], c[1][1]);
    o[1] = make_float4(c[0][2], c[1][2], c[0][3], c[1][3]);
}


// ------------------------------------------------------------------------------------------------
// Regrouping of the hits of depth >= 1.  The ray queue of a bounce depth is the compacted hit list of the depth
// before, still in the pixel order of the primary rays: after one reflection neighbouring entries have little in common,
// their beam (isect3.cuh, beam culling) covers the whole scene and every group scans every sphere.  Before k_setup the
// hits are therefore binned by (cube-map cell of the REFLECTED direction, grid cell of the hit point): k_setup walks the
// queue through the resulting permutation, and both its products — the hit records the shadow queries read and the next
// ray queue — come out in bin order, where 256 consecutive entries share a small region of space and (for the next
// closest-hit query) a narrow range of directions.  Any order gives the same image: every ray is processed on its
// own and accumulates into its own slot.  One counting sort per depth: k_key (bin + rank within the bin from a
// warp-aggregated atomic), k_binsum / k_binscan (exclusive scan of the bin counts), k_perm (scatter).
// ------------------------------------------------------------------------------------------------
struct SortDev {
    const float4* rayO; const float4* rayD; const float* camo; const unsigned long long* hit; const uint32_t* n_live;
    const GridDev* grid; uint2* keyb; uint32_t* hist; uint32_t* part; uint32_t* base; uint32_t* perm; uint32_t* n_sorted;
    int nb; int use_dir;
    unsigned long long* tl;      // diagnostic timeline slot: k_key start .. k_perm end (may be null)
};

// Bounding box of the spheres -> position grid with 2^posbits cells, bits handed to the axis with the largest cell.
__global__ void __launch_bounds__(256) k_grid(SceneDev sc, int posbits, int dn, GridDev* g, const uint32_t* cs, uint32_t gridkey) {
    if (cache_fresh(cs, kCsGrid) && cs[kCsGridKey] == gridkey) return;
    __shared__ float smin[3][256], smax[3][256];
    float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
    for (int i = threadIdx.x; i < sc.S; i += blockDim.x) {
        const float4 sp = sc.sphB[i]; const float r = sqrtf(fmaxf(sp.w, 0.f)); const float c[3] = {sp.x, sp.y, sp.z};
        for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], c[a] - r); hi[a] = fmaxf(hi[a], c[a] + r); }
    }
    for (int a = 0; a < 3; ++a) { smin[a][threadIdx.x] = lo[a]; smax[a][threadIdx.x] = hi[a]; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int a = 0; a < 3; ++a) for (int t = 1; t < 256; ++t) { lo[a] = fminf(lo[a], smin[a][t]); hi[a] = fmaxf(hi[a], smax[a][t]); }
        float ext[3]; int bits[3] = {0, 0, 0};
        for (int a = 0; a < 3; ++a) { if (!(hi[a] > lo[a])) { lo[a] = 0.f; hi[a] = 1.f; } const float pad = 0.02f * (hi[a] - lo[a]) + 1e-3f; lo[a] -= pad; hi[a] += pad; ext[a] = hi[a] - lo[a]; }
        for (int b = 0; b < posbits; ++b) {
            int m = 0; for (int a = 1; a < 3; ++a) if (ext[a] > ext[m]) m = a;
            bits[m]++; ext[m] *= 0.5f;
        }
        for (int a = 0; a < 3; ++a) { g->lo[a] = lo[a]; g->n[a] = 1 << bits[a]; g->inv[a] = (float)(1 << bits[a]) / (hi[a] - lo[a]); }
        g->dn = dn;
    }
}

__global__ void __launch_bounds__(256, 8) k_key(SceneDev sc, SortDev s) {
    if (s.tl && threadIdx.x == 0) atomicMin(s.tl, globaltimer_ns());
    const uint32_t n = *s.n_live;
    const GridDev g = *s.grid;
    const uint32_t lane = lane_id();
    const uint32_t nround = (n + 31u) & ~31u;                     // whole warps stay in the loop (match/shuffle below)
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nround; i += gridDim.x * blockDim.x) {
        uint32_t b = 0xffffffffu;
        if (i < n) {
            const unsigned long long key = s.hit[i];
            const int id = (int)(uint32_t)key;
            if (id >= 0) {
                const float4 d = s.rayD[i];
                uint32_t slot_; const V3 o = ray_origin(s.rayO, s.camo, i, mk(d.x, d.y, d.z), d.w, slot_);
                b = sort_bin(sc, g, o, mk(d.x, d.y, d.z), __uint_as_float((uint32_t)(key >> 32)), id, s.use_dir);
            }
        }
        // one atomic per distinct bin of the warp: neighbouring rays often share a bin
        const uint32_t peers = __match_any_sync(0xffffffffu, b);
        const uint32_t leader = (uint32_t)__ffs(peers) - 1u;
        uint32_t base = 0;
        if (b != 0xffffffffu && lane == leader) base = atomicAdd(&s.hist[b], (uint32_t)__popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (i < n) s.keyb[i] = make_uint2(b, base + (uint32_t)__popc(peers & ((1u << lane) - 1u)));
    }
}

// Exclusive scan of the bin counts in two launches: sums of 1024-bin blocks, then every block scans its own bins after
// adding up the block sums before it (nb / 1024 values).  The counts are zeroed for the next use.  256-thread CTAs with
// 4 bins per thread and <= 32 registers: like k_key / k_perm / k_setup they then fit on an SM NEXT TO two resident
// k_isect3<shadow> CTAs (8192 registers, 2 KB of shared memory are left), so the dependent chain does not wait for a
// persistent shadow CTA to retire before each of these small launches (profiles/r01_v8: ~250 us per depth at 1/8 frame).
__global__ void __launch_bounds__(256, 8) k_binsum(SortDev s) {
    __shared__ uint32_t sw[8];
    const uint32_t b0 = blockIdx.x * 1024u + threadIdx.x * 4u;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) v += (b0 + k < (uint32_t)s.nb) ? s.hist[b0 + k] : 0u;
    v = __reduce_add_sync(0xffffffffu, v);
    if (lane_id() == 0) sw[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t t = 0; for (int k = 0; k < 8; ++k) t += sw[k]; s.part[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(256, 8) k_binscan(SortDev s) {
    __shared__ uint32_t sw[8]; __shared__ uint32_t s_off;
    uint32_t pre = 0;
    for (uint32_t t = threadIdx.x; t < blockIdx.x; t += 256u) pre += s.part[t];
    pre = __reduce_add_sync(0xffffffffu, pre);
    if (lane_id() == 0) sw[threadIdx.x >> 5] = pre;
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t t = 0; for (int k = 0; k < 8; ++k) t += sw[k]; s_off = t; }
    __syncthreads();
    const uint32_t off = s_off;
    const uint32_t b0 = blockIdx.x * 1024u + threadIdx.x * 4u;
    uint32_t v[4], sum = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) { v[k] = (b0 + k < (uint32_t)s.nb) ? s.hist[b0 + k] : 0u; sum += v[k]; }
    uint32_t x = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if ((int)lane_id() >= d) x += y; }
    __syncthreads();
    if (lane_id() == 31) sw[threadIdx.x >> 5] = x;
    __syncthreads();
    uint32_t woff = 0;
    for (uint32_t k = 0; k < (threadIdx.x >> 5); ++k) woff += sw[k];
    uint32_t excl = off + woff + x - sum;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (b0 + k < (uint32_t)s.nb) { s.base[b0 + k] = excl; s.hist[b0 + k] = 0u; }
        excl += v[k];
        if (b0 + k == (uint32_t)s.nb - 1u) *s.n_sorted = excl;
    }
}
__global__ void __launch_bounds__(256, 8) k_perm(SortDev s) {
    const uint32_t n = *s.n_live;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint2 k = s.keyb[i];
        RT_CHECK(5, k.x == 0xffffffffu || (k.x < (uint32_t)s.nb && s.base[k.x] + k.y < n));
        if (k.x != 0xffffffffu) s.perm[s.base[k.x] + k.y] = i;
    }
    if (s.tl && threadIdx.x == 0) atomicMax(s.tl + 1, globaltimer_ns());
}

// ------------------------------------------------------------------------------------------------
// Per-hit shading kernels (trace.py:56-110, :123-131) — <1 % of the work at S=1024
// ------------------------------------------------------------------------------------------------
#ifndef B200RT_SETUP_OCC
#define B200RT_SETUP_OCC 6      // CTAs per SM the register allocation of k_setup aims at (6 -> 40 registers, no spills; 8 -> 32 with 60 B of spills: no gain measured)
#endif
#ifndef B200RT_SHADE_OCC
#define B200RT_SHADE_OCC 6
#endif
struct ShadeDev {
    const float4* rayO; const float4* rayD; const float* camo;      // this depth's ray queue (k_setup input; rt_math.cuh: ray_origin)
    const unsigned long long* hit;               // packed closest hit per ray (k_setup input)
    const uint32_t* perm;                        // k_setup walks the queue through this permutation (regrouped hits), or null
    float4* rayD2; unsigned long long* hit2;     // next depth's ray queue: directions (the origins are derived from hP); hit2 pre-set to "no hit"
    float4* hP; float4* hN;              // hit records: (biased P, slot) (N, id)
    uint8_t* occl; size_t ocap;          // [light][hit], row stride ocap (zeroed by k_setup, set by the shadow query)
    float4* acc;                         // per-slot colour accumulators (R, G, B, -): one 16-byte read-modify-write per hit
    const uint32_t* n_live;              // rays in the queue (k_setup input)
    uint32_t* n_hit;                     // hit records (k_setup output, k_shade input)
    uint32_t* shidx; size_t shcap; uint32_t* n_sh;   // per light: list of hit records to shadow-test, and its length (k_setup output)
    float amb, lamb, weight;             // weight = refl^depth_index (1 for the primary trace), trace.py:131
    int emit_reflection;                 // 0 on the last depth (k_setup)
    unsigned long long* tl;              // diagnostic timeline slot (may be null)
    unsigned long long* ostat; uint32_t* ticket; size_t ostat_stride;   // k_setup: ordered_append8 status [pass][chunk][8] (zeroed), chunk ticket (zeroed)
};

// trace.py:56-83: miss -> path ends; hit -> P = O + tD, colour/normal by object type, P += BIAS*N.
__global__ void __launch_bounds__(256, B200RT_SETUP_OCC) k_setup(SceneDev sc, ShadeDev s) {   // <= 32 registers: fits beside two resident k_isect3<shadow> CTAs (see k_binsum)
    if (s.tl && threadIdx.x == 0) atomicMin(s.tl, globaltimer_ns());
    const uint32_t n = *s.n_live;
    const uint32_t nchunks = (n + 255u) / 256u;
    __shared__ uint32_t s_chunk;
    for (;;) {
        // Chunks in ticket order (ordered_append8 spins on predecessors).  The ticket is taken only when the CTA is
        // ready to work on it: prefetching the next ticket during the current chunk tripled the frame time (every
        // successor then waits for a chunk whose owner is still busy with the one before).
        if (threadIdx.x == 0) s_chunk = atomicAdd(s.ticket, 1u);
        __syncthreads();
        const uint32_t c = s_chunk;
        if (c >= nchunks) break;
        const bool is_last = (c == nchunks - 1);
        const uint32_t i = c * 256u + threadIdx.x;
        bool hit = false; float4 o = make_float4(0, 0, 0, 0), d = o; float t = 0.f; int id = -1;
        if (i < n) {                                 // the loads go out together (the ray is read even if it missed)
            const uint32_t src = s.perm ? s.perm[i] : i;
            const unsigned long long key = s.hit[src]; d = s.rayD[src];
            uint32_t slot; const V3 O = ray_origin(s.rayO, s.camo, src, mk(d.x, d.y, d.z), d.w, slot);
            o = make_float4(O.x, O.y, O.z, __uint_as_float(slot));
            id = (int)(uint32_t)key; t = __uint_as_float((uint32_t)(key >> 32)); hit = id >= 0;
        }
        V3 Pb = mk(0, 0, 0), N = mk(0, 0, 0);
        if (hit) {
            const V3 P = madd3(mk(o.x, o.y, o.z), t, mk(d.x, d.y, d.z));            // trace.py:60
            if (id < sc.S) { const float4 c = __ldg(&sc.sphB[id]); N = normalize3(sub3(P, mk(c.x, c.y, c.z))); }   // common.py:94-101
            else           { const float4 nu = __ldg(&sc.plNu[id - sc.S]); N = mk(nu.x, nu.y, nu.z); }             // common.py:104-110
            Pb = madd3(P, kBias, N);                                                 // trace.py:82-83
        }
        // The reference runs the shadow query first and then drops the light when k = lamb * (Lv.N) <= 0
        // (trace.py:92-102): for those (hit, light) pairs the query's answer is never used, so only pairs with k > 0
        // go on light j's list.  k is computed exactly as k_shade computes it, so both kernels take the same decision.
        // One multi-queue append per pass: queue 0 = the hit list (first pass), the others = lights' lists
        // (n_hit and n_sh[] are adjacent counters).
        uint32_t h = 0;
        if (!s.shidx && s.perm) { h = i; if (is_last && threadIdx.x == 0) *s.n_hit = n; }            // regrouped: the permutation lists hits only, in their final order
        else if (!s.shidx) { uint32_t pos[8]; ordered_append8(hit ? 1u : 0u, 1, c, s.ostat, s.n_hit, is_last, pos); h = pos[0]; }   // dense shadow layout: hit list only
        else for (int j0 = 0, first = 1, pass = 0; first || j0 < sc.L; first = 0, ++pass) {
            const int nl = min(first ? 7 : 8, sc.L - j0), sh = first ? 1 : 0;
            uint32_t flags = (first && hit) ? 1u : 0u;
            if (hit) {
                for (int jj = 0; jj < nl; ++jj) {
                    const float4 l = __ldg(&sc.lights[j0 + jj]);
                    const V3 dl = sub3(mk(l.x, l.y, l.z), Pb);
                    // Only the SIGN of k_shade's k = lamb * (normalize(dl) . N) is needed.  It is the sign of dl . N
                    // whenever that is clearly away from zero (|cos| > 1e-4: the normalised dot differs from
                    // (dl.N)/|dl| by a few ulp of 1); the rare grazing cases take k_shade's exact arithmetic.
                    const float dn = dot3(dl, N), dd = dot3(dl, dl);
                    bool lit;
                    if (s.lamb > 1e-20f && dn * dn > 1e-8f * dd && dd < 1e30f) lit = dn > 0.f;
                    else lit = __fmul_rn(s.lamb, dot3(normalize3(dl), N)) > 0.f;
                    if (lit) flags |= 1u << (jj + sh);
                }
            }
            uint32_t pos[8];
            ordered_append8(flags, nl + sh, c, s.ostat + (size_t)pass * s.ostat_stride, first ? s.n_hit : s.n_sh + j0, is_last, pos);
            if (first) h = pos[0];
#pragma unroll
            for (int jj = 0; jj < 8; ++jj)
                if (jj >= sh && ((flags >> jj) & 1u)) { RT_CHECK(3, pos[jj] < s.shcap); s.shidx[(size_t)(j0 + jj - sh) * s.shcap + pos[jj]] = h; }
            j0 += nl;
        }
        if (hit) {
            RT_CHECK(2, h < s.shcap && h < n);
            s.hP[h] = make_float4(Pb.x, Pb.y, Pb.z, o.w);
            s.hN[h] = make_float4(N.x, N.y, N.z, __int_as_float(id));
            if (s.occl) for (int j = 0; j < sc.L; ++j) s.occl[(size_t)j * s.ocap + h] = 0;
            if (s.emit_reflection) {                 // the bounce ray does not depend on the shadow queries: queue it now,
                const V3 D = mk(d.x, d.y, d.z);      // the next closest-hit launch runs beside this depth's shadow launches
                const float dn = dot3(D, N);
                const V3 R = normalize3(madd3(D, -2.0f * dn, N));                              // common.py:113-120
                s.rayD2[h] = make_float4(R.x, R.y, R.z, 0.f);                                  // the hit list IS the next queue: origin = Pb + BIAS * R (trace.py:110), derived by the readers
                if (s.hit2) s.hit2[h] = kNoHit;     // (null: the next closest-hit query is a single k_isect3 launch that stores its keys)
            }
        }
    }
    if (s.tl && threadIdx.x == 0) atomicMax(s.tl + 1, globaltimer_ns());
}

// k_setup for regrouped depths on the light-grid path.  The counting sort already gives every hit its position in bin order
// (base[bin] + rank within the bin, k_key / k_binscan), and that position IS its index in the compacted hit list = next ray queue
// (misses have no bin and are not counted).  So the queue is read in its own order — coalesced — and the records are scattered to
// their final positions: no permutation array, no gather through it, no ordered compaction (nothing to compact).  Arithmetic as k_setup.
__global__ void __launch_bounds__(256, B200RT_SETUP_OCC) k_setup_scatter(SceneDev sc, ShadeDev s, const uint2* __restrict__ keyb, const uint32_t* __restrict__ base,
                                                                         const uint32_t* __restrict__ n_sorted) {
    if (s.tl && threadIdx.x == 0) atomicMin(s.tl, globaltimer_ns());
    const uint32_t n = *s.n_live;
    if (blockIdx.x == 0 && threadIdx.x == 0) *s.n_hit = *n_sorted;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint2 kb = keyb[i];
        if (kb.x == 0xffffffffu) continue;                                           // miss: the path ends (trace.py:56-58)
        const unsigned long long key = s.hit[i]; const float4 d = s.rayD[i];
        uint32_t slot; const V3 O = ray_origin(s.rayO, s.camo, i, mk(d.x, d.y, d.z), d.w, slot);
        const float4 o = make_float4(O.x, O.y, O.z, __uint_as_float(slot));
        const uint32_t h = base[kb.x] + kb.y;
        const int id = (int)(uint32_t)key; const float t = __uint_as_float((uint32_t)(key >> 32));
        V3 N;
        const V3 P = madd3(mk(o.x, o.y, o.z), t, mk(d.x, d.y, d.z));                 // trace.py:60
        if (id < sc.S) { const float4 c = __ldg(&sc.sphB[id]); N = normalize3(sub3(P, mk(c.x, c.y, c.z))); }   // common.py:94-101
        else           { const float4 nu = __ldg(&sc.plNu[id - sc.S]); N = mk(nu.x, nu.y, nu.z); }             // common.py:104-110
        const V3 Pb = madd3(P, kBias, N);                                            // trace.py:82-83
        s.hP[h] = make_float4(Pb.x, Pb.y, Pb.z, o.w);
        s.hN[h] = make_float4(N.x, N.y, N.z, __int_as_float(id));
        if (s.emit_reflection) {
            const V3 D = mk(d.x, d.y, d.z);
            const float dn = dot3(D, N);
            const V3 R = normalize3(madd3(D, -2.0f * dn, N));                        // common.py:113-120
            s.rayD2[h] = make_float4(R.x, R.y, R.z, 0.f);                            // origin = Pb + BIAS * R (trace.py:110), derived by the readers
            if (s.hit2) s.hit2[h] = kNoHit;
        }
    }
    if (s.tl && threadIdx.x == 0) atomicMax(s.tl + 1, globaltimer_ns());
}

// k_setup + k_shadow_shade in one kernel (light-grid path, single-stream frames): the thread that builds a hit's record shades it at once, from
// registers.  The record's normal / object id (16 B written, 16 B read) never go through memory, and the biased hit point is written once (it is
// the next depth's queue entry) but not read back: 48 bytes less per hit of the ~90 k_setup and k_shadow_shade move together.  Order of the hit
// list and arithmetic as in the two kernels (same device functions), so the image is the same bytes.
__global__ void __launch_bounds__(256, B200RT_SS_OCC) k_setup_shade(SceneDev sc, ShadeDev s, ShadowGridDev g, ShadowGridArgs a) {
    if (a.tl && threadIdx.x == 0) atomicMin(a.tl, globaltimer_ns());
    const uint32_t n = *s.n_live;
    const uint32_t nchunks = (n + 255u) / 256u;
    const float cmax = a.bounds[0], rmax = a.bounds[1];
    __shared__ float s_A[64];
    __shared__ uint32_t s_chunk;
    if ((int)threadIdx.x < min(sc.L, 64)) { const float4 vv = __ldg(&sc.lights[threadIdx.x]); s_A[threadIdx.x] = cmax + __fsqrt_rn(dot3(mk(vv.x, vv.y, vv.z), mk(vv.x, vv.y, vv.z))); }
    __syncthreads();
    unsigned long long n_filter = 0, n_exact = 0;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_chunk = atomicAdd(s.ticket, 1u);          // chunks in ticket order (ordered_append8 spins on predecessors)
        __syncthreads();
        const uint32_t c = s_chunk;
        if (c >= nchunks) break;
        const bool is_last = (c == nchunks - 1);
        const uint32_t i = c * 256u + threadIdx.x;
        bool hit = false; float4 d = make_float4(0, 0, 0, 0); V3 O = mk(0, 0, 0); uint32_t slot = 0; float t = 0.f; int id = -1;
        if (i < n) {
            const uint32_t src = s.perm ? s.perm[i] : i;
            const unsigned long long key = s.hit[src]; d = s.rayD[src];
            O = ray_origin(s.rayO, s.camo, src, mk(d.x, d.y, d.z), d.w, slot);
            id = (int)(uint32_t)key; t = __uint_as_float((uint32_t)(key >> 32)); hit = id >= 0;
        }
        V3 Pb = mk(0, 0, 0), N = mk(0, 0, 0);
        if (hit) {
            const V3 P = madd3(O, t, mk(d.x, d.y, d.z));                            // trace.py:60
            if (id < sc.S) { const float4 cc = __ldg(&sc.sphB[id]); N = normalize3(sub3(P, mk(cc.x, cc.y, cc.z))); }   // common.py:94-101
            else           { const float4 nu = __ldg(&sc.plNu[id - sc.S]); N = mk(nu.x, nu.y, nu.z); }                 // common.py:104-110
            Pb = madd3(P, kBias, N);                                                 // trace.py:82-83
        }
        uint32_t h = 0;
        if (s.perm) { h = i; if (is_last && threadIdx.x == 0) *s.n_hit = n; }        // regrouped: the permutation lists hits only, in their final order
        else { uint32_t pos[8]; ordered_append8(hit ? 1u : 0u, 1, c, s.ostat, s.n_hit, is_last, pos); h = pos[0]; }
        if (hit) {
            RT_CHECK(2, h < s.shcap && h < n);
            if (s.emit_reflection) {
                s.hP[h] = make_float4(Pb.x, Pb.y, Pb.z, __uint_as_float(slot));      // = the next depth's queue entry (origin Pb + BIAS * R by the readers)
                const V3 D = mk(d.x, d.y, d.z);
                const float dn = dot3(D, N);
                const V3 R = normalize3(madd3(D, -2.0f * dn, N));                    // common.py:113-120
                s.rayD2[h] = make_float4(R.x, R.y, R.z, 0.f);
                if (s.hit2) s.hit2[h] = kNoHit;
            }
            shadow_shade_hit(sc, g, a, s_A, cmax, rmax, Pb, N, id, slot, n_filter, n_exact);      // trace.py:77, 86-102, 131
        }
    }
    if (a.slow) {
#pragma unroll
        for (int dd = 16; dd > 0; dd >>= 1) { n_filter += __shfl_xor_sync(0xffffffffu, n_filter, dd); n_exact += __shfl_xor_sync(0xffffffffu, n_exact, dd); }
        if (lane_id() == 0) { if (n_filter) atomicAdd(a.slow + 2, n_filter); if (n_exact) atomicAdd(a.slow + 1, n_exact); }
    }
    if (a.tl && threadIdx.x == 0) atomicMax(a.tl + 1, globaltimer_ns());
}

// trace.py:77,86-102 and the accumulate of trace.py:131: RGB_trace = amb*col + sum_j [visible_j && k_j>0] k_j*col,
// acc[slot] += weight * RGB_trace.  (The reflection ray of trace.py:104-110 was queued by k_setup.)
__global__ void __launch_bounds__(256, B200RT_SHADE_OCC) k_shade(SceneDev sc, ShadeDev s) {
    if (s.tl && threadIdx.x == 0) atomicMin(s.tl, globaltimer_ns());
    const uint32_t n = *s.n_hit;
    for (uint32_t h = blockIdx.x * blockDim.x + threadIdx.x; h < n; h += gridDim.x * blockDim.x) {
        // Latency-bound kernel: every load whose address is known is issued before the arithmetic — the accumulators,
        // and the occlusion flags of the first 8 lights as one bit mask (a flag that was never written by a shadow
        // query is only looked at when k > 0 fails first, so reading it early is harmless).
        const float4 p = s.hP[h], nn = s.hN[h];
        const uint32_t slot = __float_as_uint(p.w);
        const float4 a0 = s.acc[slot]; const float ar = a0.x, ag = a0.y, ab = a0.z;
        uint32_t occ8 = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) if (j < sc.L && s.occl[(size_t)j * s.ocap + h]) occ8 |= 1u << j;
        const V3 Pb = mk(p.x, p.y, p.z), N = mk(nn.x, nn.y, nn.z);
        const int id = __float_as_int(nn.w);
        const float4 col = (id < sc.S) ? __ldg(&sc.sphCol[id]) : __ldg(&sc.plCol[id - sc.S]);
        float r = __fmul_rn(s.amb, col.x), g = __fmul_rn(s.amb, col.y), b = __fmul_rn(s.amb, col.z);   // trace.py:77
        for (int j = 0; j < sc.L; ++j) {                                                               // trace.py:86
            const bool occluded = j < 8 ? ((occ8 >> j) & 1u) != 0u : s.occl[(size_t)j * s.ocap + h] != 0;     // :95-96
            if (occluded) continue;
            const float4 l = __ldg(&sc.lights[j]);
            const V3 Lv = normalize3(sub3(mk(l.x, l.y, l.z), Pb));                                     // :89
            const float k = __fmul_rn(s.lamb, dot3(Lv, N));                                            // :99
            if (!(k > 0.f)) continue;                                                                  // :101
            r = fmaf(k, col.x, r); g = fmaf(k, col.y, g); b = fmaf(k, col.z, b);                       // :102
        }
        s.acc[slot] = make_float4(fmaf(s.weight, r, ar), fmaf(s.weight, g, ag), fmaf(s.weight, b, ab), 0.f);      // :131
    }
    if (s.tl && threadIdx.x == 0) atomicMax(s.tl + 1, globaltimer_ns());
}

// kernels.py:58-73: sum the samples in reference order with the G/B cross-wiring (F6), /9, quantise with
// round-half-even, swap G/B on output (F5), and store VEC pixels (consecutive y = contiguous bytes) per thread.
template <int VEC>
__global__ void k_resolve(FrameDev f, const float4* __restrict__ acc, uint8_t* __restrict__ result, int aliasing, int unswap) {
    const uint32_t nvec = f.npix / VEC;
    const size_t WH = (size_t)f.W * f.H;
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += gridDim.x * blockDim.x) {
        uint32_t out0[VEC], out1[VEC], out2[VEC];
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            const uint32_t p = v * VEC + e;
            int lc, y; slot_pixel(f, p, lc, y);
            const int x = col_x(f, f.x0 + lc);
            const float4 a0 = acc[p]; float R = a0.x, G = a0.y, B = a0.z;
            if (aliasing && x >= 1 && y >= 1) {                                        // kernels.py:29
                for (int k = 1; k < 9; ++k) {
                    const size_t sl = (size_t)k * f.npix + p;
                    const float4 ak = acc[sl];
                    R = __fadd_rn(R, ak.x); G = __fadd_rn(G, ak.z); B = __fadd_rn(B, ak.y);               // :58-60 (G += B_s, B += G_s)
                }
                R = __fdiv_rn(R, 9.0f); G = __fdiv_rn(G, 9.0f); B = __fdiv_rn(B, 9.0f);                  // :63-65
            }
            out0[e] = clip_color(R);
            out1[e] = unswap ? clip_color(G) : clip_color(B);                           // common.py:60-63
            out2[e] = unswap ? clip_color(B) : clip_color(G);
        }
        const uint32_t p0 = v * VEC;                                  // H % VEC == 0: a vector never straddles columns
        int lc0, y0; slot_pixel(f, p0, lc0, y0);
        const size_t o = (size_t)col_x(f, f.x0 + lc0) * f.H + (size_t)y0;
        uint8_t* d0 = result + o; uint8_t* d1 = result + WH + o; uint8_t* d2 = result + 2 * WH + o;
        if constexpr (VEC == 1) { d0[0] = (uint8_t)out0[0]; d1[0] = (uint8_t)out1[0]; d2[0] = (uint8_t)out2[0]; }
        else {
            uint32_t w0[VEC / 4], w1[VEC / 4], w2[VEC / 4];
#pragma unroll
            for (int e = 0; e < VEC / 4; ++e) {
                w0[e] = out0[4 * e] | (out0[4 * e + 1] << 8) | (out0[4 * e + 2] << 16) | (out0[4 * e + 3] << 24);
                w1[e] = out1[4 * e] | (out1[4 * e + 1] << 8) | (out1[4 * e + 2] << 16) | (out1[4 * e + 3] << 24);
                w2[e] = out2[4 * e] | (out2[4 * e + 1] << 8) | (out2[4 * e + 2] << 16) | (out2[4 * e + 3] << 24);
            }
            if constexpr (VEC == 16) {
                *reinterpret_cast<uint4*>(d0) = make_uint4(w0[0], w0[1], w0[2], w0[3]);
                *reinterpret_cast<uint4*>(d1) = make_uint4(w1[0], w1[1], w1[2], w1[3]);
                *reinterpret_cast<uint4*>(d2) = make_uint4(w2[0], w2[1], w2[2], w2[3]);
            } else {
                *reinterpret_cast<uint32_t*>(d0) = w0[0];
                *reinterpret_cast<uint32_t*>(d1) = w1[0];
                *reinterpret_cast<uint32_t*>(d2) = w2[0];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// FP32 CUDA-core peak probe (roofline denominator): independent FFMA / FFMA2 chains, no memory traffic
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_probe_ffma(float* out, int iters, unsigned long long* clk) {
    float a[8]; const float m = 1.0000001f, c = 1e-9f;
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = (float)(threadIdx.x + i);
    unsigned long long t0 = 0, g0 = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0)); t0 = clock64(); }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], m, c);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t1 = clock64(), g1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
        clk[0] = t1 - t0; clk[1] = g1 - g0;
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_probe_ffma2(float* out, int iters) {
    unsigned long long a[8], m, c;
    { float2 mm = make_float2(1.0000001f, 1.0000002f), cc = make_float2(1e-9f, 2e-9f);
      m = *reinterpret_cast<unsigned long long*>(&mm); c = *reinterpret_cast<unsigned long long*>(&cc); }
#pragma unroll
    for (int i = 0; i < 8; ++i) { float2 v = make_float2((float)(threadIdx.x + i), (float)i); a[i] = *reinterpret_cast<unsigned long long*>(&v); }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i]) : "l"(m), "l"(c));
    }
    float s = 0; for (int i = 0; i < 8; ++i) { float2 v = *reinterpret_cast<float2*>(&a[i]); s += v.x + v.y; }
    if (s == 12345.678f) out[0] = s;
}


// ------------------------------------------------------------------------------------------------
// Presentation step: the array transform of reference src/viewer/image.py:7-19.
// convert_array_to_image stacks (3,W,H) -> (W,H,3), and `ImageOps.mirror(im.rotate(270))` is a transpose for
// square input, i.e. out[y][x][c] = in[c][x][y]: planar, y-contiguous  ->  interleaved RGB8 rows, x-contiguous.
// 64x64 tiles through shared memory: 4-byte loads along y, 4-byte stores along x.
// ------------------------------------------------------------------------------------------------
constexpr int kImgTile = 64;
__global__ void __launch_bounds__(256) k_to_image(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int W, int H, int unswap, int vec) {
    __shared__ uint8_t tile[3][kImgTile][kImgTile + 4];                 // [c][x][y]
    const int x0 = blockIdx.x * kImgTile, y0 = blockIdx.y * kImgTile;
    const size_t WH = (size_t)W * H;
    const bool full = vec && (x0 + kImgTile <= W) && (y0 + kImgTile <= H);
    if (full) {
        for (int i = threadIdx.x; i < 3 * kImgTile * (kImgTile / 4); i += 256) {
            const int c = i / (kImgTile * (kImgTile / 4)), rem = i - c * (kImgTile * (kImgTile / 4));
            const int xx = rem / (kImgTile / 4), y4 = rem - xx * (kImgTile / 4);
            const uint32_t v = *reinterpret_cast<const uint32_t*>(in + c * WH + (size_t)(x0 + xx) * H + y0 + 4 * y4);
            *reinterpret_cast<uint32_t*>(&tile[c][xx][4 * y4]) = v;
        }
    } else {
        for (int i = threadIdx.x; i < 3 * kImgTile * kImgTile; i += 256) {
            const int c = i / (kImgTile * kImgTile), rem = i - c * (kImgTile * kImgTile);
            const int xx = rem / kImgTile, yy = rem - xx * kImgTile;
            if (x0 + xx < W && y0 + yy < H) tile[c][xx][yy] = in[c * WH + (size_t)(x0 + xx) * H + y0 + yy];
        }
    }
    __syncthreads();
    const int c1 = unswap ? 2 : 1, c2 = unswap ? 1 : 2;               // stored order is R,B,G (common.py:60-63); unswap -> R,G,B
    if (full) {
        constexpr int kWords = kImgTile * 3 / 4;                       // 48 words per output row of the tile
        for (int i = threadIdx.x; i < kImgTile * kWords; i += 256) {
            const int yy = i / kWords, wd = i - yy * kWords;
            uint32_t v = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int byte = 4 * wd + b, xx = byte / 3, ch = byte - 3 * xx;
                const int c = ch == 0 ? 0 : (ch == 1 ? c1 : c2);
                v |= (uint32_t)tile[c][xx][yy] << (8 * b);
            }
            *reinterpret_cast<uint32_t*>(out + ((size_t)(y0 + yy) * W + x0) * 3 + 4 * wd) = v;
        }
    } else {
        for (int i = threadIdx.x; i < kImgTile * kImgTile * 3; i += 256) {
            const int yy = i / (kImgTile * 3), rem = i - yy * (kImgTile * 3);
            const int xx = rem / 3, ch = rem - 3 * xx;
            const int c = ch == 0 ? 0 : (ch == 1 ? c1 : c2);
            if (x0 + xx < W && y0 + yy < H) out[((size_t)(y0 + yy) * W + x0 + xx) * 3 + ch] = tile[c][xx][yy];
        }
    }
}

}  // namespace b200rt

// ================================================================================================
// Host side: per-device context, scratch management, launch sequence, C ABI
// ================================================================================================
using namespace b200rt;

std::atomic<unsigned long long> b200rt::g_kernel_launches{0};   // declared in isect3.cuh (both modules count into it)
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
static int cuda_fail(cudaError_t e, const char* what) {
    g_err = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    cudaGetLastError();
    return (int)e;
}
#define CK(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return cuda_fail(e__, #call); } while (0)
#define CKL() do { g_kernel_launches.fetch_add(1, std::memory_order_relaxed); CK(cudaGetLastError()); } while (0)   // after every kernel launch

namespace {

std::atomic<unsigned long long> g_alloc_gen{1};      // bumped whenever any scratch buffer moves: captured graphs hold raw pointers

struct Buf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        g_alloc_gen.fetch_add(1, std::memory_order_relaxed);
        if (p) { cudaError_t e = cudaFree(p); p = nullptr; cap = 0; if (e != cudaSuccess) return e; }   // cudaFree waits for the device
        size_t want = bytes + bytes / 8;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { cudaGetLastError(); e = cudaMalloc(&p, bytes); want = bytes; }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) { cudaFree(p); g_alloc_gen.fetch_add(1, std::memory_order_relaxed); } p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

constexpr int kSets = 4;      // depths whose hit records are alive at once
constexpr int kSide = 4;      // shadow streams

constexpr int kLanes = 2;     // wavefronts of one call that may be in flight at once (each with its own queues and streams)

// Per-wavefront scratch and streams.  A call renders its columns as one wavefront (lane 0), or — for one GPU's share of a
// sharded frame, where a single wavefront leaves the SMs idle in the tails of its persistent kernels and between its ~30
// dependent small launches — as two half-width wavefronts in two lanes whose launch sequences interleave on the device.
struct Lane {
    Buf rayD[2], hitKey[2];                                        // ray queues (directions; origins are derived: rt_math.cuh ray_origin), double-buffered by depth parity
    Buf hP[kSets], hN[kSets], occl[kSets], shidx[kSets], shidx2[kSets];   // hit records / shadow state: ring of kSets depths
    Buf acc, ostat;
    Buf sortKey, sortPerm, sortHist, sortBase, sortPart;           // regrouping of bounce-depth hits (k_key .. k_perm)
    // streams (see the launch sequence in render_device): the dependent chain at high priority, one stream per light (up to
    // kSide) for the shadow queries of the scan path, one for the shading
    cudaStream_t chainStream = nullptr, lightStream[kSide] = {}, shadeStream = nullptr;
    cudaEvent_t evFork = nullptr, evJoin = nullptr, evSetup[kSets] = {}, evShade[kSets] = {}, evLight[kSide] = {};
    void release() {
        Buf* all[] = { &rayD[0], &rayD[1], &hitKey[0], &hitKey[1], &acc, &ostat, &sortKey, &sortPerm, &sortHist, &sortBase, &sortPart };
        for (Buf* b : all) b->release();
        for (int i = 0; i < kSets; ++i) { hP[i].release(); hN[i].release(); occl[i].release(); shidx[i].release(); shidx2[i].release(); }
        if (chainStream) {
            cudaStreamDestroy(chainStream); chainStream = nullptr; cudaStreamDestroy(shadeStream);
            for (int i = 0; i < kSide; ++i) { cudaStreamDestroy(lightStream[i]); cudaEventDestroy(evLight[i]); }
            for (int i = 0; i < kSets; ++i) { cudaEventDestroy(evSetup[i]); cudaEventDestroy(evShade[i]); }
            cudaEventDestroy(evFork); cudaEventDestroy(evJoin);
        }
    }
};

struct DevCtx {
    std::mutex mu;
    Lane lane[kLanes];
    bool init = false;
    int sm_count = 0;
    int isect_blocks_per_sm = 2;
    int force_nr = 0, force_nr0 = 0; bool no_pivot = false, no_conc = false, no_cull = false;                       // tuning overrides (env B200RT_NR, B200RT_NO_PIVOT)
    Buf sphA, sphB, sphCol, plO, plN, plNu, plCol, lights, cmax;   // packed scene
    Buf pivTab, pivPos, pivPerm, pivRad, pivClus;                                   // pivot tables (k_pivot), light-table order (k_rank)
    Buf stage;                                                     // host-scene / host-pixel_loc staging
    Buf ctr, work, slow, tl;
    Buf sgCells, sgTab, sgCone;                                      // per-light direction grids of the shadow queries (shadow_grid.cuh)
    Buf sortGrid;                                                    // bounding box / cell counts of the regrouping (scene-level)
    Buf cgHdr, cgStart, cgCursor, cgIds, cgBig, cgRec;                      // uniform grid of the spheres for per-ray closest-hit queries (closest_grid.cuh)
    int cgrid_ctas = 64;                                             // env B200RT_CGRID_CTAS: CTAs per SM of the walk's grid, at most
    float cgrid_density = 1.5f;                                       // env B200RT_CGRID_DENSITY: cells per sphere the grid aims at
    int cgrid_from = 1; size_t cgrid_max_paths = (size_t)1 << 40, cgrid_nosort_paths = 0;   // env B200RT_CGRID_FROM (first depth served by the grid; 99: never), B200RT_CGRID_MAX_PATHS
    int fuse_ss = 1;                                                 // env B200RT_FUSE_SS: 0 never, 1 single-stream frames (default), 2 always
    int sort_max_depth = 1 << 20;                                    // env B200RT_SORT_MAX_DEPTH: deepest depth whose hits are regrouped
    int split_num = 7, split_den = 8;                                // env B200RT_SPLIT=num/den: share of the columns in the first tile of b200rt_render_host's two-tile overlap
    bool no_fused = false;                                           // env B200RT_NO_FUSED: small scenes through the wavefront kernels
    bool no_fold_key = false;                                        // env B200RT_NO_FOLD_KEY: separate k_key launch
    bool no_cgrid_all = false;                                       // env B200RT_NO_CGRID_ALL
    bool no_lanes = false, lane_conc = false;                        // env B200RT_NO_LANES, B200RT_LANE_CONC
    bool no_sgrid = false, no_clus = false, no_sort_last = false; int ss_grid = 0; int sgrid_min_s = 1; bool no_sort = false; int sort_posbits = -1, sort_dirn = -1;     // env B200RT_NO_SORT, B200RT_SORT_POSBITS, B200RT_SORT_DIRN (-1: chosen from the queue size)
    int npre = 1, pre[6] = {256, 0, 0, 0, 0, 0};                    // leading shadow passes (env B200RT_PASSES=a,b,...; multiples of 8)
    bool timeline = false;                                        // env B200RT_TIMELINE: per-launch first-start / last-end (globaltimer) to stderr
    // resident scene (k_scene_cmp): the raw arrays the tables were built from, the version words, the per-frame camera block
    Buf rawScene, cacheState, camDev; size_t rawCap = 0;
    std::vector<float> hostScene; int hostS = -1, hostL = -1, hostPl = -1; unsigned long long hostStageGen = 0;   // host-side copy of the last staged host scene
    // captured launch sequences (CUDA graphs), keyed by everything that shapes the sequence except the camera
    struct GraphEntry { std::vector<unsigned char> key; cudaGraphExec_t exec = nullptr; unsigned long long gen = 0; unsigned long long last_use = 0; bool failed = false; unsigned kernels = 0; };
    std::vector<GraphEntry> graphs; unsigned long long graph_clock = 0; cudaStream_t captureStream = nullptr;
    bool no_graph = false, no_cache = false, no_scatter = false, graph_all = false, no_split = false;                       // env B200RT_NO_GRAPH, B200RT_NO_CACHE
    unsigned long long n_graph_launches = 0, n_plain_frames = 0;
    Buf batchScene;                                                // (unused since the resident scene; kept for release_all)
    cudaStream_t copyStream = nullptr; cudaEvent_t evRender[2] = {nullptr, nullptr}, evCopy[2] = {nullptr, nullptr};
    Buf hostResult, hostPixelLoc, imgOut;                          // b200rt_render_host / b200rt_to_image_host device copies
    cudaEvent_t done = nullptr; bool has_done = false;
    void release_all() {
        for (auto& g : graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
        graphs.clear(); hostS = hostL = hostPl = -1; hostScene.clear(); rawCap = 0;
        if (captureStream) { cudaStreamDestroy(captureStream); captureStream = nullptr; }
        rawScene.release(); cacheState.release(); camDev.release();
        Buf* all[] = { &sphA, &sphB, &sphCol, &plO, &plN, &plNu, &plCol, &lights, &cmax, &pivTab, &pivPos, &pivPerm, &pivRad, &pivClus, &stage,
                       &tl, &ctr, &work, &slow, &sortGrid, &cgHdr, &cgStart, &cgCursor, &cgIds, &cgBig, &cgRec, &sgCells, &sgTab, &sgCone, &hostResult, &hostPixelLoc, &imgOut, &batchScene };
        for (Buf* b : all) b->release();
        for (int l = 0; l < kLanes; ++l) lane[l].release();
        if (done) { cudaEventDestroy(done); done = nullptr; } has_done = false; init = false;
        if (copyStream) { cudaStreamDestroy(copyStream); copyStream = nullptr; for (int i = 0; i < 2; ++i) { cudaEventDestroy(evRender[i]); cudaEventDestroy(evCopy[i]); } }
    }
};

constexpr int kMaxDevices = 64;
DevCtx g_ctx[kMaxDevices];

size_t isect_smem_bytes(int Spad) { return (size_t)std::min(Spad, kChunkS) * (sizeof(float4) + sizeof(float)); }

size_t max_paths_per_chunk() {
    const char* e = getenv("B200RT_MAX_PATHS");
    if (e) { long long v = atoll(e); if (v >= 1024) return (size_t)v; }
    return (size_t)48 << 20;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct Timer {            // stats mode only: per-class device time from event pairs
    bool on = false; cudaStream_t st = nullptr;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev[3];
    void begin(int cls) { if (!on) return; cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, st); ev[cls].push_back({a, b}); }
    void end(int cls) { if (!on) return; cudaEventRecord(ev[cls].back().second, st); }
    float total(int cls) { float s = 0; for (auto& p : ev[cls]) { float ms = 0; cudaEventElapsedTime(&ms, p.first, p.second); s += ms; cudaEventDestroy(p.first); cudaEventDestroy(p.second); } ev[cls].clear(); return s; }
};

int validate(const b200rt_args* a) {
    if (!a) return fail(B200RT_E_NULL, "args is NULL");
    if (a->struct_size != sizeof(b200rt_args)) return fail(B200RT_E_STRUCT_SIZE, "args->struct_size does not match this library (version skew)");
    if (!a->result) return fail(B200RT_E_NULL, "args->result is NULL");
    if (a->W <= 0 || a->H <= 0) return fail(B200RT_E_SHAPE, "W and H must be positive");
    if ((long long)a->W * a->H > (1ll << 31)) return fail(B200RT_E_SHAPE, "W*H too large");
    if (a->x_begin < 0 || a->x_end > a->W || a->x_begin > a->x_end) return fail(B200RT_E_SHAPE, "column tile [x_begin,x_end) outside [0,W)");
    if (a->strip_width < 0 || (a->strip_width > 0 && (a->strip_world <= 0 || a->strip_rank < 0 || a->strip_rank >= a->strip_world)))
        return fail(B200RT_E_SHAPE, "bad strip_width/strip_rank/strip_world");
    if (a->S < 0 || a->L < 0 || a->Pl < 0) return fail(B200RT_E_SHAPE, "negative object count");
    if ((a->S > 0 && !a->spheres) || (a->L > 0 && !a->lights) || (a->Pl > 0 && !a->planes)) return fail(B200RT_E_NULL, "scene array pointer is NULL");
    if (a->refl_depth < 0 || a->refl_depth > B200RT_MAX_DEPTH) return fail(B200RT_E_DEPTH, "refl_depth out of range");
    int ndev = 0; if (cudaGetDeviceCount(&ndev) != cudaSuccess || a->device < 0 || a->device >= ndev || a->device >= kMaxDevices) { cudaGetLastError(); return fail(B200RT_E_DEVICE, "no such CUDA device"); }
    return 0;
}

template <bool ANY>
int launch_intersect(DevCtx& c, const SceneDev& sc, const QueueDev& q, cudaStream_t st) {
    const size_t smem = isect_smem_bytes(sc.Spad);
    const int grid = c.sm_count * c.isect_blocks_per_sm;
    k_intersect<ANY><<<grid, kThreads, smem, st>>>(sc, q);
    CKL();
    return 0;
}

int init_ctx(DevCtx& c, int device) {
    if (c.init) return 0;
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(B200RT_E_DEVICE, "libb200rt is built for sm_100a (B200) only; device is sm_" + std::to_string(prop.major * 10 + prop.minor));
    c.sm_count = prop.multiProcessorCount;
    const int smem_max = (int)isect_smem_bytes(kChunkS);
    CK(cudaFuncSetAttribute(k_intersect<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    CK(cudaFuncSetAttribute(k_intersect<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    CK(isect3_set_attr<0>()); CK(isect3_set_attr<1>()); CK(shadow_set_attr());
    { const char* e = getenv("B200RT_NR"); c.force_nr = e ? atoi(e) : 0; if (c.force_nr != 2 && c.force_nr != 4 && c.force_nr != 8) c.force_nr = 0; }
    { const char* e = getenv("B200RT_NR0"); c.force_nr0 = e ? atoi(e) : 0; if (c.force_nr0 != 2 && c.force_nr0 != 4 && c.force_nr0 != 8) c.force_nr0 = 0; }
    { const char* e = getenv("B200RT_NO_PIVOT"); c.no_pivot = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_CONCURRENCY"); c.no_conc = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_CULL"); c.no_cull = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_TIMELINE"); c.timeline = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_SORT"); c.no_sort = e && atoi(e) != 0; }
    // The last depth has no next bounce to regroup for, but its shading profits from the order: round 1 (separate k_key pass, ordered compaction in k_setup) measured the
    // sort 0.08 ms slower at C4; with the keys from the walk's epilogue and k_setup_shade it is 0.2 ms faster (10.03 -> 9.83 ms).  B200RT_SORT_LAST=0 skips it.
    { const char* e = getenv("B200RT_SORT_LAST"); if (e) c.no_sort_last = atoi(e) == 0; }
    { const char* e = getenv("B200RT_SS_GRID"); if (e) c.ss_grid = std::max(0, atoi(e)); }     // CTAs per SM of the k_shadow_shade grid (0: sized from the queue)
    { const char* e = getenv("B200RT_NO_CLUS"); c.no_clus = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_SGRID"); c.no_sgrid = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_GRAPH"); c.no_graph = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_CACHE"); c.no_cache = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_SCATTER"); c.no_scatter = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_GRAPH_ALL"); c.graph_all = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_SPLIT"); c.no_split = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_LANES"); c.no_lanes = !(e && atoi(e) != 0); }       // two-lane wavefronts: measured slower, opt-in
    { const char* e = getenv("B200RT_CGRID_FROM"); if (e) c.cgrid_from = std::max(1, atoi(e)); }
    { const char* e = getenv("B200RT_CGRID_CTAS"); if (e) c.cgrid_ctas = std::max(1, std::min(1024, atoi(e))); }
    { const char* e = getenv("B200RT_CGRID_DENSITY"); if (e) c.cgrid_density = std::max(0.25f, (float)atof(e)); }
    { const char* e = getenv("B200RT_CGRID_MAX_PATHS"); if (e) c.cgrid_max_paths = (size_t)std::max(0ll, atoll(e)); }
    { const char* e = getenv("B200RT_FUSE_SS"); if (e) c.fuse_ss = std::max(0, std::min(2, atoi(e))); }
    { const char* e = getenv("B200RT_SPLIT"); int n_ = 0, d_ = 0; if (e && sscanf(e, "%d/%d", &n_, &d_) == 2 && n_ > 0 && d_ > n_) { c.split_num = n_; c.split_den = d_; } }
    { const char* e = getenv("B200RT_SORT_MAX_DEPTH"); if (e) c.sort_max_depth = std::max(0, atoi(e)); }
    { const char* e = getenv("B200RT_NO_FUSED"); c.no_fused = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_FOLD_KEY"); c.no_fold_key = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_NO_CGRID_ALL"); c.no_cgrid_all = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_CGRID_NOSORT_PATHS"); if (e) c.cgrid_nosort_paths = (size_t)std::max(0ll, atoll(e)); }
    { const char* e = getenv("B200RT_LANE_CONC"); c.lane_conc = e && atoi(e) != 0; }
    { const char* e = getenv("B200RT_SGRID_MIN_S"); if (e) c.sgrid_min_s = std::max(1, atoi(e)); }
    { const char* e = getenv("B200RT_SORT_POSBITS"); if (e) c.sort_posbits = std::min(18, std::max(0, atoi(e))); }
    { const char* e = getenv("B200RT_SORT_DIRN"); if (e) c.sort_dirn = std::min(16, std::max(1, atoi(e))); }
    if (const char* e = getenv("B200RT_PASSES")) {
        c.npre = 0;
        for (const char* p = e; *p && c.npre < 6; ) { const int v = atoi(p) & ~7; if (v >= 8) c.pre[c.npre++] = v; while (*p && *p != ',') ++p; if (*p == ',') ++p; }
    }
    CK(cudaEventCreateWithFlags(&c.done, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&c.captureStream, cudaStreamNonBlocking));
    {
        int lo = 0, hi = 0; CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));         // hi = greatest priority (numerically lowest)
        for (int l = 0; l < kLanes; ++l) {
            Lane& ln = c.lane[l];
            CK(cudaStreamCreateWithPriority(&ln.chainStream, cudaStreamNonBlocking, hi));
            CK(cudaStreamCreateWithPriority(&ln.shadeStream, cudaStreamNonBlocking, hi));
            for (int i = 0; i < kSide; ++i) { CK(cudaStreamCreateWithPriority(&ln.lightStream[i], cudaStreamNonBlocking, lo)); CK(cudaEventCreateWithFlags(&ln.evLight[i], cudaEventDisableTiming)); }
            for (int i = 0; i < kSets; ++i) { CK(cudaEventCreateWithFlags(&ln.evSetup[i], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&ln.evShade[i], cudaEventDisableTiming)); }
            CK(cudaEventCreateWithFlags(&ln.evFork, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&ln.evJoin, cudaEventDisableTiming));
        }
    }
    c.init = true;
    return 0;
}

}  // namespace

namespace {

// Small scenes (S < 64, default flags): the frame is k_fused_primary, k_fused_bounce x refl_depth, k_resolve (fused_small.cuh).
// Only k_pack's arrays are needed — no pivot tables, grids or constant banks.
int render_fused_small(DevCtx& c, const b200rt_args* a, cudaStream_t st, const SceneDev& sc, int ncols) {
    const bool stats = a->stats != nullptr;
    const int L = a->L, depth = (int)a->refl_depth, ns = a->aliasing ? 9 : 1, W = a->W, H = a->H;
    const size_t per_col = (size_t)ns * H;
    size_t cols_per_chunk = std::min<size_t>(std::max<size_t>(1, max_paths_per_chunk() / per_col), (size_t)ncols);
    const bool tiled = (H % 16 == 0) && (ncols % 16 == 0) && (cols_per_chunk >= 16);
    if (tiled) cols_per_chunk &= ~(size_t)15;
    const int nchunks = (int)((ncols + cols_per_chunk - 1) / cols_per_chunk);
    const size_t cap = cols_per_chunk * per_col;
    Lane& ln = c.lane[0];
    for (int b = 0; b < 2; ++b) { CK(ln.rayD[b].ensure(sizeof(float4) * cap)); CK(ln.hP[b].ensure(sizeof(float4) * cap)); }
    CK(ln.acc.ensure(sizeof(float4) * cap));
    const int cps = 5 + L, ctr_stride = cps * (depth + 2);
    CK(c.ctr.ensure(sizeof(uint32_t) * (size_t)ctr_stride * nchunks));
    CK(cudaMemsetAsync(c.ctr.p, 0, sizeof(uint32_t) * (size_t)ctr_stride * nchunks, st));
    unsigned long long* tlb = nullptr; int tln = 0; std::vector<std::string> tlname;
    if (c.timeline) {
        CK(c.tl.ensure(sizeof(unsigned long long) * 2 * 1024)); tlb = c.tl.as<unsigned long long>();
        std::vector<unsigned long long> init(2048); for (int i = 0; i < 1024; ++i) { init[2 * i] = ~0ull; init[2 * i + 1] = 0; }
        CK(cudaMemcpyAsync(tlb, init.data(), sizeof(unsigned long long) * 2048, cudaMemcpyHostToDevice, st)); CK(cudaStreamSynchronize(st));
    }
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evA = nullptr, evB = nullptr;
    if (stats) { CK(cudaEventCreate(&ev0)); CK(cudaEventCreate(&ev1)); CK(cudaEventCreate(&evA)); CK(cudaEventCreate(&evB)); CK(cudaEventRecord(ev0, st)); }
    float fused_ms = 0.f;
    const bool diag = stats && !(a->flags & B200RT_F_TIME_ONLY);
    for (int ch = 0; ch < nchunks; ++ch) {
        FrameDev f{};
        f.pixel_loc = a->pixel_loc; f.cam = c.camDev.as<CamDev>();
        f.W = W; f.H = H; f.ns = ns;
        f.x0 = (int)(ch * cols_per_chunk); f.x1 = std::min(ncols, f.x0 + (int)cols_per_chunk);
        f.xb = a->x_begin; f.sw = a->strip_width; f.srank = a->strip_rank; f.sworld = a->strip_world;
        f.npix = (uint32_t)(f.x1 - f.x0) * (uint32_t)H; f.tiled = tiled ? 1 : 0;
        const size_t nslots = (size_t)f.npix * ns;
        uint32_t* ctr = c.ctr.as<uint32_t>() + (size_t)ch * ctr_stride;
        float4* acc = ln.acc.as<float4>();
        CK(cudaMemsetAsync(acc, 0, sizeof(float4) * nslots, st));
        // grid: ~4 paths per thread, at least 2 CTAs per SM
        const int grid = (int)std::min<size_t>((size_t)c.sm_count * 32, std::max<size_t>((size_t)c.sm_count * 2, (nslots + 1023) / 1024));
        if (stats) CK(cudaEventRecord(evA, st));
        for (int d = 0; d <= depth; ++d) {
            uint32_t* cd = ctr + cps * d;
            FusedArgs fa{};
            fa.rayO = d == 0 ? nullptr : ln.hP[(d - 1) & 1].as<float4>(); fa.rayD = ln.rayD[d & 1].as<float4>(); fa.camo = &c.camDev.as<CamDev>()->cox;
            fa.n_in = d == 0 ? nullptr : ctr + cps * (d - 1) + 3;
            fa.hP2 = ln.hP[d & 1].as<float4>(); fa.rayD2 = ln.rayD[(d + 1) & 1].as<float4>();
            fa.n_hit = cd + 3; fa.n_primary = ctr + 0; fa.acc = acc;
            fa.amb = (float)a->amb; fa.lamb = (float)a->lamb; fa.weight = d == 0 ? 1.0f : (float)pow(a->refl, (double)d);     // trace.py:131 refl ** (i+1)
            fa.emit_reflection = d < depth; fa.n_traced = diag ? cd + 4 : nullptr;
            if (tlb) { fa.tl = tlb + 2 * (size_t)tln++; tlname.push_back("fused d" + std::to_string(d)); }
            if (d == 0) k_fused_primary<<<grid, 256, 0, st>>>(sc, fa, f); else k_fused_bounce<<<grid, 256, 0, st>>>(sc, fa);
            CKL();
        }
        if (stats) { CK(cudaEventRecord(evB, st)); }
        const int unswap = (a->flags & B200RT_F_UNSWAP_GB) ? 1 : 0;
        const int small_grid = c.sm_count * 8;
        const bool al16 = (H % 16 == 0) && (((uintptr_t)a->result) % 16 == 0);
        const bool al4 = (H % 4 == 0) && (((uintptr_t)a->result) % 4 == 0);
        if (al16)      k_resolve<16><<<std::min<int>((f.npix / 16 + 255) / 256, small_grid), 256, 0, st>>>(f, acc, a->result, a->aliasing, unswap);
        else if (al4)  k_resolve<4><<<std::min<int>((f.npix / 4 + 255) / 256, small_grid), 256, 0, st>>>(f, acc, a->result, a->aliasing, unswap);
        else           k_resolve<1><<<std::min<int>((f.npix + 255) / 256, small_grid), 256, 0, st>>>(f, acc, a->result, a->aliasing, unswap);
        CKL();
        if (stats) { CK(cudaStreamSynchronize(st)); float ms = 0; CK(cudaEventElapsedTime(&ms, evA, evB)); fused_ms += ms; }
    }
    if (tlb) {
        CK(cudaStreamSynchronize(st));
        std::vector<unsigned long long> h(2 * (size_t)tln); CK(cudaMemcpy(h.data(), tlb, sizeof(unsigned long long) * h.size(), cudaMemcpyDeviceToHost));
        unsigned long long t0 = ~0ull; for (int i = 0; i < tln; ++i) t0 = std::min(t0, h[2 * i]);
        for (int i = 0; i < tln; ++i) fprintf(stderr, "[b200rt timeline] %-16s start %9.1f us  end %9.1f us  dur %8.1f us\n", tlname[i].c_str(), (h[2 * i] - t0) * 1e-3, (h[2 * i + 1] - t0) * 1e-3, (h[2 * i + 1] - h[2 * i]) * 1e-3);
    }
    if (stats) {
        CK(cudaEventRecord(ev1, st));
        CK(cudaStreamSynchronize(st));
        std::vector<uint32_t> h((size_t)ctr_stride * nchunks);
        CK(cudaMemcpy(h.data(), c.ctr.p, sizeof(uint32_t) * h.size(), cudaMemcpyDeviceToHost));
        b200rt_stats out{};
        for (int ch = 0; ch < nchunks; ++ch)
            for (int d = 0; d <= depth; ++d) {
                const uint32_t* cd = h.data() + (size_t)ch * ctr_stride + cps * d;
                for (int j = 0; j < L; ++j) out.shadow_traced += cd[4 + j];
                if (d == 0) out.primary += cd[0]; else out.bounce += (cd - cps)[3];
                out.shadow += (uint64_t)cd[3] * (uint64_t)L;
            }
        CK(cudaEventElapsedTime(&out.kernel_ms[0], ev0, ev1));
        out.kernel_ms[1] = fused_ms; out.kernel_ms[2] = 0.f; out.kernel_ms[3] = std::max(0.f, out.kernel_ms[0] - fused_ms);   // class "closest hit" = the fused trace kernels
        cudaEventDestroy(ev0); cudaEventDestroy(ev1); cudaEventDestroy(evA); cudaEventDestroy(evB);
        *a->stats = out;
    }
    return 0;
}

// The launch sequence.  `a` has DEVICE pointers for pixel_loc/result; scene pointers are host or device per flags.
int render_device(DevCtx& c, const b200rt_args* a, cudaStream_t st) {
    const bool stats = a->stats != nullptr;
    const int S = a->S, L = a->L, Pl = a->Pl;
    const int Spad = std::max(8, (S + 7) & ~7);               // k_isect3 scans 8 spheres per step
    const int depth = (int)a->refl_depth;
    const int ns = a->aliasing ? 9 : 1;
    const int W = a->W, H = a->H;
    const int span = a->x_end - a->x_begin;
    int ncols = span;                                            // columns this call renders
    if (a->strip_width > 0) {
        const int sw = a->strip_width, period = sw * a->strip_world;
        const int full = span / period, rem = span % period;
        ncols = full * sw + std::max(0, std::min(sw, rem - a->strip_rank * sw));
    }
    if (ncols == 0) { if (stats) memset(a->stats, 0, sizeof(*a->stats)); return 0; }

    // ---- scene: resident tables, rebuilt only when k_scene_cmp finds the arrays changed (see "Resident scene" above) ----
    CK(c.sphA.ensure(sizeof(float4) * Spad)); CK(c.sphB.ensure(sizeof(float4) * std::max(S, 1))); CK(c.sphCol.ensure(sizeof(float4) * std::max(S, 1)));
    CK(c.plO.ensure(sizeof(float4) * std::max(Pl, 1))); CK(c.plN.ensure(sizeof(float4) * std::max(Pl, 1)));
    CK(c.plNu.ensure(sizeof(float4) * std::max(Pl, 1))); CK(c.plCol.ensure(sizeof(float4) * std::max(Pl, 1)));
    CK(c.lights.ensure(sizeof(float4) * std::max(L, 1))); CK(c.cmax.ensure(2 * sizeof(float)));
    const float *d_sph = a->spheres, *d_lig = a->lights, *d_pla = a->planes;      // DEVICE arrays (run_frame staged a host scene)
    const size_t nraw = (size_t)7 * S + (size_t)3 * L + (size_t)9 * Pl;
    CK(c.pivTab.ensure(sizeof(float4) * (size_t)(L + 2) * Spad)); CK(c.pivPos.ensure(sizeof(float4) * (L + 1)));
    CK(c.pivPerm.ensure(sizeof(int) * (size_t)(L + 1) * Spad)); CK(c.pivRad.ensure(sizeof(float) * (size_t)(L + 2) * Spad));
    CK(c.pivClus.ensure(sizeof(float4) * (size_t)((Spad + 63) / 64)));
    uint32_t* cst = c.cacheState.as<uint32_t>();                 // (allocated and initialised by run_frame, with rawScene and camDev)
    {
        const bool small = S < 64 && !c.no_fused && !(a->flags & (B200RT_F_NO_FILTER | B200RT_F_DENSE | B200RT_F_NO_CULL | B200RT_F_NO_SGRID | B200RT_F_NO_FUSED));
        const PackOut po{ c.sphA.as<float4>(), c.sphB.as<float4>(), c.sphCol.as<float4>(), c.plO.as<float4>(), c.plN.as<float4>(), c.plNu.as<float4>(),
                          c.plCol.as<float4>(), c.lights.as<float4>(), c.cmax.as<float>() };
        if (small && nraw <= 16384) {                             // (the fused small-scene path; one launch instead of three)
            k_small_prep<<<1, 256, 0, st>>>(d_sph, S, Spad, d_lig, L, d_pla, Pl, c.rawScene.as<uint32_t>(), cst, po); CKL();
        } else {
        const int nb = (int)std::min<size_t>((nraw + 255) / 256 + 1, 64);
        k_scene_cmp<<<nb, 256, 0, st>>>(d_sph, S, d_lig, L, d_pla, Pl, c.rawScene.as<uint32_t>(), cst, c.cmax.as<float>()); CKL();
        const int nobj = std::max(std::max(Spad, Pl), L);
        k_pack<<<(nobj + 255) / 256, 256, 0, st>>>(d_sph, S, Spad, d_pla, Pl, d_lig, L, c.sphA.as<float4>(), c.sphB.as<float4>(),
                                                   c.sphCol.as<float4>(), c.plO.as<float4>(), c.plN.as<float4>(), c.plNu.as<float4>(),
                                                   c.plCol.as<float4>(), c.lights.as<float4>(), c.cmax.as<float>(), cst);
        CKL();
        if (small) { k_cache_commit<<<1, 1, 0, st>>>(cst, 0, 0, 0, 0u, 0); CKL(); }
        }
    }
    SceneDev sc{ c.sphA.as<float4>(), c.sphB.as<float4>(), c.sphCol.as<float4>(), c.plO.as<float4>(), c.plN.as<float4>(),
                 c.plNu.as<float4>(), c.plCol.as<float4>(), c.lights.as<float4>(), c.cmax.as<float>(), S, Spad, Pl, L };
    if (S < 64 && !c.no_fused && !(a->flags & (B200RT_F_NO_FILTER | B200RT_F_DENSE | B200RT_F_NO_CULL | B200RT_F_NO_SGRID | B200RT_F_NO_FUSED)))
        return render_fused_small(c, a, st, sc, ncols);
    {
        const int n = (L + 2) * (Spad / 2);
        if (L > 0) { k_rank<<<dim3((Spad + 255) / 256, L), 256, 0, st>>>(sc, c.pivPerm.as<int>(), cst); CKL(); }
        k_mrank<<<(Spad + 255) / 256, 256, 0, st>>>(sc, c.pivPerm.as<int>(), cst); CKL();
        k_clus<<<((Spad + 63) / 64 + 63) / 64, 64, 0, st>>>(sc, c.pivPerm.as<int>(), c.pivClus.as<float4>(), cst); CKL();
        k_pivot<<<(n + 255) / 256, 256, 0, st>>>(sc, c.camDev.as<CamDev>(), c.pivPerm.as<int>(), c.pivTab.as<float4>(), c.pivPos.as<float4>(), c.pivRad.as<float>(), cst);
        CKL();
    }
    const bool dense = (a->flags & B200RT_F_DENSE) != 0;          // brute force: every specified query scans every sphere
    const bool cull = !(a->flags & B200RT_F_NO_CULL) && !c.no_cull && !dense;
    const PivotDev pv{ c.pivTab.as<float4>(), c.pivPos.as<float4>(), c.cmax.as<float>(), c.pivPerm.as<int>(), cull ? c.pivRad.as<float>() : nullptr, (cull && !c.no_clus) ? c.pivClus.as<float4>() : nullptr };
    // Constant-memory residency (isect3.cuh): the closest-hit module holds the camera-pivot and origin-free tables,
    // the shadow module one table per light.  When they fit they are uploaded once per frame.
    const bool nofilter = (a->flags & B200RT_F_NO_FILTER) != 0;
    const bool closest_res = !nofilter && 2 * Spad <= kTabCap, shadow_res = !nofilter && L > 0 && L * Spad <= kTabCap;
    if (closest_res) { const int tabs[2] = { L, L + 1 }; CK(isect3_upload_resident(pv, Spad, tabs, 2, st)); }
    // Shadow queries through the per-light direction grids (shadow_grid.cuh) wherever work may be skipped at all
    // (beam culling on); k_isect3<shadow> + k_shade stay for the brute-force / debug modes.  (Also for the 6-sphere default
    // scene: C3 3.95 -> 2.58 ms, C5 1.78 -> 1.40 ms per frame against the dense [light][hit] scan.)
    bool sgrid = cull && !nofilter && S >= c.sgrid_min_s && S <= 65535 && L > 0 &&
                 sizeof(uint16_t) * (size_t)kSgTiers * L * kSgCells * kSgRec <= ((size_t)2 << 30) &&      // 12.6 MB per light; beyond 2 GB of grids (81 lights): scan path
                 !(a->flags & B200RT_F_NO_SGRID) && !c.no_sgrid;
    ShadowGridDev sg{};
    if (sgrid) {                                                 // no memory for the grids: the scan path needs none
        const cudaError_t e1 = c.sgCells.ensure(sizeof(uint16_t) * (size_t)kSgTiers * L * kSgCells * kSgRec);
        const cudaError_t e2 = e1 == cudaSuccess ? c.sgTab.ensure(sizeof(float4) * (size_t)L * S) : e1;
        const cudaError_t e3 = e2 == cudaSuccess ? c.sgCone.ensure(sizeof(float4) * (size_t)kSgTiers * L * S) : e2;
        if (e3 != cudaSuccess) { cudaGetLastError(); sgrid = false; }
    }
    if (sgrid) {
        sg = ShadowGridDev{ c.sgCells.as<uint16_t>(), c.sgTab.as<float4>(), c.sgCone.as<float4>(), c.pivPerm.as<int>(), Spad };
        k_sgrid_tab<<<(L * S + 255) / 256, 256, 0, st>>>(sc, sg, cst); CKL();
        k_sgrid_build<<<dim3(kSgCells / 256, L, kSgTiers), 256, 0, st>>>(sc, sg, cst); CKL();
    } else if (shadow_res) CK(shadow_upload_resident(pv, Spad, L, st));
    // Closest-hit queries of bounce rays through a uniform grid of the spheres, one ray per thread (closest_grid.cuh), where
    // the beam method loses its edge: small queues (one GPU's share of a sharded frame).
    const size_t frame_paths = (size_t)ncols * H * ns;
    // Beyond 512 sphere pairs (S > 1024) k_isect3's per-group pair list does not hold a launch's pairs and its groups fall back to
    // the dense scan of every sphere, primary rays included: there the grid walk serves every depth at any frame size.
    const bool cgrid_all = Spad / 2 > 512 && !c.no_cgrid_all;
    const int cgrid_from = cgrid_all ? 0 : c.cgrid_from;
    const bool cgrid = cull && !nofilter && S >= 64 && S <= 65535 && cgrid_from <= depth && (cgrid_all || (depth >= 1 && frame_paths <= c.cgrid_max_paths)) &&
                       !(a->flags & B200RT_F_NO_CGRID);
    CGridDev cg{};
    const int cg_max_cells = (int)std::min<size_t>((size_t)1 << 21, (size_t)32 * S + 4096);
    if (cgrid) {
        const size_t ids_cap = (size_t)64 * S + 4096;
        CK(c.cgHdr.ensure(sizeof(CGridHdr))); CK(c.cgStart.ensure(sizeof(uint32_t) * ((size_t)cg_max_cells + 1))); CK(c.cgCursor.ensure(sizeof(uint32_t) * (size_t)cg_max_cells));
        CK(c.cgIds.ensure(sizeof(uint16_t) * ids_cap)); CK(c.cgBig.ensure(sizeof(uint32_t) * kCgMaxBig)); CK(c.cgRec.ensure(sizeof(uint4) * (size_t)cg_max_cells));
        cg = CGridDev{ c.cgHdr.as<CGridHdr>(), c.cgStart.as<uint32_t>(), c.cgCursor.as<uint32_t>(), c.cgIds.as<uint16_t>(), c.cgRec.as<uint4>(), c.cgBig.as<uint32_t>(), ids_cap };
        k_cg_setup<<<1, 256, 0, st>>>(sc, cg, cg_max_cells, c.cgrid_density, cst, kCsCgrid); CKL();
        k_cg_bin<<<(S + 255) / 256, 256, 0, st>>>(sc, cg, 0, cst, kCsCgrid); CKL();
        k_cg_scan<<<1, 1024, 0, st>>>(cg, cg_max_cells, cst, kCsCgrid); CKL();
        k_cg_bin<<<(S + 255) / 256, 256, 0, st>>>(sc, cg, 1, cst, kCsCgrid); CKL();
        k_cg_pack<<<(cg_max_cells + 255) / 256, 256, 0, st>>>(cg, cst, kCsCgrid); CKL();
    }
    IsectCfg icfg{ c.sm_count, c.force_nr, c.force_nr0, dense ? 0 : c.npre, {0, 0, 0, 0, 0, 0} };
    for (int k = 0; k < icfg.npre; ++k) icfg.pre[k] = c.pre[k];

    // ---- chunking: whole columns, <= max_paths slots per chunk ----
    const size_t per_col = (size_t)ns * H;
    size_t cols_per_chunk = std::max<size_t>(1, max_paths_per_chunk() / per_col);
    cols_per_chunk = std::min<size_t>(cols_per_chunk, (size_t)ncols);
    // Two lanes (see struct Lane): frames that would run on the multi-stream schedule — many spheres, at most 12 Mi paths, i.e.
    // one GPU's share of a sharded frame — are cut into two half-width wavefronts that run side by side.  Measured on 1/8 of
    // C4: one wavefront on three streams 1.72 ms, two lanes 1.5x ms.  Tiles are bit-identical to the whole (tested).
    const bool lanes_ok = !stats && !c.timeline && !c.no_lanes && !(a->flags & B200RT_F_NO_FILTER) && S >= 64 && 2 * Spad <= kTabCap &&
                          (L == 0 || sgrid) && !c.no_conc && (size_t)ncols * per_col <= ((size_t)12 << 20) && (size_t)ncols * per_col >= ((size_t)1 << 20) &&
                          ncols >= 64 && cols_per_chunk == (size_t)ncols;
    const int nlanes = lanes_ok ? 2 : 1;
    if (nlanes == 2) cols_per_chunk = (((size_t)ncols + 1) / 2 + 15) & ~(size_t)15;
    const bool tiled = (H % 16 == 0) && (ncols % 16 == 0) && (cols_per_chunk >= 16);
    if (tiled) cols_per_chunk &= ~(size_t)15;
    const int nchunks = (int)((ncols + cols_per_chunk - 1) / cols_per_chunk);
    const size_t cap = cols_per_chunk * per_col;                 // slots per chunk
    const bool use_lists = S >= 32 && !(a->flags & B200RT_F_NO_FILTER) && !dense;
    const int nsets = std::min(kSets, depth + 1);                // depth d: hit-record set d % nsets
    for (int l = 0; l < nlanes; ++l) {
        Lane& ln = c.lane[l];
        for (int b = 0; b < 2; ++b) {                            // depth d: ray queue d & 1 (see the launch sequence below)
            CK(ln.rayD[b].ensure(sizeof(float4) * cap));
            CK(ln.hitKey[b].ensure(sizeof(unsigned long long) * cap));
        }
        for (int b = 0; b < nsets; ++b) {
            CK(ln.hP[b].ensure(sizeof(float4) * cap)); CK(ln.hN[b].ensure(sizeof(float4) * cap));
            if (!sgrid) CK(ln.occl[b].ensure(std::max<size_t>(cap * (size_t)L, 1)));
            if (use_lists && !sgrid) { CK(ln.shidx[b].ensure(sizeof(uint32_t) * std::max<size_t>(cap * (size_t)L, 1))); CK(ln.shidx2[b].ensure(sizeof(uint32_t) * std::max<size_t>(cap * (size_t)L, 1))); }
        }
        CK(ln.acc.ensure(sizeof(float4) * cap));
    }
    // regrouping of the hits of depth >= 1 (see k_key): worth it where beam culling is on and a query is expensive
    // (B200RT_CGRID_NOSORT_PATHS: frames up to that many paths skip it where the grid walk serves every bounce depth.  It paid while the keys
    // were a separate pass — 1/8 of C4: 1.578 -> 1.526 ms — but not since the walk computes them in its epilogue: 1.468 -> 1.446 with the sort; default 0)
    const bool regroup = cull && use_lists && S >= 64 && depth >= 1 && !(a->flags & B200RT_F_NO_SORT) && !c.no_sort &&
                         !(cgrid && cgrid_from <= 1 && frame_paths <= c.cgrid_nosort_paths);
    // Bin grid of the regrouping: ~40-170 rays per bin measured best at every frame share (C4: 13 position bits and 3x3
    // direction cells per cube face for the whole 33 M-ray frame, 10 bits / 2x2 for 1/8 of it) — bins much smaller than
    // a ray group only fragment the groups.
    int sort_posbits = c.sort_posbits, sort_dirn = c.sort_dirn;
    if (sort_posbits < 0) { sort_posbits = 6; while (sort_posbits < 15 && ((size_t)5793 << sort_posbits) < cap) ++sort_posbits; }   // round(log2(cap / 4096))
    if (sort_dirn < 0) sort_dirn = cap >= ((size_t)24 << 20) ? 3 : 2;
    const int sort_nb = 6 * sort_dirn * sort_dirn << sort_posbits, sort_nblk = (sort_nb + 1023) / 1024;
    if (regroup) {
        for (int l = 0; l < nlanes; ++l) {
            Lane& ln = c.lane[l];
            CK(ln.sortKey.ensure(sizeof(uint2) * cap)); CK(ln.sortPerm.ensure(sizeof(uint32_t) * cap));
            CK(ln.sortHist.ensure(sizeof(uint32_t) * sort_nb)); CK(ln.sortBase.ensure(sizeof(uint32_t) * sort_nb));
            CK(ln.sortPart.ensure(sizeof(uint32_t) * sort_nblk));
            CK(cudaMemsetAsync(ln.sortHist.p, 0, sizeof(uint32_t) * sort_nb, st));
        }
        CK(c.sortGrid.ensure(sizeof(GridDev)));
        k_grid<<<1, 256, 0, st>>>(sc, sort_posbits, sort_dirn, c.sortGrid.as<GridDev>(), cst, (uint32_t)(sort_posbits * 64 + sort_dirn)); CKL();
    }
    k_cache_commit<<<1, 1, 0, st>>>(cst, 1, sgrid ? 1 : 0, regroup ? 1 : 0, (uint32_t)(sort_posbits * 64 + sort_dirn), cgrid ? 1 : 0); CKL();
    // k_setup's ordered compaction: [pass][chunk of 256 rays][8 queues] status words + the chunk ticket, zeroed before every launch
    const size_t ostat_stride = ((cap + 255) / 256) * 8;
    const int ostat_passes = (use_lists && !sgrid) ? 1 + std::max(0, (L - 7 + 7) / 8) : 1;
    const size_t ostat_bytes = sizeof(unsigned long long) * (ostat_stride * ostat_passes + 1);
    for (int l = 0; l < nlanes; ++l) CK(c.lane[l].ostat.ensure(ostat_bytes));
    // Per-light shadow lists pay off when a query is expensive (many spheres); for tiny scenes the dense [light][hit]
    // layout is cheaper than building the lists.
    const int cps = 5 + L;                                       // counters per depth: n_live, n_hit, 2 debug work counters, n_sh[L], ray-fetch counter of the grid walk
    const int ctr_stride = cps * (depth + 2);
    CK(c.ctr.ensure(sizeof(uint32_t) * (size_t)ctr_stride * nchunks)); CK(c.slow.ensure(24 * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(c.ctr.p, 0, sizeof(uint32_t) * (size_t)ctr_stride * nchunks, st));
    const int nsch = (Spad + kTabCap - 1) / kTabCap;             // sphere chunks of the constant-memory table
    const bool store_keys = !nofilter && nsch == 1;              // one k_isect3 launch per closest-hit query: it stores the keys, nobody pre-sets them
    int sh_ranges[kMaxPass][2]; const int sh_np = (L > 0 && !nofilter) ? shadow_passes(icfg, Spad, sh_ranges) : 1;
    if (sh_np < 0) return fail(B200RT_E_SHAPE, "too many spheres for the shadow pass table");
    const size_t nreg = (cap + kRegion - 1) / kRegion;            // regions of the survivor lists
    // per depth: closest: group counter per chunk; shadow: group counter per (light, pass), region fill counts per (light, pass < np-1)
    const size_t work_per_depth = (size_t)nsch + (size_t)L * kMaxPass + (size_t)L * kMaxPass * (sh_np > 1 ? nreg : 0);
    const size_t work_n = work_per_depth * (depth + 1) * nchunks;
    CK(c.work.ensure(sizeof(uint32_t) * work_n));
    CK(cudaMemsetAsync(c.work.p, 0, sizeof(uint32_t) * work_n, st));
    CK(cudaMemsetAsync(c.slow.p, 0, 24 * sizeof(unsigned long long), st));

    if (a->flags & B200RT_F_NO_FILTER) {   // persistent grid of the scalar debug kernel
        int nb = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_intersect<false>, kThreads, isect_smem_bytes(Spad)));
        c.isect_blocks_per_sm = std::max(1, nb);
    }
    const int small_grid = c.sm_count * 8;
    Timer tm; tm.on = stats; tm.st = st;
    unsigned long long* tlb = nullptr; int tln = 0; std::vector<std::string> tlname;
    if (c.timeline) {
        CK(c.tl.ensure(sizeof(unsigned long long) * 2 * 1024)); tlb = c.tl.as<unsigned long long>();
        std::vector<unsigned long long> init(2048); for (int i = 0; i < 1024; ++i) { init[2 * i] = ~0ull; init[2 * i + 1] = 0; }
        CK(cudaMemcpyAsync(tlb, init.data(), sizeof(unsigned long long) * 2048, cudaMemcpyHostToDevice, st)); CK(cudaStreamSynchronize(st));
    }
    auto tl_mark = [&](const char* nm, int d, int from) { for (int i = from; i < tln; ++i) tlname.push_back(std::string(nm) + " d" + std::to_string(d) + (tln - from > 1 ? " j" + std::to_string(i - from) : "")); };
    // concurrency needs launches that do not touch shared constant memory
    const bool conc = !stats && !nofilter && S >= 64 && closest_res && (L == 0 || shadow_res || sgrid) && !c.no_conc
                      && (!sgrid || cap <= ((size_t)12 << 20))      // with the grids, big frames measured faster in one stream (C4: 14.8 vs 15.9 ms; 1/8 of it: 2.7 vs 2.5)
                      && (nlanes == 1 || c.lane_conc);              // two lanes: each lane's sequence in one stream unless B200RT_LANE_CONC=1
    const bool forked = conc || nlanes > 1;                      // the frame's launches run on internal streams, forked from / joined to `st`
    const int nls = std::max(1, std::min(L, kSide));
    if (forked) {
        CK(cudaEventRecord(c.lane[0].evFork, st));
        for (int l = 0; l < nlanes; ++l) CK(cudaStreamWaitEvent(c.lane[l].chainStream, c.lane[0].evFork, 0));
    }
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    if (stats) { CK(cudaEventCreate(&ev0)); CK(cudaEventCreate(&ev1)); CK(cudaEventRecord(ev0, st)); }

    std::vector<float> wts(depth + 1);
    for (int d = 0; d <= depth; ++d) wts[d] = d == 0 ? 1.0f : (float)pow(a->refl, (double)d);     // trace.py:131 refl ** (i+1)

    for (int ch = 0; ch < nchunks; ++ch) {
        Lane& ln = c.lane[ch % nlanes];
        cudaStream_t cs = forked ? ln.chainStream : st, ss = conc ? ln.shadeStream : cs;
        FrameDev f{};
        f.pixel_loc = a->pixel_loc; f.cam = c.camDev.as<CamDev>();
        f.W = W; f.H = H; f.ns = ns;
        f.x0 = (int)(ch * cols_per_chunk); f.x1 = std::min(ncols, f.x0 + (int)cols_per_chunk);
        f.xb = a->x_begin; f.sw = a->strip_width; f.srank = a->strip_rank; f.sworld = a->strip_world;
        f.npix = (uint32_t)(f.x1 - f.x0) * (uint32_t)H; f.tiled = tiled ? 1 : 0;
        const size_t nslots = (size_t)f.npix * ns;
        uint32_t* ctr = c.ctr.as<uint32_t>() + (size_t)ch * ctr_stride;
        float4* acc = ln.acc.as<float4>();

        tm.begin(2);
        CK(cudaMemsetAsync(acc, 0, sizeof(float4) * nslots, cs));
        if (ns > 1) CK(cudaMemsetAsync(ln.ostat.p, 0, ostat_bytes, cs));
        k_raygen<<<(int)std::min<size_t>((nslots + 255) / 256, (size_t)small_grid), 256, 0, cs>>>(f, ln.rayD[0].as<float4>(), store_keys ? nullptr : ln.hitKey[0].as<unsigned long long>(), ctr + 0,
                                                                                                         ln.ostat.as<unsigned long long>(), reinterpret_cast<uint32_t*>(ln.ostat.as<unsigned long long>() + ostat_stride * ostat_passes));
        CKL();
        tm.end(2);

        // Launch sequence.  The dependent chain closest(d) -> setup(d) -> closest(d+1) ... runs on `cs`; the shadow
        // query of light j of every depth on `ls[j % nls]`; the shading of every depth on `ss`.  setup(d) feeds the
        // shadow queries and the shading of depth d, and only the final resolve consumes the shading.  Why: a k_isect3
        // launch ends with warps idling for up to one ray group's time, and with few groups per warp (deep bounces, one
        // GPU of eight) those tails were >10 % of the frame; independent launches pending on other streams back-fill
        // them, and the chain's high priority keeps the critical path moving.  Ray queues alternate between two
        // buffers; the hit records of depth d live in set d % nsets, so setup(d + nsets) waits for shade(d).  With
        // tables that are not resident, the debug kernels, small scenes or stats (per-class timing) everything runs
        // on `st`.
        for (int d = 0; d <= depth; ++d) {
            const int b = d & 1, hs = d % nsets;
            uint32_t* cd = ctr + cps * d;           // [0] n_live (d == 0), [1] work(closest), [2] work(shadow) (scalar debug kernels), [3] n_hit, [4+j] n_sh[j]
            uint32_t* n_live = d == 0 ? ctr + 0 : ctr + cps * (d - 1) + 3;       // the hit list of depth d-1 is the ray queue of depth d
            uint32_t* wk = c.work.as<uint32_t>() + ((size_t)ch * (depth + 1) + d) * work_per_depth;      // k_isect3 group counters
            QueueDev q{};
            q.rayO = d == 0 ? nullptr : ln.hP[(d - 1) % nsets].as<float4>();     // depth >= 1: the hit records of the depth before are the queue
            q.camo = &c.camDev.as<CamDev>()->cox;
            q.rayD = ln.rayD[b].as<float4>(); q.hP = ln.hP[hs].as<float4>();
            q.hit = ln.hitKey[b].as<unsigned long long>(); q.occl = sgrid ? nullptr : ln.occl[hs].as<uint8_t>(); q.ocap = cap;
            q.slow = (stats && !(a->flags & B200RT_F_TIME_ONLY)) ? c.slow.as<unsigned long long>() : nullptr;
            ShadeDev s{};
            s.rayO = q.rayO; s.rayD = q.rayD; s.camo = q.camo; s.hit = q.hit;
            s.rayD2 = ln.rayD[b ^ 1].as<float4>(); s.hit2 = store_keys ? nullptr : ln.hitKey[b ^ 1].as<unsigned long long>();
            s.hP = ln.hP[hs].as<float4>(); s.hN = ln.hN[hs].as<float4>(); s.occl = sgrid ? nullptr : ln.occl[hs].as<uint8_t>(); s.ocap = cap;
            s.acc = acc;
            s.n_live = n_live; s.n_hit = cd + 3;
            s.shidx = (use_lists && !sgrid) ? ln.shidx[hs].as<uint32_t>() : nullptr; s.shcap = cap; s.n_sh = cd + 4;
            s.amb = (float)a->amb; s.lamb = (float)a->lamb; s.weight = wts[d]; s.emit_reflection = d < depth;

            tm.begin(0);
            bool key_folded = false;                            // the regrouping keys of this depth's hits were computed by the closest-hit kernel
            q.n_in = n_live; q.work = cd + 1;
            if (cgrid && d >= cgrid_from) {                     // per-ray grid walk (an unusable grid is handled inside the kernel)
                CGridArgs ca{ q.rayO, q.rayD, q.camo, n_live, q.hit, nullptr, q.slow, cd + 1, nullptr, nullptr, nullptr, 0 };
                const bool sort_here = regroup && d >= 1 && d <= c.sort_max_depth && (d < depth || !c.no_sort_last);
                if (sort_here && !c.no_fold_key) { ca.sgrid = c.sortGrid.as<GridDev>(); ca.hist = ln.sortHist.as<uint32_t>(); ca.keyb = ln.sortKey.as<uint2>(); ca.use_dir = d < depth ? 1 : 0; key_folded = true; }
                if (tlb) { ca.tl = tlb + 2 * (size_t)tln++; tlname.push_back("closest<grid> d" + std::to_string(d)); }
                const int cgg = (int)std::min<size_t>((size_t)c.sm_count * c.cgrid_ctas, std::max<size_t>((size_t)c.sm_count * 8, cap / (256 * 4)));
                if (key_folded) k_closest_grid_key<<<cgg, 256, 0, cs>>>(sc, cg, ca); else k_closest_grid<<<cgg, 256, 0, cs>>>(sc, cg, ca);
                CKL();
            }
            else if (nofilter)                { if (int rc = launch_intersect<false>(c, sc, q, cs)) return rc; }
            else if (d == 0 && !c.no_pivot)   { const int f0 = tln; CK(isect3_launch_closest<1>(icfg, sc, q, pv, nslots, wk, closest_res, cs, tlb, &tln)); tl_mark("closest<1>", d, f0); }
            else {
                const int f0 = tln; CK(isect3_launch_closest<0>(icfg, sc, q, pv, nslots, wk, closest_res, cs, tlb, &tln)); tl_mark("closest<0>", d, f0);
            }
            tm.end(0);
            tm.begin(2);
            bool scatter = false;
            if (regroup && d >= 1 && d <= c.sort_max_depth && (d < depth || !c.no_sort_last)) {
                SortDev so{ q.rayO, q.rayD, q.camo, q.hit, n_live, c.sortGrid.as<GridDev>(), ln.sortKey.as<uint2>(), ln.sortHist.as<uint32_t>(), ln.sortPart.as<uint32_t>(),
                            ln.sortBase.as<uint32_t>(), ln.sortPerm.as<uint32_t>(), cd + 2, sort_nb, d < depth ? 1 : 0, nullptr };
                if (tlb) { so.tl = tlb + 2 * (size_t)tln++; tlname.push_back("sort d" + std::to_string(d)); }
                if (!key_folded) { k_key<<<small_grid, 256, 0, cs>>>(sc, so); CKL(); }
                k_binsum<<<sort_nblk, 256, 0, cs>>>(so); CKL();
                k_binscan<<<sort_nblk, 256, 0, cs>>>(so); CKL();
                // (the scan path's per-light lists need the ordered append; at a full C4 frame the 16-byte scattered stores cost more
                // than the gather through the permutation: "other" 3.20 -> 3.71 ms, at 1/8 of it 0.471 -> 0.453)
                scatter = sgrid && L > 0 && !c.no_scatter && cap <= ((size_t)12 << 20);
                if (!scatter) {
                    k_perm<<<small_grid, 256, 0, cs>>>(so); CKL();
                    s.perm = so.perm; s.n_live = cd + 2;                                  // k_setup walks the hits in bin order
                }
            }
            if (conc && d >= nsets) CK(cudaStreamWaitEvent(cs, ln.evShade[hs], 0));        // set hs is free again
            if (tlb) { s.tl = tlb + 2 * (size_t)tln++; tlname.push_back("setup d" + std::to_string(d)); }
            s.ostat = ln.ostat.as<unsigned long long>(); s.ostat_stride = ostat_stride; s.ticket = reinterpret_cast<uint32_t*>(s.ostat + ostat_stride * ostat_passes);
            // light-grid path in one stream: k_setup and k_shadow_shade as one kernel (k_setup_shade); with the multi-stream schedule they stay
            // apart, the shading of depth d running beside the closest-hit query of depth d+1
            const bool fuse_ss = sgrid && L > 0 && !scatter && (c.fuse_ss == 2 || (c.fuse_ss == 1 && !conc));
            if (fuse_ss) {
                tm.end(2);
                tm.begin(1);
                ShadowGridArgs ga{ nullptr, nullptr, s.n_hit, acc, s.amb, s.lamb, s.weight, pv.bounds, q.slow ? cd + 4 : nullptr, q.slow, s.tl };
                CK(cudaMemsetAsync(ln.ostat.p, 0, ostat_bytes, cs));
                const int ssg = c.ss_grid > 0 ? c.sm_count * c.ss_grid : c.sm_count * B200RT_SS_OCC * 2;
                k_setup_shade<<<ssg, 256, 0, cs>>>(sc, s, sg, ga); CKL();
                if (conc) { CK(cudaEventRecord(ln.evSetup[hs], cs)); CK(cudaEventRecord(ln.evShade[hs], cs)); CK(cudaStreamWaitEvent(ss, ln.evShade[hs], 0)); }
                tm.end(1);
                continue;
            }
            if (scatter) { k_setup_scatter<<<small_grid, 256, 0, cs>>>(sc, s, ln.sortKey.as<uint2>(), ln.sortBase.as<uint32_t>(), cd + 2); CKL(); }
            else {
                CK(cudaMemsetAsync(ln.ostat.p, 0, ostat_bytes, cs));
                k_setup<<<small_grid, 256, 0, cs>>>(sc, s); CKL();
            }
            if (conc) CK(cudaEventRecord(ln.evSetup[hs], cs));
            tm.end(2);
            if (L > 0) {
                tm.begin(1);
                if (nofilter) { q.n_in = cd + 3; q.work = cd + 2; if (int rc = launch_intersect<true>(c, sc, q, cs)) return rc; }
                else if (sgrid) {                       // shadow queries + Lambert sum + accumulate in one kernel (shadow_grid.cuh)
                    ShadowGridArgs ga{ s.hP, s.hN, s.n_hit, acc, s.amb, s.lamb, s.weight, pv.bounds,
                                       q.slow ? cd + 4 : nullptr, q.slow, nullptr };        // (counters only in the diagnostic frame: B200RT_F_TIME_ONLY leaves shadow_traced 0)
                    if (conc) CK(cudaStreamWaitEvent(ss, ln.evSetup[hs], 0));
                    if (tlb) { ga.tl = tlb + 2 * (size_t)tln++; tlname.push_back("shadow+shade d" + std::to_string(d)); }
                    // grid: ~16 hits per thread, between 8 and 64 CTAs per SM (C4: 55 per SM measured 0.12 ms faster than 8; 1/8 of it stays at 8)
                    const int ssg = c.ss_grid > 0 ? c.sm_count * c.ss_grid
                                                  : (int)std::min<size_t>((size_t)c.sm_count * 64, std::max<size_t>((size_t)c.sm_count * 8, cap / (256 * 16)));
                    k_shadow_shade<<<ssg, 256, 0, ss>>>(sc, sg, ga); CKL();
                    if (conc) CK(cudaEventRecord(ln.evShade[hs], ss));
                }
                else {
                    ShadowPlan sp{};
                    sp.list0 = s.shidx; sp.list1 = use_lists ? ln.shidx2[hs].as<uint32_t>() : nullptr; sp.cap = cap;
                    sp.n0 = use_lists ? cd + 4 : cd + 3; sp.n0_stride = use_lists ? 1 : 0;      // per-light list lengths n_sh[j], or the hit count
                    sp.work = wk + nsch; sp.rn = wk + nsch + (size_t)L * kMaxPass; sp.nreg = sh_np > 1 ? nreg : 0;
                    if (conc) for (int k = 0; k < nls; ++k) CK(cudaStreamWaitEvent(ln.lightStream[k], ln.evSetup[hs], 0));
                    for (int j = 0; j < L; ++j) {
                        const int f0 = tln;
                        CK(shadow_launch_light(icfg, sc, q, pv, nslots, sp, j, shadow_res, conc ? ln.lightStream[j % nls] : cs, tlb, &tln));
                        for (int k = f0; k < tln; ++k) tlname.push_back("shadow d" + std::to_string(d) + " j" + std::to_string(j) + " p" + std::to_string(k - f0));
                    }
                    if (conc) for (int k = 0; k < nls; ++k) { CK(cudaEventRecord(ln.evLight[k], ln.lightStream[k])); CK(cudaStreamWaitEvent(ss, ln.evLight[k], 0)); }
                }
                tm.end(1);
            } else if (conc) CK(cudaStreamWaitEvent(ss, ln.evSetup[hs], 0));
            if (!(sgrid && L > 0 && !nofilter)) {
                tm.begin(2);
                if (tlb) { s.tl = tlb + 2 * (size_t)tln++; tlname.push_back("shade d" + std::to_string(d)); }
                k_shade<<<small_grid, 256, 0, ss>>>(sc, s); CKL();
                if (conc) CK(cudaEventRecord(ln.evShade[hs], ss));
                tm.end(2);
            }
        }
        if (conc) CK(cudaStreamWaitEvent(cs, ln.evShade[depth % nsets], 0));               // (ss is in order: every depth is shaded)

        tm.begin(2);
        const int unswap = (a->flags & B200RT_F_UNSWAP_GB) ? 1 : 0;
        const bool al16 = (H % 16 == 0) && (((uintptr_t)a->result) % 16 == 0);     // every column start is then 16-byte aligned
        const bool al4 = (H % 4 == 0) && (((uintptr_t)a->result) % 4 == 0);
        if (al16)      k_resolve<16><<<std::min<int>((f.npix / 16 + 255) / 256, small_grid), 256, 0, cs>>>(f, acc, a->result, a->aliasing, unswap);
        else if (al4)  k_resolve<4><<<std::min<int>((f.npix / 4 + 255) / 256, small_grid), 256, 0, cs>>>(f, acc, a->result, a->aliasing, unswap);
        else           k_resolve<1><<<std::min<int>((f.npix + 255) / 256, small_grid), 256, 0, cs>>>(f, acc, a->result, a->aliasing, unswap);
        CKL();
        tm.end(2);
    }
    if (forked) for (int l = 0; l < nlanes; ++l) { CK(cudaEventRecord(c.lane[l].evJoin, c.lane[l].chainStream)); CK(cudaStreamWaitEvent(st, c.lane[l].evJoin, 0)); }
    if (tlb) {
        CK(cudaStreamSynchronize(st));
        std::vector<unsigned long long> h(2 * (size_t)tln); CK(cudaMemcpy(h.data(), tlb, sizeof(unsigned long long) * h.size(), cudaMemcpyDeviceToHost));
        unsigned long long t0 = ~0ull; for (int i = 0; i < tln; ++i) t0 = std::min(t0, h[2 * i]);
        for (int i = 0; i < tln; ++i) fprintf(stderr, "[b200rt timeline] %-16s start %9.1f us  end %9.1f us  dur %8.1f us\n", tlname[i].c_str(), (h[2 * i] - t0) * 1e-3, (h[2 * i + 1] - t0) * 1e-3, (h[2 * i + 1] - h[2 * i]) * 1e-3);
    }

    if (stats) {
        CK(cudaEventRecord(ev1, st));
        CK(cudaStreamSynchronize(st));
        std::vector<uint32_t> h((size_t)ctr_stride * nchunks);
        CK(cudaMemcpy(h.data(), c.ctr.p, sizeof(uint32_t) * h.size(), cudaMemcpyDeviceToHost));
        b200rt_stats out{}; 
        for (int ch = 0; ch < nchunks; ++ch)
            for (int d = 0; d <= depth; ++d) {
                const uint32_t* cd = h.data() + (size_t)ch * ctr_stride + cps * d;
                // lists (scan path, S >= 32): their lengths; light grids: counted per light by k_shadow_shade in the diagnostic frame (left 0 by
                // B200RT_F_TIME_ONLY); dense [light][hit] layout: every pair is traced
                if (use_lists || sgrid) { for (int j = 0; j < L; ++j) out.shadow_traced += cd[4 + j]; }
                else out.shadow_traced += (uint64_t)cd[3] * (uint64_t)L;
                if (d == 0) out.primary += cd[0]; else out.bounce += (cd - cps)[3];       // queue of depth d = hit list of depth d-1
                out.shadow += (uint64_t)cd[3] * (uint64_t)L;
            }
        unsigned long long slow[24] = {}; CK(cudaMemcpy(slow, c.slow.p, sizeof(slow), cudaMemcpyDeviceToHost));
        if (getenv("B200RT_DEBUG_CULL")) fprintf(stderr, "[cgrid] bounce rays %llu cell steps %llu exact tests %llu\n", (unsigned long long)out.bounce, slow[10], slow[11]);
        if (getenv("B200RT_DEBUG_CULL")) fprintf(stderr, "[cull] shadow first-pass groups %llu, kap<0.1: %llu\n", slow[22], slow[23]);
        if (getenv("B200RT_DEBUG_CULL")) for (int m = 0; m < 3; ++m) fprintf(stderr, "[cull] mode %d groups %llu beam_ok %llu kept_pairs %llu of %llu  kap<0.1: %llu  and ro<1: %llu\n", m, slow[4 + 6 * m], slow[5 + 6 * m], slow[6 + 6 * m], slow[7 + 6 * m], slow[8 + 6 * m], slow[9 + 6 * m]);
        out.slow_path = slow[0]; out.exact_tests = slow[1]; out.tests_pivot = slow[2]; out.tests_general = slow[3];   // counted by k_isect3
        CK(cudaEventElapsedTime(&out.kernel_ms[0], ev0, ev1));
        out.kernel_ms[1] = tm.total(0); out.kernel_ms[2] = tm.total(1); out.kernel_ms[3] = tm.total(2);
        cudaEventDestroy(ev0); cudaEventDestroy(ev1);
        *a->stats = out;
    }
    return 0;
}

}  // namespace

namespace {

__global__ void k_set_cam(CamDev v, CamDev* dst) { if (threadIdx.x == 0) *dst = v; }

// Host scene arrays -> the device staging buffer, skipped when the arrays equal the ones staged last (compared on the host:
// 28 KB at C4).  Afterwards d->spheres / lights / planes are DEVICE pointers into the staging buffer.
int stage_host_scene(DevCtx& c, const b200rt_args* a, cudaStream_t st, b200rt_args* d) {
    const size_t nS = (size_t)7 * a->S, nL = (size_t)3 * a->L, nP = (size_t)9 * a->Pl, n = nS + nL + nP;
    CK(c.stage.ensure(sizeof(float) * std::max<size_t>(n, 1)));
    bool same = c.hostS == a->S && c.hostL == a->L && c.hostPl == a->Pl && c.hostScene.size() == n && c.hostStageGen == g_alloc_gen.load() && !c.no_cache;
    if (same && n) same = (!nS || !memcmp(c.hostScene.data(), a->spheres, sizeof(float) * nS)) &&
                          (!nL || !memcmp(c.hostScene.data() + nS, a->lights, sizeof(float) * nL)) &&
                          (!nP || !memcmp(c.hostScene.data() + nS + nL, a->planes, sizeof(float) * nP));
    float* sg = c.stage.as<float>();
    if (!same) {
        c.hostScene.resize(n);
        if (nS) memcpy(c.hostScene.data(), a->spheres, sizeof(float) * nS);
        if (nL) memcpy(c.hostScene.data() + nS, a->lights, sizeof(float) * nL);
        if (nP) memcpy(c.hostScene.data() + nS + nL, a->planes, sizeof(float) * nP);
        c.hostS = a->S; c.hostL = a->L; c.hostPl = a->Pl;
        if (n) CK(cudaMemcpyAsync(sg, c.hostScene.data(), sizeof(float) * n, cudaMemcpyHostToDevice, st));
        c.hostStageGen = g_alloc_gen.load();
    }
    d->spheres = sg; d->lights = sg + nS; d->planes = sg + nS + nL;
    d->flags &= ~B200RT_F_SCENE_ON_HOST;
    return 0;
}

// One frame tile: resident-scene bookkeeping, the per-frame camera block, then the launch sequence — replayed from a
// captured CUDA graph when the same sequence (same shapes, buffers, flags, shader scalars) has been run before.
int run_frame(DevCtx& c, const b200rt_args* a, cudaStream_t st) {
    b200rt_args d = *a;
    if (c.has_done) CK(cudaStreamWaitEvent(st, c.done, 0));      // serialise users of the shared scratch
    if (a->flags & B200RT_F_SCENE_ON_HOST) { if (int rc = stage_host_scene(c, a, st, &d)) return rc; }
    // resident-scene state
    const size_t nraw = (size_t)7 * a->S + (size_t)3 * a->L + (size_t)9 * a->Pl;
    const bool fresh_state = c.cacheState.p == nullptr;
    CK(c.cacheState.ensure(sizeof(uint32_t) * kCsWords)); CK(c.camDev.ensure(sizeof(CamDev)));
    const void* raw_before = c.rawScene.p;
    CK(c.rawScene.ensure(sizeof(uint32_t) * std::max<size_t>(nraw, 1)));
    if (fresh_state || c.rawScene.p != raw_before || c.no_cache || (a->flags & B200RT_F_NO_CACHE)) {
        uint32_t init[kCsWords] = {};                             // version 1, nothing built, shape unknown
        init[kCsVersion] = 1u; init[kCsS] = init[kCsL] = init[kCsPl] = 0xffffffffu;
        CK(cudaMemcpyAsync(c.cacheState.p, init, sizeof(init), cudaMemcpyHostToDevice, st));
    }
    CamDev cam{};
    cam.gf = a->grid.f; cam.gar = a->grid.ar; cam.gsy = a->grid.sy; cam.gsz = a->grid.sz;
    cam.ox = a->cam_origin[0]; cam.oy = a->cam_origin[1]; cam.oz = a->cam_origin[2];
    for (int i = 0; i < 9; ++i) cam.rot[i] = a->cam_rot[i];
    cam.cox = (float)a->cam_origin[0]; cam.coy = (float)a->cam_origin[1]; cam.coz = (float)a->cam_origin[2];
    k_set_cam<<<1, 32, 0, st>>>(cam, c.camDev.as<CamDev>()); CKL();

    // Replay pays where the frame is launch-bound and runs in one stream (the 6-sphere configs: C1 0.243 -> 0.190 ms) or is big
    // (C4: 10.85 -> 10.77 ms).  Frames on the multi-stream schedule (S >= 64, <= 12 Mi paths: one GPU's share of a sharded frame)
    // keep plain launches: stream priorities are not part of a captured graph and the replay measured slower (1.83 vs 1.78 ms).
    bool multistream = false;
    {
        int ncols = a->x_end - a->x_begin;
        if (a->strip_width > 0) { const int sw = a->strip_width, period = sw * a->strip_world, span = ncols; ncols = span / period * sw + std::max(0, std::min(sw, span % period - a->strip_rank * sw)); }
        const size_t paths = std::min<size_t>((size_t)ncols * a->H * (a->aliasing ? 9 : 1), max_paths_per_chunk());
        multistream = a->S >= 64 && !(a->flags & B200RT_F_NO_FILTER) && !c.no_conc && paths <= ((size_t)12 << 20);
    }
    const bool can_graph = !c.no_graph && !a->stats && !c.timeline && !(a->flags & (B200RT_F_NO_GRAPH | B200RT_F_NO_CACHE)) && !c.no_cache && (!multistream || c.graph_all);
    int rc = 0;
    bool done = false;
    if (can_graph) {
        // key: every argument except the camera block, plus the per-call tuning value and the allocation generation
        b200rt_args k = d;
        memset(k.cam_origin, 0, sizeof(k.cam_origin)); memset(k.cam_rot, 0, sizeof(k.cam_rot)); memset(&k.grid, 0, sizeof(k.grid));
        k.stream = nullptr; k.stats = nullptr;
        const size_t mp = max_paths_per_chunk();
        std::vector<unsigned char> key(sizeof(k) + sizeof(mp));
        memcpy(key.data(), &k, sizeof(k)); memcpy(key.data() + sizeof(k), &mp, sizeof(mp));
        DevCtx::GraphEntry* e = nullptr;
        for (auto& g : c.graphs) if (g.key == key) { e = &g; break; }
        const unsigned long long gen = g_alloc_gen.load();
        if (e && e->exec && e->gen != gen) { cudaGraphExecDestroy(e->exec); e->exec = nullptr; }      // scratch moved: pointers in the graph are stale
        if (e && !e->exec && !e->failed) {                        // second use of this sequence: capture it
            cudaGraph_t graph = nullptr;
            cudaError_t ce = cudaStreamBeginCapture(c.captureStream, cudaStreamCaptureModeRelaxed);
            if (ce == cudaSuccess) {
                const unsigned long long launches0 = g_kernel_launches.load();
                const int rcap = render_device(c, &d, c.captureStream);
                g_kernel_launches.store(launches0);                // (captured, not launched: counted per graph launch below)
                ce = cudaStreamEndCapture(c.captureStream, &graph);
                if (rcap != 0 && ce == cudaSuccess) ce = cudaErrorUnknown;
                if (ce == cudaSuccess) {
                    size_t nn = 0; cudaGraphGetNodes(graph, nullptr, &nn);
                    ce = cudaGraphInstantiate(&e->exec, graph, 0);
                    if (ce == cudaSuccess) {
                        cudaGraphNode_t* nodes = new cudaGraphNode_t[nn ? nn : 1]; cudaGraphGetNodes(graph, nodes, &nn);
                        unsigned kn = 0; for (size_t i = 0; i < nn; ++i) { cudaGraphNodeType t; if (cudaGraphNodeGetType(nodes[i], &t) == cudaSuccess && t == cudaGraphNodeTypeKernel) ++kn; }
                        delete[] nodes;
                        e->kernels = kn;                           // kernel nodes of the graph = launches per replay
                    }
                }
                if (graph) cudaGraphDestroy(graph);
            }
            if (ce != cudaSuccess) {
                cudaGetLastError();
                e->failed = true; e->exec = nullptr;
                fprintf(stderr, "[b200rt] CUDA graph capture failed (%s); this launch sequence keeps running as plain launches\n", cudaGetErrorName(ce));
            } else e->gen = g_alloc_gen.load();
        }
        if (e && e->exec) {
            CK(cudaGraphLaunch(e->exec, st));
            g_kernel_launches.fetch_add(e->kernels, std::memory_order_relaxed);
            e->last_use = ++c.graph_clock; ++c.n_graph_launches; done = true;
        } else if (!e) {                                          // first use: remember the key, run plainly (this allocates the scratch)
            if (c.graphs.size() >= 16) {
                size_t lru = 0; for (size_t i = 1; i < c.graphs.size(); ++i) if (c.graphs[i].last_use < c.graphs[lru].last_use) lru = i;
                if (c.graphs[lru].exec) cudaGraphExecDestroy(c.graphs[lru].exec);
                c.graphs.erase(c.graphs.begin() + lru);
            }
            DevCtx::GraphEntry ne; ne.key = key; ne.last_use = ++c.graph_clock; c.graphs.push_back(ne);
        }
    }
    if (!done) { rc = render_device(c, &d, st); ++c.n_plain_frames; }
    if (rc) return rc;
    CK(cudaEventRecord(c.done, st)); c.has_done = true;
    return 0;
}

}  // namespace

// ================================================================================================
// C ABI (include/b200rt.h)
// ================================================================================================
extern "C" {

B200RT_API int b200rt_version(void) { return B200RT_VERSION; }
B200RT_API const char* b200rt_last_error(void) { return g_err.c_str(); }
B200RT_API uint32_t b200rt_args_size(void) { return (uint32_t)sizeof(b200rt_args); }

B200RT_API int b200rt_device_count(void) {
    int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n;
}
B200RT_API unsigned long long b200rt_launch_count(void) { return b200rt::g_kernel_launches.load(std::memory_order_relaxed); }

B200RT_API unsigned long long b200rt_graph_replays(int device) {
    if (device < 0 || device >= kMaxDevices) return 0;
    return g_ctx[device].n_graph_launches;
}

B200RT_API unsigned long long b200rt_debug_check_failures(void) {
#ifdef B200RT_CHECK
    unsigned int h[8] = {}; unsigned long long t = b200rt::shadow_check_failures();
    if (cudaMemcpyFromSymbol(h, g_check_fail, sizeof(h)) == cudaSuccess) for (unsigned int v : h) t += v;
    return t;
#else
    return ~0ull;                                               // not a checked build
#endif
}

B200RT_API int b200rt_sm_count(int device) {
    int n = 0; if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) { cudaGetLastError(); return 0; } return n;
}

B200RT_API int b200rt_render(const b200rt_args* a) {
    if (int rc = validate(a)) return rc;
    DeviceGuard guard(a->device);
    DevCtx& c = g_ctx[a->device];
    std::lock_guard<std::mutex> lock(c.mu);
    if (int rc = init_ctx(c, a->device)) return rc;
    return run_frame(c, a, (cudaStream_t)a->stream);
}

B200RT_API int b200rt_render_host(const b200rt_args* a) {
    if (int rc = validate(a)) return rc;
    DeviceGuard guard(a->device);
    DevCtx& c = g_ctx[a->device];
    std::lock_guard<std::mutex> lock(c.mu);
    if (int rc = init_ctx(c, a->device)) return rc;
    cudaStream_t st = (cudaStream_t)a->stream;
    const size_t WH = (size_t)a->W * a->H;
    b200rt_args d = *a;
    d.flags |= B200RT_F_SCENE_ON_HOST;
    if (a->pixel_loc) {                                           // reference src/main.py:29
        CK(c.hostPixelLoc.ensure(sizeof(double) * 3 * WH));
        if (c.has_done) CK(cudaStreamWaitEvent(st, c.done, 0));
        CK(cudaMemcpyAsync(c.hostPixelLoc.p, a->pixel_loc, sizeof(double) * 3 * WH, cudaMemcpyHostToDevice, st));
        d.pixel_loc = c.hostPixelLoc.as<double>();
    }
    CK(c.hostResult.ensure(3 * WH));                              // src/main.py:32 (only the tile is written / read back)
    d.result = c.hostResult.as<uint8_t>();
    // Device-to-host copy of the rendered columns of [xb, xe) (`result.copy_to_host()`, src/main.py:51).  The strips of one channel are
    // equally spaced byte ranges (sw*H bytes every sw*world*H): one 2-D copy per channel, plus the clipped last strip if there is one.
    auto copy_back = [&](int xb, int xe, cudaStream_t cst) -> int {
        const int sw = a->strip_width > 0 ? a->strip_width : std::max(1, xe - xb);
        const int world = a->strip_width > 0 ? a->strip_world : 1, srank = a->strip_width > 0 ? a->strip_rank : 0;
        if (xe - xb == a->W && world == 1) { CK(cudaMemcpyAsync(a->result, d.result, 3 * WH, cudaMemcpyDeviceToHost, cst)); return 0; }
        const int x0 = xb + srank * sw, period = sw * world;
        int nfull = 0, xlast = -1;
        for (int xs = x0; xs < xe; xs += period) { if (xs + sw <= xe) ++nfull; else xlast = xs; }
        for (int ch = 0; ch < 3; ++ch) {
            const size_t off = ch * WH + (size_t)x0 * a->H;
            if (nfull) CK(cudaMemcpy2DAsync(a->result + off, (size_t)period * a->H, d.result + off, (size_t)period * a->H,
                                            (size_t)sw * a->H, nfull, cudaMemcpyDeviceToHost, cst));
            if (xlast >= 0) CK(cudaMemcpyAsync(a->result + ch * WH + (size_t)xlast * a->H, d.result + ch * WH + (size_t)xlast * a->H,
                                               (size_t)(xe - xlast) * a->H, cudaMemcpyDeviceToHost, cst));
        }
        return 0;
    };
    // Big frames are rendered as two column tiles, 7/8 and 1/8 of the columns: the first tile's copy (a second stream) runs
    // under the second tile's render, so only the last eighth of the framebuffer is copied after the last kernel
    // (C4: 99.5 MB = 1.9 ms of PCIe time, 12.3 -> 10.8 ms per frame; splits between 4/5 and 7/8 measured within 0.1 ms of each other).
    // Tiles are bit-identical to the whole frame (tested).
    const int span = a->x_end - a->x_begin;
    const int period = a->strip_width > 0 ? a->strip_width * a->strip_world : 16;
    const int world_ = a->strip_width > 0 ? a->strip_world : 1;
    int xsplit = a->x_begin + (int)((long long)span * c.split_num / c.split_den) / period * period;
    const bool split = !a->stats && !c.no_split && 3ull * (unsigned long long)(span / world_) * a->H >= (48ull << 20) && xsplit > a->x_begin && xsplit < a->x_end;
    if (!split) {
        if (int rc = run_frame(c, &d, st)) return rc;
        if (int rc = copy_back(a->x_begin, a->x_end, st)) return rc;
    } else {
        if (!c.copyStream) {
            CK(cudaStreamCreateWithFlags(&c.copyStream, cudaStreamNonBlocking));
            for (int i = 0; i < 2; ++i) { CK(cudaEventCreateWithFlags(&c.evRender[i], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&c.evCopy[i], cudaEventDisableTiming)); }
        }
        d.x_begin = a->x_begin; d.x_end = xsplit;
        if (int rc = run_frame(c, &d, st)) return rc;
        CK(cudaEventRecord(c.evRender[0], st));
        CK(cudaStreamWaitEvent(c.copyStream, c.evRender[0], 0));
        if (int rc = copy_back(a->x_begin, xsplit, c.copyStream)) return rc;
        d.x_begin = xsplit; d.x_end = a->x_end;
        if (int rc = run_frame(c, &d, st)) return rc;
        if (int rc = copy_back(xsplit, a->x_end, st)) return rc;
        CK(cudaStreamSynchronize(c.copyStream));
    }
    CK(cudaStreamSynchronize(st));
    return 0;
}

// Camera-pose batch (BASELINE config 5): the scene is staged once, frame k renders into one of two
// device framebuffers while frame k-1's D2H copy runs on a second stream.
B200RT_API int b200rt_render_batch_host(const b200rt_args* a, int n_frames, const double* cam_origins, const double* cam_rots,
                                        uint8_t* results) {
    if (int rc = validate(a)) return rc;
    if (n_frames < 0 || (n_frames > 0 && (!cam_origins || !cam_rots || !results))) return fail(B200RT_E_NULL, "batch arrays are NULL");
    if (a->x_begin != 0 || a->x_end != a->W || a->strip_width != 0) return fail(B200RT_E_SHAPE, "batch rendering takes whole frames");
    DeviceGuard guard(a->device);
    DevCtx& c = g_ctx[a->device];
    std::lock_guard<std::mutex> lock(c.mu);
    if (int rc = init_ctx(c, a->device)) return rc;
    cudaStream_t st = (cudaStream_t)a->stream;
    const size_t WH = (size_t)a->W * a->H, fb = 3 * WH;
    if (!c.copyStream) {
        CK(cudaStreamCreateWithFlags(&c.copyStream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) { CK(cudaEventCreateWithFlags(&c.evRender[i], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&c.evCopy[i], cudaEventDisableTiming)); }
    }
    CK(c.hostResult.ensure(2 * fb));
    b200rt_args d = *a;
    d.stats = nullptr;
    // scene: host arrays every frame; run_frame stages them once (they compare equal afterwards) and the tables stay resident
    d.flags |= B200RT_F_SCENE_ON_HOST;
    if (c.has_done) CK(cudaStreamWaitEvent(st, c.done, 0));
    if (a->pixel_loc) {
        CK(c.hostPixelLoc.ensure(sizeof(double) * 3 * WH));
        CK(cudaMemcpyAsync(c.hostPixelLoc.p, a->pixel_loc, sizeof(double) * 3 * WH, cudaMemcpyHostToDevice, st));
        d.pixel_loc = c.hostPixelLoc.as<double>();
    }
    for (int k = 0; k < n_frames; ++k) {
        const int b = k & 1;
        for (int i = 0; i < 3; ++i) d.cam_origin[i] = cam_origins[3 * k + i];
        for (int i = 0; i < 9; ++i) d.cam_rot[i] = cam_rots[9 * k + i];
        d.result = c.hostResult.as<uint8_t>() + (size_t)b * fb;
        if (k >= 2) CK(cudaStreamWaitEvent(st, c.evCopy[b], 0));                 // frame k-2 has left this framebuffer
        if (int rc = run_frame(c, &d, st)) return rc;
        CK(cudaEventRecord(c.evRender[b], st));
        CK(cudaStreamWaitEvent(c.copyStream, c.evRender[b], 0));
        CK(cudaMemcpyAsync(results + (size_t)k * fb, d.result, fb, cudaMemcpyDeviceToHost, c.copyStream));
        CK(cudaEventRecord(c.evCopy[b], c.copyStream));
    }
    CK(cudaStreamSynchronize(st));
    CK(cudaStreamSynchronize(c.copyStream));
    return 0;
}

B200RT_API int b200rt_fp32_probe(int device, double* ffma_tflops, double* ffma2_tflops, double* sm_mhz) {
    int ndev = b200rt_device_count();
    if (device < 0 || device >= ndev) return fail(B200RT_E_DEVICE, "no such CUDA device");
    DeviceGuard guard(device);
    int sms = b200rt_sm_count(device);
    float* out = nullptr; unsigned long long* clk = nullptr;
    CK(cudaMalloc(&out, 64)); CK(cudaMalloc(&clk, 16));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int grid = sms * 8, iters = 4096;                      // 8 CTAs x 256 threads per SM = full occupancy
    const double fma_per_thread = (double)iters * 16 * 8;
    double best1 = 0, best2 = 0, mhz = 0;
    for (int rep = 0; rep < 5; ++rep) {
        float ms = 0;
        CK(cudaEventRecord(e0)); k_probe_ffma<<<grid, 256>>>(out, iters, clk); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        double tf = 2.0 * fma_per_thread * grid * 256 / (ms * 1e-3) / 1e12; 
        unsigned long long hc[2]; CK(cudaMemcpy(hc, clk, 16, cudaMemcpyDeviceToHost));
        if (tf > best1) { best1 = tf; if (hc[1]) mhz = (double)hc[0] / (double)hc[1] * 1e3; }
        CK(cudaEventRecord(e0)); k_probe_ffma2<<<grid, 256>>>(out, iters); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        tf = 2.0 * 2.0 * fma_per_thread * grid * 256 / (ms * 1e-3) / 1e12;
        if (tf > best2) best2 = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out); cudaFree(clk);
    if (ffma_tflops) *ffma_tflops = best1;
    if (ffma2_tflops) *ffma2_tflops = best2;
    if (sm_mhz) *sm_mhz = mhz;
    return 0;
}

B200RT_API int b200rt_to_image(int device, void* stream, const uint8_t* planar, uint8_t* rgb, int W, int H, uint32_t flags) {
    if (!planar || !rgb) return fail(B200RT_E_NULL, "planar / rgb pointer is NULL");
    if (W <= 0 || H <= 0) return fail(B200RT_E_SHAPE, "W and H must be positive");
    int ndev = b200rt_device_count();
    if (device < 0 || device >= ndev) return fail(B200RT_E_DEVICE, "no such CUDA device");
    DeviceGuard guard(device);
    const int vec = (W % 4 == 0) && (H % 4 == 0) && ((uintptr_t)planar % 4 == 0) && ((uintptr_t)rgb % 4 == 0);
    dim3 grid((W + kImgTile - 1) / kImgTile, (H + kImgTile - 1) / kImgTile);
    k_to_image<<<grid, 256, 0, (cudaStream_t)stream>>>(planar, rgb, W, H, (flags & B200RT_F_UNSWAP_GB) ? 1 : 0, vec);
    CKL();
    return 0;
}

B200RT_API int b200rt_to_image_host(int device, const uint8_t* planar, uint8_t* rgb, int W, int H, uint32_t flags) {
    if (!planar || !rgb) return fail(B200RT_E_NULL, "planar / rgb pointer is NULL");
    if (W <= 0 || H <= 0) return fail(B200RT_E_SHAPE, "W and H must be positive");
    int ndev = b200rt_device_count();
    if (device < 0 || device >= ndev || device >= kMaxDevices) return fail(B200RT_E_DEVICE, "no such CUDA device");
    DeviceGuard guard(device);
    DevCtx& c = g_ctx[device];
    std::lock_guard<std::mutex> lock(c.mu);
    if (int rc = init_ctx(c, device)) return rc;
    const size_t n = (size_t)3 * W * H;
    CK(c.hostResult.ensure(n)); CK(c.imgOut.ensure(n));
    if (c.has_done) CK(cudaStreamWaitEvent(nullptr, c.done, 0));
    CK(cudaMemcpyAsync(c.hostResult.p, planar, n, cudaMemcpyHostToDevice, nullptr));
    if (int rc = b200rt_to_image(device, nullptr, c.hostResult.as<uint8_t>(), c.imgOut.as<uint8_t>(), W, H, flags)) return rc;
    CK(cudaMemcpyAsync(rgb, c.imgOut.p, n, cudaMemcpyDeviceToHost, nullptr));
    CK(cudaStreamSynchronize(nullptr));
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of b200rt_to_image_host (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: tight serving loops, harnesses that forbid it).
// The two device staging buffers that b200rt_to_image_host takes from cudaMalloc'd per-device scratch (DevCtx::hostResult,
// DevCtx::imgOut) live in ONE file-scope device array of fixed size instead: planar input in the first half, interleaved
// output in the second.  Every device that loads the library holds the array (6 MiB) until the process exits.
// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol, never a device address: the device
// address is fetched once per device with cudaGetSymbolAddress (that device current) and only that pointer goes to
// copies and kernels.  The array is shared by every call on the device and keeps the last call's contents; nothing in
// it has to start at zero (the H2D copy overwrites the whole input region and k_to_image writes every byte of the output
// region of a W x H frame), the output region is still cleared first so that a failed launch cannot hand back an older
// frame.  Calls that overlapped on the GPU would race on the array: this entry point therefore holds a per-device mutex,
// works on the default stream and synchronises before it returns, so two host threads on one device queue up.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[2 * 3 * (size_t)B200RT_NOALLOC_MAX_PIXELS];

namespace {
struct NoAllocCtx { std::mutex mu; unsigned char* dev_ptr = nullptr; };
NoAllocCtx g_noalloc[kMaxDevices];
}

B200RT_API int b200rt_to_image_host_noalloc(int device, const uint8_t* planar, uint8_t* rgb, int W, int H, uint32_t flags) {
    if (!planar || !rgb) return fail(B200RT_E_NULL, "planar / rgb pointer is NULL");
    if (W <= 0 || H <= 0) return fail(B200RT_E_SHAPE, "W and H must be positive");
    if ((long long)W * H > (long long)B200RT_NOALLOC_MAX_PIXELS)
        return fail(B200RT_E_SHAPE, "b200rt_to_image_host_noalloc: W*H = " + std::to_string((long long)W * H) + " pixels exceeds the static scratch of B200RT_NOALLOC_MAX_PIXELS = " +
                                    std::to_string((long long)B200RT_NOALLOC_MAX_PIXELS) + " pixels; use b200rt_to_image_host");
    int ndev = b200rt_device_count();
    if (device < 0 || device >= ndev || device >= kMaxDevices) return fail(B200RT_E_DEVICE, "no such CUDA device");
    DeviceGuard guard(device);
    NoAllocCtx& c = g_noalloc[device];
    std::lock_guard<std::mutex> lock(c.mu);
    if (!c.dev_ptr) {                                              // set-up, once per device, with that device current
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(B200RT_E_NOMEM, "cudaGetSymbolAddress returned a null address for the static scratch");
        c.dev_ptr = static_cast<unsigned char*>(p);
    }
    const size_t n = (size_t)3 * W * H;
    unsigned char* d_in = c.dev_ptr;
    unsigned char* d_out = c.dev_ptr + (size_t)3 * B200RT_NOALLOC_MAX_PIXELS;      // 256-byte aligned: 3 * 2^20
    CK(cudaMemcpyAsync(d_in, planar, n, cudaMemcpyHostToDevice, nullptr));
    CK(cudaMemsetAsync(d_out, 0, n, nullptr));
    if (int rc = b200rt_to_image(device, nullptr, d_in, d_out, W, H, flags)) return rc;
    CK(cudaMemcpyAsync(rgb, d_out, n, cudaMemcpyDeviceToHost, nullptr));
    CK(cudaStreamSynchronize(nullptr));
    return 0;
}

// Page-locked host memory for the framebuffer (the destination of `result.copy_to_host()`, reference src/main.py:51; numba's
// counterpart is cuda.pinned_array): a D2H copy into pinned memory is one DMA at PCIe rate (~52 GB/s measured for the 99.5 MB
// C4 frame), into pageable memory it is staged through a driver bounce buffer at a fraction of that.
B200RT_API int b200rt_alloc_host(size_t bytes, void** out) {
    if (!out) return fail(B200RT_E_NULL, "out is NULL");
    *out = nullptr;
    if (bytes == 0) return fail(B200RT_E_SHAPE, "bytes must be positive");
    CK(cudaHostAlloc(out, bytes, cudaHostAllocPortable));
    return 0;
}
B200RT_API int b200rt_free_host(void* p) { if (p) CK(cudaFreeHost(p)); return 0; }
// Pin memory the caller already owns (a numpy array, a shared-memory mapping that several ranks D2H their strips into: the
// "direct device-to-host gather" of SURVEY §8(e)).  The range stays pinned until b200rt_host_unregister.
B200RT_API int b200rt_host_register(void* p, size_t bytes) {
    if (!p) return fail(B200RT_E_NULL, "pointer is NULL");
    if (bytes == 0) return fail(B200RT_E_SHAPE, "bytes must be positive");
    CK(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
    return 0;
}
B200RT_API int b200rt_host_unregister(void* p) { if (p) CK(cudaHostUnregister(p)); return 0; }

B200RT_API int b200rt_shutdown(void) {
    int ndev = b200rt_device_count();
    for (int dev = 0; dev < ndev && dev < kMaxDevices; ++dev) {
        DevCtx& c = g_ctx[dev];
        std::lock_guard<std::mutex> lock(c.mu);
        if (!c.init) continue;
        DeviceGuard guard(dev);
        cudaDeviceSynchronize();
        c.release_all();
    }
    return 0;
}

}  // extern "C"
