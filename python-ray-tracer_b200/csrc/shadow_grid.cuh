// shadow_grid.cuh — shadow queries through a per-light direction grid (reference trace.py:86-96: "is anything hit by
// the ray from the biased hit point towards light j?", answered by get_intersection over every sphere and plane).
//
// Every shadow ray of light j lies — to within e <= 2^-20 |V-P|, the rounding of its normalised direction — on a line
// through the light position V.  Seen from V, a sphere (C, r) can only be met by such a line if the line's direction
// is within alpha = asin(r / |C-V|) of +-(C-V) (the reference's shadow ray does not stop at the light, SURVEY F8, so
// spheres behind the light count as well: hence both signs).  Once per frame k_sgrid_build lists, for each cell of a
// cube map of directions around V, the spheres whose cone reaches the cell; a shadow query (k_shadow_shade, which also
// does the Lambert sum the query's answer feeds) then looks at the ONE cell
// its direction falls in and runs, over that cell's few spheres, exactly what k_isect3<shadow> runs over all of them:
// the 4-FMA pivot filter with the same operands, operation order and per-ray threshold, then sphere_exact on the real
// ray for the candidates, then the planes.  The grid only removes spheres that provably cannot be hit, so the image is
// bit-identical to the brute-force path (tests/test_gpu_parity.py compares with B200RT_F_DENSE / B200RT_F_NO_FILTER):
//   * radius inflated by 2e-3 + 1e-3 r: covers e (<= 2.5e-4) and the exact test's rounding (<= 1e-4 in the perpendicular
//     distance) for |V-P| <= 256 and coordinates <= 256.  Shadow rays that start farther away (bounce rays reach the
//     ground plane up to 999 units out) use a second grid built with 16x the inflation, valid to |V-P| <= 4096; rays
//     from beyond that, lights / scenes outside the coordinate range and cells whose list overflows scan every sphere
//     instead (still through the same filter);
//   * a cell's list is in k_rank's order (decreasing solid angle seen from the light): likely occluders first, and a
//     query stops at its first hit;
//   * cell cones widened by 2e-3 rad: covers the rounding of the cell lookup and of the asin / acos used to build.
#pragma once
#include "isect3.cuh"

namespace b200rt {

#ifndef B200RT_SS_OCC
#define B200RT_SS_OCC 4                  // CTAs per SM the register allocation of k_shadow_shade aims at
#endif
#ifndef B200RT_SGN
#define B200RT_SGN 64                   // (32: shadow kernel +8 %; 48: +1.5 %; measured with the cosine-comparison build)
#endif
constexpr int kSgN = B200RT_SGN;                          // cells per edge of a cube-map face
constexpr int kSgCells = 6 * kSgN * kSgN;         // cells per light
constexpr int kSgRec = 128;                       // uint16 per cell record: [0] count (0xffff: scan every sphere), [1 ..] sphere ids: the first
                                                  // 16-byte load of a record brings the count and 7 ids, every further one 8 ids
constexpr int kSgCap = kSgRec - 1;
constexpr int kSgTiers = 2;                       // tier t serves shadow rays with |V-P| <= kSgRange * 16^t
constexpr float kSgRange = 256.f;
__host__ __device__ __forceinline__ float sg_range(int t) { return t == 0 ? kSgRange : kSgRange * 16.f; }
__host__ __device__ __forceinline__ double sg_inflate(int t) { return t == 0 ? 2e-3 : 3.2e-2; }

struct ShadowGridDev {
    uint16_t* cells;                              // [kSgTiers][L][kSgCells][kSgRec]
    float4* ltab;                                 // [L][S] (Cx, Cy, Cz, k = r^2 - |C - V_j|^2), k as in k_pivot (float64, rounded once); by sphere id
    float4* cone;                                 // [kSgTiers][L][S] (unit direction V_j -> C, alpha) by RANK (k_rank order); alpha >= pi: reaches every cell
    const int* perm; int Spad;                    // [L][Spad] rank -> sphere id (k_rank)
};

// Cube-map cell of a direction (any length, not all zero): face = 2 * major axis + (negative ? 1 : 0), (u, v) = the other
// two components over the major one's magnitude.  Approximate arithmetic: the lists carry a margin for a wrong neighbour.
__device__ __forceinline__ int sg_cell(float x, float y, float z) {
    const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
    float m, u, v; int face;
    if (ax >= ay && ax >= az) { m = ax; u = y; v = z; face = x < 0.f ? 1 : 0; }
    else if (ay >= az)        { m = ay; u = x; v = z; face = y < 0.f ? 3 : 2; }
    else                      { m = az; u = x; v = y; face = z < 0.f ? 5 : 4; }
    const float im = 0.5f * (float)kSgN * __frcp_rn(fmaxf(m, 1e-30f));
    const int iu = min(kSgN - 1, max(0, (int)(u * im + 0.5f * (float)kSgN)));
    const int iv = min(kSgN - 1, max(0, (int)(v * im + 0.5f * (float)kSgN)));
    return (face * kSgN + iu) * kSgN + iv;
}
__device__ __forceinline__ V3 sg_dir(int face, float a, float b) {       // un-normalised direction of face coordinates (a, b) in [-1, 1]
    const float s = (face & 1) ? -1.f : 1.f;
    return (face >> 1) == 0 ? mk(s, a, b) : (face >> 1) == 1 ? mk(a, s, b) : mk(a, b, s);
}

// One thread per (light j, rank pos): filter operand k of the sphere at that rank, and its cone seen from the light per tier.
__global__ void __launch_bounds__(256) k_sgrid_tab(SceneDev sc, ShadowGridDev g, const uint32_t* cs) {
    if (cs[2] == cs[0]) return;                                            // resident scene (b200rt.cu: kCsSgrid / kCsVersion): grids already built from this scene
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= sc.L * sc.S) return;
    const int j = idx / sc.S, pos = idx - j * sc.S;
    const int s = g.perm[(size_t)j * g.Spad + pos];
    const float4 v = sc.lights[j], sp = sc.sphB[s];
    const double ex = (double)sp.x - v.x, ey = (double)sp.y - v.y, ez = (double)sp.z - v.z;
    const double d2 = ex * ex + ey * ey + ez * ez;
    g.ltab[(size_t)j * sc.S + s] = make_float4(sp.x, sp.y, sp.z, (float)((double)sp.w - d2));
    const double dist = sqrt(d2), r = sqrt(fmax((double)sp.w, 0.0));
    for (int t = 0; t < kSgTiers; ++t) {
        const double reff = r * 1.001 + sg_inflate(t);
        float4 c = make_float4(1.f, 0.f, 0.f, 4.f);                        // alpha = 4 > pi: every cell
        if (dist > reff * 1.0001 && dist < 1e30) c = make_float4((float)(ex / dist), (float)(ey / dist), (float)(ez / dist), (float)asin(reff / dist));
        g.cone[((size_t)t * sc.L + j) * sc.S + pos] = c;
    }
}

// One thread per (cell, light, tier): the list, in rank order, of the spheres whose cone (either sign) reaches the cell's cone.
__global__ void __launch_bounds__(256) k_sgrid_build(SceneDev sc, ShadowGridDev g, const uint32_t* cs) {
    if (cs[2] == cs[0]) return;
    __shared__ float4 s_cone[256];                                         // (unit direction, cos alpha)
    __shared__ float s_sin[256];                                           // sin alpha; alpha + beta >= pi: cos = -2 (always listed)
    __shared__ int s_id[256];
    const int cell = blockIdx.x * 256 + threadIdx.x, j = blockIdx.y, tier = blockIdx.z;     // kSgCells % 256 == 0
    const int face = cell / (kSgN * kSgN), iu = (cell / kSgN) % kSgN, iv = cell % kSgN;
    const float h = 2.f / (float)kSgN, a0 = -1.f + h * (float)iu, b0 = -1.f + h * (float)iv;
    const V3 cdir = normalize3(sg_dir(face, a0 + 0.5f * h, b0 + 0.5f * h));
    float beta = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const V3 q = normalize3(sg_dir(face, a0 + ((k & 1) ? h : 0.f), b0 + ((k & 2) ? h : 0.f)));
        beta = fmaxf(beta, acosf(fminf(1.f, dot3(q, cdir))));
    }
    beta += 2e-3f;
    const float cb = cosf(beta), sb = sinf(beta);
    const float4 v = sc.lights[j];
    const bool in_range = sc.cmax[0] + sqrtf(v.x * v.x + v.y * v.y + v.z * v.z) <= kSgRange;    // (NaN -> false)
    uint16_t* rec = g.cells + (((size_t)tier * sc.L + j) * kSgCells + cell) * kSgRec;
    const float4* cone = g.cone + ((size_t)tier * sc.L + j) * sc.S;
    uint32_t cnt = 0;
    for (int s0 = 0; s0 < sc.S; s0 += 256) {
        __syncthreads();
        if (s0 + (int)threadIdx.x < sc.S) {
            const float4 c = cone[s0 + threadIdx.x];
            // angle(u, cell) <= alpha + beta  <=>  u . cdir >= cos(alpha + beta) = cos a cos b - sin a sin b  (alpha + beta < pi; beta <= 0.8)
            const bool all = c.w >= 3.1415927f - 0.9f;
            s_cone[threadIdx.x] = make_float4(c.x, c.y, c.z, all ? -2.f : cosf(c.w)); s_sin[threadIdx.x] = all ? 0.f : sinf(c.w);
            s_id[threadIdx.x] = g.perm[(size_t)j * g.Spad + s0 + threadIdx.x];
        }
        __syncthreads();
        const int n = min(256, sc.S - s0);
        for (int k = 0; k < n; ++k) {
            const float4 c = s_cone[k];
            const float dt = fabsf(c.x * cdir.x + c.y * cdir.y + c.z * cdir.z);
            if (dt >= fmaf(c.w, cb, -s_sin[k] * sb) - 1e-6f) { if (cnt < (uint32_t)kSgCap) rec[1 + cnt] = (uint16_t)s_id[k]; ++cnt; }
        }
    }
    const bool over = cnt > (uint32_t)kSgCap || !in_range;
    for (uint32_t k = min(cnt, (uint32_t)kSgCap); k < (uint32_t)kSgCap && ((k + 1u) & 7u); ++k) rec[1 + k] = 0;   // pad the last 16-byte group (a valid id: S >= 1)
    rec[0] = over ? (uint16_t)0xffff : (uint16_t)cnt;
}

__device__ __forceinline__ void warp_count(uint32_t* ctr) {     // += 1 per calling lane, one atomic per warp
    const uint32_t m = __activemask();
    if ((int)lane_id() == __ffs(m) - 1) atomicAdd(ctr, (uint32_t)__popc(m));
}

struct ShadowGridArgs {
    const float4* hP; const float4* hN;  // hit records: (biased hit point, slot) (normal, object id)
    const uint32_t* n_hit;               // number of hit records
    float4* acc;                         // per-slot colour accumulators (R, G, B, -)
    float amb, lamb, weight;             // weight = refl^depth (trace.py:131)
    const float* bounds;                 // max |C|, max r
    uint32_t* n_traced;                  // [L] diagnostics (b200rt_stats.shadow_traced): queries run per light; may be null
    unsigned long long* slow;            // diagnostics: [1] exact tests, [2] filter tests; may be null
    unsigned long long* tl;              // diagnostic timeline slot; may be null
};

// The lights loop of trace() for ONE hit record (biased hit point O, normal N, object id, accumulator slot): see k_shadow_shade below.  Also called by
// k_setup_shade (b200rt.cu), which shades a hit in the thread that has just built its record.
__device__ __forceinline__ void shadow_shade_hit(const SceneDev& sc, const ShadowGridDev& g, const ShadowGridArgs& a, const float* s_A, float cmax, float rmax,
                                                 V3 O, V3 N, int id, uint32_t slot, unsigned long long& n_filter, unsigned long long& n_exact) {
    const float4 a0 = a.acc[slot]; const float ar = a0.x, ag = a0.y, ab = a0.z;
    const float4 col = (id < sc.S) ? __ldg(&sc.sphCol[id]) : __ldg(&sc.plCol[id - sc.S]);
    float r = __fmul_rn(a.amb, col.x), gg = __fmul_rn(a.amb, col.y), b = __fmul_rn(a.amb, col.z);   // trace.py:77
    const float B = cmax + __fsqrt_rn(dot3(O, O));
    for (int j = 0; j < sc.L; ++j) {                                                               // trace.py:86
        const float4 vv = __ldg(&sc.lights[j]);
        const V3 Vp = mk(vv.x, vv.y, vv.z);
        // the ray and its filter threshold: the operations of k_isect3<shadow>'s prologue; D is also k_shade's Lv (common.py:84-91)
        const V3 d = sub3(Vp, O);
        const float nrmc = __fsqrt_rn(dot3(d, d));
        const V3 Dc = mk(__fdiv_rn(d.x, nrmc), __fdiv_rn(d.y, nrmc), __fdiv_rn(d.z, nrmc));
        const float kc = __fmul_rn(a.lamb, dot3(Dc, N));                                           // :99
        if (!(kc > 0.f)) continue;                                                                 // :101 (the query's answer would not be used)
        // the first 16 bytes of the cell's record: count + 7 ids (one load where count and ids used to be two dependent ones)
        const uint16_t* recc = g.cells + ((((size_t)(nrmc <= sg_range(0) ? 0 : 1) * sc.L + j) * kSgCells) + (size_t)sg_cell(Dc.x, Dc.y, Dc.z)) * kSgRec;
        RT_CHECK(7, sg_cell(Dc.x, Dc.y, Dc.z) >= 0 && sg_cell(Dc.x, Dc.y, Dc.z) < kSgCells);
        const uint4 hc = __ldg(reinterpret_cast<const uint4*>(recc));
        if (a.n_traced) warp_count(a.n_traced + j);
        const float A = j < 64 ? s_A[j] : cmax + __fsqrt_rn(dot3(Vp, Vp));
        const float e = nrmc * 9.5367432e-7f;
        const float th = -(fmaf(3.8146973e-6f, fmaf(A, A, B * B), fmaf(4.f * e, rmax, 16.f * e * e)) + 1e-12f);
        const float no = -dot3(Vp, Dc);
        const float4* ltab = g.ltab + (size_t)j * sc.S;
        const uint32_t cnt = hc.x & 0xffffu;
        const bool full = cnt == 0xffffu || !(nrmc <= sg_range(kSgTiers - 1));     // overflowed cell, origin beyond the last tier or NaN: every sphere
        bool occluded = false;
        // four candidates: their filter operands are requested together, then tested in order (a query stops at its first hit)
        auto test4 = [&](const uint32_t (&cid)[4], uint32_t nvalid) {
            float4 t4[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) t4[q] = __ldg(&ltab[cid[q]]);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if ((uint32_t)q >= nvalid || occluded) continue;
                const float bb = fmaf(t4[q].x, Dc.x, fmaf(t4[q].y, Dc.y, fmaf(t4[q].z, Dc.z, no)));   // k_isect3 qray(), pivot form
                const float qq = fmaf(bb, bb, t4[q].w);
                if (a.slow) ++n_filter;
                if (qq >= th) {
                    const float4 sp = __ldg(&sc.sphB[cid[q]]);
                    const float t = sphere_exact(O, Dc, sp.x, sp.y, sp.z, sp.w);
                    occluded = t > 0.f && t < kTMax;                                               // trace.py:26-29, :95
                    if (a.slow) ++n_exact;
                }
            }
        };
        if (full) {
            for (uint32_t c = 0; c < (uint32_t)sc.S && !occluded; c += 4) {
                const uint32_t cid[4] = { c, min(c + 1, (uint32_t)sc.S - 1u), min(c + 2, (uint32_t)sc.S - 1u), min(c + 3, (uint32_t)sc.S - 1u) };
                test4(cid, (uint32_t)sc.S - c);
            }
        } else if (cnt) {
            {   // candidates 0..6 came with the header
                const uint32_t c0[4] = { hc.x >> 16, hc.y & 0xffffu, hc.y >> 16, hc.z & 0xffffu };
                test4(c0, cnt);
                if (cnt > 4u && !occluded) { const uint32_t c1[4] = { hc.z >> 16, hc.w & 0xffffu, hc.w >> 16, hc.w >> 16 }; test4(c1, min(cnt - 4u, 3u)); }
            }
            for (uint32_t c = 7; c < cnt && !occluded; c += 8) {                                   // candidates c .. c+7 = rec[c+1 .. c+8]
                const uint4 w = __ldg(reinterpret_cast<const uint4*>(recc + 1 + c));
                const uint32_t c0[4] = { w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16 };
                test4(c0, cnt - c);
                if (cnt > c + 4u && !occluded) { const uint32_t c1[4] = { w.z & 0xffffu, w.z >> 16, w.w & 0xffffu, w.w >> 16 }; test4(c1, cnt - c - 4u); }
            }
        }
        for (int pl = 0; pl < sc.Pl && !occluded; ++pl) {                                          // trace.py:32-39
            const float4 po = __ldg(&sc.plO[pl]), pn = __ldg(&sc.plN[pl]);
            // t = num / den > 0 needs num and den of the same sign: where they differ (or num is 0) the quotient is <= 0
            // and plane_exact's division is skipped — the usual case for a point on the plane looking up at a light
            const float den = dot3(Dc, mk(pn.x, pn.y, pn.z));
            const float num = dot3(mk(__fsub_rn(po.x, O.x), __fsub_rn(po.y, O.y), __fsub_rn(po.z, O.z)), mk(pn.x, pn.y, pn.z));
            if ((num > 0.f) != (den > 0.f) || num == 0.f) continue;
            const float t = plane_exact(O, Dc, po, pn);
            occluded = (t > 0.f && t < kTMax);
        }
        if (occluded) continue;                                                                    // :95-96
        r = fmaf(kc, col.x, r); gg = fmaf(kc, col.y, gg); b = fmaf(kc, col.z, b);                  // :102
    }
    a.acc[slot] = make_float4(fmaf(a.weight, r, ar), fmaf(a.weight, gg, ag), fmaf(a.weight, b, ab), 0.f);     // :131
}

// The lights loop of trace() for one hit record per thread (trace.py:77, 86-102, accumulated as in :131):
//   RGB = amb * col;  for each light j:  Lv = normalize(light_j - P);  shadowed = query(P, Lv);  k = lamb * (Lv . N);
//   if not shadowed and k > 0:  RGB += k * col;        acc[slot] += weight * RGB
// The query is only run where its answer is used (k > 0), through light j's direction grid.  Lights in reference order
// and one fmaf per channel per light: the sums are those of k_shade.  (k_setup then builds no per-light lists and no
// occlusion flags exist: this kernel replaces k_isect3<shadow> + k_shade of the scan path.)
__global__ void __launch_bounds__(256, B200RT_SS_OCC) k_shadow_shade(SceneDev sc, ShadowGridDev g, ShadowGridArgs a) {
    if (a.tl && threadIdx.x == 0) atomicMin(a.tl, globaltimer_ns());
    const uint32_t n = *a.n_hit;
    const float cmax = a.bounds[0], rmax = a.bounds[1];
    __shared__ float s_A[64];                                  // per light: cmax + |V_j| (the filter threshold's A), first 64 lights
    if ((int)threadIdx.x < min(sc.L, 64)) { const float4 vv = __ldg(&sc.lights[threadIdx.x]); s_A[threadIdx.x] = cmax + __fsqrt_rn(dot3(mk(vv.x, vv.y, vv.z), mk(vv.x, vv.y, vv.z))); }
    __syncthreads();
    unsigned long long n_filter = 0, n_exact = 0;
    for (uint32_t h = blockIdx.x * blockDim.x + threadIdx.x; h < n; h += gridDim.x * blockDim.x) {
        const float4 p = a.hP[h], nn = a.hN[h];
        shadow_shade_hit(sc, g, a, s_A, cmax, rmax, mk(p.x, p.y, p.z), mk(nn.x, nn.y, nn.z), __float_as_int(nn.w), __float_as_uint(p.w), n_filter, n_exact);
    }
    if (a.slow) {
#pragma unroll
        for (int dd = 16; dd > 0; dd >>= 1) { n_filter += __shfl_xor_sync(0xffffffffu, n_filter, dd); n_exact += __shfl_xor_sync(0xffffffffu, n_exact, dd); }
        if (lane_id() == 0) { if (n_filter) atomicAdd(a.slow + 2, n_filter); if (n_exact) atomicAdd(a.slow + 1, n_exact); }
    }
    if (a.tl && threadIdx.x == 0) atomicMax(a.tl + 1, globaltimer_ns());
}

}  // namespace b200rt
