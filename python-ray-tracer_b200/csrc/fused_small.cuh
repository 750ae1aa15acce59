// fused_small.cuh — one kernel per reflection depth for scenes with few spheres (S < 64: the reference's own default scene, BASELINE
// configs C1, C2, C3, C5).  reference trace.py:44-112 (`trace`) for every live path of the depth, plus kernels.py:16-55 (primary ray) at depth 0.
//
// Why.  With 6 spheres the wavefront's stages (k_raygen, k_isect3, k_setup, k_shadow_shade) each move the whole path state through HBM and
// cost 15-20 us of fixed latency per launch even for the 0.5 M paths of C1; a query is 7 exact tests, there is nothing to cull, and the
// conservative filter, candidate queues, light grids and ordered compaction of the many-sphere path only add instructions.  Here one
// thread takes a path through the whole `trace` of this depth — closest hit, hit point / normal / bias, the lights loop with its shadow
// queries, the weighted accumulate, the reflected ray — with the ray in registers; per depth the only traffic is the queue entry in
// (32 B), the accumulator (2 x 16 B) and the next queue entry out (32 B).
//
// Arithmetic is that of the wavefront kernels, operation for operation (same helpers, same order), so the image is byte-identical to it and
// to brute force (tests: B200RT_F_NO_FUSED / B200RT_F_DENSE / B200RT_F_NO_FILTER):
//   closest hit   sphere_exact on every sphere, smallest t, lowest index on a tie, then planes with strict `<`   (k_isect3 / trace.py:7-41)
//   hit record    P = O + tD, N, Pb = P + BIAS N                                                                  (k_setup / trace.py:56-83)
//   lights        Lv = (V - Pb) / |V - Pb|, k = lamb (Lv.N); if k > 0: shadow query = any sphere / plane hit in (0, 999)   (k_shadow_shade / :86-102)
//   accumulate    acc[slot] += refl^d RGB;  next ray R = normalize(D - 2 (D.N) N), origin Pb + BIAS R by the reader (:104-110, :131)
// The next queue is filled through a warp-aggregated atomic (unordered: nothing downstream depends on the order — every path accumulates
// into its own slot, once per depth, in depth order).
#pragma once
#include "isect3.cuh"

namespace b200rt {

struct FusedArgs {
    const float4* rayO; const float4* rayD; const float* camo;    // this depth's queue (rt_math.cuh: ray_origin); unused at depth 0 when `primary`
    const uint32_t* n_in;                                          // queue length (depth >= 1)
    float4* hP2; float4* rayD2;                                    // next depth's queue: (Pb, slot), (R, -)
    uint32_t* n_hit;                                               // hits of this depth = length of the next queue (zeroed before the launch)
    uint32_t* n_primary;                                           // depth 0: number of samples the reference traces (stats)
    float4* acc;
    float amb, lamb, weight;
    int emit_reflection;
    uint32_t* n_traced;                                            // diagnostics: [L] shadow queries run per light; may be null
    unsigned long long* tl;                                        // diagnostic timeline slot; may be null
};

// `PRIMARY`: depth 0 — the ray of sample slot i is generated here (PrimaryFn: slot -> valid, D), no queue is read.
template <bool PRIMARY, class PrimaryFn>
__device__ __forceinline__ void fused_depth_body(const SceneDev& sc, const FusedArgs& a, uint32_t n, PrimaryFn primary) {
    const uint32_t lane = lane_id();
    uint32_t n_prim = 0;
    for (uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; i0 < n; i0 += gridDim.x * blockDim.x) {      // whole warps stay in the loop (ballots below)
        const uint32_t i = i0 + lane;
        bool live = i < n;
        V3 O = mk(0.f, 0.f, 0.f), D = mk(0.f, 0.f, 0.f); uint32_t slot = 0;
        if (live) {
            if (PRIMARY) { slot = i; live = primary(i, D); O = mk(a.camo[0], a.camo[1], a.camo[2]); if (live) ++n_prim; }
            else { const float4 d4 = a.rayD[i]; D = mk(d4.x, d4.y, d4.z); O = ray_origin(a.rayO, a.camo, i, D, d4.w, slot); }
        }
        bool hit = false; V3 Pb = mk(0.f, 0.f, 0.f), R = mk(0.f, 0.f, 0.f);
        if (live) {
            // ---- closest hit (trace.py:7-41)
            float best = kTMax; int id = -1;
            for (int s = 0; s < sc.S; ++s) {
                const float4 sp = __ldg(&sc.sphB[s]);
                const float t = sphere_exact(O, D, sp.x, sp.y, sp.z, sp.w);
                if (t > 0.f && t < best) { best = t; id = s; }                                       // :26-29 (first sphere wins a tie)
            }
            for (int pl = 0; pl < sc.Pl; ++pl) {
                const float t = plane_exact(O, D, __ldg(&sc.plO[pl]), __ldg(&sc.plN[pl]));
                if (t > 0.f && t < best) { best = t; id = sc.S + pl; }                               // :32-39
            }
            hit = id >= 0;
            if (hit) {
                // ---- hit record (trace.py:56-83)
                const V3 P = madd3(O, best, D);
                V3 N; float4 col;
                if (id < sc.S) { const float4 c = __ldg(&sc.sphB[id]); N = normalize3(sub3(P, mk(c.x, c.y, c.z))); col = __ldg(&sc.sphCol[id]); }
                else           { const float4 nu = __ldg(&sc.plNu[id - sc.S]); N = mk(nu.x, nu.y, nu.z); col = __ldg(&sc.plCol[id - sc.S]); }
                Pb = madd3(P, kBias, N);
                // ---- lights (trace.py:77, 86-102)
                float r = __fmul_rn(a.amb, col.x), g = __fmul_rn(a.amb, col.y), b = __fmul_rn(a.amb, col.z);
                for (int j = 0; j < sc.L; ++j) {
                    const float4 vv = __ldg(&sc.lights[j]);
                    const V3 dl = sub3(mk(vv.x, vv.y, vv.z), Pb);
                    // Only lights with k = lamb * (Lv . N) > 0 matter (:101).  The sign of k is the sign of dl . N whenever that is clearly away from
                    // zero (|cos| > 1e-4: the normalised dot differs from (dl.N)/|dl| by a few ulp of 1), so clearly unlit lights skip the
                    // square root and the three divisions; grazing cases take the exact arithmetic below.
                    const float dn_ = dot3(dl, N), dd_ = dot3(dl, dl);
                    if (a.lamb > 1e-20f && dn_ * dn_ > 1e-8f * dd_ && dd_ < 1e30f && !(dn_ > 0.f)) continue;
                    const float nrm = __fsqrt_rn(dd_);
                    const V3 Lv = mk(__fdiv_rn(dl.x, nrm), __fdiv_rn(dl.y, nrm), __fdiv_rn(dl.z, nrm));
                    const float k = __fmul_rn(a.lamb, dot3(Lv, N));
                    if (!(k > 0.f)) continue;                                                        // :101 (the query's answer would not be used)
                    if (a.n_traced) { const uint32_t m = __activemask(); if ((int)lane == __ffs(m) - 1) atomicAdd(a.n_traced + j, (uint32_t)__popc(m)); }
                    bool occluded = false;
                    for (int s = 0; s < sc.S && !occluded; ++s) {
                        const float4 sp = __ldg(&sc.sphB[s]);
                        const float t = sphere_exact(Pb, Lv, sp.x, sp.y, sp.z, sp.w);
                        occluded = t > 0.f && t < kTMax;                                             // :92-96
                    }
                    for (int pl = 0; pl < sc.Pl && !occluded; ++pl) {
                        const float4 po = __ldg(&sc.plO[pl]), pn = __ldg(&sc.plN[pl]);
                        // t = num / den > 0 needs num and den of the same sign: otherwise plane_exact's division is skipped (a point on the plane looking up at a light)
                        const float den = dot3(Lv, mk(pn.x, pn.y, pn.z));
                        const float num = dot3(mk(__fsub_rn(po.x, Pb.x), __fsub_rn(po.y, Pb.y), __fsub_rn(po.z, Pb.z)), mk(pn.x, pn.y, pn.z));
                        if ((num > 0.f) != (den > 0.f) || num == 0.f) continue;
                        const float t = plane_exact(Pb, Lv, po, pn);
                        occluded = t > 0.f && t < kTMax;
                    }
                    if (occluded) continue;
                    r = fmaf(k, col.x, r); g = fmaf(k, col.y, g); b = fmaf(k, col.z, b);             // :102
                }
                const float4 a0 = a.acc[slot];
                a.acc[slot] = make_float4(fmaf(a.weight, r, a0.x), fmaf(a.weight, g, a0.y), fmaf(a.weight, b, a0.z), 0.f);     // :131
                if (a.emit_reflection) { const float dn = dot3(D, N); R = normalize3(madd3(D, -2.0f * dn, N)); }             // common.py:113-120
            }
        }
        // ---- the hit list = next queue (count kept even on the last depth: ray statistics)
        const uint32_t hm = __ballot_sync(0xffffffffu, hit);
        if (hm) {
            uint32_t base = 0;
            if (lane == (uint32_t)__ffs(hm) - 1u) base = atomicAdd(a.n_hit, (uint32_t)__popc(hm));
            base = __shfl_sync(0xffffffffu, base, __ffs(hm) - 1);
            if (hit && a.emit_reflection) {
                const uint32_t h = base + (uint32_t)__popc(hm & ((1u << lane) - 1u));
                a.hP2[h] = make_float4(Pb.x, Pb.y, Pb.z, __uint_as_float(slot));
                a.rayD2[h] = make_float4(R.x, R.y, R.z, 0.f);
            }
        }
    }
    if (PRIMARY) {
        n_prim = __reduce_add_sync(0xffffffffu, n_prim);
        if (lane == 0 && n_prim) atomicAdd(a.n_primary, n_prim);
    }
}

}  // namespace b200rt
