// isect3.cuh — the production intersection kernel k_isect3 (brute-force closest-hit / any-hit queries of a ray queue
// against all spheres and planes; reference trace.py:7-41, intersections.py:6-68) and its launcher.
//
// Included by TWO translation units on purpose: b200rt.cu instantiates the closest-hit modes (0, 1) and
// isect_shadow.cu the shadow mode (2).  Without relocatable device code every .cu file is its own CUDA module with its
// own 64 KB constant bank, so the closest-hit tables (camera pivot + origin-free) and the lights' tables are resident
// at the same time: for scenes up to kTabCap/L spheres nothing is uploaded between launches, and closest-hit and
// shadow launches of different depths are independent of each other (they run on two streams, see render_device).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <algorithm>
#include <atomic>
#include "rt_math.cuh"

namespace b200rt {

extern std::atomic<unsigned long long> g_kernel_launches;   // kernels launched by this library since load (b200rt_launch_count)

// ------------------------------------------------------------------------------------------------
// Device-side scene (packed once per launch by k_pack from the reference's SoA arrays)
// ------------------------------------------------------------------------------------------------
struct SceneDev {
    const float4* sphA;    // [Spad] (Cx, Cy, Cz, r^2 - |C|^2)  filter operands; padding entries never pass
    const float4* sphB;    // [S]    (Cx, Cy, Cz, r^2)          exact-test operands; r^2 = float32 r*r (intersections.py:21)
    const float4* sphCol;  // [S]    (R, G, B, 0)
    const float4* plO;     // [Pl]   plane point
    const float4* plN;     // [Pl]   stored normal (used un-renormalised by the intersection, intersections.py:52)
    const float4* plNu;    // [Pl]   float32 re-normalised normal (common.py:104-110)
    const float4* plCol;   // [Pl]
    const float4* lights;  // [L]
    const float*  cmax;    // [2]    max |C|, max r over spheres (filter error bounds)
    int S, Spad, Pl, L;
};

constexpr unsigned long long kNoHitKey = ((unsigned long long)0x4479c000u << 32) | 0xffffffffu;   // (t = 999.0f, id = -1)
constexpr int kThreads = 256;            // threads per CTA

struct QueueDev {
    // closest-hit input: ray queue.  rayD = (D.xyz, slot bits at depth 0); rayO = the biased hit points (Pb.xyz, slot) of the depth before, from
    // which the origin is Pb + BIAS * D, or null at depth 0: origin = the camera constant camo[0..2] (rt_math.cuh: ray_origin)
    const float4* rayO; const float4* rayD; const float* camo;
    // shadow input: hit records
    const float4* hP;                    // (biased hit point, slot)
    const uint32_t* n_in;                // number of rays (closest) / entries of the light's list = input of its FIRST pass (shadow)
    uint32_t* work;                      // dynamic group counter (zeroed before the launch)
    unsigned long long* hit;             // closest-hit output, queue order: (float bits of t) << 32 | id; pre-set to kNoHit
    uint8_t* occl; size_t ocap;          // shadow output, [light][hit] with row stride ocap; pre-set to 0 (k_isect3 stores only 1s)
    const uint32_t* in_list;             // k_isect3 shadow input: hit records to query for this light (null: every hit record, identity map)
    // k_isect3 shadow passes: a query scans the light's table (spheres sorted by decreasing solid angle seen from the
    // light) in several launches over consecutive ranges; rays found occluded are done, the others are appended to
    // out_list (hit-record indices) = the next pass's input.  To keep neighbouring rays together without a global
    // ordered compaction, survivors are compacted per REGION: the kRegion consecutive entries [sg*kRegion, (sg+1)*kRegion)
    // of the first pass's list keep that slot range in every later list, filled from its start (out_rn[sg] = fill
    // count, zeroed before the frame).  A later pass (in_rn != null) walks the regions' filled parts.  Null on the last pass.
    uint32_t* out_list; uint32_t* out_rn; const uint32_t* in_rn;
    int last;                            // this launch covers the end of the sphere array: planes are tested here (set by the launcher)
    int only;                            // closest hit: this launch covers ALL spheres, so it stores the key (no pre-set "no hit", no atomic merge)
    unsigned long long* slow;            // diagnostic: thread-level slow-path entries (may be null)
    unsigned long long* tl;              // diagnostic (B200RT_TIMELINE): [0] first CTA start, [1] last CTA end, globaltimer ns (may be null)
};

// ------------------------------------------------------------------------------------------------
// The production intersection kernel (k_isect3)
// ------------------------------------------------------------------------------------------------
// What bounds this loop on B200 is not the FP32 pipe but the register file: measured (tools/ffma2_probe*.cu) an
// SMSP sustains ~1.75 register-operand reads per lane-cycle, every FFMA2 that reads two register pairs costs 2.3
// cycles, and a 3-input ALU op (FMNMX3) costs as much as an FFMA2 when its operands all come from registers.  So the
// loop is organised to read as few vector registers as possible:
//   * packed FFMA2 (fma.rn.f32x2): two IEEE FP32 FMAs per instruction.  The two lanes of every packed value are two
//     SPHERES (2i, 2i+1) tested against the same ray;
//   * the sphere operands are warp-uniform, so they never touch the vector register file: the table of the launch
//     lives in constant memory and reaches the FMA pipe as UNIFORM-register pairs (LDCU.128 -> FFMA2 ..., URx.F32x2);
//   * the ray operands are 32-bit registers in FFMA2's broadcast form (R.F32): 1 register read instead of a pair.
// Per (ray, sphere pair) the 4-FFMA pivot form reads 2+3+3+2 registers + 3 for the max — the probe's V7 variant runs
// this at 2.32 cycles/FFMA2 = 64 TFLOP/s executed, 86 % of the nominal FP32 peak, FMA-pipe bound.
//
// Filter algebra (unchanged from the scalar kernel, see k_intersect): disc >= 0 <=> q >= thr.
//   MODE 0 (any ray, bounces):     q = b^2 + 2 C.O + (r^2 - |C|^2) - (|O|^2 - margin)          8 FFMA2 / 2 tests
//   MODE 1/2 (pivot V on the line): q = b^2 + (r^2 - |C-V|^2),  b = C.D - V.D,  thr = -margin   4 FFMA2 / 2 tests
// Pivot form: disc(line, sphere) does not depend on which point of the line is the origin.  All primary rays start
// at the camera and every shadow ray towards light j lies (to within the rounding of its direction) on a line through
// that light, so c = |C-V|^2 - r^2 is a per-(pivot, sphere) constant tabulated once per frame by k_pivot in float64-
// rounded-once form.  The filter only SELECTS candidates (margin bounds its own rounding, the exact test's rounding
// and the distance between the true ray line and the pivot); sphere_exact() on the real ray decides, so the output is
// bit-identical to running the exact test on every sphere (B200RT_F_NO_FILTER, tests/test_gpu_parity.py).
constexpr int kTabCap = 3584;            // float4 entries of the constant-memory table = 3584 spheres (56 KB)
static __constant__ float4 c_tab[kTabCap + 2 * 4]; // per sphere pair i: [2i] = (cx0,cx1,cy0,cy1)  [2i+1] = (cz0,cz1,k0,k1); + prefetch overrun

struct PivotDev {
    const float4* tab;                   // [(L+2)][Spad] pair layout; table j < L: light j, j == L: camera, j == L+1: origin-free
    const float4* piv;                   // [(L+1)]       pivot positions (float32)
    const float*  bounds;                // [2]           max |C|, max r
    const int*    perm;                  // [L+1][Spad]   table position -> sphere index: rows 0..L-1 the lights' tables (k_rank), row L both closest-hit tables (k_mrank)
    const float*  rad;                   // [(L+2)][Spad] radius of the sphere at each table position (beam culling), or null = no culling
    const float4* clus;                  // [ceil(Spad/64)] bounding spheres of 64 consecutive positions of the closest-hit tables (k_clus), or null
};

__device__ __forceinline__ unsigned long long globaltimer_ns() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 bc(float a) { return make_float2(a, a); }       // FFMA2 broadcast operand (R.F32 / UR.F32)
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float d; asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d;   // FMNMX3 (sm_100+)
}

// MODE 0: closest hit, arbitrary rays (bounces)   MODE 1: closest hit, camera pivot (primary)   MODE 2: any hit, shadow
// rays of light j.  NR = rays per thread.  The launch covers the sphere chunk [c0, c0+cnt) (cnt % 8 == 0) whose table
// the host copied to c_tab[tb ...].  Work distribution: every WARP claims groups of 32*NR consecutive rays from a
// global counter (one per light) — no CTA-wide barrier anywhere.  One launch per (light, sphere chunk); with more than
// one chunk the running (t, id) / occlusion state is carried in the output arrays.
//
// Candidates (filter passes, ~1.5 per ray out of S) are not tested where they are found: a warp appends them to its
// own shared-memory queue and runs the exact test over the queue with all 32 lanes busy ("flush"), when the queue
// fills and at the end of the group.  Results meet in shared memory: closest hit = atomicMin on the packed key
// (float bits of t << 32 | sphere id) — min t, and on equal t the lowest index, which is the reference's tie-break
// (first object wins, trace.py:26; planes have ids >= S and so lose ties against spheres, trace.py:36).
#ifndef B200RT_KP
#define B200RT_KP 2
#endif
constexpr int KP = B200RT_KP;            // sphere pairs per scan step, pivot modes
#ifndef B200RT_KP0
#define B200RT_KP0 2                     // ... of the 8-FFMA2 form (MODE 0): its step is twice as long, and the unrolled step should fit the L0 instruction cache
#endif
#ifndef B200RT_CULL_E
#define B200RT_CULL_E 0                  // beam culling: sphere pairs per lane per round (code size / registers vs loads in flight; the next round is
                                         // prefetched).  0: 2 with 8 rays per thread, 1 with fewer (85 registers: C4 12.78 -> 12.69 ms; 3 and 4 spill)
#endif
#ifndef B200RT_FLUSH_UNROLL
#define B200RT_FLUSH_UNROLL 1            // 1: #pragma unroll 1 on the exact-test loop (0: the compiler unrolls it by 2 + tail at every inlined site)
#endif
constexpr uint32_t kRegion = 2048;       // entries per survivor-list region (8 ray groups of 256)
constexpr int kQCap = 768;               // candidate queue entries per warp
constexpr int kQFlush = kQCap - 32 * 16; // flush before an append (<= 16 per lane) could overflow
constexpr unsigned long long kNoHit = kNoHitKey;                                                 // (999.0f, id -1)

template <int NR> struct IsectWarpSmem {               // per warp (no CTA-level cooperation anywhere in k_isect3)
    float4 d[32 * NR];                                 // per ray of the group: D.xyz, -V.D (or -O.D)
    float4 o[32 * NR];                                 // MODE 0: 2*O.xyz, -thr   MODE 2: O.xyz, hit-record index (bits)
    float thr[32 * NR];                                // pivot modes: the ray's filter threshold
    uint32_t kt[32 * NR];                              // float bits of the closest t / occlusion flag
    uint32_t ki[32 * NR];                              // ... and the id of the closest object
    uint32_t q[kQCap];                                 // candidate queue
};
template <int NR> struct IsectSmem { IsectWarpSmem<NR> w[kThreads / 32]; };   // NR = 8: 14 KB per warp, 112 KB per CTA, 2 CTAs per SM

// DIAG: the counters / timeline behind q.slow and q.tl (b200rt_stats, B200RT_TIMELINE) exist only in the <.., true>
// instantiation, which the launcher picks when either pointer is set: in the production kernel they were ~200 cold
// instructions spread through the hot code, and this kernel's first limiter is instruction fetch (profiles/r01_v7).
template <int MODE, int NR, bool DIAG>
__global__ void __launch_bounds__(kThreads, NR == 8 ? 2 : 3)
k_isect3(SceneDev sc, QueueDev q, PivotDev pv, int c0, int cnt, int j, int tb) {
    constexpr bool ANY = (MODE == 2), PIV = (MODE != 0);
    constexpr int GW = 32 * NR;                               // rays per warp group
    constexpr int NO = PIV ? 1 : NR;                          // 2*O operands exist only in the 8-FFMA2 form
    constexpr int KPM = (MODE == 0) ? B200RT_KP0 : KP;        // sphere pairs per scan step of this mode
    constexpr uint32_t kQFl = (NR == 8) ? (uint32_t)kQFlush : 512u - 64u * NR;   // flush threshold of the candidate queue (see slist)
    extern __shared__ __align__(16) unsigned char s_raw[];
    IsectSmem<NR>& sm = *reinterpret_cast<IsectSmem<NR>*>(s_raw);
    if (DIAG && q.tl && threadIdx.x == 0) atomicMin(q.tl, globaltimer_ns());
    const bool last = q.last != 0;
    const uint32_t n_in = *q.n_in;
    const uint32_t* shl = ANY ? q.in_list : nullptr;          // list entry -> hit record
    const bool regions = ANY && q.in_rn != nullptr;
    const uint32_t ngroups = regions ? ((n_in + kRegion - 1) / kRegion) * (kRegion / GW) : (n_in + GW - 1) / GW;   // work items
    const int npair = cnt >> 1;
    const float cmax = pv.bounds[0], rmax = pv.bounds[1];
    const float kInf = __int_as_float(0x7f800000);
    const uint32_t lane = lane_id(), wid = threadIdx.x >> 5;
    IsectWarpSmem<NR>& ws = sm.w[wid];
    uint32_t* sq = ws.q;
    uint32_t* skt = ws.kt; uint32_t* ski = ws.ki;
    uint32_t slow_cnt = 0, exact_cnt = 0;
    unsigned long long tests_cnt = 0;                         // diagnostic: filter tests executed by this thread's rays (padding and inactive rays included)
    // j: pivot of this launch (light index for ANY, camera = L for MODE 1); its table is c_tab[tb + 2*pair + {0,1}].
    // tb is a kernel parameter so that the table index provably lives in uniform registers (LDCU -> UR operands).
    V3 Vp = mk(0.f, 0.f, 0.f);
    if (PIV) { const float4 v = __ldg(&pv.piv[j]); Vp = mk(v.x, v.y, v.z); }
    uint32_t* work = q.work;

    auto run_group = [&](const uint32_t g) {
        uint32_t gbase = g * (uint32_t)GW, nlim = n_in;       // this group's rays: list entries [gbase, min(gbase + GW, nlim))
        if (regions) {
            const uint32_t sg = g / (kRegion / GW), k = g % (kRegion / GW);
            gbase = sg * kRegion + k * (uint32_t)GW; nlim = sg * kRegion + q.in_rn[sg];
            if (__all_sync(0xffffffffu, gbase >= nlim)) return;
        }

        // ---- per-ray constants.  Computed ray by ray by ONE copy of the code (iteration k: ray k*32 + lane of the
        // group) and staged in shared memory: unrolled over the thread's rays this prologue was a third of the
        // kernel's instructions and its instruction-cache misses showed in short launches.  The staged copy also
        // serves the exact tests and the write-back, which therefore read no ray data from global memory. ----
        // Phase A issues every global load of the group before anything waits on one (two dependent loads per
        // shadow ray: list entry -> hit record); the raw records pass through the staging arrays to phase B.
        // The loads are unconditional (index clamped to the group's last ray) so that they are issued back to back.
        {
            const uint32_t ilast = nlim - 1u;                  // nlim > gbase here
            if (MODE == 2) {
                uint32_t hr[NR]; float4 pp[NR];
#pragma unroll
                for (int k = 0; k < NR; ++k) hr[k] = min(gbase + (uint32_t)k * 32u + lane, ilast);
                if (shl) {
#pragma unroll
                    for (int k = 0; k < NR; ++k) hr[k] = shl[hr[k]];
                }
#pragma unroll
                for (int k = 0; k < NR; ++k) pp[k] = q.hP[hr[k]];
#pragma unroll
                for (int k = 0; k < NR; ++k) ws.o[k * 32 + lane] = make_float4(pp[k].x, pp[k].y, pp[k].z, __uint_as_float(hr[k]));
            } else {
                float4 oo[NR], dd[NR];
#pragma unroll
                for (int k = 0; k < NR; ++k) {
                    const uint32_t i = min(gbase + (uint32_t)k * 32u + lane, ilast);
                    dd[k] = q.rayD[i];
                    oo[k] = q.rayO ? q.rayO[i] : make_float4(q.camo[0], q.camo[1], q.camo[2], 0.f);
                }
                if (q.rayO) {                                  // bounce origin from the hit record: Pb + BIAS * R (trace.py:110; ray_origin())
#pragma unroll
                    for (int k = 0; k < NR; ++k) { const V3 o2 = madd3(mk(oo[k].x, oo[k].y, oo[k].z), kBias, mk(dd[k].x, dd[k].y, dd[k].z)); oo[k].x = o2.x; oo[k].y = o2.y; oo[k].z = o2.z; }
                }
#pragma unroll
                for (int k = 0; k < NR; ++k) { ws.o[k * 32 + lane] = oo[k]; ws.d[k * 32 + lane] = dd[k]; }
            }
        }
#pragma unroll 1
        for (int k = 0; k < NR; ++k) {
            const uint32_t rl = (uint32_t)k * 32u + lane, i = gbase + rl;
            const bool active = i < nlim;
            V3 O = mk(0.f, 0.f, 0.f), D = mk(0.f, 0.f, 0.f);
            float th = kInf, no = 0.f;
            if (active) {
                if (MODE == 2) {
                    const float4 p = ws.o[rl];
                    O = mk(p.x, p.y, p.z);
                    const V3 d = sub3(Vp, O);                                   // common.py:84-91
                    const float nrm = __fsqrt_rn(dot3(d, d));
                    D = mk(__fdiv_rn(d.x, nrm), __fdiv_rn(d.y, nrm), __fdiv_rn(d.z, nrm));
                    const float A = cmax + __fsqrt_rn(dot3(Vp, Vp)), B = cmax + __fsqrt_rn(dot3(O, O));
                    const float e = nrm * 9.5367432e-7f;                        // 2^-20 |V-P|: ray line vs pivot
                    th = -(fmaf(3.8146973e-6f, fmaf(A, A, B * B), fmaf(4.f * e, rmax, 16.f * e * e)) + 1e-12f);
                    no = -dot3(Vp, D);
                } else {
                    const float4 o = ws.o[rl], d = ws.d[rl];
                    O = mk(o.x, o.y, o.z); D = mk(d.x, d.y, d.z);
                    const float oo = dot3(O, O);
                    const float A = cmax + __fsqrt_rn(oo);
                    const float margin = fmaf(3.8146973e-6f /* 64 * 2^-24 */, A * A, 1e-12f);
                    if (MODE == 1) { th = -2.f * margin; no = -dot3(Vp, D); }   // O == Vp for every primary ray
                    else           { th = oo - margin;   no = -dot3(O, D); }
                }
            }
            ws.d[rl] = make_float4(D.x, D.y, D.z, no);
            // MODE 0: the per-ray threshold is folded into q (q' = q - thr, one more FFMA2): candidates are q' >= 0
            // (an inactive ray has thr = +inf, so its q' is -inf and never qualifies).
            if (MODE == 0) ws.o[rl] = make_float4(2.f * O.x, 2.f * O.y, 2.f * O.z, -th);
            if (MODE == 2 && !active) ws.o[rl] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (PIV) ws.thr[rl] = th;                          // (MODE 2: ws.o keeps O.xyz and the hit-record index)
            skt[rl] = ANY ? 0u : (uint32_t)(kNoHit >> 32); ski[rl] = 0xffffffffu;
        }
        // the thread's rays r = 0..NR-1 are rays r*32 + lane (32-bit registers, broadcast operands of FFMA2)
        float dx[NR], dy[NR], dz[NR], nod[NR], o2x[NO], o2y[NO], o2z[NO], nthr[NO], thr[PIV ? NR : 1];
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            const float4 d = ws.d[r * 32 + lane];
            dx[r] = d.x; dy[r] = d.y; dz[r] = d.z; nod[r] = d.w;
            if (!PIV) { const float4 o = ws.o[r * 32 + lane]; o2x[PIV ? 0 : r] = o.x; o2y[PIV ? 0 : r] = o.y; o2z[PIV ? 0 : r] = o.z; nthr[PIV ? 0 : r] = o.w; }
            else      thr[PIV ? r : 0] = ws.thr[r * 32 + lane];
        }
        float thr_t = 0.f;                                    // one scan threshold per thread: the loosest of its rays
        if (PIV) {
            thr_t = thr[0];
#pragma unroll
            for (int r = 1; r < NR; ++r) thr_t = fminf(thr_t, thr[PIV ? r : 0]);
        }
        uint32_t qcnt = 0;                                    // candidates waiting in this warp's queue (warp-uniform)

        // q (minus thr in MODE 0) of ray r against sphere pair (u, w): lanes .x/.y = spheres 2i / 2i+1
        auto qray = [&](int r, const float4& u, const float4& w) -> float2 {
            const float2 b = ffma2(make_float2(u.x, u.y), bc(dx[r]), ffma2(make_float2(u.z, u.w), bc(dy[r]),
                                   ffma2(make_float2(w.x, w.y), bc(dz[r]), bc(nod[r]))));
            if (PIV) return ffma2(b, b, make_float2(w.z, w.w));
            const int ro = PIV ? 0 : r;
            const float2 t = ffma2(make_float2(u.x, u.y), bc(o2x[ro]), ffma2(make_float2(u.z, u.w), bc(o2y[ro]),
                                   ffma2(make_float2(w.x, w.y), bc(o2z[ro]), ffma2(make_float2(w.z, w.w), bc(1.0f), bc(nthr[ro])))));
            return ffma2(b, b, t);
        };

        // -- exact tests over the queue, one candidate per lane per iteration (dense) --
        // The closest hit of a ray is the minimum over its candidates of (t, id) — smallest t, lowest id on a tie
        // (trace.py:26).  Two native 32-bit shared atomics instead of a 64-bit atomicMin: the latter is a CAS loop, and
        // with a divergent loop inside the scan ptxas no longer keeps the table operands in uniform registers.
        auto flush = [&]() {
            __syncwarp();
#if B200RT_FLUSH_UNROLL
#pragma unroll 1
#endif
            for (uint32_t base = 0; base < qcnt; base += 32) {                // warp-uniform trip count
                const uint32_t idx = base + lane;
                bool hit = false; uint32_t rl = 0, tbits = 0; int sid = 0;
                if (idx < qcnt) {
                    const uint32_t item = sq[idx];
                    rl = item & 255u; sid = c0 + (int)(item >> 8);
                    sid = __ldg(&pv.perm[(size_t)(ANY ? j : sc.L) * sc.Spad + sid]);               // table position -> sphere
                    const float4 dd = ws.d[rl];                // the ray as staged by the prologue (2*O halves exactly)
                    const V3 D = mk(dd.x, dd.y, dd.z);
                    V3 O = Vp;
                    if (MODE != 1) { const float4 oo = ws.o[rl]; O = (MODE == 0) ? mk(0.5f * oo.x, 0.5f * oo.y, 0.5f * oo.z) : mk(oo.x, oo.y, oo.z); }
                    const float4 sp = __ldg(&sc.sphB[sid]);
                    const float t = sphere_exact(O, D, sp.x, sp.y, sp.z, sp.w);
                    hit = t > 0.f && t < kTMax;                                               // trace.py:26-29
                    tbits = __float_as_uint(t);
                }
                if (ANY) { if (hit) skt[rl] = 1u; __syncwarp(); }
                else {
                    if (hit && tbits < atomicMin(&skt[rl], tbits)) ski[rl] = 0xffffffffu;     // new minimum t: forget the old id
                    __syncwarp();
                    if (hit && skt[rl] == tbits) atomicMin(&ski[rl], (uint32_t)sid);
                    __syncwarp();
                }
            }
            if (DIAG) exact_cnt += (qcnt + 31 - lane) / 32;
            __syncwarp();
            qcnt = 0;
        };

        // -- candidates in sphere pair `pair` (rare): recompute q per ray, append the passing (ray, sphere)
        // combinations to the warp's queue (warp prefix sum of the per-lane counts) --
        auto slow_pair = [&](int pair) {
            const float4 u = c_tab[tb + 2 * pair], w = c_tab[tb + 2 * pair + 1];
            uint32_t cand = 0;
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                const float2 qv = qray(r, u, w);
                const float lim = PIV ? thr[PIV ? r : 0] : 0.f;
                if (qv.x >= lim) cand |= 1u << (2 * r);
                if (qv.y >= lim) cand |= 2u << (2 * r);
            }
            if (c0 + 2 * pair + 1 >= sc.S) cand &= (c0 + 2 * pair >= sc.S) ? 0u : 0x55555555u;     // padding spheres
            const uint32_t n = (uint32_t)__popc(cand);
            uint32_t x = n;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if ((int)lane >= d) x += y; }
            uint32_t pos = qcnt + x - n;
            qcnt += __shfl_sync(0xffffffffu, x, 31);
            while (cand) {
                const int bit = __ffs(cand) - 1;
                cand &= cand - 1;
                RT_CHECK(0, pos < (NR == 8 ? (uint32_t)kQCap : 512u));
                sq[pos++] = ((uint32_t)(bit >> 1) * 32u + lane) | ((uint32_t)(2 * pair + (bit & 1)) << 8);
            }
            if (qcnt > kQFl) flush();
        };
        auto slow_step = [&](int pb, const float (&mk4)[KPM]) {        // pairs pb .. pb+KP-1, per-pair maxima mk4
            if (DIAG) slow_cnt++;
#pragma unroll 1
            for (int k = 0; k < KPM; ++k) {
                float mv = mk4[0];
#pragma unroll
                for (int kk = 1; kk < KPM; ++kk) mv = (k == kk) ? mk4[kk] : mv;
                if (__any_sync(0xffffffffu, mv >= thr_t)) slow_pair(pb + k);
            }
        };

        // ---- beam culling.  The group's rays lie in a "beam": origins within ro of their centroid Oc, directions
        // within kap of their mean Dc, so the point at parameter t of any of them is within ro + kap*t of Oc + t*Dc.  A
        // sphere that stays clear of that swept cone for all t >= 0 cannot be hit by any ray of the group and needs no
        // test at all — for primary rays, shadow rays towards a light and rays reflected off a plane the beam of 256
        // neighbouring rays is narrow and most of the scene is outside it.  The test is conservative (radius inflated by
        // the exact test's rounding bound), so the image does not change; B200RT_F_NO_CULL switches it off.  Survivors
        // are listed per PAIR of table positions and scanned through the list (operands from the constant table with a
        // computed index instead of uniform registers); wide beams and poor cull rates take the dense scan. ----
        if (DIAG && q.slow && lane == 0) atomicAdd(q.slow + 4 + 6 * MODE, 1ull);                      // dbg: groups
        int ns = -1, nsr = 0;                                 // ns >= 0: sparse scan over slist[0..ns), the first nsr entries real
        // The pair list needs 2 B x npair (<= 1 KB).  NR = 8: an array that is free in the mode (1 KB each).  Fewer rays per
        // thread: those arrays are smaller, so the list takes the top third of the candidate queue, whose remaining
        // 512 entries still hold one slow_pair's worst case (32 lanes x 2 NR) beyond the flush threshold.
        uint16_t* slist = (NR == 8) ? reinterpret_cast<uint16_t*>(MODE == 2 ? (void*)ws.ki : MODE == 0 ? (void*)ws.thr : (void*)ws.o)
                                    : reinterpret_cast<uint16_t*>(ws.q + 512);
        if (pv.rad && npair <= 512 && npair >= 64) {
            float ax = 0.f, ay = 0.f, az = 0.f, bx = 0.f, by = 0.f, bz = 0.f, na = 0.f;
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                if (gbase + (uint32_t)r * 32u + lane >= nlim) continue;
                na += 1.f; bx += dx[r]; by += dy[r]; bz += dz[r];
                if (MODE == 0) { ax += 0.5f * o2x[PIV ? 0 : r]; ay += 0.5f * o2y[PIV ? 0 : r]; az += 0.5f * o2z[PIV ? 0 : r]; }
                if (MODE == 2) { const float4 oo = ws.o[r * 32 + lane]; ax += oo.x; ay += oo.y; az += oo.z; }
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                na += __shfl_xor_sync(0xffffffffu, na, d);
                bx += __shfl_xor_sync(0xffffffffu, bx, d); by += __shfl_xor_sync(0xffffffffu, by, d); bz += __shfl_xor_sync(0xffffffffu, bz, d);
                if (MODE != 1) { ax += __shfl_xor_sync(0xffffffffu, ax, d); ay += __shfl_xor_sync(0xffffffffu, ay, d); az += __shfl_xor_sync(0xffffffffu, az, d); }
            }
            const float bn = __fsqrt_rn(bx * bx + by * by + bz * bz);
            if (__all_sync(0xffffffffu, bn > 0.25f * na)) {    // (mean direction well defined; na >= 1 in a claimed group; votes keep the control flow provably uniform)
                const float ib = 1.f / bn, ia = 1.f / na;
                const V3 Dc = mk(bx * ib, by * ib, bz * ib);
                const V3 Oc = (MODE == 1) ? Vp : mk(ax * ia, ay * ia, az * ia);
                float ro2 = 0.f, kap2 = 0.f;
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    if (gbase + (uint32_t)r * 32u + lane >= nlim) continue;
                    const float ex = dx[r] - Dc.x, ey = dy[r] - Dc.y, ez = dz[r] - Dc.z;
                    kap2 = fmaxf(kap2, ex * ex + ey * ey + ez * ez);
                    if (MODE == 0) { const float fx = 0.5f * o2x[PIV ? 0 : r] - Oc.x, fy = 0.5f * o2y[PIV ? 0 : r] - Oc.y, fz = 0.5f * o2z[PIV ? 0 : r] - Oc.z; ro2 = fmaxf(ro2, fx * fx + fy * fy + fz * fz); }
                    if (MODE == 2) { const float4 oo = ws.o[r * 32 + lane]; const float fx = oo.x - Oc.x, fy = oo.y - Oc.y, fz = oo.z - Oc.z; ro2 = fmaxf(ro2, fx * fx + fy * fy + fz * fz); }
                }
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) { ro2 = fmaxf(ro2, __shfl_xor_sync(0xffffffffu, ro2, d)); kap2 = fmaxf(kap2, __shfl_xor_sync(0xffffffffu, kap2, d)); }
                const float kap = __fsqrt_rn(kap2) * 1.0001f + 1e-6f;
                if (DIAG && q.slow && lane == 0) { atomicAdd(q.slow + 5 + 6 * MODE, 1ull); if (ANY && !regions) { atomicAdd(q.slow + 22, 1ull); if (kap < 0.1f) atomicAdd(q.slow + 23, 1ull); } if (MODE != 2) { if (kap < 0.1f) atomicAdd(q.slow + 8 + 6 * MODE, 1ull); if (kap < 0.1f && ro2 < 1.f) atomicAdd(q.slow + 9 + 6 * MODE, 1ull); } }
                if (__all_sync(0xffffffffu, kap < 0.6f)) {
                    const float osz = cmax + __fsqrt_rn(dot3(Oc, Oc));
                    const float ro = __fsqrt_rn(ro2) * 1.0001f + 1e-6f * osz;
                    // slack of the exact test (it may report a hit up to ~16 eps B^2 outside the sphere) + this test's own rounding
                    const float m = 2.5e-5f * (osz + ro) * (osz + ro) + 1e-9f;
                    const float kk = 1.f - kap * kap;
                    const int tj = ANY ? j : (MODE == 1 ? sc.L : sc.L + 1);
                    const float4* gt = pv.tab + (size_t)tj * sc.Spad + c0;
                    const float2* gr = reinterpret_cast<const float2*>(pv.rad + (size_t)tj * sc.Spad + c0);
                    // sqrt(Rp^2 + m) <= Rp + sqrt(m): one square root per group instead of one per sphere (the bound only
                    // enters through kap * ..., i.e. it loosens the test by a few per cent of a radius for narrow beams)
                    const float sm_ = __fsqrt_rn(m) * 1.0001f;
                    auto clear_of_beam = [&](float cx, float cy, float cz, float rad) -> bool {
                        const float wx = cx - Oc.x, wy = cy - Oc.y, wz = cz - Oc.z;
                        const float w2 = wx * wx + wy * wy + wz * wz, a = wx * Dc.x + wy * Dc.y + wz * Dc.z;
                        const float Rp = rad + ro, Rp2 = (Rp * Rp + m) * 1.0002f;
                        const float tn = a + kap * (Rp + sm_) * 1.0004f;
                        const float lhs = w2 - Rp2;
                        return (tn < 0.f) ? (lhs > 0.f) : (lhs * kk > tn * tn * 1.0002f + 1e-3f);
                    };
                    uint32_t nsv = 0;
                    // CE pairs per lane per round; the loads of round r+1 are issued before round r is evaluated
                    constexpr int CE = B200RT_CULL_E ? B200RT_CULL_E : (NR == 8 ? 2 : 1);
                    float4 uu[CE], ww[CE]; float2 rr[CE];
                    auto load_round = [&](int p0) {
#pragma unroll
                        for (int e = 0; e < CE; ++e) {
                            const int p = min(p0 + e * 32 + (int)lane, npair - 1);
                            uu[e] = __ldg(gt + 2 * p); ww[e] = __ldg(gt + 2 * p + 1); rr[e] = __ldg(gr + p);
                        }
                    };
                    // Closest-hit tables are in Morton order (k_mrank): 64 consecutive positions = 32 pairs form a compact
                    // cluster with a bounding sphere (k_clus).  Lane c tests cluster c against the beam first; rounds whose
                    // clusters are all clear of it are skipped entirely (no loads, no per-sphere tests).
                    uint32_t rmask = 0xffffffffu;                              // bit r: round r (pairs [32 CE r, 32 CE (r+1))) has to be looked at
                    if (!ANY && pv.clus && c0 == 0) {
                        const int ncl = (npair + 31) >> 5;                     // <= 16 (npair <= 512)
                        bool keepc = false;
                        if ((int)lane < ncl) { const float4 cb = __ldg(&pv.clus[lane]); keepc = cb.w >= 0.f && !clear_of_beam(cb.x, cb.y, cb.z, cb.w); }
                        const uint32_t cm = __ballot_sync(0xffffffffu, keepc);
                        rmask = (CE == 1) ? cm : 0u;
                        if (CE > 1) for (int r = 0; r * CE < ncl; ++r) if ((cm >> (r * CE)) & ((1u << CE) - 1u)) rmask |= 1u << r;
                    }
                    rmask &= (npair + 32 * CE - 1) / (32 * CE) >= 32 ? 0xffffffffu : (1u << ((npair + 32 * CE - 1) / (32 * CE))) - 1u;
                    if (rmask) load_round((__ffs(rmask) - 1) * 32 * CE);
#pragma unroll 1
                    while (rmask) {
                        const int p0 = (__ffs(rmask) - 1) * 32 * CE;
                        rmask &= rmask - 1u;
                        float4 cu[CE], cw[CE]; float2 cr[CE];
#pragma unroll
                        for (int e = 0; e < CE; ++e) { cu[e] = uu[e]; cw[e] = ww[e]; cr[e] = rr[e]; }
                        if (rmask) load_round((__ffs(rmask) - 1) * 32 * CE);
#pragma unroll
                        for (int e = 0; e < CE; ++e) {
                            const int p = p0 + e * 32 + (int)lane;
                            const bool keep = p < npair &&
                                ((c0 + 2 * p < sc.S && !clear_of_beam(cu[e].x, cu[e].z, cw[e].x, cr[e].x)) ||
                                 (c0 + 2 * p + 1 < sc.S && !clear_of_beam(cu[e].y, cu[e].w, cw[e].y, cr[e].y)));
                            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
                            RT_CHECK(1, !keep || nsv + (uint32_t)__popc(bal & ((1u << lane) - 1u)) < 512u);
                            if (keep) slist[nsv + (uint32_t)__popc(bal & ((1u << lane) - 1u))] = (uint16_t)p;
                            nsv += (uint32_t)__popc(bal);
                        }
                    }
                    if (DIAG && q.slow && lane == 0) { atomicAdd(q.slow + 6 + 6 * MODE, (unsigned long long)nsv); atomicAdd(q.slow + 7 + 6 * MODE, (unsigned long long)npair); }
                    if (__all_sync(0xffffffffu, 2 * (int)nsv <= npair)) {      // worth it (the list path costs ~1.5x per test)
                        __syncwarp();
                        const uint16_t padv = nsv ? slist[nsv - 1] : (uint16_t)0;      // padding: the last pair again (never queued twice)
                        if (lane < (KPM - nsv % KPM) % KPM) slist[nsv + lane] = padv;
                        nsr = (int)nsv; ns = (int)((nsv + KPM - 1) / KPM * KPM);
                        __syncwarp();
                    }
                }
            }
        }
        if (__all_sync(0xffffffffu, ns >= 0)) {
            if (DIAG && q.slow) tests_cnt += (unsigned long long)ns * 2ull * (unsigned long long)NR;
#pragma unroll 1
            for (int pb = 0; pb < nsr; ++pb) {                 // one listed pair per iteration: small code, the lists are short
                const int pi = slist[pb];
                const float4 u = c_tab[tb + 2 * pi], w = c_tab[tb + 2 * pi + 1];
                float m = -3.0e38f;
#pragma unroll
                for (int r = 0; r < NR; ++r) { const float2 qv = qray(r, u, w); m = fmax3(m, qv.x, qv.y); }
                if (__any_sync(0xffffffffu, m >= thr_t)) { if (DIAG) slow_cnt++; slow_pair(pi); }
            }
        } else {
        if (DIAG && q.slow) tests_cnt += (unsigned long long)npair * 2ull * (unsigned long long)NR;
        // ---- scan: KP sphere pairs x NR rays per step, pair-major; the uniform registers of a pair are refilled
        // for the NEXT step as soon as the pair is done (rolling prefetch: LDCU latency never reaches the FMA pipe).
        // The candidate decision of a step is taken one step LATER (software pipelining): the vote/branch never
        // waits on the FMA pipe. ----
        float mprev[KPM];
        float4 u[KPM], w[KPM];
#pragma unroll
        for (int k = 0; k < KPM; ++k) { mprev[k] = -3.0e38f; u[k] = c_tab[tb + 2 * k]; w[k] = c_tab[tb + 2 * k + 1]; }
        int pb = 0;                                                   // first pair of the step being computed
        for (;;) {
            bool found = false;
            for (; pb < npair; pb += KPM) {
                float mall = mprev[0];
#pragma unroll
                for (int k = 1; k < KPM; ++k) mall = fmaxf(mall, mprev[k]);
                const bool fire = __any_sync(0xffffffffu, mall >= thr_t);       // step pb-KPM
                float mk4[KPM];
#pragma unroll
                for (int k = 0; k < KPM; ++k) {
                    float m = -3.0e38f;
#pragma unroll
                    for (int r = 0; r < NR; ++r) { const float2 qv = qray(r, u[k], w[k]); m = fmax3(m, qv.x, qv.y); }
                    mk4[k] = m;
                    u[k] = c_tab[tb + 2 * (pb + KPM + k)]; w[k] = c_tab[tb + 2 * (pb + KPM + k) + 1];   // next step (c_tab is over-allocated)
                }
                if (fire) {                                           // handle step pb-KP first, keep this step pending
                    slow_step(pb - KPM, mprev);
#pragma unroll
                    for (int k = 0; k < KPM; ++k) mprev[k] = mk4[k];
                    found = true; pb += KPM; break;
                }
#pragma unroll
                for (int k = 0; k < KPM; ++k) mprev[k] = mk4[k];
            }
            if (found) continue;
            float mall = mprev[0];
#pragma unroll
            for (int k = 1; k < KPM; ++k) mall = fmaxf(mall, mprev[k]);
            if (__any_sync(0xffffffffu, mall >= thr_t)) slow_step(npair - KPM, mprev);           // flush the last step
            break;
        }
        }
        flush();

        // ---- planes after the last sphere chunk (trace.py:32-39), then write back (one code copy, ray k*32 + lane) ----
        uint32_t surv = 0;                                     // ANY: bit k = ray k*32 + lane is still unoccluded
#pragma unroll 1
        for (int k = 0; k < NR; ++k) {
            const uint32_t rl = (uint32_t)k * 32u + lane, i = gbase + rl;
            if (i >= nlim) continue;
            const float4 dd = ws.d[rl];
            const V3 D = mk(dd.x, dd.y, dd.z);
            if (ANY) {
                bool occluded = skt[rl] != 0u;
                const float4 oo = ws.o[rl];
                if (last && !occluded && sc.Pl > 0) {
                    const V3 O = mk(oo.x, oo.y, oo.z);
                    for (int pl = 0; pl < sc.Pl && !occluded; ++pl) {
                        const float t = plane_exact(O, D, __ldg(&sc.plO[pl]), __ldg(&sc.plN[pl]));
                        occluded = (t > 0.f && t < kTMax);
                    }
                }
                if (occluded) q.occl[(size_t)j * q.ocap + __float_as_uint(oo.w)] = 1; else surv |= 1u << k;
            } else {
                float best = __uint_as_float(skt[rl]); int bid = (int)ski[rl];
                if (last) {
                    V3 O = Vp;
                    if (MODE == 0) { const float4 oo = ws.o[rl]; O = mk(0.5f * oo.x, 0.5f * oo.y, 0.5f * oo.z); }
                    for (int pl = 0; pl < sc.Pl; ++pl) {
                        const float t = plane_exact(O, D, __ldg(&sc.plO[pl]), __ldg(&sc.plN[pl]));
                        if (t > 0.f && t < best) { best = t; bid = sc.S + pl; }
                    }
                }
                const unsigned long long key = ((unsigned long long)__float_as_uint(best) << 32) | (uint32_t)bid;     // nothing hit: (999.0f, -1) = kNoHit
                if (q.only) q.hit[i] = key;
                else if (bid >= 0) atomicMin(&q.hit[i], key);
            }
        }
        if (ANY && q.out_list) {                              // survivors -> next pass's list: appended to their region, in group order
            const uint32_t total = __reduce_add_sync(0xffffffffu, (uint32_t)__popc(surv));
            uint32_t base = 0;
            if (lane == 0 && total) base = (gbase / kRegion) * kRegion + atomicAdd(&q.out_rn[gbase / kRegion], total);
            base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll 1
            for (int k = 0; k < NR; ++k) {
                const bool sv = (surv >> k) & 1u;
                const uint32_t bal = __ballot_sync(0xffffffffu, sv);
                RT_CHECK(4, !sv || base + (uint32_t)__popc(bal & ((1u << lane) - 1u)) < (gbase / kRegion + 1u) * kRegion);
                if (sv) q.out_list[base + (uint32_t)__popc(bal & ((1u << lane) - 1u))] = __float_as_uint(ws.o[(uint32_t)k * 32u + lane].w);
                base += (uint32_t)__popc(bal);
            }
        }
        __syncwarp();                                         // keys are re-initialised by the next group's setup
    };

    for (;;) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(work, 1u);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= ngroups) break;
        run_group(g);
    }
    if (DIAG && q.tl && threadIdx.x == 0) atomicMax(q.tl + 1, globaltimer_ns());
    if (DIAG && q.slow) {
        const uint32_t tot = __reduce_add_sync(0xffffffffu, slow_cnt), tot2 = __reduce_add_sync(0xffffffffu, exact_cnt);
        if (lane == 0 && tot) { atomicAdd(q.slow, (unsigned long long)tot); atomicAdd(q.slow + 1, (unsigned long long)tot2); }
        unsigned long long tt = tests_cnt;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) tt += __shfl_xor_sync(0xffffffffu, tt, d);
        if (lane == 0 && tt) atomicAdd(q.slow + (PIV ? 2 : 3), tt);
    }
}

// ------------------------------------------------------------------------------------------------
// Host side: table residency and launches (per translation unit = per constant bank)
// ------------------------------------------------------------------------------------------------
struct IsectCfg { int sm_count; int force_nr; int force_nr0; int npre; int pre[6]; };   // pre: sizes of the leading shadow passes

template <int MODE> static cudaError_t isect3_set_attr() {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_isect3<MODE, 8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IsectSmem<8>))) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_isect3<MODE, 4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IsectSmem<4>))) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_isect3<MODE, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IsectSmem<2>))) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_isect3<MODE, 8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IsectSmem<8>))) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_isect3<MODE, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IsectSmem<4>))) != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_isect3<MODE, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IsectSmem<2>));
}

// Whole tables tabs[0..n) (each Spad entries) into slots 0..n-1 of THIS module's c_tab: done once per frame when
// n * Spad <= kTabCap, after which launches pass resident = true and nothing is copied between them.
static inline cudaError_t isect3_upload_resident(const PivotDev& pv, int Spad, const int* tabs, int n, cudaStream_t st) {
    for (int k = 0; k < n; ++k) {
        cudaError_t e = cudaMemcpyToSymbolAsync(c_tab, pv.tab + (size_t)tabs[k] * Spad, sizeof(float4) * Spad,
                                                sizeof(float4) * (size_t)k * Spad, cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// Rays per thread: 8 for big queues, fewer for small ones (more groups than warps).
template <int MODE>
static cudaError_t isect3_launch_one(const IsectCfg& cfg, const SceneDev& sc, const QueueDev& q, const PivotDev& pv, size_t max_rays,
                                     int c0, int cnt, int j, int tb, cudaStream_t st) {
    const size_t warps = (size_t)cfg.sm_count * 2 * (kThreads / 32);
    // 4 rays per thread (groups of 128 rays, 3 CTAs = 24 warps per SM) measured best at every size once the 4-ray build
    // could cull 512 pairs (C4: 13.06 -> 12.78 ms; 1/8 of it 2.29 -> 2.17: shorter groups, shorter tails); 8 rays per
    // thread (B200RT_NR=8) amortise the uniform loads better but run 16 warps per SM.
    // Without culling (brute-force mode, B200RT_F_NO_CULL) the launch is one long dense scan and 8 rays per thread win
    // (C4 dense mode: 60.8 vs 69.5 ms per frame).
    int nr = cfg.force_nr ? cfg.force_nr : (max_rays >= warps * 128 * 4 ? ((pv.rad || max_rays < warps * 256 * 4) ? 4 : 8) : 2);
    if (MODE == 0 && cfg.force_nr0) nr = cfg.force_nr0;         // (measured: 8 rays/thread wins for the 8-FFMA2 form too)
    const int grid = cfg.sm_count * (nr == 8 ? 2 : 3);
    const bool diag = q.slow != nullptr || q.tl != nullptr;
    if (diag) {
        if (nr == 8)      k_isect3<MODE, 8, true><<<grid, kThreads, sizeof(IsectSmem<8>), st>>>(sc, q, pv, c0, cnt, j, tb);
        else if (nr == 4) k_isect3<MODE, 4, true><<<grid, kThreads, sizeof(IsectSmem<4>), st>>>(sc, q, pv, c0, cnt, j, tb);
        else              k_isect3<MODE, 2, true><<<grid, kThreads, sizeof(IsectSmem<2>), st>>>(sc, q, pv, c0, cnt, j, tb);
    } else {
        if (nr == 8)      k_isect3<MODE, 8, false><<<grid, kThreads, sizeof(IsectSmem<8>), st>>>(sc, q, pv, c0, cnt, j, tb);
        else if (nr == 4) k_isect3<MODE, 4, false><<<grid, kThreads, sizeof(IsectSmem<4>), st>>>(sc, q, pv, c0, cnt, j, tb);
        else              k_isect3<MODE, 2, false><<<grid, kThreads, sizeof(IsectSmem<2>), st>>>(sc, q, pv, c0, cnt, j, tb);
    }
    g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    return cudaGetLastError();
}

// Closest hit of a ray queue (MODE 0: origin-free table L+1, MODE 1: camera-pivot table L): one launch per sphere
// chunk of the constant table; chunks meet in q.hit (atomicMin).  resident (isect3_upload_resident with tabs = {L, L+1}):
// slot 0 = camera pivot, slot 1 = origin-free; otherwise the chunk's table is copied in front of its launch.
// `work`: one zeroed group counter per chunk.
template <int MODE>
static cudaError_t isect3_launch_closest(const IsectCfg& cfg, const SceneDev& sc, QueueDev q, const PivotDev& pv, size_t max_rays,
                                         uint32_t* work, bool resident, cudaStream_t st, unsigned long long* tlbase = nullptr, int* tlnext = nullptr) {
    static_assert(MODE == 0 || MODE == 1, "closest-hit modes");
    const int L = sc.L, chunk = std::min(sc.Spad, kTabCap), tj = (MODE == 1) ? L : L + 1;
    cudaError_t e;
    for (int c0 = 0, k = 0; c0 < sc.Spad; c0 += chunk, ++k) {
        const int cnt = std::min(chunk, sc.Spad - c0);
        if (!resident && (e = cudaMemcpyToSymbolAsync(c_tab, pv.tab + (size_t)tj * sc.Spad + c0, sizeof(float4) * cnt, 0,
                                                      cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
        q.work = work + k; q.last = (c0 + cnt >= sc.Spad) ? 1 : 0; q.only = (c0 == 0 && cnt >= sc.Spad) ? 1 : 0;
        if (tlbase) q.tl = tlbase + 2 * (size_t)((*tlnext)++);
        if ((e = isect3_launch_one<MODE>(cfg, sc, q, pv, max_rays, c0, cnt, L, resident ? (MODE == 1 ? 0 : sc.Spad) : 0, st)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ---- shadow queries: passes over the light's sorted table with compaction of the survivors ----
constexpr int kMaxPass = 32;

struct ShadowPlan {                                  // one depth's shadow state (device pointers)
    uint32_t* list0; uint32_t* list1; size_t cap;    // [L][cap] k_setup's lists (null: dense layout, every hit record) and the second buffer
    const uint32_t* n0; int n0_stride;               // list lengths from k_setup: n0[j * n0_stride]  (dense: the hit count, stride 0)
    uint32_t* work;                                  // [L][kMaxPass] zeroed group counters
    uint32_t* rn; size_t nreg;                       // [L][kMaxPass][nreg] zeroed region fill counts of the survivor lists (nreg = ceil(cap / kRegion))
};

// Sphere ranges [c0, c0+cnt) of the passes.  Small scenes: one pass.  (resident or not, a pass fits the constant table.)
static inline int shadow_passes(const IsectCfg& cfg, int Spad, int (*ranges)[2]) {
    int n = 0, c0 = 0;
    for (int k = 0; k < cfg.npre && Spad - c0 >= 2 * cfg.pre[k]; ++k) { ranges[n][0] = c0; ranges[n][1] = cfg.pre[k]; c0 += cfg.pre[k]; ++n; }
    while (c0 < Spad) {
        if (n == kMaxPass) return -1;
        const int cnt = std::min(kTabCap, Spad - c0);
        ranges[n][0] = c0; ranges[n][1] = cnt; c0 += cnt; ++n;
    }
    return n;
}

// Light j's shadow query (MODE 2).  resident: slot j of this module's c_tab holds the whole sorted table of light j.
template <int MODE>
static cudaError_t isect3_launch_shadow_light(const IsectCfg& cfg, const SceneDev& sc, QueueDev q, const PivotDev& pv, size_t max_rays,
                                              const ShadowPlan& sp, int j, bool resident, cudaStream_t st,
                                              unsigned long long* tlbase = nullptr, int* tlnext = nullptr) {
    static_assert(MODE == 2, "shadow mode");
    int ranges[kMaxPass][2];
    const int np = shadow_passes(cfg, sc.Spad, ranges);
    if (np < 0 || (np > 1 && !sp.list0)) return cudaErrorInvalidValue;
    const uint32_t* in = sp.list0 ? sp.list0 + (size_t)j * sp.cap : nullptr;
    const uint32_t* n_in = sp.n0 + (size_t)j * sp.n0_stride;
    cudaError_t e;
    for (int p = 0; p < np; ++p) {
        const int c0 = ranges[p][0], cnt = ranges[p][1];
        const bool lastp = (p == np - 1);
        q.in_list = in; q.n_in = n_in; q.work = sp.work + (size_t)j * kMaxPass + p; q.last = lastp ? 1 : 0;
        q.out_list = lastp ? nullptr : ((p & 1) ? sp.list0 : sp.list1) + (size_t)j * sp.cap;
        q.in_rn = p == 0 ? nullptr : sp.rn + ((size_t)j * kMaxPass + (p - 1)) * sp.nreg;
        q.out_rn = lastp ? nullptr : sp.rn + ((size_t)j * kMaxPass + p) * sp.nreg;
        if (!resident && (e = cudaMemcpyToSymbolAsync(c_tab, pv.tab + (size_t)j * sc.Spad + c0, sizeof(float4) * cnt, 0,
                                                      cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return e;
        if (tlbase) q.tl = tlbase + 2 * (size_t)((*tlnext)++);
        if ((e = isect3_launch_one<MODE>(cfg, sc, q, pv, max_rays, c0, cnt, j, resident ? j * sc.Spad + c0 : 0, st)) != cudaSuccess) return e;
        in = q.out_list;
    }
    return cudaSuccess;
}

// The shadow mode lives in isect_shadow.cu (its own module / constant bank):
cudaError_t shadow_set_attr();
unsigned long long shadow_check_failures();
cudaError_t shadow_upload_resident(const PivotDev& pv, int Spad, int L, cudaStream_t st);
cudaError_t shadow_launch_light(const IsectCfg& cfg, const SceneDev& sc, const QueueDev& q, const PivotDev& pv, size_t max_rays,
                                const ShadowPlan& sp, int j, bool resident, cudaStream_t st,
                                unsigned long long* tlbase = nullptr, int* tlnext = nullptr);

}  // namespace b200rt
