// rt_math.cuh — FP32 device math of the render path, written with explicit rounding intrinsics so that
// every kernel that inlines these helpers performs bit-identical arithmetic (no compiler-chosen FMA
// contraction; IEEE div/sqrt — the reference kernel is compiled without fast-math).
//
// Reference semantics restated here (all under /root/reference/src/ray_tracing/):
//   common.py:10-49   linear_comb / vector_difference / normalize / dot / matmul
//   common.py:52-63   clip_color / clip_color_vector (round-half-even, R,B,G wiring)
//   intersections.py:6-68  ray-sphere / ray-plane
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200rt {

constexpr float kBias  = 0.0002f;   // trace.py:82
constexpr float kEps   = 0.001f;    // intersections.py:46
constexpr float kTMax  = 999.0f;    // trace.py:17

// Checked build (-DB200RT_CHECK; tools/gpu_check.sh): index assertions at the places where a wrong index would write outside
// a buffer — compute-sanitizer is closed on this GPU pool (profiles/r02_sanitizer.log), so the checks are the library's own.
// Every translation unit counts into its own copy; b200rt_debug_check_failures() sums them.  Compiled out otherwise.
#ifdef B200RT_CHECK
static __device__ unsigned int g_check_fail[8];
#define RT_CHECK(slot, cond) do { if (!(cond)) atomicAdd(&g_check_fail[slot], 1u); } while (0)
#else
#define RT_CHECK(slot, cond) do { } while (0)
#endif
// slots: 0 candidate queue of k_isect3, 1 pair list of the beam culling, 2 hit list / next queue (k_setup), 3 per-light lists,
//        4 survivor lists, 5 regrouping (bins / permutation), 6 closest-hit grid (cells / ids), 7 light grids (cells)

struct V3 { float x, y, z; };

__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ float dot3(V3 a, V3 b) {
    return fmaf(a.x, b.x, fmaf(a.y, b.y, __fmul_rn(a.z, b.z)));
}
__device__ __forceinline__ V3 sub3(V3 a, V3 b) { return mk(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)); }
// a + s*b
__device__ __forceinline__ V3 madd3(V3 a, float s, V3 b) { return mk(fmaf(s, b.x, a.x), fmaf(s, b.y, a.y), fmaf(s, b.z, a.z)); }
// common.py:28-32 — divide by the norm (IEEE sqrt + div), no zero guard
__device__ __forceinline__ V3 normalize3(V3 v) {
    float n = __fsqrt_rn(dot3(v, v));
    return mk(__fdiv_rn(v.x, n), __fdiv_rn(v.y, n), __fdiv_rn(v.z, n));
}

// common.py:52-57: min(max(0, int(round(c))), 255) with round-half-to-even (cvt.rni), SURVEY F4
__device__ __forceinline__ uint32_t clip_color(float c) {
    int v = __float2int_rn(c);          // NaN -> 0, saturates at INT_MIN/INT_MAX
    v = max(0, min(255, v));
    return (uint32_t)v;
}

// Exact (robust) FP32 ray-sphere test, unit direction.  Restates intersections.py:6-38 with a == 1:
//   reference  t = (-b -/+ sqrt(b^2 - 4c)) / 2,  b = 2 (O-C).D,  c = |O-C|^2 - r^2
// evaluated as tc = (C-O).D, disc = r^2 - |(C-O) - tc D|^2 (perpendicular-distance form: no catastrophic
// cancellation between b^2 and c when the origin is far from a small sphere), t = tc -/+ sqrt(disc).
// Returns t > 0 (near root if positive, else far root if positive — origin inside the sphere), or -1.
__device__ __forceinline__ float sphere_exact(V3 O, V3 D, float cx, float cy, float cz, float r2) {
    V3 Lc = mk(__fsub_rn(cx, O.x), __fsub_rn(cy, O.y), __fsub_rn(cz, O.z));
    float tc = dot3(Lc, D);
    V3 v = mk(fmaf(-tc, D.x, Lc.x), fmaf(-tc, D.y, Lc.y), fmaf(-tc, D.z, Lc.z));
    float disc = __fsub_rn(r2, dot3(v, v));
    if (!(disc >= 0.0f)) return -1.0f;                 // intersections.py:25-26
    float sq = __fsqrt_rn(disc);
    float t = __fsub_rn(tc, sq);                       // :28-31
    if (t > 0.0f) return t;
    t = __fadd_rn(tc, sq);                             // :33-36
    return (t > 0.0f) ? t : -1.0f;
}

// intersections.py:41-68.  (ox,oy,oz) plane point, (nx,ny,nz) stored normal (not re-normalised here).
__device__ __forceinline__ float plane_exact(V3 O, V3 D, float4 po, float4 pn) {
    V3 n = mk(pn.x, pn.y, pn.z);
    float denom = dot3(D, n);
    if (fabsf(denom) < kEps) return -1.0f;
    V3 LP = mk(__fsub_rn(po.x, O.x), __fsub_rn(po.y, O.y), __fsub_rn(po.z, O.z));
    float t = __fdiv_rn(dot3(LP, n), denom);
    return (t > 0.0f) ? t : -1.0f;
}

// Ray queues store no origins (round 2: 16 bytes less written per ray and per hit, 16-32 less read).
//   depth 0: every primary ray starts at the camera; the queue entry is (D.xyz, slot bits) and the origin is the camera constant `camo`;
//   depth >= 1: the queue IS the hit list of the depth before: entry i = hit record i = (biased hit point Pb, slot) plus the reflected
//   direction R, and the bounce origin is Pb + BIAS * R (trace.py:110) — evaluated by this one function wherever a kernel needs it, so
//   every consumer sees the same bits that k_setup used to store.
__device__ __forceinline__ V3 ray_origin(const float4* __restrict__ rayO, const float* __restrict__ camo, uint32_t i, V3 D, float dw, uint32_t& slot) {
    if (rayO) { const float4 p = rayO[i]; slot = __float_as_uint(p.w); return madd3(mk(p.x, p.y, p.z), kBias, D); }
    slot = __float_as_uint(dw);
    return mk(camo[0], camo[1], camo[2]);
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// Ordered multi-queue append ("decoupled look-back" compaction).  Chunk c (256 consecutive input items, chunks claimed
// in ticket order) appends its items to up to 8 queues so that every queue keeps the INPUT order across the whole
// launch — not just within a CTA, as an atomic append per CTA would.  Neighbouring pixel tiles then stay neighbours in every
// compacted list, which is what keeps a ray group of k_isect3 a narrow beam at every depth.
// status: [chunk][8] 64-bit words, zeroed before the launch; bits 62-63: 1 = this chunk's count, 2 = inclusive prefix.
// Warp q of the CTA owns queue q: it publishes the chunk's count, sums its predecessors' (32 at a time, spinning only
// on chunks that are already running — tickets are handed out in order), then publishes the inclusive prefix.
// The chunk with is_last stores the queue lengths to counters[0..nq).  Contains __syncthreads(); 256-thread CTAs.
__device__ __forceinline__ void ordered_append8(uint32_t flags, int nq, uint32_t c, unsigned long long* status,
                                                uint32_t* counters, bool is_last, uint32_t (&pos)[8]) {
    __shared__ uint32_t s_cnt8[8][8];      // [queue][warp]
    __shared__ uint32_t s_base8[8];
    const uint32_t wid = threadIdx.x >> 5, lane = lane_id(), lt = (1u << lane) - 1u;
    uint32_t mask[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        mask[j] = (j < nq) ? __ballot_sync(0xffffffffu, (flags >> j) & 1u) : 0u;
        if (j < nq && lane == 0) s_cnt8[j][wid] = (uint32_t)__popc(mask[j]);
    }
    __syncthreads();
    if ((int)wid < nq) {
        const int qi = (int)wid;
        const uint32_t v = lane < 8 ? s_cnt8[qi][lane] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int d = 1; d < 8; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if ((int)lane >= d) x += y; }
        const uint32_t total = __shfl_sync(0xffffffffu, x, 7);
        if (lane < 8) s_cnt8[qi][lane] = x - v;                       // exclusive offset of each warp within the chunk
        unsigned long long* st = status + (size_t)c * 8 + qi;
        uint32_t excl = 0;
        if (c == 0) { if (lane == 0) atomicExch(st, (2ull << 62) | total); }
        else {
            if (lane == 0) atomicExch(st, (1ull << 62) | total);
            for (long long p = (long long)c - 1;; p -= 32) {
                const long long idx = p - (long long)lane;
                unsigned long long w = 2ull << 62;                    // before chunk 0: inclusive prefix 0
                if (idx >= 0) {
                    const volatile unsigned long long* sp = status + (size_t)idx * 8 + qi;
                    do { w = *sp; } while ((w >> 62) == 0ull);
                }
                const uint32_t incl = __ballot_sync(0xffffffffu, (w >> 62) == 2ull);
                const int first = incl ? __ffs(incl) - 1 : 32;
                excl += __reduce_add_sync(0xffffffffu, (int)lane <= first ? (uint32_t)w : 0u);
                if (incl) break;
            }
            if (lane == 0) atomicExch(st, (2ull << 62) | (unsigned long long)(excl + total));
        }
        if (lane == 0) { s_base8[qi] = excl; if (is_last) counters[qi] = excl + total; }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) pos[j] = (j < nq) ? s_base8[j] + s_cnt8[j][wid] + (uint32_t)__popc(mask[j] & lt) : 0u;
    __syncthreads();
}

}  // namespace b200rt
