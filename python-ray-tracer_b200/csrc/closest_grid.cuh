// closest_grid.cuh — closest-hit queries of incoherent (bounce) rays through a uniform 3-D grid of the spheres, one ray
// per thread (reference trace.py:7-41, called at :53 for the reflection traces of :128).
//
// Why a second closest-hit kernel.  k_isect3 (isect3.cuh) is a beam method: 128 rays share one culled sphere list, which is
// what makes primary rays cost 0.6 % of the brute-force tests.  Its cost per ray rises as the rays of a group diverge — deeper
// bounces, and above all SMALLER queues: the regrouping bins hold a fixed number of rays, so at 1/8 of the rays (one GPU of
// eight) a bin spans 8x the volume of ray space and the groups keep 1.6x the sphere pairs, 2x as many take the dense scan
// (profiles/r02_cullstats.txt; second bounce: 68 ns per ray at the full C4 frame, 149 ns at 1/8 of it).  A per-ray traversal
// costs the same per ray at any queue size.
//
// Exactness (the image stays byte-identical to brute force; tests compare with B200RT_F_DENSE / B200RT_F_NO_FILTER):
//   * every candidate is decided by sphere_exact() on the real ray — the function every other path uses — and the closest hit
//     is the minimum over (t, sphere id), i.e. "smallest t, lowest index on a tie" (trace.py:26); planes afterwards with a
//     strict `<`, so they lose ties against spheres (trace.py:36);
//   * the grid only has to offer every sphere the ray can hit, in a cell the walk visits before it stops.  A sphere is listed in
//     every cell its box, inflated by eps = 2e-3 + 1e-3 r, overlaps: eps covers the exact test's slack (it may report a hit up to
//     ~1e-4 outside the sphere while coordinates stay below 256) and the rounding of the walk (a cell grazed within ~1e-5 of a
//     corner may be skipped: whatever is hit there is also listed in the neighbouring cell that is visited);
//   * the walk stops after a cell only when the best hit so far lies clearly before the cell's exit (margin 1e-3 + 1e-5 t): every
//     sphere not yet offered can only be hit beyond that exit;
//   * spheres whose inflated box covers more than kCgBigCells cells are kept on a short list that every ray tests;
//     beyond coordinates of 256 the inflation and the stop margin grow in proportion (cg_eps_scale); a scene whose lists overflow the
//     grid's capacity (ids, big list, cells) marks the grid unusable on the device and every ray then tests every sphere in this kernel.
#pragma once
#include "isect3.cuh"

namespace b200rt {

// ---- regrouping key of a hit (see "Regrouping of the hits of depth >= 1" in b200rt.cu): bin = (cube-map cell of the REFLECTED direction, grid
// cell of the hit point).  Shared by k_key and by k_closest_grid, which computes it in its epilogue where (O, D, t, id) are still in registers
// (one launch and 40 bytes of reads per ray less).  Approximate arithmetic: this only chooses a bin.
struct GridDev { float lo[3], inv[3]; int n[3]; int dn; };      // position cells n[0] x n[1] x n[2], direction cells 6 x dn x dn
__device__ __forceinline__ uint32_t sort_bin(const SceneDev& sc, const GridDev& g, V3 o, V3 d, float t, int id, int use_dir) {
    const float px = fmaf(t, d.x, o.x), py = fmaf(t, d.y, o.y), pz = fmaf(t, d.z, o.z);
    uint32_t dc = 0;
    if (use_dir) {
        float nx, ny, nz;
        if (id < sc.S) { const float4 c = __ldg(&sc.sphB[id]); nx = px - c.x; ny = py - c.y; nz = pz - c.z; const float il = rsqrtf(fmaxf(nx * nx + ny * ny + nz * nz, 1e-30f)); nx *= il; ny *= il; nz *= il; }
        else           { const float4 nu = __ldg(&sc.plNu[id - sc.S]); nx = nu.x; ny = nu.y; nz = nu.z; }
        const float dn2 = 2.f * (d.x * nx + d.y * ny + d.z * nz);
        const float rx = d.x - dn2 * nx, ry = d.y - dn2 * ny, rz = d.z - dn2 * nz;
        const float ax = fabsf(rx), ay = fabsf(ry), az = fabsf(rz);
        float m, u, v; uint32_t face;
        if (ax >= ay && ax >= az) { m = ax; u = ry; v = rz; face = rx < 0.f ? 1u : 0u; }
        else if (ay >= az)        { m = ay; u = rx; v = rz; face = ry < 0.f ? 3u : 2u; }
        else                      { m = az; u = rx; v = ry; face = rz < 0.f ? 5u : 4u; }
        const float im = 0.5f * (float)g.dn / fmaxf(m, 1e-30f);
        const int iu = min(g.dn - 1, max(0, (int)(u * im + 0.5f * (float)g.dn)));
        const int iv = min(g.dn - 1, max(0, (int)(v * im + 0.5f * (float)g.dn)));
        dc = (face * (uint32_t)g.dn + (uint32_t)iu) * (uint32_t)g.dn + (uint32_t)iv;
    }
    const int ix = min(g.n[0] - 1, max(0, (int)((px - g.lo[0]) * g.inv[0])));
    const int iy = min(g.n[1] - 1, max(0, (int)((py - g.lo[1]) * g.inv[1])));
    const int iz = min(g.n[2] - 1, max(0, (int)((pz - g.lo[2]) * g.inv[2])));
    return ((dc * (uint32_t)g.n[0] + (uint32_t)ix) * (uint32_t)g.n[1] + (uint32_t)iy) * (uint32_t)g.n[2] + (uint32_t)iz;
}
// bin + rank within the bin (one atomic per distinct bin among the calling lanes) -> keyb[i]; b == 0xffffffff: a miss, not counted
__device__ __forceinline__ void sort_key_store(uint32_t b, uint32_t i, uint32_t* hist, uint2* keyb) {
    const uint32_t act = __activemask(), lane = lane_id();
    const uint32_t peers = __match_any_sync(act, b);
    const uint32_t leader = (uint32_t)__ffs(peers) - 1u;
    uint32_t base = 0;
    if (b != 0xffffffffu && lane == leader) base = atomicAdd(&hist[b], (uint32_t)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    keyb[i] = make_uint2(b, base + (uint32_t)__popc(peers & ((1u << lane) - 1u)));
}

constexpr int kCgMaxDim = 128;            // cells per axis, at most
constexpr int kCgBigCells = 512;          // a sphere covering more cells than this goes on the "big" list
constexpr int kCgMaxBig = 256;            // capacity of the big list; more big spheres than that: the grid is not used
constexpr float kCgRange = 256.f;

struct CGridHdr {                         // device-resident description of the grid (written by k_cg_setup)
    float lo[3], inv[3], edge[3];         // cell (x,y,z) = floor((p - lo) * inv); edge = 1 / inv
    int n[3]; int ncells;
    int nbig; int ok;                     // ok = 0: grid unusable for this scene (lists overflow its capacity): k_closest_grid tests every sphere
    uint32_t total;                       // entries of the id list
};

struct CGridDev {
    CGridHdr* hdr;
    uint32_t* start;                      // [ncells + 1] exclusive prefix of the cells' list lengths (counts before the scan)
    uint32_t* cursor;                     // [ncells] fill cursors
    uint16_t* ids;                        // [total] sphere ids, cell by cell
    uint4* rec;                           // [ncells] per cell 8 x uint16: [0] list length, [1..7] the first seven ids — one 16-byte load serves
                                          // a cell (average list: 2-3 spheres); longer lists continue in `ids`
    uint32_t* big;                        // [kCgMaxBig] ids of the big spheres
    size_t ids_cap;                       // capacity of `ids` (entries)
};

// eps = (2e-3 + 1e-3 r) * max(1, B / 256), B = max |C| + max r: the exact test's slack and the walk's rounding grow linearly with the coordinates
__device__ __forceinline__ float cg_eps_scale(const float* cmax) { return fmaxf(1.f, (cmax[0] + cmax[1]) * (1.f / kCgRange)); }

// Scene bounds -> grid dimensions.  One CTA.  n_target = cells_per_sphere * S cubic cells (1 - 2 per sphere measured within 1 % of each other, default 1.5:
// profiles/r02_cgrid*.txt), <= kCgMaxDim per axis.
__global__ void __launch_bounds__(256) k_cg_setup(SceneDev sc, CGridDev g, int max_cells, float cells_per_sphere, const uint32_t* cs, int cs_group) {
    if (cs[cs_group] == cs[0]) return;
    __shared__ int s_ncells;
    __shared__ float smin[3][256], smax[3][256];
    float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
    for (int i = threadIdx.x; i < sc.S; i += blockDim.x) {
        const float4 sp = sc.sphB[i]; const float r = sqrtf(fmaxf(sp.w, 0.f)) * (1.f + 1e-3f * cg_eps_scale(sc.cmax)) + 2e-3f * cg_eps_scale(sc.cmax); const float c[3] = {sp.x, sp.y, sp.z};
        for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], c[a] - r); hi[a] = fmaxf(hi[a], c[a] + r); }
    }
    for (int a = 0; a < 3; ++a) { smin[a][threadIdx.x] = lo[a]; smax[a][threadIdx.x] = hi[a]; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int a = 0; a < 3; ++a) for (int t = 1; t < 256; ++t) { lo[a] = fminf(lo[a], smin[a][t]); hi[a] = fmaxf(hi[a], smax[a][t]); }
        CGridHdr h{};
        float ext[3]; float vol = 1.f;
        for (int a = 0; a < 3; ++a) { if (!(hi[a] > lo[a])) { lo[a] = 0.f; hi[a] = 1.f; } const float pad = 1e-3f * (hi[a] - lo[a]) + 1e-3f; lo[a] -= pad; hi[a] += pad; ext[a] = hi[a] - lo[a]; vol *= ext[a]; }
        const float target = fminf(fmaxf(cells_per_sphere * (float)sc.S, 64.f), 262144.f);
        const float edge = cbrtf(vol / target);
        h.ncells = 1;
        for (int a = 0; a < 3; ++a) {
            h.n[a] = max(1, min(kCgMaxDim, (int)ceilf(ext[a] / fmaxf(edge, 1e-20f))));
            h.lo[a] = lo[a]; h.edge[a] = ext[a] / (float)h.n[a]; h.inv[a] = (float)h.n[a] / ext[a];
            h.ncells *= h.n[a];
        }
        h.ok = (sc.S > 0 && vol > 0.f && vol < 3e38f && h.ncells <= max_cells) ? 1 : 0;       // (NaN -> 0)
        h.nbig = 0; h.total = 0;
        *g.hdr = h;
        s_ncells = h.ok ? h.ncells : 0;
    }
    __syncthreads();
    for (int i = threadIdx.x; i <= s_ncells; i += blockDim.x) g.start[i] = 0u;       // the counts of pass 0 start from zero
}

__device__ __forceinline__ void cg_cell_range(const CGridHdr& h, float4 sp, float escale, int (&c0)[3], int (&c1)[3]) {
    const float r = sqrtf(fmaxf(sp.w, 0.f)) * (1.f + 1e-3f * escale) + 2e-3f * escale;      // r + eps, rounded generously
    const float c[3] = {sp.x, sp.y, sp.z};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        c0[a] = max(0, min(h.n[a] - 1, (int)floorf((c[a] - r - h.lo[a]) * h.inv[a] - 1e-3f)));
        c1[a] = max(0, min(h.n[a] - 1, (int)floorf((c[a] + r - h.lo[a]) * h.inv[a] + 1e-3f)));
    }
}

// pass 0: count the cells' list lengths (and collect the big spheres); pass 1: fill the lists.  One thread per sphere.
__global__ void __launch_bounds__(256) k_cg_bin(SceneDev sc, CGridDev g, int pass, const uint32_t* cs, int cs_group) {
    if (cs[cs_group] == cs[0]) return;
    const CGridHdr h = *g.hdr;
    if (!h.ok) return;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= sc.S) return;
    int c0[3], c1[3]; cg_cell_range(h, sc.sphB[s], cg_eps_scale(sc.cmax), c0, c1);
    const long long ncov = (long long)(c1[0] - c0[0] + 1) * (c1[1] - c0[1] + 1) * (c1[2] - c0[2] + 1);
    if (ncov > kCgBigCells) {
        if (pass == 0) { const int k = atomicAdd(&g.hdr->nbig, 1); if (k < kCgMaxBig) g.big[k] = (uint32_t)s; }
        return;
    }
    for (int z = c0[2]; z <= c1[2]; ++z) for (int y = c0[1]; y <= c1[1]; ++y) for (int x = c0[0]; x <= c1[0]; ++x) {
        const int cell = (z * h.n[1] + y) * h.n[0] + x;
        if (pass == 0) atomicAdd(&g.start[cell], 1u);
        else { const uint32_t k = atomicAdd(&g.cursor[cell], 1u); RT_CHECK(6, k < g.start[cell + 1]); if ((size_t)k < g.ids_cap) g.ids[k] = (uint16_t)s; }
    }
}

// Exclusive scan of the counts (one CTA; <= 2^21 cells, a few per thread) and validation.  cursor[cell] = start[cell].
__global__ void __launch_bounds__(1024) k_cg_scan(CGridDev g, int max_cells, const uint32_t* cs, int cs_group) {
    if (cs[cs_group] == cs[0]) return;
    __shared__ uint32_t swarp[32]; __shared__ uint32_t s_carry;
    const CGridHdr h = *g.hdr;
    if (!h.ok) return;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int n = min(h.ncells, max_cells);
    for (int base = 0; base < n; base += 1024) {
        const int i = base + (int)threadIdx.x;
        const uint32_t v = i < n ? g.start[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if ((int)(threadIdx.x & 31) >= d) x += y; }
        if ((threadIdx.x & 31) == 31) swarp[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = swarp[threadIdx.x], y2 = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, y2, d); if ((int)threadIdx.x >= d) y2 += y; }
            swarp[threadIdx.x] = y2 - w;                         // exclusive offset of each warp
        }
        __syncthreads();
        const uint32_t excl = s_carry + swarp[threadIdx.x >> 5] + x - v;
        if (i < n) { g.start[i] = excl; g.cursor[i] = excl; }
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        g.start[n] = s_carry;
        g.hdr->total = s_carry;
        if ((size_t)s_carry > g.ids_cap || h.nbig > kCgMaxBig || h.ncells > max_cells) g.hdr->ok = 0;      // does not fit: k_closest_grid tests every sphere
    }
}

// One thread per cell: the cell's 16-byte record from its list.
__global__ void __launch_bounds__(256) k_cg_pack(CGridDev g, const uint32_t* cs, int cs_group) {
    if (cs[cs_group] == cs[0]) return;
    const CGridHdr h = *g.hdr;
    if (!h.ok) return;
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= h.ncells) return;
    const uint32_t s0 = g.start[cell], cnt = g.start[cell + 1] - s0;
    uint32_t v[8]; v[0] = min(cnt, 0xffffu);
#pragma unroll
    for (int k = 0; k < 7; ++k) v[k + 1] = (uint32_t)k < cnt ? (uint32_t)g.ids[s0 + k] : 0u;
    g.rec[cell] = make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
}

struct CGridArgs {
    const float4* rayO; const float4* rayD; const float* camo; const uint32_t* n_in;      // ray queue (rt_math.cuh: ray_origin)
    unsigned long long* hit;                 // (float bits of t) << 32 | id per ray; written for every ray
    unsigned long long* tl;                  // diagnostic timeline slot; may be null
    unsigned long long* slow;                // diagnostics: [1] exact tests, [10] cell steps; may be null
    uint32_t* unused;
    const GridDev* sgrid; uint32_t* hist; uint2* keyb; int use_dir;      // non-null keyb: also compute the hits' regrouping keys (k_key folded in)
};

// One thread per ray: Amanatides-Woo walk through the cells, sphere_exact on every listed sphere, planes last.
// (Measured and dropped, profiles/r02_cgrid*.txt: a mailbox of the last two spheres offered, 0.70 -> 0.75 ms for 1/8 of C4; persistent
// lanes with dynamic ray fetch — one cell step per loop iteration, idle lanes refill from the queue — 0.73 -> 0.87 ms.)
template <bool KEY>      // KEY: also compute the regrouping keys of the hits (a separate kernel: the key code must not cost the plain walk registers)
__device__ __forceinline__ void closest_grid_body(const SceneDev& sc, const CGridDev& g, const CGridArgs& a) {
    if (a.tl && threadIdx.x == 0) atomicMin(a.tl, globaltimer_ns());
    const CGridHdr h = *g.hdr;
    const uint32_t n = *a.n_in;
    if (!h.ok) {            // grid unusable for this scene (capacity: see k_cg_scan): every ray tests every sphere, in order — slow, correct, rare
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
            const float4 d4 = a.rayD[i];
            const V3 D = mk(d4.x, d4.y, d4.z);
            uint32_t slot_; const V3 O = ray_origin(a.rayO, a.camo, i, D, d4.w, slot_);
            float best = kTMax; int id = -1;
            for (int sid = 0; sid < sc.S; ++sid) {
                const float4 sp = __ldg(&sc.sphB[sid]);
                const float t = sphere_exact(O, D, sp.x, sp.y, sp.z, sp.w);
                if (t > 0.f && t < best) { best = t; id = sid; }                                           // trace.py:26-29
            }
            for (int pl = 0; pl < sc.Pl; ++pl) {
                const float t = plane_exact(O, D, __ldg(&sc.plO[pl]), __ldg(&sc.plN[pl]));
                if (t > 0.f && t < best) { best = t; id = sc.S + pl; }
            }
            a.hit[i] = ((unsigned long long)__float_as_uint(best) << 32) | (uint32_t)id;
            if (KEY) sort_key_store(id >= 0 ? sort_bin(sc, *a.sgrid, O, D, best, id, a.use_dir) : 0xffffffffu, i, a.hist, a.keyb);
        }
        return;
    }
    const float mscale = cg_eps_scale(sc.cmax);                 // the walk's stop margin grows with the coordinate range like the lists' inflation
    const float marg0 = 1e-3f * mscale, marg1 = 1e-5f * mscale;
    unsigned long long n_exact = 0, n_steps = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 d4 = a.rayD[i];
        const V3 D = mk(d4.x, d4.y, d4.z);
        uint32_t slot_; const V3 O = ray_origin(a.rayO, a.camo, i, D, d4.w, slot_);
        float best = kTMax; uint32_t bid = 0xffffffffu;
        auto offer = [&](uint32_t sid) {
            const float4 sp = __ldg(&sc.sphB[sid]);
            const float t = sphere_exact(O, D, sp.x, sp.y, sp.z, sp.w);
            if (a.slow) ++n_exact;
            if (t > 0.f && (t < best || (t == best && sid < bid))) { best = t; bid = sid; }             // trace.py:26-29: smallest t, lowest index on a tie
        };
        for (int k = 0; k < h.nbig; ++k) offer(g.big[k]);
        // clip the ray to the grid's box (slab test, conservative: box grown by a cell fraction).  One reciprocal per axis serves the
        // slab test and the walk's increments: these parameters only steer the walk (margins below), the hits come from sphere_exact.
        const float o[3] = {O.x, O.y, O.z}, d[3] = {D.x, D.y, D.z};
        float inv[3];
        float t0 = 0.f, t1 = kTMax;
        bool miss = false;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const float lo = h.lo[ax] - 1e-3f * h.edge[ax], hi = h.lo[ax] + (float)h.n[ax] * h.edge[ax] + 1e-3f * h.edge[ax];
            inv[ax] = 0.f;
            if (fabsf(d[ax]) < 1e-20f) { if (o[ax] < lo || o[ax] > hi) miss = true; }
            else {
                inv[ax] = __frcp_rn(d[ax]);
                float ta = (lo - o[ax]) * inv[ax], tb = (hi - o[ax]) * inv[ax];
                if (ta > tb) { const float tt = ta; ta = tb; tb = tt; }
                t0 = fmaxf(t0, ta - 1e-4f * fabsf(ta)); t1 = fminf(t1, tb + 1e-4f * fabsf(tb));
            }
        }
        if (!miss && t0 <= t1) {
            int c[3], step[3]; float tmax[3], tdel[3];
#pragma unroll
            for (int ax = 0; ax < 3; ++ax) {
                const float p = fmaf(t0, d[ax], o[ax]);
                c[ax] = max(0, min(h.n[ax] - 1, (int)floorf((p - h.lo[ax]) * h.inv[ax])));
                if (d[ax] > 1e-20f)       { step[ax] = 1;  tdel[ax] = h.edge[ax] * inv[ax];  tmax[ax] = (h.lo[ax] + (float)(c[ax] + 1) * h.edge[ax] - o[ax]) * inv[ax]; }
                else if (d[ax] < -1e-20f) { step[ax] = -1; tdel[ax] = -h.edge[ax] * inv[ax]; tmax[ax] = (h.lo[ax] + (float)c[ax] * h.edge[ax] - o[ax]) * inv[ax]; }
                else                      { step[ax] = 0;  tdel[ax] = 3e38f; tmax[ax] = 3e38f; }
            }
            int cell = (c[2] * h.n[1] + c[1]) * h.n[0] + c[0];
            const int cstride[3] = { step[0], step[1] * h.n[0], step[2] * h.n[0] * h.n[1] };
            uint4 rec = __ldg(&g.rec[cell]);
            for (;;) {
                // where the walk goes next is known before this cell's spheres are looked at: the next cell's record is requested
                // first, so its latency overlaps the tests (record -> sphB[id] is the only dependent chain left)
                const float texit = fminf(tmax[0], fminf(tmax[1], tmax[2]));
                const int ax = (tmax[0] <= tmax[1] && tmax[0] <= tmax[2]) ? 0 : (tmax[1] <= tmax[2] ? 1 : 2);
                const int cn = (ax == 0 ? c[0] + step[0] : ax == 1 ? c[1] + step[1] : c[2] + step[2]);
                const bool inside = cn >= 0 && cn < (ax == 0 ? h.n[0] : ax == 1 ? h.n[1] : h.n[2]) && texit <= t1;
                const int ncell = cell + (ax == 0 ? cstride[0] : ax == 1 ? cstride[1] : cstride[2]);
                uint4 nrec = make_uint4(0u, 0u, 0u, 0u);
                if (inside) nrec = __ldg(&g.rec[ncell]);
                const uint32_t cnt = rec.x & 0xffffu;
                RT_CHECK(6, cell >= 0 && cell < h.ncells && (!inside || (ncell >= 0 && ncell < h.ncells)));
                if (cnt) {
                    offer(rec.x >> 16);
                    if (cnt > 1) { offer(rec.y & 0xffffu);
                    if (cnt > 2) { offer(rec.y >> 16);
                    if (cnt > 3) { offer(rec.z & 0xffffu);
                    if (cnt > 4) { offer(rec.z >> 16);
                    if (cnt > 5) { offer(rec.w & 0xffffu);
                    if (cnt > 6) { offer(rec.w >> 16);
                    if (cnt > 7) { const uint32_t s0 = __ldg(&g.start[cell]); for (uint32_t k = 7; k < cnt; ++k) offer((uint32_t)__ldg(&g.ids[s0 + k])); } } } } } } }
                }
                if (a.slow) ++n_steps;
                if (best < texit - fmaf(marg1, texit, marg0) || !inside) break;   // nothing beyond this cell can be closer / the ray leaves the grid
                if (ax == 0) { c[0] = cn; tmax[0] += tdel[0]; } else if (ax == 1) { c[1] = cn; tmax[1] += tdel[1]; } else { c[2] = cn; tmax[2] += tdel[2]; }
                cell = ncell; rec = nrec;
            }
        }
        int id = (int)bid;
        for (int pl = 0; pl < sc.Pl; ++pl) {                                                             // trace.py:32-39
            const float t = plane_exact(O, D, __ldg(&sc.plO[pl]), __ldg(&sc.plN[pl]));
            if (t > 0.f && t < best) { best = t; id = sc.S + pl; }
        }
        a.hit[i] = ((unsigned long long)__float_as_uint(best) << 32) | (uint32_t)id;                    // nothing hit: (999.0f, -1) = kNoHit
        if (KEY) sort_key_store(id >= 0 ? sort_bin(sc, *a.sgrid, O, D, best, id, a.use_dir) : 0xffffffffu, i, a.hist, a.keyb);
    }
    if (a.slow) {
#pragma unroll
        for (int dd = 16; dd > 0; dd >>= 1) { n_exact += __shfl_xor_sync(0xffffffffu, n_exact, dd); n_steps += __shfl_xor_sync(0xffffffffu, n_steps, dd); }
        if (lane_id() == 0) { if (n_exact) { atomicAdd(a.slow + 1, n_exact); atomicAdd(a.slow + 11, n_exact); } if (n_steps) atomicAdd(a.slow + 10, n_steps); }
    }
    if (a.tl && threadIdx.x == 0) atomicMax(a.tl + 1, globaltimer_ns());
}

// Both at 64 registers = 4 CTAs per SM (the walk lives on occupancy: 76 registers measured 16 % slower).  The plain walk is left to ptxas'
// own choice (64 registers, 16 B of spills), which measured 3 % faster than the explicit bound.
__global__ void __launch_bounds__(256) k_closest_grid(SceneDev sc, CGridDev g, CGridArgs a) { closest_grid_body<false>(sc, g, a); }
__global__ void __launch_bounds__(256, 4) k_closest_grid_key(SceneDev sc, CGridDev g, CGridArgs a) { closest_grid_body<true>(sc, g, a); }

}  // namespace b200rt
