/*
 * oracle/render_f64.c — CPU float64 restatement of the reference's `render` hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (python-ray-tracer_b200/) may
 * import, link or call this file.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker / the timed CPU baseline.
 *
 * Parity status: PINNED by generated goldens.  The reference (peter-seres/python-ray-tracer)
 * ships no tests or golden vectors; the pins are frames and function-level vectors produced by
 * running the unmodified reference under NUMBA_ENABLE_CUDASIM=1 (oracle/sim_harness.py,
 * tests/golden/make_golden.py), committed under tests/golden/.  tests/test_oracle_golden.py
 * checks this file against every one of them.
 *
 * Every function follows the cited reference lines operation by operation, in the same
 * left-to-right order, in double precision, with the reference's genuine float32 sites kept
 * (sphere r*r, plane-normal re-normalisation).  Compile with -ffp-contract=off so that no FMA
 * contraction happens: the numba simulator executes plain Python/NumPy scalar arithmetic.
 *
 * `numpy_weak` (args->numpy_weak != 0) additionally reproduces two artefacts that exist ONLY in
 * the simulator under NumPy >= 2 weak-scalar promotion (python float * np.float32 -> float32) and
 * not in numba's compiled kernel (where amb/BIAS are float64):
 *   (1) `ambient_int * colour` (trace.py:77) is rounded to float32 and the RGB triple stays
 *       float32 until a float64 term is added to it;
 *   (2) `BIAS * N` for a plane normal (trace.py:83, N is float32 from common.py:109-110) is a
 *       float32 product.
 * The default (0) is the compiled-kernel semantics: everything float64.
 *
 * Reference files (all under /root/reference/src/ray_tracing/): kernels.py, trace.py,
 * intersections.py, common.py.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

typedef struct { double x, y, z; } v3;

typedef struct {
    const float* spheres; int S;   /* (7,S) row-major: rows Cx,Cy,Cz,r,R,G,B   scene.py:17-23,69-76 */
    const float* lights;  int L;   /* (3,L)                                     scene.py:32-36,87-94 */
    const float* planes;  int Pl;  /* (9,Pl): rows ox,oy,oz,nx,ny,nz,R,G,B      scene.py:47-53,78-85 */
    int numpy_weak;
} scene_t;

/* RGB accumulator that remembers whether NumPy would hold it as float32 (numpy_weak mode). */
typedef struct { double c[3]; int f32; } rgb_t;

typedef struct { uint64_t primary, shadow, bounce; } ray_counts_t;

/* common.py:10-18  p = c1*a + c2*b */
static inline v3 linear_comb(v3 a, v3 b, double c1, double c2) {
    v3 p; p.x = c1 * a.x + c2 * b.x; p.y = c1 * a.y + c2 * b.y; p.z = c1 * a.z + c2 * b.z; return p;
}
/* common.py:21-25  to - from */
static inline v3 vector_difference(v3 from, v3 to) {
    v3 d; d.x = to.x - from.x; d.y = to.y - from.y; d.z = to.z - from.z; return d;
}
/* common.py:28-32 (no zero guard) */
static inline v3 normalize(v3 v) {
    double norm = sqrt(v.x * v.x + v.y * v.y + v.z * v.z);
    v3 r; r.x = v.x / norm; r.y = v.y / norm; r.z = v.z / norm; return r;
}
/* common.py:35-37 */
static inline double dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

/* common.py:52-57  min(max(0, int(round(c))), 255); round() is half-to-even */
static inline int clip_color(double c) {
    double r = nearbyint(c);      /* default rounding mode = to nearest even */
    if (!(r > 0.0)) return 0;     /* also catches NaN the way int(max(0, nan)) cannot: not reachable */
    if (r > 255.0) return 255;
    return (int)r;
}

/* intersections.py:6-38 */
static double intersect_ray_sphere(v3 O, v3 D, v3 C, float radius) {
    v3 R = normalize(D);                              /* :13 (redundant re-normalisation kept) */
    v3 L = vector_difference(C, O);                   /* :16  O - C */
    double a = dot(R, R);                             /* :19 */
    double b = 2 * dot(L, R);                         /* :20 */
    float r2 = radius * radius;                       /* :21  float32 * float32 */
    double c = dot(L, L) - (double)r2;                /* :21 */
    double disc = b * b - 4 * a * c;                  /* :23 */
    if (disc < 0.0) return -999.9;                    /* :25-26 */
    double num = -b - sqrt(disc);                     /* :28 */
    if (num > 0.0) return num / (2 * a);              /* :30-31 */
    num = -b + sqrt(disc);                            /* :33 */
    if (num > 0.0) return num / (2 * a);              /* :35-36 */
    return -999.0;                                    /* :38 */
}

/* intersections.py:41-68 */
static double intersect_ray_plane(v3 O, v3 D, v3 po, v3 pn) {
    double denom = dot(D, pn);                        /* :52 */
    if (fabs(denom) < 0.001) return -999.9;           /* :55-56 */
    v3 LP = vector_difference(O, po);                 /* :59  po - O */
    double nominator = dot(LP, pn);                   /* :61 */
    double dist = nominator / denom;                  /* :63 */
    if (dist > 0) return dist;                        /* :65-66 */
    return -999.0;                                    /* :68 */
}

static inline v3 sphere_center(const scene_t* sc, int i) {
    v3 c; c.x = sc->spheres[0 * sc->S + i]; c.y = sc->spheres[1 * sc->S + i]; c.z = sc->spheres[2 * sc->S + i]; return c;
}

/* trace.py:7-41 */
static double get_intersection(v3 O, v3 D, const scene_t* sc, int* obj_index, int* obj_type) {
    double best = 999.0; int idx = -999; int type = 404;            /* :17-19 */
    for (int i = 0; i < sc->S; ++i) {                               /* :22 */
        double d = intersect_ray_sphere(O, D, sphere_center(sc, i), sc->spheres[3 * sc->S + i]);
        if (best > d && d > 0) { best = d; idx = i; type = 0; }     /* :26-29 */
    }
    for (int i = 0; i < sc->Pl; ++i) {                              /* :32 */
        v3 po = { sc->planes[0 * sc->Pl + i], sc->planes[1 * sc->Pl + i], sc->planes[2 * sc->Pl + i] };
        v3 pn = { sc->planes[3 * sc->Pl + i], sc->planes[4 * sc->Pl + i], sc->planes[5 * sc->Pl + i] };
        double d = intersect_ray_plane(O, D, po, pn);
        if (best > d && d > 0) { best = d; idx = i; type = 1; }     /* :36-39 */
    }
    *obj_index = idx; *obj_type = type;
    return best;
}

/* ---- typed scalar helpers for numpy_weak mode: 0 = float64, 1 = float32, 2 = python float ---- */
static inline double typed_add(double a, int ta, double b, int tb, int* tout) {
    if (ta == 0 || tb == 0) { *tout = 0; return a + b; }
    if (ta == 1 || tb == 1) { *tout = 1; return (double)((float)a + (float)b); }
    *tout = 2; return a + b;
}
/* python-float weight w times a typed value */
static inline double typed_scale(double w, double v, int tv, int* tout) {
    if (tv == 1) { *tout = 1; return (double)((float)w * (float)v); }
    *tout = tv; return w * v;
}

typedef struct { rgb_t rgb; v3 P; v3 R; } trace_out_t;

/* trace.py:44-112 */
static trace_out_t trace(v3 O, v3 D, const scene_t* sc, double amb, double lamb,
                         ray_counts_t* cnt, int is_primary) {
    trace_out_t out;
    const int weak = sc->numpy_weak;
    out.rgb.c[0] = out.rgb.c[1] = out.rgb.c[2] = 0.0; out.rgb.f32 = weak ? 2 : 0;   /* :49 */
    int idx, type;
    if (is_primary) cnt->primary++; else cnt->bounce++;
    double dist = get_intersection(O, D, sc, &idx, &type);                          /* :53 */
    if (type == 404) {                                                              /* :56-57 */
        out.P.x = out.P.y = out.P.z = 404.0; out.R.x = out.R.y = out.R.z = 404.0;
        return out;
    }
    v3 P = linear_comb(O, D, 1.0, dist);                                            /* :60 */
    double col[3]; v3 N; int n_is_f32 = 0;
    if (type == 0) {                                                                /* :63-66 */
        col[0] = sc->spheres[4 * sc->S + idx]; col[1] = sc->spheres[5 * sc->S + idx]; col[2] = sc->spheres[6 * sc->S + idx];
        N = normalize(vector_difference(sphere_center(sc, idx), P));                /* common.py:94-101 */
    } else {                                                                        /* :68-71 */
        col[0] = sc->planes[6 * sc->Pl + idx]; col[1] = sc->planes[7 * sc->Pl + idx]; col[2] = sc->planes[8 * sc->Pl + idx];
        /* common.py:104-110: normalize() on three float32 values -> float32 arithmetic */
        float n0 = sc->planes[3 * sc->Pl + idx], n1 = sc->planes[4 * sc->Pl + idx], n2 = sc->planes[5 * sc->Pl + idx];
        float s0 = n0 * n0, s1 = n1 * n1, s2 = n2 * n2;
        float s01 = s0 + s1; float ss = s01 + s2;
        float norm = sqrtf(ss);
        float q0 = n0 / norm, q1 = n1 / norm, q2 = n2 / norm;
        N.x = q0; N.y = q1; N.z = q2; n_is_f32 = 1;
    }
    /* :77  RGB = 1.0*RGB + amb*colour */
    for (int k = 0; k < 3; ++k) {
        if (weak) { out.rgb.c[k] = (double)((float)0.0 + (float)amb * (float)col[k]); }
        else      { out.rgb.c[k] = 1.0 * 0.0 + amb * col[k]; }
    }
    out.rgb.f32 = weak ? 1 : 0;
    const double BIAS = 0.0002;                                                     /* :82 */
    if (weak && n_is_f32) {                                                         /* :83 (simulator artefact 2) */
        P.x = 1.0 * P.x + (double)((float)BIAS * (float)N.x);
        P.y = 1.0 * P.y + (double)((float)BIAS * (float)N.y);
        P.z = 1.0 * P.z + (double)((float)BIAS * (float)N.z);
    } else {
        P = linear_comb(P, N, 1.0, BIAS);                                           /* :83 */
    }
    for (int j = 0; j < sc->L; ++j) {                                               /* :86 */
        v3 lp = { sc->lights[0 * sc->L + j], sc->lights[1 * sc->L + j], sc->lights[2 * sc->L + j] };
        v3 Lv = normalize(vector_difference(P, lp));                                /* :89, common.py:84-91 */
        int sidx, stype;
        cnt->shadow++;
        (void)get_intersection(P, Lv, sc, &sidx, &stype);                           /* :92 (full closest-hit, F8) */
        if (stype != 404) continue;                                                 /* :95-96 */
        double k = lamb * dot(Lv, N);                                               /* :99 */
        if (k > 0) {                                                                /* :101-102 */
            for (int c = 0; c < 3; ++c) out.rgb.c[c] = 1.0 * out.rgb.c[c] + k * col[c];
            out.rgb.f32 = 0;
        }
    }
    /* :107, common.py:113-120 */
    double DdN = dot(D, N);
    v3 R = normalize(linear_comb(D, N, 1.0, -2.0 * DdN));
    P = linear_comb(P, R, 1.0, BIAS);                                               /* :110 */
    out.P = P; out.R = R;
    return out;
}

static inline int is404(v3 v) { return v.x == 404. && v.y == 404. && v.z == 404.; }

/* trace.py:115-133 */
static rgb_t sample(v3 O, v3 D, const scene_t* sc, double amb, double lamb, double refl,
                    int64_t refl_depth, ray_counts_t* cnt) {
    trace_out_t t = trace(O, D, sc, amb, lamb, cnt, 1);                             /* :120 */
    rgb_t RGB = t.rgb; v3 P = t.P, R = t.R;
    for (int64_t i = 0; i < refl_depth; ++i) {                                      /* :123 */
        if (is404(P) || is404(R)) continue;                                         /* :124-126 */
        trace_out_t r = trace(P, R, sc, amb, lamb, cnt, 0);                         /* :128 */
        P = r.P; R = r.R;
        double w = pow(refl, (double)(i + 1));                                      /* :131 */
        int tt = 0, ts = 0;
        for (int c = 0; c < 3; ++c) {
            double term = typed_scale(w, r.rgb.c[c], r.rgb.f32, &tt);
            RGB.c[c] = typed_add(1.0 * RGB.c[c], RGB.f32, term, tt, &ts);
        }
        RGB.f32 = ts;
    }
    return RGB;
}

/* ---------------- render: kernels.py:6-73 ---------------- */
typedef struct {
    uint32_t struct_size;
    int W, H;
    int x_begin, x_end, y_begin, y_end;    /* pixel region to render, half-open */
    const double* pixel_loc;               /* (3,W,H) C-order float64, or NULL -> use the affine grid below */
    double grid_f, grid_ar, grid_sy, grid_sz; /* camera.py:18-26 closed form: [0]=f, [1]=ar+x*sy, [2]=1+y*sz */
    double cam_origin[3];
    double cam_rot[9];                     /* row-major */
    const float* spheres; int S;
    const float* lights;  int L;
    const float* planes;  int Pl;
    double amb, lamb, refl;
    int64_t refl_depth;
    int aliasing;
    int numpy_weak;
    int threads;                           /* 0 = OpenMP default */
    uint8_t* result;                       /* (3,W,H) uint8; only the region is written */
    double* result_f64;                    /* optional (3,W,H): the three values handed to the clip, in output order */
    uint64_t* counts;                      /* optional [3]: primary, shadow, bounce queries */
} oracle_args;

/* GPU flat addressing of pixel_loc[c, x, y] with no bounds check: C-order offset, elements
 * past the end of the buffer are DEFINED as 0.0 (SURVEY §8(a) A1 contract). */
static inline double pl_at(const oracle_args* a, int c, int x, int y) {
    const int64_t W = a->W, H = a->H;
    int64_t flat = (int64_t)c * W * H + (int64_t)x * H + y;
    if (flat < 0 || flat >= 3 * W * H) return 0.0;
    if (a->pixel_loc) return a->pixel_loc[flat];
    int cc = (int)(flat / (W * H)); int64_t rem = flat % (W * H);
    int xx = (int)(rem / H), yy = (int)(rem % H);
    if (cc == 0) return a->grid_f;
    if (cc == 1) return xx * a->grid_sy + a->grid_ar;   /* numpy mgrid: index*step + start */
    return yy * a->grid_sz + 1.0;
}
static inline v3 pl3(const oracle_args* a, int x, int y) {
    v3 p = { pl_at(a, 0, x, y), pl_at(a, 1, x, y), pl_at(a, 2, x, y) }; return p;
}
static inline v3 matmul(const double* R, v3 p) {      /* common.py:40-49 */
    v3 b; b.x = R[0] * p.x + R[1] * p.y + R[2] * p.z;
    b.y = R[3] * p.x + R[4] * p.y + R[5] * p.z;
    b.z = R[6] * p.x + R[7] * p.y + R[8] * p.z; return b;
}

ORACLE_API int oracle_render(const oracle_args* a) {
    if (!a || a->struct_size != sizeof(oracle_args)) return -1;
    if (a->W <= 0 || a->H <= 0 || !a->result) return -2;
    scene_t sc = { a->spheres, a->S, a->lights, a->L, a->planes, a->Pl, a->numpy_weak };
    const v3 O = { a->cam_origin[0], a->cam_origin[1], a->cam_origin[2] };          /* :16 */
    const int64_t WH = (int64_t)a->W * a->H;
    uint64_t c_primary = 0, c_shadow = 0, c_bounce = 0;
#ifdef _OPENMP
    if (a->threads > 0) omp_set_num_threads(a->threads);
#endif
    #pragma omp parallel for schedule(dynamic, 1) reduction(+:c_primary, c_shadow, c_bounce)
    for (int x = a->x_begin; x < a->x_end; ++x) {
        ray_counts_t cnt = {0, 0, 0};
        for (int y = a->y_begin; y < a->y_end; ++y) {
            v3 P = pl3(a, x, y);                                                     /* :19 */
            v3 dir = normalize(matmul(a->cam_rot, P));                               /* :22-23 */
            rgb_t s = sample(O, dir, &sc, a->amb, a->lamb, a->refl, a->refl_depth, &cnt);   /* :26 */
            double R = s.c[0], G = s.c[1], B = s.c[2]; int tR = s.f32, tG = s.f32, tB = s.f32;
            /* :29 — for an in-range pixel the two upper tests are vacuous */
            if (a->aliasing && x + 1 <= a->W && x - 1 >= 0 && y + 1 <= a->H && y - 1 >= 0) {
                const int dx[8] = { -1, +1, 0, 0, -1, +1, -1, +1 };                  /* :32-41 */
                const int dy[8] = { 0, 0, +1, -1, +1, +1, -1, -1 };
                for (int k = 0; k < 8; ++k) {
                    v3 Q = linear_comb(P, pl3(a, x + dx[k], y + dy[k]), 0.5, 0.5);   /* :43-50 */
                    v3 d = normalize(matmul(a->cam_rot, Q));                         /* :54-55 */
                    rgb_t e = sample(O, d, &sc, a->amb, a->lamb, a->refl, a->refl_depth, &cnt);
                    R = typed_add(R, tR, e.c[0], e.f32, &tR);                        /* :58 */
                    G = typed_add(G, tG, e.c[2], e.f32, &tG);                        /* :59  G += B_s (F6) */
                    B = typed_add(B, tB, e.c[1], e.f32, &tB);                        /* :60  B += G_s (F6) */
                }
                if (tR == 1) { R = (double)((float)R / 9.0f); G = (double)((float)G / 9.0f); B = (double)((float)B / 9.0f); }
                else { R = R / 9; G = G / 9; B = B / 9; }                           /* :63-65 */
            }
            /* :69, common.py:60-63: (clip(R), clip(B), clip(G))  (F5) */
            const int64_t o = (int64_t)x * a->H + y;
            a->result[0 * WH + o] = (uint8_t)clip_color(R);
            a->result[1 * WH + o] = (uint8_t)clip_color(B);
            a->result[2 * WH + o] = (uint8_t)clip_color(G);
            if (a->result_f64) { a->result_f64[0 * WH + o] = R; a->result_f64[1 * WH + o] = B; a->result_f64[2 * WH + o] = G; }
        }
        c_primary += cnt.primary; c_shadow += cnt.shadow; c_bounce += cnt.bounce;
    }
    if (a->counts) { a->counts[0] = c_primary; a->counts[1] = c_shadow; a->counts[2] = c_bounce; }
    return 0;
}

/* ---------------- function-level entry points for known-answer tests ---------------- */
ORACLE_API double oracle_intersect_ray_sphere(const double* O, const double* D, const float* C, float r) {
    v3 o = { O[0], O[1], O[2] }, d = { D[0], D[1], D[2] }, c = { C[0], C[1], C[2] };
    return intersect_ray_sphere(o, d, c, r);
}
ORACLE_API double oracle_intersect_ray_plane(const double* O, const double* D, const float* po, const float* pn) {
    v3 o = { O[0], O[1], O[2] }, d = { D[0], D[1], D[2] }, p = { po[0], po[1], po[2] }, n = { pn[0], pn[1], pn[2] };
    return intersect_ray_plane(o, d, p, n);
}
ORACLE_API double oracle_get_intersection(const double* O, const double* D, const float* spheres, int S,
                                          const float* planes, int Pl, int* idx, int* type) {
    scene_t sc = { spheres, S, NULL, 0, planes, Pl, 0 };
    v3 o = { O[0], O[1], O[2] }, d = { D[0], D[1], D[2] };
    return get_intersection(o, d, &sc, idx, type);
}
/* out[0..2]=RGB, out[3..5]=P, out[6..8]=R */
ORACLE_API void oracle_trace(const double* O, const double* D, const float* spheres, int S, const float* lights, int L,
                             const float* planes, int Pl, double amb, double lamb, int numpy_weak, double* out) {
    scene_t sc = { spheres, S, lights, L, planes, Pl, numpy_weak };
    v3 o = { O[0], O[1], O[2] }, d = { D[0], D[1], D[2] };
    ray_counts_t cnt = {0, 0, 0};
    trace_out_t t = trace(o, d, &sc, amb, lamb, &cnt, 1);
    out[0] = t.rgb.c[0]; out[1] = t.rgb.c[1]; out[2] = t.rgb.c[2];
    out[3] = t.P.x; out[4] = t.P.y; out[5] = t.P.z; out[6] = t.R.x; out[7] = t.R.y; out[8] = t.R.z;
}
ORACLE_API void oracle_sample(const double* O, const double* D, const float* spheres, int S, const float* lights, int L,
                              const float* planes, int Pl, double amb, double lamb, double refl, int64_t refl_depth,
                              int numpy_weak, double* out, uint64_t* counts) {
    scene_t sc = { spheres, S, lights, L, planes, Pl, numpy_weak };
    v3 o = { O[0], O[1], O[2] }, d = { D[0], D[1], D[2] };
    ray_counts_t cnt = {0, 0, 0};
    rgb_t s = sample(o, d, &sc, amb, lamb, refl, refl_depth, &cnt);
    out[0] = s.c[0]; out[1] = s.c[1]; out[2] = s.c[2];
    if (counts) { counts[0] = cnt.primary; counts[1] = cnt.shadow; counts[2] = cnt.bounce; }
}
ORACLE_API int oracle_clip_color(double c) { return clip_color(c); }
ORACLE_API int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
ORACLE_API uint32_t oracle_args_size(void) { return (uint32_t)sizeof(oracle_args); }
