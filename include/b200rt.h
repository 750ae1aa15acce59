/*
 * b200rt.h — C ABI of libb200rt.so, the B200 (sm_100a) render path.
 *
 * This is the drop-in boundary for ONE path of peter-seres/python-ray-tracer: the numba kernel launch
 *
 *     render[blockspergrid, threadsperblock](pixel_loc, result, camera_origin, camera_rotation,
 *                                            spheres, lights, planes, amb, lamb, refl, refl_depth, aliasing)
 *
 * (reference src/main.py:41-42,46-47; kernel src/ray_tracing/kernels.py:6-73 and everything it inlines:
 * src/ray_tracing/trace.py:7-133, intersections.py:6-68, common.py:5-120).  The reference has no FFI; the
 * binding a maintainer adds is the ctypes stub shown in INTEGRATION.md (python-ray-tracer_b200/b200rt/_lib.py).
 *
 * Plain pointers and sizes only; no torch / numba types.  All device buffers are caller-owned.
 * Every entry point returns 0 on success, <0 for argument errors (B200RT_E_*), >0 for a cudaError_t.
 * b200rt_last_error() returns a thread-local human-readable message for the last failure.
 */
#ifndef B200RT_H
#define B200RT_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200RT_VERSION 1

#if defined(__GNUC__)
#define B200RT_API __attribute__((visibility("default")))
#else
#define B200RT_API
#endif

/* argument errors */
#define B200RT_E_NULL        (-1)   /* args == NULL or a required pointer is NULL          */
#define B200RT_E_STRUCT_SIZE (-2)   /* args->struct_size != sizeof(b200rt_args)            */
#define B200RT_E_SHAPE       (-3)   /* W/H/S/L/Pl/x range out of bounds                    */
#define B200RT_E_DEPTH       (-4)   /* refl_depth < 0 or > B200RT_MAX_DEPTH                */
#define B200RT_E_DEVICE      (-5)   /* no such CUDA device / not an sm_100 device           */
#define B200RT_E_NOMEM       (-6)   /* scratch allocation failed                            */

#define B200RT_MAX_DEPTH 64

/* flags */
#define B200RT_F_SCENE_ON_HOST   0x1u  /* spheres/lights/planes point to HOST memory (b200rt_render only; implied by _host) */
#define B200RT_F_NO_FILTER       0x2u  /* debug: run the exact sphere test on every sphere (must give bit-identical output) */
#define B200RT_F_UNSWAP_GB       0x4u  /* write channels as R,G,B instead of the reference's R,B,G wiring (common.py:60-63) */
#define B200RT_F_NO_CULL         0x8u  /* k_isect3 scans every sphere for every ray group (no beam culling); the image is the same */
#define B200RT_F_TIME_ONLY       0x20u /* with `stats`: fill the ray counts and the per-class device times but run the production kernels (no executed-work counters: slow_path, exact_tests, tests_*, and with the light grids shadow_traced, stay 0) */
#define B200RT_F_DENSE           0x40u /* brute force as the reference specifies it: every query (shadow queries of unlit pairs included) scans every sphere - no beam culling, regrouping, shadow passes or lit-pair lists; the image is the same.  The FP32-peak figure of bench.py is measured in this mode */
#define B200RT_F_NO_SGRID        0x80u /* shadow queries scan the light's sphere table (k_isect3, passes + beam culling) instead of looking up the light's direction grid; the image is the same */
#define B200RT_F_NO_GRAPH        0x100u /* launch the frame's kernels one by one instead of replaying the captured CUDA graph of the sequence (A/B, debugging); the image is the same */
#define B200RT_F_NO_CACHE        0x200u /* rebuild every scene-derived table in this call (the resident-scene tables are normally rebuilt only when the scene arrays change); implies NO_GRAPH; the image is the same */
#define B200RT_F_NO_CGRID        0x400u /* closest-hit queries of bounce rays always by the beam-culled scan (k_isect3), never by the per-ray walk through the uniform grid of the spheres (closest_grid.cuh); the image is the same */
#define B200RT_F_NO_FUSED        0x800u /* scenes with fewer than 64 spheres run through the wavefront kernels (k_isect3, k_setup, k_shadow_shade) instead of the one-kernel-per-depth path (fused_small.cuh); the image is the same */
#define B200RT_F_NO_SORT         0x10u /* bounce-depth hits stay in pixel order (no regrouping by direction/position bin); the image is the same */

/*
 * Closed form of Camera.generate_pixel_locations() (reference src/scene/camera.py:18-26), evaluated the way
 * numpy.mgrid does (index*step + start):
 *     pixel_loc[0][x][y] = f          pixel_loc[1][x][y] = x*sy + ar          pixel_loc[2][x][y] = y*sz + 1.0
 * Used when args->pixel_loc == NULL, so no 24*W*H-byte array has to exist (796 MB at 7680x4320).
 */
typedef struct b200rt_grid {
    double f, ar, sy, sz;
} b200rt_grid;

/* Ray = one get_intersection query (reference trace.py:53 closest-hit, trace.py:92 shadow). */
typedef struct b200rt_stats {
    uint64_t primary;   /* first query of every sample (kernels.py:26,56 -> trace.py:120 -> :53) */
    uint64_t shadow;    /* one per light per hit-trace (trace.py:92)                              */
    uint64_t bounce;    /* first query of every reflection trace (trace.py:128 -> :53)            */
    uint64_t slow_path; /* thread-level entries into the candidate handler (diagnostic)           */
    uint64_t exact_tests; /* exact (robust) ray-sphere tests run on filter candidates (diagnostic)  */
    uint64_t shadow_traced; /* shadow queries actually traced: (hit, light) pairs with lamb*(Lv.N) > 0; the others are
                               queries whose answer the reference computes and then ignores (trace.py:92-102)     */
    uint64_t tests_pivot;   /* (ray, sphere) filter tests EXECUTED in the 4-FMA pivot form (primary + shadow passes; padding included) */
    uint64_t tests_general; /* ... in the 8-FMA origin-free form (bounce rays).  Canonical brute force would be rays * S.          */
    float    kernel_ms[4]; /* device time: [0] total, [1] closest-hit kernels, [2] shadow kernels, [3] everything else */
} b200rt_stats;

typedef struct b200rt_args {
    uint32_t struct_size;          /* = sizeof(b200rt_args)                                                     */
    uint32_t flags;                /* B200RT_F_*                                                                */
    int32_t  device;               /* CUDA device ordinal                                                       */
    int32_t  reserved0;
    void*    stream;               /* cudaStream_t (NULL = default stream); the launch is asynchronous on it    */

    /* kernels.py:7 argument 1: pixel_loc float64 (3,W,H) C-order, DEVICE pointer; NULL -> use `grid`.          */
    const double* pixel_loc;
    b200rt_grid   grid;
    /* argument 2: result uint8 (3,W,H) C-order, indexed [channel][x][y] (kernels.py:71-73); DEVICE pointer.    */
    uint8_t* result;
    int32_t  W, H;                 /* = pixel_loc.shape[1], pixel_loc.shape[2]                                   */
    int32_t  x_begin, x_end;       /* column tile [x_begin, x_end) along array axis 1; whole frame = [0, W)     */

    double   cam_origin[3];        /* argument 3: camera_origin (src/scene/camera.py:14-16)                      */
    double   cam_rot[9];           /* argument 4: camera_rotation, row-major 3x3 (src/scene/rotation.py:34-43)   */

    const float* spheres; int32_t S;   /* argument 5: float32 (7,S): rows Cx,Cy,Cz,r,R,G,B (src/scene/scene.py:17-23,69-76) */
    const float* lights;  int32_t L;   /* argument 6: float32 (3,L)                       (scene.py:32-36,87-94)            */
    const float* planes;  int32_t Pl;  /* argument 7: float32 (9,Pl): ox,oy,oz,nx,ny,nz,R,G,B (scene.py:47-53,78-85)        */

    double   amb, lamb, refl;      /* arguments 8-10 (src/main.py:11)                                            */
    int64_t  refl_depth;           /* argument 11                                                                */
    int32_t  aliasing;             /* argument 12 (bool)                                                         */
    int32_t  reserved1;

    /* Multi-GPU interleaving: when strip_width > 0 the tile [x_begin,x_end) is cut into strips of
     * strip_width columns and only strips i with i % strip_world == strip_rank are rendered (one wavefront for
     * all of them).  strip_width == 0 renders the whole tile.  Pixels outside the selected strips are untouched. */
    int32_t  strip_width, strip_rank, strip_world, reserved2;

    b200rt_stats* stats;           /* optional HOST pointer; non-NULL makes the call synchronous and fills it    */
} b200rt_args;

B200RT_API int         b200rt_version(void);
B200RT_API const char* b200rt_last_error(void);
B200RT_API uint32_t    b200rt_args_size(void);

/* Render tile [x_begin,x_end) of the frame into the DEVICE buffer args->result.  Asynchronous on args->stream
 * unless args->stats != NULL.  Replaces the numba launch at reference src/main.py:41-42. */
B200RT_API int b200rt_render(const b200rt_args* args);

/* Same call with HOST buffers: args->pixel_loc (nullable), args->result, spheres/lights/planes are host
 * pointers.  Stages the inputs, renders, copies the tile of `result` back and synchronises — the
 * to_device / launch / copy_to_host sequence of reference src/main.py:19-51 in one call. */
B200RT_API int b200rt_render_host(const b200rt_args* args);

/* Camera-pose batch with HOST buffers (the caller of the path for animations; BASELINE config 5): `args` as for
 * b200rt_render_host (its cam_origin/cam_rot are ignored; result must be non-NULL, e.g. = results), cam_origins float64 (n,3), cam_rots float64 (n,9) row-major,
 * results uint8 (n,3,W,H).  The scene is staged once; frame k's D2H copy overlaps frame k+1's render.  Synchronous. */
B200RT_API int b200rt_render_batch_host(const b200rt_args* args, int n_frames, const double* cam_origins, const double* cam_rots,
                                        uint8_t* results);

/* Measured FP32 CUDA-core throughput of `device` (dependent-chain-free FFMA / FFMA2 loops), TFLOP/s, and the
 * SM clock (MHz) derived from clock64()/globaltimer during the probe.  Used as the roofline denominator. */
B200RT_API int b200rt_fp32_probe(int device, double* ffma_tflops, double* ffma2_tflops, double* sm_mhz);

/* Presentation step, reference src/viewer/image.py:7-19 (convert_array_to_image): planar uint8 (3,W,H) indexed
 * [c][x][y]  ->  interleaved RGB8 image of H rows x W pixels, out[y][x][c] = in[c][x][y] (what `mirror(rotate(270))`
 * yields for the square frames the reference renders; non-square frames come out whole instead of cropped).
 * flags: B200RT_F_UNSWAP_GB writes R,G,B instead of the stored R,B,G.  Device pointers; asynchronous on `stream`. */
B200RT_API int b200rt_to_image(int device, void* stream, const uint8_t* planar, uint8_t* rgb, int W, int H, uint32_t flags);
/* Same with HOST buffers (stages, transposes on the device, copies back, synchronises). */
B200RT_API int b200rt_to_image_host(int device, const uint8_t* planar, uint8_t* rgb, int W, int H, uint32_t flags);

/* Allocation-free variant of b200rt_to_image_host.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (tight serving loops, harnesses that forbid it).  Same
 * arguments, same result bytes; the device staging buffers are one file-scope `static __device__` array of
 * 2 * 3 * B200RT_NOALLOC_MAX_PIXELS bytes (6 MiB, held by every GPU that loads the library until exit), so the call
 * performs no cudaMalloc / cudaFree / allocating library routine.  W*H > B200RT_NOALLOC_MAX_PIXELS is rejected with
 * B200RT_E_SHAPE.  All calls on a device share the array (it keeps the last call's contents; the call clears what it
 * hands back); calls overlapping on the GPU would race on it, so the entry point serialises its callers per device
 * (mutex, default stream, synchronous). */
#define B200RT_NOALLOC_MAX_PIXELS (1 << 20)   /* 1024 x 1024; the reference's main.py frame is 1000 x 1000 (src/main.py:10) */
B200RT_API int b200rt_to_image_host_noalloc(int device, const uint8_t* planar, uint8_t* rgb, int W, int H, uint32_t flags);

/* Page-locked host memory for the framebuffer: the destination of `result.copy_to_host()` (reference src/main.py:51; numba's
 * counterpart of b200rt_alloc_host is cuda.pinned_array).  b200rt_render_host / _batch_host copy into whatever `result`
 * points at; when that memory is pinned the copy is one DMA at PCIe rate.  b200rt_host_register pins memory the caller
 * owns, e.g. a shared-memory mapping of the (3,W,H) frame that the ranks of a multi-GPU job copy their strips into — the
 * "direct device-to-host gather" of the tiles (north_star).  CUDA runtime must be usable (a GPU present). */
B200RT_API int b200rt_alloc_host(size_t bytes, void** out);
B200RT_API int b200rt_free_host(void* p);
B200RT_API int b200rt_host_register(void* p, size_t bytes);
B200RT_API int b200rt_host_unregister(void* p);

B200RT_API int b200rt_device_count(void);
B200RT_API int b200rt_sm_count(int device);
/* Kernels launched by this library since it was loaded (host-side count, all devices and threads): the difference
 * across a timed region is the measured `gpu_launches` of bench.py. */
B200RT_API unsigned long long b200rt_launch_count(void);
/* Frames on `device` that were replayed from a captured CUDA graph of their launch sequence (diagnostic; the second call with
 * the same shapes / buffers / flags captures, later ones replay; the camera pose is not part of the key). */
B200RT_API unsigned long long b200rt_graph_replays(int device);
/* Checked build only (-DB200RT_CHECK, tools/gpu_check.sh): number of failed index assertions of the library's kernels since load on the
 * current device; ~0 when the library was not built with the checks. */
B200RT_API unsigned long long b200rt_debug_check_failures(void);
/* Free the per-device scratch (packed scene, ray queues).  Safe to call at any time between renders. */
B200RT_API int b200rt_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_H */
