/*
 * include/rt_abi.h -- C ABI of the B200 render-loop library (librt_b200.so).
 *
 * This is the drop-in boundary for ONE path of kaitlyndlee/Raytracing-Programs:
 *     void raycast(json_data_t *json_struct, PPMFormat photo_data)   (C/raycast.h:8,
 *     body C/raycast.c:79-130, called once from main at C/raycast.c:57)
 * and everything below it (iterative_shoot :315, calc_color :381, find_nearest_object
 * :441, ray_intersect_{plane,sphere,quadric} :545/:576/:614, the attenuation / Phong
 * helpers :666-758 and v3math.c).  The reference has no FFI layer; its "operator API"
 * is that one function, so the entry points here are what a cgo/ctypes/JNI binding for
 * the path would bind.  Plain C types only, no exceptions, every call returns 0 on
 * success or a negative rt_status (text via rt_last_error()).
 *
 * There is NO CPU fallback: every rt_render* call fails with RT_ERR_NO_DEVICE when no
 * sm_100-class GPU is present.
 */
#ifndef RT_ABI_H
#define RT_ABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Version of this header, returned by rt_abi_version() of the library that was loaded: a binding compares the two
 * before it passes any struct.  History: 1 = round-1 surface; 2 = rt_stats grew by scan_slots, ball_tests_o,
 * ball_tests_d (structs only ever grow at the end, and only together with a version bump: a caller built against an
 * older header would hand the library a buffer that is too short), new calls rt_set_mode, rt_rank_rows_to_host,
 * rt_p3_format_device_noalloc. */
#define RT_ABI_VERSION 2

typedef enum rt_status {
  RT_OK = 0,
  RT_ERR_NO_DEVICE = -1,   /* no CUDA device / not initialised */
  RT_ERR_CUDA = -2,        /* a CUDA runtime call failed */
  RT_ERR_ARG = -3,         /* bad argument (NULL, non-positive size, unknown type) */
  RT_ERR_NO_SCENE = -4,    /* rt_render* before rt_scene_upload */
  RT_ERR_NCCL = -5,        /* NCCL missing or failed (only when n_gpus > 1) */
  RT_ERR_OOM = -6
} rt_status;

/* shape_type_t / light_type_t of the reference (C/objects.h:4-13), same values. */
enum { RT_SPHERE = 0, RT_PLANE = 1, RT_QUADRIC = 2 };
enum { RT_POINT = 0, RT_SPOTLIGHT = 1 };

/* One shape, in FILE ORDER (index = position in the reference's shapes list; the
 * nearest-hit tie-break and the self-skip use it, C/raycast.c:450,456).
 * Replaces shape_t (C/objects.h:15-48) without the union and the next pointer.
 * params: sphere {radius}; plane {normal x,y,z}; quadric {a,b,c,d,e,f,g,h,i,j}. */
typedef struct rt_shape {
  int32_t type;
  float diffuse[3];
  float specular[3];
  float position[3];      /* unused for quadrics (the reference leaves it uninitialised) */
  float reflectivity;
  float refractivity;
  float ior;              /* carried for completeness; the live path never reads it */
  float params[10];
} rt_shape;

/* Replaces light_t (C/objects.h:50-61). cos_theta is the host-side
 * cos(theta) of C/objects.c:179; theta/cos_theta/a0/direction are read for spots only. */
typedef struct rt_light {
  int32_t type;
  float position[3];
  float color[3];
  float radial[3];        /* radial-a0, a1, a2 */
  float theta;
  float cos_theta;
  float a0;               /* angular-a0 */
  float direction[3];
} rt_light;

/* Replaces json_data_t (C/parse.h:11-18) after the list->array copy of C/raycast.c:87-102. */
typedef struct rt_scene {
  float camera_width;
  float camera_height;
  int32_t num_shapes;
  int32_t num_lights;
  const rt_shape *shapes;
  const rt_light *lights;
} rt_scene;

/* Counters of one render (summed over all GPUs used). "Rays" are the
 * find_nearest_object calls the reference semantics make:
 * shadow rays are counted even when the any-hit search ends early. "tests" are the
 * intersection tests the kernels actually executed, split in origin-only ("o", shared
 * by all rays leaving one hit point) and per-direction ("d") parts. */
typedef struct rt_stats {
  uint64_t pixels;
  uint64_t rays_primary;
  uint64_t rays_reflection;
  uint64_t rays_shadow;
  uint64_t hit_points;          /* shaded hits (levels 0..6) */
  uint64_t sphere_tests_o, sphere_tests_d;
  uint64_t plane_tests_o, plane_tests_d;
  uint64_t quadric_tests_o, quadric_tests_d;
  uint64_t kernel_launches;
  double flops;                 /* algorithmic flops executed, SURVEY.md section 8a convention */
  double flops_shade;           /* the part of flops executed by the k_shade launches */
  double ms_kernels;            /* device time of the render kernels (max over GPUs) */
  double ms_primary;            /* per kernel class, CUDA events around every launch on the */
  double ms_shade;              /* launching stream, summed over launches (GPU 0 / this rank) */
  double ms_finalize;
  uint64_t launches_shade;
  double ms_gather;             /* band gather to GPU 0 */
  double ms_total;              /* the whole call, host clock */
  uint64_t scan_slots;          /* ray slots x primitives the bundle scans executed (>= the *_tests_d sum:
                                   slots of rays already occluded, or padding, still ride along) */
  uint64_t ball_tests_o, ball_tests_d;   /* bounding-ball pre-tests of the quadric scan (sphere-type tests: 10 + 8 flops);
                                            quadric_tests_* then count only the quadrics that passed them */
} rt_stats;

/* Initialise n_gpus devices (0 = all visible). Safe to call again after rt_shutdown. */
int rt_init(int n_gpus);
/* One-process-per-GPU launchers (torchrun): use exactly the device with this ordinal. */
int rt_init_device(int ordinal);
int rt_device_count(void);            /* devices in use, 0 before rt_init */
void rt_shutdown(void);
const char *rt_last_error(void);
int rt_abi_version(void);

/* Bound of the reflection loop: MAX_RECURSION of C/raycast.c:14,349 (default 7 = the C program: a pixel sees at
 * most 6 bounces).  The reference's CUDA and OptiX variants run the same loop with MAX_ITER 50
 * (CUDA/raycast.cu:13,216); rt_set_max_depth(50) is that mode.  1..64, takes effect at
 * the next render.  Env RT_MAX_DEPTH in the CLI. */
int rt_set_max_depth(int depth);

/* Which of the reference's two tracers combines the hits of a pixel.
 * RT_MODE_ITERATIVE (default): iterative_shoot, the live path (C/raycast.c:121,315-379): bounce colours weighted by
 * the running product of reflectivities, summed bounce first, primary last.
 * RT_MODE_SHOOT: the dormant recursive shoot() (C/raycast.c:152-296): out_k = colour_k + out_(k+1) * reflectivity_k
 * from the deepest hit outwards.  Its refraction subtree is traced by the reference but never added to the colour
 * (:293,295 are commented out), so the value depends on the reflection chain only: same hits, same shadow rays, a
 * different order of the float operations (pixels differ from the live path in the 1e-6 range).  The discarded
 * refraction rays are not traced here.  Takes effect at the next render.  Env RT_MODE=shoot in the CLI. */
#define RT_MODE_ITERATIVE 0
#define RT_MODE_SHOOT 1
int rt_set_mode(int mode);

/* Acceleration of the nearest-hit search, read by the next rt_scene_upload.
 * 0 (default): every ray tests every shape, as C/raycast.c:449 does.  1: the sphere part of shadow
 * queries goes through a per-light cube-map "light buffer" built at upload time, every other sphere
 * search and the searches over the ellipsoids among the quadrics through bounding-sphere hierarchies;
 * all of it is conservative with respect to the reference's float arithmetic, so the image is the
 * same, byte for byte.  Env RT_ACCEL=0/1 overrides. */
int rt_set_accel(int mode);

/* ---- diagnostics, not part of the drop-in surface (INTEGRATION.md names the calls a binding needs) -------------
 * Declared only with RT_ABI_DEBUG defined; the library always exports them (tests/test_qindex.py binds them by name). */
#ifdef RT_ABI_DEBUG
/* Diagnostics of the quadric index of the accelerated mode (csrc/rt_qindex.cuh); host arithmetic only, no GPU
 * needed.  tests/test_qindex.py drives them against the oracle's quadric intersector (C/raycast.c:614-656) to
 * check that the index can only drop quadrics the reference would not report.
 * rt_dbg_qindex_classify: q10 = a..j of one quadric; returns 1 and fills `out` if the quadric is an ellipsoid
 * the index can hold, else 0.
 * rt_dbg_qindex_prune: the traversal's node test for node4 = {centre, radius}, n_rays rays (P3[3n], d3[3n],
 * best_t[n] when nearest != 0) and consts5 as in rt_qindex_info; codes[n]: 0 enter, 1 line misses, 2 behind
 * the origin, 3 beyond best_t.  Returns 0. */
typedef struct rt_qindex_info {
  double centre[3], K, lmin, lmax;
  double delta[3];                     /* inflation of the level set for an origin at distance s: d0 + d1 s + d2 s^2 */
  float ball[4];                       /* bounding sphere {cx, cy, cz, r} the index stores */
  float consts[5];                     /* miss_rel, behind_rel, behind_abs, t_rel, t_abs */
} rt_qindex_info;
int rt_dbg_qindex_classify(const float *q10, rt_qindex_info *out);
int rt_dbg_qindex_prune(const float *node4, const float *P3, const float *d3, int n_rays, int nearest,
                        const float *best_t, const float *consts5, int *codes);
/* Builds the index over n_quadrics quadrics (q10[10 n]) exactly as rt_scene_upload does and walks it on the host
 * for n_rays rays: visited[r * n_quadrics + k] = 1 if the traversal would test quadric k for ray r (quadrics the
 * index does not hold are always tested).  nearest != 0: nearest-hit search pruned with the fixed best_t[r].
 * Returns the number of indexed quadrics, -1 on bad arguments. */
int rt_dbg_qindex_visit(const float *q10, int n_quadrics, const float *P3, const float *d3, int n_rays, int nearest,
                        const float *best_t, unsigned char *visited);
/* Sizes of the sphere light buffer rt_scene_upload would build for this scene: per_level[num_lights * 6] list
 * entries per (light, distance level), -1 for a level that is not built; returns the total, -1 if there is none. */
long long rt_dbg_light_buffer_sizes(const rt_scene *scene, long long *per_level);
long long rt_dbg_qindex_last_nodes(void);   /* hierarchy nodes tested by the last rt_dbg_qindex_visit call */
#endif /* RT_ABI_DEBUG */

/* Flatten to type-sorted SoA (+ original index) and upload to every device in use.
 * Takes the place of the list->VLA copy at C/raycast.c:87-107. The scene is copied;
 * the caller keeps ownership of its arrays. */
int rt_scene_upload(const rt_scene *scene);

/* The raycast() body (C/raycast.c:109-129): render W x H, rows interleaved in bands of
 * rt band height over the GPUs in use, bands gathered to GPU 0 (NCCL when n_gpus > 1),
 * then copied to the HOST buffer rgb_out (3*W*H bytes, row-major, row 0 = top, no
 * padding: the pixmap of PPMFormat, C/ppm.h:6-12, with size_t instead of int sizes).
 * With n_gpus > 1 and rgb_out in pinned / registered memory there is no gather: every GPU
 * copies its own bands to their places in rgb_out over its own PCIe link (the frame is then
 * not left on GPU 0; env RT_HOST_GATHER=1 keeps the gather). */
int rt_render(int width, int height, uint8_t *rgb_out, rt_stats *stats);

/* Same render, result left in device memory of GPU 0 (inputs resident, no D2H):
 * *dev_rgb_out receives a device pointer owned by the library, valid until the next
 * rt_render* / rt_shutdown. */
int rt_render_device(int width, int height, void **dev_rgb_out, rt_stats *stats);

/* One-process-per-GPU mode (torchrun ranks): render only the bands of `rank` out of
 * `nranks` on this process's device into dev_dst (caller-owned device memory of at least
 * 3*W*rt_rank_num_rows() bytes) or, when dev_dst is NULL, into a library-owned buffer;
 * the rank's rows lie back to back (band order). The caller gathers the rank
 * buffers with its own collective and de-interleaves with rt_rank_rows() / rt_assemble_device().
 * dev_dst must be 4-byte aligned (checked: k_finalize stores 32-bit words). */
int rt_render_rank(int width, int height, int rank, int nranks, void *dev_dst,
                   void **dev_rows_out, size_t *nbytes, rt_stats *stats);

/* De-interleave gathered rank buffers (rank r's rows start at dev_staging +
 * rank_offsets[r], device memory of this process's device) into the row-major frame
 * dev_frame (3*W*H bytes). What rank 0 runs after its gather collective.  When 3*W is a
 * multiple of 16 the copy moves 16 bytes at a time: dev_frame and every dev_staging +
 * rank_offsets[r] must then be 16-byte aligned (checked, RT_ERR_ARG otherwise). */
int rt_assemble_device(int width, int height, int nranks, const void *dev_staging,
                       const int64_t *rank_offsets, void *dev_frame);

/* Rows of `rank` (back to back in dev_rows, as rt_render_rank leaves them) copied straight to their places in the
 * row-major HOST frame (3*W*H bytes, ideally pinned or cudaHostRegister'ed): a band of full-width rows is contiguous
 * on both sides, so this is one strided copy per rank.  With one process per GPU and the host frame in shared
 * memory, every rank moves its own bands over its own PCIe link instead of rank 0 copying the whole frame; the same
 * routine serves rt_render when n_gpus > 1 and rgb_out is pinned.  Synchronous. */
int rt_rank_rows_to_host(int width, int height, int rank, int nranks, const void *dev_rows, uint8_t *host_frame);

/* Row ownership helpers shared by library, host and tests: band b = rows
 * [b*band_h, min((b+1)*band_h, H)) belongs to rank b % nranks. */
int rt_band_height(void);                       /* env RT_BAND_ROWS, default 32 */
int64_t rt_rank_num_rows(int height, int rank, int nranks);
/* Fills rows_out[0..n) with the image row of each row of the rank buffer; returns n. */
int64_t rt_rank_rows(int height, int rank, int nranks, int32_t *rows_out);

/* ---- P3 writer -------------------------------------------
 * The body of ppm_WriteOutP3 (C/ppm.c:168-184) formatted on the GPU: header
 * "P3\n<w> <h> \n<maxColor>\n" (the space before the second newline is the reference's),
 * then every byte of the pixmap as "%d\n", in pixmap order.  Byte-identical to the
 * reference's file for any pixmap; sizes in size_t (the reference's int indexing
 * overflows above ~26.7k^2). */
typedef struct rt_p3_stats {
  uint64_t text_bytes;          /* header + body */
  double ms_h2d;                /* upload of host_rgb (0 when the resident frame is used) */
  double ms_format;             /* k_p3_format, CUDA events on the launching stream */
  double ms_total;              /* whole call, host clock (includes the sink) */
} rt_p3_stats;

/* Upper bound of the text of a W x H frame: header + 4 bytes per value. */
size_t rt_p3_max_bytes(int width, int height);

/* Device to device: dev_rgb (3*W*H bytes, 16-byte aligned) -> dev_text (16-byte aligned,
 * capacity text_cap >= rt_p3_max_bytes); *text_bytes receives the exact length. */
int rt_p3_format_device(const void *dev_rgb, int width, int height, unsigned max_color,
                        void *dev_text, size_t text_cap, size_t *text_bytes, rt_p3_stats *stats);

/* The same formatter for callers that must not allocate device memory.  (No counterpart in the reference: its
 * writer is an fprintf loop on the host.  This entry point exists for the hosts this library gets embedded in.)
 *
 * Who needs it: a plug-in inside an application that has already handed the whole GPU to its own memory pool, where
 * a cudaMalloc from a guest library either fails or eats into a budget the guest was never given; and a per-frame
 * loop that cannot afford cudaMalloc / cudaFree, because both stall every stream of the device while they run.
 * rt_p3_format_device keeps its span table (text length and start offset of every run of 65536 values) in a buffer
 * it cudaMalloc's and grows on demand.  Here that table lives in a fixed array compiled into the library image
 * instead (about 192 KiB; every GPU that loads the library carries it until the process exits), so the call itself
 * makes no cudaMalloc, cudaFree, cudaHostAlloc or other allocating call.  rt_init is not required.
 *
 * dev_rgb (3*W*H bytes) and dev_text (capacity text_cap >= rt_p3_max_bytes(W, H)) are the caller's, both 16-byte
 * aligned, on the device that is current in the calling thread; `stream` is a cudaStream_t of that device passed
 * as a plain pointer (NULL = the default stream).  The kernels are queued on `stream`, then the call waits for
 * that stream and stores the exact text length in *text_bytes (may be NULL).  The text is byte-identical to what
 * rt_p3_format_device writes.
 *
 * Limits and rules:
 *  - 3*W*H <= RT_P3_NOALLOC_MAX_VALUES (2^30 values, enough for a 16384 x 16384 frame); larger frames are refused
 *    with RT_ERR_ARG and nothing is launched -- use rt_p3_format_device for those.
 *  - The first call on each device looks up the table's device address and raises the formatter's shared-memory
 *    limit on that device.  Make that first call during start-up, outside any stream capture.
 *  - Every call on a device uses the same table and leaves its last contents in it (the call re-zeroes the part
 *    it needs first).  Two calls that run on one GPU at the same time -- different streams or different host
 *    threads -- corrupt each other's text; run them one after the other. */
#define RT_P3_NOALLOC_MAX_VALUES (1ull << 30)
int rt_p3_format_device_noalloc(const void *dev_rgb, int width, int height, unsigned max_color,
                                void *dev_text, size_t text_cap, size_t *text_bytes, void *stream);

/* Consumer of the text, called with consecutive pieces in order; non-zero aborts. */
typedef int (*rt_p3_sink)(const void *data, size_t nbytes, void *user);

/* Host-facing writer: formats host_rgb (uploaded first) or, when host_rgb is NULL, the
 * frame the last rt_render / rt_render_device left on GPU 0 (its size must be W x H), and
 * streams the text through pinned double buffers to the sink (e.g. an fwrite). */
int rt_p3_write(const uint8_t *host_rgb, int width, int height, unsigned max_color,
                rt_p3_sink sink, void *user, rt_p3_stats *stats);

#ifdef __cplusplus
}
#endif
#endif /* RT_ABI_H */
