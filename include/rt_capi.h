/* include/rt_capi.h — C ABI of the B200 path tracer (librtb200.so).
 *
 * This is the drop-in boundary for the reference's per-pixel path-tracing path.  Every entry point names
 * the reference interface it replaces (paths relative to the reference repository).  The C++ classes in
 * include/raytracer.h, include/scene.h and include/camera.h keep the reference's signatures and sit on top
 * of exactly these functions.
 *
 * Conventions: every function returns 0 on success and a non-zero rt_status on failure; the message of the
 * last failure on the calling thread is returned by rt_last_error().  Nothing throws, nothing calls exit().
 * A context belongs to one CUDA device and is used from one host thread at a time.  Calls are synchronous
 * (the reference finishes its queue after every launch, src/raytracer.cpp:140,161) unless named *_async.
 * There is no CPU fallback: without a usable CUDA device rt_create fails with RT_ERR_CUDA.
 */
#ifndef RT_CAPI_H
#define RT_CAPI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_ctx rt_ctx;

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = 1,   /* bad argument / scene validation failed */
    RT_ERR_CUDA = 2,      /* CUDA runtime error (message has the CUDA string) */
    RT_ERR_STATE = 3,     /* call order (e.g. render before a scene was uploaded) */
    RT_ERR_UNSUPPORTED = 4
} rt_status;

/* Pointers + counts to the reference's HOST array-of-structs layouts (include/scene.h:32-81; a cl_float3 is
 * 16 bytes): Material 48 B, Sphere 32 B, Plane 48 B, Lens 64 B, vertex 16 B, uv 8 B, index 4 B, Mesh 16 B,
 * Model 12 B.  Replaces the nine cl::Buffer writes + the createScene kernel of
 * src/scene.cpp:71-108 / kernels/raytracer.cl:541-558. */
typedef struct rt_scene_desc {
    const void *materials, *spheres, *planes, *lenses, *vertices, *uvs, *indices, *meshes, *models;
    uint32_t n_materials, n_spheres, n_planes, n_lenses, n_vertices, n_uvs, n_indices, n_meshes, n_models;
} rt_scene_desc;

enum { RT_RNG_TABLE = 0,      /* the reference's hashed 400 000-float table, kernels/raytracer.cl:113-125 */
       RT_RNG_COUNTER = 1,    /* counter-based per-pixel generator — same sampling law, no table */
       RT_RNG_COUNTER_TABLE = 2 }; /* the reference's table entered at a counter-keyed uniform index instead of the
                                 direction hash: the reference's sampling law by construction, two integer hashes
                                 and one 12-byte gather per bounce */
enum { RT_ARITH_STRICT = 0,   /* unfused IEEE binary32 in the oracle's operation order (bit-exact arm) */
       RT_ARITH_FAST = 1 };   /* FMA contraction, approximate rcp/rsqrt/sqrt (throughput arm) */
enum { RT_ACCUM_REFERENCE = 0,/* image = (sqrt(running mean),1), kernels/raytracer.cl:509,524-531 */
       RT_ACCUM_SUM = 1 };    /* accum.rgb += radiance, accum.a += 1 (what the sample split reduces) */
enum { RT_MESH_FIRST_FRONT = 0, /* first front-facing face in order per mesh, kernels/raytracer.cl:291-303 */
       RT_MESH_NEAREST = 1 };   /* nearest front-facing face, ties -> lowest face (BVH-compatible rule) */
enum { RT_ACCEL_AUTO = 0, RT_ACCEL_BRUTE = 1, RT_ACCEL_BVH = 2 };
enum { RT_HEAD_ON = 0,        /* memoise the sample-independent head of every pixel's path */
       RT_HEAD_OFF = 1 };     /* trace every segment of every sample, as the reference does */

typedef struct rt_params {
    uint32_t max_depth;   /* reference: #define DEPTH 30 (kernels/raytracer.cl:5) */
    uint32_t rng_mode;    /* RT_RNG_*   */
    uint32_t seed;        /* key of the counter generator */
    uint32_t arith;       /* RT_ARITH_* */
    uint32_t mesh_mode;   /* RT_MESH_*  */
    uint32_t accel;       /* RT_ACCEL_* */
    uint32_t path_head;   /* RT_HEAD_*  */
} rt_params;

typedef struct rt_stats {
    uint64_t rays_total;      /* path segments (the reference's hitScene calls) since the last rt_render / rt_clear_accum */
    uint64_t rays_last;       /* ... of the last render call */
    uint64_t queries_last;    /* nearest-hit queries the last render call actually executed on the device: equal to
                                 rays_last with RT_HEAD_OFF; with RT_HEAD_ON the segments of a path head are traced
                                 once per pixel (when the head is built) instead of once per sample */
    double   ms_last;         /* device time of the last render call (CUDA events on the context's stream), including
                                 the path-head kernels when the call had to rebuild them */
    double   ms_trace_last;   /* ... of its render kernel (k_trace_*) alone */
    uint32_t samples_done;    /* samples accumulated in the image / accumulation buffer */
    uint32_t kernel_launches; /* kernels launched by this context since creation */
    uint32_t n_primitives;    /* spheres + planes + lenses + triangles on the device */
    uint32_t bvh_nodes;       /* 0 when the brute-force path is active */
} rt_stats;

#define RT_ID_MISS 0xFFFFFFFFu   /* primary id = kind << 28 | index; kind 1 sphere, 2 plane, 3 lens, 4 triangle */
#define RT_RANDOM_TABLE_FLOATS 400000u

const char* rt_last_error(void);
const char* rt_version(void);

/* KernelGL::KernelGL + RayTracer::RayTracer (src/kernelgl.cpp:20-27,58-93, src/raytracer.cpp:24-36): device
 * pick, stream, image and accumulation buffers (the headless replacement of the CL/GL shared texture). */
int rt_create(int width, int height, int device, rt_ctx** out);
int rt_destroy(rt_ctx* ctx);
/* Run the context's work on a caller-owned CUDA stream (cudaStream_t); NULL restores the private stream. */
int rt_set_stream(rt_ctx* ctx, void* cuda_stream);

/* SceneCreator::setupBuffers/createScene (src/scene.cpp:46-87).  Validates ids, converts to device SoA. */
int rt_upload_scene(rt_ctx* ctx, const rt_scene_desc* scene);
/* SceneCreator::loadTextures upload half (src/scene.cpp:164,173): equal-sized RGBA-f32 layers. */
int rt_upload_textures(rt_ctx* ctx, const float* rgba, int width, int height, int layers);
/* The random buffer of src/raytracer.cpp:67-93; rt_make_random_table reproduces that loop for a fixed seed. */
int rt_make_random_table(uint32_t seed, float* out_400000);
int rt_set_random_table(rt_ctx* ctx, const float* table, uint32_t n_floats);
int rt_set_params(rt_ctx* ctx, const rt_params* params);
int rt_get_params(rt_ctx* ctx, rt_params* params);
/* Restrict the render calls to the work-items [x0, x0+w) x [y0, y0+h) of the frame (a sub-NDRange with a global offset in
 * the terms of src/raytracer.cpp:137, which always enqueues the whole frame): pixels outside keep their contents, the
 * per-pixel RNG keys and primary rays are those of the full frame.  w <= 0 or h <= 0 restores the whole frame. */
int rt_set_region(rt_ctx* ctx, int x0, int y0, int w, int h);
/* Pixel-interleaved work split for several GPUs: the frame is cut into groups of four 8 x 4 pixel tiles (128 pixels, the
 * unit a block of the render kernels owns) numbered in launch order; this context renders group g iff g % n_parts == part
 * and leaves all other pixels untouched.  Every pixel keeps its full-frame primary ray, RNG keys and ALL its samples in
 * sample order, so the parts of a frame rendered by different contexts add up to the single-context picture BIT FOR BIT
 * (each pixel has exactly one non-zero contribution) in strict arithmetic; in fast arithmetic the lanes of a warp of the
 * brute-force render kernel share their samples (TraceLaunch::steal), the samples of a pixel are the same but
 * added in an order that depends on its warp, and the pictures agree to rounding (bit for bit again with RTB200_TAIL_STEAL=0).
 * n_parts <= 1 restores the whole frame. */
int rt_set_pixel_split(rt_ctx* ctx, int n_parts, int part);

/* RayTracer::render (src/raytracer.cpp:127-144): sample_counter = 0, one sample, overwrites the image. */
int rt_render(rt_ctx* ctx, const float camera[12]);
/* RayTracer::renderAgain (src/raytracer.cpp:146-165): ++sample_counter, one more sample, running mean.
 * Execution of rt_render / rt_render_again is DEFERRED: the calls validate their arguments and record the sample;
 * consecutive samples with one camera are rendered by a single launch as soon as anything observes the image or the
 * statistics or changes state they depend on (rt_read_image, rt_get_stats, rt_image_device_ptr, uploads, rt_set_params,
 * ...), or after 4096 recorded samples.  The image every observer sees is bit-identical to one launch per call — the
 * reference's own schedule, one launch + queue.finish() per sample (src/raytracer.cpp:137-140,158-161), which the
 * environment variable RTB200_EAGER=1 restores.  rt_stats.rays_last / ms_last then describe the last launch. */
int rt_render_again(rt_ctx* ctx, const float camera[12]);
/* n_samples consecutive samples [first, first+n) in ONE launch.  accum_mode RT_ACCUM_REFERENCE continues the
 * reference's sqrt(mean) chain in the image (first == 0 restarts it; bit-identical to render + renderAgain
 * x (n-1)); RT_ACCUM_SUM adds into the linear accumulation buffer (sample indices are global, so disjoint
 * ranges rendered on different GPUs can be summed). */
int rt_render_samples(rt_ctx* ctx, const float camera[12], uint32_t first, uint32_t n_samples, uint32_t accum_mode);
int rt_clear_accum(rt_ctx* ctx);
/* Forget everything derived from (scene, camera): path heads and certified primary ids are rebuilt by the next
 * render call.  The reference has no such state (it re-traces every segment of every sample). */
int rt_invalidate_caches(rt_ctx* ctx);
/* image = (sqrt(accum.rgb / spp), 1): the tail of kernels/raytracer.cl:509,531 applied to a summed buffer. */
int rt_finalize(rt_ctx* ctx, uint32_t total_spp);

/* Read-backs (the reference never reads back; its image lives in a GL texture, src/raytracer.cpp:167-174).
 * Row 0 is t = 0 (kernels/raytracer.cl:500-501), i.e. the bottom row on screen (shaders/screen.vs:7). */
int rt_read_image(rt_ctx* ctx, float* rgba);                  /* width*height*4 floats */
int rt_read_accum(rt_ctx* ctx, float* rgba);                  /* width*height*4 floats */
int rt_read_primary_ids(rt_ctx* ctx, const float camera[12], uint32_t* ids, float* t);
int rt_get_stats(rt_ctx* ctx, rt_stats* stats);
/* Checkpoint / resume of the progressive state (the reference keeps it only in the GL texture + sample_counter,
 * include/raytracer.h:21, and loses it on exit).  rt_restore_image puts back an image read with rt_read_image together
 * with the number of samples it holds, so that rt_render_again continues the sqrt(mean) chain bit for bit;
 * rt_restore_accum does the same for the linear accumulation buffer. */
int rt_restore_image(rt_ctx* ctx, const float* rgba, uint32_t samples_in_image);
int rt_restore_accum(rt_ctx* ctx, const float* rgba, uint32_t samples_in_buffer);

/* Device-side access for the multi-GPU sample split (one process per GPU; the caller reduces the buffer with
 * NCCL): pointer to the W*H float4 accumulation buffer, or make the context accumulate into a caller buffer. */
int rt_accum_device_ptr(rt_ctx* ctx, void** device_ptr, size_t* bytes);
int rt_image_device_ptr(rt_ctx* ctx, void** device_ptr, size_t* bytes);
int rt_set_accum_device_buffer(rt_ctx* ctx, void* device_ptr, size_t bytes);
/* Make rt_finalize / rt_finalize_peers* store the (sqrt(mean), 1) image at a caller-supplied DEVICE-ACCESSIBLE address
 * instead of the context's image buffer — typically page-locked host memory mapped into the device's address space
 * (cudaHostRegister / cudaHostAlloc mapped): the finalize kernel then writes the picture straight into the caller's host
 * image over PCIe, no staging copy.  With the sample / pixel split every rank maps the SAME host buffer (shared memory) and
 * stores its rows through its own PCIe link, so the read-back of a frame is spread over all links of the box.
 * NULL restores the image buffer.  The reference has no read-back at all (its image is a GL texture). */
int rt_set_finalize_target(rt_ctx* ctx, void* device_accessible_ptr, size_t bytes);

/* Fused reduce + finalize over NVLink peer memory (the exchange step of the sample split; the reference has no
 * multi-device code).  Every rank exports its accumulation buffer (and rank 0 its image) as CUDA IPC handles
 * (64 bytes each); after the handles of all ranks were exchanged (torch.distributed / any transport) and opened,
 * rt_finalize_peers on rank r sums rows [H*r/N, H*(r+1)/N) of ALL ranks' accumulation buffers through peer loads,
 * applies the tail of kernels/raytracer.cl:509,531 (divide by spp, sqrt) and stores the pixels straight into rank
 * 0's image: reduce-scatter + gather + finalize in one kernel per rank, every NVLink used, no staging buffer.
 * The caller synchronises the ranks before (all renders finished) and after (all slices written).
 * rt_set_peers_direct is the same with plain device pointers (several contexts inside one process). */
#define RT_IPC_HANDLE_BYTES 64
int rt_ipc_export(rt_ctx* ctx, void* accum_handle, void* image_handle);
int rt_ipc_open_peers(rt_ctx* ctx, int n_ranks, int my_rank, const void* accum_handles, const void* image_handle_rank0);
int rt_set_peers_direct(rt_ctx* ctx, int n_ranks, int my_rank, void* const* accum_ptrs, void* image_rank0);
int rt_finalize_peers(rt_ctx* ctx, uint32_t total_spp);
int rt_finalize_peers_async(rt_ctx* ctx, uint32_t total_spp);   /* enqueue only: ordered on the context's stream */
/* The same with the two rank synchronisations done ON THE DEVICE: arrival words in peer memory (they live behind every
 * accumulation buffer, so the IPC mapping of rt_ipc_open_peers already covers them) released and awaited by a one-warp
 * kernel before ("every rank's render is complete") and after ("every slice of the image is written") the fused
 * reduce + finalize.  No collective call, no host round trip: the whole exchange is three launches on the context's
 * stream.  All ranks must call it the same number of times.  rt_peer_barrier_status reports whether a wait ever gave up
 * (a rank that never arrived; ~10 s) instead of hanging the device. */
int rt_finalize_peers_barrier_async(rt_ctx* ctx, uint32_t total_spp);
int rt_peer_barrier_status(rt_ctx* ctx, int* timed_out);
int rt_close_peers(rt_ctx* ctx);

/* Probes used by the parity tests: nearest hit of arbitrary rays, and one bounce of the path loop.
 * rays6 = n x (origin.xyz, dir.xyz) on the host.  out_t/out_id may be NULL. */
int rt_trace_rays(rt_ctx* ctx, const float* rays6, uint32_t n, uint32_t* out_id, float* out_t);
/* rt_trace_rays for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  It has NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and results as rt_trace_rays in RT_ARITH_STRICT arithmetic (whatever rt_params.arith is),
 * but the device scratch of the probe kernel — the rays, the ids and the parameters — is a static device array inside the
 * library instead of per-call cudaMalloc: at most RT_NOALLOC_MAX_RAYS rays per call (RT_ERR_INVALID beyond that), 2 MiB held
 * by every GPU that loads the library until the process exits.  The call performs no cudaMalloc / cudaFree and therefore
 * builds nothing: if the scene's intersection structure (LBVH for large scenes) does not exist yet it fails with
 * RT_ERR_STATE — call rt_trace_rays or any render call once beforehand.  All calls share the one array (it keeps the last
 * call's contents; nothing needs to start at zero): calls that overlap on the GPU — several host threads, several contexts on
 * one device — race on it. */
#define RT_NOALLOC_MAX_RAYS 65536u
int rt_trace_rays_noalloc(rt_ctx* ctx, const float* rays6, uint32_t n, uint32_t* out_id, float* out_t);
/* In-place bounce as work-item (gx, gy): rays6 updated, col3 = throughput after the bounce from (1,1,1),
 * ended[i] = 1 when the path terminated (miss or light). */
int rt_bounce_rays(rt_ctx* ctx, float* rays6, uint32_t n, const uint32_t* sample, const uint32_t* bounce,
                   const uint32_t* gx, const uint32_t* gy, float* col3, uint8_t* ended);

/* ---- host-only helpers: the C view of the C++ host classes (no CUDA device needed) -------------------------
 * SceneCreator::loadScene + loadTextures (src/scene.cpp:145-190, 297-403): parse a .scene file in the reference's
 * grammar, import its OBJ models and PNG textures.  asset_root replaces the working directory the reference
 * resolves paths against; texture_dir (may be NULL) is where .mtl texture basenames are looked up. */
typedef struct rt_scene rt_scene;
int rt_scene_load(const char* scene_path, const char* asset_root, const char* texture_dir, rt_scene** out);
int rt_scene_get_desc(const rt_scene* scene, rt_scene_desc* out);
int rt_scene_get_textures(const rt_scene* scene, const float** rgba, int* width, int* height, int* layers);
int rt_scene_upload(const rt_scene* scene, rt_ctx* ctx);       /* SceneCreator::createScene */
int rt_scene_free(rt_scene* scene);
/* Camera(fov, aspect, pos, yaw, pitch).transferData() (src/camera.cpp:46-51, 94-118). */
int rt_camera_data(int fov_degrees, float aspect, const float pos[3], float yaw, float pitch, float out[12]);
/* The Camera controls main.cpp drives from GLFW callbacks (src/camera.cpp:53-92), applied in order to a camera built as
 * above: ops = n_ops x (code, a, b) with code 0..3 move FORWARD / BACK / LEFT / RIGHT by dt = a, 4 rotate(a, b),
 * 5 zoom(a), 6 setFasterSpeed(a != 0), 7 setSlowerSpeed(a != 0), 8 setSize(a); out = transferData() afterwards. */
int rt_camera_run(int fov_degrees, float aspect, const float pos[3], float yaw, float pitch, const float* ops, int n_ops,
                  float out[12]);
/* Uncompressed 24-bit TGA of the (sqrt(mean),1) image, layout of main.cpp:262-289 (18-byte header, BGR, bottom-up),
 * values clamped to [0,1].  The reference reads the GL front buffer; here the float image is converted. */
int rt_write_tga(const char* path, const float* rgba, int width, int height);

/* Counted work of LBVH renders (the roofline numerator of a BVH run is the work the run itself did, not the
 * brute-force operation count of kernels/raytracer.cl:322-360).  While counting is on, fast-arithmetic renders of LBVH
 * scenes go through an instrumented copy of the render kernel — same image, not to be timed — and rt_get_work_counts
 * returns, for the last render call: [0] node visits (two box tests each), leaf tests of [1] spheres, [2] lenses,
 * [3] triangles, [4] plane tests, [5] nearest-hit queries of the render kernel, [6] node visits made by flared rays
 * (|d| != 1; their box test costs about three ordinary ones), [7] flared queries. */
int rt_set_work_counting(rt_ctx* ctx, int on);
int rt_get_work_counts(rt_ctx* ctx, uint64_t out[8]);

/* Measured FP32 issue peak of the device (dependent-free FFMA loop on every SM): G lane-instructions/s. */
int rt_measure_fp32_peak(int device, double* gslots_per_s, double* sm_clock_mhz);

#ifdef __cplusplus
}
#endif
#endif /* RT_CAPI_H */
