/*
 * rt_abi.h -- C ABI of the B200 sphere ray tracer (librt_b200.so).
 *
 * This is the drop-in boundary for ONE path of tristanGIANDO/python_ray_tracer:
 * NumpyRenderer.raytrace_scene and everything it calls.  The reference is pure
 * Python/NumPy and has no FFI of its own; the entry points below are what a ctypes
 * binding inside the reference's Renderer port would bind (INTEGRATION.md shows the
 * stub).  Every entry point names the reference interface it replaces
 * (paths relative to the reference tree).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types;
 *   - every function returns RT_OK (0) or a negative rt_status; rt_last_error()
 *     gives a thread-local message for the last failure;
 *   - "_dev" pointers are CUDA device pointers owned by the caller, "_host"
 *     pointers are ordinary host memory owned by the caller;
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default
 *     stream).  Device entry points only enqueue work on `stream`; they do not
 *     synchronise.  Host entry points are synchronous;
 *   - the library keeps no scene state between calls.  The host entry points
 *     keep a per-thread, per-device staging cache (device scratch buffers, two
 *     streams, a small pinned landing zone for rt_render_png_host) that
 *     rt_release_cache() frees.  Their copies into the caller's buffers run under
 *     the rendering of the next row chunk only when those buffers are pinned
 *     (cudaHostAlloc / cudaHostRegister); into pageable memory every chunk's copy
 *     blocks the call, and the result is the same;
 *
 * Scene layout ("sphere SoA"): a field-major table of doubles,
 *     spheres[f * n_spheres + s],  f in rt_sphere_field,  s = shape index
 * in the order of Scene3D.shapes (domain.py:53-59).  Point lights: point_lights
 * [k * 3 + axis] in the order they appear in Scene3D.lights.
 */
#ifndef RT_ABI_H
#define RT_ABI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 3

/* The reference's miss sentinel (base.py:12, shader.py:10). */
#define RT_FARAWAY 1.0e39

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_BAD_ARG = -1,     /* NULL pointer, negative size, unknown enum value          */
    RT_ERR_CUDA = -2,        /* a CUDA runtime call failed; see rt_last_error()          */
    RT_ERR_NO_DEVICE = -3,   /* no CUDA device / device index out of range               */
    RT_ERR_UNSUPPORTED = -4  /* valid request this build cannot serve                    */
} rt_status;

/* Rows of the sphere SoA.  Geometry: NumpySphere(center, radius, shader)
 * (shape.py:17-26).  Material: NumpyShader(...) (shader.py:36-54); reflection_gain is
 * stored by the reference and never read, so it is not carried. */
typedef enum rt_sphere_field {
    RT_F_CX = 0,
    RT_F_CY = 1,
    RT_F_CZ = 2,
    RT_F_RADIUS = 3,
    RT_F_DIFFUSE_GAIN = 4,
    RT_F_SPECULAR_GAIN = 5,
    RT_F_SPECULAR_ROUGHNESS = 6,
    RT_F_IRIDESCENCE_GAIN = 7,
    RT_F_COLOR_R = 8,  /* Texture.color (shader.py:13-19); ignored by the checker      */
    RT_F_COLOR_G = 9,
    RT_F_COLOR_B = 10,
    RT_F_TEXTURE = 11, /* 0.0 = Texture (constant), 1.0 = TextureChecker (shader.py:22-32) */
    RT_SPHERE_FIELDS = 12
} rt_sphere_field;

#define RT_MAX_DOME_LIGHTS 8

/* Everything of Scene3D that is not a shape (domain.py:26-40, 53-59) plus the
 * trace options.  Passed by pointer, read on the host. */
typedef struct rt_env {
    double eye[3];               /* scene.camera.position: target of "direction to ray
                                    origin" at every bounce (shader.py:76)               */
    int32_t n_point_lights;      /* rows in the point-light table (>= 1)                 */
    int32_t n_dome_lights;       /* <= RT_MAX_DOME_LIGHTS; the host folds any excess
                                    into the last slot                                    */
    double dome_intensity[RT_MAX_DOME_LIGHTS]; /* DomeLight.intensity in list order
                                    (shader.py:239-242)                                   */
    double dome_color[3];        /* colour of the LAST DomeLight, (1,1,1) if none
                                    (shader.py:237-241)                                   */
    int32_t max_depth;           /* number of ray generations traced; primary rays are
                                    generation 0 and rays of generation >= max_depth
                                    return black.  <= 0: unlimited, like the reference
                                    (shader.py:143-161 has no cap)                        */
    int32_t mode;                /* rt_mode                                               */
    int32_t flags;               /* rt_flags                                              */
    int32_t reserved;
    double cutoff;               /* RT_MODE_FAST only: stop a bounce chain once its
                                    remaining contribution is provably below this
                                    (linear units).  0 = never cut.                       */
    const void* accel_dev;       /* optional: device copy of the blob rt_accel_build_host()
                                    made for THIS sphere table (a bounding-volume tree);
                                    rays then skip the sphere tests that provably miss.
                                    NULL = test every sphere, like base.py:97 /
                                    shader.py:126.  Results are identical either way.     */
    int64_t accel_bytes;         /* size of that blob                                     */
    const double* spheres_host;  /* optional: the caller's HOST copy of the sphere table the
                                    call's spheres_dev holds (same layout, same values).  The
                                    device entry points never read device memory on the host, so
                                    without it every block derives the per-scene constants itself;
                                    with it they are formed once, here, and scenes of at most
                                    RT_SMALL_SCENE_SPHERES shapes travel as a kernel parameter
                                    (no staging, sphere constants as instruction operands).
                                    Same results either way.  NULL = not available.          */
    const double* point_lights_host; /* optional: host copy of the point-light table, likewise */
} rt_env;

#define RT_SMALL_SCENE_SPHERES 4

typedef enum rt_mode {
    RT_MODE_EXACT = 0, /* all-FP64, reference operation order, no FMA contraction        */
    RT_MODE_FAST = 1   /* FP64 geometry, restructured arithmetic, throughput cutoff      */
} rt_mode;

typedef enum rt_flags {
    RT_FLAG_NONE = 0,
    RT_FLAG_ALL_POINT_LIGHTS = 1, /* extension (main.py:45 TODO): sum diffuse/specular and
                                    shadow over every point light.  Off = reference
                                    semantics: only lights[0] shades (shader.py:75).     */
    RT_FLAG_ACCEL_FAST_BUILD = 8, /* rt_accel_build_host: median splits only (about 3x faster to
                                    build, ~7 % slower to trace): what the host-buffer entry
                                    points use for the tree they build per call           */
    RT_FLAG_NO_ACCEL = 4,         /* host-buffer entry points only: never build the tree
                                    themselves (see RT_ACCEL_AUTO_SPHERES)               */
    RT_FLAG_WIDE_STORES = 16,     /* rt_render_rows_dev / rt_render_strips_dev: warps render 32x1
                                    pixel patches instead of 8x4, so each of their stores is a
                                    128-byte line (96 bytes with RT_FLAG_RGB8) instead of a
                                    32-byte sector.  For frames written into peer memory
                                    (RT_FLAG_IMAGE_ORDER): NVLink carries the wide stores at
                                    about twice the rate; tracing itself is ~4 % slower at 8K.
                                    Same pixels either way.                                */
    RT_FLAG_EVEN_STORES = 64,     /* with RT_FLAG_WIDE_STORES: patches are taken two from the
                                    bottom, one from the top of the rows of the call instead of
                                    bottom-up.  The reference camera (base.py:130-141) sees sky
                                    in the upper rows, which trace several times faster than
                                    the rest; taken last they arrive as a burst of stores
                                    that one NVLink port cannot take from seven writers.    */
    RT_FLAG_RGB8 = 32,            /* rt_render_rows_dev / rt_render_strips_dev: rgba8_dev holds
                                    3 bytes per pixel (R,G,B -- the array save_image hands to
                                    PIL, base.py:143-151) instead of 4.  With RT_FLAG_WIDE_STORES
                                    and cam->width and the pointer multiples of 4 the warp
                                    packs a patch row into 32-bit words; byte stores
                                    otherwise.                                              */
    RT_FLAG_GENERIC_KERNEL = 128, /* never use the kernel-parameter path of small scenes (see
                                    rt_env.spheres_host): for tests and measurements that compare
                                    the two paths                                             */
    RT_FLAG_IMAGE_ORDER = 2       /* rt_render_rows_dev / rt_render_strips_dev: rgb_dev and
                                    rgba8_dev are WHOLE frames (height x width) and every
                                    pixel is stored at row*width + col, whatever rows the
                                    call renders.  The buffers may be peer memory of
                                    another GPU (rt_peer_buffer_open): the shards of a
                                    frame are then written straight into rank 0's
                                    framebuffer over NVLink while the kernel runs, and no
                                    gather follows.                                       */
} rt_flags;

/* Camera(position, width, height) (domain.py:14-23). */
typedef struct rt_camera {
    double position[3];
    int32_t width;
    int32_t height;
} rt_camera;

typedef enum rt_dtype { RT_DTYPE_F32 = 0, RT_DTYPE_F64 = 1 } rt_dtype;

/* Work counters a launch adds into (device memory, 4 x uint64, caller zeroes):
 * [0] rays traced (all generations), [1] hits shaded, [2] deepest generation reached + 1,
 * [3] sphere tests evaluated. */
#define RT_STATS_WORDS 4

int rt_abi_version(void);
const char* rt_last_error(void);
int rt_device_count(int* count_out);
int rt_release_cache(void);
/* What a kernel of an earlier launch on `device` reported, if anything: device entry points only
 * enqueue work, so a launch that finds it cannot render what it was given (an rt_env.accel_dev that
 * is not a blob for its sphere table; a tree deeper than the traversal stack) sets a sticky word
 * instead.  Returns RT_OK, or RT_ERR_UNSUPPORTED with the reason in rt_last_error(), and clears
 * it.  Every device entry point makes the same check first, so such a launch never goes unnoticed;
 * call this after synchronising to learn about the launch just finished. */
int rt_device_status(int device);
/* 1 if a camera launch (rt_render_rows_dev / rt_render_strips_dev) with this env and sphere count reads the whole
 * scene from its kernel parameters (rt_env.spheres_host: at most RT_SMALL_SCENE_SPHERES shapes, fast policy, no tree,
 * RT_FLAG_GENERIC_KERNEL clear): spheres_dev and point_lights_dev are then never dereferenced -- any non-NULL
 * pointer will do -- so a caller whose scene changes every frame (an animation) need not upload it, and its host
 * thread need not wait on the stream for the copy.  0 otherwise.  Nothing in the reference corresponds to it. */
int rt_scene_travels_as_parameter(const rt_env* env, int32_t n_spheres);

/*
 * rt_measure_fma_peak -- achieved FLOP/s (2 per FMA) of a register-resident chain of
 * independent FMAs of `dtype` (rt_dtype) on every SM of `device`: the ALU roofline
 * denominator of the trace kernel, measured on the GPU it runs on.  Nothing in the
 * reference corresponds to it (the reference reports one wall-clock print,
 * main.py:55-57); it exists for the measurement contract only.  Synchronous.
 */
int rt_measure_fma_peak(int device, int32_t dtype, double* flops_out);

/*
 * rt_accel_build_host -- bounding-volume tree over the spheres, built on the host into one
 * position-independent blob the caller uploads and names in rt_env.accel_dev.  Nothing in
 * the reference corresponds to it (it intersects every ray with every shape); it exists so
 * that scenes of hundreds of spheres (BASELINE config 3) do not pay for tests that cannot
 * hit, and so that scenes too large for shared memory can be rendered at all.
 *   rt_accel_bytes(n)  capacity to allocate for n spheres (an upper bound)
 *   env                dome lights and mode of the render calls the blob will serve (it
 *                      carries the material bound of the fast policy's chain cut-off);
 *                      RT_FLAG_ACCEL_FAST_BUILD in env->flags: median splits instead of
 *                      surface-area splits
 * Returns RT_ERR_UNSUPPORTED for more than 2^26 spheres or a coordinate / radius beyond
 * 1e15 (the tree's boxes are FP32), RT_ERR_BAD_ARG if `capacity` is too small.  The tree
 * only skips sphere tests that provably miss: results with and without it are identical.
 */
#define RT_ACCEL_AUTO_SPHERES 32 /* measured break-even on B200 (profiles/r01/accel_sweep.txt):
                                   the host-buffer entry points build and upload the tree
                                   themselves for scenes at least this large when
                                   env->accel_dev is NULL and RT_FLAG_NO_ACCEL is not set */
int64_t rt_accel_bytes(int32_t n_spheres);
int rt_accel_build_host(const double* spheres_host, int32_t n_spheres, const rt_env* env,
                        void* blob_out, int64_t capacity, int64_t* bytes_out);

/*
 * rt_render_rows_dev -- camera rays for image rows [row0, row1) of cam.
 * Replaces NumpyRenderer.get_ray_directions (base.py:123-141) followed by
 * NumpyRenderer.raytrace_scene (base.py:91-121) on those pixels, including every
 * NumpyShader.create it triggers (shader.py:63-112).
 *   spheres_dev       sphere SoA, RT_SPHERE_FIELDS x n_spheres doubles
 *   point_lights_dev  env->n_point_lights x 3 doubles
 *   rgb_dev           planar linear RGB of the band: channel c of pixel p (row-major
 *                     inside the band, p = (row-row0)*width + col) is at
 *                     rgb_dev[c*plane_stride + p], element type rgb_dtype.  May be NULL.
 *   rgba8_dev         optional: 4 bytes per pixel, uint8(255*clip(v,0,1)) truncating
 *                     (base.py:143-151) in R,G,B and 255 in A -- or 3 bytes per pixel
 *                     (R,G,B) with RT_FLAG_RGB8.  May be NULL.
 *   stats_dev         optional RT_STATS_WORDS x uint64 counters, or NULL.
 */
int rt_render_rows_dev(int device, const double* spheres_dev, int32_t n_spheres,
                       const double* point_lights_dev, const rt_env* env,
                       const rt_camera* cam, int32_t row0, int32_t row1, void* rgb_dev,
                       int32_t rgb_dtype, int64_t plane_stride, uint8_t* rgba8_dev,
                       uint64_t* stats_dev, void* stream);

/*
 * rt_render_strips_dev -- the same, for an interleaved shard of the frame: the image is
 * cut into strips of strip_rows rows (strip k = rows [k*strip_rows, (k+1)*strip_rows),
 * the last one possibly shorter) and this call renders strips strip_first,
 * strip_first + strip_step, ...  The multi-GPU row-band partition (rank r of G renders
 * strips r, r+G, ...: every shard is a stratified sample of the frame, so shards cost
 * the same without a cost model).  Pixels are independent in the reference
 * (base.py:102-119 masks per pixel), so any partition is exact.
 *   rgb_dev / rgba8_dev  the shard's rows packed in strip order: local row r is image
 *                        row (strip_first + (r / strip_rows) * strip_step) * strip_rows
 *                        + r % strip_rows;  rt_strip_shard_rows() rows in total.
 */
int64_t rt_strip_shard_rows(int32_t height, int32_t strip_rows, int32_t strip_first, int32_t strip_step);
int rt_render_strips_dev(int device, const double* spheres_dev, int32_t n_spheres,
                         const double* point_lights_dev, const rt_env* env,
                         const rt_camera* cam, int32_t strip_rows, int32_t strip_first,
                         int32_t strip_step, void* rgb_dev, int32_t rgb_dtype,
                         int64_t plane_stride, uint8_t* rgba8_dev, uint64_t* stats_dev,
                         void* stream);

/*
 * Peer framebuffers (one process per GPU): the owner creates a device buffer and gets a
 * 64-byte handle to send to the other processes of the node (any byte transport, e.g.
 * torch.distributed.broadcast_object_list); they open it and pass the pointer as
 * rgb_dev / rgba8_dev of a render call with RT_FLAG_IMAGE_ORDER.  Nothing in the
 * reference corresponds to this (it is single-process); it is the
 * exchange step of the row-band sharding of BASELINE config 4.
 *   create   cudaMalloc + zero + export; the owner frees with rt_peer_buffer_destroy
 *   open     map the owner's buffer into this process (peer access over NVLink/PCIe);
 *            a process cannot open its own handle
 */
#define RT_IPC_HANDLE_BYTES 64
int rt_peer_buffer_create(int device, uint64_t bytes, void** ptr_out, uint8_t* handle_out);
int rt_peer_buffer_open(int device, const uint8_t* handle_in, void** ptr_out);
int rt_peer_buffer_close(int device, void* ptr);
int rt_peer_buffer_destroy(int device, void* ptr);

/*
 * rt_trace_rays_dev -- an arbitrary batch of n rays.
 * Replaces NumpyRenderer.raytrace_scene(ray_origin, normalized_ray_direction, scene)
 * (base.py:91-121) for callers that bring their own rays.
 *   origins_dev     3 x n doubles, planar (x plane, y plane, z plane), or NULL to
 *                   broadcast origin0 to every ray (the primary-ray case, where the
 *                   reference passes scalar origins)
 *   directions_dev  3 x n doubles, planar
 *   rgb_dev         as above with p = ray index
 */
int rt_trace_rays_dev(int device, const double* spheres_dev, int32_t n_spheres,
                      const double* point_lights_dev, const rt_env* env,
                      const double* origins_dev, const double* origin0,
                      const double* directions_dev, int64_t n_rays, void* rgb_dev,
                      int32_t rgb_dtype, int64_t plane_stride, uint64_t* stats_dev,
                      void* stream);

/*
 * rt_ray_directions_dev -- materialise the unit primary directions of rows
 * [row0,row1) as 3 planes of doubles.  Replaces NumpyRenderer.get_ray_directions
 * (base.py:123-141) for callers that want the array itself.
 */
int rt_ray_directions_dev(int device, const rt_camera* cam, int32_t row0, int32_t row1,
                          double* directions_dev, int64_t plane_stride, void* stream);

/*
 * rt_sphere_distance_dev -- distance from each ray's origin to one sphere along the
 * ray, RT_FARAWAY on a miss.  Replaces NumpySphere.intersect (shape.py:28-51).
 *   origins_dev / origin0 / directions_dev  as in rt_trace_rays_dev
 *   center        3 doubles (host)
 *   mode          rt_mode: RT_MODE_EXACT evaluates the reference's expression
 *   distance_dev  n doubles
 */
int rt_sphere_distance_dev(int device, const double* origins_dev, const double* origin0,
                           const double* directions_dev, int64_t n_rays, const double* center,
                           double radius, int32_t mode, double* distance_dev, void* stream);

/*
 * rt_quantize_dev -- uint8(255*clip(v,0,1)), truncating, planar -> interleaved.
 * Replaces the arithmetic of NumpyRenderer.save_image (base.py:143-151); the PNG
 * container is written by the host.
 *   channels_out  3 (RGB) or 4 (RGBA, A = 255)
 */
int rt_quantize_dev(int device, const void* rgb_dev, int32_t rgb_dtype, int64_t plane_stride,
                    int64_t n_pixels, uint8_t* out_dev, int32_t channels_out, void* stream);

/*
 * Host-buffer entry points: the same operations with ordinary host memory on both
 * sides; uploads, launches, downloads and synchronises internally.
 *   rgb_host    planar, 3 x n doubles or floats (plane stride = n), may be NULL
 *   rgb8_host   interleaved RGB bytes (3 per pixel), uint8(255*clip(v,0,1)) truncating of
 *               the FP64 pixel (base.py:143-151), quantised by the trace kernel; may be
 *               NULL.  With rgb_host NULL no linear planes are written at all.
 *   stats_host  RT_STATS_WORDS x uint64, may be NULL
 * RT_FLAG_IMAGE_ORDER in env->flags is ignored here (the outputs are the band's own buffers).
 */
int rt_render_frame_host(int device, const double* spheres_host, int32_t n_spheres,
                         const double* point_lights_host, const rt_env* env,
                         const rt_camera* cam, int32_t row0, int32_t row1, void* rgb_host,
                         int32_t rgb_dtype, uint8_t* rgb8_host, uint64_t* stats_host);

int rt_trace_rays_host(int device, const double* spheres_host, int32_t n_spheres,
                       const double* point_lights_host, const rt_env* env,
                       const double* origins_host, const double* origin0,
                       const double* directions_host, int64_t n_rays, void* rgb_host,
                       int32_t rgb_dtype, uint64_t* stats_host);

/*
 * rt_png_deflate_rows -- the compressed stream of the PNG save_image writes (base.py:143-151:
 * the reference hands its uint8 array to Pillow, i.e. zlib level 6, 0.45 s per 1080p frame).
 * Host code, no GPU involved; for animation batches, where the encoder bounds the frame rate.
 * Compresses `rows` rows of 8-bit pixels (pixel_bytes = 3: R,G,B; 4: R,G,B,A with A dropped;
 * rows row_stride bytes apart) into RAW deflate of the PNG-filtered rows (filter 1, Sub; runs at
 * distance 1 and 3; one dynamic-Huffman block, or stored blocks when smaller).  A band that is
 * not the `last` of its image ends byte-aligned with BFINAL clear, so bands compressed
 * concurrently concatenate into one stream: the IDAT payload is 0x78 0x01, the bands in order,
 * and the big-endian Adler-32 of all filtered bytes (combine the bands' `adler_out` with their
 * `raw_bytes_out`, as zlib's adler32_combine does).  Decoded pixels equal the input.
 *   rt_png_deflate_bound(width, rows)  capacity `out` must have
 */
int64_t rt_png_deflate_bound(int32_t width, int32_t rows);
int rt_png_deflate_rows(const uint8_t* pixels, int32_t width, int32_t rows, int32_t pixel_bytes,
                        int64_t row_stride, int32_t last, uint8_t* out, int64_t capacity,
                        int64_t* bytes_out, uint32_t* adler_out, int64_t* raw_bytes_out);

/*
 * The same PNG, encoded ON THE DEVICE (csrc/rt_png_dev.cu): where the 8-bit image has to travel -- over PCIe to
 * the host, over NVLink into rank 0's framebuffer -- it travels as the ~10x smaller file.  Sub filter, matches
 * at distance 3, one fixed Huffman code; the image is cut into bands of rt_png_rows_per_band() rows, each of
 * which becomes one IDAT chunk (with its own CRC-32), so bands are encoded independently -- by different thread
 * blocks, and by different GPUs when a frame is sharded.  Decoded pixels equal the input.
 *   rt_png_encode_bands_dev  encodes the bands of the `n_rows` rows at pixels_dev (8-bit, pixel_bytes = 3: R,G,B
 *                            or 4: R,G,B,A with A dropped) into `scratch_dev` (rt_png_scratch_bytes(); device
 *                            memory of `device` or of a peer: one slot per band of the IMAGE).  The rows are image
 *                            rows row0.. (strip_step = 1) or the interleaved shard of rt_render_strips_dev (strips
 *                            of strip_rows rows, every strip_step-th, starting at image row row0): strip_rows and
 *                            row0 must then be multiples of rt_png_rows_per_band().
 *   rt_png_pack_dev          once every band of the image is in scratch_dev: writes the file (signature, IHDR, the
 *                            IDAT chunks in order, a tail chunk with the Adler-32, IEND) to png_out and its size to
 *                            *png_bytes_out (0 if capacity was too small; rt_png_file_bound() always suffices).
 *                            Both may be device memory or device-mapped (pinned) host memory.
 *   rt_png_encode_file_dev   a WHOLE image (height rows at pixels_dev) to the finished file in ONE launch, for the
 *                            common case that one GPU holds the frame: the same bands, but every band's block publishes
 *                            its chunk size as soon as it is known, adds up the sizes of the bands before it (status
 *                            words in the scratch area) and stores its chunk at its place in the file, the last block to finish
 *                            writes the tail -- no slots, no pack launch, and a file in pinned host memory crosses
 *                            PCIe while later bands are still being encoded.  Byte for byte the file of the two calls
 *                            above.  scratch_dev: rt_png_scratch_bytes() of device memory (its control words are
 *                            cleared by a cudaMemsetAsync on `stream`).
 *                            Measured on B200, file into pinned host memory: 549 us against 589 for an 8K frame
 *                            (4 320 bands), 76 against 71 at 1080p (540 bands: the memset and the per-block epilogue
 *                            cost more than the very light pack kernel) -- worth it from RT_PNG_ONE_LAUNCH_BANDS bands on.
 * All only enqueue work on `stream`.  rt_render_png_host: rt_render_frame_host + whichever of the two is faster for
 * the frame's size; the file is all that crosses PCIe.
 */
#define RT_PNG_ONE_LAUNCH_BANDS 2048
int32_t rt_png_rows_per_band(int32_t width, int32_t height);
int64_t rt_png_scratch_bytes(int32_t width, int32_t height);
int64_t rt_png_file_bound(int32_t width, int32_t height);
int rt_png_encode_bands_dev(int device, const uint8_t* pixels_dev, int32_t pixel_bytes, int64_t row_stride,
                            int32_t n_rows, int32_t width, int32_t height, int32_t row0, int32_t strip_rows,
                            int32_t strip_step, void* scratch_dev, int64_t scratch_bytes, void* stream);
int rt_png_pack_dev(int device, int32_t width, int32_t height, const void* scratch_dev, uint8_t* png_out,
                    int64_t capacity, uint64_t* png_bytes_out, void* stream);
int rt_png_encode_file_dev(int device, const uint8_t* pixels_dev, int32_t pixel_bytes, int64_t row_stride, int32_t width,
                           int32_t height, void* scratch_dev, int64_t scratch_bytes, uint8_t* png_out, int64_t capacity,
                           uint64_t* png_bytes_out, void* stream);
int rt_render_png_host(int device, const double* spheres_host, int32_t n_spheres,
                       const double* point_lights_host, const rt_env* env, const rt_camera* cam,
                       uint8_t* png_host, int64_t capacity, int64_t* png_bytes_out);

/*
 * rt_png_encode_static_dev -- rt_png_encode_bands_dev + rt_png_pack_dev for one whole image, ALLOCATING NOTHING:
 * the band scratch is a file-scope device array of RT_PNG_STATIC_SCRATCH_BYTES inside the library instead of
 * memory the caller made.  It has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  No cudaMalloc, cudaFree or allocating library routine is called; the caller still passes the
 * input and both outputs, and the file is byte for byte the one the two calls above produce.
 *   pixels_dev            height rows of 8-bit pixels (pixel_bytes = 3 or 4, rows row_stride bytes apart)
 *   png_out, capacity     the file (device memory or device-mapped pinned host memory; rt_png_file_bound() suffices)
 *   png_bytes_out         its size, written by the pack kernel (0 if capacity was too small)
 * Limit: images with rt_png_scratch_bytes(width, height) <= RT_PNG_STATIC_SCRATCH_BYTES (8 MiB: every frame up to
 * 1920 x 1080, e.g. 2048 x 1024); larger ones return RT_ERR_UNSUPPORTED with the two sizes in rt_last_error().
 * Every GPU that loads the library holds the array until the process exits.  All calls on a device SHARE it, and it
 * keeps the last call's contents (nothing in it has to start at zero: every byte that is read was written by the
 * same call): two calls whose kernels overlap on the GPU -- different streams, or replays of two captured graphs
 * at once -- race; order them on one stream or with events.  The array's device address is looked up
 * (cudaGetSymbolAddress) on the first call per device, which like any first launch also loads the module: make
 * that call before capturing.  Only enqueues work on `stream`.
 */
#define RT_PNG_STATIC_SCRATCH_BYTES (8u << 20)
int rt_png_encode_static_dev(int device, const uint8_t* pixels_dev, int32_t pixel_bytes, int64_t row_stride,
                             int32_t width, int32_t height, uint8_t* png_out, int64_t capacity,
                             uint64_t* png_bytes_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RT_ABI_H */
