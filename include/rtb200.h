/* rtb200.h -- C ABI of the B200-native replacement for the reference's render path.
 *
 * The reference (antoinedesbois/Ray-Tracer-Rust) has no FFI or plugin interface: the hot
 * path is reached by the in-binary call `render(scene)` (src/main.rs:361 -> src/main.rs:242).
 * The entry points below are what a Rust `extern "C"` block in that crate would bind to
 * replace that call and the constructors feeding it; each one names the reference
 * interface it stands in for.  Plain pointers and sizes only; no CUDA or torch types.
 *
 * Conventions: every function returning int returns RTB_OK (0) or a negative rtb_status;
 * rtb_last_error() gives the message for the calling thread.  All host buffers are owned
 * by the caller and only borrowed for the duration of the call.  A handle is bound to
 * one CUDA device and is not re-entrant; distinct handles are independent.
 * There is no CPU fallback: without a usable CUDA device every compute call fails.
 */
#ifndef RTB200_H
#define RTB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    RTB_OK = 0,
    RTB_ERR_ARG = -1,      /* bad argument */
    RTB_ERR_CUDA = -2,     /* CUDA runtime error (message in rtb_last_error) */
    RTB_ERR_STATE = -3,    /* call order: e.g. render before rtb_scene_set_samples */
    RTB_ERR_IO = -4,       /* file could not be read / written */
    RTB_ERR_LIMIT = -5     /* scene exceeds a built-in limit (e.g. BVH depth > stack) */
} rtb_status;

/* Opaque device-resident scene: triangles (SoA float4), BVH, light, sample table.
 * Replaces `Scene` (src/tracer/utils/scene.rs:6-12) + `BoundingVolumeHierarchy`
 * (src/tracer/utils/bounding_volume_hierarchy.rs:145-148). */
typedef struct rtb_scene rtb_scene;

/* What `main` hands to Scene{..}/BoundingVolumeHierarchy::new (src/main.rs:324-358). */
typedef struct {
    uint32_t n_tris;          /* primitives in the reference's order: OBJ faces in argv order, ground last */
    const float* verts;       /* n_tris * 9: v0.xyz v1.xyz v2.xyz (Triangle::new args, triangle.rs:22) */
    const float* colors;      /* n_tris * 3: linear RGB (Color, color.rs:4-8); NULL: every primitive white, as import_obj makes them
                                 (src/main.rs:143) -- saves the 12 bytes per triangle of host array and upload for large meshes */
    const int32_t* tie_rank;  /* optional n_tris: order used ONLY to break exact-equal-distance ties between two
                                 primitives (larger wins; the reference's rule is "later leaf in its tree",
                                 bounding_volume_hierarchy.rs:123-130).  NULL: computed on the device by
                                 replaying the reference's pairing (n_tris <= RTB_TIE_RANK_MAX_TRIS), else index. */
    float light[9];           /* area-light triangle; only primitives[0] is sampled (light.rs:10-14) */
    int32_t device;           /* CUDA device ordinal */
} rtb_scene_desc;

#define RTB_TIE_RANK_MAX_TRIS 65536u

/* Tunables of a scene (all optional: pass NULL to rtb_scene_create_ex, or leave a field 0, for the default).  They were
 * environment variables in round 1; the variables (named beside each field) are still read and OVERRIDE the struct, for
 * experiments only. */
typedef struct {
    uint32_t struct_size;         /* sizeof(rtb_scene_options), so that the struct can grow */
    int32_t shadow_mode;          /* 0 auto: light grids when they apply; 1 candidate lists / BVH walks only; 2 = 0   RTB200_SHADOW=lists|grid */
    int32_t primary_mode;         /* 0 auto: camera grid when it applies; 1 BVH traversal only                       RTB200_PRIMARY=bvh */
    int32_t term_mode;            /* 0 auto (single-chunk renders); 1 always; -1 never: the term wavefront             RTB200_TERMS=1|0 */
    uint64_t chunk_rays;          /* rays per wavefront chunk, 0 = 16 Mi                                               RTB200_CHUNK_RAYS */
    uint64_t term_bytes;          /* memory per buffer set of the term wavefront, 0 = 1 GiB, UINT64_MAX = none         RTB200_TERM_MB */
    uint64_t grid_entry_cap;      /* light / camera grid entries above which the list / walk path is taken, 0 = 512 Mi */
    float grid_cell_factor;       /* light-grid cell size / log-mean primitive extent, 0 = 0.25                        RTB200_GRID_CELL */
    float cam_cell_px;            /* camera-grid cell size in pixels, 0 = 2                                            RTB200_CAM_CELL */
    int32_t cdiv_off;             /* 1: never the two-operation division by denom                                      RTB200_CDIV=0 */
    int32_t trace_chunks;         /* 1: per-chunk kernel times of profiled renders on stderr                           RTB200_TRACE_CHUNKS */
} rtb_scene_options;

/* Per-frame parameters: the compile-time constants of src/main.rs:38-40 and the literals of
 * src/main.rs:349-356 made explicit, plus the multi-GPU band ownership. */
typedef struct {
    uint32_t width, height;       /* Scene.width/height (src/main.rs:350-351) */
    uint32_t nb_ray;              /* NB_RAY, rays per pixel (src/main.rs:38) */
    uint32_t nb_light_sample;     /* NB_LIGHT_SAMPLE (src/main.rs:39) */
    float eye[3], look_at[3], up[3], distance;   /* Camera::new args (camera.rs:17-20) */
    /* Image rows are grouped in bands of band_rows rows; band b belongs to rank (b % world).
     * world == 1 (or 0) renders the whole frame.  Owned bands are stored compactly, in
     * increasing band order, in the output buffer (see rtb_owned_rows). */
    uint32_t band_rows, rank, world;
    uint32_t flags;               /* RTB_FLAG_* */
} rtb_render_params;

#define RTB_FLAG_COUNTERS 1u      /* also count AABB / triangle tests (slower instrumented kernel) */
#define RTB_FLAG_PROFILE 2u       /* per-kernel times in rtb_stats: the chunks of the frame run one after the other on one stream
                                     instead of overlapping on two (a little slower) */
#define RTB_FLAG_FRAME_ROWS 4u    /* the output buffer is the FULL frame (height * width * 3 bytes) and every owned row is written at its
                                     own position in it, not compactly.  rtb_render_device: with d_rgb8 = rank 0's frame mapped into this
                                     process (rtb_peer_open) the shading kernel stores its pixels straight into the gathered frame over
                                     NVLink: render and gather are one kernel, no staging buffer, no placement pass.  rtb_render: the HOST
                                     buffer is the full frame (e.g. one pinned allocation shared by all ranks) and each chunk's rows are
                                     copied device-to-host straight to their rows of it (strided copies, one per chunk). */
#define RTB_FLAG_GATHER_PEER 8u   /* rtb_render_multi only: devices 1.. store their rows into device 0's frame over NVLink peer access
                                     (RTB_FLAG_FRAME_ROWS) and ONE device-to-host copy follows; default: every device copies its own
                                     rows straight into the caller's host frame over its own PCIe link. */

typedef struct {
    uint64_t primary_rays, primary_hits, shadow_rays, shadow_occluded;
    uint64_t box_tests, tri_tests;   /* only with RTB_FLAG_COUNTERS, else 0 */
    float kernel_ms;                 /* device time of the trace kernels of this call (CUDA events) */
    uint32_t launches;               /* kernels launched by this call */
    float primary_ms, shadow_ms, shade_ms;   /* k_primary / the shadow kernel(s) (k_shadow_grid on the light-grid path; k_shadow_plan +
                                                k_shadow_lists + k_shadow_queue on the list / walk path) / k_shade (+ k_shade_terms);
                                                only with RTB_FLAG_PROFILE, else 0 */
    uint32_t queued_hits;            /* hit points whose shadow rays needed the BVH (shadow queue length) */
    uint32_t queued_groups;          /* (hit point, light-sample group) pairs left after the per-group bundle test */
    uint32_t queued_walks;           /* ... of which need a per-ray BVH walk (not just a short candidate list) */
    uint32_t grid_cells, grid_entries, grid_big;   /* light grids in use for this call (0 = list / walk path): cells, entries,
                                                      primitives outside the grids */
    float grid_build_ms;             /* device time of the last grid build (once per sample table, not per frame) */
    uint32_t grid_rebuilds;          /* light- or camera-grid builds this call enqueued (0: both cached) ... */
    uint32_t grid_redo;              /* ... of which failed their device-side capacity check: the frame was rendered a second time
                                        after a synchronous rebuild (see rtb_render) */
} rtb_stats;

typedef struct {
    uint32_t n_tris, n_nodes, max_depth, ploc_iterations;
    float sah_cost;                  /* surface-area-heuristic cost of the built tree (diagnostic) */
    float build_ms;
    uint32_t tie_mode;               /* how exact equal-distance ties are broken: RTB_TIE_REFERENCE (the reference's leaf order,
                                        replayed on the device), RTB_TIE_CALLER (rtb_scene_desc.tie_rank) or RTB_TIE_INDEX (larger
                                        primitive index: scenes above RTB_TIE_RANK_MAX_TRIS primitives -- on exact ties such a
                                        scene may differ from the reference's tree order) */
} rtb_scene_info;
#define RTB_TIE_REFERENCE 0u
#define RTB_TIE_CALLER 1u
#define RTB_TIE_INDEX 2u

/* --- replaces BoundingVolumeHierarchy::new + Scene{..} (src/main.rs:349-358): uploads the
 *     primitives, flattens them to SoA float4, builds the BVH on the device. */
int rtb_scene_create(const rtb_scene_desc* desc, rtb_scene** out);
int rtb_scene_create_ex(const rtb_scene_desc* desc, const rtb_scene_options* options, rtb_scene** out);
void rtb_scene_destroy(rtb_scene* scene);
int rtb_scene_get_info(const rtb_scene* scene, rtb_scene_info* info);

/* Extension (the reference reads no materials, src/main.rs:130,137, and has no specular term): ks_ns = n_tris * 4
 * floats {Ks.rgb, Ns} (HOST, e.g. from rtb_import_obj_ex) switches the shading of a lit sample from (c * |n.l|) / denom to
 * (c * |n.l| + Ks * max(0, r.v)^Ns) / denom with r = 2 |n.l| n_f - l (n_f: the normal on the light's side) and v the unit
 * vector from the hit point to the eye.  NULL switches back to the reference's Lambert (the default).  Every other rule of
 * the path is unchanged; with Ks = 0 the pixels are those of the default path bit for bit. */
int rtb_scene_set_materials(rtb_scene* scene, const float* ks_ns);

/* --- replaces the random_samples table of render() (src/main.rs:253,260-265): n_pairs (u,v)
 *     pairs in [0,1).  Must be called before any render.  The table is an explicit input
 *     because the reference draws it from an OS-seeded RNG. */
int rtb_scene_set_samples(rtb_scene* scene, const float* samples, uint32_t n_pairs);

/* Same table, uploaded asynchronously: `samples` must be PINNED host memory (cudaHostAlloc / cudaHostRegister; pageable memory
 * falls back to rtb_scene_set_samples) and must stay valid and unchanged until the next render call on this scene has returned.
 * Returns at once.  The first entries (the only ones the light points and the light grids depend on) go first on the scene's
 * stream, the rest follows on a copy stream while the next render rebuilds the light grids; the primary-ray kernel waits for it. */
int rtb_scene_set_samples_async(rtb_scene* scene, const float* samples, uint32_t n_pairs);

/* --- replaces render(scene) up to (not including) the PNG save (src/main.rs:242-310):
 *     renders the rows owned by (rank, world) and copies them to the HOST buffer rgb8
 *     (rtb_owned_rows(...) * width * 3 bytes, row-major RGB8, bands in increasing order).
 *     Blocking.  stats may be NULL. */
int rtb_render(rtb_scene* scene, const rtb_render_params* params, uint8_t* rgb8, rtb_stats* stats);

/* rtb_render in two halves, for a caller that renders frame after frame, each with its own sample table (the reference draws a
 * fresh table in every render(), src/main.rs:260-265, and its orbit loop calls render() per frame, src/main.rs:319-362).
 *   rtb_render_begin  enqueues the frame and returns; rgb8 must be PINNED host memory and stay valid until rtb_render_end.
 *   rtb_scene_set_samples_async called between the two (pinned table) is the table of the NEXT frame: it goes into the scene's
 *                     spare table set, and its upload, light tables and light grids run next to the kernels of the frame in
 *                     flight (same parameters assumed; others are rebuilt by the next begin).  The table must stay valid until
 *                     the rtb_render_end of the frame that uses it.
 *   rtb_render_end    waits for the frame: rgb8 is complete when it returns.  Should a speculative grid rebuild not have fitted,
 *                     it is redone here and the frame rendered once more before returning (stats->grid_redo = 1).
 * One frame in flight per scene; every other render / set_samples call on the scene fails with RTB_ERR_STATE until rtb_render_end.
 * Frames are bit-identical to rtb_render with the same table. */
int rtb_render_begin(rtb_scene* scene, const rtb_render_params* params, uint8_t* rgb8);
int rtb_render_end(rtb_scene* scene, rtb_stats* stats);

/* render() on several GPUs of one box from ONE host thread's point of view (the reference's render fans out to its worker threads
 * and joins them itself, src/main.rs:269-303): scenes[i] is the same scene created on device i (rtb_scene_desc.device), all with the
 * same sample table (rtb_scene_set_samples[_async] on each, or rtb_scenes_set_samples).  Band b of params->band_rows rows (0 = 4)
 * is rendered by scenes[b % n_gpus] (params->rank / world are ignored); one host thread per device runs inside the call and every
 * device copies its finished rows straight to their place in rgb8, the full HOST frame (height * width * 3; pinned memory makes
 * the copies asynchronous).  With RTB_FLAG_GATHER_PEER the rows travel over NVLink into scenes[0]'s device frame instead and one
 * device-to-host copy follows.  Blocking.  stats (may be NULL) = sums over the devices, times = the slowest device.
 * The bytes equal those of rtb_render on one GPU. */
int rtb_render_multi(rtb_scene* const* scenes, uint32_t n_gpus, const rtb_render_params* params, uint8_t* rgb8, rtb_stats* stats);
int rtb_scenes_set_samples(rtb_scene* const* scenes, uint32_t n_gpus, const float* samples, uint32_t n_pairs);

/* Frame sequence (replaces calling render() + img.save per pose, src/main.rs:242-317): renders n_poses frames that
 * differ only in the camera.  Scene, BVH, sample table and light grids stay resident; the camera grid is rebuilt per pose without
 * host synchronisation; frames alternate between two device / pinned host buffers so that the device-to-host copy and the
 * consumer of frame k overlap the kernels of frame k + 1.  Per finished frame, in pose order:
 *   frames != NULL: the frame is copied to frames + k * height * width * 3 (HOST);
 *   png_pattern != NULL: a printf pattern with one %u (e.g. "frame_%04u.png"), the frame is encoded and written by a pool of host
 *     threads that runs alongside the rendering (img.save, src/main.rs:312-315);
 *   on_frame != NULL: called from the consumer thread with the pinned host frame (valid during the call).
 * params->rank / world / band_rows select band ownership as in rtb_render (world <= 1: whole frames).  Blocking; stats may be NULL. */
typedef struct { float eye[3], look_at[3], up[3], distance; } rtb_camera_pose;
typedef void (*rtb_frame_fn)(void* user, uint32_t pose_index, const uint8_t* rgb8, uint32_t width, uint32_t rows);
typedef struct {
    double seconds;                  /* wall time of the whole call */
    float device_ms;                 /* device time from the first kernel to the last copy (CUDA events) */
    uint32_t frames, cam_grid_redo;  /* frames delivered; poses whose speculative camera-grid build had to be redone */
    uint64_t primary_rays, shadow_rays;
} rtb_sequence_stats;
int rtb_render_sequence(rtb_scene* scene, const rtb_render_params* params, const rtb_camera_pose* poses, uint32_t n_poses,
                        uint8_t* frames, const char* png_pattern, rtb_frame_fn on_frame, void* user, rtb_sequence_stats* stats);
/* The same over several GPUs, frame-parallel: pose k is rendered whole by scenes[k % n_gpus] (replicas, no exchange); outputs as
 * above, delivered per device in pose order. */
int rtb_render_sequence_multi(rtb_scene* const* scenes, uint32_t n_gpus, const rtb_render_params* params, const rtb_camera_pose* poses,
                              uint32_t n_poses, uint8_t* frames, const char* png_pattern, rtb_sequence_stats* stats);

/* Roofline denominators measured on the device: k_fp32_peak (independent FFMA chains on every SM: FP32 TFLOP/s with
 * FMA = 2 flop), k_l2_peak (read-only sweep of a 32 MiB buffer by all SMs, 128-bit loads: GB/s) and the same sweep over 2 GiB
 * (HBM read GB/s).  Any output may be NULL. */
int rtb_measure_peaks(int32_t device, float* fp32_tflops, float* l2_read_gbs, float* hbm_read_gbs);

/* Same, but the output stays on the device: d_rgb8 is a device pointer on the scene's device,
 * cuda_stream a cudaStream_t (NULL = default stream).  Asynchronous: returns after the launch.
 * This is the form the multi-GPU gather uses (the buffer is an NCCL send buffer). */
int rtb_render_device(rtb_scene* scene, const rtb_render_params* params, void* d_rgb8, void* cuda_stream);

/* --- replaces render_pixel (src/main.rs:180-240) for an explicit list of n_pix (px,py) pairs;
 *     all outputs are HOST arrays indexed by list position, any may be NULL:
 *     rgb_lin 3 floats (the returned Color), rgb8 3 bytes (Color::to_rgba, color.rs:26-33),
 *     hit_count = primary rays of the pixel that hit, prim0/t0 = primitive and distance of ray 0
 *     (-1 / 0 on a miss) i.e. HitInfo.distance (bounding_volume_hierarchy.rs:150-156). */
int rtb_render_pixels(rtb_scene* scene, const rtb_render_params* params, const uint32_t* pixels, uint64_t n_pix,
                      float* rgb_lin, uint8_t* rgb8, uint8_t* hit_count, int32_t* prim0, float* t0,
                      rtb_stats* stats);

/* --- replaces Color::to_rgba (src/tracer/utils/color.rs:26-33) on its own: n linear f32 values (HOST) -> n bytes
 *     (HOST), computed by the same device code the trace kernel's epilogue uses. */
int rtb_color_to_u8(rtb_scene* scene, const float* linear, uint64_t n, uint8_t* out);

/* --- NO counterpart in the reference: added for callers that cannot allocate (code under
 *     CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The same conversion as rtb_color_to_u8, by the
 *     same kernel, but the call allocates NOTHING (no cudaMalloc / cudaFree, no allocating library routine): its device staging --
 *     which rtb_color_to_u8 cudaMallocs per call -- is a static device array of RTB_COLOR_STATIC_MAX values (5 MiB: 4 input bytes +
 *     1 output byte per value) that every GPU loading the library holds until the process exits.  n > RTB_COLOR_STATIC_MAX returns
 *     RTB_ERR_LIMIT.  linear / out are HOST arrays as in rtb_color_to_u8.  Every call on a GPU shares that one array (it keeps the
 *     previous call's contents; a call overwrites what it reads): calls that overlap on the same GPU -- from several threads or
 *     scenes -- race and must be serialised by the caller.  Blocking. */
#define RTB_COLOR_STATIC_MAX (1u << 20)
int rtb_color_to_u8_static(rtb_scene* scene, const float* linear, uint64_t n, uint8_t* out);

/* Raw device counters of the last render on this scene (16 values; diagnostics of the RTB_FLAG_COUNTERS build, layout in
 * csrc/rtb_common.cuh). */
int rtb_scene_get_counters(rtb_scene* scene, uint64_t* out16);

/* Device self-test: the shared-reciprocal divisions of the trace kernels (csrc/rtb_trace.cu, div3_exact / div_by) against
 * the compiler's IEEE `/` on n pseudo-random operand sets; *mismatches must come back 0. */
int rtb_selftest_division(rtb_scene* scene, uint64_t seed, uint64_t n, uint64_t* mismatches);

/* The tie ranks in use (HOST array of n_tris int32): the leaf order of the reference's tree as replayed on the
 * device, or the caller's, or the primitive index (see rtb_scene_desc.tie_rank). */
int rtb_scene_get_tie_ranks(rtb_scene* scene, int32_t* out);

/* The built BVH, for inspection and structural tests (HOST array of n_nodes * 16 floats).  Node i: floats 0-5 =
 * child 0 box (lo.xyz, hi.xyz), 6-11 = child 1 box, 12-13 = child indices as int32 bit patterns (>= 0 inner node,
 * < 0 = ~primitive), 14-15 unused.  Node 0 is the root.  The tree's shape is NOT part of the results contract. */
int rtb_scene_get_nodes(rtb_scene* scene, float* out);

/* Rows owned by (rank, world) for a frame of `height` rows in bands of band_rows. */
uint32_t rtb_owned_rows(uint32_t height, uint32_t band_rows, uint32_t rank, uint32_t world);

/* Rank 0 after the gather: d_staged holds world consecutive per-rank buffers (each padded to
 * stride_bytes); writes the row-major frame to d_frame.  Device pointers, asynchronous.
 * Replaces the put_pixel placement of src/main.rs:291-295. */
int rtb_deinterleave_bands(int32_t device, const void* d_staged, uint64_t stride_bytes, void* d_frame,
                           uint32_t width, uint32_t height, uint32_t band_rows, uint32_t world, void* cuda_stream);

/* Peer frame: the put_pixel placement of src/main.rs:291-295 across GPUs without a collective.  Rank 0 allocates the frame
 * and exports a 64-byte CUDA IPC handle (rtb_peer_alloc); every other rank maps it (rtb_peer_open; peer access over
 * NVLink is enabled by the mapping) and renders with RTB_FLAG_FRAME_ROWS into the mapped pointer.  The caller orders
 * "all ranks done" -> "rank 0 reads" itself (one barrier per frame).  rtb_peer_close unmaps (owner = 0) or frees (owner = 1). */
int rtb_peer_alloc(int32_t device, uint64_t bytes, void** d_ptr_out, uint8_t handle_out[64]);
int rtb_peer_open(int32_t device, const uint8_t handle[64], void** d_ptr_out);
int rtb_peer_close(int32_t device, void* d_ptr, int32_t owner);

/* --- host-side entry points kept from main.rs --- */

/* import_obj (src/main.rs:114-149): `v`/`f` lines only, split on single spaces, bare 1-based
 * indices.  *verts_out is malloc'ed n*9 floats (free with rtb_free).  An unreadable path is
 * NOT an error for the reference (prints "Not a valid path" and yields no triangles); here it
 * returns RTB_ERR_IO with *n_out = 0 so the caller can print and continue. */
int rtb_import_obj(const char* path, float** verts_out, uint32_t* n_out);
void rtb_free(void* p);

/* Robust OBJ + MTL import (an extension: the reference skips `mtllib` / `usemtl`, src/main.rs:130,137, and panics on anything
 * but single-space `v x y z` / `f a b c`).  Any blank runs, CR LF, comments, `v x y z [w]`, corners `a`, `a/b`, `a//c`, `a/b/c`,
 * negative indices, polygons (fan around the first corner).  flags & RTB_OBJ_MTL: `mtllib` (relative to the OBJ's directory)
 * and `usemtl` are honoured; without it every triangle is white as in the reference (src/main.rs:143).
 * Outputs (malloc'ed, rtb_free; kd_out / ks_ns_out may be NULL): verts n*9, kd n*3 (the colour to pass as
 * rtb_scene_desc.colors), ks_ns n*4 {Ks.rgb, Ns} (models/big_bunny.mtl:4-10).  On the reference's own models the triangles
 * equal rtb_import_obj's. */
#define RTB_OBJ_MTL 1u
int rtb_import_obj_ex(const char* path, uint32_t flags, float** verts_out, float** kd_out, float** ks_ns_out, uint32_t* n_out);

/* import_obj (src/main.rs:114-149) for large files: rtb_import_obj_ex's grammar without materials, parsed by n_threads
 * host threads (0 = all cores) in three passes over chunks of the file (count, vertices, faces) straight into one output array.
 * flags & RTB_OBJ_PINNED: *verts_out is page-locked host memory (cudaHostAlloc: rtb_scene_create then uploads it as one DMA) and is
 * freed with rtb_free_pinned; otherwise malloc'ed, rtb_free.  The triangles equal rtb_import_obj_ex's, in the same order.
 * stats may be NULL.  An unreadable path returns RTB_ERR_IO with "Not a valid path: ..." as rtb_import_obj does. */
#define RTB_OBJ_PINNED 2u
typedef struct {
    double read_s, parse_s;          /* wall time of reading the file into memory / of the three passes (without alloc_s) */
    double alloc_s;                  /* time inside the output allocation: page-locking costs about 1.5 s per GB (cudaHostAlloc), which only
                                        pays when the array is uploaded several times (one scene per GPU); a single upload from pageable
                                        memory is faster than pinning first */
    uint64_t file_bytes;
    uint32_t threads, n_vertices;
} rtb_import_stats;
int rtb_import_obj_parallel(const char* path, uint32_t n_threads, uint32_t flags, float** verts_out, uint32_t* n_out, rtb_import_stats* stats);
/* The same with materials: flags & RTB_OBJ_MTL honours `mtllib` / `usemtl` exactly as rtb_import_obj_ex does and returns kd_out (n * 3)
 * and ks_ns_out (n * 4) beside the triangles (either may be NULL; without the flag they come back NULL and every triangle is the
 * reference's white).  All three arrays come from the same allocator (RTB_OBJ_PINNED: rtb_free_pinned, else rtb_free). */
int rtb_import_obj_parallel_ex(const char* path, uint32_t n_threads, uint32_t flags, float** verts_out, float** kd_out, float** ks_ns_out,
                               uint32_t* n_out, rtb_import_stats* stats);
void rtb_free_pinned(void* p);

/* --- the Sphere primitive and the random scene generators, dead code in the reference (src/tracer/primitives/sphere.rs is
 *     complete, but main() builds triangle scenes only; gen_random_spheres / gen_random_triangles, src/main.rs:43-99, are
 *     #[allow(dead_code)]).  Kept at the same level here: exact pieces with a batch entry, not wired into the frame kernels.
 * rtb_spheres_closest_hit: per ray (origin, direction: n_rays * 3 floats each, HOST; the direction is normalised as Ray::new does) the
 *     closest of n_spheres spheres {x, y, z, radius}: idx_out (-1 = none; an exact tie goes to the later sphere), t_out = the distance
 *     Intersectable::intersect returns (sphere.rs:50-84), normal_out = HasNormal::get_normal at origin + t * direction (sphere.rs:90-94).
 *     Bit-exact with the oracle's restatement (tests/test_gpu_parity.py).
 * rtb_spheres_bounds: HasBoundingBox (sphere.rs:31-43), on the host.
 * rtb_gen_random_spheres / _triangles: the generators' shapes and ranges from an explicit seed (the reference draws from an OS-seeded
 *     RNG): 1 sphere (4 + 3 floats) / 10 triangles (90 + 30 floats); return the primitive count. */
int rtb_spheres_closest_hit(int32_t device, const float* spheres_xyzr, uint32_t n_spheres, const float* origins, const float* dirs,
                            uint64_t n_rays, float* t_out, int32_t* idx_out, float* normal_out);
void rtb_spheres_bounds(const float* spheres_xyzr, uint32_t n, float* boxes6_out);
uint32_t rtb_gen_random_spheres(uint64_t seed, float* spheres_xyzr_out, float* colors_out);
uint32_t rtb_gen_random_triangles(uint64_t seed, float* verts_out, float* colors_out);

/* Deterministic stand-in for thread_rng (src/main.rs:260-265): SplitMix64(seed) stream, value j =
 * top 24 bits * 2^-24; out holds 2*n_pairs floats. */
void rtb_fill_samples(uint64_t seed, uint32_t n_pairs, float* out);

/* img.save(fout, image::PNG) (src/main.rs:312-315): RGB8, row-major. */
int rtb_write_png(const char* path, const uint8_t* rgb8, uint32_t width, uint32_t height);

const char* rtb_last_error(void);
const char* rtb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RTB200_H */
