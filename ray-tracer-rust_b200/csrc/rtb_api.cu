// rtb_api.cu -- the extern "C" boundary declared in include/rtb200.h.
// Device scene ownership, per-frame argument marshalling, launches, host<->device copies.
// There is deliberately no CPU implementation behind any compute entry point.
#include "rtb_common.cuh"
#include "../../include/rtb200.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>
#include <chrono>

// launchers / builders defined in the other translation units
struct RtbBuildInfo { uint32_t n_nodes, max_depth, iterations; float sah_cost, build_ms; };
cudaError_t rtb_build_scene(const float*, const float*, const int*, uint32_t, float4*, float4*, float4**, int32_t*, RtbBuildInfo*, cudaStream_t);
cudaError_t rtb_tie_rank_device(const float4* d_tris, uint32_t n, int* d_rank_out, float4* d_colors4, cudaStream_t st);
cudaError_t rtb_launch_trace(const TraceArgs&, bool, cudaStream_t, cudaEvent_t*, cudaEvent_t);
cudaError_t rtb_grid_scene_stats(RtbGrid&, const float4*, uint32_t, cudaStream_t);
cudaError_t rtb_grid_build(RtbGrid&, const float4*, uint32_t, const float4*, uint32_t, float, bool, cudaStream_t);
cudaError_t rtb_cam_grid_build(RtbGrid&, const float4*, uint32_t, float4*, float4*, const CameraBasis&, float, bool, cudaStream_t);
cudaError_t rtb_grid_validate(RtbGrid&, cudaStream_t, bool*);
cudaError_t rtb_grid_sort_on_reuse(RtbGrid&, cudaStream_t);
void rtb_grid_free(RtbGrid&);
cudaError_t rtb_launch_light_table(const float2*, uint32_t, const float*, float4*, float*, cudaStream_t);
cudaError_t rtb_launch_light_groups(const float4*, const float*, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint16_t*, float*, cudaStream_t);
cudaError_t rtb_launch_fill_f32(float*, size_t, float, cudaStream_t);
cudaError_t rtb_launch_deinterleave(const uint8_t*, uint64_t, uint8_t*, uint32_t, uint32_t, uint32_t, uint32_t, cudaStream_t);
cudaError_t rtb_upload_gamma_table(const float*);
cudaError_t rtb_launch_color_to_u8(const float*, uint64_t, uint8_t*, cudaStream_t);
cudaError_t rtb_color_to_u8_static_run(int, const float*, uint64_t, uint8_t*, cudaStream_t);
cudaError_t rtb_launch_get_ranks(const float4*, uint32_t, int*, cudaStream_t);
cudaError_t rtb_launch_div_selftest(uint64_t, uint64_t, unsigned long long*, cudaStream_t);
cudaError_t rtb_launch_cdiv_check(float, float, float, uint32_t*, cudaStream_t);

static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
    return code;
}
#define CUDA_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail(RTB_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e_)); } while (0)

struct rtb_scene {
    int device;
    uint32_t n_tris;
    float light[9];
    float4 *d_tris, *d_colors4, *d_nodes;
    float4* d_ksns;                              // {Ks.rgb, Ns} per primitive (rtb_scene_set_materials), nullptr = Lambert only
    int32_t root;
    RtbBuildInfo info;
    float2* d_samples; uint32_t n_samples;
    float4* d_light_pts; uint32_t light_pts_cap;
    uint16_t* d_light_perm; size_t light_perm_cap;
    float* d_light_gbox; size_t light_gbox_cap;
    unsigned long long* d_counters;
    bool trace_chunks;                           // RTB200_TRACE_CHUNKS=1: per-chunk kernel times of profiled renders on stderr
    float* d_light_box;                          // 6 floats, bounds of d_light_pts (k_light_bounds)
    uint8_t* d_frame; size_t frame_cap;          // staging for rtb_render (host output)
    // wavefront buffers, grown on demand and kept (see rtb_trace.cu)
    // (two sets: consecutive chunks run on two worker streams so that one chunk's kernel tails overlap the other's work)
    struct WaveBufs {
        uint2* d_records; float4* d_items; uint32_t* d_masks; uint32_t* d_queue_ctl;
        uint4* d_subitems; uint4* d_sublists; int* d_cand_pool;
        size_t rays_cap, mask_cap, sub_cap, pool_cap;
        uint32_t* d_tile_slot; uint4* d_term_tiles; float* d_terms; size_t tile_slot_cap, term_slots, term_floats;   // term wavefront
    } wb[2];
    cudaStream_t wstream[2];
    cudaEvent_t ev_fork, ev_join[2];
    uint8_t* h_pinned; size_t pinned_cap;
    cudaStream_t stream;
    cudaEvent_t ev0, ev1;
    cudaEvent_t evp[4 * 32];                     // per-chunk events around the wavefront kernels (per-kernel times in rtb_stats)
    uint32_t timed_chunks, n_chunks;
    uint64_t chunk_rays;                         // rays per wavefront chunk (RTB200_CHUNK_RAYS, default 16 Mi)
    RtbGrid grid;                                // light grids (rtb_grid.cu), rebuilt when the sample table or the light-point count changes
    uint64_t samples_version;
    int shadow_mode;                             // RTB200_SHADOW: 0 auto (grid when applicable), 1 lists/walk only, 2 grid (same as auto)
    float grid_cell_factor;                      // RTB200_GRID_CELL: cell size / log-mean primitive extent
    // camera grid (rtb_grid.cu): the light-grid structure with the eye as projection centre, for primary rays
    RtbGrid cam_grid; float4* d_cam_tris; float4* d_cam_origin; CameraBasis cam_key; bool cam_key_set;
    bool light_tables_ran;                       // the last prepare_args launched the four light-table kernels
    bool light_key_set; uint64_t light_key_samples, light_key_need; uint32_t light_key_nb_ray, light_key_nls; cudaStream_t light_key_stream;
    int primary_mode;                            // RTB200_PRIMARY: 0 auto (camera grid when it applies), 1 bvh
    float cam_cell_px;                           // RTB200_CAM_CELL: camera-grid cell size in pixels
    // two-operation division by the frame's denom (k_shade): checked exhaustively on the device once per (denom, stream)
    uint32_t* d_cdiv_bad; float cdiv_denom; bool cdiv_key_set; cudaStream_t cdiv_stream;
    bool all_grey;                               // every primitive colour has r == g == b: one term channel instead of three
    int term_mode;                               // RTB200_TERMS: 0 auto (single-chunk renders), 1 always, -1 never
    bool terms_ran;                              // the last render launched k_shade_terms
    uint64_t term_bytes;                         // RTB200_TERM_MB: memory per buffer set for the term wavefront (default 1 GiB, 0 = off)
    bool cdiv_off;                               // RTB200_CDIV=0: always the three-operation shared-reciprocal division
    uint32_t tie_mode;                           // RTB_TIE_*
    // asynchronous sample-table upload (rtb_scene_set_samples_async): the prefix the light tables need travels on `stream`,
    // the rest on `copy_stream`; renders make their stream wait for ev_prefix / ev_samples where they need the data
    cudaStream_t copy_stream; cudaEvent_t ev_prefix, ev_samples, ev_idle; bool samples_pending; uint32_t samples_prefix;
    uint32_t grid_rebuilds, grid_redo;           // grid builds enqueued / redone by the last blocking render
    bool no_refine;                              // RTB200_NO_REFINE: never rebuild the light grids finer for a table that keeps being reused
    // rtb_render_multi with RTB_FLAG_GATHER_PEER: the full frame on this device, written by every device of the call
    uint8_t* d_full; size_t full_cap;
    // rtb_render_sequence: two device frames, two pinned host frames, snapshots of the camera grid's control words
    uint8_t* d_seq[2]; uint8_t* h_seq[2]; size_t seq_cap; uint32_t* h_seq_ctl; cudaEvent_t ev_seq[2], ev_seq0, ev_seq1;
    // Everything that is derived from ONE sample table: the table, its light tables and its light grids.  The scene owns two such
    // sets -- the members above are the current one, `alt` the spare -- so that the table of the NEXT frame can be uploaded and its
    // grids built (on build_stream) while the kernels of a frame begun with rtb_render_begin still read the current set.
    struct TableSet {
        float2* d_samples; uint32_t n_samples; float4* d_light_pts; uint32_t light_pts_cap; uint16_t* d_light_perm; size_t light_perm_cap;
        float* d_light_gbox; size_t light_gbox_cap; float* d_light_box; RtbGrid grid; uint64_t samples_version;
        bool light_key_set; uint64_t light_key_samples, light_key_need; uint32_t light_key_nb_ray, light_key_nls; cudaStream_t light_key_stream;
        cudaEvent_t ev_prefix, ev_samples; bool samples_pending; uint32_t samples_prefix;
    } alt;
    bool alt_inited, alt_ready;                  // alt_ready: alt holds a prefetched table, the next rtb_render_begin takes it
    cudaStream_t build_stream; cudaEvent_t ev_built;
    uint64_t version_counter;                    // source of samples_version (unique across both sets)
    uint32_t prefetch_rebuilds;
    // rtb_render_begin .. rtb_render_end
    struct Pending {
        bool active; rtb_render_params p; uint8_t* h_rgb8; uint8_t* d_full; TraceArgs a; uint8_t* stage; bool h_frame_rows; uint32_t rebuilds;
    } pend;
};

static void swap_table_sets(rtb_scene* s) {
    rtb_scene::TableSet& t = s->alt;
    std::swap(s->d_samples, t.d_samples); std::swap(s->n_samples, t.n_samples); std::swap(s->d_light_pts, t.d_light_pts);
    std::swap(s->light_pts_cap, t.light_pts_cap); std::swap(s->d_light_perm, t.d_light_perm); std::swap(s->light_perm_cap, t.light_perm_cap);
    std::swap(s->d_light_gbox, t.d_light_gbox); std::swap(s->light_gbox_cap, t.light_gbox_cap); std::swap(s->d_light_box, t.d_light_box);
    std::swap(s->grid, t.grid); std::swap(s->samples_version, t.samples_version);
    std::swap(s->light_key_set, t.light_key_set); std::swap(s->light_key_samples, t.light_key_samples); std::swap(s->light_key_need, t.light_key_need);
    std::swap(s->light_key_nb_ray, t.light_key_nb_ray); std::swap(s->light_key_nls, t.light_key_nls); std::swap(s->light_key_stream, t.light_key_stream);
    std::swap(s->ev_prefix, t.ev_prefix); std::swap(s->ev_samples, t.ev_samples); std::swap(s->samples_pending, t.samples_pending);
    std::swap(s->samples_prefix, t.samples_prefix);
}

// ---- Color::to_rgba threshold table (color.rs:10-33) ---------------------------------------
static uint8_t host_gamma_u8(float linear) {
    const float GAMMA = 2.2f;
    float g = powf(linear, 1.0f / GAMMA) * 255.0f;
    if (!(g == g) || g <= 0.0f) return 0;
    if (g >= 255.0f) return 255;
    return (uint8_t)g;
}
// thr[k] = smallest non-negative f32 whose byte is >= k.  Positive floats order like their bit
// patterns, so this is a binary search over bit patterns; it assumes libm's powf is monotone,
// which the CPU and GPU test suites check around every threshold.
static void build_gamma_table(float* thr) {
    thr[0] = 0.0f;
    for (int k = 1; k < 256; ++k) {
        uint32_t lo = 0, hi = 0x40000000u;   // 0.0 .. 2.0
        while (lo < hi) {
            uint32_t mid = lo + (hi - lo) / 2;
            float c; memcpy(&c, &mid, 4);
            if (host_gamma_u8(c) >= k) hi = mid; else lo = mid + 1;
        }
        memcpy(&thr[k], &lo, 4);
    }
}

// ---- Camera::new (camera.rs:17-35), unfused f32 on the host ---------------------------------
static CameraBasis make_camera(const rtb_render_params* p) {
    CameraBasis c;
    const f3 eye = mk3(p->eye[0], p->eye[1], p->eye[2]), look = mk3(p->look_at[0], p->look_at[1], p->look_at[2]);
    c.w = normalize3(sub3(eye, look));
    const f3 o = normalize3(mk3(p->up[0], p->up[1], p->up[2]));
    c.u = normalize3(cross3(o, c.w));
    c.v = normalize3(cross3(c.u, c.w));
    c.eye = eye;
    c.distance = p->distance;
    return c;
}

extern "C" {

void rtb_set_error_(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg ? msg : ""); }
const char* rtb_last_error(void) { return g_err; }
const char* rtb_version(void) { return "rtb200 0.1 (sm_100a)"; }

uint32_t rtb_owned_rows(uint32_t height, uint32_t band_rows, uint32_t rank, uint32_t world) {
    if (world <= 1) return height;
    if (band_rows == 0 || rank >= world) return 0;
    const uint32_t n_bands = (height + band_rows - 1) / band_rows;
    uint32_t rows = 0;
    for (uint32_t b = rank; b < n_bands; b += world) {
        const uint32_t y0 = b * band_rows;
        rows += (height - y0 < band_rows) ? height - y0 : band_rows;
    }
    return rows;
}

void rtb_scene_destroy(rtb_scene* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    cudaFree(s->d_tris); cudaFree(s->d_colors4); cudaFree(s->d_nodes); cudaFree(s->d_samples); cudaFree(s->d_light_pts); cudaFree(s->d_light_perm); cudaFree(s->d_light_gbox);
    cudaFree(s->d_counters); cudaFree(s->d_light_box); cudaFree(s->d_frame); cudaFree(s->d_cdiv_bad); cudaFree(s->d_ksns);
    for (int k = 0; k < 2; ++k) {
        rtb_scene::WaveBufs& b = s->wb[k];
        cudaFree(b.d_records); cudaFree(b.d_items); cudaFree(b.d_masks); cudaFree(b.d_queue_ctl);
        cudaFree(b.d_subitems); cudaFree(b.d_sublists); cudaFree(b.d_cand_pool);
        cudaFree(b.d_tile_slot); cudaFree(b.d_term_tiles); cudaFree(b.d_terms);
        if (s->wstream[k]) cudaStreamDestroy(s->wstream[k]);
        if (s->ev_join[k]) cudaEventDestroy(s->ev_join[k]);
    }
    if (s->ev_fork) cudaEventDestroy(s->ev_fork);
    rtb_grid_free(s->grid); rtb_grid_free(s->cam_grid); cudaFree(s->d_cam_tris); cudaFree(s->d_cam_origin);
    cudaFree(s->d_full);
    for (int k = 0; k < 2; ++k) { cudaFree(s->d_seq[k]); if (s->h_seq[k]) cudaFreeHost(s->h_seq[k]); if (s->ev_seq[k]) cudaEventDestroy(s->ev_seq[k]); }
    if (s->h_seq_ctl) cudaFreeHost(s->h_seq_ctl);
    if (s->ev_seq0) cudaEventDestroy(s->ev_seq0);
    if (s->ev_seq1) cudaEventDestroy(s->ev_seq1);
    cudaFree(s->alt.d_samples); cudaFree(s->alt.d_light_pts); cudaFree(s->alt.d_light_perm); cudaFree(s->alt.d_light_gbox); cudaFree(s->alt.d_light_box);
    rtb_grid_free(s->alt.grid);
    if (s->alt.ev_prefix) cudaEventDestroy(s->alt.ev_prefix);
    if (s->alt.ev_samples) cudaEventDestroy(s->alt.ev_samples);
    if (s->build_stream) cudaStreamDestroy(s->build_stream);
    if (s->ev_built) cudaEventDestroy(s->ev_built);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    if (s->ev_prefix) cudaEventDestroy(s->ev_prefix);
    if (s->ev_samples) cudaEventDestroy(s->ev_samples);
    if (s->ev_idle) cudaEventDestroy(s->ev_idle);
    if (s->h_pinned) cudaFreeHost(s->h_pinned);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    for (int k = 0; k < 4 * 32; ++k) if (s->evp[k]) cudaEventDestroy(s->evp[k]);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

int rtb_scene_create(const rtb_scene_desc* desc, rtb_scene** out) { return rtb_scene_create_ex(desc, nullptr, out); }

int rtb_scene_create_ex(const rtb_scene_desc* desc, const rtb_scene_options* opt_in, rtb_scene** out) {
    if (!desc || !out) return fail(RTB_ERR_ARG, "rtb_scene_create: null argument");
    rtb_scene_options opt;
    memset(&opt, 0, sizeof(opt));
    if (opt_in) {
        if (opt_in->struct_size == 0 || opt_in->struct_size > sizeof(opt)) return fail(RTB_ERR_ARG, "rtb_scene_create_ex: options.struct_size %u (this library: %zu)", opt_in->struct_size, sizeof(opt));
        memcpy(&opt, opt_in, opt_in->struct_size);
    }
    *out = nullptr;
    if (desc->n_tris && !desc->verts) return fail(RTB_ERR_ARG, "rtb_scene_create: null vertex array");
    if (desc->n_tris > 0x3fffffffu) return fail(RTB_ERR_LIMIT, "rtb_scene_create: too many primitives");
    int n_dev = 0;
    CUDA_TRY(cudaGetDeviceCount(&n_dev));
    if (desc->device < 0 || desc->device >= n_dev) return fail(RTB_ERR_ARG, "rtb_scene_create: device %d of %d", desc->device, n_dev);
    CUDA_TRY(cudaSetDevice(desc->device));

    float thr[256];
    build_gamma_table(thr);
    CUDA_TRY(rtb_upload_gamma_table(thr));

    rtb_scene* s = new rtb_scene();
    memset(s, 0, sizeof(*s));
    s->device = desc->device; s->n_tris = desc->n_tris; s->root = RTB_CHILD_NONE;
    memcpy(s->light, desc->light, sizeof(s->light));
    const uint32_t n = desc->n_tris;
    float *d_verts = nullptr, *d_colors = nullptr; int* d_rank = nullptr;
    int rc = RTB_OK;
    cudaError_t e;
#define TRY_(x) do { e = (x); if (e != cudaSuccess) { rc = fail(RTB_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e)); goto cleanup; } } while (0)
    TRY_(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    TRY_(cudaEventCreate(&s->ev0)); TRY_(cudaEventCreate(&s->ev1));
    for (int k = 0; k < 4 * 32; ++k) TRY_(cudaEventCreate(&s->evp[k]));
    // tunables: rtb_scene_options (0 = default), then the environment variables of the same meaning as overrides
    s->chunk_rays = opt.chunk_rays ? opt.chunk_rays : 16ull << 20;
    if (const char* env = getenv("RTB200_CHUNK_RAYS")) { const long long v = atoll(env); if (v > 0) s->chunk_rays = (uint64_t)v; }
    s->shadow_mode = opt.shadow_mode == 1 ? 1 : 0; s->grid_cell_factor = opt.grid_cell_factor > 0.0f ? opt.grid_cell_factor : 0.25f;
    if (const char* env = getenv("RTB200_SHADOW")) s->shadow_mode = !strcmp(env, "lists") ? 1 : (!strcmp(env, "grid") ? 2 : 0);
    s->cdiv_off = opt.cdiv_off != 0;
    if (const char* env = getenv("RTB200_CDIV")) s->cdiv_off = atoi(env) == 0;
    s->term_mode = opt.term_mode > 0 ? 1 : (opt.term_mode < 0 ? -1 : 0);
    if (const char* env = getenv("RTB200_TERMS")) s->term_mode = !strcmp(env, "1") ? 1 : (!strcmp(env, "0") ? -1 : 0);
    s->term_bytes = opt.term_bytes == UINT64_MAX ? 0 : (opt.term_bytes ? opt.term_bytes : 1ull << 30);
    if (const char* env = getenv("RTB200_TERM_MB")) { const long long v = atoll(env); if (v >= 0) s->term_bytes = (uint64_t)v << 20; }
    s->grid.entry_cap_limit = opt.grid_entry_cap; s->cam_grid.entry_cap_limit = opt.grid_entry_cap;
    s->all_grey = true;
    for (uint32_t i = 0; desc->colors && i < desc->n_tris && s->all_grey; ++i)
        s->all_grey = desc->colors[3 * i] == desc->colors[3 * i + 1] && desc->colors[3 * i + 1] == desc->colors[3 * i + 2];
    s->trace_chunks = opt.trace_chunks != 0;
    if (const char* env = getenv("RTB200_TRACE_CHUNKS")) s->trace_chunks = atoi(env) != 0;
    s->cam_cell_px = opt.cam_cell_px > 0.0f ? opt.cam_cell_px : 2.0f;
    if (const char* env = getenv("RTB200_CAM_CELL")) { const float v = (float)atof(env); if (v > 0.0f) s->cam_cell_px = v; }
    s->primary_mode = opt.primary_mode == 1 ? 1 : 0;
    if (const char* env = getenv("RTB200_PRIMARY")) s->primary_mode = strcmp(env, "bvh") == 0 ? 1 : 0;
    TRY_(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
    TRY_(cudaEventCreateWithFlags(&s->ev_prefix, cudaEventDisableTiming));
    TRY_(cudaEventCreateWithFlags(&s->ev_samples, cudaEventDisableTiming));
    TRY_(cudaEventCreateWithFlags(&s->ev_idle, cudaEventDisableTiming));
    s->no_refine = getenv("RTB200_NO_REFINE") != nullptr;
    if (const char* env = getenv("RTB200_GRID_CELL")) { const float v = (float)atof(env); if (v > 0.0f) s->grid_cell_factor = v; }
    TRY_(cudaMalloc(&s->d_counters, RTB_NUM_CTRS * sizeof(unsigned long long)));
    TRY_(cudaMalloc(&s->d_light_box, 8 * sizeof(float)));
    for (int k = 0; k < 2; ++k) {
        TRY_(cudaMalloc(&s->wb[k].d_queue_ctl, 16 * sizeof(uint32_t)));
        TRY_(cudaStreamCreateWithFlags(&s->wstream[k], cudaStreamNonBlocking));
        TRY_(cudaEventCreateWithFlags(&s->ev_join[k], cudaEventDisableTiming));
    }
    TRY_(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
    if (n) {
        TRY_(cudaMalloc(&d_verts, (size_t)n * 9 * 4)); TRY_(cudaMalloc(&d_colors, (size_t)n * 3 * 4));
        TRY_(cudaMemcpyAsync(d_verts, desc->verts, (size_t)n * 9 * 4, cudaMemcpyHostToDevice, s->stream));
        if (desc->colors) TRY_(cudaMemcpyAsync(d_colors, desc->colors, (size_t)n * 3 * 4, cudaMemcpyHostToDevice, s->stream));
        else TRY_(rtb_launch_fill_f32(d_colors, (size_t)n * 3, 1.0f, s->stream));                 // Color(1, 1, 1), src/main.rs:143
        if (desc->tie_rank) {
            TRY_(cudaMalloc(&d_rank, (size_t)n * 4));
            TRY_(cudaMemcpyAsync(d_rank, desc->tie_rank, (size_t)n * 4, cudaMemcpyHostToDevice, s->stream));
        }
        TRY_(cudaMalloc(&s->d_tris, (size_t)n * 3 * sizeof(float4)));
        TRY_(cudaMalloc(&s->d_colors4, (size_t)n * sizeof(float4)));
        TRY_(rtb_build_scene(d_verts, d_colors, d_rank, n, s->d_tris, s->d_colors4, &s->d_nodes, &s->root, &s->info, s->stream));
        s->tie_mode = desc->tie_rank ? RTB_TIE_CALLER : RTB_TIE_INDEX;
        if (!desc->tie_rank && n > 1 && n <= RTB_TIE_RANK_MAX_TRIS) {
            TRY_(rtb_tie_rank_device(s->d_tris, n, nullptr, s->d_colors4, s->stream));
            s->tie_mode = RTB_TIE_REFERENCE;
        }
        if (!desc->tie_rank && n <= 1) s->tie_mode = RTB_TIE_REFERENCE;
        TRY_(cudaStreamSynchronize(s->stream));
        TRY_(rtb_grid_scene_stats(s->grid, s->d_tris, n, s->stream));
        if (s->info.max_depth > RTB_STACK_SIZE - 1) {
            rc = fail(RTB_ERR_LIMIT, "BVH depth %u exceeds traversal stack %d", s->info.max_depth, RTB_STACK_SIZE);
            goto cleanup;
        }
    }
cleanup:
    cudaFree(d_verts); cudaFree(d_colors); cudaFree(d_rank);
    if (rc != RTB_OK) { rtb_scene_destroy(s); return rc; }
    *out = s;
    return RTB_OK;
#undef TRY_
}

int rtb_scene_set_materials(rtb_scene* s, const float* ks_ns) {
    if (!s) return fail(RTB_ERR_ARG, "rtb_scene_set_materials: null scene");
    CUDA_TRY(cudaSetDevice(s->device));
    if (!ks_ns || s->n_tris == 0) { cudaFree(s->d_ksns); s->d_ksns = nullptr; return RTB_OK; }
    if (!s->d_ksns) CUDA_TRY(cudaMalloc(&s->d_ksns, (size_t)s->n_tris * sizeof(float4)));
    CUDA_TRY(cudaMemcpyAsync(s->d_ksns, ks_ns, (size_t)s->n_tris * sizeof(float4), cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return RTB_OK;
}

int rtb_scene_get_info(const rtb_scene* s, rtb_scene_info* info) {
    if (!s || !info) return fail(RTB_ERR_ARG, "rtb_scene_get_info: null argument");
    info->n_tris = s->n_tris; info->n_nodes = s->info.n_nodes; info->max_depth = s->info.max_depth;
    info->ploc_iterations = s->info.iterations; info->sah_cost = s->info.sah_cost; info->build_ms = s->info.build_ms;
    info->tie_mode = s->tie_mode;
    return RTB_OK;
}

int rtb_scene_set_samples(rtb_scene* s, const float* samples, uint32_t n_pairs) {
    if (!s || !samples || n_pairs == 0) return fail(RTB_ERR_ARG, "rtb_scene_set_samples: null/empty table");
    if (s->pend.active) return fail(RTB_ERR_STATE, "rtb_scene_set_samples: a frame is in flight (rtb_render_begin); use rtb_scene_set_samples_async with a pinned table");
    CUDA_TRY(cudaSetDevice(s->device));
    s->alt_ready = false;                         // a table prefetched earlier is superseded
    if (n_pairs != s->n_samples) {
        cudaFree(s->d_samples); s->d_samples = nullptr; s->n_samples = 0;
        CUDA_TRY(cudaMalloc(&s->d_samples, (size_t)n_pairs * sizeof(float2)));
        s->n_samples = n_pairs;
    }
    s->samples_pending = false;
    CUDA_TRY(cudaMemcpyAsync(s->d_samples, samples, (size_t)n_pairs * sizeof(float2), cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->samples_version = ++s->version_counter;
    return RTB_OK;
}

static int prefetch_samples(rtb_scene* s, const float* samples, uint32_t n_pairs);

#define RTB_SAMPLES_PREFIX 65536u      // pairs uploaded first (512 KB): far more than any frame's light points (nb_ray^2 + nb_light_sample)

int rtb_scene_set_samples_async(rtb_scene* s, const float* samples, uint32_t n_pairs) {
    if (!s || !samples || n_pairs == 0) return fail(RTB_ERR_ARG, "rtb_scene_set_samples_async: null/empty table");
    cudaPointerAttributes pa;
    const bool pinned = cudaPointerGetAttributes(&pa, samples) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    if (!pinned) { cudaGetLastError(); return rtb_scene_set_samples(s, samples, n_pairs); }
    CUDA_TRY(cudaSetDevice(s->device));
    if (s->pend.active) return prefetch_samples(s, samples, n_pairs);       // a frame is in flight: the table of the NEXT frame
    s->alt_ready = false;                         // a table prefetched earlier is superseded
    if (n_pairs != s->n_samples) {
        CUDA_TRY(cudaStreamSynchronize(s->stream));
        cudaFree(s->d_samples); s->d_samples = nullptr; s->n_samples = 0;
        CUDA_TRY(cudaMalloc(&s->d_samples, (size_t)n_pairs * sizeof(float2)));
        s->n_samples = n_pairs;
    }
    // the previous frame (on the scene's stream) has read the old table: the copy stream starts after it
    CUDA_TRY(cudaEventRecord(s->ev_idle, s->stream));
    CUDA_TRY(cudaStreamWaitEvent(s->copy_stream, s->ev_idle, 0));
    const uint32_t prefix = n_pairs < RTB_SAMPLES_PREFIX ? n_pairs : RTB_SAMPLES_PREFIX;
    CUDA_TRY(cudaMemcpyAsync(s->d_samples, samples, (size_t)prefix * sizeof(float2), cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaEventRecord(s->ev_prefix, s->stream));
    if (n_pairs > prefix)
        CUDA_TRY(cudaMemcpyAsync(s->d_samples + prefix, samples + 2 * (size_t)prefix, (size_t)(n_pairs - prefix) * sizeof(float2),
                                 cudaMemcpyHostToDevice, s->copy_stream));
    CUDA_TRY(cudaEventRecord(s->ev_samples, s->copy_stream));
    s->samples_pending = true; s->samples_prefix = prefix;
    s->samples_version = ++s->version_counter;
    return RTB_OK;
}

}  // extern "C"

// What a frame derives from the current sample table, (re)built on `st` where it is not cached: the light-point tables (points,
// pairs, bounds, sorted groups) and the light grids.
struct TableInfo { uint64_t need; uint32_t n_groups, group_size; bool use_perm; };
static int build_tables(rtb_scene* s, const rtb_render_params* p, cudaStream_t st, bool allow_spec, TableInfo* ti) {
    // light points used by a frame: indices (r*nb_ray + i) % n_samples, r < nb_ray, i < nb_light_sample
    uint64_t need = p->nb_ray ? (uint64_t)(p->nb_ray - 1) * p->nb_ray + p->nb_light_sample : 0;
    if (need > s->n_samples) need = s->n_samples;
    if (need > s->light_pts_cap) {
        cudaFree(s->d_light_pts); s->d_light_pts = nullptr; s->light_pts_cap = 0;
        CUDA_TRY(cudaMalloc(&s->d_light_pts, (size_t)need * 3 * sizeof(float4)));      // points + pair table (k_light_pairs)
        s->light_pts_cap = (uint32_t)need;
    }
    // the light tables depend only on the sample table and on (nb_ray, nb_light_sample): computed once per such key.  They are
    // produced asynchronously on `st`, so a call on another stream recomputes them (stream order is the only ordering here).
    const bool lights_cached = s->light_key_set && s->light_key_samples == s->samples_version && s->light_key_need == need &&
                               s->light_key_nb_ray == p->nb_ray && s->light_key_nls == p->nb_light_sample && s->light_key_stream == st;
    s->light_tables_ran = !lights_cached;
    if (s->samples_pending) {                     // asynchronous table upload: the light tables need its prefix only
        CUDA_TRY(cudaStreamWaitEvent(st, s->ev_prefix, 0));
        if (need > s->samples_prefix) CUDA_TRY(cudaStreamWaitEvent(st, s->ev_samples, 0));
    }
    if (!lights_cached) CUDA_TRY(rtb_launch_light_table(s->d_samples, (uint32_t)need, s->light, s->d_light_pts, s->d_light_box, st));
    // light-sample groups (k_light_groups): at most RTB_MAX_GROUPS runs of the Hilbert-sorted samples, ~32 samples each
#ifndef RTB_GROUP_TARGET
#define RTB_GROUP_TARGET 32
#endif
    uint32_t n_groups = (p->nb_light_sample + RTB_GROUP_TARGET - 1) / RTB_GROUP_TARGET;
    if (n_groups > RTB_MAX_GROUPS) n_groups = RTB_MAX_GROUPS;
    if (n_groups == 0) n_groups = 1;
    uint32_t group_size = (p->nb_light_sample + n_groups - 1) / n_groups;
    if (group_size == 0) group_size = 1;          // nb_light_sample == 0: no shadow rays at all
    const size_t perm_need = (size_t)p->nb_ray * p->nb_light_sample;
    const bool use_perm = perm_need > 0 && p->nb_light_sample <= 2048;
    if (use_perm && perm_need > s->light_perm_cap) {
        cudaFree(s->d_light_perm); s->d_light_perm = nullptr; s->light_perm_cap = 0;
        CUDA_TRY(cudaMalloc(&s->d_light_perm, perm_need * sizeof(uint16_t)));
        s->light_perm_cap = perm_need;
    }
    const size_t gbox_need = (size_t)p->nb_ray * n_groups * 6;
    if (gbox_need > s->light_gbox_cap) {
        cudaFree(s->d_light_gbox); s->d_light_gbox = nullptr; s->light_gbox_cap = 0;
        CUDA_TRY(cudaMalloc(&s->d_light_gbox, gbox_need * sizeof(float)));
        s->light_gbox_cap = gbox_need;
    }
    if (!lights_cached) {
        CUDA_TRY(rtb_launch_light_groups(s->d_light_pts, s->d_light_box, p->nb_ray, p->nb_light_sample, s->n_samples, n_groups, group_size,
                                         use_perm ? s->d_light_perm : nullptr, s->d_light_gbox, st));
        s->light_key_set = true; s->light_key_samples = s->samples_version; s->light_key_need = need; s->light_key_nb_ray = p->nb_ray;
        s->light_key_nls = p->nb_light_sample; s->light_key_stream = st;
    }
    // light grids: one per light point the frame addresses; cached until the table or that count changes
    if (s->shadow_mode != 1 && p->nb_ray && p->nb_light_sample &&
        !(s->grid.key_set && s->grid.key_samples == s->samples_version && s->grid.key_n_lights == (uint32_t)need)) {
        const uint32_t w = p->world ? p->world : 1;
        const uint64_t share_rays = (uint64_t)rtb_owned_rows(p->height, w > 1 ? p->band_rows : p->height, w > 1 ? p->rank : 0, w) * p->width * p->nb_ray;
        s->grid.sort_at_build = share_rays >= (48ull << 20);        // the ordering pass repays itself within this frame (rtb_grid.cu, pass E)
        CUDA_TRY(rtb_grid_build(s->grid, s->d_tris, s->n_tris, s->d_light_pts, (uint32_t)need, s->grid_cell_factor, allow_spec, st));
        s->grid_rebuilds++;
        s->grid.key_set = true; s->grid.key_samples = s->samples_version; s->grid.key_n_lights = (uint32_t)need;
        s->grid.uses = 0; s->grid.refined = false;
    } else if (s->shadow_mode != 1 && p->nb_ray && p->nb_light_sample) {
        // The table is being reused.  A build is sized for a caller that draws a new table per frame (the reference, src/main.rs:
        // 260-265): cells of 0.25 mean primitive extents, lists in two classes -- rebuilding costs about a millisecond per frame.  Smaller
        // cells and fully ordered lists make every later frame faster (4K / 16 spp: 15.6 -> 14.5 ms at 0.1 extents, distance 576:
        // 38.0 -> 33.6 ms) but cost 3 - 6 times the build: the second frame of a table orders the lists, the third rebuilds finer.
        s->grid.uses++;
        if (s->grid.uses == 2 && !s->grid.refined && !s->no_refine && s->grid.valid) {
            s->grid.sort_at_build = true; s->grid.generous = true;
            CUDA_TRY(rtb_grid_build(s->grid, s->d_tris, s->n_tris, s->d_light_pts, (uint32_t)need, s->grid_cell_factor * 0.4f, false, st));
            s->grid.generous = false;
            s->grid_rebuilds++;
            s->grid.refined = true;
            if (!s->grid.valid) {                  // the finer grids did not fit: back to the build the first frames used
                s->grid.sort_at_build = true;
                CUDA_TRY(rtb_grid_build(s->grid, s->d_tris, s->n_tris, s->d_light_pts, (uint32_t)need, s->grid_cell_factor, false, st));
                s->grid_rebuilds++;
            }
        } else {
            CUDA_TRY(rtb_grid_sort_on_reuse(s->grid, st));
        }
    }
    ti->need = need; ti->n_groups = n_groups; ti->group_size = group_size; ti->use_perm = use_perm;
    return RTB_OK;
}

// rtb_scene_set_samples_async while a frame begun with rtb_render_begin is in flight: the table goes into the spare set, and the
// light tables and light grids of the in-flight frame's parameters are built from it on build_stream, next to that frame's
// kernels (which read the current set only).  The next rtb_render_begin swaps the sets and waits for ev_built.
static int prefetch_samples(rtb_scene* s, const float* samples, uint32_t n_pairs) {
    if (!s->alt_inited) {
        rtb_scene::TableSet& t = s->alt;
        CUDA_TRY(cudaStreamCreateWithFlags(&s->build_stream, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&s->ev_built, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&t.ev_prefix, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&t.ev_samples, cudaEventDisableTiming));
        CUDA_TRY(cudaMalloc(&t.d_light_box, 8 * sizeof(float)));
        t.grid.entry_cap_limit = s->grid.entry_cap_limit;
        t.grid.y_min = s->grid.y_min; t.grid.mean_extent = s->grid.mean_extent; t.grid.stats_ok = s->grid.stats_ok;
        s->alt_inited = true;
    }
    s->alt_ready = false;
    swap_table_sets(s);                           // the members below are the SPARE set now (the frame in flight took its pointers by value)
    int rc = RTB_OK;
    cudaError_t e = cudaSuccess;
    do {
        if (n_pairs != s->n_samples) {
            cudaFree(s->d_samples); s->d_samples = nullptr; s->n_samples = 0;
            if ((e = cudaMalloc(&s->d_samples, (size_t)n_pairs * sizeof(float2))) != cudaSuccess) break;
            s->n_samples = n_pairs;
        }
        // whatever was enqueued for this set before (an earlier prefetch nobody rendered) is ahead of us on build_stream; the tail
        // copy on copy_stream starts behind it too
        if ((e = cudaEventRecord(s->ev_idle, s->build_stream)) != cudaSuccess) break;
        if ((e = cudaStreamWaitEvent(s->copy_stream, s->ev_idle, 0)) != cudaSuccess) break;
        const uint32_t prefix = n_pairs < RTB_SAMPLES_PREFIX ? n_pairs : RTB_SAMPLES_PREFIX;
        if ((e = cudaMemcpyAsync(s->d_samples, samples, (size_t)prefix * sizeof(float2), cudaMemcpyHostToDevice, s->build_stream)) != cudaSuccess) break;
        if ((e = cudaEventRecord(s->ev_prefix, s->build_stream)) != cudaSuccess) break;
        if (n_pairs > prefix &&
            (e = cudaMemcpyAsync(s->d_samples + prefix, samples + 2 * (size_t)prefix, (size_t)(n_pairs - prefix) * sizeof(float2),
                                 cudaMemcpyHostToDevice, s->copy_stream)) != cudaSuccess) break;
        if ((e = cudaEventRecord(s->ev_samples, s->copy_stream)) != cudaSuccess) break;
        s->samples_pending = true; s->samples_prefix = prefix;
        s->samples_version = ++s->version_counter;
        const uint32_t before = s->grid_rebuilds;
        s->grid_rebuilds = 0;
        TableInfo ti;
        rc = build_tables(s, &s->pend.p, s->build_stream, !getenv("RTB200_NO_SPEC"), &ti);
        s->prefetch_rebuilds = s->grid_rebuilds; s->grid_rebuilds = before;
        if (rc != RTB_OK) break;
        s->light_key_stream = s->stream;          // the frame that uses the set waits for ev_built on its own stream: same ordering
        e = cudaEventRecord(s->ev_built, s->build_stream);
    } while (0);
    swap_table_sets(s);                           // back: the current set is the in-flight frame's again
    if (rc != RTB_OK) return rc;
    if (e != cudaSuccess) return fail(RTB_ERR_CUDA, "rtb_scene_set_samples_async (prefetch): %s", cudaGetErrorString(e));
    s->alt_ready = true;
    return RTB_OK;
}

// Validates params, (re)builds the light-point table on `st`, fills TraceArgs.
static int prepare_args(rtb_scene* s, const rtb_render_params* p, cudaStream_t st, TraceArgs* a, bool allow_spec = false, bool take_prefetched = true) {
    if (!s || !p) return fail(RTB_ERR_ARG, "render: null argument");
    if (!s->d_samples) return fail(RTB_ERR_STATE, "render: rtb_scene_set_samples has not been called");
    if (s->pend.active) return fail(RTB_ERR_STATE, "render: a frame begun with rtb_render_begin is in flight (rtb_render_end first)");
    if (p->width == 0 || p->height == 0) return fail(RTB_ERR_ARG, "render: empty frame");
    const uint32_t world = p->world ? p->world : 1;
    if (world > 1 && (p->band_rows == 0 || p->rank >= world)) return fail(RTB_ERR_ARG, "render: bad band ownership");
    CUDA_TRY(cudaSetDevice(s->device));
    s->grid_rebuilds = 0;
    // a table prefetched during the previous frame (rtb_scene_set_samples_async between rtb_render_begin and rtb_render_end): its set
    // becomes the current one for whichever render call comes next.  (Not for the redo of a frame: that renders the SAME table again.)
    if (take_prefetched && s->alt_ready) {
        swap_table_sets(s);
        s->alt_ready = false;
        CUDA_TRY(cudaStreamWaitEvent(st, s->ev_built, 0));
        s->grid_rebuilds += s->prefetch_rebuilds; s->prefetch_rebuilds = 0;
    }
    TableInfo ti;
    {
        const int rc = build_tables(s, p, st, allow_spec, &ti);
        if (rc != RTB_OK) return rc;
    }
    const uint64_t need = ti.need;
    const uint32_t n_groups = ti.n_groups, group_size = ti.group_size;
    const bool use_perm = ti.use_perm;
    CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, RTB_NUM_CTRS * sizeof(unsigned long long), st));
    memset(a, 0, sizeof(*a));
    a->tris = s->d_tris; a->colors = s->d_colors4; a->ksns = s->d_ksns; a->nodes = s->d_nodes; a->samples = s->d_samples; a->light_pts = s->d_light_pts; a->light_pairs = s->d_light_pts + need; a->light_box = s->d_light_box;
    a->light_perm = use_perm ? s->d_light_perm : nullptr;
    a->light_gbox = s->d_light_gbox; a->n_groups = n_groups; a->group_size = group_size;
    if (s->shadow_mode != 1 && s->grid.valid && s->grid.key_set && s->grid.key_samples == s->samples_version && s->grid.key_n_lights == (uint32_t)need) {
        a->grid_lights = s->grid.d_lights; a->grid_cell_start = s->grid.d_cell_start; a->grid_entries = s->grid.d_entries;
        a->grid_big = s->grid.d_big; a->grid_n_big = s->grid.n_big; a->grid_entry_mask = s->grid.keyed ? 0xffffff : -1;
        a->grid_union = s->grid.d_union; a->grid_sat = s->grid.d_sat;
    }
    a->n_tris = s->n_tris; a->n_samples = s->n_samples; a->n_light_pts = (uint32_t)need; a->root = s->root;
    a->width = p->width; a->height = p->height; a->nb_ray = p->nb_ray; a->nb_light_sample = p->nb_light_sample;
    a->world = world; a->rank = world > 1 ? p->rank : 0; a->band_rows = world > 1 ? p->band_rows : p->height;
    a->owned_rows = rtb_owned_rows(p->height, a->band_rows, a->rank, world);
    a->cam = make_camera(p);
    // camera grid: cached until the camera changes (a full frame only; a pixel list is too small to pay for the build)
    if (s->primary_mode != 1 && p->nb_ray && s->n_tris) {
        if (!(s->cam_key_set && memcmp(&s->cam_key, &a->cam, sizeof(CameraBasis)) == 0)) {
            if (!s->d_cam_tris) CUDA_TRY(cudaMalloc(&s->d_cam_tris, (size_t)s->n_tris * 3 * sizeof(float4)));
            if (!s->d_cam_origin) CUDA_TRY(cudaMalloc(&s->d_cam_origin, sizeof(float4)));
            CUDA_TRY(rtb_cam_grid_build(s->cam_grid, s->d_tris, s->n_tris, s->d_cam_tris, s->d_cam_origin, a->cam, s->cam_cell_px, allow_spec, st));
            s->grid_rebuilds++;
            s->cam_key = a->cam; s->cam_key_set = true;
        } else {
            CUDA_TRY(rtb_grid_sort_on_reuse(s->cam_grid, st));
        }
        if (s->cam_grid.valid) {
            a->cam_grid = s->cam_grid.d_lights; a->cam_cell_start = s->cam_grid.d_cell_start; a->cam_entries = s->cam_grid.d_entries;
            a->cam_big = s->cam_grid.d_big; a->cam_n_big = s->cam_grid.n_big; a->cam_entry_mask = s->cam_grid.keyed ? 0xffffff : -1;
        }
    }
    a->denom = (float)(p->nb_ray * p->nb_light_sample);
    if (a->denom > 0.0f && !s->cdiv_off) {
        a->cdiv_hi = 1.0f / a->denom;
        a->cdiv_lo = (float)(1.0 / (double)a->denom - (double)a->cdiv_hi);
        if (!s->d_cdiv_bad) CUDA_TRY(cudaMalloc(&s->d_cdiv_bad, sizeof(uint32_t)));
        if (!(s->cdiv_key_set && s->cdiv_denom == a->denom && s->cdiv_stream == st)) {    // stream order is the only ordering here
            CUDA_TRY(rtb_launch_cdiv_check(a->denom, a->cdiv_hi, a->cdiv_lo, s->d_cdiv_bad, st));
            s->cdiv_denom = a->denom; s->cdiv_stream = st; s->cdiv_key_set = true;
        }
        a->cdiv_bad = s->d_cdiv_bad;
    }
    a->counters = s->d_counters;
    if (s->samples_pending) CUDA_TRY(cudaStreamWaitEvent(st, s->ev_samples, 0));     // the primary-ray kernel reads the whole table
    return RTB_OK;
}

// After the caller has synchronised: the verdict of speculative grid builds enqueued by prepare_args(allow_spec).  *redo: at
// least one did not fit its allocation -- its grids were switched off on the device, the frame is wrong; the cache keys are
// dropped so that the next prepare_args builds synchronously (with room) and the caller renders again.
static int validate_grids(rtb_scene* s, bool* redo) {
    *redo = false;
    bool ok = true;
    s->grid.fill_async = false;                  // the caller has synchronised: the fill pass is long done
    CUDA_TRY(rtb_grid_validate(s->grid, s->stream, &ok));
    if (!ok) { s->grid.key_set = false; *redo = true; }
    CUDA_TRY(rtb_grid_validate(s->cam_grid, s->stream, &ok));
    if (!ok) { s->cam_key_set = false; *redo = true; }
    return RTB_OK;
}

// Copies local rows [r0, r1) of a rank's compact device rows to the host.  frame_rows: h is the FULL frame and every band goes
// to its own rows (band lb of this rank = frame band lb * world + rank): whole bands as one strided 2-D copy.
static cudaError_t copy_rows_to_host(uint8_t* h, const uint8_t* d, uint64_t r0, uint64_t r1, const TraceArgs& a, bool frame_rows, cudaStream_t st) {
    const size_t rb = (size_t)a.width * 3;
    if (r1 <= r0) return cudaSuccess;
    if (!frame_rows || a.world <= 1) return cudaMemcpyAsync(h + r0 * rb, d + r0 * rb, (r1 - r0) * rb, cudaMemcpyDeviceToHost, st);
    const uint64_t B = a.band_rows;
    uint64_t lr = r0;
    while (lr < r1) {
        const uint64_t lb = lr / B, within = lr - lb * B;
        const uint64_t grow = (lb * a.world + a.rank) * B + within;          // frame row of local row lr
        cudaError_t e;
        if (within == 0 && r1 - lr >= B) {
            const uint64_t nb = (r1 - lr) / B;
            e = cudaMemcpy2DAsync(h + grow * rb, (size_t)a.world * B * rb, d + lr * rb, B * rb, B * rb, nb, cudaMemcpyDeviceToHost, st);
            lr += nb * B;
        } else {
            const uint64_t n = (B - within) < (r1 - lr) ? (B - within) : (r1 - lr);
            e = cudaMemcpyAsync(h + grow * rb, d + lr * rb, n * rb, cudaMemcpyDeviceToHost, st);
            lr += n;
        }
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

#define RTB_MAX_TIMED_CHUNKS 32

// Grows (never shrinks) the wavefront buffers of the scene to hold one chunk of `rays` rays: worst case every ray hits,
// is queued and keeps all its light-sample groups.
static int reserve_wavefront(rtb_scene::WaveBufs& b, uint64_t rays, uint32_t n_groups, uint32_t mask_words, bool lists,
                             uint64_t term_slots, uint64_t term_slot_floats) {
    if (term_slots) {                            // term wavefront: slot table per (ray, 32-pixel tile), slot descriptors, term rows
        if (rays / 32 > b.tile_slot_cap) {
            cudaFree(b.d_tile_slot); b.d_tile_slot = nullptr; b.tile_slot_cap = 0;
            CUDA_TRY(cudaMalloc(&b.d_tile_slot, rays / 32 * sizeof(uint32_t)));
            b.tile_slot_cap = rays / 32;
        }
        if (term_slots > b.term_slots) {
            cudaFree(b.d_term_tiles); b.d_term_tiles = nullptr; b.term_slots = 0;
            CUDA_TRY(cudaMalloc(&b.d_term_tiles, term_slots * sizeof(uint4)));
            b.term_slots = term_slots;
        }
        if (term_slots * term_slot_floats > b.term_floats) {
            cudaFree(b.d_terms); b.d_terms = nullptr; b.term_floats = 0;
            CUDA_TRY(cudaMalloc(&b.d_terms, term_slots * term_slot_floats * sizeof(float)));
            b.term_floats = term_slots * term_slot_floats;
        }
    }
    if (rays > b.rays_cap) {
        cudaFree(b.d_records); cudaFree(b.d_items); b.d_records = nullptr; b.d_items = nullptr; b.rays_cap = 0;
        CUDA_TRY(cudaMalloc(&b.d_records, rays * sizeof(uint2)));
        CUDA_TRY(cudaMalloc(&b.d_items, rays * RTB_ITEM_F4 * sizeof(float4)));
        b.rays_cap = rays;
    }
    const uint64_t sub_need = lists ? rays * n_groups : 0;     // pair queues and candidate pool: list / walk path only
    if (sub_need > b.sub_cap) {
        cudaFree(b.d_subitems); cudaFree(b.d_sublists); b.d_subitems = nullptr; b.d_sublists = nullptr; b.sub_cap = 0;
        CUDA_TRY(cudaMalloc(&b.d_subitems, (sub_need ? sub_need : 1) * sizeof(uint4)));
        CUDA_TRY(cudaMalloc(&b.d_sublists, (sub_need ? sub_need : 1) * sizeof(uint4)));
        b.sub_cap = sub_need;
    }
    // candidate pool: 16 entries per ray on average is far above anything measured (big_bunny: 2.4); a full pool is not
    // an error, the remaining pairs walk the BVH instead
    uint64_t pool_need = lists ? rays * 16 : 0;
    if (pool_need > 0xfffffff0ull) pool_need = 0xfffffff0ull;
    if (pool_need > b.pool_cap) {
        cudaFree(b.d_cand_pool); b.d_cand_pool = nullptr; b.pool_cap = 0;
        CUDA_TRY(cudaMalloc(&b.d_cand_pool, (pool_need ? pool_need : 1) * sizeof(int)));
        b.pool_cap = pool_need;
    }
    const uint64_t mask_need = rays * mask_words;
    if (mask_need > b.mask_cap) {
        cudaFree(b.d_masks); b.d_masks = nullptr; b.mask_cap = 0;
        CUDA_TRY(cudaMalloc(&b.d_masks, (mask_need ? mask_need : 1) * sizeof(uint32_t)));
        b.mask_cap = mask_need;
    }
    return RTB_OK;
}

// Runs the wavefront kernels over the whole frame (a.pixel_list == nullptr: 8x4 pixel tiles of the owned rows) or pixel
// list, in chunks of at most s->chunk_rays rays so that the intermediate buffers stay bounded (about 170 bytes per ray
// in the worst case) whatever the resolution and rays per pixel.  A chunk is a whole number of tile rows (or of 256
// listed pixels).  timed: record the per-kernel events of the first RTB_MAX_TIMED_CHUNKS chunks.
static int run_wavefront(rtb_scene* s, TraceArgs a, bool count, cudaStream_t st, bool timed, uint8_t* h_rows = nullptr, bool h_frame_rows = false) {
    s->timed_chunks = 0; s->n_chunks = 0; s->terms_ran = false;
    uint64_t unit, units_total;                       // pixels per granule, granules
    if (a.pixel_list) {
        unit = 256; units_total = (a.n_pix + 255) / 256; a.tiles_x = 0;
    } else {
        a.tiles_x = (a.width + 7) / 8;
        unit = (uint64_t)a.tiles_x * 32; units_total = (a.owned_rows + 3) / 4;
    }
    if (units_total == 0) return RTB_OK;
    const uint64_t per_pixel = a.nb_ray ? a.nb_ray : 1;
    uint64_t units_per_chunk = s->chunk_rays / (unit * per_pixel);
    if (units_per_chunk == 0) units_per_chunk = 1;
    if (units_per_chunk > units_total) units_per_chunk = units_total;
    // Two chunks overlap badly on the two worker streams (a band share at 4 GPUs: 4.2 - 4.9 ms from run to run) and pay the wave
    // quantisation of every kernel twice; one chunk of up to twice the size, with the term wavefront, is steadier and faster
    // (4.25 ms).  Three or more chunks keep the bounded-memory form.
    if (units_total > units_per_chunk && units_total <= 2 * units_per_chunk && units_total * unit * per_pixel < 0x70000000ull)
        units_per_chunk = units_total;
    const uint64_t npp_max = (units_per_chunk * unit + 255) / 256 * 256;
    if (npp_max * per_pixel >= 0x7fffffffull)
        return fail(RTB_ERR_LIMIT, "render: one tile row needs %llu rays, above the 2^31 record index", (unsigned long long)(npp_max * per_pixel));
    a.mask_words = (a.nb_light_sample + 31) / 32;
    a.magic_tiles_x = rtb_magic(a.tiles_x); a.magic_band_rows = rtb_magic(a.band_rows); a.magic_n_samples = rtb_magic(a.n_samples);
    const uint64_t n_chunks = (units_total + units_per_chunk - 1) / units_per_chunk;
    // timed (per-kernel events) or a single chunk: everything on the caller's stream with buffer set 0.  Otherwise chunk c
    // runs on worker stream c & 1 with buffer set c & 1; the workers start after the caller's stream and it waits for them.
    const bool overlap = !timed && n_chunks > 1;
    // term wavefront (k_shade_terms): as many (tile, ray) slots as fit the budget, at most one per warp of k_primary
    a.term_channels = s->all_grey ? 1u : 3u;
    a.term_stride = (a.nb_light_sample + 3u) & ~3u;
    const uint64_t term_slot_floats = (uint64_t)a.term_channels * 32 * a.term_stride;
    uint64_t term_slots = 0;
    // Only where nothing else hides the stragglers: with several chunks the other worker stream fills the SMs a few long-running
    // bunny warps leave idle, and the term kernel is then extra work (4K/16spp on one GPU: 17.7 ms without, 17.9 ms with); a
    // single-chunk render -- a band share on 8 GPUs, a small frame, a pixel list -- ends on those warps (k_shade 1.15 -> 0.94 ms).
    const bool use_terms = (s->term_mode == 1 || (s->term_mode == 0 && n_chunks == 1)) && !a.ksns;     // the specular extension lives in k_shade's general loop only
    if (a.nb_ray && a.nb_light_sample && s->term_bytes && use_terms) {
        term_slots = s->term_bytes / (term_slot_floats * sizeof(float));
        if (term_slots > npp_max * per_pixel / 32) term_slots = npp_max * per_pixel / 32;
    }
    for (int k = 0; k < (overlap ? 2 : 1); ++k) {
        const int rc = reserve_wavefront(s->wb[k], npp_max * per_pixel, a.n_groups, a.mask_words, a.grid_lights == nullptr, term_slots, term_slot_floats);
        if (rc != RTB_OK) return rc;
    }
    if (overlap) {
        CUDA_TRY(cudaEventRecord(s->ev_fork, st));
        for (int k = 0; k < 2; ++k) CUDA_TRY(cudaStreamWaitEvent(s->wstream[k], s->ev_fork, 0));
    }
    const uint64_t pix_total = a.pixel_list ? a.n_pix : units_total * unit;
    for (uint64_t u0 = 0; u0 < units_total; u0 += units_per_chunk) {
        const int k = overlap ? (int)(s->n_chunks & 1) : 0;
        const rtb_scene::WaveBufs& b = s->wb[k];
        cudaStream_t cs = overlap ? s->wstream[k] : st;
        a.records = b.d_records; a.items = b.d_items; a.masks = b.d_masks; a.queue_ctl = b.d_queue_ctl; a.subitems = b.d_subitems;
        a.sublists = b.d_sublists; a.cand_pool = b.d_cand_pool; a.pool_cap = (uint32_t)b.pool_cap;
        a.queue_ticket = (unsigned long long*)(b.d_queue_ctl + 2);
        a.tile_slot = b.d_tile_slot; a.term_tiles = b.d_term_tiles; a.terms = term_slots ? b.d_terms : nullptr; a.term_cap = (uint32_t)term_slots;
        s->terms_ran = term_slots != 0;
        const uint64_t nu = units_total - u0 < units_per_chunk ? units_total - u0 : units_per_chunk;
        const uint64_t base = u0 * unit;
        uint64_t n = nu * unit;
        if (base + n > pix_total) n = pix_total - base;
        a.pix_base = (uint32_t)base; a.n_pix_chunk = (uint32_t)n; a.n_pix_padded = (uint32_t)((nu * unit + 255) / 256 * 256);
        // k_primary: two warps per 32-pixel tile, half of the pixel's rays each, where nothing else hides the kernel's tail -- a
        // single-chunk render (band shares of 1 / 2 / 4 / 8 GPUs of the 4K / 16 spp frame: k_primary 1.84 / 0.98 / 0.42 / 0.31 ->
        // 1.70 / 0.90 / 0.42 / 0.22 ms, but with several chunks on two streams the frame does not gain: 14.54 -> 14.6 ms; four warps:
        // no further gain)
        a.primary_split = (n_chunks == 1 && a.nb_ray >= 2) ? 2u : 1u;
        CUDA_TRY(cudaMemsetAsync(b.d_queue_ctl, 0, 16 * sizeof(uint32_t), cs));
        cudaEvent_t* ev = nullptr;
        if (timed && s->timed_chunks < RTB_MAX_TIMED_CHUNKS) ev = s->evp + 4 * s->timed_chunks++;
        CUDA_TRY(rtb_launch_trace(a, count, cs, ev, s->grid.fill_async ? s->grid.ev_fill : nullptr));
        if (h_rows && !a.pixel_list && !a.frame_rows) {
            // pinned host frame (rtb_render): the rows of this chunk leave right behind its k_shade, on the chunk's own stream,
            // so the device-to-host copy of one chunk overlaps the kernels of the next
            const uint64_t r0 = u0 * 4 < a.owned_rows ? u0 * 4 : a.owned_rows, r1 = (u0 + nu) * 4 < a.owned_rows ? (u0 + nu) * 4 : a.owned_rows;
            CUDA_TRY(copy_rows_to_host(h_rows, a.rgb8, r0, r1, a, h_frame_rows, cs));
        }
        s->n_chunks++;
    }
    if (overlap) {
        for (int k = 0; k < 2; ++k) {
            CUDA_TRY(cudaEventRecord(s->ev_join[k], s->wstream[k]));
            CUDA_TRY(cudaStreamWaitEvent(st, s->ev_join[k], 0));
        }
    }
    return RTB_OK;
}

static int fetch_stats(rtb_scene* s, const rtb_render_params* p, uint64_t n_pix, float ms, uint32_t launches, rtb_stats* st) {
    float k_ms[3] = {0.0f, 0.0f, 0.0f};
    unsigned long long h[RTB_NUM_CTRS];
    CUDA_TRY(cudaMemcpyAsync(h, s->d_counters, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    if (h[RTB_CTR_OVERFLOW]) return fail(RTB_ERR_LIMIT, "traversal stack overflow");
    for (uint32_t c = 0; c < s->timed_chunks; ++c) {
        for (int k = 0; k < 3; ++k) {
            float t = 0.0f;
            CUDA_TRY(cudaEventElapsedTime(&t, s->evp[4 * c + k], s->evp[4 * c + k + 1]));
            k_ms[k] += t;
            if (s->trace_chunks) fprintf(stderr, "rtb200: chunk %u %s %.3f ms\n", c, k == 0 ? "primary" : k == 1 ? "shadow" : "shade", t);
        }
    }
    if (st) {
        st->primary_rays = n_pix * p->nb_ray; st->primary_hits = h[RTB_CTR_PRIMARY_HITS];
        st->shadow_rays = h[RTB_CTR_SHADOW]; st->shadow_occluded = h[RTB_CTR_OCCLUDED];
        st->box_tests = h[RTB_CTR_BOX]; st->tri_tests = h[RTB_CTR_TRI];
        st->kernel_ms = ms;
        const bool grid = s->shadow_mode != 1 && s->grid.valid;
        st->launches = (s->light_tables_ran ? 4 : 0) + s->n_chunks * (1 + (p->nb_ray ? 1 : 0) + (p->nb_ray && p->nb_light_sample ? (grid ? 1 : 3) + (s->terms_ran ? 1 : 0) : 0));   // light table, pairs, bounds, groups + per chunk
        st->grid_cells = grid ? s->grid.total_cells : 0; st->grid_entries = grid ? s->grid.total_entries : 0;
        st->grid_big = grid ? s->grid.n_big : 0; st->grid_build_ms = grid ? s->grid.build_ms : 0.0f;
        st->primary_ms = k_ms[0]; st->shadow_ms = k_ms[1]; st->shade_ms = k_ms[2];
        st->queued_hits = (uint32_t)h[RTB_CTR_QUEUED]; st->queued_groups = (uint32_t)h[RTB_CTR_PAIRS];
        st->queued_walks = (uint32_t)h[RTB_CTR_WALKS];
        st->grid_rebuilds = s->grid_rebuilds; st->grid_redo = s->grid_redo;
    }
    return RTB_OK;
}

extern "C" {

int rtb_render_device(rtb_scene* s, const rtb_render_params* p, void* d_rgb8, void* cuda_stream) {
    TraceArgs a;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    int rc = prepare_args(s, p, st, &a);
    if (rc != RTB_OK) return rc;
    if (!d_rgb8) return fail(RTB_ERR_ARG, "rtb_render_device: null output");
    a.rgb8 = (uint8_t*)d_rgb8;
    a.frame_rows = (p->flags & RTB_FLAG_FRAME_ROWS) ? 1u : 0u;
    return run_wavefront(s, a, (p->flags & RTB_FLAG_COUNTERS) != 0, st, false);
}

int rtb_peer_alloc(int32_t device, uint64_t bytes, void** d_ptr_out, uint8_t handle_out[64]) {
    if (!d_ptr_out || !handle_out || bytes == 0) return fail(RTB_ERR_ARG, "rtb_peer_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    CUDA_TRY(cudaSetDevice(device));
    void* p = nullptr;
    CUDA_TRY(cudaMalloc(&p, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return fail(RTB_ERR_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); }
    memcpy(handle_out, &h, 64);
    *d_ptr_out = p;
    return RTB_OK;
}

int rtb_peer_open(int32_t device, const uint8_t handle[64], void** d_ptr_out) {
    if (!d_ptr_out || !handle) return fail(RTB_ERR_ARG, "rtb_peer_open: bad argument");
    CUDA_TRY(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void* p = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *d_ptr_out = p;
    return RTB_OK;
}

int rtb_peer_close(int32_t device, void* d_ptr, int32_t owner) {
    if (!d_ptr) return RTB_OK;
    CUDA_TRY(cudaSetDevice(device));
    if (owner) CUDA_TRY(cudaFree(d_ptr)); else CUDA_TRY(cudaIpcCloseMemHandle(d_ptr));
    return RTB_OK;
}

// render(): the frame (or this rank's bands) into HOST memory.  h_rgb8: compact rows, or the full frame with RTB_FLAG_FRAME_ROWS.
// d_full != nullptr (rtb_render_multi with RTB_FLAG_GATHER_PEER): the rows go to their place in that device frame (possibly a
// peer's) instead and nothing is copied to the host here.
// The light grids / camera grid a new table / camera needs are rebuilt speculatively (rtb_grid.cu: no host synchronisation
// between the build's passes and the frame's kernels); after the frame's one synchronisation the builds are validated and, should
// one not have fitted its allocation, rebuilt synchronously and the frame rendered once more.
// render_enqueue: everything of one attempt up to (not including) the host's wait; render_finish: the wait, the verdict of the
// speculative builds (*redo) and, for a pageable full-frame destination, the placement of the staged rows.
static int render_enqueue(rtb_scene* s, const rtb_render_params* p, uint8_t* h_rgb8, uint8_t* d_full, bool first, rtb_scene::Pending* pr) {
    const bool allow_spec = first && !getenv("RTB200_NO_SPEC");
    pr->p = *p; pr->h_rgb8 = h_rgb8; pr->d_full = d_full; pr->stage = nullptr; pr->rebuilds = 0;
    TraceArgs& a = pr->a;
    int rc = prepare_args(s, p, s->stream, &a, allow_spec, first);
    if (rc != RTB_OK) return rc;
    pr->rebuilds += s->grid_rebuilds;
    const size_t bytes = (size_t)a.owned_rows * a.width * 3;
    const bool h_frame_rows = !d_full && (p->flags & RTB_FLAG_FRAME_ROWS) != 0 && a.world > 1;
    pr->h_frame_rows = h_frame_rows;
    if (!d_full && bytes > s->frame_cap) {
        cudaFree(s->d_frame); s->d_frame = nullptr; s->frame_cap = 0;
        CUDA_TRY(cudaMalloc(&s->d_frame, bytes));
        s->frame_cap = bytes;
    }
    a.rgb8 = d_full ? d_full : s->d_frame;
    a.frame_rows = d_full ? 1u : 0u;
    CUDA_TRY(cudaEventRecord(s->ev0, s->stream));
    // a pinned destination takes per-chunk copies (an asynchronous copy to pageable memory would block the host chunk by chunk)
    bool pinned = false;
    if (h_rgb8) {
        cudaPointerAttributes pa;
        pinned = cudaPointerGetAttributes(&pa, h_rgb8) == cudaSuccess && pa.type == cudaMemoryTypeHost && !(p->flags & RTB_FLAG_PROFILE);
        if (!pinned) cudaGetLastError();
    }
    rc = run_wavefront(s, a, (p->flags & RTB_FLAG_COUNTERS) != 0, s->stream, (p->flags & RTB_FLAG_PROFILE) != 0, pinned ? h_rgb8 : nullptr, h_frame_rows);
    if (rc != RTB_OK) return rc;
    if (s->grid.fill_async) CUDA_TRY(cudaStreamWaitEvent(s->stream, s->grid.ev_fill, 0));    // (frames without shadow kernels never waited for it)
    CUDA_TRY(cudaEventRecord(s->ev1, s->stream));
    if (h_rgb8 && bytes && !pinned) {
        if (h_frame_rows) {                    // pageable full frame: compact rows through the scene's pinned buffer, placed by the host in render_finish
            if (bytes > s->pinned_cap) {
                if (s->h_pinned) cudaFreeHost(s->h_pinned);
                s->h_pinned = nullptr; s->pinned_cap = 0;
                CUDA_TRY(cudaMallocHost(&s->h_pinned, bytes));
                s->pinned_cap = bytes;
            }
            pr->stage = s->h_pinned;
            CUDA_TRY(cudaMemcpyAsync(pr->stage, s->d_frame, bytes, cudaMemcpyDeviceToHost, s->stream));
        } else {
            CUDA_TRY(cudaMemcpyAsync(h_rgb8, s->d_frame, bytes, cudaMemcpyDeviceToHost, s->stream));
        }
    }
    return RTB_OK;
}

static int render_finish(rtb_scene* s, rtb_scene::Pending* pr, bool* redo) {
    const TraceArgs& a = pr->a;
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->samples_pending = false;
    const int rc = validate_grids(s, redo);
    if (rc != RTB_OK) return rc;
    if (*redo) return RTB_OK;
    if (pr->stage) {
        const size_t rb = (size_t)a.width * 3;
        for (uint32_t lr = 0; lr < a.owned_rows; ++lr) {
            const uint32_t lb = lr / a.band_rows, within = lr - lb * a.band_rows;
            memcpy(pr->h_rgb8 + (size_t)((lb * a.world + a.rank) * a.band_rows + within) * rb, pr->stage + (size_t)lr * rb, rb);
        }
    }
    return RTB_OK;
}

static int render_core(rtb_scene* s, const rtb_render_params* p, uint8_t* h_rgb8, uint8_t* d_full, rtb_stats* stats, int first_attempt = 0) {
    if (!s || !p) return fail(RTB_ERR_ARG, "render: null argument");
    if (s->pend.active) return fail(RTB_ERR_STATE, "render: a frame begun with rtb_render_begin is in flight (rtb_render_end first)");
    if (first_attempt == 0) s->grid_redo = 0;
    uint32_t rebuilds = 0;
    for (int attempt = first_attempt; ; ++attempt) {
        rtb_scene::Pending pr;
        int rc = render_enqueue(s, p, h_rgb8, d_full, attempt == 0, &pr);
        if (rc != RTB_OK) return rc;
        rebuilds += pr.rebuilds;
        bool redo = false;
        rc = render_finish(s, &pr, &redo);
        if (rc != RTB_OK) return rc;
        if (redo && attempt == 0) { s->grid_redo++; continue; }
        if (redo) return fail(RTB_ERR_STATE, "render: grid rebuild failed twice");
        float ms = 0.0f;
        CUDA_TRY(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
        s->grid_rebuilds = rebuilds;
        return fetch_stats(s, p, (uint64_t)pr.a.owned_rows * pr.a.width, ms, 8, stats);
    }
}

// rtb_render in two halves.  Between them the caller may hand over the NEXT frame's sample table (rtb_scene_set_samples_async
// with a pinned table): its upload, light tables and light grids then run next to this frame's kernels.
int rtb_render_begin(rtb_scene* s, const rtb_render_params* p, uint8_t* rgb8) {
    if (!s || !p || !rgb8) return fail(RTB_ERR_ARG, "rtb_render_begin: null argument");
    if (s->pend.active) return fail(RTB_ERR_STATE, "rtb_render_begin: a frame is already in flight");
    cudaPointerAttributes pa;
    if (!(cudaPointerGetAttributes(&pa, rgb8) == cudaSuccess && pa.type == cudaMemoryTypeHost)) {
        cudaGetLastError();
        return fail(RTB_ERR_ARG, "rtb_render_begin: the frame must be page-locked host memory (the copy is asynchronous)");
    }
    if (p->flags & RTB_FLAG_PROFILE) return fail(RTB_ERR_ARG, "rtb_render_begin: RTB_FLAG_PROFILE needs the blocking rtb_render");
    s->grid_redo = 0;
    const int rc = render_enqueue(s, p, rgb8, nullptr, true, &s->pend);
    if (rc != RTB_OK) { cudaStreamSynchronize(s->stream); return rc; }
    s->pend.active = true;
    return RTB_OK;
}

int rtb_render_end(rtb_scene* s, rtb_stats* stats) {
    if (!s) return fail(RTB_ERR_ARG, "rtb_render_end: null argument");
    if (!s->pend.active) return fail(RTB_ERR_STATE, "rtb_render_end: no frame in flight");
    CUDA_TRY(cudaSetDevice(s->device));
    bool redo = false;
    int rc = render_finish(s, &s->pend, &redo);
    s->pend.active = false;
    if (rc != RTB_OK) return rc;
    if (redo) {                                    // a speculative build did not fit: rebuilt with room, the frame rendered again (blocking)
        s->grid_redo = 1;
        const uint32_t before = s->pend.rebuilds;
        rc = render_core(s, &s->pend.p, s->pend.h_rgb8, nullptr, stats, 1);
        if (rc == RTB_OK) { s->grid_rebuilds += before; if (stats) stats->grid_rebuilds = s->grid_rebuilds; }
        return rc;
    }
    float ms = 0.0f;
    CUDA_TRY(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    s->grid_rebuilds = s->pend.rebuilds;
    return fetch_stats(s, &s->pend.p, (uint64_t)s->pend.a.owned_rows * s->pend.a.width, ms, 8, stats);
}

int rtb_render(rtb_scene* s, const rtb_render_params* p, uint8_t* rgb8, rtb_stats* stats) {
    if (!rgb8) return fail(RTB_ERR_ARG, "rtb_render: null output");
    return render_core(s, p, rgb8, nullptr, stats);
}

int rtb_render_pixels(rtb_scene* s, const rtb_render_params* p, const uint32_t* pixels, uint64_t n_pix, float* rgb_lin,
                      uint8_t* rgb8, uint8_t* hit_count, int32_t* prim0, float* t0, rtb_stats* stats) {
    TraceArgs a;
    if (n_pix && !pixels) return fail(RTB_ERR_ARG, "rtb_render_pixels: null pixel list");
    int rc = prepare_args(s, p, s ? s->stream : nullptr, &a);
    if (rc != RTB_OK) return rc;
    uint2* d_pix = nullptr; float* d_lin = nullptr; uint8_t* d_rgb8 = nullptr; uint8_t* d_hc = nullptr; int32_t* d_p0 = nullptr; float* d_t0 = nullptr;
    cudaError_t e = cudaSuccess;
    float ms = 0.0f;
#define TRY_(x) do { e = (x); if (e != cudaSuccess) { rc = fail(RTB_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e)); goto cleanup; } } while (0)
    if (n_pix) {
        TRY_(cudaMalloc(&d_pix, n_pix * sizeof(uint2)));
        TRY_(cudaMemcpyAsync(d_pix, pixels, n_pix * sizeof(uint2), cudaMemcpyHostToDevice, s->stream));
        if (rgb_lin) TRY_(cudaMalloc(&d_lin, n_pix * 12));
        if (rgb8) TRY_(cudaMalloc(&d_rgb8, n_pix * 3));
        if (hit_count) TRY_(cudaMalloc(&d_hc, n_pix));
        if (prim0) TRY_(cudaMalloc(&d_p0, n_pix * 4));
        if (t0) TRY_(cudaMalloc(&d_t0, n_pix * 4));
    }
    a.pixel_list = d_pix; a.n_pix = n_pix; a.out_rgb_lin = d_lin; a.rgb8 = d_rgb8; a.out_hit_count = d_hc; a.out_prim0 = d_p0; a.out_t0 = d_t0;
    s->timed_chunks = 0; s->n_chunks = 0;
    TRY_(cudaEventRecord(s->ev0, s->stream));
    if (n_pix) {
        rc = run_wavefront(s, a, (p->flags & RTB_FLAG_COUNTERS) != 0, s->stream, (p->flags & RTB_FLAG_PROFILE) != 0);
        if (rc != RTB_OK) goto cleanup;
    }
    TRY_(cudaEventRecord(s->ev1, s->stream));
    if (n_pix) {
        if (rgb_lin) TRY_(cudaMemcpyAsync(rgb_lin, d_lin, n_pix * 12, cudaMemcpyDeviceToHost, s->stream));
        if (rgb8) TRY_(cudaMemcpyAsync(rgb8, d_rgb8, n_pix * 3, cudaMemcpyDeviceToHost, s->stream));
        if (hit_count) TRY_(cudaMemcpyAsync(hit_count, d_hc, n_pix, cudaMemcpyDeviceToHost, s->stream));
        if (prim0) TRY_(cudaMemcpyAsync(prim0, d_p0, n_pix * 4, cudaMemcpyDeviceToHost, s->stream));
        if (t0) TRY_(cudaMemcpyAsync(t0, d_t0, n_pix * 4, cudaMemcpyDeviceToHost, s->stream));
    }
    TRY_(cudaStreamSynchronize(s->stream));
    TRY_(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    rc = fetch_stats(s, p, n_pix, ms, n_pix ? 8 : 3, stats);
cleanup:
    cudaFree(d_pix); cudaFree(d_lin); cudaFree(d_rgb8); cudaFree(d_hc); cudaFree(d_p0); cudaFree(d_t0);
    return rc;
#undef TRY_
}

int rtb_color_to_u8(rtb_scene* s, const float* linear, uint64_t n, uint8_t* out) {
    if (!s || (n && (!linear || !out))) return fail(RTB_ERR_ARG, "rtb_color_to_u8: null argument");
    if (n == 0) return RTB_OK;
    CUDA_TRY(cudaSetDevice(s->device));
    float* d_in = nullptr; uint8_t* d_out = nullptr;
    int rc = RTB_OK; cudaError_t e;
#define TRY_(x) do { e = (x); if (e != cudaSuccess) { rc = fail(RTB_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e)); goto cleanup; } } while (0)
    TRY_(cudaMalloc(&d_in, n * 4)); TRY_(cudaMalloc(&d_out, n));
    TRY_(cudaMemcpyAsync(d_in, linear, n * 4, cudaMemcpyHostToDevice, s->stream));
    TRY_(rtb_launch_color_to_u8(d_in, n, d_out, s->stream));
    TRY_(cudaMemcpyAsync(out, d_out, n, cudaMemcpyDeviceToHost, s->stream));
    TRY_(cudaStreamSynchronize(s->stream));
cleanup:
    cudaFree(d_in); cudaFree(d_out);
    return rc;
#undef TRY_
}

// Same result as rtb_color_to_u8 with no allocation of any kind: the staging lives in a static device array (rtb_trace.cu).
int rtb_color_to_u8_static(rtb_scene* s, const float* linear, uint64_t n, uint8_t* out) {
    if (!s || (n && (!linear || !out))) return fail(RTB_ERR_ARG, "rtb_color_to_u8_static: null argument");
    if (n > RTB_COLOR_STATIC_MAX)
        return fail(RTB_ERR_LIMIT, "rtb_color_to_u8_static: %llu values exceed the static scratch of %u values per call (use rtb_color_to_u8, or split the call)",
                    (unsigned long long)n, (unsigned)RTB_COLOR_STATIC_MAX);
    if (n == 0) return RTB_OK;
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(rtb_color_to_u8_static_run(s->device, linear, n, out, s->stream));
    return RTB_OK;
}

int rtb_scene_get_nodes(rtb_scene* s, float* out) {
    if (!s || !out) return fail(RTB_ERR_ARG, "rtb_scene_get_nodes: null argument");
    if (s->info.n_nodes == 0) return RTB_OK;
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaMemcpyAsync(out, s->d_nodes, (size_t)s->info.n_nodes * 64, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return RTB_OK;
}

int rtb_scene_get_tie_ranks(rtb_scene* s, int32_t* out) {
    if (!s || !out) return fail(RTB_ERR_ARG, "rtb_scene_get_tie_ranks: null argument");
    if (s->n_tris == 0) return RTB_OK;
    CUDA_TRY(cudaSetDevice(s->device));
    int* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, (size_t)s->n_tris * 4));
    cudaError_t e = rtb_launch_get_ranks(s->d_colors4, s->n_tris, d, s->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d, (size_t)s->n_tris * 4, cudaMemcpyDeviceToHost, s->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    cudaFree(d);
    if (e != cudaSuccess) return fail(RTB_ERR_CUDA, "rtb_scene_get_tie_ranks: %s", cudaGetErrorString(e));
    return RTB_OK;
}

int rtb_scene_get_counters(rtb_scene* s, uint64_t* out16) {
    if (!s || !out16) return fail(RTB_ERR_ARG, "rtb_scene_get_counters: null argument");
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaMemcpy(out16, s->d_counters, RTB_NUM_CTRS * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return RTB_OK;
}

int rtb_selftest_division(rtb_scene* s, uint64_t seed, uint64_t n, uint64_t* mismatches) {
    if (!s || !mismatches) return fail(RTB_ERR_ARG, "rtb_selftest_division: null argument");
    if (n > 0xffffffffull * 256) return fail(RTB_ERR_ARG, "rtb_selftest_division: n too large");
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, RTB_NUM_CTRS * sizeof(unsigned long long), s->stream));
    CUDA_TRY(rtb_launch_div_selftest(seed, n, s->d_counters, s->stream));
    unsigned long long h = 0;
    CUDA_TRY(cudaMemcpyAsync(&h, s->d_counters, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    *mismatches = h;
    return RTB_OK;
}

int rtb_deinterleave_bands(int32_t device, const void* d_staged, uint64_t stride_bytes, void* d_frame, uint32_t width,
                           uint32_t height, uint32_t band_rows, uint32_t world, void* cuda_stream) {
    if (!d_staged || !d_frame || band_rows == 0 || world == 0) return fail(RTB_ERR_ARG, "rtb_deinterleave_bands: bad argument");
    CUDA_TRY(cudaSetDevice(device));
    CUDA_TRY(rtb_launch_deinterleave((const uint8_t*)d_staged, stride_bytes, (uint8_t*)d_frame, width * 3, height, band_rows, world,
                                     (cudaStream_t)cuda_stream));
    return RTB_OK;
}

}  // extern "C"

// ---- parallel OBJ import into (optionally) pinned memory: the parser is host/rtb_obj_parallel.cpp, the allocator lives here -------
int rtb_obj_parallel_core(const char*, uint32_t, void* (*)(size_t), float**, uint32_t*, rtb_import_stats*, std::string*, float**, float**);
static void* alloc_pinned_(size_t bytes) { void* p = nullptr; return cudaHostAlloc(&p, bytes, cudaHostAllocDefault) == cudaSuccess ? p : nullptr; }
static void* alloc_pageable_(size_t bytes) { return malloc(bytes); }

extern "C" int rtb_import_obj_parallel_ex(const char* path, uint32_t n_threads, uint32_t flags, float** verts_out, float** kd_out, float** ks_ns_out,
                                          uint32_t* n_out, rtb_import_stats* stats) {
    if (!path || !verts_out || !n_out) return fail(RTB_ERR_ARG, "rtb_import_obj_parallel: null argument");
    if (stats) memset(stats, 0, sizeof(*stats));
    if (kd_out) *kd_out = nullptr;
    if (ks_ns_out) *ks_ns_out = nullptr;
    std::string err;
    const bool pinned = (flags & RTB_OBJ_PINNED) != 0, mtl = (flags & RTB_OBJ_MTL) != 0;
    float *kd = nullptr, *ks = nullptr;
    const int rc = rtb_obj_parallel_core(path, n_threads, pinned ? alloc_pinned_ : alloc_pageable_, verts_out, n_out, stats, &err,
                                         mtl && kd_out ? &kd : nullptr, mtl && ks_ns_out ? &ks : nullptr);
    if (rc != RTB_OK) {
        float* all[3] = {*verts_out, kd, ks};
        for (float* q : all) if (q) { if (pinned) cudaFreeHost(q); else free(q); }
        *verts_out = nullptr; *n_out = 0;
        if (pinned) cudaGetLastError();
        return fail(rc, "%s", err.c_str());
    }
    if (kd_out) *kd_out = kd;
    if (ks_ns_out) *ks_ns_out = ks;
    return RTB_OK;
}
extern "C" int rtb_import_obj_parallel(const char* path, uint32_t n_threads, uint32_t flags, float** verts_out, uint32_t* n_out, rtb_import_stats* stats) {
    return rtb_import_obj_parallel_ex(path, n_threads, flags & ~RTB_OBJ_MTL, verts_out, nullptr, nullptr, n_out, stats);
}
extern "C" void rtb_free_pinned(void* p) { if (p) cudaFreeHost(p); }

// ---- several GPUs behind one call ---------------------------------------------------------------------------------------------
// One host thread per device (the reference's render() fans out to its own worker threads and joins them, src/main.rs:269-303).
struct MultiJob { int rc; std::string err; rtb_stats st; };

template <typename F>
static int run_per_device(uint32_t n, F&& fn, std::vector<MultiJob>& jobs) {
    jobs.assign(n, MultiJob());
    std::vector<std::thread> th;
    auto body = [&](uint32_t i) {
        memset(&jobs[i].st, 0, sizeof(rtb_stats));
        jobs[i].rc = fn(i, &jobs[i].st);
        if (jobs[i].rc != RTB_OK) jobs[i].err = g_err;
    };
    for (uint32_t i = 1; i < n; ++i) th.emplace_back(body, i);
    body(0);
    for (auto& t : th) t.join();
    for (uint32_t i = 0; i < n; ++i)
        if (jobs[i].rc != RTB_OK) return fail(jobs[i].rc, "device %u: %s", i, jobs[i].err.c_str());
    return RTB_OK;
}

extern "C" int rtb_scenes_set_samples(rtb_scene* const* scenes, uint32_t n_gpus, const float* samples, uint32_t n_pairs) {
    if (!scenes || n_gpus == 0) return fail(RTB_ERR_ARG, "rtb_scenes_set_samples: no scenes");
    for (uint32_t i = 0; i < n_gpus; ++i) if (!scenes[i]) return fail(RTB_ERR_ARG, "rtb_scenes_set_samples: scene %u is null", i);
    std::vector<MultiJob> jobs;
    return run_per_device(n_gpus, [&](uint32_t i, rtb_stats*) { return rtb_scene_set_samples_async(scenes[i], samples, n_pairs); }, jobs);
}

extern "C" int rtb_render_multi(rtb_scene* const* scenes, uint32_t n_gpus, const rtb_render_params* params, uint8_t* rgb8, rtb_stats* stats) {
    if (!scenes || n_gpus == 0 || !params || !rgb8) return fail(RTB_ERR_ARG, "rtb_render_multi: null argument");
    for (uint32_t i = 0; i < n_gpus; ++i) {
        if (!scenes[i]) return fail(RTB_ERR_ARG, "rtb_render_multi: scene %u is null", i);
        for (uint32_t j = 0; j < i; ++j)
            if (scenes[j]->device == scenes[i]->device) return fail(RTB_ERR_ARG, "rtb_render_multi: scenes %u and %u share device %d", j, i, scenes[i]->device);
    }
    rtb_render_params base = *params;
    if (base.band_rows == 0) base.band_rows = 4;
    base.world = n_gpus;
    base.flags = (params->flags & ~(RTB_FLAG_GATHER_PEER)) | RTB_FLAG_FRAME_ROWS;
    const bool peer = (params->flags & RTB_FLAG_GATHER_PEER) != 0 && n_gpus > 1;
    const size_t frame_bytes = (size_t)params->width * params->height * 3;
    uint8_t* d_full = nullptr;
    if (peer) {
        rtb_scene* s0 = scenes[0];
        CUDA_TRY(cudaSetDevice(s0->device));
        if (frame_bytes > s0->full_cap) {
            cudaFree(s0->d_full); s0->d_full = nullptr; s0->full_cap = 0;
            CUDA_TRY(cudaMalloc(&s0->d_full, frame_bytes));
            s0->full_cap = frame_bytes;
        }
        d_full = s0->d_full;
        for (uint32_t i = 1; i < n_gpus; ++i) {          // NVLink peer stores from device i into device 0's frame
            int can = 0;
            CUDA_TRY(cudaDeviceCanAccessPeer(&can, scenes[i]->device, s0->device));
            if (!can) return fail(RTB_ERR_CUDA, "rtb_render_multi: device %d cannot access device %d (no peer path)", scenes[i]->device, s0->device);
            CUDA_TRY(cudaSetDevice(scenes[i]->device));
            const cudaError_t e = cudaDeviceEnablePeerAccess(s0->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(RTB_ERR_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
            cudaGetLastError();
        }
    }
    std::vector<MultiJob> jobs;
    int rc = run_per_device(n_gpus, [&](uint32_t i, rtb_stats* st) {
        rtb_render_params p = base;
        p.rank = i;
        return render_core(scenes[i], &p, peer ? nullptr : rgb8, d_full, st);
    }, jobs);
    if (rc != RTB_OK) return rc;
    if (peer) {                                          // every device has synchronised its stream: the frame is complete
        CUDA_TRY(cudaSetDevice(scenes[0]->device));
        CUDA_TRY(cudaMemcpyAsync(rgb8, d_full, frame_bytes, cudaMemcpyDeviceToHost, scenes[0]->stream));
        CUDA_TRY(cudaStreamSynchronize(scenes[0]->stream));
    }
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        for (uint32_t i = 0; i < n_gpus; ++i) {
            const rtb_stats& t = jobs[i].st;
            stats->primary_rays += t.primary_rays; stats->primary_hits += t.primary_hits; stats->shadow_rays += t.shadow_rays;
            stats->shadow_occluded += t.shadow_occluded; stats->box_tests += t.box_tests; stats->tri_tests += t.tri_tests;
            stats->launches += t.launches; stats->queued_hits += t.queued_hits; stats->queued_groups += t.queued_groups;
            stats->queued_walks += t.queued_walks; stats->grid_rebuilds += t.grid_rebuilds; stats->grid_redo += t.grid_redo;
            if (t.kernel_ms > stats->kernel_ms) stats->kernel_ms = t.kernel_ms;
            if (t.primary_ms > stats->primary_ms) stats->primary_ms = t.primary_ms;
            if (t.shadow_ms > stats->shadow_ms) stats->shadow_ms = t.shadow_ms;
            if (t.shade_ms > stats->shade_ms) stats->shade_ms = t.shade_ms;
            if (t.grid_build_ms > stats->grid_build_ms) stats->grid_build_ms = t.grid_build_ms;
            stats->grid_cells = t.grid_cells; stats->grid_entries = t.grid_entries; stats->grid_big = t.grid_big;
        }
    }
    return RTB_OK;
}

// ---- frame sequences ----------------------------------------------------------------------------------------------
// A pool of host threads that encodes finished frames to PNG while the GPU renders the next ones (img.save, src/main.rs:312-315).
struct PngPool {
    std::mutex m; std::condition_variable cv_put, cv_get;
    std::deque<std::pair<std::string, std::vector<uint8_t>>> q;
    std::vector<std::thread> th;
    uint32_t w = 0, h = 0; bool done = false; int rc = RTB_OK; std::string err;
    void start(uint32_t width, uint32_t rows, unsigned n) {
        w = width; h = rows;
        for (unsigned i = 0; i < n; ++i) th.emplace_back([this] {
            for (;;) {
                std::pair<std::string, std::vector<uint8_t>> job;
                {
                    std::unique_lock<std::mutex> lk(m);
                    cv_get.wait(lk, [this] { return done || !q.empty(); });
                    if (q.empty()) return;
                    job = std::move(q.front()); q.pop_front();
                    cv_put.notify_one();
                }
                const int r = rtb_write_png(job.first.c_str(), job.second.data(), w, h);
                if (r != RTB_OK) { std::lock_guard<std::mutex> lk(m); if (rc == RTB_OK) { rc = r; err = rtb_last_error(); } }
            }
        });
    }
    void put(const std::string& path, const uint8_t* rgb8) {
        std::vector<uint8_t> copy(rgb8, rgb8 + (size_t)w * h * 3);
        std::unique_lock<std::mutex> lk(m);
        cv_put.wait(lk, [this] { return q.size() < 2 * th.size() + 2; });          // bounded: at most a few frames of memory
        q.emplace_back(path, std::move(copy));
        cv_get.notify_one();
    }
    void finish() {
        { std::lock_guard<std::mutex> lk(m); done = true; }
        cv_get.notify_all();
        for (auto& t : th) t.join();
        th.clear();
    }
};

// Poses first, first + stride, ... of `poses` on one scene.  Frame k goes to d_seq[k & 1] -> h_seq[k & 1] (pinned); a consumer
// thread delivers it when its copy has landed, while the submitting thread is already enqueueing frame k + 1.  The camera grid of
// every pose is a speculative rebuild; its control words are snapshotted behind the frame's kernels and checked by the consumer.
static int sequence_on_scene(rtb_scene* s, const rtb_render_params* params, const rtb_camera_pose* poses, uint32_t n_poses, uint32_t first,
                             uint32_t stride, uint8_t* frames, const char* png_pattern, PngPool* pool, rtb_frame_fn on_frame, void* user,
                             rtb_sequence_stats* stats) {
    if (!s || !params || (!poses && n_poses)) return fail(RTB_ERR_ARG, "rtb_render_sequence: null argument");
    CUDA_TRY(cudaSetDevice(s->device));
    const uint32_t world = params->world ? params->world : 1;
    const uint32_t rows = rtb_owned_rows(params->height, world > 1 ? params->band_rows : params->height, world > 1 ? params->rank : 0, world);
    const size_t bytes = (size_t)rows * params->width * 3;
    if (bytes > s->seq_cap) {
        for (int k = 0; k < 2; ++k) { cudaFree(s->d_seq[k]); s->d_seq[k] = nullptr; if (s->h_seq[k]) cudaFreeHost(s->h_seq[k]); s->h_seq[k] = nullptr; }
        s->seq_cap = 0;
        for (int k = 0; k < 2; ++k) { CUDA_TRY(cudaMalloc(&s->d_seq[k], bytes ? bytes : 1)); CUDA_TRY(cudaMallocHost(&s->h_seq[k], bytes ? bytes : 1)); }
        s->seq_cap = bytes;
    }
    if (!s->h_seq_ctl) CUDA_TRY(cudaMallocHost(&s->h_seq_ctl, 2 * 8 * sizeof(uint32_t)));
    for (int k = 0; k < 2; ++k) if (!s->ev_seq[k]) CUDA_TRY(cudaEventCreateWithFlags(&s->ev_seq[k], cudaEventDisableTiming));
    if (!s->ev_seq0) { CUDA_TRY(cudaEventCreate(&s->ev_seq0)); CUDA_TRY(cudaEventCreate(&s->ev_seq1)); }

    // a pinned `frames` array takes the device-to-host copies directly (no staging frame, nothing for the consumer to copy)
    bool direct = false;
    if (frames && bytes) {
        cudaPointerAttributes pa;
        direct = cudaPointerGetAttributes(&pa, frames) == cudaSuccess && pa.type == cudaMemoryTypeHost;
        if (!direct) cudaGetLastError();
    }
    std::vector<uint32_t> mine;
    for (uint32_t k = first; k < n_poses; k += stride) mine.push_back(k);
    const uint32_t n = (uint32_t)mine.size();
    const auto t_start = std::chrono::steady_clock::now();
    // shared between the submitting thread (this one) and the consumer
    std::mutex m; std::condition_variable cv;
    uint32_t submitted = 0, consumed = 0;          // frames enqueued / delivered (local indices)
    int64_t bad = -1;                               // local index of a frame whose camera grid did not fit: to be redone synchronously
    bool fixed = false, abort_all = false;
    int consumer_rc = RTB_OK; std::string consumer_err;
    uint32_t redo_count = 0;
    uint64_t primary_rays = 0;

    std::thread consumer([&] {
        cudaSetDevice(s->device);
        for (uint32_t j = 0; j < n; ++j) {
            {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [&] { return submitted > j || abort_all; });
                if (abort_all) return;
            }
            const int slot = (int)(j & 1);
            cudaError_t e = cudaEventSynchronize(s->ev_seq[slot]);
            if (e == cudaSuccess && s->h_seq_ctl[8 * slot + 4] != 0u) {           // the pose's camera grid was switched off: wrong frame
                std::unique_lock<std::mutex> lk(m);
                bad = j; fixed = false;
                cv.notify_all();
                cv.wait(lk, [&] { return fixed || abort_all; });
                if (abort_all) return;
                bad = -1;
                e = cudaEventSynchronize(s->ev_seq[slot]);
            }
            if (e != cudaSuccess) {
                std::lock_guard<std::mutex> lk(m);
                consumer_rc = RTB_ERR_CUDA; consumer_err = cudaGetErrorString(e); abort_all = true; cv.notify_all();
                return;
            }
            const uint32_t k = mine[j];
            const uint8_t* h = direct ? frames + (size_t)k * bytes : s->h_seq[slot];
            if (frames && !direct) memcpy(frames + (size_t)k * bytes, h, bytes);
            if (png_pattern && pool) { char path[1024]; snprintf(path, sizeof(path), png_pattern, k); pool->put(path, h); }
            if (on_frame) on_frame(user, k, h, params->width, rows);
            { std::lock_guard<std::mutex> lk(m); consumed = j + 1; }
            cv.notify_all();
        }
    });

    int rc = RTB_OK;
    auto submit = [&](uint32_t j, bool spec) -> int {
        rtb_render_params p = *params;
        const rtb_camera_pose& c = poses[mine[j]];
        memcpy(p.eye, c.eye, sizeof(p.eye)); memcpy(p.look_at, c.look_at, sizeof(p.look_at)); memcpy(p.up, c.up, sizeof(p.up));
        p.distance = c.distance;
        TraceArgs a;
        int r = prepare_args(s, &p, s->stream, &a, spec);
        if (r != RTB_OK) return r;
        const int slot = (int)(j & 1);
        a.rgb8 = s->d_seq[slot]; a.frame_rows = 0;
        r = run_wavefront(s, a, false, s->stream, false);
        if (r != RTB_OK) return r;
        primary_rays += (uint64_t)a.owned_rows * a.width * a.nb_ray;
        if (bytes) CUDA_TRY(cudaMemcpyAsync(direct ? frames + (size_t)mine[j] * bytes : s->h_seq[slot], s->d_seq[slot], bytes, cudaMemcpyDeviceToHost, s->stream));
        // snapshot of the camera grid's control words (failure bits) behind this frame: the next pose's build overwrites them
        if (s->cam_grid.d_ctl && s->cam_grid.pending) CUDA_TRY(cudaMemcpyAsync(s->h_seq_ctl + 8 * slot, s->cam_grid.d_ctl, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
        else memset(s->h_seq_ctl + 8 * slot, 0, 8 * sizeof(uint32_t));
        s->cam_grid.pending = false;
        CUDA_TRY(cudaEventRecord(s->ev_seq[slot], s->stream));
        return RTB_OK;
    };

    // the light grids (and a first camera grid) synchronously, once, before the pipeline starts
    if (n) {
        TraceArgs a0;
        rtb_render_params p0 = *params;
        const rtb_camera_pose& c0 = poses[mine[0]];
        memcpy(p0.eye, c0.eye, sizeof(p0.eye)); memcpy(p0.look_at, c0.look_at, sizeof(p0.look_at)); memcpy(p0.up, c0.up, sizeof(p0.up)); p0.distance = c0.distance;
        rc = prepare_args(s, &p0, s->stream, &a0, false);
        if (rc == RTB_OK && cudaStreamSynchronize(s->stream) != cudaSuccess) rc = fail(RTB_ERR_CUDA, "rtb_render_sequence: stream error");
        s->samples_pending = false;
    }
    if (rc == RTB_OK) cudaEventRecord(s->ev_seq0, s->stream);
    const bool no_spec = getenv("RTB200_NO_SPEC") != nullptr;
    for (uint32_t j = 0; j < n && rc == RTB_OK; ++j) {
        {   // slot j & 1 is free once frame j - 2 has been delivered; a failed frame is repaired first
            std::unique_lock<std::mutex> lk(m);
            cv.wait(lk, [&] { return consumed + 2 > j || bad >= 0 || abort_all; });
            if (abort_all) break;
            if (bad >= 0) {
                const uint32_t jb = (uint32_t)bad;
                lk.unlock();
                // everything in flight finishes; pose jb again with a synchronous camera-grid build (which also makes room), then on
                if (cudaStreamSynchronize(s->stream) != cudaSuccess) { rc = fail(RTB_ERR_CUDA, "rtb_render_sequence: stream error"); break; }
                s->cam_key_set = false; s->cam_grid.spec_ok = false;
                rc = submit(jb, false);
                if (rc != RTB_OK) break;
                redo_count++;
                // frames after jb that were already enqueued used the old allocation as well: enqueue them again, in order
                for (uint32_t j2 = jb + 1; j2 < j && rc == RTB_OK; ++j2) rc = submit(j2, !no_spec);
                if (rc != RTB_OK) break;
                lk.lock();
                fixed = true; bad = -1;
                cv.notify_all();
                cv.wait(lk, [&] { return consumed + 2 > j || abort_all; });
                if (abort_all) break;
            }
        }
        rc = submit(j, !no_spec);
        if (rc != RTB_OK) break;
        { std::lock_guard<std::mutex> lk(m); submitted = j + 1; }
        cv.notify_all();
    }
    // drain: the consumer may still report a bad frame among the last two
    for (;;) {
        std::unique_lock<std::mutex> lk(m);
        if (rc != RTB_OK) { abort_all = true; cv.notify_all(); break; }
        cv.wait(lk, [&] { return consumed >= n || bad >= 0 || abort_all; });
        if (abort_all || consumed >= n) break;
        const uint32_t jb = (uint32_t)bad;
        lk.unlock();
        if (cudaStreamSynchronize(s->stream) != cudaSuccess) { rc = fail(RTB_ERR_CUDA, "rtb_render_sequence: stream error"); continue; }
        s->cam_key_set = false; s->cam_grid.spec_ok = false;
        rc = submit(jb, false);
        redo_count++;
        for (uint32_t j2 = jb + 1; j2 < n && rc == RTB_OK; ++j2) rc = submit(j2, !no_spec);
        lk.lock();
        if (rc == RTB_OK) { fixed = true; bad = -1; cv.notify_all(); }
    }
    consumer.join();
    if (rc == RTB_OK && consumer_rc != RTB_OK) rc = fail(consumer_rc, "rtb_render_sequence: %s", consumer_err.c_str());
    cudaEventRecord(s->ev_seq1, s->stream);
    cudaStreamSynchronize(s->stream);
    s->cam_grid.pending = false;
    if (stats) {
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, s->ev_seq0, s->ev_seq1) != cudaSuccess) { ms = 0.0f; cudaGetLastError(); }
        stats->device_ms = ms;
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
        stats->frames = consumed; stats->cam_grid_redo = redo_count; stats->primary_rays = primary_rays; stats->shadow_rays = 0;
    }
    return rc;
}

extern "C" int rtb_render_sequence(rtb_scene* scene, const rtb_render_params* params, const rtb_camera_pose* poses, uint32_t n_poses,
                        uint8_t* frames, const char* png_pattern, rtb_frame_fn on_frame, void* user, rtb_sequence_stats* stats) {
    if (!scene || !params) return fail(RTB_ERR_ARG, "rtb_render_sequence: null argument");
    PngPool pool;
    if (png_pattern) {
        const uint32_t world = params->world ? params->world : 1;
        const uint32_t rows = rtb_owned_rows(params->height, world > 1 ? params->band_rows : params->height, world > 1 ? params->rank : 0, world);
        unsigned nt = std::thread::hardware_concurrency(); nt = nt > 8 ? 8 : (nt < 2 ? 2 : nt);
        pool.start(params->width, rows, nt);
    }
    if (stats) memset(stats, 0, sizeof(*stats));
    int rc = sequence_on_scene(scene, params, poses, n_poses, 0, 1, frames, png_pattern, png_pattern ? &pool : nullptr, on_frame, user, stats);
    if (png_pattern) {
        pool.finish();
        if (rc == RTB_OK && pool.rc != RTB_OK) rc = fail(pool.rc, "rtb_render_sequence: %s", pool.err.c_str());
        if (stats) stats->seconds += 0.0;
    }
    return rc;
}

extern "C" int rtb_render_sequence_multi(rtb_scene* const* scenes, uint32_t n_gpus, const rtb_render_params* params, const rtb_camera_pose* poses,
                              uint32_t n_poses, uint8_t* frames, const char* png_pattern, rtb_sequence_stats* stats) {
    if (!scenes || n_gpus == 0 || !params) return fail(RTB_ERR_ARG, "rtb_render_sequence_multi: null argument");
    for (uint32_t i = 0; i < n_gpus; ++i) if (!scenes[i]) return fail(RTB_ERR_ARG, "rtb_render_sequence_multi: scene %u is null", i);
    rtb_render_params p = *params;
    p.world = 1; p.rank = 0;                              // whole frames, frame-parallel
    PngPool pool;
    if (png_pattern) {
        unsigned nt = std::thread::hardware_concurrency(); nt = nt > 16 ? 16 : (nt < 2 ? 2 : nt);
        pool.start(p.width, p.height, nt);
    }
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<rtb_sequence_stats> st(n_gpus);
    std::vector<MultiJob> jobs;
    int rc = run_per_device(n_gpus, [&](uint32_t i, rtb_stats*) {
        memset(&st[i], 0, sizeof(st[i]));
        return sequence_on_scene(scenes[i], &p, poses, n_poses, i, n_gpus, frames, png_pattern, png_pattern ? &pool : nullptr, nullptr, nullptr, &st[i]);
    }, jobs);
    if (png_pattern) {
        pool.finish();
        if (rc == RTB_OK && pool.rc != RTB_OK) rc = fail(pool.rc, "rtb_render_sequence_multi: %s", pool.err.c_str());
    }
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        for (uint32_t i = 0; i < n_gpus; ++i) {
            stats->frames += st[i].frames; stats->cam_grid_redo += st[i].cam_grid_redo; stats->primary_rays += st[i].primary_rays;
            if (st[i].device_ms > stats->device_ms) stats->device_ms = st[i].device_ms;
        }
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    return rc;
}


