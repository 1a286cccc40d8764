// rtb_common.cuh -- shared device/host declarations of the B200 ray-casting path.
//
// Arithmetic contract.  The whole library is compiled with -fmad=false (and the default
// -prec-div=true -prec-sqrt=true -ftz=false), so every `a*b+c`, `/` and sqrtf below is the
// same correctly-rounded, unfused IEEE binary32 operation the reference's rustc/nalgebra code
// performs.  The only fused operations are the explicit fmaf() calls of the conservative
// inner-node slab test (rtb_trace.cu), which never decides a result on its own.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>
#include <string.h>
#include <math.h>

#define RTB_STACK_SIZE 64          // traversal stack entries (binary BVH depth limit)
#define RTB_CHILD_NONE 0x7fffffff  // absent child (single-primitive scenes)
#ifndef RTB_MAX_CAND
#define RTB_MAX_CAND 8             // primitives a hit point's shadow rays test directly (multiple of 4)
#endif
#define RTB_ITEM_F4 2                        // float4 per queued hit point: {ray id, prim word, -, -} {p_hit}
#define RTB_MAX_GROUPS 16                    // light-sample groups per hit point (k_light_groups)

struct f3 { float x, y, z; };

__host__ __device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__host__ __device__ __forceinline__ f3 sub3(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ f3 add3(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ f3 scale3(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
// nalgebra 0.11 dot: fold from zero, left to right.  The leading "0 +" only affects the sign of
// an all-zero result, which no consumer on this path can observe, so it is dropped.
__host__ __device__ __forceinline__ float dot3(f3 a, f3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__host__ __device__ __forceinline__ f3 cross3(f3 a, f3 b) {
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__host__ __device__ __forceinline__ float norm3(f3 a) { return sqrtf(dot3(a, a)); }
// Unit::new_normalize: component-wise true division by the norm (ray.rs:15, camera.rs:22-32, triangle.rs:29)
__host__ __device__ __forceinline__ f3 normalize3(f3 a) {
    float n = norm3(a);
    return mk3(a.x / n, a.y / n, a.z / n);
}

// Camera basis (src/tracer/utils/camera.rs:17-35), computed once per frame on the host in the same
// unfused f32 arithmetic (host compilers are given -ffp-contract=off).
struct CameraBasis { f3 u, v, w, eye; float distance; };

// ---- light grids (rtb_grid.cu) --------------------------------------------------------------------------------------
#define RTB_GRID_MAX_LIGHTS 1024   // light points a frame may use for the grid path (else the list / walk path)
#define RTB_GRID_MAX_BIG 64        // primitives that cannot be rasterised and are tested by every ray

struct GridLight {                 // one per light point
    float px, py, pz, H;           // the point, H = py - y_plane
    float x0, z0, cell, inv_cell;  // grid origin and cell size in plane units
    int nx, nz;                    // cells
    uint32_t cell_base, pad;       // first cell of this grid in cell_start
};

struct GridUnion {                 // occupancy of all light grids together, as a summed-area table (k_primary's shadow-queue test)
    float x0, z0, cell, inv_cell;
    int nx, nz; float y_plane, pad;
};

struct GridArgs {                  // build-time arguments
    const float4* tris; uint32_t n_tris;
    const float4* light_pts; uint32_t n_lights;
    float y_plane, cell0;
    GridLight* lights;
    unsigned int* lbounds;         // per light: ordered-uint min x, min z, max x, max z of the projections
    uint32_t* prim_flags;          // per primitive: 1 = "big" (not rasterised, tested by every ray)
    uint32_t* cell_start;          // total_cells + 1 offsets into entries
    uint32_t* cell_fill;           // per cell count / cursor
    int* entries;
    int* big_list;
    uint32_t* ctl;                 // [0] total cells [1] big primitives [2] union nx [3] union nz [4] failure bits (RTB_GRID_FAIL_*)
                                   // [6,7] total entries as one 64-bit count (a 32-bit scan of the per-cell counts can wrap)
    GridUnion* uni; uint32_t* sat; // union occupancy: (nz + 1) x (nx + 1) summed-area table with a zero border
    uint32_t cell_cap;
    // speculative rebuild (no host synchronisation): the arrays keep the capacity of the previous build and the kernels check
    // themselves against it; a grid that does not fit is disabled on the device (every nx = 0) and the failure is reported in ctl[4]
    uint32_t speculative;          // 1: check against the *_alloc capacities below instead of reporting sizes to the host
    uint32_t cells_alloc;          // cell_start / cell_fill entries allocated (total cells + 1 must fit)
    uint32_t expect_big;           // big-list length the caller's kernel arguments assume
    uint64_t sat_alloc, entries_alloc;
    uint32_t keyed;                // 1: the fill pass leaves an 8-bit coverage key above each primitive id (n_tris < 2^24); k_grid_sort orders by it
};
#define RTB_GRID_FAIL_CELLS 1u     // more cells than allocated
#define RTB_GRID_FAIL_BIG 2u       // big list longer than RTB_GRID_MAX_BIG or different from the expected length
#define RTB_GRID_FAIL_SAT 4u       // union table larger than allocated
#define RTB_GRID_FAIL_ENTRIES 8u   // more entries than allocated (or than a 32-bit offset can address)

struct RtbGrid {                   // host-side owner of the light grids of one scene (rtb_grid.cu)
    float y_min, mean_extent; bool stats_ok;
    GridLight* d_lights; unsigned int* d_lbounds; uint32_t* d_prim_flags; uint32_t* d_cell_start; uint32_t* d_cell_fill;
    int* d_entries; int* d_big; uint32_t* d_ctl; void* d_cub; GridUnion* d_union; uint32_t* d_sat;
    size_t union_cap, sat_cap;
    size_t lights_cap, lbounds_cap, flags_cap, cell_start_cap, cell_fill_cap, entries_cap, big_cap, ctl_cap, cub_bytes;
    bool valid; uint32_t n_lights, n_big, total_cells, total_entries; float build_ms;
    // cache key: the grids depend on the sample table and on how many of its entries the frame addresses
    uint64_t key_samples; uint32_t key_n_lights; bool key_set;
    // speculative rebuilds (rtb_grid.cu: grid_build_ex): state of the last synchronous build and of a pending speculative one
    bool spec_ok; uint32_t spec_n_tris, spec_n_big_; float spec_cell0, spec_y_plane; bool spec_with_union;
    uint32_t spec_cells;           // cells of that build: a speculative build scans (and checks itself against) 1.25 x this, not the whole allocation
    float spec_req;                // the cell size the last synchronous build was ASKED for (spec_cell0: what it settled on)
    bool pending;                  // a speculative build has been enqueued and not validated yet (rtb_grid_validate)
    uint64_t entry_cap_limit;      // entries above which the grid path is not taken (0 = default 512 Mi); tests lower it
    cudaEvent_t ev0, ev1;          // build time
    // speculative build: the fill pass runs on `side` next to the caller's primary-ray kernels (which need the union table only,
    // and that comes from the COUNT pass); the shadow kernels wait for ev_fill (fill_async: not waited for yet)
    cudaStream_t side; cudaEvent_t ev_count, ev_fill; bool fill_async;
    uint32_t builds, spec_builds, spec_failures;   // statistics
    // the entries carry an 8-bit coverage key above the primitive id (readers mask it off); `sorted`: k_grid_sort has ordered the lists
    // by it -- done when a build is used for a SECOND frame (rtb_grid_sort_on_reuse), a build per frame keeps the fill pass's two classes
    bool keyed, sorted;
    // frames rendered with this build after the first (the caller counts); `refined`: rebuilt with finer cells because the table keeps
    // being reused (rtb_api.cu: build_tables)
    uint32_t uses; bool refined;
    bool generous;                 // set by the caller before a build: larger entry cap / budget (a build for a reused table)
    bool sort_at_build;            // set by the caller before a build: run k_grid_sort with it (the frame share is large enough to repay the pass at once)
};

// hit record / queue item words (rtb_trace.cu)
#define RTB_REC_MISS 0x7fffffffu     // record word: no hit
#define RTB_REC_QUEUED 0x80000000u   // record word: low 31 bits = shadow-queue slot
#define RTB_REC_SELF 0x40000000u     // prim word: the hit primitive's own leaf is reachable (tested per ray in k_shade)
#define RTB_PRIM_MASK 0x3fffffffu
#define RTB_TERM_NONE 0xffffffffu

// Everything a trace kernel needs, passed by value as a kernel parameter.
struct TraceArgs {
    const float4* __restrict__ tris;     // 3 float4 per primitive: {v0,n.x} {v1,n.y} {v2,n.z}
    const float4* __restrict__ colors;   // 1 float4 per primitive: {r,g,b, tie_rank as int bits}
    const float4* __restrict__ ksns;     // {Ks.rgb, Ns} per primitive: Phong specular extension; nullptr = the reference's Lambert
    const float4* __restrict__ nodes;    // 4 float4 per inner node (layout in rtb_bvh.cu)
    const float2* __restrict__ samples;  // n_samples (u,v) pairs
    const float4* __restrict__ light_pts;// light sample points, index (r*nb_ray+i) % n_samples, see k_light_table
    uint32_t frame_rows;                    // 1: rgb8 is the full frame, owned rows go to their own position (RTB_FLAG_FRAME_ROWS)
    const float4* __restrict__ light_pairs; // pair table of the same points (k_light_pairs), 2 float4 per point
    const float* __restrict__ light_box; // 6 floats: component-wise min / max of light_pts[0..n_light_pts)
    const uint16_t* __restrict__ light_perm; // [nb_ray][nb_light_sample] sample at sorted position j (k_light_groups), nullptr = identity
    const float* __restrict__ light_gbox;    // [nb_ray][n_groups][6] bounds of each run of group_size sorted samples
    uint32_t n_groups, group_size;
    uint32_t n_tris, n_samples, n_light_pts;
    int32_t root;                        // node index, ~prim for a single-primitive scene, RTB_CHILD_NONE if empty
    uint32_t width, height, nb_ray, nb_light_sample;
    uint32_t band_rows, rank, world, owned_rows;
    CameraBasis cam;
    float denom;                         // (NB_RAY * NB_LIGHT_SAMPLE) as f32 (src/main.rs:211)
    float cdiv_hi, cdiv_lo;              // rn(1 / denom) and rn(1 / denom - cdiv_hi): the two-operation division by denom (k_shade)
    const uint32_t* cdiv_bad;            // significands for which that form differs from `/` (k_cdiv_check); use it only if 0
    // wavefront buffers (rtb_trace.cu): hit records [nb_ray][n_pix_padded], shadow queue, occlusion masks
    uint2* records;                      // {t bits, prim | flags} per (ray r, padded pixel)
    float4* items;                       // RTB_ITEM_F4 float4 per queued hit point
    uint4* subitems;                     // {hit slot, group, entry node, -} per (hit point, light-sample group) pair that walks the BVH per ray
    uint4* sublists;                     // {hit slot, group, pool offset, n} per pair with a complete candidate list
    int* cand_pool;                      // candidate primitives of the list pairs (bump-allocated by k_shadow_plan)
    uint32_t pool_cap;
    // light grids (rtb_grid.cu); grid_lights == nullptr: list / walk path
    const GridLight* grid_lights; const uint32_t* grid_cell_start; const int* grid_entries; const int* grid_big; uint32_t grid_n_big;
    int grid_entry_mask, cam_entry_mask;   // primitive id = entry & mask (0xffffff when the entries carry coverage keys, else -1)
    // camera grid for primary rays (one GridLight, the eye); cam_grid == nullptr: BVH traversal
    const GridLight* cam_grid; const uint32_t* cam_cell_start; const int* cam_entries; const int* cam_big; uint32_t cam_n_big;
    const GridUnion* grid_union; const uint32_t* grid_sat;
    uint32_t* masks;                     // mask_words per queued hit point: bit i = light sample i occluded
    uint32_t* queue_ctl;                 // [0] hit points queued by k_primary [1] k_shadow_plan ticket [2,3] k_shadow_queue ticket [4] walk pairs
                                         // [5] of which with an entry node [6] list pairs [7] pool entries used [8,9] k_shadow_lists ticket [11] term slots [12] k_shade_terms ticket
    unsigned long long* queue_ticket;    // (= queue_ctl + 2) next unclaimed ray of k_shadow_queue
    // term wavefront (k_shade_terms): hit points on primitives whose normal is not along y have their 100 shading terms
    // computed one warp per hit point, one lane per sample, into `terms`; k_shade then only adds them in order.
    uint32_t primary_split;              // k_primary: warps per 32-pixel tile (1, 2 or 4), each with a run of the pixel's rays
    uint32_t* tile_slot;                 // [nb_ray][n_pix_padded / 32]: term slot of (ray r, 32-pixel tile) or RTB_TERM_NONE (k_primary)
    uint4* term_tiles;                   // per slot: {first padded pixel of the tile, r, lanes with such a hit point, -}
    float* terms;                        // [slot][channel < term_channels][lane][sample, row stride term_stride]; nullptr: off
    uint32_t term_cap, term_stride, term_channels;   // slots available; floats per row (multiple of 4); 1 (all colours grey) or 3
    uint32_t magic_tiles_x, magic_band_rows, magic_n_samples;   // rtb_magic(): floor(2^32 / d) for the three runtime divisors of pixel_of / primary_dir
    uint32_t n_pix_padded, tiles_x, mask_words;   // n_pix_padded: record stride of THIS chunk (multiple of 256)
    uint32_t pix_base, n_pix_chunk;      // chunk = padded pixel indices [pix_base, pix_base + n_pix_chunk)
    // outputs
    uint8_t* rgb8;                       // frame mode: owned rows, compact; list mode: 3 bytes per listed pixel
    const uint2* pixel_list;             // list mode (else nullptr)
    uint64_t n_pix;
    float* out_rgb_lin; uint8_t* out_hit_count; int32_t* out_prim0; float* out_t0;
    unsigned long long* counters;        // [0] primary hits [1] shadow rays [2] occluded [3] box tests [4] tri tests [5] stack overflow
};

// n / d for a runtime divisor without the integer-division sequence: with m = floor(2^32 / d) (0xffffffff for d = 1),
// mulhi(n, m) is floor(n / d) or one less for every 32-bit n; one compare fixes it.  rem receives n % d.
static inline uint32_t rtb_magic(uint32_t d) { return d <= 1u ? 0xffffffffu : (uint32_t)(0x100000000ull / d); }
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t rtb_udiv(uint32_t n, uint32_t d, uint32_t magic, uint32_t& rem) {
    uint32_t q = __umulhi(n, magic);
    uint32_t r = n - q * d;
    if (r >= d) { ++q; r -= d; }
    rem = r;
    return q;
}
#endif

#define RTB_CTR_PRIMARY_HITS 0
#define RTB_CTR_SHADOW 1
#define RTB_CTR_OCCLUDED 2
#define RTB_CTR_BOX 3
#define RTB_CTR_TRI 4
#define RTB_CTR_OVERFLOW 5
#ifndef RTB_NUM_CTRS
#define RTB_NUM_CTRS 16
#endif
// 6..8: diagnostics of the counters build (k_shadow_queue: node steps of occluded / free rays, occluded rays)
#define RTB_CTR_QUEUED 9        // summed over the chunks of a call by k_shade: queued hit points, (hit point, group) pairs, walk pairs
#define RTB_CTR_PAIRS 10
#define RTB_CTR_WALKS 11
