// rtb_grid.cu -- light grids: a per-light-point lookup structure for shadow rays.
//
// Every shadow ray of a frame ends in one of the FEW light sample points the frame uses: the reference indexes them
// (r * NB_RAY + i) % len (src/main.rs:194), i.e. 100 points at 1 ray/pixel, 340 at 16 -- the same points for every
// pixel.  For a fixed light point P, "which primitives can the segment o -> P meet" depends only on the direction in
// which P sees o: it is a shadow-map query.  So, per light point, every primitive is projected from P onto the
// horizontal plane y = y_plane below the scene and rasterised CONSERVATIVELY into a 2-D grid of cells, each cell
// holding the list of primitives whose (dilated) projection touches it.  A shadow ray then projects its own origin o
// from P onto the same plane, reads one cell and runs the exact tests (rtb_device.cuh) on that cell's few primitives:
// no BVH walk, no stack, a handful of candidates instead of ~20 node visits.
//
// Exactness.  The grid only proposes candidates; acceptance is decided by the same exact expressions as everywhere
// else, and a shadow ray is occluded iff SOME accepted primitive lies within the light distance (any-hit, see
// rtb_trace.cu).  What must hold is that the cell read by a ray lists every primitive the exact test could accept.
// The exact test (Moller-Trumbore in float) accepts only rays that pass within a few ulps of the triangle; the ray's
// float direction deviates from the line o -> P by at most 4u L; projection magnifies by at most 1/RTB_GRID_TAU; the
// float evaluation of the projection errs by 8u (|q| + |P|).  All of this is far below the dilation `grid_margin`
// applied when rasterising.  Primitives that cannot be projected for some light point (a vertex not clearly below it)
// or that would cover too many cells (the ground) are not rasterised at all: they go to a short "big" list that every
// ray tests.  Rays whose origin is not clearly below their light point walk the BVH instead (lane_anyhit_walk).
#include "rtb_device.cuh"
#include <cstdio>
#include <cstdlib>
#include <cub/cub.cuh>

#define RTB_GRID_TAU 0.02f            // a point is projectable from P if P.y - y > TAU * (P.y - y_plane): magnification <= 50
#define RTB_GRID_MAX_CELLS_PER_PRIM 16384   // estimated list entries of one primitive in one light point's grid; above: 'big' list
#define RTB_GRID_MAX_ROWS_PER_PRIM 2048
#define RTB_GRID_MAX_DIM 8192
#define RTB_GRID_UNION_DIM 2048

__device__ __forceinline__ unsigned int g_f2ord(float f) { const unsigned int b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float g_ord2f(unsigned int o) { return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o); }

// Projection from the light point onto y = y_plane: q = P.xz + (v - P).xz * H / (P.y - v.y), H = P.y - y_plane.
__device__ __forceinline__ bool grid_project(float px, float py, float pz, float H, f3 v, float& qx, float& qz) {
    const float dy = py - v.y;
    if (!(dy > RTB_GRID_TAU * H)) return false;                 // also false for NaN
    const float s = H / dy;
    qx = fmaf(v.x - px, s, px);
    qz = fmaf(v.z - pz, s, pz);
    return true;
}

// Dilation (plane units) that covers every difference between "the exact test may accept" and "the ideal line o -> P
// meets the ideal triangle", see the header comment.
__device__ __forceinline__ float grid_margin(float cell, float qabs, float px, float pz, float H) {
    return 0.02f * cell + 0.01f + 4e-5f * (qabs + fabsf(px) + fabsf(pz) + H);
}

// ---- scene statistics (once per scene): lowest vertex and the log-mean of the primitives' largest box extent ----------
__global__ void k_grid_scene_stats(const float4* __restrict__ tris, uint32_t n, unsigned int* __restrict__ ymin_ord, double* __restrict__ sumlog,
                                   unsigned int* __restrict__ count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 q0 = tris[3 * i], q1 = tris[3 * i + 1], q2 = tris[3 * i + 2];
    const float ymin = fminf(fminf(q0.y, q1.y), q2.y);
    if (ymin == ymin) atomicMin(ymin_ord, g_f2ord(ymin));
    const float ex = fmaxf(fmaxf(q0.x, q1.x), q2.x) - fminf(fminf(q0.x, q1.x), q2.x);
    const float ey = fmaxf(fmaxf(q0.y, q1.y), q2.y) - ymin;
    const float ez = fmaxf(fmaxf(q0.z, q1.z), q2.z) - fminf(fminf(q0.z, q1.z), q2.z);
    const float e = fmaxf(fmaxf(ex, ey), ez);
    if (e > 0.0f && e < 1e30f) { atomicAdd(sumlog, (double)logf(e)); atomicAdd(count, 1u); }
}

// ---- pass A: which primitives are "big", and the bounds of each light point's projection ---------------------------
// One thread per (light point, primitive) pair.  What the pairs report -- the entry estimate, the four bound words of their light
// point -- is reduced in the warp, then in the block (shared memory: a block of 1 024 consecutive pairs belongs to one or two light
// points unless the scene is tiny), and only then sent to the global words: a warp's worth of atomics on the same few addresses
// per 32 pairs was most of this pass on a 20 M-pair build (64 bunnies: 2.4 ms).
#define RTB_CLASSIFY_THREADS 1024
__global__ void __launch_bounds__(RTB_CLASSIFY_THREADS) k_grid_classify(GridArgs g) {
    __shared__ unsigned int sb[2][4];
    __shared__ unsigned long long s_est;
    const unsigned FULL = 0xffffffffu;
    const uint64_t t0 = (uint64_t)blockIdx.x * blockDim.x, t = t0 + threadIdx.x;
    const uint64_t total = (uint64_t)g.n_lights * g.n_tris;
    const uint32_t l0 = (uint32_t)(t0 / g.n_tris);                      // the block's first light point
    if (threadIdx.x < 8) sb[threadIdx.x >> 2][threadIdx.x & 3] = (threadIdx.x & 2) ? 0u : 0xffffffffu;
    if (threadIdx.x == 0) s_est = 0ull;
    __syncthreads();
    const bool valid = t < total;
    const uint32_t light = valid ? (uint32_t)(t / g.n_tris) : 0xffffffffu, prim = valid ? (uint32_t)(t - (uint64_t)light * g.n_tris) : 0u;
    bool in = false;
    float est_cells = 0.0f;
    unsigned int ox0 = 0xffffffffu, oz0 = 0xffffffffu, ox1 = 0u, oz1 = 0u;
    if (valid) {
        const float4 P = g.light_pts[light];
        const float H = P.y - g.y_plane;
        const float4 q0 = g.tris[3 * prim], q1 = g.tris[3 * prim + 1], q2 = g.tris[3 * prim + 2];
        float x[3], z[3];
        const bool ok = H > 0.0f && grid_project(P.x, P.y, P.z, H, mk3(q0.x, q0.y, q0.z), x[0], z[0]) &&
                        grid_project(P.x, P.y, P.z, H, mk3(q1.x, q1.y, q1.z), x[1], z[1]) &&
                        grid_project(P.x, P.y, P.z, H, mk3(q2.x, q2.y, q2.z), x[2], z[2]);
        in = ok;
        if (ok) {
            const float x0 = fminf(fminf(x[0], x[1]), x[2]), x1 = fmaxf(fmaxf(x[0], x[1]), x[2]);
            const float z0 = fminf(fminf(z[0], z[1]), z[2]), z1 = fmaxf(fmaxf(z[0], z[1]), z[2]);
            // work of the scan conversion (k_grid_raster): one clipped-edge computation per cell row, one list entry per cell the
            // dilated triangle touches -- about area / cell^2 + (width + height) / cell + 1.  (The bounding box says little for the long thin
            // streaks that distant primitives project to.)
            const float rows = (z1 - z0) / g.cell0 + 2.0f, cols = (x1 - x0) / g.cell0 + 2.0f;
            const float area = 0.5f * fabsf((x[1] - x[0]) * (z[2] - z[0]) - (x[2] - x[0]) * (z[1] - z[0]));
            const float cells = area / (g.cell0 * g.cell0) + (rows - 2.0f) + (cols - 2.0f) + 1.0f;     // expected cells a convex shape meets: A / c^2 + (w + h) / c + 1
            in = cells <= (float)RTB_GRID_MAX_CELLS_PER_PRIM && rows <= (float)RTB_GRID_MAX_ROWS_PER_PRIM;      // false for NaN
            est_cells = cells;
            if (in) { ox0 = g_f2ord(x0); oz0 = g_f2ord(z0); ox1 = g_f2ord(x1); oz1 = g_f2ord(z1); }
        }
        if (!in) atomicOr(g.prim_flags + prim, 1u);
    }
    // estimated list entries of the whole build (ctl[8,9], 64-bit): the host coarsens the cells before rasterising anything if the
    // estimate is beyond its budget (wide scenes whose distant primitives project to long streaks)
    {
        const uint32_t est = in ? (uint32_t)est_cells : 0u;
        const uint32_t tot = __reduce_add_sync(FULL, est);
        if ((threadIdx.x & 31) == 0 && tot) atomicAdd(&s_est, (unsigned long long)tot);
    }
    // the lanes of a warp are consecutive primitives of (nearly always) one light point
    const unsigned peers = __match_any_sync(FULL, light);
    ox0 = __reduce_min_sync(peers, ox0); oz0 = __reduce_min_sync(peers, oz0);
    ox1 = __reduce_max_sync(peers, ox1); oz1 = __reduce_max_sync(peers, oz1);
    if (valid && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1) && ox0 <= ox1) {
        unsigned int* b = light - l0 < 2u ? sb[light - l0] : g.lbounds + 4 * light;
        atomicMin(b + 0, ox0); atomicMin(b + 1, oz0); atomicMax(b + 2, ox1); atomicMax(b + 3, oz1);
    }
    __syncthreads();
    if (threadIdx.x < 2 && l0 + threadIdx.x < g.n_lights && sb[threadIdx.x][0] <= sb[threadIdx.x][2]) {
        unsigned int* b = g.lbounds + 4 * (l0 + threadIdx.x);
        atomicMin(b + 0, sb[threadIdx.x][0]); atomicMin(b + 1, sb[threadIdx.x][1]); atomicMax(b + 2, sb[threadIdx.x][2]); atomicMax(b + 3, sb[threadIdx.x][3]);
    }
    if (threadIdx.x == 0 && s_est) atomicAdd(reinterpret_cast<unsigned long long*>(g.ctl + 8), s_est);
}

__global__ void k_grid_biglist(GridArgs g) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n_tris || !g.prim_flags[i]) return;
    const uint32_t k = atomicAdd(g.ctl + 1, 1u);
    if (k < RTB_GRID_MAX_BIG) g.big_list[k] = (int)i;
}

// ---- pass B: grid geometry per light point (one block; n_lights <= RTB_GRID_MAX_LIGHTS) ---------------------------------
__global__ void __launch_bounds__(1024) k_grid_layout(GridArgs g) {
    __shared__ float s_scale;
    __shared__ unsigned long long s_total;
    if (threadIdx.x == 0) { s_scale = 1.0f; s_total = 0ull; }
    __syncthreads();
    for (int pass = 0; pass < 2; ++pass) {
        const float cell = g.cell0 * s_scale;
        unsigned long long mine = 0;
        for (uint32_t l = threadIdx.x; l < g.n_lights; l += blockDim.x) {
            const float4 P = g.light_pts[l];
            const unsigned int* b = g.lbounds + 4 * l;
            GridLight G;
            G.px = P.x; G.py = P.y; G.pz = P.z; G.H = P.y - g.y_plane; G.cell_base = 0; G.nx = 0; G.nz = 0;
            G.x0 = 0.0f; G.z0 = 0.0f; G.inv_cell = 0.0f; G.cell = cell; G.pad = 0;
            if (b[0] <= b[2] && b[1] <= b[3] && G.H > 0.0f) {                    // at least one primitive projected
                const float x0 = g_ord2f(b[0]), z0 = g_ord2f(b[1]), x1 = g_ord2f(b[2]), z1 = g_ord2f(b[3]);
                const float m = grid_margin(cell, fmaxf(fmaxf(fabsf(x0), fabsf(x1)), fmaxf(fabsf(z0), fabsf(z1))), P.x, P.z, G.H);
                float c = cell;
                float nx = ceilf((x1 - x0 + 2.0f * m) / c) + 1.0f, nz = ceilf((z1 - z0 + 2.0f * m) / c) + 1.0f;
                if (nx > (float)RTB_GRID_MAX_DIM || nz > (float)RTB_GRID_MAX_DIM) {  // huge footprint: coarser cells for this light
                    c = fmaxf((x1 - x0 + 2.0f * m), (z1 - z0 + 2.0f * m)) / (float)(RTB_GRID_MAX_DIM - 2);
                    nx = ceilf((x1 - x0 + 2.0f * m) / c) + 1.0f; nz = ceilf((z1 - z0 + 2.0f * m) / c) + 1.0f;
                }
                G.x0 = x0 - m; G.z0 = z0 - m; G.cell = c; G.inv_cell = 1.0f / c; G.nx = (int)nx; G.nz = (int)nz;
                mine += (unsigned long long)G.nx * (unsigned long long)G.nz;
            }
            g.lights[l] = G;
        }
        atomicAdd(&s_total, mine);
        __syncthreads();
        if (pass == 0) {
            const bool too_many = s_total > (unsigned long long)g.cell_cap;
            __syncthreads();
            if (!too_many) break;
            if (threadIdx.x == 0) { s_scale = sqrtf((float)((double)s_total / (double)g.cell_cap)) * 1.05f; s_total = 0ull; }
            __syncthreads();
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {                                                       // exclusive scan of the cell counts
        unsigned long long base = 0;
        for (uint32_t l = 0; l < g.n_lights; ++l) {
            g.lights[l].cell_base = (uint32_t)base;
            base += (unsigned long long)g.lights[l].nx * (unsigned long long)g.lights[l].nz;
        }
        g.ctl[0] = base > 0xffffffffull ? 0xffffffffu : (uint32_t)base;           // total cells (the host checks it against cell_cap)
        // the union grid: covers every light grid, at most RTB_GRID_UNION_DIM cells a side
        float ux0 = 3.0e38f, uz0 = 3.0e38f, ux1 = -3.0e38f, uz1 = -3.0e38f, cell = g.cell0;
        for (uint32_t l = 0; l < g.n_lights; ++l) {
            const GridLight G = g.lights[l];
            if (G.nx <= 0) continue;
            ux0 = fminf(ux0, G.x0); uz0 = fminf(uz0, G.z0);
            ux1 = fmaxf(ux1, G.x0 + (float)G.nx * G.cell); uz1 = fmaxf(uz1, G.z0 + (float)G.nz * G.cell);
            cell = fmaxf(cell, G.cell);
        }
        GridUnion U;
        U.y_plane = g.y_plane; U.pad = 0.0f; U.x0 = 0.0f; U.z0 = 0.0f; U.cell = cell; U.inv_cell = 0.0f; U.nx = 0; U.nz = 0;
        if (ux0 <= ux1) {
            cell = fmaxf(cell, fmaxf(ux1 - ux0, uz1 - uz0) / (float)(RTB_GRID_UNION_DIM - 2));
            U.x0 = ux0; U.z0 = uz0; U.cell = cell; U.inv_cell = 1.0f / cell;
            U.nx = (int)ceilf((ux1 - ux0) / cell) + 1; U.nz = (int)ceilf((uz1 - uz0) / cell) + 1;
        }
        g.ctl[2] = (uint32_t)U.nx; g.ctl[3] = (uint32_t)U.nz;
        // Speculative rebuild: nobody on the host looks at these sizes before the next kernels run, so they are checked here
        // against what is allocated; a build that does not fit switches every grid off (nx = 0: no kernel reads a cell, every
        // lookup finds nothing) and says so in ctl[4].  The caller validates after its own synchronisation and rebuilds.
        uint32_t fail = 0;
        if (g.speculative) {
            if (base + 1ull > (unsigned long long)g.cells_alloc) fail |= RTB_GRID_FAIL_CELLS;
            if (g.ctl[1] > RTB_GRID_MAX_BIG || g.ctl[1] != g.expect_big) fail |= RTB_GRID_FAIL_BIG;
            if (g.sat && (unsigned long long)(U.nx + 1) * (unsigned long long)(U.nz + 1) > g.sat_alloc) fail |= RTB_GRID_FAIL_SAT;   // (no table: camera grid)
            if (fail) {
                for (uint32_t l = 0; l < g.n_lights; ++l) { g.lights[l].nx = 0; g.lights[l].nz = 0; g.lights[l].cell_base = 0; }
                U.nx = 0; U.nz = 0; g.ctl[0] = 0; g.ctl[2] = 0; g.ctl[3] = 0;
            }
        }
        g.ctl[4] = fail;
        *g.uni = U;
    }
}

// After the count pass: the number of entries (64-bit sum of the per-cell counts: the 32-bit scan that produces the offsets
// wraps silently above 2^32) against what is allocated / addressable.  Too many: every grid is switched off before the fill
// pass or any lookup can use a wrapped offset.
__global__ void k_grid_validate(GridArgs g) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const unsigned long long total = *reinterpret_cast<const unsigned long long*>(g.ctl + 6);
    if (total > g.entries_alloc || total > 0xfffffff0ull) {
        g.ctl[4] |= RTB_GRID_FAIL_ENTRIES;
        for (uint32_t l = 0; l < g.n_lights; ++l) { g.lights[l].nx = 0; g.lights[l].nz = 0; }
        GridUnion U = *g.uni; U.nx = 0; U.nz = 0; *g.uni = U;
    }
}

// ---- union occupancy: which plane cells have ANY primitive in ANY light point's grid ------------------------------------
__global__ void k_grid_union_mark(GridArgs g) {
    const uint32_t total_cells = g.ctl[0];                        // device-side: a speculative build launches for the allocated capacity
    const GridUnion U = *g.uni;
    if (U.nx <= 0 || U.nz <= 0) return;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < total_cells; c += gridDim.x * blockDim.x) {
    if (g.cell_start[c + 1] == g.cell_start[c]) continue;
    uint32_t lo = 0, hi = g.n_lights - 1;                          // the light whose grid holds cell c: last cell_base <= c
    while (lo < hi) {
        const uint32_t mid = (lo + hi + 1) >> 1;
        if (g.lights[mid].cell_base <= c) lo = mid; else hi = mid - 1;
    }
    while (lo > 0 && g.lights[lo].nx <= 0) --lo;                   // empty grids share their successor's base
    const GridLight G = g.lights[lo];
    if (G.nx <= 0) continue;
    const uint32_t local = c - G.cell_base;
    const int cz = (int)(local / (uint32_t)G.nx), cx = (int)(local - (uint32_t)cz * (uint32_t)G.nx);
    const float eps = 1e-3f * G.cell;
    int ux0 = (int)floorf((G.x0 + cx * G.cell - eps - U.x0) * U.inv_cell), ux1 = (int)floorf((G.x0 + (cx + 1) * G.cell + eps - U.x0) * U.inv_cell);
    int uz0 = (int)floorf((G.z0 + cz * G.cell - eps - U.z0) * U.inv_cell), uz1 = (int)floorf((G.z0 + (cz + 1) * G.cell + eps - U.z0) * U.inv_cell);
    ux0 = max(ux0, 0); uz0 = max(uz0, 0); ux1 = min(ux1, U.nx - 1); uz1 = min(uz1, U.nz - 1);
    const uint32_t W = (uint32_t)U.nx + 1;
    for (int z = uz0; z <= uz1; ++z)
        for (int x = ux0; x <= ux1; ++x) g.sat[(uint32_t)(z + 1) * W + (uint32_t)(x + 1)] = 1u;
    }
}

// Summed-area table of the occupancy marks, in place: prefix sums along the rows (a warp per row, 32 columns a step, the carry in
// lane 31), then down the columns (a block per 32 columns; its 32 row segments are summed, offset by the segments above, and
// summed again with the offset).  One thread per row / per column walked 2 049 dependent steps each: 0.16 ms of every rebuild.
__global__ void __launch_bounds__(256) k_grid_sat_rows(uint32_t* __restrict__ sat, const GridUnion* __restrict__ uni) {
    const int nx = uni->nx, nz = uni->nz;
    const int z = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = (int)(threadIdx.x & 31u);
    if (z >= nz) return;                                               // (whole warps)
    uint32_t* row = sat + (size_t)(z + 1) * (nx + 1);
    uint32_t carry = 0;
    for (int x0 = 1; x0 <= nx; x0 += 32) {
        const int x = x0 + lane;
        uint32_t v = x <= nx ? row[x] : 0u;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += u; }
        v += carry;
        if (x <= nx) row[x] = v;
        carry = __shfl_sync(0xffffffffu, v, 31);
    }
}

__global__ void __launch_bounds__(1024) k_grid_sat_cols(uint32_t* __restrict__ sat, const GridUnion* __restrict__ uni) {
    __shared__ uint32_t part[32][33];
    const int nx = uni->nx, nz = uni->nz;
    const int cx = (int)(blockIdx.x * 32 + threadIdx.x), seg = (int)threadIdx.y;
    if ((int)(blockIdx.x * 32) >= nx) return;                          // (whole blocks)
    const int per = (nz + 31) / 32, z0 = 1 + seg * per, z1 = min(z0 + per - 1, nz);
    const size_t W = (size_t)nx + 1;
    uint32_t sum = 0;
    if (cx < nx) for (int z = z0; z <= z1; ++z) sum += sat[(size_t)z * W + (size_t)(cx + 1)];
    part[seg][threadIdx.x] = sum;
    __syncthreads();
    uint32_t acc = 0;
    for (int k = 0; k < seg; ++k) acc += part[k][threadIdx.x];
    if (cx < nx) for (int z = z0; z <= z1; ++z) { uint32_t* p = sat + (size_t)z * W + (size_t)(cx + 1); acc += *p; *p = acc; }
}

// ---- passes C / D: conservative rasterisation of every projected primitive into its light point's grid -----------------
// FILL = false: count per cell (cell_fill);  FILL = true: write the primitive at cell_start[cell] + cursor.
// Work is dealt to the lanes of a warp twice, because both levels are ragged (a projection covers 1 to 2 048 rows, a row span 1 to
// thousands of cells; a thread that walked its own rows and spans left two thirds of its warp idle -- ncu: 10 of 32 threads active,
// 13 warp instructions per listed cell):
//   1. every thread sets up ONE (light point, primitive) pair -- projection, clipped-edge coefficients, row range -- and leaves the
//      record in shared memory;
//   2. the rows of the warp's 32 pairs are concatenated (exclusive scan of the row counts; the owner of item j is found by a
//      five-step search over the lanes' offsets) and lane j of a round computes the cell span [sx0, sx1] of row item j;
//   3. every lane lists the first four cells of its own span (spans average 2.7 cells); what is left of longer spans is concatenated
//      the same way and lane j of an emission round lists cell j of the remainder.
#ifndef RTB_KEY_DEPTH
#define RTB_KEY_DEPTH 1
#endif
#define RTB_KEY_FRONT 127u               // keys >= this (coverage estimate >= 60 %, raster_key) fill a cell's list from the front, the others from the back
#ifndef RTB_KEY_FRONT_LONG
#define RTB_KEY_FRONT_LONG 62u           // ... in a list too long for k_grid_sort (more than RTB_SORT_MAX entries), which has only these two classes
#endif
#define RTB_SORT_MAX 32
#define RTB_RASTER_REC 19
#define RTB_RASTER_DIRECT 4              // cells of its own span every lane lists itself
// Coverage key of a cell: from the estimated share of the cell the triangle's projection covers, i.e. the probability that it
// occludes a ray that looks the cell up.  Within the row the projection is, up to the vertices inside it, the trapezoid between its
// cross-sections I0 = [i0lo, i0hi] at the lower and I1 at the upper edge of the row (an empty one: a tip): the cell's x overlap
// with the two, averaged, times the z overlap of row and triangle (kscale = fz / (2 cell)).  A triangle that lies within the row
// has no cross-section at either edge: its area, spread over the cells of its span (ksmall >= 0).
// Keys 128 .. 254: the estimate is 60 % or more -- such a primitive occludes nearly every ray from BELOW it that looks the cell up,
// and what decides is whether the ray starts below it: the nearer to the light point, the more hit points it shadows (a hit point
// on the far side of the bunny is occluded by the near side's primitive of its cell, never by its own neighbours).  These are
// ordered by height above the plane (depth7, 0 .. 126 = at the light point).  Keys 0 .. 127: 127 sqrt(share / 0.6).
__device__ __forceinline__ uint32_t raster_key(int cx, float Gx0, float Gcell, float i0lo, float i0hi, float i1lo, float i1hi, float kscale, float ksmall, uint32_t depth7) {
    const float xa = Gx0 + (float)cx * Gcell, xb = xa + Gcell;
    const float ov0 = fmaxf(fminf(i0hi, xb) - fmaxf(i0lo, xa), 0.0f), ov1 = fmaxf(fminf(i1hi, xb) - fmaxf(i1lo, xa), 0.0f);
    const float cov = ksmall >= 0.0f ? ksmall : kscale * (ov0 + ov1);
#if RTB_KEY_DEPTH
    if (cov >= 0.6f) return 128u + depth7;
#endif
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fminf(cov * (1.0f / 0.6f), 1.0f)));   // square-root scale: small shares (many small primitives per cell) keep their order
    return (uint32_t)(r * 127.0f);
}
// Lists U cells per lane.  Count pass: one atomic each.  Fill pass: the position comes back from an atomic, so the U of them are
// issued together -- offsets loaded, atomics issued, then the stores -- or the pass is a chain of L2 round trips per warp.
// Order within a cell (a heuristic, any order is correct): likely occluders (key >= RTB_KEY_FRONT) fill the list from the front,
// the others from the back; one counter, front count in the low and back count in the high half (cells with >= 2^16 entries are
// filled front to back only).  k_grid_sort refines the order of short lists afterwards.
template <bool FILL, int U>
__device__ __forceinline__ void raster_emit(const GridArgs& g, const uint32_t (&cell)[U], const uint32_t (&eprim)[U], const bool (&ok)[U], bool (&front)[U]) {
    if (FILL) {
        uint32_t beg[U], cnt[U], old[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            beg[u] = 0u; cnt[u] = 0u;
            if (ok[u]) { beg[u] = g.cell_start[cell[u]]; cnt[u] = g.cell_start[cell[u] + 1] - beg[u]; }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            old[u] = 0u;
            // (front[] came in as key >= RTB_KEY_FRONT; a list too long for k_grid_sort has only these two classes: a lower bar)
            front[u] = front[u] || cnt[u] >= 65536u || (cnt[u] > RTB_SORT_MAX && (eprim[u] >> 24) >= RTB_KEY_FRONT_LONG);
            if (ok[u]) old[u] = atomicAdd(g.cell_fill + cell[u], front[u] ? 1u : 0x10000u);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (ok[u]) {
                const uint32_t pos = front[u] ? beg[u] + (old[u] & (cnt[u] >= 65536u ? 0xffffffffu : 0xffffu)) : beg[u] + cnt[u] - 1u - (old[u] >> 16);
                if ((uint64_t)pos < g.entries_alloc) g.entries[pos] = (int)eprim[u];
            }
        }
    } else {
#pragma unroll
        for (int u = 0; u < U; ++u) if (ok[u]) atomicAdd(g.cell_fill + cell[u], 1u);
    }
}
#ifndef RTB_RASTER_RB
#define RTB_RASTER_RB 4096               // rows of every pair dealt at a time (all: k_grid_sort decides the order of a list, not the arrival)
#endif
template <bool FILL>
__global__ void __launch_bounds__(256) k_grid_raster(GridArgs g) {
    __shared__ float rec[RTB_RASTER_REC][256];                    // [field][thread]: lanes read distinct owners -> distinct banks
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u, wbase = threadIdx.x & ~31u;
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t mine = 0;                                            // cells this thread listed (count pass)
    bool work = t < (uint64_t)g.n_lights * g.n_tris;
    const uint32_t light = work ? (uint32_t)(t / g.n_tris) : 0u, prim = work ? (uint32_t)(t - (uint64_t)light * g.n_tris) : 0u;
    work = work && !g.prim_flags[prim];
    int n_rows = 0;
    {
        const GridLight G = g.lights[light];
        work = work && G.nx > 0 && G.nz > 0;
        float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = q0, q2 = q0;
        if (work) { q0 = g.tris[3 * prim]; q1 = g.tris[3 * prim + 1]; q2 = g.tris[3 * prim + 2]; }
        float x[3] = {0.f, 0.f, 0.f}, z[3] = {0.f, 0.f, 0.f};
        work = work && grid_project(G.px, G.py, G.pz, G.H, mk3(q0.x, q0.y, q0.z), x[0], z[0]) &&
               grid_project(G.px, G.py, G.pz, G.H, mk3(q1.x, q1.y, q1.z), x[1], z[1]) &&
               grid_project(G.px, G.py, G.pz, G.H, mk3(q2.x, q2.y, q2.z), x[2], z[2]);     // (always true here: such a primitive is flagged)
        if (work) {
            const float bx0 = fminf(fminf(x[0], x[1]), x[2]), bx1 = fmaxf(fmaxf(x[0], x[1]), x[2]);
            const float bz0 = fminf(fminf(z[0], z[1]), z[2]), bz1 = fmaxf(fmaxf(z[0], z[1]), z[2]);
            const float m = grid_margin(G.cell, fmaxf(fmaxf(fabsf(bx0), fabsf(bx1)), fmaxf(fabsf(bz0), fabsf(bz1))), G.px, G.pz, G.H);
            int cx0 = (int)floorf((bx0 - m - G.x0) * G.inv_cell), cx1 = (int)floorf((bx1 + m - G.x0) * G.inv_cell);
            int cz0 = (int)floorf((bz0 - m - G.z0) * G.inv_cell), cz1 = (int)floorf((bz1 + m - G.z0) * G.inv_cell);
            cx0 = max(cx0, 0); cz0 = max(cz0, 0); cx1 = min(cx1, G.nx - 1); cz1 = min(cz1, G.nz - 1);
            n_rows = max(cz1 - cz0 + 1, 0);
            float ext = 0.0f;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int k1 = (k + 1) % 3;
                const float dz = z[k1] - z[k];
                rec[k][threadIdx.x] = x[k]; rec[3 + k][threadIdx.x] = z[k]; rec[6 + k][threadIdx.x] = x[k1] - x[k];
                rec[9 + k][threadIdx.x] = dz != 0.0f ? 1.0f / dz : 0.0f;                    // 0: a horizontal edge
                ext = fmaxf(ext, fmaxf(fabsf(x[k]), fabsf(z[k])));
            }
            rec[12][threadIdx.x] = m;
            rec[13][threadIdx.x] = 8e-6f * (ext + (bx1 - bx0) + (bz1 - bz0)) + 1e-30f;     // eps
            rec[14][threadIdx.x] = __int_as_float(cx0); rec[15][threadIdx.x] = __int_as_float(cx1); rec[16][threadIdx.x] = __int_as_float(cz0);
            if (FILL) {
                rec[17][threadIdx.x] = 0.5f * fabsf((x[1] - x[0]) * (z[2] - z[0]) - (x[2] - x[0]) * (z[1] - z[0]));     // projected area
                const float up = (fmaxf(fmaxf(q0.y, q1.y), q2.y) - g.y_plane) / G.H;                                      // height above the plane, 1 = at the light point
                rec[18][threadIdx.x] = __uint_as_float((uint32_t)(fminf(fmaxf(up, 0.0f), 1.0f) * 126.0f));
            }
        }
    }
    __syncwarp();
    const int max_rows = __reduce_max_sync(FULL, n_rows);
    const float inf = __int_as_float(0x7f800000);
    for (int r0 = 0; r0 < max_rows; r0 += RTB_RASTER_RB) {             // RTB_RASTER_RB rows of every pair at a time
    const int c_rows = min(max(n_rows - r0, 0), RTB_RASTER_RB);
    int incl_r = c_rows;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl_r, d); if ((int)lane >= d) incl_r += v; }
    const int excl_r = incl_r - c_rows, R = __shfl_sync(FULL, incl_r, 31);
    // Scan conversion by rows.  A cell belongs to the primitive iff its m-dilated rectangle meets the projected triangle.  For the
    // row cz that is a statement about x alone: the part of the (convex) triangle inside the dilated slab Z0 <= z <= Z1 projects
    // onto an x interval [X0, X1] -- the extreme x of the three edges clipped to the slab -- and the cells of the row whose dilated
    // x range meets that interval are exactly the cells whose rectangle meets the triangle.  Three clipped edges per ROW instead of
    // three edge functions per CELL of the bounding box, and no cells beyond the triangle's corners.  Rounding (parameters in [0, 1]
    // to 2^-22, x to a few ulps) is covered by `eps`, itself far below the margin m that makes the rasterisation conservative.
    // FILL also takes the triangle's cross-sections at the two edges of the row, for the coverage keys (raster_key).
    for (int rb = 0; rb < R; rb += 32) {
        const int ri = rb + (int)lane;
        int own = 0, own_ex = 0;                                  // the last lane whose offset is <= ri (it has rows)
#pragma unroll
        for (int step = 16; step >= 1; step >>= 1) {
            const int v = __shfl_sync(FULL, excl_r, own + step);
            if (v <= ri) { own += step; own_ex = v; }
        }
        const uint32_t olight = __shfl_sync(FULL, light, own), oprim = __shfl_sync(FULL, prim, own);
        int sx0 = 0, len = 0;
        uint32_t row = 0;
        float i0lo = inf, i0hi = -inf, i1lo = inf, i1hi = -inf;      // x range of the triangle AT z = Z0 and AT z = Z1 (FILL: raster_key)
        float kscale = 0.0f, ksmall = -1.0f, kx0 = 0.0f, kcell = 0.0f;
        uint32_t kdepth = 0u;
        if (ri < R) {
            const unsigned o = wbase + (unsigned)own;
            const GridLight* __restrict__ GL = g.lights + olight;
            const float Gx0 = GL->x0, Gz0 = GL->z0, Gcell = GL->cell, Ginv = GL->inv_cell;
            const int Gnx = GL->nx;
            const float m = rec[12][o], eps = rec[13][o];
            const int cx0 = __float_as_int(rec[14][o]), cx1 = __float_as_int(rec[15][o]), cz = __float_as_int(rec[16][o]) + r0 + (ri - own_ex);
            const float Z0 = Gz0 + cz * Gcell - m - eps, Z1 = Gz0 + (cz + 1) * Gcell + m + eps;
            float X0 = inf, X1 = -inf;                               // x range of the triangle within the slab
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float xa = rec[k][o], za = rec[3 + k][o], dx = rec[6 + k][o], idz = rec[9 + k][o];
                float t0, t1;
                if (idz != 0.0f) { t0 = (Z0 - za) * idz; t1 = (Z1 - za) * idz; }
                else { const bool in = za >= Z0 && za <= Z1; t0 = in ? -inf : inf; t1 = in ? inf : inf; }      // a horizontal edge: inside the slab or not
                const float tlo = fmaxf(fminf(t0, t1), 0.0f), thi = fminf(fmaxf(t0, t1), 1.0f);
                if (tlo <= thi) {                                    // the edge meets the slab (false for NaN)
                    const float xl = fmaf(tlo, dx, xa), xh = fmaf(thi, dx, xa);
                    X0 = fminf(X0, fminf(xl, xh)); X1 = fmaxf(X1, fmaxf(xl, xh));
                }
                if (FILL && idz != 0.0f) {
                    if (t0 >= 0.0f && t0 <= 1.0f) { const float xc = fmaf(t0, dx, xa); i0lo = fminf(i0lo, xc); i0hi = fmaxf(i0hi, xc); }
                    if (t1 >= 0.0f && t1 <= 1.0f) { const float xc = fmaf(t1, dx, xa); i1lo = fminf(i1lo, xc); i1hi = fmaxf(i1hi, xc); }
                }
            }
            if (X0 <= X1) {                                          // the triangle reaches this row
                sx0 = (int)floorf((X0 - m - eps - Gx0) * Ginv);
                int sx1 = (int)floorf((X1 + m + eps - Gx0) * Ginv);
                sx0 = max(sx0, cx0); sx1 = min(sx1, cx1);
                len = max(sx1 - sx0 + 1, 0);
                row = GL->cell_base + (uint32_t)cz * (uint32_t)Gnx;
                if (FILL) {
                    const float bz0 = fminf(fminf(rec[3][o], rec[4][o]), rec[5][o]), bz1 = fmaxf(fmaxf(rec[3][o], rec[4][o]), rec[5][o]);
                    const float rz0 = Gz0 + cz * Gcell;
                    const float fz = fminf(fmaxf((fminf(bz1, rz0 + Gcell) - fmaxf(bz0, rz0)) * Ginv, 0.0f), 1.0f);
                    kscale = fz * 0.5f * Ginv; kx0 = Gx0; kcell = Gcell; kdepth = __float_as_uint(rec[18][o]);
                    if (!(i0lo <= i0hi) && !(i1lo <= i1hi))
                        ksmall = fminf(rec[17][o] * Ginv * Ginv, 1.0f) / (float)max(len, 1);
                }
            }
        }
        if (!FILL) mine += (uint32_t)len;
        // ---- the cells of the round's spans.  Spans average under three cells: every lane lists the first RTB_RASTER_DIRECT cells
        // of its own span itself (no search, nothing fetched from other lanes) ...
        {
            uint32_t cell[RTB_RASTER_DIRECT], eprim[RTB_RASTER_DIRECT];
            bool ok[RTB_RASTER_DIRECT], front[RTB_RASTER_DIRECT];
#pragma unroll
            for (int u = 0; u < RTB_RASTER_DIRECT; ++u) {
                ok[u] = u < len;
                cell[u] = row + (uint32_t)(sx0 + u);
                eprim[u] = oprim; front[u] = false;
                if (FILL) {
                    const uint32_t key = raster_key(sx0 + u, kx0, kcell, i0lo, i0hi, i1lo, i1hi, kscale, ksmall, kdepth);
                    front[u] = key > RTB_KEY_FRONT;
                    if (g.keyed) eprim[u] |= key << 24;
                }
            }
            raster_emit<FILL, RTB_RASTER_DIRECT>(g, cell, eprim, ok, front);
        }
        // ... and what is left of longer spans is concatenated and dealt to the warp
        const int rem = max(len - RTB_RASTER_DIRECT, 0);
        if (!__any_sync(FULL, rem > 0)) continue;
        int incl = rem;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if ((int)lane >= d) incl += v; }
        const int excl = incl - rem, T = __shfl_sync(FULL, incl, 31);
        constexpr int U = FILL ? 2 : 1;
        for (int jb = 0; jb < T; jb += 32 * U) {
            uint32_t cell[U], eprim[U];
            bool ok[U], front[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int j = jb + 32 * u + (int)lane;
                int so = 0, so_ex = 0;                            // the last lane whose offset is <= j (it has cells left)
#pragma unroll
                for (int step = 16; step >= 1; step >>= 1) {
                    const int v = __shfl_sync(FULL, excl, so + step);
                    if (v <= j) { so += step; so_ex = v; }
                }
                const int rel = RTB_RASTER_DIRECT + (j - so_ex), cx = __shfl_sync(FULL, sx0, so) + rel;
                cell[u] = __shfl_sync(FULL, row, so) + (uint32_t)cx;
                ok[u] = j < T;
                eprim[u] = 0u; front[u] = false;
                if (FILL) {
                    eprim[u] = __shfl_sync(FULL, oprim, so);
                    const uint32_t key = raster_key(cx, __shfl_sync(FULL, kx0, so), __shfl_sync(FULL, kcell, so), __shfl_sync(FULL, i0lo, so), __shfl_sync(FULL, i0hi, so),
                                                    __shfl_sync(FULL, i1lo, so), __shfl_sync(FULL, i1hi, so), __shfl_sync(FULL, kscale, so), __shfl_sync(FULL, ksmall, so), __shfl_sync(FULL, kdepth, so));
                    front[u] = key > RTB_KEY_FRONT;
                    if (g.keyed) eprim[u] |= key << 24;
                }
            }
            raster_emit<FILL, U>(g, cell, eprim, ok, front);
        }
    }
    }
    if (!FILL) {                                                  // 64-bit total of the counts, one atomic per warp
        const uint32_t tot = __reduce_add_sync(FULL, mine);
        if (lane == 0 && tot) atomicAdd(reinterpret_cast<unsigned long long*>(g.ctl + 6), (unsigned long long)tot);
    }
}

// ---- pass E: order every cell's list by the coverage keys of the fill pass (likeliest occluder first) ----------------------------
// The any-hit lookup of k_shadow_grid stops at the first primitive that occludes its ray; how soon that happens is decided by the
// order of the list.  Covering primitives first (round 1) was worth 20 %; the fill pass's two classes by coverage estimate another
// 4 - 17 % of the frame (config 3: 16.7 -> 15.9 ms, 64 bunnies: 16.0 -> 13.2 ms), the full order a further 2 - 4 % (15.6 ms; distance
// 576: 42.1 -> 39.3 -> 38.0 ms).  The pass costs about as much as the count pass.  It runs with the build when the frame share is
// large enough to repay it at once (RtbGrid::sort_at_build: 48 Mi rays or more), else when a build is used for a SECOND frame
// (rtb_grid_sort_on_reuse): a band share of 8 GPUs that draws a new table per frame keeps the two classes, a caller that reuses
// the table (or the camera) gets the full order from the second frame on.  The keys stay in the entries; readers mask them off.
__device__ __forceinline__ void sort_ce(uint32_t& a, uint32_t& b) { const uint32_t hi = max(a, b), lo = min(a, b); a = hi; b = lo; }
__global__ void __launch_bounds__(256) k_grid_sort(GridArgs g) {
    if (g.ctl[4]) return;                                          // a build that did not fit: the grids are off, the offsets mean nothing
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t total = g.ctl[0];
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t c0 = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); c0 < total; c0 += stride) {      // a warp: 32 consecutive cells
        const uint32_t c = c0 + lane;
        uint32_t beg = 0, cnt = 0;
        if (c < total) {
            beg = g.cell_start[c];
            const uint32_t end = g.cell_start[c + 1];
            if (end > beg && (uint64_t)end <= g.entries_alloc) cnt = end - beg;
        }
        const uint32_t cmax = __reduce_max_sync(FULL, cnt);
        if (cmax <= 1) continue;
        uint32_t* e = reinterpret_cast<uint32_t*>(g.entries) + beg;
        // Whole words are compared (key above the primitive id): the order does not depend on which atomic of the fill pass came
        // first.  Lists of up to 8 entries: in registers, all loads in flight together, a sorting network (empty slots hold 0 and
        // sink to the end); the network is chosen by the longest list of the warp's 32 cells.
        if (cmax <= 4) {
            uint32_t v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) v[i] = (uint32_t)i < cnt ? e[i] : 0u;
            sort_ce(v[0], v[1]); sort_ce(v[2], v[3]); sort_ce(v[0], v[2]); sort_ce(v[1], v[3]); sort_ce(v[1], v[2]);
#pragma unroll
            for (int i = 0; i < 4; ++i) if ((uint32_t)i < cnt) e[i] = v[i];
            continue;
        }
        if (cnt <= 8) {
            uint32_t v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = (uint32_t)i < cnt ? e[i] : 0u;
            sort_ce(v[0], v[1]); sort_ce(v[2], v[3]); sort_ce(v[4], v[5]); sort_ce(v[6], v[7]);
            sort_ce(v[0], v[2]); sort_ce(v[1], v[3]); sort_ce(v[4], v[6]); sort_ce(v[5], v[7]);
            sort_ce(v[1], v[2]); sort_ce(v[5], v[6]);
            sort_ce(v[0], v[4]); sort_ce(v[1], v[5]); sort_ce(v[2], v[6]); sort_ce(v[3], v[7]);
            sort_ce(v[2], v[4]); sort_ce(v[3], v[5]);
            sort_ce(v[1], v[2]); sort_ce(v[3], v[4]); sort_ce(v[5], v[6]);
#pragma unroll
            for (int i = 0; i < 8; ++i) if ((uint32_t)i < cnt) e[i] = v[i];
        }
        // longer lists, one at a time by the whole warp: up to 32 entries are ranked by counting (entry i goes to the number of
        // entries above it; the words are distinct), longer ones keep the two-ended order of the fill pass
        unsigned todo = __ballot_sync(FULL, cnt > 8);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const uint32_t b2 = __shfl_sync(FULL, beg, src), n2 = __shfl_sync(FULL, cnt, src);
            uint32_t* e2 = reinterpret_cast<uint32_t*>(g.entries) + b2;
            if (n2 <= RTB_SORT_MAX) {
                const uint32_t x = lane < n2 ? e2[lane] : 0u;
                uint32_t rank = 0;
                for (uint32_t k = 0; k < n2; ++k) rank += __shfl_sync(FULL, x, (int)k) > x ? 1u : 0u;
                if (lane < n2) e2[rank] = x;
            }
        }
    }
}

cudaError_t rtb_grid_sort_on_reuse(RtbGrid& G, cudaStream_t st) {
    if (!G.valid || !G.keyed || G.sorted || G.pending || !G.d_ctl || !G.d_cell_start || !G.d_entries) return cudaSuccess;
    GridArgs g; memset(&g, 0, sizeof(g));
    g.ctl = G.d_ctl; g.cell_start = G.d_cell_start; g.entries = G.d_entries; g.entries_alloc = G.entries_cap;
    k_grid_sort<<<148 * 8, 256, 0, st>>>(g);
    G.sorted = true;
    return cudaGetLastError();
}

// ---- the shadow kernel of the grid path -------------------------------------------------------------------------------
// Any-hit BVH walk of one ray (per-lane stack): the fallback for a ray whose origin is not clearly below its light point.
template <bool COUNT>
__device__ __noinline__ bool lane_anyhit_walk(const TraceArgs& a, f3 o, f3 d, float L, int self_prim, const RayCull& c,
                                              uint32_t& n_box, uint32_t& n_tri, uint32_t& overflow) {
    if (a.root == RTB_CHILD_NONE) return false;
    int stack[RTB_STACK_SIZE];
    int sp = 0, cur = a.root;
    for (;;) {
        if (cur >= 0) {
            const float4* np = a.nodes + 4 * (size_t)cur;
            const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3);
            const int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            float t0, t1;
            const bool h0 = node_test(c, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, t0);
            const bool h1 = node_test(c, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, t1) && (c1 != RTB_CHILD_NONE);
            if (COUNT) n_box += 2;
            if (h0 && h1) {
                const bool swap = t1 < t0;
                if (sp < RTB_STACK_SIZE) stack[sp++] = swap ? c0 : c1; else overflow = 1;
                cur = swap ? c1 : c0;
                continue;
            }
            if (h0) { cur = c0; continue; }
            if (h1) { cur = c1; continue; }
        } else {
            if (COUNT) n_tri += 1;
            if (prim_occludes(a.tris, ~cur, ~cur == self_prim, o, d, L, &c)) return true;
        }
        if (sp == 0) return false;
        cur = stack[--sp];
    }
}

#ifndef RTB_DIAG_GRID
#define RTB_DIAG_GRID 0      // 1: round statistics of k_shadow_grid in counters 16-30 (scripts/diag_grid2.py; needs -DRTB_NUM_CTRS=32)
#endif

// ---- the shadow kernel of the grid path: ray table and work queues in shared memory ---------------------------------------------------
// The candidate lists differ in length (most rays of a tile near a shadow edge have none, a few have several) and a ray stops at
// its first occluder, so a loop per lane leaves two thirds of the lanes idle (ncu, profiles/r1c: 11.8 of 32 threads per
// instruction).  Round 1's kernel kept every ray in the registers of the lane that set it up and dealt the tests out round by
// round, per light sample (profiles/r2a: 680 thread instructions per traced ray for 1.5 Moller-Trumbore tests, a third of them
// the dealing).  Here a ticket is (32 consecutive queued hit points) x (R consecutive light samples) = 32 R rays whose
// state lives in SHARED memory from the start, and the work is organised as three dense stages over queues of ray ids:
//   1. lookup     lane = hit point, R cell lookups; a ray with a non-empty cell gets {cursor, remaining} in the ray table and
//                 its id in the test queue (59 % of the rays of big_bunny's queued hit points; the others are finished)
//   2. normalise  lane = queued ray: the exact direction and length (Ray::new, src/main.rs:201-202) only for rays that have
//                 something to test
//   3. test       lane = next ray of the queue (or, when fewer than 17 are left, a power-of-two share of candidates per ray):
//                 one prim_mt_q; a miss advances the ray's cursor IN the table and re-queues the ray, a hit moves it to the
//                 confirm queue, whose entries run prim_confirm_q 32 at a time.
// No per-sample loops, no owner lane, no verdict round trip; the same exact expressions decide every (ray, primitive) pair.
#ifndef RTB_GRID_R
#define RTB_GRID_R 4
#endif
#ifndef RTB_GRID_BLOCKS
#define RTB_GRID_BLOCKS 4
#endif
template <bool COUNT, int R>
__global__ void __launch_bounds__(256, RTB_GRID_BLOCKS) k_shadow_grid(const TraceArgs a) {
    constexpr int NR = 32 * R;                                   // rays per ticket
    constexpr int RING = NR <= 64 ? 64 : NR <= 128 ? 128 : NR <= 256 ? 256 : 512;
    constexpr uint32_t QM = RING - 1;
    static_assert(NR <= RING, "ring too small");
    __shared__ float4 s_o[8][32];                                // {o, self primitive} per hit point
    __shared__ uint32_t s_base[8][32];                           // light-point index of sample 0 per hit point
    __shared__ float4 s_dL[8][NR];                               // {d, L} per ray
    __shared__ uint2 s_cr[8][NR];                                // {next entry, entries left} per ray
    __shared__ uint2 s_tent[8][NR];                              // a candidate that passed prim_mt_q: {primitive, t}
    __shared__ uint16_t s_q[8][RING], s_q3[8][NR];               // test queue (ring), confirm queue (stack)
    __shared__ uint32_t s_occ[8][R];                             // result: bit (hit point) of word (sample)
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5, lt = (1u << lane) - 1u;
    const uint32_t n_items = a.queue_ctl[0];
    const uint32_t groups = (a.nb_light_sample + R - 1) / R;
    const uint32_t total = ((n_items + 31) / 32) * groups;
    uint32_t n_box = 0, n_tri = 0, overflow = 0;
    for (;;) {
        uint32_t tk = 0;
        if (lane == 0) tk = atomicAdd(a.queue_ctl + 8, 1u);
        tk = __shfl_sync(0xffffffffu, tk, 0);
        if (tk >= total) break;
        const uint32_t blk = tk / groups, i0 = (tk - blk * groups) * R;
        const uint32_t it = blk * 32 + lane;
        // ---- 1. lookups ------------------------------------------------------------------------------------------------------
        if (lane < R) s_occ[w][lane] = 0u;
        uint32_t walk = 0, below = 0, queued = 0;                    // bit r: ray r of this lane
        int self_prim = -1;
        f3 o = mk3(0.0f, 0.0f, 0.0f);
        uint32_t base_idx = 0;
        if (it < n_items) {
            const float4* ip = a.items + RTB_ITEM_F4 * (size_t)it;
            const float4 it0 = __ldg(ip), it1 = __ldg(ip + 1);
            self_prim = (int)(__float_as_uint(it0.y) & RTB_PRIM_MASK);
            o = mk3(it1.x, it1.y, it1.z);
            base_idx = __float_as_uint(it0.w);
        }
        s_o[w][lane] = make_float4(o.x, o.y, o.z, __int_as_float(self_prim));
        s_base[w][lane] = base_idx;
        uint32_t tail = 0, head = 0, n3 = 0;                         // warp-uniform queue counters
#pragma unroll
        for (int r = 0; r < R; ++r) {
            uint32_t cur = 0, rem = 0;
            if (it < n_items && i0 + r < a.nb_light_sample) {
                uint32_t idx = base_idx + i0 + r;                            // (r*NB_RAY + i) % len, src/main.rs:194
                if (idx >= a.n_samples) idx %= a.n_samples;
                const float4* gp = reinterpret_cast<const float4*>(a.grid_lights + idx);
                const float4 g0 = __ldg(gp), g1 = __ldg(gp + 1), g2 = __ldg(gp + 2);  // {px,py,pz,H} {x0,z0,cell,inv_cell} {nx,nz,cell_base,-}
                const float dy = g0.y - o.y;
                if (dy > RTB_GRID_TAU * g0.w) {                               // hence d.y = rn(dy / L) >= +0
                    below |= 1u << r;
                    const float s = g0.w / dy;
                    const float fx = (fmaf(o.x - g0.x, s, g0.x) - g1.x) * g1.w, fz = (fmaf(o.z - g0.z, s, g0.z) - g1.y) * g1.w;
                    const int nx = __float_as_int(g2.x), nz = __float_as_int(g2.y);
                    if (fx >= 0.0f && fz >= 0.0f && fx < (float)nx && fz < (float)nz) {   // false for NaN
                        const uint32_t cell = __float_as_uint(g2.z) + (uint32_t)fz * (uint32_t)nx + (uint32_t)fx;
                        cur = __ldg(a.grid_cell_start + cell);
                        rem = __ldg(a.grid_cell_start + cell + 1) - cur;
                    }
                } else {
                    walk |= 1u << r;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, rem > 0u);
            if (rem > 0u) {
                const uint32_t ray = (uint32_t)r * 32u + lane;
                s_cr[w][ray] = make_uint2(cur, rem);
                s_q[w][(tail + __popc(m & lt)) & QM] = (uint16_t)ray;
                queued |= 1u << r;
            }
            tail += __popc(m);
        }
        __syncwarp();
        // ---- 2. exact directions of the queued rays ----------------------------------------------------------------------------
        for (uint32_t k = lane; k < tail; k += 32) {
            const uint32_t ray = s_q[w][k];
            const float4 ro = s_o[w][ray & 31u];
            uint32_t idx = s_base[w][ray & 31u] + i0 + (ray >> 5);
            if (idx >= a.n_samples) idx %= a.n_samples;
            const float4 lp = __ldg(a.light_pts + idx);
            float L; f3 d;
            normalize_exact(sub3(mk3(lp.x, lp.y, lp.z), mk3(ro.x, ro.y, ro.z)), L, d);    // Ray::new + distance (src/main.rs:201-202)
            s_dL[w][ray] = make_float4(d.x, d.y, d.z, L);
        }
        __syncwarp();
        if (COUNT && RTB_DIAG_GRID && lane == 0) {
            atomicAdd(a.counters + 16, 1ull); atomicAdd(a.counters + 17, (unsigned long long)tail);
        }
        // ---- 3. tests ------------------------------------------------------------------------------------------------------------
        for (;;) {
            const uint32_t n = tail - head;
            if (COUNT && RTB_DIAG_GRID && lane == 0 && n > 0u && !(n3 >= 32u)) {
                atomicAdd(a.counters + 18, 1ull); atomicAdd(a.counters + 19, (unsigned long long)n);
                atomicAdd(a.counters + (n >= 32u ? 20 : n > 16u ? 21 : n > 8u ? 22 : n > 4u ? 23 : n > 2u ? 24 : 25), 1ull);
            }
            if (n3 >= 32u || (n == 0u && n3 > 0u)) {
                // confirm: the box / distance half of the test for rays whose candidate passed Moller-Trumbore
                const uint32_t cnt = n3 < 32u ? n3 : 32u;
                if (COUNT && RTB_DIAG_GRID && lane == 0) { atomicAdd(a.counters + 29, 1ull); atomicAdd(a.counters + 30, (unsigned long long)cnt); }
                if (lane < cnt) {
                    const uint32_t ray = s_q3[w][n3 - cnt + lane];
                    const uint2 tp = s_tent[w][ray];
                    const float4 rd = s_dL[w][ray], ro = s_o[w][ray & 31u];
                    const f3 po = mk3(ro.x, ro.y, ro.z), d = mk3(rd.x, rd.y, rd.z);
                    const int prim = (int)tp.x;
                    bool occluded = prim_confirm_q(__ldg(a.tris + 3 * prim), __ldg(a.tris + 3 * prim + 1), __ldg(a.tris + 3 * prim + 2), po, d, rd.w, __uint_as_float(tp.y));
                    if (!occluded) {                                      // rejected after all (rare): the rest of the list, from the round's start
                        const uint2 cr = s_cr[w][ray];
                        for (uint32_t k = 0; k < cr.y && !occluded; ++k) {
                            const int p2 = __ldg(a.grid_entries + cr.x + k) & a.grid_entry_mask;
                            if (p2 == __float_as_int(ro.w)) continue;
                            if (COUNT) n_tri += 1;
                            occluded = prim_occludes(a.tris, p2, false, po, d, rd.w, nullptr);
                        }
                    }
                    if (occluded) atomicOr(&s_occ[w][ray >> 5], 1u << (ray & 31u));
                }
                n3 -= cnt;
                __syncwarp();
                continue;
            }
            if (n == 0u) break;
            const uint32_t sh = n > 16u ? 0u : n > 8u ? 1u : n > 4u ? 2u : n > 2u ? 3u : n > 1u ? 4u : 5u;     // share = 1 << sh candidates per ray
            const uint32_t slots = n < (32u >> sh) ? n : (32u >> sh);
            const uint32_t rank = lane >> sh, off = lane & ((1u << sh) - 1u);
            const bool active = rank < slots;
            bool hit = false; int prim = -1; float t = 0.0f; uint32_t ray = 0; uint2 cr = make_uint2(0u, 0u);
            if (active) {
                ray = s_q[w][(head + rank) & QM];
                cr = s_cr[w][ray];
                if (off < cr.y) {
                    prim = __ldg(a.grid_entries + cr.x + off) & a.grid_entry_mask;
                    const float4 ro = s_o[w][ray & 31u];
                    if (prim != __float_as_int(ro.w)) {                              // the hit primitive: k_shade's per-ray self test
                        const float4 q0 = __ldg(a.tris + 3 * prim), q1 = __ldg(a.tris + 3 * prim + 1), q2 = __ldg(a.tris + 3 * prim + 2);
                        const f3 po = mk3(ro.x, ro.y, ro.z);
                        if (!prim_below_origin(q0, q1, q2, po)) {
                            const float4 rd = s_dL[w][ray];
                            if (COUNT) n_tri += 1;
                            hit = prim_mt_q(q0, q1, q2, po, mk3(rd.x, rd.y, rd.z), t);
                        }
                    }
                }
            }
            head += slots;
            const unsigned hm = __ballot_sync(0xffffffffu, hit);
            if (COUNT && RTB_DIAG_GRID) {
                const unsigned tested = __ballot_sync(0xffffffffu, active && off < cr.y);
                const unsigned skipped = __ballot_sync(0xffffffffu, active && off < cr.y && prim >= 0 && !hit && t == 0.0f);
                if (lane == 0) { atomicAdd(a.counters + 26, (unsigned long long)__popc(tested)); atomicAdd(a.counters + 27, (unsigned long long)__popc(hm)); atomicAdd(a.counters + 28, (unsigned long long)__popc(skipped)); }
            }
            const uint32_t seg = (hm >> (rank << sh)) & (sh == 5u ? 0xffffffffu : ((1u << (1u << sh)) - 1u));     // hits among this ray's lanes
            if (hit && !(seg & ((1u << off) - 1u))) s_tent[w][ray] = make_uint2((uint32_t)prim, __float_as_uint(t));    // lowest hit lane of the ray
            const bool leader = active && off == 0u;
            bool cont = false;
            if (leader && seg == 0u) {                                   // no candidate of this round hit: advance the ray's cursor in the table
                const uint32_t take = cr.y < (1u << sh) ? cr.y : (1u << sh);
                cr.x += take; cr.y -= take;
                s_cr[w][ray] = cr;
                cont = cr.y > 0u;
            }
            const unsigned m2 = __ballot_sync(0xffffffffu, cont);
            if (cont) s_q[w][(tail + __popc(m2 & lt)) & QM] = (uint16_t)ray;
            tail += __popc(m2);
            const bool to3 = leader && seg != 0u;
            const unsigned m3 = __ballot_sync(0xffffffffu, to3);
            if (to3) s_q3[w][n3 + __popc(m3 & lt)] = (uint16_t)ray;
            n3 += __popc(m3);
            __syncwarp();
        }
        // ---- 4. what the grids do not cover: un-rasterisable primitives (the ground), origins not clearly below the light point ---
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const bool valid = it < n_items && i0 + r < a.nb_light_sample;
            if (!valid) continue;
            bool occluded = (s_occ[w][r] >> lane) & 1u;
            const bool need_big = ((below >> r) & 1u) && !occluded && a.grid_n_big > 0u;
            const bool need_walk = (walk >> r) & 1u;
            if (!need_big && !need_walk) continue;
            bool have = (queued >> r) & 1u;
            float L = 0.0f; f3 d = mk3(0.0f, 0.0f, 0.0f);
            if (have) { const float4 rd = s_dL[w][r * 32 + lane]; d = mk3(rd.x, rd.y, rd.z); L = rd.w; }
            bool extra = false;
            if (need_big) {
                for (uint32_t k = 0; k < a.grid_n_big && !extra; ++k) {
                    const int prim = __ldg(a.grid_big + k);
                    if (prim == self_prim) continue;
                    const float4 q0 = __ldg(a.tris + 3 * prim), q1 = __ldg(a.tris + 3 * prim + 1), q2 = __ldg(a.tris + 3 * prim + 2);
                    if (prim_below_origin(q0, q1, q2, o)) continue;
                    if (!have) {
                        uint32_t idx = base_idx + i0 + r;
                        if (idx >= a.n_samples) idx %= a.n_samples;
                        const float4 lp = __ldg(a.light_pts + idx);
                        normalize_exact(sub3(mk3(lp.x, lp.y, lp.z), o), L, d);
                        have = true;
                    }
                    if (COUNT) n_tri += 1;
                    float t;
                    extra = prim_mt_q(q0, q1, q2, o, d, t) && prim_confirm_q(q0, q1, q2, o, d, L, t);
                }
            }
            if (need_walk) {
                if (!have) {
                    uint32_t idx = base_idx + i0 + r;
                    if (idx >= a.n_samples) idx %= a.n_samples;
                    const float4 lp = __ldg(a.light_pts + idx);
                    normalize_exact(sub3(mk3(lp.x, lp.y, lp.z), o), L, d);
                }
                extra = lane_anyhit_walk<COUNT>(a, o, d, L, -1, make_cull(o, d), n_box, n_tri, overflow);   // -1: the hit primitive is k_shade's
            }
            if (extra) atomicOr(&s_occ[w][r], 1u << lane);
        }
        __syncwarp();
        // transposed mask: one word per (block of 32 hit points, sample), bit = hit point within the block; every word written once
        if (lane < (uint32_t)R && i0 + lane < a.nb_light_sample) a.masks[(size_t)blk * a.nb_light_sample + i0 + lane] = s_occ[w][lane];
        __syncwarp();
    }
    if (__any_sync(0xffffffffu, overflow != 0) && lane == 0) atomicAdd(a.counters + RTB_CTR_OVERFLOW, 1ull);
    if (COUNT) {
        unsigned long long nb = n_box, nt = n_tri;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) { nb += __shfl_down_sync(0xffffffffu, nb, off); nt += __shfl_down_sync(0xffffffffu, nt, off); }
        if (lane == 0) { atomicAdd(a.counters + RTB_CTR_BOX, nb); atomicAdd(a.counters + RTB_CTR_TRI, nt); }
    }
}

cudaError_t rtb_launch_shadow_grid(const TraceArgs& a, bool count, unsigned blocks, cudaStream_t st) {
    const unsigned b2 = blocks / 8 * RTB_GRID_BLOCKS;        // blocks = 8 per SM
    if (count) k_shadow_grid<true, RTB_GRID_R><<<b2, 256, 0, st>>>(a);
    else k_shadow_grid<false, RTB_GRID_R><<<b2, 256, 0, st>>>(a);
    return cudaGetLastError();
}

// ---- host side: scene statistics and the per-light-table build ---------------------------------------------------------
__global__ void k_grid_init(unsigned int* __restrict__ lbounds, uint32_t n_lights, uint32_t* __restrict__ prim_flags, uint32_t n_tris, uint32_t* __restrict__ ctl) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_lights) { lbounds[4 * i] = 0xffffffffu; lbounds[4 * i + 1] = 0xffffffffu; lbounds[4 * i + 2] = 0u; lbounds[4 * i + 3] = 0u; }
    if (i < n_tris) prim_flags[i] = 0u;
    if (i < 12) ctl[i] = 0u;
}

#define GRID_CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

template <typename T>
static cudaError_t grid_reserve(T** p, size_t* cap, size_t need) {
    if (need <= *cap) return cudaSuccess;
    cudaFree(*p); *p = nullptr; *cap = 0;
    GRID_CK(cudaMalloc(p, need * sizeof(T)));
    *cap = need;
    return cudaSuccess;
}

cudaError_t rtb_grid_scene_stats(RtbGrid& G, const float4* d_tris, uint32_t n, cudaStream_t st) {
    G.stats_ok = false;
    if (n == 0) return cudaSuccess;
    unsigned int* d_u = nullptr; double* d_d = nullptr;
    GRID_CK(cudaMalloc(&d_u, 2 * sizeof(unsigned int)));
    cudaError_t e = cudaMalloc(&d_d, sizeof(double));
    unsigned int h_u[2] = {0xffffffffu, 0u}; double h_d = 0.0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_u, h_u, sizeof(h_u), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_d, &h_d, sizeof(h_d), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) { k_grid_scene_stats<<<(n + 255) / 256, 256, 0, st>>>(d_tris, n, d_u, d_d, d_u + 1); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_u, d_u, sizeof(h_u), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h_d, d_d, sizeof(h_d), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d_u); cudaFree(d_d);
    if (e != cudaSuccess) return e;
    if (h_u[0] != 0xffffffffu && h_u[1] > 0) {
        const unsigned int o = h_u[0];
        unsigned int bits = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
        float ymin; memcpy(&ymin, &bits, 4);
        G.y_min = ymin;
        G.mean_extent = (float)exp(h_d / (double)h_u[1]);
        G.stats_ok = ymin == ymin && G.mean_extent > 0.0f;
    }
    return cudaSuccess;
}

void rtb_grid_free(RtbGrid& G) {
    cudaFree(G.d_lights); cudaFree(G.d_lbounds); cudaFree(G.d_prim_flags); cudaFree(G.d_cell_start); cudaFree(G.d_cell_fill);
    cudaFree(G.d_entries); cudaFree(G.d_big); cudaFree(G.d_ctl); cudaFree(G.d_cub); cudaFree(G.d_union); cudaFree(G.d_sat);
    if (G.ev0) cudaEventDestroy(G.ev0);
    if (G.ev1) cudaEventDestroy(G.ev1);
    if (G.ev_count) cudaEventDestroy(G.ev_count);
    if (G.ev_fill) cudaEventDestroy(G.ev_fill);
    if (G.side) cudaStreamDestroy(G.side);
    memset(&G, 0, sizeof(G));
}

// Builds the grids for the light points d_light_pts[0..n_lights).  On success G.valid says whether the grid path applies
// (false: too many light points / primitives, too many un-rasterisable primitives, or the table would be too large).
//
// Two forms.  SYNCHRONOUS (first build of a scene, or after a failed speculative one): the host reads the cell and entry
// totals between the passes (two stream synchronisations) and sizes the arrays from them, with headroom.  SPECULATIVE
// (allow_spec, and a synchronous build of the same shape exists): the same kernels are enqueued back to back with NO host
// synchronisation, launched for the allocated capacity and checking themselves against it on the device (GridArgs.speculative);
// the caller's render kernels follow on the same stream, the caller synchronises once at the end of its frame and then calls
// rtb_grid_validate: if the build did not fit, the grids had switched themselves off (no out-of-bounds access, but wrong
// pixels), and the caller rebuilds synchronously and renders the frame again.  A new sample table (light grids) or a new
// camera pose (camera grid) of the same scene practically always fits the previous capacity.
static cudaError_t grid_build_ex(RtbGrid& G, const float4* d_tris, uint32_t n_tris, const float4* d_light_pts, uint32_t n_lights, float y_plane,
                                 float cell0, bool with_union, bool allow_spec, cudaStream_t st);

cudaError_t rtb_grid_build(RtbGrid& G, const float4* d_tris, uint32_t n_tris, const float4* d_light_pts, uint32_t n_lights, float cell_factor,
                           bool allow_spec, cudaStream_t st) {
    G.valid = false; G.n_lights = n_lights; G.n_big = 0; G.total_cells = 0; G.total_entries = 0; G.build_ms = 0.0f; G.pending = false;
    if (!G.stats_ok) return cudaSuccess;
    // the plane sits a little below the lowest vertex so that ground-level origins project with a finite scale
    return grid_build_ex(G, d_tris, n_tris, d_light_pts, n_lights, G.y_min - 1.0f - 1e-3f * fabsf(G.y_min), G.mean_extent * cell_factor, true, allow_spec, st);
}

// After the stream has been synchronised by the caller: the verdict of a pending speculative build.  *ok = false: the grids
// were switched off on the device and everything rendered with them must be redone after a synchronous build.
cudaError_t rtb_grid_validate(RtbGrid& G, cudaStream_t st, bool* ok) {
    *ok = true;
    if (!G.pending) return cudaSuccess;
    G.pending = false;
    uint32_t h_ctl[8];
    GRID_CK(cudaMemcpyAsync(h_ctl, G.d_ctl, sizeof(h_ctl), cudaMemcpyDeviceToHost, st));
    GRID_CK(cudaStreamSynchronize(st));
    if (h_ctl[4]) {
        static const bool dbg = getenv("RTB200_GRID_DEBUG") != nullptr;
        if (dbg) fprintf(stderr, "rtb200 grid: speculative build failed (bits %u): synchronous rebuild\n", h_ctl[4]);
        G.valid = false; G.spec_ok = false; G.spec_failures++;
        *ok = false;
        return cudaSuccess;
    }
    unsigned long long total; memcpy(&total, h_ctl + 6, 8);
    G.total_cells = h_ctl[0]; G.n_big = h_ctl[1]; G.total_entries = (uint32_t)total;
    if (cudaEventElapsedTime(&G.build_ms, G.ev0, G.ev1) != cudaSuccess) { G.build_ms = 0.0f; cudaGetLastError(); }
    return cudaSuccess;
}

static cudaError_t grid_build_ex(RtbGrid& G, const float4* d_tris, uint32_t n_tris, const float4* d_light_pts, uint32_t n_lights, float y_plane,
                                 float cell0, bool with_union, bool allow_spec, cudaStream_t st) {
    G.valid = false; G.n_lights = n_lights; G.n_big = 0; G.total_cells = 0; G.total_entries = 0; G.build_ms = 0.0f; G.pending = false;
    const uint64_t work = (uint64_t)n_lights * n_tris;
    static const bool dbg = getenv("RTB200_GRID_DEBUG") != nullptr;
    if (!(cell0 > 0.0f) || n_lights == 0 || n_lights > RTB_GRID_MAX_LIGHTS || n_tris == 0 || work > (256ull << 20)) {
        if (dbg) fprintf(stderr, "rtb200 grid: not applicable (cell0 %g, lights %u, tris %u, work %llu)\n", cell0, n_lights, n_tris, (unsigned long long)work);
        G.spec_ok = false;
        return cudaSuccess;
    }
    const size_t CELL_CAP = 512u << 20;
    // entries a build may have / should stay below.  A build for a table that keeps being reused (G.generous, set by the caller) is
    // paid once and may spend memory: 3 Gi entries (12 GB) allowed, cells coarsened only above 2 Gi
    const uint64_t ENTRY_CAP = G.entry_cap_limit ? G.entry_cap_limit : (G.generous ? (3072ull << 20) : (512ull << 20));
    // (64 bunnies, same table: 12.6 ms at 320 Mi, 9.4 at 640 Mi, 7.5 at 1 Gi, 6.9 ms at 2 Gi -- 1.68 G entries, 487 M cells of 1.9 units)
    const uint64_t BUDGET_DEFAULT = G.generous ? (2048ull << 20) : (320ull << 20);
    if (!G.ev0) { GRID_CK(cudaEventCreate(&G.ev0)); GRID_CK(cudaEventCreate(&G.ev1)); }
    GRID_CK(grid_reserve(&G.d_lights, &G.lights_cap, (size_t)n_lights));
    GRID_CK(grid_reserve(&G.d_lbounds, &G.lbounds_cap, (size_t)n_lights * 4));
    GRID_CK(grid_reserve(&G.d_prim_flags, &G.flags_cap, (size_t)n_tris));
    GRID_CK(grid_reserve(&G.d_big, &G.big_cap, (size_t)RTB_GRID_MAX_BIG));
    GRID_CK(grid_reserve(&G.d_ctl, &G.ctl_cap, (size_t)12));
    GRID_CK(grid_reserve(&G.d_union, &G.union_cap, (size_t)1));
    GridArgs g; memset(&g, 0, sizeof(g));
    g.tris = d_tris; g.n_tris = n_tris; g.light_pts = d_light_pts; g.n_lights = n_lights;
    g.y_plane = y_plane;
    g.cell0 = cell0;
    g.lights = G.d_lights; g.lbounds = G.d_lbounds; g.prim_flags = G.d_prim_flags; g.big_list = G.d_big; g.ctl = G.d_ctl; g.uni = G.d_union;
    g.cell_cap = (uint32_t)CELL_CAP;
    g.keyed = n_tris < (1u << 24) ? 1u : 0u;
    G.keyed = g.keyed != 0; G.sorted = false;
    const unsigned nmax = n_lights > n_tris ? (n_lights > 12 ? n_lights : 12) : (n_tris > 12 ? n_tris : 12);
    const unsigned gw = (unsigned)((work + 255) / 256);
    const unsigned gr = gw;                                                    // k_grid_raster: one thread per pair
    G.builds++;

    // ---- speculative form -------------------------------------------------------------------------------------------------
    const bool spec = allow_spec && G.spec_ok && G.spec_n_tris == n_tris && G.spec_with_union == with_union && G.spec_y_plane == y_plane && G.spec_req == cell0 &&
                      G.cell_start_cap > 1 && G.entries_cap > 1 && (!with_union || G.sat_cap > 0);
    if (spec) {
        g.cell0 = G.spec_cell0;
        const bool force_fail = getenv("RTB200_SPEC_FAIL") != nullptr;             // tests: every speculative build "does not fit"
        // the arrays may be far larger than this shape needs (a finer build for a reused table came before): the passes below run over
        // `nc` cells, the capacity a build of this shape was given, and the device-side check is against the same number
        size_t nc = (size_t)G.spec_cells + (size_t)G.spec_cells / 4 + 1024;
        if (nc > G.cell_start_cap) nc = G.cell_start_cap;
        g.speculative = 1u; g.cells_alloc = force_fail ? 1u : (uint32_t)nc; g.expect_big = G.n_big = G.spec_n_big_;
        g.sat_alloc = G.sat_cap; g.entries_alloc = G.entries_cap < ENTRY_CAP ? G.entries_cap : ENTRY_CAP;
        g.cell_start = G.d_cell_start; g.cell_fill = G.d_cell_fill; g.entries = G.d_entries; g.sat = G.d_sat;
        GRID_CK(cudaEventRecord(G.ev0, st));
        k_grid_init<<<(nmax + 255) / 256, 256, 0, st>>>(G.d_lbounds, n_lights, G.d_prim_flags, n_tris, G.d_ctl);
        k_grid_classify<<<(unsigned)((work + RTB_CLASSIFY_THREADS - 1) / RTB_CLASSIFY_THREADS), RTB_CLASSIFY_THREADS, 0, st>>>(g);
        k_grid_biglist<<<(n_tris + 255) / 256, 256, 0, st>>>(g);
        k_grid_layout<<<1, 1024, 0, st>>>(g);
        // (the scan covers all nc cells: those beyond the total hold zero counts)
        GRID_CK(cudaMemsetAsync(G.d_cell_fill, 0, nc * sizeof(uint32_t), st));
        k_grid_raster<false><<<gr, 256, 0, st>>>(g);
        size_t need = 0;
        GRID_CK(cub::DeviceScan::ExclusiveSum(nullptr, need, G.d_cell_fill, G.d_cell_start, (int)nc, st));
        if (need > G.cub_bytes) { cudaFree(G.d_cub); G.d_cub = nullptr; G.cub_bytes = 0; GRID_CK(cudaMalloc(&G.d_cub, need)); G.cub_bytes = need; }
        GRID_CK(cub::DeviceScan::ExclusiveSum(G.d_cub, need, G.d_cell_fill, G.d_cell_start, (int)nc, st));
        k_grid_validate<<<1, 32, 0, st>>>(g);
        // The fill pass is what a lookup needs; the union table below needs the offsets only.  With a union table (light grids) the
        // fill pass therefore goes to a side stream and runs next to whatever the caller enqueues on `st` before it makes its first
        // kernel that reads cell lists wait for G.ev_fill (rtb_api.cu: between k_primary and the shadow kernel of every chunk).
        cudaStream_t fs = st;
        if (with_union && !getenv("RTB200_NO_FILL_OVERLAP")) {
            if (!G.side) { GRID_CK(cudaStreamCreateWithFlags(&G.side, cudaStreamNonBlocking)); GRID_CK(cudaEventCreateWithFlags(&G.ev_count, cudaEventDisableTiming)); GRID_CK(cudaEventCreateWithFlags(&G.ev_fill, cudaEventDisableTiming)); }
            GRID_CK(cudaEventRecord(G.ev_count, st));
            GRID_CK(cudaStreamWaitEvent(G.side, G.ev_count, 0));
            fs = G.side;
        }
        GRID_CK(cudaMemsetAsync(G.d_cell_fill, 0, nc * sizeof(uint32_t), fs));
        k_grid_raster<true><<<gr, 256, 0, fs>>>(g);
        if (G.sort_at_build && g.keyed) { k_grid_sort<<<148 * 8, 256, 0, fs>>>(g); G.sorted = true; }
        if (fs != st) { GRID_CK(cudaEventRecord(G.ev_fill, fs)); G.fill_async = true; }
        if (with_union) {
            GRID_CK(cudaMemsetAsync(G.d_sat, 0, G.sat_cap * sizeof(uint32_t), st));
            k_grid_union_mark<<<148 * 8, 256, 0, st>>>(g);
            k_grid_sat_rows<<<(RTB_GRID_UNION_DIM + 1 + 7) / 8, 256, 0, st>>>(G.d_sat, G.d_union);
            k_grid_sat_cols<<<(RTB_GRID_UNION_DIM + 1 + 31) / 32, dim3(32, 32), 0, st>>>(G.d_sat, G.d_union);
        }
        GRID_CK(cudaGetLastError());
        GRID_CK(cudaEventRecord(G.ev1, st));
        G.pending = true; G.valid = true; G.spec_builds++;
        return cudaSuccess;
    }

    // ---- synchronous form -------------------------------------------------------------------------------------------------
    G.spec_ok = false;
    g.speculative = 0u; g.entries_alloc = ENTRY_CAP;
    GRID_CK(cudaEventRecord(G.ev0, st));
    uint32_t h_ctl[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    const unsigned long long ENTRY_BUDGET = ENTRY_CAP < BUDGET_DEFAULT ? ENTRY_CAP : BUDGET_DEFAULT;    // list entries one build should stay below
    // A primitive that would cover more than RTB_GRID_MAX_CELLS_PER_PRIM cells of some light's grid goes to the `big` list every
    // ray tests.  That is meant for the ground; in a wide scene (64 bunnies over 1600 x 1400 units) the stretched projections of
    // distant primitives cross the limit by the thousand at the cell size the primitives' own extent suggests.  Coarser cells
    // bring them back into the grids: the classification is repeated with doubled cells while the list overflows.
    for (int attempt = 0; ; ++attempt) {
        k_grid_init<<<(nmax + 255) / 256, 256, 0, st>>>(G.d_lbounds, n_lights, G.d_prim_flags, n_tris, G.d_ctl);
        k_grid_classify<<<(unsigned)((work + RTB_CLASSIFY_THREADS - 1) / RTB_CLASSIFY_THREADS), RTB_CLASSIFY_THREADS, 0, st>>>(g);
        k_grid_biglist<<<(n_tris + 255) / 256, 256, 0, st>>>(g);
        k_grid_layout<<<1, 1024, 0, st>>>(g);
        GRID_CK(cudaGetLastError());
        GRID_CK(cudaMemcpyAsync(h_ctl, G.d_ctl, sizeof(h_ctl), cudaMemcpyDeviceToHost, st));
        GRID_CK(cudaStreamSynchronize(st));
        unsigned long long est = 0; memcpy(&est, h_ctl + 8, 8);
        const bool too_big = h_ctl[1] > RTB_GRID_MAX_BIG, too_many = est > ENTRY_BUDGET;
        if ((!too_big && !too_many) || attempt >= 12) break;
        // entries fall between 1 / cell (streaks) and 1 / cell^2 (blobs): a modest step, repeated while needed
        g.cell0 *= too_big ? 2.0f : (est > 4 * ENTRY_BUDGET ? 2.0f : 1.3f);
        if (dbg) fprintf(stderr, "rtb200 grid: %u big primitives, about %llu entries: cells enlarged to %g\n", h_ctl[1], est, g.cell0);
    }
    G.total_cells = h_ctl[0]; G.n_big = h_ctl[1];
    if (dbg) fprintf(stderr, "rtb200 grid: %u lights, %u cells, %u big primitives (max %u)\n", n_lights, G.total_cells, G.n_big, (unsigned)RTB_GRID_MAX_BIG);
    if (G.n_big > RTB_GRID_MAX_BIG || G.total_cells > CELL_CAP) return cudaSuccess;
    // headroom: the next table / camera pose of this scene should fit without a synchronous rebuild
    const size_t nc = (size_t)G.total_cells + 1, nc_alloc = nc + nc / 4 + 1024;
    GRID_CK(grid_reserve(&G.d_cell_start, &G.cell_start_cap, nc_alloc));
    GRID_CK(grid_reserve(&G.d_cell_fill, &G.cell_fill_cap, G.cell_start_cap));
    g.cell_start = G.d_cell_start; g.cell_fill = G.d_cell_fill;
    GRID_CK(cudaMemsetAsync(G.d_cell_fill, 0, nc * sizeof(uint32_t), st));
    k_grid_raster<false><<<gr, 256, 0, st>>>(g);
    GRID_CK(cudaGetLastError());
    size_t need = 0;
    GRID_CK(cub::DeviceScan::ExclusiveSum(nullptr, need, G.d_cell_fill, G.d_cell_start, (int)nc, st));
    if (need > G.cub_bytes) { cudaFree(G.d_cub); G.d_cub = nullptr; G.cub_bytes = 0; GRID_CK(cudaMalloc(&G.d_cub, need)); G.cub_bytes = need; }
    GRID_CK(cub::DeviceScan::ExclusiveSum(G.d_cub, need, G.d_cell_fill, G.d_cell_start, (int)nc, st));
    // the entry total as a 64-bit sum (k_grid_raster<false>): the 32-bit scan above wraps silently beyond 2^32 entries, and an
    // array sized from a wrapped total would be written out of bounds by the fill pass
    unsigned long long h_total = 0;
    GRID_CK(cudaMemcpyAsync(&h_total, G.d_ctl + 6, sizeof(h_total), cudaMemcpyDeviceToHost, st));
    GRID_CK(cudaStreamSynchronize(st));
    if (dbg) fprintf(stderr, "rtb200 grid: %llu entries (cap %llu)\n", h_total, (unsigned long long)ENTRY_CAP);
    if (h_total > ENTRY_CAP || h_total > 0xfffffff0ull) return cudaSuccess;           // the caller takes the list / walk path
    G.total_entries = (uint32_t)h_total;
    GRID_CK(grid_reserve(&G.d_entries, &G.entries_cap, (size_t)h_total + (size_t)h_total / 4 + 1024));
    g.entries = G.d_entries; g.entries_alloc = G.entries_cap;
    GRID_CK(cudaMemsetAsync(G.d_cell_fill, 0, nc * sizeof(uint32_t), st));
    k_grid_raster<true><<<gr, 256, 0, st>>>(g);
    if (G.sort_at_build && g.keyed) { k_grid_sort<<<148 * 8, 256, 0, st>>>(g); G.sorted = true; }
    GRID_CK(cudaGetLastError());
    if (with_union) {   // union occupancy + summed-area table
        const int unx = (int)h_ctl[2], unz = (int)h_ctl[3];
        const size_t ns = (size_t)(unx + 1) * (size_t)(unz + 1);
        GRID_CK(grid_reserve(&G.d_sat, &G.sat_cap, ns + ns / 4 + 1024));
        g.sat = G.d_sat;
        GRID_CK(cudaMemsetAsync(G.d_sat, 0, ns * sizeof(uint32_t), st));
        if (G.total_cells) k_grid_union_mark<<<148 * 8, 256, 0, st>>>(g);
        if (unx > 0 && unz > 0) {
            k_grid_sat_rows<<<(unz + 7) / 8, 256, 0, st>>>(G.d_sat, G.d_union);
            k_grid_sat_cols<<<(unx + 31) / 32, dim3(32, 32), 0, st>>>(G.d_sat, G.d_union);
        }
        GRID_CK(cudaGetLastError());
    }
    GRID_CK(cudaEventRecord(G.ev1, st));
    GRID_CK(cudaEventSynchronize(G.ev1));
    GRID_CK(cudaEventElapsedTime(&G.build_ms, G.ev0, G.ev1));
    G.valid = true;
    G.spec_ok = true; G.spec_n_tris = n_tris; G.spec_cell0 = g.cell0; G.spec_req = cell0; G.spec_cells = G.total_cells + 1; G.spec_y_plane = y_plane; G.spec_with_union = with_union; G.spec_n_big_ = G.n_big;
    return cudaSuccess;
}

// ---- camera grid: the same structure for primary rays ---------------------------------------------------------------------
// Every primary ray starts at the eye: "which primitives can the ray through screen point (A, B) meet" is the same
// query with the eye as the projection centre and the image plane as the plane.  The primitives are written once in
// camera coordinates (x = c.u, y = -depth = c.w, z = c.v with c = vertex - eye; create_rays aims at A u + B v - D w,
// src/main.rs:159-174), so that the eye is the origin, the plane is y = -D and the projection of a vertex is
// (x, z) D / depth = its screen coordinates (A, B) in pixels.  One "light point" (the origin), cells of a few pixels.
// Primitives with a vertex nearer than tau D to the eye plane or covering too many cells (the ground) go to the big list.
__global__ void k_cam_tris(const float4* __restrict__ tris, uint32_t n, CameraBasis cam, float4* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float4 q = tris[3 * i + k];
        const f3 c = sub3(mk3(q.x, q.y, q.z), cam.eye);
        out[3 * i + k] = make_float4(dot3(c, cam.u), dot3(c, cam.w), dot3(c, cam.v), 0.0f);
    }
}

cudaError_t rtb_cam_grid_build(RtbGrid& G, const float4* d_tris, uint32_t n_tris, float4* d_cam_tris, float4* d_origin, const CameraBasis& cam,
                               float cell_px, bool allow_spec, cudaStream_t st) {
    G.valid = false; G.pending = false;
    if (n_tris == 0 || !(cam.distance > 0.0f)) return cudaSuccess;
    k_cam_tris<<<(n_tris + 255) / 256, 256, 0, st>>>(d_tris, n_tris, cam, d_cam_tris);
    GRID_CK(cudaMemsetAsync(d_origin, 0, sizeof(float4), st));
    GRID_CK(cudaGetLastError());
    return grid_build_ex(G, d_cam_tris, n_tris, d_origin, 1, -cam.distance, cell_px, false, allow_spec, st);
}
