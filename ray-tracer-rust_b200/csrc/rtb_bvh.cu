// rtb_bvh.cu -- device-side scene flattening and BVH construction.
//
// Replaces Triangle::new / get_bounding_box (src/tracer/primitives/triangle.rs:22-34,45-56) and
// BoundingVolumeHierarchy::new (src/tracer/utils/bounding_volume_hierarchy.rs:173-226).  The
// reference pairs nodes by the distance between box EXTENTS (bounding_box.rs:183-189), which
// yields a tree a ray has to visit ~500 nodes of; its topology is not part of the results
// contract (see rtb_trace.cu), so this builder makes its own tree: 63-bit Morton order +
// PLOC (parallel locally-ordered clustering: every cluster looks RADIUS neighbours left and
// right in Morton order for the partner with the smallest merged surface area, mutual nearest
// neighbours merge) repeated until one cluster is left.  Inner boxes are exact component-wise
// unions of exact leaf boxes -- the property the exactness argument needs.
#include "rtb_common.cuh"
#include <cub/cub.cuh>

#define PLOC_RADIUS 16

struct BuildTemps {
    // per primitive
    float4 *box_lo, *box_hi;          // exact leaf boxes (triangle.rs:45-56)
    unsigned long long *keys, *keys_alt;
    int *vals, *vals_alt;
    int *leaf_parent;
    // clusters (double buffered)
    int *cl_id[2]; float4 *cl_lo[2], *cl_hi[2];
    int *nn;
    unsigned long long *flags, *scan;
    // temporary binary tree, creation order
    int *n_left, *n_right, *n_parent; float4 *n_lo, *n_hi;
    unsigned int* bounds;             // 6 ordered-uint centroid bounds
    double* sah;
    unsigned int* depth;
    void* cub_tmp; size_t cub_bytes;
};

__device__ __forceinline__ unsigned int f2ord(float f) {
    unsigned int b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned int o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}
__device__ __forceinline__ float tri_min(float a, float b) { return a < b ? a : b; }   // triangle.rs:37-39
__device__ __forceinline__ float tri_max(float a, float b) { return a < b ? b : a; }   // triangle.rs:41-43

// AoS host layout -> SoA float4 {v0,n.x}{v1,n.y}{v2,n.z} + {colour, tie rank} + exact leaf box.
__global__ void k_prepare_tris(const float* __restrict__ verts, const float* __restrict__ colors,
                               const int* __restrict__ rank, uint32_t n, float4* __restrict__ tris,
                               float4* __restrict__ colors4, float4* __restrict__ box_lo, float4* __restrict__ box_hi,
                               unsigned int* __restrict__ bounds) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* p = verts + 9 * (size_t)i;
    const f3 v0 = mk3(p[0], p[1], p[2]), v1 = mk3(p[3], p[4], p[5]), v2 = mk3(p[6], p[7], p[8]);
    const f3 e1 = sub3(v1, v0), e2 = sub3(v2, v0);
    const f3 nrm = normalize3(cross3(e1, e2));                              // triangle.rs:29
    tris[3 * i] = make_float4(v0.x, v0.y, v0.z, nrm.x);
    tris[3 * i + 1] = make_float4(v1.x, v1.y, v1.z, nrm.y);
    tris[3 * i + 2] = make_float4(v2.x, v2.y, v2.z, nrm.z);
    const int rk = rank ? rank[i] : (int)i;
    colors4[i] = make_float4(colors[3 * i], colors[3 * i + 1], colors[3 * i + 2], __int_as_float(rk));
    const f3 lo = mk3(tri_min(tri_min(v0.x, v1.x), v2.x), tri_min(tri_min(v0.y, v1.y), v2.y), tri_min(tri_min(v0.z, v1.z), v2.z));
    const f3 hi = mk3(tri_max(tri_max(v0.x, v1.x), v2.x), tri_max(tri_max(v0.y, v1.y), v2.y), tri_max(tri_max(v0.z, v1.z), v2.z));
    box_lo[i] = make_float4(lo.x, lo.y, lo.z, 0.0f);
    box_hi[i] = make_float4(hi.x, hi.y, hi.z, 0.0f);
    const float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y), cz = 0.5f * (lo.z + hi.z);
    if (cx == cx && cy == cy && cz == cz) {
        atomicMin(bounds + 0, f2ord(cx)); atomicMin(bounds + 1, f2ord(cy)); atomicMin(bounds + 2, f2ord(cz));
        atomicMax(bounds + 3, f2ord(cx)); atomicMax(bounds + 4, f2ord(cy)); atomicMax(bounds + 5, f2ord(cz));
    }
}

__device__ __forceinline__ unsigned long long expand21(unsigned long long v) {
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}

__global__ void k_morton(const float4* __restrict__ box_lo, const float4* __restrict__ box_hi,
                         const unsigned int* __restrict__ bounds, uint32_t n, unsigned long long* __restrict__ keys,
                         int* __restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 lo = box_lo[i], hi = box_hi[i];
    const float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
    unsigned long long code = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float mn = ord2f(bounds[a]), mx = ord2f(bounds[3 + a]);
        const float ext = mx - mn;
        float q = ext > 0.0f ? (c[a] - mn) / ext : 0.0f;
        q = fminf(fmaxf(q, 0.0f), 1.0f);          // NaN -> 0
        unsigned long long iq = (unsigned long long)(q * 2097151.0f);
        code |= expand21(iq) << (2 - a);
    }
    keys[i] = code;
    vals[i] = (int)i;
}

__global__ void k_init_clusters(const int* __restrict__ sorted_prims, const float4* __restrict__ box_lo,
                                const float4* __restrict__ box_hi, uint32_t n, int* __restrict__ cl_id,
                                float4* __restrict__ cl_lo, float4* __restrict__ cl_hi) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int p = sorted_prims[i];
    cl_id[i] = ~p;
    cl_lo[i] = box_lo[p];
    cl_hi[i] = box_hi[p];
}

__device__ __forceinline__ float union_area(float4 alo, float4 ahi, float4 blo, float4 bhi) {
    const float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x);
    const float dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y);
    const float dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return dx * dy + dy * dz + dz * dx;
}

// nearest neighbour in the Morton-ordered window; ties go to the smaller index so that the pair
// with the globally smallest cost is always mutual (guarantees progress).
__global__ void k_ploc_nn(const float4* __restrict__ cl_lo, const float4* __restrict__ cl_hi, uint32_t m, int* __restrict__ nn) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const float4 lo = cl_lo[i], hi = cl_hi[i];
    const int b = (int)i - PLOC_RADIUS < 0 ? 0 : (int)i - PLOC_RADIUS;
    const int e = (int)i + PLOC_RADIUS >= (int)m ? (int)m - 1 : (int)i + PLOC_RADIUS;
    float best = __int_as_float(0x7f800000);
    int bj = -1;
    for (int j = b; j <= e; ++j) {
        if (j == (int)i) continue;
        const float a = union_area(lo, hi, cl_lo[j], cl_hi[j]);
        if (a < best || bj < 0) { best = a; bj = j; }
    }
    nn[i] = bj;
}

// flags[i] = (keep << 32) | merge
__global__ void k_ploc_flags(const int* __restrict__ nn, uint32_t m, unsigned long long* __restrict__ flags) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int j = nn[i];
    const bool mutual = j >= 0 && nn[j] == (int)i;
    const bool merge = mutual && (int)i < j;
    const bool keep = !mutual || merge;
    flags[i] = ((unsigned long long)(keep ? 1 : 0) << 32) | (merge ? 1ull : 0ull);
}

__global__ void k_ploc_emit(const int* __restrict__ nn, const unsigned long long* __restrict__ flags,
                            const unsigned long long* __restrict__ scan, uint32_t m, uint32_t node_base,
                            const int* __restrict__ cl_id, const float4* __restrict__ cl_lo, const float4* __restrict__ cl_hi,
                            int* __restrict__ out_id, float4* __restrict__ out_lo, float4* __restrict__ out_hi,
                            int* __restrict__ n_left, int* __restrict__ n_right, int* __restrict__ n_parent,
                            float4* __restrict__ n_lo, float4* __restrict__ n_hi, int* __restrict__ leaf_parent) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const unsigned long long f = flags[i];
    if (!(f >> 32)) return;
    const uint32_t pos = (uint32_t)(scan[i] >> 32);
    if (f & 1ull) {
        const int j = nn[i];
        const uint32_t node = node_base + (uint32_t)(scan[i] & 0xffffffffull);
        const float4 alo = cl_lo[i], ahi = cl_hi[i], blo = cl_lo[j], bhi = cl_hi[j];
        // exact union (bounding_box.rs:25-96): min/max select one of the inputs, no rounding
        const float4 lo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.0f);
        const float4 hi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f);
        const int l = cl_id[i], r = cl_id[j];
        n_left[node] = l; n_right[node] = r; n_lo[node] = lo; n_hi[node] = hi; n_parent[node] = -1;
        if (l >= 0) n_parent[l] = (int)node; else leaf_parent[~l] = (int)node;
        if (r >= 0) n_parent[r] = (int)node; else leaf_parent[~r] = (int)node;
        out_id[pos] = (int)node; out_lo[pos] = lo; out_hi[pos] = hi;
    } else {
        out_id[pos] = cl_id[i]; out_lo[pos] = cl_lo[i]; out_hi[pos] = cl_hi[i];
    }
}

// Temporary tree (creation order, root last) -> traversal nodes (root first: id' = n_inner-1-id), each
// holding both children's boxes.  Layout documented in rtb_trace.cu.
__global__ void k_emit_nodes(uint32_t n_inner, const int* __restrict__ n_left, const int* __restrict__ n_right,
                             const float4* __restrict__ n_lo, const float4* __restrict__ n_hi,
                             const float4* __restrict__ box_lo, const float4* __restrict__ box_hi,
                             float4* __restrict__ nodes, double* __restrict__ sah) {
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_inner) return;
    const int l = n_left[id], r = n_right[id];
    const float4 llo = l >= 0 ? n_lo[l] : box_lo[~l], lhi = l >= 0 ? n_hi[l] : box_hi[~l];
    const float4 rlo = r >= 0 ? n_lo[r] : box_lo[~r], rhi = r >= 0 ? n_hi[r] : box_hi[~r];
    const int lc = l >= 0 ? (int)(n_inner - 1 - l) : l;
    const int rc = r >= 0 ? (int)(n_inner - 1 - r) : r;
    float4* o = nodes + 4 * (size_t)(n_inner - 1 - id);
    o[0] = make_float4(llo.x, llo.y, llo.z, lhi.x);
    o[1] = make_float4(lhi.y, lhi.z, rlo.x, rlo.y);
    o[2] = make_float4(rlo.z, rhi.x, rhi.y, rhi.z);
    o[3] = make_float4(__int_as_float(lc), __int_as_float(rc), 0.0f, 0.0f);
    // SAH diagnostic: sum of child areas (relative to the root's, divided on the host)
    const double al = (double)union_area(llo, lhi, llo, lhi), ar = (double)union_area(rlo, rhi, rlo, rhi);
    atomicAdd(sah, (l >= 0 ? 1.2 : 1.0) * al + (r >= 0 ? 1.2 : 1.0) * ar);
    if (id == n_inner - 1) atomicAdd(sah + 1, (double)union_area(n_lo[id], n_hi[id], n_lo[id], n_hi[id]));
}

__global__ void k_depth(uint32_t n, const int* __restrict__ leaf_parent, const int* __restrict__ n_parent, unsigned int* __restrict__ depth) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned int d = 0;
    int p = leaf_parent[i];
    while (p >= 0) { ++d; p = n_parent[p]; }
    atomicMax(depth, d);
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = e_; goto done; } } while (0)

struct RtbBuildInfo { uint32_t n_nodes, max_depth, iterations; float sah_cost, build_ms; };

// d_verts/d_colors/d_rank: device copies of the host AoS input.  Outputs: d_tris (3n float4),
// d_colors4 (n float4) preallocated by the caller; *d_nodes_out allocated here (4*(n-1) float4).
cudaError_t rtb_build_scene(const float* d_verts, const float* d_colors, const int* d_rank, uint32_t n, float4* d_tris,
                            float4* d_colors4, float4** d_nodes_out, int32_t* root_out, RtbBuildInfo* info, cudaStream_t st) {
    cudaError_t err = cudaSuccess;
    BuildTemps T; memset(&T, 0, sizeof(T));
    *d_nodes_out = nullptr; *root_out = RTB_CHILD_NONE;
    memset(info, 0, sizeof(*info));
    if (n == 0) return cudaSuccess;
    const uint32_t n_inner = n - 1;
    const unsigned B = 256, G = (n + B - 1) / B;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float4* d_nodes = nullptr;
    unsigned int h_bounds[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    uint32_t m = n, node_base = 0, iters = 0;
    int cur = 0;
    double h_sah[2] = {0, 0};
    unsigned int h_depth = 0;
    cub::DoubleBuffer<unsigned long long> dk; cub::DoubleBuffer<int> dv;
    size_t sort_bytes = 0, scan_bytes = 0;

    CK(cudaEventCreate(&ev0)); CK(cudaEventCreate(&ev1));
    CK(cudaEventRecord(ev0, st));
    CK(cudaMalloc(&T.box_lo, n * sizeof(float4))); CK(cudaMalloc(&T.box_hi, n * sizeof(float4)));
    CK(cudaMalloc(&T.keys, n * 8)); CK(cudaMalloc(&T.keys_alt, n * 8));
    CK(cudaMalloc(&T.vals, n * 4)); CK(cudaMalloc(&T.vals_alt, n * 4));
    CK(cudaMalloc(&T.leaf_parent, n * 4));
    for (int k = 0; k < 2; ++k) {
        CK(cudaMalloc(&T.cl_id[k], n * 4)); CK(cudaMalloc(&T.cl_lo[k], n * sizeof(float4))); CK(cudaMalloc(&T.cl_hi[k], n * sizeof(float4)));
    }
    CK(cudaMalloc(&T.nn, n * 4)); CK(cudaMalloc(&T.flags, n * 8)); CK(cudaMalloc(&T.scan, n * 8));
    CK(cudaMalloc(&T.bounds, 6 * 4)); CK(cudaMalloc(&T.sah, 2 * sizeof(double))); CK(cudaMalloc(&T.depth, 4));
    if (n_inner) {
        CK(cudaMalloc(&T.n_left, n_inner * 4)); CK(cudaMalloc(&T.n_right, n_inner * 4)); CK(cudaMalloc(&T.n_parent, n_inner * 4));
        CK(cudaMalloc(&T.n_lo, n_inner * sizeof(float4))); CK(cudaMalloc(&T.n_hi, n_inner * sizeof(float4)));
        CK(cudaMalloc(&d_nodes, (size_t)n_inner * 4 * sizeof(float4)));
    }
    CK(cudaMemcpyAsync(T.bounds, h_bounds, sizeof(h_bounds), cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(T.sah, 0, 2 * sizeof(double), st)); CK(cudaMemsetAsync(T.depth, 0, 4, st));
    CK(cudaMemsetAsync(T.leaf_parent, 0xff, n * 4, st));

    k_prepare_tris<<<G, B, 0, st>>>(d_verts, d_colors, d_rank, n, d_tris, d_colors4, T.box_lo, T.box_hi, T.bounds);
    CK(cudaGetLastError());
    if (n == 1) { *root_out = ~0; info->n_nodes = 0; goto timing; }

    k_morton<<<G, B, 0, st>>>(T.box_lo, T.box_hi, T.bounds, n, T.keys, T.vals);
    CK(cudaGetLastError());
    dk = cub::DoubleBuffer<unsigned long long>(T.keys, T.keys_alt);
    dv = cub::DoubleBuffer<int>(T.vals, T.vals_alt);
    CK(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, dk, dv, (int)n, 0, 63, st));
    CK(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, T.flags, T.scan, (int)n, st));
    T.cub_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
    CK(cudaMalloc(&T.cub_tmp, T.cub_bytes ? T.cub_bytes : 16));
    CK(cub::DeviceRadixSort::SortPairs(T.cub_tmp, sort_bytes, dk, dv, (int)n, 0, 63, st));
    k_init_clusters<<<G, B, 0, st>>>(dv.Current(), T.box_lo, T.box_hi, n, T.cl_id[0], T.cl_lo[0], T.cl_hi[0]);
    CK(cudaGetLastError());

    while (m > 1) {
        const unsigned g = (m + B - 1) / B;
        k_ploc_nn<<<g, B, 0, st>>>(T.cl_lo[cur], T.cl_hi[cur], m, T.nn);
        k_ploc_flags<<<g, B, 0, st>>>(T.nn, m, T.flags);
        CK(cub::DeviceScan::ExclusiveSum(T.cub_tmp, scan_bytes, T.flags, T.scan, (int)m, st));
        k_ploc_emit<<<g, B, 0, st>>>(T.nn, T.flags, T.scan, m, node_base, T.cl_id[cur], T.cl_lo[cur], T.cl_hi[cur],
                                     T.cl_id[cur ^ 1], T.cl_lo[cur ^ 1], T.cl_hi[cur ^ 1], T.n_left, T.n_right, T.n_parent,
                                     T.n_lo, T.n_hi, T.leaf_parent);
        CK(cudaGetLastError());
        unsigned long long last_scan = 0, last_flag = 0;
        CK(cudaMemcpyAsync(&last_scan, T.scan + (m - 1), 8, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&last_flag, T.flags + (m - 1), 8, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        const unsigned long long tot = last_scan + last_flag;
        const uint32_t kept = (uint32_t)(tot >> 32), merged = (uint32_t)(tot & 0xffffffffull);
        if (merged == 0) { err = cudaErrorUnknown; goto done; }   // cannot happen (see k_ploc_nn)
        node_base += merged; m = kept; cur ^= 1; ++iters;
    }
    k_emit_nodes<<<(n_inner + B - 1) / B, B, 0, st>>>(n_inner, T.n_left, T.n_right, T.n_lo, T.n_hi, T.box_lo, T.box_hi, d_nodes, T.sah);
    k_depth<<<G, B, 0, st>>>(n, T.leaf_parent, T.n_parent, T.depth);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(h_sah, T.sah, sizeof(h_sah), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&h_depth, T.depth, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    *root_out = 0;
    info->n_nodes = n_inner; info->iterations = iters; info->max_depth = h_depth;
    info->sah_cost = h_sah[1] > 0 ? (float)(h_sah[0] / h_sah[1]) : 0.0f;
timing:
    CK(cudaEventRecord(ev1, st)); CK(cudaEventSynchronize(ev1));
    CK(cudaEventElapsedTime(&info->build_ms, ev0, ev1));
    *d_nodes_out = d_nodes; d_nodes = nullptr;
done:
    cudaFree(T.box_lo); cudaFree(T.box_hi); cudaFree(T.keys); cudaFree(T.keys_alt); cudaFree(T.vals); cudaFree(T.vals_alt);
    cudaFree(T.leaf_parent);
    for (int k = 0; k < 2; ++k) { cudaFree(T.cl_id[k]); cudaFree(T.cl_lo[k]); cudaFree(T.cl_hi[k]); }
    cudaFree(T.nn); cudaFree(T.flags); cudaFree(T.scan); cudaFree(T.bounds); cudaFree(T.sah); cudaFree(T.depth);
    cudaFree(T.n_left); cudaFree(T.n_right); cudaFree(T.n_parent); cudaFree(T.n_lo); cudaFree(T.n_hi); cudaFree(T.cub_tmp);
    cudaFree(d_nodes);
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    return err;
}
