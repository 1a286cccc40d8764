// rtb_refpair.cu -- tie ranks: the left-to-right leaf order of the REFERENCE's tree.
//
// The only thing of BoundingVolumeHierarchy::new (bounding_volume_hierarchy.rs:173-226) that can
// reach a pixel is the order of its leaves: when two primitives are hit at exactly the same
// distance the later leaf wins (bounding_volume_hierarchy.rs:123-130).  This kernel replays the
// reference's pairing on the device -- pop the last node, pick the FIRST node minimising the
// distance between box EXTENTS (bounding_box.rs:183-189, strict `<`), Vec::swap_remove it,
// parent(left = popped, right = picked), carry an odd node to the end of the next level -- and
// converts the resulting tree to leaf ranks.  The pops are inherently sequential, so one
// 1024-thread block does the argmin of each pop in parallel (O(n^2) work per level, as in the
// reference; used for n <= RTB_TIE_RANK_MAX_TRIS only).
#include "rtb_common.cuh"

#define RP_THREADS 1024
#define RP_MAX_LEVELS 64

struct RefPairBufs {
    float4 *lo, *hi;        // boxes of all 2n-1 nodes (leaves 0..n-1, inner n..)
    int *left, *right;      // children of inner nodes
    int *count, *first;     // leaves below / rank of first leaf below
    int *levelA, *levelB;   // node lists
    float4* ext;            // extents of the current list, same positions
};

__device__ __forceinline__ void argmin_combine(float& d, int& i, float d2, int i2) {
    // smaller distance wins; equal distances keep the smaller index (first minimum, strict <)
    if (i2 >= 0 && (i < 0 || d2 < d || (d2 == d && i2 < i))) { d = d2; i = i2; }
}

__global__ void __launch_bounds__(RP_THREADS, 1)
k_ref_pairing(const float4* __restrict__ tris, uint32_t n, RefPairBufs B, float4* __restrict__ colors4) {
    __shared__ float s_d[RP_THREADS / 32];
    __shared__ int s_i[RP_THREADS / 32];
    __shared__ int s_pick;
    __shared__ int s_level_begin[RP_MAX_LEVELS + 1];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

    // leaves: exact boxes (triangle.rs:45-56)
    for (uint32_t i = tid; i < n; i += RP_THREADS) {
        const float4 q0 = tris[3 * i], q1 = tris[3 * i + 1], q2 = tris[3 * i + 2];
        const float lx = (q0.x < q1.x ? q0.x : q1.x), ly = (q0.y < q1.y ? q0.y : q1.y), lz = (q0.z < q1.z ? q0.z : q1.z);
        const float hx = (q0.x < q1.x ? q1.x : q0.x), hy = (q0.y < q1.y ? q1.y : q0.y), hz = (q0.z < q1.z ? q1.z : q0.z);
        B.lo[i] = make_float4(lx < q2.x ? lx : q2.x, ly < q2.y ? ly : q2.y, lz < q2.z ? lz : q2.z, 0.0f);
        B.hi[i] = make_float4(hx < q2.x ? q2.x : hx, hy < q2.y ? q2.y : hy, hz < q2.z ? q2.z : hz, 0.0f);
        B.count[i] = 1;
        B.levelA[i] = (int)i;
    }
    __syncthreads();

    int* level = B.levelA;
    int* merged = B.levelB;
    int len = (int)n;
    int next_node = (int)n;
    int n_levels = 0;
    while (len > 1 && n_levels < RP_MAX_LEVELS) {
        if (tid == 0) s_level_begin[n_levels] = next_node;
        for (int i = tid; i < len; i += RP_THREADS) {
            const float4 l = B.lo[level[i]], h = B.hi[level[i]];
            B.ext[i] = make_float4(h.x - l.x, h.y - l.y, h.z - l.z, 0.0f);      // bounding_box.rs:183-189
        }
        __syncthreads();
        int mc = 0;
        while (len > 1) {
            const int last = level[len - 1];
            const float4 lc = B.ext[len - 1];
            len -= 1;
            float bd = 0.0f; int bi = -1;
            for (int i = tid; i < len; i += RP_THREADS) {
                const float4 e = B.ext[i];
                const float dx = lc.x - e.x, dy = lc.y - e.y, dz = lc.z - e.z;
                const float d = sqrtf((dx * dx + dy * dy) + dz * dz);            // nalgebra distance
                if (d < FLT_MAX) argmin_combine(bd, bi, d, i);                   // `d < min_dist`, min_dist starts at f32::MAX
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const float d2 = __shfl_down_sync(0xffffffffu, bd, off);
                const int i2 = __shfl_down_sync(0xffffffffu, bi, off);
                argmin_combine(bd, bi, d2, i2);
            }
            if (lane == 0) { s_d[wid] = bd; s_i[wid] = bi; }
            __syncthreads();
            if (wid == 0) {
                bd = s_d[lane]; bi = s_i[lane];
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
                    const float d2 = __shfl_down_sync(0xffffffffu, bd, off);
                    const int i2 = __shfl_down_sync(0xffffffffu, bi, off);
                    argmin_combine(bd, bi, d2, i2);
                }
                if (lane == 0) {
                    const int pick = bi < 0 ? 0 : bi;          // the reference would panic here (all NaN / >= MAX)
                    const int closest = level[pick];
                    level[pick] = level[len - 1];              // Vec::swap_remove
                    B.ext[pick] = B.ext[len - 1];
                    const float4 alo = B.lo[last], ahi = B.hi[last], blo = B.lo[closest], bhi = B.hi[closest];
                    const int id = next_node + mc;
                    B.lo[id] = make_float4(alo.x < blo.x ? alo.x : blo.x, alo.y < blo.y ? alo.y : blo.y, alo.z < blo.z ? alo.z : blo.z, 0.0f);
                    B.hi[id] = make_float4(ahi.x > bhi.x ? ahi.x : bhi.x, ahi.y > bhi.y ? ahi.y : bhi.y, ahi.z > bhi.z ? ahi.z : bhi.z, 0.0f);
                    B.left[id] = last; B.right[id] = closest;
                    B.count[id] = B.count[last] + B.count[closest];
                    merged[mc] = id;
                    s_pick = pick;
                }
            }
            __syncthreads();
            len -= 1;
            mc += 1;
        }
        if (len == 1) { if (tid == 0) merged[mc] = level[0]; mc += 1; }
        __syncthreads();
        next_node += (len == 1) ? mc - 1 : mc;
        { int* t = level; level = merged; merged = t; }
        len = mc;
        n_levels += 1;
    }
    if (tid == 0) s_level_begin[n_levels] = next_node;
    __syncthreads();
    (void)s_pick;

    // ranks: root starts at 0; left subtree first, then right (left-to-right DFS of the leaves)
    const int root = level[0];
    if (tid == 0) B.first[root] = 0;
    __syncthreads();
    for (int lv = n_levels - 1; lv >= 0; --lv) {
        const int b = s_level_begin[lv], e = s_level_begin[lv + 1];
        for (int id = b + tid; id < e; id += RP_THREADS) {
            const int f = B.first[id], l = B.left[id], r = B.right[id];
            B.first[l] = f;
            B.first[r] = f + B.count[l];
        }
        __syncthreads();
    }
    for (uint32_t i = tid; i < n; i += RP_THREADS) colors4[i].w = __int_as_float(B.first[i]);
}

cudaError_t rtb_tie_rank_device(const float4* d_tris, uint32_t n, int* d_rank_out, float4* d_colors4, cudaStream_t st) {
    (void)d_rank_out;
    if (n < 2) return cudaSuccess;
    RefPairBufs B; memset(&B, 0, sizeof(B));
    cudaError_t err = cudaSuccess;
    const size_t nn = 2 * (size_t)n;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = e_; goto done; } } while (0)
    CK(cudaMalloc(&B.lo, nn * sizeof(float4))); CK(cudaMalloc(&B.hi, nn * sizeof(float4)));
    CK(cudaMalloc(&B.left, nn * 4)); CK(cudaMalloc(&B.right, nn * 4)); CK(cudaMalloc(&B.count, nn * 4)); CK(cudaMalloc(&B.first, nn * 4));
    CK(cudaMalloc(&B.levelA, (size_t)n * 4)); CK(cudaMalloc(&B.levelB, (size_t)n * 4)); CK(cudaMalloc(&B.ext, (size_t)n * sizeof(float4)));
    k_ref_pairing<<<1, RP_THREADS, 0, st>>>(d_tris, n, B, d_colors4);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
#undef CK
done:
    cudaFree(B.lo); cudaFree(B.hi); cudaFree(B.left); cudaFree(B.right); cudaFree(B.count); cudaFree(B.first);
    cudaFree(B.levelA); cudaFree(B.levelB); cudaFree(B.ext);
    return err;
}
