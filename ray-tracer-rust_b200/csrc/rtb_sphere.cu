// rtb_sphere.cu -- the Sphere primitive.  In the reference it is dead code: src/tracer/primitives/sphere.rs is complete
// (intersect :50-84, bounding box :31-43, normal :90-94) but main() only ever builds triangle scenes (gen_random_spheres,
// src/main.rs:43-67, is #[allow(dead_code)]).  It is kept here at the same level: the exact device function and a batch entry that
// finds the closest sphere per ray -- not wired into the frame kernels, whose grids and range proofs are written for triangles.
// Arithmetic: IEEE f32, unfused (-fmad=false), sqrtf / division correctly rounded, expression order of sphere.rs.
#include "rtb_common.cuh"
#include "rtb_device.cuh"
#include "../../include/rtb200.h"

extern "C" void rtb_set_error_(const char* msg);

// sphere.rs:50-84.  d must be the unit direction of Ray::new.  Returns the distance origin -> hit point as the reference does
// (distance(p_hit, origin), :81 -- not t0 itself).
__device__ __forceinline__ bool sphere_exact(f3 center, float radius, f3 o, f3 d, float& t_out) {
    const float radius2 = radius * radius;              // :27
    const f3 l = sub3(center, o);                       // :53
    const float tca = dot3(l, d);                       // :54
    if (tca < 0.0f) return false;                       // :55-57 (false for NaN: falls through, as in the reference)
    const float d2 = dot3(l, l) - tca * tca;            // :58
    if (d2 > radius2) return false;                     // :59-61
    const float thc = sqrtf(radius2 - d2);              // :63
    float t0 = tca - thc, t1 = tca + thc;               // :65-66
    if (t0 > t1) { const float t = t0; t0 = t1; t1 = t; }
    if (t0 < 0.0f) {                                    // :73-78
        t0 = t1;
        if (t0 < 0.0f) return false;
    }
    const f3 p_hit = add3(o, scale3(t0, d));            // :80
    t_out = norm3(sub3(p_hit, o));                      // :81
    return true;
}

// One thread per ray: Ray::new (ray.rs:12-17), every sphere tested, the smaller distance wins and an exact tie goes to the later
// sphere (the rule of the triangle path); hit point origin + t * direction (src/main.rs:190), normal sphere.rs:90-94.
__global__ void __launch_bounds__(256) k_spheres_closest(const float4* __restrict__ spheres, uint32_t n_spheres, const float* __restrict__ origins,
                                                         const float* __restrict__ dirs, uint64_t n_rays, float* __restrict__ t_out,
                                                         int32_t* __restrict__ idx_out, float* __restrict__ normal_out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rays) return;
    const f3 o = mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]);
    float L; f3 d;
    normalize_exact(mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]), L, d);
    int best = -1; float bt = 0.0f;
    for (uint32_t k = 0; k < n_spheres; ++k) {
        const float4 s = __ldg(spheres + k);
        float t;
        if (!sphere_exact(mk3(s.x, s.y, s.z), s.w, o, d, t)) continue;
        if (best < 0 || t <= bt) { best = (int)k; bt = t; }
    }
    idx_out[i] = best;
    t_out[i] = best >= 0 ? bt : 0.0f;
    f3 n = mk3(0.0f, 0.0f, 0.0f);
    if (best >= 0) {
        const float4 s = __ldg(spheres + best);
        const f3 p = add3(o, scale3(bt, d));
        float Ln;
        normalize_exact(sub3(p, mk3(s.x, s.y, s.z)), Ln, n);
    }
    normal_out[3 * i] = n.x; normal_out[3 * i + 1] = n.y; normal_out[3 * i + 2] = n.z;
}

#define SP_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { rtb_set_error_(cudaGetErrorString(e_)); rc = RTB_ERR_CUDA; goto done; } } while (0)

extern "C" int rtb_spheres_closest_hit(int32_t device, const float* spheres_xyzr, uint32_t n_spheres, const float* origins, const float* dirs,
                                       uint64_t n_rays, float* t_out, int32_t* idx_out, float* normal_out) {
    if (!origins || !dirs || !t_out || !idx_out || !normal_out || (n_spheres && !spheres_xyzr)) { rtb_set_error_("rtb_spheres_closest_hit: null argument"); return RTB_ERR_ARG; }
    if (n_rays == 0) return RTB_OK;
    int rc = RTB_OK;
    float4* d_s = nullptr; float *d_o = nullptr, *d_d = nullptr, *d_t = nullptr, *d_n = nullptr; int32_t* d_i = nullptr;
    SP_TRY(cudaSetDevice(device));
    SP_TRY(cudaMalloc(&d_s, (size_t)(n_spheres ? n_spheres : 1) * sizeof(float4)));
    SP_TRY(cudaMalloc(&d_o, n_rays * 12)); SP_TRY(cudaMalloc(&d_d, n_rays * 12)); SP_TRY(cudaMalloc(&d_t, n_rays * 4));
    SP_TRY(cudaMalloc(&d_n, n_rays * 12)); SP_TRY(cudaMalloc(&d_i, n_rays * 4));
    if (n_spheres) SP_TRY(cudaMemcpy(d_s, spheres_xyzr, (size_t)n_spheres * sizeof(float4), cudaMemcpyHostToDevice));
    SP_TRY(cudaMemcpy(d_o, origins, n_rays * 12, cudaMemcpyHostToDevice));
    SP_TRY(cudaMemcpy(d_d, dirs, n_rays * 12, cudaMemcpyHostToDevice));
    k_spheres_closest<<<(unsigned)((n_rays + 255) / 256), 256>>>(d_s, n_spheres, d_o, d_d, n_rays, d_t, d_i, d_n);
    SP_TRY(cudaGetLastError());
    SP_TRY(cudaMemcpy(t_out, d_t, n_rays * 4, cudaMemcpyDeviceToHost));
    SP_TRY(cudaMemcpy(idx_out, d_i, n_rays * 4, cudaMemcpyDeviceToHost));
    SP_TRY(cudaMemcpy(normal_out, d_n, n_rays * 12, cudaMemcpyDeviceToHost));
done:
    cudaFree(d_s); cudaFree(d_o); cudaFree(d_d); cudaFree(d_t); cudaFree(d_n); cudaFree(d_i);
    return rc;
}
