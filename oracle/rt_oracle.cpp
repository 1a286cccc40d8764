// rt_oracle.cpp -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
//
// A plain C++ restatement of the per-pixel ray-casting + shading path of
// antoinedesbois/Ray-Tracer-Rust, written from the reference's behaviour
// (citations below are file:line relative to the reference tree).  It exists
// only to check the CUDA path and to be timed as the CPU baseline.  Nothing in
// the product library links, imports or calls this file; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may use it.
//
// PARITY STATUS: "parity unpinned".  The reference ships no tests, no golden
// vectors and cannot be compiled here (no Rust toolchain, un-vendored crates:
// nalgebra 0.11.2, alga 0.5.1, image 0.12.3, rand 0.3.15 per Cargo.lock).  The
// oracle is validated by hand-derived known answers (tests/golden/) and by the
// equality of its two independent traversal modes (reference tree vs brute
// force).  nalgebra 0.11.2 arithmetic is restated from its published
// semantics: cross = (ay*bz-az*by, az*bx-ax*bz, ax*by-ay*bx); dot = fold from
// zero, left to right; norm = sqrt(dot(v,v)); normalize = per-component true
// division by the norm; distance(a,b) = norm(a-b).
//
// Build: g++ -O2 -ffp-contract=off (never -ffast-math).  All arithmetic is
// scalar IEEE binary32, no double intermediates, no fused multiply-add.

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cfloat>
#include <vector>
#include <thread>
#include <atomic>
#include <algorithm>
#include <chrono>

namespace {

struct V3 { float x, y, z; };

inline V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
inline V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 scale(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
// nalgebra dot: fold(zero, |acc,(a,b)| acc + a*b)
inline float dot(V3 a, V3 b) {
    float acc = 0.0f;
    acc = acc + a.x * b.x;
    acc = acc + a.y * b.y;
    acc = acc + a.z * b.z;
    return acc;
}
inline V3 cross(V3 a, V3 b) {
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float norm(V3 a) { return sqrtf(dot(a, a)); }
// Unit::new_normalize: n = norm(v); v / n component-wise
inline V3 normalize(V3 a) {
    float n = norm(a);
    return v3(a.x / n, a.y / n, a.z / n);
}
inline float distance(V3 a, V3 b) { return norm(sub(a, b)); }

struct Ray { V3 o, d; };
// src/tracer/utils/ray.rs:12-17 -- the constructor normalises the direction.
inline Ray make_ray(V3 o, V3 dir) { Ray r; r.o = o; r.d = normalize(dir); return r; }

// src/tracer/primitives/sphere.rs:50-84 (Intersectable for Sphere; dead code in the reference's main: SURVEY 8f-4)
inline bool sphere_intersect(V3 center, float radius, const Ray& ray, float* t_out) {
    const float radius2 = radius * radius;           // sphere.rs:27
    const V3 l = sub(center, ray.o);                 // :53
    const float tca = dot(l, ray.d);                 // :54
    if (tca < 0.0f) return false;                    // :55-57
    const float d2 = dot(l, l) - tca * tca;          // :58
    if (d2 > radius2) return false;                  // :59-61
    const float thc = sqrtf(radius2 - d2);           // :63
    float t0 = tca - thc, t1 = tca + thc;            // :65-66
    if (t0 > t1) { const float t = t0; t0 = t1; t1 = t; }   // :68-70
    if (t0 < 0.0f) {                                 // :73-78
        t0 = t1;
        if (t0 < 0.0f) return false;
    }
    const V3 p_hit = add(ray.o, scale(t0, ray.d));   // :80
    *t_out = distance(p_hit, ray.o);                 // :81
    return true;
}

struct Box { V3 mn, mx; };

struct Tri {
    V3 v0, v1, v2;     // triangle.rs:12-14
    V3 e1, e2;         // triangle.rs:23-24
    V3 normal;         // triangle.rs:29
    V3 color;          // triangle.rs:15
};

// src/tracer/primitives/triangle.rs:22-34
inline Tri make_tri(V3 v0, V3 v1, V3 v2, V3 color) {
    Tri t;
    t.v0 = v0; t.v1 = v1; t.v2 = v2;
    t.e1 = sub(v1, v0);
    t.e2 = sub(v2, v0);
    t.normal = normalize(cross(t.e1, t.e2));
    t.color = color;
    return t;
}

// triangle.rs:37-43
inline float min_float(float a, float b) { return a < b ? a : b; }
inline float max_float(float a, float b) { return a < b ? b : a; }

// triangle.rs:45-56
inline Box tri_box(const Tri& t) {
    Box b;
    b.mn = v3(min_float(min_float(t.v0.x, t.v1.x), t.v2.x),
              min_float(min_float(t.v0.y, t.v1.y), t.v2.y),
              min_float(min_float(t.v0.z, t.v1.z), t.v2.z));
    b.mx = v3(max_float(max_float(t.v0.x, t.v1.x), t.v2.x),
              max_float(max_float(t.v0.y, t.v1.y), t.v2.y),
              max_float(max_float(t.v0.z, t.v1.z), t.v2.z));
    return b;
}

// src/tracer/primitives/bounding_box.rs:25-96 (component-wise union)
inline Box box_union(const Box& l, const Box& r) {
    Box b;
    b.mn = v3(l.mn.x < r.mn.x ? l.mn.x : r.mn.x,
              l.mn.y < r.mn.y ? l.mn.y : r.mn.y,
              l.mn.z < r.mn.z ? l.mn.z : r.mn.z);
    b.mx = v3(l.mx.x > r.mx.x ? l.mx.x : r.mx.x,
              l.mx.y > r.mx.y ? l.mx.y : r.mx.y,
              l.mx.z > r.mx.z ? l.mx.z : r.mx.z);
    return b;
}

// bounding_box.rs:183-189 -- "centre" is really the extent (max - min).
inline V3 box_center(const Box& b) {
    return v3(b.mx.x - b.mn.x, b.mx.y - b.mn.y, b.mx.z - b.mn.z);
}

// bounding_box.rs:99-181.  Returns hit / no hit (the value is never used by
// the callers).  The reference's assert!(tmin <= tmax) can only fire through
// NaN; the oracle does not abort there (documented divergence) and reports it
// through *nan_assert if non-null.
inline bool box_intersect(const Box& b, const Ray& r, int* nan_assert = nullptr) {
    if (r.o.x > b.mn.x && r.o.x < b.mx.x &&
        r.o.y > b.mn.y && r.o.y < b.mx.y &&
        r.o.z > b.mn.z && r.o.z < b.mx.z) {
        return true;
    }
    const float t0 = 0.0f;
    const float t1 = FLT_MAX;
    float tmin, tmax, tymin, tymax, tzmin, tzmax;
    if (r.d.x >= 0.0f) {
        tmin = (b.mn.x - r.o.x) / r.d.x;
        tmax = (b.mx.x - r.o.x) / r.d.x;
    } else {
        tmin = (b.mx.x - r.o.x) / r.d.x;
        tmax = (b.mn.x - r.o.x) / r.d.x;
    }
    if (r.d.y >= 0.0f) {
        tymin = (b.mn.y - r.o.y) / r.d.y;
        tymax = (b.mx.y - r.o.y) / r.d.y;
    } else {
        tymin = (b.mx.y - r.o.y) / r.d.y;
        tymax = (b.mn.y - r.o.y) / r.d.y;
    }
    if (tmin > tymax || tymin > tmax) return false;
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    if (r.d.z >= 0.0f) {
        tzmin = (b.mn.z - r.o.z) / r.d.z;
        tzmax = (b.mx.z - r.o.z) / r.d.z;
    } else {
        tzmin = (b.mx.z - r.o.z) / r.d.z;
        tzmax = (b.mn.z - r.o.z) / r.d.z;
    }
    if (tmin > tzmax || tzmin > tmax) return false;
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    bool hit = tmin < t1 && tmax > t0;
    if (hit && !(tmin <= tmax) && nan_assert) *nan_assert = 1;
    return hit;
}

// triangle.rs:65-95 (Moller-Trumbore, two-sided, absolute epsilon 1e-5)
inline bool tri_intersect(const Tri& t, const Ray& r, float* out) {
    V3 pvec = cross(r.d, t.e2);
    float det = dot(t.e1, pvec);
    if (det < 0.00001f && det > -0.00001f) return false;
    float inv_det = 1.0f / det;
    V3 tvec = sub(r.o, t.v0);
    float u = dot(tvec, pvec) * inv_det;
    if (u < 0.0f || u > 1.0f) return false;
    V3 qvec = cross(tvec, t.e1);
    float v = dot(r.d, qvec) * inv_det;
    if (v < 0.0f || u + v > 1.0f) return false;
    *out = dot(t.e2, qvec) * inv_det;
    return true;
}

// triangle.rs:113-127 (note c3 uses v, not sqrt(v))
inline V3 tri_sample(const Tri& t, float u, float v) {
    float u_sqrt = sqrtf(u);
    float v_sqrt = sqrtf(v);
    float c1 = 1.0f - u_sqrt;
    float c2 = u_sqrt * (1.0f - v_sqrt);
    float c3 = v * u_sqrt;
    return v3(c1 * t.v0.x + c2 * t.v1.x + c3 * t.v2.x,
              c1 * t.v0.y + c2 * t.v1.y + c3 * t.v2.y,
              c1 * t.v0.z + c2 * t.v1.z + c3 * t.v2.z);
}

// src/tracer/utils/color.rs:10-33.  `as u8` saturates (and maps NaN to 0).
inline uint8_t gamma_u8(float linear) {
    const float GAMMA = 2.2f;
    float g = powf(linear, 1.0f / GAMMA) * 255.0f;
    if (!(g == g)) return 0;
    if (g <= 0.0f) return 0;
    if (g >= 255.0f) return 255;
    return (uint8_t)g;
}

struct Camera { V3 u, v, w, eye; float distance; };

// src/tracer/utils/camera.rs:17-35
inline Camera make_camera(V3 eye, V3 look_at, V3 up, float dist) {
    Camera c;
    c.w = normalize(sub(eye, look_at));
    V3 o = normalize(up);
    c.u = normalize(cross(o, c.w));
    c.v = normalize(cross(c.u, c.w));
    c.eye = eye;
    c.distance = dist;
    return c;
}

// ---------------------------------------------------------------------------
// Reference tree: src/tracer/utils/bounding_volume_hierarchy.rs
// ---------------------------------------------------------------------------
struct RNode {
    Box box;
    int prim;          // >= 0 for a leaf (bvh.rs:25-35), -1 for an inner node
    int left, right;   // bvh.rs:37-48
};

struct Hit { bool some; float t; int prim; };

struct Counters { uint64_t box_tests, tri_tests, nan_asserts; };

struct Scene {
    std::vector<Tri> tris;
    std::vector<Box> boxes;        // per-primitive leaf boxes
    std::vector<RNode> nodes;      // reference tree (empty if not built)
    int root;
    std::vector<int> dfs_order;    // primitive ids in left-to-right leaf order
    bool has_tree;
    std::vector<float> ksns;       // Phong extension: {Ks.rgb, Ns} per primitive; empty = the reference's shading
};

// bvh.rs:173-226.  Agglomerative pairing on the distance between box
// *extents* (bounding_box.rs:183-189), first minimum wins (strict <),
// parent(left = popped last, right = closest), odd node carried up.
void build_reference_tree(Scene& s) {
    const int n = (int)s.tris.size();
    s.nodes.clear();
    s.nodes.reserve(2 * (size_t)n);
    std::vector<int> level;
    level.reserve(n);
    for (int i = 0; i < n; ++i) {
        RNode nd; nd.box = s.boxes[i]; nd.prim = i; nd.left = nd.right = -1;
        s.nodes.push_back(nd);
        level.push_back(i);
    }
    std::vector<V3> ext;
    while (level.size() > 1) {
        std::vector<int> merged;
        ext.resize(level.size());
        for (size_t i = 0; i < level.size(); ++i) ext[i] = box_center(s.nodes[level[i]].box);
        while (level.size() > 1) {
            float min_dist = FLT_MAX;
            size_t min_idx = (size_t)-1;
            int last = level.back(); level.pop_back();
            V3 lc = ext.back(); ext.pop_back();
            for (size_t i = 0; i < level.size(); ++i) {
                float d = distance(lc, ext[i]);
                if (d < min_dist) { min_dist = d; min_idx = i; }
            }
            if (min_idx == (size_t)-1) min_idx = 0;  // reference would panic (all d >= MAX / NaN)
            int closest = level[min_idx];
            // Vec::swap_remove: move the last element into the hole
            level[min_idx] = level.back(); level.pop_back();
            ext[min_idx] = ext.back(); ext.pop_back();
            RNode nd;
            nd.box = box_union(s.nodes[last].box, s.nodes[closest].box);
            nd.prim = -1; nd.left = last; nd.right = closest;
            s.nodes.push_back(nd);
            merged.push_back((int)s.nodes.size() - 1);
        }
        if (level.size() == 1) { merged.push_back(level.back()); level.pop_back(); }
        level.swap(merged);
    }
    s.root = level.empty() ? -1 : level[0];
    s.has_tree = true;
    // left-to-right leaf order (used only for the exact-tie rule of brute force)
    s.dfs_order.clear();
    std::vector<int> stack;
    if (s.root >= 0) stack.push_back(s.root);
    while (!stack.empty()) {
        int id = stack.back(); stack.pop_back();
        const RNode& nd = s.nodes[id];
        if (nd.prim >= 0) { s.dfs_order.push_back(nd.prim); continue; }
        stack.push_back(nd.right);
        stack.push_back(nd.left);
    }
}

// bvh.rs:50-142: box test, leaf => primitive test with the t < 1.0 rejection,
// inner => both children unconditionally, left wins only if strictly closer.
Hit tree_intersect(const Scene& s, int id, const Ray& r, Counters& c) {
    const RNode& nd = s.nodes[id];
    Hit none = {false, 0.0f, -1};
    int na = 0;
    c.box_tests++;
    bool bh = box_intersect(nd.box, r, &na);
    c.nan_asserts += na;
    if (!bh) return none;
    if (nd.prim >= 0) {
        float x;
        c.tri_tests++;
        if (!tri_intersect(s.tris[nd.prim], r, &x)) return none;
        if (x < 1.0f) return none;
        Hit h = {true, x, nd.prim};
        return h;
    }
    Hit l = tree_intersect(s, nd.left, r, c);
    Hit rt = tree_intersect(s, nd.right, r, c);
    if (!l.some && !rt.some) return none;
    if (!l.some) return rt;
    if (!rt.some) return l;
    if (l.t < rt.t) return l;
    return rt;
}

// Topology-free restatement of the same result: every primitive whose own box
// passes the slab test, whose MT test passes and whose t is not < 1.0 is a
// candidate; minimum t wins, exact ties go to the later primitive in the
// reference tree's leaf order (or the higher index when no tree was built).
Hit brute_intersect(const Scene& s, const Ray& r, Counters& c) {
    Hit best = {false, 0.0f, -1};
    const int n = (int)s.tris.size();
    for (int k = 0; k < n; ++k) {
        int p = s.has_tree ? s.dfs_order[k] : k;
        c.box_tests++;
        if (!box_intersect(s.boxes[p], r)) continue;
        float x;
        c.tri_tests++;
        if (!tri_intersect(s.tris[p], r, &x)) continue;
        if (x < 1.0f) continue;
        if (!best.some || !(best.t < x)) { best.some = true; best.t = x; best.prim = p; }
    }
    return best;
}

// Any-hit form of the shadow query: occluded iff SOME accepted
// hit has ||fl(o + fl(x*d)) - o|| <= L.  Used by mode 2 to prove the early
// exit is exact.
bool brute_occluded(const Scene& s, const Ray& r, float L, Counters& c) {
    const int n = (int)s.tris.size();
    for (int p = 0; p < n; ++p) {
        float x;
        c.tri_tests++;
        if (!tri_intersect(s.tris[p], r, &x)) continue;
        if (x < 1.0f) continue;
        c.box_tests++;
        if (!box_intersect(s.boxes[p], r)) continue;
        V3 ph = add(r.o, scale(x, r.d));
        if (!(distance(r.o, ph) > L)) return true;
    }
    return false;
}

inline Hit scene_intersect(const Scene& s, const Ray& r, int mode, Counters& c) {
    if (mode == 0) {
        Hit none = {false, 0.0f, -1};
        return s.root >= 0 ? tree_intersect(s, s.root, r, c) : none;
    }
    return brute_intersect(s, r, c);
}

struct Params {
    uint32_t width, height, nb_ray, nb_light_sample;
    Camera cam;
    Tri light;
    const float* samples;   // n_samples pairs
    uint32_t n_samples;
    int mode;                // 0 reference tree, 1 brute force, 2 brute force + any-hit shadows
};

struct PixelOut {
    float rgb[3];
    uint8_t hit_count;   // primary rays (of nb_ray) that hit something
    int prim0;           // primitive hit by ray 0 (-1 = miss)
    float t0;            // its distance
    uint32_t n_shadow, n_occluded;
};

// src/main.rs:151-178 (create_rays) + src/main.rs:180-240 (render_pixel)
void render_pixel(const Scene& s, const Params& P, uint32_t px, uint32_t py, PixelOut& out, Counters& c) {
    float avg[3] = {0.0f, 0.0f, 0.0f};
    out.hit_count = 0; out.prim0 = -1; out.t0 = 0.0f; out.n_shadow = 0; out.n_occluded = 0;
    const float o_x = (float)px, o_y = (float)py;
    const float w = (float)P.width, h = (float)P.height;
    const float denom = (float)(P.nb_ray * P.nb_light_sample);
    for (uint32_t r = 0; r < P.nb_ray; ++r) {
        uint32_t k = (px * P.width + py + r) % P.n_samples;
        float A = o_x - w / 2.0f + P.samples[2 * k];
        float B = o_y - h / 2.0f + P.samples[2 * k + 1];
        V3 dir = sub(add(scale(A, P.cam.u), scale(B, P.cam.v)), scale(P.cam.distance, P.cam.w));
        Ray ray = make_ray(P.cam.eye, dir);
        Hit hit = scene_intersect(s, ray, P.mode, c);
        if (r == 0 && hit.some) { out.prim0 = hit.prim; out.t0 = hit.t; }
        if (!hit.some) continue;
        out.hit_count++;
        const Tri& T = s.tris[hit.prim];
        V3 p_hit = add(ray.o, scale(hit.t, ray.d));          // bvh.rs:69
        for (uint32_t i = 0; i < P.nb_light_sample; ++i) {
            size_t idx = ((size_t)r * P.nb_ray + i) % P.n_samples;   // main.rs:194
            V3 p = tri_sample(P.light, P.samples[2 * idx], P.samples[2 * idx + 1]);
            Ray sr = make_ray(p_hit, sub(p, p_hit));
            float L = distance(p, p_hit);
            float lnd = fabsf(dot(T.normal, sr.d));
            bool lit;
            out.n_shadow++;
            if (P.mode == 2) {
                lit = !brute_occluded(s, sr, L, c);
            } else {
                Hit h2 = scene_intersect(s, sr, P.mode, c);
                if (h2.some) {
                    V3 ph2 = add(sr.o, scale(h2.t, sr.d));
                    lit = distance(p_hit, ph2) > L;            // main.rs:220
                } else {
                    lit = true;
                }
            }
            if (lit && !s.ksns.empty()) {
                // Phong extension (NOT in the reference, which reads no materials: src/main.rs:130,137):
                // + Ks * max(0, r . v)^Ns with r = 2 |n . l| n_f - l, n_f the normal on the light's side, v towards the eye
                const float* k = &s.ksns[4 * (size_t)hit.prim];
                V3 nf = dot(T.normal, sr.d) >= 0.0f ? T.normal : V3{-T.normal.x, -T.normal.y, -T.normal.z};
                V3 rr = sub(scale(2.0f * lnd, nf), sr.d);
                Ray vr = make_ray(p_hit, sub(P.cam.eye, p_hit));
                float rv = dot(rr, vr.d);
                float sp = rv > 0.0f ? powf(rv, k[3]) : 0.0f;
                avg[0] = avg[0] + ((T.color.x * lnd + k[0] * sp) / denom);
                avg[1] = avg[1] + ((T.color.y * lnd + k[1] * sp) / denom);
                avg[2] = avg[2] + ((T.color.z * lnd + k[2] * sp) / denom);
            } else if (lit) {
                avg[0] = avg[0] + ((T.color.x * lnd) / denom);
                avg[1] = avg[1] + ((T.color.y * lnd) / denom);
                avg[2] = avg[2] + ((T.color.z * lnd) / denom);
            } else {
                out.n_occluded++;
                avg[0] = avg[0] + ((0.0f * 1.0f) / denom);
                avg[1] = avg[1] + ((0.0f * 1.0f) / denom);
                avg[2] = avg[2] + ((0.0f * 1.0f) / denom);
            }
        }
    }
    out.rgb[0] = avg[0]; out.rgb[1] = avg[1]; out.rgb[2] = avg[2];
}

}  // namespace

// ---------------------------------------------------------------------------
// C interface (loaded with ctypes by tests / bench; see oracle/oracle.py)
// ---------------------------------------------------------------------------
extern "C" {

struct orc_scene { Scene s; };

struct orc_params {
    uint32_t width, height, nb_ray, nb_light_sample;
    float eye[3], look_at[3], up[3], distance;
    float light[9];
    const float* samples;
    uint32_t n_samples;
    int32_t mode;        // 0 reference tree, 1 brute force, 2 brute + any-hit shadow
    int32_t n_threads;   // <= 0: hardware_concurrency() - 1 as src/main.rs:269 (min 1)
};

struct orc_stats {
    uint64_t primary_rays, primary_hits, shadow_rays, shadow_occluded;
    uint64_t box_tests, tri_tests, nan_asserts;
    double seconds;      // wall time of the worker fan-out/join (main.rs:273-305 span)
    int32_t threads;
};

// verts: n*9 floats (v0,v1,v2), colors: n*3 floats.  build_tree != 0 builds the
// reference tree (O(n^2) per level -- do not ask for it on huge scenes).
orc_scene* orc_scene_create(const float* verts, const float* colors, int32_t n, int32_t build_tree) {
    orc_scene* sc = new orc_scene();
    Scene& s = sc->s;
    s.root = -1; s.has_tree = false;
    s.tris.reserve(n); s.boxes.reserve(n);
    for (int i = 0; i < n; ++i) {
        const float* v = verts + 9 * (size_t)i;
        const float* c = colors + 3 * (size_t)i;
        Tri t = make_tri(v3(v[0], v[1], v[2]), v3(v[3], v[4], v[5]), v3(v[6], v[7], v[8]), v3(c[0], c[1], c[2]));
        s.tris.push_back(t);
        s.boxes.push_back(tri_box(t));
    }
    if (build_tree) build_reference_tree(s);
    return sc;
}

void orc_scene_destroy(orc_scene* sc) { delete sc; }

// Phong extension: n*4 floats {Ks.rgb, Ns}, or NULL for the reference's shading (see render_pixel)
void orc_scene_set_materials(orc_scene* sc, const float* ks_ns) {
    sc->s.ksns.clear();
    if (ks_ns) sc->s.ksns.assign(ks_ns, ks_ns + 4 * sc->s.tris.size());
}

int32_t orc_scene_num_nodes(const orc_scene* sc) { return (int32_t)sc->s.nodes.size(); }

// rank[p] = position of primitive p in the reference tree's left-to-right leaf order
int32_t orc_scene_dfs_rank(const orc_scene* sc, int32_t* rank) {
    if (!sc->s.has_tree) return -1;
    for (size_t k = 0; k < sc->s.dfs_order.size(); ++k) rank[sc->s.dfs_order[k]] = (int32_t)k;
    return 0;
}

void orc_tri_normal(const orc_scene* sc, int32_t p, float out[3]) {
    out[0] = sc->s.tris[p].normal.x; out[1] = sc->s.tris[p].normal.y; out[2] = sc->s.tris[p].normal.z;
}

void orc_camera_basis(const float eye[3], const float look_at[3], const float up[3], float u[3], float v[3], float w[3]) {
    Camera c = make_camera(v3(eye[0], eye[1], eye[2]), v3(look_at[0], look_at[1], look_at[2]), v3(up[0], up[1], up[2]), 1.0f);
    u[0] = c.u.x; u[1] = c.u.y; u[2] = c.u.z;
    v[0] = c.v.x; v[1] = c.v.y; v[2] = c.v.z;
    w[0] = c.w.x; w[1] = c.w.y; w[2] = c.w.z;
}

void orc_light_sample(const float tri[9], float a, float b, float out[3]) {
    Tri t = make_tri(v3(tri[0], tri[1], tri[2]), v3(tri[3], tri[4], tri[5]), v3(tri[6], tri[7], tri[8]), v3(1, 1, 1));
    V3 p = tri_sample(t, a, b);
    out[0] = p.x; out[1] = p.y; out[2] = p.z;
}

int32_t orc_box_intersect(const float mn[3], const float mx[3], const float o[3], const float d[3]) {
    Box b; b.mn = v3(mn[0], mn[1], mn[2]); b.mx = v3(mx[0], mx[1], mx[2]);
    Ray r; r.o = v3(o[0], o[1], o[2]); r.d = v3(d[0], d[1], d[2]);   // direction used as given
    return box_intersect(b, r) ? 1 : 0;
}

int32_t orc_tri_intersect(const float v[9], const float o[3], const float d[3], float* t) {
    Tri tr = make_tri(v3(v[0], v[1], v[2]), v3(v[3], v[4], v[5]), v3(v[6], v[7], v[8]), v3(1, 1, 1));
    Ray r; r.o = v3(o[0], o[1], o[2]); r.d = v3(d[0], d[1], d[2]);
    return tri_intersect(tr, r, t) ? 1 : 0;
}

void orc_primary_ray(const orc_params* p, uint32_t px, uint32_t py, uint32_t r, float o[3], float d[3]) {
    Camera cam = make_camera(v3(p->eye[0], p->eye[1], p->eye[2]), v3(p->look_at[0], p->look_at[1], p->look_at[2]),
                             v3(p->up[0], p->up[1], p->up[2]), p->distance);
    uint32_t k = (px * p->width + py + r) % p->n_samples;
    float A = (float)px - (float)p->width / 2.0f + p->samples[2 * k];
    float B = (float)py - (float)p->height / 2.0f + p->samples[2 * k + 1];
    V3 dir = sub(add(scale(A, cam.u), scale(B, cam.v)), scale(cam.distance, cam.w));
    Ray ray = make_ray(cam.eye, dir);
    o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z;
    d[0] = ray.d.x; d[1] = ray.d.y; d[2] = ray.d.z;
}

uint8_t orc_gamma_u8(float c) { return gamma_u8(c); }

// Renders `n_pix` pixels.  pixels == NULL: all width*height pixels, output index
// py*width+px.  Otherwise pixels = n_pix (px,py) uint32 pairs and output index =
// list position.  Worker threads take contiguous slices of the list in order,
// like src/main.rs:283-289 (the remainder, which the reference drops, is given
// to the last worker).  Any output pointer may be NULL.
int32_t orc_render(const orc_scene* sc, const orc_params* prm, const uint32_t* pixels, uint64_t n_pix,
                   float* rgb_lin, uint8_t* rgb8, uint8_t* hit_count, int32_t* prim0, float* t0,
                   orc_stats* stats) {
    const Scene& s = sc->s;
    if (prm->mode == 0 && !s.has_tree) return -1;
    if (prm->n_samples == 0 || !prm->samples) return -2;
    Params P;
    P.width = prm->width; P.height = prm->height; P.nb_ray = prm->nb_ray; P.nb_light_sample = prm->nb_light_sample;
    P.cam = make_camera(v3(prm->eye[0], prm->eye[1], prm->eye[2]), v3(prm->look_at[0], prm->look_at[1], prm->look_at[2]),
                        v3(prm->up[0], prm->up[1], prm->up[2]), prm->distance);
    const float* L = prm->light;
    P.light = make_tri(v3(L[0], L[1], L[2]), v3(L[3], L[4], L[5]), v3(L[6], L[7], L[8]), v3(1, 1, 1));
    P.samples = prm->samples; P.n_samples = prm->n_samples; P.mode = prm->mode;
    if (!pixels) n_pix = (uint64_t)prm->width * prm->height;

    int nt = prm->n_threads;
    if (nt <= 0) nt = (int)std::thread::hardware_concurrency() - 1;
    if (nt < 1) nt = 1;
    if ((uint64_t)nt > n_pix && n_pix > 0) nt = (int)n_pix;

    std::vector<Counters> ctr(nt);
    std::vector<uint64_t> hits(nt, 0), shadows(nt, 0), occl(nt, 0), prim_rays(nt, 0);
    auto worker = [&](int ti) {
        Counters c = {0, 0, 0};
        uint64_t per = n_pix / nt;
        uint64_t b = ti * per, e = (ti == nt - 1) ? n_pix : (ti + 1) * per;
        for (uint64_t j = b; j < e; ++j) {
            uint32_t px, py;
            if (pixels) { px = pixels[2 * j]; py = pixels[2 * j + 1]; }
            else { px = (uint32_t)(j % prm->width); py = (uint32_t)(j / prm->width); }
            PixelOut o;
            render_pixel(s, P, px, py, o, c);
            if (rgb_lin) { rgb_lin[3 * j] = o.rgb[0]; rgb_lin[3 * j + 1] = o.rgb[1]; rgb_lin[3 * j + 2] = o.rgb[2]; }
            if (rgb8) { rgb8[3 * j] = gamma_u8(o.rgb[0]); rgb8[3 * j + 1] = gamma_u8(o.rgb[1]); rgb8[3 * j + 2] = gamma_u8(o.rgb[2]); }
            if (hit_count) hit_count[j] = o.hit_count;
            if (prim0) prim0[j] = o.prim0;
            if (t0) t0[j] = o.t0;
            prim_rays[ti] += P.nb_ray; hits[ti] += o.hit_count; shadows[ti] += o.n_shadow; occl[ti] += o.n_occluded;
        }
        ctr[ti] = c;
    };
    auto t_start = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int i = 0; i < nt; ++i) th.emplace_back(worker, i);
    for (auto& t : th) t.join();
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        for (int i = 0; i < nt; ++i) {
            stats->primary_rays += prim_rays[i]; stats->primary_hits += hits[i];
            stats->shadow_rays += shadows[i]; stats->shadow_occluded += occl[i];
            stats->box_tests += ctr[i].box_tests; stats->tri_tests += ctr[i].tri_tests; stats->nan_asserts += ctr[i].nan_asserts;
        }
        stats->seconds = secs; stats->threads = nt;
    }
    return 0;
}

// ---- the Sphere primitive (sphere.rs), piecewise ----
// dir is normalised as Ray::new does (ray.rs:12-17); returns 1 and *t on a hit
int orc_sphere_intersect(const float* center, float radius, const float* o, const float* dir, float* t) {
    const Ray r = make_ray(v3(o[0], o[1], o[2]), v3(dir[0], dir[1], dir[2]));
    return sphere_intersect(v3(center[0], center[1], center[2]), radius, r, t) ? 1 : 0;
}
// sphere.rs:31-43
void orc_sphere_box(const float* c, float radius, float* out6) {
    out6[0] = c[0] - radius; out6[1] = c[1] - radius; out6[2] = c[2] - radius;
    out6[3] = c[0] + radius; out6[4] = c[1] + radius; out6[5] = c[2] + radius;
}
// sphere.rs:90-94
void orc_sphere_normal(const float* c, const float* p, float* out3) {
    const V3 n = normalize(sub(v3(p[0], p[1], p[2]), v3(c[0], c[1], c[2])));
    out3[0] = n.x; out3[1] = n.y; out3[2] = n.z;
}
// closest sphere of a list for one ray (the rule of the triangle path: smaller distance wins, an exact tie goes to the later
// primitive); the hit point is origin + t * direction as render_pixel forms it (src/main.rs:190), the normal sphere.rs:90-94
int orc_spheres_closest(const float* xyzr, int32_t n, const float* o, const float* dir, float* t_out, float* normal_out) {
    const Ray r = make_ray(v3(o[0], o[1], o[2]), v3(dir[0], dir[1], dir[2]));
    int best = -1; float bt = 0.0f;
    for (int i = 0; i < n; ++i) {
        float t;
        if (!sphere_intersect(v3(xyzr[4 * i], xyzr[4 * i + 1], xyzr[4 * i + 2]), xyzr[4 * i + 3], r, &t)) continue;
        if (best < 0 || t < bt || t == bt) { best = i; bt = t; }
    }
    if (best >= 0) {
        const V3 p = add(r.o, scale(bt, r.d));
        const V3 nn = normalize(sub(p, v3(xyzr[4 * best], xyzr[4 * best + 1], xyzr[4 * best + 2])));
        *t_out = bt; normal_out[0] = nn.x; normal_out[1] = nn.y; normal_out[2] = nn.z;
    }
    return best;
}

}  // extern "C"
