// rtb_generate.cpp -- the reference's random scene generators (src/main.rs:43-99, both #[allow(dead_code)] there):
// gen_random_spheres (one sphere of radius in [400, 500) at (0, 0, -1000)) and gen_random_triangles (ten triangles with x, y in
// [-500, 500) and z in [-100, -50)), colours in [0, 1).  The reference draws from rand::thread_rng (OS-seeded, not reproducible);
// here the stream is SplitMix64 from an explicit seed, one 24-bit uniform per draw, mapped as rand 0.3's Range<f32> does
// (low + (high - low) * u), in the reference's order of draws.
#include "../../include/rtb200.h"
#include <cstddef>

namespace {
struct Stream {
    uint64_t s;
    float next() {                                  // top 24 bits of SplitMix64: u in [0, 1), exactly representable
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z = z ^ (z >> 31);
        return (float)(z >> 40) * (1.0f / 16777216.0f);
    }
    float range(float lo, float hi) { return lo + (hi - lo) * next(); }
};
}  // namespace

extern "C" uint32_t rtb_gen_random_spheres(uint64_t seed, float* spheres_xyzr, float* colors) {
    if (!spheres_xyzr || !colors) return 0;
    Stream r{seed};
    for (int i = 0; i < 1; ++i) {                   // src/main.rs:51: `for _ in 0..1`
        const float radius = r.range(400.0f, 500.0f);                                   // :55 (between_0_500 is Range::new(400, 500), :49)
        spheres_xyzr[4 * i] = 0.0f; spheres_xyzr[4 * i + 1] = 0.0f; spheres_xyzr[4 * i + 2] = -1000.0f; spheres_xyzr[4 * i + 3] = radius;   // :56-58
        colors[3 * i] = r.range(0.0f, 1.0f); colors[3 * i + 1] = r.range(0.0f, 1.0f); colors[3 * i + 2] = r.range(0.0f, 1.0f);             // :59-61
    }
    return 1;
}

extern "C" uint32_t rtb_gen_random_triangles(uint64_t seed, float* verts, float* colors) {
    if (!verts || !colors) return 0;
    Stream r{seed};
    for (int i = 0; i < 10; ++i) {                  // src/main.rs:78: `for _ in 0..10`
        for (int v = 0; v < 3; ++v) {               // :81-89: x, y in [-500, 500), z in [-100, -50) per vertex
            verts[9 * i + 3 * v] = r.range(-500.0f, 500.0f);
            verts[9 * i + 3 * v + 1] = r.range(-500.0f, 500.0f);
            verts[9 * i + 3 * v + 2] = r.range(-100.0f, -50.0f);
        }
        colors[3 * i] = r.range(0.0f, 1.0f); colors[3 * i + 1] = r.range(0.0f, 1.0f); colors[3 * i + 2] = r.range(0.0f, 1.0f);             // :90-92
    }
    return 10;
}

// sphere.rs:31-43 (HasBoundingBox for Sphere): n spheres {x, y, z, radius} -> n boxes {min.xyz, max.xyz}
extern "C" void rtb_spheres_bounds(const float* spheres_xyzr, uint32_t n, float* boxes6) {
    for (uint32_t i = 0; i < n; ++i) {
        const float* s = spheres_xyzr + 4 * (size_t)i;
        float* b = boxes6 + 6 * (size_t)i;
        b[0] = s[0] - s[3]; b[1] = s[1] - s[3]; b[2] = s[2] - s[3];
        b[3] = s[0] + s[3]; b[4] = s[1] + s[3]; b[5] = s[2] + s[3];
    }
}
