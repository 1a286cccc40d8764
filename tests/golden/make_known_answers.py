"""Re-derives the numeric entries of known_answers.json WITHOUT the oracle: numpy float32 step-by-step evaluation of
the cited reference formulas (camera.rs:17-35, triangle.rs:113-127, color.rs:10-33, src/main.rs:151-178) and a
float64 cross-check.  Run: python tests/golden/make_known_answers.py  (prints the values; exits 1 on disagreement
with the committed JSON)."""
import json
import os
import sys

import numpy as np

f32 = np.float32
HERE = os.path.dirname(os.path.abspath(__file__))


def norm(v):
    s = f32(0)
    for x in v:
        s = f32(s + f32(x * x))
    return f32(np.sqrt(s))


def normalize(v):
    n = norm(v)
    return [f32(x / n) for x in v]


def cross(a, b):
    return [f32(f32(a[1] * b[2]) - f32(a[2] * b[1])), f32(f32(a[2] * b[0]) - f32(a[0] * b[2])),
            f32(f32(a[0] * b[1]) - f32(a[1] * b[0]))]


def camera(eye, look, up):
    w = normalize([f32(e) - f32(l) for e, l in zip(eye, look)])
    o = normalize([f32(x) for x in up])
    u = normalize(cross(o, w))
    v = normalize(cross(u, w))
    return u, v, w


def light_sample(tri, a, b):
    a, b = f32(a), f32(b)
    us, vs = f32(np.sqrt(a)), f32(np.sqrt(b))
    c1, c2, c3 = f32(f32(1) - us), f32(us * f32(f32(1) - vs)), f32(b * us)
    t = [f32(x) for x in tri]
    return [f32(f32(f32(c1 * t[k]) + f32(c2 * t[3 + k])) + f32(c3 * t[6 + k])) for k in range(3)]


def gamma_u8(c):
    g = float(c) ** (1.0 / float(f32(2.2))) * 255.0   # float64; the listed inputs are far from byte boundaries
    return int(min(255.0, max(0.0, g)))


def main():
    ka = json.load(open(os.path.join(HERE, "known_answers.json")))
    ok = True
    cam = ka["camera_default"]
    u, v, w = camera(cam["eye"], cam["look_at"], cam["up"])
    print("camera u v w", u, v, w)
    for name, got in (("u", u), ("v", v), ("w", w)):
        ok &= np.allclose(np.array(got, dtype=np.float64), cam[name], rtol=cam["rtol"], atol=1e-12)
    light = [-10.0, 300.0, -10.0, 10.0, 300.0, -10.0, 0.0, 300.0, 0.0]
    for e in ka["light_sample"]:
        p = light_sample(light, *e["uv"])
        print("light", e["uv"], p)
        ok &= [float(x) for x in p] == e["p"]
    for c, b in ka["gamma_u8"]:
        print("gamma", c, gamma_u8(c))
        ok &= gamma_u8(c) == b
    # ground pixel (960,1079), zero jitter: dir = 0*u + 539*v - 288*w ; hit y=0 plane from eye (0,100,200)
    A, B = f32(960 - 1920 / 2), f32(1079 - 1080 / 2)
    d = [f32(f32(f32(A * u[k]) + f32(B * v[k])) - f32(f32(288.0) * w[k])) for k in range(3)]
    d = normalize(d)
    t = -100.0 / float(d[1])
    print("ground pixel t", t, "hit z", 200 + t * float(d[2]))
    ok &= abs(t - ka["ground_pixel"]["t"]) < ka["ground_pixel"]["t_atol"]
    print("OK" if ok else "MISMATCH")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
