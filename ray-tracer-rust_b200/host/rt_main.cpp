// rt_main.cpp -- host driver with the reference's command line: every plain argument is an OBJ path
// (src/main.rs:321,328-331); the scene is OBJ triangles + ground + light + camera exactly as main()
// builds it (src/main.rs:319-358); output.png is written in the current directory
// (src/main.rs:312-315).  The render itself is the C ABI call into the CUDA library.  Optional
// --flags expose the reference's compile-time constants; their defaults are the reference's values.
#include "../../include/rtb200.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

int main(int argc, char** argv) {
    uint32_t width = 1920, height = 1080;        // src/main.rs:350-351
    uint32_t nb_ray = 1, nb_light = 100;         // src/main.rs:38-39
    uint32_t n_samples = 2000000;                // src/main.rs:40
    float distance = 288.0f;                     // src/main.rs:356
    uint64_t seed = 0x5EED0001ull;
    int device = 0;
    std::string out = "output.png";
    int n_gpus = 1;                              // --gpus N: row bands b % N over N devices of this box (rtb_render_multi)
    bool gather_peer = false;                    // --gather-peer: rows over NVLink into device 0's frame, one device-to-host copy
    int orbit = 0;                               // --orbit N: N camera poses around the scene (BASELINE config 5), frames <out stem>_%04u.png
    int instances = 1;                           // --instances K: K x K flattened copies of the OBJ triangles (BASELINE config 4)
    bool use_phong = false;                      // --phong: --mtl + the specular term Ks max(0, r.v)^Ns (rtb_scene_set_materials; extension)
    bool use_mtl = false;                        // --mtl: robust OBJ reader + Kd colours from the .mtl (extension; the reference skips materials)
    std::vector<std::string> objs;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto need = [&](const char* name) -> const char* {
            if (i + 1 >= argc) { fprintf(stderr, "%s needs a value\n", name); exit(2); }
            return argv[++i];
        };
        if (a == "--width") width = (uint32_t)atoi(need("--width"));
        else if (a == "--height") height = (uint32_t)atoi(need("--height"));
        else if (a == "--spp") nb_ray = (uint32_t)atoi(need("--spp"));
        else if (a == "--light-samples") nb_light = (uint32_t)atoi(need("--light-samples"));
        else if (a == "--distance") distance = (float)atof(need("--distance"));
        else if (a == "--seed") seed = strtoull(need("--seed"), nullptr, 0);
        else if (a == "--device") device = atoi(need("--device"));
        else if (a == "--out") out = need("--out");
        else if (a == "--gpus") n_gpus = atoi(need("--gpus"));
        else if (a == "--gather-peer") gather_peer = true;
        else if (a == "--orbit") orbit = atoi(need("--orbit"));
        else if (a == "--instances") instances = atoi(need("--instances"));
        else if (a == "--mtl") use_mtl = true;
        else if (a == "--phong") { use_mtl = true; use_phong = true; }
        else objs.push_back(a);
    }

    printf("Building scene\n");                  // src/main.rs:322
    std::vector<float> verts, colors, mats;
    for (const std::string& path : objs) {
        float* v = nullptr; float* kd = nullptr; float* ksns = nullptr; uint32_t n = 0;
        int rc = use_mtl ? rtb_import_obj_ex(path.c_str(), RTB_OBJ_MTL, &v, &kd, use_phong ? &ksns : nullptr, &n) : rtb_import_obj(path.c_str(), &v, &n);
        if (rc != RTB_OK) {
            if (n == 0 && strncmp(rtb_last_error(), "Not a valid path", 16) == 0) { printf("%s\n", rtb_last_error()); continue; }
            fprintf(stderr, "%s\n", rtb_last_error());
            return 1;
        }
        verts.insert(verts.end(), v, v + (size_t)n * 9);
        if (kd) colors.insert(colors.end(), kd, kd + (size_t)n * 3);
        else for (uint32_t k = 0; k < n; ++k) { colors.push_back(1.0f); colors.push_back(1.0f); colors.push_back(1.0f); }   // :143
        if (ksns) mats.insert(mats.end(), ksns, ksns + (size_t)n * 4);
        rtb_free(v); rtb_free(kd); rtb_free(ksns);
    }
    if (instances > 1) {                         // copy (i, j) translated by ((i - (K-1)/2) * 200, 0, -j * 200) in f32
        const size_t n0 = verts.size() / 9;
        std::vector<float> v2, c2, m2;
        for (int i = 0; i < instances; ++i)
            for (int j = 0; j < instances; ++j) {
                const float ox = (float)((i - (instances - 1) / 2.0) * 200.0), oz = (float)(-j * 200.0);
                for (size_t t = 0; t < n0; ++t)
                    for (int k = 0; k < 3; ++k) {
                        v2.push_back(verts[t * 9 + 3 * k] + ox); v2.push_back(verts[t * 9 + 3 * k + 1] + 0.0f); v2.push_back(verts[t * 9 + 3 * k + 2] + oz);
                    }
                c2.insert(c2.end(), colors.begin(), colors.end());
                m2.insert(m2.end(), mats.begin(), mats.end());
            }
        verts.swap(v2); colors.swap(c2); mats.swap(m2);
    }
    const float ground[9] = {-10000.0f, 0.0f, -10000.0f, 10000.0f, 0.0f, -10000.0f, 0.0f, 0.0f, 10000.0f};   // :102-111
    verts.insert(verts.end(), ground, ground + 9);
    colors.push_back(0.5f); colors.push_back(0.5f); colors.push_back(0.5f);

    if (n_gpus < 1) n_gpus = 1;
    rtb_scene_desc desc;
    memset(&desc, 0, sizeof(desc));
    desc.n_tris = (uint32_t)(verts.size() / 9);
    desc.verts = verts.data(); desc.colors = colors.data(); desc.tie_rank = nullptr;
    const float light[9] = {-10.0f, 300.0f, -10.0f, 10.0f, 300.0f, -10.0f, 0.0f, 300.0f, 0.0f};             // :337-343
    memcpy(desc.light, light, sizeof(light));
    if (use_phong) mats.resize((size_t)desc.n_tris * 4, 0.0f);                                              // the ground: no highlight
    std::vector<rtb_scene*> scenes;                                                                         // one replica per GPU
    for (int g = 0; g < n_gpus; ++g) {
        desc.device = device + g;
        rtb_scene* sc = nullptr;
        if (rtb_scene_create(&desc, &sc) != RTB_OK) { fprintf(stderr, "%s\n", rtb_last_error()); return 1; }
        if (use_phong && rtb_scene_set_materials(sc, mats.data()) != RTB_OK) { fprintf(stderr, "%s\n", rtb_last_error()); return 1; }
        scenes.push_back(sc);
    }
    rtb_scene* scene = scenes[0];
    std::vector<float> samples((size_t)n_samples * 2);
    rtb_fill_samples(seed, n_samples, samples.data());
    if (rtb_scenes_set_samples(scenes.data(), (uint32_t)n_gpus, samples.data(), n_samples) != RTB_OK) { fprintf(stderr, "%s\n", rtb_last_error()); return 1; }

    rtb_render_params prm;
    memset(&prm, 0, sizeof(prm));
    prm.width = width; prm.height = height; prm.nb_ray = nb_ray; prm.nb_light_sample = nb_light;
    prm.eye[0] = 0.0f; prm.eye[1] = 100.0f; prm.eye[2] = 200.0f;                                             // :353
    prm.look_at[0] = 0.0f; prm.look_at[1] = 0.0f; prm.look_at[2] = -100000.0f;                               // :354
    prm.up[0] = 0.0f; prm.up[1] = 1.0f; prm.up[2] = 0.0f;                                                    // :355
    prm.distance = distance;
    prm.world = 1;
    prm.band_rows = 4;
    if (gather_peer) prm.flags |= RTB_FLAG_GATHER_PEER;

    if (orbit > 0) {                             // BASELINE config 5: pose k on the circle of radius 200 around the y axis (pose 0 = the reference camera)
        std::vector<rtb_camera_pose> poses((size_t)orbit);
        for (int k = 0; k < orbit; ++k) {
            const double th = 2.0 * 3.14159265358979323846 * k / orbit, sn = sin(th), cs = cos(th);
            rtb_camera_pose& c = poses[(size_t)k];
            c.eye[0] = (float)(200.0 * sn); c.eye[1] = 100.0f; c.eye[2] = (float)(200.0 * cs);
            c.look_at[0] = (float)(-100000.0 * sn); c.look_at[1] = 0.0f; c.look_at[2] = (float)(-100000.0 * cs);
            c.up[0] = 0.0f; c.up[1] = 1.0f; c.up[2] = 0.0f; c.distance = distance;
        }
        std::string stem = out.size() > 4 && out.substr(out.size() - 4) == ".png" ? out.substr(0, out.size() - 4) : out;
        const std::string pattern = stem + "_%04u.png";
        printf("Rendering...\n");
        rtb_sequence_stats sq;
        const int rc = n_gpus > 1 ? rtb_render_sequence_multi(scenes.data(), (uint32_t)n_gpus, &prm, poses.data(), (uint32_t)orbit, nullptr, pattern.c_str(), &sq)
                                  : rtb_render_sequence(scene, &prm, poses.data(), (uint32_t)orbit, nullptr, pattern.c_str(), nullptr, nullptr, &sq);
        if (rc != RTB_OK) { fprintf(stderr, "%s\n", rtb_last_error()); return 1; }
        printf("Rendered %u frames in %.6f seconds (%.1f frames/s, PNG files written alongside), %u camera-grid rebuilds redone\n", sq.frames, sq.seconds,
               sq.frames / sq.seconds, sq.cam_grid_redo);
        for (rtb_scene* sc : scenes) rtb_scene_destroy(sc);
        return 0;
    }

    printf("Rendering...\n");                    // src/main.rs:360
    std::vector<uint8_t> img((size_t)width * height * 3);
    rtb_stats st;
    auto t0 = std::chrono::steady_clock::now();
    // render(scene), src/main.rs:361 -- on N GPUs the call fans out to one host thread per device and joins them itself (:269-303)
    const int rrc = n_gpus > 1 ? rtb_render_multi(scenes.data(), (uint32_t)n_gpus, &prm, img.data(), &st) : rtb_render(scene, &prm, img.data(), &st);
    if (rrc != RTB_OK) { fprintf(stderr, "%s\n", rtb_last_error()); return 1; }
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    // src/main.rs:305-310 prints whole seconds and primary rays only; sub-second frames need decimals
    printf("Rendered in %.6f seconds (kernel %.3f ms)\n", secs, st.kernel_ms);
    printf("Throughput %.1fM ray/s primary, %.1fM ray/s primary+shadow\n", (double)st.primary_rays / 1e6 / secs,
           (double)(st.primary_rays + st.shadow_rays) / 1e6 / secs);

    printf("Writting image to disk\n");          // src/main.rs:312 [sic]
    if (rtb_write_png(out.c_str(), img.data(), width, height) != RTB_OK) { fprintf(stderr, "%s\n", rtb_last_error()); return 1; }
    for (rtb_scene* sc : scenes) rtb_scene_destroy(sc);
    return 0;
}
