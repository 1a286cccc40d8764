// rtb_obj_parallel.cpp -- import_obj (src/main.rs:114-149) for large files: the same grammar as rtb_import_obj_ex without
// materials (any blank runs, CR LF, comments, `v x y z [w]`, corners `a`, `a/b`, `a//c`, `a/b/c`, negative indices, polygons as
// fans), parsed by a pool of host threads straight into ONE output array that the caller may ask to be pinned, so that
// rtb_scene_create's upload is a single DMA.  Plain C++, no CUDA: the output allocator is passed in.  With materials
// (RTB_OBJ_MTL, rtb_import_obj_parallel_ex) `mtllib` / `usemtl` are honoured as rtb_import_obj_ex does: the statements are collected per
// chunk in pass 1, resolved serially in file order into (first triangle, material) switch points, and pass 3 writes every triangle's
// Kd and {Ks, Ns} next to its vertices.
//
// The file is mapped and parsed in place (no copy of the text).
// Three passes over chunks of the file cut at line ends (one chunk per thread):
//   1. count: `v` lines, triangles (corners - 2 per `f` line), lines            -> prefix sums give every chunk its offsets
//   2. vertices: every chunk parses its `v` lines into the shared vertex array  (std::from_chars: correctly rounded, like
//      Rust's str::parse::<f32> and strtof)
//   3. faces: every chunk parses its `f` lines and writes v0 v1 v2 of each triangle at its own offset of the output; a negative
//      index counts back from the vertices read before that line (the chunk's prefix + the `v` lines met so far in the chunk).
#include "../../include/rtb200.h"

#include <charconv>
#include <memory>
#include <new>
#include <sys/mman.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

extern "C" void rtb_set_error_(const char* msg);

namespace {

struct MtlEvent { bool lib; std::string name; size_t tri, line; };    // a `mtllib` / `usemtl` line: triangles of the chunk before it, line in the chunk
struct Chunk {
    size_t b = 0, e = 0;                  // byte range [b, e): whole lines
    size_t n_v = 0, n_tri = 0, n_lines = 0;
    size_t v0 = 0, tri0 = 0, line0 = 0;   // exclusive prefix sums
    std::vector<MtlEvent> events;         // (materials wanted)
    std::string err;                      // first error of this chunk
};
struct Material { std::string name; float kd[3] = {1.0f, 1.0f, 1.0f}, ks[3] = {0.0f, 0.0f, 0.0f}, ns = 0.0f; };
struct Switch { size_t tri; Material m; };                             // from global triangle `tri` on: material m

inline bool is_blank(char c) { return c == ' ' || c == '\t' || c == '\r'; }

// next token of the line [p, e): returns false at the end of the line or at a comment
inline bool next_token(const char*& p, const char* e, const char*& tb, const char*& te) {
    while (p < e && is_blank(*p)) ++p;
    if (p >= e || *p == '#') return false;
    tb = p;
    while (p < e && !is_blank(*p)) ++p;
    te = p;
    return true;
}

inline bool parse_float(const char* b, const char* e, float* out) {
    if (b < e && *b == '+') ++b;
    if (b >= e) return false;
    const auto r = std::from_chars(b, e, *out);
    if (r.ec == std::errc::result_out_of_range) {         // strtof semantics: +-inf / +-0 (denormals are returned by from_chars)
        char tmp[64]; const size_t n = (size_t)(e - b) < sizeof(tmp) - 1 ? (size_t)(e - b) : sizeof(tmp) - 1;
        memcpy(tmp, b, n); tmp[n] = 0; char* end = nullptr; *out = strtof(tmp, &end); return end == tmp + n;
    }
    return r.ec == std::errc() && r.ptr == e;
}

// one face corner: the vertex index before the first '/', 1-based or negative (relative to the nv vertices read so far)
inline bool parse_corner(const char* b, const char* e, size_t nv, size_t* out) {
    const char* s = b;
    while (s < e && *s != '/') ++s;
    if (s == b) return false;
    bool neg = false;
    const char* q = b;
    if (*q == '-') { neg = true; ++q; } else if (*q == '+') ++q;
    if (q >= s) return false;
    unsigned long long v = 0;
    for (; q < s; ++q) {
        if (*q < '0' || *q > '9') return false;
        v = v * 10 + (unsigned)(*q - '0');
        if (v > 0xffffffffffull) return false;
    }
    if (v == 0) return false;
    if (neg) { if (v > nv) return false; *out = nv - (size_t)v; }
    else { if (v > nv) return false; *out = (size_t)v - 1; }
    return true;
}

inline bool tok_is(const char* tb, const char* te, const char* w) { const size_t n = strlen(w); return (size_t)(te - tb) == n && memcmp(tb, w, n) == 0; }

// newmtl / Kd / Ks / Ns of one material library, appended to mats (the grammar of rtb_import_obj_ex's parse_mtl)
bool parse_mtl_file(const std::string& path, std::vector<Material>* mats) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    std::string data; char buf[1 << 16]; size_t got;
    while ((got = fread(buf, 1, sizeof(buf), f)) > 0) data.append(buf, got);
    fclose(f);
    const char* p = data.data(); const char* end = p + data.size();
    while (p < end) {
        const char* eol = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!eol) eol = end;
        const char* q = p; const char *tb, *te;
        p = eol + 1;
        if (!next_token(q, eol, tb, te)) continue;
        if (tok_is(tb, te, "newmtl")) { Material m; if (next_token(q, eol, tb, te)) m.name.assign(tb, te); mats->push_back(m); continue; }
        if (mats->empty()) continue;
        Material& m = mats->back();
        const bool is_kd = tok_is(tb, te, "Kd"), is_ks = tok_is(tb, te, "Ks");
        if (is_kd || is_ks) {
            float v[3]; bool ok = true;
            for (int a = 0; a < 3 && ok; ++a) ok = next_token(q, eol, tb, te) && parse_float(tb, te, &v[a]);
            if (ok) memcpy(is_kd ? m.kd : m.ks, v, sizeof(v));
        } else if (tok_is(tb, te, "Ns")) {
            float v;
            if (next_token(q, eol, tb, te) && parse_float(tb, te, &v)) m.ns = v;
        }
    }
    return true;
}

template <typename F>
void for_each_line(const char* data, const Chunk& c, F&& fn) {
    const char* p = data + c.b; const char* end = data + c.e;
    size_t line = 0;
    while (p < end) {
        const char* eol = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!eol) eol = end;
        if (!fn(p, eol, line)) return;
        ++line;
        p = eol + 1;
    }
}

template <typename F>
void run_threads(size_t n, F&& fn) {
    std::vector<std::thread> th;
    for (size_t i = 1; i < n; ++i) th.emplace_back(fn, i);
    fn(0);
    for (auto& t : th) t.join();
}

}  // namespace

// alloc(bytes) returns the output array (pinned or pageable); on failure of the parse the caller frees it.
// kd_out / ksns_out (either may be null): per-triangle Kd (3 floats) and {Ks, Ns} (4 floats) from `mtllib` / `usemtl`, allocated with `alloc` too.
int rtb_obj_parallel_core(const char* path, uint32_t n_threads, void* (*alloc)(size_t), float** verts_out, uint32_t* n_out, rtb_import_stats* st,
                          std::string* err, float** kd_out, float** ksns_out) {
    const bool want_mtl = kd_out != nullptr || ksns_out != nullptr;
    if (kd_out) *kd_out = nullptr;
    if (ksns_out) *ksns_out = nullptr;
    *verts_out = nullptr; *n_out = 0;
    const auto t0 = std::chrono::steady_clock::now();
    FILE* f = fopen(path, "rb");
    if (!f) { *err = std::string("Not a valid path: ") + path; return RTB_ERR_IO; }          // message of src/main.rs:118
    fseek(f, 0, SEEK_END);
    const long long fsz = ftell(f);
    if (fsz < 0) { fclose(f); *err = std::string("cannot size ") + path; return RTB_ERR_IO; }
    const size_t size = (size_t)fsz;
    unsigned T = n_threads ? n_threads : std::thread::hardware_concurrency();
    if (T == 0) T = 1;
    if (T > 256) T = 256;
    if (size < (size_t)T * 4096) T = (unsigned)(size / 4096 + 1);
    // The text is parsed where the page cache holds it: the file is mapped, not copied (a private read-only mapping; the parsing
    // threads take the page faults of their own chunks).  Copying it into a buffer first, even with one reader per chunk, cost as
    // much as parsing it.  A file that cannot be mapped (a pipe, an odd file system) is read into memory instead.
    struct Text {
        const char* p = nullptr; size_t n = 0; bool mapped = false; std::unique_ptr<char[]> owned;
        ~Text() { if (mapped && p) munmap(const_cast<char*>(p), n); }
    } text;
    text.n = size;
    if (size) {
        void* m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fileno(f), 0);
        if (m != MAP_FAILED) {
            text.p = (const char*)m; text.mapped = true;
            madvise(m, size, MADV_WILLNEED);
        } else {
            text.owned.reset(new (std::nothrow) char[size]);
            if (!text.owned) { fclose(f); *err = "rtb_import_obj_parallel: out of memory"; return RTB_ERR_IO; }
            bool ok = true;
            run_threads(T, [&](size_t i) {      // every thread reads its share of the file (page cache -> buffer copies in parallel)
                size_t b = size * i / T, e = size * (i + 1) / T;
                FILE* g = i == 0 ? f : fopen(path, "rb");
                if (!g) { ok = false; return; }
                if (fseek(g, (long)b, SEEK_SET) != 0 || fread(text.owned.get() + b, 1, e - b, g) != e - b) ok = false;
                if (i != 0) fclose(g);
            });
            if (!ok) { fclose(f); *err = std::string("short read of ") + path; return RTB_ERR_IO; }
            text.p = text.owned.get();
        }
    }
    fclose(f);
    const char* const data = text.p;
    const auto t1 = std::chrono::steady_clock::now();

    std::vector<Chunk> ch(T);
    for (unsigned i = 0; i < T; ++i) {
        size_t b = size * i / T;
        if (i > 0) { while (b < size && data[b - 1] != '\n') ++b; }                          // start after a line end
        ch[i].b = b;
    }
    for (unsigned i = 0; i < T; ++i) ch[i].e = i + 1 < T ? ch[i + 1].b : size;

    // pass 1: counts
    run_threads(T, [&](size_t i) {
        Chunk& c = ch[i];
        for_each_line(data, c, [&](const char* p, const char* e, size_t) {
            ++c.n_lines;
            const char *tb, *te;
            if (!next_token(p, e, tb, te)) return true;
            if (te - tb == 1 && *tb == 'v') ++c.n_v;
            else if (te - tb == 1 && *tb == 'f') {
                size_t corners = 0;
                while (next_token(p, e, tb, te)) ++corners;
                if (corners >= 3) c.n_tri += corners - 2;                                   // (fewer: an error, reported by pass 3)
            } else if (want_mtl && (tok_is(tb, te, "usemtl") || tok_is(tb, te, "mtllib"))) {
                const bool lib = *tb == 'm';
                bool any = false;
                while (next_token(p, e, tb, te)) { c.events.push_back({lib, std::string(tb, te), c.n_tri, c.n_lines}); any = true; if (!lib) break; }
                if (!any && !lib) c.events.push_back({false, std::string(), c.n_tri, c.n_lines});      // `usemtl` without a name
            }
            return true;
        });
    });
    size_t nv = 0, ntri = 0, nl = 0;
    for (Chunk& c : ch) { c.v0 = nv; c.tri0 = ntri; c.line0 = nl; nv += c.n_v; ntri += c.n_tri; nl += c.n_lines; }
    if (ntri > 0x3fffffffull) { *err = "rtb_import_obj_parallel: too many triangles"; return RTB_ERR_LIMIT; }
    // materials: the statements in file order -> switch points (serial: a handful of lines)
    std::vector<Switch> switches;
    if (want_mtl) {
        std::string dir = path;
        const size_t slash = dir.find_last_of("/\\");
        dir = slash == std::string::npos ? std::string() : dir.substr(0, slash + 1);
        std::vector<Material> mats;
        for (const Chunk& c : ch)
            for (const MtlEvent& ev : c.events) {
                if (ev.lib) {
                    if (!parse_mtl_file(dir + ev.name, &mats)) { *err = std::string(path) + ": cannot read material library " + dir + ev.name; return RTB_ERR_IO; }
                } else {
                    bool found = false;
                    Material cur;
                    for (const Material& m : mats) if (m.name == ev.name) { cur = m; found = true; }
                    if (!found) { *err = std::string(path) + ": unknown material '" + ev.name + "' at line " + std::to_string(c.line0 + ev.line); return RTB_ERR_IO; }
                    switches.push_back({c.tri0 + ev.tri, cur});
                }
            }
    }

    // pass 2: vertices
    std::vector<float> verts(nv * 3);
    run_threads(T, [&](size_t i) {
        Chunk& c = ch[i];
        size_t k = c.v0;
        for_each_line(data, c, [&](const char* p, const char* e, size_t line) {
            const char *tb, *te;
            if (!next_token(p, e, tb, te) || te - tb != 1 || *tb != 'v') return true;
            float xyz[3];
            for (int a = 0; a < 3; ++a) {
                if (!next_token(p, e, tb, te) || !parse_float(tb, te, &xyz[a])) {
                    c.err = std::string(path) + ": malformed vertex at line " + std::to_string(c.line0 + line + 1);
                    return false;
                }
            }
            verts[3 * k] = xyz[0]; verts[3 * k + 1] = xyz[1]; verts[3 * k + 2] = xyz[2];
            ++k;
            return true;
        });
    });
    for (const Chunk& c : ch) if (!c.err.empty()) { *err = c.err; return RTB_ERR_IO; }

    // pass 3: faces -> triangles (fan around the first corner), straight into the output array
    float* out = nullptr;
    const auto ta0 = std::chrono::steady_clock::now();
    float *kd = nullptr, *ksns = nullptr;
    if (ntri) {
        out = (float*)alloc(ntri * 9 * sizeof(float));
        if (kd_out) kd = (float*)alloc(ntri * 3 * sizeof(float));
        if (ksns_out) ksns = (float*)alloc(ntri * 4 * sizeof(float));
        *verts_out = out;                                                                    // (the caller frees what was allocated, also on failure)
        if (kd_out) *kd_out = kd;
        if (ksns_out) *ksns_out = ksns;
        if (!out || (kd_out && !kd) || (ksns_out && !ksns)) { *err = "rtb_import_obj_parallel: out of memory"; return RTB_ERR_IO; }
    }
    const double alloc_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - ta0).count();
    run_threads(T, [&](size_t i) {
        Chunk& c = ch[i];
        size_t seen_v = c.v0, k = c.tri0;
        std::vector<size_t> corner;
        // the material of triangle k: the last switch point at or before it (default: white, no specular -- src/main.rs:143)
        size_t sw = 0;
        while (sw < switches.size() && switches[sw].tri <= k) ++sw;
        const Material white;
        for_each_line(data, c, [&](const char* p, const char* e, size_t line) {
            const char *tb, *te;
            if (!next_token(p, e, tb, te) || te - tb != 1) return true;
            if (*tb == 'v') { ++seen_v; return true; }
            if (*tb != 'f') return true;
            corner.clear();
            while (next_token(p, e, tb, te)) {
                size_t idx;
                if (!parse_corner(tb, te, seen_v, &idx)) {
                    c.err = std::string(path) + ": malformed face at line " + std::to_string(c.line0 + line + 1);
                    return false;
                }
                corner.push_back(idx);
            }
            if (corner.size() < 3) {
                c.err = std::string(path) + ": face with fewer than 3 corners at line " + std::to_string(c.line0 + line + 1);
                return false;
            }
            for (size_t j = 1; j + 1 < corner.size(); ++j) {
                const size_t tri[3] = {corner[0], corner[j], corner[j + 1]};
                for (int c3 = 0; c3 < 3; ++c3) memcpy(out + 9 * k + 3 * c3, verts.data() + 3 * tri[c3], 3 * sizeof(float));
                if (want_mtl) {
                    while (sw < switches.size() && switches[sw].tri <= k) ++sw;
                    const Material& m = sw ? switches[sw - 1].m : white;
                    if (kd) memcpy(kd + 3 * k, m.kd, sizeof(m.kd));
                    if (ksns) { memcpy(ksns + 4 * k, m.ks, sizeof(m.ks)); ksns[4 * k + 3] = m.ns; }
                }
                ++k;
            }
            return true;
        });
    });
    for (const Chunk& c : ch) if (!c.err.empty()) { *err = c.err; *verts_out = out; return RTB_ERR_IO; }
    const auto t2 = std::chrono::steady_clock::now();
    *verts_out = out; *n_out = (uint32_t)ntri;
    if (st) {
        st->file_bytes = size; st->threads = T; st->n_vertices = (uint32_t)(nv > 0xffffffffull ? 0xffffffffull : nv);
        st->read_s = std::chrono::duration<double>(t1 - t0).count();
        st->parse_s = std::chrono::duration<double>(t2 - t1).count() - alloc_s;
        st->alloc_s = alloc_s;
    }
    return RTB_OK;
}
