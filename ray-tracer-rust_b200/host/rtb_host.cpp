// rtb_host.cpp -- host-side entry points kept from the reference's main.rs: OBJ import
// (src/main.rs:114-149), the sample table (src/main.rs:253,260-265, made deterministic) and the
// PNG save (src/main.rs:312-315).  Plain C++, no CUDA.
#include "../../include/rtb200.h"

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

extern "C" const char* rtb_last_error(void);
// defined in rtb_api.cu (thread-local message buffer); re-declared here through a setter
extern "C" void rtb_set_error_(const char* msg);

static int hfail(int code, const std::string& msg) { rtb_set_error_(msg.c_str()); return code; }

static bool parse_f32(const std::string& tok, float* out) {
    if (tok.empty() || tok[0] == ' ' || tok[0] == '\t') return false;
    errno = 0;
    char* end = nullptr;
    float v = strtof(tok.c_str(), &end);       // correctly rounded, like str::parse::<f32>
    if (end == tok.c_str() || *end != '\0') return false;
    *out = v;
    return true;
}

static bool parse_usize(const std::string& tok, unsigned long long* out) {
    if (tok.empty()) return false;
    size_t i = 0;
    if (tok[0] == '+') i = 1;
    if (i >= tok.size()) return false;
    unsigned long long v = 0;
    for (; i < tok.size(); ++i) {
        if (tok[i] < '0' || tok[i] > '9') return false;
        v = v * 10 + (unsigned)(tok[i] - '0');
        if (v > 0xffffffffull) return false;
    }
    *out = v;
    return true;
}

extern "C" int rtb_import_obj(const char* path, float** verts_out, uint32_t* n_out) {
    if (!path || !verts_out || !n_out) return hfail(RTB_ERR_ARG, "rtb_import_obj: null argument");
    *verts_out = nullptr; *n_out = 0;
    FILE* f = fopen(path, "rb");
    if (!f) return hfail(RTB_ERR_IO, std::string("Not a valid path: ") + path);   // message of src/main.rs:118
    std::string data;
    char buf[1 << 16];
    size_t got;
    while ((got = fread(buf, 1, sizeof(buf), f)) > 0) data.append(buf, got);
    fclose(f);

    std::vector<float> vertices;   // Vec<Point3<f32>>
    std::vector<float> tris;
    size_t pos = 0, line_no = 0;
    while (pos < data.size()) {
        size_t eol = data.find('\n', pos);
        if (eol == std::string::npos) eol = data.size();
        size_t end = eol;
        if (end > pos && data[end - 1] == '\r') --end;              // BufRead::lines strips "\r\n" too
        const std::string line = data.substr(pos, end - pos);
        pos = eol + 1;
        ++line_no;
        // str::split(" "): single-space separator, empty tokens are kept
        std::vector<std::string> tok;
        size_t s = 0;
        for (;;) {
            size_t sp = line.find(' ', s);
            if (sp == std::string::npos) { tok.push_back(line.substr(s)); break; }
            tok.push_back(line.substr(s, sp - s));
            s = sp + 1;
        }
        if (tok[0] == "v") {
            float x, y, z;
            if (tok.size() < 4 || !parse_f32(tok[1], &x) || !parse_f32(tok[2], &y) || !parse_f32(tok[3], &z))
                return hfail(RTB_ERR_IO, std::string(path) + ": malformed vertex at line " + std::to_string(line_no) +
                                             " (the reference panics here)");
            vertices.push_back(x); vertices.push_back(y); vertices.push_back(z);
        } else if (tok[0] == "f") {
            unsigned long long idx[3];
            const size_t nv = vertices.size() / 3;
            for (int k = 0; k < 3; ++k) {
                if (tok.size() < 4 || !parse_usize(tok[1 + k], &idx[k]) || idx[k] < 1 || idx[k] > nv)
                    return hfail(RTB_ERR_IO, std::string(path) + ": malformed face at line " + std::to_string(line_no) +
                                                 " (the reference panics here)");
            }
            for (int k = 0; k < 3; ++k)
                for (int c = 0; c < 3; ++c) tris.push_back(vertices[3 * (idx[k] - 1) + c]);
        }
    }
    const size_t n = tris.size() / 9;
    if (n) {
        float* out = (float*)malloc(tris.size() * sizeof(float));
        if (!out) return hfail(RTB_ERR_IO, "rtb_import_obj: out of memory");
        memcpy(out, tris.data(), tris.size() * sizeof(float));
        *verts_out = out;
    }
    *n_out = (uint32_t)n;
    return RTB_OK;
}

// ---- robust OBJ + MTL import ---------------------------------------------------------------------
// The reference's import_obj (src/main.rs:114-149) panics or mis-parses on anything but `v x y z` / `f a b c` separated by
// single spaces, and skips `mtllib` / `usemtl` (src/main.rs:130,137: every OBJ triangle is white).  This entry point reads
// what exporters actually write: any run of spaces / tabs, CR LF, comments, `v x y z [w]`, face corners `a`, `a/b`, `a//c`,
// `a/b/c`, negative (relative) indices, polygons (fan-triangulated around the first corner, in corner order) and -- with
// RTB_OBJ_MTL -- `mtllib` / `usemtl`, giving each triangle the material's Kd / Ks / Ns (models/big_bunny.mtl:4-10).  On the
// reference's own two models it returns exactly the triangles of rtb_import_obj (tests/test_host.py).
namespace {
struct Material { std::string name; float kd[3] = {1.0f, 1.0f, 1.0f}, ks[3] = {0.0f, 0.0f, 0.0f}, ns = 0.0f; };

bool read_file(const std::string& path, std::string* out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    char buf[1 << 16];
    size_t got;
    while ((got = fread(buf, 1, sizeof(buf), f)) > 0) out->append(buf, got);
    fclose(f);
    return true;
}

// tokens of one line: runs of non-blank characters; a '#' starts a comment
void split_ws(const std::string& data, size_t b, size_t e, std::vector<std::string>* tok) {
    tok->clear();
    size_t i = b;
    while (i < e) {
        while (i < e && (data[i] == ' ' || data[i] == '\t' || data[i] == '\r')) ++i;
        if (i >= e || data[i] == '#') break;
        size_t j = i;
        while (j < e && data[j] != ' ' && data[j] != '\t' && data[j] != '\r') ++j;
        tok->push_back(data.substr(i, j - i));
        i = j;
    }
}

bool parse_mtl(const std::string& path, std::vector<Material>* mats) {
    std::string data;
    if (!read_file(path, &data)) return false;
    std::vector<std::string> tok;
    size_t pos = 0;
    while (pos < data.size()) {
        size_t eol = data.find('\n', pos);
        if (eol == std::string::npos) eol = data.size();
        split_ws(data, pos, eol, &tok);
        pos = eol + 1;
        if (tok.empty()) continue;
        if (tok[0] == "newmtl") { Material m; m.name = tok.size() > 1 ? tok[1] : ""; mats->push_back(m); continue; }
        if (mats->empty()) continue;
        Material& m = mats->back();
        float v[3];
        if ((tok[0] == "Kd" || tok[0] == "Ks") && tok.size() >= 4 && parse_f32(tok[1], &v[0]) && parse_f32(tok[2], &v[1]) && parse_f32(tok[3], &v[2]))
            memcpy(tok[0] == "Kd" ? m.kd : m.ks, v, sizeof(v));
        else if (tok[0] == "Ns" && tok.size() >= 2 && parse_f32(tok[1], &v[0])) m.ns = v[0];
    }
    return true;
}

// one face corner `a`, `a/b`, `a//c`, `a/b/c`: the vertex index (1-based, or negative = relative to the vertices read so far)
bool parse_corner(const std::string& tok, size_t nv, size_t* out) {
    const std::string a = tok.substr(0, tok.find('/'));
    if (a.empty()) return false;
    char* end = nullptr;
    errno = 0;
    const long long v = strtoll(a.c_str(), &end, 10);
    if (end == a.c_str() || *end != '\0' || errno) return false;
    long long idx = v > 0 ? v - 1 : (long long)nv + v;
    if (v == 0 || idx < 0 || (unsigned long long)idx >= nv) return false;
    *out = (size_t)idx;
    return true;
}
}  // namespace

extern "C" int rtb_import_obj_ex(const char* path, uint32_t flags, float** verts_out, float** kd_out, float** ks_ns_out, uint32_t* n_out) {
    if (!path || !verts_out || !n_out) return hfail(RTB_ERR_ARG, "rtb_import_obj_ex: null argument");
    *verts_out = nullptr; *n_out = 0;
    if (kd_out) *kd_out = nullptr;
    if (ks_ns_out) *ks_ns_out = nullptr;
    std::string data;
    if (!read_file(path, &data)) return hfail(RTB_ERR_IO, std::string("Not a valid path: ") + path);   // message of src/main.rs:118
    std::string dir = path;
    const size_t slash = dir.find_last_of("/\\");
    dir = slash == std::string::npos ? std::string() : dir.substr(0, slash + 1);

    std::vector<float> vertices, tris, kd, ksns;
    std::vector<Material> mats;
    Material cur;                                                  // default: white, no specular (src/main.rs:143)
    std::vector<std::string> tok;
    std::vector<size_t> corner;
    size_t pos = 0, line_no = 0;
    while (pos < data.size()) {
        size_t eol = data.find('\n', pos);
        if (eol == std::string::npos) eol = data.size();
        split_ws(data, pos, eol, &tok);
        pos = eol + 1;
        ++line_no;
        if (tok.empty()) continue;
        if (tok[0] == "v") {
            float x, y, z;
            if (tok.size() < 4 || !parse_f32(tok[1], &x) || !parse_f32(tok[2], &y) || !parse_f32(tok[3], &z))
                return hfail(RTB_ERR_IO, std::string(path) + ": malformed vertex at line " + std::to_string(line_no));
            vertices.push_back(x); vertices.push_back(y); vertices.push_back(z);
        } else if (tok[0] == "f") {
            corner.clear();
            for (size_t k = 1; k < tok.size(); ++k) {
                size_t idx;
                if (!parse_corner(tok[k], vertices.size() / 3, &idx))
                    return hfail(RTB_ERR_IO, std::string(path) + ": malformed face at line " + std::to_string(line_no));
                corner.push_back(idx);
            }
            if (corner.size() < 3) return hfail(RTB_ERR_IO, std::string(path) + ": face with fewer than 3 corners at line " + std::to_string(line_no));
            for (size_t k = 1; k + 1 < corner.size(); ++k) {       // fan around the first corner
                const size_t tri[3] = {corner[0], corner[k], corner[k + 1]};
                for (int c3 = 0; c3 < 3; ++c3)
                    for (int c = 0; c < 3; ++c) tris.push_back(vertices[3 * tri[c3] + c]);
                for (int c = 0; c < 3; ++c) kd.push_back(cur.kd[c]);
                for (int c = 0; c < 3; ++c) ksns.push_back(cur.ks[c]);
                ksns.push_back(cur.ns);
            }
        } else if ((flags & RTB_OBJ_MTL) && tok[0] == "mtllib") {
            for (size_t k = 1; k < tok.size(); ++k)
                if (!parse_mtl(dir + tok[k], &mats)) return hfail(RTB_ERR_IO, std::string(path) + ": cannot read material library " + dir + tok[k]);
        } else if ((flags & RTB_OBJ_MTL) && tok[0] == "usemtl") {
            const std::string name = tok.size() > 1 ? tok[1] : "";
            bool found = false;
            for (const Material& m : mats) if (m.name == name) { cur = m; found = true; }
            if (!found) return hfail(RTB_ERR_IO, std::string(path) + ": unknown material '" + name + "' at line " + std::to_string(line_no));
        }
    }
    const size_t n = tris.size() / 9;
    if (n > 0xffffffffull) return hfail(RTB_ERR_LIMIT, "rtb_import_obj_ex: too many triangles");
    if (n) {
        float* v = (float*)malloc(tris.size() * sizeof(float));
        float* k1 = kd_out ? (float*)malloc(kd.size() * sizeof(float)) : nullptr;
        float* k2 = ks_ns_out ? (float*)malloc(ksns.size() * sizeof(float)) : nullptr;
        if (!v || (kd_out && !k1) || (ks_ns_out && !k2)) { free(v); free(k1); free(k2); return hfail(RTB_ERR_IO, "rtb_import_obj_ex: out of memory"); }
        memcpy(v, tris.data(), tris.size() * sizeof(float));
        if (k1) memcpy(k1, kd.data(), kd.size() * sizeof(float));
        if (k2) memcpy(k2, ksns.data(), ksns.size() * sizeof(float));
        *verts_out = v;
        if (kd_out) *kd_out = k1;
        if (ks_ns_out) *ks_ns_out = k2;
    }
    *n_out = (uint32_t)n;
    return RTB_OK;
}

extern "C" void rtb_free(void* p) { free(p); }

extern "C" void rtb_fill_samples(uint64_t seed, uint32_t n_pairs, float* out) {
    uint64_t state = seed;
    for (uint64_t j = 0; j < 2ull * n_pairs; ++j) {
        state += 0x9E3779B97F4A7C15ull;                       // SplitMix64
        uint64_t z = state;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z = z ^ (z >> 31);
        out[j] = (float)(z >> 40) * (1.0f / 16777216.0f);      // 24 bits -> [0,1), exact
    }
}

// ---- PNG (RGB8, zlib "stored" blocks: valid PNG, no compression library needed) -------------
static uint32_t crc_table[256];
static bool crc_ready = false;
static void crc_init() {
    for (uint32_t n = 0; n < 256; ++n) {
        uint32_t c = n;
        for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
        crc_table[n] = c;
    }
    crc_ready = true;
}
static uint32_t crc_update(uint32_t c, const uint8_t* p, size_t n) {
    for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ p[i]) & 0xff] ^ (c >> 8);
    return c;
}
static void put_be32(uint8_t* p, uint32_t v) { p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v; }

static bool write_chunk(FILE* f, const char* type, const uint8_t* data, uint32_t len) {
    uint8_t hdr[8];
    put_be32(hdr, len); memcpy(hdr + 4, type, 4);
    uint32_t c = crc_update(0xffffffffu, hdr + 4, 4);
    if (len) c = crc_update(c, data, len);
    uint8_t crc[4]; put_be32(crc, c ^ 0xffffffffu);
    return fwrite(hdr, 1, 8, f) == 8 && (len == 0 || fwrite(data, 1, len, f) == len) && fwrite(crc, 1, 4, f) == 4;
}

extern "C" int rtb_write_png(const char* path, const uint8_t* rgb8, uint32_t width, uint32_t height) {
    if (!path || !rgb8 || width == 0 || height == 0) return hfail(RTB_ERR_ARG, "rtb_write_png: bad argument");
    if (!crc_ready) crc_init();
    const size_t row = (size_t)width * 3 + 1;
    const size_t raw_len = row * height;
    std::vector<uint8_t> raw(raw_len);
    for (uint32_t y = 0; y < height; ++y) {
        raw[y * row] = 0;   // filter: none
        memcpy(&raw[y * row + 1], rgb8 + (size_t)y * width * 3, (size_t)width * 3);
    }
    // zlib stream: header, stored deflate blocks (<= 65535 bytes each), adler32
    std::vector<uint8_t> z;
    z.reserve(raw_len + raw_len / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    size_t off = 0;
    while (off < raw_len) {
        const size_t n = raw_len - off < 65535 ? raw_len - off : 65535;
        z.push_back(off + n == raw_len ? 1 : 0);
        z.push_back(n & 0xff); z.push_back(n >> 8); z.push_back(~n & 0xff); z.push_back((~n >> 8) & 0xff);
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        for (size_t i = 0; i < n; ++i) { a += raw[off + i]; if (a >= 65521) a -= 65521; b += a; if (b >= 65521) b -= 65521; }
        off += n;
    }
    uint8_t ad[4]; put_be32(ad, (b << 16) | a);
    z.insert(z.end(), ad, ad + 4);
    if (z.size() > 0x7fffffffu) return hfail(RTB_ERR_LIMIT, "rtb_write_png: image too large for one IDAT");

    FILE* f = fopen(path, "wb");
    if (!f) return hfail(RTB_ERR_IO, std::string("cannot create ") + path);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    uint8_t ihdr[13];
    put_be32(ihdr, width); put_be32(ihdr + 4, height);
    ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;   // 8-bit, colour type 2 (RGB)
    bool ok = fwrite(sig, 1, 8, f) == 8 && write_chunk(f, "IHDR", ihdr, 13) &&
              write_chunk(f, "IDAT", z.data(), (uint32_t)z.size()) && write_chunk(f, "IEND", nullptr, 0);
    ok = (fclose(f) == 0) && ok;
    return ok ? RTB_OK : hfail(RTB_ERR_IO, std::string("short write to ") + path);
}
