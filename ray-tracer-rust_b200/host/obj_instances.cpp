// obj_instances.cpp -- K x K flattened copies of an OBJ written as ONE indexed OBJ: copy (i, j) translated by
// ((i - (K-1)/2) * 200, 0, -j * 200) in f32, as SURVEY 8d config 4 defines the instanced scene (K = 8: 317 952 triangles;
// K = 64: 20.3 M triangles, the loader-throughput case of SURVEY 8f-2).  usage: obj_instances in.obj out.obj K
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: %s in.obj out.obj K\n", argv[0]); return 2; }
    const int K = atoi(argv[3]);
    FILE* f = fopen(argv[1], "r");
    if (!f) { fprintf(stderr, "Not a valid path: %s\n", argv[1]); return 1; }
    std::vector<float> v; std::vector<long> idx;
    char line[512];
    while (fgets(line, sizeof(line), f)) {
        float x, y, z; long a, b, c;
        if (sscanf(line, "v %f %f %f", &x, &y, &z) == 3) { v.push_back(x); v.push_back(y); v.push_back(z); }
        else if (sscanf(line, "f %ld %ld %ld", &a, &b, &c) == 3) { idx.push_back(a); idx.push_back(b); idx.push_back(c); }
    }
    fclose(f);
    FILE* g = fopen(argv[2], "w");
    if (!g) { fprintf(stderr, "cannot create %s\n", argv[2]); return 1; }
    std::vector<char> buf(1 << 22);
    setvbuf(g, buf.data(), _IOFBF, buf.size());
    const long nv = (long)(v.size() / 3);
    long base = 0;
    for (int i = 0; i < K; ++i)
        for (int j = 0; j < K; ++j) {
            const float ox = (float)((i - (K - 1) / 2.0) * 200.0), oz = (float)(-j * 200.0);
            for (long k = 0; k < nv; ++k) fprintf(g, "v %.9g %.9g %.9g\n", (double)(v[3 * k] + ox), (double)(v[3 * k + 1] + 0.0f), (double)(v[3 * k + 2] + oz));
            for (size_t t = 0; t < idx.size(); t += 3) fprintf(g, "f %ld %ld %ld\n", idx[t] + base, idx[t + 1] + base, idx[t + 2] + base);
            base += nv;
        }
    fclose(g);
    printf("%d x %d copies: %ld vertices, %zu triangles\n", K, K, base, idx.size() / 3 * (size_t)K * K);
    return 0;
}
