// sanitizer harness: feeds files given on argv to rt_gltf_load_memory and walks the result
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "rtb200_host.h"
int main(int argc, char** argv) {
    for (int i = 1; i < argc; i++) {
        FILE* f = fopen(argv[i], "rb");
        if (!f) continue;
        std::vector<uint8_t> buf;
        uint8_t tmp[65536]; size_t n;
        while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + n);
        fclose(f);
        RtGltfScene* sc = nullptr; char err[256] = {0};
        int rc = rt_gltf_load_memory(buf.data(), buf.size(), "/tmp/gltf_fuzz/nonexistent", 0, &sc, err, sizeof err);
        if (rc == 0 && sc) {
            RtGltfCounts c; rt_gltf_counts(sc, &c);
            volatile uint64_t sink = 0;
            const uint8_t* p = (const uint8_t*)rt_gltf_triangles(sc);
            (void)p; sink += (uint64_t)(uintptr_t)rt_gltf_vertices(sc);
            for (uint64_t k = 0; rt_gltf_warning(sc, k); k++) sink += strlen(rt_gltf_warning(sc, k));
            rt_gltf_free(sc);
        }
    }
    return 0;
}
