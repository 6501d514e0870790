#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <random>
#include "rtb200_host.h"
int main(int argc, char** argv) {
    int seed0 = argc > 1 ? atoi(argv[1]) : 0, iters = argc > 2 ? atoi(argv[2]) : 200;
    for (int it = 0; it < iters; it++) {
        std::mt19937 rng(seed0 * 100003 + it);
        auto U = [&](int n) { return (uint32_t)(rng() % (uint32_t)n); };
        uint64_t nt = (it % 7 == 0) ? U(3) : (it % 11 == 0 ? 100000 + U(3000) : U(5000));
        uint64_t nv = 1 + U(3000);
        std::vector<RtVertex> v(nv);
        std::vector<RtTriangle> t(nt ? nt : 1);
        int mode = U(6);
        for (auto& x : v) for (int k = 0; k < 3; k++) {
            float f = (float)((int)U(2001) - 1000) * (mode == 1 ? 1e30f : 0.01f);
            if (mode == 2 && U(50) == 0) f = NAN;
            if (mode == 3 && U(50) == 0) f = INFINITY;
            if (mode == 4) f = 1.0f; // all coincident
            ((float*)&x)[k] = f;
        }
        for (uint64_t i = 0; i < nt; i++) {
            uint32_t* w = (uint32_t*)&t[i];
            for (int k = 0; k < 3; k++) w[k] = (mode == 5 && U(20) == 0) ? 0xFFFFFFFFu - U(5) : U((int)nv);
            w[3] = U(8);
        }
        for (int strat = 0; strat <= 3; strat++) {
            uint64_t cap = rt_bvh_max_nodes(nt);
            std::vector<RtBvhNode> nodes(cap + 1);
            std::vector<uint32_t> idx(nt + 1);
            uint32_t nn = 0, ni = 0;
            int rc = rt_bvh_build(t.data(), nt, v.data(), nv, strat, nodes.data(), &nn, idx.data(), &ni);
            if (rc == 0) {
                if (nn > cap) { printf("FAIL: %u nodes > cap %llu (nt=%llu strat=%d)\n", nn, (unsigned long long)cap, (unsigned long long)nt, strat); return 1; }
                if (ni > nt && !(nt == 0 && ni == 0)) { printf("FAIL: ni %u > nt %llu\n", ni, (unsigned long long)nt); return 1; }
                // every index in range, leaves within the index array
                for (uint32_t i = 0; i < ni; i++) if (idx[i] >= nt) { printf("FAIL: idx out of range\n"); return 1; }
                for (uint32_t i = 0; i < nn; i++) {
                    const uint32_t* w = (const uint32_t*)&nodes[i];
                    uint32_t l = w[8], r = w[9], s = w[10], c = w[11];
                    if (l == 0xFFFFFFFFu) { if ((uint64_t)s + c > ni) { printf("FAIL: leaf range\n"); return 1; } }
                    else if (l >= nn || (r != 0xFFFFFFFFu && r >= nn)) { printf("FAIL: child out of range\n"); return 1; }
                }
            }
        }
    }
    printf("ok\n");
    return 0;
}
