// Randomised check of datafusion_upstream_b200/csrc/bitcopy.h against a per-bit model (run by tests/test_bitcopy_cpu.py).
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "bitcopy.h"

static uint64_t rng_state = 88172645463325252ULL;
static uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }

int main() {
    for (int iter = 0; iter < 20000; iter++) {
        std::vector<int> model;                       // the bitmap under construction, one int per bit
        std::vector<uint8_t> dst(600, 0);
        for (auto &b : dst) b = (uint8_t)rnd();       // unspecified memory
        int pieces = 1 + (int)(rnd() % 8);
        for (int p = 0; p < pieces; p++) {
            int64_t n = (int64_t)(rnd() % (p % 3 == 0 ? 200 : 20));
            int64_t pos = (int64_t)model.size();
            if (pos + n > 4000) break;
            if (rnd() % 4 == 0) {
                dfgpu::bits_fill_ones(dst.data(), pos, n);
                for (int64_t i = 0; i < n; i++) model.push_back(1);
            } else {
                int64_t soff = (int64_t)(rnd() % 70);
                std::vector<uint8_t> src((size_t)((soff + n + 7) / 8) + 1);
                for (auto &b : src) b = (uint8_t)rnd();
                // exact-size copy so an over-read of the source shows up under a sanitizer / valgrind
                std::vector<uint8_t> exact(src.begin(), src.begin() + (size_t)((soff + n + 7) / 8));
                dfgpu::bits_append(dst.data(), pos, exact.data(), soff, n);
                for (int64_t i = 0; i < n; i++) model.push_back((src[(size_t)((soff + i) >> 3)] >> ((soff + i) & 7)) & 1);
            }
            int64_t total = (int64_t)model.size();
            for (int64_t i = 0; i < total; i++)
                if (((dst[(size_t)(i >> 3)] >> (i & 7)) & 1) != model[(size_t)i]) { printf("FAIL iter %d piece %d bit %lld\n", iter, p, (long long)i); return 1; }
            if (n > 0 && (total & 7)) {               // the rest of the last touched byte is zero
                if (dst[(size_t)(total >> 3)] >> (total & 7)) { printf("FAIL tail iter %d piece %d\n", iter, p); return 1; }
            }
        }
    }
    printf("OK\n");
    return 0;
}
