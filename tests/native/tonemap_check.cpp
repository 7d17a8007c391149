// TEST INFRASTRUCTURE — restatement of the reference's display transform (src/window/window.cpp:507-531 with
// src/types/vector.cpp:30-36 clampToOne and Vector3 / Vector3 = Eigen cwiseQuotient), compiled with the reference's flags
// (-O3 -msse -msse2 -mavx -mavx2, no FMA contraction, no fast-math) against the same libm the reference links.
//
//   tonemap_check apply <in: W*H float4> <out: W*H RGBA8> <tone 0|1>     what the window would put on screen (bytes R, G, B, 255)
//   tonemap_check steps <out: 256 floats>                                 the byte's step function over v = c/(1+c) in [0, 1]:
//        verifies over ALL floats of [0, 1] (2^30 + 2^23 + 1 values, threaded) that the byte never decreases as v grows,
//        and writes for every byte value b the smallest v whose byte is >= b (out[0] = 0).  Exit code 1 if not monotone.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#define INV_GAMMA 0.4545f /* window.cpp:32 */

static inline unsigned char to_byte(float c) { return (unsigned char)(c * 255); } // window.cpp:520-522

static inline unsigned char channel(float c, bool tone) {
    if (tone) {
        c = c / (1.0f + c);      // c / (oneVector + c), window.cpp:513
        c = powf(c, INV_GAMMA);  // std::powf, window.cpp:515-517
    }
    c = fminf(c, 1.0f); // clampToOne
    return to_byte(c);
}

static inline unsigned char step_byte(float v) { return to_byte(fminf(powf(v, INV_GAMMA), 1.0f)); }

int main(int argc, char **argv) {
    if (argc >= 5 && !strcmp(argv[1], "apply")) {
        FILE *f = fopen(argv[2], "rb");
        if (!f) return 2;
        fseek(f, 0, SEEK_END);
        size_t n = ftell(f) / 16;
        fseek(f, 0, SEEK_SET);
        std::vector<float> img(n * 4);
        if (fread(img.data(), 16, n, f) != n) return 2;
        fclose(f);
        const bool tone = atoi(argv[4]) != 0;
        std::vector<unsigned char> out(n * 4);
        for (size_t p = 0; p < n; ++p) {
            for (int k = 0; k < 3; ++k) out[4 * p + k] = channel(img[4 * p + k], tone);
            out[4 * p + 3] = 255;
        }
        f = fopen(argv[3], "wb");
        if (!f) return 2;
        fwrite(out.data(), 1, out.size(), f);
        fclose(f);
        return 0;
    }
    if (argc >= 3 && !strcmp(argv[1], "steps")) {
        const uint32_t last = 0x3f800000u; // 1.0f; non-negative floats order like their bit patterns
        const unsigned T = std::max(1u, std::thread::hardware_concurrency());
        std::vector<uint32_t> first(T * 256, 0xffffffffu);
        std::vector<int> bad(T, 0);
        std::vector<std::thread> th;
        for (unsigned k = 0; k < T; ++k)
            th.emplace_back([&, k]() {
                const uint64_t lo = (uint64_t)(last + 1ull) * k / T, hi = (uint64_t)(last + 1ull) * (k + 1) / T;
                int prev = -1;
                for (uint64_t u = lo; u < hi; ++u) {
                    const uint32_t bits = (uint32_t)u;
                    float v;
                    memcpy(&v, &bits, 4);
                    const int b = step_byte(v);
                    if (b < prev) bad[k] = 1;
                    if (first[k * 256 + b] == 0xffffffffu) first[k * 256 + b] = bits;
                    prev = b;
                }
            });
        for (auto &t : th) t.join();
        // across chunk borders: the last byte of chunk k must not exceed the first of chunk k+1 — covered by merging below
        float out[256];
        uint32_t bestFirst[256];
        for (int b = 0; b < 256; ++b) {
            bestFirst[b] = 0xffffffffu;
            for (unsigned k = 0; k < T; ++k) bestFirst[b] = std::min(bestFirst[b], first[k * 256 + b]);
        }
        int notMonotone = 0;
        for (unsigned k = 0; k < T; ++k) notMonotone |= bad[k];
        // smallest v with byte >= b = min over b' >= b of the first occurrence of b'; monotone <=> first occurrences ascend
        uint32_t run = 0xffffffffu;
        for (int b = 255; b >= 1; --b) {
            run = std::min(run, bestFirst[b]);
            memcpy(&out[b], &run, 4);
        }
        out[0] = 0.0f;
        uint32_t prev = 0;
        for (int b = 1; b < 256; ++b)
            if (bestFirst[b] != 0xffffffffu) {
                if (bestFirst[b] < prev) notMonotone = 1;
                prev = bestFirst[b];
            }
        FILE *f = fopen(argv[2], "wb");
        if (!f) return 2;
        fwrite(out, 4, 256, f);
        fclose(f);
        printf("checked %u floats on %u threads: %s\n", last + 1u, T, notMonotone ? "NOT monotone" : "monotone");
        return notMonotone ? 1 : 0;
    }
    fprintf(stderr, "usage: tonemap_check apply <in> <out> <tone> | steps <out>\n");
    return 2;
}
