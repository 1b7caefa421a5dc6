// asan_main.cpp — TEST INFRASTRUCTURE ONLY (tests/test_hostcheck_asan_cpu.py): the production kernel body, run on the host
// by hostcheck.cpp, under AddressSanitizer.  compute-sanitizer is not available on the GPU pool, so out-of-bounds stores of
// the body's index arithmetic (ragged frames, bands, packed strips, strips into a whole frame, AOV arrays) are looked for
// here: every output buffer is a malloc of EXACTLY the size the C ABI documents, so a store one byte outside lands in an
// ASan red zone.  Cases come from a binary file the test writes; exit code 0 = every case ran and ASan stayed silent.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

extern "C" int hc_paint(int exact, int rays_per_thread, uint32_t cmd_count, const uint32_t *words, uint32_t n_words,
                        const void *uniforms144, const void *limits12, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows,
                        uint32_t strip_rows, uint32_t strip_pitch, int dst_frame_rows, uint8_t *dst, size_t pitch,
                        uint64_t *counters4, uint8_t *aov_cls, float *aov_t);

struct Case {
    int32_t exact, rays;
    uint32_t cmd_count, n_words;
    uint32_t W, H, row0, nrows, strip_rows, strip_pitch;
    int32_t dst_frame_rows;
    uint32_t dst_rows;   // rows of the destination buffer (nrows packed, H for a whole frame)
    int32_t variant;     // 0 plain, 1 counters, 2 AOV
};

int main(int argc, char **argv) {
    if (argc < 2) return 2;
    FILE *f = fopen(argv[1], "rb");
    if (!f) return 2;
    uint32_t n_cases = 0;
    if (fread(&n_cases, 4, 1, f) != 1) return 2;
    unsigned long long checksum = 0;
    for (uint32_t i = 0; i < n_cases; ++i) {
        Case c;
        if (fread(&c, sizeof c, 1, f) != 1) return 2;
        std::vector<uint32_t> words(c.n_words);
        unsigned char uniforms[144], limits[12];
        if (c.n_words && fread(words.data(), 4, c.n_words, f) != c.n_words) return 2;
        if (fread(uniforms, 1, 144, f) != 144 || fread(limits, 1, 12, f) != 12) return 2;
        const size_t pitch = (size_t)c.W * 4, dst_bytes = pitch * c.dst_rows;
        uint8_t *dst = (uint8_t *)malloc(dst_bytes ? dst_bytes : 1);
        memset(dst, 0, dst_bytes);
        uint64_t *counters = c.variant == 1 ? (uint64_t *)calloc(8, 8) : nullptr;   // the library's counter block: 8 x u64
        const size_t n_samples = (size_t)c.W * c.nrows * 16;
        uint8_t *cls = c.variant == 2 ? (uint8_t *)malloc(n_samples ? n_samples : 1) : nullptr;
        float *t = c.variant == 2 ? (float *)malloc(n_samples ? n_samples * 4 : 4) : nullptr;
        const int rc = hc_paint(c.exact, c.rays, c.cmd_count, words.data(), c.n_words, uniforms, limits, c.W, c.H, c.row0, c.nrows,
                                c.strip_rows, c.strip_pitch, c.dst_frame_rows, dst, pitch, counters, cls, t);
        if (rc != 0) { fprintf(stderr, "case %u: hc_paint returned %d\n", i, rc); return 3; }
        for (size_t k = 0; k < dst_bytes; ++k) checksum += dst[k];
        if (cls) for (size_t k = 0; k < n_samples; ++k) checksum += cls[k];
        if (counters) checksum += counters[0] + counters[3];
        free(dst); free(counters); free(cls); free(t);
    }
    fclose(f);
    printf("asan tour ok: %u cases, checksum %llu\n", n_cases, checksum);
    return 0;
}
