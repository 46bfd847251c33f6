// Native self-test of the CPU half of libb200conv_core (csrc/core/*.cpp).
// tests/test_native_selftest.py builds it with -fsanitize=address,undefined and runs it: every
// public CPU entry point is driven through its edge cases so that out-of-bounds accesses, signed
// overflow, misaligned loads and leaks in the C++ show up without a GPU.
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "b200conv.h"

static int g_fail = 0;
#define EXPECT(cond)                                                       \
    do {                                                                   \
        if (!(cond)) {                                                     \
            fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            ++g_fail;                                                      \
        }                                                                  \
    } while (0)

static std::vector<uint8_t> noise(size_t n, unsigned seed) {
    std::vector<uint8_t> v(n);
    unsigned x = seed * 2654435761u + 1u;
    for (auto &b : v) {
        x = x * 1664525u + 1013904223u;
        b = (uint8_t)(x >> 24);
    }
    return v;
}

static void test_decomposition() {
    // (n, H, W) -> grid rows
    EXPECT(b2c_rows_division(4, 2520, 1920) == 2);
    EXPECT(b2c_rows_division(1, 2520, 1920) == 1);
    EXPECT(b2c_rows_division(16, 2520, 1920) == 4);
    EXPECT(b2c_rows_division(7, 10, 10) == 0);
    int P = -1, Q = -1;
    EXPECT(b2c_decompose(8, 32768, 32768, &P, &Q) == B2C_OK && P == 2 && Q == 4);
    EXPECT(b2c_decompose(7, 10, 10, &P, &Q) == B2C_EDIVIDE);
    EXPECT(b2c_decompose(0, 10, 10, &P, &Q) < 0);
    int nb[8];
    EXPECT(b2c_neighbours(0, 2, 2, B2C_ZERO, nb) == B2C_OK);
    EXPECT(nb[B2C_N] == -1 && nb[B2C_S] == 2 && nb[B2C_W] == -1 && nb[B2C_E] == 1 && nb[B2C_SE] == 3 && nb[B2C_NW] == -1);
    EXPECT(b2c_neighbours(0, 2, 2, B2C_TORUS, nb) == B2C_OK);
    EXPECT(nb[B2C_N] == 2 && nb[B2C_S] == 2 && nb[B2C_W] == 1 && nb[B2C_E] == 1 && nb[B2C_NW] == 3);
    EXPECT(b2c_neighbours(0, 1, 1, B2C_TORUS, nb) == B2C_OK && nb[B2C_N] == 0 && nb[B2C_SE] == 0);
    EXPECT(b2c_neighbours(5, 2, 2, B2C_ZERO, nb) < 0);
    for (int n = 1; n <= 64; ++n)  // every answer is a valid grid
        for (int h : {1, 6, 96, 2520})
            for (int w : {1, 8, 90, 1920}) {
                const int r = b2c_rows_division(n, h, w);
                if (r) EXPECT(n % r == 0 && h % r == 0 && w % (n / r) == 0);
            }
}

static void test_cli() {
    b2c_cli c;
    {
        const char *argv[] = {"prog", "img.raw", "1920", "2520", "40", "rgb"};
        EXPECT(b2c_parse_cli(6, (char **)argv, &c) == B2C_OK);
        EXPECT(c.width == 1920 && c.height == 2520 && c.repetitions == 40 && c.channels == 3 && c.boundary == B2C_ZERO);
    }
    {
        const char *argv[] = {"prog", "img.raw", "8", "8", "2", "grey", "--gpus", "4", "--boundary", "torus", "--tb", "2",
                              "--check-every", "20", "--json", "r.json", "--profile"};
        EXPECT(b2c_parse_cli(17, (char **)argv, &c) == B2C_OK);
        EXPECT(c.channels == 1 && c.gpus == 4 && c.boundary == B2C_TORUS && c.tb == 2 && c.check_every == 20 && c.profile == 1);
        EXPECT(c.json && !strcmp(c.json, "r.json"));
    }
    {
        const char *argv[] = {"prog", "img.raw", "8", "8", "2", "gray"};
        EXPECT(b2c_parse_cli(6, (char **)argv, &c) == B2C_EINVAL);
        EXPECT(b2c_parse_cli(5, (char **)argv, &c) == B2C_EINVAL);
        EXPECT(b2c_parse_cli(1, (char **)argv, &c) == B2C_EINVAL);
    }
    {
        const char *argv3[] = {"prog", "img.raw", "8", "8", "2", "rgb", "--cpu", "all"};
        EXPECT(b2c_parse_cli(8, (char **)argv3, &c) == B2C_OK && c.cpu == -1);
        const char *argv4[] = {"prog", "img.raw", "8", "8", "2", "rgb", "--cpu", "0"};
        EXPECT(b2c_parse_cli(8, (char **)argv4, &c) == B2C_EINVAL);
        const char *argv[] = {"prog", "img.raw", "8", "8", "2", "rgb", "--boundary"};  // value missing
        EXPECT(b2c_parse_cli(7, (char **)argv, &c) == B2C_EINVAL);
        const char *argv2[] = {"prog", "img.raw", "8", "8", "2", "rgb", "--no-such-flag", "1"};
        EXPECT(b2c_parse_cli(8, (char **)argv2, &c) == B2C_EINVAL);
    }
    char buf[256], tiny[8];
    EXPECT(b2c_usage_text("./prog", 0, buf, sizeof buf) == B2C_OK && strstr(buf, "Image_name width height repetitions [rgb/grey]"));
    EXPECT(b2c_usage_text("./prog", 1, buf, sizeof buf) == B2C_OK && strstr(buf, "image_name width height repetitions [rgb/grey]"));
    EXPECT(b2c_usage_text("./a-very-long-program-name", 0, tiny, sizeof tiny) == B2C_EINVAL);  // truncates, no overflow
    EXPECT(strlen(tiny) < sizeof tiny);
    EXPECT(b2c_output_name("dir/x.raw", buf, sizeof buf) >= 0 && !strcmp(buf, "blur_dir/x.raw"));
    EXPECT(b2c_output_name("a-name-that-does-not-fit", tiny, sizeof tiny) < 0);
}

static void test_rawio(const std::string &dir) {
    const int W = 37, H = 29, CH = 3;
    const std::vector<uint8_t> img = noise((size_t)W * H * CH, 1);
    const std::string path = dir + "/in.raw", out = dir + "/out.raw";
    FILE *f = fopen(path.c_str(), "wb");
    EXPECT(f != nullptr);
    if (!f) return;
    fwrite(img.data(), 1, img.size(), f);
    fclose(f);
    // block read with a padded pitch, block write into a fresh file, whole-file compare
    const int r0 = 5, c0 = 7, R = 11, C = 13;
    const int64_t pitch = C * CH + 9;
    std::vector<uint8_t> blk((size_t)(pitch * R), 0xEE);
    EXPECT(b2c_read_block(path.c_str(), blk.data(), pitch, W, H, CH, r0, c0, R, C) == (int64_t)R * C * CH);
    for (int r = 0; r < R; ++r) {
        EXPECT(!memcmp(&blk[(size_t)(r * pitch)], &img[((size_t)(r0 + r) * W + c0) * CH], (size_t)C * CH));
        EXPECT(blk[(size_t)(r * pitch + C * CH)] == 0xEE);  // pad bytes untouched
    }
    std::vector<uint8_t> whole(img.size());
    EXPECT(b2c_read_block(path.c_str(), whole.data(), (int64_t)W * CH, W, H, CH, 0, 0, H, W) == (int64_t)img.size());
    EXPECT(whole == img);
    for (int k = 0; k < 4; ++k) {  // four ranks write their quadrant-ish blocks of one file
        const int br = k / 2, bc = k % 2;
        const int rr0 = br ? 14 : 0, rr = br ? H - 14 : 14, cc0 = bc ? 20 : 0, cc = bc ? W - 20 : 20;
        EXPECT(b2c_write_block(out.c_str(), &img[((size_t)rr0 * W + cc0) * CH], (int64_t)W * CH, W, H, CH, rr0, cc0, rr, cc,
                               1) == (int64_t)rr * cc * CH);
    }
    std::vector<uint8_t> back(img.size());
    EXPECT(b2c_read_block(out.c_str(), back.data(), (int64_t)W * CH, W, H, CH, 0, 0, H, W) == (int64_t)img.size());
    EXPECT(back == img);
    // errors: outside the image, missing file, short file
    EXPECT(b2c_read_block(path.c_str(), blk.data(), pitch, W, H, CH, H - 2, 0, R, C) < 0);
    EXPECT(b2c_read_block((dir + "/nope.raw").c_str(), blk.data(), pitch, W, H, CH, 0, 0, R, C) < 0);
    EXPECT(truncate(path.c_str(), (off_t)img.size() / 2) == 0);
    EXPECT(b2c_read_block(path.c_str(), whole.data(), (int64_t)W * CH, W, H, CH, 0, 0, H, W) < 0);
    // offsets beyond 2^31 on a sparse file (the reference's int offsets overflow there)
    const std::string big = dir + "/sparse.raw";
    const int BW = 65536, BH = 40000;  // 2.6e9 bytes, grey
    std::vector<uint8_t> row = noise(64, 9), got(64);
    EXPECT(b2c_write_block(big.c_str(), row.data(), 64, BW, BH, 1, BH - 1, BW - 64, 1, 64, 1) == 64);
    EXPECT(b2c_read_block(big.c_str(), got.data(), 64, BW, BH, 1, BH - 1, BW - 64, 1, 64) == 64);
    EXPECT(got == row);
    struct stat st;
    EXPECT(stat(big.c_str(), &st) == 0 && st.st_size == (off_t)BW * BH);
    unlink(big.c_str());
}

static void test_oracles() {
    for (int ch : {1, 3})
        for (int boundary : {B2C_ZERO, B2C_TORUS, B2C_CUDA_COMPAT})
            for (int W : {1, 2, 3, 17})
                for (int H : {1, 2, 3, 12}) {
                    const std::vector<uint8_t> in = noise((size_t)W * H * ch, (unsigned)(W * 31 + H));
                    std::vector<uint8_t> out(in.size(), 0xAB);
                    for (int reps : {0, 1, 2, 5}) {
                        EXPECT(b2c_oracle_blur(in.data(), out.data(), W, H, ch, reps, boundary, B2C_GAUSSIAN) == B2C_OK);
                        if (reps == 0) EXPECT(out == in);
                    }
                }
    // a constant image stays constant on a torus; on a zero-padded plane only the border decays
    std::vector<uint8_t> flat(20 * 20, 200), out(flat.size());
    EXPECT(b2c_oracle_blur(flat.data(), out.data(), 20, 20, 1, 3, B2C_TORUS, B2C_GAUSSIAN) == B2C_OK && out == flat);
    EXPECT(b2c_oracle_blur(flat.data(), out.data(), 20, 20, 1, 3, B2C_ZERO, B2C_GAUSSIAN) == B2C_OK);
    EXPECT(out[10 * 20 + 10] == 200 && out[0] < 200);
    // BOX / EDGE take the fp32 path; cuda-compat exists for the gaussian only
    EXPECT(b2c_oracle_blur(flat.data(), out.data(), 20, 20, 1, 2, B2C_ZERO, B2C_BOX) == B2C_OK);
    EXPECT(b2c_oracle_blur(flat.data(), out.data(), 20, 20, 1, 2, B2C_ZERO, B2C_EDGE) == B2C_OK);
    EXPECT(b2c_oracle_blur(flat.data(), out.data(), 20, 20, 1, 2, B2C_CUDA_COMPAT, B2C_BOX) == B2C_EINVAL);
    EXPECT(b2c_oracle_blur(flat.data(), out.data(), 20, 20, 2, 2, B2C_ZERO, B2C_GAUSSIAN) == B2C_EINVAL);
    EXPECT(b2c_oracle_blur(nullptr, out.data(), 20, 20, 1, 2, B2C_ZERO, B2C_GAUSSIAN) == B2C_EINVAL);
    // in-place call
    std::vector<uint8_t> a = noise(24 * 24 * 3, 5), want(a.size());
    EXPECT(b2c_oracle_blur(a.data(), want.data(), 24, 24, 3, 4, B2C_ZERO, B2C_GAUSSIAN) == B2C_OK);
    EXPECT(b2c_oracle_blur(a.data(), a.data(), 24, 24, 3, 4, B2C_ZERO, B2C_GAUSSIAN) == B2C_OK && a == want);
}

static void test_host_step() {
    for (int ch : {1, 3})
        for (int filter : {B2C_GAUSSIAN, B2C_BOX, B2C_EDGE})
            for (int threads : {1, 3, 0}) {
                const int W = 21, H = 13;
                const int64_t bw = (int64_t)W * ch, pitch = bw + 2 * ch + 3;
                const std::vector<uint8_t> in = noise((size_t)(bw * H), 11);
                std::vector<uint8_t> src((size_t)(pitch * (H + 2)), 0), dst(src.size(), 0x5A), want(in.size());
                for (int r = 0; r < H; ++r) memcpy(&src[(size_t)((r + 1) * pitch + ch)], &in[(size_t)(r * bw)], (size_t)bw);
                EXPECT(b2c_stencil_step_host(&src[(size_t)(pitch + ch)], &dst[(size_t)(pitch + ch)], pitch, H, (int)bw, ch, filter,
                                             threads) == B2C_OK);
                EXPECT(b2c_oracle_blur(in.data(), want.data(), W, H, ch, 1, B2C_ZERO, filter) == B2C_OK);
                for (int r = 0; r < H; ++r)
                    EXPECT(!memcmp(&dst[(size_t)((r + 1) * pitch + ch)], &want[(size_t)(r * bw)], (size_t)bw));
                EXPECT(dst[0] == 0x5A && dst[(size_t)(pitch + ch - 1)] == 0x5A);   // frame untouched
            }
    for (int ch : {1, 3})   // whole blur on host threads, in place, against the oracle
        for (int boundary : {B2C_ZERO, B2C_TORUS, B2C_CUDA_COMPAT})
            for (int W : {1, 2, 31})
                for (int H : {1, 3, 17}) {
                    std::vector<uint8_t> img = noise((size_t)W * H * ch, (unsigned)(W + 7 * H)), want(img.size());
                    EXPECT(b2c_oracle_blur(img.data(), want.data(), W, H, ch, 4, boundary, B2C_GAUSSIAN) == B2C_OK);
                    EXPECT(b2c_blur_host(img.data(), img.data(), W, H, ch, 4, boundary, B2C_GAUSSIAN, 2) == B2C_OK);
                    EXPECT(img == want);
                }
    std::vector<uint8_t> a(64, 0);
    EXPECT(b2c_blur_host(a.data(), a.data(), 4, 4, 2, 1, B2C_ZERO, B2C_GAUSSIAN, 1) == B2C_EINVAL);
    EXPECT(b2c_stencil_step_host(a.data(), a.data(), 8, 2, 4, 1, B2C_GAUSSIAN, 1) == B2C_EINVAL);   // in place
    EXPECT(b2c_stencil_step_host(a.data(), a.data() + 32, 4, 2, 4, 1, B2C_GAUSSIAN, 1) == B2C_EINVAL);   // pitch too small
}

static void test_host_models() {
    const int W = 24, H = 24;
    for (int ch : {1, 3})
        for (int boundary : {B2C_ZERO, B2C_TORUS}) {
            const std::vector<uint8_t> in = noise((size_t)W * H * ch, 77);
            std::vector<uint8_t> want(in.size()), got(in.size());
            const int grids[][2] = {{1, 1}, {1, 2}, {2, 2}, {2, 4}, {4, 2}, {3, 3}, {24, 1}, {1, 24}};
            for (auto &g : grids)
                for (int tb : {1, 2, 4, 8}) {
                    const int reps = 2 * tb + 1;
                    EXPECT(b2c_oracle_blur(in.data(), want.data(), W, H, ch, reps, boundary, B2C_GAUSSIAN) == B2C_OK);
                    EXPECT(b2c_oracle_blur_sharded(in.data(), got.data(), W, H, ch, reps, boundary, g[0], g[1], tb) == B2C_OK);
                    EXPECT(got == want);
                    b2c_emul_opts o;
                    memset(&o, 0, sizeof o);
                    o.tb = tb, o.stall_rank = -1;
                    int iters = -1;
                    EXPECT(b2c_emulate_halo_threads(in.data(), got.data(), W, H, ch, reps, boundary, g[0], g[1], &o, &iters) ==
                           B2C_OK);
                    EXPECT(iters == reps && got == want);
                }
            EXPECT(b2c_oracle_blur_sharded(in.data(), got.data(), W, H, ch, 3, boundary, 5, 1, 1) == B2C_EINVAL);  // 24 % 5
            // resident kernel: tile planner and the CTA protocol model (24 x 24 pixels = one tile)
            EXPECT(b2c_oracle_blur(in.data(), want.data(), W, H, ch, 11, boundary, B2C_GAUSSIAN) == B2C_OK);
            EXPECT(b2c_emulate_resident_threads(in.data(), got.data(), W, H, ch, 11, boundary, 4, 148, 0, 1) == B2C_OK);
            EXPECT(got == want);
            EXPECT(b2c_emulate_resident_threads(in.data(), got.data(), W, H, ch, 11, boundary, 5, 148, 0, 1) == B2C_EINVAL);
        }
    int pl[5];
    EXPECT(b2c_resident_plan(1920, 2520, 1, 8, 148, pl) == 1 && pl[0] == 10 && pl[1] == 8 && pl[2] == 18);
    EXPECT(b2c_resident_plan(1924, 2520, 1, 8, 148, pl) == 0);
    EXPECT(b2c_resident_plan(16384, 16384, 1, 4, 148, nullptr) == 0);
}

int main(int argc, char **argv) {
    const std::string dir = argc > 1 ? argv[1] : "/tmp";
    test_decomposition();
    test_cli();
    test_rawio(dir);
    test_oracles();
    test_host_step();
    test_host_models();
    printf("core_selftest: %d failure(s)\n", g_fail);
    return g_fail ? 1 : 0;
}
