// CPU oracles.  Written from the behavioural spec, not
// from the reference's loops:
//   * GAUSSIAN: out = (sum w*p) >> 4, exact integer — equals the fp32 path of
//     ConvolutionforGrey/RGB (mpi_convolution.c:301-322) and the fp64-FMA path
//     of the CUDA kernels (cuda_convolution.cu:18-22) because every partial sum
//     is a multiple of 1/16 below 256.
//   * BOX / EDGE: fp32 weights float(w/den), row-major multiply-then-add with
//     no fused multiply-add, truncating store — the order of :304-306.
//   * ZERO / TORUS / CUDA_COMPAT boundary models of §7.1; CUDA_COMPAT follows
//     the frozen-border / wrong-buffer behaviour of cuda_convolution.cu:17,34,
//     :60-61 and :93-97 (§2.7).
//   * b2c_oracle_blur_sharded: host emulation of the engine's frame/epoch halo
//     protocol on a P x Q grid (tests decomposition independence without GPUs).
#include <algorithm>
#include <thread>
#include <vector>

#include "common.h"

namespace {

using b2c::FilterDef;

struct Weights {
    int iw[3][3];
    float fw[3][3];
    bool integer16;
};

Weights make_weights(int filter) {
    const FilterDef &f = b2c::filter_def(filter);
    Weights w;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            w.iw[i][j] = f.w[i][j];
            w.fw[i][j] = (float)(f.w[i][j] / (double)f.den);  // int / 9.0 stored to float, :97-100
        }
    w.integer16 = (f.den == 16);
    return w;
}

// One pixel byte at (r, b) of a padded buffer; taps at byte stride `ch`.
inline uint8_t tap9(const uint8_t *p, int64_t pitch, int ch, const Weights &w) {
    if (w.integer16) {
        int acc = 0;
        for (int i = -1; i <= 1; ++i)
            for (int j = -1; j <= 1; ++j) acc += w.iw[i + 1][j + 1] * p[i * pitch + j * ch];
        return (uint8_t)(acc >> 4);
    }
    volatile float acc = 0.0f;  // volatile: keep the separate round after each mul and add
    for (int i = -1; i <= 1; ++i)
        for (int j = -1; j <= 1; ++j) {
            volatile float prod = (float)p[i * pitch + j * ch] * w.fw[i + 1][j + 1];
            acc = acc + prod;
        }
    return (uint8_t)acc;
}

// ZERO / TORUS on a tight image via a 1-pixel padded double buffer.
void blur_padded(const uint8_t *in, uint8_t *out, int W, int H, int ch, int reps, bool torus,
                 const Weights &w) {
    const int64_t bw = (int64_t)W * ch, pitch = bw + 2 * ch;
    std::vector<uint8_t> a((size_t)(pitch * (H + 2)), 0), b(a.size(), 0);
    for (int r = 0; r < H; ++r) memcpy(&a[(r + 1) * pitch + ch], in + r * bw, (size_t)bw);
    uint8_t *src = a.data(), *dst = b.data();
    for (int t = 0; t < reps; ++t) {
        if (torus) {
            for (int r = 1; r <= H; ++r) {  // wrap columns
                uint8_t *row = src + r * pitch;
                memcpy(row, row + bw, (size_t)ch);
                memcpy(row + ch + bw, row + ch, (size_t)ch);
            }
            memcpy(src, src + H * pitch, (size_t)pitch);  // wrap rows (corners included)
            memcpy(src + (H + 1) * pitch, src + pitch, (size_t)pitch);
        }
        for (int r = 1; r <= H; ++r) {
            const uint8_t *s = src + r * pitch + ch;
            uint8_t *d = dst + r * pitch + ch;
            for (int64_t x = 0; x < bw; ++x) d[x] = tap9(s + x, pitch, ch, w);
        }
        std::swap(src, dst);
    }
    for (int r = 0; r < H; ++r) memcpy(out + r * bw, src + (r + 1) * pitch + ch, (size_t)bw);
}

// cuda/cuda_convolution.cu grey path: A = image, B = 0, interior-only writes,
// result = A after 2*floor(reps/2) iterations.
void blur_cuda_compat_grey(const uint8_t *in, uint8_t *out, int W, int H, int reps, const Weights &w) {
    const int64_t n = (int64_t)W * H;
    std::vector<uint8_t> a(in, in + n), b((size_t)n, 0);
    uint8_t *src = a.data(), *dst = b.data();
    const int eff = 2 * (reps / 2);
    for (int t = 0; t < eff; ++t) {
        for (int r = 1; r < H - 1; ++r)
            for (int c = 1; c < W - 1; ++c) dst[(int64_t)r * W + c] = tap9(src + (int64_t)r * W + c, W, 1, w);
        std::swap(src, dst);
    }
    memcpy(out, a.data(), (size_t)n);
}

// RGB path with the `y < 3W-3` guard (:34): a stencil over the linear pixel
// index p = x*W + y for p in [W+1, (H+1)*W-4], reading p + {-W,0,W} + {-1,0,1}.
// Out-of-bounds accesses are modelled as two extra zero-initialised rows.
void blur_cuda_compat_rgb(const uint8_t *in, uint8_t *out, int W, int H, int reps, const Weights &w) {
    const int64_t npx = (int64_t)W * (H + 2);
    std::vector<uint8_t> a((size_t)(npx * 3), 0), b((size_t)(npx * 3), 0);
    memcpy(a.data(), in, (size_t)((int64_t)W * H * 3));
    uint8_t *src = a.data(), *dst = b.data();
    const int eff = 2 * (reps / 2);
    const int64_t p_lo = (int64_t)W + 1, p_hi = (int64_t)(H + 1) * W - 4;
    const bool any = H >= 3 && 3 * (int64_t)W - 3 > 1;  // some thread passes the guard
    for (int t = 0; t < eff && any; ++t) {
        for (int64_t p = p_lo; p <= p_hi; ++p)
            for (int c = 0; c < 3; ++c) dst[p * 3 + c] = tap9(src + p * 3 + c, (int64_t)W * 3, 3, w);
        std::swap(src, dst);
    }
    memcpy(out, a.data(), (size_t)((int64_t)W * H * 3));
}

// Gaussian rows [r0, r1) of a framed buffer, separable and in 16-bit lanes so that the compiler can
// vectorise it: h(x) = p(x-ch) + 2 p(x) + p(x+ch) per row (at most 1020), out = (h_up + 2 h + h_dn) >> 4.
// Three h rows rotate through a small scratch buffer.  Same bytes as tap9 (exact integers).
#if defined(__GNUC__) && defined(__x86_64__) && !defined(__SANITIZE_THREAD__) && !defined(__SANITIZE_ADDRESS__)
__attribute__((target_clones("avx2", "default")))
#endif
void gauss_rows(const uint8_t *__restrict src, uint8_t *__restrict dst, int64_t pitch, int r0, int r1, int row_bytes,
                int ch) {
    std::vector<uint16_t> scratch((size_t)row_bytes * 3);
    uint16_t *h[3] = {scratch.data(), scratch.data() + row_bytes, scratch.data() + 2 * (size_t)row_bytes};
    auto hrow = [&](int r, uint16_t *__restrict o) {
        const uint8_t *__restrict s = src + (int64_t)r * pitch;
        for (int x = 0; x < row_bytes; ++x) o[x] = (uint16_t)(s[x - ch] + 2 * s[x] + s[x + ch]);
    };
    hrow(r0 - 1, h[0]);
    hrow(r0, h[1]);
    for (int r = r0; r < r1; ++r) {
        hrow(r + 1, h[2]);
        const uint16_t *__restrict a = h[0], *__restrict b = h[1], *__restrict c = h[2];
        uint8_t *__restrict d = dst + (int64_t)r * pitch;
        for (int x = 0; x < row_bytes; ++x) d[x] = (uint8_t)((a[x] + 2 * b[x] + c[x]) >> 4);
        uint16_t *t = h[0];
        h[0] = h[1];
        h[1] = h[2];
        h[2] = t;
    }
}

}  // namespace

namespace b2c {

// k Gaussian iterations over one rank's framed buffer (frame = F rows / fb bytes of halo on every
// side).  t0 holds the input and receives the result; t1 is scratch and ends up holding iteration
// k-1.  Bytes that lie outside the global image stay 0 in ZERO mode, which is what makes the
// valid region shrink correctly from pass to pass.  (r0, c0b) = global row / byte column of the
// block's first pixel.
void framed_iterations(std::vector<uint8_t> &t0, std::vector<uint8_t> &t1, int rows, int64_t pitch, int ch, int F,
                       int64_t fb, int r0, int64_t c0b, int H, int64_t Wb, bool torus, int k) {
    static const Weights w = make_weights(B2C_GAUSSIAN);
    t1.assign(t0.size(), 0);
    for (int s = 0; s < k; ++s) {
        for (int r = 1; r < rows - 1; ++r)
            for (int64_t x = ch; x < pitch - ch; ++x) {
                const int gr = r0 + r - F;
                const int64_t gx = c0b + x - fb;
                const bool outside = gr < 0 || gr >= H || gx < 0 || gx >= Wb;
                t1[r * pitch + x] = (!torus && outside) ? 0 : tap9(&t0[r * pitch + x], pitch, ch, w);
            }
        std::swap(t0, t1);
    }
}

}  // namespace b2c

extern "C" {

int b2c_oracle_blur(const uint8_t *in, uint8_t *out, int W, int H, int ch, int reps, int boundary,
                    int filter) {
    if (!in || !out || W <= 0 || H <= 0 || (ch != 1 && ch != 3) || reps < 0) {
        b2c::set_error("oracle_blur: bad arguments");
        return B2C_EINVAL;
    }
    Weights w = make_weights(filter);
    if (boundary == B2C_CUDA_COMPAT) {
        if (filter != B2C_GAUSSIAN) {
            b2c::set_error("cuda-compat exists for the gaussian filter only");
            return B2C_EINVAL;
        }
        if (ch == 1)
            blur_cuda_compat_grey(in, out, W, H, reps, w);
        else
            blur_cuda_compat_rgb(in, out, W, H, reps, w);
        return B2C_OK;
    }
    blur_padded(in, out, W, H, ch, reps, boundary == B2C_TORUS, w);
    return B2C_OK;
}

// One iteration on caller-framed host memory, rows split statically over `threads` threads: the
// CPU counterpart of b2c_stencil_step_device, and — with ranks exchanging the frame over
// torch.distributed/gloo (parallel/sendrecv.py) — this repo's analogue of the reference's MPI and
// MPI+OpenMP programs (Convolution(), mpi_convolution.c:288-299; the static row schedule of
// omp_convolution.c:292,297).
int b2c_stencil_step_host(const uint8_t *src, uint8_t *dst, int64_t pitch, int rows, int row_bytes, int ch,
                          int filter, int threads) {
    if (!src || !dst || src == dst || rows <= 0 || row_bytes <= 0 || (ch != 1 && ch != 3) ||
        pitch < row_bytes + 2 * ch || filter < B2C_GAUSSIAN || filter > B2C_EDGE) {
        b2c::set_error("stencil_step_host: bad arguments");
        return B2C_EINVAL;
    }
    const Weights w = make_weights(filter);
    auto band = [&](int r0, int r1) {
        if (filter == B2C_GAUSSIAN) {
            gauss_rows(src, dst, pitch, r0, r1, row_bytes, ch);
            return;
        }
        for (int r = r0; r < r1; ++r) {
            const uint8_t *s = src + (int64_t)r * pitch;
            uint8_t *d = dst + (int64_t)r * pitch;
            for (int x = 0; x < row_bytes; ++x) d[x] = tap9(s + x, pitch, ch, w);
        }
    };
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min(nt, std::min(rows, 64)));
    if (nt == 1) {
        band(0, rows);
        return B2C_OK;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; ++t)
        pool.emplace_back(band, (int)((int64_t)rows * t / nt), (int)((int64_t)rows * (t + 1) / nt));
    for (auto &t : pool) t.join();
    return B2C_OK;
}

int b2c_blur_host(const uint8_t *in, uint8_t *out, int W, int H, int ch, int reps, int boundary, int filter,
                  int threads) {
    if (boundary == B2C_CUDA_COMPAT) return b2c_oracle_blur(in, out, W, H, ch, reps, boundary, filter);
    if (!in || !out || W <= 0 || H <= 0 || (ch != 1 && ch != 3) || reps < 0 ||
        (boundary != B2C_ZERO && boundary != B2C_TORUS) || filter < B2C_GAUSSIAN || filter > B2C_EDGE) {
        b2c::set_error("blur_host: bad arguments");
        return B2C_EINVAL;
    }
    const int64_t bw = (int64_t)W * ch, pitch = b2c::round_up(bw + 2 * ch, 64);
    if (bw > (1ll << 30)) {
        b2c::set_error("blur_host: row too wide");
        return B2C_ETOOBIG;
    }
    std::vector<uint8_t> a((size_t)(pitch * (H + 2)), 0), b(a.size(), 0);
    for (int r = 0; r < H; ++r) memcpy(&a[(size_t)((r + 1) * pitch + ch)], in + r * bw, (size_t)bw);
    uint8_t *src = a.data(), *dst = b.data();
    for (int t = 0; t < reps; ++t) {
        if (boundary == B2C_TORUS) {
            for (int r = 1; r <= H; ++r) {  // wrap columns, then rows (corners included)
                uint8_t *row = src + r * pitch;
                memcpy(row, row + bw, (size_t)ch);
                memcpy(row + ch + bw, row + ch, (size_t)ch);
            }
            memcpy(src, src + (int64_t)H * pitch, (size_t)(bw + 2 * ch));
            memcpy(src + (int64_t)(H + 1) * pitch, src + pitch, (size_t)(bw + 2 * ch));
        }
        const int rc = b2c_stencil_step_host(src + pitch + ch, dst + pitch + ch, pitch, H, (int)bw, ch, filter, threads);
        if (rc != B2C_OK) return rc;
        std::swap(src, dst);
    }
    for (int r = 0; r < H; ++r) memcpy(out + r * bw, src + (int64_t)(r + 1) * pitch + ch, (size_t)bw);
    return B2C_OK;
}

int b2c_oracle_blur_sharded(const uint8_t *in, uint8_t *out, int W, int H, int ch, int reps,
                            int boundary, int P, int Q, int tb) {
    if (!in || !out || W <= 0 || H <= 0 || (ch != 1 && ch != 3) || reps < 0 || P <= 0 || Q <= 0 ||
        H % P || W % Q || boundary == B2C_CUDA_COMPAT) {
        b2c::set_error("oracle_blur_sharded: bad arguments");
        return B2C_EINVAL;
    }
    const int n = P * Q, R = H / P, C = W / Q;
    const bool torus = boundary == B2C_TORUS;
    // A k-deep halo can only come from the adjacent rank, so k <= block size.
    int F = std::max(1, tb);
    if (n > 1 || torus) F = std::min(F, std::min(R, C));
    const int64_t bw = (int64_t)C * ch, fb = (int64_t)F * ch, pitch = bw + 2 * fb;
    const int rows = R + 2 * F;
    struct Rank {
        std::vector<uint8_t> buf[2];
        int nb[8];
    };
    std::vector<Rank> ranks((size_t)n);
    for (int k = 0; k < n; ++k) {
        for (auto &b : ranks[k].buf) b.assign((size_t)(pitch * rows), 0);
        b2c_neighbours(k, P, Q, boundary, ranks[k].nb);
        const int r0 = (k / Q) * R, c0 = (k % Q) * C;
        for (int r = 0; r < R; ++r)
            memcpy(&ranks[k].buf[0][(r + F) * pitch + fb], in + ((int64_t)(r0 + r) * W + c0) * ch, (size_t)bw);
    }
    // Copy the depth-F ring of every rank's `par` buffer into its neighbours' frames.
    auto push_all = [&](int par) {
        for (int k = 0; k < n; ++k)
            for (int d = 0; d < 8; ++d) {
                int nb = ranks[k].nb[d];
                if (nb < 0) continue;
                const int dr = b2c::kDirDr[d], dc = b2c::kDirDc[d];
                // source ring region (local block coordinates, rows x byte-cols)
                int sr0 = dr < 0 ? 0 : (dr > 0 ? R - F : 0), nr = dr ? F : R;
                int64_t sc0 = dc < 0 ? 0 : (dc > 0 ? bw - fb : 0), nc = dc ? fb : bw;
                // destination in the neighbour's frame: shifted by one block
                int tr0 = sr0 - dr * R;
                int64_t tc0 = sc0 - dc * bw;
                for (int r = 0; r < nr; ++r)
                    memcpy(&ranks[nb].buf[par][(tr0 + r + F) * pitch + fb + tc0],
                           &ranks[k].buf[par][(sr0 + r + F) * pitch + fb + sc0], (size_t)nc);
            }
    };
    push_all(0);
    int par = 0, left = reps;
    std::vector<uint8_t> t0, t1;
    while (left > 0) {
        const int k = std::min(left, F);
        for (int q = 0; q < n; ++q) {
            Rank &rk = ranks[q];
            const int r0 = (q / Q) * R;
            const int64_t c0b = (int64_t)(q % Q) * bw;
            t0 = rk.buf[par];
            b2c::framed_iterations(t0, t1, rows, pitch, ch, F, fb, r0, c0b, H, (int64_t)W * ch, torus, k);
            for (int r = 0; r < R; ++r)
                memcpy(&rk.buf[par ^ 1][(r + F) * pitch + fb], &t0[(r + F) * pitch + fb], (size_t)bw);
        }
        par ^= 1;
        left -= k;
        push_all(par);
    }
    for (int k = 0; k < n; ++k) {
        const int r0 = (k / Q) * R, c0 = (k % Q) * C;
        for (int r = 0; r < R; ++r)
            memcpy(out + ((int64_t)(r0 + r) * W + c0) * ch, &ranks[k].buf[par][(r + F) * pitch + fb], (size_t)bw);
    }
    return B2C_OK;
}

}  // extern "C"
