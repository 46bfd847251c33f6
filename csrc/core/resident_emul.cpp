// Tile planner of the register-resident kernel, and a threaded host model of its CTA protocol.
//
// One thread plays one CTA of csrc/cuda/stencil_resident.cu: a private window (the registers), K
// iterations on it, the K-deep ring of the valid region written into a shared exchange image chosen
// by epoch parity (plain bytes, like the kernel's L2 stores), an epoch flag released, the (up to) 8
// neighbours' flags acquired, the halo read back.  Random sleeps skew the CTAs.  Built with
// -fsanitize=thread (tests/test_halo_model.py) it checks what the GPU run cannot show directly:
// that ring stores of epoch e+2 never overlap a neighbour's halo loads of epoch e with only two
// exchange images and no acknowledgement.  The arithmetic is the plain oracle tap; the lane-level
// arithmetic of the kernel is covered by scripts/resident_model.py.
// No counterpart in the reference (cuda/cuda_convolution.cu has one kernel launch per iteration).
#include <atomic>
#include <chrono>
#include <memory>
#include <random>
#include <thread>
#include <vector>

#include "common.h"
#include "resident_plan.h"

namespace b2c {

bool resident_plan(int bw, int H, int ch, int k, int sms, ResidentPlan *out) {
    if (k != 4 && k != 8 && k != 16) return false;
    const int HB = res_halo_bytes(ch, k), VC = kResWinB - 2 * HB;
    if (bw <= 0 || H < k || bw < HB || bw % kResSB != 0) return false;
    const int strips = (int)ceil_div(bw, VC);
    if (bw - (strips - 1) * VC < HB) return false;   // a halo must come from one neighbour only
    static const int cand[] = {4, 6, 8, 10, 12, 14, 16};
    for (int rw : cand) {
        const int VR = kResWarps * rw - 2 * k;
        if (VR < k) continue;
        const int ctarows = (int)ceil_div(H, VR);
        if (H - (ctarows - 1) * VR < k) continue;
        if ((long long)strips * ctarows > sms) continue;
        if (out) *out = ResidentPlan{rw, k, strips, ctarows, VR, VC};
        return true;
    }
    return false;
}

}  // namespace b2c

namespace {

using Clock = std::chrono::steady_clock;
using b2c::ResidentPlan;

struct Tiles {
    int H, bw, ch, reps, K, HB;
    bool torus;
    ResidentPlan pl;
    int max_delay_us;
    unsigned seed;
    std::vector<uint8_t> xbuf[2];                 // shared exchange images, H x bw
    std::unique_ptr<std::atomic<uint32_t>[]> flag;  // one epoch word per CTA
    std::atomic<bool> failed{false};
    Clock::time_point deadline;

    bool wrap(int &gr, int &gc) const {   // -> inside the image (after wrapping in torus mode)
        if (torus) {
            gr = gr < 0 ? gr + H : (gr >= H ? gr - H : gr);
            gc = gc < 0 ? gc + bw : (gc >= bw ? gc - bw : gc);
        }
        return gr >= 0 && gr < H && gc >= 0 && gc < bw;
    }

    void run_cta(int cta, const uint8_t *in, uint8_t *out) {
        std::mt19937 rng(seed * 7919u + (unsigned)cta);
        auto jitter = [&] {
            if (max_delay_us > 0)
                std::this_thread::sleep_for(std::chrono::microseconds(rng() % (unsigned)(max_delay_us + 1)));
        };
        const int bx = cta % pl.strips, by = cta / pl.strips;
        const int HWIN = b2c::kResWarps * pl.rw, WINB = b2c::kResWinB;
        const int vr0 = by * pl.vr, vr1 = std::min(vr0 + pl.vr, H);
        const int vc0 = bx * pl.vc, vc1 = std::min(vc0 + pl.vc, bw);
        const int row0 = vr0 - K, col0 = vc0 - HB;
        auto valid = [&](int gr, int gc) { return gr >= vr0 && gr < vr1 && gc >= vc0 && gc < vc1; };
        auto near = [&](int gr, int gc) { return gr >= vr0 - K && gr < vr1 + K && gc >= vc0 - HB && gc < vc1 + HB; };
        auto ring = [&](int gr, int gc) {
            return valid(gr, gc) && (gr < vr0 + K || gr >= vr1 - K || gc < vc0 + HB || gc >= vc1 - HB);
        };
        // the window sits inside a 1-cell border of zeros: rows / bytes beyond it "do not exist"
        const int P = WINB + 2 * ch;
        std::vector<uint8_t> a((size_t)P * (HWIN + 2), 0), b(a.size(), 0);
        auto at = [&](std::vector<uint8_t> &v, int r, int x) -> uint8_t & { return v[(size_t)(r + 1) * P + ch + x]; };
        for (int r = 0; r < HWIN; ++r)
            for (int x = 0; x < WINB; ++x) {
                int gr = row0 + r, gc = col0 + x;
                if (near(gr, gc) && wrap(gr, gc)) at(a, r, x) = in[(size_t)gr * bw + gc];
            }
        uint32_t epoch = 0;
        int done = 0;
        while (true) {
            const int k = std::min(K, reps - done);
            for (int it = 0; it < k; ++it) {
                for (int r = 0; r < HWIN; ++r)
                    for (int x = 0; x < WINB; ++x) {
                        const uint8_t *p = &at(a, r, x);
                        const int acc = p[-P - ch] + 2 * p[-P] + p[-P + ch] + 2 * p[-ch] + 4 * p[0] + 2 * p[ch] +
                                        p[P - ch] + 2 * p[P] + p[P + ch];
                        const int gr = row0 + r, gc = col0 + x;
                        const bool inside = torus || (gr >= 0 && gr < H && gc >= 0 && gc < bw);
                        at(b, r, x) = inside ? (uint8_t)(acc >> 4) : 0;
                    }
                a.swap(b);
            }
            done += k;
            if (done >= reps) break;
            ++epoch;
            std::vector<uint8_t> &xb = xbuf[epoch & 1u];
            jitter();
            for (int r = 0; r < HWIN; ++r)
                for (int x = 0; x < WINB; ++x) {
                    const int gr = row0 + r, gc = col0 + x;
                    if (ring(gr, gc)) xb[(size_t)gr * bw + gc] = at(a, r, x);
                }
            flag[cta].store(epoch, std::memory_order_release);
            jitter();
            for (int d = 0; d < 8; ++d) {
                int ny = by + b2c::kDirDr[d], nx = bx + b2c::kDirDc[d];
                if (torus) {
                    ny = (ny + pl.ctarows) % pl.ctarows;
                    nx = (nx + pl.strips) % pl.strips;
                } else if (ny < 0 || ny >= pl.ctarows || nx < 0 || nx >= pl.strips) {
                    continue;
                }
                std::atomic<uint32_t> &f = flag[ny * pl.strips + nx];
                while ((int32_t)(f.load(std::memory_order_acquire) - epoch) < 0) {
                    if (failed.load(std::memory_order_relaxed) || Clock::now() > deadline) {
                        failed.store(true);
                        return;
                    }
                    std::this_thread::yield();
                }
            }
            jitter();
            for (int r = 0; r < HWIN; ++r)
                for (int x = 0; x < WINB; ++x) {
                    int gr = row0 + r, gc = col0 + x;
                    if (valid(gr, gc) || !near(gr, gc)) continue;
                    at(a, r, x) = wrap(gr, gc) ? xb[(size_t)gr * bw + gc] : 0;
                }
        }
        for (int r = 0; r < HWIN; ++r)
            for (int x = 0; x < WINB; ++x) {
                const int gr = row0 + r, gc = col0 + x;
                if (valid(gr, gc)) out[(size_t)gr * bw + gc] = at(a, r, x);
            }
    }
};

}  // namespace

extern "C" {

int b2c_resident_plan(int row_bytes, int rows, int channels, int k, int sms, int out[5]) {
    ResidentPlan pl;
    if (!b2c::resident_plan(row_bytes, rows, channels, k, sms, &pl)) return 0;
    if (out) { out[0] = pl.rw; out[1] = pl.strips; out[2] = pl.ctarows; out[3] = pl.vr; out[4] = pl.vc; }
    return 1;
}

int b2c_emulate_resident_threads(const uint8_t *in, uint8_t *out, int W, int H, int ch, int reps, int boundary, int k,
                                 int sms, int max_delay_us, unsigned seed) {
    Tiles t;
    if (!in || !out || in == out || W <= 0 || H <= 0 || (ch != 1 && ch != 3) || reps < 0 ||
        (boundary != B2C_ZERO && boundary != B2C_TORUS) || !b2c::resident_plan(W * ch, H, ch, k, sms, &t.pl)) {
        b2c::set_error("emulate_resident_threads: bad arguments or the image does not fit the tile plan");
        return B2C_EINVAL;
    }
    t.H = H, t.bw = W * ch, t.ch = ch, t.reps = reps, t.K = k, t.HB = b2c::res_halo_bytes(ch, k);
    t.torus = boundary == B2C_TORUS;
    t.max_delay_us = max_delay_us, t.seed = seed;
    const int n = t.pl.strips * t.pl.ctarows;
    for (auto &x : t.xbuf) x.assign((size_t)H * t.bw, 0xCD);
    t.flag.reset(new std::atomic<uint32_t>[(size_t)n]);
    for (int i = 0; i < n; ++i) t.flag[i].store(0);
    t.deadline = Clock::now() + std::chrono::seconds(30);
    std::vector<std::thread> th;
    for (int c = 0; c < n; ++c) th.emplace_back([&t, c, in, out] { t.run_cta(c, in, out); });
    for (auto &x : th) x.join();
    if (t.failed.load()) {
        b2c::set_error("emulate_resident_threads: a CTA waited past the deadline (protocol deadlock)");
        return B2C_ETIMEOUT;
    }
    return B2C_OK;
}

}  // extern "C"
