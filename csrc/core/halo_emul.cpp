// Host model of the multi-GPU protocol ("host emulation of the halo protocol
// (threads + the same slot/flag state machine) to test epoch logic without GPUs").
//
// One thread plays one rank.  The state machine is the engine's (csrc/cuda/engine.cu,
// stencil_stream.cuh, kernels.cu: converge_kernel), restated with std::atomic:
//
//   * two framed ping-pong buffers per rank; pass t reads buf[(t-1)&1] and writes buf[t&1];
//   * passes form groups of at most F iterations with ONE halo exchange per group: the first pass
//     of a group acquire-waits on the data flag of every direction that has a neighbour; the
//     passes in between also write the part of the frame that is still valid (the block extended
//     by the iterations left in the group); the last pass stores the depth-F ring of its new block
//     straight into the frames of its (up to 8) neighbours' buf[t&1], then release-stores epoch
//     g+1 into the flag word the neighbour keeps for that direction;
//   * a group of one pass needs nothing else (the neighbour's data flag also says that it has
//     finished reading the frame about to be overwritten).  In a longer group the in-between
//     passes read and write frames of both buffers, so the second-to-last pass publishes an "ack"
//     epoch once it has read its frames, and the last pass waits for the neighbours' acks before
//     it pushes;
//   * ring results are published before the interior is written (the GPU orders ring warps first);
//   * a check pass is followed by an all-to-all exchange of one "changed" bit through per-rank
//     slots tagged with the check number and double-buffered by its parity; when nobody changed
//     at check c, every rank still runs interval c+1 (on the GPU the reduction of check c is in
//     flight on a side stream meanwhile — the image is a fixed point, so nothing changes) and
//     skips everything from interval c+2 on.
//
//   * optionally (`courier`) a rank does not store epochs into its neighbours itself: it release-stores
//     them into words of its own, and a second thread per rank — the courier warp of the GPU engine,
//     which keeps the system-scope fence out of the passes — acquires them and release-stores them
//     into the neighbours.  The frames are then ordered through two hops (rank -> courier ->
//     neighbour), which is exactly the cumulativity the GPU version relies on.
//
// Random sleeps can be injected at every point where the GPU ranks may drift apart.  Built with
// -fsanitize=thread (tests/test_halo_model.py) the model shows that the protocol has no data race
// on the frames; with `single_slot_set` it reproduces the lost-tag hang that the parity fixes.
// Nothing here is taken from the reference: mpi_convolution.c:158-234 has neither epochs, nor
// diagonal neighbours, nor a convergence test.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <memory>
#include <random>
#include <thread>
#include <vector>

#include "common.h"

namespace {

using Clock = std::chrono::steady_clock;

struct Rank {
    std::vector<uint8_t> buf[2];
    int nb[8];
    std::atomic<uint32_t> flag[8];
    std::atomic<uint32_t> ack[8];
    std::atomic<uint32_t> ready[16];     // courier mode: epochs handed over (0-7 data, 8-15 ack)
    std::atomic<bool> finished{false};   // courier mode: the rank thread has published its last epoch
    std::unique_ptr<std::atomic<uint32_t>[]> slots;  // 2 sets x n ranks
    int iters = 0;
};

enum { kWaitData = 1, kPostAck = 2, kWaitAck = 4, kPush = 8 };

struct Pass {
    int k;
    bool check;
    int iters_after;
    int ext;         // iterations left in the group after this pass
    int mode;
    bool group_end;
    int interval;    // check interval (1-based; 0 without check_every)
};

// Same rule as make_passes() in engine.cu: groups of at most F iterations (a group also ends on
// every multiple of check_every), each cut into passes of k in {K, K/2, .., 1}.
std::vector<Pass> split_passes(int reps, int K, int F, int ce) {
    std::vector<Pass> v;
    int done = 0;
    while (done < reps) {
        int glen = std::min(reps - done, F);
        if (ce > 0) glen = std::min(glen, ce - done % ce);
        std::vector<int> ks;
        for (int left = glen; left > 0;) {
            int k = K;
            while (k > left) k >>= 1;
            ks.push_back(k);
            left -= k;
        }
        const int n = (int)ks.size();
        int ext = glen;
        for (int j = 0; j < n; ++j) {
            ext -= ks[j];
            done += ks[j];
            int mode = 0;
            if (j == 0) mode |= kWaitData;
            if (n >= 2 && j == n - 2) mode |= kPostAck;
            if (j == n - 1) mode |= kPush | (n >= 2 ? kWaitAck : 0);
            v.push_back({ks[j], j == n - 1 && ce > 0 && done % ce == 0, done, ext, mode, j == n - 1,
                         ce > 0 ? (done - 1) / ce + 1 : 0});
        }
    }
    return v;
}

struct World {
    int W, H, ch, P, Q, R, C, F, n;
    int64_t bw, fb, pitch;
    int rows;
    bool torus;
    b2c_emul_opts o;
    std::vector<Pass> passes;
    std::vector<Rank> ranks;
    std::atomic<int> arrived{0};
    std::atomic<bool> failed{false};
    Clock::time_point deadline;

    // Spin until pred() or the deadline; a timeout marks the whole run as failed.
    template <class Pred>
    bool wait_for(Pred pred) {
        while (!pred()) {
            if (failed.load(std::memory_order_relaxed)) return false;
            if (Clock::now() > deadline) {
                failed.store(true);
                return false;
            }
            std::this_thread::yield();
        }
        return true;
    }

    void push_ring(int q, int par) {
        Rank &me = ranks[q];
        for (int d = 0; d < 8; ++d) {
            const int nb = me.nb[d];
            if (nb < 0) continue;
            const int dr = b2c::kDirDr[d], dc = b2c::kDirDc[d];
            const int sr0 = dr > 0 ? R - F : 0, nr = dr ? F : R;
            const int64_t sc0 = dc > 0 ? bw - fb : 0, nc = dc ? fb : bw;
            const int tr0 = sr0 - dr * R;
            const int64_t tc0 = sc0 - dc * bw;
            for (int r = 0; r < nr; ++r)
                memcpy(&ranks[nb].buf[par][(tr0 + r + F) * pitch + fb + tc0],
                       &me.buf[par][(sr0 + r + F) * pitch + fb + sc0], (size_t)nc);
        }
    }

    void publish(int q, uint32_t epoch) {
        Rank &me = ranks[q];
        for (int d = 0; d < 8; ++d) {
            if (me.nb[d] < 0) continue;
            if (o.courier) me.ready[d].store(epoch, std::memory_order_release);
            else ranks[me.nb[d]].flag[b2c::opposite_dir(d)].store(epoch, std::memory_order_release);
        }
    }

    void publish_ack(int q, uint32_t epoch) {
        Rank &me = ranks[q];
        for (int d = 0; d < 8; ++d) {
            if (me.nb[d] < 0) continue;
            if (o.courier) me.ready[8 + d].store(epoch, std::memory_order_release);
            else ranks[me.nb[d]].ack[b2c::opposite_dir(d)].store(epoch, std::memory_order_release);
        }
    }

    // The courier of rank q: forwards whatever appears in the rank's `ready` words to the neighbours
    // until the rank thread has finished and everything has been forwarded.
    void run_courier(int q) {
        Rank &me = ranks[q];
        std::mt19937 rng(o.seed * 104729u + (unsigned)q);
        uint32_t sent[16] = {0};
        for (;;) {
            const bool last_look = me.finished.load(std::memory_order_acquire);
            for (int i = 0; i < 16; ++i) {
                const int d = i & 7;
                if (me.nb[d] < 0) continue;
                const uint32_t v = me.ready[i].load(std::memory_order_acquire);
                if (v == sent[i]) continue;
                if (o.max_delay_us > 0)   // the system-scope publication can be slow (PCIe traffic)
                    std::this_thread::sleep_for(std::chrono::microseconds(rng() % (unsigned)(o.max_delay_us + 1)));
                Rank &peer = ranks[me.nb[d]];
                (i < 8 ? peer.flag : peer.ack)[b2c::opposite_dir(d)].store(v, std::memory_order_release);
                sent[i] = v;
            }
            if (last_look || failed.load(std::memory_order_relaxed)) return;
            std::this_thread::yield();
        }
    }

    void run_rank(int q, const uint8_t *in, uint8_t *out) {
        run_rank_body(q, in, out);
        ranks[q].finished.store(true, std::memory_order_release);   // the courier takes one last look and leaves
    }

    void run_rank_body(int q, const uint8_t *in, uint8_t *out) {
        Rank &me = ranks[q];
        std::mt19937 rng(o.seed * 7919u + (unsigned)q);
        auto jitter = [&] {
            if (o.max_delay_us > 0)
                std::this_thread::sleep_for(std::chrono::microseconds(rng() % (unsigned)(o.max_delay_us + 1)));
        };
        const int r0 = (q / Q) * R, c0 = (q % Q) * C;
        const int64_t c0b = (int64_t)(q % Q) * bw;

        uint32_t epoch = 0;
        int cur = 0;
        uint32_t seq = 0;
        int conv_check = 0;   // number of the check that found the fixed point
        std::vector<uint8_t> t0, t1;
        me.iters = passes.empty() ? 0 : passes.back().iters_after;
        const int nimg = std::max(1, o.images);
        // allocation: the whole framed buffers are written, so nobody may push before everyone is here
        for (auto &b : me.buf) b.assign((size_t)(pitch * rows), 0);
        arrived.fetch_add(1, std::memory_order_acq_rel);
        if (!wait_for([&] { return arrived.load(std::memory_order_acquire) >= n; })) return;

        for (int img = 0; img < nimg; ++img) {
        // "upload" of image `img` (= the input with every byte XORed by 37*img): interior of the
        // current buffer only, like the engine's 2-D copy.  From the second image on there is NO
        // barrier (ShardedPipeline): neighbours may still be in the last pass of the previous image.
        for (int r = 0; r < R; ++r) {
            uint8_t *d = &me.buf[cur][(r + F) * pitch + fb];
            const uint8_t *s = in + ((int64_t)(r0 + r) * W + c0) * ch;
            for (int64_t x = 0; x < bw; ++x) d[x] = (uint8_t)(s[x] ^ (uint8_t)(37 * img));
        }
        // initial exchange of this image: push, publish, no wait (b2c_engine_exchange)
        jitter();
        push_ring(q, cur);
        publish(q, epoch + 1);
        epoch++;

        for (const Pass &ps : passes) {
            unsigned changed = 0;
            const bool skip = conv_check != 0 && ps.interval > conv_check + 1;
            if (!skip) {
                jitter();
                if (ps.mode & kWaitData)
                    for (int d = 0; d < 8; ++d)
                        if (me.nb[d] >= 0 &&
                            !wait_for([&] { return (int32_t)(me.flag[d].load(std::memory_order_acquire) - epoch) >= 0; }))
                            return;
                if (ps.mode & kWaitAck)
                    for (int d = 0; d < 8; ++d)
                        if (me.nb[d] >= 0 &&
                            !wait_for([&] { return (int32_t)(me.ack[d].load(std::memory_order_acquire) - epoch) >= 0; }))
                            return;
                // The block is extended into the frame by the iterations left in the group, on the
                // sides that have a neighbour (whole 16-byte strips west/east, like the kernel).
                const int xn = me.nb[B2C_N] >= 0 ? ps.ext : 0, xs = me.nb[B2C_S] >= 0 ? ps.ext : 0;
                const int64_t xcol = std::min<int64_t>(fb, b2c::round_up((int64_t)ps.ext * ch, 16));
                const int64_t xw = me.nb[B2C_W] >= 0 ? xcol : 0, xe = me.nb[B2C_E] >= 0 ? xcol : 0;
                // read only what the kernel reads: the extended block plus k rows / k*ch bytes around it
                {
                    const int ra = std::max(0, F - xn - ps.k), rb = std::min(rows, F + R + xs + ps.k);
                    const int64_t ca = std::max<int64_t>(0, fb - xw - (int64_t)ps.k * ch);
                    const int64_t cb = std::min<int64_t>(pitch, fb + bw + xe + (int64_t)ps.k * ch);
                    t0.assign((size_t)(pitch * rows), 0);
                    for (int r = ra; r < rb; ++r)
                        memcpy(&t0[r * pitch + ca], &me.buf[cur][r * pitch + ca], (size_t)(cb - ca));
                }
                b2c::framed_iterations(t0, t1, rows, pitch, ch, F, fb, r0, c0b, H, (int64_t)W * ch, torus, ps.k);
                auto copy_rows = [&](int ra, int rb, int64_t xa, int64_t xb) {   // block-relative, may be negative
                    for (int r = ra; r < rb; ++r) {
                        const int64_t off = (r + F) * pitch + fb;
                        memcpy(&me.buf[cur ^ 1][off + xa], &t0[off + xa], (size_t)(xb - xa));
                        if (ps.check) changed |= memcmp(&t0[off + xa], &t1[off + xa], (size_t)(xb - xa)) != 0;
                    }
                };
                // ring first (top / bottom F rows, left / right fb bytes, plus the extension), then
                // publish, then interior
                const int Ft = std::min(F, R), Fb = std::max(Ft, R - F);
                const int64_t fl = std::min(fb, bw), fr = std::max(fl, bw - fb);
                copy_rows(-xn, Ft, -xw, bw + xe);
                copy_rows(Fb, R + xs, -xw, bw + xe);
                copy_rows(Ft, Fb, -xw, fl);
                copy_rows(Ft, Fb, fr, bw + xe);
                jitter();
                if (ps.mode & kPush) {
                    push_ring(q, cur ^ 1);
                    publish(q, epoch + 1);
                } else if (ps.mode & kPostAck) {
                    publish_ack(q, epoch);
                }
                jitter();
                copy_rows(Ft, Fb, fl, fr);
                cur ^= 1;
                if (ps.group_end) epoch++;
            }
            if (ps.check && !skip) {
                ++seq;
                const uint32_t tag = seq * 2u;
                const int set = o.single_slot_set ? 0 : (int)(seq & 1u) * n;
                for (int k = 0; k < n; ++k)
                    ranks[k].slots[set + q].store(tag | (changed ? 1u : 0u), std::memory_order_release);
                if (q == o.stall_rank && o.stall_ms > 0)
                    std::this_thread::sleep_for(std::chrono::milliseconds(o.stall_ms));
                else
                    jitter();
                unsigned any = 0;
                for (int k = 0; k < n; ++k) {
                    uint32_t v = 0;
                    if (!wait_for([&] { return ((v = me.slots[set + k].load(std::memory_order_acquire)) & ~1u) == tag; }))
                        return;
                    any |= v & 1u;
                }
                if (!any && conv_check == 0) {
                    conv_check = ps.interval;
                    me.iters = ps.iters_after;
                }
            }
        }
        // "download": this rank's block of result `img`
        for (int r = 0; r < R; ++r)
            memcpy(out + (int64_t)img * W * H * ch + ((int64_t)(r0 + r) * W + c0) * ch, &me.buf[cur][(r + F) * pitch + fb],
                   (size_t)bw);
        }   // images
    }
};

}  // namespace

extern "C" int b2c_emulate_halo_threads(const uint8_t *in, uint8_t *out, int W, int H, int ch, int reps, int boundary,
                                        int P, int Q, const b2c_emul_opts *opts, int *iterations_done) {
    if (!in || !out || !opts || W <= 0 || H <= 0 || (ch != 1 && ch != 3) || reps < 0 || P <= 0 || Q <= 0 || H % P ||
        W % Q || P * Q > 64 || (boundary != B2C_ZERO && boundary != B2C_TORUS) ||
        (opts->images > 1 && opts->check_every > 0)) {
        b2c::set_error("emulate_halo_threads: bad arguments");
        return B2C_EINVAL;
    }
    World w;
    w.W = W, w.H = H, w.ch = ch, w.P = P, w.Q = Q, w.R = H / P, w.C = W / Q, w.n = P * Q;
    w.torus = boundary == B2C_TORUS;
    w.o = *opts;
    int K = opts->tb >= 8 ? 8 : (opts->tb >= 4 ? 4 : (opts->tb >= 2 ? 2 : 1));
    while (K > 1 && K > std::min(w.R, w.C)) K >>= 1;  // a k-deep halo comes from the adjacent rank only
    int F = K;
    if (opts->halo_depth > 0) F = std::max(K, std::min({opts->halo_depth, 32, w.R, w.C}) / K * K);
    w.F = F;
    w.bw = (int64_t)w.C * ch, w.fb = (int64_t)w.F * ch, w.pitch = w.bw + 2 * w.fb;
    w.rows = w.R + 2 * w.F;
    w.passes = split_passes(reps, K, F, opts->check_every);
    w.ranks = std::vector<Rank>((size_t)w.n);
    for (int q = 0; q < w.n; ++q) {
        Rank &r = w.ranks[q];
        b2c_neighbours(q, P, Q, boundary, r.nb);
        for (auto &f : r.flag) f.store(0);
        for (auto &f : r.ack) f.store(0);
        for (auto &f : r.ready) f.store(0);
        r.slots.reset(new std::atomic<uint32_t>[2 * (size_t)w.n]);
        for (int k = 0; k < 2 * w.n; ++k) r.slots[k].store(0);
    }
    w.deadline = Clock::now() + std::chrono::milliseconds(opts->timeout_ms > 0 ? opts->timeout_ms : 20000);
    std::vector<std::thread> th;
    for (int q = 0; q < w.n; ++q) th.emplace_back([&w, q, in, out] { w.run_rank(q, in, out); });
    if (opts->courier)
        for (int q = 0; q < w.n; ++q) th.emplace_back([&w, q] { w.run_courier(q); });
    for (auto &t : th) t.join();
    if (w.failed.load()) {
        b2c::set_error("emulate_halo_threads: a rank waited past the deadline (protocol deadlock)");
        return B2C_ETIMEOUT;
    }
    for (int q = 0; q < w.n; ++q)
        if (w.ranks[q].iters != w.ranks[0].iters) {
            b2c::set_error("emulate_halo_threads: ranks disagree on the iteration count");
            return B2C_EINVAL;
        }
    if (iterations_done) *iterations_done = w.ranks[0].iters;
    return B2C_OK;
}
