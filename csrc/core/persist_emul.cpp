// Threaded host model of the persistent dataflow kernel's scheduling (csrc/cuda/stencil_stream.cuh,
// stream_persist_kernel), for one rank without neighbours:
//
//   * the iteration loop is a queue of work items (pass t, chunk c, band b), pass-major and
//     chunk-major inside a pass; worker threads (the GPU's resident warps) pull items from one
//     atomic counter and fetch the next item before they work on the current one;
//   * an item may start when the 3x3 neighbourhood of tiles of pass t-1 is done — those tiles wrote
//     what it reads, and read what it overwrites (two ping-pong buffers) — which it learns from
//     per-tile "passes completed" words (release store / acquire load);
//   * there is no barrier between passes and no limit on how far apart two workers may be.
//
// The model runs the same rule with std::atomic and real memory traffic on two shared buffers, so
// that ThreadSanitizer can see a missing dependency as a data race, a wrong one as a deadlock
// (fewer workers than items per pass, random delays), and the bytes are compared with the oracle.
// Nothing comparable exists in the reference (one kernel launch per iteration from the host,
// cuda/cuda_convolution.cu:66-87).
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

struct Queue {
    int W, H, ch, K, passes, vw, cr, nb, nc;
    int64_t bw;
    std::vector<uint8_t> buf[2];               // tight H x bw images, ping-pong
    std::unique_ptr<std::atomic<uint32_t>[]> done;   // [chunk * nb + band] = passes completed
    std::atomic<uint64_t> counter{0};
    std::atomic<bool> failed{false};
    Clock::time_point deadline;
    int max_delay_us;
    unsigned seed;

    bool wait_done(int b, int c, uint32_t need) {
        const std::atomic<uint32_t> &f = done[(size_t)c * nb + b];
        while ((int32_t)(f.load(std::memory_order_acquire) - need) < 0) {
            if (failed.load(std::memory_order_relaxed)) return false;
            if (Clock::now() > deadline) {
                failed.store(true);
                return false;
            }
            std::this_thread::yield();
        }
        return true;
    }

    // One tile: K iterations on a private copy of what it needs, written back into the other buffer.
    void tile(int t, int b, int c, std::vector<uint8_t> &a0, std::vector<uint8_t> &a1) {
        const std::vector<uint8_t> &src = buf[t & 1];
        std::vector<uint8_t> &dst = buf[(t + 1) & 1];
        const int r0 = c * cr, r1 = c == nc - 1 ? H : (c + 1) * cr;
        const int64_t c0 = (int64_t)b * vw, c1 = std::min<int64_t>(c0 + vw, bw);
        const int F = K;
        const int64_t fb = (int64_t)K * ch;
        const int rows = r1 - r0 + 2 * F;
        const int64_t pitch = c1 - c0 + 2 * fb;
        a0.assign((size_t)(rows * pitch), 0);
        for (int r = 0; r < rows; ++r) {
            const int gr = r0 - F + r;
            if (gr < 0 || gr >= H) continue;
            const int64_t ga = std::max<int64_t>(0, c0 - fb), gb = std::min<int64_t>(bw, c1 + fb);
            memcpy(&a0[r * pitch + (ga - (c0 - fb))], &src[(size_t)gr * bw + ga], (size_t)(gb - ga));
        }
        b2c::framed_iterations(a0, a1, rows, pitch, ch, F, fb, r0, c0, H, bw, false, K);
        for (int r = r0; r < r1; ++r)
            memcpy(&dst[(size_t)r * bw + c0], &a0[(size_t)(r - r0 + F) * pitch + fb], (size_t)(c1 - c0));
    }

    void worker(int id) {
        std::mt19937 rng(seed * 2654435761u + (unsigned)id);
        std::vector<uint8_t> a0, a1;
        const uint64_t per_pass = (uint64_t)nb * nc, total = per_pass * (uint64_t)passes;
        uint64_t g = counter.fetch_add(1, std::memory_order_relaxed);
        while (g < total) {
            const uint64_t gnext = counter.fetch_add(1, std::memory_order_relaxed);   // fetched ahead, like the kernel
            const int t = (int)(g / per_pass), id_in = (int)(g % per_pass);
            const int c = id_in / nb, b = id_in % nb;
            if (max_delay_us > 0 && rng() % 4 == 0)
                std::this_thread::sleep_for(std::chrono::microseconds(rng() % (unsigned)(max_delay_us + 1)));
            if (t > 0)
                for (int dc = -1; dc <= 1; ++dc)
                    for (int db = -1; db <= 1; ++db) {
                        const int pb = std::min(std::max(b + db, 0), nb - 1), pc = std::min(std::max(c + dc, 0), nc - 1);
                        if (!wait_done(pb, pc, (uint32_t)t)) return;
                    }
            tile(t, b, c, a0, a1);
            done[(size_t)c * nb + b].store((uint32_t)t + 1u, std::memory_order_release);
            g = gnext;
        }
    }
};

}  // namespace

// K iterations per pass, `vw` valid bytes per band, `chunk_rows` rows per tile, `workers` threads
// (fewer than the tiles of a pass is the interesting case).  reps must be a multiple of K.
extern "C" int b2c_emulate_persist_threads(const uint8_t *in, uint8_t *out, int W, int H, int ch, int reps, int K,
                                           int vw, int chunk_rows, int workers, int max_delay_us, unsigned seed) {
    if (!in || !out || W <= 0 || H <= 0 || (ch != 1 && ch != 3) || reps < 0 || K <= 0 || reps % K || vw < K * ch ||
        chunk_rows < K || workers <= 0) {
        b2c::set_error("emulate_persist_threads: bad arguments (reps %% K == 0, vw >= K*ch, chunk_rows >= K)");
        return B2C_EINVAL;
    }
    Queue q;
    q.W = W, q.H = H, q.ch = ch, q.K = K, q.passes = reps / K, q.vw = vw, q.cr = chunk_rows;
    q.bw = (int64_t)W * ch;
    q.nb = (int)b2c::ceil_div(q.bw, vw);
    q.nc = std::max(1, H / chunk_rows);          // the last chunk takes the remainder, like the kernel's plan
    q.buf[0].assign(in, in + (size_t)q.bw * H);
    q.buf[1].assign((size_t)q.bw * H, 0);
    q.done.reset(new std::atomic<uint32_t>[(size_t)q.nb * q.nc]);
    for (size_t i = 0; i < (size_t)q.nb * q.nc; ++i) q.done[i].store(0);
    q.max_delay_us = max_delay_us;
    q.seed = seed;
    q.deadline = Clock::now() + std::chrono::seconds(30);
    std::vector<std::thread> th;
    for (int i = 0; i < workers; ++i) th.emplace_back([&q, i] { q.worker(i); });
    for (auto &t : th) t.join();
    if (q.failed.load()) {
        b2c::set_error("emulate_persist_threads: a worker waited past the deadline (dependency deadlock)");
        return B2C_ETIMEOUT;
    }
    memcpy(out, q.buf[q.passes & 1].data(), (size_t)q.bw * H);
    return B2C_OK;
}
