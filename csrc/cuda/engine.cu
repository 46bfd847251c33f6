// Host side of the GPU engine: buffers, plan, peer wiring, pass scheduling.
//
// Reference counterparts (behaviour, not code):
//   GPU_convolution()      cuda/cuda_convolution.cu:49-102  (alloc, H2D, loop, D2H)
//   padded double buffers  mpi/mpi_convolution.c:104-124
//   halo post/wait/swap    mpi/mpi_convolution.c:156-240
// B200 design: two framed, pitched ping-pong buffers per rank; neighbours write
// straight into the frame over NVLink from inside the stencil kernel; the host
// only enqueues one kernel per K fused iterations (optionally as a CUDA graph).
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <cmath>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "common.h"
#include "kernels.h"
#include "stencil_stream.cuh"
#include "stencil_resident.h"
#include "stencil_tma.h"

#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            b2c::set_error("%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return B2C_ECUDA;                                                            \
        }                                                                                \
    } while (0)

using namespace b2c;

namespace {

// NVTX range for Nsight timelines (tracing hook); a no-op without a profiler attached.
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

// Flag pad layout, in 32-bit words.  Every flag/counter sits on its own 128 B line.
constexpr int kPadFlags = 0;          // 8 x 32
constexpr int kPadCounters = 256;     // 8 x 32
constexpr int kPadSlots = 512;        // 2 parities x kMaxRanks words
constexpr int kPadChanged = 640;
constexpr int kPadConverged = 641;
constexpr int kPadConvergedAt = 642;
constexpr int kPadWords = 1024;
static_assert(kPadSlots + 2 * kMaxRanks <= kPadChanged, "convergence slots overlap the next pad field");

struct Blob {  // what b2c_engine_export ships; must fit B2C_HANDLE_BYTES
    cudaIpcMemHandle_t buf[2];
    cudaIpcMemHandle_t pad;
    int rank, device, pid;
    long long alloc_bytes;
};
static_assert(sizeof(Blob) <= B2C_HANDLE_BYTES, "blob too large");

struct Pass {
    int k;
    bool check;
    unsigned iters_after;
};

struct BandPlan {
    int hlo, vw, nbands, nchunks, chunk_rows, nbw, nbe;
};

}  // namespace

struct b2c_engine {
    int device = 0, W = 0, H = 0, CH = 1, P = 1, Q = 1, rank = 0;
    int R = 0, C = 0, row0 = 0, col0 = 0, bw = 0;
    b2c_opts opts{};
    int nb[8];
    bool any_nb = false;
    long long pitch = 0, alloc_bytes = 0;
    uint8_t *alloc[2] = {nullptr, nullptr};
    uint8_t *base[2] = {nullptr, nullptr};
    uint32_t *pad = nullptr;
    // compat buffers (tight, +2 slack rows)
    uint8_t *cbuf[2] = {nullptr, nullptr};
    long long cbytes = 0;
    // peers
    bool connected = false;
    uint8_t *peer_base[2][8];
    uint32_t *peer_pad[8];
    uint32_t *rank_pad[kMaxRanks];
    std::vector<void *> ipc_opened;
    // state
    int cur = 0;
    uint32_t epoch = 0;      // pushes completed so far (same on every rank)
    unsigned check_seq = 0;
    int iters_done = 0;
    bool pending_converge_fixup = false;
    std::vector<Pass> last_passes;
    uint32_t run_epoch0 = 0;
    int run_cur0 = 0;
    // plan
    int kernel = B2C_KERNEL_STREAM, K = 4, SB = 16, F = 1;
    int sm_count = 148;
    long long delay = 0;
    long long launches = 0;
    std::map<std::pair<int, int>, cudaGraphExec_t> graphs;
    NaiveWeights nw{};
    TmaState *tma = nullptr;
    // resident kernel (B2C_KERNEL_RESIDENT): tile plan, exchange images, per-CTA epoch words
    ResidentPlan res_plan{};
    uint8_t *res_xbuf[2] = {nullptr, nullptr};
    long long res_xpitch = 0;
    uint32_t *res_flags = nullptr;   // sm_count x 32 words, then the error word
    uint32_t res_epoch = 0;
    mutable BandPlan plan_cache[9];
    mutable bool plan_ok[9] = {false, false, false, false, false, false, false, false, false};
};

namespace {

// Window geometry of a k-stage pass for the engine's strip width.
void band_geometry(const b2c_engine *e, int k, int *hlo, int *vw, int *nbands) {
    // window start and band pitch: strip-aligned for per-lane LDGs, 16 B for TMA box origins
    const int align = e->kernel == B2C_KERNEL_TMA ? 16 : e->SB;
    *hlo = (int)round_up((long long)k * e->CH, align);
    *vw = (int)round_down(32 * e->SB - *hlo - k * e->CH, align);
    *nbands = (int)ceil_div(e->bw, *vw);
}

// Modelled time of one k-stage pass (microseconds, relative) for `cr` rows per warp, from
// the constants measured on B200 (profiles/tuning.md):
//   * one warp needs ~64 ns per stage per row (SHFL -> add -> SHFL -> add chain), so a
//     pass cannot finish faster than waves x (cr + 2k) x k x 64 ns — what small images see;
//   * the whole chip retires ~7 T raw byte-iterations/s with 16 B strips (a little less with
//     narrower ones: more SHFL per byte), what large images see; several waves are fine
//     there (the block scheduler back-fills), one exact wave is not (no back-fill).
double plan_cost(const b2c_engine *e, int k, int cr) {
    int hlo, vw, nb;
    band_geometry(e, k, &hlo, &vw, &nb);
    int occ = e->kernel == B2C_KERNEL_TMA ? stream_tma_occupancy(k, e->SB, e->CH) : stream_occupancy(k, e->SB, e->CH);
    if (occ <= 0) occ = 4;
    const double wave = (double)e->sm_count * occ * kWarpsPerCta;
    const long long nc = ceil_div(e->R, cr);
    const double warps = (double)nb * nc;
    const double t_lat = std::ceil(warps / wave) * (cr + 2 * k) * k * 0.064;
    const double rate = e->SB == 16 ? 7.2e6 : (e->SB == 8 ? 7.0e6 : 5.5e6);  // raw byte-iterations per us
    const double util = std::min(1.0, warps / wave);       // a half-empty chip runs at half rate
    const double waves_f = std::max(1.0, warps / wave);
    // with about one wave nothing back-fills the SMs that finish early (+30 % measured)
    const double t_thr = warps * 32.0 * e->SB * (cr + 2 * k) * k / (rate * util) * (1.0 + 0.3 / waves_f);
    return std::max(t_lat, t_thr);
}

double best_chunk_rows(const b2c_engine *e, int k, int *out) {
    static const int cand[] = {8, 12, 16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512};
    double best = 1e300;
    int best_cr = std::max(1, e->R);
    for (int c : cand) {
        int cr = std::max(c, std::max(2 * k, e->F));
        cr = std::min(cr, std::max(e->R, 1));
        const double t = plan_cost(e, k, cr);
        if (t < best * 0.999) { best = t; best_cr = cr; }
        if (cr >= e->R) break;
    }
    *out = best_cr;
    return best;
}

BandPlan make_band_plan(const b2c_engine *e, int k) {
    if (k >= 1 && k <= 8 && e->plan_ok[k]) return e->plan_cache[k];
    BandPlan bp{};
    const int CH = e->CH;
    band_geometry(e, k, &bp.hlo, &bp.vw, &bp.nbands);
    int chunk_rows = e->opts.chunk_rows;
    if (chunk_rows <= 0) {
        best_chunk_rows(e, k, &chunk_rows);
        const int nc = (int)ceil_div(e->R, chunk_rows);
        chunk_rows = (int)ceil_div(e->R, nc);   // even out the last chunk
    }
    chunk_rows = std::max(chunk_rows, e->F);
    chunk_rows = std::min(chunk_rows, std::max(e->R, 1));
    chunk_rows = std::max(chunk_rows, 1);
    bp.chunk_rows = chunk_rows;
    bp.nchunks = (int)ceil_div(e->R, chunk_rows);
    const int fb = e->F * CH;
    bp.nbw = std::min(bp.nbands, (int)ceil_div(fb, bp.vw));
    bp.nbe = bp.nbands - (int)(std::max(0, e->bw - fb) / bp.vw);
    bp.nbe = std::max(1, std::min(bp.nbands, bp.nbe));
    if (k >= 1 && k <= 8) {
        e->plan_cache[k] = bp;
        e->plan_ok[k] = true;
    }
    return bp;
}

void fill_pass_params(const b2c_engine *e, int k, bool check, int src_i, PassParams &p) {
    const BandPlan bp = make_band_plan(e, k);
    memset(&p, 0, sizeof p);
    p.src = e->base[src_i];
    p.dst = e->base[src_i ^ 1];
    p.pitch = e->pitch;
    p.rows = e->R;
    p.bw = e->bw;
    p.nbands = bp.nbands;
    p.nchunks = bp.nchunks;
    p.chunk_rows = bp.chunk_rows;
    p.vw = bp.vw;
    p.hlo = bp.hlo;
    const bool zero = e->opts.boundary == B2C_ZERO;
    p.zero_n = zero && e->nb[B2C_N] < 0;
    p.zero_s = zero && e->nb[B2C_S] < 0;
    p.zero_w = zero && e->nb[B2C_W] < 0;
    p.zero_e = zero && e->nb[B2C_E] < 0;
    p.changed = check ? e->pad + kPadChanged : nullptr;
    p.converged = e->opts.check_every > 0 ? e->pad + kPadConverged : nullptr;
    HaloParams &h = p.halo;
    h.nbw = bp.nbw;
    h.nbe = bp.nbe;
    h.enabled = e->any_nb ? 1 : 0;
    if (!h.enabled) return;
    h.push = (e->bw % e->SB == 0) ? 1 : 0;
    h.depth = e->F;
    h.wait_epoch = e->epoch;
    h.post_epoch = e->epoch + 1;
    h.flags = e->pad + kPadFlags;
    h.counters = e->pad + kPadCounters;
    h.delay = e->delay;
    const int exp[8] = {bp.nbands, bp.nbands, bp.nbw * bp.nchunks, bp.nbe * bp.nchunks,
                        bp.nbw, bp.nbe, bp.nbw, bp.nbe};
    for (int d = 0; d < 8; ++d) {
        h.expected[d] = exp[d];
        if (e->nb[d] < 0) continue;
        h.peer_dst[d] = e->peer_base[src_i ^ 1][d];
        // the neighbour sees our data arriving from the opposite direction
        h.peer_flag[d] = e->peer_pad[d] + kPadFlags + opposite_dir(d) * 32;
    }
}

void fill_push_params(const b2c_engine *e, int buf_i, bool wait, PushParams &q) {
    memset(&q, 0, sizeof q);
    q.src = e->base[buf_i];
    q.pitch = e->pitch;
    q.rows = e->R;
    q.bw = e->bw;
    q.ch = e->CH;
    q.depth = e->F;
    q.flags = e->pad + kPadFlags;
    q.counters = e->pad + kPadCounters;
    q.post_epoch = e->epoch + 1;
    q.wait_epoch = e->epoch;
    q.wait_epoch_valid = wait ? 1 : 0;
    q.converged = nullptr;   // only the per-pass fallback push (enqueue_pass) follows the convergence latch
    for (int d = 0; d < 8; ++d) {
        if (e->nb[d] < 0) continue;
        q.peer_dst[d] = e->peer_base[buf_i][d];
        q.peer_flag[d] = e->peer_pad[d] + kPadFlags + opposite_dir(d) * 32;
    }
}

// Splits `reps` iterations into passes of k in {K, ..., 2, 1}; a pass ends on
// every multiple of check_every (counted from the start of this run).
std::vector<Pass> make_passes(const b2c_engine *e, int reps) {
    std::vector<Pass> v;
    const int ce = e->opts.check_every;
    int done = 0;
    while (done < reps) {
        int limit = reps - done;
        if (ce > 0) limit = std::min(limit, ce - done % ce);
        int k = e->K;
        while (k > limit) k >>= 1;
        done += k;
        v.push_back({k, ce > 0 && done % ce == 0, (unsigned)done});
    }
    return v;
}

int enqueue_pass(b2c_engine *e, const Pass &ps, cudaStream_t s) {
    if (e->kernel == B2C_KERNEL_NAIVE) {
        if (e->any_nb) {  // the naive kernel reads the frame without looking at the flags
            unsigned mask = 0;
            for (int d = 0; d < 8; ++d) mask |= (e->nb[d] >= 0 ? 1u : 0u) << d;
            CK(launch_halo_wait(e->pad + kPadFlags, mask, e->epoch, s));
            e->launches++;
        }
        CK(launch_naive(e->base[e->cur], e->base[e->cur ^ 1], e->pitch, e->R, e->bw, e->CH, e->nw, s));
        e->launches++;
        if (e->any_nb) {
            PushParams q;
            fill_push_params(e, e->cur ^ 1, false, q);
            CK(launch_halo_push(q, s));
            e->launches++;
        }
    } else {
        PassParams p;
        fill_pass_params(e, ps.k, ps.check, e->cur, p);
        if (e->kernel == B2C_KERNEL_TMA)
            CK(launch_stream_tma(e->tma, p, e->cur, ps.k, e->SB, e->CH, ps.check, s));
        else
            CK(launch_stream(p, ps.k, e->SB, e->CH, ps.check, s));
        e->launches++;
        if (e->any_nb && !p.halo.push) {
            PushParams q;
            fill_push_params(e, e->cur ^ 1, false, q);
            if (e->opts.check_every > 0) q.converged = e->pad + kPadConverged;
            CK(launch_halo_push(q, s));
            e->launches++;
        }
    }
    e->cur ^= 1;
    if (e->any_nb) e->epoch++;
    if (ps.check) {
        ConvergeParams c{};
        c.changed = e->pad + kPadChanged;
        c.slots = e->pad + kPadSlots;
        const int n = e->P * e->Q;
        for (int k = 0; k < n; ++k) c.peer_slots[k] = (n == 1 ? e->pad : e->rank_pad[k]) + kPadSlots;
        c.nranks = n;
        c.rank = e->rank;
        c.seq = ++e->check_seq;
        c.converged = e->pad + kPadConverged;
        c.converged_at = e->pad + kPadConvergedAt;
        c.iters_after = ps.iters_after;
        CK(launch_converge(c, s));
        e->launches++;
    }
    return B2C_OK;
}

void choose_plan(b2c_engine *e) {
    const b2c_opts &o = e->opts;
    e->kernel = o.kernel == B2C_KERNEL_AUTO ? B2C_KERNEL_STREAM : o.kernel;
    if (o.filter != B2C_GAUSSIAN) e->kernel = B2C_KERNEL_NAIVE;
    int K = o.tb > 0 ? o.tb : 4;
    if (K >= 8) K = 8; else if (K >= 4) K = 4; else if (K >= 2) K = 2; else K = 1;
    if (e->kernel == B2C_KERNEL_NAIVE) K = 1;
    // a k-deep halo must come from the adjacent rank only
    if (e->any_nb) while (K > 1 && K > std::min(e->R, e->C)) K >>= 1;
    e->K = K;
    e->F = e->any_nb ? K : 1;
    if (e->kernel == B2C_KERNEL_NAIVE) e->F = 1;
    int SB = o.strip_bytes;
    if (SB != 4 && SB != 8 && SB != 16) {
        // strip width with the lowest modelled pass time (see plan_cost)
        double best = 1e300;
        SB = 4;
        for (int sb = 16; sb >= 4; sb >>= 1) {
            if (K * (sb / 4) > 16) continue;
            e->SB = sb;
            int cr = 0;
            const double t = best_chunk_rows(e, K, &cr);
            if (t < best * 0.97) { best = t; SB = sb; }   // prefer the wider strip on near ties
        }
    }
    while (K * (SB / 4) > 16) SB >>= 1;
    e->SB = SB;
    e->F = e->any_nb ? K : 1;
    if (e->kernel == B2C_KERNEL_NAIVE) e->F = 1;
    // the byte-per-thread kernel has no change detector: a check would always report "unchanged"
    if (e->kernel == B2C_KERNEL_NAIVE || e->kernel == B2C_KERNEL_RESIDENT) e->opts.check_every = 0;
}

// Rows per warp for large blocks are measured, not modelled: how the block scheduler packs
// 4-6 waves of CTAs is not captured by plan_cost (160-row chunks beat 128-row ones by 4 % on
// 16384² rgb, 112 beat 128 on grey).  A handful of candidates is timed on the engine's own (still
// all-zero) buffers with the halo logic off: the kernel's time does not depend on pixel values,
// zeros blur to zeros, and no frame or flag is touched.  ~10 ms once per engine; set
// B200CONV_AUTOTUNE=0 or give chunk_rows explicitly to skip it.
int autotune_chunk_rows(b2c_engine *e) {
    static const bool off = getenv("B200CONV_AUTOTUNE") && atoi(getenv("B200CONV_AUTOTUNE")) == 0;
    if (off || e->opts.chunk_rows > 0) return B2C_OK;
    if (e->kernel != B2C_KERNEL_STREAM && e->kernel != B2C_KERNEL_TMA) return B2C_OK;
    // Multi-rank blocks keep the modelled plan: what is timed here runs without neighbours, and
    // measured on 4 and 8 GPUs the halo-free optimum (longer chunks) is up to 20 % slower once
    // ring warps have to finish their chunk before a neighbour's next pass may start.
    if (e->P * e->Q > 1) return B2C_OK;
    const double bytes = (double)e->bw * e->R;
    if (bytes < 1e6) return B2C_OK;   // tiny images: a pass is shorter than the launch path, keep the model
    static const int cand[] = {8, 12, 16, 24, 32, 48, 64, 80, 96, 112, 128, 144, 160, 192};
    const bool any_nb = e->any_nb;
    e->any_nb = false;
    cudaStream_t s = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int rc = B2C_OK, best = 0;
    float best_ms = 1e30f;
    if (cudaStreamCreate(&s) != cudaSuccess || cudaEventCreate(&ev0) != cudaSuccess ||
        cudaEventCreate(&ev1) != cudaSuccess) {
        set_error("autotune: %s", cudaGetErrorString(cudaGetLastError()));
        rc = B2C_ECUDA;
    }
    const int sb0 = e->SB;
    int best_sb = sb0;
    for (int sb = 16; sb >= (bytes < 64e6 ? 4 : 8) && rc == B2C_OK; sb >>= 1) {
        // also try the other strip width unless the caller fixed it (K*words <= 16 registers budget)
        if (sb != sb0 && (e->opts.strip_bytes > 0 || e->K * (sb / 4) > 16)) continue;
        e->SB = sb;
        for (int c : cand) {
            if (c > e->R) break;
            if (c < 2 * e->K) continue;                                   // pipeline fill would dominate
            if (c < 64 && bytes >= 256e6) continue;                       // pointless on large blocks
            if (c > 48 && bytes < 16e6) continue;                         // and long chunks starve small ones
            e->opts.chunk_rows = c;
            for (bool &ok : e->plan_ok) ok = false;
            PassParams p;
            fill_pass_params(e, e->K, false, 0, p);
            cudaError_t ce = cudaSuccess;
            const int reps = bytes < 16e6 ? 41 : (bytes < 128e6 ? 9 : 3);   // short passes: average more of them
            for (int it = 0; it < reps && ce == cudaSuccess; ++it) {
                if (it == 1) cudaEventRecord(ev0, s);
                ce = e->kernel == B2C_KERNEL_TMA ? launch_stream_tma(e->tma, p, 0, e->K, e->SB, e->CH, false, s)
                                                 : launch_stream(p, e->K, e->SB, e->CH, false, s);
            }
            cudaEventRecord(ev1, s);
            if (ce != cudaSuccess || cudaEventSynchronize(ev1) != cudaSuccess) {
                set_error("autotune: %s", cudaGetErrorString(cudaGetLastError()));
                rc = B2C_ECUDA;
                break;
            }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, ev0, ev1);
            if (ms < best_ms) { best_ms = ms; best = c; best_sb = sb; }
        }
    }
    e->SB = best_sb;
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    if (s) cudaStreamDestroy(s);
    e->any_nb = any_nb;
    e->opts.chunk_rows = best;   // 0 (model) when nothing was timed
    for (bool &ok : e->plan_ok) ok = false;
    return rc;
}

int alloc_buffers(b2c_engine *e) {
    CK(cudaSetDevice(e->device));
    if (e->opts.boundary == B2C_CUDA_COMPAT) {
        e->cbytes = (long long)e->W * e->CH * (e->H + 2);
        for (auto &b : e->cbuf) {
            CK(cudaMalloc(&b, (size_t)e->cbytes + 64));
            CK(cudaMemset(b, 0, (size_t)e->cbytes + 64));
        }
    } else {
        e->pitch = round_up(kLeftPad + (long long)e->bw + 32 + 16, 128);
        e->alloc_bytes = e->pitch * (e->R + 2 * kFrameRows) + 1024;  // tail slack: last band's window
        for (int i = 0; i < 2; ++i) {
            CK(cudaMalloc(&e->alloc[i], (size_t)e->alloc_bytes));
            CK(cudaMemset(e->alloc[i], 0, (size_t)e->alloc_bytes));
            e->base[i] = e->alloc[i] + (long long)kFrameRows * e->pitch + kLeftPad;
        }
    }
    CK(cudaMalloc(&e->pad, kPadWords * sizeof(uint32_t)));
    CK(cudaMemset(e->pad, 0, kPadWords * sizeof(uint32_t)));
    CK(cudaDeviceSynchronize());
    return B2C_OK;
}

}  // namespace

extern "C" {

void b2c_default_opts(b2c_opts *o) {
    memset(o, 0, sizeof *o);
    o->boundary = B2C_ZERO;
    o->filter = B2C_GAUSSIAN;
    o->kernel = B2C_KERNEL_AUTO;
    o->use_graph = -1;
}

b2c_engine *b2c_engine_create(int device, int W, int H, int CH, int P, int Q, int rank, const b2c_opts *opts) {
    b2c_opts o;
    if (opts) o = *opts; else b2c_default_opts(&o);
    if (W <= 0 || H <= 0 || (CH != 1 && CH != 3) || P <= 0 || Q <= 0 || rank < 0 || rank >= P * Q ||
        H % P || W % Q || P * Q > kMaxRanks) {
        set_error("engine_create: bad geometry W=%d H=%d CH=%d grid=%dx%d rank=%d", W, H, CH, P, Q, rank);
        return nullptr;
    }
    if (o.boundary == B2C_CUDA_COMPAT && (P * Q != 1 || o.filter != B2C_GAUSSIAN)) {
        set_error("cuda-compat is single-GPU, gaussian only (like cuda/cuda_convolution.cu)");
        return nullptr;
    }
    if ((long long)W / Q * CH > (1ll << 30)) {
        set_error("block row too wide");
        return nullptr;
    }
    b2c_engine *e = new b2c_engine();
    e->device = device; e->W = W; e->H = H; e->CH = CH; e->P = P; e->Q = Q; e->rank = rank;
    e->opts = o;
    e->R = H / P; e->C = W / Q;
    e->row0 = (rank / Q) * e->R;   // mpi_convolution.c:85-86
    e->col0 = (rank % Q) * e->C;
    e->bw = e->C * CH;
    b2c_neighbours(rank, P, Q, o.boundary, e->nb);
    e->any_nb = false;
    for (int d = 0; d < 8; ++d) e->any_nb |= e->nb[d] >= 0;
    memset(e->peer_base, 0, sizeof e->peer_base);
    memset(e->peer_pad, 0, sizeof e->peer_pad);
    memset(e->rank_pad, 0, sizeof e->rank_pad);
    const FilterDef &f = filter_def(o.filter);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            e->nw.iw[i][j] = f.w[i][j];
            e->nw.fw[i][j] = (float)(f.w[i][j] / (double)f.den);
        }
    e->nw.integer16 = f.den == 16;
    int smc = 148;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaDeviceGetAttribute(&smc, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) {
        set_error("cudaSetDevice(%d) failed: %s", device, cudaGetErrorString(cudaGetLastError()));
        delete e;
        return nullptr;
    }
    e->sm_count = smc;
    choose_plan(e);
    if (alloc_buffers(e) != B2C_OK) {
        b2c_engine_destroy(e);
        return nullptr;
    }
    if (e->kernel == B2C_KERNEL_RESIDENT) {
        int coop = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
        // Exchange interval K (= halo depth): an explicit tb is rounded down to 4, 8 or 16 and falls
        // back to a shorter interval when its halo does not fit; otherwise the K with the lowest
        // modelled time wins.  Measured on B200 (profiles/tuning.md step 14): an iteration costs
        // ~66.5 ns per row held by a warp, an exchange ~4 us whatever K is.
        bool fits = false;
        if (o.tb > 0) {
            for (int k = o.tb >= 16 ? 16 : (o.tb >= 8 ? 8 : 4); k >= 4 && !fits; k >>= 1)
                fits = resident_plan(e->bw, e->R, CH, k, e->sm_count, &e->res_plan);
        } else {
            double best = 1e300;
            for (int k = 16; k >= 4; k >>= 1) {
                ResidentPlan pl;
                if (!resident_plan(e->bw, e->R, CH, k, e->sm_count, &pl)) continue;
                const double us_per_iteration = pl.rw * 0.0665 + 4.0 / k;
                if (us_per_iteration < best) { best = us_per_iteration; e->res_plan = pl; fits = true; }
            }
        }
        if (P * Q != 1 || o.boundary == B2C_CUDA_COMPAT || !coop || !fits) {
            set_error("the resident kernel needs one GPU, zero/torus, row bytes %% 8 == 0 and an image that fits "
                      "the register files (%d x %d bytes does not, or cooperative launch is missing)", e->bw, e->R);
            b2c_engine_destroy(e);
            return nullptr;
        }
        e->res_xpitch = round_up(e->bw, 128);
        const size_t xbytes = (size_t)(e->res_xpitch * e->R), fbytes = ((size_t)e->sm_count * 32 + 32) * sizeof(uint32_t);
        if (cudaMalloc(&e->res_xbuf[0], xbytes) != cudaSuccess || cudaMalloc(&e->res_xbuf[1], xbytes) != cudaSuccess ||
            cudaMalloc(&e->res_flags, fbytes) != cudaSuccess || cudaMemset(e->res_flags, 0, fbytes) != cudaSuccess) {
            set_error("resident kernel buffers: %s", cudaGetErrorString(cudaGetLastError()));
            b2c_engine_destroy(e);
            return nullptr;
        }
    }
    if (e->kernel == B2C_KERNEL_TMA && o.boundary != B2C_CUDA_COMPAT) {
        e->tma = tma_create(e->alloc, e->pitch, e->R + 2 * kFrameRows, e->alloc_bytes);
        if (!e->tma) {
            b2c_engine_destroy(e);
            return nullptr;
        }
    }
    if (o.boundary != B2C_CUDA_COMPAT && autotune_chunk_rows(e) != B2C_OK) {
        b2c_engine_destroy(e);
        return nullptr;
    }
    // A torus rank that is its own neighbour in every direction needs no peers.
    bool self_only = e->any_nb;
    for (int d = 0; d < 8; ++d) self_only &= (e->nb[d] < 0 || e->nb[d] == rank);
    if (self_only && e->any_nb) {
        b2c_engine *self = e;
        b2c_engine_connect_local(&self, 1);
    }
    return e;
}

void b2c_engine_destroy(b2c_engine *e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();
    for (auto &g : e->graphs) cudaGraphExecDestroy(g.second);
    for (void *p : e->ipc_opened) cudaIpcCloseMemHandle(p);
    if (e->tma) tma_destroy(e->tma);
    for (auto &b : e->res_xbuf) if (b) cudaFree(b);
    if (e->res_flags) cudaFree(e->res_flags);
    for (auto &b : e->alloc) if (b) cudaFree(b);
    for (auto &b : e->cbuf) if (b) cudaFree(b);
    if (e->pad) cudaFree(e->pad);
    delete e;
}

int b2c_engine_block(const b2c_engine *e, int *row0, int *col0, int *rows, int *cols) {
    if (!e) return B2C_EINVAL;
    if (row0) *row0 = e->row0;
    if (col0) *col0 = e->col0;
    if (rows) *rows = e->R;
    if (cols) *cols = e->C;
    return B2C_OK;
}

int b2c_engine_export(b2c_engine *e, void *out) {
    if (!e || !out || e->opts.boundary == B2C_CUDA_COMPAT) return B2C_EINVAL;
    CK(cudaSetDevice(e->device));
    Blob b;
    memset(&b, 0, sizeof b);
    CK(cudaIpcGetMemHandle(&b.buf[0], e->alloc[0]));
    CK(cudaIpcGetMemHandle(&b.buf[1], e->alloc[1]));
    CK(cudaIpcGetMemHandle(&b.pad, e->pad));
    b.rank = e->rank;
    b.device = e->device;
    b.pid = 0;
    b.alloc_bytes = e->alloc_bytes;
    memset(out, 0, B2C_HANDLE_BYTES);
    memcpy(out, &b, sizeof b);
    return B2C_OK;
}

int b2c_engine_import(b2c_engine *e, const void *blobs, int n) {
    if (!e || !blobs || n != e->P * e->Q) {
        set_error("engine_import: need %d blobs", e ? e->P * e->Q : 0);
        return B2C_EINVAL;
    }
    if (e->connected) {
        set_error("engine_import: already connected (b2c_engine_disconnect first)");
        return B2C_EINVAL;
    }
    CK(cudaSetDevice(e->device));
    std::vector<uint8_t *> rbuf0((size_t)n, nullptr), rbuf1((size_t)n, nullptr);
    const long long org = (long long)kFrameRows * e->pitch + kLeftPad;
    for (int k = 0; k < n; ++k) {
        Blob b;
        memcpy(&b, (const uint8_t *)blobs + (size_t)k * B2C_HANDLE_BYTES, sizeof b);
        if (b.rank != k || b.alloc_bytes != e->alloc_bytes) {
            set_error("engine_import: blob %d does not match (rank %d, bytes %lld)", k, b.rank, b.alloc_bytes);
            return B2C_EINVAL;
        }
        if (k == e->rank) {
            e->rank_pad[k] = e->pad;
            rbuf0[k] = e->alloc[0];
            rbuf1[k] = e->alloc[1];
            continue;
        }
        void *p = nullptr;
        CK(cudaIpcOpenMemHandle(&p, b.pad, cudaIpcMemLazyEnablePeerAccess));
        e->ipc_opened.push_back(p);
        e->rank_pad[k] = (uint32_t *)p;
        bool is_nb = false;
        for (int d = 0; d < 8; ++d) is_nb |= e->nb[d] == k;
        if (!is_nb) continue;
        CK(cudaIpcOpenMemHandle(&p, b.buf[0], cudaIpcMemLazyEnablePeerAccess));
        e->ipc_opened.push_back(p);
        rbuf0[k] = (uint8_t *)p;
        CK(cudaIpcOpenMemHandle(&p, b.buf[1], cudaIpcMemLazyEnablePeerAccess));
        e->ipc_opened.push_back(p);
        rbuf1[k] = (uint8_t *)p;
    }
    for (int d = 0; d < 8; ++d) {
        const int k = e->nb[d];
        if (k < 0) continue;
        e->peer_base[0][d] = rbuf0[k] + org;
        e->peer_base[1][d] = rbuf1[k] + org;
        e->peer_pad[d] = e->rank_pad[k];
    }
    e->connected = true;
    return B2C_OK;
}

int b2c_engine_disconnect(b2c_engine *e) {
    if (!e) return B2C_EINVAL;
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    for (void *p : e->ipc_opened) cudaIpcCloseMemHandle(p);
    e->ipc_opened.clear();
    memset(e->peer_base, 0, sizeof e->peer_base);
    memset(e->peer_pad, 0, sizeof e->peer_pad);
    memset(e->rank_pad, 0, sizeof e->rank_pad);
    e->connected = false;
    return B2C_OK;
}

int b2c_engine_connect_local(b2c_engine **engines, int n) {
    if (!engines || n <= 0 || !engines[0] || n != engines[0]->P * engines[0]->Q) {
        set_error("connect_local: need one engine per rank");
        return B2C_EINVAL;
    }
    for (int k = 0; k < n; ++k)
        if (engines[k]->rank != k || engines[k]->pitch != engines[0]->pitch) {
            set_error("connect_local: engines must be in rank order with identical geometry");
            return B2C_EINVAL;
        }
    for (int a = 0; a < n; ++a)
        for (int b = 0; b < n; ++b) {
            if (engines[a]->device == engines[b]->device) continue;
            CK(cudaSetDevice(engines[a]->device));
            cudaError_t er = cudaDeviceEnablePeerAccess(engines[b]->device, 0);
            if (er != cudaSuccess && er != cudaErrorPeerAccessAlreadyEnabled) CK(er);
            cudaGetLastError();
        }
    for (int a = 0; a < n; ++a) {
        b2c_engine *e = engines[a];
        for (int k = 0; k < n; ++k) e->rank_pad[k] = engines[k]->pad;
        for (int d = 0; d < 8; ++d) {
            const int k = e->nb[d];
            if (k < 0) continue;
            e->peer_base[0][d] = engines[k]->base[0];
            e->peer_base[1][d] = engines[k]->base[1];
            e->peer_pad[d] = engines[k]->pad;
        }
        e->connected = true;
    }
    return B2C_OK;
}

// After a run with check_every the host's buffer index and epoch are only right once the
// device-side latch has been read back.  Every entry point that depends on them settles first, so
// a caller that skips b2c_engine_iterations_done still reads the right buffer.
static int settle(b2c_engine *e) {
    if (e->pending_converge_fixup && b2c_engine_iterations_done(e) < 0) return B2C_ECUDA;
    return B2C_OK;
}

static int copy_block(b2c_engine *e, void *user, long long upitch, cudaMemcpyKind kind, bool to_engine, void *stream) {
    if (!e || !user) return B2C_EINVAL;
    if (settle(e) != B2C_OK) return B2C_ECUDA;
    NvtxRange nvtx(to_engine ? "b2c_engine_load" : "b2c_engine_store");
    CK(cudaSetDevice(e->device));
    cudaStream_t s = (cudaStream_t)stream;
    if (e->opts.boundary == B2C_CUDA_COMPAT) {
        const long long row = (long long)e->W * e->CH;
        if (upitch <= 0) upitch = row;
        if (to_engine) {
            CK(cudaMemsetAsync(e->cbuf[0], 0, (size_t)e->cbytes, s));
            CK(cudaMemsetAsync(e->cbuf[1], 0, (size_t)e->cbytes, s));  // cudaMemset(d_dst, 0), :61
            CK(cudaMemcpy2DAsync(e->cbuf[0], (size_t)row, user, (size_t)upitch, (size_t)row, (size_t)e->H, kind, s));
        } else {
            CK(cudaMemcpy2DAsync(user, (size_t)upitch, e->cbuf[0], (size_t)row, (size_t)row, (size_t)e->H, kind, s));
        }
        return B2C_OK;
    }
    if (upitch <= 0) upitch = e->bw;
    if (to_engine)
        CK(cudaMemcpy2DAsync(e->base[e->cur], (size_t)e->pitch, user, (size_t)upitch, (size_t)e->bw, (size_t)e->R, kind, s));
    else
        CK(cudaMemcpy2DAsync(user, (size_t)upitch, e->base[e->cur], (size_t)e->pitch, (size_t)e->bw, (size_t)e->R, kind, s));
    return B2C_OK;
}

int b2c_engine_upload(b2c_engine *e, const uint8_t *host, int64_t pitch, void *stream) {
    return copy_block(e, (void *)host, pitch, cudaMemcpyHostToDevice, true, stream);
}
int b2c_engine_download(b2c_engine *e, uint8_t *host, int64_t pitch, void *stream) {
    return copy_block(e, host, pitch, cudaMemcpyDeviceToHost, false, stream);
}
int b2c_engine_load_device(b2c_engine *e, const uint8_t *dev, int64_t pitch, void *stream) {
    return copy_block(e, (void *)dev, pitch, cudaMemcpyDeviceToDevice, true, stream);
}
int b2c_engine_store_device(b2c_engine *e, uint8_t *dev, int64_t pitch, void *stream) {
    return copy_block(e, dev, pitch, cudaMemcpyDeviceToDevice, false, stream);
}

int b2c_engine_exchange(b2c_engine *e, void *stream) {
    if (!e) return B2C_EINVAL;
    NvtxRange nvtx("b2c_engine_exchange");
    if (!e->any_nb) return B2C_OK;
    if (!e->connected) {
        set_error("engine_exchange: peers are not connected (import/connect_local first)");
        return B2C_EINVAL;
    }
    if (settle(e) != B2C_OK) return B2C_ECUDA;
    CK(cudaSetDevice(e->device));
    PushParams q;
    fill_push_params(e, e->cur, false, q);
    CK(launch_halo_push(q, (cudaStream_t)stream));
    e->launches++;
    e->epoch++;
    return B2C_OK;
}

int b2c_engine_run(b2c_engine *e, int reps, void *stream) {
    if (!e || reps < 0) return B2C_EINVAL;
    NvtxRange nvtx("b2c_engine_run");
    if (settle(e) != B2C_OK) return B2C_ECUDA;
    CK(cudaSetDevice(e->device));
    cudaStream_t s = (cudaStream_t)stream;
    if (e->opts.boundary == B2C_CUDA_COMPAT) {
        const int eff = 2 * (reps / 2);  // wrong-buffer copy-back of cuda_convolution.cu:93-97
        for (int t = 0; t < eff; ++t) {
            CK(launch_compat(e->cbuf[t & 1], e->cbuf[(t & 1) ^ 1], e->W, e->H, e->CH, s));
            e->launches++;
        }
        e->iters_done = eff;
        return B2C_OK;
    }
    if (e->any_nb && !e->connected) {
        set_error("engine_run: peers are not connected");
        return B2C_EINVAL;
    }
    if (e->any_nb && e->epoch == 0) {
        set_error("engine_run: call b2c_engine_exchange after upload on every rank first");
        return B2C_EINVAL;
    }
    if (e->kernel == B2C_KERNEL_RESIDENT) {
        // one cooperative launch for the whole run; CTAs trade rings through res_xbuf every 4 iterations
        e->iters_done = reps;
        if (reps == 0) return B2C_OK;
        ResidentParams q{};
        q.src = e->base[e->cur];
        q.dst = e->base[e->cur ^ 1];
        q.src_pitch = q.dst_pitch = e->pitch;
        q.xbuf[0] = e->res_xbuf[0];
        q.xbuf[1] = e->res_xbuf[1];
        q.xpitch = e->res_xpitch;
        q.flags = e->res_flags;
        q.error = e->res_flags + (size_t)e->sm_count * 32;
        q.epoch0 = e->res_epoch;
        q.H = e->R;
        q.bw = e->bw;
        q.reps = reps;
        q.strips = e->res_plan.strips;
        q.ctarows = e->res_plan.ctarows;
        q.torus = e->opts.boundary == B2C_TORUS;
        {
            const char *dbg = getenv("B200CONV_RES_DEBUG");   // profiling only, see stencil_resident.cu
            q.debug = dbg ? atoi(dbg) : 0;
        }
        CK(launch_resident(e->res_plan, q, e->CH, s));
        e->launches++;
        e->res_epoch += (uint32_t)(ceil_div(reps, e->res_plan.k) - 1);
        e->cur ^= 1;
        return B2C_OK;
    }
    const std::vector<Pass> passes = make_passes(e, reps);
    e->iters_done = reps;
    if (e->opts.check_every > 0) {
        CK(cudaMemsetAsync(e->pad + kPadConverged, 0, 2 * sizeof(uint32_t), s));
        e->last_passes = passes;
        e->run_epoch0 = e->epoch;
        e->run_cur0 = e->cur;
        e->pending_converge_fixup = true;
    }
    bool graph = e->opts.use_graph == 1 || (e->opts.use_graph < 0 && passes.size() >= 8);
    if (e->any_nb || e->opts.check_every > 0 || s == nullptr) graph = false;  // epochs are baked into params
    if (graph) {  // a caller that is capturing this stream itself gets plain launches into its graph
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(s, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) {
            cudaGetLastError();
            graph = false;
        }
    }
    if (graph) {
        const auto key = std::make_pair(reps, e->cur);
        auto it = e->graphs.find(key);
        if (it == e->graphs.end()) {
            if (e->graphs.size() >= 16) {  // callers that vary `reps` endlessly must not hoard graphs
                for (auto &g : e->graphs) cudaGraphExecDestroy(g.second);
                e->graphs.clear();
            }
            const int cur0 = e->cur;
            const long long l0 = e->launches, g0 = total_launches();
            cudaGraph_t g;
            CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
            int rc = B2C_OK;
            for (const Pass &ps : passes)
                if ((rc = enqueue_pass(e, ps, s)) != B2C_OK) break;
            cudaError_t ce = cudaStreamEndCapture(s, &g);
            e->cur = cur0;
            e->launches = l0;
            count_launch(g0 - total_launches());  // captured, not launched
            if (rc != B2C_OK) {
                if (ce == cudaSuccess) cudaGraphDestroy(g);
                return rc;
            }
            CK(ce);
            cudaGraphExec_t ge;
            ce = cudaGraphInstantiate(&ge, g, 0);
            cudaGraphDestroy(g);
            CK(ce);
            it = e->graphs.emplace(key, ge).first;
        }
        CK(cudaGraphLaunch(it->second, s));
        e->launches += (long long)passes.size();
        count_launch((long long)passes.size());
        e->cur ^= (int)(passes.size() & 1);
        return B2C_OK;
    }
    for (const Pass &ps : passes) {
        int rc = enqueue_pass(e, ps, s);
        if (rc != B2C_OK) return rc;
    }
    return B2C_OK;
}

int b2c_engine_iterations_done(b2c_engine *e) {
    if (!e) return B2C_EINVAL;
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    if (e->res_flags) {
        uint32_t err = 0;
        CK(cudaMemcpy(&err, e->res_flags + (size_t)e->sm_count * 32, sizeof err, cudaMemcpyDeviceToHost));
        if (err) {
            set_error("resident kernel: a CTA waited too long for a neighbour's epoch");
            return B2C_ECUDA;
        }
    }
    if (e->pending_converge_fixup) {
        e->pending_converge_fixup = false;
        uint32_t v[2] = {0, 0};
        CK(cudaMemcpy(v, e->pad + kPadConverged, sizeof v, cudaMemcpyDeviceToHost));
        if (v[0]) {
            // Passes after the latch were skipped on every rank: rewind the host-side
            // buffer index and epoch to the state right after the converging pass.
            int cur = e->run_cur0;
            uint32_t ep = e->run_epoch0;
            for (const Pass &ps : e->last_passes) {
                cur ^= 1;
                if (e->any_nb) ep++;
                if (ps.iters_after >= v[1]) break;
            }
            e->cur = cur;
            e->epoch = ep;
            e->iters_done = (int)v[1];
        }
    }
    return e->iters_done;
}

int64_t b2c_engine_launches(const b2c_engine *e) { return e ? e->launches : 0; }

int b2c_engine_describe(const b2c_engine *e, char *buf, size_t buflen) {
    if (!e || !buf) return B2C_EINVAL;
    const char *kn = e->kernel == B2C_KERNEL_NAIVE ? "naive" : (e->kernel == B2C_KERNEL_TMA ? "tma" : "stream");
    if (e->opts.boundary == B2C_CUDA_COMPAT) {
        snprintf(buf, buflen, "kernel=compat W=%d H=%d CH=%d", e->W, e->H, e->CH);
        return B2C_OK;
    }
    if (e->kernel == B2C_KERNEL_RESIDENT) {
        snprintf(buf, buflen, "kernel=resident tb=%d ctas=%dx%d rows_per_warp=%d valid=%dx%dB block=%dx%d(bytes %d)",
                 e->res_plan.k, e->res_plan.ctarows, e->res_plan.strips, e->res_plan.rw, e->res_plan.vr, e->res_plan.vc, e->R, e->C,
                 e->bw);
        return B2C_OK;
    }
    const BandPlan bp = make_band_plan(e, e->K);
    snprintf(buf, buflen,
             "kernel=%s tb=%d strip=%dB block=%dx%d(bytes %d) pitch=%lld bands=%d(valid %dB) chunks=%d(x%d rows) "
             "warps=%d halo_depth=%d fused_push=%d grid=%dx%d rank=%d",
             kn, e->K, e->SB, e->R, e->C, e->bw, e->pitch, bp.nbands, bp.vw, bp.nchunks, bp.chunk_rows,
             bp.nbands * bp.nchunks, e->any_nb ? e->F : 0, (e->any_nb && e->bw % e->SB == 0) ? 1 : 0, e->P, e->Q,
             e->rank);
    return B2C_OK;
}

// Test hook (host only, no CUDA call that matters): the band/chunk plan the engine would use for one
// rank's block, so that CPU property tests can check the invariants the kernels rely on.
// out = {hlo, vw, nbands, nchunks, chunk_rows, nbw, nbe, K, SB, F}.
int b2c_debug_band_plan(int rows, int cols, int channels, int has_neighbours, int tb, int strip_bytes, int chunk_rows,
                        int kernel, int sm_count, int out[10]) {
    if (rows <= 0 || cols <= 0 || (channels != 1 && channels != 3) || !out) return B2C_EINVAL;
    if ((long long)cols * channels > (1ll << 30)) return B2C_ETOOBIG;
    b2c_engine e;
    b2c_default_opts(&e.opts);
    e.opts.tb = tb;
    e.opts.strip_bytes = strip_bytes;
    e.opts.chunk_rows = chunk_rows;
    e.opts.kernel = kernel;
    e.R = rows, e.C = cols, e.CH = channels, e.bw = cols * channels;
    e.W = cols, e.H = rows;
    e.any_nb = has_neighbours != 0;
    for (int &n : e.nb) n = has_neighbours ? 0 : -1;
    e.sm_count = sm_count > 0 ? sm_count : 148;
    choose_plan(&e);
    const BandPlan bp = make_band_plan(&e, e.K);
    const int v[10] = {bp.hlo, bp.vw, bp.nbands, bp.nchunks, bp.chunk_rows, bp.nbw, bp.nbe, e.K, e.SB, e.F};
    memcpy(out, v, sizeof v);
    return B2C_OK;
}

int b2c_engine_set_delay(b2c_engine *e, int64_t cycles) {
    if (!e) return B2C_EINVAL;
    e->delay = cycles;
    return B2C_OK;
}

int b2c_blur_device(const uint8_t *d_in, int64_t in_pitch, uint8_t *d_out, int64_t out_pitch, int W, int H,
                    int CH, int reps, const b2c_opts *opts, void *stream) {
    int dev = 0;
    CK(cudaGetDevice(&dev));
    b2c_engine *e = b2c_engine_create(dev, W, H, CH, 1, 1, 0, opts);
    if (!e) return B2C_ECUDA;
    int rc = b2c_engine_load_device(e, d_in, in_pitch, stream);
    if (rc == B2C_OK) rc = b2c_engine_exchange(e, stream);
    if (rc == B2C_OK) rc = b2c_engine_run(e, reps, stream);
    if (rc == B2C_OK && e->opts.check_every > 0) rc = b2c_engine_iterations_done(e) < 0 ? B2C_ECUDA : B2C_OK;
    if (rc == B2C_OK) rc = b2c_engine_store_device(e, d_out, out_pitch, stream);
    if (rc == B2C_OK && cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) rc = B2C_ECUDA;
    b2c_engine_destroy(e);
    return rc;
}

int b2c_stencil_step_device(const uint8_t *src, uint8_t *dst, int64_t pitch, int rows, int row_bytes, int ch,
                            int filter, void *stream) {
    if (!src || !dst || rows <= 0 || row_bytes <= 0 || (ch != 1 && ch != 3)) return B2C_EINVAL;
    NaiveWeights w{};
    const FilterDef &f = filter_def(filter);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            w.iw[i][j] = f.w[i][j];
            w.fw[i][j] = (float)(f.w[i][j] / (double)f.den);
        }
    w.integer16 = f.den == 16;
    CK(launch_naive(src, dst, pitch, rows, row_bytes, ch, w, (cudaStream_t)stream));
    return B2C_OK;
}

int64_t b2c_total_launches(void) { return total_launches(); }

}  // extern "C"

// ---------------------------------------------------------------------------------------
// Allocation-free blur (b2c_blur_device_noalloc).  No counterpart in the reference.
//
// b2c_blur_device needs two framed ping-pong buffers plus a flag pad and gets them from
// cudaMalloc on every call.  Some users of this library are not allowed to do that: a
// filter plugged into a host application that hands out all device memory up front and
// forbids further allocation; a per-frame / real-time loop where cudaMalloc and cudaFree
// stall the whole device for an unbounded time; a process that runs under a fixed device
// memory budget and must not fail half-way through because an allocation was refused.
// For them the same work runs out of a fixed scratch area that is part of the library
// image itself: the array below.  Every GPU that loads the library carries these 32 MiB
// until the process exits, whether or not the function is ever called.
//
// The scratch is shared by all calls and keeps whatever the previous call left in it, so
// each call first clears the part it is going to use (halo frames and flag words must
// start at zero).  Two calls whose GPU work overlaps — different streams, or different
// host threads on different streams — corrupt each other; calls on one stream are safe.
// ---------------------------------------------------------------------------------------
#define B2C_NOALLOC_SCRATCH_BYTES (32ll << 20)

static __device__ __align__(256) unsigned char g_noalloc_scratch[B2C_NOALLOC_SCRATCH_BYTES];

namespace {

// In host code `g_noalloc_scratch` is only a stand-in for the device symbol; the usable device
// address comes from cudaGetSymbolAddress, looked up once per device with that device current.
int noalloc_scratch_base(int dev, uint8_t **out) {
    static uint8_t *addr[64];
    static bool known[64];
    static std::mutex mu;
    if (dev < 0 || dev >= 64) {
        set_error("blur_device_noalloc: device index %d out of range", dev);
        return B2C_EINVAL;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!known[dev]) {
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        addr[dev] = static_cast<uint8_t *>(p);
        known[dev] = true;
    }
    *out = addr[dev];
    return B2C_OK;
}

}  // namespace

extern "C" int b2c_blur_device_noalloc(const uint8_t *d_in, int64_t in_pitch, uint8_t *d_out, int64_t out_pitch,
                                       int W, int H, int CH, int reps, int boundary, void *stream) {
    if (!d_in || !d_out || W <= 0 || H <= 0 || (CH != 1 && CH != 3) || reps < 0 ||
        (boundary != B2C_ZERO && boundary != B2C_TORUS)) {
        set_error("blur_device_noalloc: bad arguments (boundary must be zero or torus)");
        return B2C_EINVAL;
    }
    int dev = 0;
    CK(cudaGetDevice(&dev));
    cudaStream_t s = (cudaStream_t)stream;

    // A throw-away engine description on the host stack: same planner and pass scheduler as
    // b2c_engine_create/run, but its buffers are carved out of the static scratch.
    b2c_engine e;
    e.device = dev; e.W = W; e.H = H; e.CH = CH; e.P = 1; e.Q = 1; e.rank = 0;
    b2c_default_opts(&e.opts);
    e.opts.boundary = boundary;
    e.opts.use_graph = 0;
    e.R = H; e.C = W; e.row0 = 0; e.col0 = 0;
    if ((long long)W * CH > (1ll << 30)) {
        set_error("blur_device_noalloc: row too wide");
        return B2C_ETOOBIG;
    }
    e.bw = W * CH;
    b2c_neighbours(0, 1, 1, boundary, e.nb);
    e.any_nb = boundary == B2C_TORUS;
    memset(e.peer_base, 0, sizeof e.peer_base);
    memset(e.peer_pad, 0, sizeof e.peer_pad);
    memset(e.rank_pad, 0, sizeof e.rank_pad);
    const FilterDef &f = filter_def(B2C_GAUSSIAN);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            e.nw.iw[i][j] = f.w[i][j];
            e.nw.fw[i][j] = (float)(f.w[i][j] / (double)f.den);
        }
    e.nw.integer16 = 1;
    CK(cudaDeviceGetAttribute(&e.sm_count, cudaDevAttrMultiProcessorCount, dev));
    choose_plan(&e);

    e.pitch = round_up(kLeftPad + (long long)e.bw + 32 + 16, 128);
    e.alloc_bytes = round_up(e.pitch * (e.R + 2 * kFrameRows) + 1024, 256);
    const long long pad_bytes = kPadWords * (long long)sizeof(uint32_t);
    const long long need = 2 * e.alloc_bytes + pad_bytes;
    if (need > B2C_NOALLOC_SCRATCH_BYTES) {
        set_error("blur_device_noalloc: a %dx%dx%d image needs %lld bytes of scratch but the built-in scratch is "
                  "%lld bytes (fits up to about 1920x2520 rgb); use b2c_blur_device or an engine instead",
                  W, H, CH, need, (long long)B2C_NOALLOC_SCRATCH_BYTES);
        return B2C_ETOOBIG;
    }
    uint8_t *scratch = nullptr;
    int rc = noalloc_scratch_base(dev, &scratch);
    if (rc != B2C_OK) return rc;
    e.alloc[0] = scratch;
    e.alloc[1] = scratch + e.alloc_bytes;
    e.pad = reinterpret_cast<uint32_t *>(scratch + 2 * e.alloc_bytes);
    for (int i = 0; i < 2; ++i) e.base[i] = e.alloc[i] + (long long)kFrameRows * e.pitch + kLeftPad;

    // The last call's image, halo frames, epoch flags and counters are still in there.
    CK(cudaMemsetAsync(scratch, 0, (size_t)need, s));
    if (e.any_nb) {  // torus on one GPU: the rank is its own neighbour in all eight directions
        for (int d = 0; d < 8; ++d) {
            e.peer_base[0][d] = e.base[0];
            e.peer_base[1][d] = e.base[1];
            e.peer_pad[d] = e.pad;
        }
        e.rank_pad[0] = e.pad;
        e.connected = true;
    }
    if (in_pitch <= 0) in_pitch = e.bw;
    if (out_pitch <= 0) out_pitch = e.bw;
    CK(cudaMemcpy2DAsync(e.base[0], (size_t)e.pitch, d_in, (size_t)in_pitch, (size_t)e.bw, (size_t)e.R,
                         cudaMemcpyDeviceToDevice, s));
    if (e.any_nb) {
        PushParams q;
        fill_push_params(&e, e.cur, false, q);
        CK(launch_halo_push(q, s));
        e.epoch++;
    }
    for (const Pass &ps : make_passes(&e, reps)) {
        rc = enqueue_pass(&e, ps, s);
        if (rc != B2C_OK) return rc;
    }
    CK(cudaMemcpy2DAsync(d_out, (size_t)out_pitch, e.base[e.cur], (size_t)e.pitch, (size_t)e.bw, (size_t)e.R,
                         cudaMemcpyDeviceToDevice, s));
    return B2C_OK;
}
