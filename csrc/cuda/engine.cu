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
constexpr int kPadFlags = 0;          // 8 x 32: data epochs, index = direction the ring came from
constexpr int kPadCounters = 256;     // 8 x 32: ring-warp arrival counters
constexpr int kPadSlots = 512;        // 2 parities x kMaxRanks words (convergence bits)
constexpr int kPadChanged = 640;      // 2 words: "a byte changed", by check parity (the reduction of a check runs beside the next interval)
constexpr int kPadConverged = 642;    // number of the check that found the fixed point (0: none)
constexpr int kPadConvergedAt = 643;
constexpr int kPadEpoch = 672;        // the epoch at the start of the current run (device-side mirror of e->epoch)
constexpr int kPadBarrier = 704;      // kMaxRanks words: start-barrier sequence numbers, index = rank
constexpr int kPadBarrierSkew = 768;  // 2 words (u64): cycles this rank waited in the last start barrier
constexpr int kPadAcks = 1024;        // 8 x 32: ack epochs, index = direction of the acknowledging neighbour
constexpr int kPadProf = 1280;        // 40 x u64 wait/publish counters (B200CONV_HALO_PROF=1)
constexpr int kPadReady = 1408;       // 16 x 32: epochs handed to the courier (8 data, 8 ack)
constexpr int kProfWords64 = 40;
constexpr int kPadWords = 2048;
static_assert(kPadSlots + 2 * kMaxRanks <= kPadChanged, "convergence slots overlap the next pad field");
static_assert(kPadProf + 2 * kProfWords64 <= kPadReady, "profile counters overlap the courier words");
static_assert(kPadReady + 16 * 32 <= kPadWords, "courier words overlap the end of the pad");

struct Blob {  // what b2c_engine_export ships; must fit B2C_HANDLE_BYTES
    cudaIpcMemHandle_t buf[2];
    cudaIpcMemHandle_t pad;
    int rank, device, pid;
    long long alloc_bytes;
};
static_assert(sizeof(Blob) <= B2C_HANDLE_BYTES, "blob too large");

// One kernel launch of the iteration loop.  Passes form groups of at most F iterations; halos are
// exchanged once per group (see HaloParams in stencil_stream.cuh).
struct Pass {
    int k;                 // fused iterations
    unsigned interval;     // check interval the pass belongs to (1-based; 0 without check_every)
    bool check;            // convergence check after this pass (always the last pass of its group)
    unsigned iters_after;  // iterations completed after it, counted from the start of the run
    int ext;               // iterations left in the group after this pass = how far the block is extended
    int mode;              // kHalo* bits
    bool group_end;
};

// A run prepared for the persistent kernel: the leading passes with k == K as a device table,
// the remainder (k < K at the end of a run) as ordinary launches.
struct PersistPlan {
    PassParams *d_passes = nullptr;
    int *d_cum = nullptr;
    uint32_t *d_counters = nullptr;   // per pass 8 x 32 arrival counters (passes overlap in time)
    int npasses = 0;                  // passes in the table
    long long items = 0;
    int tstride = 0;
    long long tile_words = 0;
    uint32_t groups = 0;              // exchanges completed by the table's passes
    CourierParams *d_courier = nullptr;  // the kernel's own courier (null: the last arrivers publish themselves)
    std::vector<Pass> all;            // the whole run
};

struct BandPlan {
    int hlo, vw, nbands, nchunks, chunk_rows, nbw, nbe;
    int xn, xs, xwb, xeb;   // extension of the block into the frame: rows north/south, bytes west/east
    int rows, bw;           // extended block
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
    uint32_t epoch = 0;      // exchanges completed so far (same on every rank)
    uint32_t epoch_dev = 0;  // what pad[kPadEpoch] holds once everything enqueued so far has run
    uint32_t barrier_seq = 0;
    unsigned check_seq = 0;
    int iters_done = 0;
    bool pending_converge_fixup = false;
    std::vector<Pass> last_passes;
    uint32_t run_epoch0 = 0;
    int run_cur0 = 0;
    std::vector<Pass> staged;   // b2c_engine_plan_run / b2c_engine_run_pass
    // plan
    int kernel = B2C_KERNEL_STREAM, K = 4, SB = 16, F = 1;
    int SB0 = 0;                     // strip width of the modelled plan (coupled tuning may change SB)
    bool chunk_rows_fixed = false;   // the caller gave chunk_rows: never tuned
    int sm_count = 148;
    long long delay = 0;
    bool prof = false;
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
    int res_cx = 0;                  // cluster width of the resident kernel (0: not asked yet)
    mutable std::map<std::pair<int, int>, BandPlan> plan_cache;   // key (k, ext)
    // persistent dataflow kernel: per (reps, cur) a device table of pass parameters
    std::map<std::pair<int, int>, PersistPlan> persist;
    uint32_t *tile_done = nullptr;       // per-tile "passes completed" words
    long long tile_cap = 0;              // words allocated
    unsigned long long *work_counter = nullptr;
    unsigned long long item_base = 0;    // value of *work_counter once everything enqueued so far has run
    uint32_t pass_base = 0;              // passes run through the persistent kernel so far
    int persist_ctas = 0;
    bool persist_auto = true;            // what "auto" means for this engine (block size / autotune)
    // courier (kernels.cu): forwards halo epochs to peers on other devices from a side stream
    bool remote_peers = false;           // some neighbour lives on another device
    bool courier_on = false;             // the passes being enqueued hand their epochs to the courier
    unsigned preloaded = 0;              // bit k: the k-stage pass kernel of this plan is loaded
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_check[2] = {nullptr, nullptr};   // reduction of a convergence check done, by check parity
    bool check_pending[2] = {false, false};
    // NVLS: a small symmetric buffer with a multicast mapping, handed in by the caller (b2c_engine_set_nvls)
    uint32_t *nvls_local = nullptr, *nvls_mc = nullptr;
    uint32_t nvls_barrier_seq = 0;
};

namespace {

// Forget everything that was built from the current plan or peer wiring (CUDA graphs, the
// persistent kernel's pass tables).  The caller has made sure nothing of it is still running.
void drop_graphs(b2c_engine *e) {
    e->preloaded = 0;
    for (auto &g : e->graphs) cudaGraphExecDestroy(g.second);
    e->graphs.clear();
    if (!e->persist.empty()) {
        cudaSetDevice(e->device);
        cudaDeviceSynchronize();
        for (auto &kv : e->persist) {
            cudaFree(kv.second.d_passes);
            cudaFree(kv.second.d_cum);
            cudaFree(kv.second.d_counters);
            cudaFree(kv.second.d_courier);
        }
        e->persist.clear();
    }
}

// Window geometry of a k-stage pass for the engine's strip width over a block `bw` bytes wide.
void band_geometry(const b2c_engine *e, int k, int bw, int *hlo, int *vw, int *nbands) {
    // window start and band pitch: strip-aligned for per-lane LDGs, 16 B for TMA box origins
    const int align = e->kernel == B2C_KERNEL_TMA ? 16 : e->SB;
    *hlo = (int)round_up((long long)k * e->CH, align);
    *vw = (int)round_down(32 * e->SB - *hlo - k * e->CH, align);
    *nbands = (int)ceil_div(bw, *vw);
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
    band_geometry(e, k, e->bw, &hlo, &vw, &nb);
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

// Shortest first / last chunk a plan may have.  With neighbours the F-deep ring must sit inside
// the chunks that push it, and the chunks next to them must not reach into a frame (they do not
// wait for flags): F covers the push, 2k covers the reads of the pass that acknowledges.
int min_edge_chunk(const b2c_engine *e, int k) { return e->any_nb ? std::max(e->F, 2 * k) : 1; }

double best_chunk_rows(const b2c_engine *e, int k, int *out) {
    static const int cand[] = {8, 12, 16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512};
    double best = 1e300;
    int best_cr = std::max(1, e->R);
    for (int c : cand) {
        int cr = std::max(c, std::max(2 * k, min_edge_chunk(e, k)));
        cr = std::min(cr, std::max(e->R, 1));
        const double t = plan_cost(e, k, cr);
        if (t < best * 0.999) { best = t; best_cr = cr; }
        if (cr >= e->R) break;
    }
    *out = best_cr;
    return best;
}

// Plan of a k-stage pass whose block is extended by `ext` iterations into the frame on every side
// that has a neighbour (ext = 0: the block itself).
BandPlan make_band_plan(const b2c_engine *e, int k, int ext) {
    const auto key = std::make_pair(k, ext);
    auto it = e->plan_cache.find(key);
    if (it != e->plan_cache.end()) return it->second;
    BandPlan bp{};
    const int CH = e->CH;
    bp.xn = e->nb[B2C_N] >= 0 ? ext : 0;
    bp.xs = e->nb[B2C_S] >= 0 ? ext : 0;
    // whole 16-byte strips; the bytes past ext*CH are computed from stale frame bytes and are never
    // inside the dependency cone of a byte that is used later
    bp.xwb = e->nb[B2C_W] >= 0 ? (int)round_up((long long)ext * CH, 16) : 0;
    bp.xeb = e->nb[B2C_E] >= 0 ? (int)round_up((long long)ext * CH, 16) : 0;
    bp.rows = e->R + bp.xn + bp.xs;
    bp.bw = e->bw + bp.xwb + bp.xeb;
    band_geometry(e, k, bp.bw, &bp.hlo, &bp.vw, &bp.nbands);
    int chunk_rows = e->opts.chunk_rows;
    if (chunk_rows <= 0) {
        best_chunk_rows(e, k, &chunk_rows);
        const int nc = (int)ceil_div(e->R, chunk_rows);
        chunk_rows = (int)ceil_div(e->R, nc);   // even out the last chunk
    }
    const int min_edge = min_edge_chunk(e, k);
    chunk_rows = std::max(chunk_rows, min_edge);
    chunk_rows = std::min(chunk_rows, std::max(bp.rows, 1));
    chunk_rows = std::max(chunk_rows, 1);
    bp.chunk_rows = chunk_rows;
    bp.nchunks = (int)ceil_div(bp.rows, chunk_rows);
    // the last chunk takes whatever is left (stream_band); a remainder shorter than min_edge is
    // merged into the chunk before it
    if (bp.nchunks > 1 && bp.rows - (bp.nchunks - 1) * chunk_rows < min_edge) bp.nchunks--;
    // Bands that belong to the west / east ring: those whose stored strips hold part of the F-deep
    // ring (they push), and those whose window reaches into the frame (they wait / acknowledge).
    const int fb = e->F * CH, win = 32 * e->SB;
    const int nbw_push = (int)ceil_div(fb, bp.vw);
    const int nbw_read = (int)ceil_div(bp.xwb + bp.hlo, bp.vw);
    bp.nbw = std::max(1, std::min(bp.nbands, std::max(nbw_push, nbw_read)));
    const int nbe_push = bp.nbands - (int)(std::max(0, bp.bw - fb) / bp.vw);
    const long long reach = (long long)bp.bw - bp.xeb + bp.hlo - win;   // band b reads the frame iff b*vw > reach
    const int first_reader = reach < 0 ? 0 : (int)(reach / bp.vw) + 1;
    const int nbe_read = bp.nbands - std::min(bp.nbands, first_reader);
    bp.nbe = std::max(1, std::min(bp.nbands, std::max(nbe_push, nbe_read)));
    e->plan_cache[key] = bp;
    return bp;
}

void fill_pass_params(const b2c_engine *e, const Pass &ps, int src_i, bool device_epoch, PassParams &p) {
    const BandPlan bp = make_band_plan(e, ps.k, ps.ext);
    memset(&p, 0, sizeof p);
    const long long shift = (long long)bp.xn * e->pitch + bp.xwb;
    p.src = e->base[src_i] - shift;
    p.dst = e->base[src_i ^ 1] - shift;
    p.pitch = e->pitch;
    p.rows = bp.rows;
    p.bw = bp.bw;
    p.nbands = bp.nbands;
    p.nchunks = bp.nchunks;
    p.chunk_rows = bp.chunk_rows;
    p.vw = bp.vw;
    p.hlo = bp.hlo;
    p.org_x = kLeftPad - bp.xwb;
    p.org_y = kFrameRows - bp.xn;
    // (order 2, ring last, for passes that only wait was tried: it hides the epoch's latency under PCIe
    //  load but ends every such pass on its slowest tiles; 8 GPUs lost 2.6 % device-timed, e2e unchanged)
    p.raster = 0;
    const bool zero = e->opts.boundary == B2C_ZERO;
    p.zero_n = zero && e->nb[B2C_N] < 0;
    p.zero_s = zero && e->nb[B2C_S] < 0;
    p.zero_w = zero && e->nb[B2C_W] < 0;
    p.zero_e = zero && e->nb[B2C_E] < 0;
    p.changed = ps.check ? e->pad + kPadChanged + (ps.interval & 1u) : nullptr;
    p.converged = e->opts.check_every > 0 ? e->pad + kPadConverged : nullptr;
    p.interval = ps.interval;
#ifdef B2C_DEBUG
    p.dbg_src_lo = e->alloc[src_i];
    p.dbg_src_hi = e->alloc[src_i] + e->alloc_bytes;
    p.dbg_dst_lo = e->alloc[src_i ^ 1];
    p.dbg_dst_hi = e->alloc[src_i ^ 1] + e->alloc_bytes;
    p.dbg_peer_before = (long long)kFrameRows * e->pitch + kLeftPad;
    p.dbg_peer_after = e->alloc_bytes - p.dbg_peer_before;
    p.dbg_rank = e->rank;
#endif
    HaloParams &h = p.halo;
    h.nbw = bp.nbw;
    h.nbe = bp.nbe;
    h.enabled = e->any_nb ? 1 : 0;
    if (!h.enabled) return;
    h.mode = ps.mode;
    h.fused = (e->bw % e->SB == 0) ? 1 : 0;
    static const int flag_fence = getenv("B200CONV_FLAG_FENCE") ? atoi(getenv("B200CONV_FLAG_FENCE")) : 1;
    h.flag_fence = flag_fence;
    h.depth = e->F;
    h.epoch = device_epoch ? e->epoch - e->epoch_dev : e->epoch;
    h.epoch_base = device_epoch ? e->pad + kPadEpoch : nullptr;
    h.flags = e->pad + kPadFlags;
    h.acks = e->pad + kPadAcks;
    h.counters = e->pad + kPadCounters;
    h.prof = e->prof ? reinterpret_cast<unsigned long long *>(e->pad + kPadProf) : nullptr;
    h.ready = e->courier_on && h.fused ? e->pad + kPadReady : nullptr;
    h.delay = e->delay;
    const int exp[8] = {bp.nbands, bp.nbands, bp.nbw * bp.nchunks, bp.nbe * bp.nchunks,
                        bp.nbw, bp.nbe, bp.nbw, bp.nbe};
    for (int d = 0; d < 8; ++d) {
        h.expected[d] = exp[d];
        if (e->nb[d] < 0) continue;
        h.peer_dst[d] = e->peer_base[src_i ^ 1][d];
        // the neighbour sees our data arriving from the opposite direction
        h.peer_flag[d] = e->peer_pad[d] + kPadFlags + opposite_dir(d) * 32;
        h.peer_ack[d] = e->peer_pad[d] + kPadAcks + opposite_dir(d) * 32;
    }
}

void fill_push_params(const b2c_engine *e, int buf_i, bool device_epoch, PushParams &q) {
    memset(&q, 0, sizeof q);
    q.src = e->base[buf_i];
    q.pitch = e->pitch;
    q.rows = e->R;
    q.bw = e->bw;
    q.ch = e->CH;
    q.depth = e->F;
    q.flags = e->pad + kPadFlags;
    q.counters = e->pad + kPadCounters;
    q.post_epoch = (device_epoch ? e->epoch - e->epoch_dev : e->epoch) + 1;
    q.epoch_base = device_epoch ? e->pad + kPadEpoch : nullptr;
#ifdef B2C_DEBUG
    q.dbg_before = (long long)kFrameRows * e->pitch + kLeftPad;
    q.dbg_after = e->alloc_bytes - q.dbg_before;
    q.dbg_rank = e->rank;
#endif
    q.wait_epoch_valid = 0;   // the pass in front of a push has waited for everything the push needs
    q.converged = nullptr;    // only the per-pass fallback push (enqueue_pass) follows the convergence latch
    for (int d = 0; d < 8; ++d) {
        if (e->nb[d] < 0) continue;
        q.peer_dst[d] = e->peer_base[buf_i][d];
        q.peer_flag[d] = e->peer_pad[d] + kPadFlags + opposite_dir(d) * 32;
    }
}

// Splits `reps` iterations into groups of at most F iterations (one halo exchange each; a group
// also ends on every multiple of check_every, counted from the start of this run) and every group
// into passes of k in {K, ..., 2, 1}.  Without neighbours there is nothing to exchange and a
// "group" is just the stretch up to the next check.
std::vector<Pass> make_passes(const b2c_engine *e, int reps) {
    std::vector<Pass> v;
    const int ce = e->opts.check_every;
    int done = 0;
    while (done < reps) {
        int glen = reps - done;
        if (e->any_nb) glen = std::min(glen, std::max(e->F, 1));
        if (ce > 0) glen = std::min(glen, ce - done % ce);
        std::vector<int> ks;
        for (int left = glen; left > 0;) {
            int k = e->K;
            while (k > left) k >>= 1;
            ks.push_back(k);
            left -= k;
        }
        const int n = (int)ks.size();
        int ext = glen;
        for (int j = 0; j < n; ++j) {
            ext -= ks[j];
            done += ks[j];
            Pass ps{};
            ps.k = ks[j];
            ps.group_end = j == n - 1;
            ps.check = ps.group_end && ce > 0 && done % ce == 0;
            ps.interval = ce > 0 ? (unsigned)((done - 1) / ce + 1) : 0u;
            ps.iters_after = (unsigned)done;
            ps.ext = e->any_nb ? ext : 0;
            ps.mode = 0;
            if (e->any_nb) {
                if (j == 0) ps.mode |= kHaloWaitData;
                if (n >= 2 && j == n - 2) ps.mode |= kHaloPostAck;
                if (j == n - 1) ps.mode |= kHaloPush | (n >= 2 ? kHaloWaitAck : 0);
            }
            v.push_back(ps);
        }
    }
    return v;
}

// The stream of the passes joins the reduction of the check with this parity (if one is in flight).
int join_check(b2c_engine *e, unsigned parity, cudaStream_t s) {
    if (!e->check_pending[parity & 1u]) return B2C_OK;
    CK(cudaStreamWaitEvent(s, e->ev_check[parity & 1u], 0));
    e->check_pending[parity & 1u] = false;
    return B2C_OK;
}

int enqueue_pass(b2c_engine *e, const Pass &ps, bool device_epoch, cudaStream_t s) {
    // a pass of interval j obeys the checks up to j-2: the newest of them must have been reduced
    // (the one before it was joined an interval ago); that also frees its parity's "changed" word
    if (ps.interval >= 3 && join_check(e, ps.interval, s) != B2C_OK) return B2C_ECUDA;
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
        fill_pass_params(e, ps, e->cur, device_epoch, p);
        if (e->kernel == B2C_KERNEL_TMA)
            CK(launch_stream_tma(e->tma, p, e->cur, ps.k, e->SB, e->CH, ps.check, s));
        else
            CK(launch_stream(p, ps.k, e->SB, e->CH, ps.check, s));
        e->launches++;
        if (e->any_nb && (ps.mode & kHaloPush) && !p.halo.fused) {
            PushParams q;
            fill_push_params(e, e->cur ^ 1, device_epoch, q);
            if (e->opts.check_every > 0) { q.converged = e->pad + kPadConverged; q.interval = ps.interval; }
            CK(launch_halo_push(q, s));
            e->launches++;
        }
    }
    e->cur ^= 1;
    if (e->any_nb && ps.group_end) e->epoch++;
    if (ps.check) {
        // The reduction of this check runs on a side stream beside the next interval; the stream
        // of the passes only joins it again in front of the interval after that (top of this
        // function).  So the all-ranks handshake is off the critical path, and every rank still
        // skips from exactly the same pass on (PassParams::interval).
        ConvergeParams c{};
        c.changed = e->pad + kPadChanged + (ps.interval & 1u);
        c.slots = e->pad + kPadSlots;
        const int n = e->P * e->Q;
        for (int k = 0; k < n; ++k) c.peer_slots[k] = (n == 1 ? e->pad : e->rank_pad[k]) + kPadSlots;
        c.nranks = n;
        c.rank = e->rank;
        c.seq = ++e->check_seq;
        c.converged = e->pad + kPadConverged;
        c.converged_at = e->pad + kPadConvergedAt;
        c.iters_after = ps.iters_after;
        c.check_no = ps.interval;
        cudaStream_t cs = s;
        if (s) {   // not beside the legacy default stream: there the reduction stays in line
            if (!e->side) {
                CK(cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking));
                CK(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
                CK(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
            }
            if (!e->ev_check[0]) {
                CK(cudaEventCreateWithFlags(&e->ev_check[0], cudaEventDisableTiming));
                CK(cudaEventCreateWithFlags(&e->ev_check[1], cudaEventDisableTiming));
            }
            CK(cudaEventRecord(e->ev_fork, s));
            CK(cudaStreamWaitEvent(e->side, e->ev_fork, 0));
            cs = e->side;
        }
        if (e->nvls_mc && n > 1) {
            NvlsParams nv{e->nvls_local, e->nvls_mc, n, e->rank};
            CK(launch_converge_nvls(c, nv, cs));
        } else {
            CK(launch_converge(c, cs));
        }
        e->launches++;
        if (s) {
            CK(cudaEventRecord(e->ev_check[ps.interval & 1u], e->side));
            e->check_pending[ps.interval & 1u] = true;
        }
    }
    return B2C_OK;
}

int enqueue_epoch_sync(b2c_engine *e, bool device_epoch, cudaStream_t s);

// Peers on other devices: the flag stores need system scope, and that fence is kept out of the
// passes (halo_publish in stencil_stream.cuh) by a courier kernel on a side stream.
bool want_courier(const b2c_engine *e) {
    static const bool off = getenv("B200CONV_COURIER") && atoi(getenv("B200CONV_COURIER")) == 0;
    // not with check_every: after the convergence latch the passes publish nothing and the courier would wait for ever
    return !off && e->any_nb && e->remote_peers && e->kernel != B2C_KERNEL_NAIVE && e->bw % e->SB == 0 &&
           e->opts.check_every == 0;
}

// Forks the courier off stream s for the passes that follow (which must all be enqueued before
// courier_end).  Works inside a stream capture: the side stream joins the capture through the event.
int courier_begin(b2c_engine *e, const std::vector<Pass> &passes, bool device_epoch, cudaStream_t s) {
    e->courier_on = false;
    if (!want_courier(e) || !s) return B2C_OK;   // not beside the legacy default stream
    uint32_t groups = 0, ack_group = 0;
    bool has_ack = false;
    for (const Pass &ps : passes) {
        if (ps.mode & kHaloPostAck) { has_ack = true; ack_group = groups; }
        if (ps.group_end) groups++;
    }
    if (groups == 0) return B2C_OK;
    // every kernel that will be launched while the courier spins must be loaded by now (lazy module
    // loading waits for running kernels: kernels.cu, preload_stream_kernel)
    for (const Pass &ps : passes) {
        if (e->preloaded & (1u << ps.k)) continue;
        CK(e->kernel == B2C_KERNEL_TMA ? preload_stream_tma_kernel(ps.k, e->SB, e->CH, false)
                                       : preload_stream_kernel(ps.k, e->SB, e->CH, false));
        e->preloaded |= 1u << ps.k;
    }
    if (!e->side) {
        CK(cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
    }
    CourierParams c{};
    c.ready = e->pad + kPadReady;
    for (int d = 0; d < 8; ++d) {
        if (e->nb[d] < 0) continue;
        c.peer_flag[d] = e->peer_pad[d] + kPadFlags + opposite_dir(d) * 32;
        c.peer_ack[d] = e->peer_pad[d] + kPadAcks + opposite_dir(d) * 32;
    }
    const uint32_t rel = device_epoch ? e->epoch - e->epoch_dev : e->epoch;
    c.epoch_base = device_epoch ? e->pad + kPadEpoch : nullptr;
    c.start_epoch = rel;
    c.final_data = rel + groups;
    c.final_ack = rel + ack_group;
    c.has_ack = has_ack ? 1 : 0;
    CK(cudaEventRecord(e->ev_fork, s));
    CK(cudaStreamWaitEvent(e->side, e->ev_fork, 0));
    CK(launch_courier(c, e->side));
    e->launches++;
    e->courier_on = true;
    return B2C_OK;
}

int courier_end(b2c_engine *e, cudaStream_t s) {
    if (!e->courier_on) return B2C_OK;
    e->courier_on = false;
    CK(cudaEventRecord(e->ev_join, e->side));
    CK(cudaStreamWaitEvent(s, e->ev_join, 0));
    return B2C_OK;
}

bool persist_enabled(const b2c_engine *e) {
    static const int env = getenv("B200CONV_PERSIST") ? atoi(getenv("B200CONV_PERSIST")) : -1;
    const int want = e->opts.persistent >= 0 ? e->opts.persistent : env;
    if (want == 0 || (want < 0 && !e->persist_auto)) return false;
    if (e->kernel != B2C_KERNEL_STREAM || e->opts.check_every > 0 || e->opts.boundary == B2C_CUDA_COMPAT) return false;
    if (e->any_nb && e->bw % e->SB != 0) return false;   // the ring would need halo_push_kernel between passes
    return true;
}

// Builds (or finds) the pass table of a `reps`-iteration run that starts on buffer e->cur.
int persist_plan(b2c_engine *e, int reps, PersistPlan **out) {
    const auto key = std::make_pair(reps, e->cur);
    auto it = e->persist.find(key);
    if (it != e->persist.end()) {
        *out = &it->second;
        return B2C_OK;
    }
    if (e->persist.size() >= 16) drop_graphs(e);
    PersistPlan pl;
    pl.all = make_passes(e, reps);
    while (pl.npasses < (int)pl.all.size() && pl.all[(size_t)pl.npasses].k == e->K) pl.npasses++;
    if (pl.npasses > 0) {
        const size_t np = (size_t)pl.npasses;
        CK(cudaMalloc(&pl.d_counters, np * 256 * sizeof(uint32_t)));
        CK(cudaMemset(pl.d_counters, 0, np * 256 * sizeof(uint32_t)));
        std::vector<PassParams> tab(np);
        std::vector<int> cum(np + 1, 0);
        // walk the passes with the host state a real run would have; epochs come out relative to the start
        const int cur0 = e->cur;
        const uint32_t epoch0 = e->epoch, dev0 = e->epoch_dev;
        e->epoch_dev = e->epoch;
        const bool courier0 = e->courier_on;
        bool publishes = false;   // does any pass of the table publish an epoch?
        for (size_t t = 0; t < np; ++t) publishes |= (pl.all[t].mode & (kHaloPush | kHaloPostAck)) != 0;
        const bool courier = want_courier(e) && publishes;
        e->courier_on = courier;
        int max_nb = 1, max_nc = 1;
        for (size_t t = 0; t < np; ++t) {
            const Pass &ps = pl.all[t];
            PassParams &p = tab[t];
            fill_pass_params(e, ps, e->cur, true, p);
            p.halo.epoch_base = nullptr;   // the kernel adds the epoch at launch
            p.halo.counters = pl.d_counters + t * 256;
            p.raster = (!e->any_nb || !(ps.mode & (kHaloPush | kHaloPostAck))) ? 1 : 0;
            cum[t + 1] = cum[t] + p.nbands * p.nchunks;
            max_nb = std::max(max_nb, p.nbands);
            max_nc = std::max(max_nc, p.nchunks);
            e->cur ^= 1;
            if (e->any_nb && ps.group_end) { e->epoch++; pl.groups++; }
        }
        e->cur = cur0;
        e->epoch = epoch0;
        e->epoch_dev = dev0;
        e->courier_on = courier0;
        pl.items = cum[np];
        pl.tstride = max_nb;
        pl.tile_words = (long long)max_nb * max_nc;
        CK(cudaMalloc(&pl.d_passes, np * sizeof(PassParams)));
        CK(cudaMalloc(&pl.d_cum, (np + 1) * sizeof(int)));
        CK(cudaMemcpy(pl.d_passes, tab.data(), np * sizeof(PassParams), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(pl.d_cum, cum.data(), (np + 1) * sizeof(int), cudaMemcpyHostToDevice));
        if (courier) {
            CourierParams c{};
            c.ready = e->pad + kPadReady;
            for (int d = 0; d < 8; ++d) {
                if (e->nb[d] < 0) continue;
                c.peer_flag[d] = e->peer_pad[d] + kPadFlags + opposite_dir(d) * 32;
                c.peer_ack[d] = e->peer_pad[d] + kPadAcks + opposite_dir(d) * 32;
            }
            uint32_t g = 0;
            for (size_t t = 0; t < np; ++t) {
                if (pl.all[t].mode & kHaloPostAck) { c.has_ack = 1; c.final_ack = g; }
                if (pl.all[t].group_end) g++;
            }
            c.start_epoch = 0;          // the kernel adds the epoch at launch
            c.final_data = pl.groups;
            CK(cudaMalloc(&pl.d_courier, sizeof c));
            CK(cudaMemcpy(pl.d_courier, &c, sizeof c, cudaMemcpyHostToDevice));
        }
    }
    *out = &e->persist.emplace(key, std::move(pl)).first->second;
    return B2C_OK;
}

int run_persistent(b2c_engine *e, int reps, cudaStream_t s) {
    PersistPlan *pl = nullptr;
    int rc = persist_plan(e, reps, &pl);
    if (rc != B2C_OK) return rc;
    if (pl->npasses > 0) {
        if (!e->work_counter) {
            CK(cudaMalloc(&e->work_counter, 8 * sizeof(unsigned long long)));   // [0] counter, [1..5] profile
            CK(cudaMemset(e->work_counter, 0, 8 * sizeof(unsigned long long)));
            e->item_base = 0;
        }
        if (pl->tile_words > e->tile_cap) {
            CK(cudaDeviceSynchronize());   // nobody may still be reading the old words
            if (e->tile_done) cudaFree(e->tile_done);
            e->tile_done = nullptr;
            e->tile_cap = 0;
            CK(cudaMalloc(&e->tile_done, (size_t)pl->tile_words * sizeof(uint32_t)));
            // any value not above pass_base reads as "not done yet"
            CK(cudaMemset(e->tile_done, 0, (size_t)pl->tile_words * sizeof(uint32_t)));
            if (e->pass_base > 0) {
                std::vector<uint32_t> init((size_t)pl->tile_words, e->pass_base);
                CK(cudaMemcpy(e->tile_done, init.data(), init.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
            }
            e->tile_cap = pl->tile_words;
        }
        if (!e->persist_ctas) {
            int occ = stream_persist_occupancy(e->K, e->SB, e->CH);
            if (occ <= 0) occ = 1;
            e->persist_ctas = occ * e->sm_count;
        }
    }
    if (pl->npasses > 0) {
        PersistParams q{};
        q.courier = pl->d_courier;
        q.passes = pl->d_passes;
        q.cum = pl->d_cum;
        q.npasses = pl->npasses;
        q.counter = e->work_counter;
        q.item_base = e->item_base;
        q.tile_done = e->tile_done;
        q.tstride = pl->tstride;
        q.pass_base = e->pass_base;
        q.epoch0 = e->epoch;
        // B200CONV_PERSIST_PROF=1: where the persistent kernel's warps spend their cycles (5 words behind the counter)
        static const bool pprof = getenv("B200CONV_PERSIST_PROF") && atoi(getenv("B200CONV_PERSIST_PROF")) != 0;
        q.prof = pprof ? e->work_counter + 1 : nullptr;
        const int ctas = (int)std::min<long long>(e->persist_ctas, ceil_div(pl->items, kWarpsPerCta));
        CK(launch_stream_persist(q, e->K, e->SB, e->CH, ctas, s));
        e->launches++;
        // every warp of the grid leaves after exactly one fetch past the end
        e->item_base += (unsigned long long)pl->items + (unsigned long long)ctas * kWarpsPerCta;
        e->pass_base += (uint32_t)pl->npasses;
        e->cur ^= pl->npasses & 1;
        e->epoch += pl->groups;
    }
    for (size_t t = (size_t)pl->npasses; t < pl->all.size(); ++t)   // k < K at the end of the run
        if ((rc = enqueue_pass(e, pl->all[t], false, s)) != B2C_OK) return rc;   // these publish their epochs themselves
    return enqueue_epoch_sync(e, false, s);
}

// After the last pass of a run the device-side epoch word catches up with the host's count.
int enqueue_epoch_sync(b2c_engine *e, bool device_epoch, cudaStream_t s) {
    if (!e->any_nb || e->kernel == B2C_KERNEL_NAIVE) { e->epoch_dev = e->epoch; return B2C_OK; }
    if (device_epoch) {
        if (e->epoch != e->epoch_dev) {
            CK(launch_add_u32(e->pad + kPadEpoch, e->epoch - e->epoch_dev, s));
            e->launches++;
        }
    } else {
        CK(launch_set_u32(e->pad + kPadEpoch, e->epoch, s));
        e->launches++;
    }
    e->epoch_dev = e->epoch;
    return B2C_OK;
}

void choose_plan(b2c_engine *e) {
    const b2c_opts &o = e->opts;
    e->kernel = o.kernel == B2C_KERNEL_AUTO ? B2C_KERNEL_STREAM : o.kernel;
    // AUTO: an image that fits the register files of the chip stays there for the whole run
    // (stencil_resident.cu; 1920x2520 grey: 2.1x the streaming kernel).  Any option that only
    // the streaming kernels understand keeps them.
    static const bool res_auto = !(getenv("B200CONV_RESIDENT_AUTO") && atoi(getenv("B200CONV_RESIDENT_AUTO")) == 0);
    if (res_auto && o.kernel == B2C_KERNEL_AUTO && o.filter == B2C_GAUSSIAN && e->P * e->Q == 1 &&
        o.boundary != B2C_CUDA_COMPAT && o.check_every == 0 && o.strip_bytes == 0 && o.chunk_rows == 0 &&
        o.halo_depth == 0 && o.persistent < 0 && o.use_graph < 0) {
        for (int k = 16; k >= 4; k >>= 1)
            if (resident_plan(e->bw, e->R, e->CH, k, e->sm_count, nullptr)) e->kernel = B2C_KERNEL_RESIDENT;
    }
    if (o.filter != B2C_GAUSSIAN) e->kernel = B2C_KERNEL_NAIVE;
    int K = o.tb > 0 ? o.tb : 4;
    if (K >= 8) K = 8; else if (K >= 4) K = 4; else if (K >= 2) K = 2; else K = 1;
    if (e->kernel == B2C_KERNEL_NAIVE) K = 1;
    // a k-deep halo must come from the adjacent rank only
    if (e->any_nb) while (K > 1 && K > std::min(e->R, e->C)) K >>= 1;
    e->K = K;
    e->F = e->any_nb ? K : 1;
    if (e->kernel == B2C_KERNEL_NAIVE) e->F = 1;
    e->chunk_rows_fixed = o.chunk_rows > 0;
    // The persistent kernel needs well over two waves of work items per pass: an item may start
    // when its nine predecessors of the pass before are done, a warp whose item is not ready yet
    // just waits, and with fewer items per pass than that the slow edge tiles of consecutive passes
    // form the critical path (measured: profiles/persist_sweep.txt).  Smaller blocks keep one launch
    // per pass (single-rank engines measure both in autotune_chunk_rows instead).
    e->persist_auto = (double)e->bw * e->R >= 300e6;
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
    if (e->any_nb && e->kernel != B2C_KERNEL_NAIVE) {
        // Halo depth F = iterations between two exchanges (a multiple of K).  Every handshake with
        // the neighbours costs microseconds whatever the block size; the price of a deeper halo is
        // recomputing (F-K)/2 rows and columns of the neighbours' blocks per pass on average.
        const int cap = std::min(kMaxHaloDepth, std::min(e->R, e->C));
        int F = K;
        if (o.halo_depth > 0) {
            F = std::max(K, std::min(o.halo_depth, cap) / K * K);
        } else {
            for (int m = 8; m >= 2; m >>= 1) {
                const double extra = 0.5 * (m - 1) * K * (2.0 / e->R + 2.0 / e->C);
                if (m * K <= cap && extra <= 0.02) { F = m * K; break; }
            }
        }
        e->F = F;
    }
    // the byte-per-thread kernel has no change detector: a check would always report "unchanged"
    if (e->kernel == B2C_KERNEL_NAIVE || e->kernel == B2C_KERNEL_RESIDENT) e->opts.check_every = 0;
}

// Rows per warp for large blocks are measured, not modelled: how the block scheduler packs
// 4-6 waves of CTAs is not captured by plan_cost (160-row chunks beat 128-row ones by 4 % on
// 16384² rgb, 112 beat 128 on grey).  A handful of candidates is timed on the engine's own (still
// all-zero) buffers with the halo logic off: the kernel's time does not depend on pixel values,
// zeros blur to zeros, and no frame or flag is touched.  ~10 ms once per engine; set
// B200CONV_AUTOTUNE=0 or give chunk_rows explicitly to skip it.
int run_streaming(b2c_engine *e, int reps, cudaStream_t s);

int autotune_chunk_rows(b2c_engine *e) {
    static const bool off = getenv("B200CONV_AUTOTUNE") && atoi(getenv("B200CONV_AUTOTUNE")) == 0;
    if (off || e->opts.chunk_rows > 0) return B2C_OK;
    if (e->kernel != B2C_KERNEL_STREAM && e->kernel != B2C_KERNEL_TMA) return B2C_OK;
    // Multi-rank blocks are tuned together with their neighbours once the peers are connected
    // (b2c_engine_tune_*): what is timed here runs without neighbours, and measured on 4 and 8 GPUs
    // the halo-free optimum (longer chunks) is up to 20 % slower once ring warps have to finish their
    // chunk before a neighbour's next pass may start.
    if (e->P * e->Q > 1) return B2C_OK;
    const double bytes = (double)e->bw * e->R;
    if (bytes < 1e6) return B2C_OK;   // tiny images: a pass is shorter than the launch path, keep the model
    static const int cand[] = {8, 12, 16, 24, 32, 48, 64, 80, 96, 112, 128, 144, 160, 192};
    const bool any_nb = e->any_nb;
    e->any_nb = false;
    const int use_graph = e->opts.use_graph;
    e->opts.use_graph = 0;
    // timed without the convergence check: on the all-zero buffers the first check would end the run,
    // and its bookkeeping (a pending rewind of buffer index and epoch) has no business in a trial
    const int check_every = e->opts.check_every;
    e->opts.check_every = 0;
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
    bool best_pm = e->persist_auto;
    const bool pm_open = e->opts.persistent < 0 && !getenv("B200CONV_PERSIST") && bytes >= 32e6 && bytes < 300e6;
    for (int pm = 1; pm >= 0 && rc == B2C_OK; --pm) {
    if (pm_open) e->persist_auto = pm != 0;
    else if (pm == 0) break;
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
            e->plan_cache.clear();
            drop_graphs(e);
            // the path a real run takes (persistent kernel or one launch per pass), once to warm up
            // and build its tables, once timed
            const int passes = bytes < 16e6 ? 40 : (bytes < 128e6 ? 8 : 6);   // short passes: average more of them
            int rr = run_streaming(e, passes * e->K, s);
            cudaEventRecord(ev0, s);
            if (rr == B2C_OK) rr = run_streaming(e, passes * e->K, s);
            cudaEventRecord(ev1, s);
            if (rr != B2C_OK || cudaEventSynchronize(ev1) != cudaSuccess) {
                if (rr == B2C_OK) set_error("autotune: %s", cudaGetErrorString(cudaGetLastError()));
                rc = B2C_ECUDA;
                break;
            }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, ev0, ev1);
            if (ms < best_ms) { best_ms = ms; best = c; best_sb = sb; best_pm = e->persist_auto; }
        }
    }
    }
    e->SB = best_sb;
    e->persist_auto = best_pm;
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    if (s) cudaStreamDestroy(s);
    e->any_nb = any_nb;
    e->opts.chunk_rows = best;   // 0 (model) when nothing was timed
    e->opts.use_graph = use_graph;
    e->opts.check_every = check_every;
    e->plan_cache.clear();
    drop_graphs(e);
    e->cur = 0;
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
        e->pitch = round_up(kLeftPad + (long long)e->bw + kRightPad, 128);
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
    o->halo_depth = 0;
    o->persistent = -1;
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
    e->prof = getenv("B200CONV_HALO_PROF") && atoi(getenv("B200CONV_HALO_PROF")) != 0;
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
        if ((P * Q != 1 || o.boundary == B2C_CUDA_COMPAT || !coop || !fits) && o.kernel == B2C_KERNEL_AUTO) {
            e->kernel = B2C_KERNEL_STREAM;   // AUTO never fails on this: the streaming kernel takes anything
        } else if (P * Q != 1 || o.boundary == B2C_CUDA_COMPAT || !coop || !fits) {
            set_error("the resident kernel needs one GPU, zero/torus, row bytes %% 8 == 0 and an image that fits "
                      "the register files (%d x %d bytes does not, or cooperative launch is missing)", e->bw, e->R);
            b2c_engine_destroy(e);
            return nullptr;
        }
    }
    if (e->kernel == B2C_KERNEL_RESIDENT) {
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
    drop_graphs(e);
    if (e->tile_done) cudaFree(e->tile_done);
    if (e->work_counter) cudaFree(e->work_counter);
    if (e->side) cudaStreamDestroy(e->side);
    if (e->ev_fork) cudaEventDestroy(e->ev_fork);
    if (e->ev_join) cudaEventDestroy(e->ev_join);
    for (auto &ev : e->ev_check) if (ev) cudaEventDestroy(ev);
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
    e->remote_peers = n > 1;
    drop_graphs(e);
    return B2C_OK;
}

int b2c_engine_disconnect(b2c_engine *e) {
    if (!e) return B2C_EINVAL;
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    drop_graphs(e);
    for (void *p : e->ipc_opened) cudaIpcCloseMemHandle(p);
    e->ipc_opened.clear();
    memset(e->peer_base, 0, sizeof e->peer_base);
    memset(e->peer_pad, 0, sizeof e->peer_pad);
    memset(e->rank_pad, 0, sizeof e->rank_pad);
    e->connected = false;
    e->remote_peers = false;
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
            if (engines[k]->device != e->device) e->remote_peers = true;
        }
        e->connected = true;
        drop_graphs(e);
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
    q.epoch_word = e->pad + kPadEpoch;   // the device-side epoch follows (block (0,0) stores post_epoch)
    CK(launch_halo_push(q, (cudaStream_t)stream));
    e->launches++;
    e->epoch++;
    e->epoch_dev = e->epoch;
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
        if (!e->res_cx) e->res_cx = resident_cluster_width(e->res_plan, e->CH);
        q.cx = e->res_cx;
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
    return run_streaming(e, reps, s);
}

}  // extern "C"

namespace {

// The iteration loop of the streaming kernels: one persistent kernel, or one launch per pass
// (replayed as a CUDA graph when the run is long enough).
int run_streaming(b2c_engine *e, int reps, cudaStream_t s) {
    if (persist_enabled(e)) {
        e->iters_done = reps;
        return run_persistent(e, reps, s);
    }
    const std::vector<Pass> passes = make_passes(e, reps);
    e->iters_done = reps;
    // Epochs come from a device word plus a per-pass offset, so a run is the same sequence of
    // launches every time and can be replayed as a CUDA graph — with neighbours too.  Only the
    // convergence check (whose latch rewinds the epoch) and the byte-per-thread kernel bake
    // absolute epochs into the parameters.
    const bool device_epoch = e->opts.check_every == 0 && e->kernel != B2C_KERNEL_NAIVE;
    if (e->opts.check_every > 0) {
        CK(cudaMemsetAsync(e->pad + kPadChanged, 0, 4 * sizeof(uint32_t), s));   // changed[2], converged, converged_at
        e->last_passes = passes;
        e->run_epoch0 = e->epoch;
        e->run_cur0 = e->cur;
        e->pending_converge_fixup = true;
    }
    bool graph = e->opts.use_graph == 1 || (e->opts.use_graph < 0 && passes.size() >= 8);
    if (!device_epoch || s == nullptr) graph = false;
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
        const uint32_t epoch0 = e->epoch;
        if (it == e->graphs.end()) {
            if (e->graphs.size() >= 16) drop_graphs(e);  // callers that vary `reps` endlessly must not hoard graphs
            const int cur0 = e->cur;
            const long long l0 = e->launches, g0 = total_launches();
            cudaGraph_t g;
            CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
            int rc = courier_begin(e, passes, true, s);
            for (const Pass &ps : passes)
                if (rc != B2C_OK || (rc = enqueue_pass(e, ps, true, s)) != B2C_OK) break;
            if (rc == B2C_OK) rc = courier_end(e, s);
            if (rc == B2C_OK) rc = enqueue_epoch_sync(e, true, s);
            cudaError_t ce = cudaStreamEndCapture(s, &g);
            e->cur = cur0;
            e->epoch = e->epoch_dev = epoch0;
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
        e->launches += 1;   // one host launch; the graph's kernel nodes are counted by count_launch
        long long nodes = (long long)passes.size();
        uint32_t groups = 0;
        for (const Pass &ps : passes) {
            if (e->any_nb && ps.group_end) groups++;
            if (e->any_nb && (ps.mode & kHaloPush) && e->bw % e->SB != 0) nodes++;
        }
        if (groups) nodes += want_courier(e) ? 2 : 1;
        count_launch(nodes);
        e->cur ^= (int)(passes.size() & 1);
        e->epoch = e->epoch_dev = epoch0 + groups;
        return B2C_OK;
    }
    int rc = courier_begin(e, passes, device_epoch, s);
    for (const Pass &ps : passes)
        if (rc != B2C_OK || (rc = enqueue_pass(e, ps, device_epoch, s)) != B2C_OK) return rc;
    if ((rc = courier_end(e, s)) != B2C_OK) return rc;
    if (s && (join_check(e, 0, s) != B2C_OK || join_check(e, 1, s) != B2C_OK)) return B2C_ECUDA;   // reductions still in flight
    return enqueue_epoch_sync(e, device_epoch, s);
}

}  // namespace

extern "C" {

// Pass-by-pass variant of b2c_engine_run for callers that interleave several engines on ONE
// device (the sequential multi-rank emulation of the tests): plan_run fixes the pass list,
// run_pass enqueues pass i.  Every rank must have enqueued pass i before any rank enqueues
// pass i+1 — then every flag a kernel waits for is already set when it starts.
int b2c_engine_plan_run(b2c_engine *e, int reps) {
    if (!e || reps < 0) return B2C_EINVAL;
    if (settle(e) != B2C_OK) return B2C_ECUDA;
    if (e->opts.boundary == B2C_CUDA_COMPAT || e->kernel == B2C_KERNEL_RESIDENT || e->opts.check_every > 0) {
        set_error("engine_plan_run: streaming kernels without check_every only");
        return B2C_EINVAL;
    }
    if (e->any_nb && (!e->connected || e->epoch == 0)) {
        set_error("engine_plan_run: connect the peers and call b2c_engine_exchange first");
        return B2C_EINVAL;
    }
    e->staged = make_passes(e, reps);
    e->iters_done = reps;
    return (int)e->staged.size();
}

int b2c_engine_run_pass(b2c_engine *e, int i, void *stream) {
    if (!e || i < 0 || i >= (int)e->staged.size()) return B2C_EINVAL;
    CK(cudaSetDevice(e->device));
    const bool device_epoch = e->kernel != B2C_KERNEL_NAIVE;
    int rc = enqueue_pass(e, e->staged[(size_t)i], device_epoch, (cudaStream_t)stream);
    if (rc == B2C_OK && i + 1 == (int)e->staged.size()) {
        rc = enqueue_epoch_sync(e, device_epoch, (cudaStream_t)stream);
        e->staged.clear();
    }
    return rc;
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
            // v[0] = number c of the check that found the fixed point: every rank ran the intervals up
            // to c+1 and skipped the rest (the image did not change any more after check c)
            for (const Pass &ps : e->last_passes) {
                if (ps.interval > v[0] + 1) break;
                cur ^= 1;
                if (e->any_nb && ps.group_end) ep++;
            }
            e->cur = cur;
            e->epoch = e->epoch_dev = ep;
            e->iters_done = (int)v[1];
            if (e->any_nb) CK(cudaMemcpy(e->pad + kPadEpoch, &ep, sizeof ep, cudaMemcpyHostToDevice));
        }
    }
    return e->iters_done;
}

int64_t b2c_engine_launches(const b2c_engine *e) { return e ? e->launches : 0; }

int b2c_engine_describe(const b2c_engine *e, char *buf, size_t buflen) {
    if (!e || !buf) return B2C_EINVAL;
    const char *kn = e->kernel == B2C_KERNEL_NAIVE ? "naive" : (e->kernel == B2C_KERNEL_TMA ? "tma" :
                     (persist_enabled(e) ? "stream-persistent" : "stream"));
    if (e->opts.boundary == B2C_CUDA_COMPAT) {
        snprintf(buf, buflen, "kernel=compat W=%d H=%d CH=%d", e->W, e->H, e->CH);
        return B2C_OK;
    }
    if (e->kernel == B2C_KERNEL_RESIDENT) {
        snprintf(buf, buflen, "kernel=resident tb=%d ctas=%dx%d cluster=%dx1 rows_per_warp=%d valid=%dx%dB block=%dx%d(bytes %d)",
                 e->res_plan.k, e->res_plan.ctarows, e->res_plan.strips, e->res_cx > 0 ? e->res_cx : 1, e->res_plan.rw,
                 e->res_plan.vr, e->res_plan.vc, e->R, e->C, e->bw);
        return B2C_OK;
    }
    const BandPlan bp = make_band_plan(e, e->K, 0);
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
    e.opts.kernel = kernel == B2C_KERNEL_AUTO ? B2C_KERNEL_STREAM : kernel;   // the plan of the streaming kernels
    e.R = rows, e.C = cols, e.CH = channels, e.bw = cols * channels;
    e.W = cols, e.H = rows;
    e.any_nb = has_neighbours != 0;
    for (int &n : e.nb) n = has_neighbours ? 0 : -1;
    e.sm_count = sm_count > 0 ? sm_count : 148;
    choose_plan(&e);
    const BandPlan bp = make_band_plan(&e, e.K, 0);
    const int v[10] = {bp.hlo, bp.vw, bp.nbands, bp.nchunks, bp.chunk_rows, bp.nbw, bp.nbe, e.K, e.SB, e.F};
    memcpy(out, v, sizeof v);
    return B2C_OK;
}

// Test hook (host only): the pass list of a run and each pass's plan, for the CPU property tests of
// the grouped halo exchange and of the persistent kernel's tile dependencies.  Per pass 16 ints:
// {k, ext, mode, group_end, rows, row_bytes, nbands, nchunks, chunk_rows, last_chunk_rows, nbw, nbe,
// hlo, vw, xn, xwb} (rows / row_bytes of the extended block; xn rows and xwb bytes of it lie in the
// north / west frame).  Returns the number of passes (at most max_passes are written) or a negative
// code.  A rank with neighbours on all eight sides is modelled.
int b2c_debug_pass_plans(int rows, int cols, int channels, int has_neighbours, int tb, int strip_bytes, int chunk_rows,
                         int halo_depth, int reps, int sm_count, int *out, int max_passes) {
    if (rows <= 0 || cols <= 0 || (channels != 1 && channels != 3) || reps < 0 || !out) return B2C_EINVAL;
    if ((long long)cols * channels > (1ll << 30)) return B2C_ETOOBIG;
    b2c_engine e;
    b2c_default_opts(&e.opts);
    e.opts.tb = tb;
    e.opts.strip_bytes = strip_bytes;
    e.opts.chunk_rows = chunk_rows;
    e.opts.halo_depth = halo_depth;
    e.opts.kernel = B2C_KERNEL_STREAM;
    e.R = rows, e.C = cols, e.CH = channels, e.bw = cols * channels;
    e.W = cols, e.H = rows;
    e.any_nb = has_neighbours != 0;
    for (int &n : e.nb) n = has_neighbours ? 0 : -1;
    e.sm_count = sm_count > 0 ? sm_count : 148;
    choose_plan(&e);
    const std::vector<Pass> passes = make_passes(&e, reps);
    int n = 0;
    for (const Pass &ps : passes) {
        if (n < max_passes) {
            const BandPlan bp = make_band_plan(&e, ps.k, ps.ext);
            const int v[16] = {ps.k, ps.ext, ps.mode, ps.group_end ? 1 : 0, bp.rows, bp.bw, bp.nbands, bp.nchunks,
                               bp.chunk_rows, bp.rows - (bp.nchunks - 1) * bp.chunk_rows, bp.nbw, bp.nbe,
                               bp.hlo, bp.vw, bp.xn, bp.xwb};
            memcpy(out + 16 * n, v, sizeof v);
        }
        ++n;
    }
    return n;
}

// ---- coupled tuning of rows-per-warp (collective: every rank makes the same calls in the same order)

static const int kTuneCand[] = {24, 32, 40, 48, 56, 64, 80, 96, 112, 128, 160, 192};

// Candidate i -> (strip bytes, rows per warp); returns the number of candidates for i < 0, 0 for an
// index past the end.  Strip widths: the planned one first, then the other of {16, 8} unless the
// caller fixed it (narrower strips never won on blocks large enough to be tuned).
static int tune_candidate(const b2c_engine *e, int i, int *sb_out, int *cr_out) {
    int sbs[2] = {e->SB0 > 0 ? e->SB0 : e->SB, 0};
    int nsb = 1;
    if (e->opts.strip_bytes == 0) {
        const int other = sbs[0] == 16 ? 8 : 16;
        if (e->K * (other / 4) <= 16) sbs[nsb++] = other;
    }
    int n = 0;
    const int lo = min_edge_chunk(e, e->K);
    for (int b = 0; b < nsb; ++b)
        for (int c : kTuneCand) {
            if (c < lo || 2 * c > e->R) continue;
            if (n == i) {
                if (sb_out) *sb_out = sbs[b];
                if (cr_out) *cr_out = c;
                return c;
            }
            ++n;
        }
    return i < 0 ? n : 0;
}

int b2c_engine_tune_count(const b2c_engine *e) {
    if (!e) return B2C_EINVAL;
    static const bool off = getenv("B200CONV_AUTOTUNE") && atoi(getenv("B200CONV_AUTOTUNE")) == 0;
    if (off || e->chunk_rows_fixed || e->P * e->Q == 1 || !e->any_nb) return 0;
    if (e->kernel != B2C_KERNEL_STREAM && e->kernel != B2C_KERNEL_TMA) return 0;
    if ((double)e->bw * e->R < 4e6) return 0;   // a pass is shorter than a handshake: keep the model
    return tune_candidate(e, -1, nullptr, nullptr);
}

static void tune_apply(b2c_engine *e, int sb, int cr) {
    e->SB = sb;
    e->opts.chunk_rows = cr;
    e->persist_ctas = 0;   // occupancy depends on the strip width
    e->plan_cache.clear();
    drop_graphs(e);
}

// Times candidate i with the halo exchange live: a fresh exchange, one untimed stretch, one timed
// stretch of the iteration loop on whatever the buffers hold (kernel time does not depend on pixel
// values).  Returns milliseconds on this rank; the caller reduces with MAX over the ranks.
double b2c_engine_tune_time(b2c_engine *e, int i, void *stream) {
    int sb = 0, cr = 0;
    if (!e || tune_candidate(e, i, &sb, &cr) <= 0 || !e->connected) return -1.0;
    if (cudaSetDevice(e->device) != cudaSuccess) return -1.0;
    cudaStream_t s = (cudaStream_t)stream;
    if (e->SB0 == 0) e->SB0 = e->SB;
    tune_apply(e, sb, cr);
    const int ug = e->opts.use_graph, ce = e->opts.check_every;
    e->opts.use_graph = 0;
    e->opts.check_every = 0;   // a trial is not a run: no early exit on whatever the buffers hold, no pending rewind
    const int reps = 8 * std::max(e->F, e->K);
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float ms = -1.f;
    if (cudaEventCreate(&ev0) == cudaSuccess && cudaEventCreate(&ev1) == cudaSuccess &&
        b2c_engine_exchange(e, stream) == B2C_OK && b2c_engine_run(e, reps, stream) == B2C_OK &&
        cudaEventRecord(ev0, s) == cudaSuccess && b2c_engine_run(e, reps, stream) == B2C_OK &&
        cudaEventRecord(ev1, s) == cudaSuccess && cudaEventSynchronize(ev1) == cudaSuccess)
        cudaEventElapsedTime(&ms, ev0, ev1);
    else
        set_error("engine_tune_time: %s", cudaGetErrorString(cudaGetLastError()));
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    e->opts.use_graph = ug;
    e->opts.check_every = ce;
    return ms;
}

int b2c_engine_tune_set(b2c_engine *e, int i) {
    if (!e) return B2C_EINVAL;
    if (e->SB0 == 0) e->SB0 = e->SB;
    int sb = e->SB0, cr = 0;   // i < 0: the modelled plan
    if (i >= 0 && tune_candidate(e, i, &sb, &cr) <= 0) return B2C_EINVAL;
    tune_apply(e, sb, cr);
    return B2C_OK;
}

// ---- device-side start barrier and wait-time counters

int b2c_engine_barrier(b2c_engine *e, void *stream) {
    if (!e) return B2C_EINVAL;
    const int n = e->P * e->Q;
    if (n == 1) return B2C_OK;
    if (!e->connected) {
        set_error("engine_barrier: peers are not connected");
        return B2C_EINVAL;
    }
    CK(cudaSetDevice(e->device));
    if (e->nvls_mc) {
        NvlsParams nv{e->nvls_local, e->nvls_mc, n, e->rank};
        CK(launch_barrier_nvls(nv, ++e->nvls_barrier_seq, reinterpret_cast<unsigned long long *>(e->pad + kPadBarrierSkew),
                               (cudaStream_t)stream));
        e->launches++;
        return B2C_OK;
    }
    BarrierParams b{};
    b.slots = e->pad + kPadBarrier;
    for (int k = 0; k < n; ++k) b.peer_slots[k] = e->rank_pad[k] + kPadBarrier;
    b.nranks = n;
    b.rank = e->rank;
    b.seq = ++e->barrier_seq;
    b.skew_cycles = reinterpret_cast<unsigned long long *>(e->pad + kPadBarrierSkew);
    CK(launch_barrier(b, (cudaStream_t)stream));
    e->launches++;
    return B2C_OK;
}

// NVLS multicast (collective; the caller owns the memory): `local` is this rank's replica of a
// zero-initialised symmetric buffer of at least 64 words, `mc` its multicast address.  From then
// on the convergence OR and the start barrier are one multimem.red each.  Pass nulls to go back to
// peer stores.
int b2c_engine_set_nvls(b2c_engine *e, uint32_t *local, uint32_t *mc, int words) {
    if (!e || (local == nullptr) != (mc == nullptr) || (local && words < kNvlsWords)) {
        set_error("engine_set_nvls: need both pointers and at least %d words", kNvlsWords);
        return B2C_EINVAL;
    }
    e->nvls_local = local;
    e->nvls_mc = mc;
    e->nvls_barrier_seq = 0;
    return B2C_OK;
}

int64_t b2c_engine_barrier_wait_cycles(b2c_engine *e) {
    if (!e) return B2C_EINVAL;
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    unsigned long long v = 0;
    CK(cudaMemcpy(&v, e->pad + kPadBarrierSkew, sizeof v, cudaMemcpyDeviceToHost));
    return (int64_t)v;
}

// B200CONV_PERSIST_PROF=1: cycles summed over the persistent kernel's warps since the last call:
// {waiting for tiles, inside tiles, total, items, items that had to wait}.
int b2c_engine_persist_profile(b2c_engine *e, uint64_t out[5]) {
    if (!e || !out) return B2C_EINVAL;
    memset(out, 0, 5 * sizeof(uint64_t));
    if (!e->work_counter) return B2C_OK;
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(out, e->work_counter + 1, 5 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    CK(cudaMemset(e->work_counter + 1, 0, 5 * sizeof(uint64_t)));
    return B2C_OK;
}

int b2c_engine_set_profile(b2c_engine *e, int on) {
    if (!e) return B2C_EINVAL;
    e->prof = on != 0;
    drop_graphs(e);
    return B2C_OK;
}

int b2c_engine_halo_profile(b2c_engine *e, uint64_t out[40], int reset) {
    if (!e || !out) return B2C_EINVAL;
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(out, e->pad + kPadProf, kProfWords64 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    if (reset) CK(cudaMemset(e->pad + kPadProf, 0, kProfWords64 * sizeof(uint64_t)));
    return B2C_OK;
}

int b2c_engine_set_delay(b2c_engine *e, int64_t cycles) {
    if (!e) return B2C_EINVAL;
    e->delay = cycles;
    drop_graphs(e);   // the delay is a kernel parameter
    return B2C_OK;
}

int b2c_blur_device(const uint8_t *d_in, int64_t in_pitch, uint8_t *d_out, int64_t out_pitch, int W, int H,
                    int CH, int reps, const b2c_opts *opts, void *stream) {
    // argument errors keep their own codes; only what CUDA refuses is B2C_ECUDA
    if (!d_in || !d_out || W <= 0 || H <= 0 || (CH != 1 && CH != 3) || reps < 0) {
        set_error("blur_device: bad arguments W=%d H=%d CH=%d reps=%d", W, H, CH, reps);
        return B2C_EINVAL;
    }
    if ((long long)W * CH > (1ll << 30)) {
        set_error("blur_device: row too wide");
        return B2C_ETOOBIG;
    }
    if (opts && opts->boundary == B2C_CUDA_COMPAT && opts->filter != B2C_GAUSSIAN) {
        set_error("cuda-compat is gaussian only (like cuda/cuda_convolution.cu)");
        return B2C_EINVAL;
    }
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
// start at zero).  Two calls whose GPU work overlaps — different streams — corrupt each other.
// Enqueueing is serialised inside the function (a mutex around the memset / copy / pass
// launches), so several host threads may call it on ONE stream; the calls then run one after
// the other in the order the threads got the lock.
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
    e.opts.kernel = B2C_KERNEL_STREAM;   // the kernels that need nothing but the two framed buffers and the pad
    e.opts.persistent = 0;
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

    e.pitch = round_up(kLeftPad + (long long)e.bw + kRightPad, 128);
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

    // One call's commands go into the stream as a block: two host threads interleaving their
    // memset / copy / launches on the same stream would run each other's passes on half-cleared scratch.
    static std::mutex enqueue_mu;
    std::lock_guard<std::mutex> enqueue_lock(enqueue_mu);
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
        rc = enqueue_pass(&e, ps, false, s);   // absolute epochs: nothing is replayed
        if (rc != B2C_OK) return rc;
    }
    CK(cudaMemcpy2DAsync(d_out, (size_t)out_pitch, e.base[e.cur], (size_t)e.pitch, (size_t)e.bw, (size_t)e.R,
                         cudaMemcpyDeviceToDevice, s));
    return B2C_OK;
}
