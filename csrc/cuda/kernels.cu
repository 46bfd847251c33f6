// Kernel instantiations and launch dispatch for libb200conv_core (sm_100a).
//   stream_kernel   : csrc/cuda/stencil_stream.cuh (the product path)
//   naive_kernel    : one thread per output byte, any filter, any boundary —
//                     the "minimum end-to-end slice" of SURVEY §7.3 step 2 and
//                     the only path for BOX/EDGE (fp32, reference op order).
//   compat kernels  : cuda/cuda_convolution.cu:9-47 observable semantics.
//   halo_push_kernel: standalone ring -> neighbour-frame copy + epoch flags
//                     (initial exchange, and the fallback when the block byte
//                     width is not 16-byte aligned).
#include <cuda_runtime.h>

#include <atomic>

#include "common.h"
#include "kernels.h"
#include "stencil_stream.cuh"

namespace b2c {

static std::atomic<long long> g_launches{0};
long long total_launches() { return g_launches.load(); }
void count_launch(long long n) { g_launches.fetch_add(n); }

// ------------------------------------------------------------------ stream

// op 0: launch, op 1: occupancy query (resident CTAs per SM) returned through *occ.
template <int K, int NW, int CH>
static cudaError_t stream_op_t(int op, const PassParams *p, bool check, cudaStream_t s, int *occ) {
    auto fn = check ? stream_kernel<K, NW, CH, true> : stream_kernel<K, NW, CH, false>;
    if (op == 1) return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, fn, kWarpsPerCta * 32, 0);
    if (op == 2) {   // make sure the function is loaded (see preload_stream_kernel)
        cudaFuncAttributes fa;
        return cudaFuncGetAttributes(&fa, fn);
    }
    const int warps = p->nbands * p->nchunks;
    const int ctas = (warps + kWarpsPerCta - 1) / kWarpsPerCta;
    // Programmatic dependent launch: the next pass may start scheduling CTAs while this one
    // drains; every warp executes griddepcontrol.wait before its first global access, so the
    // prologue (work-item mapping, masks, register clearing) overlaps the predecessor's tail.
    static const bool pdl = !(getenv("B200CONV_PDL") && atoi(getenv("B200CONV_PDL")) == 0);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(ctas);
    cfg.blockDim = dim3(kWarpsPerCta * 32);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, fn, *p);
    count_launch(1);
    return e != cudaSuccess ? e : cudaGetLastError();
}

static cudaError_t stream_op(int op, const PassParams *p, int k, int strip_bytes, int ch, bool check, cudaStream_t s,
                             int *occ) {
#define B2C_CASE(K_, NW_)                                                              \
    if (k == K_ && strip_bytes == NW_ * 4)                                             \
        return ch == 1 ? stream_op_t<K_, NW_, 1>(op, p, check, s, occ) : stream_op_t<K_, NW_, 3>(op, p, check, s, occ);
    B2C_CASE(1, 1) B2C_CASE(2, 1) B2C_CASE(4, 1) B2C_CASE(8, 1)
    B2C_CASE(1, 2) B2C_CASE(2, 2) B2C_CASE(4, 2) B2C_CASE(8, 2)
    B2C_CASE(1, 4) B2C_CASE(2, 4) B2C_CASE(4, 4)
#undef B2C_CASE
    return cudaErrorInvalidValue;
}

bool stream_config_ok(int k, int strip_bytes) {
    if (strip_bytes != 4 && strip_bytes != 8 && strip_bytes != 16) return false;
    if (k != 1 && k != 2 && k != 4 && k != 8) return false;
    return k * (strip_bytes / 4) <= 16;  // register budget: 4 state registers per word per stage
}

cudaError_t launch_stream(const PassParams &p, int k, int strip_bytes, int ch, bool check, cudaStream_t s) {
    return stream_op(0, &p, k, strip_bytes, ch, check, s, nullptr);
}

// With lazy module loading (the CUDA default) the first launch of a kernel loads it, and loading
// may wait for every kernel that is running on the device.  A pass variant that is first needed
// while the courier kernel spins (k = 2 or 1 at the end of a run) would therefore never load:
// the courier waits for that pass, the load waits for the courier.  So every variant a run uses
// is loaded before the courier starts.
cudaError_t preload_stream_kernel(int k, int strip_bytes, int ch, bool check) {
    return stream_op(2, nullptr, k, strip_bytes, ch, check, nullptr, nullptr);
}

int stream_occupancy(int k, int strip_bytes, int ch) {
    int occ = 0;
    if (stream_op(1, nullptr, k, strip_bytes, ch, false, nullptr, &occ) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return occ;
}

// -------------------------------------------------------------- persistent

template <int K, int NW, int CH>
static cudaError_t persist_op_t(int op, const PersistParams *q, int ctas, cudaStream_t s, int *occ) {
    auto fn = stream_persist_kernel<K, NW, CH>;
    if (op == 1) return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, fn, kWarpsPerCta * 32, 0);
    fn<<<ctas, kWarpsPerCta * 32, 0, s>>>(*q);
    count_launch(1);
    return cudaGetLastError();
}

static cudaError_t persist_op(int op, const PersistParams *q, int k, int strip_bytes, int ch, int ctas, cudaStream_t s,
                              int *occ) {
#define B2C_CASE(K_, NW_)                                                              \
    if (k == K_ && strip_bytes == NW_ * 4)                                             \
        return ch == 1 ? persist_op_t<K_, NW_, 1>(op, q, ctas, s, occ) : persist_op_t<K_, NW_, 3>(op, q, ctas, s, occ);
    B2C_CASE(1, 1) B2C_CASE(2, 1) B2C_CASE(4, 1) B2C_CASE(8, 1)
    B2C_CASE(1, 2) B2C_CASE(2, 2) B2C_CASE(4, 2) B2C_CASE(8, 2)
    B2C_CASE(1, 4) B2C_CASE(2, 4) B2C_CASE(4, 4)
#undef B2C_CASE
    return cudaErrorInvalidValue;
}

cudaError_t launch_stream_persist(const PersistParams &q, int k, int strip_bytes, int ch, int ctas, cudaStream_t s) {
    return persist_op(0, &q, k, strip_bytes, ch, ctas, s, nullptr);
}

int stream_persist_occupancy(int k, int strip_bytes, int ch) {
    int occ = 0;
    if (persist_op(1, nullptr, k, strip_bytes, ch, 0, nullptr, &occ) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return occ;
}

// ------------------------------------------------------------------- naive

// One iteration, one thread per output byte, on the engine's framed buffers.
// ZERO/TORUS come from the frame contents (zero / pushed by the ring copy).
template <bool INTEGER>
__global__ void naive_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, long long pitch,
                             int rows, int bw, int ch, NaiveWeights w) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;  // byte column -> threadIdx.x (coalesced)
    const int r = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= bw || r >= rows) return;
    const uint8_t *p = src + (long long)r * pitch + x;
    if (INTEGER) {
        int acc = 0;
#pragma unroll
        for (int i = -1; i <= 1; ++i)
#pragma unroll
            for (int j = -1; j <= 1; ++j) acc += w.iw[i + 1][j + 1] * p[i * pitch + j * ch];
        dst[(long long)r * pitch + x] = (uint8_t)(acc >> 4);
    } else {
        // fp32, row-major, multiply then add, no contraction: mpi_convolution.c:304-306
        float acc = 0.0f;
#pragma unroll
        for (int i = -1; i <= 1; ++i)
#pragma unroll
            for (int j = -1; j <= 1; ++j)
                acc = __fadd_rn(acc, __fmul_rn((float)p[i * pitch + j * ch], w.fw[i + 1][j + 1]));
        dst[(long long)r * pitch + x] = (uint8_t)acc;  // cvt.rzi: truncation
    }
}

cudaError_t launch_naive(const uint8_t *src, uint8_t *dst, long long pitch, int rows, int bw, int ch,
                         const NaiveWeights &w, cudaStream_t s) {
    dim3 block(128, 2), grid((bw + 127) / 128, (rows + 1) / 2);
    if (w.integer16)
        naive_kernel<true><<<grid, block, 0, s>>>(src, dst, pitch, rows, bw, ch, w);
    else
        naive_kernel<false><<<grid, block, 0, s>>>(src, dst, pitch, rows, bw, ch, w);
    count_launch(1);
    return cudaGetLastError();
}

// ------------------------------------------------------------- cuda-compat

// Grey: interior pixels only (guard of cuda_convolution.cu:17); the untouched
// border of the two ping-pong buffers gives the frozen original/zero ring.
__global__ void compat_grey_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, int W, int H) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y * blockDim.y + threadIdx.y;
    if (r < 1 || r >= H - 1 || c < 1 || c >= W - 1) return;
    const uint8_t *p = src + (long long)r * W + c;
    const int acc = p[-W - 1] + 2 * p[-W] + p[-W + 1] + 2 * p[-1] + 4 * p[0] + 2 * p[1] + p[W - 1] + 2 * p[W] + p[W + 1];
    dst[(long long)r * W + c] = (uint8_t)(acc >> 4);
}

// RGB: the `y < 3W-3` guard (:34) makes it a stencil over the linear pixel
// index p in [W+1, (H+1)W-4]; buffers carry two slack rows (the reference
// reads/writes past its allocations there).
__global__ void compat_rgb_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, int W,
                                  long long b_lo, long long b_hi) {
    const long long b = b_lo + (long long)blockIdx.x * blockDim.x + threadIdx.x;  // byte index
    if (b > b_hi) return;
    const long long pw = (long long)W * 3;
    const uint8_t *p = src + b;
    const int acc = p[-pw - 3] + 2 * p[-pw] + p[-pw + 3] + 2 * p[-3] + 4 * p[0] + 2 * p[3] + p[pw - 3] + 2 * p[pw] + p[pw + 3];
    dst[b] = (uint8_t)(acc >> 4);
}

cudaError_t launch_compat(const uint8_t *src, uint8_t *dst, int W, int H, int ch, cudaStream_t s) {
    if (ch == 1) {
        if (H < 3 || W < 3) return cudaSuccess;
        dim3 block(128, 2), grid((W + 127) / 128, (H + 1) / 2);
        compat_grey_kernel<<<grid, block, 0, s>>>(src, dst, W, H);
    } else {
        if (H < 3 || W < 2) return cudaSuccess;
        const long long b_lo = ((long long)W + 1) * 3, b_hi = ((long long)(H + 1) * W - 4) * 3 + 2;
        const long long n = b_hi - b_lo + 1;
        compat_rgb_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(src, dst, W, b_lo, b_hi);
    }
    count_launch(1);
    return cudaGetLastError();
}

// --------------------------------------------------------------- halo push

// One CTA per (direction, row slab).  Byte-exact copies (no alignment
// assumptions), then the last CTA of a direction publishes the epoch.
__global__ void halo_push_kernel(const PushParams q) {
    const int d = blockIdx.y;
    // After the convergence latch every pass is skipped on every rank; the ring pushes that
    // follow them must be skipped too or they would publish epochs nobody computed.
    if (q.converged) {
        const unsigned int c = *q.converged;
        if (c != 0 && c + 1 < q.interval) return;
    }
    const uint32_t base = q.epoch_base ? __ldcg(q.epoch_base) : 0u;
    if (q.epoch_word && blockIdx.x == 0 && d == 0 && threadIdx.x == 0) *q.epoch_word = q.post_epoch;
    if (!q.peer_dst[d]) return;
    if (q.wait_epoch_valid && threadIdx.x == 0) {
        // Do not overwrite a frame the neighbour may still be reading.
        const long long t0 = clock64();
        while ((int)(ld_acquire_sys(q.flags + d * 32) - (q.wait_epoch + base)) < 0) {
            if (clock64() - t0 > (20ll << 30)) {
                printf("b200conv: halo push wait timeout dir=%d want=%u\n", d, q.wait_epoch);
                __trap();
            }
        }
    }
    __syncthreads();
    const int dr = kDirDrDev(d), dc = kDirDcDev(d);
    const int F = q.depth;
    const long long fb = (long long)F * q.ch;
    const int sr0 = dr > 0 ? q.rows - F : 0, nr = dr ? F : q.rows;
    const long long sc0 = dc > 0 ? q.bw - fb : 0, nc = dc ? fb : q.bw;
    const int slabs = gridDim.x;
    const int per = (nr + slabs - 1) / slabs;
    const int r_lo = blockIdx.x * per, r_hi = min(nr, r_lo + per);
    for (int r = r_lo; r < r_hi; ++r) {
        const uint8_t *s = q.src + (long long)(sr0 + r) * q.pitch + sc0;
        uint8_t *t = q.peer_dst[d] + (long long)(sr0 + r - dr * q.rows) * q.pitch + (sc0 - dc * (long long)q.bw);
#ifdef B2C_DEBUG
        B2C_DBG_RANGE("push source", q.dbg_rank, s, nc, q.src - q.dbg_before, q.src + q.dbg_after);
        B2C_DBG_RANGE("push target", q.dbg_rank, t, nc, q.peer_dst[d] - q.dbg_before, q.peer_dst[d] + q.dbg_after);
#endif
        const bool al = (((uintptr_t)s | (uintptr_t)t) & 15) == 0;
        if (al) {
            const long long n16 = nc / 16;
            for (long long i = threadIdx.x; i < n16; i += blockDim.x)
                __stcg(reinterpret_cast<uint4 *>(t) + i, __ldcg(reinterpret_cast<const uint4 *>(s) + i));
            for (long long i = n16 * 16 + threadIdx.x; i < nc; i += blockDim.x) t[i] = s[i];
        } else {
            for (long long i = threadIdx.x; i < nc; i += blockDim.x) t[i] = s[i];
        }
    }
    // every thread's stores are ordered before the count-in at GPU scope; only the last CTA of a
    // direction pays for the system-scope fence in front of the flag (cumulative, see halo_publish)
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned old = atomicAdd(q.counters + d * 32, 1u);
        if (old + 1 == gridDim.x) {
            q.counters[d * 32] = 0;
            st_release_sys(q.peer_flag[d], q.post_epoch + base);   // its fence is also the acquire side of the atomic
        }
    }
}

cudaError_t launch_halo_push(const PushParams &q, cudaStream_t s) {
    int slabs = (int)ceil_div(q.rows, 64);
    if (slabs > 64) slabs = 64;
    if (slabs < 1) slabs = 1;
    halo_push_kernel<<<dim3(slabs, 8), 256, 0, s>>>(q);
    count_launch(1);
    return cudaGetLastError();
}

// Blocks the stream until every neighbour has published `epoch` (used in front of kernels
// that do not wait by themselves: the byte-per-thread naive kernel).
__global__ void halo_wait_kernel(const uint32_t *flags, unsigned mask, uint32_t epoch) {
    const int d = threadIdx.x;
    if (d >= 8 || !((mask >> d) & 1u)) return;
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(flags + d * 32) - epoch) < 0) {
        if (clock64() - t0 > (20ll << 30)) {
            printf("b200conv: halo wait timeout dir=%d want=%u\n", d, epoch);
            __trap();
        }
    }
}

cudaError_t launch_halo_wait(const uint32_t *flags, unsigned mask, uint32_t epoch, cudaStream_t s) {
    halo_wait_kernel<<<1, 32, 0, s>>>(flags, mask, epoch);
    count_launch(1);
    return cudaGetLastError();
}

// ------------------------------------------------------------------- misc

__global__ void set_u32_kernel(unsigned int *p, unsigned int v) { *p = v; }

// Convergence bookkeeping after a CHECK pass: publish this rank's "changed"
// bit to every rank, wait for all ranks' bits, and latch `converged` when the
// OR is zero.  One thread; runs on the pass stream right after the CHECK pass.
//
// Slots are double-buffered by the parity of the check number.  With a single
// set a fast rank could leave check s (it has seen everybody's tag s), run the
// next pass, enter check s+1 and overwrite its word in a slow rank's array
// before that rank has read tag s there — the slow rank would then wait for a
// tag that no longer exists.  A rank can only write tag s+2 after it has seen
// every rank's tag s+1, and a rank publishes s+1 after it finished reading s,
// so two sets are enough.  (Found with the host model, csrc/core/halo_emul.cpp.)
__global__ void converge_kernel(const ConvergeParams c) {
    // One warp: lane k (+32) talks to rank k, so the N stores and the N polls each go out together
    // instead of one system-scope round trip after the other.
    const int lane = threadIdx.x;
    unsigned mine = 0;
    if (lane == 0) {
        mine = *c.changed;
        *c.changed = 0;
    }
    mine = __shfl_sync(kFull, mine, 0);
    // slot value = 2*seq + bit so that a slot from an older check never matches
    const unsigned tag = c.seq * 2u;
    const int set = (int)(c.seq & 1u) * kMaxRanks;
    for (int k = lane; k < c.nranks; k += 32) st_release_sys(c.peer_slots[k] + set + c.rank, tag | (mine ? 1u : 0u));
    unsigned any = 0;
    for (int k = lane; k < c.nranks; k += 32) {
        const long long t0 = clock64();
        unsigned v;
        while (((v = ld_relaxed_sys(c.slots + set + k)) & ~1u) != tag) {
            if (clock64() - t0 > (20ll << 30)) {
                printf("b200conv: convergence wait timeout rank=%d peer=%d\n", c.rank, k);
                __trap();
            }
        }
        any |= v & 1u;
    }
    any = __any_sync(kFull, any != 0);
    if (lane == 0 && !any && !*c.converged) {
        *c.converged_at = c.iters_after;
        __threadfence();
        *c.converged = c.check_no;
    }
}

cudaError_t launch_converge(const ConvergeParams &c, cudaStream_t s) {
    converge_kernel<<<1, 32, 0, s>>>(c);   // exactly one warp
    count_launch(1);
    return cudaGetLastError();
}

cudaError_t launch_set_u32(unsigned int *p, unsigned int v, cudaStream_t s) {
    set_u32_kernel<<<1, 1, 0, s>>>(p, v);
    count_launch(1);
    return cudaGetLastError();
}

__global__ void add_u32_kernel(unsigned int *p, unsigned int v) { *p += v; }

cudaError_t launch_add_u32(unsigned int *p, unsigned int v, cudaStream_t s) {
    add_u32_kernel<<<1, 1, 0, s>>>(p, v);
    count_launch(1);
    return cudaGetLastError();
}

// Start barrier on the device (replaces MPI_Barrier in front of the timer, mpi_convolution.c:151):
// lane k (+32) stores the sequence number into rank k's array and polls word k of its own, so a
// rank's stream passes this kernel within a few microseconds of the last rank reaching it.
__global__ void barrier_kernel(const BarrierParams b) {
    const int lane = threadIdx.x;
    const long long t0 = clock64();
    for (int k = lane; k < b.nranks; k += 32) st_release_sys(b.peer_slots[k] + b.rank, b.seq);
    for (int k = lane; k < b.nranks; k += 32) {
        while ((int)(ld_relaxed_sys(b.slots + k) - b.seq) < 0) {
            if (clock64() - t0 > (20ll << 30)) {
                printf("b200conv: start barrier timeout rank=%d peer=%d\n", b.rank, k);
                __trap();
            }
        }
        (void)ld_acquire_sys(b.slots + k);
    }
    __syncwarp();
    if (lane == 0 && b.skew_cycles) *b.skew_cycles = (unsigned long long)(clock64() - t0);
}

__global__ void courier_kernel(const CourierParams c) { courier_loop(c, threadIdx.x, 0u); }

// ---- NVLS: reductions in the switch (replaces MPI_Allreduce / MPI_Barrier, mpi_convolution.c:151,264-275)

__device__ __forceinline__ void multimem_red_add(uint32_t *mc, uint32_t v) {
    asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}

// Convergence OR over all ranks: ONE multimem.red (bit + arrival count in one word) and ONE poll of
// the local replica, instead of N peer stores and N polls.
__global__ void converge_nvls_kernel(const ConvergeParams c, const NvlsParams n) {
    if (threadIdx.x != 0) return;
    const unsigned mine = *c.changed;
    *c.changed = 0;
    const int slot = kNvlsConvergeWord + (int)(c.seq & 1u);
    multimem_red_add(n.mc + slot, 0x10000u + (mine ? 1u : 0u));
    const long long t0 = clock64();
    uint32_t v;
    while (((v = ld_relaxed_sys(n.local + slot)) >> 16) != (uint32_t)n.nranks) {
        if (clock64() - t0 > (20ll << 30)) {
            printf("b200conv: NVLS convergence wait timeout rank=%d have=%#x\n", n.rank, v);
            __trap();
        }
    }
    // every rank's contribution to this replica has arrived: clear it for check seq+2 (nobody can be
    // that far ahead — check seq+1 needs this rank's contribution, which comes after this store)
    n.local[slot] = 0;
    __threadfence_system();
    if (!(v & 0xffffu) && !*c.converged) {
        *c.converged_at = c.iters_after;
        __threadfence();
        *c.converged = c.check_no;
    }
}

cudaError_t launch_converge_nvls(const ConvergeParams &c, const NvlsParams &n, cudaStream_t s) {
    converge_nvls_kernel<<<1, 32, 0, s>>>(c, n);
    count_launch(1);
    return cudaGetLastError();
}

__global__ void barrier_nvls_kernel(const NvlsParams n, const uint32_t seq, unsigned long long *skew_cycles) {
    if (threadIdx.x != 0) return;
    const long long t0 = clock64();
    multimem_red_add(n.mc + kNvlsBarrierWord, 1u);
    const uint32_t want = seq * (uint32_t)n.nranks;
    while ((int)(ld_relaxed_sys(n.local + kNvlsBarrierWord) - want) < 0) {
        if (clock64() - t0 > (20ll << 30)) {
            printf("b200conv: NVLS start barrier timeout rank=%d\n", n.rank);
            __trap();
        }
    }
    (void)ld_acquire_sys(n.local + kNvlsBarrierWord);
    if (skew_cycles) *skew_cycles = (unsigned long long)(clock64() - t0);
}

cudaError_t launch_barrier_nvls(const NvlsParams &n, uint32_t seq, unsigned long long *skew_cycles, cudaStream_t s) {
    barrier_nvls_kernel<<<1, 32, 0, s>>>(n, seq, skew_cycles);
    count_launch(1);
    return cudaGetLastError();
}

cudaError_t launch_courier(const CourierParams &c, cudaStream_t s) {
    courier_kernel<<<1, 32, 0, s>>>(c);
    count_launch(1);
    return cudaGetLastError();
}

cudaError_t launch_barrier(const BarrierParams &b, cudaStream_t s) {
    barrier_kernel<<<1, 32, 0, s>>>(b);
    count_launch(1);
    return cudaGetLastError();
}

}  // namespace b2c

// The work-item order of the streaming kernels, callable on the host so that a CPU test can check
// that it visits every (band, chunk) exactly once for every plan (a gap or a duplicate would leave a
// per-direction arrival counter short and hang the neighbours).
extern "C" int b2c_debug_map_work(int id, int nbands, int nchunks, int nbw, int nbe, int *band, int *chunk) {
    if (id < 0 || nbands <= 0 || nchunks <= 0 || id >= nbands * nchunks || !band || !chunk) return B2C_EINVAL;
    b2c::map_work(id, nbands, nchunks, nbw, nbe, *band, *chunk);
    return B2C_OK;
}
