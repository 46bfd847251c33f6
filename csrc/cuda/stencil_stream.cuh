// Register-pipelined, temporally-blocked 3x3 binomial stencil for sm_100a.
//
// Replaces kernelConvolutionforGrey/RGB (cuda/cuda_convolution.cu:9-47) and the
// per-iteration host loop (:66-87), and — with the halo block enabled — the
// Isend/Irecv/Wait schedule of mpi/mpi_convolution.c:156-240.  Not a port: the
// reference runs one thread per pixel with fp64 FMAs; here
//
//   * one warp streams a 32*SB-byte wide column band top to bottom; each lane
//     owns an SB-byte strip (SB = 4/8/16) and keeps, for each of K fused
//     iterations, two rows of separable partial sums in registers.  One HBM
//     read + one HBM write per K iterations.
//   * arithmetic is SIMD-within-register on 16-bit lanes: even bytes and odd
//     bytes of a word live in two registers (E,O); the +-CH horizontal taps are
//     the other-parity register of the same / adjacent word (a 16-bit funnel
//     shift), so grey (CH=1) and rgb (CH=3) are one code path.
//   * exact integer: out = (sum w*p) >> 4.  Lanes alternate between scale 1 and
//     scale 16 from stage to stage so the >>4 is a single mask (even stages:
//     v & 0x0FF0 == 16*floor(v/16)) or a single byte-permute (odd stages:
//     high byte of the 16-bit lane); max lane value 16*4080 = 65280 fits.
//   * left/right neighbours come from adjacent lanes by SHFL; the 32*SB window
//     loses K*CH valid bytes per side (bands overlap by that much).
//   * ZERO boundary: cells outside the global image are re-zeroed after every
//     fused stage (warp-uniform EDGE variant, only for warps touching an edge).
//   * multi-GPU: warps that own part of the block's boundary ring store those
//     rows/strips straight into the neighbours' halo frames through peer
//     (NVLink) pointers, then the last such warp per direction publishes an
//     epoch flag with st.release.sys; consumers spin with ld.acquire.sys.
//     Ring warps are scheduled first so the interior overlaps the transfer.
#pragma once
#include <cstdint>
#include <type_traits>

#include "b200conv.h"

namespace b2c {

constexpr int kMaxHaloDepth = 32;  // deepest halo F a rank may keep (iterations between two exchanges)
constexpr int kFrameRows = 40;  // halo frame rows above/below the block: max F (32) + the 3x loop overrun + slack
constexpr int kLeftPad = 128;   // frame bytes left of the block: round_up(F*3, 16) + the widest window overhang (32)
constexpr int kRightPad = 128;  // frame bytes right of the block (before rounding the pitch up to 128)
constexpr int kWarpsPerCta = 4;
#ifndef B2C_MIN_CTAS
#define B2C_MIN_CTAS 4   // resident CTAs per SM the register allocator must allow (<= 128 regs/thread)
#endif
#ifndef B2C_FAST_UNROLL2
#define B2C_FAST_UNROLL2 0   // two row groups per trip of the interior loop: 3 % fewer instructions, measured 4 % slower
#endif
constexpr unsigned kFull = 0xffffffffu;

// What a pass does with its neighbours.  A rank keeps an F-deep halo and exchanges it once per
// *group* of passes (F iterations): the first pass of a group waits for the neighbours' rings, the
// passes in between recompute the shrinking part of the halo themselves (their block is extended
// into the frame), the last pass pushes the new F-deep ring.  Because the in-between passes write
// the frames of both buffers, a pusher must also know that the neighbour is done with the frame it
// is about to overwrite: the readers of the second-to-last pass post an "ack" epoch for that.
enum : int {
    kHaloWaitData = 1,  // ring warps wait for the neighbours' data epoch before reading a frame
    kHaloPostAck = 2,   // ring warps count in and publish the ack epoch (no stores)
    kHaloWaitAck = 4,   // ring warps wait for the neighbours' ack epoch before pushing
    kHaloPush = 8,      // ring warps store their ring into the neighbours' frames and publish the data epoch
};

struct HaloParams {
    int enabled;            // any neighbour at all
    int mode;               // kHalo* bits of this pass
    int fused;              // 1: kHaloPush stores come from this kernel (aligned widths); 0: halo_push_kernel follows
    int flag_fence;         // how the last arriver orders the ring before the flag: 0 fence.sc.sys + st.release.sys,
                            // 1 st.release.sys alone (default), 2 fence.acq_rel.gpu + st.relaxed.sys (experiment)
    int depth;              // ring depth F (rows, and F*CH byte columns)
    uint32_t epoch;         // group epoch E, relative to *epoch_base: data/ack flags must be >= E, E+1 is published
    const uint32_t *epoch_base;  // device word holding the epoch at the start of the run (null: 0)
    uint8_t *peer_dst[8];   // neighbour's destination buffer, interior origin (or null)
    uint32_t *peer_flag[8]; // neighbour's data-flag word that *we* write for direction d
    uint32_t *peer_ack[8];  // neighbour's ack word that *we* write for direction d
    uint32_t *flags;        // our data-flag words, index = direction the data came from
    uint32_t *acks;         // our ack words, index = direction of the neighbour that acknowledged
    uint32_t *counters;     // our per-direction arrival counters
    uint32_t *ready;        // when set: the last arriver only stores the epoch here (index d*32 data, (8+d)*32 ack,
                            // GPU scope) and the courier kernel forwards it to the neighbour (see kernels.cu)
    unsigned long long *prof;  // opt-in wait/publish cycle counters (null: off), layout in engine.cu
    int expected[8];        // ring warps per direction
    int nbw, nbe;           // bands holding the west / east ring
    long long delay;        // test hook: spin this many clocks before each wait
};

struct alignas(16) PassParams {
    const uint8_t *src;  // interior origin of the source buffer
    uint8_t *dst;        // interior origin of the destination buffer
    long long pitch;
    int rows;            // R
    int bw;              // block byte width C*CH
    int nbands, nchunks, chunk_rows;
    int vw;              // valid bytes per band
    int hlo;             // window bytes left of the valid band
    int zero_n, zero_s, zero_w, zero_e;  // outside that side is outside the image (ZERO mode)
    unsigned int *changed;  // CHECK: OR-accumulated "last stage changed a byte"
    const unsigned int *converged;  // check number at which the fixed point was detected (0: not yet), or null
    unsigned int interval;          // check interval this pass belongs to (1-based); skipped iff *converged + 1 < interval
    int org_x, org_y;       // position of src/dst (0,0) inside the allocation (bytes, rows): TMA coordinates
    int raster;             // work order: 0 ring-first, 1 chunk-major, 2 ring-LAST (ring-first reversed)
#ifdef B2C_DEBUG
    const uint8_t *dbg_src_lo, *dbg_src_hi;   // [lo, hi) of the source / destination allocation
    uint8_t *dbg_dst_lo, *dbg_dst_hi;
    long long dbg_peer_before, dbg_peer_after;  // bytes of a peer allocation before / after its interior origin
    int dbg_rank;
#endif
    HaloParams halo;
};

#ifdef B2C_DEBUG
// Debug build (python build.py --debug): every global address the streaming kernels form is checked
// against the allocation it must fall into, epochs are checked for monotonicity; a violation prints
// rank / what / address and traps.  Stands in for compute-sanitizer where that tool is not available.
static __device__ __noinline__ void dbg_fail(const char *what, int rank, const void *p, const void *lo, const void *hi) {
    printf("b200conv DEBUG: rank %d: %s address %p outside [%p, %p)\n", rank, what, p, lo, hi);
    __trap();
}
#define B2C_DBG_RANGE(what, rank, ptr, nbytes, lo, hi)                                              \
    do {                                                                                            \
        if ((const uint8_t *)(ptr) < (const uint8_t *)(lo) || (const uint8_t *)(ptr) + (nbytes) > (const uint8_t *)(hi)) \
            b2c::dbg_fail(what, rank, ptr, lo, hi);                                                 \
    } while (0)
#else
#define B2C_DBG_RANGE(what, rank, ptr, nbytes, lo, hi) do { } while (0)
#endif

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Poll without the L1 invalidate an acquire load drags along; one ld_acquire_sys after the value
// has shown up gives the ordering.
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Streaming loads/stores: L2 only, nothing is reused from L1.
template <int NW>
__device__ __forceinline__ void load_strip(const uint8_t *p, uint32_t (&w)[NW]) {
    if constexpr (NW == 4) {
        uint4 v = __ldcg(reinterpret_cast<const uint4 *>(p));
        w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
    } else if constexpr (NW == 2) {
        uint2 v = __ldcg(reinterpret_cast<const uint2 *>(p));
        w[0] = v.x; w[1] = v.y;
    } else {
        w[0] = __ldcg(reinterpret_cast<const uint32_t *>(p));
    }
}
template <int NW>
__device__ __forceinline__ void store_strip(uint8_t *p, const uint32_t (&w)[NW]) {
    if constexpr (NW == 4) {
        __stcg(reinterpret_cast<uint4 *>(p), make_uint4(w[0], w[1], w[2], w[3]));
    } else if constexpr (NW == 2) {
        __stcg(reinterpret_cast<uint2 *>(p), make_uint2(w[0], w[1]));
    } else {
        __stcg(reinterpret_cast<uint32_t *>(p), w[0]);
    }
}
template <int NW>
__device__ __forceinline__ void store_strip_partial(uint8_t *p, const uint32_t (&w)[NW], int nbytes) {
#pragma unroll
    for (int i = 0; i < NW; ++i)
#pragma unroll
        for (int b = 0; b < 4; ++b)
            if (i * 4 + b < nbytes) p[i * 4 + b] = (uint8_t)(w[i] >> (8 * b));
}

// Horizontal 1-2-1 over byte stride CH on (E,O) 16-bit lanes, as two pairwise sums
//   t[x] = p[x] + p[x+CH],   h[x] = t[x-CH] + t[x]
// (4 adds per word instead of 6).  CH is odd, so a tap always lives in the other-parity
// register: either that register itself or a 16-bit funnel shift of two neighbours
// (ptxas folds shift+add into one LEA.HI).  Lane edges come from adjacent lanes by SHFL.
template <int NW, int CH>
__device__ __forceinline__ void hsum(const uint32_t (&E)[NW], const uint32_t (&O)[NW],
                                     uint32_t (&hE)[NW], uint32_t (&hO)[NW]) {
    uint32_t tE[NW], tO[NW];
    if constexpr (CH == 1) {
        const uint32_t En = __shfl_down_sync(kFull, E[0], 1);
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            tE[i] = E[i] + O[i];                                                          // p0+p1, p2+p3
            tO[i] = __funnelshift_l(E[i], (i + 1 < NW) ? E[i + 1] : En, 16) + O[i];      // p1+p2, p3+p4
        }
        const uint32_t tOp = __shfl_up_sync(kFull, tO[NW - 1], 1);
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            hE[i] = __funnelshift_l(i ? tO[i - 1] : tOp, tO[i], 16) + tE[i];              // t[-1]+t[0], t[1]+t[2]
            hO[i] = tE[i] + tO[i];                                                        // t[0]+t[1], t[2]+t[3]
        }
    } else {
        const uint32_t On = __shfl_down_sync(kFull, O[0], 1);
        const uint32_t En = __shfl_down_sync(kFull, E[0], 1);
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            tE[i] = __funnelshift_l(O[i], (i + 1 < NW) ? O[i + 1] : On, 16) + E[i];      // p0+p3, p2+p5
            tO[i] = ((i + 1 < NW) ? E[i + 1] : En) + O[i];                                // p1+p4, p3+p6
        }
        const uint32_t tOp = __shfl_up_sync(kFull, tO[NW - 1], 1);
        const uint32_t tEp = __shfl_up_sync(kFull, tE[NW - 1], 1);
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            hE[i] = (i ? tO[i - 1] : tOp) + tE[i];                                        // t[-3]+t[0], t[-1]+t[2]
            hO[i] = __funnelshift_l(i ? tE[i - 1] : tEp, tE[i], 16) + tO[i];              // t[-2]+t[1], t[0]+t[3]
        }
    }
}

// Per-stage sliding window of separable partial sums.  Three register slots per
// (stage, word, parity) rotate roles every input row:
//   hp = h(r-1)            sp = h(r-2)+h(r-1)            tmp = scratch for h(r)
// In phase P the roles sit in slots (3-P)%3, +1, +2; after hp += h the slot that held
// hp holds the new sp, tmp holds the new hp, and the old sp slot is free — so with the
// row loop unrolled by three no register-to-register moves are needed at all.
template <int K, int NW>
struct PipeState {
    uint32_t E[3][K][NW], O[3][K][NW];
};

// One warp's stream.
//   GENERAL = false: interior fast path — no boundary masks, no partial strips, no halo
//                    pushes; the only per-row predicate is "does this row get stored".
//   GENERAL = true : warps that touch an image edge (ZERO boundary re-imposed after every
//                    stage), own a partial last strip, or hold part of the halo ring.
template <int K, int NW, int CH, bool GENERAL, bool CHECK, class Feed>
__device__ __forceinline__ void stream_band(const PassParams &p, Feed &feed, const int band, const int chunk,
                                            const int lane, const bool do_push,
                                            const bool top, const bool bot, const bool west, const bool east) {
    constexpr int SB = NW * 4;
    const int x0 = band * p.vw - p.hlo;      // window start, block-local byte column
    const int xs = x0 + lane * SB;           // this lane's strip
    const int out_r0 = chunk * p.chunk_rows;
    const int out_r1 = chunk == p.nchunks - 1 ? p.rows : out_r0 + p.chunk_rows;  // the last chunk takes the remainder
    const int in_r0 = out_r0 - K;
    const int in_r1 = out_r1 + K;            // exclusive

    // Which strips are stored.
    const int band_lo = band * p.vw, band_hi = min(band_lo + p.vw, p.bw);
    const bool st_any = xs >= band_lo && xs < band_hi;
    const int st_bytes = st_any ? min(SB, p.bw - xs) : 0;  // SB = full strip
    // Row r is stored (as output row r-K) iff 0 <= r - first_r < nstore; lanes that store
    // nothing get nstore = 0, so one unsigned compare per row decides everything.
    const int first_r = out_r0 + K;
    const unsigned nstore = st_any ? (unsigned)(out_r1 - out_r0) : 0u;

    // Column masks (GENERAL): lanes of bytes outside the image become 0.
    uint32_t mE[NW], mO[NW];
    if constexpr (GENERAL) {
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            uint32_t me = 0, mo = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int c = xs + i * 4 + b;
                const bool ok = (c >= 0 || !p.zero_w) && (c < p.bw || !p.zero_e);
                const uint32_t m = ok ? (0xffffu << ((b >> 1) * 16)) : 0u;
                if (b & 1) mo |= m; else me |= m;
            }
            mE[i] = me; mO[i] = mo;
        }
    }

    PipeState<K, NW> st;
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int j = 0; j < K; ++j)
#pragma unroll
            for (int i = 0; i < NW; ++i) st.E[q][j][i] = st.O[q][j][i] = 0;

    uint32_t chg = 0;
    uint32_t pE[NW], pO[NW];  // CHECK: last stage's input lanes seen one step earlier
    if constexpr (CHECK) {
#pragma unroll
        for (int i = 0; i < NW; ++i) pE[i] = pO[i] = 0;
    }

    // Output pointer of the row that input row r produces (row r-K); advanced by the loop.
    uint8_t *dptr = p.dst + (long long)(in_r0 - K) * p.pitch + xs;

    // One input row through all K stages.  PH = r mod 3 relative to the first row.
    auto row = [&](auto phase_c, const uint32_t (&w)[NW], const int r, uint8_t *const dst) {
        constexpr int PH = decltype(phase_c)::value;
        constexpr int HP = (3 - PH) % 3, SP = (HP + 1) % 3, TM = (HP + 2) % 3;
        uint32_t E[NW], O[NW];
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            E[i] = w[i] & 0x00ff00ffu;
            O[i] = __byte_perm(w[i], 0u, 0x4341);
        }
        uint32_t cE[NW], cO[NW];  // CHECK: row r-K of iteration K-1 (input of the last stage)
#pragma unroll
        for (int j = 0; j < K; ++j) {
            if constexpr (CHECK) {
                if (j == K - 1) {
#pragma unroll
                    for (int i = 0; i < NW; ++i) { cE[i] = pE[i]; cO[i] = pO[i]; pE[i] = E[i]; pO[i] = O[i]; }
                }
            }
            hsum<NW, CH>(E, O, st.E[TM][j], st.O[TM][j]);
#pragma unroll
            for (int i = 0; i < NW; ++i) {
                // v of row r-j-1 (16-bit lanes).  The even lanes use a three-input add of the old
                // values (IADD3, ALU pipe), the odd lanes add the freshly updated sum (ptxas picks
                // IMAD.IADD, FMA-heavy pipe): keeps the two integer pipes evenly loaded.
                E[i] = st.E[SP][j][i] + st.E[HP][j][i] + st.E[TM][j][i];
                st.E[HP][j][i] += st.E[TM][j][i];   // h(r-1) + h(r): next row's sp
                st.O[HP][j][i] += st.O[TM][j][i];
                O[i] = st.O[SP][j][i] + st.O[HP][j][i];
            }
            if (j + 1 < K) {
                // stage j output = row r-j-1 after iteration j+1, kept as lanes for stage j+1
#pragma unroll
                for (int i = 0; i < NW; ++i) {
                    if ((j & 1) == 0) {  // inputs had scale 1 -> lanes become 16*out
                        E[i] &= 0x0ff00ff0u;
                        O[i] &= 0x0ff00ff0u;
                    } else {             // inputs had scale 16 -> the high byte of a lane is out
                        E[i] = __byte_perm(E[i], 0u, 0x4341);
                        O[i] = __byte_perm(O[i], 0u, 0x4341);
                    }
                }
                if constexpr (GENERAL) {
                    const int srow = r - j - 1;
                    const bool rok = (srow >= 0 || !p.zero_n) && (srow < p.rows || !p.zero_s);
#pragma unroll
                    for (int i = 0; i < NW; ++i) {
                        E[i] = rok ? (E[i] & mE[i]) : 0u;
                        O[i] = rok ? (O[i] & mO[i]) : 0u;
                    }
                }
            }
        }

        // Final stage output: row r-K, packed back to bytes.
        if ((unsigned)(r - first_r) < nstore) {
            uint32_t o[NW];
#pragma unroll
            for (int i = 0; i < NW; ++i) {
                if ((K - 1) & 1) o[i] = __byte_perm(E[i], O[i], 0x7351);
                else o[i] = __byte_perm(E[i] >> 4, O[i] >> 4, 0x6240);
            }
            B2C_DBG_RANGE("local store", p.dbg_rank, dst, st_bytes, p.dbg_dst_lo, p.dbg_dst_hi);
            if (!GENERAL || st_bytes == SB) store_strip<NW>(dst, o);
            else store_strip_partial<NW>(dst, o, st_bytes);

            if constexpr (CHECK) {
#pragma unroll
                for (int i = 0; i < NW; ++i) {
                    // the last stage's input row, repacked to bytes (its lanes have scale 16 when K-1 is odd)
                    const uint32_t q = ((K - 1) & 1) ? __byte_perm(cE[i] >> 4, cO[i] >> 4, 0x6240)
                                                     : __byte_perm(cE[i], cO[i], 0x6240);
                    uint32_t d = q ^ o[i];
                    const int nb = st_bytes - i * 4;  // stored bytes of this word
                    if (nb < 4) d &= nb <= 0 ? 0u : (0xffffffffu >> (8 * (4 - nb)));
                    chg |= d;
                }
            }

            if constexpr (GENERAL) {
                if (do_push) {
                    const HaloParams &h = p.halo;
                    const int F = h.depth, orow = r - K;
                    const long long doff = dst - p.dst;
                    const long long up = (long long)p.rows * p.pitch, dn = -up;
                    const bool rt = orow < F, rb = orow >= p.rows - F;
                    const bool cw = xs < F * CH, ce = xs + SB > p.bw - F * CH;
                    auto put = [&](const int d, const long long off) {
                        uint8_t *const q = h.peer_dst[d] + off;
                        B2C_DBG_RANGE("peer store", p.dbg_rank, q, SB, h.peer_dst[d] - p.dbg_peer_before,
                                      h.peer_dst[d] + p.dbg_peer_after);
                        store_strip<NW>(q, o);
                    };
                    if (top && rt && h.peer_dst[B2C_N]) put(B2C_N, doff + up);
                    if (bot && rb && h.peer_dst[B2C_S]) put(B2C_S, doff + dn);
                    if (west && cw) {
                        if (h.peer_dst[B2C_W]) put(B2C_W, doff + p.bw);
                        if (top && rt && h.peer_dst[B2C_NW]) put(B2C_NW, doff + up + p.bw);
                        if (bot && rb && h.peer_dst[B2C_SW]) put(B2C_SW, doff + dn + p.bw);
                    }
                    if (east && ce) {
                        if (h.peer_dst[B2C_E]) put(B2C_E, doff - p.bw);
                        if (top && rt && h.peer_dst[B2C_NE]) put(B2C_NE, doff + up - p.bw);
                        if (bot && rb && h.peer_dst[B2C_SE]) put(B2C_SE, doff + dn - p.bw);
                    }
                }
            }
        }
    };

    // Rows arrive three at a time from the feed (register-prefetched LDGs or a TMA-filled
    // shared-memory ring); rows past in_r1 carry junk whose outputs are never stored.
    feed.start(x0, in_r0, in_r1);
    const long long pitch = p.pitch;
    // The feed exposes the three rows of a group as a/b/c (refill hooks let a feed reload a row right
    // after it has been unpacked; neither current feed needs them).
    auto group3 = [&](const int r) {
        feed.group(r);
        row(std::integral_constant<int, 0>{}, feed.a, r, dptr);
        feed.refill_a();
        row(std::integral_constant<int, 1>{}, feed.b, r + 1, dptr + pitch);
        feed.refill_b();
        row(std::integral_constant<int, 2>{}, feed.c, r + 2, dptr + 2 * pitch);
        feed.refill_c();
        dptr += 3 * pitch;
    };
    if constexpr (!GENERAL && !CHECK && B2C_FAST_UNROLL2) {
        // Experiment (off): two groups per trip.  ptxas closes a three-row trip with ~27 register
        // copies (the rotating window does not land in the registers it started from) plus the 12
        // copies of the prefetched rows; over six rows the same copies are paid once (1236
        // instead of 1278 instructions per six rows) — but the longer body brings 9 local-memory
        // accesses into the loop and ran 110.5 instead of 106.3 ms per 1000 iterations on 16384² rgb.
        int r = in_r0;
#pragma unroll 1
        for (; r + 3 < in_r1; r += 6) {
            group3(r);
            group3(r + 3);
        }
        if (r < in_r1) group3(r);
    } else {
#pragma unroll 1
        for (int r = in_r0; r < in_r1; r += 3) group3(r);
    }
    if constexpr (CHECK) {
        chg = __reduce_or_sync(kFull, chg);
        if (lane == 0 && chg) atomicOr(p.changed, 1u);
    }
}

// Work order of a pass.  Ring-first (passes that push or acknowledge): top chunk, bottom chunk,
// then the west/east bands of the middle chunks, then the interior — so the neighbours get their
// epoch early.  Raster (every other pass): chunk-major, which is the order in which the
// persistent kernel's tile dependencies resolve without waiting.
__host__ __device__ __forceinline__ void map_work(int id, int NB, int NC, int nbw, int nbe, int &band, int &chunk) {
    if (NC <= 2 || NB <= nbw + nbe) {  // everything is ring (or raster was asked for: nbw = NB)
        chunk = id / NB; band = id % NB; return;
    }
    if (id < NB) { chunk = 0; band = id; return; }
    id -= NB;
    if (id < NB) { chunk = NC - 1; band = id; return; }
    id -= NB;
    const int side = nbw + nbe, mid = NC - 2;
    if (id < side * mid) {
        chunk = 1 + id / side;
        const int s = id % side;
        band = s < nbw ? s : NB - nbe + (s - nbw);
        return;
    }
    id -= side * mid;
    const int inner = NB - side;
    chunk = 1 + id / inner;
    band = nbw + id % inner;
}

// Feed #1: per-lane 16/8/4-byte LDGs, three rows held in registers ahead of use.
template <int NW>
struct LdgFeed {
    const PassParams &p;
    const int lane;
    const uint8_t *src;
    int in_r1;
    uint32_t a[NW], b[NW], c[NW];      // the three rows of the current group
    uint32_t na[NW], nb[NW], nc[NW];   // the next group's rows, in flight
    __device__ __forceinline__ LdgFeed(const PassParams &p_, int lane_) : p(p_), lane(lane_) {}
    __device__ __forceinline__ void start(int x0, int in_r0, int in_r1_) {
        in_r1 = in_r1_;
        src = p.src + (long long)in_r0 * p.pitch + x0 + lane * NW * 4;
        // rows up to in_r1+2 are read without a guard: the frame has kFrameRows >= F+K+4 rows
        B2C_DBG_RANGE("feed first row", p.dbg_rank, src, NW * 4, p.dbg_src_lo, p.dbg_src_hi);
        B2C_DBG_RANGE("feed last row", p.dbg_rank, src + (long long)((in_r1 - in_r0 + 2) / 3 * 3 - 1) * p.pitch, NW * 4,
                      p.dbg_src_lo, p.dbg_src_hi);
        load_strip<NW>(src, na);
        load_strip<NW>(src + p.pitch, nb);
        load_strip<NW>(src + 2 * p.pitch, nc);
    }
    // Hand out the prefetched group and put the next three loads in flight at once (measured:
    // issuing them one by one after each row is 18 % slower).
    __device__ __forceinline__ void group(int r) {
#pragma unroll
        for (int i = 0; i < NW; ++i) { a[i] = na[i]; b[i] = nb[i]; c[i] = nc[i]; }
        src += 3 * p.pitch;
        if (r + 3 < in_r1) {   // one guard per group keeps the reads inside the frame
            load_strip<NW>(src, na);
            load_strip<NW>(src + p.pitch, nb);
            load_strip<NW>(src + 2 * p.pitch, nc);
        }
    }
    __device__ __forceinline__ void refill_a() {}
    __device__ __forceinline__ void refill_b() {}
    __device__ __forceinline__ void refill_c() {}
};


// Ring warps before they touch a halo frame: lanes 0-7 poll the data flag of direction `lane`,
// lanes 8-15 the ack word of direction `lane - 8`, all at once.
__device__ __forceinline__ void halo_wait(const HaloParams &h, const unsigned dmask, const uint32_t E, const int lane) {
    const int d = lane & 7;
    const bool ack = lane >= 8;
    const bool want = lane < 16 && ((dmask >> d) & 1u) && h.peer_flag[d] != nullptr &&
                      (h.mode & (ack ? kHaloWaitAck : kHaloWaitData));
    if (want) {
        if (h.delay > 0) { const long long t0 = clock64(); while (clock64() - t0 < h.delay) {} }
        const uint32_t *const f = (ack ? h.acks : h.flags) + d * 32;
        const long long t0 = clock64();
        uint32_t have;
        while ((int)((have = ld_relaxed_sys(f)) - E) < 0) {
            if (clock64() - t0 > (20ll << 30)) {  // ~10 s: a peer died or the protocol broke
                printf("b200conv: halo %s wait timeout dir=%d want=%u have=%u\n", ack ? "ack" : "data", d, E, have);
                __trap();
            }
        }
#ifdef B2C_DEBUG
        if ((int)(have - E) > 1) {   // a neighbour can be at most one exchange ahead
            printf("b200conv DEBUG: %s epoch dir=%d ran ahead: want=%u have=%u\n", ack ? "ack" : "data", d, E, have);
            __trap();
        }
#endif
        (void)ld_acquire_sys(f);   // orders the frame reads / remote stores below after the flag
        if (h.prof) {
            const unsigned long long dt = (unsigned long long)(clock64() - t0);
            atomicAdd(h.prof + (ack ? 8 : 0) + d, dt);
            atomicMax(h.prof + 24 + (ack ? 8 : 0) + d, dt);
            atomicAdd(h.prof + 16, 1ull);
        }
    }
    __syncwarp();
}

// Ring warps after their band: every lane's remote stores are issued; lanes 0-7 each take one
// direction, order the warp's stores before the count-in, and the last arriver of a direction
// publishes the epoch (data: E+1 into the neighbour's flag; ack: E into its ack word).
//
// Ring warps release at GPU scope (fence + relaxed atomic on the counter); the last arriver's atomic
// observes them.  A system-scope fence costs microseconds on an idle machine and ~100 us while PCIe
// copies are in flight (it also waits for the host side), and a pass cannot end before its slowest
// thread: with the fence inside the pass, 8 GPUs lost 4-6 ms per 1000 iterations end to end
// (profiles/e2e_probe.md).  So with peers on other devices the last arriver only stores the epoch
// into a local word (release, GPU scope); a one-warp courier kernel on a side stream acquires it
// and release-stores the flag into the neighbour at system scope — cumulative over everything
// that happened before in causality order, i.e. every ring warp's stores.  Without a courier
// (peers on this device, one-shot paths) the last arriver publishes itself.
__device__ __forceinline__ void halo_publish(const HaloParams &h, const unsigned dmask, const uint32_t E,
                                             const int lane, const bool data) {
    __syncwarp();
    const bool mine = lane < 8 && ((dmask >> lane) & 1u) && h.peer_flag[lane] != nullptr;
    if (mine) {
        const long long t0 = h.prof ? clock64() : 0;
        __threadfence();   // one fence per warp instruction; the barrier above makes it cover all lanes
        const unsigned old = atomicAdd(h.counters + lane * 32, 1u);
        if ((int)old + 1 == h.expected[lane]) {
            h.counters[lane * 32] = 0;
            uint32_t *const f = data ? h.peer_flag[lane] : h.peer_ack[lane];
            const uint32_t v = data ? E + 1u : E;
            if (h.ready) {
                // hand the epoch to the courier: no system-scope fence inside a pass
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(h.ready + ((data ? 0 : 8) + lane) * 32), "r"(v) : "memory");
            } else if (h.flag_fence == 2) {
                __threadfence();
                asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(f), "r"(v) : "memory");
            } else {
                // the release store's own fence is also the acquire side of the counter atomic above
                if (h.flag_fence == 0) __threadfence_system();
                st_release_sys(f, v);
            }
        }
        if (h.prof) {
            atomicAdd(h.prof + 17, (unsigned long long)(clock64() - t0));
            atomicAdd(h.prof + 18, 1ull);
        }
    }
}

// One work item (band, chunk) of a pass: wait for the neighbours' epoch if this warp touches a
// halo frame, stream the band, publish our epoch.  Shared by the per-pass kernels (LDG and TMA
// feeds) and the persistent kernel.
template <int K, int NW, int CH, bool CHECK, class MakeFeed>
__device__ __forceinline__ void stream_tile(const PassParams &p, const int band, const int chunk, const int lane,
                                            MakeFeed make_feed) {
    constexpr int SB = NW * 4;
    const int x0 = band * p.vw - p.hlo;
    const int in_r0 = chunk * p.chunk_rows - K;
    const int in_r1 = (chunk == p.nchunks - 1 ? p.rows : (chunk + 1) * p.chunk_rows) + K;
    const bool edge = (p.zero_w && x0 < 0) || (p.zero_e && x0 + 32 * SB > p.bw) ||
                      (p.zero_n && in_r0 < 0) || (p.zero_s && in_r1 > p.rows);
    const bool partial = (p.bw % SB) != 0 && (band + 1) * p.vw >= p.bw;  // owns the cut last strip

    bool top = false, bot = false, west = false, east = false;
    unsigned dmask = 0;   // directions whose ring this warp belongs to (bit = B2C_N .. B2C_SE)
    uint32_t E = 0;
    const HaloParams &h = p.halo;
    if (h.enabled) {
        // ring membership, by the same formulas the host used for h.expected[]
        top = chunk == 0; bot = chunk == p.nchunks - 1;
        west = band < h.nbw; east = band >= p.nbands - h.nbe;
        dmask = (top ? 1u : 0u) | (bot ? 2u : 0u) | (west ? 4u : 0u) | (east ? 8u : 0u) | (top && west ? 16u : 0u) |
                (top && east ? 32u : 0u) | (bot && west ? 64u : 0u) | (bot && east ? 128u : 0u);
        if (dmask) {
            E = h.epoch + (h.epoch_base ? __ldcg(h.epoch_base) : 0u);
            if (h.mode & (kHaloWaitData | kHaloWaitAck)) halo_wait(h, dmask, E, lane);
        }
    }
    const bool do_push = dmask != 0 && (h.mode & kHaloPush) && h.fused;

    auto feed = make_feed();
    if (edge || partial || do_push)
        stream_band<K, NW, CH, true, CHECK>(p, feed, band, chunk, lane, do_push, top, bot, west, east);
    else
        stream_band<K, NW, CH, false, CHECK>(p, feed, band, chunk, lane, false, false, false, false, false);

    if (dmask != 0 && (do_push || (h.mode & kHaloPostAck))) halo_publish(h, dmask, E, lane, do_push);
}

// Per-pass kernels: one warp = one work item of the pass.
template <int K, int NW, int CH, bool CHECK, class MakeFeed>
__device__ __forceinline__ void stream_warp(const PassParams &p, const int wid, const int lane, MakeFeed make_feed) {
    int band, chunk;
    // Order 2: a pass that only WAITS for the neighbours' rings (the first pass of a group) takes its
    // ring items last, so that the epoch has the whole interior's time to arrive.
    map_work(p.raster == 2 ? p.nbands * p.nchunks - 1 - wid : wid, p.nbands, p.nchunks,
             p.raster == 1 ? p.nbands : p.halo.nbw, p.halo.nbe, band, chunk);

    // Programmatic dependent launch: everything above ran while the previous pass was still
    // draining.  Let the pass after this one start scheduling, then wait until the previous
    // pass (which wrote our source buffer) has completed and flushed.  Both are no-ops for
    // launches without the programmatic-serialization attribute.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    // Fixed point detected at check c: every rank runs interval c+1 to the end (the reduction of
    // check c is still in flight on another stream meanwhile) and skips everything from c+2 on.
    if (p.converged) {
        const unsigned int c = *p.converged;
        if (c != 0 && c + 1 < p.interval) return;
    }
    stream_tile<K, NW, CH, CHECK>(p, band, chunk, lane, make_feed);
}

// The CHECK variants keep one more row of lanes (the last stage's input, to compare with its
// output): under the 128-register cap that costs 400 bytes of spills inside the row loop, so they
// trade a quarter of the occupancy for registers (one pass in five at check_every = 20).
template <int K, int NW, int CH, bool CHECK>
__global__ void __launch_bounds__(kWarpsPerCta * 32, (CHECK && K * NW >= 16) ? B2C_MIN_CTAS - 1 : B2C_MIN_CTAS)
stream_kernel(const PassParams p) {
    const int lane = threadIdx.x & 31;
    const int wid = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (wid >= p.nbands * p.nchunks) return;
    stream_warp<K, NW, CH, CHECK>(p, wid, lane, [&]() { return LdgFeed<NW>(p, lane); });
}

// Courier: one warp that lives beside a run and forwards halo epochs from local "ready" words to
// the neighbours' flag / ack words with system-scope ordering (lane d: data flag of direction d,
// lane 8+d: ack of direction d).  It leaves when the last epoch of the run has been forwarded.
// Runs as a one-warp kernel on a side stream beside per-pass launches (kernels.cu), and as warp 0
// of CTA 0 inside the persistent kernel — whose CTAs fill every SM, so that a separate kernel
// might never become resident.
struct CourierParams {
    const uint32_t *ready;        // [16][32] words
    uint32_t *peer_flag[8];
    uint32_t *peer_ack[8];
    const uint32_t *epoch_base;   // added to the three values below (null: 0)
    uint32_t start_epoch;         // data epochs up to this one are already published
    uint32_t final_data;          // last data epoch of the run
    uint32_t final_ack;           // last ack epoch of the run
    int has_ack;                  // 0: no group of the run acknowledges
};

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void courier_loop(const CourierParams &c, const int lane, const uint32_t add) {
    const int d = lane & 7;
    const bool ack = lane >= 8;
    uint32_t *const dst = ack ? c.peer_ack[d] : c.peer_flag[d];
    if (lane >= 16 || !dst || (ack && !c.has_ack)) return;
    const uint32_t base = add + (c.epoch_base ? __ldcg(c.epoch_base) : 0u);
    const uint32_t last = base + (ack ? c.final_ack : c.final_data);
    uint32_t sent = base + c.start_epoch - (ack ? 1u : 0u);   // acks carry the epoch of their own group
    const uint32_t *const src = c.ready + lane * 32;
    long long t0 = clock64();
    while ((int)(sent - last) < 0) {
        const uint32_t v = ld_acquire_gpu(src);
        if ((int)(v - sent) > 0 && (int)(v - last) <= 0) {
            // The acquire above synchronises with the last arriver's release (GPU scope); the release
            // store's own system-scope fence is cumulative over everything that happened before it,
            // i.e. every ring warp's stores.  ONE system-scope fence: a second one in front of it
            // doubles the latency the neighbours see while PCIe copies are in flight.
            st_release_sys(dst, v);
            sent = v;
            t0 = clock64();
        } else {
            __nanosleep(200);
            if (clock64() - t0 > (40ll << 30)) {
                printf("b200conv: courier timeout dir=%d %s sent=%u last=%u ready=%u\n", d, ack ? "ack" : "data", sent, last, v);
                __trap();
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Persistent dataflow kernel: ONE launch runs every pass of a run.
//
// The per-pass kernels above meet at a grid-wide barrier after every K iterations (the kernel
// boundary): a pass of 2.25 waves of warps costs three rounds, its slowest warp holds everybody
// back, and every pass is a launch the front end has to fetch (measurably late when PCIe is busy
// with the next image's upload).  None of that is needed: a work item (pass t, band b, chunk c)
// depends only on the nine items (t-1, b+-1, c+-1) — they wrote what it reads, and they read what
// it overwrites.  So resident warps pull items from one global counter, in pass order, and wait
// on per-tile "passes completed" words instead of on the whole grid.  An item's dependencies are
// always earlier in the queue, i.e. already taken by a warp that is running, so the scheme cannot
// deadlock however many CTAs are resident (no cooperative launch needed).
//
// Replaces the host loop of cuda/cuda_convolution.cu:66-87 and mpi/mpi_convolution.c:156-240.
struct PersistParams {
    const PassParams *passes;      // [npasses], device memory
    const int *cum;                // [npasses + 1] work items before pass t
    int npasses;
    unsigned long long *counter;   // global work counter (never reset; item_base = its value at launch)
    unsigned long long item_base;
    uint32_t *tile_done;           // [chunk * tstride + band]: passes completed by that tile, + pass_base
    int tstride;
    uint32_t pass_base;            // tile of pass t stores pass_base + t + 1
    uint32_t epoch0;               // added to every pass's halo epoch (the exchange count at launch)
    unsigned long long *prof;      // opt-in: [0] cycles waiting for tiles, [1] cycles in tiles, [2] total, [3] items, [4] waits
    const CourierParams *courier;  // when set: warp 0 of CTA 0 is the courier (epochs relative to epoch0)
};

__device__ __forceinline__ uint32_t ld_relaxed_gpu(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t *p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int K, int NW, int CH>
__global__ void __launch_bounds__(kWarpsPerCta * 32, B2C_MIN_CTAS)
stream_persist_kernel(const PersistParams q) {
    static_assert(sizeof(PassParams) % 16 == 0, "PassParams is copied in 16-byte pieces");
    __shared__ __align__(16) PassParams sp[kWarpsPerCta];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (q.courier && blockIdx.x == 0 && warp == 0) {   // CTA 0 is resident from the first instant
        courier_loop(*q.courier, lane, q.epoch0);
        return;
    }
    const PassParams &p = sp[warp];
    const unsigned long long total = (unsigned long long)__ldg(q.cum + q.npasses);
    int t = -1;                                   // pass whose parameters sit in sp[warp]
    unsigned long long lo = 0, hi = 0;            // its item range [lo, hi)
    int prev_nb = 1, prev_nc = 1;
    // The next item is fetched before the current one is worked on, so the atomic's round trip hides
    // behind a whole tile.  Still deadlock-free: a held item's dependencies are earlier in the queue,
    // taken by running warps (possibly this one, which finishes its current tile first).
    unsigned long long g = 0;
    if (lane == 0) g = atomicAdd(q.counter, 1ull) - q.item_base;
    g = __shfl_sync(kFull, g, 0);
    long long c_wait = 0, c_tile = 0, n_items = 0, n_waits = 0;
    const long long c_begin = q.prof ? clock64() : 0;
    while (g < total) {
        unsigned long long gnext = 0;
        if (lane == 0) gnext = atomicAdd(q.counter, 1ull) - q.item_base;
        if (g >= hi) {   // items only move forward: step to the pass that holds g (usually the next one)
            do {
                ++t;
                lo = hi;
                hi = (unsigned long long)__ldg(q.cum + t + 1);
            } while (g >= hi);
            if (t > 0) {   // tiling of the pass before (its extension may differ by one band / chunk)
                prev_nb = __ldg(&q.passes[t - 1].nbands);
                prev_nc = __ldg(&q.passes[t - 1].nchunks);
            }
            __syncwarp();
            const uint4 *src = reinterpret_cast<const uint4 *>(q.passes + t);
            uint4 *dst = reinterpret_cast<uint4 *>(&sp[warp]);
            for (int i = lane; i < (int)(sizeof(PassParams) / 16); i += 32) dst[i] = __ldg(src + i);
            __syncwarp();
            if (lane == 0) sp[warp].halo.epoch += q.epoch0;
            __syncwarp();
        }
        int band, chunk;
        map_work((int)(g - lo), p.nbands, p.nchunks, p.raster == 1 ? p.nbands : p.halo.nbw, p.halo.nbe, band, chunk);
        const long long c0 = q.prof ? clock64() : 0;
        if (t > 0) {
            // lanes 0-8: the 3x3 neighbourhood in the previous pass's tiling
            if (lane < 9) {
                int pb = band + lane % 3 - 1, pc = chunk + lane / 3 - 1;
                pb = pb < 0 ? 0 : (pb >= prev_nb ? prev_nb - 1 : pb);
                pc = pc < 0 ? 0 : (pc >= prev_nc ? prev_nc - 1 : pc);
                const uint32_t *f = q.tile_done + (long long)pc * q.tstride + pb;
                const uint32_t need = q.pass_base + (uint32_t)t;
                if ((int)(ld_acquire_gpu(f) - need) < 0) {   // usually done long ago: one round trip
                    const long long t0 = clock64();
                    ++n_waits;
                    do {
                        __nanosleep(64);
                        if (clock64() - t0 > (8ll << 30)) {
                            printf("b200conv: tile dependency timeout pass=%d band=%d chunk=%d waits for (%d,%d)\n", t,
                                   band, chunk, pb, pc);
                            __trap();
                        }
                    } while ((int)(ld_acquire_gpu(f) - need) < 0);
                }
            }
            __syncwarp();
        }
        const long long c1 = q.prof ? clock64() : 0;
        stream_tile<K, NW, CH, false>(p, band, chunk, lane, [&]() { return LdgFeed<NW>(p, lane); });
        __syncwarp();
        if (lane == 0) st_release_gpu(q.tile_done + (long long)chunk * q.tstride + band, q.pass_base + (uint32_t)t + 1u);
        if (q.prof) {
            const long long c2 = clock64();
            c_wait += c1 - c0;
            c_tile += c2 - c1;
            ++n_items;
        }
        g = __shfl_sync(kFull, gnext, 0);
    }
    if (q.prof && lane == 0) {
        atomicAdd(q.prof + 0, (unsigned long long)c_wait);
        atomicAdd(q.prof + 1, (unsigned long long)c_tile);
        atomicAdd(q.prof + 2, (unsigned long long)(clock64() - c_begin));
        atomicAdd(q.prof + 3, (unsigned long long)n_items);
        n_waits = __reduce_add_sync(1u, (unsigned)n_waits);
        atomicAdd(q.prof + 4, (unsigned long long)n_waits);
    }
}

}  // namespace b2c
