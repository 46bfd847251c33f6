// Register-resident persistent stencil for images that fit on chip (B2C_KERNEL_RESIDENT).
//
// SURVEY §7.3 step 4 / §8 row K-drv: "1920x2520 grey (4.8 MB) fits in aggregate on-chip memory ->
// persistent kernel keeping the image on chip for all iterations".  The streaming kernel
// (stencil_stream.cuh) runs such an image as ~250 launches that are each latency bound: one warp
// walks its band row by row, every row waits for the SHFL->add chain of the one before.  Here
//
//   * the image is cut into (CTA rows) x (strips) tiles, one CTA of 16 warps per tile, at most one
//     CTA per SM, launched cooperatively so that all of them are resident together;
//   * a warp keeps RW rows of a 256-byte window in registers for the whole run (8 bytes per lane
//     per row, as even/odd 16-bit lanes like the streaming kernel, same hsum<> and the same
//     scale-1 / scale-16 alternation), so one iteration is RW independent row updates per warp
//     — instruction-level parallelism instead of a dependent chain — done in place;
//   * inside a CTA neighbouring warps trade the horizontal sums of their first and last row through
//     shared memory, one __syncthreads per iteration (double-buffered by iteration parity);
//   * between CTAs the window carries a halo of K rows and K*CH bytes (rounded to the lane width;
//     K = 4, 8 or 16, chosen per shape by the engine) that is recomputed redundantly; every K
//     iterations each CTA writes the K-deep ring of its valid region to an exchange image in L2 (one
//     per epoch parity), publishes an epoch with st.release.gpu, waits for its (up to) 8 neighbours
//     with ld.acquire.gpu and reloads its halo with predicated loads issued back to back.
//     Same slot-reuse argument as the multi-GPU protocol: a CTA can write epoch e+2 only after it
//     has seen its neighbours' epoch e+1, which they publish after reading epoch e.
//     Measured: an iteration costs ~66 ns per row held by a warp, an exchange ~4 us.
//   * thread-block clusters + distributed shared memory: the CTAs of one tile row are launched as
//     clusters of up to 8 (p.cx, a divisor of the strip count).  West/east neighbours inside a
//     cluster do not go through L2 for their halo columns: the owner stores its ring columns
//     straight into the neighbour's shared-memory inbox (generic stores to a map_shared_rank
//     address = st.shared::cluster), one barrier.cluster (arrive.release / wait.acquire) replaces
//     their flag handshake, and the halo comes back with LDS.  North/south and diagonal
//     neighbours live in other clusters and keep the L2 path.
//   * ZERO boundary: cells outside the image are forced to zero by the mask / byte-permute that
//     does the >>4 anyway (no extra instruction); TORUS: halo loads wrap around the image.
//
// Nothing comparable exists in the reference (cuda/cuda_convolution.cu:66-87 launches one kernel per
// iteration from the host and re-reads the whole image from DRAM every time).
#include "stencil_resident.h"

#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <mutex>

#include "common.h"
#include "kernels.h"
#include "stencil_stream.cuh"

namespace cg = cooperative_groups;

namespace b2c {
namespace {

constexpr long long kResTimeout = 3ll << 30;  // clocks (~1.6 s) before a wait gives up


// 8 bytes at (gr, gc) of an H x bw byte image; outside the image: wrapped (torus) or zero.
// gc and bw are multiples of 8, so a strip is never split by the image edge.
__device__ __forceinline__ uint2 res_load8(const uint8_t *img, long long pitch, int gr, int gc, int H, int bw,
                                           bool torus) {
    if (torus) {
        gr = gr < 0 ? gr + H : (gr >= H ? gr - H : gr);
        gc = gc < 0 ? gc + bw : (gc >= bw ? gc - bw : gc);
    }
    if ((unsigned)gr >= (unsigned)H || (unsigned)gc >= (unsigned)bw) return make_uint2(0u, 0u);
    return __ldcg(reinterpret_cast<const uint2 *>(img + (long long)gr * pitch + gc));
}

// Predicated 8-byte L2 load (zero when `pred` is false) that ptxas cannot turn into a branch.
__device__ __forceinline__ uint2 res_load8_if(const uint8_t *ptr, bool pred) {
    uint2 v;
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\tmov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\t"
        "@p ld.global.cg.v2.u32 {%0, %1}, [%2];\n\t}"
        : "=r"(v.x), "=r"(v.y)
        : "l"(ptr), "r"((unsigned)pred)
        : "memory");
    return v;
}

__device__ __forceinline__ void res_unpack(uint2 w, uint32_t (&E)[kResNW], uint32_t (&O)[kResNW]) {
    E[0] = w.x & 0x00ff00ffu;
    O[0] = __byte_perm(w.x, 0u, 0x4341);
    E[1] = w.y & 0x00ff00ffu;
    O[1] = __byte_perm(w.y, 0u, 0x4341);
}
__device__ __forceinline__ uint2 res_pack(const uint32_t (&E)[kResNW], const uint32_t (&O)[kResNW]) {
    return make_uint2(__byte_perm(E[0], O[0], 0x6240), __byte_perm(E[1], O[1], 0x6240));
}

// Shared-memory mailbox: [iteration parity][warp][first row / last row][E / O][word][lane]
struct ResShared {
    uint32_t h[2][kResWarps][2][2][kResNW][32];
    int abort;
};


// One iteration of all RW rows of a warp, in place.  ODD = the lanes come in at scale 16 (second
// iteration of a pair) and leave at scale 1; otherwise scale 1 in, scale 16 out.
template <int RW, int CH, bool ODD>
__device__ __forceinline__ void res_iteration(uint32_t (&E)[RW][kResNW], uint32_t (&O)[RW][kResNW], ResShared &sh,
                                              const int buf, const int warp, const int lane, const int wr0,
                                              const int H, const bool lane_ok, const bool torus) {
    constexpr int NW = kResNW;
    uint32_t hcE[NW], hcO[NW], hLE[NW], hLO[NW];
    hsum<NW, CH>(E[0], O[0], hcE, hcO);
    hsum<NW, CH>(E[RW - 1], O[RW - 1], hLE, hLO);
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        sh.h[buf][warp][0][0][i][lane] = hcE[i];
        sh.h[buf][warp][0][1][i][lane] = hcO[i];
        sh.h[buf][warp][1][0][i][lane] = hLE[i];
        sh.h[buf][warp][1][1][i][lane] = hLO[i];
    }
    __syncthreads();
    // s(r-1) = h(r-1) + h(r); rows above warp 0 / below the last warp do not exist (their
    // influence stays inside the K-row halo)
    uint32_t psE[NW], psO[NW], dnE[NW], dnO[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        const uint32_t upE = warp > 0 ? sh.h[buf][warp - 1][1][0][i][lane] : 0u;
        const uint32_t upO = warp > 0 ? sh.h[buf][warp - 1][1][1][i][lane] : 0u;
        psE[i] = upE + hcE[i];
        psO[i] = upO + hcO[i];
        dnE[i] = warp + 1 < kResWarps ? sh.h[buf][warp + 1][0][0][i][lane] : 0u;
        dnO[i] = warp + 1 < kResWarps ? sh.h[buf][warp + 1][0][1][i][lane] : 0u;
    }
#pragma unroll
    for (int r = 0; r < RW; ++r) {
        uint32_t hnE[NW], hnO[NW];
        if (r + 1 < RW - 1) {
            hsum<NW, CH>(E[r + 1], O[r + 1], hnE, hnO);   // still the old row r+1: only rows <= r are rewritten
        } else {
#pragma unroll
            for (int i = 0; i < NW; ++i) {
                hnE[i] = (r + 1 == RW - 1) ? hLE[i] : dnE[i];
                hnO[i] = (r + 1 == RW - 1) ? hLO[i] : dnO[i];
            }
        }
        const bool ok = lane_ok && (torus || (unsigned)(wr0 + r) < (unsigned)H);
        const uint32_t m = ok ? 0x0ff00ff0u : 0u;       // scale 1 in: 16*floor(v/16), or 0 outside the image
        const uint32_t sel = ok ? 0x4341u : 0x4444u;    // scale 16 in: high byte of each lane, or 0
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            const uint32_t sE = hcE[i] + hnE[i], sO = hcO[i] + hnO[i];
            const uint32_t vE = psE[i] + sE, vO = psO[i] + sO;
            psE[i] = sE;
            psO[i] = sO;
            hcE[i] = hnE[i];
            hcO[i] = hnO[i];
            E[r][i] = ODD ? __byte_perm(vE, 0u, sel) : (vE & m);
            O[r][i] = ODD ? __byte_perm(vO, 0u, sel) : (vO & m);
        }
    }
}

template <int RW, int CH, int K>
__global__ void __launch_bounds__(kResWarps * 32, 1) resident_kernel(const ResidentParams p) {
    constexpr int NW = kResNW, SB = kResSB;
    constexpr int HB = res_halo_bytes(CH, K);
    constexpr int VR = kResWarps * RW - 2 * K;   // valid rows per CTA
    constexpr int VC = kResWinB - 2 * HB;        // valid bytes per CTA
    __shared__ ResShared sh;
    // inbox[epoch parity][0 = from the west neighbour, 1 = from the east][row of the CTA][8-byte slot]
    extern __shared__ __align__(16) uint2 inbox[];
    constexpr int kSlots = HB / SB, kRows = kResWarps * RW;
    cg::cluster_group cluster = cg::this_cluster();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bx = blockIdx.x % p.strips, by = blockIdx.x / p.strips;
    const bool torus = p.torus != 0;
    if (threadIdx.x == 0) sh.abort = 0;   // ordered before its first read by the barriers in between
    // West / east neighbour inside this cluster?  (clusters are cx consecutive strips of one tile row;
    // the torus wrap-around pair is in the same cluster only when the cluster is the whole row)
    const int cx = p.cx, crank = bx % (cx > 0 ? cx : 1);
    const bool whole_row = cx == p.strips && cx > 1;
    const bool w_in = cx > 1 && (crank > 0 || (torus && whole_row));
    const bool e_in = cx > 1 && (crank + 1 < cx || (torus && whole_row));
    uint2 *const w_peer = w_in ? cluster.map_shared_rank(inbox, (crank + cx - 1) % cx) : nullptr;   // its "from the east" side
    uint2 *const e_peer = e_in ? cluster.map_shared_rank(inbox, (crank + 1) % cx) : nullptr;        // its "from the west" side

    // This CTA's valid region, this lane's strip, this warp's rows — all in image coordinates.
    const int vr0 = by * VR, vr1 = min(vr0 + VR, p.H);
    const int vc0 = bx * VC, vc1 = min(vc0 + VC, p.bw);
    const int gc = vc0 - HB + lane * SB;
    const int wr0 = vr0 - K + warp * RW;
    const bool col_valid = gc >= vc0 && gc < vc1;
    const bool col_near = gc >= vc0 - HB && gc < vc1 + HB;          // valid or halo (not dead)
    const bool col_ring = col_valid && (gc < vc0 + HB || gc >= vc1 - HB);
    const bool lane_ok = torus || (gc >= 0 && gc < p.bw);

    uint32_t E[RW][NW], O[RW][NW];
#pragma unroll
    for (int r = 0; r < RW; ++r) {
        const int gr = wr0 + r;
        const bool row_near = gr >= vr0 - K && gr < vr1 + K;
        uint2 w = make_uint2(0u, 0u);
        if (row_near && col_near) w = res_load8(p.src, p.src_pitch, gr, gc, p.H, p.bw, torus);
        res_unpack(w, E[r], O[r]);
    }

    uint32_t itc = 0;            // iterations so far: parity selects the shared-memory mailbox
    uint32_t epoch = p.epoch0;
    int done = 0;
    while (true) {
        const int k = min(K, p.reps - done);
        for (int it = 0; it < k; it += 2) {
            res_iteration<RW, CH, false>(E, O, sh, (int)(itc & 1u), warp, lane, wr0, p.H, lane_ok, torus);
            ++itc;
            if (it + 1 < k) {
                res_iteration<RW, CH, true>(E, O, sh, (int)(itc & 1u), warp, lane, wr0, p.H, lane_ok, torus);
                ++itc;
            }
        }
        if (k & 1) {   // odd count: lanes are at scale 16 (multiples of 16), bring them back to scale 1
#pragma unroll
            for (int r = 0; r < RW; ++r)
#pragma unroll
                for (int i = 0; i < NW; ++i) {
                    E[r][i] >>= 4;
                    O[r][i] >>= 4;
                }
        }
        done += k;
        if (done >= p.reps) break;
        if (p.debug & 1) continue;   // profiling only (B200CONV_RES_DEBUG=1): no exchange, wrong bytes

        // ---- exchange: ring out, epoch, neighbours' epochs, halo in
        ++epoch;
        uint8_t *const xb = (epoch & 1u) ? p.xbuf[1] : p.xbuf[0];
        const int par = (int)(epoch & 1u);
#pragma unroll
        for (int r = 0; r < RW; ++r) {
            const int gr = wr0 + r;
            const bool row_valid = gr >= vr0 && gr < vr1;
            const bool row_ring = gr < vr0 + K || gr >= vr1 - K;
            if (row_valid && col_valid) {
                const bool ring_w = gc < vc0 + HB, ring_e = gc >= vc1 - HB;
                const uint2 packed = res_pack(E[r], O[r]);
                // ring columns go into the shared memory of a west / east neighbour of this cluster ...
                const int lr = warp * RW + r;
                if (ring_w && w_in) w_peer[((par * 2 + 1) * kRows + lr) * kSlots + (gc - vc0) / SB] = packed;
                if (ring_e && e_in) e_peer[((par * 2 + 0) * kRows + lr) * kSlots + (gc - (vc1 - HB)) / SB] = packed;
                // ... and into the exchange image in L2 for everybody else
                if (row_ring || (ring_w && !w_in) || (ring_e && !e_in))
                    __stcg(reinterpret_cast<uint2 *>(xb + (long long)gr * p.xpitch + gc), packed);
            }
        }
        // barrier, then ONE gpu-scope fence by the thread that publishes (the grid-sync pattern of
        // cooperative groups): the barrier orders every thread's ring stores before thread 0's fence,
        // the fence is cumulative.  A fence per thread cost 4 % of the run.
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            st_release_gpu(p.flags + (long long)blockIdx.x * 32, epoch);
        }
        // in-cluster neighbours: their ring sits in our inbox once everybody has passed this barrier
        if (cx > 1) cluster.sync();
        if (threadIdx.x < 8) {
            int ny = by + kDirDrDev(threadIdx.x), nx = bx + kDirDcDev(threadIdx.x);
            bool exists = !((threadIdx.x == B2C_W && w_in) || (threadIdx.x == B2C_E && e_in));
            if (torus) {
                ny = ny < 0 ? ny + p.ctarows : (ny >= p.ctarows ? ny - p.ctarows : ny);
                nx = nx < 0 ? nx + p.strips : (nx >= p.strips ? nx - p.strips : nx);
            } else {
                exists = exists && ny >= 0 && ny < p.ctarows && nx >= 0 && nx < p.strips;
            }
            if (exists) {
                const uint32_t *f = p.flags + (long long)(ny * p.strips + nx) * 32;
                const long long t0 = clock64();
                while ((int)(ld_acquire_gpu(f) - epoch) < 0) {
                    if (clock64() - t0 > kResTimeout) {
                        atomicExch(p.error, 1u);
                        if (cx > 1) __trap();   // cluster mates would wait for ever at their next barrier
                        sh.abort = 1;
                        break;
                    }
                }
            }
        }
        __syncthreads();
        if (sh.abort) return;   // same value for the whole CTA; the neighbours time out in turn
        // All halo loads are issued before the first one is used (predicated, no branches): one L2
        // round trip per exchange instead of one per row (ncu: the row-by-row version spent 18 % of
        // all stall samples here).
        {
            int sc = gc;
            if (torus) sc = sc < 0 ? sc + p.bw : (sc >= p.bw ? sc - p.bw : sc);
            const bool sc_ok = (unsigned)sc < (unsigned)p.bw;
            uint2 hw[RW];
            bool hwant[RW];
#pragma unroll
            for (int r = 0; r < RW; ++r) {
                const int gr = wr0 + r;
                const bool row_valid = gr >= vr0 && gr < vr1;
                const bool row_near = gr >= vr0 - K && gr < vr1 + K;
                hwant[r] = !(row_valid && col_valid) && row_near && col_near;
                int sr = gr;
                if (torus) sr = sr < 0 ? sr + p.H : (sr >= p.H ? sr - p.H : sr);
                // halo columns beside valid rows come from the inbox when that neighbour is a cluster mate
                const bool from_w = row_valid && gc < vc0 && w_in, from_e = row_valid && gc >= vc1 && e_in;
                const bool ld = hwant[r] && !from_w && !from_e && sc_ok && (unsigned)sr < (unsigned)p.H;   // else: outside the image = 0
                hw[r] = res_load8_if(xb + (long long)(ld ? sr : 0) * p.xpitch + (ld ? sc : 0), ld);
                if (hwant[r] && (from_w || from_e)) {
                    const int lr = warp * RW + r;
                    const int slot = from_w ? (gc - (vc0 - HB)) / SB : (gc - vc1) / SB;
                    hw[r] = inbox[((par * 2 + (from_w ? 0 : 1)) * kRows + lr) * kSlots + slot];
                }
            }
#pragma unroll
            for (int r = 0; r < RW; ++r) {
                uint32_t nE[NW], nO[NW];
                res_unpack(hw[r], nE, nO);
#pragma unroll
                for (int i = 0; i < NW; ++i) {
                    E[r][i] = hwant[r] ? nE[i] : E[r][i];
                    O[r][i] = hwant[r] ? nO[i] : O[r][i];
                }
            }
        }
    }

#pragma unroll
    for (int r = 0; r < RW; ++r) {
        const int gr = wr0 + r;
        if (gr >= vr0 && gr < vr1 && col_valid)
            __stcg(reinterpret_cast<uint2 *>(p.dst + (long long)gr * p.dst_pitch + gc), res_pack(E[r], O[r]));
    }
    if (cx > 1) cluster.sync();   // nobody leaves while a cluster mate may still store into its shared memory
}

template <int RW>
cudaError_t launch_rw(const ResidentParams &p, int ch, int k, int ctas, cudaStream_t s, int *cluster_cap) {
    const void *fn =
        k == 16 ? (ch == 1 ? (const void *)resident_kernel<RW, 1, 16> : (const void *)resident_kernel<RW, 3, 16>)
        : k == 8 ? (ch == 1 ? (const void *)resident_kernel<RW, 1, 8> : (const void *)resident_kernel<RW, 3, 8>)
                 : (ch == 1 ? (const void *)resident_kernel<RW, 1, 4> : (const void *)resident_kernel<RW, 3, 4>);
    // inbox: [2 parities][2 sides][16*RW rows][halo bytes]
    const size_t smem = (size_t)4 * kResWarps * RW * res_halo_bytes(ch, k);
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (p.cx > 8) {   // more than the portable eight CTAs per cluster: B200 allows sixteen on request
        e = cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(ctas);
    cfg.blockDim = dim3(kResWarps * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    // cooperative: the launch fails instead of deadlocking if the CTAs cannot all be resident
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = (unsigned)p.cx;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p.cx > 1 ? 2 : 1;
    if (cluster_cap) {   // query only: how many clusters of this size can be resident at once
        int n = 0;
        e = cudaOccupancyMaxActiveClusters(&n, fn, &cfg);
        *cluster_cap = e == cudaSuccess ? n : 0;
        return e;
    }
    ResidentParams q = p;
    void *args[] = {&q};
    return cudaLaunchKernelExC(&cfg, fn, args);
}

template <int RW>
cudaError_t dispatch(const ResidentParams &p, int ch, int k, int ctas, cudaStream_t s, int *cluster_cap) {
    return launch_rw<RW>(p, ch, k, ctas, s, cluster_cap);
}

cudaError_t by_rw(const ResidentPlan &plan, const ResidentParams &p, int ch, int ctas, cudaStream_t s, int *cluster_cap) {
    switch (plan.rw) {
        case 4: return dispatch<4>(p, ch, plan.k, ctas, s, cluster_cap);
        case 6: return dispatch<6>(p, ch, plan.k, ctas, s, cluster_cap);
        case 8: return dispatch<8>(p, ch, plan.k, ctas, s, cluster_cap);
        case 10: return dispatch<10>(p, ch, plan.k, ctas, s, cluster_cap);
        case 12: return dispatch<12>(p, ch, plan.k, ctas, s, cluster_cap);
        case 14: return dispatch<14>(p, ch, plan.k, ctas, s, cluster_cap);
        case 16: return dispatch<16>(p, ch, plan.k, ctas, s, cluster_cap);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace

cudaError_t launch_resident(const ResidentPlan &plan, const ResidentParams &p, int ch, cudaStream_t s) {
    const int ctas = plan.strips * plan.ctarows;
    cudaError_t e = by_rw(plan, p, ch, ctas, s, nullptr);
    count_launch(1);
    return e != cudaSuccess ? e : cudaGetLastError();
}

// Largest cluster width (CTAs of one tile row sharing halo columns through distributed shared
// memory) that divides the strip count, is at most 16 (8 is the portable limit, B200 takes 16 on
// request) and leaves every cluster of the grid resident at once — the kernel's CTAs wait for each other, so all of them
// have to be on the chip.  1 = no clusters.
int resident_cluster_width(const ResidentPlan &plan, int ch) {
    // Opt-in (B200CONV_RES_CLUSTER=1): measured on B200 the cluster path is 4-8 % SLOWER than flags in
    // L2 alone (profiles/resident_clusters.md) — north/south and diagonal neighbours are in other
    // clusters, so the L2 handshake stays on the critical path and barrier.cluster comes on top of it.
    static const bool on = getenv("B200CONV_RES_CLUSTER") && atoi(getenv("B200CONV_RES_CLUSTER")) != 0;
    if (!on) return 1;
    static const bool verbose = getenv("B200CONV_RES_DEBUG") && (atoi(getenv("B200CONV_RES_DEBUG")) & 4);
    const int ctas = plan.strips * plan.ctarows;
    for (int cx = 16; cx >= 2; --cx) {
        if (plan.strips % cx) continue;
        ResidentParams q{};
        q.cx = cx;
        int cap = 0;
        const cudaError_t e = by_rw(plan, q, ch, ctas, nullptr, &cap);
        if (verbose)
            fprintf(stderr, "b200conv: resident clusters of %d: %d can be resident, %d needed (%s)\n", cx, cap, ctas / cx,
                    cudaGetErrorString(e));
        if (e != cudaSuccess) {
            cudaGetLastError();
            continue;
        }
        if (cap * cx >= ctas) return cx;
    }
    return 1;
}

}  // namespace b2c
