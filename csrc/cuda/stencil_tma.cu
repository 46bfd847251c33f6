// TMA-fed variant of the streaming stencil (B2C_KERNEL_TMA).
//
// Same register pipeline as stencil_stream.cuh; only the way input rows reach the
// registers differs.  Every warp owns a ring of `depth` stages in shared memory; a stage
// is a 3-row x (32*SB)-byte tile fetched by cp.async.bulk.tensor.2d (one or two 256-byte
// wide boxes of a uint8 tensor map laid over the whole framed buffer) and signalled on
// the stage's mbarrier.  One elected lane re-arms a stage as soon as its rows sit in
// registers, so 3*depth rows (~36) are in flight per warp with zero registers spent on
// prefetch — the LDG feed can afford three.  The tensor map's out-of-bounds zero fill
// covers the rows/columns that stick out of the allocation.
//
// Replaces the per-thread byte loads of kernelConvolutionforGrey/RGB
// (cuda/cuda_convolution.cu:21,38-40).
#include <cuda.h>
#include <cuda_runtime.h>

#include <mutex>

#include "common.h"
#include "kernels.h"
#include "stencil_stream.cuh"
#include "stencil_tma.h"

namespace b2c {

constexpr int kTmaRows = 3;       // rows per stage = one trip of the 3x unrolled row loop
constexpr int kSmemBudget = 227 * 1024;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, int x, int y, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(tm), "r"(x), "r"(y), "r"(bar) : "memory");
}

// Feed #2: rows come out of the warp's TMA-filled shared-memory ring.
template <int NW>
struct TmaFeed {
    static constexpr int SB = NW * 4;
    static constexpr int kWin = 32 * SB;                  // window bytes per row
    static constexpr int kBox = kWin < 256 ? kWin : 256;  // TMA box width (<= 256 elements)
    static constexpr int kBoxes = kWin / kBox;
    static constexpr int kStage = kTmaRows * kWin;        // bytes per stage

    const CUtensorMap *tm;
    const int lane;
    const uint32_t ring, bars;
    const int depth;
    const int org_x, org_y;         // allocation coordinates of the pass origin (PassParams::org_x/org_y)
    int cx, cy, ngroups, g;
    uint32_t a[NW], b[NW], c[NW];   // the three rows of the current group

    __device__ __forceinline__ TmaFeed(const CUtensorMap *tm_, int lane_, uint32_t ring_, uint32_t bars_, int depth_,
                                       int org_x_, int org_y_)
        : tm(tm_), lane(lane_), ring(ring_), bars(bars_), depth(depth_), org_x(org_x_), org_y(org_y_) {}

    __device__ __forceinline__ void issue(int grp) {  // one lane
        const int slot = grp % depth;
        const uint32_t bar = bars + 8u * slot, dst = ring + (uint32_t)slot * kStage;
        mbar_expect_tx(bar, kStage);
#pragma unroll
        for (int b = 0; b < kBoxes; ++b) tma_load_2d(dst + b * kTmaRows * kBox, tm, cx + b * kBox, cy + kTmaRows * grp, bar);
    }

    __device__ __forceinline__ void start(int x0, int in_r0, int in_r1) {
        ngroups = (in_r1 - in_r0 + kTmaRows - 1) / kTmaRows;
        cx = org_x + x0;           // tensor coordinates are relative to the allocation
        cy = org_y + in_r0;
        g = 0;
        if (lane == 0) {
            for (int i = 0; i < depth; ++i) mbar_init(bars + 8u * i, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            // Halo frames are written by peers (and by the previous pass) with ordinary stores; the
            // tiles below are read through the async proxy.  After the flag acquire in stream_warp a
            // cross-proxy fence is needed before the first bulk copy may read those bytes.
            asm volatile("fence.proxy.async;" ::: "memory");
            const int n = ngroups < depth ? ngroups : depth;
            for (int i = 0; i < n; ++i) issue(i);
        }
        __syncwarp();
    }

    __device__ __forceinline__ void refill_a() {}
    __device__ __forceinline__ void refill_b() {}
    __device__ __forceinline__ void refill_c() {}
    __device__ __forceinline__ void group(int) {
        // The previous group's rows have been consumed from registers: its stage can be refilled.
        if (g >= 1 && g - 1 + depth < ngroups) {
            __syncwarp();
            if (lane == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue(g - 1 + depth);
            }
        }
        const int slot = g % depth;
        mbar_wait(bars + 8u * slot, (uint32_t)((g / depth) & 1));
        constexpr int kLanesPerBox = kBox / SB;
        const uint32_t sa = ring + (uint32_t)slot * kStage + (lane / kLanesPerBox) * (kTmaRows * kBox) +
                            (lane % kLanesPerBox) * SB;
        lds(sa, a);
        lds(sa + kBox, b);
        lds(sa + 2 * kBox, c);
        ++g;
    }

    __device__ __forceinline__ static void lds(uint32_t addr, uint32_t (&w)[NW]) {
        if constexpr (NW == 4)
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(addr));
        else if constexpr (NW == 2)
            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "r"(addr));
        else
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[0]) : "r"(addr));
    }
};

template <int K, int NW, int CH, bool CHECK>
__global__ void __launch_bounds__(kWarpsPerCta * 32, B2C_MIN_CTAS)
stream_tma_kernel(const __grid_constant__ CUtensorMap tm, const PassParams p, const int depth) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int wid = blockIdx.x * kWarpsPerCta + warp;
    if (wid >= p.nbands * p.nchunks) return;
    constexpr int kStage = TmaFeed<NW>::kStage;
    const uint32_t ring = smem_u32(smem) + (uint32_t)(warp * depth) * kStage;
    const uint32_t bars = smem_u32(smem) + (uint32_t)(kWarpsPerCta * depth) * kStage + (uint32_t)(warp * depth) * 8u;
    stream_warp<K, NW, CH, CHECK>(p, wid, lane, [&]() { return TmaFeed<NW>(&tm, lane, ring, bars, depth, p.org_x, p.org_y); });
}

// ------------------------------------------------------------------ host side

struct TmaState {
    CUtensorMap map[2][2];  // [buffer][box: 0 = 128 B wide, 1 = 256 B wide]
};

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                             const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeFn get_encode() {
    static EncodeFn fn = nullptr;
    if (fn) return fn;
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = (EncodeFn)p;
    return fn;
}

TmaState *tma_create(uint8_t *const alloc[2], long long pitch, int alloc_rows, long long) {
    EncodeFn enc = get_encode();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled is not available from this driver");
        return nullptr;
    }
    TmaState *t = new TmaState();
    for (int b = 0; b < 2; ++b)
        for (int w = 0; w < 2; ++w) {
            const cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)alloc_rows};
            const cuuint64_t strides[1] = {(cuuint64_t)pitch};
            const cuuint32_t box[2] = {(cuuint32_t)(w ? 256 : 128), (cuuint32_t)kTmaRows};
            const cuuint32_t estr[2] = {1, 1};
            CUresult r = enc(&t->map[b][w], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, alloc[b], dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) {
                set_error("cuTensorMapEncodeTiled failed (%d) pitch=%lld rows=%d", (int)r, pitch, alloc_rows);
                delete t;
                return nullptr;
            }
        }
    return t;
}

void tma_destroy(TmaState *t) { delete t; }

// Ring depth and shared memory for a variant: as many stages as fit next to the
// register-limited number of resident CTAs.
template <int K, int NW, int CH>
static cudaError_t tma_op_t(int op, TmaState *t, const PassParams *p, int src_index, bool check, cudaStream_t s,
                            int *occ_out) {
    auto fn = check ? stream_tma_kernel<K, NW, CH, true> : stream_tma_kernel<K, NW, CH, false>;
    constexpr int kStage = TmaFeed<NW>::kStage;
    // ring depth and the shared-memory opt-in are per (device, variant)
    static int depth_tab[64][2];
    static bool cfg_done[64][2];
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    int dev = 0;
    cudaGetDevice(&dev);
    bool &cfg = cfg_done[dev & 63][check ? 1 : 0];
    int &depth = depth_tab[dev & 63][check ? 1 : 0];
    if (!depth) {
        int by_regs = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&by_regs, fn, kWarpsPerCta * 32, 0);
        if (e != cudaSuccess) return e;
        if (by_regs < 1) by_regs = 1;
        int d = (kSmemBudget / by_regs - 1024 - 64) / (kWarpsPerCta * (kStage + 8));
        if (d > 16) d = 16;
        if (d < 2) d = 2;
        depth = d;
    }
    const size_t smem = (size_t)kWarpsPerCta * depth * (kStage + 8);
    if (!cfg) {
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        cfg = true;
    }
    if (op == 1) return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ_out, fn, kWarpsPerCta * 32, smem);
    if (op == 2) return cudaSuccess;   // cudaFuncSetAttribute above has loaded the function
    const int warps = p->nbands * p->nchunks;
    const int ctas = (warps + kWarpsPerCta - 1) / kWarpsPerCta;
    // programmatic dependent launch, as for the LDG kernel (kernels.cu)
    static const bool pdl = !(getenv("B200CONV_PDL") && atoi(getenv("B200CONV_PDL")) == 0);
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(ctas);
    lc.blockDim = dim3(kWarpsPerCta * 32);
    lc.dynamicSmemBytes = smem;
    lc.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = pdl ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&lc, fn, t->map[src_index][NW == 1 ? 0 : 1], *p, depth);
    count_launch(1);
    return e != cudaSuccess ? e : cudaGetLastError();
}

static cudaError_t tma_op(int op, TmaState *t, const PassParams *p, int src_index, int k, int strip_bytes, int ch,
                          bool check, cudaStream_t s, int *occ) {
#define B2C_CASE(K_, NW_)                                                                  \
    if (k == K_ && strip_bytes == NW_ * 4)                                                 \
        return ch == 1 ? tma_op_t<K_, NW_, 1>(op, t, p, src_index, check, s, occ)          \
                       : tma_op_t<K_, NW_, 3>(op, t, p, src_index, check, s, occ);
    B2C_CASE(1, 1) B2C_CASE(2, 1) B2C_CASE(4, 1) B2C_CASE(8, 1)
    B2C_CASE(1, 2) B2C_CASE(2, 2) B2C_CASE(4, 2) B2C_CASE(8, 2)
    B2C_CASE(1, 4) B2C_CASE(2, 4) B2C_CASE(4, 4)
#undef B2C_CASE
    return cudaErrorInvalidValue;
}

cudaError_t launch_stream_tma(TmaState *t, const PassParams &p, int src_index, int k, int strip_bytes, int ch,
                              bool check, cudaStream_t s) {
    return tma_op(0, t, &p, src_index, k, strip_bytes, ch, check, s, nullptr);
}

cudaError_t preload_stream_tma_kernel(int k, int strip_bytes, int ch, bool check) {
    return tma_op(2, nullptr, nullptr, 0, k, strip_bytes, ch, check, nullptr, nullptr);
}

int stream_tma_occupancy(int k, int strip_bytes, int ch) {
    int occ = 0;
    if (tma_op(1, nullptr, nullptr, 0, k, strip_bytes, ch, false, nullptr, &occ) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return occ;
}

}  // namespace b2c
