// Host-visible launch interface of kernels.cu / stencil_tma.cu.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace b2c {

struct PassParams;
struct PersistParams;

struct NaiveWeights {
    int iw[3][3];
    float fw[3][3];
    int integer16;
};

struct PushParams {
    const uint8_t *src;      // interior origin of the buffer whose ring is pushed
    long long pitch;
    int rows, bw, ch, depth;
    uint8_t *peer_dst[8];    // same buffer on the neighbour (interior origin) or null
    uint32_t *peer_flag[8];
    uint32_t *flags;
    uint32_t *counters;
    uint32_t post_epoch;     // published value, relative to *epoch_base when that is set
    uint32_t wait_epoch;
    int wait_epoch_valid;
    const uint32_t *epoch_base;      // device word holding the epoch at the start of the run (null: 0)
    uint32_t *epoch_word;            // when set, block (0,0) stores post_epoch there (initial exchange)
    const unsigned int *converged;   // as PassParams::converged: the push is skipped like the passes of its interval
    unsigned int interval;
#ifdef B2C_DEBUG
    long long dbg_before, dbg_after; // bytes of an allocation before / after its interior origin
    int dbg_rank;
#endif
};

constexpr int kMaxRanks = 64;

struct ConvergeParams {
    unsigned int *changed;        // this rank's OR-accumulated bit (reset here)
    uint32_t *slots;              // our slot array: 2 parities x kMaxRanks words
    uint32_t *peer_slots[kMaxRanks];  // every rank's slot array (index = rank)
    int nranks, rank;
    unsigned int seq;             // check sequence number (>= 1)
    unsigned int *converged;      // latched: number of the check (1-based within the run) that found the fixed point
    unsigned int *converged_at;   // iterations completed when it latched
    unsigned int check_no;        // this check's number within the run
    unsigned int iters_after;
};

__host__ __device__ inline int kDirDrDev(int d) { return d == 0 || d == 4 || d == 5 ? -1 : (d == 1 || d == 6 || d == 7 ? 1 : 0); }
__host__ __device__ inline int kDirDcDev(int d) { return d == 2 || d == 4 || d == 6 ? -1 : (d == 3 || d == 5 || d == 7 ? 1 : 0); }

bool stream_config_ok(int k, int strip_bytes);
int stream_occupancy(int k, int strip_bytes, int ch);
cudaError_t preload_stream_kernel(int k, int strip_bytes, int ch, bool check);   // load the variant now (lazy loading)  // resident CTAs (of kWarpsPerCta warps) per SM
cudaError_t launch_stream(const PassParams &p, int k, int strip_bytes, int ch, bool check, cudaStream_t s);
// Persistent dataflow kernel (stencil_stream.cuh): every pass of a run in one launch of `ctas` CTAs.
cudaError_t launch_stream_persist(const PersistParams &q, int k, int strip_bytes, int ch, int ctas, cudaStream_t s);
int stream_persist_occupancy(int k, int strip_bytes, int ch);
cudaError_t launch_naive(const uint8_t *src, uint8_t *dst, long long pitch, int rows, int bw, int ch,
                         const NaiveWeights &w, cudaStream_t s);
cudaError_t launch_compat(const uint8_t *src, uint8_t *dst, int W, int H, int ch, cudaStream_t s);
cudaError_t launch_halo_push(const PushParams &q, cudaStream_t s);
cudaError_t launch_halo_wait(const uint32_t *flags, unsigned mask, uint32_t epoch, cudaStream_t s);
cudaError_t launch_converge(const ConvergeParams &c, cudaStream_t s);
cudaError_t launch_set_u32(unsigned int *p, unsigned int v, cudaStream_t s);
cudaError_t launch_add_u32(unsigned int *p, unsigned int v, cudaStream_t s);
// Device-side start barrier over peer memory: every rank stores `seq` into its word of every
// rank's slot array and spins until all words of its own array carry `seq`.
struct BarrierParams {
    uint32_t *slots;                  // our array: kMaxRanks words
    uint32_t *peer_slots[kMaxRanks];  // every rank's array (index = rank)
    int nranks, rank;
    uint32_t seq;
    unsigned long long *skew_cycles;  // optional: cycles this rank spent waiting (written by lane 0)
};
cudaError_t launch_barrier(const BarrierParams &b, cudaStream_t s);

// NVLS variants (the box exposes multicast: one multimem.red lands in every rank's replica of a small
// symmetric buffer, reduced in the switch).  Word layout of that buffer, in 32-bit words:
//   [0], [1]   convergence slots by check parity: sum over ranks of (0x10000 + changed bit)
//   [32]       start-barrier arrivals (never reset: the n-th barrier waits for n * nranks)
constexpr int kNvlsConvergeWord = 0;
constexpr int kNvlsBarrierWord = 32;
constexpr int kNvlsWords = 64;
struct NvlsParams {
    uint32_t *local;   // this rank's replica
    uint32_t *mc;      // multicast address of the same buffer
    int nranks, rank;
};
cudaError_t launch_converge_nvls(const ConvergeParams &c, const NvlsParams &n, cudaStream_t s);
cudaError_t launch_barrier_nvls(const NvlsParams &n, uint32_t seq, unsigned long long *skew_cycles, cudaStream_t s);

struct CourierParams;   // stencil_stream.cuh
cudaError_t launch_courier(const CourierParams &c, cudaStream_t s);

long long total_launches();
void count_launch(long long n);

}  // namespace b2c
