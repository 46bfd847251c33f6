/*
 * b200conv — Blackwell-native iterative 3x3 image convolution engine.
 *
 * Public C ABI of libb200conv_core.so.  Everything the launcher, the PyTorch
 * binding and ctypes users call goes through this header.  The library has no
 * torch dependency.
 *
 * Reference parity notes cite files under /root/reference (the upstream
 * theopaid/Parallel-Image-Convolution-using-MPI-OPENMP-and-CUDA tree).
 */
#ifndef B200CONV_H
#define B200CONV_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ enums */

/* Boundary semantics.
 *  ZERO        : cells outside the image are 0 at every iteration
 *                == mpi/mpi_convolution.c at one rank (calloc'd halo, :104-119).
 *  TORUS       : indices wrap modulo H and W (README.md:63-66).
 *  CUDA_COMPAT : cuda/cuda_convolution.cu semantics: interior-only update,
 *                frozen 1-px border alternating original/zero, and the
 *                odd-repetition copy-back bug (:93-97) => 2*floor(reps/2)
 *                effective iterations.  RGB uses the linear-wrap model of the
 *                buggy guard at :34. */
enum { B2C_ZERO = 0, B2C_TORUS = 1, B2C_CUDA_COMPAT = 2 };

/* Filter table of mpi/mpi_convolution.c:89-102.  Only GAUSSIAN is live in the
 * reference; BOX (/9) and EDGE (/28) are the two commented-out alternatives. */
enum { B2C_GAUSSIAN = 0, B2C_BOX = 1, B2C_EDGE = 2 };

/* Neighbour directions, index into the 8-entry neighbour table. */
enum { B2C_N = 0, B2C_S = 1, B2C_W = 2, B2C_E = 3, B2C_NW = 4, B2C_NE = 5, B2C_SW = 6, B2C_SE = 7 };

/* Kernel families selectable through b2c_opts.kernel. */
enum {
    B2C_KERNEL_AUTO = 0,   /* pick by shape                                    */
    B2C_KERNEL_NAIVE = 1,  /* one pass per iteration, byte-per-thread          */
    B2C_KERNEL_STREAM = 2, /* register-pipelined temporal blocking, LDG feed   */
    B2C_KERNEL_TMA = 3,    /* same pipeline fed by TMA bulk-tensor tiles       */
    B2C_KERNEL_RESIDENT = 4 /* small images: whole image held in registers for the run
                              (one GPU, gaussian, zero/torus, row bytes % 8 == 0, no check_every);
                              AUTO picks it whenever the image fits */
};

/* Error codes (negative returns). */
enum {
    B2C_OK = 0,
    B2C_EINVAL = -1,
    B2C_ECUDA = -2,
    B2C_EIO = -3,
    B2C_ENOMEM = -4,
    B2C_ETOOBIG = -5,
    B2C_EDIVIDE = -6,
    B2C_ETIMEOUT = -7
};

const char *b2c_last_error(void);
const char *b2c_version(void);

/* ------------------------------------------------- decomposition (CPU, pure) */

/* Same rule as RowsDivision() mpi/mpi_convolution.c:350-364: smallest block
 * semi-perimeter among exact divisors, ties keep the smallest row count.
 * Returns the number of process-grid rows, 0 when nothing divides. */
int b2c_rows_division(int n_processes, int rows, int columns);

/* P x Q grid for n ranks over an H x W image; returns B2C_EDIVIDE with the
 * reference's "Cannot divide to processes" condition (:53-58). */
int b2c_decompose(int n, int height, int width, int *grid_rows, int *grid_cols);

/* 8-neighbour table of `rank` on a row-major P x Q grid (:72-73,142-149 only
 * has N/S/W/E and never wraps).  Missing neighbour = -1.  TORUS wraps (a rank
 * may be its own neighbour). */
int b2c_neighbours(int rank, int grid_rows, int grid_cols, int boundary, int out[8]);

/* ------------------------------------------------------------ CLI (CPU, pure) */

typedef struct b2c_cli {
    const char *image;  /* argv[1], not copied */
    int width, height, repetitions;
    int channels;       /* 1 = grey, 3 = rgb */
    /* optional trailing flags (extensions; absent in the reference) */
    int gpus;           /* 0 = all visible */
    int boundary;       /* B2C_ZERO default */
    int filter;         /* B2C_GAUSSIAN default */
    int tb;             /* temporal-blocking depth, 0 = auto */
    int kernel;         /* B2C_KERNEL_* */
    int check_every;    /* convergence check period, 0 = off */
    int checkpoint_every; /* write the image every N iterations, 0 = off */
    int profile;        /* per-phase timings on stderr */
    const char *json;   /* path for the JSON report, NULL = none */
    int cpu;            /* --cpu N|all: run on N host threads (-1 = all cores) instead of GPUs; 0 = GPUs */
    int halo_depth;     /* --halo-depth N: iterations between two halo exchanges (several GPUs / torus), 0 = auto */
} b2c_cli;

/* Grammar of Usage() mpi/mpi_convolution.c:328-348 and Initialization()
 * cuda/functions.c:10-29: five positionals, argv[5] in {"grey","rgb"},
 * atoi() for the numbers.  Returns B2C_OK or B2C_EINVAL. */
int b2c_parse_cli(int argc, char **argv, b2c_cli *out);

/* Writes the reference usage text.  style 0 = MPI program wording (:344),
 * style 1 = CUDA program wording (functions.c:26). */
int b2c_usage_text(const char *argv0, int style, char *buf, size_t buflen);

/* "blur_" + argv[1] (whole string, like :244-246 / cuda/main.c:36-38). */
int b2c_output_name(const char *image, char *buf, size_t buflen);

/* ------------------------------------------------------------------ raw I/O */

/* Read/write the R x C block at (row0, col0) of a headerless interleaved uint8
 * W x H x CH image (mpi_convolution.c:126-140,244-262) with 64-bit offsets.
 * `buf` rows are `pitch` bytes apart.  Returns bytes moved or a negative code.
 * Short files are an error (the reference loops forever, functions.c:39-45). */
int64_t b2c_read_block(const char *path, uint8_t *buf, int64_t pitch, int width, int height,
                       int channels, int row0, int col0, int rows, int cols);
int64_t b2c_write_block(const char *path, const uint8_t *buf, int64_t pitch, int width, int height,
                        int channels, int row0, int col0, int rows, int cols, int create);

/* ------------------------------------------------------------- CPU oracle */

/* Integer oracle out = (sum w*p) >> 4 for GAUSSIAN, and the reference's fp32
 * multiply-then-add order (mpi_convolution.c:301-322) for BOX/EDGE.  `in` and
 * `out` are tightly packed H x W*CH; they may alias. */
int b2c_oracle_blur(const uint8_t *in, uint8_t *out, int width, int height, int channels,
                    int repetitions, int boundary, int filter);

/* CPU emulation of the multi-rank halo protocol (same frames, epochs and
 * 8-neighbour pushes as the GPU engine) on a P x Q grid with k-step passes.
 * Result must equal b2c_oracle_blur on every grid (decomposition independence). */
int b2c_oracle_blur_sharded(const uint8_t *in, uint8_t *out, int width, int height, int channels,
                            int repetitions, int boundary, int grid_rows, int grid_cols, int tb);

/* Whole blur on host threads (tight H x W*CH images, `in` and `out` may alias): the no-GPU fallback of
 * the launcher (`--cpu N`) and of the Python package.  ZERO / TORUS run b2c_stencil_step_host on
 * framed ping-pong buffers (threads = 0: all cores), CUDA_COMPAT runs the oracle.  Exact. */
int b2c_blur_host(const uint8_t *in, uint8_t *out, int width, int height, int channels, int repetitions,
                  int boundary, int filter, int threads);

/* Threaded host model of the multi-GPU protocol: one thread per rank, framed
 * ping-pong buffers, per-direction epoch flags (release/acquire), ring pushes
 * into the neighbours' frames, and the tagged-slot convergence exchange —
 * the state machine of the GPU engine restated with std::atomic so that the
 * epoch logic can be tested (and run under ThreadSanitizer) without a GPU.
 * No counterpart in the reference. */
typedef struct b2c_emul_opts {
    int tb;              /* fused iterations per pass (1, 2, 4, 8)                         */
    int check_every;     /* 0 = off                                                         */
    int max_delay_us;    /* random sleep up to this long at every drift point; 0 = none     */
    unsigned seed;       /* for the sleeps                                                  */
    int timeout_ms;      /* a wait longer than this fails the run; 0 = 20 s                 */
    int single_slot_set; /* 1 = model one convergence slot set instead of two (regression) */
    int stall_rank;      /* this rank sleeps stall_ms between publishing and reading its    */
    int stall_ms;        /*   convergence slots; -1 / 0 = nobody                            */
    int images;          /* > 1: that many images back to back through the same buffers with no */
                         /*   barrier in between (image j = input XOR 37*j); needs check_every 0 */
    int halo_depth;      /* iterations between two halo exchanges (multiple of tb, <= 32); 0 = tb  */
    int courier;         /* 1 = epochs reach the neighbours through a second thread per rank (the   */
                         /*   courier warp of the GPU engine) instead of directly                   */
} b2c_emul_opts;

/* Returns B2C_OK (out = the gathered result of every image, max(1, images) x H x W*CH bytes;
 * *iterations_done = iterations run per image),
 * B2C_EINVAL, or B2C_ETIMEOUT when a rank waited past the deadline. */
int b2c_emulate_halo_threads(const uint8_t *in, uint8_t *out, int width, int height, int channels,
                             int repetitions, int boundary, int grid_rows, int grid_cols,
                             const b2c_emul_opts *opts, int *iterations_done);

/* One plain iteration on caller-framed HOST memory (`src`/`dst` point at the
 * interior origin; one halo row above/below and `channels` halo bytes left/right
 * must exist), rows split over `threads` threads (0 = all cores).  With the
 * frame exchanged between processes over torch.distributed/gloo this is the CPU
 * arm of the package (what mpi_convolution.c / omp_convolution.c are to the
 * reference); the GPU twin is b2c_stencil_step_device. */
int b2c_stencil_step_host(const uint8_t *src, uint8_t *dst, int64_t pitch, int rows, int row_bytes,
                          int channels, int filter, int threads);

/* Threaded host model of the resident kernel's CTA protocol (one thread per CTA, private window,
 * shared exchange images by epoch parity, release/acquire epoch flags): out = blurred image.
 * k = exchange interval (4, 8, 16), sms = CTA budget.  B2C_EINVAL when the image does not fit the
 * tile plan, B2C_ETIMEOUT on a deadlock.  No counterpart in the reference. */
int b2c_emulate_resident_threads(const uint8_t *in, uint8_t *out, int width, int height, int channels,
                                 int repetitions, int boundary, int k, int sms, int max_delay_us,
                                 unsigned seed);

/* Threaded host model of the persistent dataflow kernel's scheduling (one rank, ZERO boundary): a
 * queue of (pass, chunk, band) items pulled by `workers` threads, per-tile "passes completed" words,
 * an item starts when the 3x3 neighbourhood of the pass before is done, no barrier between passes.
 * reps must be a multiple of k; vw = valid bytes per band; B2C_ETIMEOUT on a deadlock.  No
 * counterpart in the reference. */
int b2c_emulate_persist_threads(const uint8_t *in, uint8_t *out, int width, int height, int channels,
                                int repetitions, int k, int vw, int chunk_rows, int workers, int max_delay_us,
                                unsigned seed);

/* --------------------------------------------------------------- GPU engine */

typedef struct b2c_opts {
    int boundary;     /* B2C_ZERO / B2C_TORUS / B2C_CUDA_COMPAT */
    int filter;       /* B2C_GAUSSIAN / B2C_BOX / B2C_EDGE       */
    int kernel;       /* B2C_KERNEL_*                            */
    int tb;           /* iterations fused per HBM pass (1,2,4,8); 0 = auto */
    int strip_bytes;  /* bytes per thread per row (4,8,16); 0 = auto       */
    int chunk_rows;   /* rows per warp stream; 0 = auto                    */
    int use_graph;    /* capture the pass loop in a CUDA graph: -1 auto, 0, 1 */
    int check_every;  /* fused convergence check period in iterations, 0 = off (ignored by the
                         naive kernel and cuda-compat, which have no change detector) */
    int halo_depth;   /* multi-rank: iterations between two halo exchanges = depth of the halo a
                         rank keeps (a multiple of tb, at most 32); 0 = auto */
    int persistent;   /* run every pass of b2c_engine_run in ONE persistent kernel whose warps pull
                         (pass, band, chunk) items from a queue and wait on per-tile flags instead of
                         on kernel boundaries: -1 auto (on where it applies), 0 off, 1 on */
} b2c_opts;

void b2c_default_opts(b2c_opts *o);

typedef struct b2c_engine b2c_engine;

/* One engine = one rank's block on one device (cudaSetDevice is the caller's
 * job only for single-rank use; the engine remembers `device`).
 * Grid 1x1/rank 0 is the single-GPU case of cuda/cuda_convolution.cu:49-102.
 * The engine owns two padded ping-pong buffers with a halo frame (the device
 * analogue of the calloc'd (R+2)x(C+2) buffers, mpi_convolution.c:104-119). */
b2c_engine *b2c_engine_create(int device, int width, int height, int channels, int grid_rows,
                              int grid_cols, int rank, const b2c_opts *opts);
void b2c_engine_destroy(b2c_engine *e);

/* Geometry of this rank's block. */
int b2c_engine_block(const b2c_engine *e, int *row0, int *col0, int *rows, int *cols);

/* Peer wiring.  export: opaque bytes (cudaIpc handles for both buffers and the
 * flag pad) to ship to the other ranks; import: all ranks' blobs in rank order.
 * For ranks living in the same process, b2c_engine_connect_local wires raw
 * pointers (used with cudaDeviceEnablePeerAccess, and by the one-GPU
 * sequential multi-rank simulation in the tests). */
#define B2C_HANDLE_BYTES 256
int b2c_engine_export(b2c_engine *e, void *blob /* B2C_HANDLE_BYTES */);
int b2c_engine_import(b2c_engine *e, const void *blobs /* n * B2C_HANDLE_BYTES */, int n);
int b2c_engine_connect_local(b2c_engine **engines, int n);
/* Unmap every imported peer buffer.  Tear-down order for cudaIpc peers (freeing exported
 * memory that a peer still has mapped is undefined behaviour): every rank disconnects ->
 * host-level barrier -> every rank destroys its engine. */
int b2c_engine_disconnect(b2c_engine *e);

/* Data movement for this rank's block (tight or pitched host/device memory). */
int b2c_engine_upload(b2c_engine *e, const uint8_t *host, int64_t pitch, void *stream);
int b2c_engine_download(b2c_engine *e, uint8_t *host, int64_t pitch, void *stream);
int b2c_engine_load_device(b2c_engine *e, const uint8_t *dev, int64_t pitch, void *stream);
int b2c_engine_store_device(b2c_engine *e, uint8_t *dev, int64_t pitch, void *stream);

/* Initial halo exchange (epoch 0) — call on every rank after upload, before run.
 * No-op for a 1x1 ZERO/CUDA_COMPAT engine. */
int b2c_engine_exchange(b2c_engine *e, void *stream);

/* Run `repetitions` iterations on `stream` (asynchronous).  No MPI/NCCL/
 * cudaMemcpyPeer inside: halos move by remote stores + release/acquire flags
 * from inside the stencil kernel (replaces mpi_convolution.c:156-240). */
int b2c_engine_run(b2c_engine *e, int repetitions, void *stream);

/* b2c_engine_run cut into its passes, for callers that interleave several engines on ONE device
 * (tests/parallel/local.py): plan_run returns the number of passes of a `repetitions` run,
 * run_pass enqueues pass i.  Every rank must have enqueued pass i before any rank enqueues i+1. */
int b2c_engine_plan_run(b2c_engine *e, int repetitions);
int b2c_engine_run_pass(b2c_engine *e, int i, void *stream);

/* Coupled tuning of the rows-per-warp of a multi-rank engine (collective: every rank makes the
 * same calls in the same order, after the peers are connected and before data is loaded).
 * tune_count: number of candidates (0: nothing to tune); tune_time: milliseconds of a short
 * coupled run with candidate i on this rank (reduce with MAX over ranks); tune_set: adopt
 * candidate i (-1 = the modelled plan).  No counterpart in the reference. */
int b2c_engine_tune_count(const b2c_engine *e);
double b2c_engine_tune_time(b2c_engine *e, int i, void *stream);
int b2c_engine_tune_set(b2c_engine *e, int i);

/* Device-side barrier over peer memory (replaces MPI_Barrier in front of the timer,
 * mpi_convolution.c:151): work enqueued on `stream` after it starts within microseconds on all
 * ranks.  Collective.  barrier_wait_cycles: SM clocks this rank spent in the last one (synchronises). */
int b2c_engine_barrier(b2c_engine *e, void *stream);
/* NVLS multicast for the two collectives (replaces MPI_Barrier / the max-reduction's role of
 * mpi_convolution.c:151,264-275): `local` = this rank's replica of a zero-initialised symmetric
 * buffer of >= 64 32-bit words that every rank of the grid has mapped, `mc` = its multicast address
 * (e.g. from torch.distributed._symmetric_memory).  Afterwards the convergence OR and
 * b2c_engine_barrier are one multimem.red.add + one poll.  Collective; the caller keeps the memory
 * alive; nulls switch back to peer stores. */
int b2c_engine_set_nvls(b2c_engine *e, uint32_t *local, uint32_t *mc, int words);
int64_t b2c_engine_barrier_wait_cycles(b2c_engine *e);

/* Opt-in wait-time counters of the halo protocol (also B200CONV_HALO_PROF=1): 40 words,
 * [0..7] cycles ring warps waited for the data epoch per direction, [8..15] same for the ack
 * epoch, [16] number of waits, [17] cycles spent in fence + count-in + publish, [18] number of
 * those, [24..31] / [32..39] longest single data / ack wait per direction. */
int b2c_engine_set_profile(b2c_engine *e, int on);
/* With B200CONV_PERSIST_PROF=1 in the environment: SM cycles summed over the persistent kernel's
 * warps since the last call {waiting for tile dependencies, inside tiles, total, items, items
 * that had to wait}.  Synchronises. */
int b2c_engine_persist_profile(b2c_engine *e, uint64_t out[5]);
int b2c_engine_halo_profile(b2c_engine *e, uint64_t out[40], int reset);

/* Iterations actually executed by the last run (differs from `repetitions`
 * only with CUDA_COMPAT odd counts or a convergence early-exit).  Synchronises. */
int b2c_engine_iterations_done(b2c_engine *e);

/* Kernel launches issued by this engine since creation. */
/* Tile plan of the resident kernel (host only, no GPU needed): 1 and out = {rows per warp,
 * strips, CTA rows, valid rows, valid bytes} when a row_bytes x rows image fits on `sms` SMs with
 * an exchange every k (4, 8, 16) iterations, else 0. */
int b2c_resident_plan(int row_bytes, int rows, int channels, int k, int sms, int out[5]);

int64_t b2c_engine_launches(const b2c_engine *e);
/* Human-readable plan (kernel, tb, strip, chunks, grid) for reports. */
int b2c_engine_describe(const b2c_engine *e, char *buf, size_t buflen);

/* Test hook: busy-wait `cycles` SM clocks before every flag
 * wait on this rank to shake out slot-reuse races. */
int b2c_engine_set_delay(b2c_engine *e, int64_t cycles);

/* ------------------------------------------------------- one-shot device API */

/* Blur a tightly-or-pitched device image in one call (allocates the engine's
 * scratch buffers internally, frees them before returning).  Single GPU. */
int b2c_blur_device(const uint8_t *d_in, int64_t in_pitch, uint8_t *d_out, int64_t out_pitch,
                    int width, int height, int channels, int repetitions, const b2c_opts *opts,
                    void *stream);

/* Same result as b2c_blur_device (gaussian; boundary B2C_ZERO or B2C_TORUS) for callers
 * that must not allocate device memory.  This entry point has no counterpart in the
 * reference, whose GPU program cudaMallocs on every run.
 *
 * Who needs it: a filter running inside a host application that owns all device memory and
 * forbids plugins to allocate; a per-frame or real-time loop that cannot afford the
 * device-wide stall and unbounded latency of cudaMalloc/cudaFree; a job with a fixed device
 * memory budget that must not fail mid-run on a refused allocation.
 *
 * The call performs no cudaMalloc/cudaFree and no allocating library call.  Its two work
 * buffers and flag words live in a fixed 32 MiB scratch array that is part of the library
 * image (every GPU that loads the library holds it until the process exits).  Limit: the
 * padded image must fit, 2 * (round_up(W*CH + 256, 128) * (H + 80) + 1024) + 8192 <= 32 MiB,
 * which covers images up to about 1920x2520 rgb; larger ones are rejected with
 * B2C_ETOOBIG and a message from b2c_last_error().
 *
 * The first call on a device looks up the scratch array's device address (with that device
 * current) and caches it; make that first call before starting any stream capture.
 * The work is enqueued on `stream` and the function returns without synchronising.  All
 * calls share the one scratch area: it is cleared at the start of each call, and calls
 * whose GPU work overlaps in time (different streams) race with each other — keep them on
 * one stream or serialise them yourself.  Host threads calling at the same moment are
 * serialised inside the function while they enqueue, so sharing one stream between threads
 * is safe.  `in_pitch` / `out_pitch` <= 0 mean tight rows. */
int b2c_blur_device_noalloc(const uint8_t *d_in, int64_t in_pitch, uint8_t *d_out, int64_t out_pitch,
                            int width, int height, int channels, int repetitions, int boundary,
                            void *stream);

/* One plain iteration on caller-owned framed memory: `src`/`dst` point at the
 * interior origin of buffers that have at least one halo row above/below and
 * `channels` halo bytes left/right (the caller fills the halo, e.g. with NCCL
 * send/recv — this is what the NCCL baseline uses).  Byte-per-thread kernel. */
int b2c_stencil_step_device(const uint8_t *src, uint8_t *dst, int64_t pitch, int rows, int row_bytes,
                            int channels, int filter, void *stream);

/* Test hook (host only): work item `id` -> (band, chunk) in the ring-first order the streaming kernels
 * use.  Lets a CPU test prove that every plan visits each (band, chunk) exactly once. */
int b2c_debug_map_work(int id, int nbands, int nchunks, int nbw, int nbe, int *band, int *chunk);

/* Test hook (host only): the streaming kernels' plan for one rank's block of rows x cols pixels.
 * out = {hlo, vw, nbands, nchunks, chunk_rows, nbw, nbe, K, strip_bytes, halo_depth}. */
int b2c_debug_band_plan(int rows, int cols, int channels, int has_neighbours, int tb, int strip_bytes,
                        int chunk_rows, int kernel, int sm_count, int out[10]);

/* Test hook (host only): the passes of a `reps`-iteration run of a rank with eight neighbours and
 * the plan of each; 16 ints per pass {k, ext, mode, group_end, rows, row_bytes, nbands, nchunks,
 * chunk_rows, last_chunk_rows, nbw, nbe, hlo, vw, xn, xwb}.  Returns the number of passes. */
int b2c_debug_pass_plans(int rows, int cols, int channels, int has_neighbours, int tb, int strip_bytes,
                         int chunk_rows, int halo_depth, int reps, int sm_count, int *out, int max_passes);

/* Whole-program equivalent of cuda/main.c:10-53 (read file, blur, write
 * blur_<name>) on one GPU; returns loop milliseconds in *loop_ms. */
int b2c_blur_file(const char *image, int width, int height, int channels, int repetitions,
                  const b2c_opts *opts, double *loop_ms, double *total_ms);

/* Total kernel launches issued by the library in this process. */
int64_t b2c_total_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* B200CONV_H */
