// b200conv — native launcher with the reference's command line.
//
//   b200conv image.raw width height repetitions grey|rgb [--gpus N] [--boundary ...] ... [--cpu N|all]
//
// Replaces `mpiexec -n N ./mpi_convolution ...` (README.md:19-23) and `./cuda_convolution ...`
// (README.md:32-34): the launcher forks one worker process per GPU (the MPI rank analogue,
// mpi_convolution.c:22-25), the workers exchange 256-byte peer handles through an anonymous
// shared mapping (instead of MPI_Init/Bcast), read their block of the .raw file, run the fused
// halo+stencil engine, write their block of blur_<image>, and report their device-timed loop;
// the parent prints the maximum as "%f\n" (mpi_convolution.c:264-275).  With one GPU it also
// prints the CUDA program's end-to-end line (cuda/main.c:47-49).
// The parent never touches CUDA (forking an initialised CUDA context is not allowed).
#include <cuda_runtime.h>
#include <sched.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "b200conv.h"

namespace {

constexpr int kMaxWorkers = 64;

struct Shared {
    std::atomic<int> arrived;
    std::atomic<int> generation;
    std::atomic<int> failed;
    unsigned char blobs[kMaxWorkers][B2C_HANDLE_BYTES];
    double loop_ms[kMaxWorkers];
    double phase_ms[kMaxWorkers][5];  // read, upload, loop, download, write
    int iterations[kMaxWorkers];
    char plan[kMaxWorkers][512];
    double tune_ms[kMaxWorkers][32];      // coupled tuning: every rank's time per candidate
    long long start_wait[kMaxWorkers];    // SM cycles each rank spent in the device-side start barrier
};

double now_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec * 1e3 + (double)ts.tv_nsec * 1e-6;
}

// Sense-reversing barrier over the shared mapping (the MPI_Barrier of :151).
bool barrier(Shared *sh, int n) {
    const int gen = sh->generation.load();
    if (sh->arrived.fetch_add(1) + 1 == n) {
        sh->arrived.store(0);
        sh->generation.fetch_add(1);
        return !sh->failed.load();
    }
    const double t0 = now_ms();
    while (sh->generation.load() == gen) {
        if (sh->failed.load()) return false;
        if (now_ms() - t0 > 120e3) return false;  // a peer died
        sched_yield();
    }
    return !sh->failed.load();
}

#define WCK(call)                                                                           \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess) {                                                            \
            fprintf(stderr, "b200conv[%d]: %s: %s\n", rank, #call, cudaGetErrorString(e_)); \
            return 1;                                                                       \
        }                                                                                   \
    } while (0)
#define WB2C(call)                                                                    \
    do {                                                                              \
        if ((call) < 0) {                                                             \
            fprintf(stderr, "b200conv[%d]: %s: %s\n", rank, #call, b2c_last_error()); \
            return 1;                                                                 \
        }                                                                             \
    } while (0)

int count_gpus_in_child() {
    int fds[2];
    if (pipe(fds) != 0) return 0;
    pid_t pid = fork();
    if (pid == 0) {
        close(fds[0]);
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess) n = 0;
        ssize_t ignored = write(fds[1], &n, sizeof n);
        (void)ignored;
        _exit(0);
    }
    close(fds[1]);
    int n = 0;
    if (read(fds[0], &n, sizeof n) != (ssize_t)sizeof n) n = 0;
    close(fds[0]);
    int st;
    waitpid(pid, &st, 0);
    return n;
}

int worker(const b2c_cli &cli, int rank, int n, int P, int Q, Shared *sh) {
    const int W = cli.width, H = cli.height, CH = cli.channels;
    b2c_opts o;
    b2c_default_opts(&o);
    o.boundary = cli.boundary;
    o.filter = cli.filter;
    o.kernel = cli.kernel;
    o.tb = cli.tb;
    o.check_every = cli.check_every;
    o.halo_depth = cli.halo_depth;
    WCK(cudaSetDevice(rank));
    b2c_engine *e = b2c_engine_create(rank, W, H, CH, P, Q, rank, &o);
    if (!e) {
        fprintf(stderr, "b200conv[%d]: %s\n", rank, b2c_last_error());
        return 1;
    }
    int row0, col0, R, C;
    b2c_engine_block(e, &row0, &col0, &R, &C);
    const long long row_bytes = (long long)C * CH;
    uint8_t *host = nullptr;
    WCK(cudaMallocHost(&host, (size_t)(row_bytes * R)));
    cudaStream_t s;
    WCK(cudaStreamCreate(&s));
    cudaEvent_t ev0, ev1;
    WCK(cudaEventCreate(&ev0));
    WCK(cudaEventCreate(&ev1));

    double t = now_ms();
    WB2C(b2c_read_block(cli.image, host, row_bytes, W, H, CH, row0, col0, R, C));   // :126-140
    sh->phase_ms[rank][0] = now_ms() - t;

    if (n > 1) {
        WB2C(b2c_engine_export(e, sh->blobs[rank]));
        if (!barrier(sh, n)) return 1;
        WB2C(b2c_engine_import(e, sh->blobs, n));
        if (!barrier(sh, n)) return 1;  // every peer mapped
        // rows per warp: a few candidates timed with the halo exchange live, all ranks together;
        // the slowest rank's time decides (the buffers hold no image yet)
        const int nc = b2c_engine_tune_count(e);
        if (nc > 0) {
            for (int i = 0; i < nc && i < 32; ++i) {
                const double ms = b2c_engine_tune_time(e, i, s);
                if (ms < 0) {
                    fprintf(stderr, "b200conv[%d]: %s\n", rank, b2c_last_error());
                    return 1;
                }
                sh->tune_ms[rank][i] = ms;
            }
            if (!barrier(sh, n)) return 1;
            int best = 0;
            double best_ms = 1e300;
            for (int i = 0; i < nc && i < 32; ++i) {
                double worst = 0;
                for (int k = 0; k < n; ++k) worst = sh->tune_ms[k][i] > worst ? sh->tune_ms[k][i] : worst;
                if (worst < best_ms) { best_ms = worst; best = i; }
            }
            WB2C(b2c_engine_tune_set(e, best));
            if (!barrier(sh, n)) return 1;
        }
    }
    t = now_ms();
    WB2C(b2c_engine_upload(e, host, 0, s));
    WCK(cudaStreamSynchronize(s));
    sh->phase_ms[rank][1] = now_ms() - t;
    if (n > 1 && !barrier(sh, n)) return 1;  // every block uploaded before a frame is written
    WB2C(b2c_engine_exchange(e, s));
    WCK(cudaStreamSynchronize(s));
    if (n > 1) {
        if (!barrier(sh, n)) return 1;       // MPI_Barrier before the timer, :151-154 ...
        WB2C(b2c_engine_barrier(e, s));      // ... finished on the device: the streams start within microseconds
    }

    char out_name[4096];
    b2c_output_name(cli.image, out_name, sizeof out_name);
    const int seg = cli.checkpoint_every > 0 ? cli.checkpoint_every : cli.repetitions;
    int done = 0, executed = 0;
    float loop_ms = 0.f;
    double dl_ms = 0, wr_ms = 0;
    do {
        const int k = cli.repetitions - done < seg ? cli.repetitions - done : seg;
        WCK(cudaEventRecord(ev0, s));
        WB2C(b2c_engine_run(e, k, s));
        WCK(cudaEventRecord(ev1, s));
        const int it = b2c_engine_iterations_done(e);
        if (it < 0) {
            fprintf(stderr, "b200conv[%d]: %s\n", rank, b2c_last_error());
            return 1;
        }
        float ms = 0.f;
        WCK(cudaEventElapsedTime(&ms, ev0, ev1));
        loop_ms += ms;
        done += k;
        executed += it;
        // the image is the whole state: writing it is the checkpoint
        t = now_ms();
        WB2C(b2c_engine_download(e, host, 0, s));
        WCK(cudaStreamSynchronize(s));
        dl_ms += now_ms() - t;
        t = now_ms();
        WB2C(b2c_write_block(out_name, host, row_bytes, W, H, CH, row0, col0, R, C, 1));  // :244-262
        wr_ms += now_ms() - t;
        if (it < k) break;  // converged: nothing changes any more
    } while (done < cli.repetitions);

    sh->loop_ms[rank] = loop_ms;
    sh->phase_ms[rank][2] = loop_ms;
    sh->phase_ms[rank][3] = dl_ms;
    sh->phase_ms[rank][4] = wr_ms;
    sh->iterations[rank] = executed;
    sh->start_wait[rank] = n > 1 ? (long long)b2c_engine_barrier_wait_cycles(e) : 0;
    b2c_engine_describe(e, sh->plan[rank], sizeof sh->plan[rank]);
    if (n > 1) {
        // cudaIpc tear-down order: all writes done -> everyone unmaps -> everyone frees
        if (!barrier(sh, n)) return 1;
        WB2C(b2c_engine_disconnect(e));
        if (!barrier(sh, n)) return 1;
    }
    b2c_engine_destroy(e);
    cudaFreeHost(host);
    return 0;
}

// File names are arbitrary bytes: quote what JSON cannot carry verbatim.
std::string json_escape(const char *in) {
    std::string out;
    for (const unsigned char *p = (const unsigned char *)in; *p; ++p) {
        if (*p == '"' || *p == '\\') {
            out += '\\';
            out += (char)*p;
        } else if (*p < 0x20) {
            char buf[8];
            snprintf(buf, sizeof buf, "\\u%04x", *p);
            out += buf;
        } else {
            out += (char)*p;
        }
    }
    return out;
}

void write_json(const b2c_cli &cli, int n, int P, int Q, const Shared *sh, double max_ms, double total_s) {
    FILE *f = fopen(cli.json, "w");
    if (!f) {
        fprintf(stderr, "cannot open %s\n", cli.json);
        return;
    }
    const int it = sh->iterations[0];
    const double pix_it = (double)cli.width * cli.height * it;
    const double gpix = max_ms > 0 ? pix_it / (max_ms * 1e-3) / 1e9 : 0;
    const int R = cli.height / P, C = cli.width / Q;
    fprintf(f, "{\"image\": \"%s\", \"width\": %d, \"height\": %d, \"channels\": %d, \"repetitions\": %d, "
               "\"iterations_done\": %d, \"gpus\": %d, \"grid\": [%d, %d], \"boundary\": %d, "
               "\"loop_seconds\": %.6f, \"total_seconds\": %.6f, \"gpix_it_per_s\": %.3f, "
               "\"hbm_bytes_per_iteration_k1\": %.0f, \"halo_bytes_per_gpu_per_iteration\": %d, "
               "\"pct_of_k1_hbm_roofline_6555GBs\": %.2f, \"plan\": \"%s\", \"per_rank_loop_ms\": [",
            json_escape(cli.image).c_str(), cli.width, cli.height, cli.channels, cli.repetitions, it, n, P, Q,
            cli.boundary,
            max_ms * 1e-3, total_s, gpix, 2.0 * cli.width * cli.height * cli.channels,
            n > 1 ? 2 * C * cli.channels + 2 * R * cli.channels + 4 * cli.channels : 0,
            gpix > 0 ? 100.0 * gpix / (6555.5 / 2 / cli.channels * n) : 0.0, json_escape(sh->plan[0]).c_str());
    for (int k = 0; k < n; ++k) fprintf(f, "%s%.4f", k ? ", " : "", sh->loop_ms[k]);
    fprintf(f, "], \"start_barrier_wait_cycles\": [");
    for (int k = 0; k < n; ++k) fprintf(f, "%s%lld", k ? ", " : "", sh->start_wait[k]);
    fprintf(f, "]}\n");
    fclose(f);
}

// --cpu N|all: the whole program on host threads, no CUDA call at all (a machine without a GPU, or a
// quick check).  Same files, same output lines; --gpus, --kernel, --tb, --check-every and
// --checkpoint-every do not apply.  This is the role the reference's MPI / MPI+OpenMP programs play.
int run_on_host(const b2c_cli &cli, double t_start) {
    const long long row = (long long)cli.width * cli.channels, bytes = row * cli.height;
    if (cli.width <= 0 || cli.height <= 0 || row > (1ll << 30)) {
        fprintf(stderr, "bad image size\n");
        return EXIT_FAILURE;
    }
    std::vector<uint8_t> img((size_t)bytes);
    double t = now_ms();
    if (b2c_read_block(cli.image, img.data(), row, cli.width, cli.height, cli.channels, 0, 0, cli.height, cli.width) < 0) {
        fprintf(stderr, "%s\n", b2c_last_error());
        return EXIT_FAILURE;
    }
    const double read_ms = now_ms() - t;
    t = now_ms();
    if (b2c_blur_host(img.data(), img.data(), cli.width, cli.height, cli.channels, cli.repetitions, cli.boundary,
                      cli.filter, cli.cpu < 0 ? 0 : cli.cpu) != B2C_OK) {
        fprintf(stderr, "%s\n", b2c_last_error());
        return EXIT_FAILURE;
    }
    const double loop_ms = now_ms() - t;
    char out_name[4096];
    b2c_output_name(cli.image, out_name, sizeof out_name);
    t = now_ms();
    if (b2c_write_block(out_name, img.data(), row, cli.width, cli.height, cli.channels, 0, 0, cli.height, cli.width, 1) < 0) {
        fprintf(stderr, "%s\n", b2c_last_error());
        return EXIT_FAILURE;
    }
    const double write_ms = now_ms() - t;
    const double total_s = (now_ms() - t_start) * 1e-3;
    printf("%f\n", loop_ms * 1e-3);                       // mpi_convolution.c:274
    printf("Execution time: %.3f sec\n", total_s);        // cuda/main.c:49
    if (cli.profile)
        fprintf(stderr, "host: read=%.3fms loop=%.3fms write=%.3fms threads=%d\n", read_ms, loop_ms, write_ms, cli.cpu);
    if (cli.json) {
        FILE *f = fopen(cli.json, "w");
        if (f) {
            fprintf(f, "{\"image\": \"%s\", \"width\": %d, \"height\": %d, \"channels\": %d, \"repetitions\": %d, "
                       "\"device\": \"host\", \"threads\": %d, \"loop_seconds\": %.6f, \"total_seconds\": %.6f, "
                       "\"gpix_it_per_s\": %.3f}\n",
                    json_escape(cli.image).c_str(), cli.width, cli.height, cli.channels, cli.repetitions, cli.cpu,
                    loop_ms * 1e-3, total_s,
                    loop_ms > 0 ? (double)cli.width * cli.height * cli.repetitions / (loop_ms * 1e-3) / 1e9 : 0.0);
            fclose(f);
        }
    }
    return EXIT_SUCCESS;
}

}  // namespace

int main(int argc, char **argv) {
    const double t_start = now_ms();
    b2c_cli cli;
    if (b2c_parse_cli(argc, argv, &cli) != B2C_OK) {
        char buf[1024];
        b2c_usage_text(argv[0], 0, buf, sizeof buf);  // wording of Usage(), mpi_convolution.c:344
        fputs(buf, stderr);
        return EXIT_FAILURE;
    }
    if (cli.cpu != 0) {
        if (access(cli.image, R_OK) != 0) {
            fprintf(stderr, "cannot open %s\n", cli.image);  // cuda/main.c:24
            return EXIT_FAILURE;
        }
        return run_on_host(cli, t_start);
    }
    int n = cli.gpus;
    int P = 0, Q = 0;
    if (n <= 0) {
        // default = every visible GPU that the image divides over (the `mpiexec -n` analogue)
        n = count_gpus_in_child();
        if (n <= 0) {
            fprintf(stderr, "%s: no CUDA device (add --cpu N to run on host threads)\n", argv[0]);
            return EXIT_FAILURE;
        }
        if (n > kMaxWorkers) n = kMaxWorkers;
        if (cli.boundary == B2C_CUDA_COMPAT) n = 1;
        while (n > 1 && b2c_decompose(n, cli.height, cli.width, &P, &Q) != B2C_OK) --n;
    }
    if (n > kMaxWorkers || b2c_decompose(n, cli.height, cli.width, &P, &Q) != B2C_OK) {
        fprintf(stderr, "%s: Cannot divide to processes\n", argv[0]);  // :55
        return EXIT_FAILURE;
    }
    if (access(cli.image, R_OK) != 0) {
        fprintf(stderr, "cannot open %s\n", cli.image);  // cuda/main.c:24
        return EXIT_FAILURE;
    }

    void *mem = mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (mem == MAP_FAILED) {
        perror("mmap");
        return EXIT_FAILURE;
    }
    Shared *sh = new (mem) Shared();  // anonymous mappings start zeroed; this also starts the atomics' lifetime

    std::vector<pid_t> pids;
    for (int k = 0; k < n; ++k) {
        pid_t pid = fork();
        if (pid < 0) {
            perror("fork");
            sh->failed.store(1);
            break;
        }
        if (pid == 0) {
            int rc = worker(cli, k, n, P, Q, sh);
            if (rc) sh->failed.store(1);
            fflush(nullptr);
            _exit(rc);
        }
        pids.push_back(pid);
    }
    // Reap; on the first failure take the whole group down (failure detection).
    int bad = sh->failed.load();
    size_t left = pids.size();
    while (left > 0) {
        int st = 0;
        pid_t p = wait(&st);
        if (p < 0) {
            if (errno == EINTR) continue;
            break;
        }
        --left;
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) {
            if (!bad) {
                bad = 1;
                sh->failed.store(1);
                for (pid_t q : pids)
                    if (q != p) kill(q, SIGTERM);
            }
        }
    }
    if (bad || pids.size() != (size_t)n) {
        fprintf(stderr, "%s: a worker failed\n", argv[0]);
        return EXIT_FAILURE;
    }
    double max_ms = 0;
    for (int k = 0; k < n; ++k)
        if (sh->loop_ms[k] > max_ms) max_ms = sh->loop_ms[k];
    const double total_s = (now_ms() - t_start) * 1e-3;
    printf("%f\n", max_ms * 1e-3);                                  // mpi_convolution.c:274
    if (n == 1) printf("Execution time: %.3f sec\n", total_s);      // cuda/main.c:49
    if (cli.profile) {
        static const char *names[5] = {"read", "upload", "loop", "download", "write"};
        for (int k = 0; k < n; ++k) {
            fprintf(stderr, "rank %d:", k);
            for (int i = 0; i < 5; ++i) fprintf(stderr, " %s=%.3fms", names[i], sh->phase_ms[k][i]);
            fprintf(stderr, " iterations=%d\n", sh->iterations[k]);
        }
        fprintf(stderr, "plan: %s\n", sh->plan[0]);
    }
    if (cli.json) write_json(cli, n, P, Q, sh, max_ms, total_s);
    return EXIT_SUCCESS;
}
