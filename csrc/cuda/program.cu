// Whole-program driver for one GPU: read .raw -> blur -> write blur_<name>.
// Behavioural counterpart of cuda/main.c:10-53 (+ GPU_convolution, :49-102):
// same file naming and byte layout; differences are pinned staging buffers,
// 64-bit sizes, CUDA-event loop timing next to the end-to-end wall clock.
#include <cuda_runtime.h>
#include <time.h>

#include <string>

#include "common.h"

#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            b2c::set_error("%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return B2C_ECUDA;                                                            \
        }                                                                                \
    } while (0)

static double now_ms() {  // micro_time(), cuda/functions.c:47-51
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

extern "C" int b2c_blur_file(const char *image, int W, int H, int CH, int reps, const b2c_opts *opts,
                             double *loop_ms, double *total_ms) {
    const double t0 = now_ms();
    if (!image || W <= 0 || H <= 0 || (CH != 1 && CH != 3) || reps < 0) {
        b2c::set_error("blur_file: bad arguments");
        return B2C_EINVAL;
    }
    const long long bytes = (long long)W * H * CH;
    uint8_t *host = nullptr;
    CK(cudaMallocHost(&host, (size_t)bytes));
    int rc = B2C_OK;
    b2c_engine *e = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaStream_t s = nullptr;
    float ms = 0.f;
    char out_name[4096];
    do {
        if (b2c_read_block(image, host, (long long)W * CH, W, H, CH, 0, 0, H, W) < 0) { rc = B2C_EIO; break; }
        int dev = 0;
        cudaGetDevice(&dev);
        e = b2c_engine_create(dev, W, H, CH, 1, 1, 0, opts);
        if (!e) { rc = B2C_ECUDA; break; }
        if (cudaStreamCreate(&s) != cudaSuccess || cudaEventCreate(&ev0) != cudaSuccess ||
            cudaEventCreate(&ev1) != cudaSuccess) { rc = B2C_ECUDA; break; }
        if ((rc = b2c_engine_upload(e, host, 0, s)) != B2C_OK) break;
        if ((rc = b2c_engine_exchange(e, s)) != B2C_OK) break;
        cudaEventRecord(ev0, s);
        if ((rc = b2c_engine_run(e, reps, s)) != B2C_OK) break;
        cudaEventRecord(ev1, s);
        if (b2c_engine_iterations_done(e) < 0) { rc = B2C_ECUDA; break; }
        if ((rc = b2c_engine_download(e, host, 0, s)) != B2C_OK) break;
        if (cudaStreamSynchronize(s) != cudaSuccess) { rc = B2C_ECUDA; b2c::set_error("%s", cudaGetErrorString(cudaGetLastError())); break; }
        cudaEventElapsedTime(&ms, ev0, ev1);
        if (b2c_output_name(image, out_name, sizeof out_name) != B2C_OK) { rc = B2C_EINVAL; break; }
        if (b2c_write_block(out_name, host, (long long)W * CH, W, H, CH, 0, 0, H, W, 1) < 0) { rc = B2C_EIO; break; }
    } while (0);
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    if (e) b2c_engine_destroy(e);
    if (s) cudaStreamDestroy(s);
    cudaFreeHost(host);
    if (loop_ms) *loop_ms = ms;
    if (total_ms) *total_ms = now_ms() - t0;
    return rc;
}
