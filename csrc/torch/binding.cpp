// Thin PyTorch binding over libb200conv_core: registers `torch.ops.b200conv.*`.
// No compute lives here — tensors in, raw pointers + the current CUDA stream down to the C ABI.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>
#include <torch/torch.h>

#include <string>

#include "b200conv.h"

namespace {

int boundary_code(const std::string &s) {
    if (s == "zero") return B2C_ZERO;
    if (s == "torus") return B2C_TORUS;
    if (s == "cuda-compat" || s == "cuda_compat") return B2C_CUDA_COMPAT;
    TORCH_CHECK(false, "boundary must be zero|torus|cuda-compat, got ", s);
}
int filter_code(const std::string &s) {
    if (s == "gaussian") return B2C_GAUSSIAN;
    if (s == "box") return B2C_BOX;
    if (s == "edge") return B2C_EDGE;
    TORCH_CHECK(false, "filter must be gaussian|box|edge, got ", s);
}
int kernel_code(const std::string &s) {
    if (s == "auto") return B2C_KERNEL_AUTO;
    if (s == "naive") return B2C_KERNEL_NAIVE;
    if (s == "stream") return B2C_KERNEL_STREAM;
    if (s == "tma") return B2C_KERNEL_TMA;
    if (s == "resident") return B2C_KERNEL_RESIDENT;
    TORCH_CHECK(false, "kernel must be auto|naive|stream|tma|resident, got ", s);
}

// blur(img[H,W] | img[H,W,3] uint8 CUDA, reps, boundary, filter, tb, kernel) -> Tensor
at::Tensor blur(const at::Tensor &img, int64_t reps, std::string boundary, std::string filter, int64_t tb,
                std::string kernel) {
    TORCH_CHECK(img.is_cuda(), "b200conv::blur needs a CUDA tensor");
    TORCH_CHECK(img.scalar_type() == at::kByte, "b200conv::blur needs uint8");
    TORCH_CHECK(img.dim() == 2 || (img.dim() == 3 && (img.size(2) == 1 || img.size(2) == 3)),
                "image must be [H,W] or [H,W,3]");
    const c10::cuda::CUDAGuard guard(img.device());
    at::Tensor in = img.contiguous();
    at::Tensor out = at::empty_like(in);
    const int H = (int)in.size(0), W = (int)in.size(1), CH = in.dim() == 2 ? 1 : (int)in.size(2);
    b2c_opts o;
    b2c_default_opts(&o);
    o.boundary = boundary_code(boundary);
    o.filter = filter_code(filter);
    o.kernel = kernel_code(kernel);
    o.tb = (int)tb;
    cudaStream_t s = at::cuda::getCurrentCUDAStream();
    const int rc = b2c_blur_device(in.data_ptr<uint8_t>(), 0, out.data_ptr<uint8_t>(), 0, W, H, CH, (int)reps, &o, s);
    TORCH_CHECK(rc == B2C_OK, "b200conv::blur failed: ", b2c_last_error());
    return out;
}

// CPU tensors take the exact CPU oracle (no GPU needed; `tb` and `kernel` do not apply).
at::Tensor blur_cpu(const at::Tensor &img, int64_t reps, std::string boundary, std::string filter, int64_t /*tb*/,
                    std::string kernel) {
    TORCH_CHECK(img.scalar_type() == at::kByte, "b200conv::blur needs uint8");
    TORCH_CHECK(img.dim() == 2 || (img.dim() == 3 && (img.size(2) == 1 || img.size(2) == 3)),
                "image must be [H,W] or [H,W,3]");
    kernel_code(kernel);   // validate the spelling even though no kernel runs
    at::Tensor in = img.contiguous();
    at::Tensor out = at::empty_like(in);
    const int H = (int)in.size(0), W = (int)in.size(1), CH = in.dim() == 2 ? 1 : (int)in.size(2);
    const int rc = b2c_oracle_blur(in.data_ptr<uint8_t>(), out.data_ptr<uint8_t>(), W, H, CH, (int)reps,
                                   boundary_code(boundary), filter_code(filter));
    TORCH_CHECK(rc == B2C_OK, "b200conv::blur (cpu) failed: ", b2c_last_error());
    return out;
}

// stencil_step(src, dst, rows, row_bytes, channels, filter, threads): one plain iteration on caller-framed
// [rows+2, pitch] uint8 buffers (block at [1:rows+1, channels:channels+row_bytes]); CUDA or CPU by device.
void stencil_step(const at::Tensor &src, at::Tensor dst, int64_t rows, int64_t row_bytes, int64_t channels,
                  std::string filter, int64_t threads) {
    TORCH_CHECK(src.scalar_type() == at::kByte && dst.scalar_type() == at::kByte, "uint8 buffers");
    TORCH_CHECK(src.dim() == 2 && dst.dim() == 2 && src.sizes() == dst.sizes() && src.stride(1) == 1 &&
                    dst.stride(1) == 1 && src.stride(0) == dst.stride(0) && src.device() == dst.device(),
                "src and dst must be equally shaped 2-D row-major buffers on one device");
    TORCH_CHECK(src.size(0) >= rows + 2 && src.stride(0) >= row_bytes + 2 * channels, "frame missing");
    const int64_t pitch = src.stride(0), off = pitch + channels;
    int rc;
    if (src.is_cuda()) {
        const c10::cuda::CUDAGuard guard(src.device());
        cudaStream_t s = at::cuda::getCurrentCUDAStream();
        rc = b2c_stencil_step_device(src.data_ptr<uint8_t>() + off, dst.data_ptr<uint8_t>() + off, pitch, (int)rows,
                                     (int)row_bytes, (int)channels, filter_code(filter), s);
    } else {
        rc = b2c_stencil_step_host(src.data_ptr<uint8_t>() + off, dst.data_ptr<uint8_t>() + off, pitch, (int)rows,
                                   (int)row_bytes, (int)channels, filter_code(filter), (int)threads);
    }
    TORCH_CHECK(rc == B2C_OK, "b200conv::stencil_step failed: ", b2c_last_error());
}

int64_t rows_division(int64_t n, int64_t rows, int64_t cols) { return b2c_rows_division((int)n, (int)rows, (int)cols); }

}  // namespace

TORCH_LIBRARY(b200conv, m) {
    m.def("blur(Tensor img, int reps, str boundary=\"zero\", str filter=\"gaussian\", int tb=0, str kernel=\"auto\") -> Tensor");
    m.def("rows_division(int n, int rows, int cols) -> int", &rows_division);
    m.def("stencil_step(Tensor src, Tensor(a!) dst, int rows, int row_bytes, int channels, str filter=\"gaussian\", "
          "int threads=0) -> ()", &stencil_step);
}
TORCH_LIBRARY_IMPL(b200conv, CUDA, m) { m.impl("blur", &blur); }
TORCH_LIBRARY_IMPL(b200conv, CPU, m) { m.impl("blur", &blur_cpu); }
