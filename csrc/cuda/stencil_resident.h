// Register-resident persistent stencil for small images (stencil_resident.cu, B2C_KERNEL_RESIDENT):
// the whole image lives in the register files of at most one CTA per SM for the entire run; CTAs
// trade K-deep rings (K = 4, 8 or 16) through L2 every K iterations.  See the .cu for the design.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace b2c {

struct ResidentPlan {
    int rw;        // rows per warp (template instance)
    int k;         // iterations between two exchanges = halo depth (4, 8 or 16)
    int strips;    // CTA columns
    int ctarows;   // CTA rows
    int vr, vc;    // valid rows / bytes per CTA
};

struct ResidentParams {
    const uint8_t *src;   // image origin (8-byte aligned)
    uint8_t *dst;
    long long src_pitch, dst_pitch;   // multiples of 8
    uint8_t *xbuf[2];     // exchange images, one per epoch parity, H rows of xpitch bytes
    long long xpitch;
    uint32_t *flags;      // one epoch word per CTA, 32 words apart
    unsigned int *error;  // set to 1 when a wait timed out
    uint32_t epoch0;      // epochs published by earlier launches
    int H, bw;            // rows, bytes per row
    int reps;
    int strips, ctarows;
    int torus;
    int debug;            // profiling switches, 0 in production (see the kernel)
};

// Smallest tile that covers a bw x H byte image with at most `sms` CTAs; false when the image is
// too large (or too awkward: bw must be a multiple of 8) for the resident kernel.
bool resident_plan(int bw, int H, int ch, int k, int sms, ResidentPlan *out);

cudaError_t launch_resident(const ResidentPlan &plan, const ResidentParams &p, int ch, cudaStream_t s);

}  // namespace b2c
