// Register-resident persistent stencil for small images (stencil_resident.cu, B2C_KERNEL_RESIDENT):
// the whole image lives in the register files of at most one CTA per SM for the entire run; CTAs
// trade K-deep rings (K = 4, 8 or 16) through L2 every K iterations.  See the .cu for the design.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "resident_plan.h"

namespace b2c {

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
    int cx;               // cluster width: CTAs of one tile row that trade halo columns through DSMEM (1 = none)
};

cudaError_t launch_resident(const ResidentPlan &plan, const ResidentParams &p, int ch, cudaStream_t s);
int resident_cluster_width(const ResidentPlan &plan, int ch);

}  // namespace b2c
