// Tile geometry of the register-resident kernel (csrc/cuda/stencil_resident.cu), free of CUDA so
// that the planner and the threaded host model of the CTA protocol (resident_emul.cpp) build and
// run on a CPU-only machine.
#pragma once

namespace b2c {

constexpr int kResWarps = 16;               // warps per CTA (one CTA per SM)
constexpr int kResNW = 2;                   // 32-bit words per lane per row
constexpr int kResSB = kResNW * 4;          // bytes per lane per row
constexpr int kResWinB = 32 * kResSB;       // window width in bytes

// K = iterations between two inter-CTA exchanges = halo depth in rows; the halo is K*CH bytes wide,
// rounded up to whole lanes.
constexpr int res_halo_bytes(int ch, int k) { return (k * ch + kResSB - 1) / kResSB * kResSB; }

struct ResidentPlan {
    int rw;        // rows per warp (template instance)
    int k;         // iterations between two exchanges = halo depth (4, 8 or 16)
    int strips;    // CTA columns
    int ctarows;   // CTA rows
    int vr, vc;    // valid rows / bytes per CTA
};

// Smallest tile that covers a bw x H byte image with at most `sms` CTAs; false when the image is
// too large (or too awkward: bw must be a multiple of 8) for the resident kernel.
bool resident_plan(int bw, int H, int ch, int k, int sms, ResidentPlan *out);

}  // namespace b2c
