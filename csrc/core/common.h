// Internal helpers shared by the CPU and CUDA halves of libb200conv_core.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "b200conv.h"

namespace b2c {

// Thread-local last-error text behind b2c_last_error().
void set_error(const char *fmt, ...);

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }   // FRACTION_CEILING, cuda/functions.h:18
inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }
inline int64_t round_down(int64_t a, int64_t b) { return (a / b) * b; }

// Integer weights and divisor of the three reference filters (mpi_convolution.c:91-93).
struct FilterDef {
    int w[3][3];
    int den;
};
const FilterDef &filter_def(int filter);

// Offsets (drow, dcol) of the 8 directions, B2C_N .. B2C_SE.
extern const int kDirDr[8];
extern const int kDirDc[8];
inline int opposite_dir(int d) {
    static const int opp[8] = {B2C_S, B2C_N, B2C_E, B2C_W, B2C_SE, B2C_SW, B2C_NE, B2C_NW};
    return opp[d];
}

// oracle.cpp: k Gaussian iterations over one rank's framed buffer (used by both host models).
void framed_iterations(std::vector<uint8_t> &t0, std::vector<uint8_t> &t1, int rows, int64_t pitch, int ch, int F,
                       int64_t fb, int r0, int64_t c0b, int H, int64_t Wb, bool torus, int k);

}  // namespace b2c
