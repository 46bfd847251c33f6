// Process-grid decomposition and neighbour topology (layer L2 of SURVEY §1).
// Behavioural spec: RowsDivision() mpi/mpi_convolution.c:350-364, the validity
// check :53-58, block origin :72-73,85-86 and the neighbour table :142-149.
#include "common.h"

namespace b2c {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

const int kDirDr[8] = {-1, 1, 0, 0, -1, -1, 1, 1};
const int kDirDc[8] = {0, 0, -1, 1, -1, 1, -1, 1};

const FilterDef &filter_def(int filter) {
    static const FilterDef box = {{{1, 1, 1}, {1, 1, 1}, {1, 1, 1}}, 9};
    static const FilterDef gauss = {{{1, 2, 1}, {2, 4, 2}, {1, 2, 1}}, 16};
    static const FilterDef edge = {{{1, 4, 1}, {4, 8, 4}, {1, 4, 1}}, 28};
    switch (filter) {
        case B2C_BOX: return box;
        case B2C_EDGE: return edge;
        default: return gauss;
    }
}

}  // namespace b2c

extern "C" {

const char *b2c_last_error(void) { return b2c::g_err; }
const char *b2c_version(void) { return "b200conv 0.1 (sm_100a)"; }

int b2c_rows_division(int n, int rows, int columns) {
    // Candidate row counts are the divisors of n that also split both image
    // sides exactly; cost = block semi-perimeter; strict '<' keeps the first
    // (smallest) row count on ties, so square images prefer cutting columns.
    int best = 0;
    long best_cost = (long)rows + columns + 1;
    for (int p = 1; p <= n; ++p) {
        if (n % p != 0 || rows % p != 0) continue;
        int q = n / p;
        if (columns % q != 0) continue;
        long cost = rows / p + columns / q;
        if (cost < best_cost) {
            best_cost = cost;
            best = p;
        }
    }
    return best;
}

int b2c_decompose(int n, int height, int width, int *grid_rows, int *grid_cols) {
    if (n <= 0 || height <= 0 || width <= 0) {
        b2c::set_error("decompose: non-positive argument");
        return B2C_EINVAL;
    }
    int p = b2c_rows_division(n, height, width);
    if (p <= 0 || height % p || n % p || width % (n / p)) {
        b2c::set_error("Cannot divide to processes");
        return B2C_EDIVIDE;
    }
    if (grid_rows) *grid_rows = p;
    if (grid_cols) *grid_cols = n / p;
    return B2C_OK;
}

int b2c_neighbours(int rank, int P, int Q, int boundary, int out[8]) {
    if (P <= 0 || Q <= 0 || rank < 0 || rank >= P * Q) {
        b2c::set_error("neighbours: bad rank/grid");
        return B2C_EINVAL;
    }
    const int pr = rank / Q, pc = rank % Q;
    for (int d = 0; d < 8; ++d) {
        int r = pr + b2c::kDirDr[d], c = pc + b2c::kDirDc[d];
        if (boundary == B2C_TORUS) {
            r = (r + P) % P;
            c = (c + Q) % Q;
            out[d] = r * Q + c;
        } else {
            out[d] = (r < 0 || r >= P || c < 0 || c >= Q) ? -1 : r * Q + c;
        }
    }
    return B2C_OK;
}

}  // extern "C"
