// Headerless interleaved-uint8 .raw block I/O (layer L3).
// Spec: per-row seek+read/write of each rank's block, mpi_convolution.c:126-140
// and :244-262; whole-file read_info/write_info, cuda/functions.c:31-45.
// Differences on purpose: 64-bit offsets (the reference uses int and overflows
// at 2^31 bytes), EOF is an error instead of an endless loop, EINTR is retried.
#include <cerrno>
#include <fcntl.h>
#include <unistd.h>

#include "common.h"

namespace {

int64_t full_pread(int fd, uint8_t *dst, int64_t n, int64_t off) {
    int64_t done = 0;
    while (done < n) {
        ssize_t k = pread(fd, dst + done, (size_t)(n - done), (off_t)(off + done));
        if (k < 0) {
            if (errno == EINTR) continue;
            return -1;
        }
        if (k == 0) return done;  // EOF
        done += k;
    }
    return done;
}

int64_t full_pwrite(int fd, const uint8_t *src, int64_t n, int64_t off) {
    int64_t done = 0;
    while (done < n) {
        ssize_t k = pwrite(fd, src + done, (size_t)(n - done), (off_t)(off + done));
        if (k < 0) {
            if (errno == EINTR) continue;
            return -1;
        }
        done += k;
    }
    return done;
}

bool block_ok(int width, int height, int ch, int row0, int col0, int rows, int cols) {
    return width > 0 && height > 0 && (ch == 1 || ch == 3) && row0 >= 0 && col0 >= 0 && rows >= 0 &&
           cols >= 0 && row0 + rows <= height && col0 + cols <= width;
}

}  // namespace

extern "C" {

int64_t b2c_read_block(const char *path, uint8_t *buf, int64_t pitch, int width, int height, int ch,
                       int row0, int col0, int rows, int cols) {
    if (!path || !buf || !block_ok(width, height, ch, row0, col0, rows, cols)) {
        b2c::set_error("read_block: bad arguments");
        return B2C_EINVAL;
    }
    int fd = open(path, O_RDONLY);
    if (fd < 0) {
        b2c::set_error("cannot open %s", path);
        return B2C_EIO;
    }
    const int64_t row_bytes = (int64_t)cols * ch, file_pitch = (int64_t)width * ch;
    int64_t total = 0;
    if (cols == width && pitch == row_bytes) {  // whole rows, tight buffer: one request
        int64_t want = row_bytes * rows;
        int64_t got = full_pread(fd, buf, want, (int64_t)row0 * file_pitch);
        if (got != want) {
            close(fd);
            b2c::set_error("%s: short read (%lld of %lld bytes)", path, (long long)got, (long long)want);
            return B2C_EIO;
        }
        total = got;
    } else {
        for (int r = 0; r < rows; ++r) {
            int64_t off = (int64_t)(row0 + r) * file_pitch + (int64_t)col0 * ch;
            int64_t got = full_pread(fd, buf + (int64_t)r * pitch, row_bytes, off);
            if (got != row_bytes) {
                close(fd);
                b2c::set_error("%s: short read at row %d", path, row0 + r);
                return B2C_EIO;
            }
            total += got;
        }
    }
    close(fd);
    return total;
}

int64_t b2c_write_block(const char *path, const uint8_t *buf, int64_t pitch, int width, int height,
                        int ch, int row0, int col0, int rows, int cols, int create) {
    if (!path || !buf || !block_ok(width, height, ch, row0, col0, rows, cols)) {
        b2c::set_error("write_block: bad arguments");
        return B2C_EINVAL;
    }
    // O_CREAT|O_WRONLY without O_TRUNC, mode 0644: cuda/main.c:39 and
    // MPI_MODE_CREATE|MPI_MODE_WRONLY (mpi_convolution.c:248) behave this way.
    int fd = open(path, (create ? O_CREAT : 0) | O_WRONLY, 0644);
    if (fd < 0) {
        b2c::set_error("cannot open-create %s", path);
        return B2C_EIO;
    }
    const int64_t row_bytes = (int64_t)cols * ch, file_pitch = (int64_t)width * ch;
    int64_t total = 0;
    if (cols == width && pitch == row_bytes) {
        total = full_pwrite(fd, buf, row_bytes * rows, (int64_t)row0 * file_pitch);
        if (total != row_bytes * rows) total = -1;
    } else {
        for (int r = 0; r < rows && total >= 0; ++r) {
            int64_t off = (int64_t)(row0 + r) * file_pitch + (int64_t)col0 * ch;
            int64_t k = full_pwrite(fd, buf + (int64_t)r * pitch, row_bytes, off);
            total = (k == row_bytes) ? total + k : -1;
        }
    }
    close(fd);
    if (total < 0) {
        b2c::set_error("%s: write failed", path);
        return B2C_EIO;
    }
    return total;
}

}  // extern "C"
