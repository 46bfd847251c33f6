// TMA-fed variant of the streaming stencil (stencil_tma.cu): the same register
// pipeline, but input rows arrive in shared memory through
// cp.async.bulk.tensor.2d tiles signalled by mbarriers instead of per-lane LDGs.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace b2c {

struct PassParams;
struct TmaState;

// `alloc` are the two framed ping-pong allocations (row 0 of the allocation,
// not the interior origin); one uint8 tensor map of pitch x rows is encoded per buffer.
TmaState *tma_create(uint8_t *const alloc[2], long long pitch, int alloc_rows, long long alloc_bytes);
void tma_destroy(TmaState *t);
cudaError_t launch_stream_tma(TmaState *t, const PassParams &p, int src_index, int k, int strip_bytes, int ch,
                              bool check, cudaStream_t s);

int stream_tma_occupancy(int k, int strip_bytes, int ch);
cudaError_t preload_stream_tma_kernel(int k, int strip_bytes, int ch, bool check);   // load the variant now (lazy loading)  // resident CTAs per SM with its shared-memory ring

}  // namespace b2c
