// TMA-fed streaming stencil — placeholder until the bulk-tensor pipeline lands.
#include "stencil_tma.h"

#include "common.h"

namespace b2c {

struct TmaState {
    int unused;
};

TmaState *tma_create(uint8_t *const[2], long long, int, long long) {
    set_error("kernel=tma is not built yet");
    return nullptr;
}
void tma_destroy(TmaState *t) { delete t; }
cudaError_t launch_stream_tma(TmaState *, const PassParams &, int, int, int, int, bool, cudaStream_t) {
    return cudaErrorNotSupported;
}

}  // namespace b2c
