// Internal (non-ABI) hooks between translation units of libdqnb200.so.
#pragma once
#include "common.cuh"

namespace dqnb {
// replay.cu: flush staged puts, upload host slots, gather into the given device buffers.
int replay_gather_into(dqnb_replay* h, const int64_t* slots_host, int64_t n, uint8_t* s_tm1, uint8_t* s_t,
                       int32_t* a, float* r, float* d, cudaStream_t st);
int replay_frame_shape(const dqnb_replay* h, int* height, int* width, int* stack);
}  // namespace dqnb
