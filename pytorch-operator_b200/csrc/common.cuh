// Shared definitions for the sm_100a kernels of libb200mnist.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200mnist {

// Flat parameter bucket: element offsets of the eight tensors (256-byte aligned).
constexpr int kNumTensors = 8;
constexpr int kOffC1W = 0, kOffC1B = 512, kOffC2W = 576, kOffC2B = 25600, kOffF1W = 25664, kOffF1B = 425664,
              kOffF2W = 426176, kOffF2B = 431232, kFlatElems = 431296;
constexpr int kTensorOff[kNumTensors] = {kOffC1W, kOffC1B, kOffC2W, kOffC2B, kOffF1W, kOffF1B, kOffF2W, kOffF2B};
constexpr int kTensorElems[kNumTensors] = {500, 20, 25000, 50, 400000, 500, 5000, 10};

// Per-image activation sizes (floats).
constexpr int kXElems = 784, kP1Elems = 2880, kP2Elems = 800, kHElems = 500, kClasses = 10;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One gradient tensor whose value is the sum of `nsplit` partial buffers
// (part + s*stride, s < nsplit), landing at flat offset `off` .. off+n.
struct GradSegment {
    int off, n, nsplit, stride;
    const float *part;
};
struct GradSegments {
    GradSegment seg[6];
    int count;
};

}  // namespace b200mnist
