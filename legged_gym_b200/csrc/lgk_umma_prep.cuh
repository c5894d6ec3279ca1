// Weight images of the tcgen05 actuator MLP (lgk_mlp_umma.cu): layout constants and the per-element work of
// building them from the fp32 blob.  Shared with lgk_control.cu, whose history kernel carries the image
// build as 64 extra CTAs (no launch of its own, no time of its own).
#pragma once
#include "lgk_common.cuh"

#ifndef ACT_LEGS
#define ACT_LEGS 4
#define ACT_LEG_DOF 3
#endif
#ifndef ACT_DOF
#define ACT_DOF 12
#endif
#ifndef ACT_IN
#define ACT_IN 30
#endif
#ifndef ACT_HID
#define ACT_HID 128
#endif

namespace um {
constexpr int K0 = 32;                                  // layer-0 reduction length, 30 padded to 4 k-steps
constexpr uint32_t IMG0 = ACT_HID * K0 * 4;             // 16 KB
constexpr uint32_t IMG = ACT_HID * ACT_HID * 4;         // 64 KB
// global image buffer: W0hi W0lo W1hi W1lo W2hi W2lo
constexpr uint32_t G_W0HI = 0, G_W0LO = IMG0, G_W1HI = 2 * IMG0, G_W1LO = G_W1HI + IMG, G_W2HI = G_W1LO + IMG,
                   G_W2LO = G_W2HI + IMG, G_TOTAL = G_W2LO + IMG;
constexpr int CBLOCK_FLOATS = 3 * ACT_HID + ACT_LEG_DOF * ACT_HID + 4;
constexpr int PREP_ELEMENTS = ACT_HID * ACT_HID;        // one thread per element of a 128 x 128 layer
}  // namespace um

__device__ __forceinline__ void um_split(float v, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(v) + 0x1000u) & 0xffffe000u;
  lo = (__float_as_uint(v - __uint_as_float(hi)) + 0x1000u) & 0xffffe000u;
}

// element (n, k) of an [128][K] K-major operand inside its no-swizzle UMMA image, in floats
__device__ __forceinline__ uint32_t um_img_index(int n, int k, int K) {
  return (uint32_t)(((n >> 3) * (K >> 2) + (k >> 2)) * 32 + (n & 7) * 4 + (k & 3));
}

// thread i of um::PREP_ELEMENTS: its elements of the six images
__device__ __forceinline__ void lgk_umma_prep_element(int i, const float* __restrict__ weights, float* __restrict__ img) {
  using namespace um;
  const float* W0 = weights;
  const float* W1 = W0 + ACT_HID * ACT_IN + ACT_HID;
  const float* W2 = W1 + ACT_HID * ACT_HID + ACT_HID;
  uint32_t* out = reinterpret_cast<uint32_t*>(img);
  if (i < ACT_HID * K0) {
    const int n = i / K0, k = i % K0;
    uint32_t hi, lo;
    um_split(k < ACT_IN ? W0[n * ACT_IN + k] : 0.f, hi, lo);
    const uint32_t at = um_img_index(n, k, K0);
    out[G_W0HI / 4 + at] = hi;
    out[G_W0LO / 4 + at] = lo;
  }
  if (i < ACT_HID * ACT_HID) {
    const int n = i >> 7, k = i & 127;
    const uint32_t at = um_img_index(n, k, ACT_HID);
    uint32_t hi, lo;
    um_split(W1[i], hi, lo);
    out[G_W1HI / 4 + at] = hi;
    out[G_W1LO / 4 + at] = lo;
    um_split(W2[i], hi, lo);
    out[G_W2HI / 4 + at] = hi;
    out[G_W2LO / 4 + at] = lo;
  }
}

