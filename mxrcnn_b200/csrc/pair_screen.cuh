// The float32 screen of the IoU test (nms.py:25-36) shared by the fused proposal kernel (proposal_fused.cuh) and the
// fused NMS kernel (nms_fused.cuh): staged boxes, one pair's decision value, 32 x 32 bit blocks.
#pragma once
#include "device_math.cuh"

namespace mxr {
namespace fz {

#ifndef FZ_PAIR_GROUP
#define FZ_PAIR_GROUP 8  // pairs per unrolled group of the IoU block loop (compile-time bit positions)
#endif
static const int kWMax = 512;               // widest stripe
static const int kMWords = kWMax / 32;      // u32 words per bit row; rows are stored word-major: M[word][row], so that a
                                            // warp's 32 rows of one word are 128 contiguous bytes (one remote transaction)

// One staged candidate / kept box as the float32 screen sees it.
struct Box32 {
  float x1, y1, x2p, y2p;  // x2p = x2 + 1, y2p = y2 + 1
  float a, t;              // thr / (1 + thr) * area, tolerance
};

// float32 screen of nms.py:25-36 for one row box against up to 32 staged columns.
// inter > thr * union  <=>  inter > c (area_i + area_j) with c = thr / (1 + thr); the areas arrive scaled by c.
// supp: bit j set when column j is suppressed for certain; und: bit j set when the pair is inside the error bound
// (also when d is NaN or a tolerance infinite) and must be decided by the float64 test.
__device__ __forceinline__ float pair_d(const Box32& r, const float4 cb, const float2 ca, float* tol) {
  const float xx1 = fmaxf(r.x1, cb.x), yy1 = fmaxf(r.y1, cb.y);
  const float xx2 = fminf(r.x2p, cb.z), yy2 = fminf(r.y2p, cb.w);
  // Only ONE extent is clamped at zero: with w >= 0 a negative h makes the product <= 0, i.e. "no overlap" decided as it
  // should be, and an h within the error bound of zero changes the product by at most w * delta, which the tolerance
  // covers (FMNMX runs at half rate: five of them per pair instead of six).
  const float w = fmaxf(xx2 - xx1, 0.0f), h = yy2 - yy1;
  *tol = r.t + ca.y;
  return __fmaf_rn(w, h, -(r.a + ca.x));  // a screen, not the reference's arithmetic: contraction is welcome
}
__device__ __forceinline__ void block_bits(const Box32& r, const float4* __restrict__ cb, const float2* __restrict__ ca, int jn,
                                           u32& supp_out, u32& und_out) {
  // groups of FZ_PAIR_GROUP columns with compile-time bit positions; columns at and beyond jn (garbage inside the staging
  // arrays, which are kWMax long) are masked off by the caller
  u32 supp = 0;
  bool any_und = false;
#pragma unroll 1
  for (int j0 = 0; j0 < jn; j0 += FZ_PAIR_GROUP) {
    u32 s8 = 0;
#pragma unroll
    for (int j = 0; j < FZ_PAIR_GROUP; ++j) {
      float tol;
      const float d = pair_d(r, cb[j0 + j], ca[j0 + j], &tol);
      s8 |= d > tol ? (1u << j) : 0u;
      any_und = any_und || !(fabsf(d) > tol);
    }
    supp |= s8 << j0;
  }
  u32 und = 0;
  if (any_und) {  // rare: find which
    for (int j = 0; j < jn; ++j) {
      float tol;
      const float d = pair_d(r, cb[j], ca[j], &tol);
      und |= (fabsf(d) > tol ? 0u : 1u) << j;
    }
  }
  supp_out = supp;
  und_out = und;
}

}  // namespace fz
}  // namespace mxr
