// cvprim.h — OpenCV-free restatement of the image primitives the reference calls.
//
// TEST INFRASTRUCTURE (oracle + timed CPU baseline).  Never linked into the product library.
// Each function restates the arithmetic of one un-vendored OpenCV primitive (OpenCV 4.x
// semantics, pinned to cv2 4.13.0 by tests/test_oracle_vs_cv2.py) at the reference's call site:
//   resize(INTER_LINEAR)            ORBextractor.cc:1165
//   copyMakeBorder(REFLECT_101)     ORBextractor.cc:1167,1172
//   FAST(9_16, nonmax=true)         ORBextractor.cc:808,827
//   GaussianBlur(7x7, sigma 2)      ORBextractor.cc:1115
//   fastAtan2                       ORBextractor.cc:101
//   BFMatcher(NORM_HAMMING).knnMatch Frame.cc:1134
//   norm(NORM_L1)                   Frame.cc:908
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace orbref {

inline int cvRoundF(float v) { return (int)__builtin_lrintf(v); }   // half-to-even (default FE mode)
inline int cvRoundD(double v) { return (int)__builtin_lrint(v); }

inline int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
    return i;
}

struct FastKp { int x, y, score; };

// dst (dw x dh, stride dstep) = bilinear 8-bit resize of src (sw x sh, stride sstep).
void resizeLinearU8(const uint8_t* src, int sw, int sh, ptrdiff_t sstep, uint8_t* dst, int dw, int dh, ptrdiff_t dstep);
// buf is a (w+2b) x (h+2b) buffer whose interior (at offset b,b) is already filled; fill the border.
void fillBorderReflect101(uint8_t* buf, int w, int h, int b, ptrdiff_t step);
// FAST-9/16 with 3x3 non-max suppression on a w x h ROI; output in row-major order.
void fast9_16(const uint8_t* img, int w, int h, ptrdiff_t step, int threshold, std::vector<FastKp>& out);
// 7x7 sigma=2 fixed-point Gaussian with reflect-101 border; src and dst may not alias.
void gaussianBlur7x7(const uint8_t* src, int w, int h, ptrdiff_t sstep, uint8_t* dst, ptrdiff_t dstep);
float fastAtan2(float y, float x);
int hammingDistance256(const uint8_t* a, const uint8_t* b);
// 2-NN brute force, ties to the lowest train index.  idx/dist are nq x 2 (-1 where absent).
void knn2Hamming(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist);

}  // namespace orbref
