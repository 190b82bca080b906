// afe_fft.h -- register-resident small FFTs used by the 512-point real transform.
//
// All loops are fully unrolled with compile-time indices, so the arrays live in registers.
// Forward transform, e^{-2*pi*i*nk/N}.  T is float (acoustic path; the reference itself is
// float32 there) or double (fbank path; the reference runs float64 and the normalised
// features are sensitive to ~1e-9 relative perturbations in empty bands -- DESIGN.md).
#pragma once
#include "afe_platform.h"

namespace afe {

// cos/sin(2*pi*m/16), m = 0..15
template <typename T> struct W16 {
  AFE_HD static T c(int m) {
    const T C1 = T(0.92387953251128675613), C2 = T(0.70710678118654752440), C3 = T(0.38268343236508977173);
    switch (m & 15) {
      case 0: return T(1);   case 1: return C1;   case 2: return C2;   case 3: return C3;
      case 4: return T(0);   case 5: return -C3;  case 6: return -C2;  case 7: return -C1;
      case 8: return T(-1);  case 9: return -C1;  case 10: return -C2; case 11: return -C3;
      case 12: return T(0);  case 13: return C3;  case 14: return C2;  default: return C1;
    }
  }
  AFE_HD static T s(int m) { return c((m + 12) & 15); }   // sin(x) = cos(x - pi/2)
};

// y = x * e^{-2*pi*i*m/16}  (compile-time m after unrolling)
template <typename T> AFE_HD void mul_w16(T& re, T& im, int m) {
  m &= 15;
  if (m == 0) return;
  if (m == 4) { T t = re; re = im; im = -t; return; }          // * (-i)
  if (m == 8) { re = -re; im = -im; return; }
  if (m == 12) { T t = re; re = -im; im = t; return; }         // * (+i)
  const T c = W16<T>::c(m), s = W16<T>::s(m);                   // w = c - i s
  const T r = re * c + im * s;
  const T i = im * c - re * s;
  re = r; im = i;
}

// 4-point DFT on (a0,a1,a2,a3), in place
template <typename T>
AFE_HD void dft4(T& r0, T& i0, T& r1, T& i1, T& r2, T& i2, T& r3, T& i3) {
  const T t0r = r0 + r2, t0i = i0 + i2, t1r = r0 - r2, t1i = i0 - i2;
  const T t2r = r1 + r3, t2i = i1 + i3, t3r = r1 - r3, t3i = i1 - i3;
  r0 = t0r + t2r; i0 = t0i + t2i;
  r2 = t0r - t2r; i2 = t0i - t2i;
  r1 = t1r + t3i; i1 = t1i - t3r;      // t1 - i*t3
  r3 = t1r - t3i; i3 = t1i + t3r;      // t1 + i*t3
}

// dft4 with a3 = 0 (pruned: the frame has 400 samples, the transform 512 points)
template <typename T>
AFE_HD void dft4_a3zero(T& r0, T& i0, T& r1, T& i1, T& r2, T& i2, T& r3, T& i3) {
  const T t0r = r0 + r2, t0i = i0 + i2, t1r = r0 - r2, t1i = i0 - i2;
  r0 = t0r + r1; i0 = t0i + i1;
  r2 = t0r - r1; i2 = t0i - i1;
  r3 = t1r - i1; i3 = t1i + r1;        // t1 + i*a1   (before r1/i1 are overwritten)
  const T nr1 = t1r + i1, ni1 = t1i - r1;      // t1 - i*a1
  r1 = nr1; i1 = ni1;
}
// dft4 with Im(a3) = 0
template <typename T>
AFE_HD void dft4_i3zero(T& r0, T& i0, T& r1, T& i1, T& r2, T& i2, T& r3, T& i3) {
  const T t0r = r0 + r2, t0i = i0 + i2, t1r = r0 - r2, t1i = i0 - i2;
  const T t2r = r1 + r3, t3r = r1 - r3;
  r0 = t0r + t2r; i0 = t0i + i1;
  r2 = t0r - t2r; i2 = t0i - i1;
  r3 = t1r - i1; i3 = t1i + t3r;
  const T nr1 = t1r + i1, ni1 = t1i - t3r;
  r1 = nr1; i1 = ni1;
}

// 16-point complex FFT, natural order in, natural order out.
//   n = 4*n1 + n2, k = k1 + 4*k2:  X[k1+4k2] = sum_n2 W16^(n2 k1) W4^(n2 k2) sum_n1 x[4n1+n2] W4^(n1 k1)
//   kTailZero: inputs 13, 14, 15 and Im(input 12) are zero (stage 1 of the 400-sample frame);
//   the first pass then skips their additions (x + 0.0 is not foldable under IEEE rules).
template <typename T, bool kTailZero = false> AFE_HD void fft16(T (&re)[16], T (&im)[16]) {
  if (kTailZero) {
    dft4_i3zero(re[0], im[0], re[4], im[4], re[8], im[8], re[12], im[12]);
#pragma unroll
    for (int n2 = 1; n2 < 4; ++n2)
      dft4_a3zero(re[n2], im[n2], re[4 + n2], im[4 + n2], re[8 + n2], im[8 + n2], re[12 + n2], im[12 + n2]);
  } else {
#pragma unroll
    for (int n2 = 0; n2 < 4; ++n2)     // a[n2][k1] stored at index 4*k1 + n2
      dft4(re[n2], im[n2], re[4 + n2], im[4 + n2], re[8 + n2], im[8 + n2], re[12 + n2], im[12 + n2]);
  }
#pragma unroll
  for (int k1 = 1; k1 < 4; ++k1)
#pragma unroll
    for (int n2 = 1; n2 < 4; ++n2) mul_w16(re[4 * k1 + n2], im[4 * k1 + n2], n2 * k1);
#pragma unroll
  for (int k1 = 0; k1 < 4; ++k1)       // DFT over n2 -> index 4*k1 + k2 holds X[k1 + 4*k2]
    dft4(re[4 * k1], im[4 * k1], re[4 * k1 + 1], im[4 * k1 + 1], re[4 * k1 + 2], im[4 * k1 + 2],
         re[4 * k1 + 3], im[4 * k1 + 3]);
  // un-transpose: slot 4*k1 + k2  ->  X[k1 + 4*k2]
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = a + 1; b < 4; ++b) {
      T t = re[4 * a + b]; re[4 * a + b] = re[4 * b + a]; re[4 * b + a] = t;
      t = im[4 * a + b]; im[4 * a + b] = im[4 * b + a]; im[4 * b + a] = t;
    }
}

// 8-point complex FFT, natural order in/out.
//   n = 2*n1 + n2, k = k1 + 4*k2:  X[k1+4k2] = sum_n2 W8^(n2 k1) (-1)^(n2 k2) sum_n1 x[2n1+n2] W4^(n1 k1)
template <typename T> AFE_HD void fft8(T (&re)[8], T (&im)[8]) {
  dft4(re[0], im[0], re[2], im[2], re[4], im[4], re[6], im[6]);    // even samples -> slots 0,2,4,6 = k1 0..3
  dft4(re[1], im[1], re[3], im[3], re[5], im[5], re[7], im[7]);    // odd samples  -> slots 1,3,5,7
#pragma unroll
  for (int k1 = 1; k1 < 4; ++k1) mul_w16(re[2 * k1 + 1], im[2 * k1 + 1], 2 * k1);   // W8^k1 = W16^(2 k1)
  T xr[8], xi[8];
#pragma unroll
  for (int k1 = 0; k1 < 4; ++k1) {
    xr[k1] = re[2 * k1] + re[2 * k1 + 1];     xi[k1] = im[2 * k1] + im[2 * k1 + 1];
    xr[k1 + 4] = re[2 * k1] - re[2 * k1 + 1]; xi[k1 + 4] = im[2 * k1] - im[2 * k1 + 1];
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) { re[k] = xr[k]; im[k] = xi[k]; }
}

}  // namespace afe
