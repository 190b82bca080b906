// afe_rfft512.h -- 512-point real FFT power spectrum of a 400-sample frame, split over
// 16 "roles" (warps) x 32 frames (lanes).  Thread (role, f) owns frame f of the tile.
//
//   N = 512 = 32 x 16,  n = 16*n1 + n2,  k = k1 + 32*k2
//   stage 1 (role = n2):  R_n2[k1] = sum_{n1<25} y[16 n1 + n2] W32^(n1 k1),  k1 = 0..16
//                         (real 32-point DFT through one complex 16-point FFT + untangling),
//                         times the twiddle W512^(n2 k1); written to the exchange buffer.
//   stage 2 (role = k1):  X[k1 + 32 k2] = sum_n2 (...) W16^(n2 k2): one complex 16-point FFT.
//                         Rows k1 = 1..15 also cover bins 512-k by conjugate symmetry, so the
//                         257 powers come out with no partner exchange.  Role 0 owns the two
//                         real rows k1 = 0 and k1 = 16 (two 8-point FFTs).
//
// The caller pre-scales the frame by 1/2 (folded into the window table) so that the
// untangling needs no halving.  Exchange layout: cplx<T> ex[(slot*16 + n2)*TF + f], slot 0
// holds the real pair (R[0], R[16]), slot k1 holds the twiddled R[k1].
// Power layout: float sP[bin*TF + f].  TF = frames per tile: 32 (one role per warp) or 16 (two
// roles per warp, lanes 0-15 / 16-31), which halves the exchange buffer.
#pragma once
#include "afe_fft.h"

namespace afe {

template <typename T> struct W32 {   // e^{-2 pi i k/32} = c(k) - i s(k), k = 0..8
  AFE_HD static T c(int k) {
    switch (k) {
      case 0: return T(1);
      case 1: return T(0.98078528040323044913);
      case 2: return T(0.92387953251128675613);
      case 3: return T(0.83146961230254523708);
      case 4: return T(0.70710678118654752440);
      case 5: return T(0.55557023301960222474);
      case 6: return T(0.38268343236508977173);
      case 7: return T(0.19509032201612826785);
      default: return T(0);
    }
  }
  AFE_HD static T s(int k) { return c(8 - k); }
};

constexpr int kTileFrames = 32;
constexpr int kRoles = 16;
constexpr int kWinRow = 26;          // window tables are role-major: w[role][n1] = window[16 n1 + role], 25 used

// the 25 windowed samples of (role, frame): global table entries fetched two per load
template <typename T, class Sample>
AFE_HD void windowed_column(const T* AFE_RESTRICT wrow, T (&r)[25], Sample sample) {
#pragma unroll
  for (int n1 = 0; n1 < 24; n1 += 2) {
    T w0, w1;
    ldg2(wrow + n1, w0, w1);
    r[n1] = w0 * sample(n1); r[n1 + 1] = w1 * sample(n1 + 1);
  }
  r[24] = ldg(wrow + 24) * sample(24);
}
// same with the window row read through win(n1) (a table that rides in the kernel parameters)
template <typename T, class Win, class Sample>
AFE_HD void windowed_column_fn(Win win, T (&r)[25], Sample sample) {
#pragma unroll
  for (int n1 = 0; n1 < 25; ++n1) r[n1] = win(n1) * sample(n1);
}

// twiddle accessors for rfft512_stage1: tw(k1, c, s) yields W512^(n2 k1) = c + i s of the caller's n2
template <typename T> struct TwGlobal {           // [16][16][2] table in global memory
  const T* AFE_RESTRICT row;                       // tw512 + n2 * 32
  AFE_HD void operator()(int k1, T& c, T& s) const { ldg2(row + k1 * 2, c, s); }
};

template <typename T, int TF = kTileFrames, class Tw>
AFE_HD void rfft512_stage1_tw(const T (&r)[25], int n2, int f, Tw tw, cplx<T>* AFE_RESTRICT ex) {
  T re[16], im[16];
#pragma unroll
  for (int m = 0; m < 12; ++m) { re[m] = r[2 * m]; im[m] = r[2 * m + 1]; }
  re[12] = r[24]; im[12] = T(0);
#pragma unroll
  for (int m = 13; m < 16; ++m) { re[m] = T(0); im[m] = T(0); }
  fft16<T, true>(re, im);
  // untangle the packed real transform (inputs were pre-scaled by 1/2)
  const T r0 = T(2) * (re[0] + im[0]);
  const T r16 = T(2) * (re[0] - im[0]);
#pragma unroll
  for (int k = 1; k < 8; ++k) {
    const int m = 16 - k;
    const T er = re[k] + re[m], ei = im[k] - im[m];
    const T dr = re[k] - re[m], di = im[k] + im[m];
    const T orr = di, oi = -dr;                              // O = -i * D
    const T c = W32<T>::c(k), s = W32<T>::s(k);
    const T tr = orr * c + oi * s, ti = oi * c - orr * s;    // W32^k * O
    re[k] = er + tr; im[k] = ei + ti;                        // R[k]
    re[m] = er - tr; im[m] = -(ei - ti);                     // R[16-k] = conj(E - T)
  }
  re[8] = T(2) * re[8]; im[8] = T(-2) * im[8];               // R[8] = 2 conj(C[8])
  ex[(0 * 16 + n2) * TF + f] = cplx<T>{r0, r16};
#pragma unroll
  for (int k1 = 1; k1 < 16; ++k1) {
    T c, s;                                                                                   // c + i s
    tw(k1, c, s);
    ex[(k1 * 16 + n2) * TF + f] = cplx<T>{re[k1] * c - im[k1] * s, re[k1] * s + im[k1] * c};
  }
}

template <typename T, int TF = kTileFrames>
AFE_HD void rfft512_stage1(const T (&r)[25], int n2, int f, const T* AFE_RESTRICT tw512,
                           cplx<T>* AFE_RESTRICT ex) {
  rfft512_stage1_tw<T, TF>(r, n2, f, TwGlobal<T>{tw512 + n2 * 32}, ex);
}

AFE_HD float rfft_mag(float p) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(p));
  return r;
#else
  return std::sqrt(p);
#endif
}

// kMag: also store the magnitude sqrt(power) to sM (same layout as sP)
// v[n2] = the exchange values ex[(role * 16 + n2) * TF + f] of (role, frame), already in registers
template <typename T, bool kMag = false, int TF = kTileFrames>
AFE_HD void rfft512_stage2_regs(int role, int f, const cplx<T> (&v)[16], float* AFE_RESTRICT sP,
                                float* AFE_RESTRICT sM = nullptr) {
#define AFE_PUT(bin, val) { const float p_ = (float)(val); sP[(bin) * TF + f] = p_; if (kMag) sM[(bin) * TF + f] = rfft_mag(p_); }
  if (role != 0) {
    T re[16], im[16];
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) { re[n2] = v[n2].re; im[n2] = v[n2].im; }
    fft16(re, im);
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) {
      const int bin = (k2 < 8) ? (role + 32 * k2) : ((32 - role) + 32 * (15 - k2));
      AFE_PUT(bin, re[k2] * re[k2] + im[k2] * im[k2])
    }
    return;
  }
  // role 0: row k1 = 0 (real input a0) and row k1 = 16 (real input a16, half-bin shifted)
  T a0[16], a16[16];
#pragma unroll
  for (int n2 = 0; n2 < 16; ++n2) { a0[n2] = v[n2].re; a16[n2] = v[n2].im; }
  {
    T re[8], im[8];
#pragma unroll
    for (int m = 0; m < 8; ++m) { re[m] = a0[2 * m]; im[m] = a0[2 * m + 1]; }
    fft8(re, im);
    const T x0 = re[0] + im[0], x8 = re[0] - im[0];
    AFE_PUT((32 * 0), x0 * x0)
    AFE_PUT((32 * 8), x8 * x8)
    AFE_PUT((32 * 4), re[4] * re[4] + im[4] * im[4])
#pragma unroll
    for (int k = 1; k < 4; ++k) {
      const int m = 8 - k;
      const T er = T(0.5) * (re[k] + re[m]), ei = T(0.5) * (im[k] - im[m]);
      const T dr = T(0.5) * (re[k] - re[m]), di = T(0.5) * (im[k] + im[m]);
      const T orr = di, oi = -dr;
      const T c = W16<T>::c(k), s = W16<T>::s(k);
      const T tr = orr * c + oi * s, ti = oi * c - orr * s;
      const T ar = er + tr, ai = ei + ti, br = er - tr, bi = ei - ti;
      AFE_PUT((32 * k), ar * ar + ai * ai)
      AFE_PUT((32 * m), br * br + bi * bi)
    }
  }
  {
    T re[8], im[8];
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      const T c = W32<T>::c(n), s = W32<T>::s(n);          // (a - i b)(c - i s)
      re[n] = a16[n] * c - a16[n + 8] * s;
      im[n] = -a16[n] * s - a16[n + 8] * c;
    }
    fft8(re, im);                                           // V[m] = Y[2m]
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int m = (k & 1) ? (15 - k) / 2 : k / 2;
      AFE_PUT((16 + 32 * k), re[m] * re[m] + im[m] * im[m])
    }
  }
}
#undef AFE_PUT

template <typename T, bool kMag = false, int TF = kTileFrames>
AFE_HD void rfft512_stage2(int role, int f, const cplx<T>* AFE_RESTRICT ex, float* AFE_RESTRICT sP,
                           float* AFE_RESTRICT sM = nullptr) {
  cplx<T> v[16];
#pragma unroll
  for (int n2 = 0; n2 < 16; ++n2) v[n2] = ex[(role * 16 + n2) * TF + f];
  rfft512_stage2_regs<T, kMag, TF>(role, f, v, sP, sM);
}

}  // namespace afe
