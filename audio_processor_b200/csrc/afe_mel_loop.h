// afe_mel_loop.h -- triangular mel banks as ONE table-driven loop shared by all roles.
//
// A phase of these block programs runs once per tile, so its code is cold every time: on the B200 a
// warp streaming its own straight-line code (the role-specialised generated banks of afe_mel_gen.h,
// ~110-200 instructions per role) is bound by instruction fetch -- ncu shows 43-59 % of those phases'
// samples as no-instruction stalls.  Here every warp runs the same few instructions over its role's
// op list (afe_tables.h MelIvOps): per bin one table load, one power load, two FMAs, and at the end of
// an interval the pair (U[i], D[i]) goes to shared memory.  Filter m is U[m] + D[m+1], formed by the
// consumer (same operands and accumulation order as the generated code, so the same bits).
#pragma once
#include "afe_platform.h"
#include "afe_tables.h"

namespace afe {

struct MelIvDev {
  const MelIvOp* ops;            // device
  const int32_t* role_op;        // [17]
};

constexpr int kIvStride = 33;    // floats per interval row of U / D: lane = frame, odd stride

AFE_HD MelIvOp ldg_op(const MelIvOp* p) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  const int4 v = __ldg(reinterpret_cast<const int4*>(p));
  MelIvOp o;
  o.wu = __int_as_float(v.x); o.wd = __int_as_float(v.y); o.bin = v.z; o.end = v.w;
  return o;
#else
  return *p;
#endif
}

// interval sums of frame f over role's intervals.  sP[bin * TF + f] -> sU / sD[interval * 33 + f]
template <int TF>
AFE_HD void mel_interval_sums(const MelIvDev& t, int role, int f, const float* AFE_RESTRICT sP,
                              float* AFE_RESTRICT sU, float* AFE_RESTRICT sD) {
  const int i0 = ldg(t.role_op + role), i1 = ldg(t.role_op + role + 1);
  const float* AFE_RESTRICT col = sP + f;
  float u = 0.f, d = 0.f;
#pragma unroll 1
  for (int i = i0; i < i1; i += 4) {
    MelIvOp o[4];
    float p[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = ldg_op(t.ops + i + k);
#pragma unroll
    for (int k = 0; k < 4; ++k) p[k] = col[o[k].bin * TF];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      u = fma_f(o[k].wu, p[k], u);
      d = fma_f(o[k].wd, p[k], d);
      if (o[k].end >= 0) {                     // uniform across the warp
        sU[o[k].end * kIvStride + f] = u; sD[o[k].end * kIvStride + f] = d;
        u = 0.f; d = 0.f;
      }
    }
  }
}

// energy of filter m in frame f
AFE_HD float mel_energy(const float* AFE_RESTRICT sU, const float* AFE_RESTRICT sD, int m, int f) {
  return sU[m * kIvStride + f] + sD[(m + 1) * kIvStride + f];
}

}  // namespace afe
