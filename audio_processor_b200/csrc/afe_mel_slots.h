// afe_mel_slots.h -- triangular mel banks in SLOT form: one straight-line routine shared by all 16 roles.
//
// Why: a phase of these block programs runs once per tile, so its code is cold every time.  The generated
// role-specialised banks (one generated code path per role, 28-33 KB of SASS) made the loop body of the tile
// kernels 50-60 KB -- more than the SM's 32 KB instruction cache -- and ncu showed 31-48 % of the mel phases'
// samples as no-instruction stalls.  Here filter (role r, slot s) is m = 16 s + (s even ? r : 15 - r) (a zigzag:
// filter widths grow with m, so every role gets the same work), every slot is padded to the widest filter of
// its 16 (zero weights), and the weights / first bins of a role are one row of a small shared-memory table.
// All warps run the same ~2.5 KB of code: loads with immediate offsets, FMAs in two chains per filter, no
// branches.  Tables: afe_tables.h make_mel_slots (checked there against the compile-time slot lengths).
#pragma once
#include "afe_platform.h"
#include "afe_tables.h"

namespace afe {

template <int TF, int L, int OFF>
AFE_HD float mel_slot(const float (&w)[kMelSlotRow], const float* AFE_RESTRICT col, int start) {
  const float* AFE_RESTRICT p = col + start * TF;
  float a0 = 0.f, a1 = 0.f;
#pragma unroll
  for (int j = 0; j + 1 < L; j += 2) {
    a0 = fma_f(w[OFF + j], p[j * TF], a0);
    a1 = fma_f(w[OFF + j + 1], p[(j + 1) * TF], a1);
  }
  if (L & 1) a0 = fma_f(w[OFF + L - 1], p[(L - 1) * TF], a0);
  return a0 + a1;
}

AFE_HD void mel_slot_row(const float* AFE_RESTRICT wrow, float (&w)[kMelSlotRow]) {
  const float4* AFE_RESTRICT w4 = reinterpret_cast<const float4*>(wrow);
#pragma unroll
  for (int i = 0; i < kMelSlotRow / 4; ++i) {
    const float4 v = w4[i];
    w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
  }
}

// Kaldi bank, 80 filters: slots of 4, 4, 6, 10, 16 bins (kKaldiSlotLen).  sP[bin * TF + f]; out(m, energy)
template <int TF, class Out>
AFE_HD void mel_slots_kaldi(int role, int f, const float* AFE_RESTRICT sP, const float* AFE_RESTRICT wtab,
                            const int32_t* AFE_RESTRICT stab, Out out) {
  float w[kMelSlotRow];
  mel_slot_row(wtab + role * kMelSlotRow, w);
  const int32_t* st = stab + role * kMelSlotStarts;
  const float* col = sP + f;
  const int up = role, dn = 15 - role;
  // all sums first, outputs (log + store) afterwards: the stores would otherwise fence the next slot's loads
  const float e0 = mel_slot<TF, 4, 0>(w, col, st[0]);
  const float e1 = mel_slot<TF, 4, 4>(w, col, st[1]);
  const float e2 = mel_slot<TF, 6, 8>(w, col, st[2]);
  const float e3 = mel_slot<TF, 10, 14>(w, col, st[3]);
  const float e4 = mel_slot<TF, 16, 24>(w, col, st[4]);
  out(up, e0); out(16 + dn, e1); out(32 + up, e2); out(48 + dn, e3); out(64 + up, e4);
}

// HTK bank, 128 filters: slots of 2, 2, 2, 3, 4, 6, 8, 11 bins (kHtkSlotLen)
template <int TF, class Out>
AFE_HD void mel_slots_htk(int role, int f, const float* AFE_RESTRICT sP, const float* AFE_RESTRICT wtab,
                          const int32_t* AFE_RESTRICT stab, Out out) {
  float w[kMelSlotRow];
  mel_slot_row(wtab + role * kMelSlotRow, w);
  const int32_t* st = stab + role * kMelSlotStarts;
  const float* col = sP + f;
  const int up = role, dn = 15 - role;
  const float e0 = mel_slot<TF, 2, 0>(w, col, st[0]);
  const float e1 = mel_slot<TF, 2, 2>(w, col, st[1]);
  const float e2 = mel_slot<TF, 2, 4>(w, col, st[2]);
  const float e3 = mel_slot<TF, 3, 6>(w, col, st[3]);
  const float e4 = mel_slot<TF, 4, 9>(w, col, st[4]);
  const float e5 = mel_slot<TF, 6, 13>(w, col, st[5]);
  const float e6 = mel_slot<TF, 8, 19>(w, col, st[6]);
  const float e7 = mel_slot<TF, 11, 27>(w, col, st[7]);
  out(up, e0); out(16 + dn, e1); out(32 + up, e2); out(48 + dn, e3);
  out(64 + up, e4); out(80 + dn, e5); out(96 + up, e6); out(112 + dn, e7);
}

}  // namespace afe
