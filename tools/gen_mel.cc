// gen_mel.cc -- emits audio_processor_b200/csrc/afe_mel_gen.h: the triangular mel banks as
// straight-line, role-specialised code with the weights as float immediates.
//   g++ -O1 -std=c++17 -Iaudio_processor_b200/csrc tools/gen_mel.cc audio_processor_b200/csrc/afe_tables.cc -o /tmp/gen_mel && /tmp/gen_mel > audio_processor_b200/csrc/afe_mel_gen.h
#include <cstdio>
#include <vector>
#include "afe_tables.h"
using namespace afe;

static std::vector<int> partition(const std::vector<int>& cost, int parts) {
  // contiguous partition minimising the maximum part cost (binary search + greedy)
  long lo = 0, hi = 0;
  for (int c : cost) { if (c > lo) lo = c; hi += c; }
  auto cuts_for = [&](long cap, std::vector<int>* cuts) {
    int used = 1; long acc = 0;
    if (cuts) { cuts->clear(); cuts->push_back(0); }
    for (size_t i = 0; i < cost.size(); ++i) {
      if (acc + cost[i] > cap) { ++used; acc = 0; if (cuts) cuts->push_back((int)i); }
      acc += cost[i];
    }
    return used;
  };
  while (lo < hi) { long mid = (lo + hi) / 2; if (cuts_for(mid, nullptr) <= parts) hi = mid; else lo = mid + 1; }
  std::vector<int> cuts; cuts_for(lo, &cuts);
  while ((int)cuts.size() < parts) cuts.push_back((int)cost.size());
  cuts.push_back((int)cost.size());
  return cuts;
}

static void emit(const SparseBank& b, const char* name) {
  const int F = b.filters;
  std::vector<int> cost(F + 1);
  for (int i = 0; i <= F; ++i) {
    int c = 2;
    for (int k = b.iv_start[i]; k < b.iv_start[i] + b.iv_count[i]; ++k) c += 1 + (b.up[k] != 0.f) + (b.dn[k] != 0.f);
    cost[i] = c;
  }
  std::vector<int> cuts = partition(cost, 16);
  printf("// %s: %d filters, %d intervals; role costs:", name, F, F + 1);
  for (int r = 0; r < 16; ++r) { int c = 0; for (int i = cuts[r]; i < cuts[r + 1]; ++i) c += cost[i]; printf(" %d", c); }
  printf("\ntemplate <int TF> AFE_HD void mel_intervals_%s(int role, int f, const float* AFE_RESTRICT sP, float* AFE_RESTRICT sU, float* AFE_RESTRICT sD) {\n", name);
  printf("  float u = 0.f, d = 0.f;\n  switch (role) {\n");
  for (int r = 0; r < 16; ++r) {
    printf("    case %d:\n", r);
    for (int i = cuts[r]; i < cuts[r + 1]; ++i) {
      printf("     ");
      for (int k = b.iv_start[i]; k < b.iv_start[i] + b.iv_count[i]; ++k) {
        if (b.up[k] != 0.f && b.dn[k] != 0.f) printf(" AFE_MEL_UD(%d, %af, %af)", k, b.up[k], b.dn[k]);
        else if (b.up[k] != 0.f) printf(" AFE_MEL_U(%d, %af)", k, b.up[k]);
        else if (b.dn[k] != 0.f) printf(" AFE_MEL_D(%d, %af)", k, b.dn[k]);
      }
      printf(" AFE_MEL_END(%d)\n", i);
    }
    printf("      break;\n");
  }
  printf("  }\n}\n\n");
}

// Fused form: role r owns the FILTERS [cuts[r], cuts[r+1]) and walks the intervals
// cuts[r] .. cuts[r+1] (the last one for its down-slopes only), keeping U of the previous
// interval in a register: filter m = U[m] + D[m+1] leaves through out(m, value) with no
// shared-memory round trip and no barrier between the bank and the log.
static void emit_fused(const SparseBank& b, const char* name) {
  const int F = b.filters;
  std::vector<int> cost(F);
  for (int m = 0; m < F; ++m) {
    int c = 4;
    for (int i = m; i <= m + 1; ++i)
      for (int k = b.iv_start[i]; k < b.iv_start[i] + b.iv_count[i]; ++k) c += 2;
    cost[m] = c;
  }
  std::vector<int> cuts = partition(cost, 16);
  printf("// %s fused: filters per role:", name);
  for (int r = 0; r < 16; ++r) printf(" %d", cuts[r + 1] - cuts[r]);
  printf("\ntemplate <int TF, class Out> AFE_HD void mel_filters_%s(int role, int f, const float* AFE_RESTRICT sP, Out out) {\n", name);
  printf("  float u = 0.f, d = 0.f, pu = 0.f;\n  switch (role) {\n");
  for (int r = 0; r < 16; ++r) {
    printf("    case %d:\n", r);
    for (int i = cuts[r]; i <= cuts[r + 1] && cuts[r] < cuts[r + 1]; ++i) {
      const bool first = (i == cuts[r]), last = (i == cuts[r + 1]);
      printf("     ");
      for (int k = b.iv_start[i]; k < b.iv_start[i] + b.iv_count[i]; ++k) {
        const bool hu = b.up[k] != 0.f && !last && i < F, hd = b.dn[k] != 0.f && !first;
        if (hu && hd) printf(" AFE_MEL_UD(%d, %af, %af)", k, b.up[k], b.dn[k]);
        else if (hu) printf(" AFE_MEL_U(%d, %af)", k, b.up[k]);
        else if (hd) printf(" AFE_MEL_D(%d, %af)", k, b.dn[k]);
      }
      if (first) printf(" AFE_MEL_FIRST()\n"); else printf(" AFE_MEL_OUT(%d)\n", i - 1);
    }
    printf("      break;\n");
  }
  printf("  }\n}\n\n");
}

// DCT-II (13 x 128, ortho) as straight-line FFMAs with immediate coefficients: warp W of the
// MFCC kernel owns mel bins [32W, 32W+32) of its lane's frame (row = 8 float4 of shared memory).
static void emit_dct() {
  const std::vector<float> d = make_dct_matrix();
  printf("#define AFE_DCT13(x");
  for (int c = 0; c < kMfcc; ++c) printf(", c%d", c);
  printf(") {");
  for (int c = 0; c < kMfcc; ++c) printf(" a[%d] = fma_f(x, c%d, a[%d]);", c, c, c);
  printf(" }\n");
  for (int w = 0; w < 4; ++w) {
    printf("AFE_HD void dct13_partial_%d(const float4* AFE_RESTRICT row, float (&a)[%d]) {\n", w, kMfcc);
    for (int q = 0; q < 8; ++q) {
      printf("  { const float4 v = row[%d];\n", q);
      for (int j = 0; j < 4; ++j) {
        const int n = 32 * w + 4 * q + j;
        printf("    AFE_DCT13(v.%c", "xyzw"[j]);
        for (int c = 0; c < kMfcc; ++c) printf(", %af", d[(size_t)c * kAcMel + n]);
        printf(")\n");
      }
      printf("  }\n");
    }
    printf("}\n");
  }
  printf("#undef AFE_DCT13\n\n");
}

int main() {
  printf("// GENERATED by tools/gen_mel.cc from afe_tables.cc -- do not edit.\n");
  printf("// Interval form of the triangular banks (afe_tables.h): a bin in interval i feeds the up-slope\n");
  printf("// of filter i and the down-slope of filter i-1; filter m = U[m] + D[m+1].  Role r (warp r)\n");
  printf("// owns a contiguous run of intervals; lanes are frames (TF per tile).  sP[bin*TF+f], sU/sD[interval*TF+f].\n");
  printf("#pragma once\n#include \"afe_platform.h\"\nnamespace afe {\n");
  printf("#define AFE_MEL_UD(k, wu, wd) { const float p = sP[(k) * TF + f]; u = fma_f(wu, p, u); d = fma_f(wd, p, d); }\n");
  printf("#define AFE_MEL_U(k, wu) { u = fma_f(wu, sP[(k) * TF + f], u); }\n");
  printf("#define AFE_MEL_D(k, wd) { d = fma_f(wd, sP[(k) * TF + f], d); }\n");
  printf("#define AFE_MEL_END(i) { sU[(i) * TF + f] = u; sD[(i) * TF + f] = d; u = 0.f; d = 0.f; }\n\n");
  emit(make_kaldi_mel_bank(), "kaldi");
  emit(make_htk_mel_bank(), "htk");
  printf("#undef AFE_MEL_UD\n#undef AFE_MEL_U\n#undef AFE_MEL_D\n#undef AFE_MEL_END\n\n");
  printf("#define AFE_MEL_UD(k, wu, wd) { const float p = sP[(k) * TF + f]; u = fma_f(wu, p, u); d = fma_f(wd, p, d); }\n");
  printf("#define AFE_MEL_U(k, wu) { u = fma_f(wu, sP[(k) * TF + f], u); }\n");
  printf("#define AFE_MEL_D(k, wd) { d = fma_f(wd, sP[(k) * TF + f], d); }\n");
  printf("#define AFE_MEL_FIRST() { pu = u; u = 0.f; d = 0.f; }\n");
  printf("#define AFE_MEL_OUT(m) { out(m, pu + d); pu = u; u = 0.f; d = 0.f; }\n\n");
  emit_fused(make_kaldi_mel_bank(), "kaldi");
  emit_fused(make_htk_mel_bank(), "htk");
  printf("#undef AFE_MEL_UD\n#undef AFE_MEL_U\n#undef AFE_MEL_D\n#undef AFE_MEL_FIRST\n#undef AFE_MEL_OUT\n\n");
  emit_dct();
  printf("}  // namespace afe\n");
  return 0;
}
