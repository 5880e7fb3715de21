// Shortest round-trip formatting of an IEEE double in CPython's repr() layout.
//
// The reference prints mem_util with '%s' % float (gpu.py:135-139), i.e. repr(float):
// the shortest decimal string that reads back to the same double, in fixed notation when
// -4 <= exp10 < 16 and d[.ddd]e+XX otherwise, with ".0" appended to integers.  This
// header produces those bytes on the device (and, for the CPU unit tests, on the host).
//
// Digit generation follows Ulf Adams' Ryu algorithm ("Ryu: fast float-to-string conversion",
// PLDI 2018; reference implementation github.com/ulfjack/ryu, dual-licensed Apache-2.0 /
// Boost-1.0): scale the three neighbours of the value by a 128-bit fixed-point power of five,
// then strip digits while they stay distinguishable.  The helper names and multiplier constants
// (pow5_bits, log10_pow2, log10_pow5, mul_shift) are Ryu's; the code was restated for this
// library (device-compatible, non-negative finite inputs only), not copied from a release.
// Tables come from gen/make_d2s_tables.py.  Only finite, non-negative inputs occur on
// this path (used*100.0/total with used,total < 2^64, total != 0).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define KGM_HD __host__ __device__ __forceinline__
#define KGM_HDM __host__ __device__ __forceinline__     // member functions
#else
#define KGM_HD static inline
#define KGM_HDM inline
#endif

#if defined(__CUDA_ARCH__)
#define KGM_TABLE_QUAL static __device__ const
#else
#define KGM_TABLE_QUAL static const
#endif
#include "kgm_d2s_tables.h"

namespace kgm {

KGM_HD uint64_t mul64(uint64_t a, uint64_t b, uint64_t *hi) {
#if defined(__CUDA_ARCH__)
  *hi = __umul64hi(a, b);
  return a * b;
#else
  unsigned __int128 p = (unsigned __int128)a * b;
  *hi = (uint64_t)(p >> 64);
  return (uint64_t)p;
#endif
}

// (m * mul) >> j for a 128-bit mul = {lo, hi}; 64 <= j < 128 + 64.
KGM_HD uint64_t mul_shift(uint64_t m, const unsigned long long *mul, int j) {
  uint64_t h0, h1;
  (void)mul64(m, mul[0], &h0);
  uint64_t l1 = mul64(m, mul[1], &h1);
  uint64_t sum = h0 + l1;
  if (sum < h0) ++h1;
  int s = j - 64;
  return s == 0 ? sum : (sum >> s) | (h1 << (64 - s));
}

KGM_HD int pow5_bits(int e) { return (int)(((uint32_t)e * 1217359u) >> 19) + 1; }
KGM_HD int log10_pow2(int e) { return (int)(((uint32_t)e * 78913u) >> 18); }
KGM_HD int log10_pow5(int e) { return (int)(((uint32_t)e * 732923u) >> 20); }

KGM_HD bool divisible_by_pow5(uint64_t v, int p) {
  int n = 0;
  while (v % 5 == 0) { v /= 5; ++n; }
  return n >= p;
}

// Shortest decimal: value = digits * 10^exp10.  Input is the raw bit pattern of a finite
// positive double.
KGM_HD void shortest_decimal(uint64_t bits, uint64_t *digits, int *exp10) {
  const uint64_t frac = bits & ((1ull << 52) - 1);
  const int bexp = (int)((bits >> 52) & 0x7ff);
  int e2;
  uint64_t m2;
  if (bexp == 0) { e2 = 1 - 1023 - 52 - 2; m2 = frac; }
  else { e2 = bexp - 1023 - 52 - 2; m2 = (1ull << 52) | frac; }
  const bool accept = (m2 & 1) == 0;
  const uint64_t mv = 4 * m2;
  const uint32_t mm_shift = (frac != 0 || bexp <= 1) ? 1u : 0u;
  uint64_t vr, vp, vm;
  int e10;
  bool vm_tz = false, vr_tz = false;
  if (e2 >= 0) {
    const int q = log10_pow2(e2) - (e2 > 3);
    e10 = q;
    const int k = KGM_POW5_INV_BITS + pow5_bits(q) - 1;
    const int i = -e2 + q + k;
    vr = mul_shift(mv, kgm_pow5_inv_split[q], i);
    vp = mul_shift(mv + 2, kgm_pow5_inv_split[q], i);
    vm = mul_shift(mv - 1 - mm_shift, kgm_pow5_inv_split[q], i);
    if (q <= 21) {
      if (mv % 5 == 0) vr_tz = divisible_by_pow5(mv, q);
      else if (accept) vm_tz = divisible_by_pow5(mv - 1 - mm_shift, q);
      else vp -= divisible_by_pow5(mv + 2, q);
    }
  } else {
    const int q = log10_pow5(-e2) - (-e2 > 1);
    e10 = q + e2;
    const int i = -e2 - q;
    const int k = pow5_bits(i) - KGM_POW5_BITS;
    const int j = q - k;
    vr = mul_shift(mv, kgm_pow5_split[i], j);
    vp = mul_shift(mv + 2, kgm_pow5_split[i], j);
    vm = mul_shift(mv - 1 - mm_shift, kgm_pow5_split[i], j);
    if (q <= 1) {
      vr_tz = true;
      if (accept) vm_tz = mm_shift == 1;
      else --vp;
    } else if (q < 63) {
      vr_tz = (mv & ((1ull << q) - 1)) == 0;
    }
  }
  int removed = 0;
  uint32_t last = 0;
  uint64_t out;
  if (vm_tz || vr_tz) {
    while (vp / 10 > vm / 10) {
      vm_tz &= vm % 10 == 0;
      vr_tz &= last == 0;
      last = (uint32_t)(vr % 10);
      vr /= 10; vp /= 10; vm /= 10;
      ++removed;
    }
    if (vm_tz) {
      while (vm % 10 == 0) {
        vr_tz &= last == 0;
        last = (uint32_t)(vr % 10);
        vr /= 10; vp /= 10; vm /= 10;
        ++removed;
      }
    }
    if (vr_tz && last == 5 && vr % 2 == 0) last = 4;   // exactly halfway: round to even
    out = vr + (((vr == vm) && (!accept || !vm_tz)) || last >= 5);
  } else {
    bool up = false;
    while (vp / 10 > vm / 10) {
      up = vr % 10 >= 5;
      vr /= 10; vp /= 10; vm /= 10;
      ++removed;
    }
    out = vr + ((vr == vm) || up);
  }
  *digits = out;
  *exp10 = e10 + removed;
}

KGM_HD int decimal_length17(uint64_t v) {            // v < 10^17
  if (v >= 10000000000000000ull) return 17;
  if (v >= 1000000000000000ull) return 16;
  if (v >= 100000000000000ull) return 15;
  if (v >= 10000000000000ull) return 14;
  if (v >= 1000000000000ull) return 13;
  if (v >= 100000000000ull) return 12;
  if (v >= 10000000000ull) return 11;
  if (v >= 1000000000ull) return 10;
  const uint32_t w = (uint32_t)v;
  if (w >= 100000000u) return 9;
  if (w >= 10000000u) return 8;
  if (w >= 1000000u) return 7;
  if (w >= 100000u) return 6;
  if (w >= 10000u) return 5;
  if (w >= 1000u) return 4;
  if (w >= 100u) return 3;
  if (w >= 10u) return 2;
  return 1;
}

// The nd decimal digits of dg (< 10^17), most significant first, using 32-bit arithmetic
// for all but one division.
KGM_HD void decimal_digits17(uint64_t dg, int nd, unsigned char *d) {
  uint32_t hi = (uint32_t)(dg / 100000000ull);       // < 10^9
  uint32_t lo = (uint32_t)(dg - (uint64_t)hi * 100000000ull);
  int i = nd - 1;
  for (int k = 0; k < 8 && i >= 0; ++k, --i) { const uint32_t q = lo / 10u; d[i] = (unsigned char)('0' + (lo - q * 10u)); lo = q; }
  for (; i >= 0; --i) { const uint32_t q = hi / 10u; d[i] = (unsigned char)('0' + (hi - q * 10u)); hi = q; }
}

// repr(float) for a finite double >= 0.  Writes at most 24 bytes, returns the length.
KGM_HD int float_repr(double x, unsigned char *out) {
  union { double d; uint64_t u; } cvt;
  cvt.d = x;
  uint64_t bits = cvt.u & 0x7fffffffffffffffull;
  if (bits == 0) { out[0] = '0'; out[1] = '.'; out[2] = '0'; return 3; }
  uint64_t dg;
  int e10;
  shortest_decimal(bits, &dg, &e10);
  const int nd = decimal_length17(dg);
  unsigned char d[17];
  decimal_digits17(dg, nd, d);
  const int decpt = nd + e10;          // value = 0.d1..dn * 10^decpt
  int n = 0;
  if (decpt > 16 || decpt < -3) {      // exponent form: d[.ddd]e+XX
    out[n++] = d[0];
    if (nd > 1) {
      out[n++] = '.';
      for (int i = 1; i < nd; ++i) out[n++] = d[i];
    }
    out[n++] = 'e';
    int e = decpt - 1;
    if (e < 0) { out[n++] = '-'; e = -e; } else out[n++] = '+';
    if (e >= 100) { out[n++] = (unsigned char)('0' + e / 100); e %= 100; out[n++] = (unsigned char)('0' + e / 10); }
    else out[n++] = (unsigned char)('0' + e / 10);
    out[n++] = (unsigned char)('0' + e % 10);
  } else if (decpt <= 0) {             // 0.000ddd
    out[n++] = '0'; out[n++] = '.';
    for (int i = 0; i < -decpt; ++i) out[n++] = '0';
    for (int i = 0; i < nd; ++i) out[n++] = d[i];
  } else if (decpt >= nd) {            // ddd000.0
    for (int i = 0; i < nd; ++i) out[n++] = d[i];
    for (int i = nd; i < decpt; ++i) out[n++] = '0';
    out[n++] = '.'; out[n++] = '0';
  } else {                             // dd.ddd
    for (int i = 0; i < decpt; ++i) out[n++] = d[i];
    out[n++] = '.';
    for (int i = decpt; i < nd; ++i) out[n++] = d[i];
  }
  return n;
}

// mem_util exactly as gpu.py:135 computes it: int * 100.0 / int in IEEE double, each
// int->double conversion and each operation correctly rounded (no contraction).
KGM_HD double mem_util(uint64_t used, uint64_t total) {
#if defined(__CUDA_ARCH__)
  return __ddiv_rn(__dmul_rn(__ull2double_rn(used), 100.0), __ull2double_rn(total));
#else
  volatile double u = (double)used;
  volatile double p = u * 100.0;
  volatile double t = (double)total;
  return p / t;
#endif
}

}  // namespace kgm
