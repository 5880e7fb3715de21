// Docker-side text on the device: what the reference reads from
// `docker inspect <container>` on every poll (gpu.py:86-88,184-199): the label gate
// (labels['mlgpu'] == 'mlgpu'), io.kubernetes.pod.name / .namespace, and the user_name split rules.
//
//     container = command(['docker','inspect',id]).replace("\n","").replace(" ","")    gpu.py:87
//     labels    = json.loads(container)[0]['Config']['Labels']                        gpu.py:88,184
//
// Per-thread code compiled twice, like the other leaf headers: by nvcc for k_inspect (kgm_replay.cu's
// neighbour kgm_docker.cu) and by g++ for tests/native/hostcheck.cpp, which holds it to kgm/fleet.py
// (json.loads and str.split themselves) on generated and damaged texts.
//
// The scanner ACCEPTS A SUBSET of what json.loads accepts and says IRREGULAR for everything else; an
// irregular container is parsed by kgm/fleet.py on the host as before (that is where every raise
// point of the reference -- JSONDecodeError, KeyError, IndexError, TypeError, AttributeError -- is
// classified).  What it accepts, it must get exactly right:
//   * ' ' and '\n' do not exist (the reference removes them from the WHOLE text first, inside
//     strings, numbers and literals too); '\t' and '\r' are white space between tokens only;
//   * strings: bytes 0x20..0x7e without '"' and '\'; numbers: the JSON grammar; true / false / null;
//     nesting up to 64; one top-level array and nothing behind it;
//   * element 0 is an object with exactly one "Config" key whose value is an object with exactly
//     one "Labels" key whose value is an object; in it "mlgpu", "io.kubernetes.pod.name" and
//     "io.kubernetes.pod.namespace" occur at most once each and have string values; a pod name
//     comes with a namespace; the two values contain no removed byte (so they are runs of the text);
//   * user_name = podname.split("turing-gpu-notebook-")[1].rsplit("-",2)[0] if the marker occurs,
//     else podname.split("-")[1], which needs a '-' (gpu.py:196-199).
#pragma once
#include <stdint.h>

#include "kgm_d2s.cuh"    // KGM_HD

namespace kgm {
namespace dk {

struct Fields {
  uint32_t status;        // 0 = taken, 1 = irregular (the host parses this container)
  uint32_t gated;         // labels['mlgpu'] == 'mlgpu'
  uint32_t has_pod;       // io.kubernetes.pod.name present (with its namespace)
  uint32_t pod_off, pod_len, ns_off, ns_len, user_off, user_len;   // runs of the container's text
};

struct Cursor {
  const uint8_t *t;
  uint32_t n, p;
};
// next byte of the squeezed text (' ' and '\n' removed), or 256 at the end; does not advance
KGM_HD uint32_t peek(Cursor &c) {
  while (c.p < c.n && (c.t[c.p] == 32u || c.t[c.p] == 10u)) ++c.p;
  return c.p < c.n ? c.t[c.p] : 256u;
}
KGM_HD uint32_t take(Cursor &c) {
  const uint32_t b = peek(c);
  if (b != 256u) ++c.p;
  return b;
}
KGM_HD void skip_ws(Cursor &c) {                       // '\t' and '\r' between tokens
  for (;;) {
    const uint32_t b = peek(c);
    if (b != 9u && b != 13u) return;
    ++c.p;
  }
}
// A string whose opening quote has been taken.  `lit` (or null): is the squeezed content this
// literal?  `off`/`len`: the raw extent of the content; `pure`: no removed byte inside it.
KGM_HD bool scan_string(Cursor &c, const char *lit, bool *is_lit, uint32_t *off, uint32_t *len, bool *pure) {
  peek(c);
  const uint32_t start = c.p;                            // (removed bytes in front of the content belong to it)
  uint32_t k = 0, kept = 0;
  bool same = lit != nullptr;
  for (;;) {
    if (c.p >= c.n) return false;
    const uint32_t b = c.t[c.p];
    if (b == 32u || b == 10u) { ++c.p; continue; }
    if (b == 34u) break;
    if (b < 32u || b >= 127u || b == 92u) return false;
    if (same) { if (lit[k] && (uint8_t)lit[k] == b) ++k; else same = false; }
    ++kept;
    ++c.p;
  }
  const uint32_t end = c.p;
  ++c.p;                                                 // closing quote
  if (is_lit) *is_lit = same && lit[k] == 0;
  // raw extent without removed bytes at either end; pure if nothing was removed in between
  uint32_t a = start, z = end;
  while (a < z && (c.t[a] == 32u || c.t[a] == 10u)) ++a;
  while (z > a && (c.t[z - 1] == 32u || c.t[z - 1] == 10u)) --z;
  if (off) *off = a;
  if (len) *len = z - a;
  if (pure) *pure = (z - a) == kept;
  return true;
}
KGM_HD bool is_digit(uint32_t b) { return b - 48u <= 9u; }
KGM_HD bool scan_number(Cursor &c) {                     // first byte not yet taken: '-' or a digit
  if (peek(c) == 45u) take(c);
  uint32_t b = take(c);
  if (!is_digit(b)) return false;
  if (b != 48u) while (is_digit(peek(c))) take(c);
  if (peek(c) == 46u) {
    take(c);
    if (!is_digit(take(c))) return false;
    while (is_digit(peek(c))) take(c);
  }
  b = peek(c);
  if (b == 101u || b == 69u) {
    take(c);
    b = peek(c);
    if (b == 43u || b == 45u) take(c);
    if (!is_digit(take(c))) return false;
    while (is_digit(peek(c))) take(c);
  }
  return true;
}
KGM_HD bool scan_literal(Cursor &c, const char *w) {
  for (uint32_t i = 0; w[i]; ++i)
    if (take(c) != (uint8_t)w[i]) return false;
  return true;
}

// text[lo, hi) contains `m`?  first occurrence at or after `from`, or hi
KGM_HD uint32_t find_sub(const uint8_t *t, uint32_t from, uint32_t hi, const char *m, uint32_t mlen) {
  for (uint32_t i = from; i + mlen <= hi; ++i) {
    uint32_t k = 0;
    while (k < mlen && t[i + k] == (uint8_t)m[k]) ++k;
    if (k == mlen) return i;
  }
  return hi;
}
// gpu.py:196-199 on the pod name text[off, off+len); false = the reference raises (IndexError)
KGM_HD bool user_name(const uint8_t *t, uint32_t off, uint32_t len, uint32_t *uo, uint32_t *ul) {
  const char *marker = "turing-gpu-notebook-";
  const uint32_t ml = 20, hi = off + len;
  const uint32_t m = find_sub(t, off, hi, marker, ml);
  if (m < hi) {
    const uint32_t a = m + ml, z = find_sub(t, a, hi, marker, ml);       // split(marker)[1]
    uint32_t cut = z, dashes = 0;                                         // rsplit("-", 2)[0]
    for (uint32_t i = z; i > a; --i)
      if (t[i - 1] == 45u) { cut = i - 1; if (++dashes == 2) break; }
    *uo = a; *ul = (dashes ? cut : z) - a;
    return true;
  }
  uint32_t d1 = off;                                                      // split("-")[1]
  while (d1 < hi && t[d1] != 45u) ++d1;
  if (d1 == hi) return false;
  uint32_t d2 = d1 + 1;
  while (d2 < hi && t[d2] != 45u) ++d2;
  *uo = d1 + 1; *ul = d2 - d1 - 1;
  return true;
}

// One container's `docker inspect` text -> Fields.
KGM_HD Fields inspect_fields(const uint8_t *text, uint32_t n) {
  Fields f;
  f.status = 1; f.gated = 0; f.has_pod = 0;
  f.pod_off = f.pod_len = f.ns_off = f.ns_len = f.user_off = f.user_len = 0;
  Cursor c{text, n, 0};
  enum { VALUE, KEY, AFTER };
  enum { W_NONE, W_CFG, W_LAB, W_MLGPU, W_POD, W_NS };
  uint64_t kinds = 0;                                    // bit d-1: the container at depth d is an object
  uint32_t depth = 0, idx0 = 0, state = VALUE, want = W_NONE;
  bool cfg_open = false, lab_open = false, gate_is_mlgpu = false, pod_pure = false, ns_pure = false;
  uint32_t n_cfg = 0, n_lab = 0, n_mlgpu = 0, n_pod = 0, n_ns = 0;
  for (uint32_t guard = 0; guard <= 4u * n + 16u; ++guard) {
    skip_ws(c);
    if (state == VALUE) {
      const uint32_t b = peek(c);
      const uint32_t w = want;
      want = W_NONE;
      if (depth == 0 && b != 91u) return f;              // not a list: [0] raises
      if (depth == 1 && idx0 == 0 && b != 123u) return f;   // element 0 is not an object
      if ((w == W_CFG || w == W_LAB) && b != 123u) return f;
      if (w >= W_MLGPU && b != 34u) return f;            // a label value that is not a string
      if (b == 123u || b == 91u) {
        take(c);
        if (depth >= 64u) return f;
        kinds = b == 123u ? kinds | (1ull << depth) : kinds & ~(1ull << depth);
        ++depth;
        if (w == W_CFG) cfg_open = true;
        if (w == W_LAB) lab_open = true;
        skip_ws(c);
        const uint32_t e = peek(c);
        if (e == (b == 123u ? 125u : 93u)) {             // empty container
          if (depth == 1) return f;                      // []: [0] raises
          take(c);
          --depth;
          if (depth == 3) lab_open = false;
          if (depth == 2) cfg_open = false;
          state = AFTER;
        } else {
          state = b == 123u ? KEY : VALUE;
        }
      } else if (b == 34u) {
        take(c);
        bool lit = false, pure = false;
        uint32_t off = 0, len = 0;
        if (!scan_string(c, w == W_MLGPU ? "mlgpu" : nullptr, &lit, &off, &len, &pure)) return f;
        if (w == W_MLGPU) gate_is_mlgpu = lit;
        if (w == W_POD) { f.pod_off = off; f.pod_len = len; pod_pure = pure; }
        if (w == W_NS) { f.ns_off = off; f.ns_len = len; ns_pure = pure; }
        state = AFTER;
      } else if (b == 45u || is_digit(b)) {
        if (!scan_number(c)) return f;
        state = AFTER;
      } else if (b == 116u) { if (!scan_literal(c, "true")) return f; state = AFTER; }
      else if (b == 102u) { if (!scan_literal(c, "false")) return f; state = AFTER; }
      else if (b == 110u) { if (!scan_literal(c, "null")) return f; state = AFTER; }
      else return f;
    } else if (state == KEY) {
      if (take(c) != 34u) return f;
      const bool in_e0 = depth == 2 && idx0 == 0, in_cfg = depth == 3 && cfg_open, in_lab = depth == 4 && lab_open;
      const uint32_t save = c.p;
      bool lit = false;
      if (!scan_string(c, in_e0 ? "Config" : in_cfg ? "Labels" : in_lab ? "mlgpu" : nullptr, &lit, nullptr, nullptr, nullptr))
        return f;
      if (in_e0 && lit) { ++n_cfg; want = W_CFG; }
      if (in_cfg && lit) { ++n_lab; want = W_LAB; }
      if (in_lab) {
        if (lit) { ++n_mlgpu; want = W_MLGPU; }
        else {                                           // the two other label keys: scan the key again
          Cursor k{text, n, save};
          bool is_pod = false, is_ns = false;
          scan_string(k, "io.kubernetes.pod.name", &is_pod, nullptr, nullptr, nullptr);
          k.p = save;
          scan_string(k, "io.kubernetes.pod.namespace", &is_ns, nullptr, nullptr, nullptr);
          if (is_pod) { ++n_pod; want = W_POD; }
          if (is_ns) { ++n_ns; want = W_NS; }
        }
      }
      skip_ws(c);
      if (take(c) != 58u) return f;
      state = VALUE;
    } else {                                             // AFTER a value
      if (depth == 0) {
        if (peek(c) != 256u) return f;                   // extra data
        break;
      }
      const uint32_t b = take(c);
      const bool obj = (kinds >> (depth - 1)) & 1ull;
      if (b == 44u) {
        if (depth == 1) ++idx0;
        state = obj ? KEY : VALUE;
      } else if (b == (obj ? 125u : 93u)) {
        --depth;
        if (depth == 3) lab_open = false;
        if (depth == 2) cfg_open = false;
      } else {
        return f;
      }
    }
  }
  if (depth != 0 || state != AFTER) return f;
  if (n_cfg != 1 || n_lab != 1 || n_mlgpu > 1 || n_pod > 1 || n_ns > 1) return f;
  if (n_pod != n_ns) return f;                           // a pod name without a namespace: KeyError; the reverse: not read, but rare -- host
  if (n_pod && (!pod_pure || !ns_pure)) return f;
  f.gated = n_mlgpu && gate_is_mlgpu;
  f.has_pod = n_pod;
  if (f.gated && f.has_pod && !user_name(text, f.pod_off, f.pod_len, &f.user_off, &f.user_len)) return f;
  f.status = 0;
  return f;
}

}  // namespace dk
}  // namespace kgm
