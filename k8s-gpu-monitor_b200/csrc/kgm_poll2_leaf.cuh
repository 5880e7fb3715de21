// Leaf functions and table layouts of the poll kernel k_poll (kgm_poll2.cuh).
//
// Like kgm_common.cuh / kgm_fast.cuh this is per-thread code compiled twice: by nvcc for the
// kernel (the word array is then the CTA's shared memory) and by g++ for the CPU tests in
// tests/native/hostcheck.cpp, which compose whole lines with it and compare them with the
// reference's bytes (gpu.py:136-143, 206-216, 240-250).
//
// Vocabulary
//   run word   one byte run inside the word array: offset << 12 | length (offset < 2^20 bytes,
//              length < 4096)
//   record     the run words of one output line's variable cells: 16 words per --query-gpu row
//              (GREC_*), 8 words per --query-compute-apps row or idle pseudo-row (PREC_*)
//   slot       the constant pieces of one node's / one container's line template, built once by
//              kgm_load_fleet and read from global memory: 48-byte header (8 run words relative
//              to the slot start, the constant byte count of a line), then the piece bytes, every
//              piece starting on a 16-byte boundary
//   segment    a line is a fixed sequence of segments, each a run taken from the slot ('T'), from
//              the table of constant strings ('K') or from the line's record ('R')
//   line writer  one lane writes one line: it streams the segments' bytes through a pending word
//              and stores aligned words only, so all lanes of a warp run the same segment sequence
#pragma once
#include <stdint.h>

#include "kgm_fast.cuh"

namespace kgm {
namespace p2 {

using namespace fast;

#if !defined(__CUDACC__)
struct alignas(16) uint4 { uint32_t x, y, z, w; };   // (host build of the leaf functions)
#endif

constexpr uint32_t KIND_NODE = 0, KIND_BUSY = 1, KIND_IDLE = 2;

KGM_HD uint32_t rw(uint32_t off, uint32_t len) { return (off << 12) | len; }
KGM_HD uint32_t rw_off(uint32_t w) { return w >> 12; }
KGM_HD uint32_t rw_len(uint32_t w) { return w & 0xfffu; }
constexpr uint32_t RW_MAX_LEN = 4095;

// ---- records -----------------------------------------------------------------------------------
constexpr uint32_t GREC_WORDS = 16, GREC_BYTES = 64;
constexpr uint32_t GREC_UTIL = 0, GREC_MU = 1, GREC_MT = 2, GREC_MF = 3, GREC_FSTR = 4, GREC_SID = 5,
                   GREC_KEY = 6, GREC_LINE_LEN = 7, GREC_HASH = 8 /* 2 words */, GREC_FSTR_BYTES = 40 /* byte offset */;
constexpr uint32_t PREC_WORDS = 8, PREC_BYTES = 32;
constexpr uint32_t PREC_PID = 0, PREC_NAME = 1 /* NAME_RUNS words */, PREC_MEM = 5, PREC_UTIL = 6, PREC_SID = 7;
static_assert(NAME_RUNS == 4, "PREC_NAME..PREC_MEM");

// ---- constant strings shared by a CTA ------------------------------------------------------------
// The ConstArea of kgm_fast.cuh (strings + decimal strings), then the table of 'K' run words, the
// segment tables and the constant part of an idle pseudo-row.
struct CA2 {
  static constexpr uint32_t NL = 120;                                   // "\n" (free bytes of ConstArea)
  static constexpr uint32_t KTAB = (ConstArea::BYTES + 15u) & ~15u;      // 16 run words
  static constexpr uint32_t SEGTAB = KTAB + 64;                          // 2 x 32 segment descriptors
  static constexpr uint32_t IDLE_REC = SEGTAB + 64;                      // 8 run words
  static constexpr uint32_t BYTES = IDLE_REC + 32;
};
constexpr uint32_t K_GPU_UTIL = 0, K_MEM_UTIL = 1, K_MEM_USED = 2, K_MEM_TOTAL = 3, K_MEM_FREE = 4,
                   K_GPU_UTIL_T = 5, K_GPU_ID_V = 6, K_USED_MEMORY = 7, K_NL = 8, K_NONE = 9, K_DASH = 10,
                   K_ZERO = 11, K_EMPTY = 12;

// ---- segment tables ----------------------------------------------------------------------------
// descriptor = type << 6 | index;  type 0 = slot run word, 1 = 'K' run word, 2 = record run word
constexpr uint32_t SEG_T = 0u << 6, SEG_K = 1u << 6, SEG_R = 2u << 6;
constexpr uint32_t NODE_SEGS = 15, POD_SEGS = 20;

// node_gpu,host=H,gpu_id=<K> gpu_util=<U>,mem_util=<F>,mem_used=<MU>,mem_total=<MT>,node="H",gid=<K>,mem_free=<MF>\n
KGM_HD constexpr uint32_t seg_node(int k) {
  return k == 0 ? SEG_T | 0 : k == 1 ? SEG_R | GREC_SID : k == 2 ? SEG_K | K_GPU_UTIL : k == 3 ? SEG_R | GREC_UTIL
       : k == 4 ? SEG_K | K_MEM_UTIL : k == 5 ? SEG_R | GREC_FSTR : k == 6 ? SEG_K | K_MEM_USED
       : k == 7 ? SEG_R | GREC_MU : k == 8 ? SEG_K | K_MEM_TOTAL : k == 9 ? SEG_R | GREC_MT : k == 10 ? SEG_T | 1
       : k == 11 ? SEG_R | GREC_SID : k == 12 ? SEG_K | K_MEM_FREE : k == 13 ? SEG_R | GREC_MF : SEG_K | K_NL;
}
// pod_gpu,host=H,namespace=NS,gpu_id=<G>,containerid=C,pid_t=<P>,gpu_util_t=<U>,user_name_t=UN procressname="<PN>",
// podname="POD",pid="<P>",gpu_id_v="<G>",node="H",user_name="UN",gpu_util=<U>,used_memory=<UM>\n
KGM_HD constexpr uint32_t seg_pod(int k) {
  return k == 0 ? SEG_T | 0 : k == 1 ? SEG_R | PREC_SID : k == 2 ? SEG_T | 1 : k == 3 ? SEG_R | PREC_PID
       : k == 4 ? SEG_K | K_GPU_UTIL_T : k == 5 ? SEG_R | PREC_UTIL : k == 6 ? SEG_T | 2
       : k == 7 ? SEG_R | (PREC_NAME + 0) : k == 8 ? SEG_R | (PREC_NAME + 1) : k == 9 ? SEG_R | (PREC_NAME + 2)
       : k == 10 ? SEG_R | (PREC_NAME + 3) : k == 11 ? SEG_T | 3 : k == 12 ? SEG_R | PREC_PID
       : k == 13 ? SEG_K | K_GPU_ID_V : k == 14 ? SEG_R | PREC_SID : k == 15 ? SEG_T | 4
       : k == 16 ? SEG_R | PREC_UTIL : k == 17 ? SEG_K | K_USED_MEMORY : k == 18 ? SEG_R | PREC_MEM : SEG_K | K_NL;
}
// The constant area is filled in three independent parts (the kernel gives them to different
// threads; the area must be zeroed first).  `abs0` = its byte offset in the word array that
// run words refer to.
KGM_HD void fill_ca2_strings(uint8_t *c) {
  fill_const_strings(c);
  c[CA2::NL] = 10u;
}
KGM_HD void fill_ca2_tables(uint8_t *c, uint32_t abs0) {
  uint32_t *kt = reinterpret_cast<uint32_t *>(c + CA2::KTAB);
  for (int i = 0; i < 16; ++i) kt[i] = rw(abs0, 0);
  kt[K_GPU_UTIL] = rw(abs0 + ConstArea::GPU_UTIL, 10);
  kt[K_MEM_UTIL] = rw(abs0 + ConstArea::MEM_UTIL, 10);
  kt[K_MEM_USED] = rw(abs0 + ConstArea::MEM_USED, 10);
  kt[K_MEM_TOTAL] = rw(abs0 + ConstArea::MEM_TOTAL, 11);
  kt[K_MEM_FREE] = rw(abs0 + ConstArea::MEM_FREE, 10);
  kt[K_GPU_UTIL_T] = rw(abs0 + ConstArea::GPU_UTIL_T, 12);
  kt[K_GPU_ID_V] = rw(abs0 + ConstArea::GPU_ID_V, 12);
  kt[K_USED_MEMORY] = rw(abs0 + ConstArea::USED_MEMORY, 13);
  kt[K_NL] = rw(abs0 + CA2::NL, 1);
  kt[K_NONE] = rw(abs0 + ConstArea::NONE, 4);
  kt[K_DASH] = rw(abs0 + ConstArea::DASH, 1);
  kt[K_ZERO] = rw(abs0 + ConstArea::ZERO, 1);
  // idle pseudo-row (gpu.py:234-239): pid "-", procressname "-", used_memory 0, gpu_util 0
  uint32_t *ir = reinterpret_cast<uint32_t *>(c + CA2::IDLE_REC);
  ir[PREC_PID] = kt[K_DASH];
  ir[PREC_NAME] = kt[K_DASH];
  ir[PREC_NAME + 1] = ir[PREC_NAME + 2] = ir[PREC_NAME + 3] = kt[K_EMPTY];
  ir[PREC_MEM] = kt[K_ZERO];
  ir[PREC_UTIL] = kt[K_ZERO];
  ir[PREC_SID] = kt[K_NONE];
}
KGM_HD void fill_ca2_segtab(uint8_t *c, int k) {                         // k in 0..31
  uint8_t *st = c + CA2::SEGTAB;
  uint32_t a = SEG_K | K_EMPTY, b = SEG_K | K_EMPTY;
  for (int i = 0; i < (int)POD_SEGS; ++i) {
    if (i == k && i < (int)NODE_SEGS) a = seg_node(i);
    if (i == k) b = seg_pod(i);
  }
  st[k] = (uint8_t)a;
  st[32 + k] = (uint8_t)b;
}
KGM_HD void fill_ca2(uint8_t *c, uint32_t abs0) {                       // the whole area, one thread (host)
  for (uint32_t i = 0; i < CA2::BYTES; ++i) c[i] = 0;
  fill_ca2_strings(c);
  for (uint32_t v = 0; v <= MAX_FAST_SID; ++v) fill_const_decimal(c, v);
  fill_ca2_tables(c, abs0);
  for (int k = 0; k < 32; ++k) fill_ca2_segtab(c, k);
}

KGM_HD uint32_t sid_run(uint32_t abs0, uint32_t sid) {                   // '%s' % shortunitids.get(uuid)
  if (sid == NONE32) return rw(abs0 + ConstArea::NONE, 4);
  return rw(abs0 + ConstArea::DEC + 4u * sid, dec_len(sid));
}

// ---- slots -------------------------------------------------------------------------------------
constexpr uint32_t SLOT_HEADER = 48;
constexpr uint32_t SLOT_FIXED = 46;          // u16: constant bytes of a line (incl. '\n')

// lengths of the 'K' strings (the host needs them for the slots' constant-byte prefixes)
KGM_HD constexpr uint32_t k_len(uint32_t idx) {
  return idx == K_MEM_TOTAL ? 11u : idx == K_GPU_UTIL_T || idx == K_GPU_ID_V ? 12u : idx == K_USED_MEMORY ? 13u
       : idx == K_NL || idx == K_DASH || idx == K_ZERO ? 1u : idx == K_NONE ? 4u : idx == K_EMPTY ? 0u : 10u;
}

// ---- measuring a line: the bytes of its variable cells -----------------------------------------
template <bool POD>
KGM_HD uint32_t line_var_bytes(const uint32_t *r) {
  uint32_t acc = 0;
  constexpr int NS = POD ? (int)POD_SEGS : (int)NODE_SEGS;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int k = 0; k < NS; ++k) {
    const uint32_t e = POD ? seg_pod(k) : seg_node(k);
    if ((e >> 6) == 2u) acc += rw_len(r[e & 63u]);
  }
  return acc;
}

// ---- composing --------------------------------------------------------------------------------
// A lane writes its line front to back.  The bytes not yet stored (fewer than 4) wait in `pend`,
// low byte first; only whole aligned words are stored.  A line that starts inside a word stores
// that word with zeros in the bytes of its predecessor; the predecessor's last bytes are OR-ed
// into it afterwards (lw_finish), when every lane of the round has written its words.
struct LineWriter {
  uint32_t *d;        // next aligned word of the destination
  uint32_t pend;      // pending bytes, low byte first; bytes >= fill are zero
  uint32_t fill;      // 0..3
};
KGM_HD uint32_t shl8(uint32_t v, uint32_t nbytes) { return nbytes >= 4u ? 0u : v << (8u * nbytes); }
KGM_HD uint32_t shr8(uint32_t v, uint32_t nbytes) { return nbytes >= 4u ? 0u : v >> (8u * nbytes); }

// r bytes (1..3), the low bytes of `cur` (its other bytes zero)
KGM_HD void lw_tail(LineWriter &w, uint32_t cur, uint32_t r) {
  const uint32_t x = w.pend | shl8(cur, w.fill), total = w.fill + r;
  if (total >= 4u) {
    *w.d++ = x;
    w.pend = shr8(cur, 4u - w.fill);
    w.fill = total - 4u;
  } else {
    w.pend = x;
    w.fill = total;
  }
}
// n bytes starting at byte `a` (0..3) of the aligned word array src[]; reads src[0 .. (a+n+3)/4]
KGM_HD void lw_put(LineWriter &w, const uint32_t *src, uint32_t a, uint32_t n) {
  const uint32_t sa = 8u * a, full = n >> 2;
  uint32_t lo = src[0];
  if (full) {
    // whole words: the stream word (source realigned) goes out behind the pending bytes
    const uint32_t sh = 8u * w.fill;
    uint32_t prev = shl8(w.pend, 4u - w.fill);       // pending bytes at the top of `prev`
    for (uint32_t j = 0; j < full; ++j) {
      const uint32_t hi = src[j + 1];
      const uint32_t cur = funnel_r(lo, hi, sa);
#if defined(__CUDA_ARCH__)
      w.d[j] = __funnelshift_l(prev, cur, sh);
#else
      w.d[j] = sh ? (prev >> (32u - sh)) | (cur << sh) : cur;
#endif
      prev = cur;
      lo = hi;
    }
    w.d += full;
    w.pend = shr8(prev, 4u - w.fill);
  }
  const uint32_t r = n & 3u;
  if (r) lw_tail(w, funnel_r(lo, src[full + 1], sa) & low_mask(r), r);
}
// n bytes of a slot piece: `g` is 16-byte aligned, the piece starts at its first byte and is
// padded to a multiple of 16 bytes
KGM_HD void lw_put_piece(LineWriter &w, const uint4 *g, uint32_t n) {
  const uint32_t sh = 8u * w.fill;
  uint32_t prev = shl8(w.pend, 4u - w.fill);
  uint32_t *d = w.d;
  uint32_t left = n;
  auto word = [&](uint32_t cur) {
#if defined(__CUDA_ARCH__)
    *d++ = __funnelshift_l(prev, cur, sh);
#else
    *d++ = sh ? (prev >> (32u - sh)) | (cur << sh) : cur;
#endif
    prev = cur;
  };
  for (; left >= 16u; left -= 16u, ++g) {
#if defined(__CUDA_ARCH__)
    const uint4 v = __ldg(g);
#else
    const uint4 v = *g;
#endif
    word(v.x); word(v.y); word(v.z); word(v.w);
  }
  uint32_t last = 0;
  if (left) {
#if defined(__CUDA_ARCH__)
    const uint4 v = __ldg(g);
#else
    const uint4 v = *g;
#endif
    const uint32_t fullw = left >> 2;
    if (fullw > 0) word(v.x);
    if (fullw > 1) word(v.y);
    if (fullw > 2) word(v.z);
    last = fullw == 0 ? v.x : fullw == 1 ? v.y : fullw == 2 ? v.z : v.w;
  }
  w.d = d;
  w.pend = shr8(prev, 4u - w.fill);
  const uint32_t r = left & 3u;
  if (r) lw_tail(w, last & low_mask(r), r);
}

// The table-driven form of compose_line (below): one loop over the segment table in the constant
// area.  Kept as the plain statement of what a line is (the CPU tests compare the two); the
// kernels call the unrolled form.
// Writes one line at byte offset `dst` of the word array B.  `ca` = byte offset of the constant
// area (CA2) in B, `slot` = the line's template slot (global memory), `rec` = byte offset of its
// record in B.  Returns the writer: its pending bytes still have to be stored (lw_finish).
KGM_HD LineWriter compose_line_table(uint32_t *B, uint32_t ca, uint32_t pod, const uint8_t *slot, uint32_t rec,
                               uint32_t dst) {
  const uint8_t *Bb = reinterpret_cast<const uint8_t *>(B);
  const uint8_t *st = Bb + ca + CA2::SEGTAB + (pod ? 32u : 0u);
  const uint32_t *sw = reinterpret_cast<const uint32_t *>(slot);
  const uint32_t nseg = pod ? POD_SEGS : NODE_SEGS;
  LineWriter w;
  w.d = B + (dst >> 2);
  w.fill = dst & 3u;
  w.pend = 0;
  for (uint32_t k = 0; k < nseg; ++k) {
    const uint32_t e = st[k], type = e >> 6;
    if (type == 0u) {
#if defined(__CUDA_ARCH__)
      const uint32_t pw = __ldg(sw + (e & 63u));
#else
      const uint32_t pw = sw[e & 63u];
#endif
      lw_put_piece(w, reinterpret_cast<const uint4 *>(slot + rw_off(pw)), rw_len(pw));
    } else {
      const uint32_t pw = B[(((type == 1u ? ca + CA2::KTAB : rec)) >> 2) + (e & 63u)];
      const uint32_t len = rw_len(pw), so = rw_off(pw);
      if (len) lw_put(w, B + (so >> 2), so & 3u, len);
    }
  }
  return w;
}
// ---- the unrolled form --------------------------------------------------------------------------
// The same line, with what is known at compile time taken out of the loop: the segment sequence
// is unrolled, the 'K' strings are immediates, a cell of at most 8 bytes (nearly all of them) is
// appended without a branch, the record sits in registers, and the slot is read as ONE stream of
// 16-byte units two units ahead of its use (the pieces lie behind each other in the slot).
KGM_HD uint32_t fsl(uint32_t lo, uint32_t hi, uint32_t s) {            // ((hi:lo) << s) >> 32, s in {0,8,16,24}
#if defined(__CUDA_ARCH__)
  return __funnelshift_l(lo, hi, s);
#else
  return s ? (lo >> (32u - s)) | (hi << s) : hi;
#endif
}
// n <= 8 bytes: v0 holds the first four, v1 the rest, bytes from n on are zero
KGM_HD void lw_put8(LineWriter &w, uint32_t v0, uint32_t v1, uint32_t n) {
  const uint32_t sh = 8u * w.fill;
  const uint32_t x0 = w.pend | (v0 << sh), x1 = fsl(v0, v1, sh), x2 = fsl(v1, 0u, sh);
  const uint32_t total = w.fill + n, nw = total >> 2;
  if (nw >= 1u) w.d[0] = x0;
  if (nw >= 2u) w.d[1] = x1;
  w.pend = nw == 0u ? x0 : nw == 1u ? x1 : x2;
  w.d += nw;
  w.fill = total & 3u;
}
// exactly 8 bytes
KGM_HD void lw_put_w2(LineWriter &w, uint32_t v0, uint32_t v1) {
  const uint32_t sh = 8u * w.fill;
  w.d[0] = w.pend | (v0 << sh);
  w.d[1] = fsl(v0, v1, sh);
  w.pend = fsl(v1, 0u, sh);
  w.d += 2;
}
// word j of an n-byte string, zero padded (compile time)
KGM_HD constexpr uint32_t kword(const char *s, uint32_t n, uint32_t j) {
  uint32_t v = 0;
  for (uint32_t i = 0; i < 4; ++i)
    if (4 * j + i < n) v |= (uint32_t)(uint8_t)s[4 * j + i] << (8 * i);
  return v;
}
template <uint32_t N>
KGM_HD void lw_put_const(LineWriter &w, uint32_t k0, uint32_t k1, uint32_t k2, uint32_t k3) {
  static_assert(N >= 1 && N <= 16, "a 'K' string");
  if (N <= 8) {
    lw_put8(w, k0, k1, N);
  } else {
    lw_put_w2(w, k0, k1);
    lw_put8(w, k2, k3, N - 8u);
  }
}
#define KGM_PUT_K(w, s)                                                                              \
  do {                                                                                               \
    constexpr uint32_t n_ = sizeof(s) - 1u, a_ = kword(s, n_, 0), b_ = kword(s, n_, 1),             \
                       c_ = kword(s, n_, 2), d_ = kword(s, n_, 3);                                  \
    lw_put_const<n_>(w, a_, b_, c_, d_);                                                             \
  } while (0)

// the cell a run word names (absolute offset in B); reads up to 11 bytes past the cell's first word
KGM_HD void lw_put_cell(LineWriter &w, const uint32_t *B, uint32_t pw) {
  const uint32_t len = rw_len(pw), so = rw_off(pw);
  if (len == 0u) return;
  if (len <= 8u) {
    const uint32_t *p = B + (so >> 2);
    const uint32_t s = 8u * (so & 3u);
    const uint32_t a = p[0], b = p[1], c = p[2];
    const uint32_t v0 = funnel_r(a, b, s) & low_mask(len < 4u ? len : 4u);
    const uint32_t v1 = funnel_r(b, c, s) & low_mask(len > 4u ? len - 4u : 0u);
    lw_put8(w, v0, v1, len);
  } else {
    lw_put(w, B + (so >> 2), so & 3u, len);
  }
}
// a cell known to be at most 4 bytes (the short id: "None" or up to three digits)
KGM_HD void lw_put_cell4(LineWriter &w, const uint32_t *B, uint32_t pw) {
  const uint32_t len = rw_len(pw), so = rw_off(pw);
  lw_put8(w, load4(B, so) & low_mask(len), 0u, len);
}

#ifndef KGM_SS_DEPTH
#define KGM_SS_DEPTH 2
#endif
constexpr int SS_DEPTH = KGM_SS_DEPTH;               // units fetched ahead of their use
struct SlotStream {                                  // the 16-byte units of a slot behind its header
  const uint4 *g;                                    // next unit to fetch
  uint4 q[SS_DEPTH];                                 // the next units to use
};
KGM_HD uint4 ldg16(const uint4 *p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}
// (reads SS_DEPTH units past the last piece: the pool ends with SLOT_SLACK spare bytes)
constexpr uint32_t SLOT_SLACK = 16u * SS_DEPTH;
KGM_HD void ss_open(SlotStream &s, const uint8_t *slot) {
  s.g = reinterpret_cast<const uint4 *>(slot + SLOT_HEADER);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < SS_DEPTH; ++i) s.q[i] = ldg16(s.g + i);
  s.g += SS_DEPTH;
}
KGM_HD uint4 ss_next(SlotStream &s) {
  const uint4 r = s.q[0];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i + 1 < SS_DEPTH; ++i) s.q[i] = s.q[i + 1];
  s.q[SS_DEPTH - 1] = ldg16(s.g++);
  return r;
}
// the next piece of the stream, n bytes
KGM_HD void lw_put_stream(LineWriter &w, SlotStream &s, uint32_t n) {
  const uint32_t sh = 8u * w.fill;
  uint32_t prev = shl8(w.pend, 4u - w.fill);
  uint32_t *d = w.d;
  uint32_t left = n;
  auto word = [&](uint32_t cur) {
    *d++ = fsl(prev, cur, sh);
    prev = cur;
  };
  for (; left >= 16u; left -= 16u) {
    const uint4 v = ss_next(s);
    word(v.x); word(v.y); word(v.z); word(v.w);
  }
  uint32_t last = 0;
  if (left) {
    const uint4 v = ss_next(s);
    const uint32_t fullw = left >> 2;
    if (fullw > 0) word(v.x);
    if (fullw > 1) word(v.y);
    if (fullw > 2) word(v.z);
    last = fullw == 0 ? v.x : fullw == 1 ? v.y : fullw == 2 ? v.z : v.w;
  }
  w.d = d;
  w.pend = shr8(prev, 4u - w.fill);
  const uint32_t r = left & 3u;
  if (r) lw_tail(w, last & low_mask(r), r);
}

template <bool POD>
KGM_HD LineWriter compose_line_t(uint32_t *B, const uint8_t *slot, uint32_t rec, uint32_t dst) {
  const uint4 *hp = reinterpret_cast<const uint4 *>(slot);
  const uint4 h0 = ldg16(hp);                        // run words of pieces 0..3 (the lengths are used)
  uint32_t h4 = 0;
  if (POD) {
#if defined(__CUDA_ARCH__)
    h4 = __ldg(reinterpret_cast<const uint32_t *>(slot) + 4);
#else
    h4 = reinterpret_cast<const uint32_t *>(slot)[4];
#endif
  }
  SlotStream s;
  ss_open(s, slot);
  uint32_t r[8];
#if defined(__CUDA_ARCH__)
  {
    const uint4 a = *reinterpret_cast<const uint4 *>(B + (rec >> 2)), b = *reinterpret_cast<const uint4 *>(B + (rec >> 2) + 4);
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w; r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
  }
#else
  for (int i = 0; i < 8; ++i) r[i] = B[(rec >> 2) + i];
#endif
  LineWriter w;
  w.d = B + (dst >> 2);
  w.fill = dst & 3u;
  w.pend = 0;
  if (!POD) {                                        // seg_node(0..14)
    lw_put_stream(w, s, rw_len(h0.x));
    lw_put_cell4(w, B, r[GREC_SID]);
    KGM_PUT_K(w, " gpu_util=");
    lw_put_cell(w, B, r[GREC_UTIL]);
    KGM_PUT_K(w, ",mem_util=");
    { const uint32_t pw = r[GREC_FSTR]; lw_put(w, B + (rw_off(pw) >> 2), rw_off(pw) & 3u, rw_len(pw)); }   // (never empty)
    KGM_PUT_K(w, ",mem_used=");
    lw_put_cell(w, B, r[GREC_MU]);
    KGM_PUT_K(w, ",mem_total=");
    lw_put_cell(w, B, r[GREC_MT]);
    lw_put_stream(w, s, rw_len(h0.y));
    lw_put_cell4(w, B, r[GREC_SID]);
    KGM_PUT_K(w, ",mem_free=");
    lw_put_cell(w, B, r[GREC_MF]);
  } else {                                           // seg_pod(0..19)
    lw_put_stream(w, s, rw_len(h0.x));
    lw_put_cell4(w, B, r[PREC_SID]);
    lw_put_stream(w, s, rw_len(h0.y));
    lw_put_cell(w, B, r[PREC_PID]);
    KGM_PUT_K(w, ",gpu_util_t=");
    lw_put_cell(w, B, r[PREC_UTIL]);
    lw_put_stream(w, s, rw_len(h0.z));
    lw_put_cell(w, B, r[PREC_NAME + 0]);
    lw_put_cell(w, B, r[PREC_NAME + 1]);
    lw_put_cell(w, B, r[PREC_NAME + 2]);
    lw_put_cell(w, B, r[PREC_NAME + 3]);
    lw_put_stream(w, s, rw_len(h0.w));
    lw_put_cell(w, B, r[PREC_PID]);
    KGM_PUT_K(w, "\",gpu_id_v=\"");
    lw_put_cell4(w, B, r[PREC_SID]);
    lw_put_stream(w, s, rw_len(h4));
    lw_put_cell(w, B, r[PREC_UTIL]);
    KGM_PUT_K(w, ",used_memory=");
    lw_put_cell(w, B, r[PREC_MEM]);
  }
  lw_put8(w, 10u, 0u, 1u);                           // '\n'
  return w;
}
// ---- the loop form ------------------------------------------------------------------------------
// One loop over the segment table with the same primitives as the unrolled form (slot stream,
// branch-free short cells): a body of a few hundred instructions instead of the unrolled form's
// 2000.  Built because 40 % of the composing warps' stall samples are instruction fetch
// (profiles/README.md, round 2) -- and measured 7 - 9 % SLOWER than the unrolled form on every
// bench configuration (it executes more instructions per line), so the kernels use the unrolled
// form; -DKGM_COMPOSE_LOOP switches, the CPU tests compare all three forms byte for byte.
// a cell of at most 16 bytes without a branch on its length; longer ones through lw_put
KGM_HD void lw_put_cell16(LineWriter &w, const uint32_t *B, uint32_t pw) {
  const uint32_t len = rw_len(pw), so = rw_off(pw);
  if (len > 16u) { lw_put(w, B + (so >> 2), so & 3u, len); return; }
  const uint32_t *p = B + (so >> 2);
  const uint32_t s = 8u * (so & 3u);
  const uint32_t a = p[0], b = p[1], c = p[2];
  if (len <= 8u) {                                   // (one branch, taken alike by nearly all lanes of a segment)
    const uint32_t v0 = funnel_r(a, b, s) & low_mask(len < 4u ? len : 4u);
    const uint32_t v1 = funnel_r(b, c, s) & low_mask(len > 4u ? len - 4u : 0u);
    lw_put8(w, v0, v1, len);
  } else {
    const uint32_t d = p[3], e = p[4];
    lw_put_w2(w, funnel_r(a, b, s), funnel_r(b, c, s));
    const uint32_t r = len - 8u;
    const uint32_t v0 = funnel_r(c, d, s) & low_mask(r < 4u ? r : 4u);
    const uint32_t v1 = funnel_r(d, e, s) & low_mask(r > 4u ? r - 4u : 0u);
    lw_put8(w, v0, v1, r);
  }
}
KGM_HD LineWriter compose_line_loop(uint32_t *B, uint32_t ca, uint32_t pod, const uint8_t *slot, uint32_t rec,
                                    uint32_t dst) {
  const uint8_t *Bb = reinterpret_cast<const uint8_t *>(B);
  const uint8_t *st = Bb + ca + CA2::SEGTAB + (pod ? 32u : 0u);
  const uint32_t *sw = reinterpret_cast<const uint32_t *>(slot);
  const uint32_t *kt = B + ((ca + CA2::KTAB) >> 2), *rw_ = B + (rec >> 2);
  const uint32_t nseg = pod ? POD_SEGS : NODE_SEGS;
  SlotStream s;
  ss_open(s, slot);
  LineWriter w;
  w.d = B + (dst >> 2);
  w.fill = dst & 3u;
  w.pend = 0;
  for (uint32_t k = 0; k < nseg; ++k) {
    const uint32_t e = st[k], type = e >> 6, idx = e & 63u;
    if (type == 0u) {
#if defined(__CUDA_ARCH__)
      const uint32_t pw = __ldg(sw + idx);
#else
      const uint32_t pw = sw[idx];
#endif
      lw_put_stream(w, s, rw_len(pw));
    } else {
      lw_put_cell16(w, B, type == 1u ? kt[idx] : rw_[idx]);
    }
  }
  return w;
}
// Writes one line at byte offset `dst` of the word array B: `ca` = byte offset of the constant area
// in B, `slot` = the line's template slot (global memory), `rec` = byte offset of its record in B
// (16-byte aligned).  Returns the writer: its pending bytes still have to be stored (lw_finish).
KGM_HD LineWriter compose_line(uint32_t *B, uint32_t ca, uint32_t pod, const uint8_t *slot, uint32_t rec,
                               uint32_t dst) {
#if defined(KGM_COMPOSE_LOOP)
  return compose_line_loop(B, ca, pod, slot, rec, dst);
#else
  (void)ca;
  return pod ? compose_line_t<true>(B, slot, rec, dst) : compose_line_t<false>(B, slot, rec, dst);
#endif
}
// The last bytes of a line, after every line of the round has been written: OR-ed into the word
// the next line has started (`has_next`), or stored as they are.
KGM_HD void lw_finish(const LineWriter &w, bool has_next) {
  if (w.fill) *w.d = has_next ? (*w.d | w.pend) : w.pend;
}

// ---- parsing: the row parsers of kgm_fast.cuh, returning runs relative to the text window -------
struct G2 {                                         // one --query-gpu row (gpu.py:104-109)
  Run key, util, mt, mu, mf;
  uint64_t hash;                                    // of lstrip(cell 0)
  uint64_t total, used;                             // int(digits) of memory.total / memory.used
};

// int(digits) of a run of at most 18 digits, two words at a time
KGM_HD uint64_t run_value2(const uint32_t *W, Run r) {
  uint64_t v = 0;
  uint32_t o = r.off, n = r.len;
  while (n >= 4) {
    const uint32_t w = load4(W, o) - 0x30303030u;                      // 4 digits, first one in the low byte
    const uint32_t t = (w * 10u + (w >> 8)) & 0x00ff00ffu;              // pairs: d0d1 in bits 0..7, d2d3 in 16..23
    v = v * 10000u + ((t & 0xffu) * 100u + (t >> 16));
    o += 4; n -= 4;
  }
  for (; n; --n, ++o) v = v * 10u + (load1(W, o) - 48u);
  return v;
}

// Same accept / decline rules as fast::parse_gpu_row (kgm_fast.cuh); false = irregular row.
KGM_HD bool parse_gpu_row2(const uint32_t *W, uint32_t s, uint32_t e, G2 &g) {
  if (e <= s || e - s > 1000) return false;
  RowShape sh;
  scan_commas<6>(W, s, e, sh);
  if (sh.irregular || sh.ncommas < 5) return false;               // fewer than 6 cells: KeyError
  const uint32_t c0 = sh.comma[0], c1 = sh.comma[1], c2 = sh.comma[2], c3 = sh.comma[3], c4 = sh.comma[4];
  const uint32_t c5 = sh.ncommas >= 6 ? sh.comma[5] : e;          // zip() drops extra cells (gpu.py:49)
  uint32_t k = s;
  while (k < c0 && load1(W, k) == 32u) ++k;                       // lstrip (gpu.py:108)
  if (c0 - k > 4u * KEY_WORDS) return false;
  g.key.off = (uint16_t)k; g.key.len = (uint16_t)(c0 - k);
  {
    uint32_t kw[KEY_WORDS];
    load_key(W, k, c0 - k, kw);
    g.hash = key_hash_words(kw, c0 - k);
  }
  if (!digit_run(W, c0 + 1, c1, g.util)) return false;
  if (!digit_run(W, c2 + 1, c3, g.mt)) return false;
  if (!digit_run(W, c3 + 1, c4, g.mu)) return false;
  if (!digit_run(W, c4 + 1, c5, g.mf)) return false;
  if (g.mt.len == 0 || g.mu.len == 0) return false;               // ValueError gpu.py:133-134
  if (g.mt.len > 18 || g.mu.len > 18) return false;
  g.total = run_value2(W, g.mt);
  g.used = run_value2(W, g.mu);
  return g.total != 0;                                            // ZeroDivisionError gpu.py:135
}

// ---- stage 3 values (window reductions; no counterpart in the reference) -----------------------
// The integers behind a line's fields, with the staged front end's rules (kgm_window.cu
// k_win_samples over kgm_common.cuh GpuRowScan / ProcRowScan): digits stripped, saturating.
// int(digits) of a run of at most 9 digits at byte offset `off` of W
KGM_HD uint32_t run_u32(const uint32_t *W, uint32_t off, uint32_t len) {
  uint32_t v = 0;
  for (uint32_t i = 0; i < len; ++i) v = v * 10u + (load1(W, off + i) - 48u);
  return v;
}
// The 7th cell of a --query-gpu row (power.draw of an extended query; zip() drops it, gpu.py:49):
// 0xffffffff = no such cell, no digit in it, or a value of 2^32 - 1 and more (the staged front
// end's 32-bit saturation value is its "no value" too), else int(its digits).  False: more than 10
// digits (the staged front end's significant-digit rule decides those).
KGM_HD bool gpu_row_power(const uint32_t *W, uint32_t s, uint32_t e, uint32_t *out) {
  *out = NONE32;
  RowShape sh;
  scan_commas<6>(W, s, e, sh);
  if (sh.irregular || sh.ncommas < 6) return !sh.irregular;
  uint64_t val = 0;
  uint32_t ndig = 0;
  for (uint32_t p = sh.comma[5] + 1u; p < e; ++p) {
    const uint32_t c = load1(W, p);
    if (c == 44u) break;
    if (c - 48u <= 9u) {
      if (++ndig > 10u) return false;
      val = val * 10u + (c - 48u);
    }
  }
  if (ndig && val < 0xffffffffull) *out = (uint32_t)val;
  return true;
}

struct P2 {                                         // one --query-compute-apps row (gpu.py:111-116)
  uint64_t pidkey;                                  // pid cell bytes (<= 7) | len << 56; ~0 if longer
  Run pid, name[NAME_RUNS], mem, key;
  uint64_t khash;                                   // of lstrip(cell 2)
};

KGM_HD bool parse_proc_row2(const uint32_t *W, uint32_t s, uint32_t e, P2 &p) {
  PRec r;
  uint32_t kw[KEY_WORDS], klen = 0;
  uint64_t kh = 0;
  uint32_t koff = 0;
  if (!parse_proc_row(W, s, e, r, kw, &klen, &kh, &koff)) return false;
  p.pidkey = r.pidkey;
  p.pid = r.pid;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < NAME_RUNS; ++i) p.name[i] = r.name[i];
  p.mem = r.mem;
  p.key.off = (uint16_t)koff; p.key.len = (uint16_t)klen;
  p.khash = kh;
  return true;
}

// The record of a gpu row: run words with absolute offsets.  `tg` = byte offset of the text window
// the runs of `g` refer to, `rec` = byte offset of the record itself (its bytes 40..63 receive
// repr(mem_util), gpu.py:135,139), `sid` = the row's short id.  Returns the variable bytes of
// its node_gpu line.
KGM_HD void store_grec(uint32_t *B, uint32_t tg, uint32_t rec, const G2 &g, uint32_t sid) {
  uint32_t *r = B + (rec >> 2);
  const uint32_t fs = rec + GREC_FSTR_BYTES;
  const uint32_t flen = (uint32_t)float_repr(mem_util(g.used, g.total), reinterpret_cast<uint8_t *>(B) + fs);
  r[GREC_UTIL] = rw(tg + g.util.off, g.util.len);
  r[GREC_MU] = rw(tg + g.mu.off, g.mu.len);
  r[GREC_MT] = rw(tg + g.mt.off, g.mt.len);
  r[GREC_MF] = rw(tg + g.mf.off, g.mf.len);
  r[GREC_FSTR] = rw(fs, flen);
  r[GREC_SID] = sid_run(0u, sid);
  r[GREC_KEY] = rw(tg + g.key.off, g.key.len);
  r[GREC_LINE_LEN] = 0;
  r[GREC_HASH] = (uint32_t)g.hash;
  r[GREC_HASH + 1] = (uint32_t)(g.hash >> 32);
}
// The record of a process row; `grec` = the record of its gpu (utilization and short id are the
// gpu's, gpu.py:202,205).
KGM_HD void store_prec(uint32_t *B, uint32_t ta, uint32_t rec, const P2 &p, const uint32_t *grec) {
  uint32_t *q = B + (rec >> 2);
  q[PREC_PID] = rw(ta + p.pid.off, p.pid.len);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < NAME_RUNS; ++i) q[PREC_NAME + i] = rw(ta + p.name[i].off, p.name[i].len);
  q[PREC_MEM] = rw(ta + p.mem.off, p.mem.len);
  q[PREC_UTIL] = grec[GREC_UTIL];
  q[PREC_SID] = grec[GREC_SID];
}
// The pseudo-row of an idle line (gpu.py:225-239): constants, and the short id or None.
KGM_HD void store_idle_rec(uint32_t *B, uint32_t ca, uint32_t rec, uint32_t sid) {
  uint32_t *q = B + (rec >> 2);
  const uint32_t *proto = B + ((ca + CA2::IDLE_REC) >> 2);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < (int)PREC_WORDS; ++i) q[i] = proto[i];
  q[PREC_SID] = sid_run(ca, sid);
}

// bytes [a, a+n) of word array A == bytes [b, b+n) of word array B  (n <= 4 * KEY_WORDS)
KGM_HD bool keys_equal(const uint32_t *A, uint32_t a, const uint32_t *Bw, uint32_t b, uint32_t n) {
  bool eq = true;
  for (uint32_t i = 0; i < n; i += 4) {
    uint32_t x = load4(A, a + i), y = load4(Bw, b + i);
    if (i + 4 > n) { const uint32_t m = low_mask(n - i); x &= m; y &= m; }
    eq &= x == y;
  }
  return eq;
}

}  // namespace p2
}  // namespace kgm
