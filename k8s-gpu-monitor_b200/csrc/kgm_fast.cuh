// Leaf functions of the fused poll kernel's fast path (k_poll, kgm_fused.cuh).
//
// The fast path handles *regular* documents entirely out of shared memory: texts made of
// printable ASCII without quotes or carriage returns, every echoed numeric cell holding one
// contiguous digit run, unique keys, no raise point of the reference reached.  Everything
// else ("irregular") is detected here, exactly or conservatively, and re-done from scratch
// by the generic per-byte code of kgm_common.cuh inside the same kernel -- so whichever
// path a document takes, the bytes are the reference's (gpu.py:104-256).
//
// Like kgm_common.cuh this header is per-thread code compiled twice: by nvcc for the
// kernel (the word pointers then point into shared memory) and by g++ for the CPU tests in
// tests/native/hostcheck.cpp.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "kgm_common.cuh"

namespace kgm {
namespace fast {

// ---- unaligned access to a little-endian word array ------------------------------------------
KGM_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t s) {       // ((hi:lo) >> s) & 0xffffffff, s in {0,8,16,24}
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, s);
#else
  return s ? (lo >> s) | (hi << (32 - s)) : lo;
#endif
}
KGM_HD uint32_t low_mask(uint32_t nbytes) {                            // nbytes in 0..4
#if defined(__CUDA_ARCH__)
  return __funnelshift_rc(0xffffffffu, 0u, 32u - 8u * nbytes);        // clamps the shift at 32
#else
  return nbytes >= 4 ? 0xffffffffu : ((1u << (8 * nbytes)) - 1u);
#endif
}
// 4 bytes starting at byte offset `o` of the word array W (reads W[o/4] and W[o/4+1])
KGM_HD uint32_t load4(const uint32_t *W, uint32_t o) {
  const uint32_t *p = W + (o >> 2);
  return funnel_r(p[0], p[1], 8u * (o & 3u));
}
KGM_HD uint32_t load1(const uint32_t *W, uint32_t o) { return (W[o >> 2] >> (8u * (o & 3u))) & 0xffu; }

// ---- SWAR byte classes (valid for words whose bytes are all < 0x80) --------------------------
// 0x80 in every byte that is < n  (n <= 0x80)
KGM_HD uint32_t lt_mask(uint32_t w, uint32_t n_rep) { return ~((w | 0x80808080u) - n_rep) & 0x80808080u; }
// 0x80 in every byte that is an ASCII digit
KGM_HD uint32_t digit_mask(uint32_t w) { return lt_mask(w ^ 0x30303030u, 0x0a0a0a0au); }
// 0x80 in every byte equal to the byte replicated in pat
KGM_HD uint32_t eq_mask(uint32_t w, uint32_t pat) { return lt_mask(w ^ pat, 0x01010101u); }
// 0x80-per-byte mask -> 4-bit mask
KGM_HD uint32_t pack4(uint32_t m) { return ((m >> 7) * 0x00204081u >> 21) & 0xfu; }

// ---- key hash: word-wise over the zero-padded key, length mixed in ---------------------------
// (the fleet's visible-uuid hashes are computed with the same function on the host)
constexpr int KEY_WORDS = 12;                       // keys up to 48 bytes are handled by the fast path
KGM_HD uint64_t key_hash_words(const uint32_t *k, uint32_t len) {
  uint32_t a = 0x9e3779b9u ^ len, b = 0x85ebca6bu + len;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < KEY_WORDS; ++i) {
    if (4u * i < len) {
      a = (a ^ k[i]) * 0x9e3779b1u;
      b = ((b + k[i]) * 0x85ebca77u) ^ (a >> 15);
    }
  }
  return hash_mix(((uint64_t)a << 32) | b);
}
// key bytes [o, o+len) of W -> k[KEY_WORDS], zero padded (len <= 4*KEY_WORDS)
KGM_HD void load_key(const uint32_t *W, uint32_t o, uint32_t len, uint32_t *k) {
  const uint32_t *p = W + (o >> 2);
  const uint32_t s = 8u * (o & 3u);
  uint32_t prev = p[0];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < KEY_WORDS; ++i) {
    const uint32_t b = 4u * i;
    uint32_t w = 0;
    if (b < len) {
      const uint32_t cur = p[i + 1];
      w = funnel_r(prev, cur, s);
      prev = cur;
      if (b + 4 > len) w &= low_mask(len - b);
    }
    k[i] = w;
  }
}
// host-side helper for plain byte strings
KGM_HD uint64_t key_hash_bytes(const uint8_t *p, uint32_t len) {
  uint32_t k[KEY_WORDS];
  for (int i = 0; i < KEY_WORDS; ++i) {
    uint32_t w = 0;
    for (int b = 0; b < 4; ++b) {
      const uint32_t idx = 4u * i + b;
      if (idx < len) w |= (uint32_t)p[idx] << (8 * b);
    }
    k[i] = w;
  }
  return key_hash_words(k, len);
}

// ---- row records (shared memory) -----------------------------------------------------------
struct Run { uint16_t off; uint16_t len; };         // a byte run inside the warp's text window

struct alignas(8) GRec {                            // one --query-gpu row (gpu.py:104-109)
  uint64_t hash;                                    // of lstrip(cell 0) = gpu_uuid
  Run key;                                          // lstrip(cell 0)
  Run util, mt, mu, mf;                             // digit runs of cells 1, 3, 4, 5
  uint8_t fstr[24];                                 // repr(mem_util)  gpu.py:135,139
  uint8_t flen;
  uint8_t pad[3];
  uint32_t line_len;                                // node_gpu line incl. '\n'
  uint32_t keyw[KEY_WORDS];                         // the key, word aligned and zero padded
};

constexpr int NAME_RUNS = 4;
struct alignas(8) PRec {                            // one --query-compute-apps row (gpu.py:111-116)
  uint64_t pidkey;                                  // pid cell bytes (<= 7) | len << 56; ~0 if longer
  Run pid;                                          // cell 0, verbatim
  Run name[NAME_RUNS];                              // cell 1 split at spaces (gpu.py:203 removes them)
  Run mem;                                          // digit run of cell 3
  uint16_t sid;                                     // short id of its gpu
  uint16_t crank;                                   // rank of the gated container that lists the pid, else 0xffff
  uint16_t pos;                                     // place among that container's busy lines
  uint16_t pad;
  uint32_t digest;                                  // of (pid, sid): duplicate-key screening
};

// ---- row scanning ----------------------------------------------------------------------------
// Walks the bytes of a row once, word-wise, and reports the positions of ',' and ' ';
// anything else that could change the reference's behaviour makes the row irregular:
// control characters (incl. '\r' and tab), '"', bytes >= 0x7f.
struct RowShape {
  uint32_t ncommas;
  uint16_t comma[6];                                // positions of the first 6 commas
  bool irregular;
};

// cells of a row [s, e) (at most 1000 bytes): the first WANT commas are recorded (more are
// counted, not stored).  Lanes meet commas in different words, so the per-comma work inside
// the word loop is kept to a shift-or into a packed register (10 bits per position).
template <int WANT>
KGM_HD void scan_commas(const uint32_t *W, uint32_t s, uint32_t e, RowShape &r) {
  r.ncommas = 0;
  r.irregular = false;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < 6; ++i) r.comma[i] = 0;
  if (s >= e) return;
  uint32_t bad = 0, n = 0;
  uint64_t packed = 0;
  auto word = [&](uint32_t wi, uint32_t w) {
    // any byte >= 0x80, < 0x20 or == '"' makes the row irregular ("any" tests, not positions)
    const uint32_t q = w ^ 0x22222222u;
    bad |= w | ((w - 0x20202020u) & ~w) | ((q - 0x01010101u) & ~q);
    const uint32_t c = (w & 0x7f7f7f7fu) ^ 0x2c2c2c2cu;              // exact ',' positions
    uint32_t cm = ~((c + 0x7f7f7f7fu) | w) & 0x80808080u;
    while (cm) {
#if defined(__CUDA_ARCH__)
      const uint32_t bit = __ffs(cm) - 1;
#else
      const uint32_t bit = __builtin_ctz(cm);
#endif
      cm &= cm - 1;
      if (n < (uint32_t)WANT) packed |= (uint64_t)((wi << 2) + (bit >> 3) - s) << (10u * n);
      ++n;
    }
  };
  // bytes outside [s, e) in the first and last word are replaced by 'a'
  const uint32_t a = s >> 2, z = (e - 1) >> 2;
  uint32_t keep_a = ~low_mask(s - (a << 2));
  const uint32_t keep_z = low_mask(e - (z << 2));
  if (a == z) keep_a &= keep_z;
  word(a, (W[a] & keep_a) | (0x61616161u & ~keep_a));
  for (uint32_t wi = a + 1; wi < z; ++wi) word(wi, W[wi]);
  if (z > a) word(z, (W[z] & keep_z) | (0x61616161u & ~keep_z));
  bad &= 0x80808080u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < WANT; ++i) r.comma[i] = (uint16_t)(s + ((uint32_t)(packed >> (10 * i)) & 0x3ffu));
  r.ncommas = n;
  r.irregular = bad != 0;
}

// The digits of cell [lo, hi) as one run.  Returns false if the cell is longer than 20
// bytes or its digits are not contiguous (then the generic path echoes them byte by byte).
// A cell without digits gives an empty run (gpu.py:138 echoes '' for "[N/A]").
KGM_HD bool digit_run(const uint32_t *W, uint32_t lo, uint32_t hi, Run &out) {
  out.off = (uint16_t)lo; out.len = 0;
  if (hi <= lo) return true;
  if (hi - lo > 20) return false;
  uint32_t mask = 0;                                 // bit i = byte lo+i is a digit
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < 5; ++i) {
    const uint32_t b = lo + 4u * i;
    if (b < hi) {
      uint32_t w = load4(W, b);
      if (b + 4 > hi) w = (w & low_mask(hi - b)) | (0x61616161u & ~low_mask(hi - b));
      mask |= pack4(digit_mask(w & 0x7f7f7f7fu)) << (4 * i);
    }
  }
  if (!mask) return true;
#if defined(__CUDA_ARCH__)
  const uint32_t first = __ffs(mask) - 1, cnt = __popc(mask);
#else
  const uint32_t first = __builtin_ctz(mask), cnt = __builtin_popcount(mask);
#endif
  if ((mask >> first) != ((1u << cnt) - 1u)) return false;        // not one contiguous run
  out.off = (uint16_t)(lo + first);
  out.len = (uint16_t)cnt;
  return true;
}

KGM_HD uint64_t run_value(const uint32_t *W, Run r) {               // int(digits), len <= 18
  uint64_t v = 0;
  for (uint32_t i = 0; i < r.len; ++i) v = v * 10 + (load1(W, r.off + i) - 48u);
  return v;
}

// One --query-gpu row [s, e) of the window W.  Returns false = irregular (generic path).
// Covers gpu.py:47-55 (cells), :108 (lstrip of the key), :133-143 (digit strip, int(),
// mem_util) for rows on which the reference raises nothing.
KGM_HD bool parse_gpu_row(const uint32_t *W, uint32_t s, uint32_t e, uint32_t node_fixed,
                          uint32_t sid, GRec &g) {
  if (e <= s || e - s > 1000) return false;
  RowShape sh;
  scan_commas<6>(W, s, e, sh);
  if (sh.irregular || sh.ncommas < 5) return false;               // fewer than 6 cells: KeyError
  const uint32_t c0 = sh.comma[0], c1 = sh.comma[1], c2 = sh.comma[2], c3 = sh.comma[3], c4 = sh.comma[4];
  const uint32_t c5 = sh.ncommas >= 6 ? sh.comma[5] : e;          // zip() drops extra cells (gpu.py:49)
  // key = lstrip(cell 0): only ' ' can be leading whitespace here (control bytes are irregular)
  uint32_t k = s;
  while (k < c0 && load1(W, k) == 32u) ++k;
  if (c0 - k > 4u * KEY_WORDS) return false;
  g.key.off = (uint16_t)k; g.key.len = (uint16_t)(c0 - k);
  load_key(W, k, c0 - k, g.keyw);
  g.hash = key_hash_words(g.keyw, c0 - k);
  if (!digit_run(W, c0 + 1, c1, g.util)) return false;
  if (!digit_run(W, c2 + 1, c3, g.mt)) return false;
  if (!digit_run(W, c3 + 1, c4, g.mu)) return false;
  if (!digit_run(W, c4 + 1, c5, g.mf)) return false;
  if (g.mt.len == 0 || g.mu.len == 0) return false;               // ValueError gpu.py:133-134
  if (g.mt.len > 18 || g.mu.len > 18) return false;
  if (run_value(W, g.mt) == 0) return false;                      // ZeroDivisionError gpu.py:135
  g.flen = 0;
  g.line_len = node_fixed + 2 * dec_len(sid);                     // completed by finish_gpu_row
  return true;
}

// Second half of a gpu row: mem_util = used * 100.0 / total and its repr (gpu.py:135,139), the
// length of the node_gpu line.  Separate so that the kernel can run it beside other work.
KGM_HD bool finish_gpu_row(const uint32_t *W, GRec &g) {
  const uint64_t total = run_value(W, g.mt), used = run_value(W, g.mu);
  const int flen = float_repr(mem_util(used, total), g.fstr);
  g.flen = (uint8_t)flen;
  g.line_len += g.util.len + (uint32_t)flen + g.mu.len + g.mt.len + g.mf.len + 1;
  return g.line_len <= MAX_LINE_BYTES;
}

// One --query-compute-apps row.  `kw` receives the lstripped gpu_uuid cell (for the
// dictionary lookup), `klen` its length.  Returns false = irregular.
KGM_HD bool parse_proc_row(const uint32_t *W, uint32_t s, uint32_t e, PRec &p, uint32_t *kw,
                           uint32_t *klen, uint64_t *khash, uint32_t *koff = nullptr) {
  if (e <= s || e - s > 1000) return false;
  RowShape sh;
  scan_commas<4>(W, s, e, sh);
  if (sh.irregular || sh.ncommas < 3) return false;               // KeyError gpu.py:53 / 204
  const uint32_t c0 = sh.comma[0], c1 = sh.comma[1], c2 = sh.comma[2];
  const uint32_t c3 = sh.ncommas >= 4 ? sh.comma[3] : e;
  // pid cell, verbatim (gpu.py:178 compares the strings)
  const uint32_t plen = c0 - s;
  p.pid.off = (uint16_t)s; p.pid.len = (uint16_t)plen;
  if (plen <= 7) {
    uint64_t lo = plen ? load4(W, s) : 0u;
    if (plen > 4) lo |= (uint64_t)load4(W, s + 4) << 32;
    lo &= plen >= 8 ? ~0ull : ((1ull << (8 * plen)) - 1ull);
    p.pidkey = lo | ((uint64_t)plen << 56);
  } else {
    p.pidkey = ~0ull;
  }
  // process name: runs between spaces (gpu.py:203 .replace(' ', '')).  The positions of up
  // to NAME_RUNS-1 spaces are packed into one register inside the word loop (lanes meet
  // their spaces in different words); the runs are cut afterwards.
  {
    uint32_t lo = c0 + 1;
    const uint32_t hi = c1;
    if (lo < hi && load1(W, lo) == 32u) ++lo;                     // the cell's usual leading space
    uint32_t sp = 0, ns = 0;
    for (uint32_t wi = lo >> 2; lo < hi && wi <= (hi - 1) >> 2; ++wi) {
      uint32_t w = W[wi];
      const uint32_t base = wi << 2;
      uint32_t keep = 0xffffffffu;
      if (base < lo) keep &= ~low_mask(lo - base);
      if (base + 4 > hi) keep &= low_mask(hi - base);
      w = (w & keep) | (0x61616161u & ~keep);
      uint32_t sm = eq_mask(w & 0x7f7f7f7fu, 0x20202020u);
      while (sm) {
#if defined(__CUDA_ARCH__)
        const uint32_t bit = __ffs(sm) - 1;
#else
        const uint32_t bit = __builtin_ctz(sm);
#endif
        sm &= sm - 1;
        if (ns < (uint32_t)NAME_RUNS - 1u) sp |= (base + (bit >> 3) - lo) << (10u * ns);
        ++ns;
      }
    }
    if (ns > (uint32_t)NAME_RUNS - 1u) return false;
    uint32_t start = lo;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < NAME_RUNS; ++i) {
      const uint32_t end = (uint32_t)i < ns ? lo + ((sp >> (10 * i)) & 0x3ffu) : hi;
      const bool live = (uint32_t)i <= ns && lo <= hi;
      p.name[i].off = (uint16_t)(live ? start : lo);
      p.name[i].len = (uint16_t)(live && end > start ? end - start : 0u);
      start = end + 1;
    }
  }
  // gpu_uuid = lstrip(cell 2) (gpu.py:115)
  uint32_t k = c1 + 1;
  while (k < c2 && load1(W, k) == 32u) ++k;
  if (c2 - k > 4u * KEY_WORDS) return false;
  *klen = c2 - k;
  if (koff) *koff = k;
  load_key(W, k, c2 - k, kw);
  *khash = key_hash_words(kw, c2 - k);
  if (!digit_run(W, c2 + 1, c3, p.mem)) return false;
  p.sid = 0xffff; p.crank = 0xffff; p.digest = 0;
  return true;
}

KGM_HD uint32_t name_len(const PRec &p) {
  uint32_t n = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < NAME_RUNS; ++i) n += p.name[i].len;
  return n;
}

// ---- line writer ---------------------------------------------------------------------------
// Copies the byte run [so, so+n) of the word array B to byte offset `o` of the same array:
// byte stores up to the next word boundary of the destination, aligned 32-bit stores
// (source realigned with a funnel shift) for the body, byte stores for the tail.  Lines of
// neighbouring lanes meet inside a word, so only byte stores ever touch a shared word.
// Reads up to 2 words past the run.
KGM_HD void copy_run(uint32_t *B, uint32_t o, uint32_t so, uint32_t n) {
  uint8_t *Bb = reinterpret_cast<uint8_t *>(B);
  if (n <= 3u) {                                     // short ids, utilization digits, "-", "0"
    Bb[o] = Bb[so];
    if (n > 1) Bb[o + 1] = Bb[so + 1];
    if (n > 2) Bb[o + 2] = Bb[so + 2];
    return;
  }
  uint32_t h = (4u - (o & 3u)) & 3u;
  if (h > n) h = n;
  if (h > 0) Bb[o] = Bb[so];
  if (h > 1) Bb[o + 1] = Bb[so + 1];
  if (h > 2) Bb[o + 2] = Bb[so + 2];
  o += h; so += h; n -= h;
  const uint32_t full = n >> 2, t = n & 3u;
  const uint32_t *p = B + (so >> 2);
  const uint32_t s = 8u * (so & 3u);
  uint32_t *d = B + (o >> 2);
  uint32_t prev = p[0];
  for (uint32_t j = 0; j < full; ++j) {
    const uint32_t cur = p[j + 1];
    d[j] = funnel_r(prev, cur, s);
    prev = cur;
  }
  if (t) {
    const uint32_t w = funnel_r(prev, p[full + 1], s);
    uint8_t *tail = reinterpret_cast<uint8_t *>(d + full);
    tail[0] = (uint8_t)w;
    if (t > 1) tail[1] = (uint8_t)(w >> 8);
    if (t > 2) tail[2] = (uint8_t)(w >> 16);
  }
}

struct Seg { uint32_t off; uint32_t len; };          // byte offset into the warp's word array

constexpr int MAX_SEGS = 19;

// Writes the line made of segs[0..MAX_SEGS) + '\n' at byte offset `o` of the word array
// `B` (sources and destination live in the same array: the CTA's shared memory).
// Returns the number of bytes written.
KGM_HD uint32_t write_line(uint32_t *B, uint32_t o, const Seg *segs) {
  uint32_t at = o;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < MAX_SEGS; ++i) {
    if (segs[i].len) copy_run(B, at, segs[i].off, segs[i].len);
    at += segs[i].len;
  }
  reinterpret_cast<uint8_t *>(B)[at++] = 10u;
  return at - o;
}

// ---- segment lists of the three line kinds ----------------------------------------------------
// Node template block ("TB", built by kgm_load_fleet, copied to shared memory per document):
//   TbHeader | TbCont[nc] | node pieces | per gated container: 5 pieces (unpadded, back to back)
struct TbHeader {
  uint32_t nc;
  uint16_t n0_off, n0_len;      // "node_gpu,host=H,gpu_id="
  uint16_t n5_off, n5_len;      // ",node=\"H\",gid="
  uint16_t fixed;               // bytes of a node_gpu line that do not depend on the row
  uint16_t pad;
};
struct TbCont {                 // one container of the node, docker stats order (gpu.py:172-174)
  uint32_t blk_off;             // first piece, relative to the block start
  uint16_t plen[5];             // P0 P1 P3 P4 P6 (see kgm_skeleton.h)
  uint16_t fixed;               // constant bytes of a pod_gpu line for this container
  uint8_t gated;                // gpu.py:185-188
  uint8_t pad[3];
  uint32_t vis_begin, nvis;     // visible uuids (global index into FleetDev::vis_*)
};

// Constant strings every CTA keeps in shared memory (offsets into the const area)
struct ConstArea {
  // " gpu_util=" ",mem_util=" ",mem_used=" ",mem_total=" ",mem_free=" ",gpu_util_t=" "\",gpu_id_v=\""
  // ",used_memory=" "-" "0" "None" and the decimal strings 0..MAX_FAST_SID
  static constexpr uint32_t GPU_UTIL = 0, MEM_UTIL = 12, MEM_USED = 24, MEM_TOTAL = 36, MEM_FREE = 48,
                            GPU_UTIL_T = 60, GPU_ID_V = 76, USED_MEMORY = 92, DASH = 108, ZERO = 112,
                            NONE = 116, DEC = 124, BYTES = 124 + 4 * 256 + 16;
};
constexpr uint32_t MAX_FAST_SID = 255;

KGM_HD void fill_const_strings(uint8_t *c) {        // the area must be zeroed first
  auto put = [&](uint32_t off, const char *s) { for (int i = 0; s[i]; ++i) c[off + i] = (uint8_t)s[i]; };
  put(ConstArea::GPU_UTIL, " gpu_util=");
  put(ConstArea::MEM_UTIL, ",mem_util=");
  put(ConstArea::MEM_USED, ",mem_used=");
  put(ConstArea::MEM_TOTAL, ",mem_total=");
  put(ConstArea::MEM_FREE, ",mem_free=");
  put(ConstArea::GPU_UTIL_T, ",gpu_util_t=");
  put(ConstArea::GPU_ID_V, "\",gpu_id_v=\"");
  put(ConstArea::USED_MEMORY, ",used_memory=");
  put(ConstArea::DASH, "-");
  put(ConstArea::ZERO, "0");
  put(ConstArea::NONE, "None");
}
KGM_HD void fill_const_decimal(uint8_t *c, uint32_t v) {
  uint8_t *d = c + ConstArea::DEC + 4 * v;
  if (v >= 100) { d[0] = (uint8_t)('0' + v / 100); d[1] = (uint8_t)('0' + v / 10 % 10); d[2] = (uint8_t)('0' + v % 10); }
  else if (v >= 10) { d[0] = (uint8_t)('0' + v / 10); d[1] = (uint8_t)('0' + v % 10); }
  else d[0] = (uint8_t)('0' + v);
}
KGM_HD void fill_const_area(uint8_t *c) {
  for (uint32_t i = 0; i < ConstArea::BYTES; ++i) c[i] = 0;
  fill_const_strings(c);
  for (uint32_t v = 0; v <= MAX_FAST_SID; ++v) fill_const_decimal(c, v);
}

KGM_HD Seg sid_seg(uint32_t cbase, uint32_t sid) {                 // '%s' % shortunitids.get(uuid)
  if (sid == NONE32) return Seg{cbase + ConstArea::NONE, 4u};
  return Seg{cbase + ConstArea::DEC + 4u * sid, dec_len(sid)};
}

// All offsets are bytes into the warp's word array: cbase = const area, tg / ta = the two
// text windows, tb = the node template block, grec_fstr = &grec.fstr.
KGM_HD void node_segs(Seg *s, uint32_t cbase, uint32_t tb, const TbHeader &h, uint32_t tg, const GRec &g,
                      uint32_t fstr_off, uint32_t sid) {
  int i = 0;
  s[i++] = Seg{tb + h.n0_off, h.n0_len};                           // node_gpu,host=H,gpu_id=
  s[i++] = sid_seg(cbase, sid);
  s[i++] = Seg{cbase + ConstArea::GPU_UTIL, 10u};                  //  gpu_util=
  s[i++] = Seg{tg + g.util.off, g.util.len};
  s[i++] = Seg{cbase + ConstArea::MEM_UTIL, 10u};                  // ,mem_util=
  s[i++] = Seg{fstr_off, g.flen};
  s[i++] = Seg{cbase + ConstArea::MEM_USED, 10u};                  // ,mem_used=
  s[i++] = Seg{tg + g.mu.off, g.mu.len};
  s[i++] = Seg{cbase + ConstArea::MEM_TOTAL, 11u};                 // ,mem_total=
  s[i++] = Seg{tg + g.mt.off, g.mt.len};
  s[i++] = Seg{tb + h.n5_off, h.n5_len};                           // ,node="H",gid=
  s[i++] = sid_seg(cbase, sid);
  s[i++] = Seg{cbase + ConstArea::MEM_FREE, 10u};                  // ,mem_free=
  s[i++] = Seg{tg + g.mf.off, g.mf.len};
  for (; i < MAX_SEGS; ++i) s[i] = Seg{cbase, 0u};
}

KGM_HD void piece_offsets(const TbCont &c, uint32_t tb, uint32_t *po) {
  uint32_t o = tb + c.blk_off;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < 5; ++i) { po[i] = o; o += c.plen[i]; }
}

KGM_HD void busy_segs(Seg *s, uint32_t cbase, uint32_t tb, const TbCont &c, uint32_t tg, uint32_t ta,
                      const PRec &p, const GRec &g, uint32_t sid) {
  uint32_t po[5];
  piece_offsets(c, tb, po);
  int i = 0;
  s[i++] = Seg{po[0], c.plen[0]};                                  // pod_gpu,host=H,namespace=NS,gpu_id=
  s[i++] = sid_seg(cbase, sid);
  s[i++] = Seg{po[1], c.plen[1]};                                  // ,containerid=C,pid_t=
  s[i++] = Seg{ta + p.pid.off, p.pid.len};
  s[i++] = Seg{cbase + ConstArea::GPU_UTIL_T, 12u};                // ,gpu_util_t=
  s[i++] = Seg{tg + g.util.off, g.util.len};
  s[i++] = Seg{po[2], c.plen[2]};                                  // ,user_name_t=UN procressname="
  for (int r = 0; r < NAME_RUNS; ++r) s[i++] = Seg{ta + p.name[r].off, p.name[r].len};
  s[i++] = Seg{po[3], c.plen[3]};                                  // ",podname="POD",pid="
  s[i++] = Seg{ta + p.pid.off, p.pid.len};
  s[i++] = Seg{cbase + ConstArea::GPU_ID_V, 12u};                  // ",gpu_id_v="
  s[i++] = sid_seg(cbase, sid);
  s[i++] = Seg{po[4], c.plen[4]};                                  // ",node="H",user_name="UN",gpu_util=
  s[i++] = Seg{tg + g.util.off, g.util.len};
  s[i++] = Seg{cbase + ConstArea::USED_MEMORY, 13u};               // ,used_memory=
  s[i++] = Seg{ta + p.mem.off, p.mem.len};
}

KGM_HD void idle_segs(Seg *s, uint32_t cbase, uint32_t tb, const TbCont &c, uint32_t sid) {
  uint32_t po[5];
  piece_offsets(c, tb, po);
  const Seg dash{cbase + ConstArea::DASH, 1u}, zero{cbase + ConstArea::ZERO, 1u}, none{cbase, 0u};
  int i = 0;
  s[i++] = Seg{po[0], c.plen[0]};
  s[i++] = sid_seg(cbase, sid);
  s[i++] = Seg{po[1], c.plen[1]};
  s[i++] = dash;                                                   // pid_t=-
  s[i++] = Seg{cbase + ConstArea::GPU_UTIL_T, 12u};
  s[i++] = zero;                                                   // gpu_util_t=0
  s[i++] = Seg{po[2], c.plen[2]};
  s[i++] = dash;                                                   // procressname="-"
  for (int r = 1; r < NAME_RUNS; ++r) s[i++] = none;
  s[i++] = Seg{po[3], c.plen[3]};
  s[i++] = dash;                                                   // pid="-"
  s[i++] = Seg{cbase + ConstArea::GPU_ID_V, 12u};
  s[i++] = sid_seg(cbase, sid);
  s[i++] = Seg{po[4], c.plen[4]};
  s[i++] = zero;                                                   // gpu_util=0
  s[i++] = Seg{cbase + ConstArea::USED_MEMORY, 13u};
  s[i++] = zero;                                                   // used_memory=0
}

KGM_HD uint32_t busy_len(const TbCont &c, const PRec &p, const GRec &g, uint32_t sid) {
  return c.fixed + 2u * dec_len(sid) + 2u * p.pid.len + 2u * g.util.len + name_len(p) + p.mem.len + 1u;
}
KGM_HD uint32_t idle_len(const TbCont &c, uint32_t sid) { return c.fixed + 2u * sid_len(sid) + 7u; }

// ---- compact pid -> container table (fast path; pids of up to 7 bytes) -------------------------
struct alignas(16) PidSlot16 {
  uint64_t key;       // pid bytes | len << 56
  uint32_t node;      // NONE32 = empty
  uint32_t crank;
};
KGM_HD uint32_t pid16_hash(uint64_t key, uint32_t node) {
  return (uint32_t)hash_mix(key ^ (0x9e3779b97f4a7c15ull * (node + 1)));
}

KGM_HD uint32_t pair_digest(uint64_t pidkey, uint32_t pid_len, uint32_t sid) {
  return (uint32_t)hash_mix(pidkey ^ ((uint64_t)pid_len << 40) ^ ((uint64_t)(sid + 1) * 0x9e3779b97f4a7c15ull));
}

}  // namespace fast
}  // namespace kgm
