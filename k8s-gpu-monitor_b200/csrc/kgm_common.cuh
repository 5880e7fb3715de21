// Shared record layouts and per-thread leaf functions of the poll pipeline.
//
// Everything in here is plain per-thread code (no warp collectives), written once and
// compiled twice: by nvcc for the sm_100a kernels and by g++ for the CPU unit tests in
// tests/native/hostcheck.cpp, which check the byte-level semantics (CSV cell splitting,
// digit stripping, pid packing, line composition) against Python without a GPU.
#pragma once
#include <stdint.h>

#include "kgm_d2s.cuh"

namespace kgm {

// ---- per-document status: must match KGM_DOC_* in include/kgm.h ------------------------
enum DocStatus : uint32_t {
  DOC_OK = 0, DOC_KEY = 1, DOC_VALUE = 2, DOC_ZERODIV = 3, DOC_ATTR = 4, DOC_INDEX = 5,
  DOC_CSV = 6, DOC_JSON = 7, DOC_TYPE = 8, DOC_OTHER = 9,
  DOC_UNSUP_QUOTE = 20, DOC_UNSUP_BYTE = 21, DOC_UNSUP_RANGE = 22,
};

constexpr uint32_t NONE32 = 0xffffffffu;
constexpr uint32_t MAX_LINE_BYTES = 8192;     // longer lines -> DOC_UNSUP_RANGE
constexpr uint32_t FSTR_STRIDE = 32;          // repr(mem_util): 24 bytes + length at [24]

// str.lstrip() / str.split() whitespace, ASCII part (gpu.py:108,115,224)
KGM_HD bool is_str_space(uint32_t c) {
  return c == 32u || (c - 9u) <= 4u || (c - 28u) <= 3u;
}
KGM_HD bool is_digit(uint32_t c) { return (c - 48u) <= 9u; }

KGM_HD uint64_t hash_init() { return 0xcbf29ce484222325ull; }
KGM_HD uint64_t hash_step(uint64_t h, uint32_t c) { return (h ^ c) * 0x100000001b3ull; }
KGM_HD uint64_t hash_mix(uint64_t x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
  x ^= x >> 27; x *= 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

KGM_HD uint32_t dec_len(uint32_t v) {
  uint32_t n = 1;
  while (v >= 10) { v /= 10; ++n; }
  return n;
}

// ---- row flags --------------------------------------------------------------------------
enum : uint8_t {
  RF_CR = 1,        // a '\r' that is not part of the line terminator (csv.Error)
  RF_QUOTE = 2,     // a cell starts with '"' (csv would parse a quoted field)
  RF_HIBYTE = 4,    // byte >= 0x80
  RF_LONG = 8,      // row longer than 65535 bytes
  RF_LONGPID = 16,  // pid cell longer than 15 bytes: cannot match any table entry
  RF_RANGE = 32,    // memory.total / memory.used with more than 19 digits
};

// One row of the --query-gpu text (gpu.py:104-109), tokenised.
struct GpuRow {
  uint32_t row_off;       // first byte of the row in gpu_csv
  uint32_t row_end;       // one past its last byte (the '\n', or end of text)
  uint16_t cell_end[6];   // end of cell i relative to row_off; cell i starts at cell_end[i-1]+1
  uint16_t key_lo;        // start of lstrip(cell 0) = gpu_uuid key (gpu.py:108)
  uint8_t  ncells;        // number of CSV cells (0 = blank line), saturating at 255
  uint8_t  flags;
  uint8_t  nd[4];         // digit counts of cells 1,3,4,5 = len(re.sub('\D','',cell))
  uint32_t util;          // int(digits(utilization.gpu)), saturating (stage 3 only)
  uint64_t key_hash;
  uint64_t mem_total;     // int(digits(memory.total))   gpu.py:133
  uint64_t mem_used;      // int(digits(memory.used))    gpu.py:134
  // written by stage 2
  uint32_t first;         // local index of the first row with the same key
  uint32_t last;          // (on first rows) local index of the last row with that key
  uint32_t sid;           // (on first rows) short id = rank in first-seen order (gpu.py:119)
  uint32_t power;         // int(digits(cell 6)) of an extended query (power.draw), saturating;
                          // NONE32 if the row has no 7th cell or it holds no digit (stage 3 only)
};

// One row of the --query-compute-apps text (gpu.py:111-116), tokenised.
struct ProcRow {
  uint32_t row_off;
  uint32_t row_end;
  uint16_t cell_end[4];
  uint16_t key_lo;        // start of lstrip(cell 2) = gpu_uuid (gpu.py:115)
  uint16_t name_len;      // len(process_name.replace(' ',''))   gpu.py:203
  uint8_t  ncells;
  uint8_t  flags;
  uint8_t  nd_mem;        // len(digits(used_memory))            gpu.py:204
  uint8_t  pid_len;       // len(pid cell), saturating at 255
  uint32_t used_mem;      // int(digits(used_memory)), saturating (stage 3 only)
  uint64_t uuid_hash;
  uint64_t pid_lo;        // pid cell bytes 0..7, little endian
  uint64_t pid_hi;        // bytes 8..14, length in the top byte
  // written by stage 2
  uint32_t sid;           // short id of its gpu (NONE32 = unknown uuid -> KeyError gpu.py:151)
  uint32_t first;         // local index of the first row with the same (pid, uuid)
  uint32_t last;          // (on first rows) local index of the last such row: its values win
  uint32_t crank;         // (on first rows) container rank of the first table match, or NONE32
  uint32_t nmatch;        // (on first rows) number of containers listing this pid
  uint32_t pad;
};

// pid -> container table slot (open addressing, linear probing; node == NONE32 is empty)
struct alignas(16) PidSlot {
  uint64_t lo, hi;
  uint32_t node;
  uint32_t crank;         // rank of the container inside its node (docker stats order)
  uint32_t pad[2];
};

struct Piece { uint32_t off; uint32_t len; };   // a constant chunk of a line, in the skeleton pool
constexpr int NODE_PIECES = 7;
constexpr int POD_PIECES = 8;

enum : uint32_t { LINE_NODE = 0, LINE_BUSY = 1, LINE_IDLE = 2 };

// One output line, as planned by stage 2 and written by stage 4.
struct alignas(16) LineDesc {
  uint32_t doc;
  uint32_t kind_ref;      // kind in the top 2 bits; node lines: 0, pod lines: global container index
  uint32_t a;             // node: global GpuRow index (values row); busy: global ProcRow index (values row);
                          // idle: short id or NONE32
  uint32_t b;             // node: short id; busy: global GpuRow index of its gpu's values row
};

KGM_HD uint64_t pid_hash(uint64_t lo, uint64_t hi, uint32_t node) {
  return hash_mix(lo ^ hash_mix(hi + 0x9e3779b97f4a7c15ull * (node + 1)));
}

// ---- byte iteration ---------------------------------------------------------------------
// Calls f.step(byte, pos) for pos in [lo, hi).  On the device the text is read in aligned
// 16-byte vectors (the text buffers are 16-byte aligned and padded), one vector per 16
// positions, so that a lane walking its own row costs one L1 wavefront per 16 bytes.
template <class F>
KGM_HD void for_each_byte(const uint8_t *text, uint32_t lo, uint32_t hi, F &f) {
#if defined(__CUDA_ARCH__)
  const uint4 *q = reinterpret_cast<const uint4 *>(text) + (lo >> 4);
  uint32_t pos = lo & ~15u;
  while (pos < hi) {
    const uint4 v = __ldg(q++);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int b = 0; b < 16; ++b, ++pos) {
      if (pos >= lo && pos < hi) f.step((w[b >> 2] >> (8 * (b & 3))) & 0xffu, pos);
    }
  }
#else
  for (uint32_t pos = lo; pos < hi; ++pos) f.step(text[pos], pos);
#endif
}

// ---- stage 1 leaf: CSV row scanners -------------------------------------------------------
// csv.reader's default dialect on these texts (gpu.py:47) reduces to: cells split on ',',
// the line's trailing "\r"s dropped, a '\r' anywhere else is an error, a cell that
// starts with '"' starts a quoted field (not supported here: flagged).
struct GpuRowScan {
  GpuRow r;
  uint32_t base;          // row_off
  uint32_t ci;            // current cell
  uint32_t len;           // row length without trailing CRs
  uint32_t pending_cr;
  uint32_t ndig;          // digits in the current cell
  uint32_t sig;           // of them, from the first non-zero one on
  uint64_t val;
  bool cell_first, lead, over;
  uint32_t quote;         // cell 0 only: 0 = not quoted, 1 = inside "...", 2 = closed, waiting for the ','

  KGM_HDM void begin(uint32_t row_off, uint32_t row_end) {
    r = GpuRow();
    r.row_off = row_off; r.row_end = row_end;
    r.key_hash = hash_init();
    r.power = NONE32;
    base = row_off; ci = 0; len = 0; pending_cr = 0; ndig = 0; sig = 0; val = 0;
    cell_first = true; lead = true; over = false; quote = 0;
  }
  KGM_HDM uint32_t sat32() const { return (over || sig > 10 || val > 0xffffffffull) ? 0xffffffffu : (uint32_t)val; }
  KGM_HDM void close_cell(uint32_t p) {
    if (p > 65535u) r.flags |= RF_LONG;
    if (ci < 6) r.cell_end[ci] = (uint16_t)p;
    if (ci == 0 && lead) r.key_lo = (uint16_t)p;
    const uint8_t nd8 = (uint8_t)(ndig > 255u ? 255u : ndig);
    if (ci == 1) { r.nd[0] = nd8; r.util = sat32(); }
    else if (ci == 3) { r.nd[1] = nd8; r.mem_total = val; if (over) r.flags |= RF_RANGE; }
    else if (ci == 4) { r.nd[2] = nd8; r.mem_used = val; if (over) r.flags |= RF_RANGE; }
    else if (ci == 5) { r.nd[3] = nd8; }
    else if (ci == 6) { r.power = ndig ? sat32() : NONE32; }
    ndig = 0; sig = 0; val = 0; over = false;
  }
  KGM_HDM void step(uint32_t c, uint32_t pos) {
    const uint32_t p = pos - base;
    if (c == '\r') { ++pending_cr; return; }
    if (pending_cr) { r.flags |= RF_CR; pending_cr = 0; }
    len = p + 1;
    if (c >= 0x80u) r.flags |= RF_HIBYTE;
    // csv.reader: a first cell "..." is the text between the quotes (gpu.py:47).  Taken here when
    // the quotes close right in front of the delimiter and hold no quote, comma or line end;
    // every other use of a leading quote stays flagged.
    if (quote == 1) {
      if (c == '"') { quote = 2; r.cell_end[0] = (uint16_t)p; if (lead) r.key_lo = (uint16_t)p; return; }
      if (c == ',') r.flags |= RF_QUOTE;             // a delimiter inside the quotes is content to csv
    } else if (quote == 2) {
      if (c != ',') r.flags |= RF_QUOTE;             // csv appends what follows the closing quote
    }
    if (c == ',' && quote != 1) {
      if (quote == 2) {                              // cell 0 ended at its closing quote
        if (p > 65535u) r.flags |= RF_LONG;
        quote = 0; ndig = 0; sig = 0; val = 0; over = false;
      } else {
        close_cell(p);
      }
      ++ci; cell_first = true; return;
    }
    if (cell_first) {
      cell_first = false;
      if (c == '"') {
        if (ci == 0) { quote = 1; return; }
        r.flags |= RF_QUOTE;
      }
    }
    if (ci == 0) {
      if (lead && is_str_space(c)) return;
      if (lead) { r.key_lo = (uint16_t)p; lead = false; }
      r.key_hash = hash_step(r.key_hash, c);
    } else if (is_digit(c)) {
      ++ndig;
      if (sig || c != 48u) ++sig;
      if (sig > 19) over = true; else val = val * 10 + (c - 48u);
    }
  }
  KGM_HDM void finish() {
    if (len == 0) { r.ncells = 0; return; }       // blank line: csv yields [] (KeyError gpu.py:51)
    if (quote == 1) r.flags |= RF_QUOTE;            // the quoted cell runs on into the next line
    if (quote == 2) { quote = 0; ndig = 0; sig = 0; val = 0; over = false; }   // row = one quoted cell
    else close_cell(len);
    r.ncells = (uint8_t)(ci + 1 > 255u ? 255u : ci + 1);
  }
};

struct ProcRowScan {
  ProcRow r;
  uint32_t base, ci, len, pending_cr, ndig, sig, name_len, pid_len;
  uint64_t val;
  bool cell_first, lead;

  KGM_HDM void begin(uint32_t row_off, uint32_t row_end) {
    r = ProcRow();
    r.row_off = row_off; r.row_end = row_end;
    r.uuid_hash = hash_init();
    r.sid = NONE32; r.crank = NONE32;
    base = row_off; ci = 0; len = 0; pending_cr = 0; ndig = 0; sig = 0; name_len = 0; pid_len = 0; val = 0;
    cell_first = true; lead = true;
  }
  KGM_HDM void close_cell(uint32_t p) {
    if (p > 65535u) r.flags |= RF_LONG;
    if (ci < 4) r.cell_end[ci] = (uint16_t)p;
    if (ci == 0) {
      if (pid_len > 15) r.flags |= RF_LONGPID;
      r.pid_len = (uint8_t)(pid_len > 255u ? 255u : pid_len);
      if (pid_len <= 15) r.pid_hi |= (uint64_t)pid_len << 56;
    } else if (ci == 1) {
      if (name_len > 65535u) r.flags |= RF_LONG;
      r.name_len = (uint16_t)name_len;
    } else if (ci == 2) {
      if (lead) r.key_lo = (uint16_t)p;
    } else if (ci == 3) {
      r.nd_mem = (uint8_t)(ndig > 255u ? 255u : ndig);
      r.used_mem = (sig > 10 || val > 0xffffffffull) ? 0xffffffffu : (uint32_t)val;
    }
    ndig = 0; sig = 0; val = 0;
  }
  KGM_HDM void step(uint32_t c, uint32_t pos) {
    const uint32_t p = pos - base;
    if (c == '\r') { ++pending_cr; return; }
    if (pending_cr) { r.flags |= RF_CR; pending_cr = 0; }
    len = p + 1;
    // process_name (cell 1) is only copied, minus its ' ' characters (gpu.py:203): its UTF-8
    // bytes pass through.  Elsewhere a non-ASCII byte could be whitespace or a digit to Python.
    if (c >= 0x80u && ci != 1) r.flags |= RF_HIBYTE;
    if (c == ',') { close_cell(p); ++ci; cell_first = true; return; }
    if (cell_first) { if (c == '"') r.flags |= RF_QUOTE; cell_first = false; }
    if (ci == 0) {
      if (pid_len < 8) r.pid_lo |= (uint64_t)c << (8 * pid_len);
      else if (pid_len < 15) r.pid_hi |= (uint64_t)c << (8 * (pid_len - 8));
      ++pid_len;
    } else if (ci == 1) {
      if (c != ' ') ++name_len;
    } else if (ci == 2) {
      if (lead && is_str_space(c)) return;
      if (lead) { r.key_lo = (uint16_t)p; lead = false; }
      r.uuid_hash = hash_step(r.uuid_hash, c);
    } else if (ci == 3 && is_digit(c)) {
      ++ndig;
      if (sig || c != 48u) ++sig;
      if (sig <= 10) val = val * 10 + (c - 48u);
    }
  }
  KGM_HDM void finish() {
    if (len == 0) { r.ncells = 0; return; }
    close_cell(len);
    r.ncells = (uint8_t)(ci + 1 > 255u ? 255u : ci + 1);
  }
};

// Parse-level verdict for a row (raised before any line of the document is emitted).
KGM_HD uint32_t parse_error_of(uint8_t flags, uint32_t ncells, uint32_t min_cells) {
  if (flags & RF_HIBYTE) return DOC_UNSUP_BYTE;
  if (flags & RF_LONG) return DOC_UNSUP_RANGE;
  if (flags & RF_QUOTE) return DOC_UNSUP_QUOTE;
  if (flags & RF_CR) return DOC_CSV;
  if (ncells < min_cells) return DOC_KEY;        // gpu.py:51 / 53
  return DOC_OK;
}

// Cell bounds helpers -----------------------------------------------------------------------
KGM_HD uint32_t gcell_lo(const GpuRow &r, int i) { return i ? (uint32_t)r.cell_end[i - 1] + 1u : 0u; }
KGM_HD uint32_t pcell_lo(const ProcRow &r, int i) { return i ? (uint32_t)r.cell_end[i - 1] + 1u : 0u; }

// Node-loop verdict for one unique gpu (its values row), gpu.py:133-143, in evaluation order.
KGM_HD uint32_t node_error_of(const GpuRow &v) {
  if (v.ncells < 4) return DOC_KEY;
  if (v.nd[1] == 0) return DOC_VALUE;
  if (v.ncells < 5) return DOC_KEY;
  if (v.nd[2] == 0) return DOC_VALUE;
  if (v.flags & RF_RANGE) return DOC_UNSUP_RANGE;
  if (v.mem_total == 0) return DOC_ZERODIV;
  if (v.ncells < 6) return DOC_KEY;
  return DOC_OK;
}

// ---- stage 4 leaf: byte sink + filtered copies ----------------------------------------------
enum CopyMode { COPY_ALL = 0, COPY_DIGITS = 1, COPY_NOSPACE = 2 };

struct ByteSink {
  uint8_t *p;
  KGM_HDM void put(uint32_t c) { *p++ = (uint8_t)c; }
};

template <int MODE>
struct CopyStep {
  ByteSink *s;
  KGM_HDM void step(uint32_t c, uint32_t) {
    if (MODE == COPY_DIGITS && !is_digit(c)) return;
    if (MODE == COPY_NOSPACE && c == ' ') return;
    s->put(c);
  }
};

template <int MODE>
KGM_HD void copy_span(ByteSink &s, const uint8_t *src, uint32_t lo, uint32_t hi) {
  CopyStep<MODE> st{&s};
  for_each_byte(src, lo, hi, st);
}

KGM_HD void put_dec(ByteSink &s, uint32_t v) {
  uint8_t tmp[10];
  int n = 0;
  do { tmp[n++] = (uint8_t)('0' + v % 10); v /= 10; } while (v);
  while (n) s.put(tmp[--n]);
}

KGM_HD void put_sid(ByteSink &s, uint32_t sid) {       // '%s' % shortunitids.get(uuid)  gpu.py:226
  if (sid == NONE32) { s.put('N'); s.put('o'); s.put('n'); s.put('e'); }
  else put_dec(s, sid);
}
KGM_HD uint32_t sid_len(uint32_t sid) { return sid == NONE32 ? 4u : dec_len(sid); }

// node_gpu line, gpu.py:136-143.  `pc` = the node's 7 constant chunks.
KGM_HD void compose_node_line(ByteSink &s, const uint8_t *skel, const Piece *pc,
                              const uint8_t *gpu_csv, const GpuRow &v, uint32_t sid,
                              const uint8_t *fstr) {
  const uint32_t o = v.row_off;
  copy_span<COPY_ALL>(s, skel, pc[0].off, pc[0].off + pc[0].len);      // node_gpu,host=H,gpu_id=
  put_dec(s, sid);
  copy_span<COPY_ALL>(s, skel, pc[1].off, pc[1].off + pc[1].len);      //  gpu_util=
  copy_span<COPY_DIGITS>(s, gpu_csv, o + gcell_lo(v, 1), o + v.cell_end[1]);
  copy_span<COPY_ALL>(s, skel, pc[2].off, pc[2].off + pc[2].len);      // ,mem_util=
  for (uint32_t i = 0; i < fstr[24]; ++i) s.put(fstr[i]);
  copy_span<COPY_ALL>(s, skel, pc[3].off, pc[3].off + pc[3].len);      // ,mem_used=
  copy_span<COPY_DIGITS>(s, gpu_csv, o + gcell_lo(v, 4), o + v.cell_end[4]);
  copy_span<COPY_ALL>(s, skel, pc[4].off, pc[4].off + pc[4].len);      // ,mem_total=
  copy_span<COPY_DIGITS>(s, gpu_csv, o + gcell_lo(v, 3), o + v.cell_end[3]);
  copy_span<COPY_ALL>(s, skel, pc[5].off, pc[5].off + pc[5].len);      // ,node="H",gid=
  put_dec(s, sid);
  copy_span<COPY_ALL>(s, skel, pc[6].off, pc[6].off + pc[6].len);      // ,mem_free=
  copy_span<COPY_DIGITS>(s, gpu_csv, o + gcell_lo(v, 5), o + v.cell_end[5]);
  s.put('\n');
}

KGM_HD uint32_t node_line_len(uint32_t fixed, const GpuRow &v, uint32_t sid, uint32_t flen) {
  return fixed + 2 * dec_len(sid) + v.nd[0] + flen + v.nd[2] + v.nd[1] + v.nd[3] + 1;
}

// pod_gpu line, busy branch, gpu.py:200-216.  `pc` = the container's 8 constant chunks.
KGM_HD void compose_busy_line(ByteSink &s, const uint8_t *skel, const Piece *pc,
                              const uint8_t *gpu_csv, const uint8_t *apps_csv,
                              const ProcRow &pr, const GpuRow &gv, uint32_t sid) {
  const uint32_t po = pr.row_off, go = gv.row_off;
  const uint32_t u_lo = go + gcell_lo(gv, 1), u_hi = go + gv.cell_end[1];
  copy_span<COPY_ALL>(s, skel, pc[0].off, pc[0].off + pc[0].len);      // pod_gpu,host=H,namespace=NS,gpu_id=
  put_dec(s, sid);
  copy_span<COPY_ALL>(s, skel, pc[1].off, pc[1].off + pc[1].len);      // ,containerid=C,pid_t=
  copy_span<COPY_ALL>(s, apps_csv, po, po + pr.cell_end[0]);
  copy_span<COPY_ALL>(s, skel, pc[2].off, pc[2].off + pc[2].len);      // ,gpu_util_t=
  copy_span<COPY_DIGITS>(s, gpu_csv, u_lo, u_hi);
  copy_span<COPY_ALL>(s, skel, pc[3].off, pc[3].off + pc[3].len);      // ,user_name_t=UN procressname="
  copy_span<COPY_NOSPACE>(s, apps_csv, po + pcell_lo(pr, 1), po + pr.cell_end[1]);
  copy_span<COPY_ALL>(s, skel, pc[4].off, pc[4].off + pc[4].len);      // ",podname="POD",pid="
  copy_span<COPY_ALL>(s, apps_csv, po, po + pr.cell_end[0]);
  copy_span<COPY_ALL>(s, skel, pc[5].off, pc[5].off + pc[5].len);      // ",gpu_id_v="
  put_dec(s, sid);
  copy_span<COPY_ALL>(s, skel, pc[6].off, pc[6].off + pc[6].len);      // ",node="H",user_name="UN",gpu_util=
  copy_span<COPY_DIGITS>(s, gpu_csv, u_lo, u_hi);
  copy_span<COPY_ALL>(s, skel, pc[7].off, pc[7].off + pc[7].len);      // ,used_memory=
  copy_span<COPY_DIGITS>(s, apps_csv, po + pcell_lo(pr, 3), po + pr.cell_end[3]);
  s.put('\n');
}

KGM_HD uint32_t busy_line_len(uint32_t fixed, const ProcRow &pr, const GpuRow &gv, uint32_t sid) {
  return fixed + 2 * dec_len(sid) + 2u * pr.cell_end[0] + 2u * gv.nd[0] + pr.name_len + pr.nd_mem + 1;
}

// pod_gpu line, idle branch, gpu.py:225-250.
KGM_HD void compose_idle_line(ByteSink &s, const uint8_t *skel, const Piece *pc, uint32_t sid) {
  copy_span<COPY_ALL>(s, skel, pc[0].off, pc[0].off + pc[0].len);
  put_sid(s, sid);
  copy_span<COPY_ALL>(s, skel, pc[1].off, pc[1].off + pc[1].len);
  s.put('-');
  copy_span<COPY_ALL>(s, skel, pc[2].off, pc[2].off + pc[2].len);
  s.put('0');
  copy_span<COPY_ALL>(s, skel, pc[3].off, pc[3].off + pc[3].len);
  s.put('-');
  copy_span<COPY_ALL>(s, skel, pc[4].off, pc[4].off + pc[4].len);
  s.put('-');
  copy_span<COPY_ALL>(s, skel, pc[5].off, pc[5].off + pc[5].len);
  put_sid(s, sid);
  copy_span<COPY_ALL>(s, skel, pc[6].off, pc[6].off + pc[6].len);
  s.put('0');
  copy_span<COPY_ALL>(s, skel, pc[7].off, pc[7].off + pc[7].len);
  s.put('0');
  s.put('\n');
}

KGM_HD uint32_t idle_line_len(uint32_t fixed, uint32_t sid) {
  return fixed + 2 * sid_len(sid) + 7;
}

}  // namespace kgm
