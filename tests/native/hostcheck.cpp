// TEST INFRASTRUCTURE: host build of the per-thread leaf functions that the CUDA kernels
// use (same headers, compiled by g++), so their byte-level semantics can be checked on a
// machine without a GPU.  Nothing in the product links or loads this.
#include <stdint.h>
#include <string.h>

#include <string>

#include "../../k8s-gpu-monitor_b200/csrc/kgm_common.cuh"
#include "../../k8s-gpu-monitor_b200/csrc/kgm_skeleton.h"
#include "../../k8s-gpu-monitor_b200/csrc/kgm_docker.cuh"

using namespace kgm;

extern "C" {

int hc_float_repr(double x, unsigned char *out) { return float_repr(x, out); }

int hc_mem_util_repr(uint64_t used, uint64_t total, unsigned char *out) {
  return float_repr(mem_util(used, total), out);
}

void hc_float_repr_many(const double *x, int64_t n, unsigned char *out /* n*24 */, int32_t *len) {
  for (int64_t i = 0; i < n; ++i) len[i] = float_repr(x[i], out + 24 * i);
}

// Tokenise one --query-gpu row.  out[]: ncells, flags, key_lo, cell_end[6], nd[4], util,
// parse_error; u64out[]: key_hash, mem_total, mem_used
void hc_scan_gpu_row(const uint8_t *text, uint32_t lo, uint32_t hi, uint32_t *out, uint64_t *u64out) {
  GpuRowScan sc;
  sc.begin(lo, hi);
  for_each_byte(text, lo, hi, sc);
  sc.finish();
  const GpuRow &r = sc.r;
  int k = 0;
  out[k++] = r.ncells; out[k++] = r.flags; out[k++] = r.key_lo;
  for (int i = 0; i < 6; ++i) out[k++] = r.cell_end[i];
  for (int i = 0; i < 4; ++i) out[k++] = r.nd[i];
  out[k++] = r.util;
  out[k++] = parse_error_of(r.flags, r.ncells, 1);
  out[k++] = node_error_of(r);
  out[k++] = r.power;
  u64out[0] = r.key_hash; u64out[1] = r.mem_total; u64out[2] = r.mem_used;
}

// out[]: ncells, flags, key_lo, cell_end[4], name_len, nd_mem, pid_len, used_mem, parse_error
void hc_scan_proc_row(const uint8_t *text, uint32_t lo, uint32_t hi, uint32_t *out, uint64_t *u64out) {
  ProcRowScan sc;
  sc.begin(lo, hi);
  for_each_byte(text, lo, hi, sc);
  sc.finish();
  const ProcRow &r = sc.r;
  int k = 0;
  out[k++] = r.ncells; out[k++] = r.flags; out[k++] = r.key_lo;
  for (int i = 0; i < 4; ++i) out[k++] = r.cell_end[i];
  out[k++] = r.name_len; out[k++] = r.nd_mem; out[k++] = r.pid_len; out[k++] = r.used_mem;
  out[k++] = parse_error_of(r.flags, r.ncells, 3);
  u64out[0] = r.uuid_hash; u64out[1] = r.pid_lo; u64out[2] = r.pid_hi;
}

// One node_gpu line from a gpu row text.  Returns the byte count written (with '\n');
// *planned receives the length stage 2 would have planned.
int hc_node_line(const char *host, const uint8_t *row, uint32_t row_len, uint32_t sid,
                 uint8_t *out, uint32_t *planned) {
  SkelPool pool;
  Piece pc[NODE_PIECES];
  const uint32_t fixed = pool.node_pieces(host, pc);
  GpuRowScan sc;
  sc.begin(0, row_len);
  for_each_byte(row, 0, row_len, sc);
  sc.finish();
  if (node_error_of(sc.r)) return -(int)node_error_of(sc.r);
  uint8_t fs[FSTR_STRIDE];
  const int flen = float_repr(mem_util(sc.r.mem_used, sc.r.mem_total), fs);
  fs[24] = (uint8_t)flen;
  *planned = node_line_len(fixed, sc.r, sid, (uint32_t)flen);
  ByteSink s{out};
  compose_node_line(s, pool.bytes.data(), pc, row, sc.r, sid, fs);
  return (int)(s.p - out);
}

// One busy pod_gpu line from a process row and its gpu's row.
int hc_busy_line(const char *host, const char *ns, const char *cid, const char *user, const char *pod,
                 const uint8_t *prow, uint32_t prow_len, const uint8_t *grow, uint32_t grow_len,
                 uint32_t sid, uint8_t *out, uint32_t *planned) {
  SkelPool pool;
  Piece pc[POD_PIECES];
  const uint32_t fixed = pool.pod_pieces(host, ns, cid, user, pod, pc);
  ProcRowScan ps;
  ps.begin(0, prow_len);
  for_each_byte(prow, 0, prow_len, ps);
  ps.finish();
  GpuRowScan gs;
  gs.begin(0, grow_len);
  for_each_byte(grow, 0, grow_len, gs);
  gs.finish();
  if (ps.r.ncells < 4) return -1;
  *planned = busy_line_len(fixed, ps.r, gs.r, sid);
  ByteSink s{out};
  compose_busy_line(s, pool.bytes.data(), pc, grow, prow, ps.r, gs.r, sid);
  return (int)(s.p - out);
}

int hc_idle_line(const char *host, const char *ns, const char *cid, const char *user, const char *pod,
                 uint32_t sid, uint8_t *out, uint32_t *planned) {
  SkelPool pool;
  Piece pc[POD_PIECES];
  const uint32_t fixed = pool.pod_pieces(host, ns, cid, user, pod, pc);
  *planned = idle_line_len(fixed, sid);
  ByteSink s{out};
  compose_idle_line(s, pool.bytes.data(), pc, sid);
  return (int)(s.p - out);
}

// ---- fast path of the fused kernel (kgm_fast.cuh), laid out like a warp's shared-memory slice:
//      const area | gpu text window | apps text window | template block | records | stage
struct FastSlice {
  std::vector<uint32_t> words;
  uint32_t tg, ta, tb, rec, stage;
  uint8_t *bytes() { return reinterpret_cast<uint8_t *>(words.data()); }
};

static FastSlice make_slice(const uint8_t *gtext, uint32_t glen, uint32_t gshift, const uint8_t *atext,
                            uint32_t alen, uint32_t ashift, const std::vector<uint8_t> &tb) {
  FastSlice sl;
  auto r16 = [](uint32_t v) { return (v + 15u) & ~15u; };
  const uint32_t cpad = r16(fast::ConstArea::BYTES);
  sl.tg = cpad;
  sl.ta = sl.tg + r16(gshift + glen) + 48;
  sl.tb = sl.ta + r16(ashift + alen) + 48;
  sl.rec = sl.tb + r16((uint32_t)tb.size()) + 32;
  sl.stage = sl.rec + 4096;
  sl.words.assign((sl.stage + 3 * MAX_LINE_BYTES) / 4, 0x5a5a5a5au);
  fast::fill_const_area(sl.bytes());
  memcpy(sl.bytes() + sl.tg + gshift, gtext, glen);
  memcpy(sl.bytes() + sl.ta + ashift, atext, alen);
  memcpy(sl.bytes() + sl.tb, tb.data(), tb.size());
  return sl;
}

static bool one_container_block(const char *host, const char *ns, const char *cid, const char *user,
                                const char *pod, std::vector<uint8_t> &tb, uint32_t *node_fixed) {
  SkelPool pool;
  Piece pc[POD_PIECES], npc[NODE_PIECES];
  *node_fixed = pool.node_pieces(host, npc);
  std::vector<TbContIn> conts(1);
  conts[0].gated = true;
  conts[0].cid = cid; conts[0].ns = ns; conts[0].pod = pod; conts[0].user = user;
  conts[0].vis_begin = 0; conts[0].nvis = 0;
  conts[0].fixed = pool.pod_pieces(host, ns, cid, user, pod, pc);
  size_t start = 0, len = 0;
  return build_template_block(host, *node_fixed, conts, tb, &start, &len);
}

// node_gpu line through the fast path.  Returns -1 if the row is irregular (generic path),
// else the bytes written at stage offset `align`; *planned = the length the fast path plans.
int hc_fast_node_line(const char *host, const uint8_t *row, uint32_t row_len, uint32_t sid, uint32_t shift,
                      uint32_t align, uint8_t *out, uint32_t *planned) {
  std::vector<uint8_t> tb;
  uint32_t node_fixed;
  if (!one_container_block(host, "ns", "cid", "user", "pod", tb, &node_fixed)) return -2;
  FastSlice sl = make_slice(row, row_len, shift, row, 0, 0, tb);
  fast::GRec g;
  if (!fast::parse_gpu_row(sl.words.data() + sl.tg / 4, shift, shift + row_len, node_fixed, sid, g)) return -1;
  if (!fast::finish_gpu_row(sl.words.data() + sl.tg / 4, g)) return -1;
  fast::GRec *gr = reinterpret_cast<fast::GRec *>(sl.bytes() + sl.rec);
  *gr = g;
  *planned = g.line_len;
  const fast::TbHeader &hd = *reinterpret_cast<const fast::TbHeader *>(sl.bytes() + sl.tb);
  fast::Seg segs[fast::MAX_SEGS];
  fast::node_segs(segs, 0u, sl.tb, hd, sl.tg, *gr, sl.rec + (uint32_t)offsetof(fast::GRec, fstr), sid);
  const uint32_t n = fast::write_line(sl.words.data(), sl.stage + 64 + align, segs);
  memcpy(out, sl.bytes() + sl.stage + 64 + align - 8, n + 16);      // 8 guard bytes either side
  return (int)n;
}

int hc_fast_busy_line(const char *host, const char *ns, const char *cid, const char *user, const char *pod,
                      const uint8_t *prow, uint32_t prow_len, const uint8_t *grow, uint32_t grow_len,
                      uint32_t sid, uint32_t shift, uint32_t align, uint8_t *out, uint32_t *planned) {
  std::vector<uint8_t> tb;
  uint32_t node_fixed;
  if (!one_container_block(host, ns, cid, user, pod, tb, &node_fixed)) return -2;
  FastSlice sl = make_slice(grow, grow_len, shift, prow, prow_len, (shift * 7 + 3) & 15, tb);
  const uint32_t ashift = (shift * 7 + 3) & 15;
  fast::GRec g;
  if (!fast::parse_gpu_row(sl.words.data() + sl.tg / 4, shift, shift + grow_len, node_fixed, sid, g)) return -1;
  if (!fast::finish_gpu_row(sl.words.data() + sl.tg / 4, g)) return -1;
  fast::PRec p;
  uint32_t kw[fast::KEY_WORDS], klen;
  uint64_t kh;
  if (!fast::parse_proc_row(sl.words.data() + sl.ta / 4, ashift, ashift + prow_len, p, kw, &klen, &kh)) return -1;
  const fast::TbCont *tc = reinterpret_cast<const fast::TbCont *>(sl.bytes() + sl.tb + sizeof(fast::TbHeader));
  p.sid = (uint16_t)sid; p.crank = 0;
  *planned = fast::busy_len(tc[0], p, g, sid);
  fast::Seg segs[fast::MAX_SEGS];
  fast::busy_segs(segs, 0u, sl.tb, tc[0], sl.tg, sl.ta, p, g, sid);
  const uint32_t n = fast::write_line(sl.words.data(), sl.stage + 64 + align, segs);
  memcpy(out, sl.bytes() + sl.stage + 64 + align - 8, n + 16);
  return (int)n;
}

int hc_fast_idle_line(const char *host, const char *ns, const char *cid, const char *user, const char *pod,
                      uint32_t sid, uint32_t align, uint8_t *out, uint32_t *planned) {
  std::vector<uint8_t> tb;
  uint32_t node_fixed;
  if (!one_container_block(host, ns, cid, user, pod, tb, &node_fixed)) return -2;
  const uint8_t dummy[4] = {0, 0, 0, 0};
  FastSlice sl = make_slice(dummy, 0, 0, dummy, 0, 0, tb);
  const fast::TbCont *tc = reinterpret_cast<const fast::TbCont *>(sl.bytes() + sl.tb + sizeof(fast::TbHeader));
  *planned = fast::idle_len(tc[0], sid);
  fast::Seg segs[fast::MAX_SEGS];
  fast::idle_segs(segs, 0u, sl.tb, tc[0], sid);
  const uint32_t n = fast::write_line(sl.words.data(), sl.stage + 64 + align, segs);
  memcpy(out, sl.bytes() + sl.stage + 64 + align - 8, n + 16);
  return (int)n;
}

// Fields the fast gpu-row parser extracts: out = key off/len, util off/len, mt, mu, mf off/len
int hc_fast_gpu_row(const uint8_t *row, uint32_t row_len, uint32_t shift, uint32_t *out, uint64_t *hash) {
  std::vector<uint8_t> tb(16, 0);
  FastSlice sl = make_slice(row, row_len, shift, row, 0, 0, tb);
  fast::GRec g;
  if (!fast::parse_gpu_row(sl.words.data() + sl.tg / 4, shift, shift + row_len, 100, 0, g)) return 0;
  const fast::Run runs[5] = {g.key, g.util, g.mt, g.mu, g.mf};
  for (int i = 0; i < 5; ++i) { out[2 * i] = runs[i].off - shift; out[2 * i + 1] = runs[i].len; }
  *hash = g.hash;
  return 1;
}

// out = pid off/len, 4 name runs off/len, mem off/len, key len; key bytes -> key[48]
int hc_fast_proc_row(const uint8_t *row, uint32_t row_len, uint32_t shift, uint32_t *out, uint8_t *key,
                     uint64_t *u64out) {
  std::vector<uint8_t> tb(16, 0);
  FastSlice sl = make_slice(row, 0, 0, row, row_len, shift, tb);
  fast::PRec p;
  uint32_t kw[fast::KEY_WORDS], klen;
  uint64_t kh;
  if (!fast::parse_proc_row(sl.words.data() + sl.ta / 4, shift, shift + row_len, p, kw, &klen, &kh)) return 0;
  int k = 0;
  out[k++] = p.pid.off - shift; out[k++] = p.pid.len;
  for (int i = 0; i < fast::NAME_RUNS; ++i) { out[k++] = p.name[i].off - shift; out[k++] = p.name[i].len; }
  out[k++] = p.mem.off - shift; out[k++] = p.mem.len;
  out[k++] = klen;
  memcpy(key, kw, 4 * fast::KEY_WORDS);
  u64out[0] = kh; u64out[1] = p.pidkey;
  return 1;
}

// ---- k_poll2 (kgm_poll2_leaf.cuh): a line written as 4 or 8 independent jobs, laid out like a
//      warp's slice: constant area | gpu text window | apps text window | slot | records | stage
struct P2Mem {
  std::vector<uint32_t> words;
  uint32_t tg, ta, slot, grec, prec, stage;
  uint8_t *bytes() { return reinterpret_cast<uint8_t *>(words.data()); }
};

static P2Mem p2_make(const uint8_t *gtext, uint32_t glen, uint32_t gshift, const uint8_t *atext, uint32_t alen,
                     uint32_t ashift, const std::vector<uint8_t> &pool, size_t slot_start, size_t slot_len) {
  P2Mem m;
  auto r16 = [](uint32_t v) { return (v + 15u) & ~15u; };
  m.tg = r16(p2::CA2::BYTES);
  m.ta = m.tg + r16(gshift + glen) + 48;
  m.slot = m.ta + r16(ashift + alen) + 48;
  m.grec = m.slot + r16((uint32_t)slot_len) + 16;
  m.prec = m.grec + p2::GREC_BYTES;
  m.stage = m.prec + p2::PREC_BYTES + 16;
  m.words.assign((m.stage + 3 * MAX_LINE_BYTES) / 4 + 4096, 0x5a5a5a5au);
  p2::fill_ca2(m.bytes(), 0u);
  memcpy(m.bytes() + m.tg + gshift, gtext, glen);
  memcpy(m.bytes() + m.ta + ashift, atext, alen);
  memcpy(m.bytes() + m.slot, pool.data() + slot_start, slot_len);
  return m;
}

// One line through the line writer, at byte offset `dst`; a second copy of it is written right
// behind (before the first one is finished), as the next lane of a round would, so that the
// shared word between two lines is exercised.  Returns the line's length.
static uint32_t p2_write(P2Mem &m, uint32_t pod, uint32_t rec, uint32_t dst, uint32_t var) {
  const uint32_t len = *reinterpret_cast<const uint16_t *>(m.bytes() + m.slot + p2::SLOT_FIXED) + var;
  // the slot is read as "global memory": give the writer its own aligned copy
  static std::vector<p2::uint4> slotcopy;
  slotcopy.assign(1024, p2::uint4{0, 0, 0, 0});
  memcpy(slotcopy.data(), m.bytes() + m.slot, 8192);
  const uint8_t *slot = reinterpret_cast<const uint8_t *>(slotcopy.data());
  p2::LineWriter w2 = p2::compose_line(m.words.data(), 0u, pod, slot, rec, dst + len);
  p2::LineWriter w1 = p2::compose_line(m.words.data(), 0u, pod, slot, rec, dst);
  p2::lw_finish(w1, true);
  p2::lw_finish(w2, false);
  // the table-driven form of the same line, further up in the stage at the same alignment
  const uint32_t dst2 = ((dst + 2 * len + 64) & ~3u) + 64 + (dst & 3u);
  p2::LineWriter w3 = p2::compose_line_table(m.words.data(), 0u, pod, slot, rec, dst2);
  p2::lw_finish(w3, false);
  if (w1.fill != w3.fill || memcmp(m.bytes() + dst, m.bytes() + dst2, len) != 0) return 0xffffffffu;
  // ... and the loop form, further up again
  const uint32_t dst3 = ((dst2 + len + 64) & ~3u) + 64 + (dst & 3u);
  p2::LineWriter w4 = p2::compose_line_loop(m.words.data(), 0u, pod, slot, rec, dst3);
  p2::lw_finish(w4, false);
  if (w1.fill != w4.fill || memcmp(m.bytes() + dst, m.bytes() + dst3, len) != 0) return 0xffffffffu;
  return len;
}

int hc_p2_node_line(const char *host, const uint8_t *row, uint32_t row_len, uint32_t sid, uint32_t shift,
                    uint32_t align, uint32_t jobs, uint8_t *out, uint32_t *planned) {
  std::vector<uint8_t> pool;
  size_t st = 0, ln = 0;
  if (!build_node_slot(host, pool, &st, &ln)) return -2;
  P2Mem m = p2_make(row, row_len, shift, row, 0, 0, pool, st, ln);
  p2::G2 g;
  if (!p2::parse_gpu_row2(m.words.data() + m.tg / 4, shift, shift + row_len, g)) return -1;
  p2::store_grec(m.words.data(), m.tg, m.grec, g, sid);
  const uint32_t var = p2::line_var_bytes<false>(m.words.data() + m.grec / 4);
  const uint32_t n = p2_write(m, 0u, m.grec, m.stage + 64 + align, var);
  if (n == 0xffffffffu) return -9;                 // unrolled and table-driven forms differ
  *planned = n;
  memcpy(out, m.bytes() + m.stage + 64 + align - 8, 2 * n + 16);  // 8 guard bytes in front, the second copy behind
  (void)jobs;
  return (int)n;
}

int hc_p2_busy_line(const char *host, const char *ns, const char *cid, const char *user, const char *pod,
                    const uint8_t *prow, uint32_t prow_len, const uint8_t *grow, uint32_t grow_len, uint32_t sid,
                    uint32_t shift, uint32_t align, uint32_t jobs, uint8_t *out, uint32_t *planned) {
  std::vector<uint8_t> pool;
  size_t st = 0, ln = 0;
  if (!build_pod_slot(host, ns, cid, user, pod, pool, &st, &ln)) return -2;
  const uint32_t ashift = (shift * 7 + 3) & 15;
  P2Mem m = p2_make(grow, grow_len, shift, prow, prow_len, ashift, pool, st, ln);
  p2::G2 g;
  if (!p2::parse_gpu_row2(m.words.data() + m.tg / 4, shift, shift + grow_len, g)) return -1;
  p2::store_grec(m.words.data(), m.tg, m.grec, g, sid);
  p2::P2 p;
  if (!p2::parse_proc_row2(m.words.data() + m.ta / 4, ashift, ashift + prow_len, p)) return -1;
  p2::store_prec(m.words.data(), m.ta, m.prec, p, m.words.data() + m.grec / 4);
  const uint32_t var = p2::line_var_bytes<true>(m.words.data() + m.prec / 4);
  const uint32_t n = p2_write(m, 1u, m.prec, m.stage + 64 + align, var);
  if (n == 0xffffffffu) return -9;                 // unrolled and table-driven forms differ
  *planned = n;
  memcpy(out, m.bytes() + m.stage + 64 + align - 8, 2 * n + 16);
  (void)jobs;
  return (int)n;
}

int hc_p2_idle_line(const char *host, const char *ns, const char *cid, const char *user, const char *pod,
                    uint32_t sid, uint32_t align, uint32_t jobs, uint8_t *out, uint32_t *planned) {
  std::vector<uint8_t> pool;
  size_t st = 0, ln = 0;
  if (!build_pod_slot(host, ns, cid, user, pod, pool, &st, &ln)) return -2;
  const uint8_t dummy[4] = {0, 0, 0, 0};
  P2Mem m = p2_make(dummy, 0, 0, dummy, 0, 0, pool, st, ln);
  p2::store_idle_rec(m.words.data(), 0u, m.prec, sid);
  const uint32_t var = p2::line_var_bytes<true>(m.words.data() + m.prec / 4);
  const uint32_t n = p2_write(m, 1u, m.prec, m.stage + 64 + align, var);
  if (n == 0xffffffffu) return -9;                 // unrolled and table-driven forms differ
  *planned = n;
  memcpy(out, m.bytes() + m.stage + 64 + align - 8, 2 * n + 16);
  (void)jobs;
  return (int)n;
}

// the fields parse_gpu_row2 extracts: out = key off/len, util, mt, mu, mf off/len; u64 = hash, total, used
int hc_p2_gpu_row(const uint8_t *row, uint32_t row_len, uint32_t shift, uint32_t *out, uint64_t *u64out) {
  std::vector<uint8_t> pool(64, 0);
  P2Mem m = p2_make(row, row_len, shift, row, 0, 0, pool, 0, 16);
  p2::G2 g;
  if (!p2::parse_gpu_row2(m.words.data() + m.tg / 4, shift, shift + row_len, g)) return 0;
  const fast::Run runs[5] = {g.key, g.util, g.mt, g.mu, g.mf};
  for (int i = 0; i < 5; ++i) { out[2 * i] = runs[i].off - shift; out[2 * i + 1] = runs[i].len; }
  u64out[0] = g.hash; u64out[1] = g.total; u64out[2] = g.used;
  return 1;
}

// kgm_docker.cuh: one container's `docker inspect` text -> {status, gated, has_pod, pod off/len, ns off/len, user off/len}
void hc_inspect_fields(const uint8_t *t, uint32_t n, uint32_t *out) {
  const dk::Fields f = dk::inspect_fields(t, n);
  out[0] = f.status; out[1] = f.gated; out[2] = f.has_pod; out[3] = f.pod_off; out[4] = f.pod_len;
  out[5] = f.ns_off; out[6] = f.ns_len; out[7] = f.user_off; out[8] = f.user_len;
}

// kgm_poll2_leaf.cuh gpu_row_power (stage 3's fused front end): 1 + the value in *out, or 0 = declined
int hc_p2_gpu_power(const uint8_t *row, uint32_t row_len, uint32_t shift, uint32_t *out) {
  std::vector<uint8_t> pool(64, 0);
  P2Mem m = p2_make(row, row_len, shift, row, 0, 0, pool, 0, 16);
  return p2::gpu_row_power(m.words.data() + m.tg / 4, shift, shift + row_len, out) ? 1 : 0;
}

uint64_t hc_key_hash_bytes(const uint8_t *p, uint32_t n) { return fast::key_hash_bytes(p, n); }

uint64_t hc_hash_bytes(const uint8_t *p, uint32_t n) {
  uint64_t h = hash_init();
  for (uint32_t i = 0; i < n; ++i) h = hash_step(h, p[i]);
  return h;
}

}  // extern "C"
