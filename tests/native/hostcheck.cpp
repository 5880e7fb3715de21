// TEST INFRASTRUCTURE: host build of the per-thread leaf functions that the CUDA kernels
// use (same headers, compiled by g++), so their byte-level semantics can be checked on a
// machine without a GPU.  Nothing in the product links or loads this.
#include <stdint.h>
#include <string.h>

#include <string>

#include "../../k8s-gpu-monitor_b200/csrc/kgm_common.cuh"
#include "../../k8s-gpu-monitor_b200/csrc/kgm_skeleton.h"

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

uint64_t hc_hash_bytes(const uint8_t *p, uint32_t n) {
  uint64_t h = hash_init();
  for (uint32_t i = 0; i < n; ++i) h = hash_step(h, p[i]);
  return h;
}

}  // extern "C"
