// sm_100a kernels of the poll pipeline (see DESIGN.md for the data flow).
//
//   stage 1  k_tokenize : nvidia-smi CSV text -> row tables        (gpu.py:40-55,62-73,104-116)
//   stage 2  k_join     : uuid dictionary, pid->container->pod join, line plan
//                                                                   (gpu.py:119-135,151,172-205,220-239)
//            k_scan     : device-wide exclusive scan of per-document sizes
//   stage 4  k_encode   : line plan -> line-protocol bytes          (gpu.py:136-143,206-216,240-250)
//
// All of it is HBM-bound byte/integer work: no tensor cores.  Stages 1 and 2 run one warp
// per document; stage 4 runs one thread per output line, staging a block's lines in shared
// memory and streaming them out with aligned 16-byte stores.
#include <cuda_runtime.h>

#include "../../include/kgm.h"
#include "kgm_device.cuh"

namespace kgm {
namespace {

constexpr uint32_t FULL = 0xffffffffu;

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t lanemask_lt() { return (1u << lane_id()) - 1u; }

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ uint32_t warp_min(uint32_t v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = min(v, __shfl_xor_sync(FULL, v, o));
  return v;
}
__device__ __forceinline__ uint32_t warp_max(uint32_t v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(FULL, v, o));
  return v;
}
__device__ __forceinline__ uint64_t warp_min64(uint64_t v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const uint64_t t = __shfl_xor_sync(FULL, v, o);
    v = t < v ? t : v;
  }
  return v;
}
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
  const uint32_t l = lane_id();
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(FULL, v, o);
    if (l >= (uint32_t)o) v += t;
  }
  return v;
}

// bit i set <=> byte i of the 16-byte vector equals the byte replicated in `pat`
__device__ __forceinline__ uint32_t eq_mask16(const uint4 &v, uint32_t pat) {
  auto m4 = [pat](uint32_t w) -> uint32_t {
    return (((__vcmpeq4(w, pat) & 0x01010101u) * 0x00204081u) >> 21) & 0xfu;
  };
  return m4(v.x) | (m4(v.y) << 4) | (m4(v.z) << 8) | (m4(v.w) << 12);
}

// bits of a 16-byte vector at `pos` whose positions fall inside [lo, hi)
__device__ __forceinline__ uint32_t range_mask16(uint32_t pos, uint32_t lo, uint32_t hi) {
  uint32_t m = 0xffffu;
  if (pos < lo) m &= (lo - pos >= 16u) ? 0u : (0xffffu << (lo - pos));
  if (pos + 16u > hi) m &= (hi <= pos) ? 0u : ((1u << (hi - pos)) - 1u);
  return m & 0xffffu;
}

// Every lane calls this; lanes with `has` learn whether a[0..len) == b[0..len).  The
// comparison itself is done by the whole warp, one candidate pair at a time, with
// coalesced byte loads (pairs are rare: only hash-equal keys get here).
__device__ bool warp_verify_equal(bool has, const uint8_t *a, const uint8_t *b, uint32_t len) {
  uint32_t todo = __ballot_sync(FULL, has);
  bool eq = false;
  const uint32_t l = lane_id();
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const uint8_t *pa = reinterpret_cast<const uint8_t *>(__shfl_sync(FULL, (unsigned long long)a, src));
    const uint8_t *pb = reinterpret_cast<const uint8_t *>(__shfl_sync(FULL, (unsigned long long)b, src));
    const uint32_t n = __shfl_sync(FULL, len, src);
    bool ne = false;
    for (uint32_t i = l; i < n; i += 32) ne |= pa[i] != pb[i];
    const bool any_ne = __any_sync(FULL, ne);
    if ((int)l == src) eq = !any_ne;
  }
  return eq;
}

// ======================================================================================
// stage 1: tokenizer
// ======================================================================================
template <class RowT, class ScanT, uint32_t MIN_CELLS>
__device__ void tokenize_text(const uint8_t *text, uint32_t lo, uint32_t hi, RowT *rows,
                              uint32_t cap, uint32_t *cursor, uint32_t *ovf, uint32_t ovf_bit,
                              uint32_t *row0_out, uint32_t *nrows_out, uint32_t *err_out) {
  const uint32_t l = lane_id();
  // pass 1: count line feeds and find the end of the header line (gpu.py:67-68)
  uint32_t cnt = 0, first = NONE32;
  for (uint32_t pos = (lo & ~15u) + l * 16u; pos < hi; pos += 512u) {
    const uint4 v = __ldg(reinterpret_cast<const uint4 *>(text + pos));
    const uint32_t m = eq_mask16(v, 0x0a0a0a0au) & range_mask16(pos, lo, hi);
    cnt += __popc(m);
    if (m && first == NONE32) first = pos + __ffs(m) - 1;
  }
  cnt = warp_sum(cnt);
  first = warp_min(first);
  uint32_t nrows = 0, body = hi;
  if (first != NONE32) {
    body = first + 1;
    nrows = cnt - 1;
    if (body < hi && text[hi - 1] != '\n') ++nrows;       // last row without terminator
  }
  uint32_t row0 = 0;
  if (l == 0 && nrows) row0 = atomicAdd(cursor, nrows);
  row0 = __shfl_sync(FULL, row0, 0);
  if ((uint64_t)row0 + nrows > cap) {
    if (l == 0) atomicOr(ovf, ovf_bit);
    *row0_out = 0; *nrows_out = 0; *err_out = 0;
    return;
  }
  rows += row0;
  // pass 2: row ends
  uint32_t running = 0;
  for (uint32_t it = body & ~15u; it < hi; it += 512u) {
    const uint32_t pos = it + l * 16u;
    uint32_t m = 0;
    if (pos < hi) {
      const uint4 v = __ldg(reinterpret_cast<const uint4 *>(text + pos));
      m = eq_mask16(v, 0x0a0a0a0au) & range_mask16(pos, body, hi);
    }
    const uint32_t c = __popc(m);
    const uint32_t incl = warp_incl_scan(c);
    uint32_t r = running + incl - c;
    while (m) {
      const uint32_t bit = __ffs(m) - 1;
      m &= m - 1;
      rows[r++].row_end = pos + bit;
    }
    running += __shfl_sync(FULL, incl, 31);
  }
  if (l == 0 && running < nrows) rows[nrows - 1].row_end = hi;
  __syncwarp();
  // pass 3: one lane per row
  uint32_t err = NONE32;
  for (uint32_t c0 = 0; c0 < nrows; c0 += 32) {
    const uint32_t r = c0 + l;
    const bool valid = r < nrows;
    uint32_t s = 0, e = 0;
    if (valid) {
      s = r ? rows[r - 1].row_end + 1 : body;
      e = rows[r].row_end;
    }
    __syncwarp();
    if (valid) {
      ScanT sc;
      sc.begin(s, e);
      for_each_byte(text, s, e, sc);
      sc.finish();
      sc.r.first = r;
      sc.r.last = r;
      rows[r] = sc.r;
      const uint32_t pe = parse_error_of(sc.r.flags, sc.r.ncells, MIN_CELLS);
      if (pe) err = min(err, (r << 8) | pe);
    }
    __syncwarp();
  }
  err = warp_min(err);
  *row0_out = row0;
  *nrows_out = nrows;
  *err_out = err == NONE32 ? 0u : (err & 0xffu);
}

constexpr int TOK_THREADS = 128;

__global__ void __launch_bounds__(TOK_THREADS) k_tokenize(BatchDev b, WorkDev w) {
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  for (uint32_t d = warp; d < b.n_docs; d += warps) {
    uint32_t g0, ng, gerr, p0, np, perr;
    tokenize_text<GpuRow, GpuRowScan, 1>(b.gpu_csv, (uint32_t)b.gpu_off[d], (uint32_t)b.gpu_off[d + 1],
                                         w.grow, w.grow_cap, w.counters + CTR_GROW,
                                         w.counters + CTR_OVERFLOW, OVF_GROW, &g0, &ng, &gerr);
    tokenize_text<ProcRow, ProcRowScan, 3>(b.apps_csv, (uint32_t)b.apps_off[d], (uint32_t)b.apps_off[d + 1],
                                           w.prow, w.prow_cap, w.counters + CTR_PROW,
                                           w.counters + CTR_OVERFLOW, OVF_PROW, &p0, &np, &perr);
    if (lane_id() == 0) {
      w.doc_g0[d] = g0; w.doc_ng[d] = ng;
      w.doc_p0[d] = p0; w.doc_np[d] = np;
      w.doc_perr[d] = gerr ? gerr : perr;     // the gpu text is parsed first (gpu.py:104,111)
    }
  }
}

// ======================================================================================
// stage 2: join + line plan
// ======================================================================================
constexpr int JOIN_THREADS = 128;

// where the reference would raise inside the container loop: (container rank, line
// position + 1 inside the container, exception class); smaller = earlier
__device__ __forceinline__ uint64_t raise_key(uint32_t c, uint32_t pos_p1, uint32_t cls) {
  return ((uint64_t)c << 40) | ((uint64_t)pos_p1 << 8) | cls;
}

__device__ void join_doc(uint32_t d, const BatchDev &b, const FleetDev &f, const WorkDev &w,
                         uint32_t *cnt, uint32_t *start, uint32_t *cur) {
  const uint32_t l = lane_id();
  const uint32_t n = b.doc_node[d];
  const uint32_t g0 = w.doc_g0[d], ng = w.doc_ng[d], p0 = w.doc_p0[d], np = w.doc_np[d];
  const uint32_t cb = f.cont_begin[n], nc = f.cont_begin[n + 1] - cb;
  GpuRow *G = w.grow + g0;
  ProcRow *P = w.prow + p0;
  uint32_t *uniq = w.uniq + g0;

  uint32_t status = f.node_err_parse[n];         // docker ps/stats parse comes first (gpu.py:92-103)
  if (!status) status = w.doc_perr[d];
  uint32_t n_node = 0, total_pod = 0, nu = 0;
  bool pods = false;
  uint64_t rk = ~0ull;

  if (!status) {
    // ---- uuid dictionary: first-seen order, last row's values (gpu.py:54,119-128) ----
    for (uint32_t c0 = 0; c0 < ng; c0 += 32) {
      const uint32_t r = c0 + l;
      const bool valid = r < ng;
      uint64_t h = 0;
      uint32_t klen = 0;
      const uint8_t *kp = b.gpu_csv;
      if (valid) {
        h = G[r].key_hash;
        klen = (uint32_t)G[r].cell_end[0] - G[r].key_lo;
        kp = b.gpu_csv + G[r].row_off + G[r].key_lo;
      }
      uint32_t first = r;
      const uint32_t jmax = min(c0 + 32u, ng);
      for (uint32_t j = 0; j + 1 < jmax; ++j) {
        const uint64_t hj = G[j].key_hash;
        const uint32_t lj = (uint32_t)G[j].cell_end[0] - G[j].key_lo;
        const bool cand = valid && j < r && first == r && hj == h && lj == klen;
        const bool eq = warp_verify_equal(cand, kp, b.gpu_csv + G[j].row_off + G[j].key_lo, klen);
        if (cand && eq) first = j;
      }
      if (valid) {
        G[r].first = first;
        if (first != r) atomicMax(&G[first].last, r);
      }
      const bool isf = valid && first == r;
      const uint32_t bal = __ballot_sync(FULL, isf);
      if (isf) {
        const uint32_t sid = nu + __popc(bal & lanemask_lt());
        G[r].sid = sid;
        uniq[sid] = r;
      }
      nu += __popc(bal);
      __syncwarp();
    }
    __threadfence_block();
    __syncwarp();

    // ---- node loop verdicts (gpu.py:128 renamekeys, then 132-143 per gpu) ----
    bool rename_fail = false;
    uint32_t fail = NONE32;
    for (uint32_t c0 = 0; c0 < nu; c0 += 32) {
      const uint32_t k = c0 + l;
      if (k < nu) {
        const uint32_t vr = __ldcg(&G[uniq[k]].last);
        const GpuRow &v = G[vr];
        if (v.ncells < 2) rename_fail = true;
        uint32_t e = node_error_of(v);
        uint8_t *fs = w.fstr + (size_t)(g0 + vr) * FSTR_STRIDE;
        if (!e) {
          const int flen = float_repr(mem_util(v.mem_used, v.mem_total), fs);
          fs[24] = (uint8_t)flen;
          const uint32_t len = node_line_len(f.node_fixed[n], v, k, (uint32_t)flen);
          *reinterpret_cast<uint32_t *>(fs + 28) = len;
          if (len > MAX_LINE_BYTES) e = DOC_UNSUP_RANGE;
        }
        if (e) fail = min(fail, (k << 8) | e);
      }
    }
    rename_fail = __any_sync(FULL, rename_fail);
    fail = warp_min(fail);
    if (rename_fail) {
      status = DOC_KEY;                            // gpu.py:82: nothing emitted yet
    } else if (fail != NONE32) {
      n_node = fail >> 8;
      status = fail & 0xffu;
    } else {
      n_node = nu;
    }
  }

  if (!status) {
    // ---- process rows: uuid -> short id (gpu.py:151) ----
    bool unknown = false;
    for (uint32_t c0 = 0; c0 < np; c0 += 32) {
      const uint32_t r = c0 + l;
      const bool valid = r < np;
      uint64_t h = 0;
      uint32_t klen = 0;
      const uint8_t *kp = b.apps_csv;
      if (valid) {
        h = P[r].uuid_hash;
        klen = (uint32_t)P[r].cell_end[2] - P[r].key_lo;
        kp = b.apps_csv + P[r].row_off + P[r].key_lo;
      }
      uint32_t sid = NONE32;
      for (uint32_t k = 0; k < nu; ++k) {
        const GpuRow &u = G[uniq[k]];
        const uint32_t lk = (uint32_t)u.cell_end[0] - u.key_lo;
        const bool cand = valid && sid == NONE32 && u.key_hash == h && lk == klen;
        const bool eq = warp_verify_equal(cand, kp, b.gpu_csv + u.row_off + u.key_lo, klen);
        if (cand && eq) sid = k;
      }
      if (valid) {
        P[r].sid = sid;
        if (sid == NONE32) unknown = true;
      }
    }
    unknown = __any_sync(FULL, unknown);
    __syncwarp();
    if (unknown) status = DOC_KEY;
    else status = f.node_err_join[n];              // gpu.py:172
  }

  if (!status) {
    pods = true;
    // ---- (pid, gpu) dictionary: first position, last values (gpu.py:54,151) ----
    for (uint32_t c0 = 0; c0 < np; c0 += 32) {
      const uint32_t r = c0 + l;
      const bool valid = r < np;
      uint64_t lo = 0, hi = 0;
      uint32_t sid = 0;
      bool longpid = true;
      if (valid) { lo = P[r].pid_lo; hi = P[r].pid_hi; sid = P[r].sid; longpid = (P[r].flags & RF_LONGPID) != 0; }
      uint32_t first = r;
      const uint32_t jmax = min(c0 + 32u, np);
      for (uint32_t j = 0; j + 1 < jmax; ++j) {
        const ProcRow &q = P[j];
        if (valid && !longpid && j < r && first == r && !(q.flags & RF_LONGPID) &&
            q.pid_lo == lo && q.pid_hi == hi && q.sid == sid)
          first = j;
      }
      if (valid) {
        P[r].first = first;
        if (first != r) atomicMax(&P[first].last, r);
      }
    }
    for (uint32_t c = l; c < nc; c += 32) { cnt[c] = 0; cur[c] = 0; }
    __threadfence_block();
    __syncwarp();

    // ---- probe the resident pid table (gpu.py:177-178) ----
    for (uint32_t r = l; r < np; r += 32) {
      const ProcRow &q = P[r];
      if (q.first != r || (q.flags & RF_LONGPID)) continue;
      uint32_t nm = 0, c_first = NONE32;
      uint32_t h = (uint32_t)pid_hash(q.pid_lo, q.pid_hi, n) & f.pid_mask;
      for (;;) {
        const PidSlot s = f.pid_table[h];
        if (s.node == NONE32) break;
        if (s.node == n && s.lo == q.pid_lo && s.hi == q.pid_hi) {
          if (!nm) c_first = s.crank;
          ++nm;
          atomicAdd(&cnt[s.crank], 1u);
        }
        h = (h + 1) & f.pid_mask;
      }
      P[r].crank = c_first;
      P[r].nmatch = nm;
    }
    __syncwarp();

    // ---- per container: line count, first line, container-level raise points ----
    uint32_t running = 0;
    for (uint32_t c0 = 0; c0 < nc; c0 += 32) {
      const uint32_t c = c0 + l;
      uint32_t lines = 0;
      if (c < nc) {
        const uint32_t gc = cb + c;
        const bool gate = f.gated[gc] != 0;
        const uint32_t k = cnt[c];
        const uint32_t nv = f.vis_begin[gc + 1] - f.vis_begin[gc];
        lines = gate ? (k ? k : nv) : 0u;
        const uint32_t ei = f.err_inspect[gc], ee = f.err_emit[gc];
        if (ei) rk = min(rk, raise_key(c, 0, ei));                     // gpu.py:184-187
        else if (gate && lines && ee) rk = min(rk, raise_key(c, 1, ee));   // gpu.py:193-199 / 227-233
      }
      const uint32_t incl = warp_incl_scan(lines);
      if (c < nc) start[c] = running + incl - lines;
      running += __shfl_sync(FULL, incl, 31);
    }
    total_pod = running;
    __syncwarp();
  }

  // ---- allocate the document's slice of the line plan ----
  const uint32_t n_alloc = n_node + total_pod;
  uint32_t desc0 = 0;
  if (l == 0 && n_alloc) desc0 = atomicAdd(w.counters + CTR_DESC, n_alloc);
  desc0 = __shfl_sync(FULL, desc0, 0);
  if ((uint64_t)desc0 + n_alloc > w.desc_cap) {
    if (l == 0) {
      atomicOr(w.counters + CTR_OVERFLOW, OVF_DESC);
      w.doc_desc0[d] = 0; w.doc_nlines[d] = 0; w.doc_nbytes[d] = 0; w.doc_status[d] = (int32_t)status;
    }
    return;
  }

  // node lines (gpu.py:132-146)
  for (uint32_t k = l; k < n_node; k += 32) {
    const uint32_t vr = __ldcg(&G[uniq[k]].last);
    const uint32_t idx = desc0 + k;
    w.desc[idx] = LineDesc{d, LINE_NODE << 30, g0 + vr, k};
    w.line_len[idx] = *reinterpret_cast<const uint32_t *>(w.fstr + (size_t)(g0 + vr) * FSTR_STRIDE + 28);
  }

  if (pods) {
    const uint32_t pod0 = desc0 + n_node;
    // ---- busy lines, container-major, first-insertion order inside (gpu.py:178,190) ----
    auto emit_busy = [&](uint32_t r, uint32_t c, uint32_t pos) {
      const uint32_t gc = cb + c;
      if (!f.gated[gc]) return;
      const uint32_t pvr = __ldcg(&P[r].last);
      const ProcRow &pv = P[pvr];
      const uint32_t sid = pv.sid;
      const uint32_t gvr = __ldcg(&G[uniq[sid]].last);
      const uint32_t len = busy_line_len(f.cont_fixed[gc], pv, G[gvr], sid);
      const uint32_t idx = pod0 + start[c] + pos;
      w.desc[idx] = LineDesc{d, (LINE_BUSY << 30) | gc, p0 + pvr, g0 + gvr};
      w.line_len[idx] = len;
      if (pv.ncells < 4) rk = min(rk, raise_key(c, pos + 1, DOC_KEY));                // gpu.py:204
      else if (len > MAX_LINE_BYTES) rk = min(rk, raise_key(c, pos + 1, DOC_UNSUP_RANGE));
    };
    for (uint32_t c0 = 0; c0 < np; c0 += 32) {
      const uint32_t r = c0 + l;
      uint32_t nm = 0;
      uint64_t lo = 0, hi = 0;
      if (r < np && P[r].first == r && !(P[r].flags & RF_LONGPID)) {
        nm = P[r].nmatch; lo = P[r].pid_lo; hi = P[r].pid_hi;
      }
      const uint32_t maxm = warp_max(nm);
      if (maxm == 1) {
        // common case: each pid is listed by at most one container.  Lanes are in entry
        // order, so the rank among the lanes that hit the same container is the position.
        const uint32_t c = nm ? P[r].crank : NONE32;
        const uint32_t act = __ballot_sync(FULL, c != NONE32);
        if (c != NONE32) {
          const uint32_t peers = __match_any_sync(act, c);
          const uint32_t rank = __popc(peers & lanemask_lt());
          const int leader = __ffs(peers) - 1;
          uint32_t base = 0;
          if ((int)l == leader) { base = cur[c]; cur[c] = base + __popc(peers); }
          base = __shfl_sync(peers, base, leader);
          emit_busy(r, c, base + rank);
        }
      } else if (maxm > 1) {
        // a pid listed by several containers (gpu.py:178 emits it under each): keep entry
        // order by letting the lanes of this chunk take their positions one after another
        for (uint32_t src = 0; src < 32; ++src) {
          if (l == src && nm) {
            uint32_t h = (uint32_t)pid_hash(lo, hi, n) & f.pid_mask;
            for (;;) {
              const PidSlot s = f.pid_table[h];
              h = (h + 1) & f.pid_mask;
              if (s.node == NONE32) break;
              if (s.node == n && s.lo == lo && s.hi == hi) {
                const uint32_t pos = cur[s.crank];
                cur[s.crank] = pos + 1;
                emit_busy(r, s.crank, pos);
              }
            }
          }
          __syncwarp();
        }
      }
      __syncwarp();
    }
    // ---- idle lines: gated container without a matched process (gpu.py:220-253) ----
    for (uint32_t c = 0; c < nc; ++c) {
      const uint32_t gc = cb + c;
      if (!f.gated[gc] || cnt[c]) continue;
      const uint32_t vb = f.vis_begin[gc], ve = f.vis_begin[gc + 1];
      for (uint32_t v0 = vb; v0 < ve; v0 += 32) {
        const uint32_t v = v0 + l;
        const bool valid = v < ve;
        uint64_t h = 0;
        uint32_t vl = 0;
        const uint8_t *vp = f.vis_bytes;
        if (valid) { h = f.vis_hash[v]; vl = f.vis_off[v + 1] - f.vis_off[v]; vp = f.vis_bytes + f.vis_off[v]; }
        uint32_t sid = NONE32;
        for (uint32_t k = 0; k < nu; ++k) {
          const GpuRow &u = G[uniq[k]];
          const uint32_t lk = (uint32_t)u.cell_end[0] - u.key_lo;
          const bool cand = valid && sid == NONE32 && u.key_hash == h && lk == vl;
          const bool eq = warp_verify_equal(cand, vp, b.gpu_csv + u.row_off + u.key_lo, vl);
          if (cand && eq) sid = k;
        }
        if (valid) {
          const uint32_t pos = v - vb;
          const uint32_t len = idle_line_len(f.cont_fixed[gc], sid);
          const uint32_t idx = pod0 + start[c] + pos;
          w.desc[idx] = LineDesc{d, (LINE_IDLE << 30) | gc, sid, 0};
          w.line_len[idx] = len;
          if (len > MAX_LINE_BYTES) rk = min(rk, raise_key(c, pos + 1, DOC_UNSUP_RANGE));
        }
      }
    }
    rk = warp_min64(rk);
  }

  uint32_t n_emit = n_alloc;
  if (rk != ~0ull) {
    const uint32_t c = (uint32_t)(rk >> 40);
    const uint32_t pos_p1 = (uint32_t)(rk >> 8);
    n_emit = n_node + start[c] + (pos_p1 ? pos_p1 - 1 : 0);
    status = (uint32_t)(rk & 0xffu);
  }
  __threadfence_block();
  __syncwarp();

  // ---- byte offset of each emitted line inside the document ----
  uint32_t off = 0;
  for (uint32_t i0 = 0; i0 < n_emit; i0 += 32) {
    const uint32_t i = i0 + l;
    const uint32_t len = i < n_emit ? w.line_len[desc0 + i] : 0u;
    const uint32_t incl = warp_incl_scan(len);
    if (i < n_emit) w.line_loff[desc0 + i] = off + incl - len;
    off += __shfl_sync(FULL, incl, 31);
  }
  if (l == 0) {
    w.doc_desc0[d] = desc0;
    w.doc_nlines[d] = n_emit;
    w.doc_nbytes[d] = off;
    w.doc_status[d] = (int32_t)status;
  }
}

__global__ void __launch_bounds__(JOIN_THREADS) k_join(BatchDev b, FleetDev f, WorkDev w) {
  extern __shared__ uint32_t join_smem[];
  const uint32_t wib = threadIdx.x >> 5;
  uint32_t *cnt = join_smem + (size_t)wib * 3 * f.max_nc;
  uint32_t *start = cnt + f.max_nc;
  uint32_t *cur = start + f.max_nc;
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  for (uint32_t d = warp; d < b.n_docs; d += warps) {
    join_doc(d, b, f, w, cnt, start, cur);
    __syncwarp();
  }
}

// ======================================================================================
// device-wide exclusive scan over documents (lines, bytes)
// ======================================================================================
constexpr int SCAN_THREADS = 1024;

__global__ void __launch_bounds__(SCAN_THREADS) k_scan(uint32_t n_docs, WorkDev w) {
  __shared__ uint64_t sl[32], sb[32];
  __shared__ uint64_t carry_l, carry_b;
  const uint32_t t = threadIdx.x, l = t & 31u, wi = t >> 5;
  if (t == 0) { carry_l = 0; carry_b = 0; }
  __syncthreads();
  for (uint32_t base = 0; base < n_docs; base += SCAN_THREADS) {
    const uint32_t i = base + t;
    uint64_t nl = i < n_docs ? w.doc_nlines[i] : 0, nb = i < n_docs ? w.doc_nbytes[i] : 0;
    uint64_t il = nl, ib = nb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t tl = __shfl_up_sync(FULL, il, o), tb = __shfl_up_sync(FULL, ib, o);
      if (l >= (uint32_t)o) { il += tl; ib += tb; }
    }
    if (l == 31) { sl[wi] = il; sb[wi] = ib; }
    __syncthreads();
    if (wi == 0) {
      uint64_t xl = sl[l], xb = sb[l];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint64_t tl = __shfl_up_sync(FULL, xl, o), tb = __shfl_up_sync(FULL, xb, o);
        if (l >= (uint32_t)o) { xl += tl; xb += tb; }
      }
      sl[l] = xl; sb[l] = xb;                     // inclusive over warps
    }
    __syncthreads();
    const uint64_t wl = wi ? sl[wi - 1] : 0, wb = wi ? sb[wi - 1] : 0;
    if (i < n_docs) {
      w.doc_line_off[i] = carry_l + wl + il - nl;
      w.doc_byte_off[i] = carry_b + wb + ib - nb;
    }
    __syncthreads();
    if (t == 0) { carry_l += sl[31]; carry_b += sb[31]; }
    __syncthreads();
  }
  if (t == 0) {
    w.doc_line_off[n_docs] = carry_l;
    w.doc_byte_off[n_docs] = carry_b;
    w.totals[0] = carry_l;
    w.totals[1] = carry_b;
    write_out_header(w, carry_l, carry_b);
  }
}

// ======================================================================================
// stage 4: encoder (write pass)
// ======================================================================================
constexpr int ENC_THREADS = 128;
constexpr uint32_t STAGE_BYTES = 48 * 1024;

__device__ __forceinline__ void copy_out(uint8_t *out, const uint8_t *stage, uint64_t sbase,
                                         uint64_t lo, uint64_t hi) {
  const uint32_t t = threadIdx.x;
  const uint64_t a_lo = (lo + 15) & ~15ull, a_hi = hi & ~15ull;
  if (a_lo >= a_hi) {
    for (uint64_t g = lo + t; g < hi; g += ENC_THREADS) out[g] = stage[g - sbase];
    return;
  }
  if (lo + t < a_lo) out[lo + t] = stage[lo + t - sbase];
  for (uint64_t v = a_lo + 16ull * t; v < a_hi; v += 16ull * ENC_THREADS)
    *reinterpret_cast<uint4 *>(out + v) = *reinterpret_cast<const uint4 *>(stage + (v - sbase));
  if (a_hi + t < hi) out[a_hi + t] = stage[a_hi + t - sbase];
}

__global__ void __launch_bounds__(ENC_THREADS) k_encode(BatchDev b, FleetDev f, WorkDev w) {
  extern __shared__ __align__(16) uint8_t stage[];
  __shared__ unsigned long long s_g0, s_g1, s_cut;
  const uint64_t total_lines = w.totals[0];
  if (w.totals[1] > w.out_cap) {
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(w.counters + CTR_OVERFLOW, OVF_OUT);
    return;
  }
  const uint32_t t = threadIdx.x;
  for (uint64_t L0 = (uint64_t)blockIdx.x * ENC_THREADS; L0 < total_lines;
       L0 += (uint64_t)gridDim.x * ENC_THREADS) {
    const uint64_t L = L0 + t;
    const bool valid = L < total_lines;
    LineDesc ds = LineDesc{0, 0, 0, 0};
    uint32_t len = 0;
    uint64_t goff = 0;
    if (valid) {
      uint32_t lo = 0, hi = b.n_docs;           // document holding line L
      while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (w.doc_line_off[mid + 1] <= L) lo = mid + 1; else hi = mid;
      }
      const uint32_t di = w.doc_desc0[lo] + (uint32_t)(L - w.doc_line_off[lo]);
      ds = w.desc[di];
      len = w.line_len[di];
      goff = w.doc_byte_off[lo] + w.line_loff[di];
      w.line_off[L] = goff;
      if (L + 1 == total_lines) w.line_off[total_lines] = goff + len;
      if (t == 0) s_g0 = goff;
      if (t == ENC_THREADS - 1 || L + 1 == total_lines) s_g1 = goff + len;
    }
    __syncthreads();
    uint64_t curp = s_g0;
    const uint64_t g1 = s_g1;
    bool done = !valid;
    while (curp < g1) {
      const uint64_t sbase = curp & ~15ull;
      if (t == 0) s_cut = 0;
      __syncthreads();
      if (!done && goff >= curp && goff + len <= sbase + STAGE_BYTES) {
        ByteSink s{stage + (goff - sbase)};
        const uint32_t kind = ds.kind_ref >> 30, ref = ds.kind_ref & 0x3fffffffu;
        if (kind == LINE_NODE) {
          const uint32_t n = b.doc_node[ds.doc];
          compose_node_line(s, f.skel, f.node_piece + (size_t)n * NODE_PIECES, b.gpu_csv,
                            w.grow[ds.a], ds.b, w.fstr + (size_t)ds.a * FSTR_STRIDE);
        } else if (kind == LINE_BUSY) {
          const ProcRow &pr = w.prow[ds.a];
          compose_busy_line(s, f.skel, f.cont_piece + (size_t)ref * POD_PIECES, b.gpu_csv,
                            b.apps_csv, pr, w.grow[ds.b], pr.sid);
        } else {
          compose_idle_line(s, f.skel, f.cont_piece + (size_t)ref * POD_PIECES, ds.a);
        }
        done = true;
        atomicMax(&s_cut, (unsigned long long)(goff + len));
      }
      __syncthreads();
      const uint64_t cut = s_cut;
      if (cut <= curp) break;                   // cannot happen: every line fits the stage
      copy_out(w.out, stage, sbase, curp, cut);
      __syncthreads();
      curp = cut;
    }
    __syncthreads();
  }
}

__global__ void k_flush(uint4 *buf, uint64_t n16) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride)
    buf[i] = make_uint4((uint32_t)i, 1u, 2u, 3u);
}

}  // namespace
}  // namespace kgm

#include "kgm_fused.cuh"
#include "kgm_poll2.cuh"

namespace kgm {

void launch_tokenize(const BatchDev &b, const WorkDev &w, int sms, cudaStream_t s) {
  const int warps_per_block = TOK_THREADS / 32;
  int blocks = (int)((b.n_docs + warps_per_block - 1) / warps_per_block);
  blocks = blocks < 1 ? 1 : (blocks > sms * 16 ? sms * 16 : blocks);
  k_tokenize<<<blocks, TOK_THREADS, 0, s>>>(b, w);
}

void launch_join(const BatchDev &b, const FleetDev &f, const WorkDev &w, int sms, cudaStream_t s) {
  const int warps_per_block = JOIN_THREADS / 32;
  const size_t smem = (size_t)warps_per_block * 3 * f.max_nc * sizeof(uint32_t);
  int blocks = (int)((b.n_docs + warps_per_block - 1) / warps_per_block);
  blocks = blocks < 1 ? 1 : (blocks > sms * 16 ? sms * 16 : blocks);
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(k_join, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_join<<<blocks, JOIN_THREADS, smem, s>>>(b, f, w);
}

void launch_scan(const BatchDev &b, const WorkDev &w, cudaStream_t s) {
  k_scan<<<1, SCAN_THREADS, 0, s>>>(b.n_docs, w);
}

int encode_smem_bytes() { return (int)STAGE_BYTES + 16; }

void launch_encode(const BatchDev &b, const FleetDev &f, const WorkDev &w, int sms, int device, cudaStream_t s) {
  static bool attr_set[64];                          // the shared-memory opt-in is per device
  if (device < 0 || device >= 64 || !attr_set[device]) {
    cudaFuncSetAttribute(k_encode, cudaFuncAttributeMaxDynamicSharedMemorySize, encode_smem_bytes());
    if (device >= 0 && device < 64) attr_set[device] = true;
  }
  k_encode<<<sms * 4, ENC_THREADS, encode_smem_bytes(), s>>>(b, f, w);
}

void launch_flush(uint8_t *buf, uint64_t bytes, int sms, cudaStream_t s) {
  k_flush<<<sms * 8, 256, 0, s>>>(reinterpret_cast<uint4 *>(buf), bytes / 16);
}

}  // namespace kgm

// ======================================================================================
// Rank-major packing of fixed-stride shard slots (kgm_concat_shards, include/kgm.h)
// ======================================================================================
namespace kgm {
namespace {

__global__ void k_concat_shards(const uint8_t *slots, uint64_t stride, const long long *counts, uint32_t world,
                                uint8_t *out, uint64_t out_cap, unsigned long long *totals_out) {
  // every block handles every shard in turn (grid-stride over 16-byte units); shard sizes
  // are read from device memory, the destination offset is their running sum
  uint64_t base = 0, lines = 0;
  bool bad = false;
  for (uint32_t r = 0; r < world; ++r) {
    const long long *hdr = counts ? counts + 2 * r : reinterpret_cast<const long long *>(slots + (uint64_t)r * stride);
    uint64_t n = (uint64_t)hdr[0];
    const uint8_t *src = slots + (uint64_t)r * stride + (counts ? 0 : 16);
    // a size the slot cannot hold (a stale or damaged header) or that does not fit the output:
    // the shard is left out and the totals say so
    if (n > stride - (counts ? 0u : 16u) || base + n > out_cap) { bad = true; n = 0; }
    if (n) {
      uint8_t *dst = out + base;
      const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
      // bytes up to the destination's 16-byte boundary, realigned vectors, tail bytes
      const uint64_t head = min(n, (uint64_t)((16u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u)) & 15u));
      if (tid < head) dst[tid] = src[tid];
      const uint64_t nvec = (n - head) >> 4;
      const uint32_t sh = 8u * (uint32_t)(head & 3u);
      const uint32_t *sw = reinterpret_cast<const uint32_t *>(src + (head & ~3ull));   // slots are 16-byte aligned
      for (uint64_t v = tid; v < nvec; v += nthreads) {
        const uint32_t *q = sw + 4 * v;
        const uint32_t a0 = q[0], a1 = q[1], a2 = q[2], a3 = q[3], a4 = sh ? q[4] : 0u;
        uint4 o;
        o.x = __funnelshift_r(a0, a1, sh); o.y = __funnelshift_r(a1, a2, sh);
        o.z = __funnelshift_r(a2, a3, sh); o.w = __funnelshift_r(a3, a4, sh);
        *reinterpret_cast<uint4 *>(dst + head + 16 * v) = o;
      }
      const uint64_t done = head + 16 * nvec;
      if (tid < n - done) dst[done + tid] = src[done + tid];
    }
    base += n;
    lines += (uint64_t)hdr[1];
  }
  if (totals_out && blockIdx.x == 0 && threadIdx.x == 0) {
    totals_out[0] = bad ? ~0ull : base;
    totals_out[1] = bad ? ~0ull : lines;
  }
}

}  // namespace
}  // namespace kgm

extern "C" int kgm_concat_shards(const uint8_t *slots, uint64_t stride, const int64_t *counts, uint32_t world,
                                 uint8_t *out, uint64_t out_cap, uint64_t *totals_out, void *stream) {
  if (!slots || !out || !world || (stride & 15) || (reinterpret_cast<uintptr_t>(slots) & 15))
    return KGM_E_INVALID;
  static int sm_count[64];                           // per device
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return KGM_E_CUDA;
  int sms = dev >= 0 && dev < 64 ? sm_count[dev] : 0;
  if (!sms) {
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) return KGM_E_CUDA;
    if (dev >= 0 && dev < 64) sm_count[dev] = sms;
  }
  kgm::k_concat_shards<<<sms * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      slots, stride, reinterpret_cast<const long long *>(counts), world, out, out_cap,
      reinterpret_cast<unsigned long long *>(totals_out));
  return cudaGetLastError() == cudaSuccess ? KGM_OK : KGM_E_CUDA;
}

// ======================================================================================
// Stream-ordered flags between GPUs (kgm_stream_wait_flags / kgm_stream_signal_flags)
// ======================================================================================
// The ranks of a multi-GPU job tell each other "my shard of step k has landed in your memory"
// and "slot set k is free again" through 32-bit counters in peer-mapped memory instead of a
// collective per step: a one-warp kernel stores the values behind the poll kernel in stream
// order, another one-warp kernel holds a stream until the counters it names have been reached.
namespace kgm {
namespace {

struct FlagTargets { uint32_t *p[KGM_MAX_FLAG_TARGETS]; };

__global__ void k_signal_flags(FlagTargets t, uint32_t n, uint32_t value) {
  if (threadIdx.x < n) {
    __threadfence_system();                          // everything this stream wrote before is visible first
    // max, not store: the steps of one rank run on several streams and may finish out of order,
    // and a counter must never go back
    asm volatile("red.release.sys.global.max.u32 [%0], %1;" ::"l"(t.p[threadIdx.x]), "r"(value) : "memory");
  }
}

__global__ void k_wait_flags(const uint32_t *flags, uint32_t n, uint32_t value) {
  if (threadIdx.x < n) {
    for (;;) {
      uint32_t v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
      if ((int32_t)(v - value) >= 0) break;          // counters only grow (wrap-safe comparison)
      __nanosleep(200);
    }
  }
}

}  // namespace
}  // namespace kgm

// KGM_FLAG_MEMOPS=1: stream memory operations (cuStreamWriteValue32 / cuStreamWaitValue32) instead of the
// one-warp kernels: no SM is needed, so the flags do not wait for room beside the persistent poll kernel.
// A write is a plain store, not a max: the caller then has to give every stream that signals its own
// counter (kgm/shard.py does: one per rank and context).  Whatever the driver declines falls back to
// the kernels.
#include <cuda.h>
namespace kgm {
namespace {
typedef CUresult (*MemOp32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
struct MemOps {
  MemOp32 write = nullptr, wait = nullptr;
  bool on = false;
  MemOps() {
    const char *env = getenv("KGM_FLAG_MEMOPS");
    if (!env || atoi(env) == 0) return;
    cudaDriverEntryPointQueryResult q;
    void *w = nullptr, *v = nullptr;
    if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &w, cudaEnableDefault, &q) != cudaSuccess || !w) return;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &v, cudaEnableDefault, &q) != cudaSuccess || !v) return;
    write = reinterpret_cast<MemOp32>(w); wait = reinterpret_cast<MemOp32>(v);
    on = true;
  }
};
const MemOps &memops() { static MemOps m; return m; }
}  // namespace
}  // namespace kgm

extern "C" int kgm_stream_signal_flags(void *stream, uint32_t *const *targets, uint32_t n, uint32_t value) {
  if (!targets || n == 0 || n > KGM_MAX_FLAG_TARGETS) return KGM_E_INVALID;
  if (kgm::memops().on) {
    bool all = true;
    for (uint32_t i = 0; i < n && all; ++i)
      all = targets[i] && kgm::memops().write(static_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(targets[i]), value,
                                               CU_STREAM_WRITE_VALUE_DEFAULT) == CUDA_SUCCESS;
    if (all) return KGM_OK;
    cudaGetLastError();
  }
  kgm::FlagTargets t{};
  for (uint32_t i = 0; i < n; ++i) {
    if (!targets[i] || (reinterpret_cast<uintptr_t>(targets[i]) & 3)) return KGM_E_INVALID;
    t.p[i] = targets[i];
  }
  kgm::k_signal_flags<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(t, n, value);
  return cudaGetLastError() == cudaSuccess ? KGM_OK : KGM_E_CUDA;
}

extern "C" int kgm_stream_wait_flags(void *stream, const uint32_t *flags, uint32_t n, uint32_t value) {
  if (!flags || n == 0 || n > 32 || (reinterpret_cast<uintptr_t>(flags) & 3)) return KGM_E_INVALID;
  if (kgm::memops().on) {
    bool all = true;
    for (uint32_t i = 0; i < n && all; ++i)                // (int32_t)(*addr - value) >= 0, like k_wait_flags
      all = kgm::memops().wait(static_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(flags + i), value,
                               CU_STREAM_WAIT_VALUE_GEQ) == CUDA_SUCCESS;
    if (all) return KGM_OK;
    cudaGetLastError();
  }
  kgm::k_wait_flags<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(flags, n, value);
  return cudaGetLastError() == cudaSuccess ? KGM_OK : KGM_E_CUDA;
}

// ======================================================================================
// Allocation-free tokenizer entry point (kgm_tokenize_gpu_noalloc, include/kgm.h)
// ======================================================================================
// Some processes that want the per-GPU samples of a batch are not allowed to take device
// memory of their own: a monitoring sidecar sharing a GPU whose HBM a training job has
// reserved to the last byte (cudaMalloc fails there, or stalls the tenant while the driver
// synchronises the device), or a collector running under a memory manager that owns every
// allocation.  For them the stage-1 tokenizer is offered once more with nothing to
// allocate: the caller brings the input text and the output arrays, and the kernel's own
// working storage -- the row cursor, the overflow flag and the row table it builds between
// its passes -- lives in the fixed array below, which every GPU that loads this library
// carries until the process exits.
namespace kgm {
namespace {

constexpr uint32_t NOALLOC_MAX_ROWS = KGM_NOALLOC_MAX_ROWS;
constexpr size_t NOALLOC_HEADER = 256;          // [0] row cursor, [1] overflow flag
constexpr size_t NOALLOC_BYTES = NOALLOC_HEADER + (size_t)NOALLOC_MAX_ROWS * sizeof(GpuRow);

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
static void *g_noalloc_ptr[NOALLOC_MAX_DEVICES];   // device address of the array, per device

__global__ void __launch_bounds__(TOK_THREADS)
k_tokenize_gpu_noalloc(const uint8_t *text, const uint64_t *off, uint32_t n_docs,
                       unsigned char *scratch, uint32_t cap, uint32_t *doc_row0,
                       uint32_t *doc_nrows, uint32_t *doc_err, uint32_t *util,
                       uint64_t *mem_used, uint64_t *mem_total) {
  uint32_t *counters = reinterpret_cast<uint32_t *>(scratch);
  GpuRow *rows = reinterpret_cast<GpuRow *>(scratch + NOALLOC_HEADER);
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  for (uint32_t d = warp; d < n_docs; d += warps) {
    uint32_t g0, ng, gerr;
    tokenize_text<GpuRow, GpuRowScan, 1>(text, (uint32_t)off[d], (uint32_t)off[d + 1], rows, cap,
                                         counters, counters + 1, 1u, &g0, &ng, &gerr);
    __syncwarp();
    for (uint32_t r = lane_id(); r < ng; r += 32) {
      const GpuRow &g = rows[g0 + r];
      util[g0 + r] = g.util;
      mem_used[g0 + r] = g.mem_used;
      mem_total[g0 + r] = g.mem_total;
    }
    if (lane_id() == 0) { doc_row0[d] = g0; doc_nrows[d] = ng; doc_err[d] = gerr; }
  }
}

}  // namespace
}  // namespace kgm

extern "C" int kgm_tokenize_gpu_noalloc(const uint8_t *gpu_csv, const uint64_t *gpu_off,
                                        uint32_t n_docs, uint64_t n_bytes, uint32_t max_rows,
                                        uint32_t *doc_row0, uint32_t *doc_nrows, uint32_t *doc_err,
                                        uint32_t *util, uint64_t *mem_used, uint64_t *mem_total,
                                        void *stream) {
  using namespace kgm;
  if (!gpu_off || !doc_row0 || !doc_nrows || !doc_err || !util || !mem_used || !mem_total ||
      (n_bytes && !gpu_csv) || (reinterpret_cast<uintptr_t>(gpu_csv) & 15))
    return KGM_E_INVALID;
  if (max_rows > NOALLOC_MAX_ROWS || n_bytes >= 0xfff00000ull) return KGM_E_LIMIT;   // beyond the fixed array
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NOALLOC_MAX_DEVICES) return KGM_E_CUDA;
  if (!g_noalloc_ptr[dev]) {
    // first call on this device: look the array's device address up.  `g_noalloc_scratch`
    // as written in host code is only a handle; the kernel needs the address behind it.
    void *p = nullptr;
    if (cudaGetSymbolAddress(&p, g_noalloc_scratch) != cudaSuccess || !p) { cudaGetLastError(); return KGM_E_CUDA; }
    g_noalloc_ptr[dev] = p;
  }
  unsigned char *scratch = static_cast<unsigned char *>(g_noalloc_ptr[dev]);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // the cursor and the overflow flag must start at zero; the row table is fully rewritten
  if (cudaMemsetAsync(scratch, 0, NOALLOC_HEADER, s) != cudaSuccess) { cudaGetLastError(); return KGM_E_CUDA; }
  if (n_docs) {
    int blocks = (int)((n_docs + TOK_THREADS / 32 - 1) / (TOK_THREADS / 32));
    blocks = blocks > 148 * 16 ? 148 * 16 : blocks;
    k_tokenize_gpu_noalloc<<<blocks, TOK_THREADS, 0, s>>>(gpu_csv, gpu_off, n_docs, scratch, max_rows,
                                                          doc_row0, doc_nrows, doc_err, util, mem_used,
                                                          mem_total);
    if (cudaGetLastError() != cudaSuccess) return KGM_E_CUDA;     // a launch that did not happen writes nothing
  }
  uint32_t head[2] = {0, 0};
  if (cudaMemcpyAsync(head, scratch, sizeof(head), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
      cudaStreamSynchronize(s) != cudaSuccess) {
    cudaGetLastError();
    return KGM_E_CUDA;
  }
  return head[1] ? KGM_E_LIMIT : KGM_OK;        // more rows in the text than max_rows
}
