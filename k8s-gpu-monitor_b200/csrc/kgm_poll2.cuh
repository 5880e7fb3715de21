// k_poll2: the whole poll path of one batch in ONE persistent kernel, one WARP per document
// (included by kgm_kernels.cu after kgm_fused.cuh, whose look-back and generic-path functions it
// uses).
//
// Documents are handed out in batch order by a ticket counter, 16 consecutive ones to a CTA at a
// time (one per warp; the warps of a CTA meet again before the next 16, so they run through the
// same code at about the same time -- the hot path is 128 KB of instructions).  A regular document
// lives in its warp's slice of shared memory (12 - 18 KB, 16 documents in flight per SM):
//   text windows  <- cp.async of the two nvidia-smi texts (16-byte units)     (gpu.py:104-116)
//   rows          <- SWAR line-feed scan; one lane per row: cells, digit runs, keys, repr(mem_util)
//   join          <- uuid -> short id against the gpu rows (key hashes by shuffle, bytes compared
//                    once), pid -> container probe of the HBM-resident table, container-major
//                    placement; an unknown uuid = the KeyError of gpu.py:151: node lines + status
//                                                                          (gpu.py:119-128,151,172-178)
//   records       <- per row the run words of its variable cells; per line: record, template
//                    slot, length
//   device-wide exclusive scan over documents by decoupled look-back (kgm_fused.cuh)
//   rounds        <- up to 32 lines of one kind at a time (pod rounds first), one lane per line: the
//                    unrolled composer of kgm_poll2_leaf.cuh writes the line through a pending word
//                    into the stage (template pieces from the slot pool in global memory as one
//                    prefetched stream of 16-byte loads, cells from the text windows, constant
//                    strings as immediates); the warp then streams the stage out with aligned
//                    16-byte streaming stores, realigned to the document's position in the output
//                                                                          (gpu.py:136-143,206-216,240-250)
// A document the fast path declines (any other raise point of the reference, unusual bytes,
// duplicate keys, sizes beyond the slice) is re-done from scratch by the generic per-byte code, out
// of line, and joins the same look-back chain: same bytes either way.
#pragma once

#include "kgm_poll2_leaf.cuh"
#include "kgm_window.cuh"

namespace kgm {
namespace {

using namespace p2;

// Shared memory holds 16 - 17 documents of the bench configurations, so the kernel is compiled for
// 16 warps and 128 registers (no spills); 20 warps x 96 registers measured the same or slower.
#ifndef KGM_P2_MAX_WARPS
#define KGM_P2_MAX_WARPS 16
#endif
constexpr int P2_MAX_THREADS = 32 * KGM_P2_MAX_WARPS;

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void *g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

struct P2Slice {                                     // byte offsets from the CTA's shared memory base
  uint32_t tg, ta, grec, prec, cnt, cstart, plan, loff, toff, stage, gstart, pstart, cp, digest;
};

// The 16-byte aligned window covering text[lo, hi) -> shared memory, asynchronously.
__device__ __forceinline__ void window_issue(uint32_t sdst, const uint8_t *text, uint64_t lo, uint64_t hi) {
  const uint64_t w0 = lo & ~15ull;
  const uint32_t n = (uint32_t)(((hi + 15ull) & ~15ull) - w0);
  for (uint32_t i = lane_id() * 16u; i < n; i += 512u) cp_async16(sdst + i, text + w0 + i);
}
// Bytes of the neighbouring documents (before s, after e) and the slack that word-wise readers
// may touch are zeroed: no line feed, comma or digit outside the text.
__device__ __forceinline__ void window_trim(uint8_t *dst, uint64_t lo, uint64_t hi, uint32_t *s, uint32_t *e) {
  *s = (uint32_t)(lo & 15ull);
  *e = *s + (uint32_t)(hi - lo);
  if (lane_id() < *s) dst[lane_id()] = 0;
  dst[*e + lane_id()] = 0;
  if (lane_id() < 16) dst[*e + 32u + lane_id()] = 0;
}

constexpr uint32_t CP_NONE = 0x7fffu;                // cp word: screened << 31 | crank << 16 | pos

// One regular document, planned and written by one warp.  Returns false (nothing published or
// written yet) if the generic path has to take it.  SAMPLES: the planning half only, feeding stage 3
// -- every line the document would emit becomes a (key, order, values) sample at wo->s[d * stride +
// line] instead of bytes (kgm_window.cu; same values as the staged front end's k_win_samples).
template <bool SAMPLES>
__device__ bool p2_fast_doc(uint32_t d, const BatchDev &b, const FleetDev &f, const WorkDev &w, const FusedParams &P,
                            const Poll2Params &Q, uint32_t *S, uint32_t s0, const P2Slice &sl, const WinFront *wo) {
  const uint32_t l = lane_id();
  // (SAMPLES) where the document's samples go, its window, and the short id of every process row
  const uint64_t sbase = SAMPLES ? (uint64_t)d * wo->stride : 0ull;
  const uint32_t tick = SAMPLES ? b.doc_tick[d] : 0u, window = SAMPLES ? tick / wo->window_ticks : 0u;
  uint8_t *sidv = reinterpret_cast<uint8_t *>(S) + sl.stage + Q.stage_bytes - Q.cap_np;
  auto put_sample = [&](uint32_t idx, uint64_t key, uint32_t util, uint32_t mem, uint32_t power) {
    const uint64_t i = sbase + idx;
    wo->s.key[i] = key;
    wo->s.order[i] = win_order(tick, d, idx);
    wo->s.val[0][i] = util; wo->s.val[1][i] = mem; wo->s.val[2][i] = power;
    wo->s.node[i] = b.doc_node[d];
  };
  uint8_t *Sb = reinterpret_cast<uint8_t *>(S);
  const uint32_t half = 0;                           // (stage clocks: see STAGE_MARK in kgm_fused.cuh)
  long long stage_t = 0;
#ifdef KGM_PROFILE_STAGES
  stage_t = clock64();
#endif
  (void)half;
  const uint32_t n = b.doc_node[d];
  if (!f.node_fast2[n]) return false;
  const uint64_t g_lo = b.gpu_off[d], g_hi = b.gpu_off[d + 1], a_lo = b.apps_off[d], a_hi = b.apps_off[d + 1];
  if (((g_hi + 15ull) & ~15ull) - (g_lo & ~15ull) + 48ull > Q.cap_g) return false;
  if (((a_hi + 15ull) & ~15ull) - (a_lo & ~15ull) + 48ull > Q.cap_a) return false;
  if (g_hi == g_lo || a_hi == a_lo) return false;
  const uint32_t cb = f.cont_begin[n], nc = f.cont_begin[n + 1] - cb;
  if (nc > Q.cap_c) return false;

  // ---- windows, rows ----
  window_issue(s0 + sl.tg, b.gpu_csv, g_lo, g_hi);
  window_issue(s0 + sl.ta, b.apps_csv, a_lo, a_hi);
  cp_async_wait_all();
  __syncwarp();
  uint32_t gs, ge, as, ae;
  window_trim(Sb + sl.tg, g_lo, g_hi, &gs, &ge);
  window_trim(Sb + sl.ta, a_lo, a_hi, &as, &ae);
  __syncwarp();
  const uint32_t *TG = S + (sl.tg >> 2), *TA = S + (sl.ta >> 2);
  uint16_t *gstart = reinterpret_cast<uint16_t *>(Sb + sl.gstart);
  uint16_t *pstart = reinterpret_cast<uint16_t *>(Sb + sl.pstart);
  uint32_t ng = 0, np = 0;
  if (!find_rows(TG, gs, ge, gstart, Q.cap_ng, &ng) || ng > 32u) return false;
  if (!find_rows(TA, as, ae, pstart, Q.cap_np, &np)) return false;
  STAGE_MARK(1);                                     // windows + rows

  // ---- gpu rows: one lane per row (gpu.py:104-109,119-143) ----
  uint32_t *grec = S + (sl.grec >> 2);
  uint64_t my_hash = 0;                              // lane j: the key hash of gpu row j
  {
    bool rok = true;
    if (l < ng) {
      G2 g;
      rok = parse_gpu_row2(TG, gstart[l], gstart[l + 1] - 1u, g);
      if (rok) {
        store_grec(S, sl.tg, sl.grec + GREC_BYTES * l, g, l);   // unique keys: the row index is the short id (gpu.py:119-120)
        my_hash = g.hash;
      }
      if (SAMPLES && rok) {                          // the node line's sample; the row's power for its process lines
        uint32_t power = NONE32;
        rok = g.util.len <= 9u && gpu_row_power(TG, gstart[l], gstart[l + 1] - 1u, &power);
        if (rok) {
          grec[GREC_WORDS * l + GREC_LINE_LEN] = power;
          put_sample(l, win_make_key(0u, n, l, window), g.util.len ? run_u32(TG, g.util.off, g.util.len) : NONE32,
                     g.used > (uint64_t)WIN_SAT ? WIN_SAT : (uint32_t)g.used, power);
        }
      }
    }
    bool dup = false;                                // equal keys: last row wins (gpu.py:54) -> generic path
    for (uint32_t j = 0; j + 1 < ng; ++j) {
      const uint64_t hj = __shfl_sync(FULL, my_hash, j);
      dup |= l < ng && j < l && hj == my_hash;
    }
    if (!__all_sync(FULL, rok) || __any_sync(FULL, dup)) return false;
  }
  STAGE_MARK(2);                                     // gpu rows
  uint16_t *cnt = reinterpret_cast<uint16_t *>(Sb + sl.cnt), *cstart = reinterpret_cast<uint16_t *>(Sb + sl.cstart);
  for (uint32_t c = l; c < nc; c += 32) cnt[c] = 0;
  __syncwarp();

  // ---- process rows: uuid -> short id (gpu.py:151), pid -> container (gpu.py:177-178), place
  //      inside the container ----
  uint32_t *prec = S + (sl.prec >> 2);
  uint32_t *cp = S + (sl.cp >> 2), *digest = S + (sl.digest >> 2);
  bool doc_key = false;
  for (uint32_t r0 = 0; r0 < np; r0 += 32) {
    const uint32_t r = r0 + l;
    const bool valid = r < np;
    bool ok = true;
    P2 p;
    p.pidkey = ~0ull;
    p.khash = 0;
    uint32_t nm = 0, c_first = NONE32, sid = NONE32, emit_c = NONE32, dg = 0;
    if (valid) {
      ok = parse_proc_row2(TA, pstart[r], pstart[r + 1] - 1u, p);
      if (ok && p.pidkey != ~0ull) pid_probe2(f, n, p.pidkey, &nm, &c_first);
    }
    // first gpu row with this key hash (the hashes come from the lanes that parsed the gpu rows);
    // length and bytes are compared once, afterwards
    for (uint32_t j = 0; j < ng; ++j) {
      const uint64_t hj = __shfl_sync(FULL, my_hash, j);
      if (sid == NONE32 && hj == p.khash) sid = j;
    }
    if (valid && ok && sid == NONE32) doc_key = true;  // no gpu row has this uuid: KeyError gpu.py:151 once every row is read
    if (valid && ok && sid != NONE32) {
      const uint32_t gk = grec[GREC_WORDS * sid + GREC_KEY];
      ok = rw_len(gk) == p.key.len && keys_equal(TA, p.key.off, S, rw_off(gk), p.key.len);   // (a hash collision declines)
      if (ok && p.pidkey != ~0ull) {
        if (nm > 1) ok = false;                      // a pid listed by several containers
        if (nm == 1 && c_first < nc && f.gated[cb + c_first]) emit_c = c_first;
        dg = pair_digest(p.pidkey, p.pid.len, sid);
      }
    }
    if (!__all_sync(FULL, ok)) return false;
    // duplicate (pid, gpu) keys keep the first row's place and the last row's values (gpu.py:54):
    // screened by digest, any hit goes to the generic path
    const bool screened = valid && p.pidkey != ~0ull && sid != NONE32;
    const uint32_t act = __ballot_sync(FULL, screened);
    bool dup = false;
    if (screened) {
      dup = __popc(__match_any_sync(act, dg)) > 1;
      for (uint32_t j = 0; j < r0 && !dup; ++j) dup = (cp[j] >> 31) && digest[j] == dg;
    }
    if (__any_sync(FULL, dup)) return false;
    // place inside the container: rank among the earlier rows of the same container
    const uint32_t emitting = __ballot_sync(FULL, emit_c != NONE32);
    uint32_t pos = 0;
    if (emit_c != NONE32) {
      const uint32_t peers = __match_any_sync(emitting, emit_c);
      const int leader = __ffs(peers) - 1;
      uint32_t base = 0;
      if ((int)l == leader) { base = cnt[emit_c]; cnt[emit_c] = (uint16_t)(base + __popc(peers)); }
      base = __shfl_sync(peers, base, leader);
      pos = base + __popc(peers & lanemask_lt());
    }
    if (valid) {
      if (sid != NONE32) store_prec(S, sl.ta, sl.prec + PREC_BYTES * r, p, grec + GREC_WORDS * sid);
      if (SAMPLES) sidv[r] = (uint8_t)sid;
      cp[r] = (screened ? 0x80000000u : 0u) | ((emit_c == NONE32 ? CP_NONE : emit_c) << 16) | pos;
      digest[r] = dg;
    }
    __syncwarp();
  }

  STAGE_MARK(3);                                     // process rows
  // A process row that names no gpu of this poll: every table has been read and the node lines have
  // gone out when gpu.py:151 raises, so the document is its node lines and the KeyError.
  doc_key = __any_sync(FULL, doc_key);
  // ---- containers in docker stats order: busy lines, or idle lines (gpu.py:174-253) ----
  uint32_t total_pod = 0, total_idle = 0;
  for (uint32_t c0 = 0; c0 < (doc_key ? 0u : nc); c0 += 32) {
    const uint32_t c = c0 + l;
    uint32_t lines = 0, idle = 0;
    if (c < nc && f.gated[cb + c]) {
      const uint32_t k = cnt[c];
      if (k) lines = k;
      else lines = idle = f.vis_begin[cb + c + 1] - f.vis_begin[cb + c];
    }
    const uint32_t incl = warp_incl_scan(lines), iincl = warp_incl_scan(idle);
    if (c < nc) {
      cstart[c] = (uint16_t)(total_pod + incl - lines);
      if (idle) cnt[c] = (uint16_t)(0x8000u | (total_idle + iincl - idle));    // first idle pseudo-row of the container
    }
    total_pod += __shfl_sync(FULL, incl, 31);
    total_idle += __shfl_sync(FULL, iincl, 31);
  }
  const uint32_t nl = ng + total_pod;
  if (nl > Q.cap_lines || np + total_idle > Q.cap_rec || total_pod >= 0x8000u) return false;
  __syncwarp();

  // ---- per line: record, slot, length, variable bytes in front of the job boundaries ----
  uint16_t *plan = reinterpret_cast<uint16_t *>(Sb + sl.plan);
  uint32_t *loff = S + (sl.loff >> 2), *toff = S + (sl.toff >> 2);
  bool too_long = false;
  if (l < ng) {                                      // node lines (gpu.py:132-146)
    plan[l] = (uint16_t)((KIND_NODE << 14) | l);
    toff[l] = f.nslot_off[n];
    loff[l] = f.node_fixed[n] + 1u + line_var_bytes<false>(grec + GREC_WORDS * l);
  }
  for (uint32_t c = l; c < (doc_key ? 0u : nc); c += 32) {   // idle containers: one line per visible uuid (gpu.py:221-250)
    const uint32_t cw = cnt[c];
    if (!(cw & 0x8000u)) continue;
    const uint32_t gc = cb + c, nvis = f.vis_begin[gc + 1] - f.vis_begin[gc];
    for (uint32_t v = 0; v < nvis; ++v) {
      const uint32_t gv = f.vis_begin[gc] + v;
      const uint64_t vh = f.vis_hash2[gv];
      const uint32_t vo = f.vis_off[gv], vlen = f.vis_off[gv + 1] - vo;
      uint32_t sid = NONE32;
      for (uint32_t j = 0; j < ng && sid == NONE32; ++j) {
        const uint32_t *gr = grec + GREC_WORDS * j;
        if (gr[GREC_HASH] == (uint32_t)vh && gr[GREC_HASH + 1] == (uint32_t)(vh >> 32) && rw_len(gr[GREC_KEY]) == vlen) {
          bool eq = true;
          for (uint32_t i = 0; i < vlen; ++i) eq &= f.vis_bytes[vo + i] == Sb[rw_off(gr[GREC_KEY]) + i];
          if (eq) sid = j;
        }
      }
      const uint32_t ir = np + (cw & 0x7fffu) + v, idx = ng + cstart[c] + v;
      store_idle_rec(S, 0u, sl.prec + PREC_BYTES * ir, sid);
      if (SAMPLES)                                   // gpu.py:234-250: gpu_util=0, used_memory=0
        put_sample(idx, win_make_key(1u, gc, sid == NONE32 ? 0xffffu : sid, window), 0u, 0u, NONE32);
      plan[idx] = (uint16_t)((KIND_IDLE << 14) | ir);
      toff[idx] = f.tslot_off[gc];
      loff[idx] = f.cont_fixed[gc] + 1u + line_var_bytes<true>(prec + PREC_WORDS * ir);
    }
  }
  for (uint32_t r = l; r < (doc_key ? 0u : np); r += 32) {   // busy lines (gpu.py:190-219)
    const uint32_t c = (cp[r] >> 16) & 0x7fffu;
    if (c == CP_NONE) continue;
    const uint32_t gc = cb + c, idx = ng + cstart[c] + (cp[r] & 0xffffu);
    if (SAMPLES) {
      const uint32_t sid = sidv[r], *gr = grec + GREC_WORDS * sid, mw = prec[PREC_WORDS * r + PREC_MEM];
      const uint32_t uw = gr[GREC_UTIL];
      too_long |= rw_len(mw) > 9u;                   // (declines: the staged front end's saturation rules decide)
      put_sample(idx, win_make_key(1u, gc, sid, window), rw_len(uw) ? run_u32(S, rw_off(uw), rw_len(uw)) : NONE32,
                 rw_len(mw) && rw_len(mw) <= 9u ? run_u32(S, rw_off(mw), rw_len(mw)) : NONE32, gr[GREC_LINE_LEN]);
    }
    plan[idx] = (uint16_t)((KIND_BUSY << 14) | r);
    toff[idx] = f.tslot_off[gc];
    loff[idx] = f.cont_fixed[gc] + 1u + line_var_bytes<true>(prec + PREC_WORDS * r);
  }
  __syncwarp();
  // lengths -> offsets inside the document
  uint32_t nb = 0;
  for (uint32_t i0 = 0; i0 < nl; i0 += 32) {
    const uint32_t i = i0 + l;
    const uint32_t len = i < nl ? loff[i] : 0u;
    too_long |= len + 64u > Q.stage_bytes;
    const uint32_t incl = warp_incl_scan(len);
    if (i < nl) loff[i] = nb + incl - len;
    nb += __shfl_sync(FULL, incl, 31);
  }
  if (__any_sync(FULL, too_long)) return false;
  if (l == 0) loff[nl] = nb;
  __syncwarp();
  if (SAMPLES) {                                     // the planning half ends here: unused places are invalid
    for (uint32_t i = nl + l; i < wo->stride; i += 32) wo->s.key[sbase + i] = WIN_KEY_INVALID;
    return true;
  }

  STAGE_MARK(4);                                     // containers, lines, lengths
  // ---- from here on the document is this warp's: publish its size, write its lines ----
  scan_publish(P, d, b.n_docs, nl, nb);
  STAGE_MARK(5);                                     // publish (block aggregators: + their look-back)

  uint64_t line0 = 0, byte0 = 0;
  bool resolved = false, fits_out = true;
  // The rounds start with the pod lines and end with the node lines (the output order is the
  // documents' either way): the look-back is resolved after the first round has been composed,
  // and the longer that takes, the fewer documents wait for their predecessors there.
  uint32_t i0 = (Q.pods_first && nl > ng) ? ng : 0u, left = nl;
  do {
    // a round: up to 32 lines of one kind whose bytes fit the stage
    uint32_t i1 = min(i0 < ng ? ng : nl, i0 + 32u), r_lo = 0;
    if (nl) {
      r_lo = loff[i0];
      while (i1 > i0 + 1 && loff[i1] - r_lo + 64u > Q.stage_bytes) --i1;
      STAGE_MARK(6);
      const uint32_t i = i0 + l;
      LineWriter lw;
      lw.fill = 0;
      if (i < i1) {
        const uint32_t pl = plan[i], pod = (pl >> 14) != KIND_NODE ? 1u : 0u, ref = pl & 0x3fffu;
        lw = compose_line(S, 0u, pod, f.slot_pool + 16ull * toff[i],
                          pod ? sl.prec + PREC_BYTES * ref : sl.grec + GREC_BYTES * ref, sl.stage + (loff[i] - r_lo));
      }
      __syncwarp();
      if (i < i1) lw_finish(lw, i + 1 < i1);
      __syncwarp();
      STAGE_MARK(7);                                 // compose
    }
    if (!resolved) {
      scan_resolve(P, d, nl, nb, &line0, &byte0);
      if (P.carry) { line0 += P.carry[0]; byte0 += P.carry[1]; }
      if (l == 0) {
        w.doc_line_off[d] = line0;
        w.doc_byte_off[d] = byte0;
        w.doc_status[d] = doc_key ? (int32_t)DOC_KEY : 0; w.doc_nlines[d] = nl; w.doc_nbytes[d] = nb;
        if (d + 1 == b.n_docs) {
          w.doc_line_off[b.n_docs] = line0 + nl;
          w.doc_byte_off[b.n_docs] = byte0 + nb;
          w.totals[0] = line0 + nl;
          w.totals[1] = byte0 + nb;
          post_totals(P, line0 + nl, byte0 + nb);
          write_out_header(w, line0 + nl, byte0 + nb);
          w.line_off[min(line0 + nl, (uint64_t)w.desc_cap)] = byte0 + nb;
        }
        if (byte0 + nb > w.out_cap || line0 + nl > w.desc_cap)
          atomicOr(w.counters + CTR_OVERFLOW, (byte0 + nb > w.out_cap ? OVF_OUT : 0u) |
                                                  (line0 + nl > w.desc_cap ? OVF_DESC : 0u));
      }
      resolved = true;
      fits_out = byte0 + nb <= w.out_cap && line0 + nl <= w.desc_cap;
      STAGE_MARK(8);                                 // look-back incl. waiting
    }
    if (!fits_out || nl == 0) break;
    if (i0 + l < i1) w.line_off[line0 + i0 + l] = byte0 + loff[i0 + l];
    // stage -> global
    const uint32_t nbytes = loff[i1] - r_lo;
    uint8_t *g = w.out + byte0 + r_lo;
    const uint8_t *st = Sb + sl.stage;
    const uint32_t head = min(nbytes, (uint32_t)((16u - ((byte0 + r_lo) & 15ull)) & 15u));
    if (l < head) g[l] = st[l];
    const uint32_t nvec = (nbytes - head) >> 4;
    {
      const uint32_t *sw = S + ((sl.stage + head) >> 2);     // stage word holding byte `head`
      const uint32_t sh = 8u * (head & 3u);
      for (uint32_t v = l; v < nvec; v += 32u) {
        const uint32_t *q = sw + 4u * v;
        const uint32_t a0 = q[0], a1 = q[1], a2 = q[2], a3 = q[3], a4 = q[4];
        uint4 o;
        o.x = __funnelshift_r(a0, a1, sh); o.y = __funnelshift_r(a1, a2, sh);
        o.z = __funnelshift_r(a2, a3, sh); o.w = __funnelshift_r(a3, a4, sh);
        __stcs(reinterpret_cast<uint4 *>(g + head + 16u * v), o);      // written once, read by the copy engine: streaming
      }
    }
    const uint32_t done = head + 16u * nvec;
    if (l < nbytes - done) g[done + l] = st[done + l];
    left -= i1 - i0;
    i0 = i1 == nl ? 0u : i1;
    __syncwarp();                                    // the stage may be overwritten
    STAGE_MARK(9);                                   // copy-out
  } while (left);
  return true;
}

// The generic path for one document: the per-byte device functions of the staged kernels, one
// warp, token tables and line plan in the global workspace.
__device__ __noinline__ void p2_slow_doc(uint32_t d, const BatchDev &b, const FleetDev &f, const WorkDev &w,
                                         const FusedParams &P, uint32_t *scratch) {
  const uint32_t l = lane_id();
  uint32_t nl = 0, nb = 0;
  slow_plan(d, b, f, w, scratch, &nl, &nb);
  if (l == 0) atomicAdd(w.counters + CTR_SLOW, 1u);
  scan_publish(P, d, b.n_docs, nl, nb);
  uint64_t line0, byte0;
  scan_resolve(P, d, nl, nb, &line0, &byte0);
  if (P.carry) { line0 += P.carry[0]; byte0 += P.carry[1]; }
  if (l == 0) {
    w.doc_line_off[d] = line0;
    w.doc_byte_off[d] = byte0;
    if (d + 1 == b.n_docs) {
      w.doc_line_off[b.n_docs] = line0 + nl;
      w.doc_byte_off[b.n_docs] = byte0 + nb;
      w.totals[0] = line0 + nl;
      w.totals[1] = byte0 + nb;
      post_totals(P, line0 + nl, byte0 + nb);
      write_out_header(w, line0 + nl, byte0 + nb);
      w.line_off[min(line0 + nl, (uint64_t)w.desc_cap)] = byte0 + nb;
    }
  }
  if (byte0 + nb > w.out_cap || line0 + nl > w.desc_cap) {
    if (l == 0) atomicOr(w.counters + CTR_OVERFLOW, (byte0 + nb > w.out_cap ? OVF_OUT : 0u) |
                                                     (line0 + nl > w.desc_cap ? OVF_DESC : 0u));
  } else {
    slow_emit(d, b, f, w, nl, line0, byte0);
  }
  __syncwarp();
}

// (the parameters are __grid_constant__: the generic path takes them by reference, which would
// otherwise make every thread copy them to its local memory)
__global__ void __launch_bounds__(P2_MAX_THREADS, 1)
k_poll2(const __grid_constant__ BatchDev b, const __grid_constant__ FleetDev f, const __grid_constant__ WorkDev w,
        const __grid_constant__ FusedParams P, const __grid_constant__ Poll2Params Q) {
  extern __shared__ __align__(16) uint32_t S[];
  // constant strings and tables: the image kgm_load_fleet built (same for every CTA)
  for (uint32_t i = threadIdx.x; i < ((CA2::BYTES + 15u) & ~15u) / 16; i += blockDim.x)
    reinterpret_cast<uint4 *>(S)[i] = __ldg(f.ca2 + i);
  // the counters of the NEXT launch start from zero (this launch uses the other set)
  if (blockIdx.x == 0 && threadIdx.x < CTR_COUNT) P.next_counters[threadIdx.x] = 0;
  __syncthreads();

  const uint32_t warp = threadIdx.x >> 5, l = lane_id();
  uint32_t base = ((CA2::BYTES + 15u) & ~15u) + warp * Q.slice_bytes;
  asm volatile("" : "+r"(base));                     // one register, not recomputed from the warp index at every use
  P2Slice sl;
  sl.tg = base + Q.off_tg; sl.ta = base + Q.off_ta; sl.grec = base + Q.off_grec; sl.prec = base + Q.off_prec;
  sl.cnt = base + Q.off_cnt; sl.cstart = base + Q.off_cstart; sl.plan = base + Q.off_plan; sl.loff = base + Q.off_loff;
  sl.toff = base + Q.off_toff; sl.stage = base + Q.off_stage;
  sl.gstart = base + Q.off_gstart; sl.pstart = base + Q.off_pstart; sl.cp = base + Q.off_cp; sl.digest = base + Q.off_digest;
  const uint32_t s0 = (uint32_t)__cvta_generic_to_shared(S);

  __shared__ unsigned long long cta_base;
  for (;;) {
    unsigned long long t = 0;
    if (Q.cta_tickets) {
      // the warps of a CTA start their documents together, `warps` consecutive ones at a time, so
      // that they run through the same code at about the same time
      __syncthreads();
      if (threadIdx.x == 0) cta_base = atomicAdd(P.ticket, (unsigned long long)Q.warps) - P.ticket_base;
      __syncthreads();
      if (cta_base >= b.n_docs) break;
      t = cta_base + warp;
      if (t >= b.n_docs) continue;
    } else {
      if (l == 0) t = atomicAdd(P.ticket, 1ull) - P.ticket_base;
      t = __shfl_sync(FULL, t, 0);
      if (t >= b.n_docs) break;
    }
    const uint32_t d = (uint32_t)t;
    if (P.allow_fast && p2_fast_doc<false>(d, b, f, w, P, Q, S, s0, sl, nullptr)) continue;
    cp_async_wait_all();
    __syncwarp();
    p2_slow_doc(d, b, f, w, P, S + (sl.stage >> 2));
  }
}

// Stage 3's front end: the planning half of k_poll2 over every document, samples instead of bytes
// (documents by a fixed stride, no ticket and no look-back: nothing is ordered here).  A document
// the fast path declines is counted; the host then lets the staged front end take the batch.
__global__ void __launch_bounds__(P2_MAX_THREADS, 1)
k_poll2_samples(const __grid_constant__ BatchDev b, const __grid_constant__ FleetDev f, const __grid_constant__ WorkDev w,
                const __grid_constant__ FusedParams P, const __grid_constant__ Poll2Params Q,
                const __grid_constant__ WinFront wo) {
  extern __shared__ __align__(16) uint32_t S[];
  for (uint32_t i = threadIdx.x; i < ((CA2::BYTES + 15u) & ~15u) / 16; i += blockDim.x)
    reinterpret_cast<uint4 *>(S)[i] = __ldg(f.ca2 + i);
  __syncthreads();
  const uint32_t warp = threadIdx.x >> 5, l = lane_id();
  const uint32_t base = ((CA2::BYTES + 15u) & ~15u) + warp * Q.slice_bytes;
  P2Slice sl;
  sl.tg = base + Q.off_tg; sl.ta = base + Q.off_ta; sl.grec = base + Q.off_grec; sl.prec = base + Q.off_prec;
  sl.cnt = base + Q.off_cnt; sl.cstart = base + Q.off_cstart; sl.plan = base + Q.off_plan; sl.loff = base + Q.off_loff;
  sl.toff = base + Q.off_toff; sl.stage = base + Q.off_stage;
  sl.gstart = base + Q.off_gstart; sl.pstart = base + Q.off_pstart; sl.cp = base + Q.off_cp; sl.digest = base + Q.off_digest;
  const uint32_t s0 = (uint32_t)__cvta_generic_to_shared(S);
  for (uint32_t d = blockIdx.x * Q.warps + warp; d < b.n_docs; d += gridDim.x * Q.warps) {
    if (p2_fast_doc<true>(d, b, f, w, P, Q, S, s0, sl, &wo)) continue;
    cp_async_wait_all();
    __syncwarp();
    for (uint32_t i = l; i < wo.stride; i += 32) wo.s.key[(uint64_t)d * wo.stride + i] = WIN_KEY_INVALID;
    if (l == 0) atomicAdd(wo.declined, 1u);
  }
}

}  // namespace

void launch_poll2_samples(const BatchDev &b, const FleetDev &f, const WorkDev &w, const FusedParams &P,
                          const Poll2Params &Q, const WinFront &wo, int blocks, int device, cudaStream_t s) {
  static int attr_bytes[64];                         // the opt-in is per device
  const int need = poll2_smem_bytes(Q);
  if (device < 0 || device >= 64 || need > attr_bytes[device]) {
    cudaFuncSetAttribute(k_poll2_samples, cudaFuncAttributeMaxDynamicSharedMemorySize, need);
    if (device >= 0 && device < 64) attr_bytes[device] = need;
  }
  k_poll2_samples<<<blocks, 32 * Q.warps, need, s>>>(b, f, w, P, Q, wo);
}

int poll2_max_warps() { return KGM_P2_MAX_WARPS; }
int poll2_smem_bytes(const Poll2Params &Q) { return (int)(((CA2::BYTES + 15u) & ~15u) + Q.warps * Q.slice_bytes); }

void launch_poll2(const BatchDev &b, const FleetDev &f, const WorkDev &w, const FusedParams &P, const Poll2Params &Q,
                  int blocks, int device, cudaStream_t s) {
  static int attr_bytes[64];                         // the opt-in is per device
  const int need = poll2_smem_bytes(Q);
  if (device < 0 || device >= 64 || need > attr_bytes[device]) {
    cudaFuncSetAttribute(k_poll2, cudaFuncAttributeMaxDynamicSharedMemorySize, need);
    if (device >= 0 && device < 64) attr_bytes[device] = need;
  }
  k_poll2<<<blocks, 32 * Q.warps, need, s>>>(b, f, w, P, Q);
}

}  // namespace kgm
