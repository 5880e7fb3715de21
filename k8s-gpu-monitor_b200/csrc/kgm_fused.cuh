// k_poll: the whole poll path of one batch in ONE persistent kernel (included by
// kgm_kernels.cu, after the staged kernels whose device functions it reuses).
//
// A team of two warps owns one document at a time (documents are handed out in batch order
// by a ticket counter).  For a regular document everything happens in the team's slice of
// shared memory:
//   text windows  <- coalesced 16-byte loads of the two nvidia-smi texts      (gpu.py:104-116)
//   rows          <- SWAR line-feed scan, one lane per row: cells, digit runs, keys
//   join          <- uuid dictionary in registers/shared memory, pid -> container probe of
//                    the HBM-resident table, container-major placement     (gpu.py:119-128,151,172-178)
//   lengths       <- per line, warp scan -> offsets inside the document
//   device-wide exclusive scan over documents by decoupled look-back (lines and bytes)
//   bytes         <- one lane per line appends its segments with aligned 32-bit stores into
//                    the stage, the warp streams the stage out with 16-byte stores
//                                                                          (gpu.py:136-143,206-216,240-250)
// so the texts are read from HBM once and the line protocol is written once, with no
// intermediate tables.  A document the fast path cannot take (any raise point of the
// reference, unusual bytes, duplicate keys, sizes beyond the team's slice) is re-done from
// scratch by the generic per-byte code of the staged pipeline by one warp of the team, and
// joins the same look-back chain -- the output is identical either way.
#pragma once

namespace kgm {
namespace {

using namespace fast;

constexpr int POLL_THREADS = 512;                   // at most 8 teams of 2 warps per CTA (128 registers per thread)
constexpr uint32_t CONST_PAD = (ConstArea::BYTES + 15u) & ~15u;
constexpr uint32_t KIND_NODE = 0, KIND_BUSY = 1, KIND_IDLE = 2;
constexpr uint32_t IDLE_NONE = 0xfffu;

// ---- optional stage clocks (build with -DKGM_PROFILE_STAGES; read with kgm_debug_stage_clocks) --
#ifdef KGM_PROFILE_STAGES
__device__ unsigned long long g_stage_clk[64];       // [half * 32 + stage]: SM clocks summed over documents
#define STAGE_MARK(id)                                                                      \
  do {                                                                                      \
    if (lane_id() == 0) {                                                                   \
      const long long now__ = clock64();                                                    \
      atomicAdd(&g_stage_clk[half * 32 + (id)], (unsigned long long)(now__ - stage_t));     \
      stage_t = now__;                                                                      \
    }                                                                                       \
  } while (0)
#else
#define STAGE_MARK(id) do { (void)stage_t; } while (0)
#endif

// ---- look-back state: two tagged 64-bit words per document -----------------------------------
// tag = epoch (16 bits) | status (2 bits); value in the low 46 bits
constexpr uint64_t ST_AGG = 1, ST_INC = 2;
__device__ __forceinline__ uint64_t st_tag(uint32_t epoch, uint64_t status) {
  return ((uint64_t)(epoch & 0xffffu) << 48) | (status << 46);
}
__device__ __forceinline__ void st_relaxed(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_release(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ uint64_t warp_sum64(uint64_t v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

// Device-wide exclusive scan over documents: decoupled look-back on two levels.  Level 1
// holds every document's (lines, bytes) as an aggregate; level 2 holds, per block of 32
// consecutive documents, the block's aggregate and then its inclusive prefix.  The team
// that owns the last document of a block sums the block and looks back over the BLOCK
// entries (one 32-wide window covers 1024 documents, more than are ever in flight), so a
// block's inclusive prefix appears a few memory round trips after the block's documents
// have been planned.  A document's exclusive prefix is then one window of level 1 (its own
// block) plus one level-2 word pair.  All waits are on values that teams with EARLIER
// tickets publish right after planning, never on their later stages, so the chain cannot
// deadlock; scan_resolve is called after the first round of lines has been composed, when
// those values have normally long been there.
struct ScanWords { uint64_t l, b; bool inc; };

__device__ __forceinline__ ScanWords scan_wait(const unsigned long long *st, uint64_t agg, uint64_t inc,
                                               bool need_inc) {
  const uint64_t VAL = (1ull << 46) - 1ull;
  // a waiting lane shares its scheduler with warps that have work: back off quickly
  for (uint32_t ns = 20;; ns = min(ns + 20u, 200u)) {
    const uint64_t w0 = ld_acquire(st);
    const uint64_t t0 = w0 & ~VAL;
    if (t0 == inc || (!need_inc && t0 == agg)) {
      const uint64_t w1 = ld_relaxed(st + 1);
      if ((w1 & ~VAL) == t0) return ScanWords{w0 & VAL, w1 & VAL, t0 == inc};
    }
    __nanosleep(ns);
  }
}
__device__ __forceinline__ void scan_store(unsigned long long *st, uint64_t tag, uint64_t l, uint64_t b) {
  st_relaxed(st + 1, tag | b);
  st_release(st, tag | l);
}

// Every lane of the (half 0) warp calls this right after the document has been planned.
__device__ void scan_publish(const FusedParams &P, uint32_t d, uint32_t n_docs, uint64_t nl, uint64_t nb) {
  const uint32_t l = lane_id();
  const uint64_t agg = st_tag(P.epoch, ST_AGG), inc = st_tag(P.epoch, ST_INC);
  if (l == 0) scan_store(P.scan_state + 2ull * d, agg, nl, nb);
  if ((d & 31u) != 31u) return;
  (void)n_docs;
  // last document of block k: the block's aggregate, then its inclusive prefix
  const uint32_t k = d >> 5;
  uint64_t vl = nl, vb = nb;
  if (l != 31u) {
    const ScanWords w = scan_wait(P.scan_state + 2ull * (32u * k + l), agg, inc, false);
    vl = w.l; vb = w.b;
  }
  const uint64_t bl = warp_sum64(vl), bb = warp_sum64(vb);
  unsigned long long *blk = P.scan_blocks;
  if (l == 0) scan_store(blk + 2ull * k, agg, bl, bb);
  uint64_t ex_l = 0, ex_b = 0;
  long long j = (long long)k - 1;
  while (j >= 0) {
    const long long idx = j - (long long)l;
    uint64_t xl = 0, xb = 0;
    bool is_inc = true;                              // before block 0: prefix 0
    if (idx >= 0) {
      const ScanWords w = scan_wait(blk + 2ull * idx, agg, inc, false);
      xl = w.l; xb = w.b; is_inc = w.inc;
    }
    const uint32_t incs = __ballot_sync(FULL, is_inc);
    const uint32_t first = incs ? (uint32_t)__ffs(incs) - 1u : 32u;
    if (l > first) { xl = 0; xb = 0; }
    ex_l += warp_sum64(xl);
    ex_b += warp_sum64(xb);
    if (incs) break;
    j -= 32;
  }
  if (l == 0) scan_store(blk + 2ull * k, inc, ex_l + bl, ex_b + bb);
}

__device__ void scan_resolve(const FusedParams &P, uint32_t d, uint64_t nl, uint64_t nb, uint64_t *line0,
                             uint64_t *byte0) {
  (void)nl; (void)nb;
  const uint32_t l = lane_id();
  const uint64_t agg = st_tag(P.epoch, ST_AGG), inc = st_tag(P.epoch, ST_INC);
  const uint32_t k = d >> 5, p = d & 31u;
  uint64_t vl = 0, vb = 0;
  if (l < p) {                                       // earlier documents of the same block
    const ScanWords w = scan_wait(P.scan_state + 2ull * (32u * k + l), agg, inc, false);
    vl = w.l; vb = w.b;
  } else if (l == 31u && k > 0) {                    // everything before the block
    const ScanWords w = scan_wait(P.scan_blocks + 2ull * (k - 1u), agg, inc, true);
    vl = w.l; vb = w.b;
  }
  *line0 = warp_sum64(vl);
  *byte0 = warp_sum64(vb);
}

// ---- fast path -------------------------------------------------------------------------------
struct WarpSlice {                                   // byte offsets from the CTA's shared memory base
  uint32_t tg, ta, gstart, pstart, grec, prec, cnt, cstart, plan, loff, tb, stage, mbox;
};

// Coalesced copy of the 16-byte aligned window that covers text[lo, hi) into shared memory.
// The chunked end-to-end path: the host waits for the totals of a launch in mapped host
// memory (no copy engine, no stream synchronisation) so that it can queue the copy of this
// launch's lines while the previous copy is still running.  Each value is one 8-byte store
// carrying the launch's 16-bit tag, so no system-scope fence is needed (such a fence would
// wait behind the result copy that saturates the same PCIe direction).
__device__ __forceinline__ void post_totals(const FusedParams &P, uint64_t lines, uint64_t bytes) {
  if (!P.post) return;
  volatile unsigned long long *q = P.post;
  q[0] = (P.post_seq << 48) | lines;
  q[1] = (P.post_seq << 48) | bytes;
}

__device__ __forceinline__ void load_window(const uint8_t *text, uint64_t lo, uint64_t hi, uint8_t *dst,
                                            uint32_t *s, uint32_t *e) {
  const uint64_t w0 = lo & ~15ull;
  const uint32_t n = (uint32_t)(((hi + 15ull) & ~15ull) - w0);
  for (uint32_t i = lane_id() * 16u; i < n; i += 512u)
    *reinterpret_cast<uint4 *>(dst + i) = __ldg(reinterpret_cast<const uint4 *>(text + w0 + i));
  *s = (uint32_t)(lo - w0);
  *e = *s + (uint32_t)(hi - lo);
  __syncwarp();
  // bytes of the neighbouring documents (before s, after e) and the slack that word-wise
  // readers may touch are zeroed: no line feed, comma or digit outside the text
  if (lane_id() < *s) dst[lane_id()] = 0;
  dst[*e + lane_id()] = 0;
  if (lane_id() < 16) dst[*e + 32u + lane_id()] = 0;
}

// Line feeds of the window bytes [s, e) -> row_start[0..nrows]; row r = [row_start[r],
// row_start[r+1]-1).  The first line is the header (gpu.py:67-68).  False if the text has no
// line feed, does not end with one, or holds more than `cap` rows.
__device__ bool find_rows(const uint32_t *T, uint32_t s, uint32_t e, uint16_t *row_start, uint32_t cap,
                          uint32_t *nrows) {
  const uint32_t l = lane_id();
  uint32_t running = 0;
  for (uint32_t it = s & ~31u; it < e; it += 1024u) {
    const uint32_t pos = it + 32u * l;              // this lane's 32 bytes = 8 words
    uint32_t m[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) m[k] = 0;
    if (pos < e) {
      const uint4 v0 = *reinterpret_cast<const uint4 *>(T + (pos >> 2));
      const uint4 v1 = *reinterpret_cast<const uint4 *>(T + (pos >> 2) + 4);
      const uint32_t w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t z = w[k] ^ 0x0a0a0a0au;                         // exact zero-byte test
        m[k] = ~(((z & 0x7f7f7f7fu) + 0x7f7f7f7fu) | z) & 0x80808080u;
      }
    }
    uint32_t bits = 0;                              // bit i = byte pos+i is a line feed
#pragma unroll
    for (int k = 0; k < 8; ++k) bits |= pack4(m[k]) << (4 * k);
    const uint32_t c = __popc(bits);
    const uint32_t incl = warp_incl_scan(c);
    uint32_t idx = running + incl - c;
    while (bits) {
      const uint32_t bit = __ffs(bits) - 1;
      bits &= bits - 1;
      if (idx <= cap) row_start[idx] = (uint16_t)(pos + bit + 1u);
      ++idx;
    }
    running += __shfl_sync(FULL, incl, 31);
  }
  __syncwarp();
  if (running == 0 || running - 1 > cap) return false;
  if (row_start[running - 1] != e) return false;                        // last row without terminator
  *nrows = running - 1;
  return true;
}

// ---- a document is worked on by a TEAM of two warps -----------------------------------------
// half 0 owns the --query-gpu side (window, rows, float formatting, template block, container
// scan, look-back), half 1 the --query-compute-apps side (window, rows, uuid / pid lookups);
// both compose half of the lines and copy half of the stage out.  The halves meet at named
// barriers; what they tell each other goes through a small mailbox in the team's slice.
// Every decision that ends the fast path is taken by both halves from the same mailbox
// words after a barrier, so the two warps always leave together.
struct Mailbox {
  uint32_t ok0, ok1;            // the half's verdict of the current stage (1 = regular so far)
  uint32_t ng, np;
  uint32_t nl, nb;
  uint32_t total0;              // bytes of the lines half 0 measured
  uint32_t okf;                 // half 0: the floats formatted and the node lines fit
  unsigned long long line0, byte0;
  unsigned long long ticket;
};

__device__ __forceinline__ void team_sync(uint32_t team) {
  asm volatile("bar.sync %0, 64;" ::"r"(team + 1u) : "memory");
}

// pid -> container: how many containers of node n list this pid string, and the first one
__device__ __forceinline__ void pid_probe(const FleetDev &f, uint32_t n, uint64_t pidkey, uint32_t *nm,
                                          uint32_t *c_first) {
  uint32_t h = pid16_hash(pidkey, n) & f.pid16_mask;
  for (;;) {
    const uint4 raw = __ldg(reinterpret_cast<const uint4 *>(f.pid16 + h));
    if (raw.z == NONE32) break;
    const uint64_t key = ((uint64_t)raw.y << 32) | raw.x;
    if (raw.z == n && key == pidkey) { if (!*nm) *c_first = raw.w; ++*nm; }
    h = (h + 1) & f.pid16_mask;
  }
}
// The same, two slots per round trip: at a load of at most 0.5 a probe is usually one hit and the
// empty slot behind it (k_poll2; in k_poll the second load in flight costs more than it saves).
__device__ __forceinline__ void pid_probe2(const FleetDev &f, uint32_t n, uint64_t pidkey, uint32_t *nm,
                                           uint32_t *c_first) {
  uint32_t h = pid16_hash(pidkey, n) & f.pid16_mask;
  for (;;) {
    const uint32_t h1 = (h + 1) & f.pid16_mask;
    const uint4 raw = __ldg(reinterpret_cast<const uint4 *>(f.pid16 + h));
    const uint4 nxt = __ldg(reinterpret_cast<const uint4 *>(f.pid16 + h1));
    if (raw.z == NONE32) break;
    if (raw.z == n && (((uint64_t)raw.y << 32) | raw.x) == pidkey) { if (!*nm) *c_first = raw.w; ++*nm; }
    if (nxt.z == NONE32) break;
    if (nxt.z == n && (((uint64_t)nxt.y << 32) | nxt.x) == pidkey) { if (!*nm) *c_first = nxt.w; ++*nm; }
    h = (h1 + 1) & f.pid16_mask;
  }
}

// Plans a regular document.  On success the team's slice holds the row records, the line
// plan and the in-document offsets, and the mailbox its line and byte counts.
__device__ bool fast_plan(uint32_t d, const BatchDev &b, const FleetDev &f, const FusedParams &P, uint32_t *S,
                          uint8_t *Sb, const WarpSlice &ws, uint32_t team, uint32_t half, long long &stage_t) {
  const uint32_t l = lane_id();
  volatile Mailbox *mb = reinterpret_cast<volatile Mailbox *>(Sb + ws.mbox);
  const uint32_t n = b.doc_node[d];
  if (!f.node_fast[n]) return false;                 // (both halves read the same words: same verdict)
  const uint64_t g_lo = b.gpu_off[d], g_hi = b.gpu_off[d + 1], a_lo = b.apps_off[d], a_hi = b.apps_off[d + 1];
  const uint32_t tb_len = f.tb_len[n];
  if (((g_hi + 15ull) & ~15ull) - (g_lo & ~15ull) + 32ull > P.cap_g) return false;
  if (((a_hi + 15ull) & ~15ull) - (a_lo & ~15ull) + 32ull > P.cap_a) return false;
  if (tb_len + 16u > P.cap_t || g_hi == g_lo || a_hi == a_lo) return false;

  const uint32_t *TG = S + (ws.tg >> 2), *TA = S + (ws.ta >> 2);
  const TbHeader &hd = *reinterpret_cast<const TbHeader *>(Sb + ws.tb);
  const TbCont *tc = reinterpret_cast<const TbCont *>(Sb + ws.tb + sizeof(TbHeader));
  uint16_t *gstart = reinterpret_cast<uint16_t *>(Sb + ws.gstart);
  uint16_t *pstart = reinterpret_cast<uint16_t *>(Sb + ws.pstart);
  GRec *grec = reinterpret_cast<GRec *>(Sb + ws.grec);
  PRec *prec = reinterpret_cast<PRec *>(Sb + ws.prec);
  uint32_t *cnt = S + (ws.cnt >> 2), *cstart = S + (ws.cstart >> 2);
  uint32_t *plan = S + (ws.plan >> 2), *loff = S + (ws.loff >> 2);

  // ---- stage A: windows -> shared memory, rows; half 0 also parses the gpu rows ----
  if (half == 0) {
    uint32_t gs, ge, ng = 0;
    load_window(b.gpu_csv, g_lo, g_hi, Sb + ws.tg, &gs, &ge);
    {
      const uint8_t *src = f.tb_pool + f.tb_off[n];
      for (uint32_t i = l * 16u; i < tb_len; i += 512u)
        *reinterpret_cast<uint4 *>(Sb + ws.tb + i) = __ldg(reinterpret_cast<const uint4 *>(src + i));
    }
    __syncwarp();
    bool ok = hd.nc <= P.cap_c && find_rows(TG, gs, ge, gstart, P.cap_ng, &ng) && ng <= 32u;
    if (ok) {                                        // one lane per gpu row (gpu.py:104-109,119-143)
      uint64_t my_hash = 0;
      bool rok = true;
      if (l < ng) {
        GRec g;
        rok = parse_gpu_row(TG, gstart[l], gstart[l + 1] - 1u, hd.fixed, l, g);
        if (rok) { grec[l] = g; my_hash = g.hash; }
      }
      bool dup = false;                              // equal keys: last row wins (gpu.py:54) -> generic path
      for (uint32_t j = 0; j + 1 < ng; ++j) {
        const uint64_t hj = __shfl_sync(FULL, my_hash, j);
        dup |= l < ng && j < l && hj == my_hash;
      }
      ok = __all_sync(FULL, rok) && !__any_sync(FULL, dup);
      for (uint32_t c = l; c < hd.nc && ok; c += 32) cnt[c] = 0;
    }
    if (l == 0) { mb->ok0 = ok ? 1u : 0u; mb->ng = ng; }
  }
  // half 1: window, rows, and the parse of the first 32 process rows -- none of it needs
  // the gpu side, so it runs while half 0 is still formatting floats
  PRec p0;                                           // (half 1) row `l` of the first chunk
  uint32_t kw0[KEY_WORDS], klen0 = 0, np_local = 0;
  uint64_t kh0 = 0;
  uint32_t nm0 = 0, cfirst0 = NONE32;
  if (half == 1) {
    uint32_t as, ae;
    load_window(b.apps_csv, a_lo, a_hi, Sb + ws.ta, &as, &ae);
    bool ok = find_rows(TA, as, ae, pstart, P.cap_np, &np_local);
    if (ok) {
      bool rok = true;
      if (l < np_local) {
        rok = parse_proc_row(TA, pstart[l], pstart[l + 1] - 1u, p0, kw0, &klen0, &kh0);
        if (rok && p0.pidkey != ~0ull) pid_probe(f, n, p0.pidkey, &nm0, &cfirst0);
      }
      ok = __all_sync(FULL, rok);
    }
    if (l == 0) { mb->ok1 = ok ? 1u : 0u; mb->np = np_local; }
  }
  STAGE_MARK(1);                                     // stage A work
  team_sync(team);
  if (!(mb->ok0 & mb->ok1)) return false;
  const uint32_t ng = mb->ng, np = mb->np, nc = hd.nc;
  team_sync(team);                                   // the mailbox may be rewritten now
  STAGE_MARK(2);                                     // waiting for the other half

  // ---- stage B (half 1): uuid -> short id (gpu.py:151), pid -> container (gpu.py:177-178),
  //      place inside the container; further chunks of 32 rows are parsed here as well ----
  if (half == 1) {
    bool all_ok = true;
    for (uint32_t r0 = 0; r0 < np && all_ok; r0 += 32) {
      const uint32_t r = r0 + l;
      const bool valid = r < np;
      bool ok = true;
      PRec p;
      uint32_t kw[KEY_WORDS], klen = 0, nm = 0, c_first = NONE32;
      uint64_t kh = 0;
      if (r0 == 0) {
        p = p0; klen = klen0; kh = kh0; nm = nm0; c_first = cfirst0;
#pragma unroll
        for (int i = 0; i < KEY_WORDS; ++i) kw[i] = kw0[i];
      } else if (valid) {
        ok = parse_proc_row(TA, pstart[r], pstart[r + 1] - 1u, p, kw, &klen, &kh);
        if (ok && p.pidkey != ~0ull) pid_probe(f, n, p.pidkey, &nm, &c_first);
      }
      uint32_t emit_c = NONE32;
      if (valid && ok) {
        // first gpu row with this hash and length; the bytes are compared once, afterwards
        uint32_t sid = NONE32;
        for (uint32_t j = 0; j < ng && sid == NONE32; ++j)
          if (grec[j].hash == kh && grec[j].key.len == klen) sid = j;
        bool eq = sid != NONE32;                    // no such uuid: KeyError gpu.py:151
        if (eq) {
#pragma unroll
          for (int i = 0; i < KEY_WORDS; ++i) eq &= grec[sid].keyw[i] == kw[i];
        }
        ok = eq;                                    // (a hash collision also goes to the generic path)
        p.sid = (uint16_t)sid;
        if (ok && p.pidkey != ~0ull) {
          if (nm > 1) ok = false;                   // a pid listed by several containers
          if (nm == 1 && c_first < nc && tc[c_first].gated) emit_c = c_first;
          p.digest = pair_digest(p.pidkey, p.pid.len, sid);
        }
      }
      all_ok = __all_sync(FULL, ok);
      if (!all_ok) break;
      // duplicate (pid, gpu) keys keep the first row's place and the last row's values
      // (gpu.py:54): screened by digest, any hit goes to the generic path
      const bool screened = valid && p.pidkey != ~0ull;
      const uint32_t act = __ballot_sync(FULL, screened);
      bool dup = false;
      if (screened) {
        dup = __popc(__match_any_sync(act, p.digest)) > 1;
        for (uint32_t j = 0; j < r0 && !dup; ++j) dup = prec[j].pidkey != ~0ull && prec[j].digest == p.digest;
      }
      if (__any_sync(FULL, dup)) { all_ok = false; break; }
      // place inside the container: rank among the earlier rows of the same container
      const uint32_t emitting = __ballot_sync(FULL, emit_c != NONE32);
      uint32_t pos = 0;
      if (emit_c != NONE32) {
        const uint32_t peers = __match_any_sync(emitting, emit_c);
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if ((int)l == leader) { base = cnt[emit_c]; cnt[emit_c] = base + __popc(peers); }
        base = __shfl_sync(peers, base, leader);
        pos = base + __popc(peers & lanemask_lt());
      }
      if (valid) {
        p.crank = (uint16_t)(emit_c == NONE32 ? 0xffffu : emit_c);
        p.pos = (uint16_t)pos;
        prec[r] = p;
      }
      __syncwarp();
    }
    if (l == 0) mb->ok1 = all_ok ? 1u : 0u;
  } else {
    // half 0 meanwhile: repr(mem_util) and the node line lengths (gpu.py:135-143)
    bool fok = true;
    if (l < ng) fok = finish_gpu_row(TG, grec[l]);
    fok = __all_sync(FULL, fok);
    if (l == 0) mb->okf = fok ? 1u : 0u;
  }
  STAGE_MARK(3);                                     // stage B work (half 1) / floats (half 0)
  team_sync(team);
  if (!(mb->okf & mb->ok1)) return false;
  STAGE_MARK(4);

  // ---- stage C (half 0): containers in docker stats order: busy lines, or idle lines
  //      (gpu.py:174-253); node entries and idle entries of the plan ----
  if (half == 0) {
    uint32_t total_pod = 0;
    for (uint32_t c0 = 0; c0 < nc; c0 += 32) {
      const uint32_t c = c0 + l;
      uint32_t lines = 0;
      if (c < nc && tc[c].gated) lines = cnt[c] ? cnt[c] : tc[c].nvis;
      const uint32_t incl = warp_incl_scan(lines);
      if (c < nc) cstart[c] = total_pod + incl - lines;
      total_pod += __shfl_sync(FULL, incl, 31);
    }
    const uint32_t nl = ng + total_pod;
    if (l == 0) { mb->nl = nl; mb->ok0 = nl <= P.cap_lines ? 1u : 0u; }
    __syncwarp();
    if (nl <= P.cap_lines) {
      if (l < ng) plan[l] = (KIND_NODE << 30) | l;
      for (uint32_t c = l; c < nc; c += 32) {
        if (!tc[c].gated || cnt[c]) continue;
        for (uint32_t v = 0; v < tc[c].nvis; ++v) {  // uuids visible inside an idle container (gpu.py:221-226)
          const uint32_t gv = tc[c].vis_begin + v;
          const uint64_t vh = f.vis_hash2[gv];
          const uint32_t vo = f.vis_off[gv], vlen = f.vis_off[gv + 1] - vo;
          uint32_t sid = IDLE_NONE;
          for (uint32_t j = 0; j < ng && sid == IDLE_NONE; ++j) {
            if (grec[j].hash == vh && grec[j].key.len == vlen) {
              bool eq = true;
              for (uint32_t i = 0; i < vlen; ++i) eq &= f.vis_bytes[vo + i] == load1(TG, grec[j].key.off + i);
              if (eq) sid = j;
            }
          }
          plan[ng + cstart[c] + v] = (KIND_IDLE << 30) | (c << 12) | sid;
        }
      }
    }
  }
  STAGE_MARK(5);                                     // stage C work (half 0)
  team_sync(team);
  if (!mb->ok0) return false;
  const uint32_t nl = mb->nl;
  STAGE_MARK(6);
  if (half == 1) {                                   // busy entries (needs cstart)
    for (uint32_t r = l; r < np; r += 32) {
      const PRec &p = prec[r];
      if (p.crank != 0xffffu) plan[ng + cstart[p.crank] + p.pos] = (KIND_BUSY << 30) | r;
    }
  }
  team_sync(team);

  STAGE_MARK(7);                                     // busy entries + barrier
  // ---- stage D: line lengths -> offsets inside the document; each half measures half of
  //      the lines, half 1 then shifts its offsets by half 0's bytes ----
  const uint32_t mid = (nl + 1u) >> 1;
  const uint32_t lo_i = half ? mid : 0u, hi_i = half ? nl : mid;
  uint32_t nb = 0;
  bool too_long = false;
  for (uint32_t i0 = lo_i; i0 < hi_i; i0 += 32) {
    const uint32_t i = i0 + l;
    uint32_t len = 0;
    if (i < hi_i) {
      const uint32_t pl = plan[i], kind = pl >> 30;
      if (kind == KIND_NODE) {
        len = grec[pl & 0xffffu].line_len;
      } else if (kind == KIND_BUSY) {
        const PRec &p = prec[pl & 0x3fffffffu];
        len = busy_len(tc[p.crank], p, grec[p.sid], p.sid);
      } else {
        const uint32_t sid = pl & 0xfffu;
        len = idle_len(tc[(pl >> 12) & 0x3ffffu], sid == IDLE_NONE ? NONE32 : sid);
      }
      too_long |= len > MAX_LINE_BYTES;
    }
    const uint32_t incl = warp_incl_scan(len);
    if (i < hi_i) loff[i] = nb + incl - len;
    nb += __shfl_sync(FULL, incl, 31);
  }
  too_long = __any_sync(FULL, too_long);
  if (l == 0) {
    if (half == 0) { mb->total0 = nb; mb->ok0 = too_long ? 0u : 1u; }
    else { mb->nb = nb; mb->ok1 = too_long ? 0u : 1u; }
  }
  team_sync(team);
  if (!(mb->ok0 & mb->ok1)) return false;
  const uint32_t total0 = mb->total0, total = total0 + mb->nb;
  if (half == 1) {
    for (uint32_t i = lo_i + l; i < hi_i; i += 32) loff[i] += total0;
    if (l == 0) loff[nl] = total;
  }
  team_sync(team);
  STAGE_MARK(8);                                     // stage D
  if (half == 0 && l == 0) { mb->nb = total; }
  team_sync(team);
  return true;
}

// Writes the planned lines of a regular document.  Lines are composed into the stage from
// its start (whatever the global alignment will be), half of a round's lines by each warp
// of the team; the copy to global memory realigns them: byte stores up to the first 16-byte
// boundary of the destination, then 16-byte stores assembled from five shared-memory words
// each, byte stores for the tail.  The document's position in the output is only needed for
// that copy, so the look-back that yields it (scan_resolve, half 0) runs after the first
// round of lines has been composed.
__device__ void fast_emit(const FusedParams &P, const WorkDev &w, uint32_t *S, uint8_t *Sb, const WarpSlice &ws,
                          uint32_t team, uint32_t half, uint32_t d, uint32_t n_docs, uint32_t ng, uint32_t nl,
                          uint32_t nb, long long &stage_t) {
  const uint32_t l = lane_id();
  volatile Mailbox *mb = reinterpret_cast<volatile Mailbox *>(Sb + ws.mbox);
  const TbHeader &hd = *reinterpret_cast<const TbHeader *>(Sb + ws.tb);
  const TbCont *tc = reinterpret_cast<const TbCont *>(Sb + ws.tb + sizeof(TbHeader));
  const GRec *grec = reinterpret_cast<const GRec *>(Sb + ws.grec);
  const PRec *prec = reinterpret_cast<const PRec *>(Sb + ws.prec);
  const uint32_t *plan = S + (ws.plan >> 2), *loff = S + (ws.loff >> 2);
  // rounds of equal size: as few as the stage allows, at most 64 lines (32 per warp) each
  const uint32_t rounds = max((nl + 63u) / 64u, (nb + P.stage_bytes - 64u - 1u) / (P.stage_bytes - 64u));
  const uint32_t per = rounds ? (nl + rounds - 1u) / rounds : 0u;
  uint64_t line0 = 0, byte0 = 0;
  bool resolved = false, fits_out = true;
  uint32_t i0 = 0;
  do {
    uint32_t i1 = min(nl, i0 + per);
    uint32_t r_lo = 0;
    if (nl) {
      r_lo = loff[i0];
      // the next `per` lines, fewer if their bytes exceed the stage (both warps compute the same)
      while (i1 > i0 + 1 && loff[i1] - r_lo + 64u > P.stage_bytes) i1 = i0 + ((i1 - i0) >> 1);
      // split of the round between the two warps: the node lines (a different segment
      // list, so mixing them with pod lines in one warp costs both code paths) go to half 0
      // alone when the pod lines of the round fit one warp; otherwise half and half
      const uint32_t cnt = i1 - i0;
      uint32_t cut = (cnt + 1u) >> 1;
      if (i0 < ng && i1 > ng && ng - i0 <= 32u && i1 - ng <= 32u) cut = ng - i0;
      const uint32_t first = half ? cut : 0u, last = half ? cnt : cut;
      const uint32_t i = i0 + first + l;
      if (i < i0 + last) {
        const uint32_t pl = plan[i], kind = pl >> 30;
        Seg segs[MAX_SEGS];
        if (kind == KIND_NODE) {
          const uint32_t r = pl & 0xffffu;
          node_segs(segs, 0u, ws.tb, hd, ws.tg, grec[r],
                    ws.grec + r * (uint32_t)sizeof(GRec) + (uint32_t)offsetof(GRec, fstr), r);
        } else if (kind == KIND_BUSY) {
          const PRec &p = prec[pl & 0x3fffffffu];
          busy_segs(segs, 0u, ws.tb, tc[p.crank], ws.tg, ws.ta, p, grec[p.sid], p.sid);
        } else {
          const uint32_t sid = pl & 0xfffu;
          idle_segs(segs, 0u, ws.tb, tc[(pl >> 12) & 0x3ffffu], sid == IDLE_NONE ? NONE32 : sid);
        }
        write_line(S, ws.stage + (loff[i] - r_lo), segs);
      }
    }
    STAGE_MARK(9);                                   // segments + compose
    if (!resolved) {
      if (half == 0) {
        scan_resolve(P, d, nl, nb, &line0, &byte0);
        if (P.carry) { line0 += P.carry[0]; byte0 += P.carry[1]; }
        if (l == 0) {
          mb->line0 = line0; mb->byte0 = byte0;
          w.doc_line_off[d] = line0;
          w.doc_byte_off[d] = byte0;
          w.doc_status[d] = 0; w.doc_nlines[d] = nl; w.doc_nbytes[d] = nb;
          if (d + 1 == n_docs) {
            w.doc_line_off[n_docs] = line0 + nl;
            w.doc_byte_off[n_docs] = byte0 + nb;
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
      }
      team_sync(team);                               // lines composed, position known
      line0 = mb->line0; byte0 = mb->byte0;
      resolved = true;
      fits_out = byte0 + nb <= w.out_cap && line0 + nl <= w.desc_cap;
    } else {
      team_sync(team);                               // lines of this round composed
    }
    STAGE_MARK(10);                                  // look-back + barrier
    if (!fits_out || nl == 0) return;
    for (uint32_t i = i0 + 32u * half + l; i < i1; i += 64u) w.line_off[line0 + i] = byte0 + loff[i];
    // stage -> global, the vectors dealt alternately to the two warps
    const uint32_t nbytes = loff[i1] - r_lo;
    uint8_t *g = w.out + byte0 + r_lo;
    const uint8_t *st = Sb + ws.stage;
    const uint32_t head = min(nbytes, (uint32_t)((16u - ((byte0 + r_lo) & 15ull)) & 15u));
    if (half == 0 && l < head) g[l] = st[l];
    const uint32_t nvec = (nbytes - head) >> 4;
    {
      const uint32_t *sw = S + ((ws.stage + head) >> 2);     // stage word holding byte `head`
      const uint32_t sh = 8u * (head & 3u);
      for (uint32_t v = 32u * half + l; v < nvec; v += 64u) {
        const uint32_t *q = sw + 4u * v;
        const uint32_t a0 = q[0], a1 = q[1], a2 = q[2], a3 = q[3], a4 = q[4];
        uint4 o;
        o.x = __funnelshift_r(a0, a1, sh); o.y = __funnelshift_r(a1, a2, sh);
        o.z = __funnelshift_r(a2, a3, sh); o.w = __funnelshift_r(a3, a4, sh);
        *reinterpret_cast<uint4 *>(g + head + 16u * v) = o;
      }
    }
    const uint32_t done = head + 16u * nvec;
    if (half == 1 && l < nbytes - done) g[done + l] = st[done + l];
    STAGE_MARK(11);                                  // line_off + copy-out
    i0 = i1;
    team_sync(team);                                 // the stage may be overwritten
  } while (i0 < nl);
}

// ---- generic path for one document (device functions of the staged kernels) -------------------
__device__ void slow_plan(uint32_t d, const BatchDev &b, const FleetDev &f, const WorkDev &w, uint32_t *scratch,
                          uint32_t *nl, uint32_t *nb) {
  uint32_t g0, ng, gerr, p0, np, perr;
  tokenize_text<GpuRow, GpuRowScan, 1>(b.gpu_csv, (uint32_t)b.gpu_off[d], (uint32_t)b.gpu_off[d + 1], w.grow,
                                       w.grow_cap, w.counters + CTR_GROW, w.counters + CTR_OVERFLOW, OVF_GROW,
                                       &g0, &ng, &gerr);
  tokenize_text<ProcRow, ProcRowScan, 3>(b.apps_csv, (uint32_t)b.apps_off[d], (uint32_t)b.apps_off[d + 1], w.prow,
                                         w.prow_cap, w.counters + CTR_PROW, w.counters + CTR_OVERFLOW, OVF_PROW,
                                         &p0, &np, &perr);
  if (lane_id() == 0) {
    w.doc_g0[d] = g0; w.doc_ng[d] = ng;
    w.doc_p0[d] = p0; w.doc_np[d] = np;
    w.doc_perr[d] = gerr ? gerr : perr;
  }
  __threadfence_block();
  __syncwarp();
  join_doc(d, b, f, w, scratch, scratch + f.max_nc, scratch + 2 * f.max_nc);
  __threadfence_block();
  __syncwarp();
  *nl = w.doc_nlines[d];
  *nb = w.doc_nbytes[d];
}

__device__ void slow_emit(uint32_t d, const BatchDev &b, const FleetDev &f, const WorkDev &w, uint32_t nl,
                          uint64_t line0, uint64_t byte0) {
  const uint32_t desc0 = w.doc_desc0[d];
  for (uint32_t i = lane_id(); i < nl; i += 32) {
    const uint32_t di = desc0 + i;
    const LineDesc ds = w.desc[di];
    const uint64_t goff = byte0 + w.line_loff[di];
    ByteSink s{w.out + goff};
    const uint32_t kind = ds.kind_ref >> 30, ref = ds.kind_ref & 0x3fffffffu;
    if (kind == LINE_NODE) {
      compose_node_line(s, f.skel, f.node_piece + (size_t)b.doc_node[ds.doc] * NODE_PIECES, b.gpu_csv,
                        w.grow[ds.a], ds.b, w.fstr + (size_t)ds.a * FSTR_STRIDE);
    } else if (kind == LINE_BUSY) {
      const ProcRow &pr = w.prow[ds.a];
      compose_busy_line(s, f.skel, f.cont_piece + (size_t)ref * POD_PIECES, b.gpu_csv, b.apps_csv, pr,
                        w.grow[ds.b], pr.sid);
    } else {
      compose_idle_line(s, f.skel, f.cont_piece + (size_t)ref * POD_PIECES, ds.a);
    }
    w.line_off[line0 + i] = goff;
  }
}

__global__ void __launch_bounds__(POLL_THREADS, 1) k_poll(BatchDev b, FleetDev f, WorkDev w, FusedParams P) {
  extern __shared__ __align__(16) uint32_t S[];
  uint8_t *Sb = reinterpret_cast<uint8_t *>(S);
  // constant strings (same for every CTA)
  for (uint32_t i = threadIdx.x; i < CONST_PAD / 4; i += blockDim.x) S[i] = 0;
  __syncthreads();
  if (threadIdx.x == 0) fill_const_strings(Sb);
  for (uint32_t v = threadIdx.x; v <= MAX_FAST_SID; v += blockDim.x) fill_const_decimal(Sb, v);
  // the counters of the NEXT launch start from zero (this launch uses the other set)
  if (blockIdx.x == 0 && threadIdx.x < CTR_COUNT) P.next_counters[threadIdx.x] = 0;
  __syncthreads();

  const uint32_t team = threadIdx.x >> 6, half = (threadIdx.x >> 5) & 1u, l = lane_id();
  const uint32_t base = CONST_PAD + team * P.warp_bytes;
  WarpSlice ws;
  ws.tg = base + P.off_tg; ws.ta = base + P.off_ta; ws.gstart = base + P.off_gstart;
  ws.pstart = base + P.off_pstart; ws.grec = base + P.off_grec; ws.prec = base + P.off_prec;
  ws.cnt = base + P.off_cnt; ws.cstart = base + P.off_cstart; ws.plan = base + P.off_plan;
  ws.loff = base + P.off_loff; ws.tb = base + P.off_tb; ws.stage = base + P.off_stage;
  ws.mbox = base + P.off_mbox;
  volatile Mailbox *mb = reinterpret_cast<volatile Mailbox *>(Sb + ws.mbox);

  for (;;) {
    if (half == 0 && l == 0) mb->ticket = atomicAdd(P.ticket, 1ull) - P.ticket_base;
    team_sync(team);
    const unsigned long long t = mb->ticket;
    if (t >= b.n_docs) break;
    const uint32_t d = (uint32_t)t;
    long long stage_t = 0;
#ifdef KGM_PROFILE_STAGES
    stage_t = clock64();
#endif
    const bool fastdoc = P.allow_fast && fast_plan(d, b, f, P, S, Sb, ws, team, half, stage_t);
    if (fastdoc) {
      const uint32_t ng = mb->ng, nl = mb->nl, nb = mb->nb;
      team_sync(team);
      if (half == 0) scan_publish(P, d, b.n_docs, nl, nb);
      fast_emit(P, w, S, Sb, ws, team, half, d, b.n_docs, ng, nl, nb, stage_t);
      team_sync(team);
      STAGE_MARK(12);                                // last barrier
      continue;
    }
    team_sync(team);                                 // both halves have left the fast path
    if (half == 0) {                                 // the generic path is one warp's work
      uint32_t nl = 0, nb = 0;
      slow_plan(d, b, f, w, S + (ws.stage >> 2), &nl, &nb);
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
    team_sync(team);
  }
}

}  // namespace

int debug_stage_clocks(unsigned long long *out64) {
#ifdef KGM_PROFILE_STAGES
  return cudaMemcpyFromSymbol(out64, g_stage_clk, sizeof(g_stage_clk)) == cudaSuccess ? 64 : -1;
#else
  (void)out64;
  return 0;
#endif
}

int poll_smem_bytes(const FusedParams &P) { return (int)(CONST_PAD + P.teams * P.warp_bytes); }

void launch_poll(const BatchDev &b, const FleetDev &f, const WorkDev &w, const FusedParams &P, int blocks,
                 int device, cudaStream_t s) {
  static int attr_bytes[64];                         // the shared-memory opt-in is per device
  const int need = poll_smem_bytes(P);
  if (device < 0 || device >= 64 || need > attr_bytes[device]) {
    cudaFuncSetAttribute(k_poll, cudaFuncAttributeMaxDynamicSharedMemorySize, need);
    if (device >= 0 && device < 64) attr_bytes[device] = need;
  }
  k_poll<<<blocks, 64 * P.teams, need, s>>>(b, f, w, P);
}

}  // namespace kgm
