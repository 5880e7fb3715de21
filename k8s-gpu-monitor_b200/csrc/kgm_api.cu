// C ABI of libkgm.so (include/kgm.h): context, fleet tables, batch plumbing, launches.
//
// Host code only; every byte of the poll path itself is produced by the kernels in
// kgm_kernels.cu.  There is deliberately no CPU implementation behind these entry points.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <new>
#include <string>
#include <vector>

#include "../../include/kgm.h"
#include "kgm_device.cuh"
#include "kgm_skeleton.h"
#include "kgm_window.cuh"

using namespace kgm;

namespace {

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    bytes = (bytes + 255) & ~(size_t)255;
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes + 256);     // +256: aligned 16-byte over-reads stay inside
    if (e == cudaSuccess) cap = bytes;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct HostBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    bytes = (bytes + 4095) & ~(size_t)4095;
    cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocMapped);      // kernels may write it (unified addressing)
    if (e == cudaSuccess) cap = bytes;
    return e;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
  template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

enum { EV_START = 0, EV_H2D, EV_TOK, EV_JOIN, EV_SCAN, EV_ENC, EV_D2H, EV_COUNT };
constexpr int MAX_CHUNKS = 8;
constexpr size_t OUT_HEADER = 16;          // bytes in front of the lines: {n_bytes, n_lines} of the last run
constexpr uint32_t CTR_RING = 16;        // counter sets: launch e uses set e % CTR_RING and zeroes the next one,
                                         // so the sets of the last CTR_RING - 1 launches can still be read

}  // namespace

struct kgm_ctx {
  int device = 0;
  int sms = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[EV_COUNT] = {};
  std::string cuda_err;

  // fleet
  bool fleet_loaded = false;
  FleetDev fleet{};
  std::vector<uint64_t> node_idle_cap;     // per node: total visible uuids of gated containers
  DevBuf f_cont_begin, f_nerr_parse, f_nerr_join, f_gated, f_err_inspect, f_err_emit, f_vis_begin,
      f_vis_off, f_vis_bytes, f_vis_hash, f_node_piece, f_node_fixed, f_cont_piece, f_cont_fixed,
      f_skel, f_pid_table;

  // batch
  bool batch_resident = false;
  BatchDev batch{};
  uint64_t gpu_bytes = 0, apps_bytes = 0, idle_cap = 0;
  DevBuf b_gpu, b_gpu_off, b_apps, b_apps_off, b_node, b_tick;

  // workspace
  WorkDev work{};
  DevBuf w_grow, w_prow, w_uniq, w_fstr, w_desc, w_len, w_loff, w_counters, w_doc32, w_status,
      w_doc_line_off, w_doc_byte_off, w_totals, w_line_off, w_out, w_flush;
  uint32_t grow_cap = 0, prow_cap = 0, desc_cap = 0;
  uint64_t out_cap = 0;

  // fused kernel (k_poll)
  int pipeline = 0;                        // KGM_PIPELINE_*
  bool fused_ok = false;                   // the current batch fits the fused kernel's shared-memory slices
  FusedParams fused{};
  int fused_ctas_per_sm = 2;
  uint32_t epoch = 0;
  bool counters_clean = false;             // the counter set of the next fused launch is known to be zero
  unsigned long long ticket_next = 0;      // value of the ticket counter when the next launch starts
  DevBuf w_ticket, w_scan;
  uint64_t scan_docs_cap = 0;              // documents the level-1 scan state has room for
  DevBuf f_node_fast, f_tb_off, f_tb_len, f_tb_pool, f_pid16, f_vis_hash2;
  uint32_t fast_max_nc = 1, fast_max_tb = 0, fast_max_idle = 0, any_fast = 0;
  std::vector<uint8_t> h_node_fast;        // host copies of FleetDev::node_fast / tb_len
  bool all_nodes_fast = false;
  std::vector<uint32_t> h_tb_len;
  uint32_t slow_docs = 0;
  bool last_fused = false;
  bool last_poll2 = false;
  bool totals_external = false;            // chunked path: every chunk has its own totals slot
  bool chunk_launch = false;               // chunked path: no timing events around the single launches
  // chunked end-to-end path: copies and kernels of consecutive chunks overlap
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaEvent_t ev_in[MAX_CHUNKS] = {}, ev_k[MAX_CHUNKS] = {};
  cudaEvent_t ev_enc_done = nullptr, ev_copy_done = nullptr;   // kgm_push_copy_step
  bool copy_pending = false;
  HostBuf h_chunk;                         // per chunk: posted totals {lines, bytes, seq} + counters
  HostBuf h_docarr;                        // staging of the per-document arrays
  uint64_t post_seq = 0;
  uint32_t last_chunks = 1;
  // k_poll2 (one warp per document)
  bool poll2_ok = false;                   // the current batch fits k_poll2's slices
  Poll2Params poll2{};
  DevBuf f_node_fast2, f_nslot_off, f_tslot_off, f_slot_pool, f_ca2;
  uint32_t p2_max_nc = 1, p2_max_slot = 64, p2_max_fixed = 0, p2_max_idle = 0, p2_any = 0;
  std::vector<uint8_t> h_node_fast2;
  bool all_nodes_fast2 = false;
  uint32_t force_lines = 0;                // KGM_POLL2_LINES=4..32: lines per round the stage is sized for (experiments)
  uint32_t force_warps = 0;                // KGM_POLL2_WARPS: documents in flight per SM (experiments)
  uint32_t force_teams = 0;                // KGM_FUSED_TEAMS=4..8: documents in flight per SM (experiments)
  int reserve_ctas = 0;                    // KGM_RESERVE_CTAS: CTA slots k_poll leaves free (for NCCL kernels)
  // window reductions
  uint32_t max_tick = 0;
  bool have_ticks = false;
  DevBuf w_win_samples, w_win_groups, w_win_temp, w_win_scal;
  bool win_front_fused = false;            // the last kgm_window_reduce took its samples from k_poll2_samples
  HostBuf h_win_groups;

  // results
  bool have_result = false;
  uint8_t *ext_out = nullptr;              // kgm_set_output_buffer: the caller's buffer instead of w_out
  uint64_t ext_out_bytes = 0;
  uint64_t peers_enabled = 0;              // devices cudaDeviceEnablePeerAccess has been called for
  bool async_pending = false;              // kgm_encode_resident_async without its kgm_encode_finish
  uint64_t n_lines = 0, n_bytes = 0;
  HostBuf h_small, h_lines, h_line_off, h_doc_line_off, h_status;
  kgm_timing timing{};
  bool timed_h2d = false, timed_d2h = false;
};

namespace {

int cuda_fail(kgm_ctx *c, cudaError_t e, const char *what) {
  c->cuda_err = std::string(what) + ": " + cudaGetErrorString(e);
  cudaGetLastError();
  return e == cudaErrorMemoryAllocation ? KGM_E_NOMEM : KGM_E_CUDA;
}

#define CU(call)                                              \
  do {                                                        \
    cudaError_t e__ = (call);                                 \
    if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call); \
  } while (0)

template <class T>
int upload(kgm_ctx *ctx, DevBuf &buf, const T *src, size_t n) {
  CU(buf.ensure(std::max<size_t>(n, 1) * sizeof(T)));
  if (n) CU(cudaMemcpyAsync(buf.p, src, n * sizeof(T), cudaMemcpyDefault, ctx->stream));   // (host or device source: unified addressing)
  return KGM_OK;
}

uint64_t hash_bytes(const uint8_t *p, uint32_t n) {
  uint64_t h = hash_init();
  for (uint32_t i = 0; i < n; ++i) h = hash_step(h, p[i]);
  return h;
}

bool offsets_ok(const uint32_t *off, uint32_t n) {
  if (!off) return false;
  for (uint32_t i = 0; i < n; ++i)
    if (off[i + 1] < off[i]) return false;
  return true;
}

std::string str_of(const kgm_strings &s, uint32_t i) {
  return std::string(reinterpret_cast<const char *>(s.bytes) + s.off[i], s.off[i + 1] - s.off[i]);
}

int ensure_workspace(kgm_ctx *ctx) {
  const uint32_t n = ctx->batch.n_docs;
  WorkDev &w = ctx->work;
  CU(ctx->w_grow.ensure((size_t)ctx->grow_cap * sizeof(GpuRow)));
  CU(ctx->w_uniq.ensure((size_t)ctx->grow_cap * sizeof(uint32_t)));
  CU(ctx->w_fstr.ensure((size_t)ctx->grow_cap * FSTR_STRIDE));
  CU(ctx->w_prow.ensure((size_t)ctx->prow_cap * sizeof(ProcRow)));
  CU(ctx->w_desc.ensure((size_t)ctx->desc_cap * sizeof(LineDesc)));
  CU(ctx->w_len.ensure((size_t)ctx->desc_cap * sizeof(uint32_t)));
  CU(ctx->w_loff.ensure((size_t)ctx->desc_cap * sizeof(uint32_t)));
  CU(ctx->w_line_off.ensure(((size_t)ctx->desc_cap + 1) * sizeof(uint64_t)));
  {
    const size_t before = ctx->w_counters.cap;
    CU(ctx->w_counters.ensure(CTR_RING * CTR_SET * sizeof(uint32_t)));
    if (ctx->w_counters.cap != before) {
      CU(cudaMemsetAsync(ctx->w_counters.p, 0, CTR_RING * CTR_SET * sizeof(uint32_t), ctx->stream));
      ctx->counters_clean = true;
    }
    const size_t tb = ctx->w_ticket.cap;
    CU(ctx->w_ticket.ensure(64));
    if (ctx->w_ticket.cap != tb) {
      CU(cudaMemsetAsync(ctx->w_ticket.p, 0, 64, ctx->stream));
      ctx->ticket_next = 0;
    }
    const size_t sb = ctx->w_scan.cap;
    CU(ctx->w_scan.ensure(((size_t)std::max(n, 1u) + (size_t)n / 32 + 2) * 16));
    if (ctx->w_scan.cap != sb) CU(cudaMemsetAsync(ctx->w_scan.p, 0, ctx->w_scan.cap, ctx->stream));
    ctx->scan_docs_cap = std::max<uint64_t>(ctx->scan_docs_cap, std::max(n, 1u));
  }
  CU(ctx->w_doc32.ensure((size_t)std::max(n, 1u) * 8 * sizeof(uint32_t)));
  CU(ctx->w_status.ensure((size_t)std::max(n, 1u) * sizeof(int32_t)));
  CU(ctx->w_doc_line_off.ensure(((size_t)n + 1) * sizeof(uint64_t)));
  CU(ctx->w_doc_byte_off.ensure(((size_t)n + 1) * sizeof(uint64_t)));
  CU(ctx->w_totals.ensure(2 * (MAX_CHUNKS + 1) * sizeof(uint64_t)));
  if (!ctx->ext_out) CU(ctx->w_out.ensure((size_t)ctx->out_cap + 16 + OUT_HEADER));
  w.grow = ctx->w_grow.as<GpuRow>(); w.grow_cap = ctx->grow_cap;
  w.prow = ctx->w_prow.as<ProcRow>(); w.prow_cap = ctx->prow_cap;
  w.uniq = ctx->w_uniq.as<uint32_t>();
  w.fstr = ctx->w_fstr.as<uint8_t>();
  w.desc = ctx->w_desc.as<LineDesc>(); w.desc_cap = ctx->desc_cap;
  w.line_len = ctx->w_len.as<uint32_t>();
  w.line_loff = ctx->w_loff.as<uint32_t>();
  w.counters = ctx->w_counters.as<uint32_t>();   // launch_pipeline picks the set
  uint32_t *d32 = ctx->w_doc32.as<uint32_t>();
  const size_t st = std::max(n, 1u);
  w.doc_g0 = d32; w.doc_ng = d32 + st; w.doc_p0 = d32 + 2 * st; w.doc_np = d32 + 3 * st;
  w.doc_perr = d32 + 4 * st; w.doc_desc0 = d32 + 5 * st; w.doc_nlines = d32 + 6 * st;
  w.doc_nbytes = d32 + 7 * st;
  w.doc_status = ctx->w_status.as<int32_t>();
  w.doc_line_off = ctx->w_doc_line_off.as<uint64_t>();
  w.doc_byte_off = ctx->w_doc_byte_off.as<uint64_t>();
  w.totals = ctx->w_totals.as<uint64_t>();           // launch_pipeline points it into the counter set
  w.line_off = ctx->w_line_off.as<uint64_t>();
  w.out = ctx->w_out.as<uint8_t>() + OUT_HEADER; w.out_cap = ctx->out_cap;   // {bytes, lines} go in front of the lines
  if (ctx->ext_out) { w.out = ctx->ext_out + OUT_HEADER; w.out_cap = ctx->ext_out_bytes - OUT_HEADER; }
  return KGM_OK;
}

// The batch goes through the fused kernel unless it was switched off or does not fit.
// Two fused kernels: k_poll (two-warp teams, 7 - 8 documents in flight per SM, whole document
// staged) has the shorter path per document and takes the small batches; k_poll2 (one warp per
// document, 16 in flight) takes every batch that fills the GPU at least once (measured on the
// configs[1] batch, 3840 documents, and the configs[2] batch, 40 000: profiles/README.md).
bool use_poll2(const kgm_ctx *ctx) {
  if (ctx->pipeline == KGM_PIPELINE_STAGED || ctx->pipeline == KGM_PIPELINE_FUSED_V1 || !ctx->poll2_ok) return false;
  if (ctx->pipeline == KGM_PIPELINE_FUSED_V2 || !ctx->fused_ok) return true;
  return ctx->batch.n_docs >= (uint32_t)ctx->sms * 16u;          // at least one full wave of k_poll2
}
bool use_fused(const kgm_ctx *ctx) {
  return use_poll2(ctx) || (ctx->pipeline != KGM_PIPELINE_STAGED && ctx->fused_ok);
}

// Launches the pipeline on the context's stream: the fused kernel (one launch), or the four
// staged kernels (events around each).
int launch_pipeline(kgm_ctx *ctx) {
  uint32_t *cbase = ctx->w_counters.as<uint32_t>();
  if (use_fused(ctx)) {
    ++ctx->epoch;
    if ((ctx->epoch & 0xffffu) == 0) {             // tags wrapped: stale state could alias
      ++ctx->epoch;
      ctx->counters_clean = false;
      CU(cudaMemsetAsync(ctx->w_scan.p, 0, ctx->w_scan.cap, ctx->stream));
    }
    if (!ctx->counters_clean)
      CU(cudaMemsetAsync(cbase, 0, CTR_RING * CTR_SET * sizeof(uint32_t), ctx->stream));
    const uint32_t cur = ctx->epoch % CTR_RING;
    ctx->work.counters = cbase + cur * CTR_SET;
    if (!ctx->totals_external) ctx->work.totals = reinterpret_cast<uint64_t *>(ctx->work.counters + CTR_TOTALS);
    const bool v2 = use_poll2(ctx);
    FusedParams P = ctx->fused;
    P.epoch = ctx->epoch;
    P.allow_fast = ctx->pipeline == KGM_PIPELINE_FUSED_GENERIC ? 0u : (v2 ? ctx->p2_any : ctx->any_fast);
    P.ticket = ctx->w_ticket.as<unsigned long long>();
    P.ticket_base = ctx->ticket_next;
    P.scan_state = ctx->w_scan.as<unsigned long long>();
    P.scan_blocks = P.scan_state + 2ull * ctx->scan_docs_cap;
    P.next_counters = cbase + ((cur + 1u) % CTR_RING) * CTR_SET;
    const uint32_t n = ctx->batch.n_docs;
    const int wpb = v2 ? (int)ctx->poll2.warps : (int)P.teams;   // documents in flight per CTA
    int blocks = (int)((n + wpb - 1) / wpb);
    blocks = std::max(1, std::min(blocks, ctx->sms * (v2 ? 1 : ctx->fused_ctas_per_sm) - ctx->reserve_ctas));
    if (v2 && ctx->poll2.cta_tickets)
      ctx->ticket_next += ((unsigned long long)(n + wpb - 1) / wpb) * wpb + (unsigned long long)blocks * wpb;
    else
      ctx->ticket_next += (unsigned long long)n + (unsigned long long)blocks * wpb;
    if (!ctx->chunk_launch) CU(cudaEventRecord(ctx->ev[EV_SCAN], ctx->stream));
    if (v2) launch_poll2(ctx->batch, ctx->fleet, ctx->work, P, ctx->poll2, blocks, ctx->device, ctx->stream);
    else launch_poll(ctx->batch, ctx->fleet, ctx->work, P, blocks, ctx->device, ctx->stream);
    ctx->last_poll2 = v2;
    if (!ctx->chunk_launch) CU(cudaEventRecord(ctx->ev[EV_ENC], ctx->stream));
    CU(cudaGetLastError());
    ctx->counters_clean = true;                     // k_poll zeroed the other set
    ctx->last_fused = true;
    return KGM_OK;
  }
  ctx->work.counters = cbase;
  ctx->work.totals = reinterpret_cast<uint64_t *>(cbase + CTR_TOTALS);
  ctx->counters_clean = false;
  ctx->last_fused = false;
  CU(cudaMemsetAsync(ctx->work.counters, 0, CTR_COUNT * sizeof(uint32_t), ctx->stream));
  CU(cudaEventRecord(ctx->ev[EV_H2D], ctx->stream));
  launch_tokenize(ctx->batch, ctx->work, ctx->sms, ctx->stream);
  CU(cudaEventRecord(ctx->ev[EV_TOK], ctx->stream));
  launch_join(ctx->batch, ctx->fleet, ctx->work, ctx->sms, ctx->stream);
  CU(cudaEventRecord(ctx->ev[EV_JOIN], ctx->stream));
  launch_scan(ctx->batch, ctx->work, ctx->stream);
  CU(cudaEventRecord(ctx->ev[EV_SCAN], ctx->stream));
  launch_encode(ctx->batch, ctx->fleet, ctx->work, ctx->sms, ctx->device, ctx->stream);
  CU(cudaEventRecord(ctx->ev[EV_ENC], ctx->stream));
  CU(cudaGetLastError());
  return KGM_OK;
}

// Grows the guesses an overflow word names.
void grow_after_overflow(kgm_ctx *ctx, const uint32_t *hc) {
  const uint32_t ovf = hc[CTR_OVERFLOW];
  const uint64_t *ht = reinterpret_cast<const uint64_t *>(hc + CTR_TOTALS);
  if (ovf & OVF_GROW) ctx->grow_cap = std::max(ctx->grow_cap * 2, hc[CTR_GROW] + 64);
  if (ovf & OVF_PROW) ctx->prow_cap = std::max(ctx->prow_cap * 2, hc[CTR_PROW] + 64);
  if (ovf & OVF_DESC) ctx->desc_cap = std::max(ctx->desc_cap * 2, hc[CTR_DESC] + 64);
  if (ovf & OVF_OUT) ctx->out_cap = ht[1] + ht[1] / 8 + 4096;
}

// Launches the pipeline on the resident batch and queues the read-back of its counters;
// does not wait.
int launch_resident(kgm_ctx *ctx, bool record_start) {
  if (!ctx->fleet_loaded || !ctx->batch_resident) return KGM_E_STATE;
  CU(ctx->h_small.ensure(128));
  int rc = ensure_workspace(ctx);
  if (rc) return rc;
  if (record_start) CU(cudaEventRecord(ctx->ev[EV_START], ctx->stream));
  if (ctx->batch.n_docs == 0)                      // no kernel writes the size header of an empty batch
    CU(cudaMemsetAsync(ctx->work.out - OUT_HEADER, 0, OUT_HEADER, ctx->stream));
  rc = launch_pipeline(ctx);
  if (rc) return rc;
  CU(cudaMemcpyAsync(ctx->h_small.p, ctx->work.counters, CTR_SET * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  return KGM_OK;
}

// Waits for launch_resident's work; re-runs with a larger workspace while a capacity guess
// was too low.
int finish_resident(kgm_ctx *ctx) {
  for (int attempt = 0; attempt < 6; ++attempt) {
    CU(cudaStreamSynchronize(ctx->stream));
    const uint32_t *hc = ctx->h_small.as<uint32_t>();
    const uint64_t *ht = reinterpret_cast<const uint64_t *>(hc + CTR_TOTALS);
    ctx->slow_docs = ctx->last_fused ? hc[CTR_SLOW] : 0;
    if (!hc[CTR_OVERFLOW]) {
      ctx->n_lines = ht[0];
      ctx->n_bytes = ht[1];
      ctx->have_result = true;
      return KGM_OK;
    }
    if ((hc[CTR_OVERFLOW] & OVF_OUT) && ctx->ext_out) return KGM_E_LIMIT;    // the caller's buffer is too small
    grow_after_overflow(ctx, hc);
    int rc = launch_resident(ctx, true);
    if (rc) return rc;
  }
  return KGM_E_LIMIT;
}

// Runs the pipeline, growing the workspace and re-running when a capacity guess was low.
int run_resident(kgm_ctx *ctx, bool with_h2d_event) {
  int rc = launch_resident(ctx, !with_h2d_event);
  return rc ? rc : finish_resident(ctx);
}

void fill_timing(kgm_ctx *ctx, bool h2d, bool d2h) {
  kgm_timing &t = ctx->timing;
  t = kgm_timing{};
  if (ctx->last_fused) {
    if (h2d) cudaEventElapsedTime(&t.h2d_ms, ctx->ev[EV_START], ctx->ev[EV_SCAN]);
  } else {
    if (h2d) cudaEventElapsedTime(&t.h2d_ms, ctx->ev[EV_START], ctx->ev[EV_H2D]);
    cudaEventElapsedTime(&t.tokenize_ms, ctx->ev[EV_H2D], ctx->ev[EV_TOK]);
    cudaEventElapsedTime(&t.join_ms, ctx->ev[EV_TOK], ctx->ev[EV_JOIN]);
    cudaEventElapsedTime(&t.scan_ms, ctx->ev[EV_JOIN], ctx->ev[EV_SCAN]);
  }
  cudaEventElapsedTime(&t.encode_ms, ctx->ev[EV_SCAN], ctx->ev[EV_ENC]);
  if (d2h) cudaEventElapsedTime(&t.d2h_ms, ctx->ev[EV_ENC], ctx->ev[EV_D2H]);
  cudaEventElapsedTime(&t.total_ms, ctx->ev[EV_START], ctx->ev[d2h ? EV_D2H : EV_ENC]);
  t.launches = ctx->last_fused ? 1 : 4;
  t.fused = ctx->last_fused ? 1 : 0;
  t.slow_docs = ctx->slow_docs;
  t.fused_ctas_per_sm = ctx->last_fused ? (ctx->last_poll2 ? 1u : (uint32_t)ctx->fused_ctas_per_sm) : 0u;
  t.docs_in_flight_per_sm = ctx->last_fused ? (ctx->last_poll2 ? ctx->poll2.warps : ctx->fused.teams * ctx->fused_ctas_per_sm) : 0u;
  cudaGetLastError();
}

void fill_result_dev(kgm_ctx *ctx, kgm_result *out) {
  out->lines = ctx->work.out;
  out->n_bytes = ctx->n_bytes;
  out->line_off = ctx->work.line_off;
  out->n_lines = ctx->n_lines;
  out->doc_line_off = ctx->work.doc_line_off;
  out->doc_status = ctx->work.doc_status;
  out->n_docs = ctx->batch.n_docs;
}

void fill_result_host(kgm_ctx *ctx, kgm_result *out) {
  out->lines = ctx->h_lines.as<uint8_t>();
  out->n_bytes = ctx->n_bytes;
  out->line_off = ctx->h_line_off.as<uint64_t>();
  out->n_lines = ctx->n_lines;
  out->doc_line_off = ctx->h_doc_line_off.as<uint64_t>();
  out->doc_status = ctx->h_status.as<int32_t>();
  out->n_docs = ctx->batch.n_docs;
}

int fetch(kgm_ctx *ctx, kgm_result *out) {
  const uint32_t n = ctx->batch.n_docs;
  CU(ctx->h_lines.ensure(ctx->n_bytes + 16));
  CU(ctx->h_line_off.ensure((ctx->n_lines + 1) * sizeof(uint64_t)));
  CU(ctx->h_doc_line_off.ensure(((size_t)n + 1) * sizeof(uint64_t)));
  CU(ctx->h_status.ensure(std::max<size_t>(n, 1) * sizeof(int32_t)));
  if (ctx->n_bytes)
    CU(cudaMemcpyAsync(ctx->h_lines.p, ctx->work.out, ctx->n_bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(ctx->h_line_off.p, ctx->work.line_off, (ctx->n_lines + 1) * sizeof(uint64_t),
                     cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(ctx->h_doc_line_off.p, ctx->work.doc_line_off, ((size_t)n + 1) * sizeof(uint64_t),
                     cudaMemcpyDeviceToHost, ctx->stream));
  if (n)
    CU(cudaMemcpyAsync(ctx->h_status.p, ctx->work.doc_status, (size_t)n * sizeof(int32_t),
                       cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaEventRecord(ctx->ev[EV_D2H], ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  if (ctx->n_lines == 0) ctx->h_line_off.as<uint64_t>()[0] = 0;
  out->lines = ctx->h_lines.as<uint8_t>();
  out->n_bytes = ctx->n_bytes;
  out->line_off = ctx->h_line_off.as<uint64_t>();
  out->n_lines = ctx->n_lines;
  out->doc_line_off = ctx->h_doc_line_off.as<uint64_t>();
  out->doc_status = ctx->h_status.as<int32_t>();
  out->n_docs = n;
  return KGM_OK;
}

// Lays out a warp's shared-memory slice for this batch (FusedParams) and decides whether the
// fused kernel can run it.  Slices are sized for two CTAs of four warps per SM; the text
// windows, the template block and the per-document tables are clamped to that budget, and a
// document (or a node) that does not fit takes the generic path inside the kernel.  If
// fewer than 90 % of the documents fit that way, one CTA per SM with slices twice as large
// is used when that covers more of them.
struct FusedFit { FusedParams P; uint64_t fits; bool ok; };

FusedFit fused_layout(kgm_ctx *ctx, const kgm_batch *b, int ctas, uint32_t teams, uint64_t max_g, uint64_t max_a) {
  auto r16 = [](uint64_t v) { return (uint32_t)((v + 15) & ~15ull); };
  FusedFit f{};
  const uint32_t const_pad = (fast::ConstArea::BYTES + 15u) & ~15u;
  const uint32_t dyn_max = ctas == 2 ? 115712u : 232448u;
  const uint32_t warp_bytes = ((dyn_max - const_pad) / teams) & ~15u;
  const uint64_t need_stage = std::max<uint64_t>(MAX_LINE_BYTES + 64, 12ull * ctx->fleet.max_nc + 16);
  if (need_stage + 4096 > warp_bytes) return f;
  const uint64_t budget = warp_bytes - need_stage;            // texts + template block + tables
  // start from what the batch needs (bounded by what a slice could ever hold); the per-row
  // tables follow the text size; texts and template block are shrunk until everything fits
  uint64_t tb = std::min<uint64_t>(r16(ctx->fast_max_tb) + 32, budget * 60 / 100);
  uint64_t g = std::min<uint64_t>(max_g, 8192), a = std::min<uint64_t>(max_a, budget * 60 / 100);
  FusedParams &P = f.P;
  uint32_t o = 0;
  for (int round = 0; round < 12; ++round) {
    P = FusedParams{};
    P.cap_g = r16(g + 15) + 48;
    P.cap_a = r16(a + 15) + 48;
    P.cap_ng = (uint32_t)std::min<uint64_t>(32, std::max<uint64_t>(8, g / 64 + 2));
    P.cap_np = (uint32_t)std::min<uint64_t>(2048, std::max<uint64_t>(8, a / 64 + 2));
    P.cap_c = std::min(512u, std::max(1u, ctx->fast_max_nc));
    P.cap_t = r16(tb);
    P.cap_lines = P.cap_ng + P.cap_np + std::min(ctx->fast_max_idle, 512u);
    o = 0;
    P.off_mbox = o; o += 64;
    P.off_tg = o; o += P.cap_g;
    P.off_ta = o; o += P.cap_a;
    P.off_gstart = o; o += r16(2ull * (P.cap_ng + 2));
    P.off_pstart = o; o += r16(2ull * (P.cap_np + 2));
    P.off_grec = o; o += r16(sizeof(fast::GRec) * (uint64_t)P.cap_ng);
    P.off_prec = o; o += r16(sizeof(fast::PRec) * (uint64_t)P.cap_np);
    P.off_cnt = o; o += r16(4ull * P.cap_c);
    P.off_cstart = o; o += r16(4ull * P.cap_c);
    P.off_plan = o; o += r16(4ull * P.cap_lines);
    P.off_loff = o; o += r16(4ull * (P.cap_lines + 1));
    P.off_tb = o; o += P.cap_t;
    P.off_stage = o;
    if ((uint64_t)o + need_stage <= warp_bytes) break;
    const uint64_t over = (uint64_t)o + need_stage - warp_bytes;
    const uint64_t cut_a = std::min<uint64_t>(a > 1024 ? a - 1024 : 0, over * 2 / 5 + 64);
    const uint64_t cut_t = std::min<uint64_t>(tb > 1024 ? tb - 1024 : 0, over * 2 / 5 + 64);
    if (!cut_a && !cut_t) break;
    a -= cut_a; tb -= cut_t;
  }
  if ((uint64_t)o + need_stage > warp_bytes) return f;
  P.warp_bytes = warp_bytes;
  P.stage_bytes = warp_bytes - o;
  P.teams = teams;
  // how many documents can take the fast path with these caps (same tests as fast_plan);
  // no need to look at them one by one when even the longest texts and blocks fit
  if (ctx->all_nodes_fast && max_g + 30 + 32 <= P.cap_g && max_a + 30 + 32 <= P.cap_a &&
      (uint64_t)ctx->fast_max_tb + 16 <= P.cap_t) {
    f.fits = b->n_docs;
    f.ok = true;
    return f;
  }
  for (uint32_t d = 0; d < b->n_docs; ++d) {
    const uint32_t nd = b->doc_node[d];
    const uint64_t g_lo = b->gpu_off[d], g_hi = b->gpu_off[d + 1], a_lo = b->apps_off[d], a_hi = b->apps_off[d + 1];
    if (!ctx->h_node_fast[nd]) continue;
    if (((g_hi + 15) & ~15ull) - (g_lo & ~15ull) + 32 > P.cap_g) continue;
    if (((a_hi + 15) & ~15ull) - (a_lo & ~15ull) + 32 > P.cap_a) continue;
    if (ctx->h_tb_len[nd] + 16 > P.cap_t) continue;
    ++f.fits;
  }
  f.ok = true;
  return f;
}

void plan_fused(kgm_ctx *ctx, const kgm_batch *b, uint64_t max_g, uint64_t max_a) {
  ctx->fused_ok = false;
  if (b->n_docs == 0) return;
  // Candidates: (CTAs per SM, teams per CTA) = 2x4, 2x3, 1x7, 1x6, 1x5, 1x4, i.e. 8 .. 4
  // documents in flight per SM with slices of 28 .. 56 KB.  A candidate is usable if at
  // least 90 % of the documents fit its slices; preferred is the one whose stage holds a
  // whole document (lines are ~3.4x the text: otherwise the two warps of a team compose
  // half-empty rounds) and that needs the fewest waves of documents -- the look-back keeps
  // the teams roughly in step, so a batch takes ceil(documents / teams) waves.
  static const int cand[][2] = {{2, 4}, {1, 7}, {2, 3}, {1, 6}, {1, 5}, {1, 4}};
  const uint64_t est_out = (max_g + max_a) * 17 / 5 + 256;
  FusedFit best{};
  int best_ctas = 0;
  double best_cost = 0;
  uint64_t most_fits = 0;
  FusedFit fallback{};
  int fallback_ctas = 0;
  for (const auto &c : cand) {
    if (ctx->force_teams && (uint32_t)(c[0] * c[1]) != ctx->force_teams) continue;
    const FusedFit f = fused_layout(ctx, b, c[0], (uint32_t)c[1], max_g, max_a);
    if (!f.ok) continue;
    if (f.fits > most_fits || !fallback.ok) { most_fits = f.fits; fallback = f; fallback_ctas = c[0]; }
    if (f.fits * 10 < (uint64_t)b->n_docs * 9) continue;
    const uint64_t teams = (uint64_t)ctx->sms * c[0] * c[1];
    const double waves = (double)((b->n_docs + teams - 1) / teams);
    const double rounds = f.P.stage_bytes >= est_out ? 1.0 : 1.35;       // half-empty rounds cost
    const double cost = waves * rounds * (1.0 + 0.03 * (c[0] * c[1] - 4));   // more in flight: a little slower each
    if (!best.ok || cost < best_cost) { best = f; best_ctas = c[0]; best_cost = cost; }
  }
  if (!best.ok) { best = fallback; best_ctas = fallback_ctas; }
  if (!best.ok) return;
  ctx->fused = best.P;
  ctx->fused_ctas_per_sm = best_ctas;
  ctx->fused_ok = true;
}

// Lays out a warp's shared-memory slice of k_poll2 for this batch.  Persistent part: the two text
// windows, the row records, the per-container and per-line tables; behind them the template
// slots and the stage of a round, which double as the parse-time tables (row starts, join
// words) and as the generic path's scratch.  Text windows are clamped so that at least a few
// documents stay in flight per SM; a document that does not fit takes the generic path.
bool poll2_layout(kgm_ctx *ctx, uint32_t lines, uint64_t max_g, uint64_t max_a, Poll2Params *out) {
  auto r16 = [](uint64_t v) { return (uint32_t)((v + 15) & ~15ull); };
  const uint32_t ca = (p2::CA2::BYTES + 15u) & ~15u;
  const uint32_t dyn_max = 232448u - ca;
  uint64_t g = std::min<uint64_t>(max_g, 6144), a = std::min<uint64_t>(max_a, 36864);
  Poll2Params Q{};
  for (int round = 0; round < 8; ++round) {
    Q = Poll2Params{};
    Q.cap_g = r16(g + 30) + 48;
    Q.cap_a = r16(a + 30) + 48;
    // rows of the two texts are rarely shorter than 80 / 72 bytes (uuid 40 + five cells; pid, name,
    // uuid 40, memory); a document with more rows than that takes the generic path
    Q.cap_ng = (uint32_t)std::min<uint64_t>(32, std::max<uint64_t>(8, g / 80 + 2));
    Q.cap_np = (uint32_t)std::min<uint64_t>(4096, std::max<uint64_t>(8, a / 72 + 2));
    const uint32_t cap_idle = std::min(ctx->p2_max_idle, 1024u);
    Q.cap_rec = Q.cap_np + cap_idle;
    Q.cap_c = std::min(512u, std::max(1u, ctx->p2_max_nc));
    Q.cap_lines = Q.cap_ng + Q.cap_np + cap_idle;
    const uint32_t est_line = ctx->p2_max_fixed + 80u;         // constant bytes + typical variable cells
    Q.stage_bytes = r16((uint64_t)lines * est_line + 64);
    uint32_t o = 0;
    Q.off_tg = o; o += Q.cap_g;
    Q.off_ta = o; o += Q.cap_a;
    Q.off_grec = o; o += p2::GREC_BYTES * Q.cap_ng;
    Q.off_prec = o; o += p2::PREC_BYTES * Q.cap_rec;
    Q.off_cnt = o; o += r16(2ull * Q.cap_c);
    Q.off_cstart = o; o += r16(2ull * Q.cap_c);
    Q.off_plan = o; o += r16(2ull * Q.cap_lines);
    Q.off_loff = o; o += r16(4ull * (Q.cap_lines + 1));
    Q.off_toff = o; o += r16(4ull * Q.cap_lines);
    Q.off_stage = o;
    // parse-time tables and the generic path's scratch share the stage
    uint32_t t = o;
    Q.off_gstart = t; t += r16(2ull * (Q.cap_ng + 2));
    Q.off_pstart = t; t += r16(2ull * (Q.cap_np + 2));
    Q.off_cp = t; t += r16(4ull * Q.cap_np);
    Q.off_digest = t; t += r16(4ull * Q.cap_np);
    const uint32_t need = std::max<uint32_t>(t - o, r16(12ull * ctx->fleet.max_nc + 16));
    Q.stage_bytes = std::max(Q.stage_bytes, need);
    o += Q.stage_bytes;
    Q.slice_bytes = r16(o);
    if (Q.slice_bytes * 4u <= dyn_max) break;                  // at least 4 documents in flight per SM
    if (a > 4096) a = a * 3 / 4; else if (g > 1024) g = g * 3 / 4; else return false;
  }
  if (Q.slice_bytes * 4u > dyn_max) return false;
  Q.warps = std::min<uint32_t>((uint32_t)poll2_max_warps(), dyn_max / Q.slice_bytes);
  if (ctx->force_warps) Q.warps = std::max(1u, std::min(Q.warps, ctx->force_warps));
  *out = Q;
  return true;
}

void plan_poll2(kgm_ctx *ctx, const kgm_batch *b, uint64_t max_g, uint64_t max_a) {
  ctx->poll2_ok = false;
  if (b->n_docs == 0 || !ctx->p2_any) return;
  // lines per round: the stage holds that many typical lines (one lane writes one line)
  const uint32_t lines = ctx->force_lines ? ctx->force_lines : 16u;
  Poll2Params q{};
  if (!poll2_layout(ctx, lines, max_g, max_a, &q)) return;
  // Rounds: the pod lines first, the node lines last (the output order is the document's either
  // way): the look-back is resolved after the first round, and the longer that round, the fewer
  // documents wait there for their predecessors' sizes (measured on every bench configuration).
  q.pods_first = 1u;
  if (const char *env = getenv("KGM_POLL2_PODS_FIRST")) q.pods_first = atoi(env) ? 1u : 0u;
  // tickets per CTA (its warps start `warps` consecutive documents together and so run through the
  // same code at about the same time: fewer instruction-cache misses) -- equal or faster on every
  // bench configuration; KGM_POLL2_CTA_TICKETS=0 gives every warp its own ticket
  q.cta_tickets = 1u;
  if (const char *env = getenv("KGM_POLL2_CTA_TICKETS")) q.cta_tickets = atoi(env) ? 1u : 0u;
  ctx->poll2 = q;
  ctx->poll2_ok = true;
}

// Validates a batch, sizes the device buffers and the workspace guesses for it and points
// ctx->batch at them.  Nothing is copied.
int prepare_batch(kgm_ctx *ctx, const kgm_batch *b, uint64_t *gb_out, uint64_t *ab_out) {
  if (!ctx->fleet_loaded) return KGM_E_STATE;
  if (!b || (b->n_docs && (!b->gpu_off || !b->apps_off || !b->doc_node))) return KGM_E_INVALID;
  const uint32_t n = b->n_docs;
  static const uint64_t zero_off[1] = {0};
  const uint64_t *goff = n ? b->gpu_off : zero_off, *aoff = n ? b->apps_off : zero_off;
  if (goff[0] != 0 || aoff[0] != 0) return KGM_E_INVALID;
  uint64_t idle_cap = 0, max_g = 0, max_a = 0;
  for (uint32_t d = 0; d < n; ++d) {
    if (goff[d + 1] < goff[d] || aoff[d + 1] < aoff[d]) return KGM_E_INVALID;
    if (b->doc_node[d] >= ctx->fleet.n_nodes) return KGM_E_INVALID;
    idle_cap += ctx->node_idle_cap[b->doc_node[d]];
    max_g = std::max(max_g, goff[d + 1] - goff[d]);
    max_a = std::max(max_a, aoff[d + 1] - aoff[d]);
  }
  plan_fused(ctx, b, max_g, max_a);
  plan_poll2(ctx, b, max_g, max_a);
  ctx->have_ticks = b->doc_tick != nullptr;
  ctx->max_tick = 0;
  if (b->doc_tick) for (uint32_t d = 0; d < n; ++d) ctx->max_tick = std::max(ctx->max_tick, b->doc_tick[d]);
  const uint64_t gb = goff[n], ab = aoff[n];
  if (gb >= 0xfff00000ull || ab >= 0xfff00000ull) return KGM_E_LIMIT;
  if ((gb && !b->gpu_csv) || (ab && !b->apps_csv)) return KGM_E_INVALID;
  ctx->batch_resident = false;
  ctx->have_result = false;
  if (ctx->async_pending) { CU(cudaStreamSynchronize(ctx->stream)); ctx->async_pending = false; }
  CU(ctx->b_gpu.ensure(std::max<uint64_t>(gb, 1)));
  CU(ctx->b_apps.ensure(std::max<uint64_t>(ab, 1)));
  CU(ctx->b_gpu_off.ensure(((size_t)n + 1) * sizeof(uint64_t)));
  CU(ctx->b_apps_off.ensure(((size_t)n + 1) * sizeof(uint64_t)));
  CU(ctx->b_node.ensure(std::max<size_t>(n, 1) * sizeof(uint32_t)));
  if (b->doc_tick) CU(ctx->b_tick.ensure(std::max<size_t>(n, 1) * sizeof(uint32_t)));
  ctx->batch.n_docs = n;
  ctx->batch.gpu_csv = ctx->b_gpu.as<uint8_t>();
  ctx->batch.gpu_off = ctx->b_gpu_off.as<uint64_t>();
  ctx->batch.apps_csv = ctx->b_apps.as<uint8_t>();
  ctx->batch.apps_off = ctx->b_apps_off.as<uint64_t>();
  ctx->batch.doc_node = ctx->b_node.as<uint32_t>();
  ctx->batch.doc_tick = b->doc_tick ? ctx->b_tick.as<uint32_t>() : nullptr;
  ctx->gpu_bytes = gb; ctx->apps_bytes = ab; ctx->idle_cap = idle_cap;
  // capacity guesses; a kernel that runs out says so and the pipeline is re-run larger
  ctx->grow_cap = std::max<uint64_t>(ctx->grow_cap, gb / 48 + n + 64);
  ctx->prow_cap = std::max<uint64_t>(ctx->prow_cap, ab / 40 + n + 64);
  ctx->desc_cap = std::max<uint64_t>(ctx->desc_cap, gb / 48 + ab / 40 + idle_cap + n + 64);
  ctx->out_cap = std::max<uint64_t>(ctx->out_cap, 4 * (gb + ab) + 400 * idle_cap + (1u << 20));
  *gb_out = gb; *ab_out = ab;
  return KGM_OK;
}

// Copies the small per-document arrays of a prepared batch (offsets, node ids, ticks).
int upload_doc_arrays(kgm_ctx *ctx, const kgm_batch *b) {
  const uint32_t n = b->n_docs;
  static const uint64_t zero_off[1] = {0};
  int rc;
  if ((rc = upload(ctx, ctx->b_gpu_off, n ? b->gpu_off : zero_off, (size_t)n + 1))) return rc;
  if ((rc = upload(ctx, ctx->b_apps_off, n ? b->apps_off : zero_off, (size_t)n + 1))) return rc;
  if ((rc = upload(ctx, ctx->b_node, b->doc_node, n))) return rc;
  if (b->doc_tick && (rc = upload(ctx, ctx->b_tick, b->doc_tick, n))) return rc;
  return KGM_OK;
}

int upload_batch(kgm_ctx *ctx, const kgm_batch *b) {
  uint64_t gb = 0, ab = 0;
  int rc = prepare_batch(ctx, b, &gb, &ab);
  if (rc) return rc;
  CU(cudaEventRecord(ctx->ev[EV_START], ctx->stream));
  if ((rc = upload(ctx, ctx->b_gpu, b->gpu_csv, gb))) return rc;
  if ((rc = upload(ctx, ctx->b_apps, b->apps_csv, ab))) return rc;
  if ((rc = upload_doc_arrays(ctx, b))) return rc;
  ctx->batch_resident = true;
  return KGM_OK;
}

// End-to-end path for a large batch in host memory: the batch is cut into up to
// MAX_CHUNKS runs of documents (small ones first and last, so that the copies start early
// and the tail is short); the text copy of chunk c+1, the fused kernel of chunk c and the
// result copy of chunk c-1 run at the same time on three streams.  Every kernel takes the
// running {lines, bytes} of its predecessors from device memory, so the chunks write one
// contiguous result, identical to the one-launch result, and posts its own running totals
// into mapped host memory: the host learns the size of a chunk's lines without a stream
// synchronisation or a small copy queued behind the large ones, and queues the copy of
// chunk c while that of chunk c-1 is still running.  Returns 1 if a workspace guess was too
// small (the caller then runs the plain path on the resident batch).
int encode_pipelined(kgm_ctx *ctx, const kgm_batch *b, uint32_t chunks, kgm_result *out) {
  uint64_t gb = 0, ab = 0;
  static const bool trace = getenv("KGM_TRACE_HOST") != nullptr;
  auto now = []() { return std::chrono::steady_clock::now(); };
  auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b2) {
    return std::chrono::duration<double, std::micro>(b2 - a).count();
  };
  const auto t_begin = now();
  int rc = prepare_batch(ctx, b, &gb, &ab);
  const auto t_prepared = now();
  if (rc) return rc;
  // (the lines land in one pinned, mapped buffer of the size of the device's: beyond 2 GiB of
  // page-locked memory the plain path, which copies n_bytes only, takes the batch)
  if (!use_fused(ctx) || b->n_docs < 4 * chunks || ctx->out_cap + 16 > (2ull << 30)) {
    rc = upload_batch(ctx, b);
    return rc ? rc : 1;
  }
  const uint32_t n = b->n_docs;
  if ((rc = ensure_workspace(ctx))) return rc;
  constexpr size_t POST_WORDS = 8;                 // one 64-byte line per chunk
  CU(ctx->h_chunk.ensure(MAX_CHUNKS * (POST_WORDS * sizeof(uint64_t) + CTR_COUNT * sizeof(uint32_t))));
  CU(ctx->h_lines.ensure(ctx->out_cap + 16));
  CU(ctx->h_line_off.ensure(((size_t)ctx->desc_cap + 1) * sizeof(uint64_t)));
  CU(ctx->h_doc_line_off.ensure(((size_t)n + 1) * sizeof(uint64_t)));
  CU(ctx->h_status.ensure(std::max<size_t>(n, 1) * sizeof(int32_t)));
  volatile uint64_t *h_post = ctx->h_chunk.as<uint64_t>();
  uint32_t *h_cnt = reinterpret_cast<uint32_t *>(ctx->h_chunk.as<uint64_t>() + POST_WORDS * MAX_CHUNKS);
  void *d_post_v = nullptr;
  CU(cudaHostGetDevicePointer(&d_post_v, ctx->h_chunk.p, 0));
  unsigned long long *d_post = static_cast<unsigned long long *>(d_post_v);
  uint64_t *d_tot = ctx->w_totals.as<uint64_t>();

  // chunk boundaries: weights 1, 2, 3, 3, ..., 3, 2, 1
  uint32_t c0[MAX_CHUNKS + 1];
  {
    uint32_t wsum = 0, acc = 0;
    for (uint32_t c = 0; c < chunks; ++c) wsum += std::min({c + 1, chunks - c, 3u});
    c0[0] = 0;
    for (uint32_t c = 0; c < chunks; ++c) {
      acc += std::min({c + 1, chunks - c, 3u});
      c0[c + 1] = (uint32_t)((uint64_t)n * acc / wsum);
    }
  }

  CU(cudaEventRecord(ctx->ev[EV_START], ctx->stream));
  CU(cudaStreamWaitEvent(ctx->s_h2d, ctx->ev[EV_START], 0));      // earlier work on the buffers is done
  {
    // the small per-document arrays go through one pinned staging buffer (the caller's may
    // be pageable, and a pageable copy blocks the host), ahead of the first chunk's text
    const size_t o_g = 0, o_a = o_g + ((size_t)n + 1) * 8, o_n = o_a + ((size_t)n + 1) * 8, o_t = o_n + (size_t)n * 4;
    CU(ctx->h_docarr.ensure(o_t + (size_t)n * 4));
    uint8_t *st = ctx->h_docarr.as<uint8_t>();
    memcpy(st + o_g, b->gpu_off, ((size_t)n + 1) * 8);
    memcpy(st + o_a, b->apps_off, ((size_t)n + 1) * 8);
    memcpy(st + o_n, b->doc_node, (size_t)n * 4);
    if (b->doc_tick) memcpy(st + o_t, b->doc_tick, (size_t)n * 4);
    CU(cudaMemcpyAsync(ctx->b_gpu_off.p, st + o_g, ((size_t)n + 1) * 8, cudaMemcpyHostToDevice, ctx->s_h2d));
    CU(cudaMemcpyAsync(ctx->b_apps_off.p, st + o_a, ((size_t)n + 1) * 8, cudaMemcpyHostToDevice, ctx->s_h2d));
    CU(cudaMemcpyAsync(ctx->b_node.p, st + o_n, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->s_h2d));
    if (b->doc_tick) CU(cudaMemcpyAsync(ctx->b_tick.p, st + o_t, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->s_h2d));
  }
  // The text goes up in few, growing pieces (chunk 0 alone, then two chunks, then the rest):
  // while the result copy saturates the other direction every copy command is slow to
  // start, and the text of all chunks must be in long before the last kernels need it.
  uint32_t piece_of[MAX_CHUNKS];
  {
    const uint32_t piece_end[3] = {1u, std::min(3u, chunks), chunks};
    uint32_t from = 0;
    for (uint32_t p = 0; p < 3 && from < chunks; ++p) {
      const uint32_t to = piece_end[p];
      if (to <= from) continue;
      const uint64_t g0 = b->gpu_off[c0[from]], g1 = b->gpu_off[c0[to]];
      const uint64_t a0 = b->apps_off[c0[from]], a1 = b->apps_off[c0[to]];
      if (g1 > g0) CU(cudaMemcpyAsync(ctx->b_gpu.as<uint8_t>() + g0, b->gpu_csv + g0, g1 - g0, cudaMemcpyHostToDevice, ctx->s_h2d));
      if (a1 > a0) CU(cudaMemcpyAsync(ctx->b_apps.as<uint8_t>() + a0, b->apps_csv + a0, a1 - a0, cudaMemcpyHostToDevice, ctx->s_h2d));
      CU(cudaEventRecord(ctx->ev_in[p], ctx->s_h2d));
      for (uint32_t c = from; c < to; ++c) piece_of[c] = p;
      from = to;
    }
  }
  const BatchDev whole = ctx->batch;
  const WorkDev wfull = ctx->work;
  auto restore = [&]() {
    ctx->batch = whole; ctx->work = wfull; ctx->totals_external = false; ctx->chunk_launch = false;
    ctx->fused.carry = nullptr; ctx->fused.post = nullptr;
  };
  struct Restore {                                 // every way out of the chunk loop puts the whole batch back
    decltype(restore) &f;
    bool armed = true;
    ~Restore() { if (armed) f(); }
  } guard{restore};
  uint64_t seq[MAX_CHUNKS];
  const uint32_t *d_cnt[MAX_CHUNKS];
  static_assert(CTR_RING > MAX_CHUNKS + 1, "the counter sets of all chunks must survive the call");
  for (uint32_t c = 0; c < chunks; ++c) {
    const uint32_t lo = c0[c], cn = c0[c + 1] - c0[c];
    BatchDev &bd = ctx->batch;
    bd = whole;
    bd.n_docs = cn; bd.gpu_off = whole.gpu_off + lo; bd.apps_off = whole.apps_off + lo;
    bd.doc_node = whole.doc_node + lo; bd.doc_tick = whole.doc_tick ? whole.doc_tick + lo : nullptr;
    WorkDev &wd = ctx->work;
    wd = wfull;
    wd.doc_g0 += lo; wd.doc_ng += lo; wd.doc_p0 += lo; wd.doc_np += lo; wd.doc_perr += lo;
    wd.doc_desc0 += lo; wd.doc_nlines += lo; wd.doc_nbytes += lo; wd.doc_status += lo;
    wd.doc_line_off += lo; wd.doc_byte_off += lo;
    wd.totals = d_tot + 2 * (c + 1);
    ctx->totals_external = true;
    ctx->chunk_launch = true;
    ctx->fused.carry = c ? reinterpret_cast<const unsigned long long *>(d_tot + 2 * c) : nullptr;
    seq[c] = 1 + (ctx->post_seq++ % 65535u);
    h_post[POST_WORDS * c] = 0; h_post[POST_WORDS * c + 1] = 0;
    ctx->fused.post = d_post + POST_WORDS * c;
    ctx->fused.post_seq = seq[c];
    if (c == 0 || piece_of[c] != piece_of[c - 1]) CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_in[piece_of[c]], 0));
    if (c == 0) CU(cudaEventRecord(ctx->ev[EV_H2D], ctx->stream));
    rc = launch_pipeline(ctx);                     // cn > 0; sets wd.counters to this launch's set
    if (rc) return rc;
    d_cnt[c] = wd.counters;
    CU(cudaEventRecord(ctx->ev_k[c], ctx->stream));
  }
  restore();
  guard.armed = false;
  CU(cudaEventRecord(ctx->ev[EV_ENC], ctx->stream));
  const auto t_queued = now();

  uint64_t lines_done = 0, bytes_done = 0;
  bool overflow = false;
  for (uint32_t c = 0; c < chunks && !overflow; ++c) {
    volatile uint64_t *post = h_post + POST_WORDS * c;
    auto posted = [&]() { return (post[0] >> 48) == seq[c] && (post[1] >> 48) == seq[c]; };
    for (uint32_t spins = 0; !posted(); ++spins) {
      if ((spins & 1023u) == 1023u) {              // a failed launch never posts
        const cudaError_t q = cudaEventQuery(ctx->ev_k[c]);
        if (q == cudaSuccess) break;
        if (q != cudaErrorNotReady) return cuda_fail(ctx, q, "k_poll (chunked)");
      }
#if defined(__x86_64__) || defined(__i386__)
      __builtin_ia32_pause();
#endif
    }
    if (!posted()) { overflow = true; break; }     // finished without totals: take the plain path
    const uint64_t nl = post[0] & 0xffffffffffffull, nb = post[1] & 0xffffffffffffull;
    if (nb > ctx->work.out_cap || nl > ctx->desc_cap) { overflow = true; break; }   // (work.out_cap: the caller's buffer, if one is set)
    CU(cudaStreamWaitEvent(ctx->s_d2h, ctx->ev_k[c], 0));
    if (nb > bytes_done)
      CU(cudaMemcpyAsync(ctx->h_lines.as<uint8_t>() + bytes_done, ctx->work.out + bytes_done, nb - bytes_done,
                         cudaMemcpyDeviceToHost, ctx->s_d2h));
    lines_done = nl; bytes_done = nb;
  }
  if (!overflow) {
    // the line offsets (8 bytes per line) go in one copy at the end: under load every copy
    // command costs more than its bytes
    if (lines_done)
      CU(cudaMemcpyAsync(ctx->h_line_off.p, ctx->work.line_off, lines_done * sizeof(uint64_t), cudaMemcpyDeviceToHost,
                         ctx->s_d2h));
    for (uint32_t c = 0; c < chunks; ++c)
      CU(cudaMemcpyAsync(h_cnt + CTR_COUNT * c, d_cnt[c], CTR_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->s_d2h));
    CU(cudaMemcpyAsync(ctx->h_doc_line_off.p, ctx->work.doc_line_off, ((size_t)n + 1) * sizeof(uint64_t),
                       cudaMemcpyDeviceToHost, ctx->s_d2h));
    CU(cudaMemcpyAsync(ctx->h_status.p, ctx->work.doc_status, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost,
                       ctx->s_d2h));
    CU(cudaEventRecord(ctx->ev[EV_D2H], ctx->s_d2h));
  }
  const auto t_copies = now();
  CU(cudaStreamSynchronize(ctx->s_d2h));
  CU(cudaStreamSynchronize(ctx->s_h2d));
  CU(cudaStreamSynchronize(ctx->stream));
  if (trace)
    fprintf(stderr, "kgm host us: prepare %.1f, queue %.1f, posts+copies %.1f, final sync %.1f\n",
            us(t_begin, t_prepared), us(t_prepared, t_queued), us(t_queued, t_copies), us(t_copies, now()));
  ctx->batch_resident = true;                      // every text byte has been copied
  uint32_t slow = 0;
  if (!overflow) {
    for (uint32_t c = 0; c < chunks; ++c) {
      const uint32_t *cnt = h_cnt + CTR_COUNT * c;
      if (cnt[CTR_OVERFLOW]) overflow = true;
      slow += cnt[CTR_SLOW];
    }
  }
  if (overflow) {
    // the plain path re-runs the batch and grows whichever guess was too small
    return 1;
  }
  ctx->h_line_off.as<uint64_t>()[lines_done] = bytes_done;
  ctx->n_lines = lines_done; ctx->n_bytes = bytes_done;
  ctx->slow_docs = slow;
  ctx->have_result = true;
  ctx->last_chunks = chunks;
  out->lines = ctx->h_lines.as<uint8_t>();
  out->n_bytes = bytes_done;
  out->line_off = ctx->h_line_off.as<uint64_t>();
  out->n_lines = lines_done;
  out->doc_line_off = ctx->h_doc_line_off.as<uint64_t>();
  out->doc_status = ctx->h_status.as<int32_t>();
  out->n_docs = n;
  return KGM_OK;
}

}  // namespace

extern "C" {

int kgm_abi_version(void) { return KGM_ABI_VERSION; }

const char *kgm_strerror(int s) {
  switch (s) {
    case KGM_OK: return "ok";
    case KGM_E_INVALID: return "invalid argument";
    case KGM_E_CUDA: return "CUDA error";
    case KGM_E_NOMEM: return "out of memory";
    case KGM_E_LIMIT: return "input beyond a documented limit";
    case KGM_E_STATE: return "call order: fleet or batch not loaded";
    case KGM_DOC_KEY_ERROR: return "KeyError";
    case KGM_DOC_VALUE_ERROR: return "ValueError";
    case KGM_DOC_ZERO_DIVISION: return "ZeroDivisionError";
    case KGM_DOC_ATTRIBUTE_ERROR: return "AttributeError";
    case KGM_DOC_INDEX_ERROR: return "IndexError";
    case KGM_DOC_CSV_ERROR: return "Error";
    case KGM_DOC_JSON_ERROR: return "JSONDecodeError";
    case KGM_DOC_TYPE_ERROR: return "TypeError";
    case KGM_DOC_OTHER_ERROR: return "Exception";
    case KGM_DOC_UNSUPPORTED_QUOTE: return "unsupported: quoted CSV cell";
    case KGM_DOC_UNSUPPORTED_BYTE: return "unsupported: non-ASCII byte";
    case KGM_DOC_UNSUPPORTED_RANGE: return "unsupported: value or line out of range";
    default: return "unknown status";
  }
}

int kgm_create(kgm_ctx **out, int device) {
  if (!out) return KGM_E_INVALID;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) {
    cudaGetLastError();
    return KGM_E_CUDA;
  }
  kgm_ctx *ctx = new (std::nothrow) kgm_ctx();
  if (!ctx) return KGM_E_NOMEM;
  ctx->device = device;
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  for (int i = 0; i < EV_COUNT && e == cudaSuccess; ++i) e = cudaEventCreate(&ctx->ev[i]);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->s_h2d, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->s_d2h, cudaStreamNonBlocking);
  for (int i = 0; i < MAX_CHUNKS && e == cudaSuccess; ++i) {
    e = cudaEventCreate(&ctx->ev_in[i]);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev_k[i]);
  }
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_enc_done, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_copy_done, cudaEventDisableTiming);
  cudaDeviceProp prop;
  if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) { kgm_destroy(ctx); cudaGetLastError(); return KGM_E_CUDA; }
  ctx->sms = prop.multiProcessorCount;
  if (const char *env = getenv("KGM_RESERVE_CTAS")) ctx->reserve_ctas = std::max(0, std::min(atoi(env), 128));
  if (const char *env = getenv("KGM_FUSED_TEAMS")) {
    ctx->force_teams = (uint32_t)std::max(0, std::min(atoi(env), 8));
  }
  if (const char *env = getenv("KGM_POLL2_LINES")) ctx->force_lines = (uint32_t)std::max(4, std::min(atoi(env), 32));
  if (const char *env = getenv("KGM_POLL2_WARPS")) ctx->force_warps = (uint32_t)std::max(0, std::min(atoi(env), 20));
  if (const char *env = getenv("KGM_PIPELINE")) {
    if (!strcmp(env, "staged")) ctx->pipeline = KGM_PIPELINE_STAGED;
    else if (!strcmp(env, "fused-generic")) ctx->pipeline = KGM_PIPELINE_FUSED_GENERIC;
    else if (!strcmp(env, "fused-v1")) ctx->pipeline = KGM_PIPELINE_FUSED_V1;
    else if (!strcmp(env, "fused-v2")) ctx->pipeline = KGM_PIPELINE_FUSED_V2;
  }
  *out = ctx;
  return KGM_OK;
}

void kgm_destroy(kgm_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->s_d2h) cudaStreamSynchronize(ctx->s_d2h);
  DevBuf *dev[] = {&ctx->f_cont_begin, &ctx->f_nerr_parse, &ctx->f_nerr_join, &ctx->f_gated,
                   &ctx->f_err_inspect, &ctx->f_err_emit, &ctx->f_vis_begin, &ctx->f_vis_off,
                   &ctx->f_vis_bytes, &ctx->f_vis_hash, &ctx->f_node_piece, &ctx->f_node_fixed,
                   &ctx->f_cont_piece, &ctx->f_cont_fixed, &ctx->f_skel, &ctx->f_pid_table,
                   &ctx->b_gpu, &ctx->b_gpu_off, &ctx->b_apps, &ctx->b_apps_off, &ctx->b_node,
                   &ctx->b_tick, &ctx->w_grow, &ctx->w_prow, &ctx->w_uniq, &ctx->w_fstr,
                   &ctx->w_desc, &ctx->w_len, &ctx->w_loff, &ctx->w_counters, &ctx->w_doc32,
                   &ctx->w_status, &ctx->w_doc_line_off, &ctx->w_doc_byte_off, &ctx->w_totals,
                   &ctx->w_line_off, &ctx->w_out, &ctx->w_flush, &ctx->w_ticket, &ctx->w_scan,
                   &ctx->f_node_fast, &ctx->f_tb_off, &ctx->f_tb_len, &ctx->f_tb_pool, &ctx->f_pid16,
                   &ctx->f_vis_hash2, &ctx->f_node_fast2, &ctx->f_nslot_off, &ctx->f_tslot_off, &ctx->f_slot_pool, &ctx->f_ca2,
                   &ctx->w_win_samples, &ctx->w_win_groups, &ctx->w_win_temp, &ctx->w_win_scal};
  for (DevBuf *b : dev) b->release();
  HostBuf *host[] = {&ctx->h_small, &ctx->h_lines, &ctx->h_line_off, &ctx->h_doc_line_off, &ctx->h_status,
                     &ctx->h_chunk, &ctx->h_docarr, &ctx->h_win_groups};
  for (HostBuf *b : host) b->release();
  for (int i = 0; i < EV_COUNT; ++i) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
  for (int i = 0; i < MAX_CHUNKS; ++i) {
    if (ctx->ev_in[i]) cudaEventDestroy(ctx->ev_in[i]);
    if (ctx->ev_k[i]) cudaEventDestroy(ctx->ev_k[i]);
  }
  if (ctx->ev_enc_done) cudaEventDestroy(ctx->ev_enc_done);
  if (ctx->ev_copy_done) cudaEventDestroy(ctx->ev_copy_done);
  if (ctx->s_h2d) cudaStreamDestroy(ctx->s_h2d);
  if (ctx->s_d2h) cudaStreamDestroy(ctx->s_d2h);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  cudaGetLastError();
  delete ctx;
}

const char *kgm_last_cuda_error(kgm_ctx *ctx) { return ctx ? ctx->cuda_err.c_str() : ""; }

int kgm_load_fleet(kgm_ctx *ctx, const kgm_fleet *fl) {
  if (!ctx || !fl) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  const uint32_t nn = fl->n_nodes, nc = fl->n_containers;
  if (!offsets_ok(fl->node_name.off, nn) || !fl->cont_begin || !fl->node_err_parse || !fl->node_err_join)
    return KGM_E_INVALID;
  if (fl->cont_begin[0] != 0 || fl->cont_begin[nn] != nc) return KGM_E_INVALID;
  for (uint32_t i = 0; i < nn; ++i)
    if (fl->cont_begin[i + 1] < fl->cont_begin[i]) return KGM_E_INVALID;
  if (nc && (!fl->gated || !fl->err_inspect || !fl->err_emit)) return KGM_E_INVALID;
  if (!offsets_ok(fl->cid.off, nc) || !offsets_ok(fl->ns.off, nc) || !offsets_ok(fl->pod.off, nc) ||
      !offsets_ok(fl->user.off, nc) || !fl->pid_begin || !fl->vis_begin)
    return KGM_E_INVALID;
  for (uint32_t c = 0; c < nc; ++c)
    if (fl->pid_begin[c + 1] < fl->pid_begin[c] || fl->vis_begin[c + 1] < fl->vis_begin[c]) return KGM_E_INVALID;
  if (fl->pid_begin[0] != 0 || fl->vis_begin[0] != 0) return KGM_E_INVALID;
  const uint32_t npid = fl->pid_begin[nc], nvis = fl->vis_begin[nc];
  if (!offsets_ok(fl->pid.off, npid) || !offsets_ok(fl->vis.off, nvis)) return KGM_E_INVALID;
  if (nc >= (1u << 30)) return KGM_E_LIMIT;

  // ---- constant chunks of the two line templates (gpu.py:136-143, 206-216) ----
  SkelPool pool;
  std::vector<Piece> node_piece((size_t)nn * NODE_PIECES), cont_piece((size_t)nc * POD_PIECES);
  std::vector<uint32_t> node_fixed(nn), cont_fixed(nc);
  ctx->node_idle_cap.assign(nn, 0);
  uint32_t max_nc = 1;
  for (uint32_t n = 0; n < nn; ++n) {
    const std::string host = str_of(fl->node_name, n);
    node_fixed[n] = pool.node_pieces(host, &node_piece[(size_t)n * NODE_PIECES]);
    const uint32_t cb = fl->cont_begin[n], ce = fl->cont_begin[n + 1];
    max_nc = std::max(max_nc, ce - cb);
    for (uint32_t c = cb; c < ce; ++c) {
      const std::string cid = str_of(fl->cid, c), ns = str_of(fl->ns, c), pod = str_of(fl->pod, c),
                        user = str_of(fl->user, c);
      cont_fixed[c] = pool.pod_pieces(host, ns, cid, user, pod, &cont_piece[(size_t)c * POD_PIECES]);
      if (fl->gated[c]) ctx->node_idle_cap[n] += fl->vis_begin[c + 1] - fl->vis_begin[c];
    }
    if (pool.bytes.size() >= 0xf0000000ull) return KGM_E_LIMIT;
  }
  if (max_nc > 4096) return KGM_E_LIMIT;
  while (pool.bytes.size() & 15) pool.bytes.push_back(0);

  // ---- pid -> container table (gpu.py:177-178: `proc in pids`, exact string match) ----
  uint32_t tsize = 16;
  while (tsize < 2ull * npid + 1) tsize <<= 1;
  std::vector<PidSlot> table(tsize);
  for (PidSlot &s : table) { s.lo = s.hi = 0; s.node = NONE32; s.crank = 0; s.pad[0] = s.pad[1] = 0; }
  for (uint32_t n = 0; n < nn; ++n) {
    for (uint32_t c = fl->cont_begin[n]; c < fl->cont_begin[n + 1]; ++c) {
      for (uint32_t k = fl->pid_begin[c]; k < fl->pid_begin[c + 1]; ++k) {
        const uint32_t len = fl->pid.off[k + 1] - fl->pid.off[k];
        if (len > 15) return KGM_E_LIMIT;
        const uint8_t *s = fl->pid.bytes + fl->pid.off[k];
        uint64_t lo = 0, hi = (uint64_t)len << 56;
        for (uint32_t i = 0; i < len; ++i) {
          if (i < 8) lo |= (uint64_t)s[i] << (8 * i);
          else hi |= (uint64_t)s[i] << (8 * (i - 8));
        }
        const uint32_t crank = c - fl->cont_begin[n];
        uint32_t h = (uint32_t)pid_hash(lo, hi, n) & (tsize - 1);
        bool dup = false;
        while (table[h].node != NONE32) {
          if (table[h].node == n && table[h].lo == lo && table[h].hi == hi && table[h].crank == crank) { dup = true; break; }
          h = (h + 1) & (tsize - 1);
        }
        if (!dup) { table[h].lo = lo; table[h].hi = hi; table[h].node = n; table[h].crank = crank; }
      }
    }
  }
  std::vector<uint64_t> vis_hash(std::max(nvis, 1u)), vis_hash2(std::max(nvis, 1u));
  for (uint32_t v = 0; v < nvis; ++v) {
    vis_hash[v] = hash_bytes(fl->vis.bytes + fl->vis.off[v], fl->vis.off[v + 1] - fl->vis.off[v]);
    vis_hash2[v] = fast::key_hash_bytes(fl->vis.bytes + fl->vis.off[v], fl->vis.off[v + 1] - fl->vis.off[v]);
  }

  // ---- tables of the fused kernel's fast path: per node one template block (header,
  //      container directory, the variable pieces of its line templates) and a compact
  //      pid table.  A node with any raise point on the docker side, a pid string longer
  //      than 7 bytes or an oversized block is left to the generic path. ----
  std::vector<uint8_t> node_fast(std::max(nn, 1u), 0), tb_pool;
  std::vector<uint32_t> tb_off(std::max(nn, 1u), 0), tb_len(std::max(nn, 1u), 0);
  uint32_t fast_max_nc = 1, fast_max_tb = 0, fast_max_idle = 0, any_fast = 0;
  uint64_t n_short_pids = 0;
  for (uint32_t n = 0; n < nn; ++n) {
    const uint32_t cb = fl->cont_begin[n], ce = fl->cont_begin[n + 1], ncn = ce - cb;
    bool ok = !fl->node_err_parse[n] && !fl->node_err_join[n] && ncn <= 512;
    for (uint32_t c = cb; c < ce && ok; ++c) {
      if (fl->err_inspect[c] || fl->err_emit[c]) ok = false;
      for (uint32_t k = fl->pid_begin[c]; k < fl->pid_begin[c + 1] && ok; ++k)
        if (fl->pid.off[k + 1] - fl->pid.off[k] > 7) ok = false;
    }
    if (ctx->node_idle_cap[n] > 2048) ok = false;
    if (!ok) continue;
    std::vector<TbContIn> conts(ncn);
    for (uint32_t c = cb; c < ce; ++c) {
      TbContIn &in = conts[c - cb];
      in.gated = fl->gated[c] != 0;
      in.cid = str_of(fl->cid, c); in.ns = str_of(fl->ns, c); in.pod = str_of(fl->pod, c);
      in.user = str_of(fl->user, c);
      in.vis_begin = fl->vis_begin[c];
      in.nvis = fl->vis_begin[c + 1] - fl->vis_begin[c];
      in.fixed = cont_fixed[c];
    }
    size_t start = 0, total = 0;
    if (!build_template_block(str_of(fl->node_name, n), node_fixed[n], conts, tb_pool, &start, &total)) continue;
    if (tb_pool.size() >= 0xf0000000ull) return KGM_E_LIMIT;
    node_fast[n] = 1; any_fast = 1;
    tb_off[n] = (uint32_t)start; tb_len[n] = (uint32_t)total;
    fast_max_nc = std::max(fast_max_nc, ncn);
    fast_max_tb = std::max<uint32_t>(fast_max_tb, (uint32_t)total);
    fast_max_idle = std::max<uint32_t>(fast_max_idle, (uint32_t)ctx->node_idle_cap[n]);
    for (uint32_t c = cb; c < ce; ++c) n_short_pids += fl->pid_begin[c + 1] - fl->pid_begin[c];
  }
  for (int i = 0; i < 64; ++i) tb_pool.push_back(0);            // over-read slack
  uint32_t t16 = 16;
  while (t16 < 2ull * n_short_pids + 1) t16 <<= 1;
  std::vector<fast::PidSlot16> table16(t16);
  for (fast::PidSlot16 &s16 : table16) { s16.key = 0; s16.node = NONE32; s16.crank = 0; }
  for (uint32_t n = 0; n < nn; ++n) {
    if (!node_fast[n]) continue;
    for (uint32_t c = fl->cont_begin[n]; c < fl->cont_begin[n + 1]; ++c) {
      for (uint32_t k = fl->pid_begin[c]; k < fl->pid_begin[c + 1]; ++k) {
        const uint32_t len = fl->pid.off[k + 1] - fl->pid.off[k];
        const uint8_t *ps = fl->pid.bytes + fl->pid.off[k];
        uint64_t key = (uint64_t)len << 56;
        for (uint32_t i = 0; i < len; ++i) key |= (uint64_t)ps[i] << (8 * i);
        const uint32_t crank = c - fl->cont_begin[n];
        uint32_t h = fast::pid16_hash(key, n) & (t16 - 1);
        bool dup = false;
        while (table16[h].node != NONE32) {
          if (table16[h].node == n && table16[h].key == key && table16[h].crank == crank) { dup = true; break; }
          h = (h + 1) & (t16 - 1);
        }
        if (!dup) { table16[h].key = key; table16[h].node = n; table16[h].crank = crank; }
      }
    }
  }

  // ---- template slots of k_poll2: one per node, one per gated container.  A node with any raise
  //      point on the docker side or a pid string longer than 7 bytes is left to the generic path. ----
  std::vector<uint8_t> node_fast2(std::max(nn, 1u), 0), slot_pool;
  std::vector<uint32_t> nslot_off(std::max(nn, 1u), 0), tslot_off(std::max(nc, 1u), 0);
  uint32_t p2_max_nc = 1, p2_max_slot = 64, p2_max_fixed = 0, p2_max_idle = 0, p2_any = 0;
  for (uint32_t n = 0; n < nn; ++n) {
    const uint32_t cb = fl->cont_begin[n], ce = fl->cont_begin[n + 1], ncn = ce - cb;
    bool ok = !fl->node_err_parse[n] && !fl->node_err_join[n] && ncn <= 512 && ctx->node_idle_cap[n] <= 1024;
    for (uint32_t c = cb; c < ce && ok; ++c) {
      if (fl->err_inspect[c] || fl->err_emit[c]) ok = false;
      for (uint32_t k = fl->pid_begin[c]; k < fl->pid_begin[c + 1] && ok; ++k)
        if (fl->pid.off[k + 1] - fl->pid.off[k] > 7) ok = false;
    }
    if (!ok || !node_fast[n]) continue;                // (the compact pid table holds the v1-fast nodes' pids)
    const size_t mark = slot_pool.size();
    const std::string host = str_of(fl->node_name, n);
    size_t start = 0, len = 0;
    uint32_t max_slot = 0, max_fixed = node_fixed[n] + 1;
    ok = build_node_slot(host, slot_pool, &start, &len);
    if (ok) { nslot_off[n] = (uint32_t)(start >> 4); max_slot = (uint32_t)len; }
    for (uint32_t c = cb; c < ce && ok; ++c) {
      if (!fl->gated[c]) continue;
      ok = build_pod_slot(host, str_of(fl->ns, c), str_of(fl->cid, c), str_of(fl->user, c), str_of(fl->pod, c),
                          slot_pool, &start, &len);
      if (ok) {
        tslot_off[c] = (uint32_t)(start >> 4);
        max_slot = std::max<uint32_t>(max_slot, (uint32_t)len);
        max_fixed = std::max(max_fixed, cont_fixed[c] + 1);
      }
    }
    if (!ok || max_slot > 4096 || slot_pool.size() >= 0xf0000000ull) { slot_pool.resize(mark); continue; }
    node_fast2[n] = 1; p2_any = 1;
    p2_max_nc = std::max(p2_max_nc, ncn);
    p2_max_slot = std::max(p2_max_slot, max_slot);
    p2_max_fixed = std::max(p2_max_fixed, max_fixed);
    p2_max_idle = std::max<uint32_t>(p2_max_idle, (uint32_t)ctx->node_idle_cap[n]);
  }
  slot_pool.resize(slot_pool.size() + p2_max_slot + 64, 0);      // every slot can be read at the largest slot's size

  ctx->fleet_loaded = false;
  ctx->batch_resident = false;
  ctx->have_result = false;
  int rc;
  if ((rc = upload(ctx, ctx->f_node_fast2, node_fast2.data(), node_fast2.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_nslot_off, nslot_off.data(), nslot_off.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_tslot_off, tslot_off.data(), tslot_off.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_slot_pool, slot_pool.data(), slot_pool.size()))) return rc;
  {
    std::vector<uint8_t> ca2((p2::CA2::BYTES + 15u) & ~15u, 0);   // the constant area every CTA of k_poll2 starts from
    p2::fill_ca2(ca2.data(), 0u);
    if ((rc = upload(ctx, ctx->f_ca2, ca2.data(), ca2.size()))) return rc;
  }
  if ((rc = upload(ctx, ctx->f_cont_begin, fl->cont_begin, (size_t)nn + 1))) return rc;
  if ((rc = upload(ctx, ctx->f_nerr_parse, fl->node_err_parse, nn))) return rc;
  if ((rc = upload(ctx, ctx->f_nerr_join, fl->node_err_join, nn))) return rc;
  if ((rc = upload(ctx, ctx->f_gated, fl->gated, nc))) return rc;
  if ((rc = upload(ctx, ctx->f_err_inspect, fl->err_inspect, nc))) return rc;
  if ((rc = upload(ctx, ctx->f_err_emit, fl->err_emit, nc))) return rc;
  if ((rc = upload(ctx, ctx->f_vis_begin, fl->vis_begin, (size_t)nc + 1))) return rc;
  if ((rc = upload(ctx, ctx->f_vis_off, fl->vis.off, (size_t)nvis + 1))) return rc;
  if ((rc = upload(ctx, ctx->f_vis_bytes, fl->vis.bytes, fl->vis.off[nvis]))) return rc;
  if ((rc = upload(ctx, ctx->f_vis_hash, vis_hash.data(), nvis))) return rc;
  if ((rc = upload(ctx, ctx->f_node_piece, node_piece.data(), node_piece.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_node_fixed, node_fixed.data(), nn))) return rc;
  if ((rc = upload(ctx, ctx->f_cont_piece, cont_piece.data(), cont_piece.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_cont_fixed, cont_fixed.data(), nc))) return rc;
  if ((rc = upload(ctx, ctx->f_skel, pool.bytes.data(), pool.bytes.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_pid_table, table.data(), table.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_node_fast, node_fast.data(), node_fast.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_tb_off, tb_off.data(), tb_off.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_tb_len, tb_len.data(), tb_len.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_tb_pool, tb_pool.data(), tb_pool.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_pid16, table16.data(), table16.size()))) return rc;
  if ((rc = upload(ctx, ctx->f_vis_hash2, vis_hash2.data(), nvis))) return rc;
  CU(cudaStreamSynchronize(ctx->stream));        // the host vectors above go out of scope
  FleetDev &f = ctx->fleet;
  f.n_nodes = nn; f.n_containers = nc; f.max_nc = max_nc;
  f.cont_begin = ctx->f_cont_begin.as<uint32_t>();
  f.node_err_parse = ctx->f_nerr_parse.as<uint8_t>();
  f.node_err_join = ctx->f_nerr_join.as<uint8_t>();
  f.gated = ctx->f_gated.as<uint8_t>();
  f.err_inspect = ctx->f_err_inspect.as<uint8_t>();
  f.err_emit = ctx->f_err_emit.as<uint8_t>();
  f.vis_begin = ctx->f_vis_begin.as<uint32_t>();
  f.vis_off = ctx->f_vis_off.as<uint32_t>();
  f.vis_bytes = ctx->f_vis_bytes.as<uint8_t>();
  f.vis_hash = ctx->f_vis_hash.as<uint64_t>();
  f.node_piece = ctx->f_node_piece.as<Piece>();
  f.node_fixed = ctx->f_node_fixed.as<uint32_t>();
  f.cont_piece = ctx->f_cont_piece.as<Piece>();
  f.cont_fixed = ctx->f_cont_fixed.as<uint32_t>();
  f.skel = ctx->f_skel.as<uint8_t>();
  f.pid_table = ctx->f_pid_table.as<PidSlot>();
  f.pid_mask = tsize - 1;
  f.node_fast = ctx->f_node_fast.as<uint8_t>();
  f.tb_off = ctx->f_tb_off.as<uint32_t>();
  f.tb_len = ctx->f_tb_len.as<uint32_t>();
  f.tb_pool = ctx->f_tb_pool.as<uint8_t>();
  f.pid16 = ctx->f_pid16.as<fast::PidSlot16>();
  f.pid16_mask = t16 - 1;
  f.vis_hash2 = ctx->f_vis_hash2.as<uint64_t>();
  f.node_fast2 = ctx->f_node_fast2.as<uint8_t>();
  f.nslot_off = ctx->f_nslot_off.as<uint32_t>();
  f.tslot_off = ctx->f_tslot_off.as<uint32_t>();
  f.slot_pool = ctx->f_slot_pool.as<uint8_t>();
  f.ca2 = ctx->f_ca2.as<uint4>();
  ctx->p2_max_nc = p2_max_nc; ctx->p2_max_slot = p2_max_slot; ctx->p2_max_fixed = p2_max_fixed;
  ctx->p2_max_idle = p2_max_idle; ctx->p2_any = p2_any;
  ctx->h_node_fast2 = node_fast2;
  ctx->all_nodes_fast2 = std::all_of(node_fast2.begin(), node_fast2.end(), [](uint8_t v) { return v != 0; });
  ctx->poll2_ok = false;
  ctx->fast_max_nc = fast_max_nc; ctx->fast_max_tb = fast_max_tb; ctx->fast_max_idle = fast_max_idle;
  ctx->any_fast = any_fast;
  ctx->h_node_fast = node_fast;
  ctx->all_nodes_fast = std::all_of(node_fast.begin(), node_fast.end(), [](uint8_t v) { return v != 0; });
  ctx->h_tb_len = tb_len;
  ctx->fused_ok = false;
  ctx->fleet_loaded = true;
  return KGM_OK;
}

int kgm_upload_batch(kgm_ctx *ctx, const kgm_batch *b) {
  if (!ctx) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  int rc = upload_batch(ctx, b);
  if (rc) return rc;
  CU(cudaStreamSynchronize(ctx->stream));
  return KGM_OK;
}

int kgm_encode_resident(kgm_ctx *ctx, kgm_result *out) {
  if (!ctx || !out) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  ctx->async_pending = false;
  int rc = run_resident(ctx, false);
  if (rc) return rc;
  fill_timing(ctx, false, false);
  fill_result_dev(ctx, out);
  return KGM_OK;
}

int kgm_encode_resident_async(kgm_ctx *ctx, kgm_result *out) {
  if (!ctx || !out) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  ctx->have_result = false;
  int rc = launch_resident(ctx, true);
  if (rc) return rc;
  ctx->async_pending = true;
  ctx->n_lines = 0; ctx->n_bytes = 0;
  fill_result_dev(ctx, out);                        // the buffers are known, the sizes are not
  return KGM_OK;
}

int kgm_encode_finish(kgm_ctx *ctx, kgm_result *out) {
  if (!ctx || !out) return KGM_E_INVALID;
  if (!ctx->async_pending) return KGM_E_STATE;
  CU(cudaSetDevice(ctx->device));
  ctx->async_pending = false;
  int rc = finish_resident(ctx);
  if (rc) return rc;
  fill_timing(ctx, false, false);
  fill_result_dev(ctx, out);
  return KGM_OK;
}

int kgm_set_output_buffer(kgm_ctx *ctx, void *dev_ptr, uint64_t bytes) {
  if (!ctx) return KGM_E_INVALID;
  if (!dev_ptr) { ctx->ext_out = nullptr; ctx->ext_out_bytes = 0; return KGM_OK; }
  if (((uintptr_t)dev_ptr & 15u) || bytes < 64) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  cudaPointerAttributes attr{};
  if (cudaPointerGetAttributes(&attr, dev_ptr) != cudaSuccess) { cudaGetLastError(); return KGM_E_INVALID; }
  if (attr.type == cudaMemoryTypeDevice && attr.device != ctx->device && attr.device >= 0 && attr.device < 64 &&
      !((ctx->peers_enabled >> attr.device) & 1ull)) {
    // a peer GPU's memory: this context's kernels need peer access to it
    const cudaError_t e = cudaDeviceEnablePeerAccess(attr.device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(ctx, e, "cudaDeviceEnablePeerAccess");
    cudaGetLastError();
    ctx->peers_enabled |= 1ull << attr.device;
  }
  ctx->ext_out = static_cast<uint8_t *>(dev_ptr);
  ctx->ext_out_bytes = bytes & ~15ull;
  return KGM_OK;
}

int kgm_push_step(kgm_ctx *ctx, void *slot, uint64_t slot_bytes, const uint32_t *wait_flag, uint32_t wait_value,
                  uint32_t *signal_flag, uint32_t signal_value, kgm_result *out) {
  if (!ctx || !out || !slot) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  int rc;
  if (slot != ctx->ext_out || (slot_bytes & ~15ull) != ctx->ext_out_bytes) {
    if ((rc = kgm_set_output_buffer(ctx, slot, slot_bytes))) return rc;
  }
  if (wait_flag && (rc = kgm_stream_wait_flags(ctx->stream, wait_flag, 1, wait_value))) return rc;
  ctx->have_result = false;
  if ((rc = launch_resident(ctx, true))) return rc;
  ctx->async_pending = true;
  ctx->n_lines = 0; ctx->n_bytes = 0;
  fill_result_dev(ctx, out);
  if (signal_flag) {
    uint32_t *t[1] = {signal_flag};
    if ((rc = kgm_stream_signal_flags(ctx->stream, t, 1, signal_value))) return rc;
  }
  return KGM_OK;
}

int kgm_push_copy_step(kgm_ctx *ctx, void *slot, uint64_t slot_bytes, const uint32_t *wait_flag, uint32_t wait_value,
                       uint32_t *signal_flag, uint32_t signal_value, kgm_result *out) {
  if (!ctx || !out || !slot || (reinterpret_cast<uintptr_t>(slot) & 15)) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  int rc;
  if (ctx->ext_out && (rc = kgm_set_output_buffer(ctx, nullptr, 0))) return rc;    // the context's own buffer
  if (ctx->copy_pending) CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_copy_done, 0));   // ... once its last copy has left it
  ctx->have_result = false;
  if ((rc = launch_resident(ctx, true))) return rc;
  ctx->async_pending = true;
  ctx->n_lines = 0; ctx->n_bytes = 0;
  fill_result_dev(ctx, out);
  CU(cudaEventRecord(ctx->ev_enc_done, ctx->stream));
  CU(cudaStreamWaitEvent(ctx->s_d2h, ctx->ev_enc_done, 0));
  if (wait_flag && (rc = kgm_stream_wait_flags(ctx->s_d2h, wait_flag, 1, wait_value))) return rc;
  const uint64_t have = ((ctx->out_cap + 16 + OUT_HEADER) & ~15ull);
  CU(cudaMemcpyAsync(slot, ctx->work.out - OUT_HEADER, std::min<uint64_t>(slot_bytes & ~15ull, have), cudaMemcpyDefault,
                     ctx->s_d2h));
  if (signal_flag) {
    uint32_t *t[1] = {signal_flag};
    if ((rc = kgm_stream_signal_flags(ctx->s_d2h, t, 1, signal_value))) return rc;
  }
  CU(cudaEventRecord(ctx->ev_copy_done, ctx->s_d2h));
  ctx->copy_pending = true;
  return KGM_OK;
}

void *kgm_stream(kgm_ctx *ctx) { return ctx ? static_cast<void *>(ctx->stream) : nullptr; }

uint64_t kgm_output_capacity(kgm_ctx *ctx) {
  if (!ctx || !ctx->work.out) return 0;
  return ctx->ext_out ? ctx->ext_out_bytes - OUT_HEADER : ctx->out_cap + 16;
}

int kgm_fetch_result(kgm_ctx *ctx, kgm_result *out) {
  if (!ctx || !out) return KGM_E_INVALID;
  if (!ctx->have_result) return KGM_E_STATE;
  CU(cudaSetDevice(ctx->device));
  return fetch(ctx, out);
}

int kgm_encode_batch(kgm_ctx *ctx, const kgm_batch *b, int want_host_copy, kgm_result *out) {
  if (!ctx || !out) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  ctx->last_chunks = 1;
  // large host batches: chunked, copies overlapped with the fused kernel
  if (want_host_copy && b && ctx->fleet_loaded && ctx->pipeline != KGM_PIPELINE_STAGED && b->n_docs >= 1024 &&
      b->gpu_off && b->apps_off && b->doc_node &&
      b->gpu_off[b->n_docs] + b->apps_off[b->n_docs] >= (4u << 20)) {
    const uint64_t total = b->gpu_off[b->n_docs] + b->apps_off[b->n_docs];
    uint32_t chunks = (uint32_t)std::min<uint64_t>(MAX_CHUNKS, std::max<uint64_t>(2, total / (1u << 20)));
    if (const char *env = getenv("KGM_CHUNKS")) chunks = (uint32_t)std::max(2, std::min(atoi(env), MAX_CHUNKS));
    int rc = encode_pipelined(ctx, b, chunks, out);
    if (rc <= 0) {
      if (rc == KGM_OK) {
        ctx->timing = kgm_timing{};
        cudaEventElapsedTime(&ctx->timing.total_ms, ctx->ev[EV_START], ctx->ev[EV_D2H]);   // copies and kernels overlap
        // the three phases overlap; these are the spans that do not: until the first
        // chunk's text is in, from there to the last kernel, and the copy tail after it
        cudaEventElapsedTime(&ctx->timing.h2d_ms, ctx->ev[EV_START], ctx->ev[EV_H2D]);
        cudaEventElapsedTime(&ctx->timing.encode_ms, ctx->ev[EV_H2D], ctx->ev[EV_ENC]);
        cudaEventElapsedTime(&ctx->timing.d2h_ms, ctx->ev[EV_ENC], ctx->ev[EV_D2H]);
        ctx->timing.launches = ctx->last_chunks;
        ctx->timing.fused = 1;
        ctx->timing.slow_docs = ctx->slow_docs;
        ctx->timing.fused_ctas_per_sm = ctx->last_poll2 ? 1u : (uint32_t)ctx->fused_ctas_per_sm;
        ctx->timing.docs_in_flight_per_sm = ctx->last_poll2 ? ctx->poll2.warps : ctx->fused.teams * ctx->fused_ctas_per_sm;
        cudaGetLastError();
      }
      return rc;
    }
    // a workspace guess was too small (or the batch does not fit the fused kernel): the
    // batch is resident now, run it the plain way
    rc = run_resident(ctx, false);
    if (rc) return rc;
    rc = fetch(ctx, out);
    if (rc) return rc;
    fill_timing(ctx, false, true);
    return KGM_OK;
  }
  int rc = upload_batch(ctx, b);          // records EV_START, copies are asynchronous
  if (rc) return rc;
  rc = run_resident(ctx, true);
  if (rc) return rc;
  if (want_host_copy) {
    rc = fetch(ctx, out);
    if (rc) return rc;
  } else {
    fill_result_dev(ctx, out);
  }
  fill_timing(ctx, true, want_host_copy != 0);
  return KGM_OK;
}

extern "C++" {
namespace {

// Runs the staged tokenizer and join only (token tables + line plan stay in the workspace).
int run_plan_only(kgm_ctx *ctx, uint32_t *n_desc) {
  if (!ctx->fleet_loaded || !ctx->batch_resident) return KGM_E_STATE;
  CU(ctx->h_small.ensure(64));
  for (int attempt = 0; attempt < 6; ++attempt) {
    int rc = ensure_workspace(ctx);
    if (rc) return rc;
    ctx->work.counters = ctx->w_counters.as<uint32_t>();
    ctx->counters_clean = false;
    CU(cudaMemsetAsync(ctx->work.counters, 0, CTR_COUNT * sizeof(uint32_t), ctx->stream));
    launch_tokenize(ctx->batch, ctx->work, ctx->sms, ctx->stream);
    launch_join(ctx->batch, ctx->fleet, ctx->work, ctx->sms, ctx->stream);
    CU(cudaGetLastError());
    uint32_t *hc = ctx->h_small.as<uint32_t>();
    CU(cudaMemcpyAsync(hc, ctx->work.counters, CTR_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    const uint32_t ovf = hc[CTR_OVERFLOW];
    if (!ovf) { *n_desc = hc[CTR_DESC]; return KGM_OK; }
    if (ovf & OVF_GROW) ctx->grow_cap = std::max(ctx->grow_cap * 2, hc[CTR_GROW] + 64);
    if (ovf & OVF_PROW) ctx->prow_cap = std::max(ctx->prow_cap * 2, hc[CTR_PROW] + 64);
    if (ovf & OVF_DESC) ctx->desc_cap = std::max(ctx->desc_cap * 2, hc[CTR_DESC] + 64);
  }
  return KGM_E_LIMIT;
}

// Carves 256-byte aligned arrays out of one allocation.
struct Carver {
  uint8_t *base;
  size_t off = 0;
  template <class T> T *take(size_t n) {
    T *p = reinterpret_cast<T *>(base + off);
    off += (n * sizeof(T) + 255) & ~(size_t)255;
    return p;
  }
};

WinGroups carve_groups(uint8_t *base, size_t n, size_t *bytes) {
  Carver c{base};
  WinGroups g{};
  g.kind = c.take<uint8_t>(n);
  g.node = c.take<uint32_t>(n); g.container = c.take<uint32_t>(n); g.gpu_id = c.take<uint32_t>(n);
  g.window = c.take<uint32_t>(n); g.count = c.take<uint32_t>(n);
  for (int m = 0; m < 3; ++m) {
    g.m_count[m] = c.take<uint32_t>(n); g.m_min[m] = c.take<uint32_t>(n); g.m_max[m] = c.take<uint32_t>(n);
    g.m_last[m] = c.take<uint32_t>(n); g.m_sum[m] = c.take<uint64_t>(n); g.m_mean[m] = c.take<double>(n);
  }
  g.key = c.take<uint64_t>(n);
  for (int m = 0; m < 3; ++m) g.m_order[m] = c.take<uint64_t>(n);
  *bytes = c.off;
  return g;
}

}  // namespace
}  // extern "C++"

int kgm_window_reduce(kgm_ctx *ctx, uint32_t window_ticks, int want_host_copy, kgm_windows *out) {
  if (!ctx || !out || window_ticks == 0) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  if (!ctx->fleet_loaded || !ctx->batch_resident) return KGM_E_STATE;
  if (!ctx->have_ticks) return KGM_E_INVALID;                  // the batch was uploaded without doc_tick
  if (ctx->max_tick / window_ticks > 0xffffu || ctx->max_tick >= (1u << 20) || ctx->batch.n_docs >= (1u << 26) ||
      ctx->fleet.n_containers >= 0x7fffffffu)
    return KGM_E_LIMIT;
  *out = kgm_windows{};
  uint32_t *h32 = ctx->h_small.as<uint32_t>();
  auto carve_samples = [&](size_t n, WinSamples *s) -> int {
    const size_t sample_bytes = 6 * ((n * 8 + 255) & ~(size_t)255) + 4096;
    CU(ctx->w_win_samples.ensure(sample_bytes));
    Carver c{ctx->w_win_samples.as<uint8_t>()};
    s->key = c.take<uint64_t>(n); s->order = c.take<uint64_t>(n);
    s->idx = c.take<uint32_t>(n);
    for (int m = 0; m < 3; ++m) s->val[m] = c.take<uint32_t>(n);
    s->node = c.take<uint32_t>(n);
    return KGM_OK;
  };
  // scal: the batch's largest short id and its window range (they size the index space); [3] = documents
  // the fused front end declined
  CU(ctx->w_win_scal.ensure(256));
  uint32_t *scal = ctx->w_win_scal.as<uint32_t>();
  auto reset_scal = [&]() -> int {
    h32[0] = 0; h32[1] = 0xffffffffu; h32[2] = 0; h32[3] = 0;
    CU(cudaMemcpyAsync(scal, h32, 16, cudaMemcpyHostToDevice, ctx->stream));
    return KGM_OK;
  };
  int rc;
  uint32_t n_desc = 0;                                         // sample places
  WinSamples s{};
  // Front end 1: the planning half of k_poll2 writes the samples itself (one launch over the resident
  // texts).  Any document it declines -> front end 2 takes the whole batch.
  bool fused_front = ctx->poll2_ok && ctx->p2_any && ctx->pipeline != KGM_PIPELINE_STAGED &&
                     ctx->pipeline != KGM_PIPELINE_FUSED_GENERIC && !getenv("KGM_WINDOW_STAGED_FRONT") &&
                     (uint64_t)ctx->batch.n_docs * ctx->poll2.cap_lines < (1ull << 31);
  ctx->win_front_fused = false;
  if (fused_front) {
    const uint32_t stride = ctx->poll2.cap_lines;
    n_desc = ctx->batch.n_docs * stride;
    if ((rc = carve_samples(n_desc, &s)) || (rc = reset_scal())) return rc;
    WinFront wo{};
    wo.s = s; wo.stride = stride; wo.window_ticks = window_ticks; wo.declined = scal + 3;
    const int wpb = (int)ctx->poll2.warps;
    const int blocks = std::max(1, std::min((int)((ctx->batch.n_docs + wpb - 1) / wpb), ctx->sms));
    launch_poll2_samples(ctx->batch, ctx->fleet, ctx->work, ctx->fused, ctx->poll2, wo, blocks, ctx->device, ctx->stream);
    CU(cudaGetLastError());
    CU(launch_window_scal(s, n_desc, scal, ctx->stream));
    CU(cudaMemcpyAsync(h32, scal, 16, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (h32[3] != 0) fused_front = false;
    ctx->win_front_fused = fused_front;
  }
  // Front end 2: the staged tokenizer + join leave the token tables and the line plan; k_win_samples
  // reads them.
  if (!fused_front) {
    if ((rc = run_plan_only(ctx, &n_desc))) return rc;
    if (n_desc == 0) return KGM_OK;
    if ((rc = carve_samples(n_desc, &s)) || (rc = reset_scal())) return rc;
    CU(launch_window_samples(ctx->batch, ctx->work, n_desc, window_ticks, s, scal, ctx->stream));
    CU(cudaMemcpyAsync(h32, scal, 16, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
  }
  if (h32[1] == 0xffffffffu) return KGM_OK;                    // no line was emitted
  WinIndex x{};
  x.n_nodes = ctx->fleet.n_nodes;
  x.sid_slots = h32[0] + 2u;                                   // ids 0 .. max, and "None"
  x.min_win = h32[1];
  x.n_win = h32[2] - h32[1] + 1u;
  const uint64_t size64 = ((uint64_t)ctx->fleet.n_nodes + ctx->fleet.n_containers) * x.sid_slots * x.n_win;
  if (size64 > (1ull << 28)) return KGM_E_LIMIT;               // (1 GiB of flags: a batch that spans too many windows)
  x.size = (uint32_t)size64;
  const uint32_t tiles = window_scan_tiles(x.size);
  const size_t flag_bytes = (((size_t)x.size + 1) * 4 + 255) & ~(size_t)255;
  CU(ctx->w_win_temp.ensure(flag_bytes + ((size_t)tiles + 1) * 4 + 256));
  uint32_t *flags = ctx->w_win_temp.as<uint32_t>();
  uint32_t *tops = reinterpret_cast<uint32_t *>(ctx->w_win_temp.as<uint8_t>() + flag_bytes);
  CU(cudaMemsetAsync(flags, 0, flag_bytes, ctx->stream));
  CU(launch_window_index(s, n_desc, x, flags, tops, ctx->stream));
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(h32, flags + x.size, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  const uint32_t n_groups = h32[0];
  const uint32_t *seg_incl = flags;                            // index -> group
  if (n_groups == 0) return KGM_OK;
  size_t group_bytes = 0;
  (void)carve_groups(nullptr, n_groups, &group_bytes);
  CU(ctx->w_win_groups.ensure(group_bytes));
  const WinGroups g = carve_groups(ctx->w_win_groups.as<uint8_t>(), n_groups, &group_bytes);
  CU(launch_window_reduce(s, seg_incl, n_desc, g, n_groups, ctx->stream));
  WinGroups r = g;
  if (want_host_copy) {
    CU(ctx->h_win_groups.ensure(group_bytes));
    CU(cudaMemcpyAsync(ctx->h_win_groups.p, ctx->w_win_groups.p, group_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    r = carve_groups(ctx->h_win_groups.as<uint8_t>(), n_groups, &group_bytes);
  }
  CU(cudaStreamSynchronize(ctx->stream));
  out->n = n_groups;
  out->kind = r.kind; out->node = r.node; out->container = r.container; out->gpu_id = r.gpu_id;
  out->window = r.window; out->count = r.count;
  for (int m = 0; m < 3; ++m) {
    out->m_count[m] = r.m_count[m]; out->m_min[m] = r.m_min[m]; out->m_max[m] = r.m_max[m];
    out->m_last[m] = r.m_last[m]; out->m_sum[m] = r.m_sum[m]; out->m_mean[m] = r.m_mean[m];
  }
  return KGM_OK;
}

int kgm_window_front(kgm_ctx *ctx) {
  if (!ctx) return KGM_E_INVALID;
  return ctx->win_front_fused ? 1 : 0;
}

int kgm_debug_stage_clocks(uint64_t *out64) {
  return out64 ? debug_stage_clocks(reinterpret_cast<unsigned long long *>(out64)) : KGM_E_INVALID;
}

int kgm_debug_chunk_times(kgm_ctx *ctx, float *text_in_ms, float *kernel_done_ms, uint32_t cap) {
  if (!ctx || !text_in_ms || !kernel_done_ms) return KGM_E_INVALID;
  if (ctx->last_chunks <= 1) return 1;
  for (uint32_t c = 0; c < ctx->last_chunks && c < cap; ++c) {
    cudaEventElapsedTime(text_in_ms + c, ctx->ev[EV_START], ctx->ev_in[c == 0 ? 0 : c < 3 ? 1 : 2]);
    cudaEventElapsedTime(kernel_done_ms + c, ctx->ev[EV_START], ctx->ev_k[c]);
  }
  cudaGetLastError();
  return (int)ctx->last_chunks;
}

int kgm_set_pipeline(kgm_ctx *ctx, int mode) {
  if (!ctx || mode < KGM_PIPELINE_AUTO || mode > KGM_PIPELINE_FUSED_V2) return KGM_E_INVALID;
  ctx->pipeline = mode;
  return KGM_OK;
}

int kgm_get_timing(kgm_ctx *ctx, kgm_timing *out) {
  if (!ctx || !out) return KGM_E_INVALID;
  *out = ctx->timing;
  return KGM_OK;
}

int kgm_last_kernel(kgm_ctx *ctx) {
  if (!ctx) return KGM_E_INVALID;
  return ctx->last_fused ? (ctx->last_poll2 ? 2 : 1) : 0;
}

int kgm_flush_l2(kgm_ctx *ctx, uint64_t bytes) {
  if (!ctx) return KGM_E_INVALID;
  CU(cudaSetDevice(ctx->device));
  bytes = (bytes + 15) & ~15ull;
  CU(ctx->w_flush.ensure(bytes));
  launch_flush(ctx->w_flush.as<uint8_t>(), bytes, ctx->sms, ctx->stream);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(ctx->stream));
  return KGM_OK;
}

void *kgm_host_alloc(uint64_t bytes) {
  void *p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}

void kgm_host_free(void *p) {
  if (p) cudaFreeHost(p);
}

}  // extern "C"
