// Stage 3: per-series time-window reductions over the lines of a batch.
//
// No counterpart in the reference (gpu.py emits raw per-poll lines and leaves aggregation to
// InfluxQL / Grafana): this is the north-star's additive stage.  It never changes a byte of
// node_gpu / pod_gpu output.  A *sample* is one emitted line; its series is
//     node_gpu line -> (host, gpu_id)                    the line's tag set, gpu.py:137
//     pod_gpu line  -> (host, container, gpu_id)         "per (node, pod, gpu)"
// its window is doc_tick / window_ticks (ticks < 2^20, windows < 2^16, documents < 2^26), and its values are the integers behind the line's
// fields: gpu_util, mem (mem_used of a node line, used_memory of a pod line) and, when the
// --query-gpu text carries a 7th column (power.draw), that cell's digits.  Per (series,
// window): count, and per value min / max / sum (-> mean) / last (by tick, then line order).
//
// Data flow: the staged tokenizer + join leave the token tables and the line plan in the
// workspace; k_win_samples turns every planned line into (key, order, values) and finds the
// batch's largest short id and its window range.  A key is then a DIRECT index: series slot
// (node ordinal, or nodes + container ordinal) x short-id slot x window, so no sort is needed:
// k_win_mark flags the indices that occur, an exclusive scan over the flags (k_scan_*, below)
// numbers the groups in ascending key order, k_win_accum adds every sample to its group with
// atomics, k_win_last / k_win_final fill in the "last" values, the means and the unpacked keys.
#include <cuda_runtime.h>

#include "kgm_device.cuh"
#include "kgm_window.cuh"

namespace kgm {
namespace {

constexpr uint32_t FULLMASK = 0xffffffffu;
constexpr uint64_t KEY_INVALID = WIN_KEY_INVALID;
constexpr int WIN_THREADS = 256;

__device__ __forceinline__ uint64_t make_key(uint32_t kind, uint32_t entity, uint32_t sid, uint32_t window) {
  return win_make_key(kind, entity, sid, window);
}

// One planned line -> one sample.  Lines past a document's raise point are planned but not
// emitted (doc_nlines): they get the invalid key and sort to the end.
// scal[0] = largest short id of an emitted line (0xffff = None not counted), scal[1] = smallest window, scal[2] =
// largest window (the host sizes the index space from them).
__global__ void k_win_samples(BatchDev b, WorkDev w, uint32_t n_desc, uint32_t window_ticks, WinSamples s, uint32_t *scal) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_desc) return;
  const LineDesc ds = w.desc[i];
  const uint32_t d = ds.doc;
  uint64_t key = KEY_INVALID;
  uint32_t util = NONE32, mem = NONE32, power = NONE32, node = 0;
  if (d < b.n_docs && i >= w.doc_desc0[d] && i - w.doc_desc0[d] < w.doc_nlines[d]) {
    node = b.doc_node[d];
    const uint32_t window = b.doc_tick[d] / window_ticks;
    const uint32_t kind = ds.kind_ref >> 30, ref = ds.kind_ref & 0x3fffffffu;
    if (kind == LINE_NODE) {
      const GpuRow &g = w.grow[ds.a];
      if (g.nd[0]) util = min(g.util, WIN_SAT);
      mem = g.mem_used > WIN_SAT ? WIN_SAT : (uint32_t)g.mem_used;
      if (g.power != NONE32) power = min(g.power, WIN_SAT);
      key = make_key(0, node, ds.b, window);
    } else if (kind == LINE_BUSY) {
      const ProcRow &p = w.prow[ds.a];
      const GpuRow &g = w.grow[ds.b];
      if (g.nd[0]) util = min(g.util, WIN_SAT);
      if (p.nd_mem) mem = min(p.used_mem, WIN_SAT);
      if (g.power != NONE32) power = min(g.power, WIN_SAT);
      key = make_key(1, ref, p.sid, window);
    } else {
      util = 0; mem = 0;                             // gpu.py:234-250: gpu_util=0, used_memory=0
      key = make_key(1, ref, ds.a == NONE32 ? 0xffffu : ds.a, window);
    }
  }
  s.key[i] = key;
  {
    const bool ok = key != KEY_INVALID;
    const uint32_t sid = (uint32_t)(key >> 16) & 0xffffu, win = (uint32_t)key & 0xffffu;
    const uint32_t am = __activemask();
    const uint32_t mx_sid = __reduce_max_sync(am, ok && sid != 0xffffu ? sid : 0u);
    const uint32_t mn_win = __reduce_min_sync(am, ok ? win : 0xffffffffu);
    const uint32_t mx_win = __reduce_max_sync(am, ok ? win : 0u);
    if ((threadIdx.x & 31u) == (uint32_t)__ffs(am) - 1u) {
      atomicMax(scal + 0, mx_sid);
      atomicMin(scal + 1, mn_win);
      atomicMax(scal + 2, mx_win);
    }
  }
  {
    const uint32_t tick = d < b.n_docs ? b.doc_tick[d] : 0u;
    const uint32_t pos = d < b.n_docs ? i - w.doc_desc0[d] : 0u;
    s.order[i] = win_order(tick, d, pos);
  }
  s.val[0][i] = util; s.val[1][i] = mem; s.val[2][i] = power;
  s.node[i] = node;
}

__global__ void k_win_init(WinGroups g, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  g.count[i] = 0;
  g.key[i] = 0;
  g.node[i] = 0;
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    g.m_count[m][i] = 0; g.m_min[m][i] = 0xffffffffu; g.m_max[m][i] = 0; g.m_sum[m][i] = 0;
    g.m_order[m][i] = 0; g.m_last[m][i] = 0;
  }
}

// scal[] (see k_win_samples) of samples that another kernel wrote (k_poll2_samples)
__global__ void k_win_scal(const uint64_t *key, uint32_t n, uint32_t *scal) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = key[i];
  const bool ok = k != KEY_INVALID;
  const uint32_t sid = (uint32_t)(k >> 16) & 0xffffu, win = (uint32_t)k & 0xffffu;
  const uint32_t am = __activemask();
  const uint32_t mx_sid = __reduce_max_sync(am, ok && sid != 0xffffu ? sid : 0u);
  const uint32_t mn_win = __reduce_min_sync(am, ok ? win : 0xffffffffu);
  const uint32_t mx_win = __reduce_max_sync(am, ok ? win : 0u);
  if ((threadIdx.x & 31u) == (uint32_t)__ffs(am) - 1u) {
    atomicMax(scal + 0, mx_sid);
    atomicMin(scal + 1, mn_win);
    atomicMax(scal + 2, mx_win);
  }
}

// key -> direct index (ascending in the key): see WinIndex in kgm_window.cuh
__device__ __forceinline__ uint32_t index_of(uint64_t key, const WinIndex &x) {
  const uint32_t kind = (uint32_t)(key >> 63), entity = (uint32_t)(key >> 32) & 0x7fffffffu;
  const uint32_t sid = (uint32_t)(key >> 16) & 0xffffu, win = (uint32_t)key & 0xffffu;
  const uint32_t slot = kind ? x.n_nodes + entity : entity;
  const uint32_t sslot = sid == 0xffffu ? x.sid_slots - 1u : sid;
  return (slot * x.sid_slots + sslot) * x.n_win + (win - x.min_win);
}

// flags[index] = 1 for every index that occurs; idx[i] = the sample's index (NONE32: not emitted)
__global__ void k_win_mark(WinSamples s, uint32_t n, WinIndex x, uint32_t *flags) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t key = s.key[i];
  uint32_t ix = NONE32;
  if (key != KEY_INVALID) {
    ix = index_of(key, x);
    flags[ix] = 1u;
  }
  s.idx[i] = ix;
}

// ---- exclusive scan of a uint32 array, in place: tiles of SCAN_TILE, their totals, the add-back ----
constexpr int SCAN_THREADS = 1024, SCAN_PER = 4, SCAN_TILE = SCAN_THREADS * SCAN_PER;

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *total) {   // every thread of the block
  __shared__ uint32_t warp_sum[32];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(FULLMASK, incl, o);
    if (lane >= (uint32_t)o) incl += u;
  }
  if (lane == 31u) warp_sum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t ws = warp_sum[lane], wi = ws;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(FULLMASK, wi, o);
      if (lane >= (uint32_t)o) wi += u;
    }
    warp_sum[lane] = wi - ws;                        // exclusive over the warps
    if (lane == 31u) *total = wi;
  }
  __syncthreads();
  const uint32_t r = warp_sum[warp] + incl - v;
  __syncthreads();                                   // warp_sum may be reused by the caller's next round
  return r;
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tiles(uint32_t *a, uint32_t n, uint32_t *tops) {
  __shared__ uint32_t total;
  const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_PER;
  uint32_t v[SCAN_PER], sum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_PER; ++k) { v[k] = base + k < n ? a[base + k] : 0u; sum += v[k]; }
  uint32_t ex = block_excl_scan(sum, &total);
#pragma unroll
  for (int k = 0; k < SCAN_PER; ++k) { if (base + k < n) a[base + k] = ex; ex += v[k]; }
  if (threadIdx.x == 0) tops[blockIdx.x] = total;
}
// one block: tops[0 .. nt) -> exclusive, tops[nt] = the grand total
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tops(uint32_t *tops, uint32_t nt) {
  __shared__ uint32_t total;
  uint32_t carry = 0;
  for (uint32_t i0 = 0; i0 < nt; i0 += SCAN_THREADS) {
    const uint32_t i = i0 + threadIdx.x;
    const uint32_t v = i < nt ? tops[i] : 0u;
    const uint32_t ex = block_excl_scan(v, &total);
    if (i < nt) tops[i] = carry + ex;
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) tops[nt] = carry;
}
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_add(uint32_t *a, uint32_t n, const uint32_t *tops, uint32_t nt) {
  const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_PER, add = tops[blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_PER; ++k)
    if (base + k < n) a[base + k] += add;
  if (blockIdx.x == 0 && threadIdx.x == 0) a[n] = tops[nt];          // a[] has n + 1 entries
}

// Every emitted sample is added to its group: group = offs[index].
__global__ void __launch_bounds__(WIN_THREADS) k_win_accum(WinSamples s, const uint32_t *offs, uint32_t n, WinGroups g) {
  const uint32_t i = blockIdx.x * WIN_THREADS + threadIdx.x;
  if (i >= n) return;
  const uint32_t ix = s.idx[i];
  if (ix == NONE32) return;
  const uint32_t my = offs[ix];
  const uint64_t order = s.order[i];
  atomicAdd(&g.count[my], 1u);
  g.key[my] = s.key[i];                              // (the same value from every sample of the group)
  g.node[my] = s.node[i];
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    const uint32_t v = s.val[m][i];
    if (v == NONE32) continue;
    atomicAdd(&g.m_count[m][my], 1u);
    atomicMin(&g.m_min[m][my], v);
    atomicMax(&g.m_max[m][my], v);
    atomicAdd(reinterpret_cast<unsigned long long *>(&g.m_sum[m][my]), (unsigned long long)v);
    atomicMax(reinterpret_cast<unsigned long long *>(&g.m_order[m][my]), (unsigned long long)order);
  }
}

// The value of the sample that holds the group's greatest (tick, line) order.
__global__ void k_win_last(WinSamples s, const uint32_t *offs, uint32_t n, WinGroups g) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t ix = s.idx[i];
  if (ix == NONE32) return;
  const uint32_t seg = offs[ix];
  const uint64_t order = s.order[i];
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    const uint32_t v = s.val[m][i];
    if (v != NONE32 && g.m_order[m][seg] == order) g.m_last[m][seg] = v;
  }
}

__global__ void k_win_final(WinGroups g, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = g.key[i];
  g.kind[i] = (uint8_t)(k >> 63);
  const uint32_t entity = (uint32_t)(k >> 32) & 0x7fffffffu;
  g.container[i] = (k >> 63) ? entity : NONE32;
  g.gpu_id[i] = (uint32_t)(k >> 16) & 0xffffu;
  g.window[i] = (uint32_t)k & 0xffffu;
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    const uint32_t c = g.m_count[m][i];
    g.m_mean[m][i] = c ? (double)g.m_sum[m][i] / (double)c : 0.0;     // exact integer sum, one IEEE division
    if (!c) { g.m_min[m][i] = 0; g.m_max[m][i] = 0; g.m_last[m][i] = 0; }
  }
}

}  // namespace

cudaError_t launch_window_scal(const WinSamples &s, uint32_t n, uint32_t *scal, cudaStream_t st) {
  if (!n) return cudaSuccess;
  k_win_scal<<<(n + WIN_THREADS - 1) / WIN_THREADS, WIN_THREADS, 0, st>>>(s.key, n, scal);
  return cudaGetLastError();
}

uint32_t window_scan_tiles(uint64_t n) { return (uint32_t)((n + SCAN_TILE - 1) / SCAN_TILE); }

cudaError_t launch_window_samples(const BatchDev &b, const WorkDev &w, uint32_t n_desc, uint32_t window_ticks,
                                  const WinSamples &s, uint32_t *scal, cudaStream_t st) {
  if (!n_desc) return cudaSuccess;
  const int blocks = (int)((n_desc + WIN_THREADS - 1) / WIN_THREADS);
  k_win_samples<<<blocks, WIN_THREADS, 0, st>>>(b, w, n_desc, window_ticks, s, scal);
  return cudaGetLastError();
}

// flags[0 .. x.size] (zeroed by the caller) -> offs: offs[index] = the group of an index that occurs,
// offs[x.size] = the number of groups.  tops: window_scan_tiles(x.size) + 1 words.
cudaError_t launch_window_index(const WinSamples &s, uint32_t n_desc, const WinIndex &x, uint32_t *flags,
                                uint32_t *tops, cudaStream_t st) {
  if (!n_desc || !x.size) return cudaSuccess;
  const int blocks = (int)((n_desc + WIN_THREADS - 1) / WIN_THREADS);
  k_win_mark<<<blocks, WIN_THREADS, 0, st>>>(s, n_desc, x, flags);
  const uint32_t nt = window_scan_tiles(x.size);
  k_scan_tiles<<<nt, SCAN_THREADS, 0, st>>>(flags, x.size, tops);
  k_scan_tops<<<1, SCAN_THREADS, 0, st>>>(tops, nt);
  k_scan_add<<<nt, SCAN_THREADS, 0, st>>>(flags, x.size, tops, nt);
  return cudaGetLastError();
}

cudaError_t launch_window_reduce(const WinSamples &s, const uint32_t *offs, uint32_t n_desc, const WinGroups &g,
                                 uint32_t n_groups, cudaStream_t st) {
  if (!n_desc || !n_groups) return cudaSuccess;
  const int gb = (int)((n_groups + WIN_THREADS - 1) / WIN_THREADS);
  const int sb = (int)((n_desc + WIN_THREADS - 1) / WIN_THREADS);
  k_win_init<<<gb, WIN_THREADS, 0, st>>>(g, n_groups);
  k_win_accum<<<sb, WIN_THREADS, 0, st>>>(s, offs, n_desc, g);
  k_win_last<<<sb, WIN_THREADS, 0, st>>>(s, offs, n_desc, g);
  k_win_final<<<gb, WIN_THREADS, 0, st>>>(g, n_groups);
  return cudaGetLastError();
}

}  // namespace kgm
