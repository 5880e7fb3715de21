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
// workspace; k_win_samples turns every planned line into (key, order, values); the keys are
// radix-sorted (CUB, plumbing); k_win_reduce stages a tile of sorted samples in shared
// memory and reduces runs of equal keys with warp shuffles, one atomic per (warp, run);
// k_win_last / k_win_final fill in the "last" values, the means and the unpacked keys.
#include <cuda_runtime.h>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "kgm_device.cuh"
#include "kgm_window.cuh"

namespace kgm {
namespace {

constexpr uint32_t FULLMASK = 0xffffffffu;
constexpr uint64_t KEY_INVALID = ~0ull;
constexpr int WIN_THREADS = 256;
constexpr uint32_t WIN_SAT = 0xfffffffeu;            // values saturate here; 0xffffffff = "no value"

__device__ __forceinline__ uint64_t make_key(uint32_t kind, uint32_t entity, uint32_t sid, uint32_t window) {
  return ((uint64_t)kind << 63) | ((uint64_t)(entity & 0x7fffffffu) << 32) | ((uint64_t)(sid & 0xffffu) << 16) |
         (window & 0xffffu);
}

// One planned line -> one sample.  Lines past a document's raise point are planned but not
// emitted (doc_nlines): they get the invalid key and sort to the end.
__global__ void k_win_samples(BatchDev b, WorkDev w, uint32_t n_desc, uint32_t window_ticks, WinSamples s) {
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
  s.idx[i] = i;
  {
    // "last" = greatest (tick, document, position in the document): 20 | 26 | 18 bits
    const uint32_t tick = d < b.n_docs ? b.doc_tick[d] : 0u;
    const uint32_t pos = d < b.n_docs ? min(i - w.doc_desc0[d], 0x3ffffu) : 0u;
    s.order[i] = ((uint64_t)tick << 44) | ((uint64_t)(d & 0x3ffffffu) << 18) | pos;
  }
  s.val[0][i] = util; s.val[1][i] = mem; s.val[2][i] = power;
  s.node[i] = node;
}

__global__ void k_win_heads(const uint64_t *key, uint32_t n, uint32_t *head) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = key[i];
  head[i] = (k != KEY_INVALID && (i == 0 || key[i - 1] != k)) ? 1u : 0u;
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

struct Agg { uint32_t cnt, mn, mx; uint64_t sum, order; };

__device__ __forceinline__ Agg agg_of(uint32_t v, uint64_t order) {
  Agg a;
  const bool has = v != NONE32;
  a.cnt = has ? 1u : 0u; a.mn = has ? v : 0xffffffffu; a.mx = has ? v : 0u; a.sum = has ? v : 0ull;
  a.order = has ? order : 0ull;
  return a;
}
__device__ __forceinline__ Agg agg_merge(const Agg &a, const Agg &b) {
  Agg r;
  r.cnt = a.cnt + b.cnt; r.mn = min(a.mn, b.mn); r.mx = max(a.mx, b.mx); r.sum = a.sum + b.sum;
  r.order = a.order > b.order ? a.order : b.order;
  return r;
}
__device__ __forceinline__ Agg agg_shfl_up(const Agg &a, int delta) {
  Agg r;
  r.cnt = __shfl_up_sync(FULLMASK, a.cnt, delta); r.mn = __shfl_up_sync(FULLMASK, a.mn, delta);
  r.mx = __shfl_up_sync(FULLMASK, a.mx, delta); r.sum = __shfl_up_sync(FULLMASK, a.sum, delta);
  r.order = __shfl_up_sync(FULLMASK, a.order, delta);
  return r;
}

// Sorted samples -> groups.  A block stages its tile (gathered through the sort
// permutation) in shared memory; each warp then runs a segmented inclusive scan over its 32
// samples with shuffles, and the last lane of every run of equal groups adds the run's
// aggregate to the group with atomics (a run continues across warps and blocks).
__global__ void __launch_bounds__(WIN_THREADS) k_win_reduce(WinSamples s, const uint32_t *seg_incl, uint32_t n,
                                                            WinGroups g) {
  __shared__ uint32_t sh_seg[WIN_THREADS];
  __shared__ uint32_t sh_val[3][WIN_THREADS];
  __shared__ uint64_t sh_order[WIN_THREADS];
  __shared__ uint64_t sh_key[WIN_THREADS];
  __shared__ uint32_t sh_node[WIN_THREADS];
  const uint32_t t = threadIdx.x, i = blockIdx.x * WIN_THREADS + t;
  uint32_t seg = NONE32;
  if (i < n && s.key_sorted[i] != KEY_INVALID) {
    const uint32_t src = s.idx_sorted[i];
    seg = seg_incl[i] - 1u;
    sh_val[0][t] = s.val[0][src]; sh_val[1][t] = s.val[1][src]; sh_val[2][t] = s.val[2][src];
    sh_order[t] = s.order[src];
    sh_key[t] = s.key_sorted[i];
    sh_node[t] = s.node[src];
  }
  sh_seg[t] = seg;
  __syncthreads();
  const uint32_t lane = t & 31u;
  const uint32_t my = sh_seg[t];
  const uint32_t prev = lane ? sh_seg[t - 1] : NONE32 - 1u;          // lane 0 starts a run inside the warp
  const bool head = my != prev;
  const uint32_t next = (lane == 31u) ? NONE32 - 2u : sh_seg[t + 1];
  const bool tail = my != next;
  uint32_t cnt = my != NONE32 ? 1u : 0u;
  Agg a[3];
#pragma unroll
  for (int m = 0; m < 3; ++m) a[m] = my != NONE32 ? agg_of(sh_val[m][t], sh_order[t]) : agg_of(NONE32, 0);
  // distance to the run's head inside the warp
  const uint32_t heads = __ballot_sync(FULLMASK, head);
  const uint32_t below = heads & (0xffffffffu >> (31u - lane));       // heads at or below this lane
  const uint32_t run_start = 31u - __clz(below);
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t c2 = __shfl_up_sync(FULLMASK, cnt, o);
    Agg up[3];
#pragma unroll
    for (int m = 0; m < 3; ++m) up[m] = agg_shfl_up(a[m], o);
    if (lane >= run_start + (uint32_t)o) {
      cnt += c2;
#pragma unroll
      for (int m = 0; m < 3; ++m) a[m] = agg_merge(a[m], up[m]);
    }
  }
  if (tail && my != NONE32) {
    atomicAdd(&g.count[my], cnt);
    g.key[my] = sh_key[t];
    g.node[my] = sh_node[t];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
      if (!a[m].cnt) continue;
      atomicAdd(&g.m_count[m][my], a[m].cnt);
      atomicMin(&g.m_min[m][my], a[m].mn);
      atomicMax(&g.m_max[m][my], a[m].mx);
      atomicAdd(reinterpret_cast<unsigned long long *>(&g.m_sum[m][my]), (unsigned long long)a[m].sum);
      atomicMax(reinterpret_cast<unsigned long long *>(&g.m_order[m][my]), (unsigned long long)a[m].order);
    }
  }
}

// The value of the sample that holds the group's greatest (tick, line) order.
__global__ void k_win_last(WinSamples s, const uint32_t *seg_incl, uint32_t n, WinGroups g) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || s.key_sorted[i] == KEY_INVALID) return;
  const uint32_t src = s.idx_sorted[i], seg = seg_incl[i] - 1u;
  const uint64_t order = s.order[src];
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    const uint32_t v = s.val[m][src];
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

size_t window_sort_temp_bytes(uint32_t n) {
  size_t a = 0, b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, a, (const uint64_t *)nullptr, (uint64_t *)nullptr,
                                  (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n);
  cub::DeviceScan::InclusiveSum(nullptr, b, (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n);
  return (a > b ? a : b) + 256;
}

cudaError_t launch_window_samples(const BatchDev &b, const WorkDev &w, uint32_t n_desc, uint32_t window_ticks,
                                  const WinSamples &s, void *temp, size_t temp_bytes, uint32_t *head,
                                  uint32_t *seg_incl, cudaStream_t st) {
  if (!n_desc) return cudaSuccess;
  const int blocks = (int)((n_desc + WIN_THREADS - 1) / WIN_THREADS);
  k_win_samples<<<blocks, WIN_THREADS, 0, st>>>(b, w, n_desc, window_ticks, s);
  cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, s.key, s.key_sorted, s.idx, s.idx_sorted,
                                                  (int)n_desc, 0, 64, st);
  if (e != cudaSuccess) return e;
  k_win_heads<<<blocks, WIN_THREADS, 0, st>>>(s.key_sorted, n_desc, head);
  return cub::DeviceScan::InclusiveSum(temp, temp_bytes, head, seg_incl, (int)n_desc, st);
}

cudaError_t launch_window_reduce(const WinSamples &s, const uint32_t *seg_incl, uint32_t n_desc, const WinGroups &g,
                                 uint32_t n_groups, cudaStream_t st) {
  if (!n_desc || !n_groups) return cudaSuccess;
  const int gb = (int)((n_groups + WIN_THREADS - 1) / WIN_THREADS);
  const int sb = (int)((n_desc + WIN_THREADS - 1) / WIN_THREADS);
  k_win_init<<<gb, WIN_THREADS, 0, st>>>(g, n_groups);
  k_win_reduce<<<sb, WIN_THREADS, 0, st>>>(s, seg_incl, n_desc, g);
  k_win_last<<<sb, WIN_THREADS, 0, st>>>(s, seg_incl, n_desc, g);
  k_win_final<<<gb, WIN_THREADS, 0, st>>>(g, n_groups);
  return cudaGetLastError();
}

}  // namespace kgm
