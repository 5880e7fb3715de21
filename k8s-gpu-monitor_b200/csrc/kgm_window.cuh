// Device-side arrays of the window reductions (kgm_window.cu), shared with kgm_api.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace kgm {

struct BatchDev;
struct WorkDev;

struct WinSamples {               // one entry per planned line
  uint64_t *key;                  // kind | entity | gpu_id | window; ~0 = line not emitted
  uint32_t *idx;                  // the key as a direct index (WinIndex); 0xffffffff = line not emitted
  uint64_t *order;                // tick | document | position: picks "last"
  uint32_t *val[3];               // gpu_util, mem, power; 0xffffffff = the line has none
  uint32_t *node;
};

// The index space of a batch: (series slot x short-id slot x window), ascending in the key.
// Series slot = node ordinal for node_gpu lines, n_nodes + container ordinal for pod_gpu lines;
// short-id slot = the id, the last slot for "None".
struct WinIndex {
  uint32_t n_nodes, sid_slots, min_win, n_win;
  uint32_t size;                  // (n_nodes + n_containers) * sid_slots * n_win
};

struct WinGroups {                // one entry per (series, window), ascending key order
  uint8_t *kind;
  uint32_t *node, *container, *gpu_id, *window, *count;
  uint32_t *m_count[3], *m_min[3], *m_max[3], *m_last[3];
  uint64_t *m_sum[3];
  double *m_mean[3];
  uint64_t *key, *m_order[3];     // internal
};

// The front end that k_poll2's planning half feeds directly (kgm_poll2.cuh, k_poll2_samples): sample
// (d, line) sits at d * stride + line; unused places carry the invalid key.
struct WinFront {
  WinSamples s;
  uint32_t stride;                // places per document (the kernel's lines-per-document capacity)
  uint32_t window_ticks;
  uint32_t *declined;             // documents the fast path declined (then the staged front end takes the batch)
};

constexpr uint64_t WIN_KEY_INVALID = ~0ull;
constexpr uint32_t WIN_SAT = 0xfffffffeu;            // values saturate here; 0xffffffff = "no value"
#if defined(__CUDACC__)
__device__ __forceinline__ uint64_t win_make_key(uint32_t kind, uint32_t entity, uint32_t sid, uint32_t window) {
  return ((uint64_t)kind << 63) | ((uint64_t)(entity & 0x7fffffffu) << 32) | ((uint64_t)(sid & 0xffffu) << 16) |
         (window & 0xffffu);
}
// "last" = greatest (tick, document, position in the document): 20 | 26 | 18 bits
__device__ __forceinline__ uint64_t win_order(uint32_t tick, uint32_t d, uint32_t pos) {
  return ((uint64_t)tick << 44) | ((uint64_t)(d & 0x3ffffffu) << 18) | (pos < 0x3ffffu ? pos : 0x3ffffu);
}
#endif

uint32_t window_scan_tiles(uint64_t n);
cudaError_t launch_window_scal(const WinSamples &s, uint32_t n, uint32_t *scal, cudaStream_t st);
cudaError_t launch_window_samples(const BatchDev &b, const WorkDev &w, uint32_t n_desc, uint32_t window_ticks,
                                  const WinSamples &s, uint32_t *scal, cudaStream_t st);
cudaError_t launch_window_index(const WinSamples &s, uint32_t n_desc, const WinIndex &x, uint32_t *flags,
                                uint32_t *tops, cudaStream_t st);
cudaError_t launch_window_reduce(const WinSamples &s, const uint32_t *offs, uint32_t n_desc, const WinGroups &g,
                                 uint32_t n_groups, cudaStream_t st);

}  // namespace kgm
