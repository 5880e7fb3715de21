// Device-side arrays of the window reductions (kgm_window.cu), shared with kgm_api.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace kgm {

struct BatchDev;
struct WorkDev;

struct WinSamples {               // one entry per planned line
  uint64_t *key, *key_sorted;     // kind | entity | gpu_id | window; ~0 = line not emitted
  uint32_t *idx, *idx_sorted;     // sort permutation
  uint64_t *order;                // tick << 32 | line index: picks "last"
  uint32_t *val[3];               // gpu_util, mem, power; 0xffffffff = the line has none
  uint32_t *node;
};

struct WinGroups {                // one entry per (series, window), ascending key order
  uint8_t *kind;
  uint32_t *node, *container, *gpu_id, *window, *count;
  uint32_t *m_count[3], *m_min[3], *m_max[3], *m_last[3];
  uint64_t *m_sum[3];
  double *m_mean[3];
  uint64_t *key, *m_order[3];     // internal
};

size_t window_sort_temp_bytes(uint32_t n);
cudaError_t launch_window_samples(const BatchDev &b, const WorkDev &w, uint32_t n_desc, uint32_t window_ticks,
                                  const WinSamples &s, void *temp, size_t temp_bytes, uint32_t *head,
                                  uint32_t *seg_incl, cudaStream_t st);
cudaError_t launch_window_reduce(const WinSamples &s, const uint32_t *seg_incl, uint32_t n_desc, const WinGroups &g,
                                 uint32_t n_groups, cudaStream_t st);

}  // namespace kgm
