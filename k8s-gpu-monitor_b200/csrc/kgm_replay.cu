// Device-side replay generator: the two nvidia-smi texts of a fleet's polls, written straight into
// HBM (every value is f(seed, node, gpu-or-process, tick, field) of a
// counter-based mixer, "so numpy and a device generator agree byte-for-byte").
//
// This is bench / test plumbing, not part of the reference's path: there is no nvidia-smi on the GPU
// box, so fleet-scale input is replayed, and for the streamed configurations (BASELINE configs[3]:
// 12 500 hosts x 3600 ticks) generating the texts in host Python would dominate everything.  The
// CPU statement of the same texts is kgm/replay.py make_texts(); tests/test_replay_gpu.py compares
// the two byte for byte.  Row shapes:
//   <uuid>, <util> %, <umem> %, <total> MiB, <used> MiB, <free> MiB          (MIG parent: [N/A], [N/A])
//   <pid>, <process name>, <uuid>, <mem> MiB                                   (mig == 2: MIG-<uuid[4:]>)
// Documents are host-major (all ticks of host 0, then host 1, ...), like replay.make_replay.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/kgm.h"

namespace kgm {
namespace {

constexpr uint32_t FULL = 0xffffffffu;
constexpr int UUID_BYTES = 40;

__device__ __forceinline__ uint64_t mix64(uint64_t x) {                 // replay.py mix64 (splitmix64 finaliser)
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__device__ __forceinline__ uint64_t h_step(uint64_t x, uint64_t c) { return mix64(x ^ (c + 0x9E3779B97F4A7C15ull)); }
__device__ __forceinline__ uint64_t h_seed(uint64_t seed) { return mix64(seed + 0x9E3779B97F4A7C15ull); }

__device__ __forceinline__ uint32_t dec_digits(uint32_t v) {
  return v < 10u ? 1u : v < 100u ? 2u : v < 1000u ? 3u : v < 10000u ? 4u : v < 100000u ? 5u
       : v < 1000000u ? 6u : v < 10000000u ? 7u : v < 100000000u ? 8u : v < 1000000000u ? 9u : 10u;
}
__device__ __forceinline__ uint8_t *put_dec(uint8_t *p, uint32_t v) {
  const uint32_t n = dec_digits(v);
  for (uint32_t i = n; i-- > 0;) { p[i] = (uint8_t)('0' + v % 10u); v /= 10u; }
  return p + n;
}
__device__ __forceinline__ uint8_t *put_lit(uint8_t *p, const char *s) {
  while (*s) *p++ = (uint8_t)*s++;
  return p;
}
__device__ __forceinline__ uint8_t *put_bytes(uint8_t *p, const uint8_t *s, uint32_t n) {
  for (uint32_t i = 0; i < n; ++i) p[i] = s[i];
  return p + n;
}

__device__ const char GPU_HEADER[] =
    "uuid, utilization.gpu [%], utilization.memory [%], memory.total [MiB], memory.used [MiB], memory.free [MiB]\n";
__device__ const char APPS_HEADER[] = "pid, process_name, gpu_uuid, used_gpu_memory [MiB]\n";
constexpr uint32_t GPU_HEADER_BYTES = sizeof(GPU_HEADER) - 1, APPS_HEADER_BYTES = sizeof(APPS_HEADER) - 1;

struct GpuVals { uint32_t util, umem, used; };
__device__ __forceinline__ GpuVals gpu_vals(uint64_t base, uint32_t g, uint32_t total) {
  GpuVals v;
  v.util = (uint32_t)(mix64(base ^ (4ull * g + 1ull)) % 101ull);
  v.umem = (uint32_t)(mix64(base ^ (4ull * g + 2ull)) % 101ull);
  v.used = (uint32_t)(mix64(base ^ (4ull * g + 3ull)) % (uint64_t)(total + 1u));
  return v;
}
__device__ __forceinline__ uint32_t proc_mem(uint64_t hs, uint32_t node_id, uint32_t tick, uint32_t ci, uint32_t j) {
  uint64_t x = h_step(hs, 11);
  x = h_step(x, node_id); x = h_step(x, tick); x = h_step(x, ci); x = h_step(x, j);
  return 100u + (uint32_t)(x % 39901ull);
}

// bytes of one gpu row / one process row
__device__ __forceinline__ uint32_t gpu_row_bytes(const GpuVals &v, uint32_t total, bool mig) {
  const uint32_t head = mig ? 5u + 2u + 5u                                     // "[N/A]" ", " "[N/A]"
                            : dec_digits(v.util) + 2u + 2u + dec_digits(v.umem) + 2u;   // "<u> %" ", " "<m> %"
  return UUID_BYTES + 2u + head + 2u + dec_digits(total) + 4u + 2u + dec_digits(v.used) + 4u + 2u +
         dec_digits(total - v.used) + 4u + 1u;
}
__device__ __forceinline__ uint32_t proc_row_bytes(uint32_t pid, uint32_t name_len, uint32_t mem) {
  return dec_digits(pid) + 2u + name_len + 2u + UUID_BYTES + 2u + dec_digits(mem) + 4u + 1u;
}

// One warp per document: the sizes of its two texts.
__global__ void k_replay_sizes(kgm_replay_spec s, uint32_t tick0, uint32_t n_ticks, uint32_t *gpu_len, uint32_t *apps_len) {
  const uint32_t d = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, l = threadIdx.x & 31u;
  if (d >= s.n_nodes * n_ticks) return;
  const uint32_t n = d / n_ticks, tick = tick0 + d % n_ticks, node_id = s.node_id[n];
  const uint64_t hs = h_seed(s.seed);
  const uint64_t base = h_step(h_step(h_step(hs, 10), node_id), tick);
  const bool mig = s.mig[n] != 0;
  uint32_t gb = 0, ab = 0;
  const uint32_t g0 = s.gpu_begin[n], ng = s.gpu_begin[n + 1] - g0;
  for (uint32_t g = l; g < ng; g += 32) gb += gpu_row_bytes(gpu_vals(base, g, s.mem_total_mib), s.mem_total_mib, mig);
  const uint32_t p0 = s.proc_begin[n], np = s.proc_begin[n + 1] - p0;
  for (uint32_t r = l; r < np; r += 32) {
    const uint32_t q = p0 + r, nm = s.proc_name[q];
    ab += proc_row_bytes(s.proc_pid[q], s.name_off[nm + 1] - s.name_off[nm],
                         proc_mem(hs, node_id, tick, s.proc_ci[q], s.proc_j[q]));
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) { gb += __shfl_xor_sync(FULL, gb, o); ab += __shfl_xor_sync(FULL, ab, o); }
  if (l == 0) { gpu_len[d] = GPU_HEADER_BYTES + gb; apps_len[d] = APPS_HEADER_BYTES + ab; }
}

// One block: exclusive 64-bit offsets of n 32-bit lengths (off[n] = total).
__global__ void __launch_bounds__(1024) k_replay_offsets(const uint32_t *len, uint32_t n, unsigned long long *off) {
  __shared__ unsigned long long warp_sum[32];
  __shared__ unsigned long long carry_s;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (uint32_t i0 = 0; i0 < n; i0 += 1024) {
    const uint32_t i = i0 + threadIdx.x;
    const unsigned long long v = i < n ? len[i] : 0ull;
    unsigned long long incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long u = __shfl_up_sync(FULL, incl, o);
      if (lane >= (uint32_t)o) incl += u;
    }
    if (lane == 31u) warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const unsigned long long ws = warp_sum[lane];
      unsigned long long wi = ws;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long u = __shfl_up_sync(FULL, wi, o);
        if (lane >= (uint32_t)o) wi += u;
      }
      warp_sum[lane] = wi - ws;
    }
    __syncthreads();
    const unsigned long long carry = carry_s;
    if (i < n) off[i] = carry + warp_sum[warp] + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + warp_sum[31] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) off[n] = carry_s;
}

// One warp per document: lane 0 writes the headers, every lane its rows.  A row's position is the
// sum of the rows before it (recomputed from the same mixer: sizes are cheap, a second pass over
// memory is not needed).
__global__ void k_replay_fill(kgm_replay_spec s, uint32_t tick0, uint32_t n_ticks, const unsigned long long *gpu_off,
                              const unsigned long long *apps_off, uint8_t *gpu_csv, uint8_t *apps_csv) {
  const uint32_t d = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, l = threadIdx.x & 31u;
  if (d >= s.n_nodes * n_ticks) return;
  const uint32_t n = d / n_ticks, tick = tick0 + d % n_ticks, node_id = s.node_id[n];
  const uint64_t hs = h_seed(s.seed);
  const uint64_t base = h_step(h_step(h_step(hs, 10), node_id), tick);
  const uint32_t mig = s.mig[n], total = s.mem_total_mib;
  uint8_t *gp = gpu_csv + gpu_off[d], *ap = apps_csv + apps_off[d];
  if (l == 0) { put_lit(gp, GPU_HEADER); put_lit(ap, APPS_HEADER); }
  // gpu rows
  const uint32_t g0 = s.gpu_begin[n], ng = s.gpu_begin[n + 1] - g0;
  uint32_t done = GPU_HEADER_BYTES;
  for (uint32_t r0 = 0; r0 < ng; r0 += 32) {
    const uint32_t g = r0 + l;
    GpuVals v{0, 0, 0};
    uint32_t len = 0;
    if (g < ng) { v = gpu_vals(base, g, total); len = gpu_row_bytes(v, total, mig != 0); }
    uint32_t incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(FULL, incl, o);
      if (l >= (uint32_t)o) incl += u;
    }
    if (g < ng) {
      uint8_t *p = gp + done + incl - len;
      p = put_bytes(p, s.uuid + (size_t)UUID_BYTES * (g0 + g), UUID_BYTES);
      if (mig) {
        p = put_lit(p, ", [N/A], [N/A], ");
      } else {
        p = put_lit(p, ", "); p = put_dec(p, v.util); p = put_lit(p, " %, ");
        p = put_dec(p, v.umem); p = put_lit(p, " %, ");
      }
      p = put_dec(p, total); p = put_lit(p, " MiB, ");
      p = put_dec(p, v.used); p = put_lit(p, " MiB, ");
      p = put_dec(p, total - v.used); p = put_lit(p, " MiB\n");
    }
    done += __shfl_sync(FULL, incl, 31);
  }
  // process rows (GPU-major, as nvidia-smi lists them)
  const uint32_t p0 = s.proc_begin[n], np = s.proc_begin[n + 1] - p0;
  done = APPS_HEADER_BYTES;
  for (uint32_t r0 = 0; r0 < np; r0 += 32) {
    const uint32_t r = r0 + l, q = p0 + r;
    uint32_t len = 0, pid = 0, mem = 0, nm = 0, nlen = 0;
    if (r < np) {
      pid = s.proc_pid[q]; nm = s.proc_name[q]; nlen = s.name_off[nm + 1] - s.name_off[nm];
      mem = proc_mem(hs, node_id, tick, s.proc_ci[q], s.proc_j[q]);
      len = proc_row_bytes(pid, nlen, mem);
    }
    uint32_t incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(FULL, incl, o);
      if (l >= (uint32_t)o) incl += u;
    }
    if (r < np) {
      uint8_t *p = ap + done + incl - len;
      p = put_dec(p, pid); p = put_lit(p, ", ");
      p = put_bytes(p, s.names + s.name_off[nm], nlen); p = put_lit(p, ", ");
      const uint8_t *u = s.uuid + (size_t)UUID_BYTES * (g0 + s.proc_gpu[q]);
      if (mig == 2) { p = put_lit(p, "MIG-"); p = put_bytes(p, u + 4, UUID_BYTES - 4); }   // the process rows name MIG instances
      else p = put_bytes(p, u, UUID_BYTES);
      p = put_lit(p, ", "); p = put_dec(p, mem); p = put_lit(p, " MiB\n");
    }
    done += __shfl_sync(FULL, incl, 31);
  }
}

}  // namespace
}  // namespace kgm

extern "C" int kgm_replay_offsets(const kgm_replay_spec *spec, uint32_t tick0, uint32_t n_ticks, uint32_t *len_scratch,
                                  uint64_t *gpu_off, uint64_t *apps_off, void *stream) {
  if (!spec || !len_scratch || !gpu_off || !apps_off || !n_ticks || !spec->n_nodes) return KGM_E_INVALID;
  const uint64_t n_docs = (uint64_t)spec->n_nodes * n_ticks;
  if (n_docs >= (1ull << 26)) return KGM_E_LIMIT;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const uint32_t n = (uint32_t)n_docs;
  kgm::k_replay_sizes<<<(n + 7) / 8, 256, 0, st>>>(*spec, tick0, n_ticks, len_scratch, len_scratch + n);
  kgm::k_replay_offsets<<<1, 1024, 0, st>>>(len_scratch, n, reinterpret_cast<unsigned long long *>(gpu_off));
  kgm::k_replay_offsets<<<1, 1024, 0, st>>>(len_scratch + n, n, reinterpret_cast<unsigned long long *>(apps_off));
  return cudaGetLastError() == cudaSuccess ? KGM_OK : KGM_E_CUDA;
}

extern "C" int kgm_replay_fill(const kgm_replay_spec *spec, uint32_t tick0, uint32_t n_ticks, const uint64_t *gpu_off,
                               const uint64_t *apps_off, uint8_t *gpu_csv, uint8_t *apps_csv, void *stream) {
  if (!spec || !gpu_off || !apps_off || !gpu_csv || !apps_csv || !n_ticks || !spec->n_nodes) return KGM_E_INVALID;
  const uint64_t n_docs = (uint64_t)spec->n_nodes * n_ticks;
  if (n_docs >= (1ull << 26)) return KGM_E_LIMIT;
  const uint32_t n = (uint32_t)n_docs;
  kgm::k_replay_fill<<<(n + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      *spec, tick0, n_ticks, reinterpret_cast<const unsigned long long *>(gpu_off),
      reinterpret_cast<const unsigned long long *>(apps_off), gpu_csv, apps_csv);
  return cudaGetLastError() == cudaSuccess ? KGM_OK : KGM_E_CUDA;
}
