// kgm_inspect_labels: the `docker inspect` texts of a fleet's containers scanned on the device, one
// thread per container (kgm_docker.cuh has the scanner and what it accepts).  Docker-side work is per
// container, not per sample: this is for fleets whose container count
// makes json.loads in host Python the slow part of kgm_load_fleet's preparation.
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "../../include/kgm.h"
#include "kgm_docker.cuh"

namespace kgm {
namespace {

__global__ void k_inspect(const uint8_t *texts, const unsigned long long *off, uint32_t n, dk::Fields *out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long lo = off[i], hi = off[i + 1];
  dk::Fields f;
  if (hi < lo || hi - lo > 0xfff00000ull) {              // (a text of 4 GiB: the host's)
    f.status = 1; f.gated = f.has_pod = 0;
    f.pod_off = f.pod_len = f.ns_off = f.ns_len = f.user_off = f.user_len = 0;
  } else {
    f = dk::inspect_fields(texts + lo, (uint32_t)(hi - lo));
  }
  out[i] = f;
}

struct DevMem {
  void *p = nullptr;
  ~DevMem() { if (p) cudaFree(p); }
};

}  // namespace
}  // namespace kgm

extern "C" int kgm_inspect_labels(int device, const uint8_t *texts, const uint64_t *off, uint32_t n,
                                  const kgm_inspect_out *out) {
  if (!off || !out || !out->status || !out->gated || !out->has_pod || !out->pod_off || !out->pod_len ||
      !out->ns_off || !out->ns_len || !out->user_off || !out->user_len)
    return KGM_E_INVALID;
  if (n == 0) return KGM_OK;
  const uint64_t bytes = off[n];
  if (off[0] != 0 || (bytes && !texts)) return KGM_E_INVALID;
  for (uint32_t i = 0; i < n; ++i)
    if (off[i + 1] < off[i]) return KGM_E_INVALID;
  if (cudaSetDevice(device) != cudaSuccess) return KGM_E_CUDA;
  kgm::DevMem d_text, d_off, d_out;
  if (cudaMalloc(&d_text.p, bytes + 16) != cudaSuccess || cudaMalloc(&d_off.p, ((size_t)n + 1) * 8) != cudaSuccess ||
      cudaMalloc(&d_out.p, (size_t)n * sizeof(kgm::dk::Fields)) != cudaSuccess) {
    cudaGetLastError();
    return KGM_E_NOMEM;
  }
  if ((bytes && cudaMemcpy(d_text.p, texts, bytes, cudaMemcpyHostToDevice) != cudaSuccess) ||
      cudaMemcpy(d_off.p, off, ((size_t)n + 1) * 8, cudaMemcpyHostToDevice) != cudaSuccess)
    return KGM_E_CUDA;
  kgm::k_inspect<<<(n + 127) / 128, 128>>>(static_cast<const uint8_t *>(d_text.p),
                                           static_cast<const unsigned long long *>(d_off.p), n,
                                           static_cast<kgm::dk::Fields *>(d_out.p));
  if (cudaGetLastError() != cudaSuccess) return KGM_E_CUDA;
  std::vector<kgm::dk::Fields> h(n);
  if (cudaMemcpy(h.data(), d_out.p, (size_t)n * sizeof(kgm::dk::Fields), cudaMemcpyDeviceToHost) != cudaSuccess)
    return KGM_E_CUDA;
  for (uint32_t i = 0; i < n; ++i) {
    const kgm::dk::Fields &f = h[i];
    out->status[i] = (uint8_t)f.status; out->gated[i] = (uint8_t)f.gated; out->has_pod[i] = (uint8_t)f.has_pod;
    out->pod_off[i] = f.pod_off; out->pod_len[i] = f.pod_len; out->ns_off[i] = f.ns_off; out->ns_len[i] = f.ns_len;
    out->user_off[i] = f.user_off; out->user_len[i] = f.user_len;
  }
  return KGM_OK;
}
