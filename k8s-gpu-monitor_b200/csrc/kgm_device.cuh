// Device-side parameter blocks shared by the kernels (kgm_kernels.cu) and the C ABI
// implementation (kgm_api.cu).
#pragma once
#include "kgm_common.cuh"
#include "kgm_fast.cuh"

namespace kgm {

struct BatchDev {
  uint32_t n_docs;
  const uint8_t *gpu_csv;        // 16-byte aligned, padded
  const uint64_t *gpu_off;       // n_docs + 1
  const uint8_t *apps_csv;
  const uint64_t *apps_off;
  const uint32_t *doc_node;
  const uint32_t *doc_tick;      // may be null
};

// Fleet tables resident in HBM (built by kgm_load_fleet).
struct FleetDev {
  uint32_t n_nodes, n_containers, max_nc;
  const uint32_t *cont_begin;    // n_nodes + 1
  const uint8_t *node_err_parse, *node_err_join;
  const uint8_t *gated, *err_inspect, *err_emit;
  const uint32_t *vis_begin;     // n_containers + 1
  const uint32_t *vis_off;       // n_vis + 1 into vis_bytes
  const uint8_t *vis_bytes;
  const uint64_t *vis_hash;      // same hash as GpuRow::key_hash
  const Piece *node_piece;       // NODE_PIECES per node
  const uint32_t *node_fixed;    // sum of the node's piece lengths
  const Piece *cont_piece;       // POD_PIECES per container
  const uint32_t *cont_fixed;
  const uint8_t *skel;           // skeleton pool, pieces 16-byte aligned
  const PidSlot *pid_table;
  uint32_t pid_mask;             // table size - 1 (power of two)
  // tables of the fused kernel's fast path (kgm_fast.cuh)
  const uint8_t *node_fast;      // 1 = no raise point and nothing unusual on the docker side of this node
  const uint32_t *tb_off;        // per node: its template block in tb_pool (16-byte aligned)
  const uint32_t *tb_len;
  const uint8_t *tb_pool;
  const fast::PidSlot16 *pid16;  // pids of up to 7 bytes
  uint32_t pid16_mask;
  const uint64_t *vis_hash2;     // fast::key_hash_bytes of every visible uuid
  // tables of k_poll2 (kgm_poll2_leaf.cuh): one template slot per node and per gated container
  const uint8_t *node_fast2;     // 1 = the node's documents may take k_poll2's fast path
  const uint32_t *nslot_off;     // per node: its slot in slot_pool, in 16-byte units
  const uint32_t *tslot_off;     // per container: its slot (0 for containers that are not gated)
  const uint8_t *slot_pool;      // 16-byte aligned slots, padded so that slot_bytes can be read from any of them
  const uint4 *ca2;              // image of the k_poll2 constant area (p2::CA2, run words relative to offset 0)
};

enum : int { CTR_GROW = 0, CTR_PROW = 1, CTR_DESC = 2, CTR_OVERFLOW = 3, CTR_SLOW = 4, CTR_COUNT = 8,
             CTR_TOTALS = 8,   // {lines, bytes} as two u64 right behind the counters: one read-back
             CTR_SET = 16 };   // u32 words per counter set
enum : uint32_t { OVF_GROW = 1, OVF_PROW = 2, OVF_DESC = 4, OVF_OUT = 8 };

// Per-batch workspace ("scratch"): token tables, join state, line plan.
struct WorkDev {
  GpuRow *grow;   uint32_t grow_cap;
  ProcRow *prow;  uint32_t prow_cap;
  uint32_t *uniq;                // [grow_cap] k-th unique gpu of a doc -> local row index
  uint8_t *fstr;                 // [grow_cap * FSTR_STRIDE] repr(mem_util) per values row
  LineDesc *desc; uint32_t desc_cap;
  uint32_t *line_len;            // [desc_cap]
  uint32_t *line_loff;           // [desc_cap] byte offset of the line inside its document
  uint32_t *counters;            // [CTR_COUNT]
  uint32_t *doc_g0, *doc_ng, *doc_p0, *doc_np, *doc_perr;   // stage 1 -> stage 2
  uint32_t *doc_desc0, *doc_nlines, *doc_nbytes;            // stage 2 -> scan / stage 4
  int32_t *doc_status;
  uint64_t *doc_line_off;        // n_docs + 1 (result)
  uint64_t *doc_byte_off;        // n_docs + 1
  uint64_t *totals;              // [0] lines, [1] bytes
  uint64_t *line_off;            // [desc_cap + 1] (result)
  uint8_t *out;                  // encoded bytes (result)
  uint64_t out_cap;
};

// The 16 bytes in front of the lines hold {n_bytes, n_lines} of the run that wrote them, so
// that device-side consumers (the shard gather) need no host round trip for the sizes.
__device__ __forceinline__ void write_out_header(const WorkDev &w, uint64_t lines, uint64_t bytes) {
  uint64_t *h = reinterpret_cast<uint64_t *>(w.out) - 2;
  h[0] = bytes;
  h[1] = lines;
}

// Launch parameters of k_poll (kgm_fused.cuh): the layout of a warp's shared-memory slice
// (byte offsets from the slice start, all 16-byte aligned) and the scan / ticket state.
struct FusedParams {
  uint32_t cap_g, cap_a;         // text windows (bytes, incl. slack)
  uint32_t cap_ng, cap_np;       // rows per document
  uint32_t cap_c;                // containers per node
  uint32_t cap_t;                // template block
  uint32_t cap_lines;            // lines per document
  uint32_t warp_bytes, stage_bytes;   // bytes of a team's slice / of its stage
  uint32_t teams;                // teams (documents in flight) per CTA: 3 or 4
  uint32_t off_tg, off_ta, off_gstart, off_pstart, off_grec, off_prec, off_cnt, off_cstart, off_plan,
      off_loff, off_tb, off_stage, off_mbox;
  uint32_t epoch;                // tags the look-back state of this launch
  uint32_t allow_fast;           // 0 = every document takes the generic path (testing)
  unsigned long long ticket_base;
  unsigned long long *ticket;    // monotonic document ticket counter
  unsigned long long *scan_state;   // 2 words per document
  unsigned long long *scan_blocks;  // 2 words per block of 32 documents
  uint32_t *next_counters;       // the counter set of the next launch (zeroed by this one)
  const unsigned long long *carry;  // {lines, bytes} emitted by earlier launches of the same batch
  unsigned long long *post;         // mapped host memory (or null): {tag:16 | lines:48, tag:16 | bytes:48} once the totals are known
  unsigned long long post_seq;      // the tag (1 .. 65535)
                                    // (chunked end-to-end path), or null
};

// Launch parameters of k_poll2 (kgm_poll2.cuh): the layout of a warp's shared-memory slice (byte
// offsets from the slice start, all 16-byte aligned).  The scan / ticket state is FusedParams'.
struct Poll2Params {
  uint32_t warps;                // warps (documents in flight) per CTA
  uint32_t slice_bytes;
  uint32_t cap_g, cap_a;         // text windows (bytes, incl. slack)
  uint32_t cap_ng, cap_np;       // rows per document
  uint32_t cap_rec;              // process rows + idle pseudo-rows
  uint32_t cap_c;                // containers per node
  uint32_t cap_lines;            // lines per document
  uint32_t stage_bytes;          // incl. 64 bytes of slack
  uint32_t off_tg, off_ta, off_grec, off_prec, off_cnt, off_cstart, off_plan, off_loff, off_toff, off_stage;
  uint32_t off_gstart, off_pstart, off_cp, off_digest;   // parse-time tables, inside the stage
  uint32_t pods_first;           // compose the pod rounds before the node round
  uint32_t cta_tickets;          // tickets per CTA (`warps` documents at a time) instead of per warp
};

void launch_poll(const BatchDev &b, const FleetDev &f, const WorkDev &w, const FusedParams &P, int blocks,
                 int device, cudaStream_t s);
void launch_poll2(const BatchDev &b, const FleetDev &f, const WorkDev &w, const FusedParams &P, const Poll2Params &Q,
                  int blocks, int device, cudaStream_t s);
int poll2_smem_bytes(const Poll2Params &Q);
struct WinFront;
void launch_poll2_samples(const BatchDev &b, const FleetDev &f, const WorkDev &w, const FusedParams &P,
                          const Poll2Params &Q, const WinFront &wo, int blocks, int device, cudaStream_t s);
int poll2_max_warps();                          // warps per CTA k_poll2 was compiled for
int poll_smem_bytes(const FusedParams &P);
int debug_stage_clocks(unsigned long long *out64);   // 0 unless built with -DKGM_PROFILE_STAGES

void launch_tokenize(const BatchDev &b, const WorkDev &w, int sms, cudaStream_t s);
void launch_join(const BatchDev &b, const FleetDev &f, const WorkDev &w, int sms, cudaStream_t s);
void launch_scan(const BatchDev &b, const WorkDev &w, cudaStream_t s);
void launch_encode(const BatchDev &b, const FleetDev &f, const WorkDev &w, int sms, int device, cudaStream_t s);
void launch_flush(uint8_t *buf, uint64_t bytes, int sms, cudaStream_t s);
int encode_smem_bytes();

}  // namespace kgm
