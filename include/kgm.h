/* kgm.h -- C ABI of libkgm.so: the B200-native replacement for the body of the reference's
 * poll function.
 *
 * The reference (ma2331550908/k8s-gpu-monitor) has no plugin or FFI surface: it is one
 * script whose `main()` (gpu.py:90-256) turns subprocess texts into InfluxDB line
 * protocol and POSTs it.  This library replaces exactly the part of `main()` between
 * "texts acquired" and "lines POSTed" for a whole batch of polls ("documents") at once.
 * The Python shim a maintainer would bind it with is shown in INTEGRATION.md.
 *
 * Everything here is plain C: opaque context, pointers and sizes, no C++ or torch types.
 * All functions return KGM_OK (0) or a negative KGM_E_* code; they never throw, abort or
 * print.  One host thread per context at a time (the reference is single-threaded,
 * gpu.py:258-265).  There is no CPU implementation behind this ABI: without a CUDA
 * device kgm_create fails with KGM_E_CUDA.
 */
#ifndef KGM_H
#define KGM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KGM_ABI_VERSION 3

/* ---- call status ---------------------------------------------------------------- */
#define KGM_OK          0
#define KGM_E_INVALID  (-1)  /* bad argument (NULL, inconsistent offsets, unknown node id)  */
#define KGM_E_CUDA     (-2)  /* CUDA runtime / driver error; kgm_last_cuda_error has detail */
#define KGM_E_NOMEM    (-3)  /* host or device allocation failed                            */
#define KGM_E_LIMIT    (-4)  /* input beyond a documented limit (see DESIGN.md "Limits")    */
#define KGM_E_STATE    (-5)  /* call order: no fleet loaded / no batch resident             */

/* ---- per-document status ------------------------------------------------------------
 * The reference has no try/except: a malformed input raises and kills the poll, and the
 * lines POSTed before the raise stay written.  The library reproduces
 * that per document: it emits exactly the lines the reference would have POSTed and
 * reports the class of the exception here instead of dying. */
#define KGM_DOC_OK              0
#define KGM_DOC_KEY_ERROR       1   /* KeyError            gpu.py:51,53,82,133-143,151,172,204 */
#define KGM_DOC_VALUE_ERROR     2   /* ValueError          gpu.py:133-134 (int(''))            */
#define KGM_DOC_ZERO_DIVISION   3   /* ZeroDivisionError   gpu.py:135                          */
#define KGM_DOC_ATTRIBUTE_ERROR 4   /* AttributeError      gpu.py:186 (Labels: null)           */
#define KGM_DOC_INDEX_ERROR     5   /* IndexError          gpu.py:199,233                      */
#define KGM_DOC_CSV_ERROR       6   /* _csv.Error          gpu.py:48 (stray CR in a row)       */
#define KGM_DOC_JSON_ERROR      7   /* JSONDecodeError     gpu.py:88                           */
#define KGM_DOC_TYPE_ERROR      8   /* TypeError           gpu.py:88,184                       */
#define KGM_DOC_OTHER_ERROR     9   /* any other exception class raised host-side             */
/* Inputs outside the supported dialect: no lines are emitted for the document from the
 * point of detection on, and the status says why (documented divergences). */
#define KGM_DOC_UNSUPPORTED_QUOTE   20  /* a CSV cell starts with '"' (quoted-field syntax)   */
#define KGM_DOC_UNSUPPORTED_BYTE    21  /* byte >= 0x80 in an nvidia-smi text                 */
#define KGM_DOC_UNSUPPORTED_RANGE   22  /* >19-digit memory value, row > 65535 B, line > 8 KiB */

typedef struct kgm_ctx kgm_ctx;

/* n strings stored back to back: string i is bytes[off[i] .. off[i+1]). */
typedef struct kgm_strings {
  const uint8_t  *bytes;
  const uint32_t *off;        /* n + 1 entries */
} kgm_strings;

/* ---- fleet state: the docker-side inputs of the join ---------------------------------
 * Slow-changing dimension data, built by the host shim with the reference's own parsing
 * of `docker ps/stats/top/inspect` and `nvidia-smi -L` text (gpu.py:86-103,172-199,
 * 221-233) and kept resident in HBM.  A "node" here is one snapshot of one host's docker
 * state; documents refer to it by index.  Containers of node i are
 * [cont_begin[i], cont_begin[i+1]) in `docker stats` row order (gpu.py:172,174).       */
typedef struct kgm_fleet {
  uint32_t        n_nodes;
  kgm_strings     node_name;      /* NODE_NAME as printed by '%s' (gpu.py:131; "None" if unset) */
  const uint32_t *cont_begin;     /* n_nodes + 1                                                */
  const uint8_t  *node_err_parse; /* KGM_DOC_* raised while parsing docker ps/stats (gpu.py:92-103), else 0 */
  const uint8_t  *node_err_join;  /* KGM_DOC_* raised at gpu.py:172 (stats id not in ps), else 0 */

  uint32_t        n_containers;
  const uint8_t  *gated;          /* labels['mlgpu'] == 'mlgpu'          (gpu.py:185-188)      */
  const uint8_t  *err_inspect;    /* KGM_DOC_* raised at gpu.py:184-187 for this container, else 0 */
  const uint8_t  *err_emit;       /* KGM_DOC_* raised at gpu.py:193-199 / 227-233 when its first
                                     line is about to be formatted, else 0                      */
  kgm_strings     cid;            /* container id as in docker stats                            */
  kgm_strings     ns;             /* namespace tag value  ("-" without k8s labels)              */
  kgm_strings     pod;            /* podname field value  ("-" ...)                             */
  kgm_strings     user;           /* user_name derived per gpu.py:196-199 ("-" ...)             */

  const uint32_t *pid_begin;      /* n_containers + 1: pid strings of container c               */
  kgm_strings     pid;            /* lines 2..n-1 of `docker top c -eo pid` (gpu.py:177); each
                                     at most 15 bytes (KGM_E_LIMIT otherwise)                   */
  const uint32_t *vis_begin;      /* n_containers + 1: uuids visible inside container c         */
  kgm_strings     vis;            /* whitespace-split output of gpu.py:221-224                  */
} kgm_fleet;

/* ---- one batch of documents ------------------------------------------------------------
 * Document d = one poll of one node: the two nvidia-smi texts of gpu.py:104-116, stored
 * concatenated.  gpu_off / apps_off have n_docs + 1 entries; each text total < 4 GiB.   */
typedef struct kgm_batch {
  uint32_t        n_docs;
  const uint8_t  *gpu_csv;
  const uint64_t *gpu_off;
  const uint8_t  *apps_csv;
  const uint64_t *apps_off;
  const uint32_t *doc_node;       /* n_docs: index into the loaded fleet's nodes               */
  const uint32_t *doc_tick;       /* n_docs or NULL: tick index, used only by kgm_window_*     */
} kgm_batch;

/* Result of one encode.  `lines` is every emitted line followed by '\n', documents in
 * batch order, lines of a document in the reference's POST order.  All pointers belong to
 * the context and stay valid until the next kgm_upload_batch / kgm_encode_* / kgm_destroy
 * on it.  They are HOST pointers from kgm_encode_batch(..., want_host_copy=1) and DEVICE
 * pointers otherwise.                                                                     */
typedef struct kgm_result {
  const uint8_t  *lines;
  uint64_t        n_bytes;
  const uint64_t *line_off;       /* n_lines + 1 byte offsets into `lines`                     */
  uint64_t        n_lines;
  const uint64_t *doc_line_off;   /* n_docs + 1: first line of each document                   */
  const int32_t  *doc_status;     /* n_docs: KGM_DOC_*                                         */
  uint32_t        n_docs;
} kgm_result;

/* Device time of the last encode, measured with CUDA events on the context's stream.   */
typedef struct kgm_timing {
  float h2d_ms;        /* host->device copies of the batch (0 for a resident batch)          */
  float tokenize_ms;   /* stage 1 kernel                                                     */
  float join_ms;       /* stage 2 kernel                                                     */
  float scan_ms;       /* device-wide exclusive scan of per-document sizes                   */
  float encode_ms;     /* stage 4 byte-writing kernel                                        */
  float d2h_ms;        /* device->host copies of the result (0 without host copy)            */
  float total_ms;      /* first event to last event                                          */
  uint32_t launches;   /* kernels launched by the library in that region                     */
  uint32_t fused;      /* 1: the batch went through the fused single-pass kernel; its time is
                          reported as encode_ms and tokenize/join/scan are 0                  */
  uint32_t slow_docs;  /* fused kernel: documents that took the generic per-byte path inside it */
  uint32_t fused_ctas_per_sm;  /* fused kernel: resident CTAs per SM its shared-memory slices allowed */
  uint32_t docs_in_flight_per_sm;  /* fused kernel: documents its CTAs of one SM work on at a time     */
} kgm_timing;

int  kgm_abi_version(void);
const char *kgm_strerror(int status);            /* static string for KGM_E_* / KGM_DOC_*  */

/* Replaces the process set-up of gpu.py:258-260.  `device` is a CUDA ordinal.           */
int  kgm_create(kgm_ctx **out, int device);
void kgm_destroy(kgm_ctx *ctx);
const char *kgm_last_cuda_error(kgm_ctx *ctx);

/* Replaces the docker-side reads of gpu.py:92-103,172,177,184-199,221-233 with a
 * resident table (see kgm_fleet).  Copies everything; the caller's arrays may be freed. */
int  kgm_load_fleet(kgm_ctx *ctx, const kgm_fleet *fleet);

/* Replaces gpu.py:104-256 (parse both nvidia-smi texts, node lines, join, pod lines) for
 * a batch held in HOST memory: copies the batch to the device, runs the pipeline and, if
 * want_host_copy != 0, copies lines / offsets / statuses back into pinned host buffers.
 * A large batch is cut into chunks whose text copies, kernels and result copies overlap;
 * the result is the same bytes.  The batch is resident afterwards.                        */
int  kgm_encode_batch(kgm_ctx *ctx, const kgm_batch *batch, int want_host_copy, kgm_result *out);

/* The same in two steps, for callers that keep a batch resident in HBM and run it more
 * than once (bench.py's device-timed region).  kgm_encode_resident returns device
 * pointers.  For kgm_upload_batch the two texts (gpu_csv, apps_csv) may also be DEVICE
 * pointers (texts produced on the GPU, e.g. by kgm_replay_fill); the offset and per-document
 * arrays are always host memory (they are validated on the host).                          */
int  kgm_upload_batch(kgm_ctx *ctx, const kgm_batch *batch);
int  kgm_encode_resident(kgm_ctx *ctx, kgm_result *out);

/* kgm_encode_resident in two halves (no counterpart in the reference), for callers that
 * keep several contexts busy or hand the lines to device-side consumers: _async queues the
 * kernels on the context's stream and returns at once -- `out` holds the DEVICE pointers the
 * result will be at, with n_bytes = n_lines = 0; kgm_encode_finish waits, re-runs the batch if
 * a workspace guess was too small (the pointers may then differ) and fills in the sizes.
 * In every device result the 16 bytes in front of `lines` hold {uint64 n_bytes, uint64
 * n_lines}, written by the kernels in stream order, so a consumer queued behind the
 * context's stream needs no host round trip for the sizes.  One async call per context at
 * a time; any other encode/upload call on the context completes it first.                */
int  kgm_encode_resident_async(kgm_ctx *ctx, kgm_result *out);
int  kgm_encode_finish(kgm_ctx *ctx, kgm_result *out);

/* Where the encodes (kgm_encode_resident, _async, and the device side of kgm_encode_batch) write
 * their result (no counterpart in the reference): `dev_ptr` is any 16-byte aligned memory the context's
 * device can write -- its own, or a peer GPU's mapped into this process (CUDA IPC, peer
 * access), which is how the shards of a multi-GPU job land on the gathering GPU while the
 * kernel that encodes them is still running.  The 16-byte {n_bytes, n_lines} header goes
 * to dev_ptr, the lines to dev_ptr + 16 (kgm_result.lines); `bytes` counts in whole
 * 16-byte units.  A result that does not fit fails with KGM_E_LIMIT.  NULL switches back to the context's own buffer.                */
int  kgm_set_output_buffer(kgm_ctx *ctx, void *dev_ptr, uint64_t bytes);

/* The CUDA stream (cudaStream_t) the context queues its work on, for ordering a caller's
 * own streams against it with events.                                                     */
void *kgm_stream(kgm_ctx *ctx);

/* Bytes that may be read at a device result's `lines` (>= n_bytes rounded up to 16): the
 * size of the context's output buffer, for consumers that move fixed-size slots.          */
uint64_t kgm_output_capacity(kgm_ctx *ctx);

/* Copies the last result of kgm_encode_resident to pinned host memory.                  */
int  kgm_fetch_result(kgm_ctx *ctx, kgm_result *out_host);

int  kgm_get_timing(kgm_ctx *ctx, kgm_timing *out);
/* Which kernel the last run went through: 0 = the four staged kernels, 1 = k_poll (two-warp
 * teams), 2 = k_poll2 (one warp per document); negative = KGM_E_*.                        */
int  kgm_last_kernel(kgm_ctx *ctx);

/* ---- stage 3: time-window reductions -------------------------------------------------
 * No counterpart in the reference: gpu.py:136-143,206-216 emit raw per-poll lines with no
 * timestamp and leave aggregation to InfluxQL / Grafana.  This is an ADDITIVE output; it
 * never changes a byte of the node_gpu / pod_gpu lines.  Every emitted line of the resident
 * batch is one sample of a series
 *     kind 0  node_gpu line: (host, gpu_id)              -- its tag set, gpu.py:137
 *     kind 1  pod_gpu line:  (host, container, gpu_id)   -- "per (node, pod, gpu)"
 * in window doc_tick / window_ticks.  Values are the integers behind the line's fields
 * (digit strip as in gpu.py:133-134,138-143,204-205, saturating at 2^32-2): metric 0 =
 * gpu_util, 1 = mem_used (node line) / used_memory (pod line), 2 = the digits of a 7th
 * --query-gpu column (power.draw of an extended query; absent in the reference's query).  A
 * line whose field has no digits contributes no sample to that metric (m_count).  Groups
 * are returned in ascending (kind, node|container, gpu_id, window) order; gpu_id 0xffff is
 * the idle branch's "None" (gpu.py:226).  last = the value of the line with the greatest
 * (tick, position in the batch).  mean = sum / m_count in IEEE double.
 * The batch must have been uploaded with doc_tick.  Pointers belong to the context and
 * stay valid until the next call on it (host pointers iff want_host_copy != 0).          */
typedef struct kgm_windows {
  uint64_t        n;
  const uint8_t  *kind;
  const uint32_t *node;          /* node index of the series                                */
  const uint32_t *container;     /* global container index (kind 1), 0xffffffff (kind 0)     */
  const uint32_t *gpu_id;
  const uint32_t *window;
  const uint32_t *count;         /* lines in the group                                       */
  const uint32_t *m_count[3];
  const uint32_t *m_min[3];
  const uint32_t *m_max[3];
  const uint32_t *m_last[3];
  const uint64_t *m_sum[3];
  const double   *m_mean[3];
} kgm_windows;
int  kgm_window_reduce(kgm_ctx *ctx, uint32_t window_ticks, int want_host_copy, kgm_windows *out);
/* Which front end the last kgm_window_reduce took its samples from: 1 = the planning half of k_poll2
 * (one launch over the resident texts), 0 = the staged tokenizer + join (any document the fast path
 * declines sends the whole batch there; KGM_WINDOW_STAGED_FRONT=1 forces it); negative = KGM_E_*.  */
int  kgm_window_front(kgm_ctx *ctx);

/* Debug aid (no counterpart in the reference): when the library was built with
 * -DKGM_PROFILE_STAGES, copies the fused kernel's per-stage SM-clock sums (64 values: 32
 * stages for each of a team's two warps, accumulated since load) to `out64` and returns 64;
 * returns 0 in a normal build.                                                            */
int  kgm_debug_stage_clocks(uint64_t *out64);

/* Debug aid (no counterpart in the reference): after a chunked end-to-end kgm_encode_batch
 * call, the times (ms since the call's first device operation) at which each chunk's text
 * had arrived (`text_in_ms`) and its kernel had finished (`kernel_done_ms`); both arrays take
 * up to `cap` values.  Returns the number of chunks of the last call (1: not chunked).      */
int  kgm_debug_chunk_times(kgm_ctx *ctx, float *text_in_ms, float *kernel_done_ms, uint32_t cap);

/* Which kernels run a batch (no counterpart in the reference; results are identical):
 *   AUTO           one fused single-pass kernel when the batch fits its shared-memory slices
 *                  (k_poll for batches of a few waves, k_poll2 for large ones), the staged
 *                  kernels otherwise (default)
 *   STAGED         always tokenize -> join -> scan -> encode as four kernels
 *   FUSED_GENERIC  the fused kernel with its fast path switched off: every document takes
 *                  the generic per-byte path inside it (testing)
 * The environment variable KGM_PIPELINE=staged|fused-generic sets the initial mode.        */
#define KGM_PIPELINE_AUTO          0
#define KGM_PIPELINE_STAGED        1
#define KGM_PIPELINE_FUSED_GENERIC 2
#define KGM_PIPELINE_FUSED_V1      3   /* always the team kernel k_poll (two warps per document)            */
#define KGM_PIPELINE_FUSED_V2      4   /* always the warp kernel k_poll2 (one warp per document)            */
int  kgm_set_pipeline(kgm_ctx *ctx, int mode);

/* Writes `bytes` of device scratch on the context's stream so that the L2 holds none of
 * the batch afterwards (timing hygiene between benchmark iterations).                    */
int  kgm_flush_l2(kgm_ctx *ctx, uint64_t bytes);

/* Page-locked host memory for batch texts, so that the host->device copies inside
 * kgm_encode_batch run at full PCIe rate (plain malloc'ed memory also works, slower).     */
void *kgm_host_alloc(uint64_t bytes);
void  kgm_host_free(void *p);

/* Multi-GPU assembly (no counterpart in the reference, whose instances never talk to each
 * other: gpu_monitor.yaml:2 runs one per node).  After the ranks' {bytes, lines} have been
 * all-gathered and their encoded shards have arrived in fixed-stride slots on one GPU, this
 * packs the slots into the rank-major concatenation without the host ever learning the
 * sizes: shard r is slots[r * stride .. + counts[2 * r]).  All pointers are DEVICE pointers
 * on the current device; `counts` holds world x {int64 bytes, int64 lines}; totals_out (may
 * be NULL) receives {bytes, lines} of the concatenation.  With counts == NULL every slot
 * carries its own sizes instead: {int64 bytes, int64 lines} in its first 16 bytes, the
 * shard behind them (so that sizes and bytes arrive in one transfer).  A shard whose size
 * exceeds its slot or the room left in `out` is left out, and totals_out then receives
 * {UINT64_MAX, UINT64_MAX}: the output must not be used.  Asynchronous on `stream`.          */
int  kgm_concat_shards(const uint8_t *slots, uint64_t stride, const int64_t *counts, uint32_t world,
                       uint8_t *out, uint64_t out_cap, uint64_t *totals_out, void *stream);

/* Docker-side text on the device: what the reference reads from
 * `docker inspect <container>` on every poll -- json.loads(text.replace("\n","").replace(" ",""))
 * [0]['Config']['Labels'], the gate labels['mlgpu'] == 'mlgpu', io.kubernetes.pod.name / .namespace
 * and the user_name split rules (gpu.py:86-88,184-199) -- for n containers at once, one thread per
 * container.  `texts` holds the n texts back to back, off[n + 1] their byte offsets (host memory).
 * The scanner takes a subset of what json.loads accepts (csrc/kgm_docker.cuh says which) and is exact
 * on it; for every other container status[i] = 1 and the caller parses that text itself
 * (kgm/fleet.py does, with json.loads: that is where the reference's raise points are classified).
 * For status[i] = 0: gated[i]; has_pod[i]; pod / ns / user = runs of container i's text (user only
 * for a gated container with a pod name).  Synchronous; allocates and frees its own device memory. */
typedef struct {
  uint8_t  *status, *gated, *has_pod;                 /* [n] each                                    */
  uint32_t *pod_off, *pod_len, *ns_off, *ns_len, *user_off, *user_len;   /* [n] each, relative to the container's text */
} kgm_inspect_out;
int  kgm_inspect_labels(int device, const uint8_t *texts, const uint64_t *off, uint32_t n, const kgm_inspect_out *out);

/* Replay generator on the device (bench / test plumbing; no counterpart in the reference, which
 * reads live `nvidia-smi` output, gpu.py:104-116): writes the two nvidia-smi texts of n_nodes x
 * n_ticks polls (host-major: all ticks of host 0, then host 1, ...) straight into device memory,
 * byte for byte what k8s-gpu-monitor_b200/kgm/replay.py make_texts() produces for the same
 * coordinates (every value is a counter-based hash of (seed, host, gpu or process, tick)).
 * All pointers in the spec and in the calls are DEVICE pointers on the current device.
 *   kgm_replay_offsets  fills gpu_off / apps_off (n_docs + 1 each: exclusive byte offsets, the last
 *                       entry = the total) from the sizes; len_scratch: 2 x n_docs words
 *   kgm_replay_fill     writes the texts at those offsets
 * Both are asynchronous on `stream`.                                                          */
typedef struct {
  uint64_t seed;
  uint32_t n_nodes;
  uint32_t mem_total_mib;        /* what the memory.total column prints                          */
  const uint32_t *node_id;       /* [n_nodes] the host's number in the fleet (a hash coordinate) */
  const uint8_t  *mig;           /* [n_nodes] 0 plain, 1 MIG parents ([N/A] utilization), 2 also MIG uuids in process rows */
  const uint32_t *gpu_begin;     /* [n_nodes + 1] first gpu of a host in `uuid`                  */
  const uint8_t  *uuid;          /* 40 bytes per gpu ("GPU-8-4-4-4-12"), no terminator           */
  const uint32_t *proc_begin;    /* [n_nodes + 1] first process row of a host, rows GPU-major    */
  const uint32_t *proc_pid;
  const uint16_t *proc_ci;       /* container ordinal and process ordinal inside it (hash coordinates) */
  const uint8_t  *proc_j;
  const uint8_t  *proc_gpu;      /* gpu ordinal on the host                                      */
  const uint8_t  *proc_name;     /* index into the process-name pool                             */
  const uint8_t  *names;         /* process-name pool                                            */
  const uint32_t *name_off;      /* [names + 1]                                                  */
} kgm_replay_spec;
int  kgm_replay_offsets(const kgm_replay_spec *spec, uint32_t tick0, uint32_t n_ticks, uint32_t *len_scratch,
                        uint64_t *gpu_off, uint64_t *apps_off, void *stream);
int  kgm_replay_fill(const kgm_replay_spec *spec, uint32_t tick0, uint32_t n_ticks, const uint64_t *gpu_off,
                     const uint64_t *apps_off, uint8_t *gpu_csv, uint8_t *apps_csv, void *stream);

/* Stream-ordered counters between the GPUs of one job (no counterpart in the reference).  The
 * ranks tell each other "my shard of step k has landed in your slot" and "that slot set is free
 * again" through 32-bit counters that only grow, kept in memory every rank has mapped (CUDA IPC,
 * peer access), instead of running a collective per step.
 *   kgm_stream_signal_flags  behind everything queued on `stream` so far, raises each of the
 *                            n <= KGM_MAX_FLAG_TARGETS counters to `value` (system-scope atomic max
 *                            with release: signals from different streams may arrive out of order)
 *   kgm_stream_wait_flags    holds `stream` until each of the n <= 32 consecutive counters at `flags`
 *                            has reached `value` (wrap-safe comparison; a one-warp kernel that polls)
 * All pointers are device pointers valid on the current device.  With KGM_FLAG_MEMOPS=1 in the
 * environment both use stream memory operations (cuStreamWriteValue32 / cuStreamWaitValue32) instead of
 * kernels; a signal is then a plain store, so every signalling stream needs a counter of its own.  */
#define KGM_MAX_FLAG_TARGETS 16
int  kgm_stream_signal_flags(void *stream, uint32_t *const *targets, uint32_t n, uint32_t value);
int  kgm_stream_wait_flags(void *stream, const uint32_t *flags, uint32_t n, uint32_t value);

/* One step of a multi-GPU job in one call (no counterpart in the reference): on the context's
 * stream, wait until the counter at `wait_flag` has reached `wait_value` (skipped if NULL), encode
 * the resident batch into `slot` (as kgm_set_output_buffer + kgm_encode_resident_async), then
 * store `signal_value` to `signal_flag` (skipped if NULL).  `slot` is typically this rank's
 * fixed-stride slot on the step's gathering GPU.  Finish with kgm_encode_finish.             */
int  kgm_push_step(kgm_ctx *ctx, void *slot, uint64_t slot_bytes, const uint32_t *wait_flag, uint32_t wait_value,
                   uint32_t *signal_flag, uint32_t signal_value, kgm_result *out);

/* The same step with the shard carried by a copy engine instead of the poll kernel's own
 * stores (no counterpart in the reference): the context encodes into its own buffer; on a second
 * stream, behind the kernel, wait for `wait_flag`, copy {header, lines} (`slot_bytes`, at most
 * the context's output capacity) to `slot` -- typically peer memory -- and store `signal_value`
 * to `signal_flag`.  The kernel never waits for another GPU and writes local memory only; the
 * copy overlaps the next step's kernel.  The context's buffer is reused only after its last copy
 * has left it.                                                                               */
int  kgm_push_copy_step(kgm_ctx *ctx, void *slot, uint64_t slot_bytes, const uint32_t *wait_flag,
                        uint32_t wait_value, uint32_t *signal_flag, uint32_t signal_value, kgm_result *out);

/* Stage-1 tokenizer for callers that may not allocate device memory.  (No counterpart in
 * the reference; it exists for deployment reasons only.)
 *
 * Who needs it: a collector that shares a GPU with a job that has reserved all of its HBM,
 * or one that runs under a memory manager which owns every allocation -- there a
 * cudaMalloc either fails or stalls the other tenant while the driver synchronises the
 * device.  This entry point calls no cudaMalloc/cudaFree and no allocating library routine:
 * the caller passes every input and output as DEVICE pointers on the current device, and
 * the kernel's working storage (row cursor, overflow flag, row table) is a fixed array
 * inside the library, sized for KGM_NOALLOC_MAX_ROWS rows (about 1.3 MB; every GPU that
 * loads the library holds it until the process exits).
 *
 * It tokenizes the --query-gpu text of a batch (gpu.py:104-109 semantics, as
 * kgm_encode_batch does internally) and writes, for the rows of document d at
 * [doc_row0[d], doc_row0[d] + doc_nrows[d]) (documents land in no particular order):
 * util = int(digits(utilization.gpu)), mem_used / mem_total = int(digits(memory.used/total)),
 * and per document doc_err = KGM_DOC_* of a parse-level failure (0 if none).
 *
 *   gpu_csv   16-byte aligned, with at least 16 readable bytes after n_bytes
 *   gpu_off   n_docs + 1 offsets;  n_bytes = gpu_off[n_docs] < 4 GiB
 *   max_rows  capacity of util / mem_used / mem_total; at most KGM_NOALLOC_MAX_ROWS,
 *             larger values are refused with KGM_E_LIMIT before anything runs; KGM_E_LIMIT
 *             is also returned if the text turns out to hold more rows than max_rows
 *   stream    a cudaStream_t (NULL = default stream); the call returns after the work
 *             on it has finished
 *
 * The first call on a device looks up the device address of the fixed array
 * (cudaGetSymbolAddress); make that first call before any stream capture begins.  Every call
 * copies the overflow flag back and waits for the stream, so it returns with the outputs
 * complete.  All
 * calls on a device share the one array and it keeps the previous call's contents: the
 * entry point clears what has to start at zero itself, but two calls whose GPU work
 * overlaps (different streams or threads) race on it -- serialise them.                  */
#define KGM_NOALLOC_MAX_ROWS 16384
int  kgm_tokenize_gpu_noalloc(const uint8_t *gpu_csv, const uint64_t *gpu_off, uint32_t n_docs,
                              uint64_t n_bytes, uint32_t max_rows, uint32_t *doc_row0,
                              uint32_t *doc_nrows, uint32_t *doc_err, uint32_t *util,
                              uint64_t *mem_used, uint64_t *mem_total, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* KGM_H */
