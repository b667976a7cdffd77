/*
 * mcp_b200.h -- C ABI of the B200-native retrieval + level-scheduling hot path.
 *
 * The reference (control_plane.py) has no plugin / operator / FFI interface; its seams are
 * Python methods.  This header is the boundary a maintainer of the
 * reference would bind with ctypes (INTEGRATION.md shows the stub).  Each entry point names
 * the reference interface it stands behind.
 *
 * Conventions
 *   - every function returns 0 on success, a negative MCP_E* code on failure; the message is
 *     available from mcp_last_error() (thread-local, valid until the next failing call on the
 *     same thread).  No exception, no torch type, crosses this boundary.
 *   - the caller owns every buffer it passes; the library owns the handle and all device
 *     memory it allocates.  "host" pointers are ordinary process memory, "device" pointers are
 *     CUDA device memory on the index's device; `stream` is a cudaStream_t passed as void*
 *     (NULL = the legacy default stream).
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails.
 *   - an index may be used from any thread; calls on one index serialise on an internal lock.
 *   - ids are int32 service ids = id_base + row number; ranking key is (score, id) ascending,
 *     NaN scores (zero vector under cosine) rank last; when k exceeds the number of rows the
 *     tail is id -1 / score +inf.
 */
#ifndef MCP_B200_H
#define MCP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MCP_METRIC_COSINE 0 /* pgvector `<=>`: 1 - clamp(q.e / sqrt(|q|^2 |e|^2), -1, 1) */
#define MCP_METRIC_L2 1     /* pgvector `<->`: sqrt(sum (q_i - e_i)^2)                    */

#define MCP_OK 0
#define MCP_EINVAL (-1)   /* bad argument (NULL, size out of range, non-finite input)    */
#define MCP_ECUDA (-2)    /* a CUDA runtime / driver call failed                          */
#define MCP_ESTATE (-3)   /* call order (search before load, penalty before load ...)     */
#define MCP_ENOMEM (-4)   /* device or host allocation failed                              */

#define MCP_MAX_K 256       /* largest k a search accepts                                   */
#define MCP_MAX_DAG_NODES 64
#define MCP_MAX_WIDE_DAG_NODES 255 /* mcp_dag_levels_wide: single oversized plans                */

/* which coarse-scan kernel a search uses (mcp_index_set_option "path") */
#define MCP_PATH_AUTO 0
#define MCP_PATH_STREAM_FP32 1 /* K1: fp32 CUDA-core streaming scan                        */
#define MCP_PATH_GEMM_TF32 2   /* K2: tcgen05 / TMEM TF32 distance GEMM                    */
#define MCP_PATH_EXACT_FP64 3  /* full fp64 scan (the certificate's fallback, also a checker) */
#define MCP_PATH_GEMM_F16 4    /* K2 over the fp16 shadow table: kind::f16 MMAs, half the bytes, twice the tensor rate.
                                  AUTO takes it whenever the index holds a shadow (for fewer than 8 queries: from 8192 rows
                                  on; below that the streaming kernel)                                                   */

typedef struct mcp_index mcp_index;

/* counters since creation; see mcp_index_get_stats */
typedef struct mcp_stats {
    int64_t searches;          /* search calls                                             */
    int64_t queries;           /* queries answered                                         */
    int64_t rows_scanned;      /* sum over searches of n_rows (coarse pass)                */
    int64_t bytes_scanned;     /* rows_scanned * dim * 4                                    */
    int64_t candidates_rescored;
    int64_t certificate_failures; /* queries the exact-rescoring certificate sent to the fp64 scan */
    int64_t exact_scans;       /* queries answered by the full fp64 scan                    */
    int64_t kernel_launches;   /* kernels this library launched for searches               */
    int64_t coarse_timed;      /* coarse-scan stages timed (option "time_coarse")           */
    double coarse_ms;          /* sum of their CUDA-event durations, milliseconds           */
    int32_t last_path;         /* MCP_PATH_* of the last search                            */
    int32_t last_kprime;       /* oversampled candidate count of the last search           */
} mcp_stats;

/* ---- lifetime ------------------------------------------------------------------------------
 * Stands where GraphPlanner.__init__ opens its pgvector connection (control_plane.py:46-49):
 * the handle is the "connection" to the embedding table, now resident in HBM.              */
int mcp_index_create(mcp_index **out, int device_id, int dim, int metric);
int mcp_index_destroy(mcp_index *idx);   /* waits for the device; must not race with any other call on the same handle */
/* The same "connection" over a table too large (or too hot) for one GPU: the reference is ONE process whose planner makes
 * plain method calls (control_plane.py:135-138, :57-75), so the sharding of SURVEY.md section 8e has to fit behind one
 * handle.  The handle owns one single-device index per entry of device_ids (contiguous row ranges, global ids) and one
 * worker thread per shard; mcp_index_load / _set_penalty / _update_rows split their host arrays by rows, mcp_index_search
 * fans the queries out, every GPU answers them on its shard (certified exact, like a single-device index), the packed
 * (score, id) lists are copied device-to-device to device_ids[0] and merged there with the same (score, id) key, so the
 * answer is identical to the unsharded search.  Kernels only ever read their own GPU's memory.  Entry points that take
 * device pointers (adopt, search_device, ...) belong to one GPU and return MCP_ESTATE on such a handle.  A device may be
 * listed more than once (several shards on one GPU).                                                                   */
int mcp_index_create_multi(mcp_index **out, const int *device_ids, int n_devices, int dim, int metric);

/* ---- table load ----------------------------------------------------------------------------
 * Replaces GraphPlanner._fetch_embeddings_metadata (control_plane.py:51-55), the only read of
 * service_schemas.input_schema_vector: rows are the N x dim fp32 vectors in table order.
 * load copies host rows to the device; adopt uses caller-owned device memory in place
 * (row pitch `ld` floats, ld >= dim, ld % 4 == 0, base 16-byte aligned; the caller keeps it
 * alive and unmodified until destroy or the next load/adopt).  The adopted range, like every
 * other device pointer handed to an index (queries, penalty), must be ordinary device memory of
 * the index's own GPU from its first byte to its last: peer allocations, IPC imports, managed
 * memory and virtual ranges with a mapping on another GPU are refused with MCP_EINVAL.  (The
 * kernels and the TMA tensor maps K2 builds over these buffers read this GPU's memory only; a
 * sharded registry gives every GPU its own index over its own rows.)  Both compute per-row fp64
 * squared norms and reject NaN/Inf elements (pgvector rejects them at input).
 * id_base is the service id of row 0 (a row shard of a larger table passes its offset).     */
int mcp_index_load(mcp_index *idx, const float *rows_host, int64_t n_rows, int64_t id_base);
int mcp_index_adopt_device(mcp_index *idx, const float *rows_device, int64_t n_rows, int64_t ld,
                           int64_t id_base, void *stream);
/* The same read from a flat table file instead of Postgres rows: 64-byte header (int64 x 8:
 * magic "MCPTBL01", dim, n_rows, id_base, metric, 0, 0, 0) followed by n_rows x dim fp32, row-major.  Rows [first_row,
 * first_row + n_rows) are loaded (n_rows = -1: to the end; a rank of a sharded registry reads only its own range) through a
 * ring of pinned chunks: two reader threads, one cudaMemcpyAsync per chunk, reads and copies overlapped.  ids are the
 * file's id_base + row number in the file.                                                                              */
int mcp_index_load_file(mcp_index *idx, const char *path, int64_t first_row, int64_t n_rows);

/* Registry refresh without a reload: rows [first_row, first_row + n_rows) are replaced by
 * rows_host (tables made by mcp_index_load), or -- rows_host = NULL -- were changed in place by the owner of an adopted
 * table; their norms are recomputed (that in-place variant blocks until searches in flight have finished).           */
int mcp_index_update_rows(mcp_index *idx, int64_t first_row, const float *rows_host, int64_t n_rows);
/* The same for scattered rows, without stopping queries: rows_host [n_rows x dim] replace the table rows row_ids[i] (row
 * numbers, not service ids; tables made by mcp_index_load).  The rows are validated on the host, staged in pinned memory and
 * applied by one kernel on the index's own stream -- copy, exact norms, shadow rows -- so the call returns once the work is
 * enqueued: searches in flight finish on the old rows, later searches see the new ones (searches on another stream wait on
 * the device, not the host).  mcp_index_update_rows with rows_host != NULL takes the same route.  A failed validation
 * (NaN / Inf) changes nothing.                                                                                          */
int mcp_index_upsert_rows(mcp_index *idx, const int64_t *row_ids, const float *rows_host, int64_t n_rows);

/* ---- telemetry fold --------------------------------------------------------------------------
 * One fp32 penalty per service, added to the distance before ranking.  The reference names the
 * fields (cost_profile: control_plane.py:31, README.md:93; latency / error rate: README.md:43-48)
 * but holds no formula; the host layer folds them (telemetry.py) and hands the result here.
 * NULL clears the penalty.  May be called between searches at any time.                      */
int mcp_index_set_penalty(mcp_index *idx, const float *penalty_host);
int mcp_index_set_penalty_device(mcp_index *idx, const float *penalty_device, void *stream);

/* ---- search ----------------------------------------------------------------------------------
 * The candidate-retrieval step that belongs between control_plane.py:58 and :65 in
 * GraphPlanner.plan: ORDER BY input_schema_vector <op> query (+ penalty) LIMIT k, exact.
 * out_ids [n_queries*k], out_scores [n_queries*k] fp32 (the fp64 score rounded once);
 * out_scores64 may be NULL.  The host variant copies in/out and returns finished results.
 * The device variant only enqueues work on `stream`; queries that fail the exactness
 * certificate are re-answered on the device by the fp64 scan inside the same call, however
 * many there are.  mcp_index_sync_status reports the rare leftovers (n_unresolved: queries
 * whose fallback list overflowed); mcp_index_repair_device re-answers those (the host variant
 * does all of this itself).                                                                  */
int mcp_index_search(mcp_index *idx, const float *queries_host, int n_queries, int k,
                     int32_t *out_ids, float *out_scores, double *out_scores64);
int mcp_index_search_device(mcp_index *idx, const float *queries_device, int n_queries, int k,
                            int32_t *out_ids_device, float *out_scores_device,
                            double *out_scores64_device, void *stream);
/* blocks until every search_device issued on this index has finished and reports the totals since
 * the previous sync_status; n_* may be NULL.  n_unresolved > 0 means that many queries still hold
 * uncertified results: call mcp_index_repair_device.  Consecutive
 * search_device calls on the same stream queue behind one another without host waits; a call on
 * a different stream waits for the previous one first (they share the index's workspace).     */
int mcp_index_sync_status(mcp_index *idx, int32_t *n_certificate_failures, int32_t *n_unresolved);

/* After mcp_index_sync_status reported n_unresolved > 0 for ONE search_device call: the same arguments again.  Exactly the
 * unresolved queries (their fallback list overflowed: thousands of rows within the coarse error band of the k-th score) are
 * re-answered in place by the dense fp64 scan; every other row of the outputs is left alone.  Blocks until done.  (The host
 * variant mcp_index_search does this itself.)  With several search_device calls between two sync_status calls the list of
 * unresolved queries mixes them: sync after each call whose answers you mean to repair.                                  */
int mcp_index_repair_device(mcp_index *idx, const float *queries_device, int n_queries, int k,
                            int32_t *out_ids_device, float *out_scores_device,
                            double *out_scores64_device, void *stream);

/* validation aid: the coarse score (distance + penalty, fp32) of every (query,row) pair as one coarse
 * kernel computes it, out [n_queries * n_rows]; path = MCP_PATH_STREAM_FP32 or MCP_PATH_GEMM_TF32.
 * The tests use it to check the error bounds the certificate assumes.  No reference counterpart. */
int mcp_index_coarse_scores_device(mcp_index *idx, int path, const float *queries_device, int n_queries,
                                   float *out_scores_device, void *stream);

/* ---- exact search that allocates nothing ----------------------------------------------------------
 * (No counterpart in the reference.)  For callers that must not allocate device memory at request
 * time: a planner thread serving /plan while another stream of the process is mid-scan (cudaMalloc /
 * cudaFree serialise against the whole device and would stall it), or a process whose HBM is fully
 * budgeted, where a late allocation can fail.  Runs the fp64 full scan -- the same answer as
 * mcp_index_search_device with option "path" = MCP_PATH_EXACT_FP64 -- taking all scratch from one fixed
 * array inside the library: MCP_NOALLOC_SCRATCH_BYTES of device memory held by every GPU that loads the
 * library until the process exits.  Limits: the index holds at most MCP_NOALLOC_MAX_ROWS rows and
 * n_queries <= MCP_NOALLOC_MAX_QUERIES; anything larger returns MCP_EINVAL.  The caller still passes
 * queries and outputs (device pointers); the table must already be loaded (that is where memory is
 * allocated).  The array is shared by every call on a device and keeps the previous call's contents; each
 * call rewrites what it reads, but calls whose GPU work overlaps (other streams, other indexes on the same
 * device) race on it -- keep them ordered.  The first call on a device looks up the array's device address
 * (one-time set-up): make that first call before starting any stream capture.  Ordering: mcp_index_update_rows,
 * _set_penalty, _load and _adopt_device wait for searches issued through the ordinary entry points, but NOT for a scan
 * issued through this one (it leaves no mark on the index) -- order those calls behind it yourself.                   */
#define MCP_NOALLOC_MAX_ROWS (1 << 20)
#define MCP_NOALLOC_MAX_QUERIES 64
#define MCP_NOALLOC_SCRATCH_BYTES ((8u << 20) + 4096u)
int mcp_index_search_exact_noalloc_device(mcp_index *idx, const float *queries_device, int n_queries, int k,
                                          int32_t *out_ids_device, float *out_scores_device,
                                          double *out_scores64_device, void *stream);

/* options: "path" (MCP_PATH_*), "kprime" (0 = automatic oversampling, else candidates kept per
 * query before exact rescoring, >= k; automatic starts at max(1.5 k, k + 32) and doubles, up to 4x, while more than 1 in 200
 * queries need the fp64 scan -- clustered registries -- and steps back after 64 clean status reads), "max_fallback" (queries one pass of the fp64 scan serves at a time, 1 .. 6; default 6),
 * "shadow" (default 1: load / adopt also build an fp16 image of the table (each row scaled by a power of two), +50 % device memory, which the batched scan
 * reads instead of the fp32 rows; the answer stays exact -- candidates are rescored in fp64 from the fp32 rows and certified
 * with a bound computed from the actual rounding residuals.  0 = no shadow, the batched scan reads the fp32 rows as tf32.
 * Set it before the table is loaded.  When the shadow does not fit in device memory the index quietly goes without),
 * "gemm_l2_prefetch" (1 = K2 with a per-row 1 KB L2 bulk prefetch; default 0, measured slower with either operand kind),
 * "gemm_balance" (1 = K2 tile height shrunk so that every CTA gets the same number of tiles; default 0 = 256-row tiles, measured equal),
 * "gemm_cluster" (CTAs per cluster that share a row tile: 0 = automatic -- 2 when the batch has an even number of
 * 128-query blocks, else 1 -- or 1 / 2 / 4),
 * "gemm_pair" (default 1 = a cluster of 2 runs as one tcgen05 CTA pair, cta_group::2 with M = 256; 0 = two M = 128
 * CTAs that multicast the row tile to each other -- on tables made by mcp_index_load only: over adopted memory the
 * non-pair shapes run as single CTAs with plain loads, multicast being kept to memory the library allocated itself),
 * "gemm_sync" (default 1: the CTAs of K2 that stream the same row tiles stay within this many tiles of each other, so the
 * tiles are served from L2 instead of being re-read from HBM by every query block; 0 = off),
 * "gemm_coop" (default 1; 0 = launch K2 without the cooperative attribute -- for Nsight Compute captures only: co-residency
 * of the CTAs K2's spin-waits rely on is then no longer guaranteed, so it is refused unless "gemm_debug" holds bit 4096),
 * "gemm_debug" (bit mask for timing experiments that switch off parts of K2 / K4 -- results are WRONG; default 0),
 * "gemm_bootstrap" (0 = K2 without the cross-CTA threshold bootstrap; default 1),
 * "time_coarse" (1 = bracket every coarse scan with CUDA events on the caller's stream; the sums
 * appear in mcp_stats after mcp_index_sync_status).                                            */
int mcp_index_set_option(mcp_index *idx, const char *name, int64_t value);
int mcp_index_get_stats(mcp_index *idx, mcp_stats *out);
int64_t mcp_index_rows(const mcp_index *idx);

/* ---- shard merge -----------------------------------------------------------------------------
 * Row-sharded registries: every shard answers the same queries, the per-shard sorted lists
 * [n_shards][n_queries][k] (as gathered by ncclAllGather) are merged with the same (score, id)
 * key.  No counterpart in the reference (it is single-process); north_star's exchange step.  */
int mcp_merge_topk_device(const double *scores64_device, const int32_t *ids_device, int n_shards,
                          int n_queries, int k, int32_t *out_ids_device, float *out_scores_device,
                          double *out_scores64_device, void *stream);

/* The same merge over ONE gathered buffer: shard s contributes a block of mcp_merge_packed_block_bytes(n_queries, k)
 * bytes laid out [n_queries*k fp64 scores][n_queries*k int32 ids][pad to 16], blocks back to back -- what one
 * ncclAllGather of each rank's block leaves behind (a search_device call can write its outputs straight into the
 * rank's block).  One collective per search instead of two.                                                   */
int64_t mcp_merge_packed_block_bytes(int n_queries, int k);
int mcp_merge_topk_packed_device(const void *packed_device, int n_shards, int n_queries, int k,
                                 int32_t *out_ids_device, float *out_scores_device,
                                 double *out_scores64_device, void *stream);

/* ---- DAG level scheduling --------------------------------------------------------------------
 * Batched replacement for the ordering step of Orchestrator.execute (control_plane.py:104,
 * nx.topological_sort -> networkx topological_generations, dag.py:166-240) on graphs built as
 * control_plane.py:94-100 builds them.  DAG b has n_nodes[b] <= 64 nodes numbered in insertion
 * order and owns the (src,dst) byte pairs edge_off[b] .. edge_off[b+1] of `edges`, in insertion
 * order (duplicates allowed: a DiGraph keeps one).
 *   out_level   [n_dags*64]  generation of each node; 0xFF = no such node / never emitted (cycle)
 *   out_order   [n_dags*64]  nodes in the order nx.topological_sort yields them; may be NULL
 *   out_n_levels[n_dags]     generations yielded
 *   out_status  [n_dags]     0 ok, 1 = NetworkXUnfeasible (cycle), 2 = more nodes than the kernel takes (device variants;
 *                            the host variants refuse such a batch with MCP_EINVAL)
 * The host variant copies in/out on `device_id` through a per-device workspace that only grows and a stream of its own:
 * once the workspace has seen a size a call allocates nothing, so the one-graph-per-request use in Orchestrator.execute
 * does not stall scans running on other streams.  A host batch of 262,144 DAGs or more is piped through in chunks (pinned
 * slots, two streams, three helper threads for the duration of the call), so staging, both copy directions and the kernel overlap.  The device
 * variant enqueues on `stream`.
 * Two kernels stand behind the <= 64-node pair, with identical bytes out: one warp per DAG (any batch size), and from
 * 65,536 DAGs up, when the order is asked for, one thread per DAG (2.7 x faster at 1 M DAGs; it needs `edges`,
 * `out_level` and `out_order` 16-byte aligned -- otherwise the warp kernel runs -- and reads `edges` in aligned 16-byte pieces,
 * so the buffer must be readable up to the 16-byte boundary after its last edge, as any cudaMalloc'd buffer is).  Environment variable
 * MCP_DAG_KERNEL = warp | fifo forces one of them (measurement only).
 * The _wide pair is the same for plans of 65 .. MCP_MAX_WIDE_DAG_NODES nodes (the reference's topological_sort takes any
 * size): one CTA per DAG, outputs with a pitch of 256 entries per DAG instead of 64.                                    */
int mcp_dag_levels(int device_id, const uint8_t *n_nodes, const uint32_t *edge_off,
                   const uint8_t *edges, int64_t n_dags, uint8_t *out_level, uint8_t *out_order,
                   uint8_t *out_n_levels, uint8_t *out_status);
int mcp_dag_levels_device(const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges,
                          int64_t n_dags, uint8_t *out_level, uint8_t *out_order,
                          uint8_t *out_n_levels, uint8_t *out_status, void *stream);
int mcp_dag_levels_wide(int device_id, const uint8_t *n_nodes, const uint32_t *edge_off,
                        const uint8_t *edges, int64_t n_dags, uint8_t *out_level, uint8_t *out_order,
                        uint8_t *out_n_levels, uint8_t *out_status);
int mcp_dag_levels_wide_device(const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges,
                               int64_t n_dags, uint8_t *out_level, uint8_t *out_order,
                               uint8_t *out_n_levels, uint8_t *out_status, void *stream);

/* Launches so far in this process of the kernel behind the <= 64-node pair: which = 0 the warp-per-DAG kernel, 1 the
 * thread-per-DAG kernel, 2 the pass over the DAGs it handed back (launched with it every time; it finds work only when some
 * DAG's edge list exceeded the thread kernel's pool).  For tests and benchmarks that must know which kernel answered.     */
int64_t mcp_dag_launch_count(int which);

/* ---- misc ------------------------------------------------------------------------------------ */
const char *mcp_last_error(void);
int mcp_device_count(void);          /* 0 when no CUDA device is usable */
const char *mcp_build_info(void);    /* arch, nvcc version, kernel list */

#ifdef __cplusplus
}
#endif
#endif /* MCP_B200_H */
