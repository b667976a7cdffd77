// mcp_abi.cu -- the C ABI (include/mcp_b200.h) and the host-side orchestration of the kernels.
//
//   K0 prep.cuh        row / query norms, fp16 shadow rows and rounding residuals, row upsert, penalty statistics
//   K1 scan_fp32.cuh   fp32 streaming distance + telemetry score            (small batches, tables without a shadow)
//   K2 gemm_tf32.cuh   tcgen05 / TMEM distance GEMM over the fp16 shadow (kind::f16) or the fp32 rows (kind::tf32),
//                      fused score + top-k' filter with a shared, tightening threshold
//   K3 select.cuh      radix top-k' selection / list reduction
//   K4 rescore.cuh     exact fp64 rescoring, final sort, certificate; exact_scan.cuh: the fp64 scan behind it
//   K5 merge.cuh       shard merge
//   K6 kahn.cuh        batched Kahn levels (<= 64 nodes: per warp, or per thread for large batches; 65..255 nodes per CTA)
//      multi.cuh       one handle over several GPUs (worker thread per shard)
//
// There is no CPU implementation of any of these in the library: without a device every entry
// point fails with MCP_ECUDA.
#include <cuda_runtime.h>
#include <fcntl.h>
#include <stdarg.h>
#include <unistd.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mcp_b200.h"

#define MCP_MAX_K_POW2 256
#define MCP_STR2(x) #x
#define MCP_STR(x) MCP_STR2(x)
#include "common.cuh"
#include "prep.cuh"
#include "scan_fp32.cuh"
#include "select.cuh"
#include "rescore.cuh"
#include "merge.cuh"
#include "kahn.cuh"
#include "gemm_tf32.cuh"
#define XS_SMEM_LIMIT (224u * 1024u)      // dynamic shared memory exact_scan2_kernel is launched with
#include "exact_scan.cuh"

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(e_ == cudaErrorMemoryAllocation ? MCP_ENOMEM : MCP_ECUDA, "%s:%d %s -> %s", \
                        __FILE__, __LINE__, #call, cudaGetErrorString(e_));                        \
    } while (0)
#define KCHECK() CU(cudaGetLastError())
#define KL() do { idx->stats.kernel_launches++; KCHECK(); } while (0)

extern "C" const char *mcp_last_error(void) { return g_err.c_str(); }

// launch as a programmatic dependent of the previous kernel in the stream (common.cuh: pdl_wait / pdl_trigger)
template <class... KArgs, class... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

extern "C" int mcp_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" const char *mcp_build_info(void) {
    return "mcp_b200 sm_100a nvcc " MCP_STR(__CUDACC_VER_MAJOR__) "." MCP_STR(__CUDACC_VER_MINOR__)
           " kernels: prep_rows prep_queries scan_fp32 gemm_tf32_tcgen05 (cta_group::1 / cta_group::2) select_from_scores"
           " select_from_lists rescore_finalize exact_scan2 exact_pick exact_select shadow_rows merge_shards kahn_levels kahn_fifo";
}

// ---------------------------------------------------------------------------------------------
// device buffer that only grows
// ---------------------------------------------------------------------------------------------
struct Buf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes) {
        if (bytes <= cap) return MCP_OK;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        CU(cudaMalloc(&p, want));
        cap = want;
        return MCP_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

// Every byte of [p, p + bytes) must be plain device memory that lives on `device`: the kernels of an index -- and the
// TMA tensor maps K2 builds over the table and the queries -- only ever read this GPU's memory.  (TMA multicast loads here
// read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.)  Peer
// allocations, IPC imports and virtual ranges whose mappings reach another GPU report that GPU as their device; mappings
// are made of >= 2 MiB granules, so one probe per 2 MiB boundary sees every mapping a range is stitched from.
static int require_local(const void *p, size_t bytes, int device, const char *what) {
    if (!p || !bytes) return MCP_OK;
    const uintptr_t lo = (uintptr_t)p, hi = lo + bytes - 1, step = (uintptr_t)2 << 20;
    for (uintptr_t a = lo;;) {
        cudaPointerAttributes at;
        cudaError_t e = cudaPointerGetAttributes(&at, (const void *)a);
        if (e != cudaSuccess) { cudaGetLastError(); return fail(MCP_EINVAL, "%s: cudaPointerGetAttributes(%p) -> %s", what, (const void *)a, cudaGetErrorString(e)); }
        if (at.type != cudaMemoryTypeDevice || at.device != device)
            return fail(MCP_EINVAL, "%s: byte %llu of the buffer is not device memory of GPU %d (memory type %d, device %d); "
                        "an index reads only its own GPU's memory -- copy remote shards in, or give each GPU its own index",
                        what, (unsigned long long)(a - lo), device, (int)at.type, at.device);
        if (a == hi) break;
        const uintptr_t next = (a & ~(step - 1)) + step;
        a = next > hi ? hi : next;
    }
    return MCP_OK;
}

struct mcp_multi;
struct mcp_index {
    mcp_multi *multi = nullptr;   // set on a handle made by mcp_index_create_multi: the table is row-sharded over sub-indexes
    int device = 0, dim = 0, metric = 0, n_sm = 148;
    int64_t n_rows = 0, ld = 0, id_base = 0;
    const float *d_rows = nullptr;
    Buf own_rows, norm2, inv_norm, penalty, consts;
    Buf shadow;                 // fp16 image of the table (rows scaled by powers of two), pitch ldh elements: K2's fp16 operand; empty when option "shadow" is 0
    Buf shadow_scale;           // [n_rows] fp32: 1 / scale of each shadow row
    int64_t ldh = 0;
    bool has_shadow = false;
    int opt_shadow = 1;
    int kp_boost = 0, clean_looks = 0;   // automatic oversampling: k' = base << kp_boost (note_outcome)
    int64_t q_since_look = 0;
    bool has_penalty = false;
    TableConsts h_consts{};
    // options
    int opt_path = MCP_PATH_AUTO, opt_kprime = 0, opt_max_fallback = XS_MAXS, opt_time_coarse = 0;
    // workspace
    Buf q_copy, nq2, inv_nq, qflag, scores, listsA, listsB, cntA, cntB, flags,
        exact_all, out_ids, gemm_ws, q_pad, nq2f, drop_min, rho_q, q_op, q_scale;
    int *h_status = nullptr;   // pinned + mapped: [0] certificate failures, [1] unresolved; written by the fallback kernel's status fold
    int *h_status_dev = nullptr;   // the same words as the device addresses them
    Buf sp_keys, sp_rows, sp_cnt, bound_key, bound_row;   // the fallback's sparse lists [slot][XS_SPARSE_CAP] and per-query bounds
    Buf unres;                 // [0] count, [1..] queries whose sparse list overflowed (host entry point re-answers them)
    unsigned char *h_stage = nullptr;   // pinned staging for the host-buffer search: results + status in one copy
    size_t h_stage_cap = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    bool pending = false;
    cudaStream_t last_stream = nullptr;
    // row upserts are enqueued on `stream`, after whatever search is in flight and ahead of the next one
    unsigned char *h_up = nullptr;   // pinned staging: rows, then ids
    size_t h_up_cap = 0;
    Buf d_up;                        // the same on the device
    cudaEvent_t up_copied = nullptr; // the staging buffer has been read (it may be refilled)
    cudaEvent_t up_done = nullptr;   // the upsert kernel has run: searches on other streams wait for it on the device
    bool up_pending = false;
    std::vector<cudaEvent_t> ev;   // pairs bracketing coarse scans when opt_time_coarse
    int ev_used = 0;
    GemmPlan gemm;
    std::mutex mu;
    mcp_stats stats{};
};

// multi-device handles (multi.cuh, included at the end of this file)
static int multi_destroy(mcp_index *idx);
static int multi_load(mcp_index *idx, const float *rows_host, int64_t n_rows, int64_t id_base);
static int multi_set_penalty(mcp_index *idx, const float *penalty_host);
static int multi_update_rows(mcp_index *idx, int64_t first_row, const float *rows_host, int64_t n_rows);
static int multi_upsert_rows(mcp_index *idx, const int64_t *row_ids, const float *rows_host, int64_t n_rows);
static int multi_load_file(mcp_index *idx, const char *path, int64_t first_row, int64_t n_rows);
static int multi_set_option(mcp_index *idx, const char *name, int64_t value);
static int multi_get_stats(mcp_index *idx, mcp_stats *out);
static int multi_search(mcp_index *idx, const float *queries_host, int nq, int k, int32_t *out_ids, float *out_scores, double *out_scores64);
#define NOT_ON_MULTI(idx, what)                                                                                                  \
    do { if ((idx)->multi) return fail(MCP_ESTATE, what " is not available on a multi-device handle (device pointers belong to one GPU): " \
                                       "use the host-buffer entry points, or one index per GPU"); } while (0)

#define NOALLOC_MAX_DEVICES_ 64
#define SCORE_BUDGET_BYTES (512ull << 20)
#define UNRES_MAX 4096

static int check_idx(const mcp_index *idx) { return idx ? MCP_OK : fail(MCP_EINVAL, "index handle is NULL"); }

extern "C" int mcp_index_create(mcp_index **out, int device_id, int dim, int metric) {
    if (!out) return fail(MCP_EINVAL, "out is NULL");
    if (dim < 1 || dim > 8192) return fail(MCP_EINVAL, "dim %d outside [1, 8192]", dim);
    if (metric != MCP_METRIC_COSINE && metric != MCP_METRIC_L2) return fail(MCP_EINVAL, "unknown metric %d", metric);
    int n = 0;
    CU(cudaGetDeviceCount(&n));
    if (device_id < 0 || device_id >= n) return fail(MCP_EINVAL, "device %d of %d", device_id, n);
    CU(cudaSetDevice(device_id));
    mcp_index *idx = new mcp_index();
    idx->device = device_id; idx->dim = dim; idx->metric = metric;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device_id));
    idx->n_sm = prop.multiProcessorCount;
    if (prop.major != 10) { delete idx; return fail(MCP_ECUDA, "device %d is sm_%d%d; this library holds sm_100a code only", device_id, prop.major, prop.minor); }
    CU(cudaStreamCreateWithFlags(&idx->stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&idx->done, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&idx->up_copied, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&idx->up_done, cudaEventDisableTiming));
    CU(cudaHostAlloc(&idx->h_status, 4 * sizeof(int), cudaHostAllocMapped));
    idx->h_status[0] = idx->h_status[1] = 0;
    CU(cudaHostGetDevicePointer(&idx->h_status_dev, idx->h_status, 0));
    CU(cudaFuncSetAttribute(rescore_finalize_kernel<5, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RF_SMEM_LIMIT));
    CU(cudaFuncSetAttribute(rescore_finalize_kernel<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RF_SMEM_LIMIT));
    CU(cudaFuncSetAttribute(exact_scan2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)XS_SMEM_LIMIT));
    int rc = idx->consts.ensure(sizeof(TableConsts));
    if (!rc) rc = idx->flags.ensure(sizeof(int) * 4096);
    if (!rc) rc = idx->unres.ensure(sizeof(int) * (1 + UNRES_MAX));
    if (rc) { delete idx; return rc; }
    CU(cudaMemset(idx->flags.p, 0, sizeof(int) * 8));
    CU(cudaMemset(idx->unres.p, 0, sizeof(int)));
    *out = idx;
    return MCP_OK;
}

extern "C" int mcp_index_destroy(mcp_index *idx) {
    if (!idx) return MCP_OK;
    if (idx->multi) return multi_destroy(idx);
    cudaSetDevice(idx->device);
    cudaDeviceSynchronize();
    Buf *bufs[] = {&idx->own_rows, &idx->norm2, &idx->inv_norm, &idx->penalty, &idx->consts, &idx->q_copy,
                   &idx->nq2, &idx->inv_nq, &idx->qflag, &idx->scores, &idx->listsA, &idx->listsB, &idx->cntA,
                   &idx->cntB, &idx->flags, &idx->exact_all, &idx->out_ids,
                   &idx->gemm_ws, &idx->q_pad, &idx->nq2f, &idx->drop_min, &idx->sp_keys, &idx->sp_rows, &idx->sp_cnt, &idx->bound_key, &idx->bound_row, &idx->unres, &idx->shadow, &idx->shadow_scale, &idx->rho_q, &idx->q_op, &idx->q_scale};
    for (Buf *b : bufs) b->release();
    for (cudaEvent_t e : idx->ev) cudaEventDestroy(e);
    if (idx->h_status) cudaFreeHost(idx->h_status);
    if (idx->h_stage) cudaFreeHost(idx->h_stage);
    if (idx->h_up) cudaFreeHost(idx->h_up);
    idx->d_up.release();
    if (idx->up_copied) cudaEventDestroy(idx->up_copied);
    if (idx->up_done) cudaEventDestroy(idx->up_done);
    if (idx->done) cudaEventDestroy(idx->done);
    if (idx->stream) cudaStreamDestroy(idx->stream);
    delete idx;
    return MCP_OK;
}

extern "C" int64_t mcp_index_rows(const mcp_index *idx) { return idx ? idx->n_rows : -1; }

// K0 over the adopted / loaded rows; synchronous (load time).
static int prepare_table(mcp_index *idx, cudaStream_t st) {
    int rc;
    idx->gemm.valid = false;
    // a cluster multicasts its row tiles only over a table this library allocated on this device itself
    idx->gemm.mc_ok = idx->d_rows == idx->own_rows.as<float>();
    if ((rc = idx->norm2.ensure(sizeof(double) * (size_t)idx->n_rows))) return rc;
    if ((rc = idx->inv_norm.ensure(sizeof(float) * (size_t)idx->n_rows))) return rc;
    CU(cudaMemsetAsync(idx->consts.p, 0, sizeof(TableConsts), st));
    unsigned blocks = (unsigned)((idx->n_rows + PREP_ROWS - 1) / PREP_ROWS);
    prep_rows_kernel<false><<<blocks, PREP_ROWS, 0, st>>>(idx->d_rows, idx->n_rows, idx->dim, idx->ld,
                                                          idx->norm2.as<double>(), idx->inv_norm.as<float>(),
                                                          nullptr, idx->consts.as<TableConsts>());
    KCHECK();
    // fp16 shadow (K2's fp16 operand: half the bytes, twice the tensor rate) + the rounding residuals the certificate needs.
    // A table too large for a shadow beside it still works: the tf32 scan reads the fp32 rows themselves.
    idx->ldh = (idx->dim + 7) & ~7;
    idx->has_shadow = false;
    if (idx->opt_shadow) {
        rc = idx->shadow.ensure(sizeof(__half) * (size_t)idx->n_rows * (size_t)idx->ldh);
        if (!rc) rc = idx->shadow_scale.ensure(sizeof(float) * (size_t)idx->n_rows);
        if (rc == MCP_ENOMEM) { cudaGetLastError(); idx->shadow.release(); idx->shadow_scale.release(); }
        else if (rc) return rc;
        else idx->has_shadow = true;
    } else { idx->shadow.release(); idx->shadow_scale.release(); }
    shadow_rows_kernel<<<(unsigned)((idx->n_rows + 7) / 8), 256, 0, st>>>(idx->d_rows, idx->n_rows, idx->dim, idx->ld, idx->norm2.as<double>(),
                                                                     idx->has_shadow ? idx->shadow.as<__half>() : nullptr, idx->ldh,
                                                                     idx->has_shadow ? idx->shadow_scale.as<float>() : nullptr,
                                                                     idx->consts.as<TableConsts>());
    KCHECK();
    CU(cudaMemcpyAsync(&idx->h_consts, idx->consts.p, sizeof(TableConsts), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    idx->has_penalty = false;
    if (idx->h_consts.bad_rows)
        return fail(MCP_EINVAL, "%u rows hold NaN/Inf elements (pgvector rejects these at input)", idx->h_consts.bad_rows);
    return MCP_OK;
}

extern "C" int mcp_index_load(mcp_index *idx, const float *rows_host, int64_t n_rows, int64_t id_base) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (idx->multi) return multi_load(idx, rows_host, n_rows, id_base);
    if (!rows_host || n_rows < 1) return fail(MCP_EINVAL, "rows NULL or n_rows < 1");
    if (n_rows >= (1ll << 31) || id_base < 0 || id_base + n_rows > (1ll << 31)) return fail(MCP_EINVAL, "ids must fit int32");
    std::lock_guard<std::mutex> lk(idx->mu);
    CU(cudaSetDevice(idx->device));
    if (idx->pending) { CU(cudaEventSynchronize(idx->done)); }      // searches in flight read the old table and its norms
    int64_t ld = (idx->dim + 3) & ~3;                       // 16-byte rows; pad columns are zero
    int rc = idx->own_rows.ensure(sizeof(float) * (size_t)n_rows * (size_t)ld);
    if (rc) return rc;
    if (ld != idx->dim) CU(cudaMemsetAsync(idx->own_rows.p, 0, sizeof(float) * (size_t)n_rows * (size_t)ld, idx->stream));
    CU(cudaMemcpy2DAsync(idx->own_rows.p, sizeof(float) * ld, rows_host, sizeof(float) * idx->dim,
                         sizeof(float) * idx->dim, (size_t)n_rows, cudaMemcpyHostToDevice, idx->stream));
    idx->d_rows = idx->own_rows.as<float>();
    idx->n_rows = n_rows; idx->ld = ld; idx->id_base = id_base;
    return prepare_table(idx, idx->stream);
}

// ---- table file -> HBM at storage speed (stands where control_plane.py:54 fetches the rows) ----------
// File: 64-byte header (int64 x 8: magic "MCPTBL01", dim, n_rows, id_base, metric, 0, 0, 0) + n_rows x dim fp32, row-major.
// Rows [first_row, first_row + n_rows) are read (a rank of a sharded registry reads only its own range) by two reader threads
// into a ring of pinned chunks while the previous chunks are on their way to the device: one cudaMemcpyAsync per chunk, the
// reads, the copies and the page-cache / NVMe traffic overlap.  K0 (norms, shadow) then runs over the resident rows as after
// mcp_index_load; it costs 0.6 ms per GB, which is why there is no norms sidecar to skip it.
#define LF_CHUNK_BYTES ((size_t)32 << 20)
#define LF_NBUF 4
struct TableFileHeader { int64_t magic, dim, n_rows, id_base, metric, r0, r1, r2; };

static int load_file_locked(mcp_index *idx, const char *path, int64_t first_row, int64_t n_rows) {
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(MCP_EINVAL, "cannot open %s", path);
    TableFileHeader h;
    int64_t magic;
    memcpy(&magic, "MCPTBL01", 8);
    if (pread(fd, &h, sizeof h, 0) != (ssize_t)sizeof h || h.magic != magic) { close(fd); return fail(MCP_EINVAL, "%s is not an MCPTBL01 table file", path); }
    if (h.dim != idx->dim) { close(fd); return fail(MCP_EINVAL, "%s holds dim %lld, the index has dim %d", path, (long long)h.dim, idx->dim); }
    if (n_rows < 0) n_rows = h.n_rows - first_row;
    if (first_row < 0 || n_rows < 1 || first_row + n_rows > h.n_rows) { close(fd); return fail(MCP_EINVAL, "rows [%lld, %lld) outside the file's %lld rows", (long long)first_row, (long long)(first_row + n_rows), (long long)h.n_rows); }
    const int64_t id_base = h.id_base + first_row;
    if (id_base < 0 || id_base + n_rows > (1ll << 31)) { close(fd); return fail(MCP_EINVAL, "ids must fit int32"); }
    cudaError_t ce = cudaSetDevice(idx->device);
    if (ce != cudaSuccess) { close(fd); return fail(MCP_ECUDA, "cudaSetDevice: %s", cudaGetErrorString(ce)); }
    if (idx->pending) { cudaEventSynchronize(idx->done); }
    const int dim = idx->dim;
    const int64_t ld = (dim + 3) & ~3;
    int rc = idx->own_rows.ensure(sizeof(float) * (size_t)n_rows * (size_t)ld);
    if (rc) { close(fd); return rc; }
    cudaStream_t st = idx->stream;
    if (ld != dim) cudaMemsetAsync(idx->own_rows.p, 0, sizeof(float) * (size_t)n_rows * (size_t)ld, st);
    const size_t row_bytes = sizeof(float) * (size_t)dim;
    const int64_t rows_per_chunk = std::max<int64_t>(1, (int64_t)(LF_CHUNK_BYTES / row_bytes));
    const int64_t n_chunks = (n_rows + rows_per_chunk - 1) / rows_per_chunk;
    unsigned char *buf[LF_NBUF] = {nullptr};
    cudaEvent_t freed[LF_NBUF] = {nullptr};
    for (int b = 0; b < LF_NBUF && ce == cudaSuccess; ++b) {
        ce = cudaMallocHost(&buf[b], (size_t)rows_per_chunk * row_bytes);
        if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&freed[b], cudaEventDisableTiming);
    }
    std::mutex mu;
    std::condition_variable cv;
    std::vector<int> state(n_chunks, 0);             // 0 not read, 1 read into its buffer, 2 copy issued
    int64_t copied = 0;                              // chunks whose copy has been issued (buffer c % LF_NBUF is reusable once freed[] fires)
    bool failed = ce != cudaSuccess;
    std::string io_err;
    auto reader = [&](int tid) {
        for (int64_t c = tid; c < n_chunks; c += 2) {
            {   // the buffer's previous tenant (chunk c - LF_NBUF) must have been handed to the copy engine
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return failed || copied > c - LF_NBUF; });
                if (failed) return;
            }
            if (c >= LF_NBUF) cudaEventSynchronize(freed[c % LF_NBUF]);      // ... and the copy must have read it
            const int64_t r0 = c * rows_per_chunk, nr = std::min(rows_per_chunk, n_rows - r0);
            size_t want = (size_t)nr * row_bytes, got = 0;
            const off_t off = (off_t)(sizeof h + (size_t)(first_row + r0) * row_bytes);
            while (got < want) {
                const ssize_t k = pread(fd, buf[c % LF_NBUF] + got, want - got, off + (off_t)got);
                if (k <= 0) break;
                got += (size_t)k;
            }
            std::lock_guard<std::mutex> lk(mu);
            if (got != want) { failed = true; io_err = "short read"; }
            state[c] = 1;
            cv.notify_all();
        }
    };
    std::thread t0, t1;
    if (!failed) { t0 = std::thread(reader, 0); t1 = std::thread(reader, 1); }
    for (int64_t c = 0; c < n_chunks && !failed; ++c) {
        {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return failed || state[c] == 1; });
            if (failed) break;
        }
        const int64_t r0 = c * rows_per_chunk, nr = std::min(rows_per_chunk, n_rows - r0);
        float *dst = idx->own_rows.as<float>() + r0 * ld;
        ce = (ld == dim) ? cudaMemcpyAsync(dst, buf[c % LF_NBUF], (size_t)nr * row_bytes, cudaMemcpyHostToDevice, st)
                         : cudaMemcpy2DAsync(dst, sizeof(float) * ld, buf[c % LF_NBUF], row_bytes, row_bytes, (size_t)nr, cudaMemcpyHostToDevice, st);
        if (ce == cudaSuccess) ce = cudaEventRecord(freed[c % LF_NBUF], st);
        std::lock_guard<std::mutex> lk(mu);
        if (ce != cudaSuccess) failed = true;
        state[c] = 2; copied = c + 1;
        cv.notify_all();
    }
    { std::lock_guard<std::mutex> lk(mu); if (ce != cudaSuccess) failed = true; }
    cv.notify_all();
    if (t0.joinable()) t0.join();
    if (t1.joinable()) t1.join();
    cudaStreamSynchronize(st);
    for (int b = 0; b < LF_NBUF; ++b) { if (buf[b]) cudaFreeHost(buf[b]); if (freed[b]) cudaEventDestroy(freed[b]); }
    close(fd);
    if (failed) {
        idx->n_rows = 0;
        if (ce != cudaSuccess) return fail(ce == cudaErrorMemoryAllocation ? MCP_ENOMEM : MCP_ECUDA, "load_file: %s", cudaGetErrorString(ce));
        return fail(MCP_EINVAL, "load_file: %s reading %s", io_err.c_str(), path);
    }
    idx->d_rows = idx->own_rows.as<float>();
    idx->n_rows = n_rows; idx->ld = ld; idx->id_base = id_base;
    return prepare_table(idx, st);
}

extern "C" int mcp_index_load_file(mcp_index *idx, const char *path, int64_t first_row, int64_t n_rows) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (!path) return fail(MCP_EINVAL, "path is NULL");
    if (idx->multi) return multi_load_file(idx, path, first_row, n_rows);
    std::lock_guard<std::mutex> lk(idx->mu);
    return load_file_locked(idx, path, first_row, n_rows);
}

extern "C" int mcp_index_adopt_device(mcp_index *idx, const float *rows_device, int64_t n_rows, int64_t ld,
                                      int64_t id_base, void *stream) {
    if (check_idx(idx)) return MCP_EINVAL;
    NOT_ON_MULTI(idx, "adopt_device");
    if (!rows_device || n_rows < 1) return fail(MCP_EINVAL, "rows NULL or n_rows < 1");
    if (ld < idx->dim || (ld & 3) || (idx->dim & 3)) return fail(MCP_EINVAL, "adopt needs dim %% 4 == 0 and ld %% 4 == 0, ld >= dim");
    if (((uintptr_t)rows_device) & 15) return fail(MCP_EINVAL, "rows must be 16-byte aligned");
    if (n_rows >= (1ll << 31) || id_base < 0 || id_base + n_rows > (1ll << 31)) return fail(MCP_EINVAL, "ids must fit int32");
    std::lock_guard<std::mutex> lk(idx->mu);
    CU(cudaSetDevice(idx->device));
    int lrc = require_local(rows_device, sizeof(float) * ((size_t)(n_rows - 1) * (size_t)ld + (size_t)idx->dim), idx->device, "adopt_device rows");
    if (lrc) return lrc;
    if (idx->pending) { CU(cudaEventSynchronize(idx->done)); }      // searches in flight read the old table and its norms
    idx->own_rows.release();
    idx->d_rows = rows_device; idx->n_rows = n_rows; idx->ld = ld; idx->id_base = id_base;
    return prepare_table(idx, (cudaStream_t)stream);
}

static int upsert_locked(mcp_index *idx, const int64_t *row_ids, int64_t first_row, const float *rows_host, int64_t n);

// Registry refresh without a reload: rows [first_row, first_row + n_rows) changed.
extern "C" int mcp_index_update_rows(mcp_index *idx, int64_t first_row, const float *rows_host, int64_t n_rows) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (idx->multi) return multi_update_rows(idx, first_row, rows_host, n_rows);
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->n_rows) return fail(MCP_ESTATE, "update_rows before load");
    if (first_row < 0 || n_rows < 1 || first_row + n_rows > idx->n_rows) return fail(MCP_EINVAL, "row range [%lld, %lld) outside the table", (long long)first_row, (long long)(first_row + n_rows));
    if (rows_host && idx->d_rows == idx->own_rows.as<float>()) return upsert_locked(idx, nullptr, first_row, rows_host, n_rows);   // stream-ordered, no host wait
    CU(cudaSetDevice(idx->device));
    if (idx->pending) { CU(cudaEventSynchronize(idx->done)); }      // searches in flight read the rows
    cudaStream_t st = idx->stream;
    idx->last_stream = st;
    if (rows_host) {
        if (idx->d_rows != idx->own_rows.as<float>()) return fail(MCP_ESTATE, "the table is adopted device memory: change it in place and pass rows_host = NULL");
        CU(cudaMemcpy2DAsync(idx->own_rows.as<float>() + first_row * idx->ld, sizeof(float) * idx->ld, rows_host, sizeof(float) * idx->dim,
                             sizeof(float) * idx->dim, (size_t)n_rows, cudaMemcpyHostToDevice, st));
    }
    // norms of the changed rows; bad_rows is counted afresh, the maximum norm only ever grows (a larger bound stays valid)
    TableConsts *c = idx->consts.as<TableConsts>();
    CU(cudaMemsetAsync(&c->bad_rows, 0, sizeof(unsigned int), st));
    unsigned blocks = (unsigned)((n_rows + PREP_ROWS - 1) / PREP_ROWS);
    prep_rows_kernel<false><<<blocks, PREP_ROWS, 0, st>>>(idx->d_rows + first_row * idx->ld, n_rows, idx->dim, idx->ld,
                                                          idx->norm2.as<double>() + first_row, idx->inv_norm.as<float>() + first_row,
                                                          nullptr, c);
    KCHECK();
    // shadow rows and residual maxima of the changed rows (the maxima, like the maximum norm, only ever grow)
    shadow_rows_kernel<<<(unsigned)((n_rows + 7) / 8), 256, 0, st>>>(idx->d_rows + first_row * idx->ld, n_rows, idx->dim, idx->ld,
                                                                idx->norm2.as<double>() + first_row,
                                                                idx->has_shadow ? idx->shadow.as<__half>() + first_row * idx->ldh : nullptr,
                                                                idx->ldh, idx->has_shadow ? idx->shadow_scale.as<float>() + first_row : nullptr, c);
    KCHECK();
    CU(cudaMemcpyAsync(&idx->h_consts, c, sizeof(TableConsts), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (idx->h_consts.bad_rows) return fail(MCP_EINVAL, "%u updated rows hold NaN/Inf elements; the table now contains them -- update again", idx->h_consts.bad_rows);
    return MCP_OK;
}

// Registry ingest that does not stop queries.  Rows are checked on the host (non-finite values,
// norms the coarse kernels cannot bound), staged in pinned memory and handed to ONE kernel on the index's own stream: copy into
// the table, exact norms, shadow rows, residual maxima.  Nothing is read back, so the call returns as soon as the work is
// enqueued; searches already in flight finish on the old rows, later ones see the new rows (a search on another stream is made
// to wait on the device, not on the host).
static int upsert_locked(mcp_index *idx, const int64_t *row_ids, int64_t first_row, const float *rows_host, int64_t n) {
    if (!idx->n_rows) return fail(MCP_ESTATE, "upsert before load");
    if (!rows_host || n < 1) return fail(MCP_EINVAL, "rows NULL or n < 1");
    if (idx->d_rows != idx->own_rows.as<float>()) return fail(MCP_ESTATE, "the table is adopted device memory: change it in place and call mcp_index_update_rows with rows_host = NULL");
    const int dim = idx->dim;
    unsigned int odd = 0;
    for (int64_t i = 0; i < n; ++i) {
        const int64_t r = row_ids ? row_ids[i] : first_row + i;
        if (r < 0 || r >= idx->n_rows) return fail(MCP_EINVAL, "row %lld outside the table of %lld rows", (long long)r, (long long)idx->n_rows);
        const float *v = rows_host + i * dim;
        double n2 = 0.0;
        for (int c = 0; c < dim; ++c) n2 += (double)v[c] * (double)v[c];
        if (!std::isfinite(n2)) return fail(MCP_EINVAL, "row %lld holds NaN/Inf elements (pgvector rejects these at input); nothing was changed", (long long)i);
        if (n2 != 0.0 && (n2 < 1e-30 || n2 > 1e30)) ++odd;
    }
    CU(cudaSetDevice(idx->device));
    const size_t row_bytes = sizeof(float) * (size_t)n * dim, id_bytes = row_ids ? sizeof(int64_t) * (size_t)n : 0;
    const size_t ids_off = (row_bytes + 15) & ~(size_t)15, total = ids_off + id_bytes;
    if (total > idx->h_up_cap) {
        if (idx->up_pending) CU(cudaEventSynchronize(idx->up_copied));
        if (idx->h_up) cudaFreeHost(idx->h_up);
        idx->h_up = nullptr; idx->h_up_cap = 0;
        CU(cudaMallocHost(&idx->h_up, total * 2));
        idx->h_up_cap = total * 2;
    } else if (idx->up_pending) CU(cudaEventSynchronize(idx->up_copied));   // the previous upsert's copy has left the staging buffer
    int rc = idx->d_up.ensure(total);
    if (rc) return rc;
    memcpy(idx->h_up, rows_host, row_bytes);
    if (row_ids) memcpy(idx->h_up + ids_off, row_ids, id_bytes);
    cudaStream_t st = idx->stream;
    if (idx->pending && idx->last_stream != st) CU(cudaStreamWaitEvent(st, idx->done, 0));   // behind the search in flight, on the device
    CU(cudaMemcpyAsync(idx->d_up.p, idx->h_up, total, cudaMemcpyHostToDevice, st));
    CU(cudaEventRecord(idx->up_copied, st));
    upsert_rows_kernel<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(idx->d_up.as<float>(), dim, row_ids ? reinterpret_cast<const int64_t *>(idx->d_up.as<unsigned char>() + ids_off) : nullptr,
                                                               first_row, n, idx->own_rows.as<float>(), dim, idx->ld, idx->norm2.as<double>(), idx->inv_norm.as<float>(),
                                                               idx->has_shadow ? idx->shadow.as<__half>() : nullptr, idx->ldh,
                                                               idx->has_shadow ? idx->shadow_scale.as<float>() : nullptr, idx->consts.as<TableConsts>());
    KCHECK();
    CU(cudaEventRecord(idx->up_done, st));
    idx->up_pending = true;
    idx->h_consts.odd_rows += odd;                   // rows of extreme norm send every later search to the fp64 scan (choose_path)
    return MCP_OK;
}

extern "C" int mcp_index_upsert_rows(mcp_index *idx, const int64_t *row_ids, const float *rows_host, int64_t n_rows) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (!row_ids) return fail(MCP_EINVAL, "row_ids is NULL");
    if (idx->multi) return multi_upsert_rows(idx, row_ids, rows_host, n_rows);
    std::lock_guard<std::mutex> lk(idx->mu);
    return upsert_locked(idx, row_ids, 0, rows_host, n_rows);
}

static int penalty_stats(mcp_index *idx, cudaStream_t st) {
    TableConsts *c = idx->consts.as<TableConsts>();
    CU(cudaMemsetAsync(&c->pen_max, 0, 3 * sizeof(unsigned int), st));   // pen_max, pen_absmax, bad_pen
    penalty_stats_kernel<<<idx->n_sm * 4, 256, 0, st>>>(idx->penalty.as<float>(), idx->n_rows, c);
    KCHECK();
    penalty_stats_finish_kernel<<<1, 1, 0, st>>>(c);
    KCHECK();
    CU(cudaMemcpyAsync(&idx->h_consts, c, sizeof(TableConsts), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (idx->h_consts.bad_pen) {
        idx->has_penalty = false;
        return fail(MCP_EINVAL, "penalty holds %u NaN/Inf values", idx->h_consts.bad_pen);
    }
    idx->has_penalty = true;
    return MCP_OK;
}

static int clear_penalty(mcp_index *idx, cudaStream_t st) {
    idx->has_penalty = false;
    TableConsts *c = idx->consts.as<TableConsts>();
    CU(cudaMemsetAsync(&c->pen_max, 0, 3 * sizeof(unsigned int), st));
    CU(cudaStreamSynchronize(st));
    idx->h_consts.pen_max = idx->h_consts.pen_absmax = 0.f;
    return MCP_OK;
}

extern "C" int mcp_index_set_penalty(mcp_index *idx, const float *penalty_host) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (idx->multi) return multi_set_penalty(idx, penalty_host);
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->n_rows) return fail(MCP_ESTATE, "set_penalty before load");
    CU(cudaSetDevice(idx->device));
    if (idx->pending) { CU(cudaEventSynchronize(idx->done)); }
    idx->last_stream = idx->stream;
    if (!penalty_host) return clear_penalty(idx, idx->stream);
    int rc = idx->penalty.ensure(sizeof(float) * (size_t)idx->n_rows);
    if (rc) return rc;
    CU(cudaMemcpyAsync(idx->penalty.p, penalty_host, sizeof(float) * (size_t)idx->n_rows, cudaMemcpyHostToDevice, idx->stream));
    return penalty_stats(idx, idx->stream);
}

extern "C" int mcp_index_set_penalty_device(mcp_index *idx, const float *penalty_device, void *stream) {
    if (check_idx(idx)) return MCP_EINVAL;
    NOT_ON_MULTI(idx, "set_penalty_device");
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->n_rows) return fail(MCP_ESTATE, "set_penalty before load");
    CU(cudaSetDevice(idx->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (idx->pending && idx->last_stream != st) { CU(cudaEventSynchronize(idx->done)); }
    if (!penalty_device) return clear_penalty(idx, st);
    int rc = require_local(penalty_device, sizeof(float) * (size_t)idx->n_rows, idx->device, "set_penalty_device");
    if (rc) return rc;
    rc = idx->penalty.ensure(sizeof(float) * (size_t)idx->n_rows);
    if (rc) return rc;
    CU(cudaMemcpyAsync(idx->penalty.p, penalty_device, sizeof(float) * (size_t)idx->n_rows, cudaMemcpyDeviceToDevice, st));
    return penalty_stats(idx, st);
}

extern "C" int mcp_index_set_option(mcp_index *idx, const char *name, int64_t value) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (!name) return fail(MCP_EINVAL, "option name is NULL");
    if (idx->multi) return multi_set_option(idx, name, value);
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!strcmp(name, "path")) {
        if (value < MCP_PATH_AUTO || value > MCP_PATH_GEMM_F16) return fail(MCP_EINVAL, "path %lld", (long long)value);
        idx->opt_path = (int)value;
    } else if (!strcmp(name, "kprime")) {
        if (value < 0 || value > FINAL_MAX) return fail(MCP_EINVAL, "kprime %lld outside [0, %d]", (long long)value, FINAL_MAX);
        idx->opt_kprime = (int)value;
    } else if (!strcmp(name, "max_fallback")) {
        if (value < 1 || value > XS_MAXS) return fail(MCP_EINVAL, "max_fallback %lld outside [1, %d]", (long long)value, XS_MAXS);
        idx->opt_max_fallback = (int)value;
    } else if (!strcmp(name, "gemm_balance")) {
        idx->gemm.balance_tiles = value != 0;
    } else if (!strcmp(name, "gemm_cluster")) {
        if (value != 0 && value != 1 && value != 2 && value != 4) return fail(MCP_EINVAL, "gemm_cluster must be 0, 1, 2 or 4");
        idx->gemm.cluster_opt = (int)value;
    } else if (!strcmp(name, "gemm_sync")) {
        if (value < 0 || value > 64) return fail(MCP_EINVAL, "gemm_sync must be in [0, 64]");
        idx->gemm.sync_tiles = (int)value;
    } else if (!strcmp(name, "gemm_coop")) {
        // Without the cooperative attribute nothing guarantees that the CTAs K2's two spin-waits depend on are co-resident: on a
        // busy GPU the kernel can hang.  It exists so that Nsight Compute can replay the launch, and is honoured only together
        // with the profiling bit of gemm_debug.
        if (value == 0 && !(idx->gemm.debug & 4096))
            return fail(MCP_EINVAL, "gemm_coop = 0 is for profiler captures only (co-residency is no longer guaranteed: K2 may hang on a busy GPU); set gemm_debug = 4096 first");
        idx->gemm.coop_attr = value != 0;
    } else if (!strcmp(name, "gemm_pair")) {
        idx->gemm.allow_pair = value != 0;
    } else if (!strcmp(name, "gemm_debug")) {
        idx->gemm.debug = (int)value;
    } else if (!strcmp(name, "gemm_l2_prefetch")) {
        idx->gemm.l2_prefetch = value != 0;
    } else if (!strcmp(name, "shadow")) {
        if (idx->n_rows) return fail(MCP_ESTATE, "option shadow must be set before the table is loaded");
        idx->opt_shadow = value != 0;
    } else if (!strcmp(name, "gemm_bootstrap")) {
        idx->gemm.allow_boot = value != 0;
    } else if (!strcmp(name, "time_coarse")) {
        idx->opt_time_coarse = value != 0;
    } else
        return fail(MCP_EINVAL, "unknown option '%s'", name);
    return MCP_OK;
}

extern "C" int mcp_index_get_stats(mcp_index *idx, mcp_stats *out) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (!out) return fail(MCP_EINVAL, "out is NULL");
    if (idx->multi) return multi_get_stats(idx, out);
    std::lock_guard<std::mutex> lk(idx->mu);
    *out = idx->stats;
    return MCP_OK;
}

// ---------------------------------------------------------------------------------------------
// search
// ---------------------------------------------------------------------------------------------
__global__ void flag_all_kernel(int *flag_list, int *n_flag, int q0, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag_list[i] = q0 + i;
    if (i == 0) *n_flag = n;
}
__global__ void clear_flag_kernel(int *n_flag) { *n_flag = 0; }
__global__ void pad_queries_kernel(const float *q, int dim, int dimp, float *out) {
    for (int i = threadIdx.x; i < dimp; i += blockDim.x)
        out[(int64_t)blockIdx.x * dimp + i] = i < dim ? q[(int64_t)blockIdx.x * dim + i] : 0.f;
}

// add finished coarse-scan event pairs to the stats (blocks on the last one)
static int harvest_events(mcp_index *idx) {
    for (int i = 0; i + 1 < idx->ev_used; i += 2) {
        CU(cudaEventSynchronize(idx->ev[i + 1]));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, idx->ev[i], idx->ev[i + 1]));
        idx->stats.coarse_ms += ms;
        idx->stats.coarse_timed++;
    }
    idx->ev_used = 0;
    return MCP_OK;
}
static int coarse_event(mcp_index *idx, cudaStream_t st) {
    if (!idx->opt_time_coarse) return MCP_OK;
    if (idx->ev_used == (int)idx->ev.size()) {
        if (idx->ev.size() >= 8192 && (idx->ev_used & 1) == 0) { int rc = harvest_events(idx); if (rc) return rc; }
        else { cudaEvent_t e; CU(cudaEventCreate(&e)); idx->ev.push_back(e); }
    }
    CU(cudaEventRecord(idx->ev[idx->ev_used++], st));
    return MCP_OK;
}
// the workspace is shared by all searches of an index: same-stream calls queue, others wait
static int order_after_previous(mcp_index *idx, cudaStream_t st) {
    if (idx->up_pending && st != idx->stream) CU(cudaStreamWaitEvent(st, idx->up_done, 0));   // a row upsert ahead of this call: device-side wait
    if (idx->pending && idx->last_stream != st) CU(cudaEventSynchronize(idx->done));
    idx->last_stream = st;
    return MCP_OK;
}

static int pow2_at_least(int v) { int p = 1; while (p < v) p <<= 1; return p; }

template <int QT, int NJ>
static cudaError_t launch_scan(int metric, int grid, size_t smem, cudaStream_t st, const float *rows, int64_t n, int dim,
                               int64_t ld, const float *q, const float *inv_nq, const float *inv_ne, const float *pen,
                               float *scores, int64_t stride, int nvalid) {
    if (metric == 0) {
        cudaFuncSetAttribute(scan_fp32_kernel<QT, NJ, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        scan_fp32_kernel<QT, NJ, 0><<<grid, SCAN_THREADS, smem, st>>>(rows, n, dim, ld, q, inv_nq, inv_ne, pen, scores, stride, nvalid);
    } else {
        cudaFuncSetAttribute(scan_fp32_kernel<QT, NJ, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        scan_fp32_kernel<QT, NJ, 1><<<grid, SCAN_THREADS, smem, st>>>(rows, n, dim, ld, q, inv_nq, inv_ne, pen, scores, stride, nvalid);
    }
    return cudaGetLastError();
}

// K1 over one pass of up to 8 queries
static int run_scan_pass(mcp_index *idx, cudaStream_t st, const float *q, int dim, const float *inv_nq, int nvalid,
                         float *scores, int64_t stride) {
    int qt = nvalid <= 1 ? 1 : nvalid <= 2 ? 2 : nvalid <= 4 ? 4 : 8;
    if (nvalid > qt) return fail(MCP_EINVAL, "internal: scan pass of %d > %d", nvalid, qt);
    size_t smem = (size_t)qt * dim * 4;
    if (smem > 200 * 1024) return fail(MCP_EINVAL, "internal: scan pass needs %zu bytes of shared memory", smem);
    int per_sm = qt >= 8 ? 2 : 4;
    int64_t want = (idx->n_rows + (SCAN_THREADS / 32) - 1) / (SCAN_THREADS / 32);
    int grid = (int)std::min<int64_t>((int64_t)idx->n_sm * per_sm, want);
    const float *pen = idx->has_penalty ? idx->penalty.as<float>() : nullptr;
    const float *inv_ne = idx->inv_norm.as<float>();
    // rows of 512 / 768 / 1024 / 1536 / 2048 / 3072 floats get all their loads issued up front (NJ float4 per lane);
    // any other dim % 4 == 0 takes the generic loop (4 loads in flight per lane)
    const int nj = (dim % 128 == 0) ? dim / 128 : 0;
    cudaError_t e;
#define SCAN_ARGS idx->metric, grid, smem, st, idx->d_rows, idx->n_rows, dim, idx->ld, q, inv_nq, inv_ne, pen, scores, stride, nvalid
#define SCAN_CASE(QT)                                                     \
    switch (nj) {                                                         \
        case 4:  e = launch_scan<QT, 4>(SCAN_ARGS); break;                \
        case 6:  e = launch_scan<QT, 6>(SCAN_ARGS); break;                \
        case 8:  e = launch_scan<QT, 8>(SCAN_ARGS); break;                \
        case 12: e = launch_scan<QT, 12>(SCAN_ARGS); break;               \
        case 16: e = launch_scan<QT, 16>(SCAN_ARGS); break;               \
        case 24: e = launch_scan<QT, 24>(SCAN_ARGS); break;               \
        default: e = launch_scan<QT, 0>(SCAN_ARGS); break;                \
    }
    switch (qt) {
        case 1: SCAN_CASE(1); break;
        case 2: SCAN_CASE(2); break;
        case 4: SCAN_CASE(4); break;
        default: SCAN_CASE(8); break;
    }
#undef SCAN_ARGS
#undef SCAN_CASE
    if (e != cudaSuccess) return fail(MCP_ECUDA, "scan_fp32 launch: %s", cudaGetErrorString(e));
    idx->stats.kernel_launches++;
    return MCP_OK;
}

// reduce [nq][L][cap] lists until one list per query is left; returns the final buffers
static int reduce_lists(mcp_index *idx, cudaStream_t st, int nq, int kprime, u64 *in, int *in_cnt, int L, int cap,
                        bool in_is_A, u64 **out_keys, int **out_cnt, int *out_stride, bool raw_in = false,
                        const u64 *tau_pub = nullptr /* K2's final shared bound per query, or NULL */, uint32_t epoch = 0) {
    int rc;
    bool first = true;
    while (L > 1 || first) {
        if (L == 1 && cap <= kprime && !raw_in) break;   // already one list of at most k' ordered keys
        first = false;
        int G = std::min(L, SEL_THREADS);          // one CTA per query unless there are more than SEL_THREADS lists
        // Few queries, many lists (one query over 1 M rows leaves ~300 lists and ~15 K keys): one CTA per query would chew
        // through them 4096 keys at a time (55 us measured); groups of >= 16 lists go to CTAs of their own and a second round
        // reduces the group results, so that about as many CTAs as there are SMs work at once.
        // (Measured with K2's final bound filtering the keys as they are gathered: two rounds 0.573 ms, one CTA per query 0.592 ms
        // for one query over 1 M rows.)
        if (nq < idx->n_sm && L > 32) G = std::min(L, std::max(16, (int)((int64_t)L * nq / idx->n_sm + 1)));
        int L_out = (L + G - 1) / G;
        Buf &ob = in_is_A ? idx->listsB : idx->listsA;
        Buf &oc = in_is_A ? idx->cntB : idx->cntA;
        if ((rc = ob.ensure(sizeof(u64) * (size_t)nq * L_out * kprime))) return rc;
        if ((rc = oc.ensure(sizeof(int) * (size_t)nq * L_out))) return rc;
        CU(launch_pdl(select_from_lists_kernel, dim3(L_out, nq), dim3(SEL_THREADS), 0, st, in, in_cnt, L, cap, G, kprime, ob.as<u64>(),
                      oc.as<int>(), kprime, raw_in ? 1 : 0, raw_in ? tau_pub : (const u64 *)nullptr, epoch));
        KL();
        in = ob.as<u64>(); in_cnt = oc.as<int>(); L = L_out; cap = kprime; in_is_A = !in_is_A; raw_in = false;
    }
    *out_keys = in; *out_cnt = in_cnt; *out_stride = cap;
    return MCP_OK;
}

static int choose_kprime(const mcp_index *idx, int k) {
    // automatic: half as many again as asked for, at least 32 more (the residual-based certificate bound is a quarter of the
    // worst-case one this used to be sized for, when it was 2k)
    int kp = idx->opt_kprime ? std::max(idx->opt_kprime, k) : std::max(k + k / 2, k + 32) << idx->kp_boost;
    kp = (kp + 31) & ~31;
    return std::min(kp, FINAL_MAX);
}

// Automatic oversampling follows the data: when more than 1 in 200 queries since the last look needed the fp64 scan (tight
// clusters, near-duplicates: the gap between rank k and rank k' is inside the coarse error band), the next searches keep twice
// as many candidates (up to 4x, within the kernels' limits) -- rescoring 2k' rows costs microseconds, a full scan a table
// read.  64 clean looks in a row step it back down.  Exactness never depends on this; only how often the fallback runs does.
static void note_outcome(mcp_index *idx, int failures) {
    const int64_t queries = idx->q_since_look;
    idx->q_since_look = 0;
    if (queries <= 0) return;
    if ((int64_t)failures * 200 > queries) { if (idx->kp_boost < 2) idx->kp_boost++; idx->clean_looks = 0; }
    else if (failures == 0 && idx->kp_boost > 0 && ++idx->clean_looks >= 64) { idx->kp_boost--; idx->clean_looks = 0; }
}

static int choose_path(const mcp_index *idx, int nq) {
    if (idx->h_consts.odd_rows) return MCP_PATH_EXACT_FP64;   // rows the fp32 coarse pass cannot bound
    if (idx->opt_path == MCP_PATH_GEMM_F16 && !idx->has_shadow) return MCP_PATH_GEMM_TF32;   // no shadow: the fp32 rows themselves
    if (idx->opt_path != MCP_PATH_AUTO) return idx->opt_path;
    // With a shadow the fp16 scan reads half the bytes, which wins at every batch size once the table is large enough for the
    // scan (not the launches) to be the cost: 1 query over 1 M rows 0.64 ms against 0.90 ms for the fp32 streaming kernel.
    if (idx->has_shadow && (nq >= 8 || idx->n_rows >= 8192)) return MCP_PATH_GEMM_F16;
    if (nq < 8) return MCP_PATH_STREAM_FP32;
    return MCP_PATH_GEMM_TF32;
}
static bool is_gemm_path(int path) { return path == MCP_PATH_GEMM_TF32 || path == MCP_PATH_GEMM_F16; }

// slots one pass of the fp64 scan serves: what fits in shared memory beside the row ring, at most opt_max_fallback
static int scan_slots(const mcp_index *idx) {
    const size_t per = sizeof(double) * (size_t)((idx->dim + 3) & ~3);
    const int fit = (int)((XS_SMEM_LIMIT - XS_TILE_BYTES) / per);
    return std::max(1, std::min(fit, idx->opt_max_fallback));
}
static int scan_grid(const mcp_index *idx) {
    const int64_t tiles = (idx->n_rows + XS_ROWS - 1) / XS_ROWS;
    return (int)std::max<int64_t>(1, std::min<int64_t>(idx->n_sm, tiles));
}
// the fp64 scan's TMA view of the fp32 table: boxes of XS_ROWS rows x 32 columns (host-side encode, no allocation)
static int scan_tensor_map(mcp_index *idx, CUtensorMap *tm) {
    if (!gemm_supported(0, idx->d_rows, idx->ld)) return fail(MCP_EINVAL, "the fp64 scan needs a 16-byte aligned table with ld %% 4 == 0");
    if (gemm_encode_2d(idx->gemm, tm, 0, idx->d_rows, idx->n_rows, idx->dim, idx->ld, XS_ROWS)) return fail(MCP_ECUDA, "%s", idx->gemm.err);
    return MCP_OK;
}

// path = exact_fp64: flag_list holds `c` queries (flag_all_kernel); dense fp64 scan S slots at a time + exact select
static int run_exact_dense(mcp_index *idx, cudaStream_t st, const float *d_q, const double *nq2, int k, int c,
                           int32_t *out_ids, float *out_scores, double *out_scores64) {
    int rc;
    const int S = scan_slots(idx);
    if ((rc = idx->exact_all.ensure(sizeof(u64) * (size_t)S * (size_t)idx->n_rows))) return rc;
    int *flag_list = idx->flags.as<int>() + 8, *n_flag = idx->flags.as<int>();
    const float *pen = idx->has_penalty ? idx->penalty.as<float>() : nullptr;
    CUtensorMap tm;
    if ((rc = scan_tensor_map(idx, &tm))) return rc;
    for (int s0 = 0; s0 < c; s0 += S) {
        const int sc = std::min(S, c - s0);
        exact_scan2_kernel<<<scan_grid(idx), XS_THREADS, XS_SMEM_BYTES(idx->dim, S), st>>>(
            tm, idx->n_rows, idx->dim, d_q, nq2, idx->norm2.as<double>(), pen, idx->metric, flag_list, n_flag, s0, S,
            idx->exact_all.as<u64>(), XsSparse{});
        KL();
        exact_select_kernel<<<sc, EXSEL_THREADS, 0, st>>>(idx->exact_all.as<u64>(), idx->n_rows, k, idx->id_base, flag_list, n_flag, s0,
                                                          out_ids, out_scores, out_scores64);
        KL();
    }
    return MCP_OK;
}

// The certificate's fallback, device-driven: one sparse fp64 scan over however many queries were flagged (none, normally:
// the kernel then exits at once) and one pick kernel that writes their exact top k and folds the status words.
static int run_fallback(mcp_index *idx, cudaStream_t st, const float *d_q, const double *nq2, int k, int cq, int q_base,
                        int32_t *out_ids, float *out_scores, double *out_scores64) {
    int rc;
    const size_t need = (size_t)cq * XS_SPARSE_CAP;
    if (sizeof(int) * (size_t)cq > idx->sp_cnt.cap) {
        if ((rc = idx->sp_cnt.ensure(sizeof(int) * (size_t)std::max(cq, 1024)))) return rc;
        CU(cudaMemsetAsync(idx->sp_cnt.p, 0, idx->sp_cnt.cap, st));
    }
    if ((rc = idx->sp_keys.ensure(sizeof(u64) * need))) return rc;
    if ((rc = idx->sp_rows.ensure(sizeof(int) * need))) return rc;
    int *flag_list = idx->flags.as<int>() + 8, *n_flag = idx->flags.as<int>();
    const float *pen = idx->has_penalty ? idx->penalty.as<float>() : nullptr;
    XsSparse sp{idx->bound_key.as<u64>(), idx->bound_row.as<int>(), idx->sp_keys.as<u64>(), idx->sp_rows.as<int>(), idx->sp_cnt.as<int>(), XS_SPARSE_CAP};
    const int S = scan_slots(idx);
    CUtensorMap tm;
    if ((rc = scan_tensor_map(idx, &tm))) return rc;
    CU(launch_pdl(exact_scan2_kernel, dim3(scan_grid(idx)), dim3(XS_THREADS), XS_SMEM_BYTES(idx->dim, S), st,
                  tm, idx->n_rows, idx->dim, d_q, nq2, idx->norm2.as<double>(), pen, idx->metric, flag_list, n_flag, 0, S,
                  (u64 *)nullptr, sp));
    KL();
    CU(launch_pdl(exact_pick_kernel, dim3(std::min(cq, 16)), dim3(256), 0, st, flag_list, n_flag, sp, k, idx->id_base, q_base,
                  out_ids, out_scores, out_scores64, n_flag + 1, idx->h_status_dev, idx->unres.as<int>(), UNRES_MAX));
    KL();
    return MCP_OK;
}

static int search_locked(mcp_index *idx, const float *d_q, int nq, int k, int32_t *out_ids, float *out_scores,
                         double *out_scores64, cudaStream_t st) {
    int rc;
    const int dim = idx->dim;
    const int64_t n = idx->n_rows;
    const int path = choose_path(idx, nq);
    int kprime = choose_kprime(idx, k);
    if (is_gemm_path(path)) kprime = std::min(kprime, gemm_max_kprime());   // >= k since k <= 256
    const float *pen = idx->has_penalty ? idx->penalty.as<float>() : nullptr;

    bool q_copied = false;                       // the tensor-core scan reads the library's operand image of the queries, not the caller's buffer
    if ((rc = idx->nq2.ensure(sizeof(double) * nq))) return rc;
    if ((rc = idx->inv_nq.ensure(sizeof(float) * nq))) return rc;
    if ((rc = idx->qflag.ensure(nq))) return rc;
    if ((size_t)(nq + 8) * sizeof(int) > idx->flags.cap) return fail(MCP_EINVAL, "n_queries %d exceeds the flag buffer", nq);
    // flags: [0] n_flag (zero between kernels) [1] certificate failures [2] unresolved (both since last sync_status) [8..] list
    int *n_flag = idx->flags.as<int>(), *flag_list = n_flag + 8;
    if ((rc = idx->drop_min.ensure(sizeof(unsigned int) * nq))) return rc;
    if ((rc = idx->bound_key.ensure(sizeof(u64) * nq))) return rc;
    if ((rc = idx->bound_row.ensure(sizeof(int) * nq))) return rc;
    if ((rc = idx->nq2f.ensure(sizeof(float) * nq))) return rc;
    // coarse kernels only need fp32 norms; the exact fp64 norm of each query is computed inside the fused rescoring
    // kernel (as one more chain), so nothing serial sits in front of the scan
    if (path == MCP_PATH_EXACT_FP64) {
        prep_queries_kernel<<<nq, PREPQ_THREADS, sizeof(float) * ((dim + 3) & ~3), st>>>(
            d_q, dim, idx->nq2.as<double>(), idx->inv_nq.as<float>(), idx->qflag.as<unsigned char>(), idx->nq2f.as<float>());
        KL();
    } else {
        // the tensor-core scans read an operand image of the queries: scaled fp16 (pitch ldh), or -- tf32 with dim % 4 != 0 -- an fp32
        // copy whose pitch a tensor map accepts; otherwise the caller's queries are the operand
        // (also when the batch is not a multiple of 8 queries: the image is padded with zero rows, see the kernel)
        void *q_out = nullptr; int ldq = dim, as_f16 = 0;
        const int nq_pad = (nq + 7) & ~7;
        if (path == MCP_PATH_GEMM_F16) { ldq = (int)idx->ldh; as_f16 = 1; }
        else if (path == MCP_PATH_GEMM_TF32) ldq = (dim + 3) & ~3;
        q_copied = as_f16 || (path == MCP_PATH_GEMM_TF32 && (ldq != dim || nq_pad != nq));
        if (q_copied) {
            if ((rc = idx->q_op.ensure((size_t)nq_pad * ldq * (as_f16 ? 2 : 4)))) return rc;
            q_out = idx->q_op.p;
        }
        if ((rc = idx->rho_q.ensure(sizeof(float) * 2 * (size_t)nq))) return rc;
        if ((rc = idx->q_scale.ensure(sizeof(float) * (size_t)nq))) return rc;
        prep_queries_fast_kernel<<<((q_copied ? nq_pad : nq) + 3) / 4, 128, 0, st>>>(d_q, nq, dim, idx->inv_nq.as<float>(), idx->nq2f.as<float>(), n_flag,
                                                               idx->drop_min.as<unsigned int>(), idx->rho_q.as<float>(), q_out, ldq, as_f16,
                                                               idx->q_scale.as<float>(), q_copied ? nq_pad : nq);
        KL();
    }

    idx->stats.searches++; idx->stats.queries += nq; idx->stats.last_path = path; idx->stats.last_kprime = kprime;
    if (path != MCP_PATH_EXACT_FP64) idx->q_since_look += nq;

    if (path == MCP_PATH_EXACT_FP64) {
        const int slots = 256;                     // flag list entries per round (the scan takes them scan_slots() at a time)
        for (int q0 = 0; q0 < nq; q0 += slots) {
            int c = std::min(slots, nq - q0);
            flag_all_kernel<<<(c + 127) / 128, 128, 0, st>>>(flag_list, n_flag, q0, c);
            KL();
            if ((rc = run_exact_dense(idx, st, d_q, idx->nq2.as<double>(), k, c, out_ids, out_scores, out_scores64))) return rc;
        }
        clear_flag_kernel<<<1, 1, 0, st>>>(n_flag);
        KL();
        idx->stats.exact_scans += nq;
        return MCP_OK;
    }

    // query chunking keeps the K1 score matrix within budget; K2 needs no score matrix
    int chunk = std::min(nq, G_MAX_QBLOCKS * G_BM);
    if (path == MCP_PATH_STREAM_FP32)
        chunk = (int)std::max<int64_t>(1, std::min<int64_t>(nq, (int64_t)(SCORE_BUDGET_BYTES / (4ull * (size_t)n))));
    const int kc = (int)std::min<int64_t>(kprime, n);

    CertParams cp;
    cp.metric = idx->metric; cp.path = path; cp.complete = (n <= kprime);
    const double u = 5.9604644775390625e-08;   // 2^-24
    cp.kind = 0;
    if (path == MCP_PATH_STREAM_FP32) cp.rel = 2.0 * (dim / 32.0 + 32.0) * u;   // + the fp32 tree-summed query norm
    else {
        // operand rounding enters per query through rho_q and the table's rho_e (rescore.cuh); what is left is the fp32
        // accumulation inside the tensor core and the rounding of the two fp32 inverse norms (the query's is a tree sum)
        cp.rel = dim * 4.0 * u + 128.0 * u;
        cp.kind = path == MCP_PATH_GEMM_F16 ? 2 : 1;     // the table's rho_e is read on the device (TableConsts)
    }
    cp.abs_pen = 8.0 * u;
    cp.debug = idx->gemm.debug >> 8;

    for (int q0 = 0; q0 < nq; q0 += chunk) {
        const int cq = std::min(chunk, nq - q0);
        const float *dq = d_q + (int64_t)q0 * dim;
        u64 *cand; int *cand_cnt; int cand_stride;
        if (q0 > 0) { clear_flag_kernel<<<1, 1, 0, st>>>(n_flag); KL(); }   // the previous chunk's flags have been served
        if ((rc = coarse_event(idx, st))) return rc;
        if (path == MCP_PATH_STREAM_FP32) {
            // the scan reads 128-bit words: rows are padded to ld by load(); queries get the same zero padding
            const int dimp = (dim + 3) & ~3;
            const float *dqs = dq;
            if (dimp != dim) {
                if ((rc = idx->q_pad.ensure(sizeof(float) * (size_t)cq * dimp))) return rc;
                pad_queries_kernel<<<cq, 128, 0, st>>>(dq, dim, dimp, idx->q_pad.as<float>());
                KL();
                dqs = idx->q_pad.as<float>();
            }
            if ((rc = idx->scores.ensure(sizeof(float) * (size_t)cq * (size_t)n))) return rc;
            int qt_max = 8;
            while (qt_max > 1 && (size_t)qt_max * dimp * 4 > 160 * 1024) qt_max >>= 1;
            for (int p = 0; p < cq; p += qt_max) {
                int nv = std::min(qt_max, cq - p);
                if ((rc = run_scan_pass(idx, st, dqs + (int64_t)p * dimp, dimp, idx->inv_nq.as<float>() + q0 + p, nv,
                                        idx->scores.as<float>() + (int64_t)p * n, n)))
                    return rc;
            }
            if ((rc = coarse_event(idx, st))) return rc;
            int P = (int)((n + SEL_MAX - 1) / SEL_MAX);
            if ((rc = idx->listsA.ensure(sizeof(u64) * (size_t)cq * P * kprime))) return rc;
            if ((rc = idx->cntA.ensure(sizeof(int) * (size_t)cq * P))) return rc;
            select_from_scores_kernel<<<dim3(P, cq), SEL_THREADS, 0, st>>>(idx->scores.as<float>(), n, n, kprime,
                                                                          idx->listsA.as<u64>(), idx->cntA.as<int>(), kprime);
            KL();
            if ((rc = reduce_lists(idx, st, cq, kprime, idx->listsA.as<u64>(), idx->cntA.as<int>(), P, kprime, true,
                                   &cand, &cand_cnt, &cand_stride)))
                return rc;
        } else {
            int L, cap; u64 *lists; int *cnts;
            const bool bf = path == MCP_PATH_GEMM_F16;
            const int ldq = bf ? (int)idx->ldh : (dim + 3) & ~3;
            const void *qop = bf ? (const void *)(idx->q_op.as<__half>() + (int64_t)q0 * ldq)
                                 : q_copied ? (const void *)(idx->q_op.as<float>() + (int64_t)q0 * ldq) : (const void *)dq;
            if ((rc = gemm_tf32_run(idx->gemm, idx->gemm_ws, st, idx->n_sm, bf ? 1 : 0, bf ? (const void *)idx->shadow.p : (const void *)idx->d_rows,
                                    n, dim, bf ? idx->ldh : idx->ld, qop, ldq, q_copied ? (cq + 7) & ~7 : cq, cq,
                                    idx->inv_nq.as<float>() + q0, idx->nq2f.as<float>() + q0, idx->inv_norm.as<float>(),
                                    idx->norm2.as<double>(), pen, idx->metric, kprime, &lists, &cnts, &L, &cap, nullptr,
                                    idx->drop_min.as<unsigned int>() + q0, bf ? idx->q_scale.as<float>() + q0 : nullptr,
                                    bf ? idx->shadow_scale.as<float>() : nullptr)))
                return fail(rc, "%s", idx->gemm.err);
            idx->stats.kernel_launches++;
            if ((rc = coarse_event(idx, st))) return rc;
            // the GEMM's lists live in gemm_ws; reductions ping-pong between listsA / listsB
            // K2 lists are raw (fp32 bits, row) pairs: always pass them through one reduction, which builds the ordered keys
            if ((rc = reduce_lists(idx, st, cq, kprime, lists, cnts, L, cap, false, &cand, &cand_cnt, &cand_stride, true,
                                   idx->gemm.last_tau_pub, idx->gemm.last_epoch)))
                return rc;
        }
        {
            const int G = (std::min(cand_stride, kprime) + RS_ROWS - 1) / RS_ROWS;
            int GP = std::max(1, cq <= idx->n_sm ? std::min(G, RF_MAXG) : std::min(G, 2));   // few queries: one pass; many: more CTAs per SM
            const int npow2 = pow2_at_least(std::max(std::min(cand_stride, kprime), 2));
            int stages = 5;
            // Many queries, three groups of candidates (k' = 96): two passes of two groups leave a half-empty second pass.  A
            // shallower ring (3 stages) takes all three groups in one pass and still fits two CTAs per SM.
            const bool shallow = cq > idx->n_sm && G == 3 && RF_SMEM_BYTES(dim, 3, 3, npow2) <= 112 * 1024 && !(idx->gemm.debug & 16384);
            if (shallow) { GP = 3; stages = 3; }
            while (GP > 1 && RF_SMEM_BYTES(dim, GP, stages, npow2) > RF_SMEM_LIMIT) --GP;
            auto rf_kernel = shallow ? rescore_finalize_kernel<3, 2> : rescore_finalize_kernel<5, 3>;
            CU(launch_pdl(rf_kernel, dim3(cq), dim3(32 * (GP + RF_NPROD + 1)), RF_SMEM_BYTES(dim, GP, stages, npow2), st,
                          idx->d_rows, dim, idx->ld, dq, idx->norm2.as<double>(), pen, idx->metric, cand, cand_cnt, cand_stride,
                          npow2, k, idx->id_base, idx->consts.as<TableConsts>(), cp,
                          idx->nq2.as<double>() + q0, idx->qflag.as<unsigned char>() + q0, out_ids + (int64_t)q0 * k,
                          out_scores + (int64_t)q0 * k, out_scores64 ? out_scores64 + (int64_t)q0 * k : nullptr, flag_list, n_flag,
                          is_gemm_path(path) ? idx->drop_min.as<unsigned int>() + q0 : nullptr,
                          is_gemm_path(path) ? idx->rho_q.as<float>() + 2 * q0 : nullptr,
                          idx->bound_key.as<u64>(), idx->bound_row.as<int>()));
            KL();
        }
        // certificate failures: fp64 scan on the device, flag_list holds chunk-local query numbers
        if ((rc = run_fallback(idx, st, dq, idx->nq2.as<double>() + q0, k, cq, q0, out_ids + (int64_t)q0 * k, out_scores + (int64_t)q0 * k,
                               out_scores64 ? out_scores64 + (int64_t)q0 * k : nullptr)))
            return rc;
        // flags[1] failures, [2] unresolved, [3] ticket: folded by exact_pick_kernel's last CTA
        idx->stats.rows_scanned += n;
        idx->stats.bytes_scanned += path == MCP_PATH_GEMM_F16 ? n * idx->ldh * 2 : n * (int64_t)dim * 4;
        idx->stats.candidates_rescored += (int64_t)cq * kc;
    }
    return MCP_OK;
}

static int validate_search(const mcp_index *idx, const void *q, int nq, int k, const void *ids, const void *sc) {
    if (!idx->n_rows) return fail(MCP_ESTATE, "search before load");
    if (!q || !ids || !sc) return fail(MCP_EINVAL, "NULL buffer");
    if (nq < 1 || nq > 4000) return fail(MCP_EINVAL, "n_queries %d outside [1, 4000]", nq);
    if (k < 1 || k > MCP_MAX_K) return fail(MCP_EINVAL, "k %d outside [1, %d]", k, MCP_MAX_K);
    return MCP_OK;
}

// Re-answer the queries listed in idx->unres with the dense fp64 scan (one at a time: this is the rare path behind the rare
// path) and clear the list.  Synchronises the stream.
static int resolve_leftovers(mcp_index *idx, const float *d_q, int nq, int k, int32_t *out_ids, float *out_scores,
                             double *out_scores64, cudaStream_t st) {
    std::vector<int> h(1 + UNRES_MAX);
    CU(cudaMemcpyAsync(h.data(), idx->unres.p, sizeof(int) * (1 + UNRES_MAX), cudaMemcpyDeviceToHost, st));
    CU(cudaMemsetAsync(idx->unres.p, 0, sizeof(int), st));
    CU(cudaStreamSynchronize(st));
    const int n = h[0];
    const int saved = idx->opt_path;
    idx->opt_path = MCP_PATH_EXACT_FP64;
    int rc = MCP_OK;
    if (n > UNRES_MAX) {                                  // more than the list holds: the whole batch
        rc = search_locked(idx, d_q, nq, k, out_ids, out_scores, out_scores64, st);
    } else {
        for (int i = 0; i < n && !rc; ++i) {
            const int q = h[1 + i];
            if (q < 0 || q >= nq) continue;
            rc = search_locked(idx, d_q + (int64_t)q * idx->dim, 1, k, out_ids + (int64_t)q * k, out_scores + (int64_t)q * k,
                               out_scores64 ? out_scores64 + (int64_t)q * k : nullptr, st);
        }
    }
    idx->opt_path = saved;
    return rc;
}

extern "C" int mcp_index_search_device(mcp_index *idx, const float *queries_device, int n_queries, int k,
                                       int32_t *out_ids_device, float *out_scores_device,
                                       double *out_scores64_device, void *stream) {
    if (check_idx(idx)) return MCP_EINVAL;
    NOT_ON_MULTI(idx, "search_device");
    std::lock_guard<std::mutex> lk(idx->mu);
    int rc = validate_search(idx, queries_device, n_queries, k, out_ids_device, out_scores_device);
    if (rc) return rc;
    CU(cudaSetDevice(idx->device));
    if ((rc = require_local(queries_device, sizeof(float) * (size_t)n_queries * idx->dim, idx->device, "search_device queries"))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if ((rc = order_after_previous(idx, st))) return rc;
    rc = search_locked(idx, queries_device, n_queries, k, out_ids_device, out_scores_device, out_scores64_device, st);
    if (rc) return rc;
    // the status words reach idx->h_status through the fallback kernel's fold (host-mapped memory): no copy node
    CU(cudaEventRecord(idx->done, st));
    idx->pending = true;
    return MCP_OK;
}

extern "C" int mcp_index_sync_status(mcp_index *idx, int32_t *n_certificate_failures, int32_t *n_unresolved) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (idx->multi) {                                  // host-buffer searches only: nothing is ever pending
        if (n_certificate_failures) *n_certificate_failures = 0;
        if (n_unresolved) *n_unresolved = 0;
        return MCP_OK;
    }
    std::lock_guard<std::mutex> lk(idx->mu);
    CU(cudaSetDevice(idx->device));
    int fails = 0, unresolved = 0;
    if (idx->pending) {
        CU(cudaEventSynchronize(idx->done));
        idx->pending = false;
        fails = idx->h_status[0]; unresolved = idx->h_status[1];
        idx->h_status[0] = idx->h_status[1] = 0;      // the device rewrites them from its own (now cleared) counters
        idx->stats.certificate_failures += fails;
        idx->stats.exact_scans += fails - unresolved;
        note_outcome(idx, fails);
        CU(cudaMemsetAsync(idx->flags.as<int>() + 1, 0, 2 * sizeof(int), idx->last_stream));
        CU(cudaStreamSynchronize(idx->last_stream));
    }
    int rc = harvest_events(idx);
    if (rc) return rc;
    if (n_certificate_failures) *n_certificate_failures = fails;
    if (n_unresolved) *n_unresolved = unresolved;
    return MCP_OK;
}

extern "C" int mcp_index_repair_device(mcp_index *idx, const float *queries_device, int n_queries, int k,
                                       int32_t *out_ids_device, float *out_scores_device, double *out_scores64_device,
                                       void *stream) {
    if (check_idx(idx)) return MCP_EINVAL;
    NOT_ON_MULTI(idx, "repair_device");
    std::lock_guard<std::mutex> lk(idx->mu);
    int rc = validate_search(idx, queries_device, n_queries, k, out_ids_device, out_scores_device);
    if (rc) return rc;
    CU(cudaSetDevice(idx->device));
    if ((rc = require_local(queries_device, sizeof(float) * (size_t)n_queries * idx->dim, idx->device, "repair_device queries"))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if ((rc = order_after_previous(idx, st))) return rc;
    if (idx->pending) { CU(cudaEventSynchronize(idx->done)); idx->pending = false; }
    rc = resolve_leftovers(idx, queries_device, n_queries, k, out_ids_device, out_scores_device, out_scores64_device, st);
    if (rc) return rc;
    CU(cudaStreamSynchronize(st));
    return MCP_OK;
}

extern "C" int mcp_index_search(mcp_index *idx, const float *queries_host, int n_queries, int k, int32_t *out_ids,
                                float *out_scores, double *out_scores64) {
    if (check_idx(idx)) return MCP_EINVAL;
    if (idx->multi) return multi_search(idx, queries_host, n_queries, k, out_ids, out_scores, out_scores64);
    std::unique_lock<std::mutex> lk(idx->mu);
    int rc = validate_search(idx, queries_host, n_queries, k, out_ids, out_scores);
    if (rc) return rc;
    CU(cudaSetDevice(idx->device));
    cudaStream_t st = idx->stream;
    if (idx->pending) {   // fold an unfinished search_device into the stats first: this call reports only its own status
        CU(cudaEventSynchronize(idx->done));
        idx->pending = false;
        idx->stats.certificate_failures += idx->h_status[0];
        idx->stats.exact_scans += idx->h_status[0] - idx->h_status[1];
        note_outcome(idx, idx->h_status[0]);
        CU(cudaMemsetAsync(idx->flags.as<int>() + 1, 0, 2 * sizeof(int), st));
    }
    idx->h_status[0] = idx->h_status[1] = 0;      // device counters are zero here (cleared after every host-buffer search)
    idx->last_stream = st;
    // Results come back in ONE device->host copy into a pinned staging buffer, after ONE synchronisation
    // (device layout [scores64 f64 nk | ids i32 nk | scores f32 nk]); the status words arrive in host-mapped memory.
    const size_t nk = (size_t)n_queries * k;
    const size_t off_ids = sizeof(double) * nk, off_sc = off_ids + sizeof(int32_t) * nk;
    const size_t total = off_sc + sizeof(float) * nk;
    if ((rc = idx->q_copy.ensure(sizeof(float) * (size_t)n_queries * idx->dim))) return rc;
    if ((rc = idx->out_ids.ensure(total))) return rc;                 // one buffer holds all three outputs
    if (total > idx->h_stage_cap) {
        if (idx->h_stage) cudaFreeHost(idx->h_stage);
        idx->h_stage = nullptr; idx->h_stage_cap = 0;
        CU(cudaMallocHost(&idx->h_stage, total + total / 4));
        idx->h_stage_cap = total + total / 4;
    }
    unsigned char *dbase = idx->out_ids.as<unsigned char>();
    double *d_s64 = reinterpret_cast<double *>(dbase);
    int32_t *d_ids = reinterpret_cast<int32_t *>(dbase + off_ids);
    float *d_sc = reinterpret_cast<float *>(dbase + off_sc);
    CU(cudaMemcpyAsync(idx->q_copy.p, queries_host, sizeof(float) * (size_t)n_queries * idx->dim, cudaMemcpyHostToDevice, st));
    rc = search_locked(idx, idx->q_copy.as<float>(), n_queries, k, d_ids, d_sc, d_s64, st);
    if (rc) return rc;
    CU(cudaMemcpyAsync(idx->h_stage, dbase, total, cudaMemcpyDeviceToHost, st));
    CU(cudaMemsetAsync(idx->flags.as<int>() + 1, 0, 2 * sizeof(int), st));
    CU(cudaStreamSynchronize(st));
    const int hs[2] = {idx->h_status[0], idx->h_status[1]};
    idx->h_status[0] = idx->h_status[1] = 0;
    idx->stats.certificate_failures += hs[0];
    idx->stats.exact_scans += hs[0] - hs[1];
    note_outcome(idx, hs[0]);
    if (hs[1] > 0) {
        // Queries whose fallback list overflowed (thousands of rows inside the coarse error band of the k-th score): exactly
        // those are re-answered by the dense fp64 scan, each into its own rows of the result; nothing else is recomputed.
        rc = resolve_leftovers(idx, idx->q_copy.as<float>(), n_queries, k, d_ids, d_sc, d_s64, st);
        if (rc) return rc;
        CU(cudaMemcpyAsync(idx->h_stage, dbase, total, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    memcpy(out_ids, idx->h_stage + off_ids, sizeof(int32_t) * nk);
    memcpy(out_scores, idx->h_stage + off_sc, sizeof(float) * nk);
    if (out_scores64) memcpy(out_scores64, idx->h_stage, sizeof(double) * nk);
    return MCP_OK;
}

// coarse scores of every (query,row) pair from one coarse kernel: lets tests check the error bounds the
// certificate assumes, and localises a wrong tensor-core descriptor to the element.
extern "C" int mcp_index_coarse_scores_device(mcp_index *idx, int path, const float *queries_device, int n_queries,
                                              float *out_scores_device, void *stream) {
    if (check_idx(idx)) return MCP_EINVAL;
    NOT_ON_MULTI(idx, "coarse_scores_device");
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->n_rows) return fail(MCP_ESTATE, "coarse_scores before load");
    if (!queries_device || !out_scores_device || n_queries < 1 || n_queries > G_MAX_QBLOCKS * G_BM) return fail(MCP_EINVAL, "bad arguments");
    CU(cudaSetDevice(idx->device));
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if ((rc = require_local(queries_device, sizeof(float) * (size_t)n_queries * idx->dim, idx->device, "coarse_scores queries"))) return rc;
    if ((rc = order_after_previous(idx, st))) return rc;
    const int dim = idx->dim;
    if ((rc = idx->nq2.ensure(sizeof(double) * n_queries))) return rc;
    if ((rc = idx->inv_nq.ensure(sizeof(float) * n_queries))) return rc;
    if ((rc = idx->qflag.ensure(n_queries))) return rc;
    if ((rc = idx->nq2f.ensure(sizeof(float) * n_queries))) return rc;
    // the same query norms a search feeds its coarse kernel
    const bool bf = path == MCP_PATH_GEMM_F16;
    if (bf && !idx->has_shadow) return fail(MCP_ESTATE, "coarse_scores: the index holds no fp16 shadow table");
    const int ldq = bf ? (int)idx->ldh : (dim + 3) & ~3;
    const int nq_pad = (n_queries + 7) & ~7;
    void *q_out = nullptr;
    if (bf || (path == MCP_PATH_GEMM_TF32 && (ldq != dim || nq_pad != n_queries))) {
        if ((rc = idx->q_op.ensure((size_t)nq_pad * ldq * (bf ? 2 : 4)))) return rc;
        q_out = idx->q_op.p;
    }
    if ((rc = idx->q_scale.ensure(sizeof(float) * (size_t)n_queries))) return rc;
    prep_queries_fast_kernel<<<((q_out ? nq_pad : n_queries) + 3) / 4, 128, 0, st>>>(queries_device, n_queries, dim, idx->inv_nq.as<float>(),
                                                                 idx->nq2f.as<float>(), nullptr, nullptr, nullptr, q_out, ldq, bf ? 1 : 0,
                                                                 idx->q_scale.as<float>(), q_out ? nq_pad : n_queries);
    KCHECK();
    const float *pen = idx->has_penalty ? idx->penalty.as<float>() : nullptr;
    if (path == MCP_PATH_STREAM_FP32) {
        if (dim & 3) return fail(MCP_EINVAL, "coarse_scores: dim %% 4 != 0");
        for (int p = 0; p < n_queries; p += 8) {
            int nv = std::min(8, n_queries - p);
            if ((rc = run_scan_pass(idx, st, queries_device + (int64_t)p * dim, dim, idx->inv_nq.as<float>() + p, nv,
                                    out_scores_device + (int64_t)p * idx->n_rows, idx->n_rows)))
                return rc;
        }
    } else if (is_gemm_path(path)) {
        u64 *l; int *c; int L, cap;
        if ((rc = gemm_tf32_run(idx->gemm, idx->gemm_ws, st, idx->n_sm, bf ? 1 : 0, bf ? (const void *)idx->shadow.p : (const void *)idx->d_rows,
                                idx->n_rows, dim, bf ? idx->ldh : idx->ld, q_out ? (const void *)q_out : (const void *)queries_device, ldq,
                                q_out ? nq_pad : n_queries, n_queries, idx->inv_nq.as<float>(), idx->nq2f.as<float>(), idx->inv_norm.as<float>(),
                                idx->norm2.as<double>(), pen, idx->metric, 64, &l, &c, &L, &cap, out_scores_device, nullptr,
                                bf ? idx->q_scale.as<float>() : nullptr, bf ? idx->shadow_scale.as<float>() : nullptr)))
            return fail(rc, "%s", idx->gemm.err);
    } else
        return fail(MCP_EINVAL, "coarse_scores: path must be STREAM_FP32, GEMM_TF32 or GEMM_F16");
    CU(cudaEventRecord(idx->done, st));
    idx->pending = true;
    return MCP_OK;
}

// ---------------------------------------------------------------------------------------------
// shard merge
// ---------------------------------------------------------------------------------------------
static int merge_launch(const double *scores64, const int32_t *ids, int64_t stride_s, int64_t stride_i, int n_shards, int n_queries,
                        int k, int32_t *out_ids, float *out_scores, double *out_scores64, void *stream) {
    if (!scores64 || !ids || !out_ids || !out_scores) return fail(MCP_EINVAL, "NULL buffer");
    if (n_shards < 1 || n_queries < 1 || k < 1 || k > MCP_MAX_K) return fail(MCP_EINVAL, "bad sizes");
    if ((int64_t)n_shards * k > MERGE_MAX) return fail(MCP_EINVAL, "n_shards * k = %d exceeds %d", n_shards * k, MERGE_MAX);
    int npow2 = pow2_at_least(std::max(n_shards * k, 2));
    size_t smem = (size_t)npow2 * (sizeof(u64) + sizeof(int));
    CU(cudaFuncSetAttribute(merge_shards_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    merge_shards_kernel<<<n_queries, MERGE_THREADS, smem, (cudaStream_t)stream>>>(
        scores64, ids, n_shards, stride_s, stride_i, n_queries, k, npow2, out_ids, out_scores, out_scores64);
    KCHECK();
    return MCP_OK;
}

extern "C" int mcp_merge_topk_device(const double *scores64_device, const int32_t *ids_device, int n_shards,
                                     int n_queries, int k, int32_t *out_ids_device, float *out_scores_device,
                                     double *out_scores64_device, void *stream) {
    const int64_t qk = (int64_t)n_queries * k;
    return merge_launch(scores64_device, ids_device, qk, qk, n_shards, n_queries, k, out_ids_device, out_scores_device,
                        out_scores64_device, stream);
}

extern "C" int64_t mcp_merge_packed_block_bytes(int n_queries, int k) {
    const int64_t qk = (int64_t)n_queries * k;
    return (qk * 12 + 15) & ~(int64_t)15;
}

extern "C" int mcp_merge_topk_packed_device(const void *packed_device, int n_shards, int n_queries, int k,
                                            int32_t *out_ids_device, float *out_scores_device, double *out_scores64_device,
                                            void *stream) {
    if (!packed_device) return fail(MCP_EINVAL, "NULL buffer");
    if (((uintptr_t)packed_device) & 15) return fail(MCP_EINVAL, "packed buffer must be 16-byte aligned");
    const int64_t qk = (int64_t)n_queries * k, block = mcp_merge_packed_block_bytes(n_queries, k);
    const unsigned char *base = static_cast<const unsigned char *>(packed_device);
    return merge_launch(reinterpret_cast<const double *>(base), reinterpret_cast<const int32_t *>(base + qk * 8), block / 8, block / 4,
                        n_shards, n_queries, k, out_ids_device, out_scores_device, out_scores64_device, stream);
}

// ---------------------------------------------------------------------------------------------
// DAG levels
// ---------------------------------------------------------------------------------------------
static std::atomic<int64_t> g_dag_launches[3];
extern "C" int64_t mcp_dag_launch_count(int which) { return which >= 0 && which < 3 ? g_dag_launches[which].load() : -1; }

static int dag_launch(int pitch, const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges, int64_t n_dags,
                      uint8_t *out_level, uint8_t *out_order, uint8_t *out_n_levels, uint8_t *out_status, cudaStream_t st) {
    if (!n_nodes || !edge_off || !out_level || !out_n_levels || !out_status) return fail(MCP_EINVAL, "NULL buffer");
    if (n_dags < 0) return fail(MCP_EINVAL, "n_dags < 0");
    if (n_dags == 0) return MCP_OK;
    if (((uintptr_t)edges) & 1) return fail(MCP_EINVAL, "edges must be 2-byte aligned");
    int dev = 0, n_sm = 148;
    CU(cudaGetDevice(&dev));
    CU(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    // Batches from KF_MIN_DAGS up go to the thread-per-DAG kernel (and whatever it hands back to the warp kernel); below that
    // there are not enough DAGs to fill the machine with threads.  MCP_DAG_KERNEL = warp | fifo overrides the choice, and
    // MCP_DAG_SUB the number of DAGs a CTA sorts at a time (measurement only).
    const char *env = getenv("MCP_DAG_KERNEL");
    const int dag_choice = !env ? 0 : !strcmp(env, "warp") ? 1 : !strcmp(env, "fifo") ? 2 : 0;
    const bool aligned16 = !((((uintptr_t)edges) | ((uintptr_t)out_level) | ((uintptr_t)out_order)) & 15);
    if (pitch == 64 && aligned16 && dag_choice != 1 && (dag_choice == 2 || (out_order && n_dags >= KF_MIN_DAGS))) {
        static std::mutex attr_mu;
        static int ctas_per_sm[NOALLOC_MAX_DEVICES_] = {};
        const int smem = (int)sizeof(KfSmem);
        if (dev < 0 || dev >= NOALLOC_MAX_DEVICES_) return fail(MCP_EINVAL, "device %d out of range", dev);
        {
            std::lock_guard<std::mutex> lk(attr_mu);
            if (!ctas_per_sm[dev]) {
                CU(cudaFuncSetAttribute(kahn_fifo_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
                CU(cudaFuncSetAttribute(kahn_fifo_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
                CU(cudaFuncSetAttribute(kahn_fifo_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
                CU(cudaFuncSetAttribute(kahn_fifo_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
                int occ = 0;
                CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kahn_fifo_kernel<true>, KF_THREADS, smem));
                ctas_per_sm[dev] = std::max(1, occ);
            }
        }
        // every CTA resident at once, each with an equal share of the batch, taken in equal pieces of at most KF_MAXC
        const int64_t slots = (int64_t)n_sm * ctas_per_sm[dev];
        const int grid = (int)std::min<int64_t>(slots, (n_dags + 31) / 32);
        const int64_t per_cta = (n_dags + grid - 1) / grid;
        int64_t max_sub = KF_MAXC;
        if (const char *e2 = getenv("MCP_DAG_SUB")) max_sub = std::min<int64_t>(KF_MAXC, std::max<int64_t>(32, atoi(e2)));
        const int64_t pieces = (per_cta + max_sub - 1) / max_sub;
        const int sub = (int)((((per_cta + pieces - 1) / pieces) + 31) & ~(int64_t)31);
        if (out_order)
            kahn_fifo_kernel<true><<<grid, KF_THREADS, smem, st>>>(n_nodes, edge_off, edges, n_dags, per_cta, sub, out_level, out_order, out_n_levels, out_status);
        else
            kahn_fifo_kernel<false><<<grid, KF_THREADS, smem, st>>>(n_nodes, edge_off, edges, n_dags, per_cta, sub, out_level, nullptr, out_n_levels, out_status);
        KCHECK();
        ++g_dag_launches[1];
        ++g_dag_launches[2];
        const int g2 = (int)std::min<int64_t>((int64_t)n_sm * 2, ((n_dags + 31) / 32 + KAHN_WARPS - 1) / KAHN_WARPS);
        if (out_order)
            kahn_levels_deferred_kernel<true><<<g2, KAHN_THREADS, 0, st>>>(n_nodes, edge_off, edges, n_dags, out_level, out_order, out_n_levels, out_status);
        else
            kahn_levels_deferred_kernel<false><<<g2, KAHN_THREADS, 0, st>>>(n_nodes, edge_off, edges, n_dags, out_level, nullptr, out_n_levels, out_status);
    } else if (pitch == 64) {
        ++g_dag_launches[0];
        int64_t want = (n_dags + KAHN_WARPS - 1) / KAHN_WARPS;
        int grid = (int)std::min<int64_t>((int64_t)n_sm * 8, want);
        if (out_order)
            kahn_levels_kernel<true><<<grid, KAHN_THREADS, 0, st>>>(n_nodes, edge_off, edges, n_dags, out_level, out_order, out_n_levels, out_status);
        else
            kahn_levels_kernel<false><<<grid, KAHN_THREADS, 0, st>>>(n_nodes, edge_off, edges, n_dags, out_level, nullptr, out_n_levels, out_status);
    } else {
        int grid = (int)std::min<int64_t>((int64_t)n_sm * 4, n_dags);
        if (out_order)
            kahn_levels_wide_kernel<true><<<grid, KAHNW_THREADS, 0, st>>>(n_nodes, edge_off, edges, n_dags, out_level, out_order, out_n_levels, out_status);
        else
            kahn_levels_wide_kernel<false><<<grid, KAHNW_THREADS, 0, st>>>(n_nodes, edge_off, edges, n_dags, out_level, nullptr, out_n_levels, out_status);
    }
    KCHECK();
    return MCP_OK;
}

extern "C" int mcp_dag_levels_device(const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges,
                                     int64_t n_dags, uint8_t *out_level, uint8_t *out_order, uint8_t *out_n_levels,
                                     uint8_t *out_status, void *stream) {
    return dag_launch(64, n_nodes, edge_off, edges, n_dags, out_level, out_order, out_n_levels, out_status, (cudaStream_t)stream);
}
extern "C" int mcp_dag_levels_wide_device(const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges,
                                          int64_t n_dags, uint8_t *out_level, uint8_t *out_order, uint8_t *out_n_levels,
                                          uint8_t *out_status, void *stream) {
    return dag_launch(256, n_nodes, edge_off, edges, n_dags, out_level, out_order, out_n_levels, out_status, (cudaStream_t)stream);
}

// The host variants keep one workspace per device: device buffers and a pinned staging buffer that only grow, and a
// non-blocking stream of their own.  A call therefore allocates nothing once the workspace has seen its size (Orchestrator.execute
// sends one small graph per request: cudaMalloc / cudaFree there would serialise against a retrieval scan running on another
// stream) and never touches the legacy default stream.  Calls on one device take turns (the lock is held for the call).
struct DagWorkspace {
    std::mutex mu;
    cudaStream_t st = nullptr, st2 = nullptr;
    Buf in, out;
    unsigned char *h = nullptr;
    size_t h_cap = 0;
    // large batches: pinned slots of the chunk pipeline (below)
    unsigned char *hp = nullptr;
    size_t hp_cap = 0;
    cudaEvent_t done[4] = {nullptr, nullptr, nullptr, nullptr};
};
static DagWorkspace g_dag_ws[NOALLOC_MAX_DEVICES_];
#define DAG_PINNED_MAX (8u << 20)      // larger batches are copied straight from / to the caller's memory, or piped in chunks
#define DAG_PIPE_CHUNK 131072          // DAGs per chunk of the pipeline (>= KF_MIN_DAGS: a full chunk takes the thread-per-DAG kernel)
#define DAG_PIPE_SLOTS 4               // two per lane
static_assert(DAG_PIPE_CHUNK >= KF_MIN_DAGS && DAG_PIPE_CHUNK % 16 == 0, "chunk size");

// Large host batches (two chunks or more): the caller's arrays are pageable, so a plain cudaMemcpy of 300 MB each way is
// what the call costs (1 M DAGs: 47 ms around a 1.2 ms kernel).  Here the batch goes through in chunks of DAG_PIPE_CHUNK
// DAGs in two lanes (even chunks / odd chunks), each with two pinned slots, a stream of its own and two threads: one copies
// chunk i into its slot and enqueues H2D, kernel, D2H, the other waits for chunk i's event and copies its results out; so
// the staging copies (two cores each way), both DMA directions and the kernels overlap.  A chunk keeps ABSOLUTE edge
// offsets: its edges sit in the slot at the same address modulo 16 as in the caller's array and the kernel gets the base
// pointer that makes edges[edge_off[b]] land on them (the thread-per-DAG kernel wants that base 16-byte aligned).
static int dag_levels_host_piped(DagWorkspace &w, int pitch, int device_id, const uint8_t *n_nodes, const uint32_t *edge_off,
                                 const uint8_t *edges, int64_t n_dags, uint8_t *out_level, uint8_t *out_order,
                                 uint8_t *out_n_levels, uint8_t *out_status) {
    const int64_t n_chunks = (n_dags + DAG_PIPE_CHUNK - 1) / DAG_PIPE_CHUNK;
    const int n_out = out_order ? 2 : 1;
    size_t in_cap = 0;
    for (int64_t c = 0; c < n_chunks; ++c) {
        const int64_t b0 = c * DAG_PIPE_CHUNK, b1 = std::min<int64_t>(n_dags, b0 + DAG_PIPE_CHUNK);
        in_cap = std::max<size_t>(in_cap, 2 * (size_t)(edge_off[b1] - edge_off[b0]));
    }
    const size_t off_bytes = ((sizeof(uint32_t) * (DAG_PIPE_CHUNK + 1)) + 15) & ~(size_t)15;
    const size_t ed_cap = (in_cap + 16 + 32 + 15) & ~(size_t)15;                  // shift (< 16) + the bytes + slack for the last aligned load
    const size_t in_slot = (off_bytes + ed_cap + DAG_PIPE_CHUNK + 255) & ~(size_t)255;
    const size_t out_slot = ((size_t)DAG_PIPE_CHUNK * pitch * n_out + 2 * (size_t)DAG_PIPE_CHUNK + 255) & ~(size_t)255;
    int rc;
    if ((rc = w.in.ensure(in_slot * DAG_PIPE_SLOTS))) return rc;
    if ((rc = w.out.ensure(out_slot * DAG_PIPE_SLOTS))) return rc;
    if ((in_slot + out_slot) * DAG_PIPE_SLOTS > w.hp_cap) {
        if (w.hp) cudaFreeHost(w.hp);
        w.hp = nullptr; w.hp_cap = 0;
        CU(cudaMallocHost(&w.hp, (in_slot + out_slot) * DAG_PIPE_SLOTS));
        w.hp_cap = (in_slot + out_slot) * DAG_PIPE_SLOTS;
    }
    if (!w.st2) CU(cudaStreamCreateWithFlags(&w.st2, cudaStreamNonBlocking));
    for (int i = 0; i < DAG_PIPE_SLOTS; ++i)
        if (!w.done[i]) CU(cudaEventCreateWithFlags(&w.done[i], cudaEventDisableTiming));
    unsigned char *h_in = w.hp, *h_out = w.hp + in_slot * DAG_PIPE_SLOTS;

    struct LaneResult { int rc = MCP_OK; std::string msg; };
    auto lane_fn = [&](const int lane, LaneResult &res) {
        cudaSetDevice(device_id);
        cudaStream_t st = lane ? w.st2 : w.st;
        const int64_t mine = (n_chunks - lane + 1) / 2;                       // chunks lane, lane + 2, ...
        std::mutex mu;
        std::condition_variable cv;
        int64_t enqueued = 0, drained = 0;   // of this lane's chunks: work on the stream / results in the caller's arrays
        bool failed = false;
        std::thread drain([&] {
            cudaSetDevice(device_id);
            for (int64_t k = 0; k < mine; ++k) {
                {
                    std::unique_lock<std::mutex> lk(mu);
                    cv.wait(lk, [&] { return enqueued > k || failed; });
                    if (enqueued <= k) return;                                // the producer gave up before this chunk
                }
                const int slot = lane * 2 + (int)(k & 1);
                const bool ok = cudaEventSynchronize(w.done[slot]) == cudaSuccess;
                if (ok) {
                    const int64_t b0 = (lane + 2 * k) * (int64_t)DAG_PIPE_CHUNK, cnt = std::min<int64_t>(n_dags, b0 + DAG_PIPE_CHUNK) - b0;
                    const unsigned char *ho = h_out + out_slot * slot;
                    const size_t lv = (size_t)cnt * pitch;
                    memcpy(out_level + (size_t)b0 * pitch, ho, lv);
                    if (out_order) memcpy(out_order + (size_t)b0 * pitch, ho + lv, lv);
                    memcpy(out_n_levels + b0, ho + lv * n_out, (size_t)cnt);
                    memcpy(out_status + b0, ho + lv * n_out + cnt, (size_t)cnt);
                }
                {
                    std::lock_guard<std::mutex> lk(mu);
                    if (!ok) failed = true;
                    drained = k + 1;
                }
                cv.notify_all();
            }
        });
        for (int64_t k = 0; k < mine && res.rc == MCP_OK; ++k) {
            const int slot = lane * 2 + (int)(k & 1);
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return drained + 2 > k || failed; });       // the slot's previous chunk has left it
                if (failed) break;
            }
            const int64_t b0 = (lane + 2 * k) * (int64_t)DAG_PIPE_CHUNK, cnt = std::min<int64_t>(n_dags, b0 + DAG_PIPE_CHUNK) - b0;
            const size_t e_first = edge_off[b0], e_bytes = 2 * (size_t)(edge_off[b0 + cnt] - e_first), shift = (2 * e_first) & 15;
            unsigned char *hi = h_in + in_slot * slot, *di = w.in.as<unsigned char>() + in_slot * slot, *d_o = w.out.as<unsigned char>() + out_slot * slot;
            memcpy(hi, edge_off + b0, sizeof(uint32_t) * (size_t)(cnt + 1));
            if (e_bytes) memcpy(hi + off_bytes + shift, edges + 2 * e_first, e_bytes);
            memcpy(hi + off_bytes + ed_cap, n_nodes + b0, (size_t)cnt);
            const size_t lv = (size_t)cnt * pitch;
            cudaError_t e = cudaMemcpyAsync(di, hi, off_bytes + ed_cap + (size_t)cnt, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) {
                // edges[edge_off[b]] for b in this chunk must land on di + off_bytes + shift + 2 * (edge_off[b] - e_first)
                const unsigned char *ed_base = reinterpret_cast<const unsigned char *>((uintptr_t)(di + off_bytes + shift) - 2 * (uintptr_t)e_first);
                res.rc = dag_launch(pitch, di + off_bytes + ed_cap, reinterpret_cast<const uint32_t *>(di), ed_base, cnt, d_o, out_order ? d_o + lv : nullptr,
                                    d_o + lv * n_out, d_o + lv * n_out + cnt, st);
                if (res.rc != MCP_OK) res.msg = mcp_last_error();
                if (res.rc == MCP_OK) e = cudaMemcpyAsync(h_out + out_slot * slot, d_o, lv * n_out + 2 * (size_t)cnt, cudaMemcpyDeviceToHost, st);
                if (res.rc == MCP_OK && e == cudaSuccess) e = cudaEventRecord(w.done[slot], st);
            }
            if (res.rc == MCP_OK && e != cudaSuccess) { res.rc = MCP_ECUDA; res.msg = std::string("DAG pipeline: ") + cudaGetErrorString(e); }
            {
                std::lock_guard<std::mutex> lk(mu);
                if (res.rc == MCP_OK) enqueued = k + 1;
                else failed = true;
            }
            cv.notify_all();
        }
        {
            std::lock_guard<std::mutex> lk(mu);
            if (res.rc != MCP_OK) failed = true;
        }
        cv.notify_all();
        drain.join();
        cudaStreamSynchronize(st);
        if (res.rc == MCP_OK && failed) { res.rc = MCP_ECUDA; res.msg = "DAG pipeline: a chunk failed on the device"; }
    };
    LaneResult r0, r1;
    std::thread other(lane_fn, 1, std::ref(r1));
    lane_fn(0, r0);
    other.join();
    if (r0.rc != MCP_OK) return fail(r0.rc, "%s", r0.msg.c_str());
    if (r1.rc != MCP_OK) return fail(r1.rc, "%s", r1.msg.c_str());
    return MCP_OK;
}

static int dag_levels_host(int pitch, int max_nodes, int device_id, const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges,
                           int64_t n_dags, uint8_t *out_level, uint8_t *out_order, uint8_t *out_n_levels, uint8_t *out_status) {
    if (!n_nodes || !edge_off || !out_level || !out_n_levels || !out_status) return fail(MCP_EINVAL, "NULL buffer");
    if (n_dags < 0) return fail(MCP_EINVAL, "n_dags < 0");
    if (n_dags == 0) return MCP_OK;
    if (device_id < 0 || device_id >= NOALLOC_MAX_DEVICES_) return fail(MCP_EINVAL, "device %d out of range", device_id);
    for (int64_t b = 0; b < n_dags; ++b) {
        if (n_nodes[b] > max_nodes) return fail(MCP_EINVAL, "DAG %lld has %d nodes (max %d)", (long long)b, n_nodes[b], max_nodes);
        if (edge_off[b + 1] < edge_off[b]) return fail(MCP_EINVAL, "edge_off not monotone at %lld", (long long)b);
    }
    const size_t n_edges = edge_off[n_dags];
    if (n_edges && !edges) return fail(MCP_EINVAL, "edges is NULL");
    CU(cudaSetDevice(device_id));
    DagWorkspace &w = g_dag_ws[device_id];
    std::lock_guard<std::mutex> lk(w.mu);
    if (!w.st) CU(cudaStreamCreateWithFlags(&w.st, cudaStreamNonBlocking));
    if (pitch == 64 && n_dags >= 2 * (int64_t)DAG_PIPE_CHUNK)
        return dag_levels_host_piped(w, pitch, device_id, n_nodes, edge_off, edges, n_dags, out_level, out_order, out_n_levels, out_status);
    // device layout in: [edge_off | edges | n_nodes], out: [level | order | n_levels | status]
    // (edges start on a 16-byte boundary: the thread-per-DAG kernel reads them 16 bytes at a time)
    const size_t B = (size_t)n_dags, off_len = sizeof(uint32_t) * (B + 1), off_bytes = (off_len + 15) & ~(size_t)15;
    const size_t ed_bytes = (2 * n_edges + 3) & ~(size_t)3;
    const size_t in_bytes = off_bytes + ed_bytes + B, lv_bytes = B * (size_t)pitch;
    const size_t out_bytes = lv_bytes * (out_order ? 2 : 1) + 2 * B;
    int rc;
    if ((rc = w.in.ensure(in_bytes + 16))) return rc;
    if ((rc = w.out.ensure(out_bytes + 16))) return rc;
    unsigned char *di = w.in.as<unsigned char>(), *d_o = w.out.as<unsigned char>();
    uint8_t *d_level = d_o, *d_order = out_order ? d_o + lv_bytes : nullptr, *d_nl = d_o + lv_bytes * (out_order ? 2 : 1), *d_st = d_nl + B;
    const bool staged = in_bytes + out_bytes <= DAG_PINNED_MAX;
    if (staged && in_bytes + out_bytes > w.h_cap) {
        if (w.h) cudaFreeHost(w.h);
        w.h = nullptr; w.h_cap = 0;
        const size_t want = std::max<size_t>(in_bytes + out_bytes, 64u << 10) * 2;
        CU(cudaMallocHost(&w.h, std::min<size_t>(want, DAG_PINNED_MAX)));
        w.h_cap = std::min<size_t>(want, DAG_PINNED_MAX);
    }
    if (staged) {
        memcpy(w.h, edge_off, off_len);
        if (n_edges) memcpy(w.h + off_bytes, edges, 2 * n_edges);
        memcpy(w.h + off_bytes + ed_bytes, n_nodes, B);
        CU(cudaMemcpyAsync(di, w.h, in_bytes, cudaMemcpyHostToDevice, w.st));
    } else {
        CU(cudaMemcpyAsync(di, edge_off, off_len, cudaMemcpyHostToDevice, w.st));
        if (n_edges) CU(cudaMemcpyAsync(di + off_bytes, edges, 2 * n_edges, cudaMemcpyHostToDevice, w.st));
        CU(cudaMemcpyAsync(di + off_bytes + ed_bytes, n_nodes, B, cudaMemcpyHostToDevice, w.st));
    }
    if ((rc = dag_launch(pitch, di + off_bytes + ed_bytes, reinterpret_cast<const uint32_t *>(di), di + off_bytes, n_dags, d_level, d_order, d_nl, d_st, w.st))) return rc;
    if (staged) {
        unsigned char *ho = w.h + in_bytes;
        CU(cudaMemcpyAsync(ho, d_o, out_bytes, cudaMemcpyDeviceToHost, w.st));
        CU(cudaStreamSynchronize(w.st));
        memcpy(out_level, ho, lv_bytes);
        if (out_order) memcpy(out_order, ho + lv_bytes, lv_bytes);
        memcpy(out_n_levels, ho + lv_bytes * (out_order ? 2 : 1), B);
        memcpy(out_status, ho + lv_bytes * (out_order ? 2 : 1) + B, B);
    } else {
        CU(cudaMemcpyAsync(out_level, d_level, lv_bytes, cudaMemcpyDeviceToHost, w.st));
        if (out_order) CU(cudaMemcpyAsync(out_order, d_order, lv_bytes, cudaMemcpyDeviceToHost, w.st));
        CU(cudaMemcpyAsync(out_n_levels, d_nl, B, cudaMemcpyDeviceToHost, w.st));
        CU(cudaMemcpyAsync(out_status, d_st, B, cudaMemcpyDeviceToHost, w.st));
        CU(cudaStreamSynchronize(w.st));
    }
    return MCP_OK;
}

extern "C" int mcp_dag_levels(int device_id, const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges,
                              int64_t n_dags, uint8_t *out_level, uint8_t *out_order, uint8_t *out_n_levels,
                              uint8_t *out_status) {
    return dag_levels_host(64, MCP_MAX_DAG_NODES, device_id, n_nodes, edge_off, edges, n_dags, out_level, out_order, out_n_levels, out_status);
}
extern "C" int mcp_dag_levels_wide(int device_id, const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges,
                                   int64_t n_dags, uint8_t *out_level, uint8_t *out_order, uint8_t *out_n_levels,
                                   uint8_t *out_status) {
    return dag_levels_host(256, MCP_MAX_WIDE_DAG_NODES, device_id, n_nodes, edge_off, edges, n_dags, out_level, out_order, out_n_levels, out_status);
}

// ---------------------------------------------------------------------------------------------
// Exact search that allocates nothing (no counterpart in the reference)
// ---------------------------------------------------------------------------------------------
// Who this is for.  The ordinary search entry points grow the index's workspace with cudaMalloc the first
// time a larger batch or table shows up.  Some callers of this library must not do that at request time:
//   * a planner thread answering /plan while another stream of the same process is in the middle of a
//     table scan -- cudaMalloc / cudaFree serialise against the whole device and would stall that scan;
//   * a process whose HBM is fully budgeted (a 61 GB registry shard next to other tenants, or a fixed
//     memory pool handed out at start-up), where an allocation in the request path can fail with
//     out-of-memory long after start-up checks passed.
// For them this entry point runs the fp64 full scan (the same kernels as path = MCP_PATH_EXACT_FP64) with
// every byte of scratch taken from one fixed array that is part of the library image.
//
// The array: MCP_NOALLOC_SCRATCH_BYTES = 8 MiB + 4 KiB of device memory, resident on every GPU that loads
// the library, from load until the process exits.  It bounds the call: n_rows <= MCP_NOALLOC_MAX_ROWS
// (1,048,576) and n_queries <= MCP_NOALLOC_MAX_QUERIES (64); larger requests are refused with MCP_EINVAL.
//
// Sharing.  There is ONE array per device and every call on that device uses it.  It still holds whatever the
// previous call left there; the call rewrites each part before reading it (query norms, the flag word and
// list, one score per table row), so stale contents are never consumed.  Two calls whose GPU work overlaps
// (different streams, or different indexes on the same device) race on the array and corrupt each other:
// keep such calls on one stream or otherwise ordered.
//
// Set-up.  `g_noalloc_scratch` names the array only inside device code; on the host the symbol is a stand-in,
// not an address the GPU can use.  The first call on each device therefore asks the runtime for the real
// device address (cudaGetSymbolAddress, with that device current) and remembers it.  Make that first call
// before starting any stream capture.
#define NOALLOC_SCORES_BYTES ((size_t)MCP_NOALLOC_MAX_ROWS * sizeof(u64))   // one fp64 score image per row
#define NOALLOC_TAIL_BYTES 4096                                               // norms, flags
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCORES_BYTES + NOALLOC_TAIL_BYTES];
static_assert(NOALLOC_SCORES_BYTES + NOALLOC_TAIL_BYTES == MCP_NOALLOC_SCRATCH_BYTES, "the header's scratch size and the array drifted apart");
static_assert(MCP_NOALLOC_MAX_QUERIES * (sizeof(double) + sizeof(float) + 1) <= 1024 && (8 + MCP_NOALLOC_MAX_QUERIES) * sizeof(int) <= NOALLOC_TAIL_BYTES - 1024,
              "the tail of the scratch array holds the query norms and the flag list");

#define NOALLOC_MAX_DEVICES NOALLOC_MAX_DEVICES_
static std::mutex g_noalloc_mu;
static unsigned char *g_noalloc_addr[NOALLOC_MAX_DEVICES] = {nullptr};

// one-time, per device: turn the symbol into a device address (device must be current)
static int noalloc_scratch_for(int device, unsigned char **out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(MCP_EINVAL, "device %d out of range", device);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_addr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(MCP_ECUDA, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d -> %s", device, cudaGetErrorString(e));
        g_noalloc_addr[device] = static_cast<unsigned char *>(p);
    }
    *out = g_noalloc_addr[device];
    return MCP_OK;
}

extern "C" int mcp_index_search_exact_noalloc_device(mcp_index *idx, const float *queries_device, int n_queries, int k,
                                                     int32_t *out_ids_device, float *out_scores_device,
                                                     double *out_scores64_device, void *stream) {
    if (check_idx(idx)) return MCP_EINVAL;
    NOT_ON_MULTI(idx, "search_exact_noalloc_device");
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->n_rows) return fail(MCP_ESTATE, "search before load");
    if (!queries_device || !out_ids_device || !out_scores_device) return fail(MCP_EINVAL, "NULL buffer");
    if (k < 1 || k > MCP_MAX_K) return fail(MCP_EINVAL, "k %d outside [1, %d]", k, MCP_MAX_K);
    if (n_queries < 1 || n_queries > MCP_NOALLOC_MAX_QUERIES)
        return fail(MCP_EINVAL, "n_queries %d outside [1, %d]: the fixed scratch array holds norms for at most %d queries",
                    n_queries, MCP_NOALLOC_MAX_QUERIES, MCP_NOALLOC_MAX_QUERIES);
    if (idx->n_rows > MCP_NOALLOC_MAX_ROWS)
        return fail(MCP_EINVAL, "table has %lld rows; the fixed scratch array holds one score per row for at most %d rows",
                    (long long)idx->n_rows, MCP_NOALLOC_MAX_ROWS);
    CU(cudaSetDevice(idx->device));
    unsigned char *base = nullptr;
    int rc = noalloc_scratch_for(idx->device, &base);
    if (rc) return rc;
    if ((rc = require_local(queries_device, sizeof(float) * (size_t)n_queries * idx->dim, idx->device, "search_exact_noalloc queries"))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // carve the array: [scores | nq2 (64 f64) | inv_nq (64 f32) | qflag (64 B) | n_flag + flag list (ints)]
    u64 *scores = reinterpret_cast<u64 *>(base);
    unsigned char *tail = base + NOALLOC_SCORES_BYTES;
    double *nq2 = reinterpret_cast<double *>(tail);
    float *inv_nq = reinterpret_cast<float *>(tail + 512);
    unsigned char *qflag = tail + 768;
    int *n_flag = reinterpret_cast<int *>(tail + 1024);
    int *flag_list = n_flag + 8;
    const int dim = idx->dim;
    const float *pen = idx->has_penalty ? idx->penalty.as<float>() : nullptr;
    // every query's norm is rewritten here; nothing from an earlier call is read
    prep_rows_kernel<true><<<1, PREP_ROWS, 0, st>>>(queries_device, n_queries, dim, dim, nq2, inv_nq, qflag, nullptr);
    KCHECK();
    const int gx = scan_grid(idx);
    CUtensorMap tm;                                                        // host-side encode of the table's TMA view: no allocation
    if ((rc = scan_tensor_map(idx, &tm))) return rc;
    for (int q = 0; q < n_queries; ++q) {
        flag_all_kernel<<<1, 32, 0, st>>>(flag_list, n_flag, q, 1);        // (re)sets the flag word and the one-entry list
        KCHECK();
        exact_scan2_kernel<<<gx, XS_THREADS, XS_SMEM_BYTES(dim, 1), st>>>(tm, idx->n_rows, dim, queries_device,
                                                                         nq2, idx->norm2.as<double>(), pen, idx->metric, flag_list,
                                                                         n_flag, 0, 1, scores, XsSparse{});   // writes scores[0 .. n_rows)
        KCHECK();
        exact_select_kernel<<<1, EXSEL_THREADS, 0, st>>>(scores, idx->n_rows, k, idx->id_base, flag_list, n_flag, 0,
                                                          out_ids_device, out_scores_device, out_scores64_device);
        KCHECK();
    }
    return MCP_OK;
}

#include "multi.cuh"
