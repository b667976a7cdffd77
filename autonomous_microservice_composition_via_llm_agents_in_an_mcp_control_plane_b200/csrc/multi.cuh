// multi.cuh -- one handle over a registry row-sharded across the GPUs of a box, inside ONE process.
//
// The reference is a single process with module-level singletons (control_plane.py:135-138) whose planner makes one plain
// method call per request (:57-75); a drop-in for it cannot ask the caller to run one process per GPU.  A handle made by
// mcp_index_create_multi owns one ordinary single-device index per shard (contiguous row ranges) and
// one worker thread per shard.  A host-buffer search fans out: every worker copies the queries to its GPU, runs the same
// certified local search as a single-device index (so every per-shard list is the exact local top-k), and pushes its packed
// (score fp64, id) block into a gather buffer on the first shard's GPU with a peer copy (copy engine, device to device);
// the caller's thread then runs K5 (merge_shards_kernel) there and copies the merged answer out.  Kernels and TMA tensor
// maps only ever read their own GPU's memory: a shard's rows, norms, shadow and workspace all live on the shard's device.
#pragma once

struct MultiWorker {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<int()> job;
    bool has_job = false, done = false, quit = false;
    int rc = 0;
    std::string err;
    MultiWorker() { th = std::thread([this] { run(); }); }
    ~MultiWorker() {
        { std::lock_guard<std::mutex> lk(m); quit = true; }
        cv.notify_all();
        if (th.joinable()) th.join();
    }
    void run() {
        for (;;) {
            std::function<int()> f;
            {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [this] { return has_job || quit; });
                if (quit) return;
                f = std::move(job);
                has_job = false;
            }
            int r = f();
            {
                std::lock_guard<std::mutex> lk(m);
                rc = r;
                err = r ? g_err : std::string();          // the message lives in this thread's g_err: hand it over
                done = true;
            }
            cv.notify_all();
        }
    }
    void submit(std::function<int()> f) {
        { std::lock_guard<std::mutex> lk(m); job = std::move(f); has_job = true; done = false; }
        cv.notify_all();
    }
    int wait() {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [this] { return done; });
        return rc;
    }
};

struct mcp_multi {
    std::vector<mcp_index *> sub;                    // one single-device index per shard
    std::vector<int64_t> lo;                         // shard g holds rows [lo[g], lo[g+1]) of the table
    std::vector<std::unique_ptr<MultiWorker>> workers;
    std::vector<Buf> block;                          // per shard, on its device: the packed (score, id) block + fp32 scores behind it
    int active = 0;                                  // shards that hold rows (a table shorter than the shard count leaves some empty)
    Buf gather, merged;                              // on sub[0]'s device
    unsigned char *h_out = nullptr;                  // pinned staging for the merged answer
    size_t h_out_cap = 0;
    cudaStream_t stream0 = nullptr;
};

// run f(g) on every active shard's worker, wait for all, report the first failure
static int multi_fanout(mcp_multi *m, int n, const std::function<int(int)> &f) {
    for (int g = 0; g < n; ++g) m->workers[g]->submit([g, &f] { return f(g); });
    int rc = MCP_OK;
    std::string msg;
    for (int g = 0; g < n; ++g) {
        const int r = m->workers[g]->wait();
        if (r && !rc) { rc = r; msg = m->workers[g]->err; }
    }
    return rc ? fail(rc, "shard: %s", msg.c_str()) : MCP_OK;
}

extern "C" int mcp_index_create_multi(mcp_index **out, const int *device_ids, int n_devices, int dim, int metric) {
    if (!out || !device_ids) return fail(MCP_EINVAL, "NULL argument");
    if (n_devices < 1 || n_devices > 16) return fail(MCP_EINVAL, "n_devices %d outside [1, 16]", n_devices);
    mcp_index *idx = new mcp_index();
    mcp_multi *m = new mcp_multi();
    idx->multi = m;
    idx->dim = dim; idx->metric = metric; idx->device = device_ids[0];
    for (int g = 0; g < n_devices; ++g) {
        mcp_index *s = nullptr;
        int rc = mcp_index_create(&s, device_ids[g], dim, metric);
        if (rc) { std::string keep = g_err; mcp_index_destroy(idx); return fail(rc, "shard %d on device %d: %s", g, device_ids[g], keep.c_str()); }
        m->sub.push_back(s);
        m->workers.emplace_back(new MultiWorker());
    }
    m->block.resize(n_devices);
    m->lo.assign(n_devices + 1, 0);
    cudaError_t e = cudaSetDevice(device_ids[0]);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&m->stream0, cudaStreamNonBlocking);
    if (e != cudaSuccess) { mcp_index_destroy(idx); return fail(MCP_ECUDA, "stream on device %d: %s", device_ids[0], cudaGetErrorString(e)); }
    *out = idx;
    return MCP_OK;
}

static int multi_destroy(mcp_index *idx) {
    mcp_multi *m = idx->multi;
    m->workers.clear();                              // joins the threads
    for (size_t g = 0; g < m->sub.size(); ++g) {
        cudaSetDevice(m->sub[g]->device);
        if (g < m->block.size()) m->block[g].release();
        mcp_index_destroy(m->sub[g]);
    }
    cudaSetDevice(idx->device);
    m->gather.release(); m->merged.release();
    if (m->h_out) cudaFreeHost(m->h_out);
    if (m->stream0) cudaStreamDestroy(m->stream0);
    delete m;
    delete idx;
    return MCP_OK;
}

static int multi_load(mcp_index *idx, const float *rows_host, int64_t n_rows, int64_t id_base) {
    mcp_multi *m = idx->multi;
    if (!rows_host || n_rows < 1) return fail(MCP_EINVAL, "rows NULL or n_rows < 1");
    std::lock_guard<std::mutex> lk(idx->mu);
    const int S = (int)m->sub.size();
    const int64_t per = (n_rows + S - 1) / S;
    m->active = 0;
    for (int g = 0; g <= S; ++g) m->lo[g] = std::min<int64_t>((int64_t)g * per, n_rows);
    for (int g = 0; g < S; ++g) if (m->lo[g + 1] > m->lo[g]) m->active = g + 1;
    const int dim = idx->dim;
    int rc = multi_fanout(m, m->active, [&](int g) {
        return mcp_index_load(m->sub[g], rows_host + m->lo[g] * dim, m->lo[g + 1] - m->lo[g], id_base + m->lo[g]);
    });
    if (rc) { m->active = 0; idx->n_rows = 0; return rc; }
    idx->n_rows = n_rows; idx->id_base = id_base;
    return MCP_OK;
}

// every shard reads its own row range of the file, all at once
static int multi_load_file(mcp_index *idx, const char *path, int64_t first_row, int64_t n_rows) {
    mcp_multi *m = idx->multi;
    std::lock_guard<std::mutex> lk(idx->mu);
    TableFileHeader h;
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(MCP_EINVAL, "cannot open %s", path);
    const bool ok = pread(fd, &h, sizeof h, 0) == (ssize_t)sizeof h;
    close(fd);
    if (!ok) return fail(MCP_EINVAL, "%s is not an MCPTBL01 table file", path);
    if (n_rows < 0) n_rows = h.n_rows - first_row;
    if (first_row < 0 || n_rows < 1 || first_row + n_rows > h.n_rows) return fail(MCP_EINVAL, "rows [%lld, %lld) outside the file's %lld rows", (long long)first_row, (long long)(first_row + n_rows), (long long)h.n_rows);
    const int S = (int)m->sub.size();
    const int64_t per = (n_rows + S - 1) / S;
    m->active = 0;
    for (int g = 0; g <= S; ++g) m->lo[g] = std::min<int64_t>((int64_t)g * per, n_rows);
    for (int g = 0; g < S; ++g) if (m->lo[g + 1] > m->lo[g]) m->active = g + 1;
    int rc = multi_fanout(m, m->active, [&](int g) { return mcp_index_load_file(m->sub[g], path, first_row + m->lo[g], m->lo[g + 1] - m->lo[g]); });
    if (rc) { m->active = 0; idx->n_rows = 0; return rc; }
    idx->n_rows = n_rows; idx->id_base = h.id_base + first_row;
    return MCP_OK;
}

static int multi_set_penalty(mcp_index *idx, const float *penalty_host) {
    mcp_multi *m = idx->multi;
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->n_rows) return fail(MCP_ESTATE, "set_penalty before load");
    return multi_fanout(m, m->active, [&](int g) { return mcp_index_set_penalty(m->sub[g], penalty_host ? penalty_host + m->lo[g] : nullptr); });
}

static int multi_update_rows(mcp_index *idx, int64_t first_row, const float *rows_host, int64_t n_rows) {
    mcp_multi *m = idx->multi;
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->n_rows) return fail(MCP_ESTATE, "update_rows before load");
    if (!rows_host) return fail(MCP_EINVAL, "a multi-device table is library memory: pass the new rows");
    if (first_row < 0 || n_rows < 1 || first_row + n_rows > idx->n_rows) return fail(MCP_EINVAL, "row range [%lld, %lld) outside the table", (long long)first_row, (long long)(first_row + n_rows));
    const int dim = idx->dim;
    return multi_fanout(m, m->active, [&](int g) {
        const int64_t a = std::max(first_row, m->lo[g]), b = std::min(first_row + n_rows, m->lo[g + 1]);
        if (a >= b) return (int)MCP_OK;
        return mcp_index_update_rows(m->sub[g], a - m->lo[g], rows_host + (a - first_row) * dim, b - a);
    });
}

static int multi_upsert_rows(mcp_index *idx, const int64_t *row_ids, const float *rows_host, int64_t n_rows) {
    mcp_multi *m = idx->multi;
    std::lock_guard<std::mutex> lk(idx->mu);
    if (!idx->n_rows) return fail(MCP_ESTATE, "upsert before load");
    if (!rows_host || n_rows < 1) return fail(MCP_EINVAL, "rows NULL or n < 1");
    const int dim = idx->dim;
    for (int64_t i = 0; i < n_rows; ++i)
        if (row_ids[i] < 0 || row_ids[i] >= idx->n_rows) return fail(MCP_EINVAL, "row %lld outside the table of %lld rows", (long long)row_ids[i], (long long)idx->n_rows);
    // split by owner: every shard gets its rows gathered into a contiguous host block with shard-local ids
    return multi_fanout(m, m->active, [&](int g) {
        std::vector<int64_t> ids;
        std::vector<float> rows;
        for (int64_t i = 0; i < n_rows; ++i)
            if (row_ids[i] >= m->lo[g] && row_ids[i] < m->lo[g + 1]) {
                ids.push_back(row_ids[i] - m->lo[g]);
                rows.insert(rows.end(), rows_host + i * dim, rows_host + (i + 1) * dim);
            }
        if (ids.empty()) return (int)MCP_OK;
        return mcp_index_upsert_rows(m->sub[g], ids.data(), rows.data(), (int64_t)ids.size());
    });
}

static int multi_set_option(mcp_index *idx, const char *name, int64_t value) {
    for (mcp_index *s : idx->multi->sub) { int rc = mcp_index_set_option(s, name, value); if (rc) return rc; }
    return MCP_OK;
}

static int multi_get_stats(mcp_index *idx, mcp_stats *out) {
    mcp_stats t{};
    bool first = true;
    for (mcp_index *s : idx->multi->sub) {
        mcp_stats x;
        int rc = mcp_index_get_stats(s, &x);
        if (rc) return rc;
        t.rows_scanned += x.rows_scanned; t.bytes_scanned += x.bytes_scanned; t.candidates_rescored += x.candidates_rescored;
        t.certificate_failures += x.certificate_failures; t.exact_scans += x.exact_scans; t.kernel_launches += x.kernel_launches;
        t.coarse_timed += x.coarse_timed; t.coarse_ms += x.coarse_ms;
        if (first) { t.searches = x.searches; t.queries = x.queries; t.last_path = x.last_path; t.last_kprime = x.last_kprime; first = false; }
    }
    *out = t;
    return MCP_OK;
}

// one shard's part of a host-buffer search, on its worker thread: certified local top-k -> packed block -> gather buffer
static int multi_shard_search(mcp_multi *m, int g, const float *queries_host, int nq, int k, unsigned char *gather0, int dev0, int64_t blk) {
    mcp_index *s = m->sub[g];
    std::lock_guard<std::mutex> lk(s->mu);
    CU(cudaSetDevice(s->device));
    cudaStream_t st = s->stream;
    int rc;
    const size_t qk = (size_t)nq * k;
    if ((rc = s->q_copy.ensure(sizeof(float) * (size_t)nq * s->dim))) return rc;
    if ((rc = m->block[g].ensure((size_t)blk + sizeof(float) * qk))) return rc;
    unsigned char *b = m->block[g].as<unsigned char>();
    double *d_s64 = reinterpret_cast<double *>(b);
    int32_t *d_ids = reinterpret_cast<int32_t *>(b + qk * 8);
    float *d_sc = reinterpret_cast<float *>(b + blk);
    if (s->pending) { CU(cudaEventSynchronize(s->done)); s->pending = false; }
    s->h_status[0] = s->h_status[1] = 0;
    s->last_stream = st;
    CU(cudaMemcpyAsync(s->q_copy.p, queries_host, sizeof(float) * (size_t)nq * s->dim, cudaMemcpyHostToDevice, st));
    if ((rc = search_locked(s, s->q_copy.as<float>(), nq, k, d_ids, d_sc, d_s64, st))) return rc;
    CU(cudaMemsetAsync(s->flags.as<int>() + 1, 0, 2 * sizeof(int), st));
    CU(cudaStreamSynchronize(st));
    const int hs[2] = {s->h_status[0], s->h_status[1]};
    s->h_status[0] = s->h_status[1] = 0;
    s->stats.certificate_failures += hs[0];
    s->stats.exact_scans += hs[0] - hs[1];
    note_outcome(s, hs[0]);
    if (hs[1] > 0 && (rc = resolve_leftovers(s, s->q_copy.as<float>(), nq, k, d_ids, d_sc, d_s64, st))) return rc;
    CU(cudaMemcpyPeerAsync(gather0 + (size_t)g * blk, dev0, b, s->device, (size_t)blk, st));
    CU(cudaStreamSynchronize(st));
    return MCP_OK;
}

static int multi_search(mcp_index *idx, const float *queries_host, int nq, int k, int32_t *out_ids, float *out_scores, double *out_scores64) {
    mcp_multi *m = idx->multi;
    std::lock_guard<std::mutex> lk(idx->mu);
    int rc = validate_search(idx, queries_host, nq, k, out_ids, out_scores);
    if (rc) return rc;
    if ((int64_t)m->active * k > MERGE_MAX) return fail(MCP_EINVAL, "shards * k = %d exceeds %d", m->active * k, MERGE_MAX);
    const int64_t blk = mcp_merge_packed_block_bytes(nq, k);
    const size_t nk = (size_t)nq * k;
    const size_t off_ids = sizeof(double) * nk, off_sc = off_ids + sizeof(int32_t) * nk, total = off_sc + sizeof(float) * nk;
    CU(cudaSetDevice(idx->device));
    if ((rc = m->gather.ensure((size_t)blk * m->active))) return rc;
    if ((rc = m->merged.ensure(total))) return rc;
    if (total > m->h_out_cap) {
        if (m->h_out) cudaFreeHost(m->h_out);
        m->h_out = nullptr; m->h_out_cap = 0;
        CU(cudaMallocHost(&m->h_out, total + total / 4));
        m->h_out_cap = total + total / 4;
    }
    unsigned char *gather0 = m->gather.as<unsigned char>();
    const int dev0 = idx->device;
    if ((rc = multi_fanout(m, m->active, [&](int g) { return multi_shard_search(m, g, queries_host, nq, k, gather0, dev0, blk); }))) return rc;
    CU(cudaSetDevice(idx->device));
    unsigned char *db = m->merged.as<unsigned char>();
    rc = merge_launch(reinterpret_cast<const double *>(gather0), reinterpret_cast<const int32_t *>(gather0 + nk * 8), blk / 8, blk / 4, m->active, nq, k,
                      reinterpret_cast<int32_t *>(db + off_ids), reinterpret_cast<float *>(db + off_sc), reinterpret_cast<double *>(db), m->stream0);
    if (rc) return rc;
    CU(cudaMemcpyAsync(m->h_out, db, total, cudaMemcpyDeviceToHost, m->stream0));
    CU(cudaStreamSynchronize(m->stream0));
    memcpy(out_ids, m->h_out + off_ids, sizeof(int32_t) * nk);
    memcpy(out_scores, m->h_out + off_sc, sizeof(float) * nk);
    if (out_scores64) memcpy(out_scores64, m->h_out, sizeof(double) * nk);
    return MCP_OK;
}
