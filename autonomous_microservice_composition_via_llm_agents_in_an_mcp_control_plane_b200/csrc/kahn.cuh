// kahn.cuh -- K6: batched, bit-exact Kahn level scheduling (integer only).
//
// Stands for nx.topological_generations (networkx dag.py:218-240) on the DiGraph that
// control_plane.py:94-100 builds.  One warp per DAG (<= 64 nodes): lane l owns nodes l and l+32.
//   build   pred[v] = 64-bit mask of v's distinct predecessors (atomicOr on its 32-bit half: a duplicate edge sets
//           the same bit again, which is exactly DiGraph's "same edge collapses"; a self loop
//           sets bit v of pred[v], so v can never become ready -> cycle)
//   levels  generation L = { v not done : pred[v] & ~done == 0 }, found with two ballots; no
//           per-node counters, no queues.
//   order   (optional) the order nx.topological_sort yields: generation 0 in node order, later
//           generations by (position of the parent whose decrement zeroed the node, index of
//           the first edge parent->node), which is the order dag.py:224-234 appends them in.
// HBM traffic per DAG: 1 + 8 + 2E bytes in, 64 (+64) + 2 bytes out; everything else is
// shared memory and registers.
// Three kernels, identical bytes out: kahn_levels_kernel (a warp per DAG, any batch), kahn_fifo_kernel
// (a thread per DAG, large batches with the order; further down) and kahn_levels_wide_kernel (65..255 nodes).
#pragma once
#include "common.cuh"

#define KAHN_THREADS 256
#define KAHN_WARPS (KAHN_THREADS / 32)

#define KAHN_FE 4
struct KahnWarp {
    u64 pred[64];
    u64 fm[64];            // frontier mask of each generation
    u64 key[64];
    uint32_t firste[64];
    uint32_t fe[64][KAHN_FE];  // first edge index from each of a node's (up to KAHN_FE) parents in the previous generation
    unsigned char lvl[64], pos[64], ustar[64], order[64];
    int fe_overflow;           // some node has more such parents than the table holds: this DAG rescans its edges per generation
};

// One DAG, one warp.
template <bool WITH_ORDER>
__device__ __forceinline__ void kahn_warp_dag(KahnWarp &w, const int lane, const int64_t b,
                                              const unsigned char *__restrict__ n_nodes, const uint32_t *__restrict__ edge_off,
                                              const uchar2 *__restrict__ ed,
                                              unsigned char *__restrict__ out_level, unsigned char *__restrict__ out_order,
                                              unsigned char *__restrict__ out_n_levels, unsigned char *__restrict__ out_status) {
    {
        const int n = n_nodes[b];
        if (n > 64) {                                    // not this kernel's DAG (kahn_levels_wide_kernel takes up to 255 nodes): say so
            out_level[b * 64 + lane] = 0xFF; out_level[b * 64 + 32 + lane] = 0xFF;
            if (WITH_ORDER) { out_order[b * 64 + lane] = 0xFF; out_order[b * 64 + 32 + lane] = 0xFF; }
            if (lane == 0) { out_n_levels[b] = 0; out_status[b] = 2; }
            return;
        }
        const uint32_t e0 = edge_off[b], e1 = edge_off[b + 1];
        w.pred[lane] = 0; w.pred[lane + 32] = 0;
        __syncwarp();
        for (uint32_t e = e0 + lane; e < e1; e += 32) {
            uchar2 uv = ed[e];
            // (a 32-bit OR on the half that holds the bit: native; the 64-bit form compiles to a compare-and-swap loop)
            if (uv.x < n && uv.y < n) atomicOr(reinterpret_cast<unsigned *>(&w.pred[uv.y]) + (uv.x >> 5), 1u << (uv.x & 31));
        }
        __syncwarp();
        const u64 p0 = w.pred[lane], p1 = w.pred[lane + 32];
        const bool ok0 = lane < n, ok1 = lane + 32 < n;
        u64 done = 0;
        int level = 0;
        unsigned char l0 = 0xFF, l1 = 0xFF;
        if (n <= 32) {                                   // (warp-uniform) every node and every predecessor bit in the low half
            const unsigned q0 = (unsigned)p0;
            unsigned done32 = 0;
            for (;;) {
                const bool r0 = ok0 && !((done32 >> lane) & 1u) && ((q0 & ~done32) == 0u);
                const unsigned f = __ballot_sync(MCP_FULL, r0);
                if (f == 0u) break;
                if (r0) l0 = (unsigned char)level;
                if (WITH_ORDER && lane == 0) w.fm[level] = (u64)f;
                done32 |= f;
                ++level;
            }
            done = (u64)done32;
        } else
        for (;;) {
            bool r0 = ok0 && !((done >> lane) & 1ull) && ((p0 & ~done) == 0ull);
            bool r1 = ok1 && !((done >> (lane + 32)) & 1ull) && ((p1 & ~done) == 0ull);
            u64 f = (u64)__ballot_sync(MCP_FULL, r0) | ((u64)__ballot_sync(MCP_FULL, r1) << 32);
            if (f == 0ull) break;
            if (r0) l0 = (unsigned char)level;
            if (r1) l1 = (unsigned char)level;
            if (WITH_ORDER && lane == 0) w.fm[level] = f;
            done |= f;
            ++level;
        }
        out_level[b * 64 + lane] = l0;
        out_level[b * 64 + 32 + lane] = l1;
        if (lane == 0) {
            out_n_levels[b] = (unsigned char)level;
            out_status[b] = (unsigned char)(__popcll(done) != n);
        }
        if (WITH_ORDER) {
            w.lvl[lane] = l0; w.lvl[lane + 32] = l1;
            w.order[lane] = 0xFF; w.order[lane + 32] = 0xFF;
#pragma unroll
            for (int j = 0; j < 2 * KAHN_FE; ++j) (&w.fe[0][0])[j * 32 + lane] = 0xFFFFFFFFu;
            if (lane == 0) w.fe_overflow = 0;
            __syncwarp();
            // One pass over the edges for the tie-break among siblings: the order of a generation needs, per node, the index
            // of the first edge from the parent that released it -- one of its parents in the previous generation.  Which one
            // is only known once that generation is ordered, so the first edge from EACH such parent is recorded here
            // (slot = rank of the parent among them), instead of rescanning the edge list generation after generation.
            for (uint32_t e = e0 + lane; e < e1; e += 32) {
                const uchar2 uv = ed[e];
                if (uv.x < n && uv.y < n) {
                    const int lv = w.lvl[uv.y];
                    if (lv != 0xFF && lv > 0 && w.lvl[uv.x] == lv - 1) {
                        const u64 cand = w.pred[uv.y] & w.fm[lv - 1];
                        const int slot = __popcll(cand & ((1ull << uv.x) - 1ull));
                        if (slot < KAHN_FE) atomicMin(&w.fe[uv.y][slot], e - e0);
                        else w.fe_overflow = 1;
                    }
                }
            }
            __syncwarp();
            const bool rescan = w.fe_overflow != 0;
            int base = 0;
            for (int L = 0; L < level; ++L) {
                const u64 f = w.fm[L];
                const bool in0 = (f >> lane) & 1ull, in1 = (f >> (lane + 32)) & 1ull;
                int rk0 = 0, rk1 = 0;
                if (L == 0) {
                    // generation 0 comes in node order: the rank is a population count
                    rk0 = __popcll(f & ((1ull << lane) - 1ull));
                    rk1 = __popcll(f & ((1ull << (lane + 32)) - 1ull));
                } else if ((f & (f - 1ull)) != 0ull) {           // a generation of one node needs no ordering: rank 0
                    const u64 prev = w.fm[L - 1];
                    int bp0 = -1, bp1 = -1, bu0 = 0, bu1 = 0;
                    if (in0) {
                        u64 c = p0 & prev;
                        while (c) { int u = __ffsll((long long)c) - 1; c &= c - 1; int p = w.pos[u]; if (p > bp0) { bp0 = p; bu0 = u; } }
                    }
                    if (in1) {
                        u64 c = p1 & prev;
                        while (c) { int u = __ffsll((long long)c) - 1; c &= c - 1; int p = w.pos[u]; if (p > bp1) { bp1 = p; bu1 = u; } }
                    }
                    // Positions of the zeroing parents as a bit set.  When no two nodes of the generation share a parent
                    // (most generations) the position alone orders them -- the rank is a population count and the edge list
                    // need not be consulted; the first-edge index only breaks ties between siblings.
                    const u64 m0 = in0 ? 1ull << bp0 : 0ull, m1 = in1 ? 1ull << bp1 : 0ull;
                    const u64 mm = m0 | m1;
                    const u64 BP = (u64)__reduce_or_sync(MCP_FULL, (unsigned)mm) | ((u64)__reduce_or_sync(MCP_FULL, (unsigned)(mm >> 32)) << 32);
                    if (__popcll(BP) == __popcll(f)) {
                        rk0 = in0 ? __popcll(BP & (m0 - 1ull)) : 0;
                        rk1 = in1 ? __popcll(BP & (m1 - 1ull)) : 0;
                    } else {
                        uint32_t fe0 = 0, fe1 = 0;
                        if (rescan) {
                            if (in0) { w.ustar[lane] = (unsigned char)bu0; w.firste[lane] = 0xFFFFFFFFu; }
                            if (in1) { w.ustar[lane + 32] = (unsigned char)bu1; w.firste[lane + 32] = 0xFFFFFFFFu; }
                            __syncwarp();
                            for (uint32_t e = e0 + lane; e < e1; e += 32) {
                                uchar2 uv = ed[e];
                                if (uv.x < n && uv.y < n && w.lvl[uv.y] == L && w.ustar[uv.y] == uv.x)
                                    atomicMin(&w.firste[uv.y], e - e0);
                            }
                            __syncwarp();
                            if (in0) fe0 = w.firste[lane];
                            if (in1) fe1 = w.firste[lane + 32];
                        } else {
                            if (in0) fe0 = w.fe[lane][__popcll((p0 & prev) & ((1ull << bu0) - 1ull))];
                            if (in1) fe1 = w.fe[lane + 32][__popcll((p1 & prev) & ((1ull << bu1) - 1ull))];
                        }
                        u64 k0 = 0, k1 = 0;
                        if (in0) { k0 = ((u64)bp0 << 32) | fe0; w.key[lane] = k0; }
                        if (in1) { k1 = ((u64)bp1 << 32) | fe1; w.key[lane + 32] = k1; }
                        __syncwarp();
                        for (u64 m = f; m; m &= m - 1) {
                            u64 kx = w.key[__ffsll((long long)m) - 1];
                            rk0 += kx < k0; rk1 += kx < k1;
                        }
                    }
                }
                if (in0) { w.pos[lane] = (unsigned char)(base + rk0); w.order[base + rk0] = (unsigned char)lane; }
                if (in1) { w.pos[lane + 32] = (unsigned char)(base + rk1); w.order[base + rk1] = (unsigned char)(lane + 32); }
                __syncwarp();
                base += __popcll(f);
            }
            out_order[b * 64 + lane] = w.order[lane];
            out_order[b * 64 + 32 + lane] = w.order[lane + 32];
            __syncwarp();
        }
    }
}

template <bool WITH_ORDER>
__global__ void __launch_bounds__(KAHN_THREADS)
kahn_levels_kernel(const unsigned char *__restrict__ n_nodes, const uint32_t *__restrict__ edge_off,
                   const unsigned char *__restrict__ edges, int64_t n_dags,
                   unsigned char *__restrict__ out_level, unsigned char *__restrict__ out_order,
                   unsigned char *__restrict__ out_n_levels, unsigned char *__restrict__ out_status) {
    __shared__ KahnWarp sw[KAHN_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uchar2 *ed = reinterpret_cast<const uchar2 *>(edges);
    for (int64_t b = (int64_t)blockIdx.x * KAHN_WARPS + warp; b < n_dags; b += (int64_t)gridDim.x * KAHN_WARPS)
        kahn_warp_dag<WITH_ORDER>(sw[warp], lane, b, n_nodes, edge_off, ed, out_level, out_order, out_n_levels, out_status);
}

// The DAGs kahn_fifo_kernel (below) handed back (status KAHN_DEFERRED): found by reading the status bytes 32 at a time.
#define KAHN_DEFERRED 3
template <bool WITH_ORDER>
__global__ void __launch_bounds__(KAHN_THREADS)
kahn_levels_deferred_kernel(const unsigned char *__restrict__ n_nodes, const uint32_t *__restrict__ edge_off,
                            const unsigned char *__restrict__ edges, int64_t n_dags,
                            unsigned char *__restrict__ out_level, unsigned char *__restrict__ out_order,
                            unsigned char *__restrict__ out_n_levels, unsigned char *__restrict__ out_status) {
    __shared__ KahnWarp sw[KAHN_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uchar2 *ed = reinterpret_cast<const uchar2 *>(edges);
    const int64_t n_groups = (n_dags + 31) / 32;
    for (int64_t g = (int64_t)blockIdx.x * KAHN_WARPS + warp; g < n_groups; g += (int64_t)gridDim.x * KAHN_WARPS) {
        const int64_t b = g * 32 + lane;
        unsigned m = __ballot_sync(MCP_FULL, b < n_dags && out_status[b] == KAHN_DEFERRED);
        for (; m; m &= m - 1)
            kahn_warp_dag<WITH_ORDER>(sw[warp], lane, g * 32 + (__ffs((int)m) - 1), n_nodes, edge_off, ed, out_level, out_order, out_n_levels, out_status);
    }
}


// ---- large batches: one THREAD per DAG ---------------------------------------------------------------------------------
// The warp kernel above spends ~2.6 K warp instructions per DAG on the order (profiles/r2e_summary.md): it is issue-bound,
// and what it issues is mostly per-generation fixed cost.  For a batch large enough to fill the machine with threads the
// literal algorithm is cheaper (470 instructions per DAG, profiles/r2k_summary.md): each thread runs dag.py:218-240 on its
// own DAG -- bucket the edges by source in first-seen order (DiGraph's adjacency order; a repeated edge is dropped there),
// count distinct parents, then a FIFO: the queue IS topological_sort's order, and a node's level is its releasing parent's
// + 1.  About (4 E + 3 n) short steps per DAG.
//   * the grid is resident at once and every CTA owns an equal, contiguous share of the batch (no CTA is left a chunk
//     behind its neighbours); it takes the share in pieces of at most KF_MAXC DAGs;
//   * a warp costs its longest lane, so the CTA counting-sorts a piece by the work each DAG is (4 E + 3 n) in shared memory
//     and deals it out in groups of 32 neighbours of that order, heaviest group first; each warp pulls its next group
//     from a CTA counter -- no barrier until the piece is done;
//   * per-thread state lives in shared memory, word-interleaved across the lanes of a warp (word j of lane t at [j][t]):
//     every access, whatever the index, is bank-conflict free; the bucketed children of a group share the warp's byte
//     pool, and a group whose edges exceed it runs in several passes over fewer lanes;
//   * edges are read twice from global memory, eight at a time (one 16-byte load per lane); levels go straight to the
//     output row (one byte per release: the thread wrote the row's 0xFF fill itself, so program order holds);
//   * counters are plain load-add-store on the lane's own words: shared-memory atomics in their place measured 5.7 x
//     slower for the whole kernel (r2k);
//   * a DAG whose edge list alone exceeds the pool (only possible with repeated edges) is handed back with status
//     KAHN_DEFERRED for kahn_levels_deferred_kernel.
#define KF_THREADS 160
#define KF_WARPS (KF_THREADS / 32)
#define KF_POOL 13424         // bytes of bucketed children per warp (what 2 CTAs of 5 warps per SM leave after the per-thread state and the sort arrays)
#define KF_BINS 1280
#define KF_MAXC 3520          // DAGs a CTA sorts at a time, at most (1 M DAGs over 148 x 2 CTAs: one piece each)
#define KF_MIN_DAGS 65536     // below ~50 K the warp kernel wins: one thread walks its heaviest DAG for ~0.17 ms whatever the batch
struct KfWarp {
    uint32_t start[32][32];    // 64 x u16 per DAG: out-degree -> list begin -> list end
    uint32_t indeg[16][32];    // 64 x u8: distinct parents not yet done
    uint32_t order[16][32];    // 64 x u8: the queue
    unsigned char pool[KF_POOL + 4];   // + 4: the queue reads one entry ahead
};
struct KfSmem {
    KfWarp w[KF_WARPS];
    unsigned short perm[KF_MAXC];   // position in the sorted order -> DAG of the chunk (its edge count is read again from edge_off)
    uint32_t wsum[KF_WARPS];
    int next_group;
};
static_assert(KF_BINS * sizeof(uint32_t) <= KF_POOL, "the sort's histogram borrows warp 0's pool");
static_assert(KF_MAXC % KF_THREADS == 0 && KF_MAXC <= 65535, "perm is 16-bit; each thread sorts KF_MAXC / KF_THREADS entries");
static_assert(KF_BINS % KF_THREADS == 0 && KF_POOL <= 32767, "each thread scans KF_BINS / KF_THREADS bins; list offsets are 16-bit");
static_assert(2 * (sizeof(KfSmem) + 1024) <= 232448, "two CTAs per SM");

// Byte i / half-word i of a lane's column, given the column's base (&a[0][lane]): word i >> 2 (i >> 1) sits 128 bytes on.
__device__ __forceinline__ unsigned char *kf_b(unsigned char *col, int i) { return col + (i & 0xFC) * 31 + i; }
__device__ __forceinline__ unsigned short *kf_h(unsigned char *col, int i) {
    return reinterpret_cast<unsigned short *>(col + (i & 0xFE) * 62 + 2 * i);
}

// f(u, v) for every edge of [e0, e1) with both ends below n, eight edges (one aligned 16-byte load) at a time; the first and
// the last load of a DAG may cover neighbours' edges, which are masked off.  `edges` is 16-byte aligned (the launcher checks), so
// a load never leaves the 16-byte granule that holds one of the DAG's own edges.
template <class F>
__device__ __forceinline__ void kf_edges(const unsigned char *__restrict__ edges, uint32_t e0, uint32_t e1, int n, F f) {
    if (e0 >= e1) return;
    const uint32_t E = e1 - e0;
    const uint4 *e128 = reinterpret_cast<const uint4 *>(edges);
    uint32_t e = e0 & ~7u;
    uint4 q = __ldg(e128 + (e >> 3));
    for (;;) {
        const uint32_t en = e + 8;
        const bool more = en < e1;
        uint4 qn = q;
        if (more) qn = __ldg(e128 + (en >> 3));
        const uint32_t ws[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int u0 = ws[j] & 0xFF, v0 = (ws[j] >> 8) & 0xFF, u1 = (ws[j] >> 16) & 0xFF, v1 = ws[j] >> 24;
            const uint32_t ra = e - e0 + 2 * j;                       // edge index within the DAG; wraps for a neighbour's edge in front
            if (ra < E && u0 < n && v0 < n) f(u0, v0);
            if (ra + 1 < E && u1 < n && v1 < n) f(u1, v1);
        }
        if (!more) break;
        e = en;
        q = qn;
    }
}

// One DAG, one thread (the lane's columns of w; its children at pl).  Returns the status; *nlev_out = generations found.
// A pool byte is a child id (6 bits) | 0x40 on the first entry of a source's list | 0x80 on a repeat of an earlier entry.
__device__ __forceinline__ int kf_one_dag(KfWarp &w, const int lane, unsigned char *pl, const unsigned char *__restrict__ edges,
                                          const uint32_t e0, const uint32_t e1, const int n,
                                          unsigned char *__restrict__ lvl_row, int *nlev_out) {
    unsigned char *const cs = reinterpret_cast<unsigned char *>(&w.start[0][lane]);
    unsigned char *const ci = reinterpret_cast<unsigned char *>(&w.indeg[0][lane]);
    unsigned char *const co = reinterpret_cast<unsigned char *>(&w.order[0][lane]);
    // 1: out-degree per source (repeats included)
    kf_edges(edges, e0, e1, n, [&](int u, int) { ++*kf_h(cs, u); });
    // list begins
    uint32_t acc = 0;
    for (int j = 0; 2 * j < n; ++j) {
        const uint32_t c = w.start[j][lane];
        const uint32_t c0 = c & 0xFFFFu, c1 = c >> 16;
        w.start[j][lane] = acc | ((acc + c0) << 16);
        acc += c0 + c1;
    }
    // 2: children bucketed by source, in edge order; start[u] ends as the END of u's list
    kf_edges(edges, e0, e1, n, [&](int u, int v) {
        unsigned short *p = kf_h(cs, u);
        const uint32_t s = *p;
        *p = (unsigned short)(s + 1);
        pl[s] = (unsigned char)v;
    });
    // mark where each source's list begins
    {
        uint32_t beg = 0;
        for (int j = 0; 2 * j < n; ++j) {
            const uint32_t c = w.start[j][lane];
            const uint32_t x0 = c & 0xFFFFu, x1 = c >> 16;
            if (x0 > beg) pl[beg] |= 0x40;
            if (x1 > x0) pl[x0] |= 0x40;
            beg = x1;
        }
    }
    // 3: one pass over the bucketed children: flag repeats within a list, count distinct parents
    {
        u64 seen = 0;
        uint32_t c = pl[0];
        for (uint32_t r = 0; r < acc; ++r) {
            const uint32_t cn = pl[r + 1];                 // (the pool has room for the read past the last entry)
            if (c & 0x40) seen = 0;
            const int v = c & 0x3F;
            const u64 bit = 1ull << v;
            if (seen & bit) pl[r] = (unsigned char)(c | 0x80);
            else ++*kf_b(ci, v);
            seen |= bit;
            c = cn;
        }
    }
    // 4: the queue.  Generation g + 1 starts where the queue's tail stood when the first node of generation g was taken.
    int tail = 0;
    for (int j = 0; 4 * j < n; ++j) {
        const uint32_t wd = w.indeg[j][lane];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int v = 4 * j + k;
            if (v < n && ((wd >> (8 * k)) & 0xFFu) == 0) { *kf_b(co, tail++) = (unsigned char)v; lvl_row[v] = 0; }
        }
    }
    int head = 0, lu = 0, bound = tail, nlev = 0;
    uint32_t r = 0, rend = 0, cur = 0;
    for (;;) {
        if (r == rend) {                                   // take the next node off the queue
            if (head == tail) break;
            if (head == bound) { ++lu; bound = tail; }
            const int u = *kf_b(co, head++);
            nlev = lu + 1;
            r = u ? *kf_h(cs, u - 1) : 0u;
            rend = *kf_h(cs, u);
            cur = pl[r];
        }
        if (r < rend) {                                    // one child of the node in hand (the next one is already on its way)
            const uint32_t c = cur;
            cur = pl[++r];
            if (!(c & 0x80u)) {
                const int v = c & 0x3F;
                unsigned char *pd = kf_b(ci, v);
                const int d = *pd - 1;
                *pd = (unsigned char)d;
                if (d == 0) { *kf_b(co, tail++) = (unsigned char)v; lvl_row[v] = (unsigned char)(lu + 1); }
            }
        }
    }
    *nlev_out = nlev;
    return tail != n;
}

// CTA i owns DAGs [i * per_cta, (i + 1) * per_cta) and takes them `sub` at a time (sub % 32 == 0, sub <= KF_MAXC): the launcher
// splits the batch evenly over a grid that is resident at once, so no CTA is left with a chunk more than its neighbours.
template <bool WITH_ORDER>
__global__ void __launch_bounds__(KF_THREADS)
kahn_fifo_kernel(const unsigned char *__restrict__ n_nodes, const uint32_t *__restrict__ edge_off,
                 const unsigned char *__restrict__ edges, int64_t n_dags, int64_t per_cta, int sub,
                 unsigned char *__restrict__ out_level, unsigned char *__restrict__ out_order,
                 unsigned char *__restrict__ out_n_levels, unsigned char *__restrict__ out_status) {
    extern __shared__ __align__(16) unsigned char kf_raw[];
    KfSmem &S = *reinterpret_cast<KfSmem *>(kf_raw);
    constexpr int PER = KF_MAXC / KF_THREADS;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    KfWarp &w = S.w[warp];
    uint32_t *hist = reinterpret_cast<uint32_t *>(S.w[0].pool);
    const int64_t b_end = min(n_dags, ((int64_t)blockIdx.x + 1) * per_cta);
  for (int64_t b0 = (int64_t)blockIdx.x * per_cta; b0 < b_end; b0 += sub) {
    const int cs = (int)min((int64_t)sub, b_end - b0), GROUPS = (cs + 31) >> 5, padded = GROUPS * 32;

    // ---- counting sort of the chunk's DAGs by the work they are (clipped at KF_BINS - 1; those DAGs are neighbours anyway) ----
#pragma unroll
    for (int j = 0; j < KF_BINS / KF_THREADS; ++j) hist[t + j * KF_THREADS] = 0;
    __syncthreads();
    uint32_t key[PER], within[PER];
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const int i = t + j * KF_THREADS;
        uint32_t E = 0, nn = 0;
        if (i < cs) {
            const uint32_t a0 = edge_off[b0 + i], a1 = edge_off[b0 + i + 1];
            E = a1 < a0 ? 0xFFFFu : min(a1 - a0, 0xFFFFu);
            nn = n_nodes[b0 + i];
        }
        key[j] = min((4u * E + 3u * nn) >> 1, (uint32_t)KF_BINS - 1u);      // ~ the steps the DAG will take
        within[j] = 0;
        if (i < padded) within[j] = atomicAdd(&hist[key[j]], 1u);
    }
    __syncthreads();
    {
        constexpr int BPT = KF_BINS / KF_THREADS;
        uint32_t h[BPT], sum = 0;
#pragma unroll
        for (int j = 0; j < BPT; ++j) { h[j] = hist[t * BPT + j]; sum += h[j]; }
        uint32_t inc = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t x = __shfl_up_sync(MCP_FULL, inc, d); if (lane >= d) inc += x; }
        if (lane == 31) S.wsum[warp] = inc;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < KF_WARPS; ++j) inc += j < warp ? S.wsum[j] : 0u;
        uint32_t run = inc - sum;
#pragma unroll
        for (int j = 0; j < BPT; ++j) { hist[t * BPT + j] = run; run += h[j]; }
    }
    if (t == 0) S.next_group = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const int i = t + j * KF_THREADS;
        if (i < padded) {
            const uint32_t pos = hist[key[j]] + within[j];
            S.perm[pos] = (unsigned short)i;
        }
    }
    __syncthreads();                                       // the histogram is done with: warp 0's pool is a pool again

    // ---- groups of 32 neighbours, heaviest first ----
    for (;;) {
        int g = 0;
        if (lane == 0) g = atomicAdd(&S.next_group, 1);
        g = __shfl_sync(MCP_FULL, g, 0);
        if (g >= GROUPS) break;
        const int pos = (GROUPS - 1 - g) * 32 + lane;
        const int mine = S.perm[pos];
        const int64_t b = b0 + mine;
        const bool valid = mine < cs;
        uint32_t E16 = 0, e_first = 0;
        if (valid) {
            e_first = edge_off[b];
            const uint32_t e_last = edge_off[b + 1];
            E16 = e_last < e_first ? 0xFFFFu : min(e_last - e_first, 0xFFFFu);
        }
        int n = valid ? n_nodes[b] : 0;
        const bool wide = n > 64, over = !wide && E16 > KF_POOL;
        const bool run = valid && !wide && !over;
        const uint32_t e0 = run ? e_first : 0u;
        const uint32_t E = run ? E16 : 0u;
        if (!run) n = 0;
        uint32_t inc = E;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t x = __shfl_up_sync(MCP_FULL, inc, d); if (lane >= d) inc += x; }
#pragma unroll
        for (int j = 0; j < 16; ++j) { w.indeg[j][lane] = 0; w.order[j][lane] = 0xFFFFFFFFu; }
#pragma unroll
        for (int j = 0; j < 32; ++j) w.start[j][lane] = 0;
        unsigned char *lvl_row = out_level + b * 64;
        if (valid) {
            uint4 *ol = reinterpret_cast<uint4 *>(lvl_row);
            const uint4 ff = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
#pragma unroll
            for (int i = 0; i < 4; ++i) ol[i] = ff;
        }
        int nlev = 0, status = wide ? 2 : (over ? KAHN_DEFERRED : 0);
        int first = 0;
        while (first < 32) {                               // one pass unless the group's edges exceed the pool
            const uint32_t pbase = __shfl_sync(MCP_FULL, inc - E, first);
            const bool fit = lane >= first && inc - pbase <= KF_POOL;
            if (run && fit) status = kf_one_dag(w, lane, w.pool + (inc - E - pbase), edges, e0, e0 + E, n, lvl_row, &nlev);
            __syncwarp();
            first += __popc(__ballot_sync(MCP_FULL, fit));
        }
        if (valid) {
            if (WITH_ORDER) {
                uint4 *oo = reinterpret_cast<uint4 *>(out_order + b * 64);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    oo[i] = make_uint4(w.order[4 * i][lane], w.order[4 * i + 1][lane], w.order[4 * i + 2][lane], w.order[4 * i + 3][lane]);
            }
            out_n_levels[b] = (unsigned char)nlev;
            out_status[b] = (unsigned char)status;
        }
        __syncwarp();
    }
    __syncthreads();                                       // the next chunk's sort reuses the arrays
  }
}


// ---- plans with 65 .. 255 nodes: one CTA per DAG, thread = node, 256-bit predecessor masks ----------------------------
// Same rules as the warp kernel above (duplicate edges collapse, a self loop never becomes ready, generation order by
// (position of the zeroing parent, index of the first edge parent -> node)); outputs have a pitch of 256 per DAG.  These
// are single oversized plans on the /execute path, not the batched million-DAG case, so the kernel is written for
// clarity: a block barrier per step, the edge list rescanned once per generation for the order.
#define KAHNW_THREADS 256
struct KahnWide {
    u64 pred[256][4];
    u64 done[4], fm[256][4];
    u64 key[256];
    uint32_t firste[256];
    uint32_t piece[8];
    unsigned char lvl[256], pos[256], ustar[256], order[256];
    int any;
};

template <bool WITH_ORDER>
__global__ void __launch_bounds__(KAHNW_THREADS)
kahn_levels_wide_kernel(const unsigned char *__restrict__ n_nodes, const uint32_t *__restrict__ edge_off,
                        const unsigned char *__restrict__ edges, int64_t n_dags,
                        unsigned char *__restrict__ out_level, unsigned char *__restrict__ out_order,
                        unsigned char *__restrict__ out_n_levels, unsigned char *__restrict__ out_status) {
    __shared__ KahnWide w;
    const int v = threadIdx.x, lane = v & 31, warp = v >> 5;
    const uchar2 *ed = reinterpret_cast<const uchar2 *>(edges);
    for (int64_t b = blockIdx.x; b < n_dags; b += gridDim.x) {
        const int n = n_nodes[b];
        const uint32_t e0 = edge_off[b], e1 = edge_off[b + 1];
        __syncthreads();                                  // the previous DAG has left shared memory
#pragma unroll
        for (int j = 0; j < 4; ++j) w.pred[v][j] = 0;
        if (v < 4) w.done[v] = 0;
        w.lvl[v] = 0xFF; w.order[v] = 0xFF;
        __syncthreads();
        for (uint32_t e = e0 + v; e < e1; e += KAHNW_THREADS) {
            const uchar2 uv = ed[e];
            if (uv.x < n && uv.y < n) atomicOr(reinterpret_cast<unsigned *>(&w.pred[uv.y][0]) + (uv.x >> 5), 1u << (uv.x & 31));
        }
        __syncthreads();
        u64 p[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) p[j] = w.pred[v][j];
        int level = 0, n_done = 0;
        unsigned char mylvl = 0xFF;
        for (;;) {
            const bool isdone = (w.done[v >> 6] >> (v & 63)) & 1ull;
            const bool ready = v < n && !isdone && !((p[0] & ~w.done[0]) | (p[1] & ~w.done[1]) | (p[2] & ~w.done[2]) | (p[3] & ~w.done[3]));
            const uint32_t bal = __ballot_sync(MCP_FULL, ready);
            if (lane == 0) w.piece[warp] = bal;
            __syncthreads();
            uint32_t anyb = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) anyb |= w.piece[j];
            if (anyb == 0) break;                         // block-uniform
            if (ready) mylvl = (unsigned char)level;
            __syncthreads();                              // everyone has read `done` and the pieces
            if (v < 4) {
                const u64 f = (u64)w.piece[2 * v] | ((u64)w.piece[2 * v + 1] << 32);
                w.done[v] |= f;
                if (WITH_ORDER) w.fm[level][v] = f;
            }
            for (int j = 0; j < 8; ++j) n_done += __popc(w.piece[j]);
            ++level;
            __syncthreads();
        }
        out_level[b * 256 + v] = mylvl;
        if (v == 0) { out_n_levels[b] = (unsigned char)level; out_status[b] = (unsigned char)(n_done != n); }
        if (WITH_ORDER) {
            w.lvl[v] = mylvl;
            __syncthreads();
            int base = 0;
            for (int L = 0; L < level; ++L) {
                const bool in = mylvl == L;
                u64 k = (u64)v;                           // generation 0: node order
                if (L > 0) {
                    int bp = -1, bu = 0;
                    if (in) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            u64 c = p[j] & w.fm[L - 1][j];
                            while (c) { const int u = j * 64 + __ffsll((long long)c) - 1; c &= c - 1; const int q = w.pos[u]; if (q > bp) { bp = q; bu = u; } }
                        }
                        w.ustar[v] = (unsigned char)bu; w.firste[v] = 0xFFFFFFFFu;
                    }
                    __syncthreads();
                    for (uint32_t e = e0 + v; e < e1; e += KAHNW_THREADS) {
                        const uchar2 uv = ed[e];
                        if (uv.x < n && uv.y < n && w.lvl[uv.y] == L && w.ustar[uv.y] == uv.x) atomicMin(&w.firste[uv.y], e - e0);
                    }
                    __syncthreads();
                    if (in) k = ((u64)bp << 32) | w.firste[v];
                }
                if (in) w.key[v] = k;
                __syncthreads();
                int rk = 0, cnt = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    u64 m = w.fm[L][j];
                    cnt += __popcll(m);
                    if (in) for (; m; m &= m - 1) rk += w.key[j * 64 + __ffsll((long long)m) - 1] < k;
                }
                __syncthreads();                          // keys read before pos (same bytes are not shared, but order[] below is)
                if (in) { w.pos[v] = (unsigned char)(base + rk); w.order[base + rk] = (unsigned char)v; }
                base += cnt;
                __syncthreads();
            }
            out_order[b * 256 + v] = w.order[v];
        }
    }
}
