// kahn.cuh -- K6: batched, bit-exact Kahn level scheduling (integer only).
//
// Stands for nx.topological_generations (networkx dag.py:218-240) on the DiGraph that
// control_plane.py:94-100 builds.  One warp per DAG (<= 64 nodes): lane l owns nodes l and l+32.
//   build   pred[v] = 64-bit mask of v's distinct predecessors (atomicOr: a duplicate edge sets
//           the same bit again, which is exactly DiGraph's "same edge collapses"; a self loop
//           sets bit v of pred[v], so v can never become ready -> cycle)
//   levels  generation L = { v not done : pred[v] & ~done == 0 }, found with two ballots; no
//           per-node counters, no queues.
//   order   (optional) the order nx.topological_sort yields: generation 0 in node order, later
//           generations by (position of the parent whose decrement zeroed the node, index of
//           the first edge parent->node), which is the order dag.py:224-234 appends them in.
// HBM traffic per DAG: 1 + 8 + 2E bytes in, 64 (+64) + 2 bytes out; everything else is
// shared memory and registers.
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

template <bool WITH_ORDER>
__global__ void __launch_bounds__(KAHN_THREADS)
kahn_levels_kernel(const unsigned char *__restrict__ n_nodes, const uint32_t *__restrict__ edge_off,
                   const unsigned char *__restrict__ edges, int64_t n_dags,
                   unsigned char *__restrict__ out_level, unsigned char *__restrict__ out_order,
                   unsigned char *__restrict__ out_n_levels, unsigned char *__restrict__ out_status) {
    __shared__ KahnWarp sw[KAHN_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    KahnWarp &w = sw[warp];
    const uchar2 *ed = reinterpret_cast<const uchar2 *>(edges);
    for (int64_t b = (int64_t)blockIdx.x * KAHN_WARPS + warp; b < n_dags; b += (int64_t)gridDim.x * KAHN_WARPS) {
        const int n = n_nodes[b];
        if (n > 64) {                                    // not this kernel's DAG (kahn_levels_wide_kernel takes up to 255 nodes): say so
            out_level[b * 64 + lane] = 0xFF; out_level[b * 64 + 32 + lane] = 0xFF;
            if (WITH_ORDER) { out_order[b * 64 + lane] = 0xFF; out_order[b * 64 + 32 + lane] = 0xFF; }
            if (lane == 0) { out_n_levels[b] = 0; out_status[b] = 2; }
            continue;
        }
        const uint32_t e0 = edge_off[b], e1 = edge_off[b + 1];
        w.pred[lane] = 0; w.pred[lane + 32] = 0;
        __syncwarp();
        for (uint32_t e = e0 + lane; e < e1; e += 32) {
            uchar2 uv = ed[e];
            if (uv.x < n && uv.y < n) atomicOr(&w.pred[uv.y], 1ull << uv.x);
        }
        __syncwarp();
        const u64 p0 = w.pred[lane], p1 = w.pred[lane + 32];
        const bool ok0 = lane < n, ok1 = lane + 32 < n;
        u64 done = 0;
        int level = 0;
        unsigned char l0 = 0xFF, l1 = 0xFF;
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
            if (uv.x < n && uv.y < n) atomicOr(&w.pred[uv.y][uv.x >> 6], 1ull << (uv.x & 63));
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
