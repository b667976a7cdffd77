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

struct KahnWarp {
    u64 pred[64];
    u64 fm[64];            // frontier mask of each generation
    u64 key[64];
    uint32_t firste[64];
    unsigned char lvl[64], pos[64], ustar[64], order[64];
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
            __syncwarp();
            int base = 0;
            for (int L = 0; L < level; ++L) {
                const u64 f = w.fm[L];
                const bool in0 = (f >> lane) & 1ull, in1 = (f >> (lane + 32)) & 1ull;
                u64 k0 = (u64)lane, k1 = (u64)(lane + 32);       // generation 0: node order
                if (L > 0) {
                    const u64 prev = w.fm[L - 1];
                    int bp0 = -1, bp1 = -1;
                    if (in0) {
                        u64 c = p0 & prev; int bu = 0;
                        while (c) { int u = __ffsll((long long)c) - 1; c &= c - 1; int p = w.pos[u]; if (p > bp0) { bp0 = p; bu = u; } }
                        w.ustar[lane] = (unsigned char)bu; w.firste[lane] = 0xFFFFFFFFu;
                    }
                    if (in1) {
                        u64 c = p1 & prev; int bu = 0;
                        while (c) { int u = __ffsll((long long)c) - 1; c &= c - 1; int p = w.pos[u]; if (p > bp1) { bp1 = p; bu = u; } }
                        w.ustar[lane + 32] = (unsigned char)bu; w.firste[lane + 32] = 0xFFFFFFFFu;
                    }
                    __syncwarp();
                    for (uint32_t e = e0 + lane; e < e1; e += 32) {
                        uchar2 uv = ed[e];
                        if (uv.x < n && uv.y < n && w.lvl[uv.y] == L && w.ustar[uv.y] == uv.x)
                            atomicMin(&w.firste[uv.y], e - e0);
                    }
                    __syncwarp();
                    if (in0) k0 = ((u64)bp0 << 32) | w.firste[lane];
                    if (in1) k1 = ((u64)bp1 << 32) | w.firste[lane + 32];
                }
                if (in0) w.key[lane] = k0;
                if (in1) w.key[lane + 32] = k1;
                __syncwarp();
                int rk0 = 0, rk1 = 0;
                for (u64 m = f; m; m &= m - 1) {
                    u64 kx = w.key[__ffsll((long long)m) - 1];
                    rk0 += kx < k0; rk1 += kx < k1;
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
