// gemm_tf32.cuh -- K2: tcgen05 / TMEM distance GEMM with the telemetry score and the
// top-k' filter fused into the epilogue (large query batches).  Two operand kinds, one kernel:
//   kind 0 (tf32): the fp32 table rows themselves, consumed by kind::tf32 MMAs straight from shared memory;
//   kind 1 (f16):  the fp16 shadow of the table (prep.cuh shadow_rows_kernel: every row scaled by a power of two into fp16's
//                  range, 11 bits kept -- one more than tf32 keeps) and fp16 images of the queries, consumed by kind::f16
//                  MMAs at twice the tensor rate and half the bytes.  A 128-byte k-block then holds 64 elements instead
//                  of 32; tile bytes, swizzle, descriptors and the +32 B step per MMA (K = 16 instead of 8) are the same,
//                  so the two kinds share everything but the instruction and the element count per k-block.  The scales
//                  are undone in the epilogue (folded into the inverse norms; one more multiply for L2).
//
//   D[128 queries x 256 rows] (fp32, TMEM) += Q[128 x 32] (smem, K-major) * E[256 x 32]^T (smem, K-major)
//
// Queries are the M operand so that one TMEM lane = one query: an epilogue thread owns a query,
// keeps its running threshold in a register and sees that query's scores against 256 table rows
// per tile.  Table rows are the N operand and are streamed once per query block; fp32 rows land
// in shared memory by TMA (128-byte swizzle) and are consumed by kind::tf32 MMAs directly -- no
// conversion pass, no shadow table, the bytes read are the fp32 table.
//
// Warp roles (320 threads, 1 CTA / SM, persistent over row tiles):
//   warp 0   TMA producer: per k-block one 16 KB query tile + one 32 KB row tile into a 4-stage ring
//   warp 1   TMEM allocator + MMA issuer: 4 x tcgen05.mma (M128 N256 K8) per stage, tcgen05.commit
//            frees the stage; the last commit of a tile publishes the accumulator
//   warps 2-9  epilogue (two warps per scheduler, each thread = one query x half of the tile's rows):
//            tcgen05.ld the accumulator (double-buffered in TMEM: 2 x 256 columns),
//            score = 1 - dot * inv|q| * inv|e| + penalty (or sqrt(|q|^2+|e|^2-2 dot) + penalty),
//            compare with the thread's threshold, append survivors to the (CTA, query) list in
//            global memory; a list that fills is cut back to k' by a register-only warp radix
//            select, which also tightens the threshold.
//
// Threshold bootstrap.  A fresh list accepts everything, and with ~70 lists per query the start-up transient
// (fill, prune, fill, prune ...) used to dominate the kernel.  Instead every CTA first reads its first tile only
// to find, per query, its m smallest keys (m = ceil(k'/L), registers), publishes the m-th, and waits until the L
// CTAs that share its queries have done the same (cooperative launch: all CTAs are co-resident).  The largest
// published key t0 has at least L*m >= k' keys of the union below it, so it bounds the final k'-th key from
// above and every list starts with it; the first tile is then read from TMEM a second time through the filter.
//
// Coarse error (certificate, rescore.cuh): kind::tf32 reads the top 19 bits of each fp32 operand,
// i.e. truncates 13 mantissa bits: relative error < 2^-10 per operand, < 2^-9 (1 + 2^-11) per
// product, so |dot_tf32 - dot| <= 2^-9 * 1.001 * sum|q_i e_i| + fp32 accumulation (<= dim * 2^-22 |q||e|).
#pragma once
#include <cuda.h>

#include "common.cuh"

#define G_BM 128
#define G_BN 256
#define G_BK 32
#define G_A_BYTES (G_BM * G_BK * 4)
#define G_B_BYTES (G_BN * G_BK * 4)
#define G_MAX_QBLOCKS 8

struct GemmArgs {
    int64_t n_rows;
    int dim, nq, nqb, kprime, cap;
    const float *inv_nq;       // [nq]
    const float *nq2f;         // [nq] fp32 squared norms (L2 metric)
    const float *inv_ne;       // [n_rows]
    const double *ne2;         // [n_rows]
    const float *penalty;      // [n_rows] or NULL
    u64 *lists;                // [nq][L][cap]
    int *list_cnt;             // [nq][L]
    float *dump;               // debug: [nq][n_rows] coarse scores instead of lists, or NULL
    // cross-CTA bootstrap of the per-query threshold (cooperative launch only; boot_m == 0 disables it)
    u64 *pub;                  // [nq][L]  each (CTA, query)'s boot_m-th smallest key of its first tile
    unsigned int *arrive;      // [nqb]    CTAs of a query block that have published (zero when the launch starts)
    unsigned int *arrive_next; // [G_MAX_QBLOCKS] the other slot of the pair: this launch clears it for the next one
    u64 *tau_pub;              // [nq] (launch epoch << 32 | float bits of the shared bound), posted by list 0 of each query as the scan goes on
    unsigned int *drop_min;    // [nq] ordered image of the best coarse score the bootstrap tile may have dropped (all-ones: none); never NULL when boot_m > 0
    u64 *progress;             // [ngroups][nqb] (launch epoch << 32 | tile index each CTA has started), or NULL: keeps the CTAs that
                               // walk the same row tiles within `G_DRIFT` tiles of each other so that the tiles stay in L2
    int drift;                 // tiles a CTA may run ahead of the slowest CTA of its group
    uint32_t epoch;            // this launch's number: older entries in `progress` read as "not started"
    int boot_early;            // the bootstrap tile hands its accumulator back before the wait (large tables only, see gemm_tf32_run)
    int boot_m, boot_lists;    // lists 0 .. boot_lists-1 publish their boot_m-th smallest key (boot_lists even)
    const unsigned char *rows_b;   // operand table base, row pitch and row length in bytes, for the L2 row prefetch
    int64_t row_pitch_b;
    int row_len_b;
    int l2_prefetch;
    int bn;                    // rows per tile (<= 256, a multiple of the cluster size): chosen so that every CTA gets the same number of tiles
    int cluster;               // CTAs per cluster (1, 2 or 4): the query-block CTAs that share a row tile load a slice each and multicast it
    uint32_t idesc_pair;       // pair mode: instruction descriptor with N = bn
    int pair;                  // cluster == 2 only: the two CTAs run ONE tcgen05.mma.cta_group::2 (M = 256) instead of two multicast-fed M = 128 MMAs
    int kind;                  // 0: fp32 operands read as tf32 (kind::tf32, K = 8 per MMA); 1: fp16 operands (kind::f16, K = 16 per MMA)
    int ek;                    // elements per 128-byte k-block: 32 (tf32) or 64 (f16)
    const float *q_scale_inv;  // kind 1: [nq] 1 / (power-of-two scale of the query's fp16 image)
    const float *row_scale_inv;// kind 1: [n_rows] the same per shadow row
    uint32_t idesc_single;     // instruction descriptor of the single-CTA MMA (M = 128, N = 256) for this kind
    int debug;                 // timing experiments only (results are wrong): 1 skip epilogue math, 2 skip row-tile loads, 4 skip query-tile loads, 8 skip MMAs
};
#define G_BOOT_MAX 4
#define G_DRIFT 1
#define G_MAX_GROUPS 160


// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *tm, int c0, int c1, u64 *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// the same box, delivered to the same shared-memory offset (and counted on the same mbarrier offset) of every CTA in ctaMask
__device__ __forceinline__ void tma_load_2d_mc(void *dst, const CUtensorMap *tm, int c0, int c1, u64 *bar, uint16_t mask) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(u64 *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// arrive on the mbarrier at this offset in every CTA of the cluster named in mask, once the MMAs issued so far retire
__device__ __forceinline__ void tc_commit_mc(u64 *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
// ---- CTA pair (cta_group::2) ----
// a load issued by either CTA of the pair into its own shared memory, its bytes counted on the LEADER's (even rank) mbarrier:
// clearing bit 24 of a shared::cluster address names the same offset in the even CTA (cute: Sm100MmaPeerBitMask)
__device__ __forceinline__ void tma_load_2d_pair(void *dst, const CUtensorMap *tm, int c0, int c1, u64 *bar) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_commit_pair(u64 *bar) {     // same offset in both CTAs, once the pair's MMAs so far retire
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32_pair(uint32_t d_tmem, u64 adesc, u64 bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_mma_f16_pair(uint32_t d_tmem, u64 adesc, u64 bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, u64 adesc, u64 bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on the mbarrier at this offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(u64 *bar, uint32_t rank) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}"
        ::"r"(smem_u32(bar)), "r"(rank) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], kind::tf32, issued by one thread
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, u64 adesc, u64 bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// 32 lanes x 32 consecutive columns: thread i of the warp gets TMEM lane (base lane + i)
__device__ __forceinline__ void tc_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128-byte swizzle shared-memory operand descriptor (cute::UMMA::SmemDescriptor layout):
// [0,14) start >> 4, [16,30) LBO >> 4 (unused for swizzled K-major: 1), [32,46) SBO >> 4 = 1024 B between
// 8-row groups, [46,48) version = 1, [61,64) layout = 2 (SWIZZLE_128B).
__device__ __forceinline__ u64 umma_desc_k128(const void *tile) {
    return (u64)((smem_u32(tile) & 0x3FFFF) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): c F32 (1<<4), a/b format at [7,10) / [10,13) (TF32 = 2 under
// kind::tf32, F16 = 0 under kind::f16), both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
static inline uint32_t g_idesc(int kind, int m, int n) {
    const uint32_t fmt = kind ? 0u : 2u;
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// K2 lists store raw pairs (fp32 score bits << 32 | row); selection works on the ordered image
__device__ __forceinline__ u64 raw_to_key(u64 raw) { return ((u64)f32_ord(__uint_as_float((uint32_t)(raw >> 32))) << 32) | (raw & 0xFFFFFFFFull); }
__device__ __forceinline__ u64 key_to_raw(u64 key) { return ((u64)__float_as_uint(ord_f32((uint32_t)(key >> 32))) << 32) | (key & 0xFFFFFFFFull); }

// ---- register-only warp radix select --------------------------------------------------------------
// The warp cuts lane `src`'s list (n keys in global memory, n > kprime) back to its kprime smallest
// keys and returns the largest kept key (the new threshold).  Keys are unique.
template <int KPL>
__device__ __forceinline__ u64 warp_prune_list(u64 *lp, int n, int kprime, int lane) {
    u64 kk[KPL];
    uint32_t act = 0;
#pragma unroll
    for (int i = 0; i < KPL; ++i) {
        int idx = i * 32 + lane;
        kk[i] = idx < n ? raw_to_key(lp[idx]) : ~0ull;
        act |= (idx < n ? 1u : 0u) << i;
    }
    int need = kprime, tot = n;
    for (int bit = 63; bit >= 0 && tot != need; --bit) {
        uint32_t z = 0;
        int c0 = 0;
#pragma unroll
        for (int i = 0; i < KPL; ++i) {
            bool zero = ((act >> i) & 1u) && !((kk[i] >> bit) & 1ull);
            z |= (zero ? 1u : 0u) << i;
            c0 += __popc(__ballot_sync(MCP_FULL, zero));
        }
        if (need <= c0) { act = z; tot = c0; }
        else { act &= ~z; need -= c0; tot -= c0; }
    }
    // every remaining active key is kept; threshold = their maximum
    u64 mx = 0;
#pragma unroll
    for (int i = 0; i < KPL; ++i) if ((act >> i) & 1u) mx = kk[i] > mx ? kk[i] : mx;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        u64 t = __shfl_xor_sync(MCP_FULL, mx, o);
        mx = t > mx ? t : mx;
    }
    int base = 0;
#pragma unroll
    for (int i = 0; i < KPL; ++i) {
        bool keep = kk[i] <= mx;
        uint32_t b = __ballot_sync(MCP_FULL, keep);
        if (keep) lp[base + __popc(b & ((1u << lane) - 1u))] = key_to_raw(kk[i]);
        base += __popc(b);
    }
    return mx;
}

// MT = query blocks (of 128) per CTA.  Only MT = 1 is instantiated: D[128 x 256], TMEM double-buffered (2 x 256 columns), so the
// epilogue of tile i overlaps the MMAs of tile i+1.  (MT = 2 -- two accumulators sharing the row tile, all 512 TMEM columns,
// single-buffered -- measured slower in round 1 and was retired; the template parameter stays for the shared-memory layout.)
template <int MT, int PAIR = 0> struct GemmCfg {
    static constexpr int B_ROWS = PAIR ? G_BN / 2 : G_BN;      // rows of the tile this CTA's shared memory holds
    static constexpr int STAGES = PAIR ? 6 : (MT == 1 ? 4 : 3);   // pair: 32 KB per stage, 6 stages = 5 k-blocks (~1.8 us of MMA time) in flight
    static constexpr int NBUF = MT == 1 ? 2 : 1;
    static constexpr int A_BYTES = MT * G_BM * G_BK * 4;
    static constexpr int EPI_THREADS = 256 * MT;
    static constexpr int THREADS = 64 + EPI_THREADS;
};
template <int MT, int PAIR = 0> struct GemmSmemT {
    float a[GemmCfg<MT, PAIR>::STAGES][MT * G_BM * G_BK];
    float b[GemmCfg<MT, PAIR>::STAGES][GemmCfg<MT, PAIR>::B_ROWS * G_BK];
    float2 meta[2][G_BN];      // per row of the tile: (inv|e| or |e|^2, penalty)
    float mscale[2][G_BN];     // kind 1, L2 metric: 1 / (the shadow row's power-of-two scale)
    u64 full[GemmCfg<MT, PAIR>::STAGES], empty[GemmCfg<MT, PAIR>::STAGES], tfull[2], tempty[2];
    uint32_t tmem_base;
};

template <int MT, int KPL, int METRIC, int PAIR>
__global__ void __launch_bounds__(GemmCfg<MT>::THREADS, 1)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmArgs a) {
    using Cfg = GemmCfg<MT, (MT == 1 && PAIR) ? 1 : 0>;
    using Smem = GemmSmemT<MT, (MT == 1 && PAIR) ? 1 : 0>;
    extern __shared__ unsigned char g_smem_raw[];
    // 1024-byte alignment for the 128-byte swizzle; plain pointer arithmetic on the __shared__ array keeps the
    // accesses in the shared address space (a uintptr_t round trip turns them into generic loads)
    Smem &s = *reinterpret_cast<Smem *>(g_smem_raw + ((1024u - (smem_u32(g_smem_raw) & 1023u)) & 1023u));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qb = blockIdx.x % a.nqb, group = blockIdx.x / a.nqb, ngroups = gridDim.x / a.nqb;
    const int BN = a.bn;                                       // rows per tile; the MMA is always N = 256, columns >= BN are ignored
    const int n_tiles = (int)((a.n_rows + BN - 1) / BN);
    const int my_tiles = group < n_tiles ? (n_tiles - group + ngroups - 1) / ngroups : 0;
    const int EK = a.ek;                                       // elements per k-block (128 bytes of a row)
    const int KB = (a.dim + EK - 1) / EK;
    constexpr int CAP = KPL * 32;

    // Cluster of C query-block CTAs that work on the same row tiles: each loads 256 / C rows of the tile and multicasts
    // them to all C, so the tile leaves L2 once per cluster instead of once per CTA (and the CTAs cannot drift apart,
    // which used to turn L2 hits into DRAM re-reads).  A stage slot is refilled only when all C consumers released it.
    //
    // Pair mode (a.pair, C == 2, MT == 1): the two CTAs are one tcgen05 CTA pair.  Each loads its own 128 queries (A) and
    // HALF of the row tile (B), the leader (even rank) issues tcgen05.mma.cta_group::2 with M = 256: its tensor core
    // and the peer's each produce their own 128 x 256 accumulator, reading B half from either shared memory.  Per SM that
    // is 32 KB of TMA fill and 32 KB of operand reads per k-block instead of 48 + 48 -- the multicast variant saves L2
    // traffic but not shared-memory bandwidth, which is what held the M = 128 kernel at 0.70 of the TF32 peak.
    const int C = a.cluster;
    const uint32_t crank = C > 1 ? cluster_ctarank() : 0u;
    const uint16_t cmask = (uint16_t)((1u << C) - 1u);
    constexpr bool pair = MT == 1 && PAIR != 0;     // compile time: a kernel that holds cta_group::2 code can only be launched in clusters
    const bool leader = !pair || crank == 0;
    if (threadIdx.x == 0) {
        // pair: `full` and `tempty` live in the leader only (the peer's loads and epilogue arrive there); `empty` and
        // `tfull` get one multicast commit each from the leader's MMA thread
        for (int i = 0; i < Cfg::STAGES; ++i) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], pair ? 1u : (uint32_t)C); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s.tfull[i], 1); mbar_init(&s.tempty[i], Cfg::EPI_THREADS * (pair ? 2 : 1)); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        if constexpr (pair) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (C > 1) cluster_sync_all();                             // peers' barriers exist before anything remote touches them
    tc_fence_after();
    const uint32_t tmem = s.tmem_base;

    if (warp == 0) {
        // ===== TMA producer (whole warp walks the loop, lane 0 issues) =====
        if (blockIdx.x == 0 && lane < G_MAX_QBLOCKS && a.arrive_next) a.arrive_next[lane] = 0;
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        }
        int stage = 0;
        uint32_t phase = 0;
        for (int i = 0; i < my_tiles; ++i) {
            const int t = group + i * ngroups;
            if (a.progress) {
                // The nqb CTAs of a group stream the same row tiles.  Left alone they drift apart by more tiles than L2 holds
                // and each re-fetches the tile from HBM (measured: 21.6 GB read for 7.7 GB of table at config 3).  So each
                // announces the tile it starts and waits until none of the others is more than a.drift tiles behind.
                u64 *pg = a.progress + (size_t)group * a.nqb;
                if (lane == 0) *(volatile u64 *)(pg + qb) = ((u64)a.epoch << 32) | (u64)(uint32_t)i;
                if (i >= a.drift) {
                    const u64 want = ((u64)a.epoch << 32) | (u64)(uint32_t)(i - a.drift);
                    for (;;) {
                        const u64 v = lane < a.nqb ? *(volatile u64 *)(pg + lane) : ~0ull;
                        if (__all_sync(MCP_FULL, v >= want)) break;
                        __nanosleep(100);
                    }
                }
            }
            for (int kb = 0; kb < KB; ++kb) {
                // DRAM locality: a TMA box touches 256 rows x 128 B, i.e. 256 different DRAM pages, and comes back to
                // each page one k-block later, long after other tiles have closed it.  Every 8 k-blocks the query
                // block 0 CTA therefore pulls the *next* 1 KB of each of the tile's rows into L2 with one bulk prefetch
                // per row (8 rows per lane): DRAM sees 1 KB bursts, the TMA boxes that follow hit L2.
                if (a.l2_prefetch && qb == 0 && (kb & 7) == 0) {
                    int pt = t, pk = kb + 8;
                    if (pk >= KB) { pt = t + ngroups; pk = 0; }
                    if (pt < n_tiles) {
                        const int bytes = min(1024, a.row_len_b - pk * 128);       // a k-block is 128 bytes of a row in either kind
                        const int64_t prow0 = (int64_t)pt * BN;
#pragma unroll
                        for (int u = 0; u < G_BN / 32; ++u) {
                            const int64_t r = prow0 + u * 32 + lane;
                            if (r < a.n_rows && bytes >= 16)
                                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.rows_b + r * a.row_pitch_b + pk * 128), "r"(bytes & ~15) : "memory");
                        }
                    }
                }
                mbar_wait(&s.empty[stage], phase ^ 1u);
                if (lane == 0) {
                    const bool la = !(a.debug & 4) || (i == 0 && kb < Cfg::STAGES), lb = !(a.debug & 2) || (i == 0 && kb < Cfg::STAGES);
                    if constexpr (pair) {
                        // both halves of the pair report to the leader's barrier, which expects the bytes of both
                        if (leader) mbar_expect_tx(&s.full[stage], 2 * ((la ? Cfg::A_BYTES : 0) + (lb ? (BN / 2) * G_BK * 4 : 0)));
                        if (la) tma_load_2d_pair(s.a[stage], &tmA, kb * EK, qb * G_BM, &s.full[stage]);
                        if (lb) tma_load_2d_pair(s.b[stage], &tmB, kb * EK, t * BN + (int)crank * (BN / 2), &s.full[stage]);
                    } else {
                    mbar_expect_tx(&s.full[stage], (la ? Cfg::A_BYTES : 0) + (lb ? BN * G_BK * 4 : 0));
                    if (la) tma_load_2d(s.a[stage], &tmA, kb * EK, qb * MT * G_BM, &s.full[stage]);
                    if (lb) {
                        if (C == 1) tma_load_2d(s.b[stage], &tmB, kb * EK, t * BN, &s.full[stage]);
                        else {                                 // my slice of the tile, to everyone in the cluster
                            const int srows = BN / C;
                            tma_load_2d_mc(s.b[stage] + (size_t)crank * srows * G_BK, &tmB, kb * EK, t * BN + (int)crank * srows,
                                           &s.full[stage], cmask);
                        }
                    }
                    }
                }
                __syncwarp();
                if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
            }
        }
        if (a.progress && lane == 0) *(volatile u64 *)(a.progress + (size_t)group * a.nqb + qb) = ((u64)a.epoch << 32) | 0xFFFFFFFFull;   // done: nobody waits for me
    } else if (warp == 1) {
        // ===== MMA issuer =====
        int stage = 0;
        uint32_t phase = 0;
        for (int i = 0; i < (leader ? my_tiles : 0); ++i) {   // pair: the leader's thread drives both tensor cores
            const int buf = i % Cfg::NBUF;
            const uint32_t tphase = (uint32_t)(i / Cfg::NBUF) & 1u;
            mbar_wait(&s.tempty[buf], tphase ^ 1u);           // epilogue has drained this accumulator (pair: both epilogues)
            tc_fence_after();
            const uint32_t d_tmem = tmem + (uint32_t)(buf * G_BN);
            for (int kb = 0; kb < KB; ++kb) {
                mbar_wait(&s.full[stage], phase);             // TMA bytes have landed
                tc_fence_after();
                if (lane == 0) {
                    const u64 ad = umma_desc_k128(s.a[stage]), bd = umma_desc_k128(s.b[stage]);
#pragma unroll
                    for (int k = 0; k < G_BK / 8; ++k) {      // +32 bytes along K inside the swizzle atom = +2 in the address field
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt) {     // query block mt: A rows 128*mt.. (16 KB further), accumulator columns 256*mt..
                            if ((a.debug & 8) && k != 0) continue;
                            const uint32_t accum = (uint32_t)((kb | k) != 0);
                            if constexpr (pair) {
                                if (a.kind) tc_mma_f16_pair(d_tmem, ad + (u64)(2 * k), bd + (u64)(2 * k), a.idesc_pair, accum);
                                else tc_mma_tf32_pair(d_tmem, ad + (u64)(2 * k), bd + (u64)(2 * k), a.idesc_pair, accum);
                            } else {
                                const uint32_t dd = d_tmem + (uint32_t)(mt * G_BN);
                                const u64 aa = ad + (u64)(2 * k) + (u64)(mt * (G_A_BYTES >> 4));
                                if (a.kind) tc_mma_f16(dd, aa, bd + (u64)(2 * k), a.idesc_single, accum);
                                else tc_mma_tf32(dd, aa, bd + (u64)(2 * k), a.idesc_single, accum);
                            }
                        }
                    }
                    if constexpr (pair) tc_commit_pair(&s.empty[stage]); // frees the stage in both CTAs
                    else if (C == 1) tc_commit(&s.empty[stage]);   // frees the smem stage when these MMAs retire
                    else tc_commit_mc(&s.empty[stage], cmask); // ... in every CTA of the cluster: their producers write my copy too
                    if (kb == KB - 1) { if constexpr (pair) tc_commit_pair(&s.tfull[buf]); else tc_commit(&s.tfull[buf]); }
                }
                __syncwarp();
                if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else {
        // ===== epilogue: 8 warps per query block; a thread = one query (TMEM lane) x one half of the tile's 256 rows =====
        const int quarter = warp & 3;                          // a warp may only touch TMEM lanes 32*(warp%4)..+31
        const int egrp = (warp - 2) >> 2;                      // 0 .. 2*MT-1
        const int mt = egrp >> 1, half = egrp & 1;             // query block within the CTA, columns [half*128, +128)
        const int m = quarter * 32 + lane;
        const int qg = (qb * MT + mt) * G_BM + m;
        const bool qvalid = qg < a.nq;
        // kind 1: the accumulator is (s_q q^).(s_r e^); 1 / s_q is folded into the query's inverse norm here, 1 / s_r into the
        // row's as the tile's constants are staged (powers of two: exact); the L2 form undoes both on the dot product
        const float qsi = (qvalid && a.kind) ? a.q_scale_inv[qg] : 1.f;
        const float inq = qvalid ? a.inv_nq[qg] * qsi : 0.f;
        const float nq2f = qvalid ? a.nq2f[qg] : 0.f;
        const float m2q = -2.0f * qsi;
        const int et = threadIdx.x - 64;                       // 0 .. EPI_THREADS-1
        const int L = 2 * ngroups, li = 2 * group + half;      // one list per (CTA, query, half)
        u64 *mylist = a.lists + ((int64_t)(qvalid ? qg : 0) * L + li) * CAP;
        int cnt = 0;
        float tau_f = qvalid ? CUDART_INF_F : -CUDART_INF_F;   // float image of tau: the common-case reject is one FSETP
        // hands accumulator `buf` back to the MMA thread; in pair mode that thread lives in the leader CTA
        auto release_acc = [&](int buf) {
            if constexpr (pair) { if (leader) mbar_arrive(&s.tempty[buf]); else mbar_arrive_remote(&s.tempty[buf], 0u); }
            else mbar_arrive(&s.tempty[buf]);
        };
        auto score = [&](float acc, int mb, int r) -> float {
            const float2 mm = s.meta[mb][r];
            if (METRIC == 0) return fmaf(-(acc * inq), mm.x, 1.0f) + mm.y;
            if (a.kind) acc *= s.mscale[mb][r];
            // sqrt.approx (one MUFU, max relative error 2^-23) instead of the IEEE sqrtf sequence: every lane scores every
            // element, so the epilogue's instruction count is what the L2 metric pays for; the extra ulp is inside the
            // 8-ulp allowance the certificate already makes for the square root and the final add (CertParams::abs_pen)
            float rt;
            asm("sqrt.approx.f32 %0, %1;" : "=f"(rt) : "f"(fmaxf(fmaf(m2q, acc, nq2f + mm.x), 0.f)));
            return rt + mm.y;
        };
        // my four smallest keys so far: found by the bootstrap pass over the first tile and, when the bootstrap hands its tile back
        // early (large tables), kept current by the filter pass, so that the shared bound can be tightened as the scan goes on
        u64 b0 = ~0ull, b1 = ~0ull, b2 = ~0ull, b3 = ~0ull;
        // ... and, from then on, four small scores of rows this list has seen, ascending (floats; +inf = none).  They need not be
        // THE four smallest: the maximum of any four seen scores bounds the list's fourth-smallest from above, which is all the
        // shared bound asks for.  Only the lists that publish (the first boot_lists, a warp-uniform property) keep them current.
        float f0 = CUDART_INF_F, f1 = CUDART_INF_F, f2 = CUDART_INF_F, f3 = CUDART_INF_F;
        u64 tp = 0;                                            // tau_pub word read one tile ago (epoch 0 never matches)
        const bool refresh = a.boot_m > 0 && a.boot_early && !(a.debug & 64);
        // per-row constants of a tile, fetched one tile ahead so that their global-load latency hides under the previous
        // tile's epilogue instead of sitting in front of the barrier below.  Rows past the table get +inf.
        auto fetch_meta = [&](int ii, float &m0, float &m1, float &rs) {
            m0 = 0.f; m1 = CUDART_INF_F; rs = 1.f;
            const int64_t row = (int64_t)(group + ii * ngroups) * BN + et;
            if (et < BN && ii < my_tiles && row < a.n_rows) {
                if (a.kind) rs = a.row_scale_inv[row];
                m0 = METRIC == 0 ? a.inv_ne[row] * rs : (float)a.ne2[row];
                m1 = a.penalty ? a.penalty[row] : 0.f;
            }
        };
        float nm0, nm1, nrs;
        fetch_meta(0, nm0, nm1, nrs);
        for (int i = 0; i < my_tiles; ++i) {
            const int t = group + i * ngroups, buf = i % Cfg::NBUF, mbuf = i & 1;
            const uint32_t tphase = (uint32_t)(i / Cfg::NBUF) & 1u;
            const int64_t row0 = (int64_t)t * BN;
            const int rows_here = (int)min((int64_t)BN, a.n_rows - row0);
            // meta[mbuf] was last read two tiles ago; the named barrier of the tile in between ordered those reads before
            // these writes
            if (et < G_BN) {
                s.meta[mbuf][et] = make_float2(nm0, nm1);
                if (METRIC == 1) s.mscale[mbuf][et] = nrs;
            }
            fetch_meta(i + 1, nm0, nm1, nrs);
            if (refresh && qvalid) {
                // Tighten the shared bound as the scan goes on.  At tiles 1, 2, 4, 8 ... each publishing list re-publishes its
                // boot_m-th smallest key (it only ever decreases); one tile later the thread of list 0 (one CTA per query
                // block) takes the maximum of what is published and posts it, tagged with this launch's epoch, in tau_pub;
                // every list reads that one word per tile, a tile ahead of its use.  Any such maximum has at least
                // boot_lists * boot_m >= k' keys of the union at or below it, whichever mix of older and newer values it was
                // taken from, so nobody waits for anybody.  After n rows per list the bound sits near the
                // (2 boot_m / n) quantile: a list collects a few dozen rows over the whole scan instead of filling up and
                // being pruned again and again (the prunes -- one warp selecting while seven wait at the tile barrier -- were
                // 60 % of the epilogue).  (First version: every thread read all the published keys itself; 38 K threads
                // hitting the same 64 KB at the same moment cost ~20 us per round.)
                if ((uint32_t)(tp >> 32) == a.epoch) tau_f = fminf(tau_f, __uint_as_float((uint32_t)tp));
                tp = __ldcg(a.tau_pub + qg);
                u64 *pq = a.pub + (int64_t)qg * a.boot_lists;
                if (i > 0 && (i & (i - 1)) == 0 && li < a.boot_lists) {
                    const float fm = a.boot_m == 1 ? f0 : a.boot_m == 2 ? f1 : a.boot_m == 3 ? f2 : f3;
                    // a key at or above every key with that score: "at least boot_m of my keys are at or below this one"
                    pq[li] = fm < CUDART_INF_F ? (((u64)f32_ord(fm) << 32) | 0xFFFFFFFFull) : ~0ull;
                }
                if (li == 0 && i > 1 && ((i - 1) & (i - 2)) == 0) {
                    const ulonglong2 *pp = reinterpret_cast<const ulonglong2 *>(pq);
                    u64 t0 = 0;
                    for (int l0 = 0; l0 < a.boot_lists / 2; l0 += 16) {          // 32 published keys per round of loads (64 lists: two rounds)
                        ulonglong2 x[16];
#pragma unroll
                        for (int l = 0; l < 16; ++l) x[l] = l0 + l < a.boot_lists / 2 ? __ldcg(pp + l0 + l) : make_ulonglong2(0ull, 0ull);
#pragma unroll
                        for (int l = 0; l < 16; ++l) { t0 = x[l].x > t0 ? x[l].x : t0; t0 = x[l].y > t0 ? x[l].y : t0; }
                    }
                    const float tf = ord_f32((uint32_t)(t0 >> 32));
                    if (t0 != ~0ull && tf == tf) a.tau_pub[qg] = ((u64)a.epoch << 32) | (u64)__float_as_uint(tf);
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(Cfg::EPI_THREADS) : "memory");
            mbar_wait(&s.tfull[buf], tphase);
            tc_fence_after();
            const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * G_BN + mt * G_BN + half * 128);
            const int cbase = half * 128;
            if (a.debug & 1) { tc_fence_before(); release_acc(buf); continue; }
            if (a.debug & 32) {                                // timing experiment: read the accumulator, score nothing
                uint32_t x = 0;
                for (int c = 0; c < 128 && cbase + c < rows_here; c += 32) {
                    uint32_t v[32];
                    tc_ld_32x32(taddr + (uint32_t)c, v);
                    tc_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) x ^= v[j];
                }
                if (x == 0x12345u) cnt++;
                tc_fence_before(); release_acc(buf); continue;
            }
            if (a.dump) {
                for (int c = 0; c < 128 && cbase + c < rows_here; c += 32) {
                    uint32_t v[32];
                    tc_ld_32x32(taddr + (uint32_t)c, v);
                    tc_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int r = cbase + c + j;
                        if (qvalid && r < rows_here)
                            a.dump[(int64_t)qg * a.n_rows + row0 + r] = score(__uint_as_float(v[j]), mbuf, r);
                    }
                }
                tc_fence_before();
                release_acc(buf);
                continue;
            }
            if (i == 0 && a.boot_m > 0) {
                // ---- bootstrap pass over the first tile: my m smallest keys, publish, wait, take the max ----
                // four smallest keys in scalars (an indexed array ends up in local memory)
                float best_f = CUDART_NAN_F;                   // NaN: nothing is rejected until 4 keys are held
                for (int c = 0; c < 128 && cbase + c < rows_here; c += 32) {
                    uint32_t v[32];
                    tc_ld_32x32(taddr + (uint32_t)c, v);
                    tc_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int r = cbase + c + j;
                        const float sc = score(__uint_as_float(v[j]), mbuf, r);
                        if (!(sc > best_f)) {
                            const u64 key = coarse_key(sc, (uint32_t)(row0 + r));
                            if (r < rows_here && key < b3) {
                                b3 = key;
                                if (b3 < b2) { const u64 x = b3; b3 = b2; b2 = x; }
                                if (b2 < b1) { const u64 x = b2; b2 = b1; b1 = x; }
                                if (b1 < b0) { const u64 x = b1; b1 = b0; b0 = x; }
                                best_f = ord_f32((uint32_t)(b3 >> 32));
                            }
                        }
                    }
                }
                // The accumulator is handed back right here: what this thread may still need from the tile is in b0..b3.
                // (Holding it through the publish / wait below kept the MMA thread out of this buffer for longer than
                // the MMAs of the next tile take -- a measured 14 K-cycle stall, 7 % of the kernel.)
                if (a.boot_early) { tc_fence_before(); release_acc(buf); }
                const u64 mine = a.boot_m == 1 ? b0 : a.boot_m == 2 ? b1 : a.boot_m == 3 ? b2 : b3;
                // only the first boot_lists lists (CTAs with group < boot_lists/2) publish and are waited for:
                // boot_lists * boot_m >= k' keys below the maximum is all the bound needs, and fewer lists -> lower maximum
                if (qvalid && li < a.boot_lists) a.pub[(int64_t)qg * a.boot_lists + li] = mine;
                __threadfence();
                asm volatile("bar.sync 1, %0;" ::"n"(Cfg::EPI_THREADS) : "memory");
                if (et == 0) {
                    const unsigned int want = (unsigned int)min(ngroups, (a.boot_lists + 1) / 2);
                    if (2 * group < a.boot_lists) atomicAdd(&a.arrive[qb], 1u);
                    while (*(volatile unsigned int *)&a.arrive[qb] < want) __nanosleep(32);
                    __threadfence();
                }
                asm volatile("bar.sync 1, %0;" ::"n"(Cfg::EPI_THREADS) : "memory");
                if (qvalid) {
                    u64 t0 = 0;
                    const ulonglong2 *pp = reinterpret_cast<const ulonglong2 *>(a.pub + (int64_t)qg * a.boot_lists);
                    for (int l = 0; l < a.boot_lists / 2; ++l) {
                        const ulonglong2 x = __ldcg(pp + l);
                        t0 = x.x > t0 ? x.x : t0; t0 = x.y > t0 ? x.y : t0;
                    }
                    tau_f = t0 == ~0ull ? CUDART_INF_F : ord_f32((uint32_t)(t0 >> 32));   // scores up to and including the bound pass
                }
                if (a.boot_early && qvalid) {
                    // This tile's rows at or below the bound, straight from the registers.  Four are held; if even
                    // the fourth passes, rows behind it (key > b3, so score >= its score) may have passed too and are
                    // gone: their best possible score goes to drop_min, which the certificate takes into its threshold
                    // (rescore_finalize_kernel), so a dropped row that mattered sends the query to the fp64 scan.
                    const u64 bb[4] = {b0, b1, b2, b3};
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (bb[u] != ~0ull && !(ord_f32((uint32_t)(bb[u] >> 32)) > tau_f)) mylist[cnt++] = key_to_raw(bb[u]);
                    if (b3 != ~0ull && !(ord_f32((uint32_t)(b3 >> 32)) > tau_f)) atomicMin(&a.drop_min[qg], (uint32_t)(b3 >> 32));
                }
                if (refresh) {                                 // the four smallest of tile 0 seed the float set (NaN keys: none)
                    f0 = b0 == ~0ull ? CUDART_INF_F : ord_f32((uint32_t)(b0 >> 32)); f1 = b1 == ~0ull ? CUDART_INF_F : ord_f32((uint32_t)(b1 >> 32));
                    f2 = b2 == ~0ull ? CUDART_INF_F : ord_f32((uint32_t)(b2 >> 32)); f3 = b3 == ~0ull ? CUDART_INF_F : ord_f32((uint32_t)(b3 >> 32));
                    if (!(f0 == f0)) f0 = CUDART_INF_F;
                    if (!(f1 == f1)) f1 = CUDART_INF_F;
                    if (!(f2 == f2)) f2 = CUDART_INF_F;
                    if (!(f3 == f3)) f3 = CUDART_INF_F;
                }
                if (a.boot_early) continue;                    // otherwise: the filter pass below re-reads the tile
            }
            // ---- filter pass ----
            // Lists hold raw (fp32 score bits, row) pairs: the ordered image is only built when a list is read
            // (prune, select), so an append is a predicate, one store and one add.  With ~4 % of the scores
            // passing per lane, some lane of the warp passes most of the time, so this path must stay short.
            for (int c = 0; c < 128 && cbase + c < rows_here; c += 32) {   // chunks past the tile's rows hold nothing
                uint32_t v[32];
                tc_ld_32x32(taddr + (uint32_t)c, v);
                tc_ld_wait();
                // score all 32 first and only then look for survivors: once the shared bound has tightened, whole chunks pass
                // without one, and the loop below (a branch, a store and the four-smallest insertion per element) is skipped
                bool any = false;
                const bool ins = refresh && li < a.boot_lists;  // warp-uniform: only the lists that publish keep the set current
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float sc = score(__uint_as_float(v[j]), mbuf, cbase + c + j);
                    v[j] = __float_as_uint(sc);
                    any |= !(sc > tau_f);                      // at or below the threshold, or NaN
                }
                if (any || (a.debug & 128)) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int r = cbase + c + j;
                        const float sc = __uint_as_float(v[j]);
                        if (!(sc > tau_f) && qvalid && r < rows_here) {
                            mylist[cnt++] = ((u64)v[j] << 32) | (u64)(uint32_t)(row0 + r);
                            if (ins && sc < f3) {              // branch-free insertion into the ascending four (never NaN: sc < f3)
                                f3 = fminf(f3, fmaxf(f2, sc)); f2 = fminf(f2, fmaxf(f1, sc));
                                f1 = fminf(f1, fmaxf(f0, sc)); f0 = fminf(f0, sc);
                            }
                        }
                    }
                }
                // a list within 32 of its capacity is cut back to k' (every chunk appends at most 32)
                uint32_t full_mask = __ballot_sync(MCP_FULL, cnt > CAP - 32);
                while (full_mask) {
                    const int src = __ffs(full_mask) - 1;
                    full_mask &= full_mask - 1;
                    u64 *lp = (u64 *)__shfl_sync(MCP_FULL, (u64)(uintptr_t)mylist, src);
                    const int n = __shfl_sync(MCP_FULL, cnt, src);
                    __syncwarp();
                    const u64 thr = warp_prune_list<KPL>(lp, n, a.kprime, lane);   // ordered key of the worst survivor
                    if (lane == src) { cnt = a.kprime; tau_f = ord_f32((uint32_t)(thr >> 32)); }
                    __syncwarp();
                }
            }
            tc_fence_before();
            release_acc(buf);
        }
        if (qvalid && !a.dump) a.list_cnt[(int64_t)qg * L + li] = cnt;
    }
    tc_fence_before();
    __syncthreads();
    if (C > 1) cluster_sync_all();                             // nobody leaves while a peer may still write its smem / barriers
    if (warp == 1) {
        tc_fence_after();
        if constexpr (pair) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

// ---- host side ---------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct GemmPlan {
    bool valid = false;
    CUtensorMap tmB;           // box = 256 / tmB_cluster rows
    int tmB_cluster = 0, tmB_bn = 0;
    bool balance_tiles = false; // option "gemm_balance" (measured: no gain -- at Q = 256 the tensor pipe is a co-bottleneck and shorter tiles waste MMA columns)
    int cluster_opt = 0;       // option "gemm_cluster": 0 = automatic, else force 1 / 2 / 4
    bool coop_attr = true;     // option "gemm_coop"
    int sync_tiles = G_DRIFT;  // option "gemm_sync": keep the CTAs of a group within this many tiles of each other (0 = off)
    uint32_t epoch = 0;
    bool allow_pair = true;    // option "gemm_pair": clusters of 2 run as one tcgen05 CTA pair (cta_group::2)
    bool mc_ok = false;        // the row table is memory this library allocated on this device: only such a table is read with
                               // multicast loads (TMA multicast loads here read only this GPU's memory, since on B200 a TMA
                               // multicast load from another GPU's memory hung the GPU); any other table takes plain loads
    const void *rows = nullptr;
    int64_t n_rows = 0, ld = 0;
    int dim = 0, kind = -1;
    PFN_encodeTiled encode = nullptr;
    int coop = -1;             // device supports cooperative launch (-1 = not asked yet)
    const void *arrive_ws = nullptr;   // workspace (pointer, capacity) whose arrive-counter pair is known to be initialised
    size_t arrive_ws_cap = 0;
    int arrive_slot = 0;       // which counter set the next bootstrap launch uses (it clears the other one)
    bool allow_boot = true;    // option "gemm_bootstrap"
    bool l2_prefetch = false;  // option "gemm_l2_prefetch"
    int debug = 0;             // option "gemm_debug" (timing experiments, wrong results)
    const u64 *last_tau_pub = nullptr;   // the last launch's shared-bound words (NULL when that launch did not refresh the bound)
    uint32_t last_epoch = 0;
    char err[256] = {0};
};

// a TMA tensor map needs a 16-byte aligned base and row pitch; the innermost extent (dim) is free -- columns past it read as zero
static inline bool gemm_supported(int kind, const void *base, int64_t ld) {
    const int64_t pitch = ld * (kind ? 2 : 4);
    return base && (pitch % 16) == 0 && (((uintptr_t)base) & 15) == 0;
}

static int gemm_encode_2d(GemmPlan &p, CUtensorMap *tm, int kind, const void *base, int64_t rows, int dim, int64_t ld, int box_rows) {
    if (!p.encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
            snprintf(p.err, sizeof p.err, "cuTensorMapEncodeTiled entry point unavailable (%s)", cudaGetErrorString(e));
            return -2;
        }
        p.encode = (PFN_encodeTiled)fn;
    }
    cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)ld * (kind ? 2 : 4)};
    cuuint32_t box[2] = {(cuuint32_t)(kind ? 2 * G_BK : G_BK), (cuuint32_t)box_rows};     // 128 bytes of a row x box_rows rows
    cuuint32_t estr[2] = {1, 1};
    CUresult r = p.encode(tm, kind ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)base, gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(p.err, sizeof p.err, "cuTensorMapEncodeTiled failed with CUresult %d (rows %lld dim %d ld %lld)", (int)r,
                 (long long)rows, dim, (long long)ld);
        return -2;
    }
    return 0;
}

static inline int gemm_kpl_for(int kprime) { return kprime <= 64 ? 4 : kprime <= 128 ? 8 : 16; }
static inline int gemm_max_kprime() { return 256; }

static const void *gemm_kernel_ptr(int mt, int kpl, int metric, bool pair) {
#define G_PICK(MTV, KPL, PR) (metric == 0 ? (const void *)gemm_tf32_kernel<MTV, KPL, 0, PR> : (const void *)gemm_tf32_kernel<MTV, KPL, 1, PR>)
    (void)mt;                                        // 256-query CTAs (MT = 2) measured slower and are no longer instantiated
    if (pair) return kpl == 4 ? G_PICK(1, 4, 1) : kpl == 8 ? G_PICK(1, 8, 1) : G_PICK(1, 16, 1);
    return kpl == 4 ? G_PICK(1, 4, 0) : kpl == 8 ? G_PICK(1, 8, 0) : G_PICK(1, 16, 0);
#undef G_PICK
}

// One launch over <= G_MAX_QBLOCKS * 128 queries.  Lists come back as [nq][L][cap] in `ws`.
// Returns 0, or a negative MCP_E* code with the message in plan.err.
// kind 0: rows / queries are fp32 with pitches ld / ldq floats; kind 1: scaled fp16 images with pitches in fp16 elements
// and their inverse scales in q_scale_inv [nq] / row_scale_inv [n_rows].
template <class BufT>
static int gemm_tf32_run(GemmPlan &p, BufT &ws, cudaStream_t st, int n_sm, int kind, const void *rows, int64_t n_rows, int dim,
                         int64_t ld, const void *queries, int64_t ldq, int nq_map /* rows of the query operand: >= nq, zero rows past nq */,
                         int nq, const float *inv_nq, const float *nq2f,
                         const float *inv_ne, const double *ne2, const float *penalty, int metric, int kprime,
                         u64 **out_lists, int **out_cnt, int *out_L, int *out_cap, float *dump = nullptr,
                         unsigned int *drop_min = nullptr /* [nq], all-ones on entry: enables the cross-CTA bootstrap */,
                         const float *q_scale_inv = nullptr, const float *row_scale_inv = nullptr) {
    if (!gemm_supported(kind, rows, ld) || !gemm_supported(kind, queries, ldq)) {
        snprintf(p.err, sizeof p.err, "gemm scan needs 16-byte aligned operands and row pitches (kind %d, ld %lld, ldq %lld)", kind, (long long)ld, (long long)ldq);
        return -1;
    }
    if (kprime > gemm_max_kprime()) { snprintf(p.err, sizeof p.err, "gemm_tf32 keeps at most %d candidates (asked %d)", gemm_max_kprime(), kprime); return -1; }
    if (nq > G_MAX_QBLOCKS * G_BM) { snprintf(p.err, sizeof p.err, "gemm_tf32 launch takes at most %d queries", G_MAX_QBLOCKS * G_BM); return -1; }
    const int MTsel = 1;                                         // 128-query CTAs
    CUtensorMap tmA;
    int rc = gemm_encode_2d(p, &tmA, kind, queries, nq_map, dim, ldq, MTsel * G_BM);
    if (rc) return rc;
    const int nqb = (nq + MTsel * G_BM - 1) / (MTsel * G_BM);
    int n_tiles = (int)((n_rows + G_BN - 1) / G_BN);
    int ngroups = std::max(1, std::min(n_sm / nqb, n_tiles));
    const int kpl = gemm_kpl_for(kprime), cap = kpl * 32;
    size_t smem = sizeof(GemmSmemT<1>) + 1024;
    const int threads = GemmCfg<1>::THREADS;
    cudaError_t e = cudaSuccess;
    // cluster size: pairs of query-block CTAs share (multicast) their row tiles.  Measured on config 3's per-GPU share:
    // C = 1 7.37 ms, C = 2 6.78 ms, C = 4 6.88 ms (clusters of 4 leave SMs idle: a cluster must fit inside one GPC)
    int C = p.cluster_opt ? p.cluster_opt : (nqb % 2 == 0 ? 2 : 1);
    if (nqb % C != 0 || dump || (p.debug & 2)) C = 1;
    // clusters of two run as one tcgen05 CTA pair (option gemm_pair, on by default) unless tiles are being balanced
    bool pair = C == 2 && MTsel == 1 && p.allow_pair;
    if (C > 1 && !pair && !p.mc_ok) C = 1;           // the non-pair cluster shape multicasts the row tile: library-owned tables only
    const void *func = gemm_kernel_ptr(MTsel, kpl, metric, pair);
    if (pair) smem = sizeof(GemmSmemT<1, 1>) + 1024;
    e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { snprintf(p.err, sizeof p.err, "gemm_tf32 smem attribute: %s", cudaGetErrorString(e)); return -2; }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attrs[2];
    cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = st; cfg.attrs = attrs;
    if (C > 1) {
        // a cluster lives inside one GPC, so fewer CTAs may be co-resident than there are SMs: size the grid to what fits
        attrs[0].id = cudaLaunchAttributeClusterDimension;
        attrs[0].val.clusterDim.x = C; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
        cfg.numAttrs = 1; cfg.gridDim = dim3(nqb * ngroups);
        int maxc = 0;
        if (cudaOccupancyMaxActiveClusters(&maxc, func, &cfg) == cudaSuccess && maxc > 0) {
            if (maxc * C < nqb * ngroups) ngroups = maxc * C / nqb;
        } else (void)cudaGetLastError();
        if (ngroups < 1) {
            C = 1; ngroups = std::max(1, std::min(n_sm / nqb, n_tiles));
            if (pair) {
                pair = false;
                func = gemm_kernel_ptr(MTsel, kpl, metric, false);
                smem = sizeof(GemmSmemT<1>) + 1024; cfg.dynamicSmemBytes = smem;
                e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) { snprintf(p.err, sizeof p.err, "gemm_tf32 smem attribute: %s", cudaGetErrorString(e)); return -2; }
            }
        }
    }
    // Balanced tiles (option gemm_balance, off by default): shrink the tile height until every group runs the same number of
    // rounds, BN = ceil(N / (rounds * groups)).  Measured on config 2: 155.8 vs 154.1 us -- no gain: the MMA is N = 256 either
    // way, and at 256 queries the tensor pipe (a TF32 M128 N256 K8 MMA retires every ~171 cycles here, not 128) is as much the
    // bound as HBM is, so shorter tiles only add tensor work.
    const int rounds = (n_tiles + ngroups - 1) / ngroups;
    int bn = (int)((n_rows + (int64_t)rounds * ngroups - 1) / ((int64_t)rounds * ngroups));
    const int bq = pair ? 16 : 2 * C;                // pair: the M = 256 MMA takes N in steps of 16, and each CTA holds N / 2 rows
    bn = std::min(G_BN, std::max(bq, (bn + bq - 1) / bq * bq));
    if (!p.balance_tiles) bn = G_BN;
    n_tiles = (int)((n_rows + bn - 1) / bn);
    ngroups = std::max(1, std::min(ngroups, n_tiles));
    if (!p.valid || p.rows != rows || p.n_rows != n_rows || p.ld != ld || p.dim != dim || p.tmB_cluster != C || p.tmB_bn != bn || p.kind != kind) {
        if ((rc = gemm_encode_2d(p, &p.tmB, kind, rows, n_rows, dim, ld, bn / C))) return rc;
        p.rows = rows; p.n_rows = n_rows; p.ld = ld; p.dim = dim; p.tmB_cluster = C; p.tmB_bn = bn; p.kind = kind; p.valid = true;
    }
    const int L = 2 * ngroups;                       // one list per (CTA, query, half tile)
    // workspace: [arrive counters, two sets][lists][list counts][published bounds].  The counters sit first so that their
    // place does not move with the launch shape: a bootstrap launch uses one set (zero on entry) and clears the other for
    // the launch after it, which saves a memset node per search; only a new workspace is cleared from the host.
    // header: arrive counter pair, the progress words, the shared-bound words.  All three are only ever matched against this
    // launch's epoch (or are zero), so they sit at fixed offsets -- a place that moved with the launch shape could hold another
    // launch's list data that happens to look like a tagged word -- and a new workspace is cleared once.
    const size_t tau_off = 256 + sizeof(u64) * G_MAX_GROUPS * G_MAX_QBLOCKS;
    const size_t hdr = tau_off + sizeof(u64) * G_MAX_QBLOCKS * G_BM;
    static_assert(2 * G_MAX_QBLOCKS * sizeof(unsigned int) <= 256, "arrive counter pair");
    const size_t list_bytes = sizeof(u64) * (size_t)nq * L * cap;
    const size_t cnt_off = (hdr + list_bytes + 255) & ~(size_t)255;
    const size_t pub_off = (cnt_off + sizeof(int) * (size_t)nq * L + 255) & ~(size_t)255;
    const size_t end_off = (pub_off + sizeof(u64) * (size_t)nq * 64 + 255) & ~(size_t)255;
    if ((rc = ws.ensure(end_off))) { snprintf(p.err, sizeof p.err, "workspace allocation failed"); return rc; }
    unsigned char *wsb = reinterpret_cast<unsigned char *>(ws.p);
    GemmArgs a;
    a.n_rows = n_rows; a.dim = dim; a.nq = nq; a.nqb = nqb; a.kprime = kprime; a.cap = cap;
    a.inv_nq = inv_nq; a.nq2f = nq2f; a.inv_ne = inv_ne; a.ne2 = ne2; a.penalty = penalty;
    a.lists = reinterpret_cast<u64 *>(wsb + hdr);
    a.list_cnt = reinterpret_cast<int *>(wsb + cnt_off);
    a.pub = reinterpret_cast<u64 *>(wsb + pub_off);
    a.tau_pub = reinterpret_cast<u64 *>(wsb + tau_off);
    a.arrive = nullptr; a.arrive_next = nullptr;
    a.dump = dump;
    a.rows_b = (const unsigned char *)rows; a.row_pitch_b = ld * (kind ? 2 : 4); a.row_len_b = dim * (kind ? 2 : 4);
    a.l2_prefetch = p.l2_prefetch ? 1 : 0; a.debug = p.debug; a.cluster = C; a.bn = bn;
    a.pair = pair ? 1 : 0;
    a.kind = kind; a.ek = kind ? 2 * G_BK : G_BK;
    a.q_scale_inv = q_scale_inv; a.row_scale_inv = row_scale_inv;
    if (kind && (!q_scale_inv || !row_scale_inv)) { snprintf(p.err, sizeof p.err, "fp16 scan without operand scales"); return -1; }
    a.idesc_single = g_idesc(kind, G_BM, G_BN);
    a.idesc_pair = g_idesc(kind, 2 * G_BM, bn);
    // bootstrap needs every CTA of a query block resident at once: cooperative launch, grid <= #SMs (1 CTA / SM)
    // lists that publish: 32 when four keys each cover k' (fewer lists -> lower maximum), else up to 64 (k' up to 256)
    int boot_lists = std::min(L, 32) & ~1;                       // even: both halves of a CTA
    if (boot_lists && (kprime + boot_lists - 1) / boot_lists > G_BOOT_MAX) boot_lists = std::min(L, 64) & ~1;
    const int m = boot_lists ? (kprime + boot_lists - 1) / boot_lists : G_BOOT_MAX + 1;
    if (p.coop < 0) {
        int dev = 0, v = 0;
        cudaGetDevice(&dev);
        p.coop = (cudaDeviceGetAttribute(&v, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess && v) ? 1 : 0;
    }
    a.drop_min = drop_min;
    // progress words: only when several clusters share a group's tiles, every CTA is co-resident (the wait spins) and the
    // table is large enough for drift to matter
    a.progress = nullptr; a.drift = G_DRIFT;
    if (++p.epoch == 0) p.epoch = 1;                 // never 0: stale words of other launches (or fresh memory) must not match
    a.epoch = p.epoch;
    a.boot_m = (!dump && drop_min && !(p.debug & (1 | 16 | 32)) && p.coop == 1 && m <= G_BOOT_MAX && nqb * ngroups <= n_sm && p.allow_boot) ? m : 0;
    a.boot_lists = boot_lists;
    // Early hand-back keeps only four rows per (query, half tile) of the first tile and lets the certificate account for the
    // rest (drop_min).  That is a good trade only when the bound sits far above the final top-k', i.e. when the rows the
    // bound was estimated from (boot_lists x 128) are a small part of the table; small tables re-read the tile instead.
    a.boot_early = (a.boot_m && n_rows >= (int64_t)16 * boot_lists * 128) ? 1 : 0;
    if (a.boot_m && p.sync_tiles && nqb > C && ngroups <= G_MAX_GROUPS && n_rows * (int64_t)dim * (kind ? 2 : 4) > (int64_t)64 << 20) {
        a.progress = reinterpret_cast<u64 *>(wsb + 256);
        a.drift = p.sync_tiles;
    }
    if (a.boot_m) {
        if (p.arrive_ws != (const void *)wsb || p.arrive_ws_cap != (size_t)ws.cap) {
            e = cudaMemsetAsync(wsb, 0, hdr, st);
            p.arrive_ws = wsb; p.arrive_ws_cap = (size_t)ws.cap; p.arrive_slot = 0;
        }
        unsigned int *pair = reinterpret_cast<unsigned int *>(wsb);
        a.arrive = pair + p.arrive_slot * G_MAX_QBLOCKS;
        a.arrive_next = pair + (1 - p.arrive_slot) * G_MAX_QBLOCKS;
        p.arrive_slot ^= 1;
    }
    int nattr = C > 1 ? 1 : 0;
    // option gemm_coop = 0 drops the cooperative attribute (profiling only: Nsight Compute cannot replay the cluster + cooperative
    // launch; with one CTA per SM on an otherwise idle GPU the CTAs are co-resident anyway, but nothing guarantees it)
    if (a.boot_m && p.coop_attr) { attrs[nattr].id = cudaLaunchAttributeCooperative; attrs[nattr].val.cooperative = 1; ++nattr; }
    cfg.numAttrs = nattr; cfg.gridDim = dim3(nqb * ngroups);
    void *kargs[3] = {(void *)&tmA, (void *)&p.tmB, (void *)&a};
    if (e == cudaSuccess) {
        // plain / cooperative launches keep their classic entry points: Nsight Compute (2025.2, driver 580) profiles those
        // fine, while the same kernel launched through cudaLaunchKernelExC with cluster + cooperative attributes faulted
        // under it (never outside it): it is the cooperative attribute ncu trips over, so captures use option gemm_coop = 0 (profiles/r1i_summary.md)
        if (C > 1) e = cudaLaunchKernelExC(&cfg, func, kargs);
        else if (a.boot_m && p.coop_attr) e = cudaLaunchCooperativeKernel(func, dim3(nqb * ngroups), dim3(threads), kargs, smem, st);
        else e = cudaLaunchKernel(func, dim3(nqb * ngroups), dim3(threads), kargs, smem, st);
    }
    if (e != cudaSuccess) { snprintf(p.err, sizeof p.err, "gemm_tf32 launch (cluster %d, grid %d): %s", C, nqb * ngroups, cudaGetErrorString(e)); return -2; }
    *out_lists = a.lists; *out_cnt = a.list_cnt; *out_L = L; *out_cap = cap;
    // the list selection that follows may drop keys above the final shared bound -- only when this launch kept one
    const bool refreshed = a.boot_m > 0 && a.boot_early && !(p.debug & 64);
    p.last_tau_pub = refreshed ? a.tau_pub : nullptr;
    p.last_epoch = a.epoch;
    return 0;
}
