// float_top3.cu — A1 (float descriptor mutual-NN, superglue.py:259-260, :268-278) as ONE fp16 tensor-core pass per frame
// pair with certified results.
//
// Why: the split-fp16 kernel (float_tc.cu) needs three MMAs per k-step to be fp32-grade and a cross-lane column
// reduction in its epilogue; both cap it at ~20 % of the algorithmic tensor roofline.  Here the contraction runs ONCE
// on the hi halves (fp16 x fp16 -> fp32 in TMEM) and the epilogue keeps, per row, the TOP-3 packed keys, and per column
// (single-pass form, k_float_top3_tc<1>) the best / runner-up 16-bit value of every slab of 16 rows.  A rigorous bound
// delta on |exact - tensor-core score| (float_ops.cuh) then decides what the approximate keys already prove:
//   rows   : best - second > 2 delta: the arg-max is certain; best - third > 2 delta: it is one of two columns -> both
//            are rescored exactly (fp32, hi + lo); otherwise the row is listed and rescored from the slab bound
//            (k_float_slabscan), or, in the two-direction form, rescanned over all columns (k_float_fullscan);
//   columns: is the row the arg-max of its best column?  k_float_colmerge / k_float_colcheck (see there).
// The same interval test guards the threshold and the ratio decisions.  Index sets are therefore those of the exact
// score, never of the fp16 one; score values are exact where a decision depended on them and within delta elsewhere.
//
// History (measured on B200, 256 pairs x 2048^2).  The two-direction form (k_float_top3_tc<0>: rows of a against b, then
// rows of b against a, so that the column side is a row problem and the epilogue never leaves the thread) was bound by
// the tensor pipe doing twice the algorithmic work: 0.955 ms.  A first single-pass attempt that reduced the columns with
// redux.sync was slower (1.01 ms: CREDUX + uniform->vector moves; a 32-bit shuffle butterfly needs ~5.3 ALU operations
// per score).  What works is 16-bit packing: the top 16 of the 20 live key bits, two columns per register, transposed
// through a per-warp shared-memory tile and reduced with VIMNMX.U16x2 (1.2 ALU operations per score): 0.527 ms.
//
// k_float_top3_tc — one CTA = ([direction,] frame pair, 256-row block of the "row" frame), 18 warps:
//   warp 16   : TMA producer.  The row block (2 x 128 rows x 256 fp16 = 128 KB) is loaded once and stays resident;
//               the other frame streams through a 4-stage ring of [128 x 64] boxes (16 KB, 128-byte swizzle);
//   warp 17   : MMA issuer: per 128-column tile 2 x 16 tcgen05.mma (128x128x16, kind::f16) into two of four TMEM
//               accumulators (2 row halves x double buffer = 512 columns); tcgen05.commit frees ring stages / publishes;
//   warps 0-15: epilogue, thread = row, 32 columns at a time: tcgen05.ld 32x32b, key = bits(fma(v, sc, bias)) * 4096 +
//               (4095 - col) (FMA pipe), tournament top-2 of the chunk + top-3 of the chunk maxima (ALU pipe); single
//               pass: pack, staging tile, transposed read-back, packed tournament, slab results to global memory.
#include <cuda.h>
#include <cuda_fp16.h>
#include "float_ops.cuh"
#include "sm100_ptx.cuh"

namespace {

using namespace sm100;

constexpr int T3_BN = 128, T3_BK = 64, T3_KCH = FS_FLOAT_DIM / T3_BK;  // 4 k-chunks of 64
constexpr int T3_ROWS = 256;                                          // rows per CTA (two MMA halves of 128)
#ifndef T3_STAGES_N
#define T3_STAGES_N 4
#endif
constexpr int T3_STAGES = T3_STAGES_N;  // one tile; (6 stages with run-time stage indexing measured SLOWER: the MMA issue thread is the
                              // critical path — one tile of MMAs takes as long to issue as to execute — so it gets no extra work)
// Epilogue warps: T3_CSPLIT warps share each 32-row TMEM lane quarter of each half and split the tile's 128 columns
// (CSPLIT = 2: 16 warps, four per scheduler: the epilogue is a chain of dependent phases (tcgen05.ld -> keys -> staging
// tile -> transposed read -> tournaments), and two warps per scheduler could not cover each other's latencies).
#ifndef T3_CSPLIT
#define T3_CSPLIT 2
#endif
constexpr int T3_EPI_WARPS = 8 * T3_CSPLIT;
constexpr int T3_THREADS = (T3_EPI_WARPS + 2) * 32;
// (registers: warps are allocated in groups of four, so 18 warps cost 20: 65536 / 640 = 102 -> 96 per thread;
// __maxnreg__(112) compiles and then fails to launch)
constexpr int T3_TMA_WARP = T3_EPI_WARPS, T3_MMA_WARP = T3_EPI_WARPS + 1;  // warps below are the epilogue
constexpr int T3_CW = T3_BN / T3_CSPLIT;     // columns of a tile per epilogue warp
// CTA pair (tcgen05 cta_group::2): the two CTAs of a cluster own neighbouring 256-row blocks and stream the SAME frame, so
// one tcgen05.mma with M = 256 serves a 128-row half of BOTH: each CTA keeps only its 64 columns of every streamed box and
// its tensor core reads 4 KB (own rows) + 2 KB (own columns) per instruction instead of 4 + 4 KB.  The kernel is bound by
// shared-memory bandwidth (operand reads were half of it), and the ring shrinks to 32 KB, which pays for a second staging
// tile per epilogue warp.  Rank 0 issues every MMA; TMA loads of either CTA signal rank 0's barriers (.cta_group::2).
#ifndef T3_2CTA
#define T3_2CTA 1
#endif
constexpr int T3_B_STAGE_BYTES = T3_2CTA ? 64 * T3_BK * 2 : 128 * T3_BK * 2;  // streamed-frame bytes per ring stage per CTA
#ifndef T3_NBUF_N
#define T3_NBUF_N (T3_CSPLIT == 1 ? 2 : 1)  // (CTA pair, 16 warps: a second tile measured slower, 0.563 vs 0.533 ms)
#endif
constexpr int T3_NBUF = T3_NBUF_N;  // staging tiles per warp (single pass)
constexpr uint32_t T3_IDESC2 = make_idesc_f16(256, T3_BN, /*F16*/ 0);
constexpr int T3_BOX_BYTES = 128 * T3_BK * 2;               // 16 KB: one [128 x 64] fp16 box
constexpr int T3_A_BYTES = 2 * T3_KCH * T3_BOX_BYTES;       // 128 KB
constexpr int T3_TMEM_COLS = 512;
constexpr uint32_t T3_IDESC = make_idesc_f16(128, T3_BN, /*F16*/ 0);

struct T3Barriers {
    uint64_t a_full;  // CTA pair: rank 0's collects both CTAs' row blocks
    uint64_t full[T3_STAGES], empty[T3_STAGES];
    uint64_t tmem_full[2], tmem_empty[2];
    uint32_t tmem_base;
    uint32_t pad;
};
constexpr int T3_SMEM_BYTES = 1024 /*alignment slack*/ + T3_A_BYTES + T3_STAGES * T3_B_STAGE_BYTES + 256 /*barriers*/;
// single-pass variant: per epilogue warp T3_NBUF staging tiles of 32 rows x 32 packed 16-bit scores (2 KB each)
constexpr int T3_STAGE_BYTES = 32 * 64;
constexpr int T3_SP_SMEM_BYTES = T3_SMEM_BYTES + T3_EPI_WARPS * T3_NBUF * T3_STAGE_BYTES;

// (best, runner-up) of two disjoint key sets from their (best, runner-up): 3 ops
__device__ __forceinline__ void top2_merge(uint32_t& h, uint32_t& s, uint32_t h2, uint32_t s2) {
    s = __vimax3_u32(min(h, h2), s, s2);
    h = max(h, h2);
}

// Tournament over the 32 keys of one chunk: its best key (with column) and its runner-up key.  77 min/max ops per 32
// columns, against 128 for a streaming top-3: the ALU pipe is what bounds this kernel's epilogue.
__device__ __forceinline__ void chunk_top2(const uint32_t (&e)[32], uint32_t& best, uint32_t& second) {
    uint32_t h[16], s[16];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        h[i] = max(e[2 * i], e[2 * i + 1]);
        s[i] = min(e[2 * i], e[2 * i + 1]);
    }
#pragma unroll
    for (int w = 8; w >= 1; w >>= 1)
#pragma unroll
        for (int i = 0; i < w; i++) top2_merge(h[i], s[i], h[i + w], s[i + w]);
    best = h[0];
    second = s[0];
}

__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(v) : "r"(addr) : "memory");
    return v;
}

// insert one key into the running sorted triple (k1 >= k2 >= k3): 5 ops
__device__ __forceinline__ void top3_insert(uint32_t& k1, uint32_t& k2, uint32_t& k3, uint32_t e) {
    const uint32_t t1 = min(k1, e);
    k1 = max(k1, e);
    const uint32_t t2 = min(k2, t1);
    k2 = max(k2, t1);
    k3 = max(k3, t2);
}

// per-thread state of the epilogue across the tiles of a row
struct T3Epi {
    // k1 >= k2 >= k3: the three largest CHUNK maxima (each from a different 32-column chunk, column in the low bits);
    // ks: the largest chunk runner-up.  Every column other than those of k1, k2, k3 scores <= max(k3, ks).
    uint32_t k1, k2, k3, ks;
    // single-pass column side: staging addresses (bytes, shared window) and the slab's output row
    uint32_t st_base, st_wr[4], st_rd[4], rowmask, rowbits;
    bool warp_partial;
    uint32_t* col_out;
};

// one chunk of 32 columns of one row (r: its 32 accumulator values).  MASK: columns >= n1 may exist in this chunk, or
// (single pass) rows >= n0 in this warp's 32-row group.  c: chunk counter (selects the staging tile when there are two).
// MODE: 0 = rows only (two-direction form), 1 = single pass, float scores (column side = top 16 bits of the keys),
// 2 = single pass, Hamming (+-1 operands: the accumulator is the exact integer dot = 256 - 2 distance; the column side's
// 16-bit value is (129 + dot / 2) << 4 | 15 - row-in-slab, so the slab maximum IS the slab's best row, lowest row first
// on equal distance, and no certification is needed).
// after_keys(): called once the accumulator values r[] are dead (the caller prefetches the next chunk into them).
template <int MODE, bool MASK, class F>
__device__ __forceinline__ void t3_chunk(const uint32_t (&r)[32], int col0, int c, int n1, float sc, uint32_t key_mul,
                                         T3Epi& es, F&& after_keys) {
    constexpr bool SP = MODE != 0;
    // keys of this chunk carry (4095 - j), a compile-time constant per register; the chunk's best key is shifted to the
    // global column afterwards
    const int nvalid = n1 - col0;
    if (MASK && nvalid <= 0) {  // (warp-uniform) the slab results of an empty chunk are never read
        after_keys();
        return;
    }
    uint32_t e[32];
#pragma unroll
    for (int j = 0; j < 32; j++)
        e[j] = __float_as_uint(fmaf(__uint_as_float(r[j]), sc, fs::KEY_BIAS)) * key_mul + (uint32_t)(4095 - j);
    if (MASK) {
#pragma unroll
        for (int j = 0; j < 32; j++) e[j] = (j < nvalid) ? e[j] : 0u;
    }
    uint32_t x[16];
    if constexpr (SP) {
        // top 16 bits of the keys (= top 16 of the 20 live score bits), two columns per register
        uint32_t pk[16];
        if constexpr (MODE == 2) {
            // 2^23 + 129 + dot / 2: the low mantissa bits are the integer 129 + dot / 2 in [1, 257]; * 16 + row bits on the
            // FMA pipe (key_mul >> 8 = 16 is a run-time value, so the multiply-add stays an IMAD)
            const uint32_t mul16 = key_mul >> 8;
#pragma unroll
            for (int j = 0; j < 16; j++) {
                uint32_t t0 = __float_as_uint(fmaf(__uint_as_float(r[2 * j]), 0.5f, 8388737.f)) * mul16 + es.rowbits;
                uint32_t t1 = __float_as_uint(fmaf(__uint_as_float(r[2 * j + 1]), 0.5f, 8388737.f)) * mul16 + es.rowbits;
                if (MASK) {
                    t0 = (2 * j < nvalid) ? t0 : 0u;
                    t1 = (2 * j + 1 < nvalid) ? t1 : 0u;
                }
                pk[j] = __byte_perm(t0, t1, 0x5410);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 16; j++) pk[j] = __byte_perm(e[2 * j], e[2 * j + 1], 0x7632);
        }
        if (MASK) {  // rows >= n0 of the last, partially valid 32-row group
#pragma unroll
            for (int j = 0; j < 16; j++) pk[j] &= es.rowmask;
        }
        // two tiles: the reads of this one (two chunks ago) are ordered before these writes by the __syncwarp of the chunk
        // in between; one tile: by the __syncwarp here
        const uint32_t sbuf = es.st_base + (c & (T3_NBUF - 1)) * T3_STAGE_BYTES;
        if (T3_NBUF == 1) __syncwarp();
#pragma unroll
        for (int cc = 0; cc < 4; cc++)
            sts128(sbuf + es.st_wr[cc], pk[4 * cc], pk[4 * cc + 1], pk[4 * cc + 2], pk[4 * cc + 3]);
        after_keys();
        __syncwarp();
        // transposed read-back; the row tournament below runs while these loads are in flight
#pragma unroll
        for (int m = 0; m < 16; m++) x[m] = lds32(sbuf + es.st_rd[m & 3] + m * 128);
    }
    if constexpr (!SP) after_keys();
    uint32_t cb, cs;
    chunk_top2(e, cb, cs);
    // (4095 - j) - col0 = 4095 - (col0 + j): no borrow out of the low 12 bits; an empty chunk stays 0
    cb -= cb ? (uint32_t)col0 : 0u;
    top3_insert(es.k1, es.k2, es.k3, cb);
    es.ks = max(es.ks, cs);
    if constexpr (SP) {
        uint32_t h[8], s2[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            h[i] = __vmaxu2(x[2 * i], x[2 * i + 1]);
            s2[i] = __vminu2(x[2 * i], x[2 * i + 1]);
        }
#pragma unroll
        for (int w = 4; w >= 1; w >>= 1)
#pragma unroll
            for (int i = 0; i < w; i++) {
                s2[i] = __vimax3_u16x2(__vminu2(h[i], h[i + w]), s2[i], s2[i + w]);
                h[i] = __vmaxu2(h[i], h[i + w]);
            }
        // (best | second << 16) of column 2 cp and of column 2 cp + 1
        const uint2 o2 = make_uint2(__byte_perm(h[0], s2[0], 0x5410), __byte_perm(h[0], s2[0], 0x7632));
        *reinterpret_cast<uint2*>(es.col_out + col0) = o2;
    }
}

// Two CTAs that own neighbouring 256-row blocks of the same (direction, pair) form a cluster and share the streamed
// frame: each loads HALF of every [128 x 64] box (64 rows) and multicasts it into both shared memories.  L2 -> SM traffic
// halves (measured before: with MMAs and epilogue switched off the kernel ran at 13.7 TB/s, the L2 cap, so the real
// kernel queued behind its own loads).  A ring stage is free when BOTH CTAs' MMAs have read it: tcgen05.commit
// arrives on both CTAs' empty barriers (count 2).
//
// SP = true is the SINGLE-PASS form (fs_float_match with the mutual check): only direction 0 runs (half the MMAs), and the
// column side of superglue.py:268 is reduced inside the epilogue.  Every epilogue warp owns a 32 x 32 block of scores per
// 32-column chunk (thread = row); it packs the top 16 bits of its 32 keys two to a register (PRMT), writes them to a
// private 2 KB staging tile (STS.128, swizzled, conflict free), and reads the tile back TRANSPOSED (LDS.32: lane = column
// pair x row parity, 16 rows each).  A tournament of packed 16-bit min / max (VIMNMX.U16x2, VIMNMX3.U16x2: two columns
// per instruction, 1.2 ALU operations per score) gives, per column, the best and the runner-up VALUE over a "slab" of 16
// rows (the rows of one parity inside a 32-row group).  Slab results go to global memory as they are
// (colraw [pair][slab][column] = best16 | second16 << 16, 8-byte coalesced stores); k_float_colmerge folds the slabs of a
// column into (three largest slab maxima with slab ids, largest slab runner-up), and k_float_colcheck decides for every
// row whether it is the arg-max of its best column — from those keys alone when the gaps prove it, by exact rescoring of
// at most two slabs (32 rows) when they do not, and by an exact column scan otherwise.
template <int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(T3_THREADS, 1)
k_float_top3_tc(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_half,
                const void* __restrict__ operands, int n_max, const int32_t* __restrict__ counts,
                const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b, int n_pairs_total, int pair0,
                uint32_t key_mul, int probe, long long* __restrict__ trace, uint4* __restrict__ keys,
                uint32_t* __restrict__ colraw, int cpitch, int nslab) {
    // colraw [pair][nslab][cpitch]: cpitch = n_max rounded up to the 128-column tile, nslab = 2 per 32-row group
    constexpr bool SP = MODE != 0;
#define T3_TRACE(slot)                                                                   \
    do {                                                                                 \
        if (trace && blockIdx.x == 0 && blockIdx.z == 0 && (int)blockIdx.y == (probe >> 8)) \
            trace[slot] = clock64();                                                     \
    } while (0)
    // probe (measurement only, fs_float_top3_keys): bit 0 = issue no MMAs, bit 1 = skip the top-3 network
    // key_mul = 4096 comes in as an argument so that bits * key_mul + idx stays an IMAD (FMA pipe) instead of being
    // strength-reduced to a shift-add on the ALU pipe, which the top-3 network already saturates
    if (threadIdx.x == 0) T3_TRACE(0);
    const int p = pair0 + blockIdx.y, dir = SP ? 0 : blockIdx.z;
    const int fa = dir ? pair_b[p] : pair_a[p], fb = dir ? pair_a[p] : pair_b[p];
    const int n0 = min(counts[fa], n_max), n1 = min(counts[fb], n_max);
    const int m0 = blockIdx.x * T3_ROWS;
    if ((int)(blockIdx.x & ~1u) * T3_ROWS >= n0) return;  // the whole cluster is past the last row
    const bool ghost = m0 >= n0;  // no rows of its own: still loads its half of every box and frees ring stages
    const uint32_t rank = cluster_ctarank();
    const int ntiles = (n1 + T3_BN - 1) / T3_BN;
    const int nhalf = (m0 + 128 < n0) ? 2 : 1;  // the second 128-row half may be empty

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                 // [half][kc] boxes of 16 KB
    uint8_t* sB = smem + T3_A_BYTES;    // [stage]: boxes of 16 KB, or (CTA pair) this CTA's 64 columns of them, 8 KB
    T3Barriers& bar = *reinterpret_cast<T3Barriers*>(sB + T3_STAGES * T3_B_STAGE_BYTES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        prefetch_tensormap(&tmap);
        prefetch_tensormap(&tmap_half);
        mbar_init(&bar.a_full, 1);
        for (int s = 0; s < T3_STAGES; s++) {
            mbar_init(&bar.full[s], 1);
            mbar_init(&bar.empty[s], T3_2CTA ? 1 : 2);
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(&bar.tmem_full[b], 1);
            // every epilogue warp arrives once per tile, whether it has rows or not (CTA pair: both CTAs' warps on rank 0's)
            mbar_init(&bar.tmem_empty[b], (T3_2CTA ? 2 : 1) * T3_EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp == T3_MMA_WARP) {
        if (T3_2CTA) {
            tmem_alloc2(&bar.tmem_base, T3_TMEM_COLS);
            tmem_relinquish2();
        } else {
            tmem_alloc(&bar.tmem_base, T3_TMEM_COLS);
            tmem_relinquish();
        }
    }
    tc_fence_before();
    cluster_sync_all();  // both CTAs' barriers exist before either signals the other's
    tc_fence_after();
    const uint32_t tmem_base = bar.tmem_base;
    if (threadIdx.x == 0) T3_TRACE(1);

    // The TMA and the MMA warp run their loops CONVERGED (all 32 lanes wait on the barriers) and elect one lane only for the
    // asynchronous instructions themselves: in warp-uniform control flow the compiler keeps the loop state, addresses and
    // descriptors in uniform registers, so a tcgen05.mma is one instruction.  (Under `if (lane == 0)` every UTCHMMA /
    // UTMALDG sat in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop of ~13 dependent instructions: the MMA warp took as long
    // to issue a tile as the tensor pipe took to execute it, and any competition for its issue slots slowed the pipe.)
#if T3_2CTA
    // rows of the peer (rank 1 sits 256 rows further down; rank 0 always has at least as many valid halves)
    const int peer_m0 = (int)(blockIdx.x ^ 1u) * T3_ROWS;
    const int peer_nhalf = peer_m0 >= n0 ? 0 : ((peer_m0 + 128 < n0) ? 2 : 1);
    const int own_nhalf = ghost ? 0 : nhalf;
    if (warp == T3_TMA_WARP) {
        // ===================================================== TMA producer (both CTAs; completions go to rank 0)
        if (ntiles > 0) {
            const uint32_t lead_a_full = mapa_u32(smem_u32(&bar.a_full), 0);
            if (elect_one()) {
                if (rank == 0) mbar_arrive_expect_tx(&bar.a_full, (own_nhalf + peer_nhalf) * T3_KCH * T3_BOX_BYTES);
                for (int h = 0; h < own_nhalf; h++)
                    for (int kc = 0; kc < T3_KCH; kc++)
                        tma_load_3d_2sm(sA + (h * T3_KCH + kc) * T3_BOX_BYTES, &tmap, lead_a_full, kc * T3_BK, m0 + h * 128, fa);
            }
            __syncwarp();
            int it = 0;
            for (int t = 0; t < ntiles; t++) {
#pragma unroll
                for (int kc = 0; kc < T3_KCH; kc++, it++) {
                    const int s = it % T3_STAGES;
                    mbar_wait(&bar.empty[s], ((it / T3_STAGES) & 1) ^ 1);
                    if (elect_one()) {
                        if (rank == 0) mbar_arrive_expect_tx(&bar.full[s], 2 * T3_B_STAGE_BYTES);  // own + the peer's columns
                        tma_load_3d_2sm(sB + s * T3_B_STAGE_BYTES, &tmap_half, mapa_u32(smem_u32(&bar.full[s]), 0), kc * T3_BK,
                                        t * T3_BN + (int)rank * 64, fb);
                    }
                    __syncwarp();
                }
                if (lane == 0 && t < 20) T3_TRACE(10 + t);
            }
        }
    } else if (warp == T3_MMA_WARP) {
        // ===================================================== MMA issuer: rank 0 only, for both CTAs.  This warp has the
        // highest warp id of the CTA: the sub-partition's arbiter prefers it over the epilogue warps it shares issue slots
        // with.
        if (ntiles > 0 && rank == 0) {
            mbar_wait(&bar.a_full, 0);
            tc_fence_after();
            if (lane == 0) T3_TRACE(3);
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);  // (tells the compiler it is warp-uniform)
            const uint64_t a_desc0 = make_desc_k_sw128(smem_u32(sA)), b_desc0 = make_desc_k_sw128(smem_u32(sB));
            int it = 0;
            for (int t = 0; t < ntiles; t++) {
                const int buf = t & 1;
                mbar_wait(&bar.tmem_empty[buf], ((t >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d0 = tmem_u + (buf * 2) * T3_BN;
                if (lane == 0 && t < 20) T3_TRACE(30 + t);
#pragma unroll
                for (int kc = 0; kc < T3_KCH; kc++, it++) {
                    const int s = it % T3_STAGES;
                    mbar_wait(&bar.full[s], (it / T3_STAGES) & 1);
                    tc_fence_after();
                    const uint64_t bs = b_desc0 + (uint64_t)((s * T3_B_STAGE_BYTES) >> 4);
                    if (elect_one()) {
#pragma unroll
                        for (int ks = 0; ks < T3_BK / 16; ks++) {
                            // descriptor start-address field is in 16-byte units: +2 per K = 16 step inside the swizzle row
                            const uint64_t bd = bs + (uint64_t)((ks * 32) >> 4);
                            const uint64_t ad = a_desc0 + (uint64_t)((kc * T3_BOX_BYTES + ks * 32) >> 4);
                            if (probe & 1) continue;
                            umma_f16_2cta(d0, ad, bd, T3_IDESC2, (kc | ks) != 0);
                            if (nhalf == 2)
                                umma_f16_2cta(d0 + T3_BN, ad + (uint64_t)((T3_KCH * T3_BOX_BYTES) >> 4), bd, T3_IDESC2, (kc | ks) != 0);
                        }
                        umma_commit2_multicast(&bar.empty[s], (uint16_t)3);  // both CTAs' ring stages have been read
                    }
                    __syncwarp();
                }
                if (elect_one()) umma_commit2_multicast(&bar.tmem_full[buf], (uint16_t)3);  // this tile's accumulators, both CTAs
                __syncwarp();
                if (lane == 0 && t < 20) T3_TRACE(50 + t);
            }
        }
#else
    if (warp == T3_TMA_WARP) {
        // ===================================================== TMA producer
        if (ntiles > 0) {
            if (!ghost && elect_one()) {
                mbar_arrive_expect_tx(&bar.a_full, nhalf * T3_KCH * T3_BOX_BYTES);
                for (int h = 0; h < nhalf; h++)
                    for (int kc = 0; kc < T3_KCH; kc++)
                        tma_load_3d(sA + (h * T3_KCH + kc) * T3_BOX_BYTES, &tmap, &bar.a_full, kc * T3_BK, m0 + h * 128, fa);
            }
            __syncwarp();
            int it = 0;
            for (int t = 0; t < ntiles; t++) {
#pragma unroll
                for (int kc = 0; kc < T3_KCH; kc++, it++) {
                    const int s = it % T3_STAGES;
                    mbar_wait(&bar.empty[s], ((it / T3_STAGES) & 1) ^ 1);
                    if (elect_one()) {
                        mbar_arrive_expect_tx(&bar.full[s], T3_BOX_BYTES);  // own half + the peer's half
                        tma_load_3d_multicast(sB + s * T3_BOX_BYTES + rank * (T3_BOX_BYTES / 2), &tmap_half, &bar.full[s],
                                              kc * T3_BK, t * T3_BN + (int)rank * 64, fb, (uint16_t)3);
                    }
                    __syncwarp();
                }
                if (lane == 0 && t < 20) T3_TRACE(10 + t);
            }
        }
    } else if (warp == T3_MMA_WARP) {
        // ===================================================== MMA issuer.  This warp has the highest warp id of the CTA:
        // the sub-partition's arbiter prefers it over the epilogue warps it shares issue slots with, so the tensor pipe is
        // never starved by the ALU-bound epilogue (measured: with the issuer as warp 1 the MMAs and the epilogue of
        // consecutive tiles ran back to back instead of overlapped).
        if (ntiles > 0 && ghost) {
            for (int it = 0; it < ntiles * T3_KCH; it++) {
                const int s = it % T3_STAGES;
                mbar_wait(&bar.full[s], (it / T3_STAGES) & 1);
                if (elect_one()) umma_commit_multicast(&bar.empty[s], (uint16_t)3);
                __syncwarp();
            }
        } else if (ntiles > 0) {
            mbar_wait(&bar.a_full, 0);
            tc_fence_after();
            if (lane == 0) T3_TRACE(3);
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);  // (tells the compiler it is warp-uniform)
            const uint64_t a_desc0 = make_desc_k_sw128(smem_u32(sA)), b_desc0 = make_desc_k_sw128(smem_u32(sB));
            int it = 0;
            for (int t = 0; t < ntiles; t++) {
                const int buf = t & 1;
                mbar_wait(&bar.tmem_empty[buf], ((t >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d0 = tmem_u + (buf * 2) * T3_BN;
                if (lane == 0 && t < 20) T3_TRACE(30 + t);
#pragma unroll
                for (int kc = 0; kc < T3_KCH; kc++, it++) {
                    const int s = it % T3_STAGES;
                    mbar_wait(&bar.full[s], (it / T3_STAGES) & 1);
                    tc_fence_after();
                    const uint64_t bs = b_desc0 + (uint64_t)((s * T3_BOX_BYTES) >> 4);
                    if (elect_one()) {
#pragma unroll
                        for (int ks = 0; ks < T3_BK / 16; ks++) {
                            // descriptor start-address field is in 16-byte units: +2 per K = 16 step inside the swizzle row
                            const uint64_t bd = bs + (uint64_t)((ks * 32) >> 4);
                            const uint64_t ad = a_desc0 + (uint64_t)((kc * T3_BOX_BYTES + ks * 32) >> 4);
                            if (probe & 1) continue;
                            umma_f16(d0, ad, bd, T3_IDESC, (kc | ks) != 0);
                            if (nhalf == 2)
                                umma_f16(d0 + T3_BN, ad + (uint64_t)((T3_KCH * T3_BOX_BYTES) >> 4), bd, T3_IDESC, (kc | ks) != 0);
                        }
                        umma_commit_multicast(&bar.empty[s], (uint16_t)3);  // ring stage read (here); the peer adds its own
                    }
                    __syncwarp();
                }
                if (elect_one()) umma_commit(&bar.tmem_full[buf]);  // both accumulators of this tile complete
                __syncwarp();
                if (lane == 0 && t < 20) T3_TRACE(50 + t);
            }
        }
#endif
    } else {
        // ===================================================== epilogue: warps 0..7, thread = row
        const int half = (warp >> 2) & 1;
        const int quarter = warp & 3;  // the TMEM lane quarter this warp may read
        const int csp = warp >> 3;     // which T3_CW columns of every tile
        {
            const int row = m0 + half * 128 + quarter * 32 + lane;
            // hands an accumulator buffer back to the MMA issuer (CTA pair: rank 0's barrier collects both CTAs' warps)
            const uint32_t rel_bar0 = T3_2CTA ? mapa_u32(smem_u32(&bar.tmem_empty[0]), 0) : smem_u32(&bar.tmem_empty[0]);
            auto release = [&](int buf) {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (T3_2CTA) mbar_arrive_cluster(rel_bar0 + buf * 8);
                    else mbar_arrive(&bar.tmem_empty[buf]);
                }
            };
            const fs::PairScale ps = fs::pair_scale(operands, n_max, fa, fb);
            const float sc = ps.sc;
            T3Epi es;
            es.k1 = es.k2 = es.k3 = es.ks = 0;
            es.st_base = 0;
            es.rowmask = 0xFFFFFFFFu;
            es.rowbits = 0;
            es.warp_partial = false;
            es.col_out = nullptr;
            if constexpr (SP) {
                // write: row `lane`, 16-byte chunk cc at lane * 64 + ((cc ^ ((lane >> 1) & 3)) << 4)
                // read : lane = parity * 16 + column pair cp; row 2 m + parity, word cp: chunk (cp >> 2) ^ (m & 3)
                es.st_base = smem_u32(sB + T3_STAGES * T3_B_STAGE_BYTES + 256) + warp * (T3_NBUF * T3_STAGE_BYTES);
                const int cp = lane & 15, par = lane >> 4;
#pragma unroll
                for (int x = 0; x < 4; x++) {
                    es.st_wr[x] = lane * 64 + ((x ^ ((lane >> 1) & 3)) << 4);
                    es.st_rd[x] = par * 64 + ((((cp >> 2) ^ x) << 4) | ((cp & 3) << 2));
                }
                es.rowmask = (row < n0) ? 0xFFFFFFFFu : 0u;
                es.rowbits = 15u - (uint32_t)(lane >> 1);
                es.warp_partial = (m0 + half * 128 + quarter * 32 + 31 >= n0);
                const int slab = ((m0 + half * 128 + quarter * 32) >> 5) * 2 + par;
                es.col_out = colraw + ((size_t)p * nslab + slab) * cpitch + 2 * cp;
            }
            // a warp without rows (beyond n0, or an empty half) only keeps the accumulator hand-shake going
            const bool warp_dead = ghost || half >= nhalf || m0 + half * 128 + quarter * 32 >= n0;
            for (int t = 0; t < ntiles; t++) {
                const int buf = t & 1;
                mbar_wait(&bar.tmem_full[buf], (t >> 1) & 1);
                tc_fence_after();
                if (threadIdx.x == 0 && t < 20) T3_TRACE(70 + t);
                if (warp_dead) {
                    release(buf);
                    continue;
                }
                const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (buf * 2 + half) * T3_BN + csp * T3_CW;
                // the ragged last tile (n1 not a multiple of 128) and the last, partially valid row group run a second copy
                // of the chunk code with the column (and, single pass, row) masks; the hot copy has none (if-converted
                // masks cost 2.5 ALU operations per score inline)
                const bool masked = (t == ntiles - 1 && (n1 & (T3_BN - 1))) || es.warp_partial;
                const int tcol0 = t * T3_BN + csp * T3_CW;
                if constexpr (T3_CSPLIT == 1) {
                    // the whole 128-column accumulator row goes to registers at once, and the TMEM buffer is handed back
                    // to the MMA warp BEFORE the top-3 network runs
                    uint32_t rr[T3_CW / 32][32];
#pragma unroll
                    for (int c = 0; c < T3_CW / 32; c++) tmem_ld_32x32b_x32(taddr + c * 32, rr[c]);
                    tmem_ld_wait();
                    release(buf);
                    if (threadIdx.x == 0 && t < 20) T3_TRACE(110 + t);
                    if (probe & 2) {
                        es.k1 = max(es.k1, rr[0][0] ^ rr[T3_CW / 32 - 1][31]);
                        continue;
                    }
#pragma unroll
                    for (int c = 0; c < T3_CW / 32; c++) {
                        if (masked) t3_chunk<MODE, true>(rr[c], tcol0 + c * 32, c, n1, sc, key_mul, es, [] {});
                        else t3_chunk<MODE, false>(rr[c], tcol0 + c * 32, c, n1, sc, key_mul, es, [] {});
                    }
                } else {
                    // chunk by chunk (the register budget of 18 warps holds one): the buffer goes back after the last
                    // chunk's load.  (Fetching the second chunk into the first one's registers as soon as its keys exist,
                    // so that the buffer returns while the first chunk's tournaments still run, measured SLOWER: 0.560
                    // against 0.533 ms; so did a second staging tile per warp, 0.563 ms.)
                    uint32_t rr[32];
#pragma unroll
                    for (int c = 0; c < T3_CW / 32; c++) {
                        tmem_ld_32x32b_x32(taddr + c * 32, rr);
                        tmem_ld_wait();
                        if (c == T3_CW / 32 - 1) {
                            release(buf);
                            if (threadIdx.x == 0 && t < 20) T3_TRACE(110 + t);
                        }
                        if (probe & 2) {
                            es.k1 = max(es.k1, rr[0] ^ rr[31]);
                            continue;
                        }
                        if (masked) t3_chunk<MODE, true>(rr, tcol0 + c * 32, c, n1, sc, key_mul, es, [] {});
                        else t3_chunk<MODE, false>(rr, tcol0 + c * 32, c, n1, sc, key_mul, es, [] {});
                    }
                }
                if (threadIdx.x == 0 && t < 20) T3_TRACE(90 + t);
            }
            if constexpr (T3_CSPLIT == 1) {
                if (row < n0) keys[((size_t)dir * n_pairs_total + p) * n_max + row] = make_uint4(es.k1, es.k2, es.k3, es.ks);
            } else {
                // the warps of a row group merge their column ranges: the upper ones park their keys in shared memory
                uint4* park = reinterpret_cast<uint4*>(sA) + (warp & 7) * 32 + lane;  // the row block is dead: every MMA is done
                if (csp > 0) *park = make_uint4(es.k1, es.k2, es.k3, es.ks);
                named_bar_sync(1 + (warp & 7), 64);
                if (csp == 0) {
                    const uint4 o = *park;
                    top3_insert(es.k1, es.k2, es.k3, o.x);
                    top3_insert(es.k1, es.k2, es.k3, o.y);
                    top3_insert(es.k1, es.k2, es.k3, o.z);
                    es.ks = max(es.ks, o.w);
                    if (row < n0) keys[((size_t)dir * n_pairs_total + p) * n_max + row] = make_uint4(es.k1, es.k2, es.k3, es.ks);
                }
            }
        }
    }
    if (threadIdx.x == 0) T3_TRACE(4);
    tc_fence_before();
    cluster_sync_all();  // the peer may still be signalling this CTA's barriers
    if (warp == T3_MMA_WARP) {
        tc_fence_after();
        if (T3_2CTA) tmem_dealloc2(tmem_base, T3_TMEM_COLS);
        else tmem_dealloc(tmem_base, T3_TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------ resolve
// Lowe ratio decision exactly as k_float_finalize evaluates it (float_fp32.cu): monotone in both arguments
__device__ __forceinline__ bool ratio_pass(float b, float s, float ratio) {
    const float db = __fsqrt_rn(fmaxf(0.f, __fsub_rn(2.f, __fmul_rn(2.f, b))));
    const float ds = __fsqrt_rn(fmaxf(0.f, __fsub_rn(2.f, __fmul_rn(2.f, s))));
    return db < __fmul_rn(ratio, ds);
}

struct T3Out {
    float* best;     // [2][P][n_max]  direction 0 = rows of a, direction 1 = rows of b
    float* second;
    int32_t* idx;
    uint32_t* list;  // rows that need the exact full scan: (dir << 31) | (p * n_max + row)
    int32_t* counters;  // [0] full scans, [1] rows with rescored candidates
};

// thread = (direction, pair, row): decode the top-3 keys, decide what they prove, rescore what they do not
__global__ void __launch_bounds__(256)
k_float_resolve(const void* __restrict__ operands, int n_max, const int32_t* __restrict__ counts,
                const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b, int n_pairs_total, int pair0,
                const uint4* __restrict__ keys, float scale, float threshold, float ratio, T3Out out,
                uint32_t* __restrict__ rowkey) {
    const int p = pair0 + blockIdx.y, dir = blockIdx.z;
    const int row = blockIdx.x * 256 + threadIdx.x, lane = threadIdx.x & 31;
    const int fa = dir ? pair_b[p] : pair_a[p], fb = dir ? pair_a[p] : pair_b[p];
    const int n0 = min(counts[fa], n_max), n1 = min(counts[fb], n_max);
    if ((row & ~31) >= n0 || n1 <= 0) return;  // (warp-uniform)
    const bool active = row < n0;
    const fs::PairScale ps = fs::pair_scale(operands, n_max, fa, fb);
    const size_t o = ((size_t)dir * n_pairs_total + p) * n_max + row;
    float v1 = -INFINITY, v2 = -INFINITY, v3 = -INFINITY, v2s = -INFINITY;
    int i1 = 0, i2 = 0;
    bool amb_idx = false, want1 = false, want2 = false, full = false;
    uint32_t kx = 0, ky = 0;
    if (active) {
        const uint4 k = keys[o];
        kx = k.x;
        ky = k.y;
        v1 = fs::key_score(k.x, ps);
        v2 = fs::key_score(k.y, ps);
        // k.x >= k.y >= k.z are chunk maxima, k.w the largest chunk runner-up.  Every column other than i1 and i2
        // scores <= v3 := max(third chunk maximum, best chunk runner-up); the approximate runner-up VALUE of the row is
        // exactly max(v2, vs) (if vs comes from another chunk than the best's, that chunk's maximum is <= k.y).
        const float vs = fs::key_score(k.w, ps);
        v3 = fmaxf(fs::key_score(k.z, ps), vs);
        i1 = fs::key_col(k.x);
        i2 = fs::key_col(k.y);
        const float d2 = 2.f * ps.delta;
        amb_idx = !(v1 - fmaxf(v2, v3) > d2);
        const bool second_sure = (v2 - v3 > d2);  // the runner-up is certainly column i2
        if (amb_idx) {
            if (v1 - v3 > d2) want1 = want2 = true;  // the arg-max is i1 or i2
            else full = true;
        }
        v2s = fmaxf(v2, vs);
        if (dir == 0) {
            if (isfinite(threshold) && fabsf(v1 * scale - threshold) <= ps.delta * fabsf(scale) * 1.001f + 1e-12f)
                want1 = true;
            if (ratio > 0.f && (k.y | k.w) != 0u) {
                const float sa = fmaxf(v2, vs);
                const bool undecided = amb_idx || (ratio_pass(v1 + ps.delta, sa - ps.delta, ratio) !=
                                                   ratio_pass(v1 - ps.delta, sa + ps.delta, ratio));
                if (undecided) {
                    if (second_sure) want1 = want2 = true;
                    else full = true;
                }
            }
        }
    }
    // warp-cooperative exact rescoring of the flagged rows' candidates (float_ops.cuh: warp_dot)
    const __half* bank_a = static_cast<const __half*>(operands) + (size_t)fa * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
    const __half* bank_b = static_cast<const __half*>(operands) + (size_t)fb * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
    float ex1 = 0.f, ex2 = 0.f;
    unsigned jobs = __ballot_sync(0xffffffffu, active && !full && (want1 || want2));
    if (jobs && lane == 0) atomicAdd(out.counters + 1, __popc(jobs));
    while (jobs) {
        const int l = __ffs(jobs) - 1;
        jobs &= jobs - 1;
        const int r_ = __shfl_sync(0xffffffffu, row, l);
        const int c1 = __shfl_sync(0xffffffffu, i1, l), c2 = __shfl_sync(0xffffffffu, i2, l);
        const bool w2 = __shfl_sync(0xffffffffu, (int)want2, l) != 0;
        const fs::OpRow8 a = fs::load_op_row(bank_a + (size_t)r_ * fs::OP_ROW_HALFS, lane);
        const float e1 = fs::warp_dot(a, fs::load_op_row(bank_b + (size_t)c1 * fs::OP_ROW_HALFS, lane));
        float e2 = 0.f;
        if (w2) e2 = fs::warp_dot(a, fs::load_op_row(bank_b + (size_t)c2 * fs::OP_ROW_HALFS, lane));
        if (lane == l) {
            ex1 = e1;
            ex2 = e2;
        }
    }
    if (!active) return;
    float b = v1, s = v2s;
    int ib = i1;
    if (!full) {
        if (want1) b = ex1;
        if (want2) {
            s = ex2;
            if (amb_idx && fs::better(ex2, i2, ex1, i1)) {
                b = ex2;
                ib = i2;
                s = ex1;
            }
        }
    } else {
        const int pos = atomicAdd(out.counters, 1);
        out.list[pos] = ((uint32_t)dir << 31) | (uint32_t)((size_t)p * n_max + row);
    }
    out.best[o] = b;
    out.second[o] = s;
    out.idx[o] = ib;
    // single-pass form: the tensor-core key of the chosen column (0 = unknown: the row goes through the exact scan)
    if (rowkey) rowkey[o] = full ? 0u : (ib == i1 ? kx : ky);
}

// ------------------------------------------------------------------------------------------------ single pass: columns
// colraw layout (single pass): [pair][nslab][cpitch] u32
__host__ __device__ inline int sp_cpitch(int n_max) { return (n_max + T3_BN - 1) / T3_BN * T3_BN; }
__host__ __device__ inline int sp_nslab(int n_max) { return (n_max + 31) / 32 * 2; }

// 16-bit column value (top 16 of the 20 live score bits) -> lower edge of the score interval [v, v + step)
__device__ __forceinline__ float dec16(uint32_t v16, const fs::PairScale& ps) {
    if (v16 == 0u) return -INFINITY;
    return (__uint_as_float(0x3F800000u | (v16 << 4)) - fs::KEY_BIAS) * ps.inv_sc;
}

// thread = (pair, column): fold the slab results of the column into the three largest slab maxima
// (value << 16 | 0xFFFF - slab: larger value, then lower slab) and the largest slab runner-up value
constexpr int CM_UNROLL = 8;
__global__ void __launch_bounds__(256)
k_float_colmerge(int n_max, const int32_t* __restrict__ counts, const int32_t* __restrict__ pair_a,
                 const int32_t* __restrict__ pair_b, int pair0, const uint32_t* __restrict__ colraw,
                 uint4* __restrict__ colkeys) {
    const int p = pair0 + blockIdx.y, j = blockIdx.x * 256 + threadIdx.x;
    const int n0 = min(counts[pair_a[p]], n_max), n1 = min(counts[pair_b[p]], n_max);
    if (j >= n1) return;
    const int nsl = ((n0 + 31) >> 5) * 2;
    const int cpitch = sp_cpitch(n_max);
    const uint32_t* src = colraw + (size_t)p * sp_nslab(n_max) * cpitch + j;
    uint32_t k1 = 0, k2 = 0, k3 = 0, ks = 0;
    for (int s0 = 0; s0 < nsl; s0 += CM_UNROLL) {
        uint32_t x[CM_UNROLL];
#pragma unroll
        for (int u = 0; u < CM_UNROLL; u++) x[u] = (s0 + u < nsl) ? __ldg(src + (size_t)(s0 + u) * cpitch) : 0u;
#pragma unroll
        for (int u = 0; u < CM_UNROLL; u++) {
            const uint32_t h = x[u] & 0xFFFFu;
            top3_insert(k1, k2, k3, h ? ((h << 16) | (uint32_t)(0xFFFF - (s0 + u))) : 0u);
            ks = max(ks, x[u] >> 16);
        }
    }
    colkeys[(size_t)p * n_max + j] = make_uint4(k1, k2, k3, ks);
}

// thread = (pair, row of frame a): is the row the arg-max of its best column?  Writes idx1[pair][column] = row when it is
// (idx1 was filled with -1; at most one row of a column can be its arg-max, so there is one writer per element).
//   certain yes : the row's own 16-bit value IS the column's largest slab maximum, in the row's slab, and that maximum
//                 beats every other slab maximum and every slab runner-up by more than 2 delta + one 16-bit step;
//   certain no  : the column's largest value exceeds what the row can possibly score;
//   otherwise   : the rows of the two best slabs (and the row itself) are rescored exactly; every other row scores below
//                 the third slab maximum + step + delta; if the exact winner clears that, it decides; else every other
//                 slab whose maximum could still reach the winner is rescored as well (counters[3]).
__global__ void __launch_bounds__(256)
k_float_colcheck(const void* __restrict__ operands, int n_max, const int32_t* __restrict__ counts,
                 const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b, int n_pairs_total, int pair0,
                 const uint4* __restrict__ colkeys, const uint32_t* __restrict__ rowkey,
                 const uint32_t* __restrict__ colraw, T3Out out, int32_t* __restrict__ idx1) {
    const int p = pair0 + blockIdx.y;
    const int row = blockIdx.x * 256 + threadIdx.x, lane = threadIdx.x & 31;
    const int fa = pair_a[p], fb = pair_b[p];
    const int n0 = min(counts[fa], n_max), n1 = min(counts[fb], n_max);
    if ((row & ~31) >= n0 || n1 <= 0) return;  // (warp-uniform)
    const bool active = row < n0;
    const fs::PairScale ps = fs::pair_scale(operands, n_max, fa, fb);
    const float step = ps.inv_sc * (1.f / 524288.f);  // 16 key units of 2^-23
    const float dl = ps.delta * 1.0005f;
    const size_t o = (size_t)p * n_max + row;
    int j = 0, sl1 = 0, sl2 = -1, state = 0;  // 0 no, 1 yes, 2 ambiguous
    float c3 = -INFINITY;
    if (active) {
        j = out.idx[o];
        const float b = out.best[o];
        const uint32_t rk = rowkey[o];
        const uint4 K = colkeys[(size_t)p * n_max + j];
        const float c1 = dec16(K.x >> 16, ps), c2 = dec16(K.y >> 16, ps), cS = dec16(K.w, ps);
        c3 = dec16(K.z >> 16, ps);
        sl1 = 0xFFFF - (int)(K.x & 0xFFFFu);
        if (K.y) sl2 = 0xFFFF - (int)(K.y & 0xFFFFu);
        const bool known = rk != 0u;
        const float a_i = fs::key_score(rk, ps);
        const float ex_hi = (known && b == a_i) ? a_i + dl : b;  // b is exact unless it still is the key's value
        const int slab_i = (row >> 5) * 2 + (row & 1);
        if (known && (rk >> 16) == (K.x >> 16) && slab_i == sl1 && c1 - (fmaxf(c2, cS) + step) > 2.f * dl) state = 1;
        else if (ex_hi < c1 - dl) state = 0;
        else state = 2;
        if (state == 1) idx1[(size_t)p * n_max + j] = row;
    }
    unsigned jobs = __ballot_sync(0xffffffffu, active && state == 2);
    if (!jobs) return;
    if (lane == 0) atomicAdd(out.counters + 2, __popc(jobs));
    const __half* bank_a = static_cast<const __half*>(operands) + (size_t)fa * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
    const __half* bank_b = static_cast<const __half*>(operands) + (size_t)fb * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
    while (jobs) {
        const int l = __ffs(jobs) - 1;
        jobs &= jobs - 1;
        const int r_ = __shfl_sync(0xffffffffu, row, l), j_ = __shfl_sync(0xffffffffu, j, l);
        const int a1 = __shfl_sync(0xffffffffu, sl1, l), a2 = __shfl_sync(0xffffffffu, sl2, l);
        const float bound = __shfl_sync(0xffffffffu, c3, l);
        const fs::OpRow8 bj = fs::load_op_row(bank_b + (size_t)j_ * fs::OP_ROW_HALFS, lane);
        float eb = -INFINITY;
        int ib = 0x7fffffff;
        // exact scores of the 16 rows of slab sl (the rows of one parity of a 32-row group), eight rows in flight
        auto rescore_slab = [&](int sl) {
#pragma unroll 1
            for (int c0 = 0; c0 < 16; c0 += 8) {
                int cand[8];
                fs::OpRow8 ar[8];
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int r = (sl >> 1) * 32 + 2 * (c0 + u) + (sl & 1);
                    cand[u] = r < n0 ? r : -1;
                    ar[u] = fs::load_op_row(bank_a + (size_t)min(r, n0 - 1) * fs::OP_ROW_HALFS, lane);
                }
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const float e = fs::warp_dot(ar[u], bj);
                    if (cand[u] >= 0 && fs::better(e, cand[u], eb, ib)) {
                        eb = e;
                        ib = cand[u];
                    }
                }
            }
        };
        {
            const float e = fs::warp_dot(fs::load_op_row(bank_a + (size_t)r_ * fs::OP_ROW_HALFS, lane), bj);
            eb = e;
            ib = r_;
        }
        rescore_slab(a1);
        if (a2 >= 0) rescore_slab(a2);
        // every row outside the rescored slabs scores below (third slab maximum + step + delta); if the exact winner does
        // not clear that, the other slabs whose maximum could still reach it are rescored too (their bound comes from
        // the slab results of this column: 4 B per slab instead of an exact scan of the whole column)
        if (!(eb > bound + step + dl)) {  // (warp-uniform: eb is the same in every lane)
            if (lane == 0) atomicAdd(out.counters + 3, 1);
            const int nsl = ((n0 + 31) >> 5) * 2;
            const int cpitch = sp_cpitch(n_max);
            const uint32_t* src = colraw + (size_t)p * sp_nslab(n_max) * cpitch + j_;
            for (int s0 = 0; s0 < nsl; s0 += 32) {
                const int sl = s0 + lane;
                const uint32_t h = (sl < nsl) ? (__ldg(src + (size_t)sl * cpitch) & 0xFFFFu) : 0u;
                const bool c = sl < nsl && sl != a1 && sl != a2 && h != 0u && (dec16(h, ps) + step + dl > eb);
                unsigned m = __ballot_sync(0xffffffffu, c);
                while (m) {
                    rescore_slab(s0 + __ffs(m) - 1);
                    m &= m - 1;
                }
            }
        }
        if (lane == 0 && ib == r_) idx1[(size_t)p * n_max + j_] = r_;
    }
}

// exact top-2 over ALL columns for the rows the keys could not decide; one CTA (16 warps) per listed row.  The scan is
// bound by DRAM round trips (the rows are scattered over the whole bank), so every warp keeps 16 x 16-byte loads per
// lane in flight: eight columns per trip.
constexpr int FSCAN_WARPS = 16, FSCAN_UNROLL = 8;
__global__ void __launch_bounds__(FSCAN_WARPS * 32)
k_float_fullscan(const void* __restrict__ operands, int n_max, const int32_t* __restrict__ counts,
                 const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b, int n_pairs_total, T3Out out,
                 const uint32_t* __restrict__ list, const int32_t* __restrict__ list_count) {
    __shared__ float s_best[FSCAN_WARPS], s_second[FSCAN_WARPS];
    __shared__ int s_idx[FSCAN_WARPS];
    const int count = *list_count;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int e = blockIdx.x; e < count; e += gridDim.x) {
        const uint32_t ent = list[e];
        const int dir = ent >> 31;
        const uint32_t lin = ent & 0x7FFFFFFFu;
        const int p = lin / n_max, row = lin % n_max;
        const int fa = dir ? pair_b[p] : pair_a[p], fb = dir ? pair_a[p] : pair_b[p];
        const int n1 = min(counts[fb], n_max);
        const __half* bank_a = static_cast<const __half*>(operands) + (size_t)fa * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
        const __half* bank_b = static_cast<const __half*>(operands) + (size_t)fb * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
        const fs::OpRow8 a = fs::load_op_row(bank_a + (size_t)row * fs::OP_ROW_HALFS, lane);
        fs::Top2 t;
        t.init();
        // columns in ascending order within the warp: c0 + FSCAN_WARPS * u
        for (int c0 = warp; c0 < n1; c0 += FSCAN_WARPS * FSCAN_UNROLL) {
            fs::OpRow8 b[FSCAN_UNROLL];
#pragma unroll
            for (int u = 0; u < FSCAN_UNROLL; u++)
                b[u] = fs::load_op_row(bank_b + (size_t)min(c0 + FSCAN_WARPS * u, n1 - 1) * fs::OP_ROW_HALFS, lane);
#pragma unroll
            for (int u = 0; u < FSCAN_UNROLL; u++) {
                const float v = fs::warp_dot(a, b[u]);
                if (c0 + FSCAN_WARPS * u < n1) t.push(v, c0 + FSCAN_WARPS * u);
            }
        }
        __syncthreads();  // the previous entry's merge is done with the staging arrays
        if (lane == 0) {
            s_best[warp] = t.best;
            s_second[warp] = t.second;
            s_idx[warp] = t.idx;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            fs::Top2 m;
            m.init();
            for (int w = 0; w < FSCAN_WARPS; w++) m.merge(s_best[w], s_second[w], s_idx[w]);
            const size_t o = ((size_t)dir * n_pairs_total + p) * n_max + row;
            out.best[o] = m.best;
            out.second[o] = m.second;
            out.idx[o] = m.idx;
        }
    }
}

// Single-pass form: the exact top-2 of a listed ROW without reading the whole other frame.  The slab results bound every
// tensor-core score of the row from above: S~(row, j) < dec16(best16 of (slab(row), j)) + step.  Two distinct columns are
// known to score >= t2 := max(second chunk maximum, best chunk runner-up), so a column whose bound is <= t2 - 2 delta
// cannot be among the exact top two; only the others (a few dozen of thousands, even on random descriptors) are rescored
// exactly.  One warp per row: 4 B per column read instead of 1 KB.
__global__ void __launch_bounds__(256)
k_float_slabscan(const void* __restrict__ operands, int n_max, const int32_t* __restrict__ counts,
                 const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b, const uint4* __restrict__ keys,
                 const uint32_t* __restrict__ colraw, T3Out out, const uint32_t* __restrict__ list,
                 const int32_t* __restrict__ list_count) {
    const int count = *list_count;
    const int lane = threadIdx.x & 31;
    const int nwarps = gridDim.x * 8;
    for (int e = blockIdx.x * 8 + (threadIdx.x >> 5); e < count; e += nwarps) {
        const uint32_t lin = list[e] & 0x7FFFFFFFu;  // direction 0 only
        const int p = lin / n_max, row = lin % n_max;
        const int fa = pair_a[p], fb = pair_b[p];
        const int n1 = min(counts[fb], n_max);
        const fs::PairScale ps = fs::pair_scale(operands, n_max, fa, fb);
        const float step = ps.inv_sc * (1.f / 524288.f), dl = ps.delta * 1.0005f;
        const uint4 K = keys[(size_t)p * n_max + row];
        const float thr = fmaxf(fs::key_score(K.y, ps), fs::key_score(K.w, ps)) - 2.f * dl - step;  // -inf: one column
        const __half* bank_a = static_cast<const __half*>(operands) + (size_t)fa * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
        const __half* bank_b = static_cast<const __half*>(operands) + (size_t)fb * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
        const fs::OpRow8 a = fs::load_op_row(bank_a + (size_t)row * fs::OP_ROW_HALFS, lane);
        const uint32_t* src = colraw + ((size_t)p * sp_nslab(n_max) + ((row >> 5) * 2 + (row & 1))) * sp_cpitch(n_max);
        fs::Top2 t;
        t.init();
        for (int j0 = 0; j0 < n1; j0 += 128) {
            uint32_t x[4];
#pragma unroll
            for (int u = 0; u < 4; u++) x[u] = (j0 + u * 32 + lane < n1) ? __ldg(src + j0 + u * 32 + lane) : 0u;
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const bool cand = (j0 + u * 32 + lane < n1) && !(dec16(x[u] & 0xFFFFu, ps) <= thr);
                unsigned m = __ballot_sync(0xffffffffu, cand);
                while (m) {  // ascending columns
                    const int j = j0 + u * 32 + __ffs(m) - 1;
                    m &= m - 1;
                    const float v = fs::warp_dot(a, fs::load_op_row(bank_b + (size_t)j * fs::OP_ROW_HALFS, lane));
                    t.push(v, j);
                }
            }
        }
        if (lane == 0) {
            const size_t o = (size_t)p * n_max + row;
            out.best[o] = t.best;
            out.second[o] = t.second;
            out.idx[o] = t.idx;
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------ C ABI
extern "C" size_t fs_float_operand_bytes(int n_frames, int n_max) {
    if (n_frames <= 0 || n_max <= 0) return 0;
    return (size_t)n_frames * fs::op_rows(n_max) * fs::OP_ROW_BYTES;
}

// FS_TUNE_FLOAT_SINGLE_PASS=0 keeps fs_float_match on the two-direction kernel (A/B measurements, and the GPU test that
// compares the two forms bit for bit); read on every call (host side only)
static bool single_pass_enabled() {
    const char* e = getenv("FS_TUNE_FLOAT_SINGLE_PASS");
    return !(e && e[0] == '0');
}

extern "C" size_t fs_float_match_tc_workspace_bytes(int n_pairs, int n_max) {
    if (n_pairs <= 0 || n_max <= 0) return 256;
    const size_t n = (size_t)n_pairs * n_max, n2 = 2 * n;
    // keys (16 B) + best, second, idx, list (4 B each) for both directions + counters
    size_t bytes = fs::align_up(n2 * 16, 256) + 4 * fs::align_up(n2 * 4, 256) + 256 + 256;
    // single-pass form: row keys of the chosen columns + slab results [pair][n_max / 16 slabs][n_max] (4 B each)
    bytes += fs::align_up(n * 4, 256) + fs::align_up(fs::float_top3_colraw_elems(n_pairs, n_max) * 4, 256);
    return bytes;
}

// the tensor-core pass alone: top-3 keys of every row of both directions (also the first stage of the tensor-core
// Hamming matcher, hamming_tc.cu: +-1 operands make the same kernel an exact integer scorer)
size_t fs::float_top3_colraw_elems(int n_pairs, int n_max) {
    return (size_t)n_pairs * sp_nslab(n_max) * sp_cpitch(n_max);
}

int fs::launch_float_top3(const CUtensorMap& tmap, const CUtensorMap& tmap_half, const void* operands, int n_max, const int32_t* counts,
                          const int32_t* pair_a, const int32_t* pair_b, int n_pairs, uint4* keys, cudaStream_t stream,
                          int probe, long long* trace, uint32_t* colraw, bool hamming, int ndirs) {
    static int granted[64], granted_sp[64], granted_hm[64];
    const int cpitch = sp_cpitch(n_max), nslab = sp_nslab(n_max);
    const int st_ = !colraw ? fs::ensure_dyn_smem(k_float_top3_tc<0>, granted, (size_t)(T3_SMEM_BYTES))
                    : hamming ? fs::ensure_dyn_smem(k_float_top3_tc<2>, granted_hm, (size_t)(T3_SP_SMEM_BYTES))
                              : fs::ensure_dyn_smem(k_float_top3_tc<1>, granted_sp, (size_t)(T3_SP_SMEM_BYTES));
    if (st_ != FS_OK) return st_;
    for (int p0 = 0; p0 < n_pairs; p0 += 32768) {
        const int np = min(32768, n_pairs - p0);
        dim3 grid(((n_max + T3_ROWS - 1) / T3_ROWS + 1) / 2 * 2, np, colraw ? 1 : ndirs);  // clusters of two along x
        if (colraw && hamming)
            k_float_top3_tc<2><<<grid, T3_THREADS, T3_SP_SMEM_BYTES, stream>>>(tmap, tmap_half, operands, n_max, counts, pair_a, pair_b,
                                                                               n_pairs, p0, 4096u, probe, trace, keys, colraw, cpitch, nslab);
        else if (colraw)
            k_float_top3_tc<1><<<grid, T3_THREADS, T3_SP_SMEM_BYTES, stream>>>(tmap, tmap_half, operands, n_max, counts, pair_a, pair_b,
                                                                               n_pairs, p0, 4096u, probe, trace, keys, colraw, cpitch, nslab);
        else
            k_float_top3_tc<0><<<grid, T3_THREADS, T3_SMEM_BYTES, stream>>>(tmap, tmap_half, operands, n_max, counts, pair_a, pair_b,
                                                                            n_pairs, p0, 4096u, probe, trace, keys, nullptr, 0, 0);
        FS_LAUNCH_CHECK("k_float_top3_tc");
    }
    return FS_OK;
}

// Measurement hook (bench.py roofline): ONLY the tensor-core kernel of fs_float_match, same grid, keys into `workspace`
extern "C" int fs_float_top3_keys(const void* operands, int n_frames, int n_max, const int32_t* counts,
                                  const int32_t* pair_a, const int32_t* pair_b, int n_pairs, int probe,
                                  long long* trace, void* workspace, size_t workspace_bytes, fs_stream_t stream_) {
    FS_REQUIRE(n_pairs > 0 && n_frames > 0 && n_max > 0 && n_max <= fs::KEY_MAX_COLS, "bad size");
    FS_REQUIRE(operands && counts && pair_a && pair_b, "null pointer");
    if (!workspace || workspace_bytes < fs_float_match_tc_workspace_bytes(n_pairs, n_max)) {
        fs::set_error("fs_float_top3_keys: workspace too small");
        return FS_ERR_WORKSPACE;
    }
    CUtensorMap tmap, tmap_half;
    int st = fs::encode_operand_map(operands, n_frames, n_max, &tmap);
    if (st == FS_OK) st = fs::encode_operand_map(operands, n_frames, n_max, &tmap_half, 64);
    if (st != FS_OK) return st;
    fs::Carver ws(workspace);
    const size_t n = (size_t)n_pairs * n_max, n2 = 2 * n;
    uint4* keys = ws.take<uint4>(n2);
    for (int i = 0; i < 4; i++) ws.take<float>(n2);
    ws.take<int32_t>(64);
    ws.take<uint32_t>(n);
    uint32_t* colraw = ws.take<uint32_t>(fs::float_top3_colraw_elems(n_pairs, n_max));
    // probe bit 2: the two-direction kernel even where fs_float_match would run the single-pass form
    const bool sp = single_pass_enabled() && n_max % 128 == 0 && !(probe & 4);
    return fs::launch_float_top3(tmap, tmap_half, operands, n_max, counts, pair_a, pair_b, n_pairs, keys,
                                 (cudaStream_t)stream_, probe & 3, trace, sp ? colraw : nullptr);
}

static int float_match_impl(const void* operands, int n_frames, int n_max, const int32_t* counts,
                            const int32_t* pair_a, const int32_t* pair_b, int n_pairs, float scale, float threshold,
                            float ratio, int cross_check, int32_t* matches0, float* score0, float* second0,
                            int32_t* matches1, int32_t* n_matches, int32_t* diag, void* workspace,
                            size_t workspace_bytes, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_frames >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(operands && counts && pair_a && pair_b && matches0, "null pointer");
    FS_REQUIRE(((uintptr_t)operands & 15) == 0, "operands must be 16-byte aligned");
    FS_REQUIRE(scale > 0.f, "scale must be positive");
    if (n_max > fs::KEY_MAX_COLS) {
        fs::set_error("fs_float_match: n_max %d > %d (use fs_float_match_split3)", n_max, fs::KEY_MAX_COLS);
        return FS_ERR_UNSUPPORTED;
    }
    FS_REQUIRE((size_t)n_pairs * n_max < (1ull << 31), "n_pairs * n_max too large for one call");
    const size_t need = fs_float_match_tc_workspace_bytes(n_pairs, n_max);
    if (!workspace || workspace_bytes < need) {
        fs::set_error("fs_float_match: workspace %zu < %zu", workspace_bytes, need);
        return FS_ERR_WORKSPACE;
    }
    CUtensorMap tmap, tmap_half;
    int st = fs::encode_operand_map(operands, n_frames, n_max, &tmap);
    if (st == FS_OK) st = fs::encode_operand_map(operands, n_frames, n_max, &tmap_half, 64);
    if (st != FS_OK) return st;
    fs::Carver ws(workspace);
    const size_t n = (size_t)n_pairs * n_max, n2 = 2 * n;
    uint4* keys = ws.take<uint4>(n2);
    T3Out out;
    out.best = ws.take<float>(n2);
    out.second = ws.take<float>(n2);
    out.idx = ws.take<int32_t>(n2);
    out.list = ws.take<uint32_t>(n2);
    out.counters = ws.take<int32_t>(64);
    uint32_t* rowkey = ws.take<uint32_t>(n);
    uint32_t* colraw = ws.take<uint32_t>(fs::float_top3_colraw_elems(n_pairs, n_max));
    // single pass: direction 0 on the tensor cores, the column side from the slab keys (see k_float_top3_tc<true>)
    const bool sp = cross_check && single_pass_enabled() && n_max % 128 == 0;
    FS_CUDA(cudaMemsetAsync(out.counters, 0, 64 * sizeof(int32_t), stream));
    st = fs::launch_float_top3(tmap, tmap_half, operands, n_max, counts, pair_a, pair_b, n_pairs, keys, stream, 0,
                               nullptr, sp ? colraw : nullptr, false, cross_check ? 2 : 1);
    if (st != FS_OK) return st;
    uint4* colkeys = keys + n;       // direction 1's key slots
    int32_t* idx1 = out.idx + n;     // arg-max row of every column (-1: no row has this column as its best)
    if (sp) FS_CUDA(cudaMemsetAsync(idx1, 0xFF, n * sizeof(int32_t), stream));
    for (int p0 = 0; p0 < n_pairs; p0 += 32768) {
        const int np = min(32768, n_pairs - p0);
        if (sp) {
            k_float_colmerge<<<dim3((n_max + 255) / 256, np), 256, 0, stream>>>(n_max, counts, pair_a, pair_b, p0, colraw, colkeys);
            FS_LAUNCH_CHECK("k_float_colmerge");
        }
        dim3 rgrid((n_max + 255) / 256, np, (sp || !cross_check) ? 1 : 2);  // knn use: direction 0 only
        k_float_resolve<<<rgrid, 256, 0, stream>>>(operands, n_max, counts, pair_a, pair_b, n_pairs, p0, keys, scale,
                                                   threshold, ratio, out, sp ? rowkey : nullptr);
        FS_LAUNCH_CHECK("k_float_resolve");
    }
    if (sp) {
        k_float_slabscan<<<fs::sm_count() * 4, 256, 0, stream>>>(operands, n_max, counts, pair_a, pair_b, keys, colraw, out,
                                                                 out.list, out.counters);
        FS_LAUNCH_CHECK("k_float_slabscan");
    } else {
        k_float_fullscan<<<fs::sm_count() * 4, FSCAN_WARPS * 32, 0, stream>>>(operands, n_max, counts, pair_a, pair_b, n_pairs,
                                                                             out, out.list, out.counters);
        FS_LAUNCH_CHECK("k_float_fullscan");
    }
    if (sp) {
        for (int p0 = 0; p0 < n_pairs; p0 += 32768) {
            const int np = min(32768, n_pairs - p0);
            k_float_colcheck<<<dim3((n_max + 255) / 256, np), 256, 0, stream>>>(operands, n_max, counts, pair_a, pair_b, n_pairs, p0,
                                                                                colkeys, rowkey, colraw, out, idx1);
            FS_LAUNCH_CHECK("k_float_colcheck");
        }
    }
    if (diag) FS_CUDA(cudaMemcpyAsync(diag, out.counters, 4 * sizeof(int32_t), cudaMemcpyDeviceToDevice, stream));
    fs::FloatRowTop2 d0{out.best, out.second, out.idx};
    return fs::launch_float_finalize(n_max, counts, pair_a, pair_b, n_pairs, d0, out.idx + n, scale, threshold, ratio,
                                     matches0, score0, second0, matches1, n_matches, stream, cross_check);
}

extern "C" int fs_float_match(const void* operands, int n_frames, int n_max, const int32_t* counts,
                              const int32_t* pair_a, const int32_t* pair_b, int n_pairs, float scale, float threshold,
                              float ratio, int32_t* matches0, float* score0, float* second0, int32_t* matches1,
                              int32_t* n_matches, int32_t* diag, void* workspace, size_t workspace_bytes,
                              fs_stream_t stream_) {
    return float_match_impl(operands, n_frames, n_max, counts, pair_a, pair_b, n_pairs, scale, threshold, ratio, 1,
                            matches0, score0, second0, matches1, n_matches, diag, workspace, workspace_bytes, stream_);
}

// BFMatcher(NORM_L2).knnMatch(k = 2) + Lowe ratio WITHOUT the mutual check (ORB_SLAM3/src/Frame.cc:1144-1151 semantics for
// float descriptors): matches0[q] = best column of row q if it passes threshold / ratio; matches1 is not produced.
extern "C" int fs_float_match_knn(const void* operands, int n_frames, int n_max, const int32_t* counts,
                                  const int32_t* pair_a, const int32_t* pair_b, int n_pairs, float scale,
                                  float threshold, float ratio, int32_t* matches0, float* score0, float* second0,
                                  int32_t* n_matches, int32_t* diag, void* workspace, size_t workspace_bytes,
                                  fs_stream_t stream_) {
    return float_match_impl(operands, n_frames, n_max, counts, pair_a, pair_b, n_pairs, scale, threshold, ratio, 0,
                            matches0, score0, second0, nullptr, n_matches, diag, workspace, workspace_bytes, stream_);
}
