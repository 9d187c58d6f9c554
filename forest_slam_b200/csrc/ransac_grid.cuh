// ransac_grid.cuh — batch-wide RANSAC engine with OpenCV's exact serial-loop semantics
// (calib3d/ptsetreg.cpp RANSACPointSetRegistrator::run), shared by
//   A4  cv2.findEssentialMat(..., method=cv2.RANSAC)   ros_ws/src/mono_slam.py:111      (ransac_essential.cu)
//   A7  cv2.solvePnPRansac(...)                        ros_ws/src/stereo_slam.py:294    (ransac_pnp.cu)
//
// OpenCV runs   for (iter = 0; iter < niters; iter++) { subset; models = solve(subset); for each model: count inliers;
//               if (count > max(best, modelPoints-1)) { best = count; niters = RANSACUpdateNumIters(...); } }
// one iteration after the other.  The minimal solvers are long serial fp64 chains (a five-point solve is ~0.25 ms of
// dependent latency on one thread) while scoring is wide, so a whole BATCH of frame pairs is advanced in rounds; each
// round is three launches whose work is spread over the whole GPU, not over one CTA per pair:
//   select : one warp per pair.  Replays the serial loop over the previous round's counts (strictly-greater update,
//            adaptive cap) — the outcome is the one the serial loop would have produced; speculative work past the
//            cap is never observed.  Pairs still running draw the next subsets with cv::RNG(2^64-1) (sequential by
//            nature: draw order = sample order) into slots of a per-round pool.
//   solve  : one THREAD per slot, slots dealt round-robin over one warp per SM sub-partition so that every warp carries
//            as few latency-bound solves as possible; a policy may suspend a rare long-running solve and hand it to
//   tail   : one WARP per suspended solve (warp-cooperative continuation, same arithmetic).
//   score  : one CTA per slab of Pol::SLAB consecutive slots of one pair: the pair's correspondences are staged in shared
//            memory once, then ONE WARP PER HYPOTHESIS counts inliers (ballot-free: per-lane count + redux).
// The first round speculates FIRST_CHUNK iterations per pair; later rounds share the pool between the pairs still
// running (a straggler that needs hundreds of iterations gets them in one round).  The number of rounds launched is
// fixed by max_iters (worst case); rounds with nothing left to do exit at their first instruction, which keeps the
// launch sequence data independent and therefore capturable in a CUDA graph.
//
// Determinism: every slot is written by exactly one thread; the only atomics are integer slot allocations whose order
// changes where a pair's slots live, never what they contain.
#pragma once
#include "common.cuh"
#include <stdlib.h>
#include "cvmath.cuh"

namespace rg {

// Per-policy tuning (Pol::FIRST_CHUNK speculative iterations per pair in round 0, Pol::SOLVE_WARPS_PER_SM solver warps
// per SM); FS_TUNE_* environment variables override them for measurement runs (read on every call: host side only).
inline int tune(const char* name, int dflt) {
    const char* e = getenv(name);
    if (!e || !*e) return dflt;
    const int v = atoi(e);
    return v > 0 ? v : dflt;
}
constexpr int SCORE_THREADS = 256;
constexpr int SELECT_THREADS = 256;
constexpr int MAX_ROUNDS = 40;

template <int MD>
struct PairState {
    unsigned long long rng;
    int n;          // correspondences of the pair (clamped to n_max)
    int niters;     // current adaptive cap
    int best;       // best inlier count so far (0 = none)
    int done;       // iterations consumed by the serial loop
    int slot_base;  // this round's slots
    int chunk;
    int base_it;    // iteration index of slot_base
    int active;     // serial loop still running
    int all_points; // n == model points: the solver runs once on all points, every point is an inlier
    int status;     // FS_PAIR_*
    double best_model[MD];
};

struct Control {
    int n_used;    // slots allocated in this round
    int n_active;  // pairs that drew subsets in this round
    int n_tail;    // solves deferred to the warp-cooperative tail in this round
    int pad;
};

constexpr int TAIL_CAP = 1024;  // deferred solves per round kept in the workspace; beyond that they finish in place

struct Plan {
    int slots_per_pair;
    int rounds;
    size_t pool;
    int first_chunk;
    int solve_warps_per_sm;
};

template <class Pol>
inline Plan make_plan(int n_pairs, int max_iters) {
    Plan pl;
    const int mi = max_iters < 1 ? 1 : max_iters;
    int first = tune(Pol::TUNE_FIRST, 0);
    if (first <= 0) {
        // Round 0 runs one solve per thread, and its latency is that of the fullest scheduler: up to ~4.4 solver warps
        // per SM the sub-partitions hold one warp each (a few hold two), beyond that every solve slows down.  Large
        // batches therefore speculate a little less per pair (measured, mono 512 pairs per step: first chunk 24 / 32 /
        // 40 / 48 / 56 -> 95k / 102k / 106k / 100k / 89k pairs/s); the pairs that need more get them in round 1.
        first = Pol::FIRST_CHUNK;
        const long long lanes = (long long)(4.4 * 32 * fs::sm_count());
        while (first > Pol::FIRST_CHUNK_MIN && (long long)n_pairs * first > lanes) first -= Pol::SLAB;
    }
    const int FIRST_CHUNK = min(256, max(Pol::SLAB, first / Pol::SLAB * Pol::SLAB));
    pl.first_chunk = FIRST_CHUNK;
    pl.solve_warps_per_sm = tune(Pol::TUNE_WARPS, Pol::SOLVE_WARPS_PER_SM);
    int spp = 256;
    // keep the launch sequence short for very large iteration caps
    while ((mi - FIRST_CHUNK + spp - 1) / spp + 1 > MAX_ROUNDS - 1) spp *= 2;
    pl.slots_per_pair = spp;
    pl.rounds = 1 + (mi > FIRST_CHUNK ? (mi - FIRST_CHUNK + spp - 1) / spp : 0);
    pl.pool = (size_t)(n_pairs > 0 ? n_pairs : 1) * spp;
    return pl;
}

template <class Pol>
struct Workspace {
    PairState<Pol::MD>* st;
    Control* ctl;  // [MAX_ROUNDS + 1]
    int* slot_pair;
    int* slot_subset;   // [pool][MP]
    int* slot_nmodels;  // [pool]
    int* slot_counts;   // [pool][MAXM]
    double* slot_models;  // [pool][MAXM][MD]
    typename Pol::Tail* tail;  // [TAIL_CAP]
    double* slot_state;  // [pool][Pol::STATE_DOUBLES]: what a two-launch solver hands from its first half to its second
    static size_t bytes(int n_pairs, const Plan& pl) {
        size_t b = 0;
        auto add = [&](size_t x) { b = fs::align_up(b, 256) + x; };
        add((size_t)(n_pairs > 0 ? n_pairs : 1) * sizeof(PairState<Pol::MD>));
        add((MAX_ROUNDS + 1) * sizeof(Control));
        add(pl.pool * sizeof(int));
        add(pl.pool * Pol::MP * sizeof(int));
        add(pl.pool * sizeof(int));
        add(pl.pool * Pol::MAXM * sizeof(int));
        add(pl.pool * Pol::MAXM * Pol::MD * sizeof(double));
        add((size_t)TAIL_CAP * sizeof(typename Pol::Tail));
        add(pl.pool * (size_t)Pol::STATE_DOUBLES * sizeof(double));
        return fs::align_up(b, 256);
    }
    void carve(fs::Carver& c, int n_pairs, const Plan& pl) {
        st = c.take<PairState<Pol::MD>>((size_t)(n_pairs > 0 ? n_pairs : 1));
        ctl = c.take<Control>(MAX_ROUNDS + 1);
        slot_pair = c.take<int>(pl.pool);
        slot_subset = c.take<int>(pl.pool * Pol::MP);
        slot_nmodels = c.take<int>(pl.pool);
        slot_counts = c.take<int>(pl.pool * Pol::MAXM);
        slot_models = c.take<double>(pl.pool * Pol::MAXM * Pol::MD);
        tail = c.take<typename Pol::Tail>(TAIL_CAP);
        slot_state = c.take<double>(pl.pool * (size_t)Pol::STATE_DOUBLES);
    }
};

// ------------------------------------------------------------------------------------------------ select
template <class Pol>
__global__ void __launch_bounds__(SELECT_THREADS)
k_rg_select(int round, int n_pairs, const int32_t* __restrict__ n_corr, int n_max, int min_points, int max_iters,
            double confidence, int pool, int first_chunk, Workspace<Pol> ws) {
    constexpr int MP = Pol::MP, MAXM = Pol::MAXM, MD = Pol::MD;
    const int lane = threadIdx.x & 31;
    const int p = blockIdx.x * (SELECT_THREADS / 32) + (threadIdx.x >> 5);
    if (p >= n_pairs) return;
    if (round > 0 && ws.ctl[round - 1].n_used == 0) return;
    PairState<MD>& S = ws.st[p];
    int n, nit, best, done, active, allp;
    unsigned long long rng_state;
    if (round == 0) {
        n = min(n_corr[p], n_max);
        nit = max(max_iters, 1);
        best = 0;
        done = 0;
        active = n >= max(min_points, MP);
        allp = active && n == MP;
        rng_state = 0xFFFFFFFFFFFFFFFFull;  // RNG rng((uint64)-1) inside run()
        if (lane == 0) {
            S.n = n;
            S.status = active ? FS_PAIR_NO_MODEL : FS_PAIR_TOO_FEW;
            S.chunk = 0;
        }
    } else {
        active = S.active;
        if (!active) return;
        n = S.n;
        nit = S.niters;
        best = S.best;
        allp = S.all_points;
        rng_state = S.rng;
        const int base_it = S.base_it, chunk = S.chunk, slot_base = S.slot_base;
        // ---- replay the serial loop over the previous round's slots (all lanes run the same uniform code)
        int last_update = -1;
        bool stop = false;
        for (int s0 = 0; s0 < chunk && !stop; s0 += 32) {
            if (base_it + s0 >= nit) break;
            int mymax = -1;
            if (s0 + lane < chunk) {
                const int slot = slot_base + s0 + lane;
                const int nm = ws.slot_nmodels[slot];
                for (int k = 0; k < nm; k++) mymax = max(mymax, ws.slot_counts[slot * MAXM + k]);
            }
            unsigned cand = __ballot_sync(0xffffffffu, mymax > max(best, MP - 1));
            while (cand) {
                const int l = __ffs(cand) - 1;
                cand &= cand - 1;
                const int it = base_it + s0 + l;
                if (it >= nit) {
                    stop = true;
                    break;
                }
                const int slot = slot_base + s0 + l;
                const int nm = ws.slot_nmodels[slot];
                for (int k = 0; k < nm; k++) {
                    const int c = ws.slot_counts[slot * MAXM + k];
                    if (c > max(best, MP - 1)) {
                        best = c;
                        if (lane < MD) S.best_model[lane] = ws.slot_models[((size_t)slot * MAXM + k) * MD + lane];
                        // confidence > 1: adaptive cap off, every one of max_iters hypotheses is solved and scored
                        // (BASELINE config 4 sweep; cv2 has no such mode)
                        if (confidence <= 1.0) nit = cvm::ransac_update_num_iters(confidence, (double)(n - c) / n, MP, nit);
                        last_update = it;
                    }
                }
            }
        }
        done = max(last_update + 1, min(base_it + chunk, nit));
        if (allp || done >= nit) active = 0;
    }
    // ---- next round's slots
    int chunk = 0, slot_base = 0;
    if (active) {
        if (allp) {
            chunk = 1;
        } else if (round == 0) {
            chunk = min(first_chunk, nit);
        } else {
            const int prev_active = max(ws.ctl[round - 1].n_active, 1);
            const int share = max((pool / prev_active) & ~(Pol::SLAB - 1), Pol::SLAB);
            chunk = min(nit - done, share);
        }
        const int alloc = (chunk + Pol::SLAB - 1) / Pol::SLAB * Pol::SLAB;
        if (lane == 0) {
            slot_base = atomicAdd(&ws.ctl[round].n_used, alloc);
            atomicAdd(&ws.ctl[round].n_active, 1);
        }
        slot_base = __shfl_sync(0xffffffffu, slot_base, 0);
        if (slot_base + alloc > pool) {  // cannot happen (shares sum to <= pool); never write out of bounds
            chunk = max(0, min(chunk, pool - slot_base));
        }
        for (int s = lane; s < alloc && slot_base + s < pool; s += 32) ws.slot_pair[slot_base + s] = s < chunk ? p : -1;
        if (lane == 0) {
            // getSubset: draw until MP distinct indices, draw order kept
            cvm::RNG rng(1);
            rng.state = rng_state;
            for (int s = 0; s < chunk; s++) {
                int* idx = ws.slot_subset + (size_t)(slot_base + s) * MP;
                if (allp) {
                    for (int i = 0; i < MP; i++) idx[i] = i;
                    continue;
                }
                int loc[MP];
                for (int i = 0; i < MP; i++) {
                    int v;
                    bool dup;
                    do {
                        v = rng.uniform(0, n);
                        dup = false;
                        for (int k = 0; k < i; k++) dup |= loc[k] == v;
                    } while (dup);
                    loc[i] = v;
                }
                for (int i = 0; i < MP; i++) idx[i] = loc[i];
            }
            rng_state = rng.state;
        }
    }
    if (lane == 0) {
        S.rng = rng_state;
        S.niters = nit;
        S.best = best;
        S.done = done;
        S.active = active;
        S.all_points = allp;
        S.slot_base = slot_base;
        S.chunk = chunk;
        S.base_it = done;
        if (best > 0) S.status = FS_PAIR_OK;
    }
}

// ------------------------------------------------------------------------------------------------ solve
template <class Pol>
__global__ void __launch_bounds__(Pol::SOLVE_THREADS)
k_rg_solve(int round, typename Pol::Params prm, Workspace<Pol> ws) {
    extern __shared__ double rg_solve_smem[];  // Pol::SOLVE_SMEM_DOUBLES per thread, interleaved by thread
    const int used = ws.ctl[round].n_used;
    if (used == 0) return;
    constexpr int SOLVE_THREADS = Pol::SOLVE_THREADS;
    const int nwarps = gridDim.x * (SOLVE_THREADS / 32);
    const int w = blockIdx.x * (SOLVE_THREADS / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
#pragma unroll 1  // one copy of the (fully inlined, tens of thousands of instructions) solver: instruction cache
    for (int slot = lane * nwarps + w; slot < used; slot += 32 * nwarps) {
        const int p = ws.slot_pair[slot];
        if (p < 0) continue;
        int sub[Pol::MP];
#pragma unroll
        for (int i = 0; i < Pol::MP; i++) sub[i] = ws.slot_subset[(size_t)slot * Pol::MP + i];
        const int nm = Pol::solve(prm, p, sub, ws.slot_models + (size_t)slot * Pol::MAXM * Pol::MD, slot, ws.tail,
                                  &ws.ctl[round].n_tail, rg_solve_smem + threadIdx.x,
                                  ws.slot_state + (size_t)slot * Pol::STATE_DOUBLES);
        if (nm >= 0) ws.slot_nmodels[slot] = nm;  // < 0: deferred to k_rg_tail / finished by k_rg_solve2
    }
}

// second half of a two-launch solver (Pol::STATE_DOUBLES > 0): one thread per slot, no shared memory, so that it runs
// at the occupancy the registers allow while the first half is held to the warps its per-thread matrices leave room for
template <class Pol>
__global__ void __launch_bounds__(Pol::SOLVE2_THREADS)
k_rg_solve2(int round, typename Pol::Params prm, Workspace<Pol> ws) {
    const int used = ws.ctl[round].n_used;
    const int slot = blockIdx.x * Pol::SOLVE2_THREADS + threadIdx.x;
    if (slot >= used || ws.slot_pair[slot] < 0) return;
    ws.slot_nmodels[slot] = Pol::solve2(prm, ws.slot_state + (size_t)slot * Pol::STATE_DOUBLES,
                                        ws.slot_models + (size_t)slot * Pol::MAXM * Pol::MD);
}

// solves the solve pass suspended (rare, long-running): one WARP per deferred solve.  One warp per CTA so that the
// entry index is a function of blockIdx only: control flow is then provably warp-uniform and the shuffles of the
// cooperative continuation need no convergence sequence around them.
template <class Pol>
__global__ void __launch_bounds__(32)
k_rg_tail(int round, typename Pol::Params prm, Workspace<Pol> ws) {
    const int n_tail = min(ws.ctl[round].n_tail, TAIL_CAP);
    for (int t = blockIdx.x; t < n_tail; t += gridDim.x) {
        const int slot = ws.tail[t].slot;
        const int nm = Pol::solve_tail(prm, ws.tail[t], ws.slot_models + (size_t)slot * Pol::MAXM * Pol::MD);
        if (threadIdx.x == 0) ws.slot_nmodels[slot] = nm;
    }
}

// ------------------------------------------------------------------------------------------------ score
template <class Pol>
__global__ void __launch_bounds__(SCORE_THREADS)
k_rg_score(int round, typename Pol::Params prm, Workspace<Pol> ws) {
    extern __shared__ __align__(16) unsigned char rg_smem[];
    __shared__ int s_off[Pol::SLAB + 1];
    const int used = ws.ctl[round].n_used;
    const int slab0 = blockIdx.x * Pol::SLAB;
    if (slab0 >= used) return;
    const int p = ws.slot_pair[slab0];
    if (p < 0) return;
    const int n = ws.st[p].n;
    Pol::stage(prm, p, n, rg_smem);
    if (threadIdx.x == 0) {
        int o = 0;
        for (int s = 0; s < Pol::SLAB; s++) {
            s_off[s] = o;
            o += (ws.slot_pair[slab0 + s] == p) ? ws.slot_nmodels[slab0 + s] : 0;
        }
        s_off[Pol::SLAB] = o;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int H = s_off[Pol::SLAB];
    for (int h = warp; h < H; h += SCORE_THREADS / 32) {
        int s = 0;
#pragma unroll
        for (int j = 1; j < Pol::SLAB; j++) s += (s_off[j] <= h) ? 1 : 0;
        const int k = h - s_off[s];
        const int slot = slab0 + s;
        typename Pol::Scorer sc(prm, ws.slot_models + ((size_t)slot * Pol::MAXM + k) * Pol::MD);
        int cnt = 0;
        for (int i = lane; i < n; i += 32) cnt += sc.inlier(rg_smem, prm, i) ? 1 : 0;
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (lane == 0) ws.slot_counts[slot * Pol::MAXM + k] = cnt;
    }
}

// launches the whole round sequence on `stream`
template <class Pol>
int run_rounds(int n_pairs, const int32_t* n_corr, int n_max, int min_points, int max_iters, double confidence,
               const typename Pol::Params& prm, const Plan& pl, Workspace<Pol>& ws, size_t stage_smem,
               cudaStream_t stream) {
    FS_CUDA(cudaMemsetAsync(ws.ctl, 0, (MAX_ROUNDS + 1) * sizeof(Control), stream));
    {
        static int granted[64];
        const int st_ = fs::ensure_dyn_smem(k_rg_score<Pol>, granted, stage_smem);
        if (st_ != FS_OK) return st_;
    }
    {
        // solvers that keep their matrices in local memory ask for the largest L1 the SM can give; a policy with
        // SOLVE_SMEM_DOUBLES > 0 keeps its hot matrix in shared memory instead (EPnP's 12x12 Jacobi SVD).  Sticky per device and set during warm-up, never under graph capture.
        static bool hinted[64];
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
        if (!hinted[dev]) {
            if (Pol::SOLVE_SMEM_DOUBLES == 0 && tune("FS_TUNE_CARVEOUT", 1) == 1)
                FS_CUDA(cudaFuncSetAttribute(k_rg_solve<Pol>, cudaFuncAttributePreferredSharedMemoryCarveout, 0));
            if (Pol::STATE_DOUBLES > 0 && tune("FS_TUNE_CARVEOUT", 1) == 1)
                FS_CUDA(cudaFuncSetAttribute(k_rg_solve2<Pol>, cudaFuncAttributePreferredSharedMemoryCarveout, 0));
            hinted[dev] = true;
        }
    }
    constexpr int SOLVE_THREADS = Pol::SOLVE_THREADS;
    const size_t solve_smem = (size_t)Pol::SOLVE_SMEM_DOUBLES * SOLVE_THREADS * sizeof(double);
    if (solve_smem > 0) {
        static int granted_solve[64];
        const int st_ = fs::ensure_dyn_smem(k_rg_solve<Pol>, granted_solve, solve_smem);
        if (st_ != FS_OK) return st_;
    }
    const int pool = (int)pl.pool;
    const int sel_grid = (n_pairs + SELECT_THREADS / 32 - 1) / (SELECT_THREADS / 32);
    // solve: one warp per 32 pool slots at most
    int solve_warps = (pool + 31) / 32;
    // E (five-point, fp64-pipe heavy): beyond one solver warp per SM sub-partition the fp64 issue slots are shared and
    // every solve gets slower; P (EPnP: 12x12 Jacobi SVD in local memory, latency bound at 255 registers) takes two
    // ... unless round 0 alone has more solves than that many warps have lanes: a thread that runs two solves back to
    // back doubles the round's latency, a sub-partition that hosts a second warp stretches it by ~1.3x (measured, mono
    // 512 pairs per step: 4.2 ms with 4 warps per SM, 2.9 ms with 6), so the policy may allow more warps then
    const int want = (int)(((long long)n_pairs * pl.first_chunk + 31) / 32);
    const int per_sm = max(pl.solve_warps_per_sm,
                           min(Pol::SOLVE_WARPS_PER_SM_MAX, (want + fs::sm_count() - 1) / fs::sm_count()));
    solve_warps = min(solve_warps, fs::sm_count() * (tune(Pol::TUNE_WARPS, 0) > 0 ? pl.solve_warps_per_sm : per_sm));
    const int solve_grid = (solve_warps + SOLVE_THREADS / 32 - 1) / (SOLVE_THREADS / 32);
    const int score_grid = (pool + Pol::SLAB - 1) / Pol::SLAB;
    for (int r = 0; r <= pl.rounds; r++) {
        k_rg_select<Pol><<<sel_grid, SELECT_THREADS, 0, stream>>>(r, n_pairs, n_corr, n_max, min_points, max_iters,
                                                                  confidence, pool, pl.first_chunk, ws);
        FS_LAUNCH_CHECK("k_rg_select");
        if (r == pl.rounds) break;  // the last select only consumes the last round's counts
        k_rg_solve<Pol><<<solve_grid, SOLVE_THREADS, solve_smem, stream>>>(r, prm, ws);
        FS_LAUNCH_CHECK("k_rg_solve");
        if (Pol::HAS_TAIL) {
            k_rg_tail<Pol><<<256, 32, 0, stream>>>(r, prm, ws);
            FS_LAUNCH_CHECK("k_rg_tail");
        }
        if constexpr (Pol::STATE_DOUBLES > 0) {
            // round 0 uses first_chunk slots per pair at most; later rounds the whole pool
            const int slots = (r == 0) ? min(pool, n_pairs * ((pl.first_chunk + Pol::SLAB - 1) / Pol::SLAB * Pol::SLAB)) : pool;
            k_rg_solve2<Pol><<<(slots + Pol::SOLVE2_THREADS - 1) / Pol::SOLVE2_THREADS, Pol::SOLVE2_THREADS, 0, stream>>>(r, prm, ws);
            FS_LAUNCH_CHECK("k_rg_solve2");
        }
        // round 0 uses first_chunk slots per pair at most
        const int sg = (r == 0) ? min(score_grid, (n_pairs * pl.first_chunk + Pol::SLAB - 1) / Pol::SLAB) : score_grid;
        k_rg_score<Pol><<<sg, SCORE_THREADS, stage_smem, stream>>>(r, prm, ws);
        FS_LAUNCH_CHECK("k_rg_score");
    }
    return FS_OK;
}

}  // namespace rg
