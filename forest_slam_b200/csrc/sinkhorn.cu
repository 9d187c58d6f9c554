// sinkhorn.cu — SURVEY.md §8f rank 2: the SuperGlue score head AFTER the contraction, as the reference runs it:
//   log_optimal_transport(scores, bin_score, iters=20)    SuperGluePretrainedNetwork/models/superglue.py:143-172
//   mutual arg-max + exp + match_threshold + -1 fill     SuperGluePretrainedNetwork/models/superglue.py:268-285
//   hyper-parameters                                     mono_slam.py:59-63 (sinkhorn_iterations 20, threshold 0.2)
//
// Arithmetic (fp32, like the reference): couplings = [[S, alpha], [alpha, alpha]] of size (m+1) x (n+1);
//   u <- log_mu - logsumexp_j(Z + v);  v <- log_nu - logsumexp_i(Z + u)   (20 times, u = v = 0 at start)
//   Zf = ((Z + u) + v) - norm;  row / column arg-max over the m x n block;  mutual;  exp;  threshold.
// The dust-bin row and column are all alpha, so they are never stored: their terms enter analytically.
//
// Design (B200): the score matrix (4 MB per pair at 1024 x 1024) is the only large object and is streamed ONCE per
// Sinkhorn iteration, not twice: a CTA owns a band of R rows, keeps it in shared memory (96 KB => two CTAs per SM, so
// one CTA's load phase overlaps the other's exp phase), computes the band's u from registers (a warp per row, the row
// lives in registers between the max and the exp-sum), then — with u known — the band's partial column log-sum-exp
// (max, sum) from shared memory.  The last CTA of a pair to finish (threadfence + counter) folds the partials into v
// and updates the two dust-bin scalars.  Pairs are processed in groups whose matrices fit in the 126 MB L2, all 20
// iterations of a group back to back, so after the first iteration the stream comes from L2, not HBM.
// Bounds: 2 MUFU.EX2 per matrix element per iteration (16 lanes/clk/SM) and 4 B per element per iteration from L2/HBM.
// The final pass reuses the same kernel (mode 1): row arg-max with v, partial column arg-max with u, and the last CTA
// resolves mutual / threshold — the reference's elementwise expression ((Z + u) + v) - norm is evaluated literally so
// that ties fall exactly where torch.max puts them (lowest index) for identical u, v.
#include <math_constants.h>
#include "common.cuh"

namespace {

constexpr int SK_THREADS = 256;
constexpr int SK_WARPS = SK_THREADS / 32;
constexpr int SK_BAND_FLOATS = 24576;  // 96 KB of scores per CTA
constexpr int SK_MAX_N = 4096;

struct SkGeom {
    int n_pad, R, bands_max;
    size_t smem;
};

inline SkGeom sk_geom(int m_max, int n_max) {
    SkGeom g;
    g.n_pad = (int)fs::align_up((size_t)n_max, 128);
    g.R = SK_BAND_FLOATS / g.n_pad;
    if (g.R < 1) g.R = 1;
    if (g.R > m_max && m_max > 0) g.R = m_max;
    g.bands_max = m_max > 0 ? (m_max + g.R - 1) / g.R : 0;
    g.smem = ((size_t)g.R * g.n_pad + g.n_pad + fs::align_up((size_t)g.R, 4)) * sizeof(float);
    return g;
}

// workspace layout (per pair strides)
struct SkWs {
    float* u;          // [P, m_max + 1]   (slot m_max = dust-bin row)
    float* v;          // [P, n_pad + 4]   (slot n_max = dust-bin column)
    float2* colpart;   // [P, bands_max, n_pad]
    float* rb_val;     // [P, m_max]
    int32_t* rb_idx;   // [P, m_max]
    unsigned* counter; // [P]
    int32_t* i1;       // [P, n_pad]  column arg-max of the final pass (resident kernel)
    float2* upart;     // [P, bands_max] (max, sum) of exp(u) over a band's rows       (resident kernel)
    float2* vpart;     // [P, bands_max] (max, sum) of exp(v) over a band's column slice
    size_t bytes;
};

inline SkWs sk_carve(void* base, int P, int m_max, int n_max) {
    SkGeom g = sk_geom(m_max, n_max);
    fs::Carver c(base);
    SkWs w;
    w.u = c.take<float>((size_t)P * (m_max + 1));
    w.v = c.take<float>((size_t)P * (g.n_pad + 4));
    w.colpart = c.take<float2>((size_t)P * g.bands_max * g.n_pad);
    w.rb_val = c.take<float>((size_t)P * m_max);
    w.rb_idx = c.take<int32_t>((size_t)P * m_max);
    w.counter = c.take<unsigned>((size_t)P);
    w.i1 = c.take<int32_t>((size_t)P * g.n_pad);
    w.upart = c.take<float2>((size_t)P * g.bands_max);
    w.vpart = c.take<float2>((size_t)P * g.bands_max);
    w.bytes = fs::align_up(c.off, 256);
    return w;
}

__device__ __forceinline__ float warp_max(float x) {
#pragma unroll
    for (int o = 16; o; o >>= 1) x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, o));
    return x;
}
__device__ __forceinline__ float warp_sum(float x) {
#pragma unroll
    for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    return x;
}
__device__ __forceinline__ float block_max(float x, float* red) {
    x = warp_max(x);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = x;
    __syncthreads();
    float r = red[0];
#pragma unroll
    for (int w = 1; w < SK_WARPS; ++w) r = fmaxf(r, red[w]);
    return r;
}
__device__ __forceinline__ float block_sum(float x, float* red) {
    x = warp_sum(x);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = x;
    __syncthreads();
    float r = red[0];
#pragma unroll
    for (int w = 1; w < SK_WARPS; ++w) r += red[w];  // fixed order: deterministic
    return r;
}

// u = v = 0 at the start (superglue.py:145): v is zeroed, and the dust-bin row's first u, which only depends on v = 0,
// is log_mu_bin - logsumexp_{j<=n}(alpha) = log_mu_bin - (alpha + log(n + 1)).  Outputs get their "no match" defaults.
__global__ void k_sk_init(int P, int m_max, int n_max, int n_pad, const int32_t* __restrict__ mcnt,
                          const int32_t* __restrict__ ncnt, float alpha, float* __restrict__ u_all,
                          float* __restrict__ v_all, unsigned* __restrict__ counters, int32_t* __restrict__ matches0,
                          int32_t* __restrict__ matches1, float* __restrict__ ms0, float* __restrict__ ms1) {
    const int p = blockIdx.x;
    const int m = min(mcnt[p], m_max), n = min(ncnt[p], n_max);
    float* v = v_all + (size_t)p * (n_pad + 4);
    for (int c = threadIdx.x; c < n_pad + 4; c += blockDim.x) v[c] = 0.f;
    for (int i = threadIdx.x; i < m_max; i += blockDim.x) u_all[(size_t)p * (m_max + 1) + i] = 0.f;  // iters == 0
    for (int i = threadIdx.x; i < m_max; i += blockDim.x) {
        matches0[(size_t)p * m_max + i] = -1;
        ms0[(size_t)p * m_max + i] = 0.f;
    }
    for (int j = threadIdx.x; j < n_max; j += blockDim.x) {
        matches1[(size_t)p * n_max + j] = -1;
        ms1[(size_t)p * n_max + j] = 0.f;
    }
    if (threadIdx.x == 0) {
        counters[p] = 0;
        float ub = 0.f;
        if (m > 0 && n > 0) {
            const float norm = -logf((float)m + (float)n);
            ub = (logf((float)n) + norm) - (alpha + logf((float)n + 1.f));
        }
        u_all[(size_t)p * (m_max + 1) + m_max] = ub;
    }
}

template <int NPL4>
__global__ void __launch_bounds__(SK_THREADS, 2)
k_sk_band(const float* __restrict__ scores, int m_max, int n_max, const int32_t* __restrict__ mcnt,
          const int32_t* __restrict__ ncnt, int pair0, float alpha, int R, int n_pad, int bands_max, int final_pass,
          float thr, float* u_all, float* v_all, float2* colpart_all, float* rb_val_all, int32_t* rb_idx_all,
          unsigned* counters, int32_t* __restrict__ matches0, int32_t* __restrict__ matches1,
          float* __restrict__ ms0, float* __restrict__ ms1, float* __restrict__ u_out, float* __restrict__ v_out) {
    extern __shared__ __align__(16) float sk_smem[];
    float* zs = sk_smem;                    // [R, n_pad] raw scores of the band (-inf outside the n valid columns)
    float* v_s = zs + (size_t)R * n_pad;    // [n_pad]
    float* u_s = v_s + n_pad;               // [R]
    __shared__ float red[SK_WARPS];
    __shared__ int s_last;

    const int p = pair0 + blockIdx.y;
    const int m = min(mcnt[p], m_max), n = min(ncnt[p], n_max);
    if (m <= 0 || n <= 0) return;
    const int bands = (m + R - 1) / R;
    const int band = blockIdx.x;
    if (band >= bands) return;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float NEG = -CUDART_INF_F;
    const float norm = -logf((float)m + (float)n);
    const float log_mu_bin = logf((float)n) + norm, log_nu_bin = logf((float)m) + norm;
    float* u = u_all + (size_t)p * (m_max + 1);
    float* v = v_all + (size_t)p * (n_pad + 4);
    const float* Z = scores + (size_t)p * m_max * n_max;

    for (int c = tid; c < n_pad; c += SK_THREADS) v_s[c] = (c < n) ? v[c] : 0.f;
    const float v_bin = v[n_max];
    __syncthreads();

    const int r0 = band * R;
    const int Rv = min(R, m - r0);  // valid rows of this band

    // ---------------------------------------------------------------- row phase: a warp per row, row in registers
    for (int rr = warp; rr < Rv; rr += SK_WARPS) {
        const int i = r0 + rr;
        const float* zrow = Z + (size_t)i * n_max;
        // rows up to 1024 columns live in registers between the passes; wider rows are re-read from shared memory
        constexpr bool kInRegs = (NPL4 <= 8);
        float4 z[kInRegs ? NPL4 : 1];
        float* zsrow = zs + (size_t)rr * n_pad;
        const float ui_final = final_pass ? __ldcg(u + i) : 0.f;
        float mx = NEG;   // iteration: row max of z + v;  final pass: best value
        int bj = 0x7fffffff;
#pragma unroll
        for (int q = 0; q < NPL4; ++q) {
            const int c = (q * 32 + lane) * 4;
            if (c >= n_pad) continue;
            float4 t = make_float4(NEG, NEG, NEG, NEG);
            if (c < n) {  // n_max % 4 == 0, so a float4 never straddles the row end
                t = __ldg(reinterpret_cast<const float4*>(zrow + c));
                if (c + 1 >= n) t.y = NEG;
                if (c + 2 >= n) t.z = NEG;
                if (c + 3 >= n) t.w = NEG;
            }
            *reinterpret_cast<float4*>(zsrow + c) = t;
            const float4 vv = *reinterpret_cast<const float4*>(v_s + c);
            if (!final_pass) {
                t.x += vv.x; t.y += vv.y; t.z += vv.z; t.w += vv.w;
                mx = fmaxf(fmaxf(mx, fmaxf(t.x, t.y)), fmaxf(t.z, t.w));
                if constexpr (kInRegs) z[q] = t;
            } else {
                const float a0 = ((t.x + ui_final) + vv.x) - norm, a1 = ((t.y + ui_final) + vv.y) - norm;
                const float a2 = ((t.z + ui_final) + vv.z) - norm, a3 = ((t.w + ui_final) + vv.w) - norm;
                if (a0 > mx) { mx = a0; bj = c; }
                if (a1 > mx) { mx = a1; bj = c + 1; }
                if (a2 > mx) { mx = a2; bj = c + 2; }
                if (a3 > mx) { mx = a3; bj = c + 3; }
            }
        }
        if (!final_pass) {
            const float b = alpha + v_bin;
            mx = fmaxf(warp_max(mx), b);
            float s = 0.f;
#pragma unroll
            for (int q = 0; q < NPL4; ++q) {
                const int c = (q * 32 + lane) * 4;
                if (c >= n_pad) continue;
                float4 t;
                if constexpr (kInRegs) {
                    t = z[q];
                } else {
                    t = *reinterpret_cast<const float4*>(zsrow + c);  // own writes: no barrier needed
                    const float4 vv = *reinterpret_cast<const float4*>(v_s + c);
                    t.x += vv.x; t.y += vv.y; t.z += vv.z; t.w += vv.w;
                }
                s += (__expf(t.x - mx) + __expf(t.y - mx)) + (__expf(t.z - mx) + __expf(t.w - mx));
            }
            s = warp_sum(s) + __expf(b - mx);
            const float ui = norm - (mx + logf(s));  // log_mu of an ordinary row is `norm`
            if (lane == 0) {
                u_s[rr] = ui;
                u[i] = ui;
            }
        } else {
            const float ui = ui_final;
            float best = mx;
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
                if (ob > best || (ob == best && oj < bj)) { best = ob; bj = oj; }
            }
            if (lane == 0) {
                u_s[rr] = ui;
                rb_val_all[(size_t)p * m_max + i] = best;
                rb_idx_all[(size_t)p * m_max + i] = bj;
            }
        }
    }
    __syncthreads();

    // ---------------------------------------------------------------- column phase: a thread per column, from smem
    float2* cp = colpart_all + ((size_t)p * bands_max + band) * n_pad;
    for (int c = tid; c < n; c += SK_THREADS) {
        const float* zc = zs + c;
        if (!final_pass) {
            float cmax = NEG;
            for (int rr = 0; rr < Rv; ++rr) cmax = fmaxf(cmax, zc[(size_t)rr * n_pad] + u_s[rr]);
            float s = 0.f;
            for (int rr = 0; rr < Rv; ++rr) s += __expf((zc[(size_t)rr * n_pad] + u_s[rr]) - cmax);
            cp[c] = make_float2(cmax, s);
        } else {
            const float vj = v_s[c];
            float best = NEG;
            int bi = 0x7fffffff;
            for (int rr = 0; rr < Rv; ++rr) {
                const float a = ((zc[(size_t)rr * n_pad] + u_s[rr]) + vj) - norm;
                if (a > best) { best = a; bi = r0 + rr; }
            }
            cp[c] = make_float2(best, __int_as_float(bi));
        }
    }

    // ---------------------------------------------------------------- the last CTA of the pair folds the partials
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned prev = atomicAdd(&counters[p], 1u);
        s_last = (prev == (unsigned)(bands - 1));
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid == 0) counters[p] = 0;  // ready for the next launch

    const float2* cpp = colpart_all + (size_t)p * bands_max * n_pad;
    if (!final_pass) {
        const float u_bin = __ldcg(u + m_max);
        const float br = alpha + u_bin;  // dust-bin ROW entry of every ordinary column
        for (int c = tid; c < n; c += SK_THREADS) {
            float M = br;
            for (int b = 0; b < bands; ++b) M = fmaxf(M, __ldcg(&cpp[(size_t)b * n_pad + c]).x);
            float S = __expf(br - M);
            for (int b = 0; b < bands; ++b) {
                const float2 t = __ldcg(&cpp[(size_t)b * n_pad + c]);
                S += t.y * __expf(t.x - M);
            }
            const float vj = norm - (M + logf(S));  // log_nu of an ordinary column is `norm`
            v[c] = vj;
            v_s[c] = vj;
        }
        // dust-bin column: every entry is alpha, rows 0..m-1 and the bin row
        float mxu = NEG;
        for (int i = tid; i < m; i += SK_THREADS) mxu = fmaxf(mxu, __ldcg(u + i));
        mxu = fmaxf(block_max(mxu, red), u_bin);
        float su = 0.f;
        for (int i = tid; i < m; i += SK_THREADS) su += expf(__ldcg(u + i) - mxu);
        su = block_sum(su, red) + expf(u_bin - mxu);
        const float v_bin_new = log_nu_bin - (alpha + (mxu + logf(su)));
        // dust-bin row of the NEXT iteration: u_bin = log_mu_bin - logsumexp_{j<=n}(alpha + v_j)
        float mxv = NEG;
        for (int c = tid; c < n; c += SK_THREADS) mxv = fmaxf(mxv, v_s[c]);
        mxv = fmaxf(block_max(mxv, red), v_bin_new);
        float sv = 0.f;
        for (int c = tid; c < n; c += SK_THREADS) sv += expf(v_s[c] - mxv);
        sv = block_sum(sv, red) + expf(v_bin_new - mxv);
        if (tid == 0) {
            v[n_max] = v_bin_new;
            u[m_max] = log_mu_bin - (alpha + (mxv + logf(sv)));
        }
    } else {
        int* i1_s = reinterpret_cast<int*>(zs);  // the band is no longer needed
        __syncthreads();
        for (int c = tid; c < n; c += SK_THREADS) {
            float best = NEG;
            int bi = 0x7fffffff;
            for (int b = 0; b < bands; ++b) {  // ascending bands + strict '>' keep the lowest row on ties
                const float2 t = __ldcg(&cpp[(size_t)b * n_pad + c]);
                if (t.x > best) { best = t.x; bi = __float_as_int(t.y); }
            }
            i1_s[c] = bi;
        }
        __syncthreads();
        const float* rbv = rb_val_all + (size_t)p * m_max;
        const int32_t* rbi = rb_idx_all + (size_t)p * m_max;
        for (int i = tid; i < m; i += SK_THREADS) {
            const int j = __ldcg(rbi + i);
            const bool mutual = (j >= 0 && j < n) && (i1_s[j] == i);  // j is out of range only for NaN rows
            const float sc = mutual ? expf(__ldcg(rbv + i)) : 0.f;
            ms0[(size_t)p * m_max + i] = sc;
            matches0[(size_t)p * m_max + i] = (mutual && sc > thr) ? j : -1;
        }
        for (int j = tid; j < n; j += SK_THREADS) {
            const int i = i1_s[j];
            const bool mutual1 = (i >= 0 && i < m) && (__ldcg(rbi + i) == j);  // then i1[i0[i]] == i as well
            const float sc = mutual1 ? expf(__ldcg(rbv + i)) : 0.f;
            ms1[(size_t)p * n_max + j] = sc;
            matches1[(size_t)p * n_max + j] = (mutual1 && sc > thr) ? i : -1;
        }
        if (u_out) {
            // ordinary rows only: the m x n block of the result never reads the dust-bin row's u
            for (int i = tid; i < m; i += SK_THREADS) u_out[(size_t)p * m_max + i] = __ldcg(u + i);
        }
        if (v_out) {
            for (int j = tid; j < n; j += SK_THREADS) v_out[(size_t)p * n_max + j] = __ldcg(v + j);
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// (Measured and dropped: 44 KB bands, 64 registers, four CTAs per SM: 3.15 vs 3.43 ms per 64 pairs of 1024^2 — the
// iteration is bound by its barrier / L2 latency chain, not by occupancy — while the streaming fallback got 1.6x slower.)
// Resident variant (the default): the whole group of pairs lives in shared memory for ALL iterations, so the score
// matrix is read from HBM exactly once.  One cooperative launch per group (co-residency guaranteed by the launch
// mode); the CTAs of a pair meet at a pair-wide barrier (arrival counter in global memory, release / acquire) twice
// per iteration: A = "column partials and u written", B = "v written".  The fold of the column partials is spread
// over the pair's CTAs (CTA b folds the b-th slice of columns, 8 lanes per column), so nothing serial sits between
// the barriers; band 0 additionally owns the two dust-bin scalars.
// exp(x - m) for the resident kernel's inner loops as ONE FFMA + ONE MUFU.EX2: ex2(x * log2(e) - m * log2(e)), flushing
// denormal results (they vanish against the maximal term 1 of every sum).  __expf without -ftz expands into a
// five-instruction denormal-preserving sequence, which made these loops issue-bound (ncu: 39 instructions per element).
__device__ __forceinline__ float ex2_ftz(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
constexpr float SK_L2E = 1.4426950408889634f;
__device__ __forceinline__ float exp_sub(float x, float m_l2e) { return ex2_ftz(fmaf(x, SK_L2E, -m_l2e)); }

// (M1, S1) (+) (M2, S2) for sums of exponentials kept as S * exp(M)
__device__ __forceinline__ void lse_merge(float& M, float& S, float M2, float S2) {
    const float nm = fmaxf(M, M2);
    if (nm == -CUDART_INF_F) return;  // both empty
    S = S * ex2_ftz((M - nm) * SK_L2E) + S2 * ex2_ftz((M2 - nm) * SK_L2E);
    M = nm;
}

__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void pair_barrier(unsigned* ctr, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        while (ld_acquire(ctr) < target) {
        }
    }
    __syncthreads();
}
// all lanes of a warp: merge the per-band (max, sum) partials of a pair and return log-sum-exp including `extra`
__device__ __forceinline__ float warp_lse_partials(const float2* part, int bands, float extra, int lane) {
    float M = -CUDART_INF_F, S = 0.f;
    for (int b = lane; b < bands; b += 32) {
        const float2 t = __ldcg(part + b);
        lse_merge(M, S, t.x, t.y);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const float M2 = __shfl_xor_sync(0xffffffffu, M, o), S2 = __shfl_xor_sync(0xffffffffu, S, o);
        lse_merge(M, S, M2, S2);
    }
    lse_merge(M, S, extra, 1.f);
    return M + logf(S);
}
template <int NPL4>
__global__ void __launch_bounds__(SK_THREADS, 2)
k_sk_resident(const float* __restrict__ scores, int m_max, int n_max, const int32_t* __restrict__ mcnt,
              const int32_t* __restrict__ ncnt, int pair0, float alpha, int R, int n_pad, int bands_max, int iters,
              float thr, float* u_all, float* v_all, float2* colpart_all, float* rb_val_all, int32_t* rb_idx_all,
              int32_t* i1_all, float2* upart_all, float2* vpart_all, unsigned* counters,
              int32_t* __restrict__ matches0, int32_t* __restrict__ matches1, float* __restrict__ ms0,
              float* __restrict__ ms1, float* __restrict__ u_out, float* __restrict__ v_out) {
    extern __shared__ __align__(16) float sk_smem[];
    float* zs = sk_smem;                    // [R, n_pad] raw scores of the band (-inf outside the n valid columns)
    float* v_s = zs + (size_t)R * n_pad;    // [n_pad]
    float* u_s = v_s + n_pad;               // [R]
    __shared__ float red[SK_WARPS];
    __shared__ float s_bins[2];             // [0] v_bin, [1] u_bin of the current iteration

    const int p = pair0 + blockIdx.y;
    const int m = min(mcnt[p], m_max), n = min(ncnt[p], n_max);
    if (m <= 0 || n <= 0) return;
    const int bands = (m + R - 1) / R;
    const int band = blockIdx.x;
    if (band >= bands) return;  // never takes part in the pair's barriers (their targets count `bands`)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float NEG = -CUDART_INF_F;
    const float norm = -logf((float)m + (float)n);
    const float log_mu_bin = logf((float)n) + norm, log_nu_bin = logf((float)m) + norm;
    float* u = u_all + (size_t)p * (m_max + 1);
    float* v = v_all + (size_t)p * (n_pad + 4);
    const float* Z = scores + (size_t)p * m_max * n_max;
    float2* cpp = colpart_all + (size_t)p * bands_max * n_pad;
    float2* cp = cpp + (size_t)band * n_pad;
    int32_t* i1 = i1_all + (size_t)p * n_pad;
    float2* upart = upart_all + (size_t)p * bands_max;
    float2* vpart = vpart_all + (size_t)p * bands_max;
    unsigned* ctr = counters + p;
    unsigned target = 0;
    float v_bin = 0.f;  // v = 0 before the first iteration (superglue.py:145)

    const int r0 = band * R;
    const int Rv = min(R, m - r0);
    const int cw = (n + bands - 1) / bands;             // this CTA folds columns [c_lo, c_hi)
    const int c_lo = min(band * cw, n), c_hi = min(c_lo + cw, n);
    constexpr bool kInRegs = (NPL4 <= 8);

    // the band: HBM -> shared memory, once
    for (int rr = warp; rr < Rv; rr += SK_WARPS) {
        const float* zrow = Z + (size_t)(r0 + rr) * n_max;
        float* zsrow = zs + (size_t)rr * n_pad;
#pragma unroll
        for (int q = 0; q < NPL4; ++q) {
            const int c = (q * 32 + lane) * 4;
            if (c >= n_pad) continue;
            float4 t = make_float4(NEG, NEG, NEG, NEG);
            if (c < n) {
                t = __ldcs(reinterpret_cast<const float4*>(zrow + c));
                if (c + 1 >= n) t.y = NEG;
                if (c + 2 >= n) t.z = NEG;
                if (c + 3 >= n) t.w = NEG;
            }
            *reinterpret_cast<float4*>(zsrow + c) = t;
        }
    }

    for (int it = 0; it <= iters; ++it) {
        const bool final_pass = (it == iters);
        for (int c = tid; c < n_pad; c += SK_THREADS) v_s[c] = (c < n) ? __ldcg(v + c) : 0.f;
        // dust-bin row of this iteration, u_bin = log_mu_bin - logsumexp_{j<=n}(alpha + v_j): every CTA derives it
        // from the per-slice partials of v (nothing serial between the barriers); v = 0 at the start.
        if (warp == 0) {
            float lse_v;
            if (it == 0) lse_v = logf((float)n + 1.f);
            else lse_v = warp_lse_partials(vpart, bands, v_bin, lane);
            if (lane == 0) s_bins[1] = log_mu_bin - (alpha + lse_v);
        }
        __syncthreads();
        const float u_bin = s_bins[1];

        // ------------------------------------------------------------ row phase: a warp per row
        for (int rr = warp; rr < Rv; rr += SK_WARPS) {
            const int i = r0 + rr;
            const float* zsrow = zs + (size_t)rr * n_pad;
            float4 z[kInRegs ? NPL4 : 1];
            const float ui_final = final_pass ? __ldcg(u + i) : 0.f;
            float mx = NEG;
            int bj = 0x7fffffff;
#pragma unroll
            for (int q = 0; q < NPL4; ++q) {
                const int c = (q * 32 + lane) * 4;
                if (c >= n_pad) continue;
                float4 t = *reinterpret_cast<const float4*>(zsrow + c);
                const float4 vv = *reinterpret_cast<const float4*>(v_s + c);
                if (!final_pass) {
                    t.x += vv.x; t.y += vv.y; t.z += vv.z; t.w += vv.w;
                    mx = fmaxf(fmaxf(mx, fmaxf(t.x, t.y)), fmaxf(t.z, t.w));
                    if constexpr (kInRegs) z[q] = t;
                } else {
                    const float a0 = ((t.x + ui_final) + vv.x) - norm, a1 = ((t.y + ui_final) + vv.y) - norm;
                    const float a2 = ((t.z + ui_final) + vv.z) - norm, a3 = ((t.w + ui_final) + vv.w) - norm;
                    if (a0 > mx) { mx = a0; bj = c; }
                    if (a1 > mx) { mx = a1; bj = c + 1; }
                    if (a2 > mx) { mx = a2; bj = c + 2; }
                    if (a3 > mx) { mx = a3; bj = c + 3; }
                }
            }
            if (!final_pass) {
                const float b = alpha + v_bin;
                mx = fmaxf(warp_max(mx), b);
                const float mxl = mx * SK_L2E;
                float s = 0.f;
#pragma unroll
                for (int q = 0; q < NPL4; ++q) {
                    const int c = (q * 32 + lane) * 4;
                    if (c >= n_pad) continue;
                    float4 t;
                    if constexpr (kInRegs) {
                        t = z[q];
                    } else {
                        t = *reinterpret_cast<const float4*>(zsrow + c);
                        const float4 vv = *reinterpret_cast<const float4*>(v_s + c);
                        t.x += vv.x; t.y += vv.y; t.z += vv.z; t.w += vv.w;
                    }
                    s += (exp_sub(t.x, mxl) + exp_sub(t.y, mxl)) + (exp_sub(t.z, mxl) + exp_sub(t.w, mxl));
                }
                s = warp_sum(s) + exp_sub(b, mxl);
                const float ui = norm - (mx + logf(s));
                if (lane == 0) {
                    u_s[rr] = ui;
                    u[i] = ui;
                }
            } else {
                float best = mx;
#pragma unroll
                for (int o = 16; o; o >>= 1) {
                    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                    const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
                    if (ob > best || (ob == best && oj < bj)) { best = ob; bj = oj; }
                }
                if (lane == 0) {
                    u_s[rr] = ui_final;
                    rb_val_all[(size_t)p * m_max + i] = best;
                    rb_idx_all[(size_t)p * m_max + i] = bj;
                }
            }
        }
        __syncthreads();
        if (warp == 0 && !final_pass) {  // (max, sum) of exp(u) over this band's rows, for the dust-bin column
            float M = NEG, S = 0.f;
            for (int rr = lane; rr < Rv; rr += 32) lse_merge(M, S, u_s[rr], 1.f);
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const float M2 = __shfl_xor_sync(0xffffffffu, M, o), S2 = __shfl_xor_sync(0xffffffffu, S, o);
                lse_merge(M, S, M2, S2);
            }
            if (lane == 0) upart[band] = make_float2(M, S);
        }

        // ------------------------------------------------------------ column phase: a thread per 4 columns
        for (int c = tid * 4; c < n; c += SK_THREADS * 4) {
            const float* zc = zs + c;
            if (!final_pass) {
                float4 cm = make_float4(NEG, NEG, NEG, NEG);
#pragma unroll 4
                for (int rr = 0; rr < Rv; ++rr) {
                    const float4 z4 = *reinterpret_cast<const float4*>(zc + (size_t)rr * n_pad);
                    const float ur = u_s[rr];
                    cm.x = fmaxf(cm.x, z4.x + ur); cm.y = fmaxf(cm.y, z4.y + ur);
                    cm.z = fmaxf(cm.z, z4.z + ur); cm.w = fmaxf(cm.w, z4.w + ur);
                }
                const float4 cl = make_float4(cm.x * SK_L2E, cm.y * SK_L2E, cm.z * SK_L2E, cm.w * SK_L2E);
                float4 sm = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
                for (int rr = 0; rr < Rv; ++rr) {
                    const float4 z4 = *reinterpret_cast<const float4*>(zc + (size_t)rr * n_pad);
                    const float ur = u_s[rr];
                    sm.x += exp_sub(z4.x + ur, cl.x); sm.y += exp_sub(z4.y + ur, cl.y);
                    sm.z += exp_sub(z4.z + ur, cl.z); sm.w += exp_sub(z4.w + ur, cl.w);
                }
                // columns >= n hold -inf: (max, sum) = (-inf, NaN) there; they are never folded (c < n only)
                *reinterpret_cast<float4*>(cp + c) = make_float4(cm.x, sm.x, cm.y, sm.y);
                *reinterpret_cast<float4*>(cp + c + 2) = make_float4(cm.z, sm.z, cm.w, sm.w);
            } else {
                const float4 vj = *reinterpret_cast<const float4*>(v_s + c);
                float4 best = make_float4(NEG, NEG, NEG, NEG);
                int b0 = 0x7fffffff, b1 = 0x7fffffff, b2 = 0x7fffffff, b3 = 0x7fffffff;
                for (int rr = 0; rr < Rv; ++rr) {
                    const float4 z4 = *reinterpret_cast<const float4*>(zc + (size_t)rr * n_pad);
                    const float ur = u_s[rr];
                    const float a0 = ((z4.x + ur) + vj.x) - norm, a1 = ((z4.y + ur) + vj.y) - norm;
                    const float a2 = ((z4.z + ur) + vj.z) - norm, a3 = ((z4.w + ur) + vj.w) - norm;
                    if (a0 > best.x) { best.x = a0; b0 = r0 + rr; }
                    if (a1 > best.y) { best.y = a1; b1 = r0 + rr; }
                    if (a2 > best.z) { best.z = a2; b2 = r0 + rr; }
                    if (a3 > best.w) { best.w = a3; b3 = r0 + rr; }
                }
                *reinterpret_cast<float4*>(cp + c) = make_float4(best.x, __int_as_float(b0), best.y, __int_as_float(b1));
                *reinterpret_cast<float4*>(cp + c + 2) = make_float4(best.z, __int_as_float(b2), best.w, __int_as_float(b3));
            }
        }
        target += (unsigned)bands;
        pair_barrier(ctr, target);  // A: every band's column partials and u partials are visible

        // ------------------------------------------------------------ fold this CTA's slice of columns
        const int l8 = tid & 7;
        float vM = NEG, vS = 0.f;  // (max, sum) of exp(v) over the slice (lanes with l8 == 0 contribute)
        for (int cb = c_lo; cb < c_hi; cb += SK_THREADS / 8) {
            const int c = cb + (tid >> 3);
            const bool live = c < c_hi;
            if (!final_pass) {
                float M = NEG, S = 0.f;
                for (int b0 = 0; b0 < bands; b0 += 64) {  // 8 independent loads in flight per lane
                    float2 t[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int b = b0 + k * 8 + l8;
                        t[k] = (live && b < bands) ? __ldcg(&cpp[(size_t)b * n_pad + c]) : make_float2(NEG, 0.f);
                    }
#pragma unroll
                    for (int k = 0; k < 8; ++k) lse_merge(M, S, t[k].x, t[k].y);
                }
#pragma unroll
                for (int o = 4; o; o >>= 1) {
                    const float M2 = __shfl_xor_sync(0xffffffffu, M, o), S2 = __shfl_xor_sync(0xffffffffu, S, o);
                    lse_merge(M, S, M2, S2);
                }
                lse_merge(M, S, alpha + u_bin, 1.f);  // dust-bin ROW entry of every ordinary column
                if (live && l8 == 0) {
                    const float vj = norm - (M + logf(S));
                    v[c] = vj;
                    lse_merge(vM, vS, vj, 1.f);
                }
            } else {
                float best = NEG;
                int bi = 0x7fffffff;
                if (live)
                    for (int b = l8; b < bands; b += 8) {
                        const float2 t = __ldcg(&cpp[(size_t)b * n_pad + c]);
                        if (t.x > best) { best = t.x; bi = __float_as_int(t.y); }
                    }
#pragma unroll
                for (int o = 4; o; o >>= 1) {
                    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
                }
                if (live && l8 == 0) i1[c] = bi;
            }
        }
        if (!final_pass) {
            // slice partial of v for the next iteration's u_bin (fixed order: lanes, then warps)
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const float M2 = __shfl_xor_sync(0xffffffffu, vM, o), S2 = __shfl_xor_sync(0xffffffffu, vS, o);
                lse_merge(vM, vS, M2, S2);
            }
            __shared__ float2 s_vp[SK_WARPS];
            if (lane == 0) s_vp[warp] = make_float2(vM, vS);
            __syncthreads();
            if (warp == 0) {
                // dust-bin column: v_bin = log_nu_bin - (alpha + logsumexp(u_0..u_{m-1}, u_bin)), from the band partials
                const float lse_u = warp_lse_partials(upart, bands, u_bin, lane);
                if (tid == 0) {
                    float M = NEG, S = 0.f;
                    for (int w = 0; w < SK_WARPS; ++w) lse_merge(M, S, s_vp[w].x, s_vp[w].y);
                    vpart[band] = make_float2(M, S);
                    s_bins[0] = log_nu_bin - (alpha + lse_u);
                }
            }
        }
        target += (unsigned)bands;
        pair_barrier(ctr, target);  // B: v and its slice partials (or the column arg-max) are complete
        if (!final_pass) v_bin = s_bins[0];
    }

    // ---------------------------------------------------------------- mutual / exp / threshold on this CTA's slices
    const float* rbv = rb_val_all + (size_t)p * m_max;
    const int32_t* rbi = rb_idx_all + (size_t)p * m_max;
    for (int rr = tid; rr < Rv; rr += SK_THREADS) {
        const int i = r0 + rr;
        const int j = __ldcg(rbi + i);
        const bool mutual = (j >= 0 && j < n) && (__ldcg(i1 + j) == i);  // j is out of range only for NaN rows
        const float sc = mutual ? expf(__ldcg(rbv + i)) : 0.f;
        ms0[(size_t)p * m_max + i] = sc;
        matches0[(size_t)p * m_max + i] = (mutual && sc > thr) ? j : -1;
        if (u_out) u_out[(size_t)p * m_max + i] = u_s[rr];
    }
    for (int j = c_lo + tid; j < c_hi; j += SK_THREADS) {
        const int i = __ldcg(i1 + j);
        const bool mutual1 = (i >= 0 && i < m) && (__ldcg(rbi + i) == j);
        const float sc = mutual1 ? expf(__ldcg(rbv + i)) : 0.f;
        ms1[(size_t)p * n_max + j] = sc;
        matches1[(size_t)p * n_max + j] = (mutual1 && sc > thr) ? i : -1;
        if (v_out) v_out[(size_t)p * n_max + j] = v_s[j];
    }
}

using ResidentKernel = void (*)(const float*, int, int, const int32_t*, const int32_t*, int, float, int, int, int, int,
                                float, float*, float*, float2*, float*, int32_t*, int32_t*, float2*, float2*, unsigned*,
                                int32_t*, int32_t*, float*, float*, float*, float*);
struct ResidentChoice {
    ResidentKernel fn;
    int* granted;
    int* ctas_per_sm;  // [64] per device, 0 = not asked yet
};
ResidentChoice sk_pick_resident(int n_pad) {
    static int g2[64], g4[64], g8[64], g16[64], g32[64];
    static int o2[64], o4[64], o8[64], o16[64], o32[64];
    const int q = n_pad / 128;
    if (q <= 2) return {k_sk_resident<2>, g2, o2};
    if (q <= 4) return {k_sk_resident<4>, g4, o4};
    if (q <= 8) return {k_sk_resident<8>, g8, o8};
    if (q <= 16) return {k_sk_resident<16>, g16, o16};
    return {k_sk_resident<32>, g32, o32};
}

using BandKernel = void (*)(const float*, int, int, const int32_t*, const int32_t*, int, float, int, int, int, int,
                            float, float*, float*, float2*, float*, int32_t*, unsigned*, int32_t*, int32_t*, float*,
                            float*, float*, float*);

struct BandChoice {
    BandKernel fn;
    int* granted;
};

BandChoice sk_pick(int n_pad) {
    static int g2[64], g4[64], g8[64], g16[64], g32[64];
    const int q = n_pad / 128;
    if (q <= 2) return {k_sk_band<2>, g2};
    if (q <= 4) return {k_sk_band<4>, g4};
    if (q <= 8) return {k_sk_band<8>, g8};
    if (q <= 16) return {k_sk_band<16>, g16};
    return {k_sk_band<32>, g32};
}

}  // namespace

extern "C" {

size_t fs_sinkhorn_workspace_bytes(int n_pairs, int m_max, int n_max) {
    if (n_pairs <= 0 || m_max <= 0 || n_max <= 0) return 256;
    return sk_carve(nullptr, n_pairs, m_max, n_max).bytes;
}

int fs_sinkhorn_match(const float* scores, int n_pairs, int m_max, int n_max, const int32_t* m_counts,
                      const int32_t* n_counts, float bin_score, int iters, float match_threshold,
                      int pairs_per_group, int32_t* matches0, int32_t* matches1, float* mscores0, float* mscores1,
                      float* u_out, float* v_out, void* workspace, size_t workspace_bytes, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && m_max >= 0 && n_max >= 0 && iters >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(m_counts && n_counts && matches0 && matches1 && mscores0 && mscores1, "null pointer");
    FS_REQUIRE(n_max <= SK_MAX_N, "n_max above 4096");
    if (m_max == 0 || n_max == 0) return FS_OK;  // nothing to write: outputs have zero columns
    FS_REQUIRE(scores && workspace, "null pointer");
    FS_REQUIRE(n_max % 4 == 0, "n_max must be a multiple of 4 (16-byte aligned score rows)");
    FS_REQUIRE((reinterpret_cast<uintptr_t>(scores) & 15) == 0, "scores must be 16-byte aligned");
    const SkGeom g = sk_geom(m_max, n_max);
    const SkWs w = sk_carve(workspace, n_pairs, m_max, n_max);
    if (workspace_bytes < w.bytes) {
        fs::set_error("fs_sinkhorn_match: workspace %zu < %zu", workspace_bytes, w.bytes);
        return FS_ERR_WORKSPACE;
    }
    const BandChoice k = sk_pick(g.n_pad);
    int s = fs::ensure_dyn_smem(k.fn, k.granted, g.smem);
    if (s != FS_OK) return s;

    k_sk_init<<<n_pairs, 256, 0, stream>>>(n_pairs, m_max, n_max, g.n_pad, m_counts, n_counts, bin_score, w.u, w.v,
                                           w.counter, matches0, matches1, mscores0, mscores1);
    FS_LAUNCH_CHECK("k_sk_init");

    // Resident path: one cooperative launch per group of pairs whose CTAs are all co-resident.
    const ResidentChoice rk = sk_pick_resident(g.n_pad);
    s = fs::ensure_dyn_smem(rk.fn, rk.granted, g.smem);
    if (s != FS_OK) return s;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (rk.ctas_per_sm[dev] == 0) {
        int nb = 0;
        FS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, rk.fn, SK_THREADS, g.smem));
        rk.ctas_per_sm[dev] = nb > 0 ? nb : -1;
    }
    const int capacity = (rk.ctas_per_sm[dev] > 0 ? rk.ctas_per_sm[dev] : 0) * fs::sm_count();
    int Gr = g.bands_max > 0 ? capacity / g.bands_max : 0;
    if (pairs_per_group > 0 && pairs_per_group < Gr) Gr = pairs_per_group;
    if (Gr >= 1 && pairs_per_group >= 0) {
        for (int p0 = 0; p0 < n_pairs; p0 += Gr) {
            const int gp = (n_pairs - p0 < Gr) ? (n_pairs - p0) : Gr;
            const dim3 grid(g.bands_max, gp);
            const float* a_scores = scores;
            int a_m = m_max, a_n = n_max, a_p0 = p0, a_R = g.R, a_np = g.n_pad, a_bm = g.bands_max, a_it = iters;
            float a_alpha = bin_score, a_thr = match_threshold;
            SkWs ww = w;
            void* args[] = {&a_scores, &a_m, &a_n, &m_counts, &n_counts, &a_p0, &a_alpha, &a_R, &a_np, &a_bm, &a_it,
                            &a_thr, &ww.u, &ww.v, &ww.colpart, &ww.rb_val, &ww.rb_idx, &ww.i1, &ww.upart, &ww.vpart,
                            &ww.counter,
                            &matches0, &matches1, &mscores0, &mscores1, &u_out, &v_out};
            FS_CUDA(cudaLaunchCooperativeKernel((const void*)rk.fn, grid, dim3(SK_THREADS), args, g.smem, stream));
            FS_LAUNCH_CHECK("k_sk_resident");
        }
        return FS_OK;
    }

    // Streaming fallback (one pair's bands exceed the co-resident capacity): a launch per iteration, the group's
    // matrices kept in L2 between launches.
    int G = pairs_per_group;
    if (G <= 0) {  // matrices of one group stay in L2 across its iterations
        const double per_pair = (double)m_max * n_max * sizeof(float);
        G = (int)(96.0e6 / per_pair);
    }
    if (G < 1) G = 1;
    if (G > n_pairs) G = n_pairs;
    for (int p0 = 0; p0 < n_pairs; p0 += G) {
        const int gp = (n_pairs - p0 < G) ? (n_pairs - p0) : G;
        const dim3 grid(g.bands_max, gp);
        for (int it = 0; it <= iters; ++it) {
            const int final_pass = (it == iters);
            k.fn<<<grid, SK_THREADS, g.smem, stream>>>(scores, m_max, n_max, m_counts, n_counts, p0, bin_score, g.R,
                                                        g.n_pad, g.bands_max, final_pass, match_threshold, w.u, w.v,
                                                        w.colpart, w.rb_val, w.rb_idx, w.counter, matches0, matches1,
                                                        mscores0, mscores1, u_out, v_out);
            FS_LAUNCH_CHECK("k_sk_band");
        }
    }
    return FS_OK;
}

}  // extern "C"
