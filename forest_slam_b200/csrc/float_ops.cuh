// float_ops.cuh — layout of the tensor-core operand bank of the A1 float matcher and the pieces both tensor-core
// matchers (float_tc.cu: split-fp16 3-MMA kernel, float_top3.cu: single-pass kernel with certification) share.
//
// Operand bank (written by fs_float_prepare, read by TMA):  [n_frames][n_max + 1][512] fp16
//   row n < n_max : [hi(256) | lo(256)],  x = hi + lo to 2^-25 (hi = fp16(x), lo = fp16(x - hi)); rows >= count are zero
//   row n_max     : frame statistics, first 16 bytes = FrameStats (the rest of the row is zero)
// Keeping the statistics inside the bank makes the tensor self-describing (clone / slice by frame keeps them).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include "float_common.cuh"

namespace fs {

constexpr int OP_ROW_HALFS = 512;                 // fp16 elements per operand row
constexpr int OP_ROW_BYTES = OP_ROW_HALFS * 2;    // 1 KB
__host__ __device__ inline int op_rows(int n_max) { return n_max + 1; }

// max over the frame's rows of the squared L2 norm of the hi part / of the lo part, as the bits of a non-negative
// float (so that atomicMax on the unsigned pattern is the float max: order independent => deterministic)
struct FrameStats {
    unsigned int hi2_bits, lo2_bits, pad0, pad1;
};

__device__ __forceinline__ const FrameStats* frame_stats(const void* operands, int n_max, int f) {
    return reinterpret_cast<const FrameStats*>(static_cast<const unsigned char*>(operands) +
                                               ((size_t)f * op_rows(n_max) + n_max) * OP_ROW_BYTES);
}

// ---- score keys of the single-pass kernel ------------------------------------------------------------------
// A raw tensor-core score v (|v| <= vmax = Ha * Hb) is mapped to v' = fma(v, sc, KEY_BIAS) in [1, 1.125): the sign, the
// exponent and the top three mantissa bits of v' are then constant, so (bits(v') << 12) keeps all 20 live mantissa bits
// and frees the low 12 bits for (4095 - column): ONE unsigned compare orders by (score, lowest column).
constexpr float KEY_BIAS = 1.0625f;
constexpr float KEY_HALF_RANGE = 0.0619f;  // < 0.0625 so that rounding can never leave [1, 1.125)
constexpr int KEY_MAX_COLS = 4096;

struct PairScale {
    float sc;      // v' = v * sc + KEY_BIAS
    float inv_sc;  // v  = (v' - KEY_BIAS) * inv_sc
    float delta;   // |exact split score - decoded tensor-core score| <= delta for every entry of the pair
};

// Identical in the matcher kernel and in the resolver (both must use the same sc).  Cauchy-Schwarz on the dropped
// terms: |sum lo_a hi_b + hi_a lo_b + lo_a lo_b| <= La Hb + Ha Lb + La Lb; the tensor core accumulates 256 exact fp16
// products in fp32 (16 MMAs of K = 16 = 272 additions, each rounded or truncated to 2^-23 relative): worst case
// 272 * 2^-23 = 3.3e-5 * vmax (derived, not fitted; the observed error is ~4x smaller);
// the key itself rounds v' to 2^-24, i.e. 2^-24 / sc in score units.
__device__ __forceinline__ PairScale pair_scale(const void* operands, int n_max, int fa, int fb) {
    const FrameStats sa = *frame_stats(operands, n_max, fa), sb = *frame_stats(operands, n_max, fb);
    const float slack = 1.0001f;  // the fp32 sums of squares behind the statistics are themselves rounded
    const float Ha = sqrtf(__uint_as_float(sa.hi2_bits)) * slack, La = sqrtf(__uint_as_float(sa.lo2_bits)) * slack;
    const float Hb = sqrtf(__uint_as_float(sb.hi2_bits)) * slack, Lb = sqrtf(__uint_as_float(sb.lo2_bits)) * slack;
    const float vmax = fmaxf(Ha * Hb, 1e-20f);
    PairScale ps;
    ps.sc = KEY_HALF_RANGE / vmax;
    ps.inv_sc = vmax / KEY_HALF_RANGE;
    ps.delta = (La * Hb + Ha * Lb + La * Lb) + 3.3e-5f * vmax + 1.2e-7f * ps.inv_sc;
    return ps;
}

__device__ __forceinline__ float key_score(unsigned int key, const PairScale& ps) {
    if (key == 0u) return -INFINITY;  // empty slot
    const float vp = __uint_as_float(0x3F800000u | (key >> 12));
    return (vp - KEY_BIAS) * ps.inv_sc;
}
__device__ __forceinline__ int key_col(unsigned int key) { return 4095 - (int)(key & 0xFFFu); }

// ---- the EXACT score of the tensor-core paths ----------------------------------------------------------------
// S(r, c) = sum_k a_k b_k with a_k = fp32(hi + lo), one warp per dot product: lane l owns k = 8l .. 8l+7 (one 16-byte
// load of hi and of lo per row), a sequential fma chain over its 8 terms, then a xor-butterfly sum.  Every place that
// needs an exact score (candidate rescoring, full-row scan, both directions) calls THIS function, so equal inputs give
// equal bits; a * b commutes, so S(r, c) is the same number whichever frame is the "row" side.
struct OpRow8 {
    float v[8];
};
__device__ __forceinline__ OpRow8 load_op_row(const __half* row, int lane) {
    const uint4 h = *reinterpret_cast<const uint4*>(row + lane * 8);
    const uint4 l = *reinterpret_cast<const uint4*>(row + 256 + lane * 8);
    const __half2* hh = reinterpret_cast<const __half2*>(&h);
    const __half2* ll = reinterpret_cast<const __half2*>(&l);
    OpRow8 r;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float2 a = __half22float2(hh[i]), b = __half22float2(ll[i]);
        r.v[2 * i] = __fadd_rn(a.x, b.x);
        r.v[2 * i + 1] = __fadd_rn(a.y, b.y);
    }
    return r;
}
__device__ __forceinline__ float warp_dot(const OpRow8& a, const OpRow8& b) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s = __fmaf_rn(a.v[i], b.v[i], s);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) s = __fadd_rn(s, __shfl_xor_sync(0xffffffffu, s, o));
    return s;
}

// TMA descriptor of the operand bank: dims {512, n_max + 1, n_frames} fp16, box {64, box_rows, 1}, 128-byte swizzle
int encode_operand_map(const void* operands, int n_frames, int n_max, CUtensorMap* out, int box_rows = 128);

// k_float_top3_tc (float_top3.cu) over both directions of every pair: keys [2][n_pairs][n_max] uint4 =
// (three largest 32-column-chunk maxima, largest chunk runner-up), each key = score bits << 12 | (4095 - column)
int launch_float_top3(const CUtensorMap& tmap, const CUtensorMap& tmap_half, const void* operands, int n_max,
                      const int32_t* counts, const int32_t* pair_a, const int32_t* pair_b, int n_pairs, uint4* keys,
                      cudaStream_t stream, int probe = 0, long long* trace = nullptr, uint32_t* colraw = nullptr,
                      bool hamming = false, int ndirs = 2);
// single-pass forms (colraw != nullptr): direction 0 only, slab results of the column side into
// colraw [n_pairs][2 * ceil(n_max / 32)][n_max rounded up to 128] u32 (float_top3.cu); element count:
size_t float_top3_colraw_elems(int n_pairs, int n_max);

}  // namespace fs
