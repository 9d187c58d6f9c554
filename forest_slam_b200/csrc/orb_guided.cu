// orb_guided.cu — guided (windowed / row-band) ORB matching, SURVEY 8f rank 4 second half and rank 1 (sparse L<->R):
//   ORBmatcher::SearchForInitialization   ros_ws/src/ORB_SLAM3/src/ORBmatcher.cc:648-762   -> fs_orb_search_for_initialization
//   Frame::ComputeStereoMatches           ros_ws/src/ORB_SLAM3/src/Frame.cc:811-981        -> fs_orb_stereo_matches
// Both are restated operation by operation (oracle: oracle/orb_guided.c, the same reference lines in C).
//
// SearchForInitialization is a SEQUENTIAL greedy assignment: query i1's candidates are filtered by the distances earlier
// queries matched with (vMatchedDistance) and a later query may steal an earlier one's partner, so the outer loop over
// i1 cannot be reordered.  One warp owns one frame pair and walks the queries in order; the parallelism is over the
// frame pairs of the batch and, inside a query, over the grid cells of its search window (one cell per lane, cells in
// the reference's (ix, iy) order; the packed key (distance, cell order, position in cell) makes "first minimum in visiting
// order" a plain warp minimum).  The candidate set of a query is a 2*window square out of the 64 x 48 frame grid
// (Frame::GetFeaturesInArea), ~140 of 2000 keypoints at window 100: evaluating only those with __popc is ~15x less work
// than the dense 2000 x 2000 tensor-core pass, so the dense kernel is not used here.
//
// ComputeStereoMatches: one warp per left keypoint scans the right keypoints whose row band covers its row (ascending
// index, so the first minimum wins as in the reference), then the 11 x 11 SAD sliding window (+-5) on the image pair and
// the parabola fit, in the reference's float arithmetic; a second kernel applies the median test on the SAD distances.
#include <cuda_runtime.h>
#include <limits.h>
#include "common.cuh"
#include "../../include/forestslam_b200.h"

namespace {

constexpr int GRID_COLS = 64, GRID_ROWS = 48, GRID_CELLS = GRID_COLS * GRID_ROWS;  // include/Frame.h:44-45
constexpr int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;                       // ORBmatcher.cc:35-37
constexpr int ORB_MAX_KPTS = 4096;

__device__ __forceinline__ int desc_distance(const uint4 a0, const uint4 a1, const uint4* __restrict__ b) {
    const uint4 b0 = b[0], b1 = b[1];
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

struct OrbInitParams {
    const float* kpts;      // [n_frames, n_max, 2]
    const float* angle;     // [n_frames, n_max] degrees
    const int32_t* octave;  // [n_frames, n_max]
    const uint8_t* desc;    // [n_frames, n_max, 32]
    const int32_t* counts;  // [n_frames]
    const int32_t* pair_a;
    const int32_t* pair_b;
    float* prev_matched;    // [n_pairs, n_max, 2] in/out
    int32_t* matches12;     // [n_pairs, n_max]
    int32_t* n_matches;     // [n_pairs]
    int n_max, window, check_orientation;
    float nn_ratio, minx, miny, winv, hinv;
};

// one warp = one frame pair.  Everything the sequential loop touches lives in shared memory: the second frame's
// keypoints, octaves and descriptors (staged once: 42 bytes per keypoint), its grid (CSR) and the match state — the loop
// is a chain of dependent look-ups (cell -> index -> position -> descriptor) and from global memory each link costs a
// DRAM/L2 round trip (measured: 66 ms per 256 pairs from global memory).
__global__ void __launch_bounds__(32)
k_orb_search_init(OrbInitParams P) {
    extern __shared__ __align__(16) unsigned char orb_smem[];
    __shared__ unsigned short s_start[GRID_CELLS + 1];
    __shared__ unsigned short s_fill[GRID_CELLS];
    __shared__ int s_hist[HISTO_LENGTH];
    const int NM = P.n_max;
    uint4* s_desc = reinterpret_cast<uint4*>(orb_smem);                                  // [NM][2]   descriptors of frame 2
    float2* s_kp2 = reinterpret_cast<float2*>(orb_smem + (size_t)NM * 32);               // [NM]
    unsigned short* s_idx = reinterpret_cast<unsigned short*>(orb_smem + (size_t)NM * 40);   // mGrid cell lists (CSR)
    unsigned short* s_mdist = s_idx + NM;                                                // vMatchedDistance (0xFFFF = INT_MAX)
    short* s_m21 = reinterpret_cast<short*>(s_mdist + NM);                               // vnMatches21
    unsigned short* s_push = reinterpret_cast<unsigned short*>(s_m21 + NM);              // partner of query i1 when it was
    //                                                        pushed into rotHist (0xFFFF: never); the bin is computed after the loop
    unsigned char* s_oct2 = reinterpret_cast<unsigned char*>(s_push + NM);               // octave of frame 2 (clamped to 255)
    const int p = blockIdx.x, lane = threadIdx.x;
    const unsigned full = 0xffffffffu;
    const int fa = P.pair_a[p], fb = P.pair_b[p];
    const int n1 = min(P.counts[fa], P.n_max), n2 = min(P.counts[fb], P.n_max);
    const float2* kp2 = s_kp2;
    const int32_t* oct1 = P.octave + (size_t)fa * P.n_max;
    const uint4* d1 = reinterpret_cast<const uint4*>(P.desc) + (size_t)fa * P.n_max * 2;
    const uint4* d2 = s_desc;
    float2* prev = reinterpret_cast<float2*>(P.prev_matched) + (size_t)p * P.n_max;
    int32_t* m12 = P.matches12 + (size_t)p * P.n_max;
    {
        const float2* g_kp2 = reinterpret_cast<const float2*>(P.kpts) + (size_t)fb * P.n_max;
        const int32_t* g_oct2 = P.octave + (size_t)fb * P.n_max;
        const uint4* g_d2 = reinterpret_cast<const uint4*>(P.desc) + (size_t)fb * P.n_max * 2;
        for (int i = lane; i < n2; i += 32) {
            s_kp2[i] = g_kp2[i];
            s_oct2[i] = (unsigned char)min(max(g_oct2[i], 0), 255);
        }
        for (int i = lane; i < 2 * n2; i += 32) s_desc[i] = g_d2[i];
    }

    // ---- Frame::AssignFeaturesToGrid (Frame.cc:385-416) for the second frame: stable counting sort by cell
    for (int c = lane; c <= GRID_CELLS; c += 32) s_start[c] = 0;
    for (int c = lane; c < GRID_CELLS; c += 32) s_fill[c] = 0;
    for (int i = lane; i < P.n_max; i += 32) {
        s_mdist[i] = 0xFFFF;
        s_m21[i] = -1;
        s_push[i] = 0xFFFF;
        if (i < n1) m12[i] = -1;
    }
    __syncwarp();
    if (lane < HISTO_LENGTH) s_hist[lane] = 0;
    __syncwarp();
    auto cell_of = [&](int i) {
        const float2 k = kp2[i];
        const int px = (int)roundf((k.x - P.minx) * P.winv), py = (int)roundf((k.y - P.miny) * P.hinv);  // PosInGrid :727-728
        return (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) ? -1 : px * GRID_ROWS + py;
    };
    for (int i0 = 0; i0 < n2; i0 += 32) {
        const int i = i0 + lane;
        const int c = i < n2 ? cell_of(i) : -1;
        // lanes of one cell count themselves once: the leader (lowest lane of the group) adds the group's size
        const unsigned grp = __match_any_sync(full, c);
        if (c >= 0 && (grp & ((1u << lane) - 1)) == 0) s_start[c + 1] += (unsigned short)__popc(grp);
        __syncwarp();
    }
    if (lane == 0) {
        unsigned acc = 0;
        for (int c = 0; c <= GRID_CELLS; c++) {
            acc += s_start[c];
            s_start[c] = (unsigned short)acc;
        }
    }
    __syncwarp();
    for (int i0 = 0; i0 < n2; i0 += 32) {
        const int i = i0 + lane;
        const int c = i < n2 ? cell_of(i) : -1;
        const unsigned grp = __match_any_sync(full, c);
        const int rank = __popc(grp & ((1u << lane) - 1));
        if (c >= 0) s_idx[s_start[c] + s_fill[c] + rank] = (unsigned short)i;
        __syncwarp();
        if (c >= 0 && rank == 0) s_fill[c] += (unsigned short)__popc(grp);
        __syncwarp();
    }

    // ---- the sequential loop of ORBmatcher.cc:663-729.  The next query's inputs (search centre, octave, descriptor) are
    // fetched from global memory one iteration ahead; the rotation histogram is filled after the loop (its counts do not
    // depend on the order of the pushes), so that no global round trip sits on the loop's dependency chain.
    const float r = (float)P.window;
    const float factor = 1.0f / HISTO_LENGTH;  // :656, the reference's factor
    float2 nq = make_float2(0.f, 0.f);
    int noct = 1;
    uint4 na0 = make_uint4(0, 0, 0, 0), na1 = na0;
    if (n1 > 0) {
        nq = prev[0];
        noct = oct1[0];
        na0 = d1[0];
        na1 = d1[1];
    }
    for (int i1 = 0; i1 < n1; i1++) {
        const float2 q = nq;   // vbPrevMatched[i1]
        const int level1 = noct;
        const uint4 a0 = na0, a1 = na1;
        if (i1 + 1 < n1) {
            nq = prev[i1 + 1];
            noct = oct1[i1 + 1];
            na0 = d1[2 * i1 + 2];
            na1 = d1[2 * i1 + 3];
        }
        if (level1 > 0) continue;  // :666
        // Frame::GetFeaturesInArea(x, y, windowSize, 0, 0)   Frame.cc:657-723
        int cx0 = (int)floorf((q.x - P.minx - r) * P.winv);
        cx0 = max(cx0, 0);
        if (cx0 >= GRID_COLS) continue;
        int cx1 = min(GRID_COLS - 1, (int)ceilf((q.x - P.minx + r) * P.winv));
        if (cx1 < 0) continue;
        int cy0 = max(0, (int)floorf((q.y - P.miny - r) * P.hinv));
        if (cy0 >= GRID_ROWS) continue;
        int cy1 = min(GRID_ROWS - 1, (int)ceilf((q.y - P.miny + r) * P.hinv));
        if (cy1 < 0) continue;
        // The cells (ix, cy0 .. cy1) of one grid column are consecutive in the CSR (cell = ix * ROWS + iy), so a column of the
        // window is ONE contiguous candidate range, visited in the reference's (iy, insertion) order.  One column per
        // lane; the columns' candidates are then dealt evenly over the lanes (a cell on a trunk or on the horizon holds
        // dozens of keypoints, most none).  key = dist << 20 | column << 12 | position in the column's range: unique,
        // so the two smallest keys are (bestDist with the FIRST index reaching it, bestDist2) of the strict-less scan.
        const int ncx = cx1 - cx0 + 1;
        unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;
        int idx = -1;
        for (int c0 = 0; c0 < ncx; c0 += 32) {
            const int k = c0 + lane;
            int cs = 0, cn = 0;
            if (k < ncx) {
                const int cell0 = (cx0 + k) * GRID_ROWS + cy0;
                cs = s_start[cell0];
                cn = s_start[cell0 + (cy1 - cy0) + 1] - cs;
            }
            int incl = cn;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(full, incl, o);
                if (lane >= o) incl += v;
            }
            const int total = __shfl_sync(full, incl, 31);
            for (int t0 = 0; t0 < total; t0 += 32) {
                const int t = t0 + lane;
                // owner = first lane whose inclusive count exceeds t (binary search over the warp's scan)
                int lo = 0;
#pragma unroll
                for (int step = 16; step >= 1; step >>= 1) {
                    const int v = __shfl_sync(full, incl, lo + step - 1);
                    if (v <= t) lo += step;
                }
                const int owner = min(lo, 31);
                const int o_incl = __shfl_sync(full, incl, owner), o_cn = __shfl_sync(full, cn, owner);
                const int o_cs = __shfl_sync(full, cs, owner);
                if (t < total) {
                    const int pos = t - (o_incl - o_cn);
                    const int i2 = s_idx[o_cs + pos];
                    if (s_oct2[i2] == 0) {  // kpUn.octave < minLevel or > maxLevel, minLevel = maxLevel = 0
                        const float2 tp = kp2[i2];
                        if (fabsf(tp.x - q.x) < r && fabsf(tp.y - q.y) < r) {
                            const int dist = desc_distance(a0, a1, d2 + 2 * i2);
                            if (!((int)s_mdist[i2] <= dist)) {  // vMatchedDistance[i2] <= dist -> skip (0xFFFF plays INT_MAX)
                                const unsigned key = ((unsigned)dist << 20) | ((unsigned)(c0 + owner) << 12) | (unsigned)pos;
                                if (key < k1) {
                                    k2 = k1;
                                    k1 = key;
                                    idx = i2;
                                } else if (key < k2) {
                                    k2 = key;
                                }
                            }
                        }
                    }
                }
            }
        }
        // warp top-2 of the unique keys
        const unsigned w1 = __reduce_min_sync(full, k1);
        const unsigned mine2 = (k1 == w1) ? k2 : k1;
        const unsigned w2 = __reduce_min_sync(full, mine2);
        const int src = __ffs(__ballot_sync(full, k1 == w1 && w1 != 0xFFFFFFFFu)) - 1;
        const int bestIdx2 = src >= 0 ? __shfl_sync(full, idx, src) : -1;
        const int bestDist = w1 == 0xFFFFFFFFu ? INT_MAX : (int)(w1 >> 20);
        const int bestDist2 = w2 == 0xFFFFFFFFu ? INT_MAX : (int)(w2 >> 20);
        if (bestDist <= TH_LOW && (float)bestDist < (float)bestDist2 * P.nn_ratio) {  // :706-708
            if (lane == 0) {
                const int old = s_m21[bestIdx2];
                if (old >= 0) m12[old] = -1;
                m12[i1] = bestIdx2;
                s_m21[bestIdx2] = (short)i1;
                s_mdist[bestIdx2] = (unsigned short)bestDist;
                s_push[i1] = (unsigned short)bestIdx2;   // rotHist[bin].push_back(i1), bin from this partner (:719-729)
            }
            __syncwarp();
        }
    }
    __syncwarp();
    // ---- rotation consistency (:731-755) with ORBmatcher::ComputeThreeMaxima (:2012-2053)
    if (P.check_orientation) {
        const float* ang1 = P.angle + (size_t)fa * P.n_max;
        const float* ang2 = P.angle + (size_t)fb * P.n_max;
        for (int i = lane; i < n1; i += 32) {
            const int partner = s_push[i];
            if (partner == 0xFFFF) continue;
            float rot = ang1[i] - ang2[partner];
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * factor);
            if (bin == HISTO_LENGTH) bin = 0;
            atomicAdd(&s_hist[bin], 1);
            s_push[i] = (unsigned short)bin;   // from here on: the bin of query i
        }
        __syncwarp();
        int ind1 = -1, ind2 = -1, ind3 = -1;
        int max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = s_hist[i];
            if (s > max1) {
                max3 = max2; max2 = max1; max1 = s;
                ind3 = ind2; ind2 = ind1; ind1 = i;
            } else if (s > max2) {
                max3 = max2; max2 = s;
                ind3 = ind2; ind2 = i;
            } else if (s > max3) {
                max3 = s;
                ind3 = i;
            }
        }
        if ((float)max2 < 0.1f * (float)max1) {
            ind2 = -1;
            ind3 = -1;
        } else if ((float)max3 < 0.1f * (float)max1) {
            ind3 = -1;
        }
        for (int i = lane; i < n1; i += 32) {
            const int b = s_push[i];
            if (b != 0xFFFF && b != ind1 && b != ind2 && b != ind3) m12[i] = -1;
        }
        __syncwarp();
    }
    // ---- update vbPrevMatched (:757-760), count
    int cnt = 0;
    for (int i = lane; i < n1; i += 32) {
        const int m = m12[i];
        if (m >= 0) {
            prev[i] = kp2[m];
            cnt++;
        }
    }
    cnt = __reduce_add_sync(full, cnt);
    if (lane == 0) P.n_matches[p] = cnt;  // = the reference's running nmatches
}


// ------------------------------------------------------------------------------------------------ ComputeStereoMatches
struct OrbStereoParams {
    const float* kpL;       // [n_frames, n_max, 2]
    const int32_t* octL;    // [n_frames, n_max]
    const uint8_t* descL;   // [n_frames, n_max, 32]
    const int32_t* countL;  // [n_frames]
    const float* kpR;
    const int32_t* octR;
    const uint8_t* descR;
    const int32_t* countR;
    const uint8_t* imL;     // [n_frames, rows, cols] u8 (pyramid level 0)
    const uint8_t* imR;
    float* u_right;         // [n_frames, n_max]   mvuRight (-1 = none)
    float* depth;           // [n_frames, n_max]   mvDepth  (-1 = none)
    int32_t* sad;           // [n_frames, n_max]   correlation distance kept for the median test (-1 = none)
    int n_max, rows, cols;
    float mb, mbf;
};

constexpr int STEREO_WARPS = 8;

// one warp = one left keypoint (Frame.cc:845-962)
__global__ void __launch_bounds__(STEREO_WARPS * 32)
k_orb_stereo(OrbStereoParams P) {
    const int f = blockIdx.y, lane = threadIdx.x & 31;
    const int iL = blockIdx.x * STEREO_WARPS + (threadIdx.x >> 5);
    const int n = min(P.countL[f], P.n_max), nr = min(P.countR[f], P.n_max);
    if (iL >= n) return;
    const unsigned full = 0xffffffffu;
    const size_t o = (size_t)f * P.n_max;
    const float2* kpL = reinterpret_cast<const float2*>(P.kpL) + o;
    const float2* kpR = reinterpret_cast<const float2*>(P.kpR) + o;
    const uint4* dLp = reinterpret_cast<const uint4*>(P.descL) + (o + iL) * 2;
    const uint4* dR = reinterpret_cast<const uint4*>(P.descR) + o * 2;
    if (lane == 0) {
        P.u_right[o + iL] = -1.0f;
        P.depth[o + iL] = -1.0f;
        P.sad[o + iL] = -1;
    }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const float minZ = P.mb, minD = 0.0f, maxD = P.mbf / minZ;
    const float2 kl = kpL[iL];
    const int levelL = P.octL[o + iL];
    const float uL = kl.x, vL = kl.y;
    const int row = (int)vL;  // vRowIndices[vL]
    const float minU = uL - maxD, maxU = uL - minD;
    if (maxU < 0) return;
    const uint4 a0 = dLp[0], a1 = dLp[1];
    // candidates in ascending iR; key = dist << 16 | iR keeps the first minimum (strict '<' of :889)
    unsigned best = ((unsigned)TH_HIGH << 16);  // only dist < TH_HIGH replaces it
    for (int iR = lane; iR < nr; iR += 32) {
        const float2 kr = kpR[iR];
        const float rr = 2.0f * 1.0f;  // 2 * mvScaleFactors[octave], level 0
        const int maxr = (int)ceilf(kr.y + rr), minr = (int)floorf(kr.y - rr);
        if (row < minr || row > maxr) continue;
        const int oR = P.octR[o + iR];
        if (oR < levelL - 1 || oR > levelL + 1) continue;
        if (kr.x >= minU && kr.x <= maxU) {
            const int dist = desc_distance(a0, a1, dR + 2 * iR);
            if (dist < TH_HIGH) best = min(best, ((unsigned)dist << 16) | (unsigned)iR);
        }
    }
    best = __reduce_min_sync(full, best);
    const int bestDist = (int)(best >> 16);
    if (!(bestDist < thOrbDist)) return;
    const int bestIdxR = (int)(best & 0xFFFFu);
    // ---- subpixel match by correlation (:899-961), pyramid level 0 (scale factor 1)
    const float uR0 = kpR[bestIdxR].x;
    const float scaleduL = roundf(uL), scaledvL = roundf(vL), scaleduR0 = roundf(uR0);
    const int w = 5, L = 5;
    const int cy = (int)scaledvL, cxl = (int)scaleduL, cxr = (int)scaleduR0;
    // a patch leaving the image makes the reference's cv::Mat::rowRange / colRange throw: no match here
    if (cy - w < 0 || cy + w + 1 > P.rows || cxl - w < 0 || cxl + w + 1 > P.cols) return;
    const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
    if (iniu < 0 || endu >= (float)P.cols) return;
    if (cxr - L - w < 0) return;
    const uint8_t* IL = P.imL + (size_t)f * P.rows * P.cols;
    const uint8_t* IR = P.imR + (size_t)f * P.rows * P.cols;
    int s[11];
#pragma unroll
    for (int k = 0; k < 11; k++) s[k] = 0;
    for (int pix = lane; pix < 121; pix += 32) {
        const int dy = pix / 11 - w, dx = pix % 11 - w;
        const int a = IL[(size_t)(cy + dy) * P.cols + cxl + dx];
        const uint8_t* rp = IR + (size_t)(cy + dy) * P.cols + cxr + dx - L;
#pragma unroll
        for (int k = 0; k < 11; k++) s[k] += abs(a - (int)rp[k]);
    }
#pragma unroll
    for (int k = 0; k < 11; k++) s[k] = __reduce_add_sync(full, s[k]);
    if (lane != 0) return;
    int bestS = INT_MAX, bestinc = 0;
#pragma unroll
    for (int k = 0; k < 11; k++) {
        if ((float)s[k] < (float)bestS) {  // float dist < int bestDist
            bestS = s[k];
            bestinc = k - L;
        }
    }
    if (bestinc == -L || bestinc == L) return;
    float dist1 = 0.f, dist2 = 0.f, dist3 = 0.f;
#pragma unroll
    for (int k = 1; k < 10; k++)
        if (k == L + bestinc) {
            dist1 = (float)s[k - 1];
            dist2 = (float)s[k];
            dist3 = (float)s[k + 1];
        }
    const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
    if (deltaR < -1 || deltaR > 1) return;
    float bestuR = 1.0f * (scaleduR0 + (float)bestinc + deltaR);
    float disparity = uL - bestuR;
    if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) {
            disparity = 0.01;
            bestuR = (float)((double)uL - 0.01);
        }
        P.depth[o + iL] = P.mbf / disparity;
        P.u_right[o + iL] = bestuR;
        P.sad[o + iL] = bestS;
    }
}

// the median test of :964-980: matches whose correlation distance is >= 1.5 * 1.4 * median are dropped.  One CTA per
// frame; the median (element size/2 of the ascending distances) comes from a counting pass over the distance values.
__global__ void __launch_bounds__(256)
k_orb_stereo_median(OrbStereoParams P) {
    __shared__ int s_hist[256];
    __shared__ int s_sel[2];
    const int f = blockIdx.x;
    const int n = min(P.countL[f], P.n_max);
    const size_t o = (size_t)f * P.n_max;
    // SAD of an 11 x 11 u8 patch <= 121 * 255 = 30855 < 2^15: two-level radix select (high 8 bits, then low 7 bits)
    int nd = 0;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int v = P.sad[o + i];
        if (v >= 0) {
            atomicAdd(&s_hist[v >> 7], 1);
            nd++;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int tot = 0;
        for (int b = 0; b < 256; b++) tot += s_hist[b];
        int k = tot / 2, b = 0;   // vDistIdx[size / 2]
        s_sel[0] = -1;
        if (tot > 0) {
            while (k >= s_hist[b]) k -= s_hist[b++];
            s_sel[0] = b;
        }
        s_sel[1] = k;
    }
    __syncthreads();
    const int hb = s_sel[0], k = s_sel[1];
    if (hb < 0) return;
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int v = P.sad[o + i];
        if (v >= 0 && (v >> 7) == hb) atomicAdd(&s_hist[v & 127], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int kk = k, b = 0;
        while (kk >= s_hist[b]) kk -= s_hist[b++];
        s_sel[0] = (hb << 7) | b;
    }
    __syncthreads();
    const float median = (float)s_sel[0];
    const float thDist = 1.5f * 1.4f * median;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int v = P.sad[o + i];
        if (v >= 0 && !((float)v < thDist)) {
            P.u_right[o + i] = -1.0f;
            P.depth[o + i] = -1.0f;
        }
    }
}

}  // namespace

extern "C" int fs_orb_search_for_initialization(const float* kpts, const float* angle, const int32_t* octave,
                                                const uint8_t* desc, int n_frames, int n_max, const int32_t* counts,
                                                const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                                                float* prev_matched, int window_size, float nn_ratio,
                                                int check_orientation, float min_x, float max_x, float min_y,
                                                float max_y, int32_t* matches12, int32_t* n_matches,
                                                fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_frames >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(kpts && octave && desc && counts && pair_a && pair_b && prev_matched && matches12 && n_matches,
               "null pointer");
    FS_REQUIRE(angle || !check_orientation, "check_orientation needs the keypoint angles");
    FS_REQUIRE(max_x > min_x && max_y > min_y && window_size > 0, "bad image bounds / window");
    if (n_max > ORB_MAX_KPTS) {
        fs::set_error("fs_orb_search_for_initialization: n_max %d > %d", n_max, ORB_MAX_KPTS);
        return FS_ERR_UNSUPPORTED;
    }
    OrbInitParams P;
    P.kpts = kpts; P.angle = angle; P.octave = octave; P.desc = desc; P.counts = counts;
    P.pair_a = pair_a; P.pair_b = pair_b; P.prev_matched = prev_matched; P.matches12 = matches12;
    P.n_matches = n_matches; P.n_max = n_max; P.window = window_size; P.check_orientation = check_orientation;
    P.nn_ratio = nn_ratio; P.minx = min_x; P.miny = min_y;
    P.winv = (float)GRID_COLS / (max_x - min_x);   // Frame.cc:159-160
    P.hinv = (float)GRID_ROWS / (max_y - min_y);
    // frame 2 staged in shared memory: 32 B descriptor + 8 B position + 2+2+2+2+1 B of state per keypoint
    const size_t smem = (size_t)n_max * 49 + 64;
    static int granted[64];
    const int st_ = fs::ensure_dyn_smem(k_orb_search_init, granted, smem);
    if (st_ != FS_OK) return st_;
    k_orb_search_init<<<n_pairs, 32, smem, (cudaStream_t)stream>>>(P);
    FS_LAUNCH_CHECK("k_orb_search_init");
    return FS_OK;
}

extern "C" int fs_orb_stereo_matches(const float* kpts_l, const int32_t* octave_l, const uint8_t* desc_l,
                                     const int32_t* counts_l, const float* kpts_r, const int32_t* octave_r,
                                     const uint8_t* desc_r, const int32_t* counts_r, int n_frames, int n_max,
                                     const uint8_t* image_l, const uint8_t* image_r, int rows, int cols, float mb,
                                     float mbf, float* u_right, float* depth, int32_t* sad, fs_stream_t stream) {
    FS_REQUIRE(n_frames >= 0 && n_max >= 0 && rows > 0 && cols > 0, "bad size");
    if (n_frames == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(kpts_l && octave_l && desc_l && counts_l && kpts_r && octave_r && desc_r && counts_r && image_l &&
                   image_r && u_right && depth && sad,
               "null pointer");
    FS_REQUIRE(mb > 0 && mbf > 0, "baseline / fx * baseline must be positive");
    if (n_max > 65535) {
        fs::set_error("fs_orb_stereo_matches: n_max %d > 65535", n_max);
        return FS_ERR_UNSUPPORTED;
    }
    OrbStereoParams P;
    P.kpL = kpts_l; P.octL = octave_l; P.descL = desc_l; P.countL = counts_l;
    P.kpR = kpts_r; P.octR = octave_r; P.descR = desc_r; P.countR = counts_r;
    P.imL = image_l; P.imR = image_r; P.u_right = u_right; P.depth = depth; P.sad = sad;
    P.n_max = n_max; P.rows = rows; P.cols = cols; P.mb = mb; P.mbf = mbf;
    dim3 grid((n_max + STEREO_WARPS - 1) / STEREO_WARPS, n_frames);
    k_orb_stereo<<<grid, STEREO_WARPS * 32, 0, (cudaStream_t)stream>>>(P);
    FS_LAUNCH_CHECK("k_orb_stereo");
    k_orb_stereo_median<<<n_frames, 256, 0, (cudaStream_t)stream>>>(P);
    FS_LAUNCH_CHECK("k_orb_stereo_median");
    return FS_OK;
}
