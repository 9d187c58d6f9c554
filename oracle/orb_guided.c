/* orb_guided.c — TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's guided ORB matchers.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load this; the product path never does.
 * ORB-SLAM3 itself cannot be compiled here (no Eigen / OpenCV C++ headers / Pangolin / boost), so these are
 * line-by-line restatements of
 *   ORBmatcher::SearchForInitialization   ros_ws/src/ORB_SLAM3/src/ORBmatcher.cc:648-762
 *   ORBmatcher::ComputeThreeMaxima        ros_ws/src/ORB_SLAM3/src/ORBmatcher.cc:2012-2053
 *   ORBmatcher::DescriptorDistance        ros_ws/src/ORB_SLAM3/src/ORBmatcher.cc:2058-2074
 *   Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea   src/Frame.cc:385-416, 725-735, 657-723
 *   Frame::ComputeStereoMatches           ros_ws/src/ORB_SLAM3/src/Frame.cc:811-981
 * The reference holds no vectors for them: parity unpinned w.r.t. reference-held fixtures.
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (see __graft_entry__._build_oracle). */
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define FRAME_GRID_ROWS 48 /* include/Frame.h:44 */
#define FRAME_GRID_COLS 64 /* include/Frame.h:45 */
#define TH_HIGH 100        /* ORBmatcher.cc:35 */
#define TH_LOW 50          /* ORBmatcher.cc:36 */
#define HISTO_LENGTH 30    /* ORBmatcher.cc:37 */

/* ORBmatcher.cc:2058-2074 (the SWAR popcount, kept literally) */
static int descriptor_distance(const uint8_t* a, const uint8_t* b) {
    const int32_t* pa = (const int32_t*)a;
    const int32_t* pb = (const int32_t*)b;
    int dist = 0;
    for (int i = 0; i < 8; i++, pa++, pb++) {
        unsigned int v = *pa ^ *pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}
int orb_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

/* ORBmatcher.cc:2012-2053 */
static void compute_three_maxima(const int* histo_size, int L, int* ind1, int* ind2, int* ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = histo_size[i];
        if (s > max1) {
            max3 = max2; max2 = max1; max1 = s;
            *ind3 = *ind2; *ind2 = *ind1; *ind1 = i;
        } else if (s > max2) {
            max3 = max2; max2 = s;
            *ind3 = *ind2; *ind2 = i;
        } else if (s > max3) {
            max3 = s;
            *ind3 = i;
        }
    }
    if (max2 < 0.1f * (float)max1) {
        *ind2 = -1;
        *ind3 = -1;
    } else if (max3 < 0.1f * (float)max1) {
        *ind3 = -1;
    }
}

/* The frame grid of Frame.cc:385-416 (mGrid[ix][iy] = indices in insertion order) as CSR: cell = ix * ROWS + iy */
typedef struct {
    int start[FRAME_GRID_COLS * FRAME_GRID_ROWS + 1];
    int* idx;
} Grid;

static void build_grid(Grid* g, int n, const float* kp, float minx, float miny, float winv, float hinv) {
    const int nc = FRAME_GRID_COLS * FRAME_GRID_ROWS;
    int* cell = (int*)malloc(sizeof(int) * (n > 0 ? n : 1));
    memset(g->start, 0, sizeof(g->start));
    for (int i = 0; i < n; i++) {
        /* Frame.cc:727-728 PosInGrid: round() of a float expression, assigned to int */
        const int px = (int)roundf((kp[2 * i] - minx) * winv);
        const int py = (int)roundf((kp[2 * i + 1] - miny) * hinv);
        cell[i] = (px < 0 || px >= FRAME_GRID_COLS || py < 0 || py >= FRAME_GRID_ROWS) ? -1 : px * FRAME_GRID_ROWS + py;
        if (cell[i] >= 0) g->start[cell[i] + 1]++;
    }
    for (int c = 0; c < nc; c++) g->start[c + 1] += g->start[c];
    g->idx = (int*)malloc(sizeof(int) * (n > 0 ? n : 1));
    int* fill = (int*)calloc(nc, sizeof(int));
    for (int i = 0; i < n; i++)
        if (cell[i] >= 0) g->idx[g->start[cell[i]] + fill[cell[i]]++] = i;
    free(fill);
    free(cell);
}

/* ORBmatcher.cc:648-762.  kp = undistorted keypoints (x, y); angle in degrees; octave; desc 32 bytes per keypoint.
 * prev_matched [n1][2] in/out (vbPrevMatched); matches12 [n1] out.  Image bounds mnMinX.. as Frame.cc:151-160.
 * Returns nmatches. */
int orb_search_for_initialization(int n1, const float* kp1, const float* angle1, const int* octave1, const uint8_t* desc1,
                                  int n2, const float* kp2, const float* angle2, const int* octave2, const uint8_t* desc2,
                                  float* prev_matched, int window_size, float nn_ratio, int check_orientation,
                                  float minx, float maxx, float miny, float maxy, int* matches12) {
    int nmatches = 0;
    const float winv = (float)FRAME_GRID_COLS / (maxx - minx); /* Frame.cc:159-160 */
    const float hinv = (float)FRAME_GRID_ROWS / (maxy - miny);
    Grid g;
    build_grid(&g, n2, kp2, minx, miny, winv, hinv);
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    int rot_size[HISTO_LENGTH];
    memset(rot_size, 0, sizeof(rot_size));
    int* rot_bin = (int*)malloc(sizeof(int) * (n1 > 0 ? n1 : 1)); /* rotHist[bin].push_back(i1): each i1 at most once */
    for (int i = 0; i < n1; i++) rot_bin[i] = -1;
    const float factor = 1.0f / HISTO_LENGTH; /* :656 (the reference's factor, kept as it is) */
    int* matched_distance = (int*)malloc(sizeof(int) * (n2 > 0 ? n2 : 1));
    int* matches21 = (int*)malloc(sizeof(int) * (n2 > 0 ? n2 : 1));
    for (int i = 0; i < n2; i++) {
        matched_distance[i] = INT_MAX;
        matches21[i] = -1;
    }
    for (int i1 = 0; i1 < n1; i1++) {
        const int level1 = octave1[i1];
        if (level1 > 0) continue; /* :666-667 */
        /* GetFeaturesInArea(x, y, windowSize, level1, level1)   Frame.cc:657-723 */
        const float x = prev_matched[2 * i1], y = prev_matched[2 * i1 + 1];
        const float r = (float)window_size;
        int nMinCellX = (int)floorf((x - minx - r) * winv);
        if (nMinCellX < 0) nMinCellX = 0;
        if (nMinCellX >= FRAME_GRID_COLS) continue;
        int nMaxCellX = (int)ceilf((x - minx + r) * winv);
        if (nMaxCellX > FRAME_GRID_COLS - 1) nMaxCellX = FRAME_GRID_COLS - 1;
        if (nMaxCellX < 0) continue;
        int nMinCellY = (int)floorf((y - miny - r) * hinv);
        if (nMinCellY < 0) nMinCellY = 0;
        if (nMinCellY >= FRAME_GRID_ROWS) continue;
        int nMaxCellY = (int)ceilf((y - miny + r) * hinv);
        if (nMaxCellY > FRAME_GRID_ROWS - 1) nMaxCellY = FRAME_GRID_ROWS - 1;
        if (nMaxCellY < 0) continue;
        /* bCheckLevels = (minLevel > 0) || (maxLevel >= 0) = true for level1 == 0 */
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        const uint8_t* d1 = desc1 + 32 * (size_t)i1;
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const int c = ix * FRAME_GRID_ROWS + iy;
                for (int j = g.start[c]; j < g.start[c + 1]; j++) {
                    const int i2 = g.idx[j];
                    if (octave2[i2] < level1) continue;
                    if (octave2[i2] > level1) continue;
                    const float distx = kp2[2 * i2] - x, disty = kp2[2 * i2 + 1] - y;
                    if (!(fabsf(distx) < r && fabsf(disty) < r)) continue;
                    /* :684-704 */
                    const int dist = descriptor_distance(d1, desc2 + 32 * (size_t)i2);
                    if (matched_distance[i2] <= dist) continue;
                    if (dist < bestDist) {
                        bestDist2 = bestDist;
                        bestDist = dist;
                        bestIdx2 = i2;
                    } else if (dist < bestDist2) {
                        bestDist2 = dist;
                    }
                }
            }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nn_ratio) {
                if (matches21[bestIdx2] >= 0) {
                    matches12[matches21[bestIdx2]] = -1;
                    nmatches--;
                }
                matches12[i1] = bestIdx2;
                matches21[bestIdx2] = i1;
                matched_distance[bestIdx2] = bestDist;
                nmatches++;
                if (check_orientation) {
                    float rot = angle1[i1] - angle2[bestIdx2];
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)roundf(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rot_size[bin]++;
                    rot_bin[i1] = bin;
                }
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        compute_three_maxima(rot_size, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i1 = 0; i1 < n1; i1++) {
            const int b = rot_bin[i1];
            if (b < 0 || b == ind1 || b == ind2 || b == ind3) continue;
            if (matches12[i1] >= 0) {
                matches12[i1] = -1;
                nmatches--;
            }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)
        if (matches12[i1] >= 0) {
            prev_matched[2 * i1] = kp2[2 * matches12[i1]];
            prev_matched[2 * i1 + 1] = kp2[2 * matches12[i1] + 1];
        }
    free(g.idx);
    free(rot_bin);
    free(matched_distance);
    free(matches21);
    return nmatches;
}

static int cmp_pair(const void* a, const void* b) {
    const int* x = (const int*)a;
    const int* y = (const int*)b;
    if (x[0] != y[0]) return x[0] < y[0] ? -1 : 1;
    return x[1] < y[1] ? -1 : (x[1] > y[1] ? 1 : 0);
}

/* Frame.cc:811-981 on pyramid level 0 (scale factor 1: every keypoint's octave is 0 in the synthetic frames; the
 * octave gates of :876 are kept).  imL / imR: [rows][cols] u8.  mb = baseline (m), mbf = fx * baseline.
 * Out: u_right [n], depth [n] (-1 = no match), sad [n] (the correlation distance kept for the median test, -1 = none).
 * A patch that would leave the image (the reference's cv::Mat::rowRange would throw) counts as "no match". */
void orb_compute_stereo_matches(int n, const float* kpL, const int* octL, const uint8_t* descL, int nr, const float* kpR,
                                const int* octR, const uint8_t* descR, const uint8_t* imL, const uint8_t* imR, int rows,
                                int cols, float mb, float mbf, float* u_right, float* depth, int* sad) {
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    for (int i = 0; i < n; i++) {
        u_right[i] = -1.0f;
        depth[i] = -1.0f;
        sad[i] = -1;
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    int (*dist_idx)[2] = (int (*)[2])malloc(sizeof(int) * 2 * (n > 0 ? n : 1));
    int nd = 0;
    for (int iL = 0; iL < n; iL++) {
        const int levelL = octL[iL];
        const float vL = kpL[2 * iL + 1], uL = kpL[2 * iL];
        const int row = (int)vL; /* vRowIndices[vL]: float -> size_t */
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        int bestIdxR = 0;
        const uint8_t* dL = descL + 32 * (size_t)iL;
        int any = 0;
        for (int iR = 0; iR < nr; iR++) {
            /* the row table of :825-836: iR sits in rows floor(kpY - r) .. ceil(kpY + r), r = 2 * scale[octave] */
            const float kpY = kpR[2 * iR + 1];
            const float r = 2.0f * 1.0f;
            const int maxr = (int)ceilf(kpY + r), minr = (int)floorf(kpY - r);
            if (row < minr || row > maxr) continue;
            any = 1;
            if (octR[iR] < levelL - 1 || octR[iR] > levelL + 1) continue;
            const float uR = kpR[2 * iR];
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptor_distance(dL, descR + 32 * (size_t)iR);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdxR = iR;
                }
            }
        }
        if (!any) continue;
        if (bestDist < thOrbDist) {
            const float uR0 = kpR[2 * bestIdxR];
            const float scaleFactor = 1.0f;
            const float scaleduL = roundf(uL * scaleFactor), scaledvL = roundf(vL * scaleFactor);
            const float scaleduR0 = roundf(uR0 * scaleFactor);
            const int w = 5, L = 5;
            const int cy = (int)scaledvL, cxl = (int)scaleduL, cxr = (int)scaleduR0;
            if (cy - w < 0 || cy + w + 1 > rows || cxl - w < 0 || cxl + w + 1 > cols) continue; /* would throw */
            int best = INT_MAX, bestincR = 0;
            float vDists[11];
            const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= cols) continue;
            if (cxr - L - w < 0) continue; /* would throw */
            for (int incR = -L; incR <= +L; incR++) {
                int s = 0;
                for (int dy = -w; dy <= w; dy++)
                    for (int dx = -w; dx <= w; dx++)
                        s += abs((int)imL[(size_t)(cy + dy) * cols + cxl + dx] - (int)imR[(size_t)(cy + dy) * cols + cxr + incR + dx]);
                const float dist = (float)s; /* cv::norm(IL, IR, NORM_L1) */
                if (dist < best) {
                    best = (int)dist;
                    bestincR = incR;
                }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = 1.0f * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) {
                    disparity = 0.01;
                    bestuR = uL - 0.01;
                }
                depth[iL] = mbf / disparity;
                u_right[iL] = bestuR;
                sad[iL] = best;
                dist_idx[nd][0] = best;
                dist_idx[nd][1] = iL;
                nd++;
            }
        }
    }
    if (nd > 0) {
        qsort(dist_idx, nd, sizeof(dist_idx[0]), cmp_pair);
        const float median = (float)dist_idx[nd / 2][0];
        const float thDist = 1.5f * 1.4f * median;
        for (int i = nd - 1; i >= 0; i--) {
            if (dist_idx[i][0] < thDist) break;
            u_right[dist_idx[i][1]] = -1;
            depth[dist_idx[i][1]] = -1;
        }
    }
    free(dist_idx);
}
