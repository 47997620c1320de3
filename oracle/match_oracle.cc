// oracle/match_oracle.cc -- TEST INFRASTRUCTURE (CPU oracle). Never linked into / called by the product path.
//
// Flat-array restatements of the reference's Hamming-matching loops (plain arrays in, plain arrays out, so the CUDA path
// and the oracle can be compared on the same seeded inputs at any size).  PINNED against the reference's own code:
// oracle/Makefile `ref` compiles src/ORBmatcher.cc, Frame.cc, KeyFrame.cc, MapPoint.cc, Map.cc, KeyFrameDatabase.cc and the
// vendored DBoW2 UNMODIFIED against the mini-cv shim (oracle/_ref/libmatchref.so, harness oracle/matchref_wrap.cc), and
// tests/test_oracle_matcher.py requires every function below to return what the reference's member function did to real
// Frame / KeyFrame / MapPoint objects built from the same arrays; tests/golden/matcher_ref.npz freezes those outputs for
// the GPU box (tests/test_gpu_matcher_ref.py, tests/test_gpu_adapter_objects.py).
//
//   orbo_descriptor_distance   <- ORBmatcher::DescriptorDistance            src/ORBmatcher.cc:1983-2003
//   orbo_bruteforce            <- sequential best/second scan with that distance (north_star config 5)
//   orbo_search_by_projection  <- ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th)
//                                 src/ORBmatcher.cc:61-183 + Frame::GetFeaturesInArea src/Frame.cc:488-548
//   orbo_stereo_rows           <- coarse part of Frame::ComputeStereoMatches  src/Frame.cc:684-765
//   orbo_search_by_bow         <- ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) src/ORBmatcher.cc:228-392
//                                 + ComputeThreeMaxima :1935-1977
//   orbo_assign_grid           <- Frame::AssignFeaturesToGrid / PosInGrid    src/Frame.cc:351-370, 558-567
//   orbo_search_for_initialization / orbo_search_best (Fuse x2, SearchBySim3) / orbo_search_for_triangulation
//                              <- src/ORBmatcher.cc:539-683, 1067-1357, 1361-1602, 850-1056 (see the headers below)
#include <cstdint>
#include <cstddef>
#include <cstring>
#include <cmath>
#include <vector>
#include <algorithm>
#include <thread>
#include "cvprim.h"

namespace {

const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;
const int GRID_COLS = 64, GRID_ROWS = 48;

// the reference's SWAR popcount, word by word (src/ORBmatcher.cc:1992-2000)
inline int hamming256(const uint8_t* a, const uint8_t* b)
{
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t x, y;
        std::memcpy(&x, a + 4 * i, 4);
        std::memcpy(&y, b + 4 * i, 4);
        uint32_t v = x ^ y;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

}  // namespace

extern "C" {

int orbo_descriptor_distance(const uint8_t* a, const uint8_t* b) { return hamming256(a, b); }

// For each query: best distance, index of the first database row attaining it, and the second-smallest
// distance of the multiset (sequential scan: `if d < best {second = best; best = d} else if d < second`).
// Empty database: best = second = 256, idx = -1.
void orbo_bruteforce(const uint8_t* q, int nq, const uint8_t* db, int ndb, int32_t* best, int32_t* best_idx,
                     int32_t* second, int threads)
{
    if (threads < 1) threads = 1;
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([=]() {
            for (int i = t; i < nq; i += threads) {
                int b1 = 256, b2 = 256, bi = -1;
                for (int j = 0; j < ndb; j++) {
                    int d = hamming256(q + 32 * (size_t)i, db + 32 * (size_t)j);
                    if (d < b1) { b2 = b1; b1 = d; bi = j; }
                    else if (d < b2) b2 = d;
                }
                best[i] = b1; best_idx[i] = bi; second[i] = b2;
            }
        });
    for (auto& x : th) x.join();
}

// Frame::AssignFeaturesToGrid: cell (ix, iy) flattened as ix*48+iy; CSR output, items in keypoint order.
void orbo_assign_grid(const float* kx, const float* ky, int n, float minX, float minY, float wInv, float hInv,
                      int32_t* cell_start /*64*48+1*/, int32_t* cell_items /*n*/)
{
    std::vector<std::vector<int>> g(GRID_COLS * GRID_ROWS);
    for (int i = 0; i < n; i++) {
        int px = (int)std::round((kx[i] - minX) * wInv);
        int py = (int)std::round((ky[i] - minY) * hInv);
        if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) continue;
        g[px * GRID_ROWS + py].push_back(i);
    }
    int off = 0;
    for (int c = 0; c < GRID_COLS * GRID_ROWS; c++) {
        cell_start[c] = off;
        for (int v : g[c]) cell_items[off++] = v;
    }
    cell_start[GRID_COLS * GRID_ROWS] = off;
}

// Frame::UndistortKeyPoints (src/Frame.cc:586-628): mvKeysUn = mvKeys when mDistCoef(0) == 0, else
// cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK).  OpenCV is not vendored by the reference; this restates
// cvUndistortPointsInternal of OpenCV 4.x for that call (R = identity, P = K, default criteria = 5 iterations, no
// tilt terms): everything in double, K and the coefficients widened from float, result narrowed to float.
// Pinned bit-for-bit against cv2 4.13 (tests/test_oracle_primitives.py, tests/golden/undistort_cv2.npz).
void orbo_undistort_points(const float* px, const float* py, int n, float fxf, float fyf, float cxf, float cyf,
                           const float* dist, int ndist, float* ox, float* oy)
{
    if (ndist <= 0 || dist[0] == 0.0f) {  // src/Frame.cc:589-592
        for (int i = 0; i < n; i++) { ox[i] = px[i]; oy[i] = py[i]; }
        return;
    }
    orbo::undistort_points(px, py, n, fxf, fyf, cxf, cyf, dist, ndist, ox, oy);
}

// SearchByProjection(Frame&, vector<MapPoint*>&, th).  out_match[idx] = index of the map point written into
// F.mvpMapPoints[idx] by this call (the last writer), or -1 when untouched.  Returns nmatches.
int orbo_search_by_projection(
    // frame
    int nkp, const float* kx, const float* ky, const int32_t* koct, const float* uright, const uint8_t* kdesc,
    const uint8_t* occupied, const int32_t* cell_start, const int32_t* cell_items, float minX, float minY,
    float wInv, float hInv, const float* scaleFactors,
    // map points
    int nmp, const float* px, const float* py, const float* pxr, const int32_t* plevel, const float* pviewcos,
    const uint8_t* pvalid, const uint8_t* pobs, const uint8_t* pdesc,
    float th, float nnratio, int32_t* out_match)
{
    (void)nkp;
    std::vector<uint8_t> occ(occupied, occupied + nkp);
    for (int i = 0; i < nkp; i++) out_match[i] = -1;
    int nmatches = 0;
    const bool bFactor = th != 1.0f;
    for (int m = 0; m < nmp; m++) {
        if (!pvalid[m]) continue;
        const int level = plevel[m];
        float r = pviewcos[m] > 0.998 ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        const float rr = r * scaleFactors[level];
        const int minLevel = level - 1, maxLevel = level;
        // GetFeaturesInArea(x, y, rr, minLevel, maxLevel)
        const float x = px[m], y = py[m];
        const int c0 = std::max(0, (int)std::floor((x - minX - rr) * wInv));
        if (c0 >= GRID_COLS) continue;
        const int c1 = std::min(GRID_COLS - 1, (int)std::ceil((x - minX + rr) * wInv));
        if (c1 < 0) continue;
        const int r0 = std::max(0, (int)std::floor((y - minY - rr) * hInv));
        if (r0 >= GRID_ROWS) continue;
        const int r1 = std::min(GRID_ROWS - 1, (int)std::ceil((y - minY + rr) * hInv));
        if (r1 < 0) continue;
        const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        bool any = false;
        for (int ix = c0; ix <= c1; ix++)
            for (int iy = r0; iy <= r1; iy++) {
                const int c = ix * GRID_ROWS + iy;
                for (int k = cell_start[c]; k < cell_start[c + 1]; k++) {
                    const int idx = cell_items[k];
                    if (checkLevels) {
                        if (koct[idx] < minLevel) continue;
                        if (maxLevel >= 0 && koct[idx] > maxLevel) continue;
                    }
                    const float dx = kx[idx] - x, dy = ky[idx] - y;
                    if (!(std::fabs(dx) < rr && std::fabs(dy) < rr)) continue;
                    any = true;
                    // candidate loop body (src/ORBmatcher.cc:113-151)
                    if (occ[idx]) continue;
                    if (uright && uright[idx] > 0) {
                        const float er = std::fabs(pxr[m] - uright[idx]);
                        if (er > rr) continue;
                    }
                    const int dist = hamming256(pdesc + 32 * (size_t)m, kdesc + 32 * (size_t)idx);
                    if (dist < bestDist) {
                        bestDist2 = bestDist; bestDist = dist;
                        bestLevel2 = bestLevel; bestLevel = koct[idx];
                        bestIdx = idx;
                    } else if (dist < bestDist2) {
                        bestLevel2 = koct[idx];
                        bestDist2 = dist;
                    }
                }
            }
        if (!any) continue;
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            out_match[bestIdx] = m;
            occ[bestIdx] = pobs[m] ? 1 : 0;
            nmatches++;
        }
    }
    return nmatches;
}

// Coarse stereo search.  For each left keypoint: best right index (or -1) and its distance.
// minZ = mb = mbf/fx (decision D5), minD = 0, maxD = mbf/minZ.
void orbo_stereo_rows(int nl, const float* lx, const float* ly, const int32_t* loct, const uint8_t* ldesc,
                      int nr, const float* rx, const float* ry, const int32_t* roct, const uint8_t* rdesc,
                      const float* scaleFactors, int nRows, float minD, float maxD,
                      int32_t* best_r, int32_t* best_dist)
{
    std::vector<std::vector<int>> rows(nRows);
    for (int iR = 0; iR < nr; iR++) {
        const float r = 2.0f * scaleFactors[roct[iR]];
        const int maxr = (int)std::ceil(ry[iR] + r);
        const int minr = (int)std::floor(ry[iR] - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) rows[yi].push_back(iR);
    }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    for (int iL = 0; iL < nl; iL++) {
        best_r[iL] = -1; best_dist[iL] = TH_HIGH;
        const int row = (int)ly[iL];
        if (row < 0 || row >= nRows) continue;
        const std::vector<int>& cand = rows[row];
        if (cand.empty()) continue;
        const float minU = lx[iL] - maxD, maxU = lx[iL] - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH, bestIdx = 0;
        for (int iR : cand) {
            if (roct[iR] < loct[iL] - 1 || roct[iR] > loct[iL] + 1) continue;
            const float uR = rx[iR];
            if (uR >= minU && uR <= maxU) {
                const int d = hamming256(ldesc + 32 * (size_t)iL, rdesc + 32 * (size_t)iR);
                if (d < bestDist) { bestDist = d; bestIdx = iR; }
            }
        }
        best_dist[iL] = bestDist;
        if (bestDist < thOrbDist) best_r[iL] = bestIdx;
    }
}

// SearchByBoW(KeyFrame*, Frame&, matches).  Feature vectors are CSR: node ids ascending, features per node.
// kf_has_mp[i] = vpMapPointsKF[i] != NULL && !isBad().  out_match[f] = KF feature index matched to frame
// feature f, or -1.  Returns nmatches (after the orientation filter).
int orbo_search_by_bow(
    int nkf, const uint8_t* kf_desc, const float* kf_angle, const uint8_t* kf_has_mp,
    int n_nodes_kf, const int32_t* kf_node_id, const int32_t* kf_node_start, const int32_t* kf_node_feat,
    int nf, const uint8_t* f_desc, const float* f_angle,
    int n_nodes_f, const int32_t* f_node_id, const int32_t* f_node_start, const int32_t* f_node_feat,
    float nnratio, int checkOrientation, int32_t* out_match)
{
    (void)nkf;
    for (int i = 0; i < nf; i++) out_match[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    int a = 0, b = 0;
    while (a < n_nodes_kf && b < n_nodes_f) {
        if (kf_node_id[a] == f_node_id[b]) {
            for (int ik = kf_node_start[a]; ik < kf_node_start[a + 1]; ik++) {
                const int realIdxKF = kf_node_feat[ik];
                if (!kf_has_mp[realIdxKF]) continue;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                for (int jf = f_node_start[b]; jf < f_node_start[b + 1]; jf++) {
                    const int realIdxF = f_node_feat[jf];
                    if (out_match[realIdxF] >= 0) continue;
                    const int dist = hamming256(kf_desc + 32 * (size_t)realIdxKF, f_desc + 32 * (size_t)realIdxF);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 <= TH_LOW) {
                    if ((float)bestDist1 < nnratio * (float)bestDist2) {
                        out_match[bestIdxF] = realIdxKF;
                        if (checkOrientation) {
                            float rot = kf_angle[realIdxKF] - f_angle[bestIdxF];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(bestIdxF);
                        }
                        nmatches++;
                    }
                }
            }
            a++; b++;
        } else if (kf_node_id[a] < f_node_id[b]) {
            a = (int)(std::lower_bound(kf_node_id + a, kf_node_id + n_nodes_kf, f_node_id[b]) - kf_node_id);
        } else {
            b = (int)(std::lower_bound(f_node_id + b, f_node_id + n_nodes_f, kf_node_id[a]) - f_node_id);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        int max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int f : rotHist[i]) { out_match[f] = -1; nmatches--; }
        }
    }
    return nmatches;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// Full Frame::ComputeStereoMatches (src/Frame.cc:673-885): coarse row search (orbo_stereo_rows above) + SAD refinement
// on the pyramid levels with sub-pixel parabola fit (:767-866) + removal of matches whose SAD distance is >= 2.1 x median
// (:868-884).  Decision D7: this reference executes the outlier block INSIDE the per-keypoint loop (upstream ORB-SLAM2 has
// it after the loop); outlier_mode 0 reproduces that placement and is what the compiled reference is compared with
// (tests/test_oracle_matcher.py): the block runs at the end of every iteration that reaches it, reads element 0 of a
// still-empty vector before the first match (never used: its loop does not run) and rejects against the running median.
// outlier_mode 1 = upstream placement, kept for callers that link against upstream semantics.  Decision D5: minZ = mb.
// Level planes are given as pointers to the interior pixel (0,0) of the padded pyramid planes.
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int orbo_stereo_match(
    int nl, const float* lx, const float* ly, const int32_t* loct, const uint8_t* ldesc,
    int nr, const float* rx, const float* ry, const int32_t* roct, const uint8_t* rdesc,
    int nlevels, const float* scaleFactors, const float* invScaleFactors,
    const uint8_t* const* lplane, const uint8_t* const* rplane, const int32_t* lstep, const int32_t* rstep, const int32_t* rcols,
    int nRows, float mbf, float mb, int outlier_mode, float* uRight, float* depth)
{
    (void)nlevels;
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    std::vector<int32_t> bestR(nl > 0 ? nl : 1), bestDistCoarse(nl > 0 ? nl : 1);
    orbo_stereo_rows(nl, lx, ly, loct, ldesc, nr, rx, ry, roct, rdesc, scaleFactors, nRows, minD, maxD, bestR.data(), bestDistCoarse.data());
    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < nl; iL++) { uRight[iL] = -1.0f; depth[iL] = -1.0f; }
    for (int iL = 0; iL < nl; iL++) {
        if (bestR[iL] < 0) continue;
        const int oct = loct[iL];
        const float uL = lx[iL];
        const float uR0 = rx[bestR[iL]];
        const float scaleFactor = invScaleFactors[oct];
        const float scaleduL = std::round(lx[iL] * scaleFactor);
        const float scaledvL = std::round(ly[iL] * scaleFactor);
        const float scaleduR0 = std::round(uR0 * scaleFactor);
        const int w = 5, L = 5;
        const uint8_t* PL = lplane[oct];
        const uint8_t* PR = rplane[oct];
        const int cvL = (int)scaledvL, cuL = (int)scaleduL;
        const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
        if (iniu < 0 || endu >= rcols[oct]) continue;
        int bestDist = INT32_MAX, bestincR = 0;
        float vDists[2 * 5 + 1];
        const float centerL = (float)PL[(ptrdiff_t)cvL * lstep[oct] + cuL];
        for (int incR = -L; incR <= L; incR++) {
            const int cuR = (int)(scaleduR0 + incR);
            const float centerR = (float)PR[(ptrdiff_t)cvL * rstep[oct] + cuR];
            float dist = 0;  // cv::norm(IL, IR, NORM_L1): integer-valued, exact in float
            for (int dy = -w; dy <= w; dy++)
                for (int dx = -w; dx <= w; dx++) {
                    const float a = (float)PL[(ptrdiff_t)(cvL + dy) * lstep[oct] + cuL + dx] - centerL;
                    const float b = (float)PR[(ptrdiff_t)(cvL + dy) * rstep[oct] + cuR + dx] - centerR;
                    dist += std::fabs(a - b);
                }
            if (dist < bestDist) { bestDist = (int)dist; bestincR = incR; }
            vDists[L + incR] = dist;
        }
        if (bestincR == -L || bestincR == L) continue;
        const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
        const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
        if (deltaR < -1 || deltaR > 1) continue;
        float bestuR = scaleFactors[oct] * ((float)scaleduR0 + (float)bestincR + deltaR);
        float disparity = (uL - bestuR);
        if (disparity >= minD && disparity < maxD) {
            if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
            depth[iL] = mbf / disparity;
            uRight[iL] = bestuR;
            vDistIdx.push_back(std::pair<int, int>(bestDist, iL));
        }
        if (outlier_mode == 0 && !vDistIdx.empty()) {  // src/Frame.cc:868-884, inside the loop
            std::sort(vDistIdx.begin(), vDistIdx.end());
            const float median = vDistIdx[vDistIdx.size() / 2].first;
            const float thDist = 1.5f * 1.4f * median;
            for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
                if (vDistIdx[i].first < thDist) break;
                uRight[vDistIdx[i].second] = -1;
                depth[vDistIdx[i].second] = -1;
            }
        }
    }
    if (outlier_mode == 1 && !vDistIdx.empty()) {
        std::sort(vDistIdx.begin(), vDistIdx.end());
        const float median = vDistIdx[vDistIdx.size() / 2].first;
        const float thDist = 1.5f * 1.4f * median;
        for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
            if (vDistIdx[i].first < thDist) break;
            uRight[vDistIdx[i].second] = -1;
            depth[vDistIdx[i].second] = -1;
        }
    }
    return (int)vDistIdx.size();
}

// ---------------------------------------------------------------------------------------------------------------------
// DBoW2 vocabulary descent: TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1259) with F::distance = FORB::distance (FORB.cpp:81-101, the same
// bit-hack Hamming).  The tree is flat: children of node i are child_ids[child_start[i] .. child_start[i+1]), a node
// without children is a leaf carrying word_id / weight.  First minimum in child order wins (strict <).
// ---------------------------------------------------------------------------------------------------------------------
extern "C" void orbo_bow_transform(int n_nodes, int L, const int32_t* child_start, const int32_t* child_ids,
                                   const uint8_t* node_desc, const int32_t* word_id, const double* weight,
                                   const uint8_t* desc, int n, int levelsup, int32_t* out_word, int32_t* out_node,
                                   double* out_weight)
{
    (void)n_nodes;
    const int nid_level = L - levelsup;
    for (int f = 0; f < n; f++) {
        int nid = 0;
        int final_id = 0, current_level = 0;
        do {
            ++current_level;
            const int cs = child_start[final_id], ce = child_start[final_id + 1];
            int best = child_ids[cs];
            double best_d = hamming256(desc + 32 * (size_t)f, node_desc + 32 * (size_t)best);
            for (int c = cs + 1; c < ce; c++) {
                const int id = child_ids[c];
                const double d = hamming256(desc + 32 * (size_t)f, node_desc + 32 * (size_t)id);
                if (d < best_d) { best_d = d; best = id; }
            }
            final_id = best;
            if (current_level == nid_level) nid = final_id;
        } while (child_start[final_id + 1] > child_start[final_id]);
        out_word[f] = word_id[final_id];
        out_weight[f] = weight[final_id];
        out_node[f] = nid;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Windowed search shared by the remaining SearchByProjection overloads ("next" row):
//   ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)              src/ORBmatcher.cc:1620-1789
//   ORBmatcher::SearchByProjection(Frame& Cur, KeyFrame*, sAlreadyFound, th, ORBdist)     src/ORBmatcher.cc:1792-1924
// after their projection step (done by the caller with the reference's own cv::Mat arithmetic): per source feature a
// projected position (u, v), window radius, the level arguments handed to Frame::GetFeaturesInArea, the predicted right
// coordinate `ur` and the source keypoint angle.  Candidate loop, acceptance `bestDist <= maxDist`, claims, rotation
// histogram and ComputeThreeMaxima exactly as in those functions.  ratio_mode 1 adds the level-aware ratio test of the
// (Frame, MapPoints) overload.  out_match: -1 untouched, -2 set to NULL by the orientation filter, else query index.
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int orbo_search_window(
    int nkp, const float* kx, const float* ky, const int32_t* koct, const float* uright, const uint8_t* kdesc,
    const uint8_t* occupied, const int32_t* cell_start, const int32_t* cell_items, float minX, float minY, float wInv,
    float hInv, const float* kangle,
    int nq, const float* qx, const float* qy, const float* qradius, const int32_t* qmin, const int32_t* qmax,
    const float* qxr, const float* qangle, const uint8_t* qvalid, const uint8_t* qblocks, const uint8_t* qdesc,
    int maxDist, int ratioMode, float nnratio, int checkOrientation, int32_t* out_match)
{
    std::vector<uint8_t> occ(occupied, occupied + nkp);
    for (int i = 0; i < nkp; i++) out_match[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    for (int m = 0; m < nq; m++) {
        if (!qvalid[m]) continue;
        const float x = qx[m], y = qy[m], rr = qradius[m];
        const int minLevel = qmin[m], maxLevel = qmax[m];
        const int c0 = std::max(0, (int)std::floor((x - minX - rr) * wInv));
        if (c0 >= GRID_COLS) continue;
        const int c1 = std::min(GRID_COLS - 1, (int)std::ceil((x - minX + rr) * wInv));
        if (c1 < 0) continue;
        const int r0 = std::max(0, (int)std::floor((y - minY - rr) * hInv));
        if (r0 >= GRID_ROWS) continue;
        const int r1 = std::min(GRID_ROWS - 1, (int)std::ceil((y - minY + rr) * hInv));
        if (r1 < 0) continue;
        const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int ix = c0; ix <= c1; ix++)
            for (int iy = r0; iy <= r1; iy++) {
                const int c = ix * GRID_ROWS + iy;
                for (int k = cell_start[c]; k < cell_start[c + 1]; k++) {
                    const int idx = cell_items[k];
                    if (checkLevels) {
                        if (koct[idx] < minLevel) continue;
                        if (maxLevel >= 0 && koct[idx] > maxLevel) continue;
                    }
                    const float dx = kx[idx] - x, dy = ky[idx] - y;
                    if (!(std::fabs(dx) < rr && std::fabs(dy) < rr)) continue;
                    if (occ[idx]) continue;
                    if (uright && uright[idx] > 0) {
                        const float er = std::fabs(qxr[m] - uright[idx]);
                        if (er > rr) continue;
                    }
                    const int dist = hamming256(qdesc + 32 * (size_t)m, kdesc + 32 * (size_t)idx);
                    if (dist < bestDist) {
                        bestDist2 = bestDist; bestDist = dist;
                        bestLevel2 = bestLevel; bestLevel = koct[idx];
                        bestIdx = idx;
                    } else if (dist < bestDist2) {
                        bestLevel2 = koct[idx];
                        bestDist2 = dist;
                    }
                }
            }
        if (bestIdx < 0 || bestDist > maxDist) continue;
        if (ratioMode == 1 && bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
        out_match[bestIdx] = m;
        occ[bestIdx] = qblocks[m] ? 1 : 0;
        nmatches++;
        if (checkOrientation) {
            float rot = qangle[m] - kangle[bestIdx];
            if (rot < 0.0) rot += 360.0f;
            int bin = (int)std::round(rot * factor);
            if (bin == HISTO_LENGTH) bin = 0;
            rotHist[bin].push_back(bestIdx);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int f : rotHist[i]) { out_match[f] = -2; nmatches--; }
        }
    }
    return nmatches;
}

// ---------------------------------------------------------------------------------------------------------------------
// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:295-360), batched over map points: for every set of
// observation descriptors the index of the descriptor with the least median distance to the others
// (median = sorted row element [0.5*(N-1)], first minimum wins).  Sets are CSR: rows set_start[s] .. set_start[s+1].
// ---------------------------------------------------------------------------------------------------------------------
extern "C" void orbo_distinctive_descriptors(int n_sets, const int32_t* set_start, const uint8_t* desc, int32_t* best_idx)
{
    for (int s = 0; s < n_sets; s++) {
        const int b = set_start[s], N = set_start[s + 1] - b;
        if (N <= 0) { best_idx[s] = -1; continue; }
        std::vector<std::vector<float>> D(N, std::vector<float>(N, 0));
        for (int i = 0; i < N; i++)
            for (int j = i + 1; j < N; j++) {
                const int d = hamming256(desc + 32 * (size_t)(b + i), desc + 32 * (size_t)(b + j));
                D[i][j] = (float)d; D[j][i] = (float)d;
            }
        int bestMedian = INT32_MAX, bestIdx = 0;
        for (int i = 0; i < N; i++) {
            std::vector<int> v(D[i].begin(), D[i].end());
            std::sort(v.begin(), v.end());
            const int median = v[(size_t)(0.5 * (N - 1))];
            if (median < bestMedian) { bestMedian = median; bestIdx = i; }
        }
        best_idx[s] = bestIdx;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12)  src/ORBmatcher.cc:696-850.
// has_mp1/2[i] = feature i of that key frame has a non-bad MapPoint.  out12[i1] = matched feature of KF2 or -1.
// Differences to the (KeyFrame, Frame) overload: both sides need map points, KF2 features are blocked through
// vbMatched2, acceptance is bestDist1 < TH_LOW (strict), the rotation histogram stores idx1.
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int orbo_search_by_bow_kf(
    int n1, const uint8_t* desc1, const float* angle1, const uint8_t* has_mp1,
    int n_nodes1, const int32_t* node_id1, const int32_t* node_start1, const int32_t* node_feat1,
    int n2, const uint8_t* desc2, const float* angle2, const uint8_t* has_mp2,
    int n_nodes2, const int32_t* node_id2, const int32_t* node_start2, const int32_t* node_feat2,
    float nnratio, int checkOrientation, int32_t* out12)
{
    for (int i = 0; i < n1; i++) out12[i] = -1;
    std::vector<uint8_t> matched2(n2 > 0 ? n2 : 1, 0);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    int nmatches = 0, a = 0, b = 0;
    while (a < n_nodes1 && b < n_nodes2) {
        if (node_id1[a] == node_id2[b]) {
            for (int i1 = node_start1[a]; i1 < node_start1[a + 1]; i1++) {
                const int idx1 = node_feat1[i1];
                if (!has_mp1[idx1]) continue;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (int i2 = node_start2[b]; i2 < node_start2[b + 1]; i2++) {
                    const int idx2 = node_feat2[i2];
                    if (matched2[idx2] || !has_mp2[idx2]) continue;
                    const int dist = hamming256(desc1 + 32 * (size_t)idx1, desc2 + 32 * (size_t)idx2);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 < TH_LOW) {
                    if ((float)bestDist1 < nnratio * (float)bestDist2) {
                        out12[idx1] = bestIdx2;
                        matched2[bestIdx2] = 1;
                        if (checkOrientation) {
                            float rot = angle1[idx1] - angle2[bestIdx2];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(idx1);
                        }
                        nmatches++;
                    }
                }
            }
            a++; b++;
        } else if (node_id1[a] < node_id2[b]) {
            a = (int)(std::lower_bound(node_id1 + a, node_id1 + n_nodes1, node_id2[b]) - node_id1);
        } else {
            b = (int)(std::lower_bound(node_id2 + b, node_id2 + n_nodes2, node_id1[a]) - node_id2);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int f : rotHist[i]) { out12[f] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// ---------------------------------------------------------------------------------------------------------------------
// Shared by the four remaining matcher overloads below: ComputeThreeMaxima (src/ORBmatcher.cc:1935-1977) on bin sizes.
// ---------------------------------------------------------------------------------------------------------------------
namespace {
void three_maxima(const std::vector<int>* rotHist, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < HISTO_LENGTH; i++) {
        const int s = (int)rotHist[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) ind3 = -1;
}
inline int rot_bin(float a1, float a2)
{
    const float factor = HISTO_LENGTH / 360.0f;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}
}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)  src/ORBmatcher.cc:539-683.
// Frame 2 is given as in orbo_search_window (keypoints + grid); frame 1 as octave / angle / descriptor and the
// vbPrevMatched positions.  Only level-0 features of F1 search, inside F2.GetFeaturesInArea(x, y, windowSize, 0, 0); a
// candidate is skipped when an earlier F1 feature already holds it with a distance <= this one (vMatchedDistance), a
// later better match steals it (vnMatches21 / vnMatches12 reset); the stolen match stays in the rotation histogram.
// out12[i1] = matched F2 feature or -1; returns nmatches.
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int orbo_search_for_initialization(
    int n1, const int32_t* oct1, const float* angle1, const uint8_t* desc1, const float* prevx, const float* prevy,
    int n2, const float* kx, const float* ky, const int32_t* koct, const float* kangle, const uint8_t* kdesc,
    const int32_t* cell_start, const int32_t* cell_items, float minX, float minY, float wInv, float hInv,
    int windowSize, float nnratio, int checkOrientation, int32_t* out12)
{
    int nmatches = 0;
    for (int i = 0; i < n1; i++) out12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vMatchedDistance(n2 > 0 ? n2 : 1, 0x7fffffff), vnMatches21(n2 > 0 ? n2 : 1, -1);
    for (int i1 = 0; i1 < n1; i1++) {
        const int level1 = oct1[i1];
        if (level1 > 0) continue;
        const float x = prevx[i1], y = prevy[i1], rr = (float)windowSize;
        const int c0 = std::max(0, (int)std::floor((x - minX - rr) * wInv));
        if (c0 >= GRID_COLS) continue;
        const int c1 = std::min(GRID_COLS - 1, (int)std::ceil((x - minX + rr) * wInv));
        if (c1 < 0) continue;
        const int r0 = std::max(0, (int)std::floor((y - minY - rr) * hInv));
        if (r0 >= GRID_ROWS) continue;
        const int r1 = std::min(GRID_ROWS - 1, (int)std::ceil((y - minY + rr) * hInv));
        if (r1 < 0) continue;
        int bestDist = 0x7fffffff, bestDist2 = 0x7fffffff, bestIdx2 = -1;
        for (int ix = c0; ix <= c1; ix++)
            for (int iy = r0; iy <= r1; iy++) {
                const int c = ix * GRID_ROWS + iy;
                for (int k = cell_start[c]; k < cell_start[c + 1]; k++) {
                    const int i2 = cell_items[k];
                    if (koct[i2] < level1) continue;   // bCheckLevels: minLevel = maxLevel = level1 = 0
                    if (koct[i2] > level1) continue;
                    const float dx = kx[i2] - x, dy = ky[i2] - y;
                    if (!(std::fabs(dx) < rr && std::fabs(dy) < rr)) continue;
                    const int dist = hamming256(desc1 + 32 * (size_t)i1, kdesc + 32 * (size_t)i2);
                    if (vMatchedDistance[i2] <= dist) continue;
                    if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
            }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { out12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                out12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOrientation) rotHist[rot_bin(angle1[i1], kangle[bestIdx2])].push_back(i1);
            }
        }
    }
    if (checkOrientation) {
        int ind1, ind2, ind3;
        three_maxima(rotHist, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (out12[idx1] >= 0) { out12[idx1] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// ---------------------------------------------------------------------------------------------------------------------
// Claim-free windowed search: the candidate loop shared by
//   ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th)                      src/ORBmatcher.cc:1067-1222  (gate 1)
//   ORBmatcher::Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint)                 src/ORBmatcher.cc:1225-1357  (gate 0)
//   the two projection passes of ORBmatcher::SearchBySim3                          src/ORBmatcher.cc:1361-1602  (gate 0)
// after the caller's projection (u, v, ur, radius = th * scale[predicted level], levels [pred-1, pred]).  Candidates come
// from KeyFrame::GetFeaturesInArea(u, v, radius) (src/KeyFrame.cc:623-658), best = first minimum (strict <).  gate 1 is
// Fuse's reprojection test: stereo keypoints (mvuRight >= 0) e2 = ex^2+ey^2+er^2, e2*invSigma2 > 7.8 rejects, monocular
// ones e2 = ex^2+ey^2, > 5.99 rejects.  best_idx[q] = -1 when there is no candidate or bestDist > maxDist.
// ---------------------------------------------------------------------------------------------------------------------
extern "C" void orbo_search_best(
    int nkp, const float* kx, const float* ky, const int32_t* koct, const float* uright, const uint8_t* kdesc,
    const int32_t* cell_start, const int32_t* cell_items, float minX, float minY, float wInv, float hInv,
    int nq, const float* qx, const float* qy, const float* qradius, const int32_t* qmin, const int32_t* qmax,
    const float* qxr, const uint8_t* qvalid, const uint8_t* qdesc, int gate, const float* invLevelSigma2, int maxDist,
    int32_t* best_idx, int32_t* best_dist)
{
    (void)nkp;
    for (int m = 0; m < nq; m++) {
        best_idx[m] = -1;
        best_dist[m] = 256;
        if (!qvalid[m]) continue;
        const float u = qx[m], v = qy[m], rr = qradius[m];
        const int c0 = std::max(0, (int)std::floor((u - minX - rr) * wInv));
        if (c0 >= GRID_COLS) continue;
        const int c1 = std::min(GRID_COLS - 1, (int)std::ceil((u - minX + rr) * wInv));
        if (c1 < 0) continue;
        const int r0 = std::max(0, (int)std::floor((v - minY - rr) * hInv));
        if (r0 >= GRID_ROWS) continue;
        const int r1 = std::min(GRID_ROWS - 1, (int)std::ceil((v - minY + rr) * hInv));
        if (r1 < 0) continue;
        int bestDist = 256, bestIdx = -1;
        for (int ix = c0; ix <= c1; ix++)
            for (int iy = r0; iy <= r1; iy++) {
                const int c = ix * GRID_ROWS + iy;
                for (int k = cell_start[c]; k < cell_start[c + 1]; k++) {
                    const int idx = cell_items[k];
                    const float dx = kx[idx] - u, dy = ky[idx] - v;
                    if (!(std::fabs(dx) < rr && std::fabs(dy) < rr)) continue;
                    const int kpLevel = koct[idx];
                    if (kpLevel < qmin[m] || kpLevel > qmax[m]) continue;
                    if (gate == 1) {
                        if (uright && uright[idx] >= 0) {
                            const float ex = u - kx[idx], ey = v - ky[idx], er = qxr[m] - uright[idx];
                            const float e2 = ex * ex + ey * ey + er * er;
                            if (e2 * invLevelSigma2[kpLevel] > 7.8) continue;
                        } else {
                            const float ex = u - kx[idx], ey = v - ky[idx];
                            const float e2 = ex * ex + ey * ey;
                            if (e2 * invLevelSigma2[kpLevel] > 5.99) continue;
                        }
                    }
                    const int dist = hamming256(qdesc + 32 * (size_t)m, kdesc + 32 * (size_t)idx);
                    if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
                }
            }
        best_dist[m] = bestDist;
        if (bestIdx >= 0 && bestDist <= maxDist) best_idx[m] = bestIdx;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)  src/ORBmatcher.cc:850-1056 with
// CheckDistEpipolarLine :186-213.  Feature vectors in CSR form; has_mp = GetMapPoint(idx) != NULL; ur = mvuRight (>= 0 is
// stereo); (ex, ey) = epipole of KF1's centre in KF2 (computed by the caller, :858-865); F12 row-major 3x3.
// Candidate rule: `dist > TH_LOW || dist > bestDist` skips, so among the candidates passing the gates the minimum
// distance wins and, on ties, the LAST one in node order.  out12[idx1] = idx2 or -1; returns nmatches.
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int orbo_search_for_triangulation(
    int n1, const float* x1, const float* y1, const float* angle1, const float* ur1, const uint8_t* desc1, const uint8_t* has_mp1,
    int n_nodes1, const int32_t* node_id1, const int32_t* node_start1, const int32_t* node_feat1,
    int n2, const float* x2, const float* y2, const int32_t* oct2, const float* angle2, const float* ur2, const uint8_t* desc2,
    const uint8_t* has_mp2, int n_nodes2, const int32_t* node_id2, const int32_t* node_start2, const int32_t* node_feat2,
    const float* F12, float ex, float ey, const float* scaleFactors2, const float* levelSigma2_2, int onlyStereo,
    int checkOrientation, int32_t* out12)
{
    for (int i = 0; i < n1; i++) out12[i] = -1;
    std::vector<uint8_t> matched2(n2 > 0 ? n2 : 1, 0);
    std::vector<int> rotHist[HISTO_LENGTH];
    int nmatches = 0, a = 0, b = 0;
    while (a < n_nodes1 && b < n_nodes2) {
        if (node_id1[a] == node_id2[b]) {
            for (int i1 = node_start1[a]; i1 < node_start1[a + 1]; i1++) {
                const int idx1 = node_feat1[i1];
                if (has_mp1[idx1]) continue;
                const bool bStereo1 = ur1[idx1] >= 0;
                if (onlyStereo && !bStereo1) continue;
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (int i2 = node_start2[b]; i2 < node_start2[b + 1]; i2++) {
                    const int idx2 = node_feat2[i2];
                    if (matched2[idx2] || has_mp2[idx2]) continue;
                    const bool bStereo2 = ur2[idx2] >= 0;
                    if (onlyStereo && !bStereo2) continue;
                    const int dist = hamming256(desc1 + 32 * (size_t)idx1, desc2 + 32 * (size_t)idx2);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - x2[idx2], distey = ey - y2[idx2];
                        if (distex * distex + distey * distey < 100 * scaleFactors2[oct2[idx2]]) continue;
                    }
                    // CheckDistEpipolarLine
                    const float la = x1[idx1] * F12[0] + y1[idx1] * F12[3] + F12[6];
                    const float lb = x1[idx1] * F12[1] + y1[idx1] * F12[4] + F12[7];
                    const float lc = x1[idx1] * F12[2] + y1[idx1] * F12[5] + F12[8];
                    const float num = la * x2[idx2] + lb * y2[idx2] + lc;
                    const float den = la * la + lb * lb;
                    if (den == 0) continue;
                    const float dsqr = num * num / den;
                    if (dsqr < 3.84 * levelSigma2_2[oct2[idx2]]) { bestIdx2 = idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    out12[idx1] = bestIdx2;
                    matched2[bestIdx2] = 1;
                    nmatches++;
                    if (checkOrientation) rotHist[rot_bin(angle1[idx1], angle2[bestIdx2])].push_back(idx1);
                }
            }
            a++; b++;
        } else if (node_id1[a] < node_id2[b]) {
            a = (int)(std::lower_bound(node_id1 + a, node_id1 + n_nodes1, node_id2[b]) - node_id1);
        } else {
            b = (int)(std::lower_bound(node_id2 + b, node_id2 + n_nodes2, node_id1[a]) - node_id2);
        }
    }
    if (checkOrientation) {
        int ind1, ind2, ind3;
        three_maxima(rotHist, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i]) { out12[idx1] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// ---------------------------------------------------------------------------------------------------------------------
// Tail of ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1585-1602): keep i1 -> idx2 only when the reverse search agrees.
// vnMatch1 / vnMatch2 are the results of orbo_search_best for the two projection passes (maxDist = TH_HIGH).
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int orbo_sim3_cross_check(int n1, const int32_t* vnMatch1, int n2, const int32_t* vnMatch2, int32_t* out12)
{
    int nFound = 0;
    for (int i1 = 0; i1 < n1; i1++) {
        out12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0 && idx2 < n2 && vnMatch2[idx2] == i1) { out12[i1] = idx2; nFound++; }
    }
    return nFound;
}
