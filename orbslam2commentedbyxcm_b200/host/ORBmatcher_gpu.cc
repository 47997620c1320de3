// host/ORBmatcher_gpu.cc -- GPU definitions of the reference's ORBmatcher, to be compiled INSIDE the ORB-SLAM2 tree IN PLACE OF
// src/ORBmatcher.cc.  The class declaration stays the reference's include/ORBmatcher.h (unchanged); this
// file defines every member it declares:
//
//   ORBmatcher::ORBmatcher, TH_LOW / TH_HIGH / HISTO_LENGTH                                  src/ORBmatcher.cc:36-50
//   SearchByProjection(Frame&, const vector<MapPoint*>&, th)                                 :61-173
//   RadiusByViewingCos, CheckDistEpipolarLine                                                :176-213
//   SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)                                       :228-392
//   SearchByProjection(KeyFrame*, cv::Mat Scw, vpPoints, vpMatched, th)                      :398-520
//   SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)                  :539-683
//   SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)                                           :696-850
//   SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)                      :850-1056
//   Fuse(KeyFrame*, vpMapPoints, th)                                                         :1067-1222
//   Fuse(KeyFrame*, cv::Mat Scw, vpPoints, th, vpReplacePoint)                               :1225-1357
//   SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th)                                 :1361-1602
//   SearchByProjection(Frame& Current, const Frame& Last, th, bMono)                         :1620-1789
//   SearchByProjection(Frame& Current, KeyFrame*, sAlreadyFound, th, ORBdist)                :1792-1924
//   ComputeThreeMaxima, DescriptorDistance                                                   :1935-2003
// and, for src/Frame.cc,
//   orbx_adapter::StereoCoarseSearch(...)    coarse part of Frame::ComputeStereoMatches      src/Frame.cc:684-765
//   orbx_adapter::ComputeStereoMatches(...)  the whole of Frame::ComputeStereoMatches        src/Frame.cc:673-885
//
// Shape of every search: (1) walk the reference's objects through their own accessors (same locks) and do the projection
// step with the same cv::Mat expressions the reference uses, so the floats handed to the search are the reference's;
// (2) ONE call into the C-ABI of include/orbx.h (candidate windows, Hamming distances, ratio / orientation tests and the
// order-dependent claims run on the GPU); (3) apply the result to the objects in the reference's order.
// tests/test_gpu_adapter_objects.py runs this file, linked with the reference's own Frame / KeyFrame / MapPoint / Map /
// DBoW2 sources (oracle/_ref/libmatchgpu.so), through the same object-level harness as the reference's ORBmatcher.cc.
#include "ORBmatcher.h"

#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>
#include <vector>

#include "orbx.h"
#include "ORBextractor.h"

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = 100;
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

namespace {

// one matcher context per calling thread (Tracking, LocalMapping and LoopClosing call concurrently)
orbm_context* ThreadContext()
{
    static thread_local orbm_context* ctx = NULL;
    if (!ctx) {
        const char* dev = std::getenv("ORBX_DEVICE");
        if (orbm_create(dev ? std::atoi(dev) : 0, &ctx) != ORBX_OK) {
            std::fprintf(stderr, "ORBmatcher: orbm_create failed: %s\n", orbx_last_error());
            ctx = NULL;
        }
    }
    return ctx;
}

bool Failed(int rc, const char* what)
{
    if (rc == ORBX_OK) return false;
    std::fprintf(stderr, "ORBmatcher::%s: %s\n", what, orbx_last_error());
    return true;
}

void FlattenFeatureVector(const DBoW2::FeatureVector& fv, std::vector<int32_t>& ids, std::vector<int32_t>& start,
                          std::vector<int32_t>& feat)
{
    ids.clear(); start.clear(); feat.clear();
    start.push_back(0);
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
        ids.push_back((int32_t)it->first);
        for (size_t k = 0; k < it->second.size(); k++) feat.push_back((int32_t)it->second[k]);
        start.push_back((int32_t)feat.size());
    }
}

void CopyDescriptors(const cv::Mat& m, std::vector<uint8_t>& out)
{
    out.resize((size_t)m.rows * 32);
    for (int i = 0; i < m.rows; i++) std::memcpy(&out[(size_t)i * 32], m.ptr(i), 32);
}

// ---- the searched side: a Frame or a KeyFrame as orbm_frame_view -------------------------------------------------
struct Features {
    std::vector<float> x, y, angle, uRight;
    std::vector<int32_t> octave, cellStart, cellItems;
    std::vector<uint8_t> desc, occupied;
    orbm_frame_view v;

    void Keypoints(const std::vector<cv::KeyPoint>& keysUn, const std::vector<float>& right, const cv::Mat& descriptors)
    {
        const int n = (int)keysUn.size();
        x.resize(n); y.resize(n); angle.resize(n); octave.resize(n);
        for (int i = 0; i < n; i++) {
            x[i] = keysUn[i].pt.x; y[i] = keysUn[i].pt.y; angle[i] = keysUn[i].angle; octave[i] = keysUn[i].octave;
        }
        uRight.assign(right.begin(), right.end());
        CopyDescriptors(descriptors, desc);
        occupied.assign(n > 0 ? n : 1, 0);
        v.n = n;
    }
    void Bind(float minX, float minY, float wInv, float hInv, const std::vector<float>& scaleFactors)
    {
        v.x = x.data(); v.y = y.data(); v.octave = octave.data();
        v.u_right = (int)uRight.size() == v.n && v.n > 0 ? uRight.data() : NULL;
        v.desc = desc.data(); v.occupied = occupied.data();
        v.cell_start = cellStart.data(); v.cell_items = cellItems.data();
        v.min_x = minX; v.min_y = minY; v.grid_w_inv = wInv; v.grid_h_inv = hInv;
        v.scale_factors = scaleFactors.data(); v.nlevels = (int)scaleFactors.size();
    }
};

// Frame::mGrid is public (include/Frame.h:290): flatten it, cell (ix, iy) at ix * FRAME_GRID_ROWS + iy
void FromFrame(const Frame& F, Features& A)
{
    A.Keypoints(F.mvKeysUn, F.mvuRight, F.mDescriptors);
    A.cellStart.resize(FRAME_GRID_COLS * FRAME_GRID_ROWS + 1);
    A.cellItems.clear();
    A.cellItems.reserve(A.v.n > 0 ? A.v.n : 1);
    for (int ix = 0; ix < FRAME_GRID_COLS; ix++)
        for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) {
            A.cellStart[ix * FRAME_GRID_ROWS + iy] = (int32_t)A.cellItems.size();
            const std::vector<std::size_t>& cell = F.mGrid[ix][iy];
            for (size_t k = 0; k < cell.size(); k++) A.cellItems.push_back((int32_t)cell[k]);
        }
    A.cellStart[FRAME_GRID_COLS * FRAME_GRID_ROWS] = (int32_t)A.cellItems.size();
    if (A.cellItems.empty()) A.cellItems.push_back(0);
    A.Bind(Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv, F.mvScaleFactors);
}

// KeyFrame::mGrid is protected (include/KeyFrame.h:353).  It is a copy of the Frame's grid (src/KeyFrame.cc:53-59), which
// Frame::AssignFeaturesToGrid fills in keypoint order with the Frame's FLOAT bounds (Frame::mnMinX / mnMinY, public statics,
// src/Frame.cc:654): rebuilding it from mvKeysUn with orbm_assign_grid (the same PosInGrid arithmetic, parity-tested against
// the reference) and those float bounds gives the same cells in the same order.  KeyFrame::mnMinX / mnMinY are the bounds
// truncated to int (include/KeyFrame.h:331-332) and only the window arithmetic of KeyFrame::GetFeaturesInArea uses them
// (src/KeyFrame.cc:630-640): the view is bound to those.  The two differ on a distorted camera (non-integral bounds).
bool FromKeyFrame(orbm_context* ctx, const KeyFrame* pKF, Features& A)
{
    A.Keypoints(pKF->mvKeysUn, pKF->mvuRight, pKF->mDescriptors);
    A.cellStart.assign(FRAME_GRID_COLS * FRAME_GRID_ROWS + 1, 0);
    A.cellItems.assign(A.v.n > 0 ? A.v.n : 1, 0);
    if (A.v.n > 0 &&
        Failed(orbm_assign_grid(ctx, A.x.data(), A.y.data(), A.v.n, Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv,
                                Frame::mfGridElementHeightInv, A.cellStart.data(), A.cellItems.data(), 0), "orbm_assign_grid"))
        return false;
    A.Bind((float)pKF->mnMinX, (float)pKF->mnMinY, pKF->mfGridElementWidthInv, pKF->mfGridElementHeightInv, pKF->mvScaleFactors);
    return true;
}

// ---- the projected side: one entry per source map point ----------------------------------------------------------
struct Queries {
    std::vector<float> x, y, radius, xr, angle;
    std::vector<int32_t> minLevel, maxLevel;
    std::vector<uint8_t> valid, blocks, desc;
    explicit Queries(int n)
        : x(n, 0.f), y(n, 0.f), radius(n, 0.f), xr(n, -1.f), angle(n, 0.f), minLevel(n, -1), maxLevel(n, -1), valid(n, 0), blocks(n, 1),
          desc((size_t)(n > 0 ? n : 1) * 32, 0)
    {
    }
    void Descriptor(int i, const cv::Mat& d) { std::memcpy(&desc[(size_t)i * 32], d.ptr(), 32); }
    orbm_window_query_view View(bool withRight, bool withAngle) const
    {
        orbm_window_query_view q;
        q.n = (int)x.size();
        q.x = x.data(); q.y = y.data(); q.radius = radius.data(); q.min_level = minLevel.data(); q.max_level = maxLevel.data();
        q.xr = withRight ? xr.data() : NULL;
        q.angle = withAngle ? angle.data() : NULL;
        q.valid = valid.data(); q.claims_block = blocks.data(); q.desc = desc.data();
        return q;
    }
};

// camera coordinates of a world point, with the reference's cv::Mat arithmetic
struct Cam { float x, y, z; };
Cam ToCamera(const cv::Mat& Rcw, const cv::Mat& tcw, const cv::Mat& p3Dw)
{
    const cv::Mat p = Rcw * p3Dw + tcw;
    Cam c = {p.at<float>(0), p.at<float>(1), p.at<float>(2)};
    return c;
}

// Sim3 [sR | st] -> R, t of the unscaled pose and the camera centre (src/ORBmatcher.cc:406-412, :1234-1239)
void SplitSim3(const cv::Mat& Scw, cv::Mat& Rcw, cv::Mat& tcw, cv::Mat& Ow)
{
    const cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
    const float scw = std::sqrt(sRcw.row(0).dot(sRcw.row(0)));
    Rcw = sRcw / scw;
    tcw = Scw.rowRange(0, 3).col(3) / scw;
    Ow = -Rcw.t() * tcw;
}

// The scale-invariance and viewing-angle gates shared by Fuse and the Sim3 projection search, and the window they lead to.
// Returns false when the reference skips the point.
bool WindowInKeyFrame(MapPoint* pMP, const KeyFrame* pKF, const cv::Mat& p3Dw, const cv::Mat& Ow, float th, bool checkNormal, Queries& Q,
                      int i)
{
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    const cv::Mat PO = p3Dw - Ow;
    const float dist3D = cv::norm(PO);
    if (dist3D < minDistance || dist3D > maxDistance) return false;
    if (checkNormal) {
        const cv::Mat Pn = pMP->GetNormal();
        if (PO.dot(Pn) < 0.5 * dist3D) return false;
    }
    const int nPredictedLevel = pMP->PredictScale(dist3D, const_cast<KeyFrame*>(pKF));
    Q.radius[i] = th * pKF->mvScaleFactors[nPredictedLevel];
    Q.minLevel[i] = nPredictedLevel - 1;
    Q.maxLevel[i] = nPredictedLevel;
    return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b)
{
    return orbm_descriptor_distance(a.ptr(), b.ptr());
}

float ORBmatcher::RadiusByViewingCos(const float& viewCos)
{
    return viewCos > 0.998 ? 2.5f : 4.0f;
}

bool ORBmatcher::CheckDistEpipolarLine(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& F12, const KeyFrame* pKF2)
{
    // line l = x1' F12 = (a, b, c) in image 2; squared distance of kp2 to it against the chi-square bound of its level
    float l[3];
    for (int k = 0; k < 3; k++) l[k] = kp1.pt.x * F12.at<float>(0, k) + kp1.pt.y * F12.at<float>(1, k) + F12.at<float>(2, k);
    const float num = l[0] * kp2.pt.x + l[1] * kp2.pt.y + l[2];
    const float den = l[0] * l[0] + l[1] * l[1];
    if (den == 0) return false;
    return num * num / den < 3.84 * pKF2->mvLevelSigma2[kp2.octave];
}

void ORBmatcher::ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3)
{
    int top[3] = {0, 0, 0};
    int* ind[3] = {&ind1, &ind2, &ind3};
    ind1 = ind2 = ind3 = -1;
    for (int b = 0; b < L; b++) {
        const int s = (int)histo[b].size();
        int k = 3;
        while (k > 0 && s > top[k - 1]) k--;  // strictly larger moves up: the first of equal bins keeps its rank
        if (k == 3) continue;
        for (int j = 2; j > k; j--) { top[j] = top[j - 1]; *ind[j] = *ind[j - 1]; }
        top[k] = s; *ind[k] = b;
    }
    if (top[1] < 0.1f * (float)top[0]) { ind2 = -1; ind3 = -1; }
    else if (top[2] < 0.1f * (float)top[0]) ind3 = -1;
}

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th)
{
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const int N = F.N;
    const int M = (int)vpMapPoints.size();
    Features A;
    FromFrame(F, A);
    for (int i = 0; i < N; i++) {
        MapPoint* p = F.mvpMapPoints[i];
        A.occupied[i] = (p && p->Observations() > 0) ? 1 : 0;  // :117-119
    }
    std::vector<float> px(M), py(M), pxr(M), pcos(M);
    std::vector<int32_t> plevel(M);
    std::vector<uint8_t> pvalid(M), pobs(M), pdesc((size_t)(M > 0 ? M : 1) * 32);
    for (int m = 0; m < M; m++) {
        MapPoint* p = vpMapPoints[m];
        pvalid[m] = (p->mbTrackInView && !p->isBad()) ? 1 : 0;  // :71-75
        px[m] = p->mTrackProjX; py[m] = p->mTrackProjY; pxr[m] = p->mTrackProjXR;
        plevel[m] = p->mnTrackScaleLevel; pcos[m] = p->mTrackViewCos;
        pobs[m] = p->Observations() > 0 ? 1 : 0;
        if (pvalid[m]) {
            const cv::Mat d = p->GetDescriptor();  // :100
            std::memcpy(&pdesc[(size_t)m * 32], d.ptr(), 32);
        }
    }
    orbm_mappoint_view mv;
    mv.n = M;
    mv.proj_x = px.data(); mv.proj_y = py.data(); mv.proj_xr = pxr.data(); mv.level = plevel.data();
    mv.view_cos = pcos.data(); mv.valid = pvalid.data(); mv.obs_pos = pobs.data(); mv.desc = pdesc.data();
    std::vector<int32_t> match(N > 0 ? N : 1, -1);
    int nmatches = 0;
    if (Failed(orbm_search_by_projection(ctx, &A.v, &mv, th, mfNNratio, match.data(), &nmatches), "SearchByProjection")) return 0;
    for (int i = 0; i < N; i++)
        if (match[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[match[i]];  // :167
    return nmatches;
}

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches)
{
    const std::vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const int nkf = (int)vpMapPointsKF.size();
    std::vector<uint8_t> kfDesc, fDesc, hasMp(nkf > 0 ? nkf : 1);
    std::vector<float> kfAngle(nkf > 0 ? nkf : 1), fAngle(F.N > 0 ? F.N : 1);
    CopyDescriptors(pKF->mDescriptors, kfDesc);
    CopyDescriptors(F.mDescriptors, fDesc);
    for (int i = 0; i < nkf; i++) {
        MapPoint* p = vpMapPointsKF[i];
        hasMp[i] = (p && !p->isBad()) ? 1 : 0;  // :284-289
        kfAngle[i] = pKF->mvKeysUn[i].angle;
    }
    for (int i = 0; i < F.N; i++) fAngle[i] = F.mvKeys[i].angle;
    std::vector<int32_t> kId, kStart, kFeat, fId, fStart, fFeat;
    FlattenFeatureVector(pKF->mFeatVec, kId, kStart, kFeat);
    FlattenFeatureVector(F.mFeatVec, fId, fStart, fFeat);
    orbm_featvec_view kv = {(int)kId.size(), kId.data(), kStart.data(), kFeat.data()};
    orbm_featvec_view fvv = {(int)fId.size(), fId.data(), fStart.data(), fFeat.data()};
    std::vector<int32_t> match(F.N > 0 ? F.N : 1, -1);
    int nmatches = 0;
    if (Failed(orbm_search_by_bow(ctx, nkf, kfDesc.data(), kfAngle.data(), hasMp.data(), &kv, F.N, fDesc.data(), fAngle.data(), &fvv,
                                  mfNNratio, mbCheckOrientation ? 1 : 0, match.data(), &nmatches), "SearchByBoW"))
        return 0;
    for (int i = 0; i < F.N; i++)
        if (match[i] >= 0) vpMapPointMatches[i] = vpMapPointsKF[match[i]];
    return nmatches;
}

int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12)
{
    const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
    const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
    const int n1 = (int)vpMapPoints1.size(), n2 = (int)vpMapPoints2.size();
    vpMatches12 = std::vector<MapPoint*>(n1, static_cast<MapPoint*>(NULL));
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    std::vector<uint8_t> d1, d2, has1(n1 > 0 ? n1 : 1), has2(n2 > 0 ? n2 : 1);
    std::vector<float> a1(n1 > 0 ? n1 : 1), a2(n2 > 0 ? n2 : 1);
    CopyDescriptors(pKF1->mDescriptors, d1);
    CopyDescriptors(pKF2->mDescriptors, d2);
    for (int i = 0; i < n1; i++) {
        MapPoint* p = vpMapPoints1[i];
        has1[i] = (p && !p->isBad()) ? 1 : 0;
        a1[i] = pKF1->mvKeysUn[i].angle;
    }
    for (int i = 0; i < n2; i++) {
        MapPoint* p = vpMapPoints2[i];
        has2[i] = (p && !p->isBad()) ? 1 : 0;
        a2[i] = pKF2->mvKeysUn[i].angle;
    }
    std::vector<int32_t> id1, st1, ft1, id2, st2, ft2;
    FlattenFeatureVector(pKF1->mFeatVec, id1, st1, ft1);
    FlattenFeatureVector(pKF2->mFeatVec, id2, st2, ft2);
    orbm_featvec_view v1 = {(int)id1.size(), id1.data(), st1.data(), ft1.data()};
    orbm_featvec_view v2 = {(int)id2.size(), id2.data(), st2.data(), ft2.data()};
    std::vector<int32_t> match(n1 > 0 ? n1 : 1, -1);
    int nmatches = 0;
    if (Failed(orbm_search_by_bow_kf(ctx, n1, d1.data(), a1.data(), has1.data(), &v1, n2, d2.data(), a2.data(), has2.data(), &v2, mfNNratio,
                                     mbCheckOrientation ? 1 : 0, match.data(), &nmatches), "SearchByBoW(KF, KF)"))
        return 0;
    for (int i = 0; i < n1; i++)
        if (match[i] >= 0) vpMatches12[i] = vpMapPoints2[match[i]];
    return nmatches;
}

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, std::vector<MapPoint*>& vpMatched,
                                   int th)
{
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
    cv::Mat Rcw, tcw, Ow;
    SplitSim3(Scw, Rcw, tcw, Ow);
    std::set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(static_cast<MapPoint*>(NULL));

    Features A;
    if (!FromKeyFrame(ctx, pKF, A)) return 0;
    for (int i = 0; i < A.v.n; i++) A.occupied[i] = vpMatched[i] ? 1 : 0;  // :490
    const int M = (int)vpPoints.size();
    Queries Q(M);
    for (int i = 0; i < M; i++) {
        MapPoint* pMP = vpPoints[i];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
        const cv::Mat p3Dw = pMP->GetWorldPos();
        const Cam c = ToCamera(Rcw, tcw, p3Dw);
        if (c.z < 0.0) continue;
        const float invz = 1 / c.z;
        const float x = c.x * invz, y = c.y * invz;
        const float u = fx * x + cx, v = fy * y + cy;
        if (!pKF->IsInImage(u, v)) continue;
        if (!WindowInKeyFrame(pMP, pKF, p3Dw, Ow, (float)th, true, Q, i)) continue;
        Q.x[i] = u; Q.y[i] = v;
        Q.Descriptor(i, pMP->GetDescriptor());
        Q.valid[i] = 1;
    }
    const orbm_window_query_view qv = Q.View(false, false);
    std::vector<int32_t> match(A.v.n > 0 ? A.v.n : 1, -1);
    int nmatches = 0;
    if (Failed(orbm_search_window(ctx, &A.v, NULL, &qv, TH_LOW, 0, mfNNratio, 0, match.data(), &nmatches), "SearchByProjection(KF, Scw)"))
        return 0;
    for (int i = 0; i < A.v.n; i++)
        if (match[i] >= 0) vpMatched[i] = vpPoints[match[i]];
    return nmatches;
}

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                        int windowSize)
{
    const int n1 = (int)F1.mvKeysUn.size();
    vnMatches12 = std::vector<int>(n1, -1);
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    Features A2;
    FromFrame(F2, A2);
    std::vector<int32_t> oct1(n1 > 0 ? n1 : 1);
    std::vector<float> ang1(n1 > 0 ? n1 : 1), px(n1 > 0 ? n1 : 1), py(n1 > 0 ? n1 : 1);
    std::vector<uint8_t> d1;
    CopyDescriptors(F1.mDescriptors, d1);
    for (int i = 0; i < n1; i++) {
        oct1[i] = F1.mvKeysUn[i].octave;
        ang1[i] = F1.mvKeysUn[i].angle;
        px[i] = vbPrevMatched[i].x; py[i] = vbPrevMatched[i].y;
    }
    std::vector<int32_t> match(n1 > 0 ? n1 : 1, -1);
    int nmatches = 0;
    if (Failed(orbm_search_for_initialization(ctx, n1, oct1.data(), ang1.data(), d1.data(), px.data(), py.data(), &A2.v, A2.angle.data(),
                                              windowSize, mfNNratio, mbCheckOrientation ? 1 : 0, match.data(), &nmatches),
               "SearchForInitialization"))
        return 0;
    for (int i = 0; i < n1; i++) {
        vnMatches12[i] = match[i];
        if (match[i] >= 0) vbPrevMatched[i] = F2.mvKeysUn[match[i]].pt;  // :677-680
    }
    return nmatches;
}

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, std::vector<pair<size_t, size_t> >& vMatchedPairs,
                                       const bool bOnlyStereo)
{
    orbm_context* ctx = ThreadContext();
    vMatchedPairs.clear();
    if (!ctx) return 0;
    // epipole: camera centre of KF1 seen from KF2 (:858-865)
    const cv::Mat Cw = pKF1->GetCameraCenter();
    const Cam C2 = ToCamera(pKF2->GetRotation(), pKF2->GetTranslation(), Cw);
    const float invz = 1.0f / C2.z;
    const float ex = pKF2->fx * C2.x * invz + pKF2->cx;
    const float ey = pKF2->fy * C2.y * invz + pKF2->cy;

    struct Side {
        std::vector<float> x, y, angle, uRight;
        std::vector<int32_t> octave, id, start, feat;
        std::vector<uint8_t> desc, hasMp;
        orbm_featvec_view fv;
        orbm_kf_view v;
        void Fill(KeyFrame* kf)
        {
            const int n = kf->N;
            x.resize(n); y.resize(n); angle.resize(n); octave.resize(n); hasMp.assign(n > 0 ? n : 1, 0);
            for (int i = 0; i < n; i++) {
                const cv::KeyPoint& kp = kf->mvKeysUn[i];
                x[i] = kp.pt.x; y[i] = kp.pt.y; angle[i] = kp.angle; octave[i] = kp.octave;
                hasMp[i] = kf->GetMapPoint(i) ? 1 : 0;  // :894-898, :916-918
            }
            uRight.assign(kf->mvuRight.begin(), kf->mvuRight.end());
            CopyDescriptors(kf->mDescriptors, desc);
            FlattenFeatureVector(kf->mFeatVec, id, start, feat);
            fv.n_nodes = (int)id.size(); fv.node_id = id.data(); fv.node_start = start.data(); fv.node_feat = feat.data();
            v.n = n; v.x = x.data(); v.y = y.data(); v.octave = octave.data(); v.angle = angle.data(); v.u_right = uRight.data();
            v.desc = desc.data(); v.has_mp = hasMp.data(); v.fv = &fv;
        }
    } s1, s2;
    s1.Fill(pKF1);
    s2.Fill(pKF2);
    float F[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) F[3 * r + c] = F12.at<float>(r, c);
    std::vector<int32_t> match(pKF1->N > 0 ? pKF1->N : 1, -1);
    int nmatches = 0;
    if (Failed(orbm_search_for_triangulation(ctx, &s1.v, &s2.v, F, ex, ey, pKF2->mvScaleFactors.data(), pKF2->mvLevelSigma2.data(),
                                             (int)pKF2->mvScaleFactors.size(), bOnlyStereo ? 1 : 0, mbCheckOrientation ? 1 : 0,
                                             match.data(), &nmatches), "SearchForTriangulation"))
        return 0;
    vMatchedPairs.reserve(nmatches);
    for (int i = 0; i < pKF1->N; i++)
        if (match[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)match[i]));
    return nmatches;
}

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th)
{
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const cv::Mat Rcw = pKF->GetRotation();
    const cv::Mat tcw = pKF->GetTranslation();
    const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy, bf = pKF->mbf;
    const cv::Mat Ow = pKF->GetCameraCenter();
    Features A;
    if (!FromKeyFrame(ctx, pKF, A)) return 0;
    const int M = (int)vpMapPoints.size();
    Queries Q(M);
    for (int i = 0; i < M; i++) {
        MapPoint* pMP = vpMapPoints[i];
        if (!pMP) continue;
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
        const cv::Mat p3Dw = pMP->GetWorldPos();
        const Cam c = ToCamera(Rcw, tcw, p3Dw);
        if (c.z < 0.0f) continue;
        const float invz = 1 / c.z;
        const float x = c.x * invz, y = c.y * invz;
        const float u = fx * x + cx, v = fy * y + cy;
        if (!pKF->IsInImage(u, v)) continue;
        if (!WindowInKeyFrame(pMP, pKF, p3Dw, Ow, th, true, Q, i)) continue;
        Q.x[i] = u; Q.y[i] = v;
        Q.xr[i] = u - bf * invz;
        Q.Descriptor(i, pMP->GetDescriptor());
        Q.valid[i] = 1;
    }
    const orbm_window_query_view qv = Q.View(true, false);
    std::vector<int32_t> best(M > 0 ? M : 1, -1);
    if (Failed(orbm_search_best(ctx, &A.v, &qv, 1, pKF->mvInvLevelSigma2.data(), TH_LOW, best.data(), NULL), "Fuse")) return 0;
    // apply in the reference's order (:1198-1217).  The search does not read map-point state, and an earlier iteration
    // can only change the skip test of a later one when the same point is listed twice: re-evaluate it here.
    int nFused = 0;
    for (int i = 0; i < M; i++) {
        if (best[i] < 0) continue;
        MapPoint* pMP = vpMapPoints[i];
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
        MapPoint* pMPinKF = pKF->GetMapPoint(best[i]);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) {
                if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                else pMPinKF->Replace(pMP);
            }
        } else {
            pMP->AddObservation(pKF, best[i]);
            pKF->AddMapPoint(pMP, best[i]);
        }
        nFused++;
    }
    return nFused;
}

int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th, vector<MapPoint*>& vpReplacePoint)
{
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
    cv::Mat Rcw, tcw, Ow;
    SplitSim3(Scw, Rcw, tcw, Ow);
    const std::set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();
    Features A;
    if (!FromKeyFrame(ctx, pKF, A)) return 0;
    const int M = (int)vpPoints.size();
    Queries Q(M);
    for (int i = 0; i < M; i++) {
        MapPoint* pMP = vpPoints[i];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
        const cv::Mat p3Dw = pMP->GetWorldPos();
        const Cam c = ToCamera(Rcw, tcw, p3Dw);
        if (c.z < 0.0f) continue;
        const float invz = 1.0 / c.z;
        const float x = c.x * invz, y = c.y * invz;
        const float u = fx * x + cx, v = fy * y + cy;
        if (!pKF->IsInImage(u, v)) continue;
        if (!WindowInKeyFrame(pMP, pKF, p3Dw, Ow, th, true, Q, i)) continue;
        Q.x[i] = u; Q.y[i] = v;
        Q.Descriptor(i, pMP->GetDescriptor());
        Q.valid[i] = 1;
    }
    const orbm_window_query_view qv = Q.View(false, false);
    std::vector<int32_t> best(M > 0 ? M : 1, -1);
    if (Failed(orbm_search_best(ctx, &A.v, &qv, 0, pKF->mvInvLevelSigma2.data(), TH_LOW, best.data(), NULL), "Fuse(Scw)")) return 0;
    int nFused = 0;
    for (int i = 0; i < M; i++) {
        if (best[i] < 0) continue;
        MapPoint* pMP = vpPoints[i];
        MapPoint* pMPinKF = pKF->GetMapPoint(best[i]);  // sees the points added by earlier iterations, as in the reference
        if (pMPinKF) {
            if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
        } else {
            pMP->AddObservation(pKF, best[i]);
            pKF->AddMapPoint(pMP, best[i]);
        }
        nFused++;
    }
    return nFused;
}

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                             const cv::Mat& t12, const float th)
{
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const float fx = pKF1->fx, fy = pKF1->fy, cx = pKF1->cx, cy = pKF1->cy;
    const cv::Mat R1w = pKF1->GetRotation(), t1w = pKF1->GetTranslation();
    const cv::Mat R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
    const cv::Mat sR12 = s12 * R12;
    const cv::Mat sR21 = (1.0 / s12) * R12.t();
    const cv::Mat t21 = -sR21 * t12;
    const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
    const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {
        MapPoint* pMP = vpMatches12[i];
        if (!pMP) continue;
        vbAlreadyMatched1[i] = true;
        const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
        if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
    }
    Features A1, A2;
    if (!FromKeyFrame(ctx, pKF1, A1) || !FromKeyFrame(ctx, pKF2, A2)) return 0;

    // one direction: the map points of `from` moved into the camera of `to` by (sR, t) after their own pose (Rw, tw)
    struct Direction {
        static void Fill(const std::vector<MapPoint*>& pts, const std::vector<bool>& skip, const cv::Mat& Rw, const cv::Mat& tw,
                         const cv::Mat& sR, const cv::Mat& t, const KeyFrame* to, float fx, float fy, float cx, float cy, float th,
                         Queries& Q)
        {
            for (int i = 0; i < (int)pts.size(); i++) {
                MapPoint* pMP = pts[i];
                if (!pMP || skip[i]) continue;
                if (pMP->isBad()) continue;
                const cv::Mat p3Dw = pMP->GetWorldPos();
                const cv::Mat own = Rw * p3Dw + tw;
                const cv::Mat other = sR * own + t;
                if (other.at<float>(2) < 0.0) continue;
                const float invz = 1.0 / other.at<float>(2);
                const float x = other.at<float>(0) * invz, y = other.at<float>(1) * invz;
                const float u = fx * x + cx, v = fy * y + cy;
                if (!to->IsInImage(u, v)) continue;
                const float maxDistance = pMP->GetMaxDistanceInvariance();
                const float minDistance = pMP->GetMinDistanceInvariance();
                const float dist3D = cv::norm(other);
                if (dist3D < minDistance || dist3D > maxDistance) continue;
                const int nPredictedLevel = pMP->PredictScale(dist3D, const_cast<KeyFrame*>(to));
                Q.radius[i] = th * to->mvScaleFactors[nPredictedLevel];
                Q.minLevel[i] = nPredictedLevel - 1; Q.maxLevel[i] = nPredictedLevel;
                Q.x[i] = u; Q.y[i] = v;
                Q.Descriptor(i, pMP->GetDescriptor());
                Q.valid[i] = 1;
            }
        }
    };
    Queries Q1in2(N1), Q2in1(N2);
    Direction::Fill(vpMapPoints1, vbAlreadyMatched1, R1w, t1w, sR21, t21, pKF2, fx, fy, cx, cy, th, Q1in2);
    Direction::Fill(vpMapPoints2, vbAlreadyMatched2, R2w, t2w, sR12, t12, pKF1, fx, fy, cx, cy, th, Q2in1);
    const orbm_window_query_view q12 = Q1in2.View(false, false), q21 = Q2in1.View(false, false);
    std::vector<int32_t> match(N1 > 0 ? N1 : 1, -1);
    int nFound = 0;
    if (Failed(orbm_search_by_sim3(ctx, &A1.v, &q21, &A2.v, &q12, match.data(), &nFound), "SearchBySim3")) return 0;
    for (int i1 = 0; i1 < N1; i1++)
        if (match[i1] >= 0) vpMatches12[i1] = vpMapPoints2[match[i1]];
    return nFound;
}

// ------------------------------------------------------------------------------------------------------------------
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono)
{
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat twc = -Rcw.t() * tcw;
    const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat tlc = Rlw * twc + tlw;
    const bool bForward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
    const bool bBackward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;

    Features A;
    FromFrame(CurrentFrame, A);
    for (int i = 0; i < CurrentFrame.N; i++) {
        MapPoint* p = CurrentFrame.mvpMapPoints[i];
        A.occupied[i] = (p && p->Observations() > 0) ? 1 : 0;  // :1707-1709
    }
    const int M = LastFrame.N;
    Queries Q(M);
    for (int i = 0; i < M; i++) {
        MapPoint* pMP = LastFrame.mvpMapPoints[i];
        if (!pMP || LastFrame.mvbOutlier[i]) continue;
        const Cam c = ToCamera(Rcw, tcw, pMP->GetWorldPos());
        const float invzc = 1.0 / c.z;
        if (invzc < 0) continue;
        const float u = CurrentFrame.fx * c.x * invzc + CurrentFrame.cx;
        const float v = CurrentFrame.fy * c.y * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        const int nLastOctave = LastFrame.mvKeys[i].octave;
        Q.radius[i] = th * CurrentFrame.mvScaleFactors[nLastOctave];
        if (bForward) { Q.minLevel[i] = nLastOctave; Q.maxLevel[i] = -1; }
        else if (bBackward) { Q.minLevel[i] = 0; Q.maxLevel[i] = nLastOctave; }
        else { Q.minLevel[i] = nLastOctave - 1; Q.maxLevel[i] = nLastOctave + 1; }
        Q.x[i] = u; Q.y[i] = v;
        Q.xr[i] = u - CurrentFrame.mbf * invzc;
        Q.angle[i] = LastFrame.mvKeysUn[i].angle;
        Q.blocks[i] = pMP->Observations() > 0 ? 1 : 0;  // a slot given to this point stops later points only then
        Q.Descriptor(i, pMP->GetDescriptor());
        Q.valid[i] = 1;
    }
    const orbm_window_query_view qv = Q.View(true, true);
    std::vector<int32_t> match(CurrentFrame.N > 0 ? CurrentFrame.N : 1, -1);
    int nmatches = 0;
    if (Failed(orbm_search_window(ctx, &A.v, A.angle.data(), &qv, TH_HIGH, 0, mfNNratio, mbCheckOrientation ? 1 : 0, match.data(), &nmatches),
               "SearchByProjection(Current, Last)"))
        return 0;
    for (int i = 0; i < CurrentFrame.N; i++) {
        if (match[i] >= 0) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[match[i]];
        else if (match[i] == -2) CurrentFrame.mvpMapPoints[i] = static_cast<MapPoint*>(NULL);  // rotation-histogram reject (:1773-1776)
    }
    return nmatches;
}

int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                                   const int ORBdist)
{
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat Ow = -Rcw.t() * tcw;
    Features A;
    FromFrame(CurrentFrame, A);
    for (int i = 0; i < CurrentFrame.N; i++) A.occupied[i] = CurrentFrame.mvpMapPoints[i] ? 1 : 0;  // :1869-1870
    const std::vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
    const int M = (int)vpMPs.size();
    Queries Q(M);
    for (int i = 0; i < M; i++) {
        MapPoint* pMP = vpMPs[i];
        if (!pMP) continue;
        if (pMP->isBad() || sAlreadyFound.count(pMP)) continue;
        const cv::Mat x3Dw = pMP->GetWorldPos();
        const Cam c = ToCamera(Rcw, tcw, x3Dw);
        const float invzc = 1.0 / c.z;
        const float u = CurrentFrame.fx * c.x * invzc + CurrentFrame.cx;
        const float v = CurrentFrame.fy * c.y * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        const cv::Mat PO = x3Dw - Ow;
        const float dist3D = cv::norm(PO);
        const float maxDistance = pMP->GetMaxDistanceInvariance();
        const float minDistance = pMP->GetMinDistanceInvariance();
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const int nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame);
        Q.radius[i] = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
        Q.minLevel[i] = nPredictedLevel - 1; Q.maxLevel[i] = nPredictedLevel + 1;
        Q.x[i] = u; Q.y[i] = v;
        Q.angle[i] = pKF->mvKeysUn[i].angle;
        Q.Descriptor(i, pMP->GetDescriptor());
        Q.valid[i] = 1;
    }
    const orbm_window_query_view qv = Q.View(false, true);
    std::vector<int32_t> match(CurrentFrame.N > 0 ? CurrentFrame.N : 1, -1);
    int nmatches = 0;
    if (Failed(orbm_search_window(ctx, &A.v, A.angle.data(), &qv, ORBdist, 0, mfNNratio, mbCheckOrientation ? 1 : 0, match.data(), &nmatches),
               "SearchByProjection(Current, KF)"))
        return 0;
    for (int i = 0; i < CurrentFrame.N; i++) {
        if (match[i] >= 0) CurrentFrame.mvpMapPoints[i] = vpMPs[match[i]];
        else if (match[i] == -2) CurrentFrame.mvpMapPoints[i] = static_cast<MapPoint*>(NULL);
    }
    return nmatches;
}

}  // namespace ORB_SLAM2

namespace orbx_adapter {

// Coarse stereo search for Frame::ComputeStereoMatches: replaces the row-table construction and the per-left-keypoint
// Hamming loop (src/Frame.cc:684-765).  vBestIdxR[iL] = right keypoint index or -1, vBestDist[iL] = its distance.
// The caller continues with the reference's SAD refinement (src/Frame.cc:767-866) for every iL with vBestIdxR >= 0.
int StereoCoarseSearch(const std::vector<cv::KeyPoint>& vKeysLeft, const cv::Mat& descLeft,
                       const std::vector<cv::KeyPoint>& vKeysRight, const cv::Mat& descRight,
                       const std::vector<float>& vScaleFactors, int nRows, float minD, float maxD,
                       std::vector<int>& vBestIdxR, std::vector<int>& vBestDist)
{
    static thread_local orbm_context* ctx = NULL;
    if (!ctx && orbm_create(std::getenv("ORBX_DEVICE") ? std::atoi(std::getenv("ORBX_DEVICE")) : 0, &ctx) != ORBX_OK) return -1;
    const int nl = (int)vKeysLeft.size(), nr = (int)vKeysRight.size();
    std::vector<float> lx(nl), ly(nl), rx(nr), ry(nr);
    std::vector<int32_t> lo(nl), ro(nr);
    std::vector<uint8_t> ld((size_t)nl * 32), rd((size_t)nr * 32);
    for (int i = 0; i < nl; i++) {
        lx[i] = vKeysLeft[i].pt.x; ly[i] = vKeysLeft[i].pt.y; lo[i] = vKeysLeft[i].octave;
        std::memcpy(&ld[(size_t)i * 32], descLeft.ptr(i), 32);
    }
    for (int i = 0; i < nr; i++) {
        rx[i] = vKeysRight[i].pt.x; ry[i] = vKeysRight[i].pt.y; ro[i] = vKeysRight[i].octave;
        std::memcpy(&rd[(size_t)i * 32], descRight.ptr(i), 32);
    }
    vBestIdxR.assign(nl, -1);
    vBestDist.assign(nl, 100);
    std::vector<int32_t> br(nl > 0 ? nl : 1), bd(nl > 0 ? nl : 1);
    const int rc = orbm_stereo_rows(ctx, nl, lx.data(), ly.data(), lo.data(), ld.data(), nr, rx.data(), ry.data(), ro.data(),
                                    rd.data(), vScaleFactors.data(), (int)vScaleFactors.size(), nRows, minD, maxD, br.data(),
                                    bd.data());
    if (rc != ORBX_OK) {
        std::fprintf(stderr, "StereoCoarseSearch: %s\n", orbx_last_error());
        return rc;
    }
    for (int i = 0; i < nl; i++) { vBestIdxR[i] = br[i]; vBestDist[i] = bd[i]; }
    return 0;
}

// The whole of Frame::ComputeStereoMatches on the GPU: coarse search + SAD refinement on the two extractors' device
// pyramids + outlier removal.  Body replacement for src/Frame.cc:673-885:
//     orbx_adapter::ComputeStereoMatches(mpORBextractorLeft, mpORBextractorRight, mvKeys, mDescriptors, mvKeysRight,
//                                        mDescriptorsRight, mbf, mb, mvuRight, mvDepth);
int ComputeStereoMatches(ORB_SLAM2::ORBextractor* pLeft, ORB_SLAM2::ORBextractor* pRight, const std::vector<cv::KeyPoint>& vKeysLeft,
                         const cv::Mat& descLeft, const std::vector<cv::KeyPoint>& vKeysRight, const cv::Mat& descRight, float mbf,
                         float mb, std::vector<float>& vuRight, std::vector<float>& vDepth)
{
    static thread_local orbm_context* ctx = NULL;
    if (!ctx && orbm_create(std::getenv("ORBX_DEVICE") ? std::atoi(std::getenv("ORBX_DEVICE")) : 0, &ctx) != ORBX_OK) return -1;
    const int nl = (int)vKeysLeft.size(), nr = (int)vKeysRight.size();
    vuRight.assign(nl, -1.0f);
    vDepth.assign(nl, -1.0f);
    if (nl == 0 || !pLeft->NativeHandle() || !pRight->NativeHandle()) return 0;
    std::vector<float> lx(nl), ly(nl), rx(nr), ry(nr);
    std::vector<int32_t> lo(nl), ro(nr);
    std::vector<uint8_t> ld((size_t)nl * 32), rd((size_t)nr * 32);
    for (int i = 0; i < nl; i++) {
        lx[i] = vKeysLeft[i].pt.x; ly[i] = vKeysLeft[i].pt.y; lo[i] = vKeysLeft[i].octave;
        std::memcpy(&ld[(size_t)i * 32], descLeft.ptr(i), 32);
    }
    for (int i = 0; i < nr; i++) {
        rx[i] = vKeysRight[i].pt.x; ry[i] = vKeysRight[i].pt.y; ro[i] = vKeysRight[i].octave;
        std::memcpy(&rd[(size_t)i * 32], descRight.ptr(i), 32);
    }
    int nMatched = 0;
    const int rc = orbm_stereo_match(ctx, pLeft->NativeHandle(), 0, pRight->NativeHandle(), 0, nl, lx.data(), ly.data(), lo.data(),
                                     ld.data(), nr, rx.data(), ry.data(), ro.data(), rd.data(), mbf, mb, vuRight.data(), vDepth.data(),
                                     &nMatched);
    if (rc != ORBX_OK) {
        std::fprintf(stderr, "ComputeStereoMatches: %s\n", orbx_last_error());
        return rc;
    }
    return nMatched;
}

}  // namespace orbx_adapter
