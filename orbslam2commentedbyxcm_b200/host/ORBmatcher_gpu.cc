// host/ORBmatcher_gpu.cc -- GPU definitions of the reference's Hamming searches, to be compiled INSIDE the ORB-SLAM2
// tree in place of the same-named definitions of src/ORBmatcher.cc.  The class declaration stays
// the reference's include/ORBmatcher.h (unchanged); this file only provides
//
//   ORBmatcher::SearchByProjection(Frame&, const std::vector<MapPoint*>&, const float th)   src/ORBmatcher.cc:61-173
//   ORBmatcher::SearchByBoW(KeyFrame*, Frame&, std::vector<MapPoint*>&)                     src/ORBmatcher.cc:228-392
//   ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&)                          src/ORBmatcher.cc:1983-2003
//   orbx_adapter::StereoCoarseSearch(...)   coarse part of Frame::ComputeStereoMatches      src/Frame.cc:684-765
//   orbx_adapter::ComputeStereoMatches(...) the whole of Frame::ComputeStereoMatches        src/Frame.cc:673-885
//
// Each function gathers flat views from Frame / KeyFrame / MapPoint (under the same locks the reference takes through
// the objects' accessors), calls the C-ABI of include/orbx.h, and applies the result to the objects in the reference's
// order.  It needs the reference's Frame.h / KeyFrame.h / MapPoint.h (and therefore OpenCV + DBoW2), which are not
// available in the authoring image: tests/cpp compiles it against declaration-only stubs of those headers.
#include "ORBmatcher.h"

#include <cstdio>
#include <vector>

#include "orbx.h"
#include "ORBextractor.h"

namespace ORB_SLAM2 {

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

}  // namespace

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b)
{
    return orbm_descriptor_distance(a.ptr(), b.ptr());
}

int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th)
{
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const int N = F.N;
    const int M = (int)vpMapPoints.size();
    // ---- frame view ----
    std::vector<float> kx(N), ky(N);
    std::vector<int32_t> koct(N);
    std::vector<uint8_t> occupied(N), kdesc;
    for (int i = 0; i < N; i++) {
        kx[i] = F.mvKeysUn[i].pt.x;
        ky[i] = F.mvKeysUn[i].pt.y;
        koct[i] = F.mvKeysUn[i].octave;
        MapPoint* p = F.mvpMapPoints[i];
        occupied[i] = (p && p->Observations() > 0) ? 1 : 0;  // src/ORBmatcher.cc:117-119
    }
    CopyDescriptors(F.mDescriptors, kdesc);
    std::vector<int32_t> cellStart(FRAME_GRID_COLS * FRAME_GRID_ROWS + 1), cellItems;
    cellItems.reserve(N);
    for (int ix = 0; ix < FRAME_GRID_COLS; ix++)
        for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) {
            cellStart[ix * FRAME_GRID_ROWS + iy] = (int32_t)cellItems.size();
            const std::vector<std::size_t>& cell = F.mGrid[ix][iy];
            for (size_t k = 0; k < cell.size(); k++) cellItems.push_back((int32_t)cell[k]);
        }
    cellStart[FRAME_GRID_COLS * FRAME_GRID_ROWS] = (int32_t)cellItems.size();
    orbm_frame_view fv;
    fv.n = N;
    fv.x = kx.data(); fv.y = ky.data(); fv.octave = koct.data();
    fv.u_right = F.mvuRight.empty() ? NULL : F.mvuRight.data();
    fv.desc = kdesc.data(); fv.occupied = occupied.data();
    fv.cell_start = cellStart.data(); fv.cell_items = cellItems.data();
    fv.min_x = Frame::mnMinX; fv.min_y = Frame::mnMinY;
    fv.grid_w_inv = Frame::mfGridElementWidthInv; fv.grid_h_inv = Frame::mfGridElementHeightInv;
    fv.scale_factors = F.mvScaleFactors.data(); fv.nlevels = (int)F.mvScaleFactors.size();
    // ---- map point view ----
    std::vector<float> px(M), py(M), pxr(M), pcos(M);
    std::vector<int32_t> plevel(M);
    std::vector<uint8_t> pvalid(M), pobs(M), pdesc((size_t)M * 32);
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
    const int rc = orbm_search_by_projection(ctx, &fv, &mv, th, mfNNratio, match.data(), &nmatches);
    if (rc != ORBX_OK) {
        std::fprintf(stderr, "ORBmatcher::SearchByProjection: %s\n", orbx_last_error());
        return 0;
    }
    for (int i = 0; i < N; i++)
        if (match[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[match[i]];  // :167
    return nmatches;
}

int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches)
{
    const std::vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
    orbm_context* ctx = ThreadContext();
    if (!ctx) return 0;
    const int nkf = (int)vpMapPointsKF.size();
    std::vector<uint8_t> kfDesc, fDesc, hasMp(nkf);
    std::vector<float> kfAngle(nkf), fAngle(F.N);
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
    const int rc = orbm_search_by_bow(ctx, nkf, kfDesc.data(), kfAngle.data(), hasMp.data(), &kv, F.N, fDesc.data(),
                                      fAngle.data(), &fvv, mfNNratio, mbCheckOrientation ? 1 : 0, match.data(), &nmatches);
    if (rc != ORBX_OK) {
        std::fprintf(stderr, "ORBmatcher::SearchByBoW: %s\n", orbx_last_error());
        return 0;
    }
    for (int i = 0; i < F.N; i++)
        if (match[i] >= 0) vpMapPointMatches[i] = vpMapPointsKF[match[i]];
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
//                                        mDescriptorsRight, mbf, mbf / fx, mvuRight, mvDepth);
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
