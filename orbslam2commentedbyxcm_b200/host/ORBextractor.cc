// host/ORBextractor.cc -- marshalling layer between the reference's ORBextractor interface and liborbx.so.
// Replaces src/ORBextractor.cc of the reference in the ORB_SLAM2 library target.
#include "ORBextractor.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace ORB_SLAM2 {

static const int EDGE_THRESHOLD = ORBX_EDGE;

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mpHandle(NULL), mbDownloadPyramid(true), mpPyramidPinned(NULL), mnPyramidBytes(0)
{
    orbx_config cfg;
    cfg.nfeatures = _nfeatures;
    cfg.scale_factor = _scaleFactor;
    cfg.nlevels = _nlevels;
    cfg.ini_th_fast = _iniThFAST;
    cfg.min_th_fast = _minThFAST;
    const char* dev = std::getenv("ORBX_DEVICE");
    const int rc = orbx_create(&cfg, dev ? std::atoi(dev) : 0, &mpHandle);
    if (rc != ORBX_OK) {
        // the reference constructor cannot fail; keep the object alive and report (operator() will return nothing)
        std::fprintf(stderr, "ORBextractor: orbx_create failed (%d): %s\n", rc, orbx_last_error());
        mpHandle = NULL;
    }
    mvScaleFactor.resize(nlevels);
    mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    mvImagePyramid.resize(nlevels);
    mvPaddedLevels.resize(nlevels);
    if (mpHandle)
        orbx_get_tables(mpHandle, &mvScaleFactor[0], &mvInvScaleFactor[0], &mvLevelSigma2[0], &mvInvLevelSigma2[0],
                        &mnFeaturesPerLevel[0]);
}

ORBextractor::~ORBextractor()
{
    mvImagePyramid.clear();
    mvPaddedLevels.clear();
    orbx_free_pinned(mpPyramidPinned);
    if (mpHandle) orbx_destroy(mpHandle);
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                              cv::OutputArray _descriptors)
{
    if (_image.empty()) return;  // src/ORBextractor.cc:1517-1518
    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);

    _keypoints.clear();
    if (!mpHandle) {
        _descriptors.release();
        return;
    }
    const int cap = orbx_max_keypoints(mpHandle, image.rows, image.cols);
    if (cap < 0) {
        std::fprintf(stderr, "ORBextractor: %s\n", orbx_last_error());
        _descriptors.release();
        return;
    }
    if ((int)mvKpBuffer.size() < cap) {
        mvKpBuffer.resize(cap);
        mvDescBuffer.resize((size_t)cap * ORBX_DESC_BYTES);
    }
    int n = 0;
    const int rc = orbx_extract(mpHandle, image.data, image.rows, image.cols, (size_t)image.step, &mvKpBuffer[0],
                                &mvDescBuffer[0], cap, &n);
    if (rc != ORBX_OK) {
        std::fprintf(stderr, "ORBextractor: orbx_extract failed (%d): %s\n", rc, orbx_last_error());
        _descriptors.release();
        return;
    }
    if (n == 0) {
        _descriptors.release();  // src/ORBextractor.cc:1552
    } else {
        _descriptors.create(n, ORBX_DESC_BYTES, CV_8U);  // :1556
        cv::Mat descriptors = _descriptors.getMat();
        for (int i = 0; i < n; i++)
            std::memcpy(descriptors.ptr(i), &mvDescBuffer[(size_t)i * ORBX_DESC_BYTES], ORBX_DESC_BYTES);
    }
    _keypoints.reserve(n);
    for (int i = 0; i < n; i++) {
        const orbx_keypoint& k = mvKpBuffer[i];
        _keypoints.push_back(cv::KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, -1));
    }
    if (mbDownloadPyramid) {
        // one page-locked staging buffer holds all padded planes (64-byte aligned rows); one call, one wait on the stream
        std::vector<size_t> off(nlevels), step(nlevels);
        std::vector<int> rows(nlevels), cols(nlevels);
        size_t total = 0;
        for (int l = 0; l < nlevels; l++) {
            if (orbx_get_level_size(mpHandle, l, &rows[l], &cols[l]) != ORBX_OK) return;
            step[l] = ((size_t)cols[l] + 2 * EDGE_THRESHOLD + 63) / 64 * 64;
            off[l] = total;
            total += step[l] * (size_t)(rows[l] + 2 * EDGE_THRESHOLD);
        }
        if (total > mnPyramidBytes) {
            orbx_free_pinned(mpPyramidPinned);
            mpPyramidPinned = (unsigned char*)orbx_alloc_pinned(total);
            mnPyramidBytes = mpPyramidPinned ? total : 0;
        }
        if (!mpPyramidPinned || orbx_get_pyramid_all(mpHandle, 0, mpPyramidPinned, &off[0], &step[0]) != ORBX_OK) {
            std::fprintf(stderr, "ORBextractor: %s\n", orbx_last_error());
            return;
        }
        for (int l = 0; l < nlevels; l++) {
            mvPaddedLevels[l] = cv::Mat(rows[l] + 2 * EDGE_THRESHOLD, cols[l] + 2 * EDGE_THRESHOLD, CV_8UC1, mpPyramidPinned + off[l], step[l]);
            mvImagePyramid[l] = mvPaddedLevels[l](cv::Rect(EDGE_THRESHOLD, EDGE_THRESHOLD, cols[l], rows[l]));
        }
    }
}

}  // namespace ORB_SLAM2
