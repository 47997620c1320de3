// host/ORBextractor.h -- drop-in replacement for the reference's include/ORBextractor.h.
//
// Same namespace, class name, constructor, call operator, accessors and public mvImagePyramid member as
// ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:76-220), so Frame.cc / Tracking.cc compile and link against
// it unchanged.  All computation happens in liborbx.so (hand-written sm_100a CUDA kernels) through the C-ABI of
// include/orbx.h; this class only marshals cv::Mat / cv::KeyPoint to flat arrays.  No OpenCV-CUDA, no CPU fallback:
// if the GPU path fails the call reports the error on stderr and returns no keypoints.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <list>
#include <vector>
#include <opencv/cv.h>

#include "orbx.h"

namespace ORB_SLAM2 {

// Kept only because the reference header exposes it (include/ORBextractor.h:41-68); the quad-tree now lives in
// k_octree (csrc/orbx_extract.cu) and nothing in Frame / Tracking / LocalMapping touches this type.
class ExtractorNode {
public:
    ExtractorNode() : bNoMore(false) {}
    std::vector<cv::KeyPoint> vKeys;
    cv::Point2i UL, UR, BL, BR;
    std::list<ExtractorNode>::iterator lit;
    bool bNoMore;
};

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // include/ORBextractor.h:93 -- creates the device context on CUDA device ORBX_DEVICE (environment, default 0)
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    ~ORBextractor();

    // include/ORBextractor.h:111 / src/ORBextractor.cc:1513-1629.  `mask` is ignored, as in the reference.
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints,
                    cv::OutputArray descriptors);

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // include/ORBextractor.h:162.  Level l is an ROI at (19,19) of a padded host buffer, valid until the next call on this
    // object, exactly like the reference (Frame::ComputeStereoMatches reads patches from it, src/Frame.cc:782-831).
    std::vector<cv::Mat> mvImagePyramid;

    // Extension: monocular / RGB-D callers never read mvImagePyramid; switching the download off saves the largest D2H
    // transfer of the call (default on = reference behaviour).
    void SetPyramidDownload(bool on) { mbDownloadPyramid = on; }
    // Extension: the C-ABI handle, so that orbx_adapter::ComputeStereoMatches can read the device-resident pyramid.
    orbx_handle* NativeHandle() { return mpHandle; }

protected:
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

private:
    ORBextractor(const ORBextractor&);
    ORBextractor& operator=(const ORBextractor&);
    orbx_handle* mpHandle;
    bool mbDownloadPyramid;
    std::vector<orbx_keypoint> mvKpBuffer;
    std::vector<unsigned char> mvDescBuffer;
    std::vector<cv::Mat> mvPaddedLevels;  // headers over the padded pyramid planes in mpPyramidPinned
    unsigned char* mpPyramidPinned;       // page-locked staging buffer of all padded planes (valid until the next operator())
    size_t mnPyramidBytes;
};

}  // namespace ORB_SLAM2

#endif
