// tests/cpp/stubs/ORBmatcher.h -- DECLARATION-ONLY stand-ins so host/ORBmatcher_gpu.cc can be compile-checked in an image
// without OpenCV / DBoW2 / the ORB-SLAM2 tree.  Member names and types are the ones the reference declares in
// include/ORBmatcher.h:48-264, include/Frame.h:219-324, include/KeyFrame.h:186-317, include/MapPoint.h:78-154 and
// Thirdparty/DBoW2/DBoW2/FeatureVector.h:21-22.  Nothing here is product code.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>
#include <opencv/cv.h>  // oracle/shim mini-cv

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

namespace DBoW2 {
typedef unsigned int NodeId;
class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {};
}

namespace ORB_SLAM2 {
class MapPoint {
public:
    int Observations();
    bool isBad();
    cv::Mat GetDescriptor();
    float mTrackProjX, mTrackProjY, mTrackProjXR;
    int mnTrackScaleLevel;
    float mTrackViewCos;
    bool mbTrackInView;
};
class Frame {
public:
    int N;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    std::vector<float> mvuRight;
    DBoW2::FeatureVector mFeatVec;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<MapPoint*> mvpMapPoints;
    static float mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    std::vector<float> mvScaleFactors;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
};
class KeyFrame {
public:
    std::vector<MapPoint*> GetMapPointMatches();
    const std::vector<cv::KeyPoint> mvKeysUn;
    const cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
};
class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true);
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3);
    int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);
    static const int TH_LOW, TH_HIGH, HISTO_LENGTH;
protected:
    float mfNNratio;
    bool mbCheckOrientation;
};
}
