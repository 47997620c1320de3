// oracle/ref_wrap.cc -- TEST INFRASTRUCTURE.
// C entry points around the reference's own ORB_SLAM2::ORBextractor, compiled UNMODIFIED from
// /root/reference/src/ORBextractor.cc against oracle/shim (see oracle/Makefile, target `_ref`).
// Output goes to oracle/_ref/liborbref.so (git-ignored, travels to the GPU box).
// Used to (a) validate oracle/orb_oracle.cc and (b) optionally as the timed CPU baseline.
#include "ORBextractor.h"  // from /root/reference/include via -I
#include <cstdint>
#include <thread>
#include <vector>

struct RefKp { float x, y, size, angle, response; int octave; };

struct RefHandle {
    ORB_SLAM2::ORBextractor* ex;
    std::vector<cv::KeyPoint> kps;
    cv::Mat desc;
};

extern "C" {

void* orbref_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh)
{
    RefHandle* h = new RefHandle;
    h->ex = new ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    return h;
}

void orbref_destroy(void* p)
{
    RefHandle* h = (RefHandle*)p;
    delete h->ex;
    delete h;
}

int orbref_extract(void* p, const uint8_t* img, int rows, int cols, size_t step)
{
    RefHandle* h = (RefHandle*)p;
    cv::Mat im(rows, cols, CV_8UC1, (void*)img, step);
    cv::Mat mask;
    h->desc = cv::Mat();
    (*h->ex)(im, mask, h->kps, h->desc);
    return (int)h->kps.size();
}

void orbref_get_keypoints(void* p, RefKp* out)
{
    RefHandle* h = (RefHandle*)p;
    for (size_t i = 0; i < h->kps.size(); i++) {
        const cv::KeyPoint& k = h->kps[i];
        out[i] = RefKp{k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave};
    }
}

void orbref_get_descriptors(void* p, uint8_t* out)
{
    RefHandle* h = (RefHandle*)p;
    for (int i = 0; i < h->desc.rows; i++) std::memcpy(out + 32 * (size_t)i, h->desc.ptr(i), 32);
}

int orbref_get_levels(void* p) { return ((RefHandle*)p)->ex->GetLevels(); }

void orbref_get_scale_factors(void* p, float* out)
{
    std::vector<float> v = ((RefHandle*)p)->ex->GetScaleFactors();
    for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
}

void orbref_get_inv_level_sigma2(void* p, float* out)
{
    std::vector<float> v = ((RefHandle*)p)->ex->GetInverseScaleSigmaSquares();
    for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
}

int orbref_get_level_size(void* p, int level, int* w, int* h)
{
    const cv::Mat& m = ((RefHandle*)p)->ex->mvImagePyramid[level];
    *w = m.cols; *h = m.rows;
    return 0;
}

// copies the padded plane (w+38) x (h+38), tight, using the ROI's surrounding memory like Frame.cc does
void orbref_get_pyramid(void* p, int level, uint8_t* out)
{
    const cv::Mat& m = ((RefHandle*)p)->ex->mvImagePyramid[level];
    const int pad = 19;
    for (int y = -pad; y < m.rows + pad; y++)
        std::memcpy(out + (size_t)(y + pad) * (m.cols + 2 * pad), m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - pad, m.cols + 2 * pad);
}

// Frame-parallel batch for the timed CPU baseline: one ORBextractor object per worker thread, exactly how the
// reference runs two extractors from two std::threads in stereo (src/Frame.cc:127-131).  Returns total keypoints.
long orbref_extract_batch(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, const uint8_t* imgs, int n,
                          int rows, int cols, int threads, int* out_counts)
{
    if (threads < 1) threads = 1;
    std::vector<long> tot(threads, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([&, t]() {
            ORB_SLAM2::ORBextractor ex(nfeatures, scaleFactor, nlevels, iniTh, minTh);
            std::vector<cv::KeyPoint> kps;
            for (int i = t; i < n; i += threads) {
                cv::Mat im(rows, cols, CV_8UC1, (void*)(imgs + (size_t)i * rows * cols), (size_t)cols);
                cv::Mat mask, desc;
                ex(im, mask, kps, desc);
                if (out_counts) out_counts[i] = (int)kps.size();
                tot[t] += (long)kps.size();
            }
        });
    for (auto& x : th) x.join();
    long s = 0;
    for (long v : tot) s += v;
    return s;
}

// Same frame-parallel batch, returning everything: keypoint records and descriptor rows of frame i go to
// out_kps[i * cap ..] / out_desc[(i * cap ..) * 32] (at most cap per frame, out_counts[i] = the true count).  Used by the
// bulk GPU parity sweep and by bench.py's parity hash of the timed batch.
long orbref_extract_batch_full(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, const uint8_t* imgs,
                               int n, int rows, int cols, int threads, int cap, RefKp* out_kps, uint8_t* out_desc,
                               int* out_counts)
{
    if (threads < 1) threads = 1;
    std::vector<long> tot(threads, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([&, t]() {
            ORB_SLAM2::ORBextractor ex(nfeatures, scaleFactor, nlevels, iniTh, minTh);
            std::vector<cv::KeyPoint> kps;
            for (int i = t; i < n; i += threads) {
                cv::Mat im(rows, cols, CV_8UC1, (void*)(imgs + (size_t)i * rows * cols), (size_t)cols);
                cv::Mat mask, desc;
                ex(im, mask, kps, desc);
                out_counts[i] = (int)kps.size();
                tot[t] += (long)kps.size();
                const int k = std::min((int)kps.size(), cap);
                for (int j = 0; j < k; j++) {
                    const cv::KeyPoint& kp = kps[j];
                    out_kps[(size_t)i * cap + j] = RefKp{kp.pt.x, kp.pt.y, kp.size, kp.angle, kp.response, kp.octave};
                    std::memcpy(out_desc + ((size_t)i * cap + j) * 32, desc.ptr(j), 32);
                }
            }
        });
    for (auto& x : th) x.join();
    long s = 0;
    for (long v : tot) s += v;
    return s;
}

}  // extern "C"
