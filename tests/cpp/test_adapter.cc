// tests/cpp/test_adapter.cc -- uses the drop-in ORB_SLAM2::ORBextractor (host/ORBextractor.{h,cc} over liborbx.so) exactly
// like Frame::ExtractORB does (src/Frame.cc:377-386) and compares keypoints, descriptors, accessors and mvImagePyramid with
// the reference's own class, reached through oracle/_ref/liborbref.so's C entry points (both classes have the same
// name, so the reference lives behind dlopen).  Exit code 0 = identical.  Needs a GPU (pytest -m gpu runs it).
#include <dlfcn.h>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ORBextractor.h"

struct RefKp { float x, y, size, angle, response; int octave; };

static unsigned lcg(unsigned& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

int main(int argc, char** argv)
{
    const char* refPath = argc > 1 ? argv[1] : "oracle/_ref/liborbref.so";
    void* ref = dlopen(refPath, RTLD_NOW | RTLD_LOCAL);
    if (!ref) { std::fprintf(stderr, "cannot open %s: %s\n", refPath, dlerror()); return 2; }
    auto r_create = (void* (*)(int, float, int, int, int))dlsym(ref, "orbref_create");
    auto r_extract = (int (*)(void*, const unsigned char*, int, int, size_t))dlsym(ref, "orbref_extract");
    auto r_kps = (void (*)(void*, RefKp*))dlsym(ref, "orbref_get_keypoints");
    auto r_desc = (void (*)(void*, unsigned char*))dlsym(ref, "orbref_get_descriptors");
    auto r_sf = (void (*)(void*, float*))dlsym(ref, "orbref_get_scale_factors");
    auto r_pyr = (void (*)(void*, int, unsigned char*))dlsym(ref, "orbref_get_pyramid");
    auto r_size = (int (*)(void*, int, int*, int*))dlsym(ref, "orbref_get_level_size");

    const int rows = 240, cols = 320;
    cv::Mat im(rows, cols, CV_8UC1);
    unsigned seed = 7;
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) {
            int block = ((x / 8) * 31 + (y / 8) * 17) % 7;
            im.at<uchar>(y, x) = (uchar)(40 + 30 * block + (lcg(seed) % 5));
        }
    ORB_SLAM2::ORBextractor ex(500, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> kps;
    cv::Mat desc;
    ex(im, cv::Mat(), kps, desc);

    void* rh = r_create(500, 1.2f, 8, 20, 7);
    const int n = r_extract(rh, im.data, rows, cols, im.step);
    std::vector<RefKp> rk(n > 0 ? n : 1);
    std::vector<unsigned char> rd((size_t)(n > 0 ? n : 1) * 32);
    r_kps(rh, rk.data());
    r_desc(rh, rd.data());
    int bad = 0;
    if ((int)kps.size() != n || desc.rows != n) { std::fprintf(stderr, "count %zu vs %d\n", kps.size(), n); return 1; }
    for (int i = 0; i < n; i++) {
        const cv::KeyPoint& k = kps[i];
        if (k.pt.x != rk[i].x || k.pt.y != rk[i].y || k.size != rk[i].size || std::memcmp(&k.angle, &rk[i].angle, 4) ||
            k.response != rk[i].response || k.octave != rk[i].octave || k.class_id != -1)
            bad++;
        if (std::memcmp(desc.ptr(i), &rd[(size_t)i * 32], 32)) bad++;
    }
    std::vector<float> sf = ex.GetScaleFactors(), rsf(8);
    r_sf(rh, rsf.data());
    for (int l = 0; l < 8; l++) bad += sf[l] != rsf[l];
    for (int l = 0; l < ex.GetLevels(); l++) {
        int w, h;
        r_size(rh, l, &w, &h);
        const cv::Mat& m = ex.mvImagePyramid[l];
        if (m.cols != w || m.rows != h) { bad++; continue; }
        std::vector<unsigned char> rp((size_t)(w + 38) * (h + 38));
        r_pyr(rh, l, rp.data());
        // compare including the 19-px frame reachable by adjustROI
        for (int y = -19; y < h + 19; y++)
            if (std::memcmp(m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - 19, &rp[(size_t)(y + 19) * (w + 38)], w + 38)) { bad++; break; }
    }
    std::printf("adapter vs reference: %d keypoints, %d mismatches\n", n, bad);
    // latency of the call a stock Frame constructor makes (src/Frame.cc:377-386) at config 1 (640 x 480, 1000 features), with the
    // pyramid download that Frame::ComputeStereoMatches needs switched on (default) and off
    {
        const int R = 480, C = 640;
        cv::Mat big(R, C, CV_8UC1);
        for (int y = 0; y < R; y++)
            for (int x = 0; x < C; x++) {
                int block = ((x / 8) * 31 + (y / 8) * 17) % 7;
                big.at<uchar>(y, x) = (uchar)(40 + 30 * block + (lcg(seed) % 5));
            }
        ORB_SLAM2::ORBextractor ex1(1000, 1.2f, 8, 20, 7);
        for (int mode = 0; mode < 2; mode++) {
            ex1.SetPyramidDownload(mode == 0);
            std::vector<double> t;
            for (int it = 0; it < 220; it++) {
                std::vector<cv::KeyPoint> k1;
                cv::Mat d1;
                const auto t0 = std::chrono::steady_clock::now();
                ex1(big, cv::Mat(), k1, d1);
                const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
                if (it >= 20) t.push_back(ms);
            }
            std::sort(t.begin(), t.end());
            std::printf("adapter latency 640x480/1000 pyramid_%s: median %.4f ms, p90 %.4f ms\n", mode == 0 ? "on" : "off", t[t.size() / 2],
                        t[t.size() * 9 / 10]);
        }
    }
    return bad ? 1 : 0;
}
