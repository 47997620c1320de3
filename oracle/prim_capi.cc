// oracle/prim_capi.cc -- TEST INFRASTRUCTURE. C entry points over cvprim.h so pytest (ctypes) can pin the
// restated OpenCV primitives against cv2 4.13.0.  See cvprim.h for what each restates.
#include "cvprim.h"
#include "sincosf_glibc.h"
#include <algorithm>
#include <thread>
#include <vector>

extern "C" {

void orbo_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstep, uint8_t* dst, int dw, int dh,
                           size_t dstep)
{
    orbo::resize_linear_u8(src, sw, sh, sstep, dst, dw, dh, dstep);
}

// src: w x h tight; dst: (w+2*pad) x (h+2*pad) tight
void orbo_copy_make_border101(const uint8_t* src, int w, int h, uint8_t* dst, int pad)
{
    size_t dstep = (size_t)w + 2 * pad;
    uint8_t* interior = dst + (size_t)pad * dstep + pad;
    for (int y = 0; y < h; y++) std::memcpy(interior + (size_t)y * dstep, src + (size_t)y * w, (size_t)w);
    orbo::border_reflect101_inplace(interior, w, h, dstep, pad);
}

// returns number of keypoints, writes up to cap triples (x, y, score)
int orbo_fast9_16(const uint8_t* img, int w, int h, size_t step, int threshold, int nms, int* out_xys, int cap)
{
    std::vector<orbo::FastPoint> v;
    orbo::fast9_16(img, w, h, step, threshold, nms != 0, v);
    int n = (int)v.size();
    for (int i = 0; i < n && i < cap; i++) {
        out_xys[3 * i] = v[i].x;
        out_xys[3 * i + 1] = v[i].y;
        out_xys[3 * i + 2] = v[i].score;
    }
    return n;
}

int orbo_fast9_16_simple(const uint8_t* img, int w, int h, size_t step, int threshold, int nms, int* out_xys, int cap)
{
    std::vector<orbo::FastPoint> v;
    orbo::fast9_16_simple(img, w, h, step, threshold, nms != 0, v);
    int n = (int)v.size();
    for (int i = 0; i < n && i < cap; i++) {
        out_xys[3 * i] = v[i].x;
        out_xys[3 * i + 1] = v[i].y;
        out_xys[3 * i + 2] = v[i].score;
    }
    return n;
}

void orbo_gaussian_blur_7x7_s2(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep)
{
    orbo::gaussian_blur_7x7_s2(src, w, h, sstep, dst, dstep);
}

void orbo_fast_atan2(const float* y, const float* x, float* out, int n)
{
    for (int i = 0; i < n; i++) out[i] = orbo::fast_atan2(y[i], x[i]);
}

int orbo_cv_round_f(float v) { return orbo::cv_round(v); }

// ---------------------------------------------------------------------------------------------------------------------
// Domain sweeps (checker side of orbx_debug_sincosf_checksum / orbx_debug_fast_atan2_checksum, include/orbx.h): the same
// generators and the same position-weighted checksum over oracle/sincosf_glibc.h (exhaustively pinned against the host
// libm by oracle/_ref/sweep_sincosf) and orbo::fast_atan2 (pinned against cv2.fastAtan2).  One thread per chunk stripe.
// ---------------------------------------------------------------------------------------------------------------------
static inline void sweep_atan2_pair(int mode, uint64_t i, int& y, int& x)
{
    if (mode == 0) {
        const uint32_t W = 10001u;
        y = (int)(i / W) - 5000;
        x = (int)(i % W) - 5000;
    } else {
        uint64_t z = i + 0x9E3779B97F4A7C15ull * (uint64_t)mode;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        y = (int)(z & 0x3fffffu) - 0x200000;
        x = (int)((z >> 32) & 0x3fffffu) - 0x200000;
    }
}

static void sweep_run(int what, int mode, uint64_t first, uint64_t total, uint64_t chunk, uint64_t* sumA, uint64_t* sumB, int threads)
{
    const uint64_t nchunks = (total + chunk - 1) / chunk;
    if (threads < 1) threads = 1;
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([=]() {
            for (uint64_t c = (uint64_t)t; c < nchunks; c += (uint64_t)threads) {
                const uint64_t lo = c * chunk, hi = std::min(lo + chunk, total);
                uint64_t a = 0, b = 0;
                for (uint64_t i = lo; i < hi; i++) {
                    const uint64_t w = 2 * (i - lo) + 1;
                    uint32_t u;
                    if (what == 0) {
                        const uint32_t bits = (uint32_t)(first + i);
                        float deg;
                        std::memcpy(&deg, &bits, 4);
                        volatile float angle = deg * factorPI;
                        const float s = orbo::glibc_sinf(angle), co = orbo::glibc_cosf(angle);
                        std::memcpy(&u, &s, 4); a += (uint64_t)u * w;
                        std::memcpy(&u, &co, 4); b += (uint64_t)u * w;
                    } else {
                        int y, x;
                        sweep_atan2_pair(mode, first + i, y, x);
                        const float v = orbo::fast_atan2((float)y, (float)x);
                        std::memcpy(&u, &v, 4); a += (uint64_t)u * w;
                    }
                }
                sumA[c] = a;
                if (sumB) sumB[c] = b;
            }
        });
    for (auto& x : th) x.join();
}

void orbo_sincosf_checksum(uint32_t first_bits, uint64_t count, uint64_t chunk, uint64_t* sum_sin, uint64_t* sum_cos, int threads)
{
    sweep_run(0, 0, first_bits, count, chunk, sum_sin, sum_cos, threads);
}
void orbo_fast_atan2_checksum(int mode, uint64_t first, uint64_t count, uint64_t chunk, uint64_t* sum, int threads)
{
    sweep_run(1, mode, first, count, chunk, sum, nullptr, threads);
}
void orbo_sweep_sincosf_values(uint32_t first_bits, int n, float* out_sin, float* out_cos)
{
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    for (int i = 0; i < n; i++) {
        const uint32_t bits = first_bits + (uint32_t)i;
        float deg;
        std::memcpy(&deg, &bits, 4);
        volatile float angle = deg * factorPI;
        out_sin[i] = orbo::glibc_sinf(angle);
        out_cos[i] = orbo::glibc_cosf(angle);
    }
}
void orbo_sweep_fast_atan2_values(int mode, uint64_t first, int n, float* out)
{
    for (int i = 0; i < n; i++) {
        int y, x;
        sweep_atan2_pair(mode, first + (uint64_t)i, y, x);
        out[i] = orbo::fast_atan2((float)y, (float)x);
    }
}

}  // extern "C"
