// oracle/prim_capi.cc -- TEST INFRASTRUCTURE. C entry points over cvprim.h so pytest (ctypes) can pin the
// restated OpenCV primitives against cv2 4.13.0.  See cvprim.h for what each restates.
#include "cvprim.h"

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

}  // extern "C"
