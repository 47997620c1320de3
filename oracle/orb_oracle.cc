// oracle/orb_oracle.cc -- TEST INFRASTRUCTURE (CPU oracle). Never linked into / called by the product path.
//
// A from-scratch CPU restatement of the reference's ORB extractor
// (/root/reference/src/ORBextractor.cc), on flat arrays, with the OpenCV primitives of cvprim.h.
// It is validated against the reference's own code compiled unmodified (oracle/_ref/liborbref.so,
// tests/test_oracle_vs_ref.py) and is what the CUDA path is compared with bit-for-bit.
//
// Oracle decisions where the reference is ill-defined:
//   D1  quad-tree phase-2 tie-break: stable sort on node size only, processed from the back
//       (the reference sorts pair<int, pointer>, i.e. by heap address -- src/ORBextractor.cc:906).
//   D2  OpenCV 4.13 arithmetic for resize / blur / FAST / fastAtan2 (cvprim.h).
//   D3  cosf/sinf = glibc 2.39 x86-64 FMA variant, restated in sincosf_glibc.h (exhaustively pinned).
//   D4  no FP contraction: build with -ffp-contract=off.
//   D6  level-0 border: the input is treated as isolated.
#include "cvprim.h"
#include "sincosf_glibc.h"
#include <list>
#include <thread>

namespace {

const int kPatch = 31, kHalfPatch = 15, kEdge = 19;

const signed char kPattern[1024] = {
#include "../orbslam2commentedbyxcm_b200/csrc/orb_pattern.inc"
};

struct Kp { float x, y, size, angle, response; int octave; };
struct Cand { float x, y; float response; };  // x,y relative to (minBorderX, minBorderY)

struct Plane {
    int w = 0, h = 0;
    size_t step = 0;
    std::vector<uint8_t> buf;
    void alloc(int _w, int _h) { w = _w; h = _h; step = (size_t)w + 2 * kEdge; buf.assign(step * ((size_t)h + 2 * kEdge), 0); }
    uint8_t* interior() { return buf.data() + (size_t)kEdge * step + kEdge; }
    const uint8_t* interior() const { return buf.data() + (size_t)kEdge * step + kEdge; }
};

struct Oracle {
    int nfeatures, nlevels, iniTh, minTh;
    float scaleFactor;
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> quota;
    int umax[kHalfPatch + 1];

    std::vector<Plane> pyr;
    std::vector<std::vector<uint8_t>> blurred;   // tight w*h per level (empty when the level has no keypoints)
    std::vector<std::vector<Cand>> cands;
    std::vector<std::vector<Kp>> levelKps;
    std::vector<Kp> kps;
    std::vector<uint8_t> desc;

    // reference: ORBextractor::ORBextractor, src/ORBextractor.cc:438-550
    Oracle(int nf, float sf, int nl, int ini, int mn) : nfeatures(nf), nlevels(nl), iniTh(ini), minTh(mn), scaleFactor(sf)
    {
        scale.resize(nl); invScale.resize(nl); sigma2.resize(nl); invSigma2.resize(nl); quota.resize(nl);
        scale[0] = 1.f; sigma2[0] = 1.f;
        // the reference keeps scaleFactor as a double member (include/ORBextractor.h): float * double -> float
        const double sfd = (double)sf;
        for (int i = 1; i < nl; i++) { scale[i] = (float)(scale[i - 1] * sfd); sigma2[i] = scale[i] * scale[i]; }
        for (int i = 0; i < nl; i++) { invScale[i] = 1.f / scale[i]; invSigma2[i] = 1.f / sigma2[i]; }
        float factor = (float)(1.0f / sfd);
        float want = nf * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
        int sum = 0;
        for (int l = 0; l < nl - 1; l++) { quota[l] = orbo::cv_round(want); sum += quota[l]; want *= factor; }
        quota[nl - 1] = std::max(nf - sum, 0);
        // circular patch half-widths (:519-549)
        int vmax = orbo::cv_floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
        int vmin = orbo::cv_ceil(kHalfPatch * std::sqrt(2.f) / 2);
        const double hp2 = kHalfPatch * kHalfPatch;
        for (int v = 0; v <= vmax; v++) umax[v] = orbo::cv_round(std::sqrt(hp2 - v * v));
        for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
    }

    // reference: ComputePyramid, :1635-1694
    void pyramid(const uint8_t* img, int rows, int cols, size_t step)
    {
        pyr.resize(nlevels);
        for (int l = 0; l < nlevels; l++) {
            float s = invScale[l];
            int w = orbo::cv_round((float)cols * s), h = orbo::cv_round((float)rows * s);
            pyr[l].alloc(w, h);
            if (l == 0) {
                for (int y = 0; y < h; y++) std::memcpy(pyr[0].interior() + (size_t)y * pyr[0].step, img + (size_t)y * step, (size_t)w);
            } else {
                orbo::resize_linear_u8(pyr[l - 1].interior(), pyr[l - 1].w, pyr[l - 1].h, pyr[l - 1].step,
                                       pyr[l].interior(), w, h, pyr[l].step);
            }
            orbo::border_reflect101_inplace(pyr[l].interior(), w, h, pyr[l].step, kEdge);
        }
    }

    // reference: cell loop of ComputeKeyPointsOctTree, :1017-1122
    void detect_level(int l)
    {
        const Plane& P = pyr[l];
        std::vector<Cand>& out = cands[l];
        out.clear();
        const int minBX = kEdge - 3, minBY = minBX;
        const int maxBX = P.w - kEdge + 3, maxBY = P.h - kEdge + 3;
        const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
        const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f);
        const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
        std::vector<orbo::FastPoint> cell;
        for (int i = 0; i < nRows; i++) {
            const float iniY = (float)(minBY + i * hCell);
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBY - 3) continue;
            if (maxY > maxBY) maxY = (float)maxBY;
            for (int j = 0; j < nCols; j++) {
                const float iniX = (float)(minBX + j * wCell);
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBX - 6) continue;
                if (maxX > maxBX) maxX = (float)maxBX;
                const int x0 = (int)iniX, x1 = (int)maxX, y0 = (int)iniY, y1 = (int)maxY;
                const uint8_t* roi = P.interior() + (ptrdiff_t)y0 * (ptrdiff_t)P.step + x0;
                orbo::fast9_16(roi, x1 - x0, y1 - y0, P.step, iniTh, true, cell);
                if (cell.empty()) orbo::fast9_16(roi, x1 - x0, y1 - y0, P.step, minTh, true, cell);
                for (const auto& p : cell)
                    out.push_back(Cand{(float)p.x + (float)(j * wCell), (float)p.y + (float)(i * hCell), (float)p.score});
            }
        }
    }

    // reference: ExtractorNode::DivideNode :581-653 and ORBextractor::DistributeOctTree :667-1013
    struct Node {
        int ulx, uly, brx, bry;     // UL and BR corners (UR = (brx, uly), BL = (ulx, bry))
        std::vector<int> keys;      // candidate indices, in candidate order
        bool noMore = false;
        std::list<Node>::iterator self;
    };

    static void split(const Node& n, const std::vector<Cand>& c, Node ch[4])
    {
        const int halfX = (int)std::ceil((float)(n.brx - n.ulx) / 2);
        const int halfY = (int)std::ceil((float)(n.bry - n.uly) / 2);
        const int mx = n.ulx + halfX, my = n.uly + halfY;
        ch[0].ulx = n.ulx; ch[0].uly = n.uly; ch[0].brx = mx;    ch[0].bry = my;
        ch[1].ulx = mx;    ch[1].uly = n.uly; ch[1].brx = n.brx; ch[1].bry = my;
        ch[2].ulx = n.ulx; ch[2].uly = my;    ch[2].brx = mx;    ch[2].bry = n.bry;
        ch[3].ulx = mx;    ch[3].uly = my;    ch[3].brx = n.brx; ch[3].bry = n.bry;
        for (int k : n.keys) {
            const bool left = c[k].x < (float)mx, up = c[k].y < (float)my;
            ch[left ? (up ? 0 : 2) : (up ? 1 : 3)].keys.push_back(k);
        }
        for (int q = 0; q < 4; q++) ch[q].noMore = (ch[q].keys.size() == 1);
    }

    void distribute(const std::vector<Cand>& c, int minX, int maxX, int minY, int maxY, int N, std::vector<int>& result)
    {
        result.clear();
        const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));
        const float hX = (float)(maxX - minX) / nIni;
        std::list<Node> nodes;
        std::vector<Node*> roots(nIni);
        for (int i = 0; i < nIni; i++) {
            Node n;
            n.ulx = (int)(hX * (float)i); n.uly = 0;
            n.brx = (int)(hX * (float)(i + 1)); n.bry = maxY - minY;
            nodes.push_back(n);
            roots[i] = &nodes.back();
        }
        for (size_t k = 0; k < c.size(); k++) roots[(size_t)(c[k].x / hX)]->keys.push_back((int)k);
        for (auto it = nodes.begin(); it != nodes.end();) {
            if (it->keys.size() == 1) { it->noMore = true; ++it; }
            else if (it->keys.empty()) it = nodes.erase(it);
            else ++it;
        }
        typedef std::pair<int, Node*> SizedNode;
        std::vector<SizedNode> expandable;
        auto push_children = [&](Node ch[4], int* nToExpand) {
            for (int q = 0; q < 4; q++) {
                if (ch[q].keys.empty()) continue;
                nodes.push_front(ch[q]);
                if (ch[q].keys.size() > 1) {
                    if (nToExpand) ++*nToExpand;
                    expandable.push_back(SizedNode((int)ch[q].keys.size(), &nodes.front()));
                    nodes.front().self = nodes.begin();
                }
            }
        };
        bool finish = false;
        while (!finish) {
            int prevSize = (int)nodes.size();
            int nToExpand = 0;
            expandable.clear();
            for (auto it = nodes.begin(); it != nodes.end();) {
                if (it->noMore) { ++it; continue; }
                Node ch[4];
                split(*it, c, ch);
                push_children(ch, &nToExpand);
                it = nodes.erase(it);
            }
            if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
                finish = true;
            } else if ((int)nodes.size() + nToExpand * 3 > N) {
                while (!finish) {
                    prevSize = (int)nodes.size();
                    std::vector<SizedNode> prev = expandable;
                    expandable.clear();
                    // D1: canonical tie-break (stable, size only)
                    std::stable_sort(prev.begin(), prev.end(), [](const SizedNode& a, const SizedNode& b) { return a.first < b.first; });
                    for (int j = (int)prev.size() - 1; j >= 0; j--) {
                        Node ch[4];
                        split(*prev[j].second, c, ch);
                        push_children(ch, nullptr);
                        nodes.erase(prev[j].second->self);
                        if ((int)nodes.size() >= N) break;
                    }
                    if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
                }
            }
        }
        for (const Node& n : nodes) {
            int best = n.keys[0];
            float bestR = c[best].response;
            for (size_t k = 1; k < n.keys.size(); k++)
                if (c[n.keys[k]].response > bestR) { best = n.keys[k]; bestR = c[best].response; }
            result.push_back(best);
        }
    }

    // reference: IC_Angle :59-106
    float ic_angle(const Plane& P, float px, float py) const
    {
        const uint8_t* center = P.interior() + (ptrdiff_t)orbo::cv_round(py) * (ptrdiff_t)P.step + orbo::cv_round(px);
        int m01 = 0, m10 = 0;
        for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * center[u];
        const ptrdiff_t step = (ptrdiff_t)P.step;
        for (int v = 1; v <= kHalfPatch; ++v) {
            int vsum = 0, d = umax[v];
            for (int u = -d; u <= d; ++u) {
                int plus = center[u + v * step], minus = center[u - v * step];
                vsum += plus - minus;
                m10 += u * (plus + minus);
            }
            m01 += v * vsum;
        }
        return orbo::fast_atan2((float)m01, (float)m10);
    }

    // reference: computeOrbDescriptor :118-172
    static void describe(const uint8_t* img, size_t step, float px, float py, float angleDeg, uint8_t* out)
    {
        const float factorPI = (float)(CV_PI_ / 180.f);
        float angle = angleDeg * factorPI;
        float a = orbo::glibc_cosf(angle), b = orbo::glibc_sinf(angle);
        const uint8_t* center = img + (ptrdiff_t)orbo::cv_round(py) * (ptrdiff_t)step + orbo::cv_round(px);
        const signed char* pat = kPattern;
        auto sample = [&](int idx) -> int {
            int x = pat[2 * idx], y = pat[2 * idx + 1];
            float fy = x * b + y * a;
            float fx = x * a - y * b;
            return center[(ptrdiff_t)orbo::cv_round(fy) * (ptrdiff_t)step + orbo::cv_round(fx)];
        };
        for (int i = 0; i < 32; ++i, pat += 32) {
            int val = 0;
            for (int k = 0; k < 8; k++) val |= (sample(2 * k) < sample(2 * k + 1)) << k;
            out[i] = (uint8_t)val;
        }
    }
    static constexpr double CV_PI_ = 3.1415926535897932384626433832795;

    // reference: operator() :1513-1629 and ComputeKeyPointsOctTree :1124-1164
    int extract(const uint8_t* img, int rows, int cols, size_t step)
    {
        kps.clear(); desc.clear();
        if (!img || rows <= 0 || cols <= 0) return 0;
        pyramid(img, rows, cols, step);
        cands.assign(nlevels, {});
        levelKps.assign(nlevels, {});
        blurred.assign(nlevels, {});
        std::vector<int> picked;
        for (int l = 0; l < nlevels; l++) {
            detect_level(l);
            const int minBX = kEdge - 3, minBY = minBX;
            const int maxBX = pyr[l].w - kEdge + 3, maxBY = pyr[l].h - kEdge + 3;
            distribute(cands[l], minBX, maxBX, minBY, maxBY, quota[l], picked);
            const int patchSize = (int)(kPatch * scale[l]);
            for (int k : picked) {
                Kp kp;
                kp.x = cands[l][k].x + minBX; kp.y = cands[l][k].y + minBY;
                kp.size = (float)patchSize; kp.angle = -1; kp.response = cands[l][k].response; kp.octave = l;
                levelKps[l].push_back(kp);
            }
        }
        for (int l = 0; l < nlevels; l++)
            for (Kp& kp : levelKps[l]) kp.angle = ic_angle(pyr[l], kp.x, kp.y);
        size_t total = 0;
        for (int l = 0; l < nlevels; l++) total += levelKps[l].size();
        desc.assign(total * 32, 0);
        size_t off = 0;
        for (int l = 0; l < nlevels; l++) {
            if (levelKps[l].empty()) continue;
            const Plane& P = pyr[l];
            // clone of the interior, then blur with REFLECT_101 on that clone (:1587-1595)
            std::vector<uint8_t> tight((size_t)P.w * P.h);
            for (int y = 0; y < P.h; y++) std::memcpy(&tight[(size_t)y * P.w], P.interior() + (size_t)y * P.step, (size_t)P.w);
            blurred[l].resize(tight.size());
            orbo::gaussian_blur_7x7_s2(tight.data(), P.w, P.h, (size_t)P.w, blurred[l].data(), (size_t)P.w);
            for (Kp& kp : levelKps[l]) {
                describe(blurred[l].data(), (size_t)P.w, kp.x, kp.y, kp.angle, &desc[off * 32]);
                off++;
                if (l != 0) { kp.x *= scale[l]; kp.y *= scale[l]; }
                kps.push_back(kp);
            }
        }
        return (int)kps.size();
    }
};

}  // namespace

extern "C" {

void* orbo_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh)
{
    return new Oracle(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void orbo_destroy(void* h) { delete (Oracle*)h; }
int orbo_extract(void* h, const uint8_t* img, int rows, int cols, size_t step) { return ((Oracle*)h)->extract(img, rows, cols, step); }
void orbo_get_keypoints(void* h, Kp* out) { Oracle* o = (Oracle*)h; std::copy(o->kps.begin(), o->kps.end(), out); }
void orbo_get_descriptors(void* h, uint8_t* out) { Oracle* o = (Oracle*)h; std::copy(o->desc.begin(), o->desc.end(), out); }
int orbo_get_levels(void* h) { return ((Oracle*)h)->nlevels; }
void orbo_get_scale_factors(void* h, float* out) { Oracle* o = (Oracle*)h; std::copy(o->scale.begin(), o->scale.end(), out); }
void orbo_get_inv_scale_factors(void* h, float* out) { Oracle* o = (Oracle*)h; std::copy(o->invScale.begin(), o->invScale.end(), out); }
void orbo_get_level_sigma2(void* h, float* out) { Oracle* o = (Oracle*)h; std::copy(o->sigma2.begin(), o->sigma2.end(), out); }
void orbo_get_inv_level_sigma2(void* h, float* out) { Oracle* o = (Oracle*)h; std::copy(o->invSigma2.begin(), o->invSigma2.end(), out); }
void orbo_get_quotas(void* h, int* out) { Oracle* o = (Oracle*)h; std::copy(o->quota.begin(), o->quota.end(), out); }
void orbo_get_umax(void* h, int* out) { Oracle* o = (Oracle*)h; std::copy(o->umax, o->umax + 16, out); }
int orbo_get_level_size(void* h, int level, int* w, int* hh)
{
    Oracle* o = (Oracle*)h;
    *w = o->pyr[level].w; *hh = o->pyr[level].h;
    return 0;
}
// padded plane, (w+38) x (h+38), tight
void orbo_get_pyramid(void* h, int level, uint8_t* out)
{
    Oracle* o = (Oracle*)h;
    std::copy(o->pyr[level].buf.begin(), o->pyr[level].buf.end(), out);
}
// blurred interior, w x h tight; returns 0 when the level was not blurred (no keypoints)
int orbo_get_blurred(void* h, int level, uint8_t* out)
{
    Oracle* o = (Oracle*)h;
    if (o->blurred[level].empty()) return 0;
    std::copy(o->blurred[level].begin(), o->blurred[level].end(), out);
    return 1;
}
// pointer to interior pixel (0,0) of a padded level (valid until the next extract on this handle) and its row stride
const uint8_t* orbo_get_level_ptr(void* h, int level, int* step)
{
    Oracle* o = (Oracle*)h;
    *step = (int)o->pyr[level].step;
    return o->pyr[level].interior();
}
int orbo_get_num_candidates(void* h, int level) { return (int)((Oracle*)h)->cands[level].size(); }
// candidates in reference order, as (x, y, response) in level pixel coordinates (minBorder already added)
void orbo_get_candidates(void* h, int level, int* out_xyr)
{
    Oracle* o = (Oracle*)h;
    const auto& c = o->cands[level];
    for (size_t i = 0; i < c.size(); i++) {
        out_xyr[3 * i] = (int)c[i].x + 16;
        out_xyr[3 * i + 1] = (int)c[i].y + 16;
        out_xyr[3 * i + 2] = (int)c[i].response;
    }
}
int orbo_get_num_level_keypoints(void* h, int level) { return (int)((Oracle*)h)->levelKps[level].size(); }

// Stand-alone stage entry points (used by the parity tests to check one CUDA stage at a time).
// Quad-tree on an arbitrary candidate list: xyr = n x (x, y, response) relative to the cell-grid origin.
int orbo_distribute(const int* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int* out_idx)
{
    Oracle o(1000, 1.2f, 8, 20, 7);
    std::vector<Cand> c(n);
    for (int i = 0; i < n; i++) c[i] = Cand{(float)xyr[3 * i], (float)xyr[3 * i + 1], (float)xyr[3 * i + 2]};
    std::vector<int> picked;
    o.distribute(c, minX, maxX, minY, maxY, N, picked);
    std::copy(picked.begin(), picked.end(), out_idx);
    return (int)picked.size();
}

void orbo_sincosf(const float* angle, float* s, float* c, int n)
{
    for (int i = 0; i < n; i++) { s[i] = orbo::glibc_sinf(angle[i]); c[i] = orbo::glibc_cosf(angle[i]); }
}
void orbo_libm_sincosf(const float* angle, float* s, float* c, int n)
{
    for (int i = 0; i < n; i++) { s[i] = sinf(angle[i]); c[i] = cosf(angle[i]); }
}

// Frame-parallel batch for the timed CPU baseline: n images of identical shape, tight rows, `threads` workers,
// one Oracle per worker (mirrors "one extractor object per thread", src/Frame.cc:127-131).  Returns total keypoints.
long orbo_extract_batch(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh,
                        const uint8_t* imgs, int n, int rows, int cols, int threads, int* out_counts)
{
    if (threads < 1) threads = 1;
    std::vector<long> tot(threads, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([&, t]() {
            Oracle o(nfeatures, scaleFactor, nlevels, iniTh, minTh);
            for (int i = t; i < n; i += threads) {
                int k = o.extract(imgs + (size_t)i * rows * cols, rows, cols, (size_t)cols);
                if (out_counts) out_counts[i] = k;
                tot[t] += k;
            }
        });
    for (auto& x : th) x.join();
    long s = 0;
    for (long v : tot) s += v;
    return s;
}

}  // extern "C"
