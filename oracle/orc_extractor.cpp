// orc_extractor.cpp — ORACLE (test infrastructure only; see orb_oracle.h).
//
// CPU restatement of ORB_SLAM2::ORBextractor (reference src/ORBextractor.cc).  Structure is
// deliberately array/index based (no std::list, no cv::Mat) so that the one allocator-
// dependent spot of the reference — the (count, pointer) sort at ORBextractor.cc:728 — is
// pinned to "pointer order == node creation order" (SURVEY.md §7 hard part 1).  Validated
// against the unmodified reference TU in oracle/_ref (tests/test_oracle_vs_ref.py).
#include "orb_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <utility>
#include <vector>

namespace {

const int PATCH_SIZE = 31;        // ORBextractor.cc:73
const int HALF_PATCH_SIZE = 15;   // :74
const int EDGE_THRESHOLD = 19;    // :75

const signed char kPattern[1024] = {
#include "../include/orbfe_pattern.inc"
};

struct Level {
    int w = 0, h = 0;
    std::vector<uint8_t> img;      // interior pixels only, stride == w (border is never read)
    std::vector<uint8_t> blur;
    std::vector<int> raw;          // x,y,score relative to minBorder, vToDistributeKeys order
};

}  // namespace

struct orc_extractor {
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> per_level;
    int umax[HALF_PATCH_SIZE + 1];
    std::vector<Level> lv;
};

// ---------------------------------------------------------------------------------------
// ORBextractor::ORBextractor — ORBextractor.cc:415-486
extern "C" orc_extractor* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels,
                                               int iniThFAST, int minThFAST) {
    orc_extractor* e = new orc_extractor;
    e->nfeatures = nfeatures; e->nlevels = nlevels; e->iniTh = iniThFAST; e->minTh = minThFAST;
    e->scaleFactor = scaleFactor;                       // double member set from a float (:417)
    e->scale.resize(nlevels); e->sigma2.resize(nlevels);
    e->inv_scale.resize(nlevels); e->inv_sigma2.resize(nlevels);
    e->scale[0] = 1.0f; e->sigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {                 // :424-428
        e->scale[i] = (float)(e->scale[i - 1] * e->scaleFactor);
        e->sigma2[i] = e->scale[i] * e->scale[i];
    }
    for (int i = 0; i < nlevels; i++) {                 // :432-436
        e->inv_scale[i] = 1.0f / e->scale[i];
        e->inv_sigma2[i] = 1.0f / e->sigma2[i];
    }
    e->per_level.resize(nlevels);
    float factor = (float)(1.0f / e->scaleFactor);      // :441
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {             // :445-450
        e->per_level[l] = orc_cv_round_f(nDesired);
        sum += e->per_level[l];
        nDesired *= factor;
    }
    e->per_level[nlevels - 1] = std::max(nfeatures - sum, 0);
    // umax — :460-485
    int v, v0;
    int vmax = (int)std::floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    int vmin = (int)std::ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    std::memset(e->umax, 0, sizeof(e->umax));
    for (v = 0; v <= vmax; ++v) e->umax[v] = orc_cv_round_d(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (e->umax[v0] == e->umax[v0 + 1]) ++v0;
        e->umax[v] = v0;
        ++v0;
    }
    e->lv.resize(nlevels);
    return e;
}

extern "C" void orc_extractor_destroy(orc_extractor* e) { delete e; }

extern "C" void orc_extractor_tables(const orc_extractor* e, float* scale, float* inv_scale,
                                     float* sigma2, float* inv_sigma2, int* per_level,
                                     int* umax16) {
    for (int i = 0; i < e->nlevels; i++) {
        if (scale) scale[i] = e->scale[i];
        if (inv_scale) inv_scale[i] = e->inv_scale[i];
        if (sigma2) sigma2[i] = e->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = e->inv_sigma2[i];
        if (per_level) per_level[i] = e->per_level[i];
    }
    if (umax16) for (int i = 0; i < 16; i++) umax16[i] = e->umax[i];
}

// ---------------------------------------------------------------------------------------
// DistributeOctTree — ORBextractor.cc:566-808 with DivideNode :498-558.
namespace {
struct QNode {
    int x0, y0, x1, y1;             // UL.x, UL.y, UR.x(=BR.x), BL.y(=BR.y)
    std::vector<int> keys;          // indices into the input, vKeys order
    bool noMore = false;
    int prev = -1, next = -1;       // std::list links
};

struct QList {                      // index-linked list emulating std::list<ExtractorNode>
    std::vector<QNode> nodes;       // index == creation sequence number
    int head = -1, tail = -1, size = 0;
    int push_back(QNode&& n) {
        int id = (int)nodes.size(); nodes.push_back(std::move(n));
        nodes[id].prev = tail; nodes[id].next = -1;
        if (tail >= 0) nodes[tail].next = id; else head = id;
        tail = id; size++; return id;
    }
    int push_front(QNode&& n) {
        int id = (int)nodes.size(); nodes.push_back(std::move(n));
        nodes[id].prev = -1; nodes[id].next = head;
        if (head >= 0) nodes[head].prev = id; else tail = id;
        head = id; size++; return id;
    }
    int erase(int id) {             // returns next
        int p = nodes[id].prev, nx = nodes[id].next;
        if (p >= 0) nodes[p].next = nx; else head = nx;
        if (nx >= 0) nodes[nx].prev = p; else tail = p;
        size--; return nx;
    }
};

// DivideNode :498-558
void divide(const QNode& p, const int* xys, QNode c[4]) {
    const int halfX = (int)std::ceil(static_cast<float>(p.x1 - p.x0) / 2);
    const int halfY = (int)std::ceil(static_cast<float>(p.y1 - p.y0) / 2);
    const int mx = p.x0 + halfX, my = p.y0 + halfY;
    c[0] = QNode{p.x0, p.y0, mx, my};
    c[1] = QNode{mx, p.y0, p.x1, my};
    c[2] = QNode{p.x0, my, mx, p.y1};
    c[3] = QNode{mx, my, p.x1, p.y1};
    for (int k : p.keys) {
        const float px = (float)xys[3 * k], py = (float)xys[3 * k + 1];
        if (px < mx) { if (py < my) c[0].keys.push_back(k); else c[2].keys.push_back(k); }
        else if (py < my) c[1].keys.push_back(k);
        else c[3].keys.push_back(k);
    }
    for (int i = 0; i < 4; i++) if (c[i].keys.size() == 1) c[i].noMore = true;
}
}  // namespace

extern "C" int orc_distribute_octtree(const int* xys, int n, int minX, int maxX, int minY,
                                      int maxY, int N, int* out_idx, int cap) {
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));  // :570
    const float hX = static_cast<float>(maxX - minX) / nIni;                             // :572
    QList L;
    std::vector<int> ini(nIni);
    for (int i = 0; i < nIni; i++) {                                                     // :581-592
        QNode ni;
        ni.x0 = (int)(hX * static_cast<float>(i));
        ni.x1 = (int)(hX * static_cast<float>(i + 1));
        ni.y0 = 0; ni.y1 = maxY - minY;
        ini[i] = L.push_back(std::move(ni));
    }
    for (int i = 0; i < n; i++)                                                          // :595-599
        L.nodes[ini[(size_t)((float)xys[3 * i] / hX)]].keys.push_back(i);
    for (int it = L.head; it >= 0;) {                                                    // :604-615
        QNode& nd = L.nodes[it];
        if (nd.keys.size() == 1) { nd.noMore = true; it = nd.next; }
        else if (nd.keys.empty()) it = L.erase(it);
        else it = nd.next;
    }
    bool finish = false;
    std::vector<std::pair<int, int>> sizeAndNode;   // (count, node id); id order == pointer order
    auto add_children = [&](QNode c[4], int* nToExpand) {
        for (int k = 0; k < 4; k++) {
            if (c[k].keys.empty()) continue;
            int cnt = (int)c[k].keys.size();
            int id = L.push_front(std::move(c[k]));
            if (cnt > 1) { if (nToExpand) (*nToExpand)++; sizeAndNode.push_back({cnt, id}); }
        }
    };
    while (!finish) {                                                                    // :625
        int prevSize = L.size;
        int nToExpand = 0;
        sizeAndNode.clear();
        for (int it = L.head; it >= 0;) {                                                // :639-699
            if (L.nodes[it].noMore) { it = L.nodes[it].next; continue; }
            QNode c[4];
            divide(L.nodes[it], xys, c);
            add_children(c, &nToExpand);
            it = L.erase(it);
        }
        if (L.size >= N || L.size == prevSize) finish = true;                            // :707
        else if (L.size + nToExpand * 3 > N) {                                           // :714
            while (!finish) {
                prevSize = L.size;
                std::vector<std::pair<int, int>> prev = sizeAndNode;
                sizeAndNode.clear();
                std::sort(prev.begin(), prev.end());                                     // :728
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    QNode c[4];
                    divide(L.nodes[prev[j].second], xys, c);
                    add_children(c, nullptr);
                    L.erase(prev[j].second);
                    if (L.size >= N) break;                                              // :774
                }
                if (L.size >= N || L.size == prevSize) finish = true;                    // :778
            }
        }
    }
    int cnt = 0;
    for (int it = L.head; it >= 0; it = L.nodes[it].next) {                              // :789-805
        const std::vector<int>& keys = L.nodes[it].keys;
        int best = keys[0];
        float maxResp = (float)xys[3 * best + 2];
        for (size_t k = 1; k < keys.size(); k++)
            if ((float)xys[3 * keys[k] + 2] > maxResp) { best = keys[k]; maxResp = (float)xys[3 * best + 2]; }
        if (cnt < cap) out_idx[cnt] = best;
        cnt++;
    }
    return cnt;
}

// ---------------------------------------------------------------------------------------
// IC_Angle — ORBextractor.cc:78-106
static float ic_angle(const uint8_t* img, int stride, int px, int py, const int* umax) {
    int m_01 = 0, m_10 = 0;
    const uint8_t* center = img + (ptrdiff_t)py * stride + px;
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0;
        int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            int val_plus = center[u + v * stride], val_minus = center[u - v * stride];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return orc_fast_atan2((float)m_01, (float)m_10);
}

// computeOrbDescriptor — ORBextractor.cc:111-152 (float op order, no FMA: -ffp-contract=off;
// cos/sin resolve to the float overloads = glibc cosf/sinf, :68,:116)
static void orb_descriptor(const uint8_t* img, int stride, int px, int py, float angle_deg,
                           uint8_t* desc) {
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float angle = angle_deg * factorPI;
    float a = cosf(angle), b = sinf(angle);
    const uint8_t* center = img + (ptrdiff_t)py * stride + px;
    const signed char* pat = kPattern;
    for (int i = 0; i < 32; ++i, pat += 32) {
        int val = 0;
        for (int j = 0; j < 8; j++) {
            const float x0 = pat[4 * j], y0 = pat[4 * j + 1], x1 = pat[4 * j + 2], y1 = pat[4 * j + 3];
            int t0 = center[orc_cv_round_f(x0 * b + y0 * a) * stride + orc_cv_round_f(x0 * a - y0 * b)];
            int t1 = center[orc_cv_round_f(x1 * b + y1 * a) * stride + orc_cv_round_f(x1 * a - y1 * b)];
            val |= (t0 < t1) << j;
        }
        desc[i] = (uint8_t)val;
    }
}

// ---------------------------------------------------------------------------------------
// operator() — ORBextractor.cc:1119-1197 (ComputePyramid :1203-1234,
// ComputeKeyPointsOctTree :815-922)
extern "C" int orc_extract(orc_extractor* e, const uint8_t* img, int w, int h, ptrdiff_t stride,
                           orc_keypoint* kps, uint8_t* desc, int cap) {
    if (!img || w <= 0 || h <= 0) return 0;
    // ComputePyramid: level sizes from the ORIGINAL size (:1208), resize source = level-1.
    for (int l = 0; l < e->nlevels; l++) {
        Level& L = e->lv[l];
        float s = e->inv_scale[l];
        L.w = orc_cv_round_f((float)w * s);
        L.h = orc_cv_round_f((float)h * s);
        L.img.assign((size_t)L.w * L.h, 0);
        if (l == 0) {
            for (int y = 0; y < h; y++) std::memcpy(&L.img[(size_t)y * w], img + y * stride, w);
        } else {
            const Level& P = e->lv[l - 1];
            orc_resize_linear_u8(P.img.data(), P.w, P.h, P.w, L.img.data(), L.w, L.h, L.w);
        }
    }
    // ComputeKeyPointsOctTree
    const float W = 30;
    struct KP { int x, y, score; float angle; };
    std::vector<std::vector<KP>> all(e->nlevels);
    std::vector<int> cellbuf;
    for (int l = 0; l < e->nlevels; l++) {
        Level& L = e->lv[l];
        const int minBX = EDGE_THRESHOLD - 3, minBY = minBX;
        const int maxBX = L.w - EDGE_THRESHOLD + 3, maxBY = L.h - EDGE_THRESHOLD + 3;
        L.raw.clear();
        const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
        const int nCols = (int)(width / W), nRows = (int)(height / W);
        if (nCols > 0 && nRows > 0) {
            const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
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
                    const int cx = (int)iniX, cy = (int)iniY, cw = (int)maxX - cx, ch = (int)maxY - cy;
                    cellbuf.resize((size_t)3 * cw * ch + 3);
                    const uint8_t* cell = &L.img[(size_t)cy * L.w + cx];
                    int n = orc_fast9_16(cell, cw, ch, L.w, e->iniTh, 1, cellbuf.data(), cw * ch);
                    if (n == 0) n = orc_fast9_16(cell, cw, ch, L.w, e->minTh, 1, cellbuf.data(), cw * ch);
                    for (int k = 0; k < n; k++) {
                        L.raw.push_back(cellbuf[3 * k] + j * wCell);
                        L.raw.push_back(cellbuf[3 * k + 1] + i * hCell);
                        L.raw.push_back(cellbuf[3 * k + 2]);
                    }
                }
            }
        }
        const int nraw = (int)L.raw.size() / 3;
        std::vector<int> sel(nraw + 4);
        int nsel = nraw ? orc_distribute_octtree(L.raw.data(), nraw, minBX, maxBX, minBY, maxBY,
                                                 e->per_level[l], sel.data(), nraw + 4)
                        : 0;
        for (int k = 0; k < nsel; k++)
            all[l].push_back(KP{L.raw[3 * sel[k]] + minBX, L.raw[3 * sel[k] + 1] + minBY,
                                L.raw[3 * sel[k] + 2], 0.f});
    }
    for (int l = 0; l < e->nlevels; l++)      // computeOrientation :920-921
        for (KP& k : all[l]) k.angle = ic_angle(e->lv[l].img.data(), e->lv[l].w, k.x, k.y, e->umax);

    int total = 0;
    for (int l = 0; l < e->nlevels; l++) total += (int)all[l].size();
    if (total > cap) return -1;
    int off = 0;
    for (int l = 0; l < e->nlevels; l++) {
        Level& L = e->lv[l];
        L.blur.clear();
        if (all[l].empty()) continue;
        L.blur.resize((size_t)L.w * L.h);
        orc_gaussian_blur_7x7_s2(L.img.data(), L.w, L.h, L.w, L.blur.data(), L.w);   // :1169-1175
        const int scaledPatch = (int)(PATCH_SIZE * e->scale[l]);                      // :905
        for (const KP& k : all[l]) {
            orb_descriptor(L.blur.data(), L.w, k.x, k.y, k.angle, desc + (size_t)32 * off);
            orc_keypoint& o = kps[off];
            o.x = (float)k.x; o.y = (float)k.y;
            if (l != 0) { o.x *= e->scale[l]; o.y *= e->scale[l]; }                   // :1187-1193
            o.size = (float)scaledPatch; o.angle = k.angle; o.response = (float)k.score;
            o.octave = l; o.class_id = -1;
            off++;
        }
    }
    return total;
}

extern "C" int orc_level_size(const orc_extractor* e, int l, int* w, int* h) {
    if (l < 0 || l >= e->nlevels) return -1;
    *w = e->lv[l].w; *h = e->lv[l].h; return 0;
}
extern "C" const uint8_t* orc_level_image(const orc_extractor* e, int l, ptrdiff_t* stride) {
    *stride = e->lv[l].w; return e->lv[l].img.data();
}
extern "C" const uint8_t* orc_level_blurred(const orc_extractor* e, int l, ptrdiff_t* stride) {
    *stride = e->lv[l].w; return e->lv[l].blur.empty() ? nullptr : e->lv[l].blur.data();
}
extern "C" int orc_level_raw(const orc_extractor* e, int l, const int** xys) {
    *xys = e->lv[l].raw.data(); return (int)e->lv[l].raw.size() / 3;
}
