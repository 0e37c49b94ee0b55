// oracle/ref_bridge.cpp — TEST INFRASTRUCTURE ONLY.
// C entry points over the UNMODIFIED reference class ORB_SLAM2::ORBextractor
// (/root/reference/include/ORBextractor.h:45-112, /root/reference/src/ORBextractor.cc), which oracle/Makefile compiles
// from where it lies against oracle/shim (a stand-in for the OpenCV headers that are not installed here).
// Output: oracle/_ref/libref_orbextractor.so (git-ignored). Used to validate the restatement in orb_oracle.cpp and
// as the "reference" CPU baseline in bench.py.
#include <cstdint>
#include <cstring>
#include <vector>

#include "ORBextractor.h"

namespace {
struct Probe : public ORB_SLAM2::ORBextractor {
    using ORB_SLAM2::ORBextractor::ORBextractor;
    const std::vector<int>& perLevel() const { return mnFeaturesPerLevel; }
    const std::vector<int>& um() const { return umax; }
    int edge() const { return edgeThreshold; }
};
}  // namespace

extern "C" {

void* ref_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int patchSize, int halfPatch,
                 int edge) {
    return new Probe(nfeatures, scaleFactor, nlevels, iniTh, minTh, patchSize, halfPatch, edge);
}
void ref_destroy(void* h) { delete (Probe*)h; }

void ref_tables(void* h, float* out, int* perLevel, int* umax) {
    Probe& e = *(Probe*)h;
    int n = e.GetLevels();
    std::vector<float> a = e.GetScaleFactors(), b = e.GetInverseScaleFactors(), c = e.GetScaleSigmaSquares(),
                       d = e.GetInverseScaleSigmaSquares();
    for (int i = 0; i < n; i++) {
        out[i] = a[i];
        out[n + i] = b[i];
        out[2 * n + i] = c[i];
        out[3 * n + i] = d[i];
        perLevel[i] = e.perLevel()[i];
    }
    for (size_t v = 0; v < e.um().size(); v++) umax[v] = e.um()[v];
}

// Runs operator(); returns keypoint count in the reference's own order (level-major, list order), -1 if cap too small.
int ref_extract(void* h, const uint8_t* img, int w, int hgt, size_t stride, void* kps_out, uint8_t* desc, int cap) {
    Probe& e = *(Probe*)h;
    cv::Mat image(hgt, w, CV_8UC1, (void*)img, stride);
    std::vector<cv::KeyPoint> kps;
    cv::Mat d;
    e(image, cv::Mat(), kps, d);
    if ((int)kps.size() > cap) return -1;
    static_assert(sizeof(cv::KeyPoint) == 28, "KeyPoint layout");
    if (!kps.empty()) {
        std::memcpy(kps_out, kps.data(), kps.size() * sizeof(cv::KeyPoint));
        for (int i = 0; i < d.rows; i++) std::memcpy(desc + (size_t)i * 32, d.ptr(i), 32);
    }
    return (int)kps.size();
}

void ref_level_size(void* h, int l, int* w, int* hgt) {
    Probe& e = *(Probe*)h;
    *w = e.mvImagePyramid[l].cols;
    *hgt = e.mvImagePyramid[l].rows;
}
// with_border: copies the whole padded buffer the level's ROI lives in ((w+2e) x (h+2e)).
void ref_level_copy(void* h, int l, uint8_t* dst, int with_border) {
    Probe& e = *(Probe*)h;
    const cv::Mat& m = e.mvImagePyramid[l];
    if (with_border) {
        int E = e.edge();
        const uint8_t* base = m.data - (size_t)E * m.step - E;
        int W = m.cols + 2 * E, H = m.rows + 2 * E;
        for (int y = 0; y < H; y++) std::memcpy(dst + (size_t)y * W, base + (size_t)y * m.step, W);
    } else {
        for (int y = 0; y < m.rows; y++) std::memcpy(dst + (size_t)y * m.cols, m.ptr(y), m.cols);
    }
}

}  // extern "C"
