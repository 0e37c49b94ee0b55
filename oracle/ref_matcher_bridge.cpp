// oracle/ref_matcher_bridge.cpp — TEST INFRASTRUCTURE ONLY.
//
// Flat C entry points over the UNMODIFIED reference code compiled into oracle/_ref/libref_orbmatcher.so
// (src/ORBmatcher.cc whole; Frame::ComputeStereoMatches, MapPoint::ComputeDistinctiveDescriptors and the small
// Frame / KeyFrame / MapPoint members ORBmatcher.cc calls, cut out by oracle/extract_ref_members.py), for the checks
// that need no scene: tests/test_cpu_refmatcher.py compares them with the oracle's restatements (orc_*).
// The method-level pin (the eleven searches on whole scenes) goes through tests/cpp/scene_run.cpp instead.
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "ORBmatcher.h"

using namespace ORB_SLAM2;

namespace {
struct RawKp { float x, y, size, angle, response; int octave, class_id; };
std::vector<cv::KeyPoint> keypoints(const RawKp* r, int n) {
    std::vector<cv::KeyPoint> k(n);
    for (int i = 0; i < n; i++) {
        k[i].pt.x = r[i].x; k[i].pt.y = r[i].y; k[i].size = r[i].size; k[i].angle = r[i].angle;
        k[i].response = r[i].response; k[i].octave = r[i].octave; k[i].class_id = r[i].class_id;
    }
    return k;
}
struct OpenMatcher : ORBmatcher {  // ComputeThreeMaxima is protected (include/ORBmatcher.h:96)
    using ORBmatcher::ComputeThreeMaxima;
};
}  // namespace

extern "C" {

// ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1643-1659
int refm_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    cv::Mat A(1, 32, CV_8U, (void*)a), B(1, 32, CV_8U, (void*)b);
    return ORBmatcher::DescriptorDistance(A, B);
}

// ORBmatcher::ComputeThreeMaxima, src/ORBmatcher.cc:1597-1638: counts[i] = histo[i].size(); out = ind1, ind2, ind3
void refm_three_maxima(const int* counts, int L, int* out3) {
    std::vector<std::vector<int> > histo(L);
    for (int i = 0; i < L; i++) histo[i].assign(counts[i], 0);
    OpenMatcher m;
    int i1 = -1, i2 = -1, i3 = -1;
    m.ComputeThreeMaxima(histo.data(), L, i1, i2, i3);
    out3[0] = i1; out3[1] = i2; out3[2] = i3;
}

// Frame::ComputeStereoMatches, src/Frame.cc:464-642. Pyramid levels are passed as interior pointers (un-blurred) with a
// pitch, like orc_stereo_match takes them.
void refm_stereo_match(const void* kL, int NL, const uint8_t* dL, const void* kR, int NR, const uint8_t* dR,
                       const uint8_t* const* pyrL, const uint8_t* const* pyrR, const int* lw, const int* lh,
                       const int* pitchL, const int* pitchR, int nlevels, const float* scaleF, const float* invScaleF,
                       float mbf, float mb, float* uRight, float* depth) {
    ORBextractor exL, exR;
    for (int l = 0; l < nlevels; l++) {
        exL.mvImagePyramid.push_back(cv::Mat(lh[l], lw[l], CV_8U, (void*)pyrL[l], (size_t)pitchL[l]));
        exR.mvImagePyramid.push_back(cv::Mat(lh[l], lw[l], CV_8U, (void*)pyrR[l], (size_t)pitchR[l]));
    }
    Frame F;
    F.mpORBextractorLeft = &exL; F.mpORBextractorRight = &exR;
    F.N = NL;
    F.mvKeys = keypoints((const RawKp*)kL, NL);
    F.mvKeysRight = keypoints((const RawKp*)kR, NR);
    F.mDescriptors = cv::Mat(NL, 32, CV_8U, (void*)dL);
    F.mDescriptorsRight = cv::Mat(NR, 32, CV_8U, (void*)dR);
    F.mvScaleFactors.assign(scaleF, scaleF + nlevels);
    F.mvInvScaleFactors.assign(invScaleF, invScaleF + nlevels);
    F.mbf = mbf; F.mb = mb;
    F.ComputeStereoMatches();
    for (int i = 0; i < NL; i++) { uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
}

// MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:242-307, for npoints map points: point p is observed in
// off[p+1]-off[p] keyframes whose descriptor row for it is desc[off[p] + k]; out[p] = the chosen descriptor (zeros and
// chosen[p] = 0 when the point has no observation: the reference returns before writing mDescriptor).
void refm_distinctive_descriptors(const uint8_t* desc, const int32_t* off, int npoints, uint8_t* out, int32_t* chosen) {
    int maxObs = 0;
    for (int p = 0; p < npoints; p++) maxObs = std::max(maxObs, off[p + 1] - off[p]);
    std::unique_ptr<KeyFrame[]> kfs(new KeyFrame[maxObs > 0 ? maxObs : 1]);  // contiguous: map<KeyFrame*,..> iterates k = 0, 1, ...
    for (int k = 0; k < maxObs; k++) {
        kfs[k].mDescriptors = cv::Mat(1, 32, CV_8U);
        kfs[k].mvuRight.assign(1, -1.0f);
    }
    for (int p = 0; p < npoints; p++) {
        const int n = off[p + 1] - off[p];
        for (int k = 0; k < n; k++) memcpy(kfs[k].mDescriptors.data, desc + (size_t)(off[p] + k) * 32, 32);
        MapPoint M;
        for (int k = 0; k < n; k++) M.AddObservation(&kfs[k], 0);
        M.ComputeDistinctiveDescriptors();
        chosen[p] = !M.mDescriptor.empty();
        if (chosen[p]) memcpy(out + (size_t)p * 32, M.mDescriptor.data, 32);
        else memset(out + (size_t)p * 32, 0, 32);
    }
}

}  // extern "C"

// The cv::Mat expression forms the reference sources use, evaluated through oracle/shim_matcher/cvmat_shim.hpp, so that
// tests/test_cpu_refmatcher.py can compare the shim with the real OpenCV (tests/golden/cv2_matops.npz, made with cv2).
//   op 0: A*B   1: A*B + C   2: A.t()*B   3: -A.t()*B   4: -A*B   5: norm(A)   6: norm(A, B, NORM_L1)   7: A.dot(B)
// out receives the result matrix (row major) or out[0] for the scalar forms; returns the number of floats written.
extern "C" int refm_mat_eval(int op, const float* a, int ar, int ac, const float* b, int br, int bc, const float* c,
                             float* out, double* scalar) {
    cv::Mat A(ar, ac, CV_32F, (void*)a), B, C;
    if (b) B = cv::Mat(br, bc, CV_32F, (void*)b);
    cv::Mat D;
    switch (op) {
    case 0: D = A * B; break;
    case 1: C = cv::Mat(ar, bc, CV_32F, (void*)c); D = A * B + C; break;
    case 2: D = A.t() * B; break;
    case 3: D = -A.t() * B; break;
    case 4: D = -A * B; break;
    case 5: *scalar = cv::norm(A); return 0;
    case 6: *scalar = cv::norm(A, B, cv::NORM_L1); return 0;
    case 7: *scalar = A.dot(B); return 0;
    default: return -1;
    }
    for (int i = 0; i < D.rows; i++)
        for (int j = 0; j < D.cols; j++) out[i * D.cols + j] = D.at<float>(i, j);
    return D.rows * D.cols;
}
