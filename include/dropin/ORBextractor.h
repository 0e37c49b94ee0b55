// include/dropin/ORBextractor.h — header-only C++ drop-in for ORB_SLAM2::ORBextractor over the liborbx C ABI (orbx.h).
//
// Put this directory ahead of the reference's include/ on the include path and link liborbx.so: Frame.cc / Tracking.cc
// compile unchanged (same constructor, operator(), getters, public mvImagePyramid). Replaces
// /root/reference/include/ORBextractor.h:45-112 and /root/reference/src/ORBextractor.cc:405-947.
// All compute runs in the CUDA kernels behind orbx_extract; this file only marshals buffers. There is no CPU path:
// when the library reports an error the wrapper throws std::runtime_error (the reference would cv::Exception).
#ifndef ORBX_DROPIN_ORBEXTRACTOR_H
#define ORBX_DROPIN_ORBEXTRACTOR_H

#include <algorithm>
#include <cassert>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv2/core/core.hpp>

#include "../orbx.h"

namespace ORB_SLAM2 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // 8-argument form of this fork (include/ORBextractor.h:51); zeros select 31/15/19 like src/Tracking.cc:118-123.
    ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_, int patchSize_,
                 int halfPatchSize_, int edgeThreshold_, int cudaDevice = 0)
        : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_),
          minThFAST(minThFAST_), patchSize(patchSize_ ? patchSize_ : 31),
          halfPatchSize(halfPatchSize_ ? halfPatchSize_ : 15), edgeThreshold(edgeThreshold_ ? edgeThreshold_ : 19) {
        orbx_params p = {nfeatures, scaleFactor_, nlevels, iniThFAST, minThFAST, patchSize, halfPatchSize, edgeThreshold};
        check(orbx_create(&p, cudaDevice, &h_), "orbx_create");
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        mnFeaturesPerLevel.resize(nlevels); umax.resize(halfPatchSize + 1);
        check(orbx_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                          mvInvLevelSigma2.data(), mnFeaturesPerLevel.data(), umax.data()), "orbx_tables");
        mvImagePyramid.resize(nlevels);
    }
    // upstream ORB-SLAM2 5-argument form (what BASELINE.json's north_star writes)
    ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_)
        : ORBextractor(nfeatures_, scaleFactor_, nlevels_, iniThFAST_, minThFAST_, 31, 15, 19) {}
    ~ORBextractor() { orbx_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // When false, operator() skips the device->host copy of the pyramid (callers that run stereo matching on the device
    // through ComputeStereoMatchesDevice do not need it). Default true = exact drop-in behaviour.
    bool keepHostPyramid = true;

    // src/ORBextractor.cc:858-919. mask is ignored as in the reference.
    void operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                    cv::OutputArray _descriptors) {
        if (_image.empty()) return;  // outputs untouched (:861-862)
        cv::Mat image = _image.getMat();
        assert(inputFormat_ != ORBX_FMT_GRAY8 || image.type() == CV_8UC1);
        int cap = 0;
        check(orbx_max_keypoints(h_, outW_ ? outW_ : image.cols, outW_ ? outH_ : image.rows, &cap), "orbx_max_keypoints");
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_kp), "cv::KeyPoint must be 28 bytes");
        kp_.resize(cap);
        desc_.resize((size_t)cap * 32);
        int n = 0;
        check(orbx_extract(h_, image.data, image.cols, image.rows, image.step, kp_.data(), desc_.data(), cap, &n),
              "orbx_extract");
        lastN_ = n;
        _keypoints.clear();
        if (n == 0) {
            _descriptors.release();
        } else {
            _descriptors.create(n, 32, CV_8U);
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; i++) std::memcpy(d.ptr(i), desc_.data() + (size_t)i * 32, 32);
            _keypoints.resize(n);
            std::memcpy((void*)_keypoints.data(), kp_.data(), (size_t)n * sizeof(orbx_kp));
        }
        if (keepHostPyramid) refreshPyramid();
    }

    // ---- upstream image prep on the device (what Tracking::GrabImage* and the EuRoC replay do before this call) ----
    // Colour images: pass the CV_8UC3 / CV_8UC4 frame to operator() instead of calling cv::cvtColor first
    // (src/Tracking.cc:269-294, 308-321, 337-350); fmt = ORBX_FMT_{RGB8,BGR8,RGBA8,BGRA8} chosen from mbRGB / channels().
    void SetInputFormat(int fmt) {
        check(orbx_set_input_format(h_, fmt), "orbx_set_input_format");
        inputFormat_ = fmt;
    }
    // Unrectified images: hand over M1, M2 of cv::initUndistortRectifyMap(.., CV_32F, M1, M2) once and pass the raw
    // frames; replaces cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) (Test/Replay/Stereo/stereo_euroc.cc:136-137).
    void SetRectifyMaps(const cv::Mat& M1, const cv::Mat& M2) {
        if (M1.empty()) {
            check(orbx_set_rectify_maps(h_, nullptr, nullptr, 0, 0), "orbx_set_rectify_maps");
            outW_ = outH_ = 0;
            return;
        }
        cv::Mat a = M1.clone(), b = M2.clone();  // continuous rows
        check(orbx_set_rectify_maps(h_, (const float*)a.data, (const float*)b.data, M1.cols, M1.rows), "orbx_set_rectify_maps");
        outW_ = M1.cols;
        outH_ = M1.rows;
    }
    // Frame::UndistortKeyPoints (src/Frame.cc:402-432) for the keypoints of the last operator() call.
    void UndistortKeyPoints(const cv::Mat& K, const cv::Mat& distCoef, std::vector<cv::KeyPoint>& keysUn) {
        const float K4[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
        float d[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const int nd = std::min(8, distCoef.rows * distCoef.cols);
        for (int i = 0; i < nd; i++) d[i] = distCoef.rows > 1 ? distCoef.at<float>(i, 0) : distCoef.at<float>(0, i);
        std::vector<orbx_kp> un(kp_.size());
        check(orbx_undistort_keypoints(h_, K4, d, nd, un.data(), (int)un.size()), "orbx_undistort_keypoints");
        keysUn.resize(lastN_);
        if (lastN_) std::memcpy((void*)keysUn.data(), un.data(), (size_t)lastN_ * sizeof(orbx_kp));
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }
    // north_star's accessor (not in this fork): the pyramid of the last image
    const std::vector<cv::Mat>& GetImagePyramid() { return mvImagePyramid; }

    // Public data member read by Frame::ComputeStereoMatches (src/Frame.cc:471,565,577,582): level interiors as ROIs
    // at (edgeThreshold, edgeThreshold) of padded BORDER_REFLECT_101 buffers, exactly like src/ORBextractor.cc:927-942.
    std::vector<cv::Mat> mvImagePyramid;

    orbx_t* handle() { return h_; }  // for ComputeStereoMatchesDevice / batch calls on the C ABI

protected:
    // mvImagePyramid after operator(): every level as an ROI at (edgeThreshold, edgeThreshold) of its padded buffer. One
    // export kernel and one device-to-host copy for all levels (orbx_pyramid_all) into one host block.
    void refreshPyramid() {
        size_t bytes = 0;
        check(orbx_pyramid_bytes(h_, 1, &bytes), "orbx_pyramid_bytes");
        if (pyr_.empty() || (size_t)pyr_.rows * pyr_.cols < bytes) pyr_.create(1, (int)bytes, CV_8UC1);
        std::vector<size_t> off(nlevels);
        check(orbx_pyramid_all(h_, 0, pyr_.data, bytes, 1, off.data()), "orbx_pyramid_all");
        for (int l = 0; l < nlevels; l++) {
            int w = 0, hgt = 0;
            check(orbx_level_size(h_, l, &w, &hgt), "orbx_level_size");
            const int W = w + 2 * edgeThreshold, H = hgt + 2 * edgeThreshold;
            cv::Mat padded(H, W, CV_8UC1, pyr_.data + off[l], (size_t)W);
            mvImagePyramid[l] = padded(cv::Rect(edgeThreshold, edgeThreshold, w, hgt));
        }
    }
    static void check(int rc, const char* what) {
        if (rc != ORBX_OK) throw std::runtime_error(std::string(what) + ": " + orbx_last_error_string());
    }

    int inputFormat_ = ORBX_FMT_GRAY8, outW_ = 0, outH_ = 0, lastN_ = 0;
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    int patchSize;
    int halfPatchSize;
    int edgeThreshold;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<int> umax;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

private:
    orbx_t* h_ = nullptr;
    std::vector<orbx_kp> kp_;
    std::vector<uint8_t> desc_;
    cv::Mat pyr_;  // all padded levels back to back (mvImagePyramid[l] are ROIs into it)
};

// Frame::ComputeStereoMatches (src/Frame.cc:464-642) on the device state of two extractors that have just processed the
// left and right image. mvuRight / mvDepth are resized to the left keypoint count and filled (-1 = no match).
inline void ComputeStereoMatchesDevice(ORBextractor& left, ORBextractor& right, int nLeft, float mbf, float minD,
                                       float maxD, std::vector<float>& mvuRight, std::vector<float>& mvDepth) {
    mvuRight.assign(nLeft, -1.0f);
    mvDepth.assign(nLeft, -1.0f);
    if (nLeft == 0) return;
    if (orbx_stereo_match(left.handle(), right.handle(), 0, mbf, minD, maxD, mvuRight.data(), mvDepth.data()) != ORBX_OK)
        throw std::runtime_error(std::string("orbx_stereo_match: ") + orbx_last_error_string());
}

}  // namespace ORB_SLAM2

#endif
