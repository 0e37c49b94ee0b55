// include/dropin/ORBmatcher.h — C++ wrapper of the descriptor-distance surface of ORB_SLAM2::ORBmatcher
// (/root/reference/include/ORBmatcher.h:38-102) over the liborbx C ABI (orbx.h).
//
// What moves to the device is the part of every search that is data parallel: DescriptorDistance
// (src/ORBmatcher.cc:1643-1659) and the "scan candidates -> best / second best with strict '<'" loop that all twelve
// searches share. The geometric gating around it (projection, frustum, epipolar tests, greedy
// bookkeeping) stays in the caller, which hands candidate lists in CSR form to ScanCandidates and applies its own
// accept rule with the helpers below. The per-thread matcher handle makes the class re-entrant like the reference's
// stack-allocated ORBmatcher objects (src/Tracking.cc:862, src/LocalMapping.cc:215, src/LoopClosing.cc:275).
#ifndef ORBX_DROPIN_ORBMATCHER_H
#define ORBX_DROPIN_ORBMATCHER_H

#include <stdexcept>
#include <string>
#include <vector>

#include <opencv2/core/core.hpp>

#include "../orbx.h"

namespace ORB_SLAM2 {

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

    static const int TH_LOW = 50;        // src/ORBmatcher.cc:35-37
    static const int TH_HIGH = 100;
    static const int HISTO_LENGTH = 30;

    // Hamming distance of two 1x32 CV_8U descriptors (src/ORBmatcher.cc:1643-1659).
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
        uint16_t d = 0;
        check(orbm_distance(tls(), a.ptr(), b.ptr(), 1, &d), "orbm_distance");
        return (int)d;
    }
    // Batched form: d[i] = distance(a.row(i), b.row(i)); a and b are continuous N x 32 CV_8U.
    static void DescriptorDistance(const cv::Mat& a, const cv::Mat& b, std::vector<uint16_t>& d) {
        d.resize(a.rows);
        if (a.rows) check(orbm_distance(tls(), a.ptr(), b.ptr(), a.rows, d.data()), "orbm_distance");
    }

    struct Scan {
        std::vector<int32_t> bestIdx, secondIdx;   // -1 = none
        std::vector<uint16_t> bestDist, secondDist;  // 256 = none
    };
    // The shared inner loop: query a scans train rows candIdx[candOff[a] .. candOff[a+1]) in list order, skipping
    // entries with candSkip[c] != 0. First candidate wins ties; secondDist counts multiplicity (else-if rule).
    static void ScanCandidates(const cv::Mat& queries, const std::vector<int32_t>& candOff,
                               const std::vector<int32_t>& candIdx, const std::vector<uint8_t>* candSkip,
                               const cv::Mat& train, Scan& out) {
        const int nq = queries.rows;
        out.bestIdx.resize(nq); out.secondIdx.resize(nq); out.bestDist.resize(nq); out.secondDist.resize(nq);
        if (!nq) return;
        check(orbm_window_search(tls(), queries.ptr(), nq, candOff.data(), candIdx.data(),
                                 candSkip ? candSkip->data() : nullptr, train.ptr(), train.rows, out.bestIdx.data(),
                                 out.bestDist.data(), out.secondDist.data(), out.secondIdx.data()),
              "orbm_window_search");
    }
    // Brute-force 2-NN (candidate set = the whole train matrix), BASELINE config 4.
    static void Knn2(const cv::Mat& queries, const cv::Mat& train, std::vector<int32_t>& i1, std::vector<uint16_t>& d1,
                     std::vector<uint16_t>& d2) {
        const int nq = queries.rows;
        i1.resize(nq); d1.resize(nq); d2.resize(nq);
        if (nq) check(orbm_knn2(tls(), queries.ptr(), nq, train.ptr(), train.rows, i1.data(), d1.data(), d2.data()), "orbm_knn2");
    }

    // Accept rules of the reference, evaluated in fp32 like mfNNratio (include/ORBmatcher.h:100).
    bool AcceptRatio(int best, int second, int th) const {  // SearchByFboW(KF,F): src/ORBmatcher.cc:225-229
        return best <= th && (float)best < mfNNratio * (float)second;
    }
    bool AcceptThreshold(int best, int th) const { return best <= th; }  // e.g. src/ORBmatcher.cc:1422

    // Rotation-consistency helper, src/ORBmatcher.cc:1597-1638 (host side: operates on the accepted matches).
    static void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {
        int max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < L; i++) {
            const int s = (int)histo[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
    }

protected:
    float mfNNratio;
    bool mbCheckOrientation;

private:
    struct Tls {
        orbm_t* m = nullptr;
        ~Tls() { orbm_destroy(m); }
    };
    static orbm_t* tls() {
        static thread_local Tls t;
        if (!t.m) check(orbm_create(0, &t.m), "orbm_create");
        return t.m;
    }
    static void check(int rc, const char* what) {
        if (rc != ORBX_OK) throw std::runtime_error(std::string(what) + ": " + orbx_last_error_string());
    }
};

}  // namespace ORB_SLAM2

#endif
