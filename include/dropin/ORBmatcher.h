// include/dropin/ORBmatcher.h — C++ wrapper of the descriptor-distance surface of ORB_SLAM2::ORBmatcher
// (/root/reference/include/ORBmatcher.h:38-102) over the liborbx C ABI (orbx.h).
//
// What moves to the device is the part of every search that is data parallel: DescriptorDistance
// (src/ORBmatcher.cc:1643-1659) and the "scan candidates -> best / second best with strict '<'" loop that all twelve
// searches share. The geometric gating around it (projection, frustum, epipolar tests, greedy
// bookkeeping) stays in the caller, which hands candidate lists in CSR form to ScanCandidates and applies its own
// accept rule with the helpers below — except for SearchByProjection(Frame, Frame), the per-frame tracking search,
// which runs end to end on the device (grid, windows, scan, greedy claims, rotation filter). The per-thread matcher handle makes the class re-entrant like the reference's
// stack-allocated ORBmatcher objects (src/Tracking.cc:862, src/LocalMapping.cc:215, src/LoopClosing.cc:275).
#ifndef ORBX_DROPIN_ORBMATCHER_H
#define ORBX_DROPIN_ORBMATCHER_H

#include <cmath>
#include <cstdint>
#include <cstring>
#include <set>
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
    // One pair stays on the host (eight 32-bit popcounts, the reference's own bit trick): the reference calls this
    // static inside O(N^2) host loops (MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:280-290), where a
    // device round trip per pair would be orders of magnitude slower. The batched overload below runs on the device.
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
        const unsigned char *pa = a.ptr(), *pb = b.ptr();
        int dist = 0;
        for (int i = 0; i < 8; i++) {
            uint32_t x, y;
            std::memcpy(&x, pa + 4 * i, 4);
            std::memcpy(&y, pb + 4 * i, 4);
            uint32_t v = x ^ y;
            v = v - ((v >> 1) & 0x55555555u);
            v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
            dist += (int)((((v + (v >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
        }
        return dist;
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

    // Batched MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307): vDescriptors of many map points back to
    // back (point p owns rows off[p] .. off[p+1]-1, 1x32 CV_8U each, in the reference's iteration order);
    // bestIdx[p] = the row (relative to off[p]) whose clone becomes pMP->mDescriptor, -1 if the point has no rows.
    static void ComputeDistinctiveDescriptors(const std::vector<cv::Mat>& vDescriptors, const std::vector<int32_t>& off,
                                              std::vector<int32_t>& bestIdx) {
        const int np = (int)off.size() - 1;
        bestIdx.assign(np > 0 ? np : 0, -1);
        if (np <= 0) return;
        std::vector<uint8_t> flat(vDescriptors.size() * 32);
        for (size_t r = 0; r < vDescriptors.size(); r++)
            for (int k = 0; k < 32; k++) flat[r * 32 + k] = vDescriptors[r].ptr()[k];
        check(orbm_distinctive_descriptors(tls(), flat.data(), off.data(), np, bestIdx.data(), nullptr),
              "orbm_distinctive_descriptors");
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

    // ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono)
    // (src/ORBmatcher.cc:1324-1466, called by Tracking::TrackWithMotionModel, src/Tracking.cc:985,991) on the device:
    // the current frame's grid, GetFeaturesInArea, the Hamming scan, the greedy claims and the rotation filter all run
    // in orbm_search_by_projection; this wrapper only flattens the Frame / MapPoint objects into arrays and writes the
    // resulting mvpMapPoints back. FrameT is ORB_SLAM2::Frame (a template so this header needs none of its includes).
    template <class FrameT>
    int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono) {
        orbm_proj_params P = orbm_proj_params();
        P.fx = CurrentFrame.fx; P.fy = CurrentFrame.fy; P.cx = CurrentFrame.cx; P.cy = CurrentFrame.cy;
        P.bf = CurrentFrame.mbf;
        P.minX = CurrentFrame.mnMinX; P.maxX = CurrentFrame.mnMaxX; P.minY = CurrentFrame.mnMinY; P.maxY = CurrentFrame.mnMaxY;
        P.grid_w_inv = CurrentFrame.mfGridElementWidthInv; P.grid_h_inv = CurrentFrame.mfGridElementHeightInv;
        P.nlevels = (int)CurrentFrame.mvScaleFactors.size();
        for (int i = 0; i < P.nlevels && i < ORBX_MAX_LEVELS; i++) P.scale[i] = CurrentFrame.mvScaleFactors[i];
        const cv::Mat& Tcw = CurrentFrame.mTcw;
        const cv::Mat& Tlw = LastFrame.mTcw;
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) P.Rcw[3 * r + c] = Tcw.template at<float>(r, c);
            P.tcw[r] = Tcw.template at<float>(r, 3);
        }
        // twc = -Rcw.t()*tcw; tlc = Rlw*twc + tlw (src/ORBmatcher.cc:1334-1341). cv::gemm evaluates a transposed-operand
        // product with a double accumulator and a plain 3x3 * 3x1 product in float (see row3 / gemm_add below).
        float twc[3], tlcz;
        for (int r = 0; r < 3; r++) {
            double a = 0;
            for (int k = 0; k < 3; k++) a += (double)(-Tcw.template at<float>(k, r)) * (double)P.tcw[k];
            twc[r] = (float)a;
        }
        {
            const float Rl2[3] = {Tlw.template at<float>(2, 0), Tlw.template at<float>(2, 1), Tlw.template at<float>(2, 2)};
            tlcz = gemm_add(row3(Rl2, twc), Tlw.template at<float>(2, 3));
        }
        P.forward = (tlcz > CurrentFrame.mb && !bMono) ? 1 : 0;
        P.backward = (-tlcz > CurrentFrame.mb && !bMono) ? 1 : 0;
        P.th = th; P.th_high = TH_HIGH; P.check_orientation = mbCheckOrientation ? 1 : 0;

        const int nL = LastFrame.N, nC = CurrentFrame.N;
        std::vector<float> wp((size_t)nL * 3, 0.f), ang(nL, 0.f);
        std::vector<uint8_t> flags(nL, 0), ld((size_t)nL * 32, 0), taken(nC, 0);
        std::vector<int32_t> oct(nL, 0);
        for (int i = 0; i < nL; i++) {
            auto* pMP = LastFrame.mvpMapPoints[i];
            oct[i] = LastFrame.mvKeys[i].octave;
            ang[i] = LastFrame.mvKeysUn[i].angle;
            if (!pMP || LastFrame.mvbOutlier[i]) continue;
            flags[i] = (uint8_t)(1 | (pMP->Observations() > 0 ? 2 : 0));
            const cv::Mat x3Dw = pMP->GetWorldPos();
            for (int k = 0; k < 3; k++) wp[3 * i + k] = x3Dw.template at<float>(k, 0);
            const cv::Mat d = pMP->GetDescriptor();
            for (int k = 0; k < 32; k++) ld[(size_t)i * 32 + k] = d.ptr()[k];
        }
        for (int i = 0; i < nC; i++)
            if (CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0) taken[i] = 1;
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_kp), "cv::KeyPoint layout");
        orbm_last_frame L = {nL, wp.data(), flags.data(), oct.data(), ang.data(), ld.data()};
        orbm_cur_frame Cf = {nC, (const orbx_kp*)CurrentFrame.mvKeysUn.data(),
                             CurrentFrame.mvuRight.empty() ? nullptr : CurrentFrame.mvuRight.data(),
                             CurrentFrame.mDescriptors.ptr(), taken.data()};
        std::vector<int32_t> curMatch(nC), lastMatch(nL);
        std::vector<uint8_t> lastKeep(nL);
        int nmatches = 0;
        check(orbm_search_by_projection(tls(), &P, &L, &Cf, curMatch.data(), lastMatch.data(), lastKeep.data(), &nmatches),
              "orbm_search_by_projection");
        // write-back: claims first (a later claim overwrites), then the rotation filter's NULLs — cur_match already
        // holds the net effect per keypoint, except that a filtered claim also clears a pointer that was there on entry
        for (int i = 0; i < nL; i++)
            if (lastMatch[i] >= 0 && !lastKeep[i]) CurrentFrame.mvpMapPoints[lastMatch[i]] = nullptr;
        for (int i2 = 0; i2 < nC; i2++)
            if (curMatch[i2] >= 0) CurrentFrame.mvpMapPoints[i2] = LastFrame.mvpMapPoints[curMatch[i2]];
        return nmatches;
    }

    // ORBmatcher::SearchByProjection(Frame &F, const std::vector<MapPoint*> &vpMapPoints, const float th = 3)
    // (src/ORBmatcher.cc:43-127, Tracking::SearchLocalPoints, src/Tracking.cc:1290) on the device.
    template <class FrameT, class MapPointT>
    int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3) {
        orbm_proj_params P = orbm_proj_params();
        P.minX = F.mnMinX; P.maxX = F.mnMaxX; P.minY = F.mnMinY; P.maxY = F.mnMaxY;
        P.grid_w_inv = F.mfGridElementWidthInv; P.grid_h_inv = F.mfGridElementHeightInv;
        P.nlevels = (int)F.mvScaleFactors.size();
        for (int i = 0; i < P.nlevels && i < ORBX_MAX_LEVELS; i++) P.scale[i] = F.mvScaleFactors[i];
        P.th = th; P.th_high = TH_HIGH;
        const int n = (int)vpMapPoints.size(), nC = F.N;
        std::vector<float> px(n, 0.f), py(n, 0.f), pxr(n, 0.f), vc(n, 0.f);
        std::vector<int32_t> lvl(n, 0);
        std::vector<uint8_t> flags(n, 0), desc((size_t)n * 32, 0), taken(nC, 0);
        for (int i = 0; i < n; i++) {
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP->mbTrackInView || pMP->isBad()) continue;
            flags[i] = (uint8_t)(1 | (pMP->Observations() > 0 ? 2 : 0));
            px[i] = pMP->mTrackProjX; py[i] = pMP->mTrackProjY; pxr[i] = pMP->mTrackProjXR;
            lvl[i] = pMP->mnTrackScaleLevel; vc[i] = pMP->mTrackViewCos;
            const cv::Mat d = pMP->GetDescriptor();
            for (int k = 0; k < 32; k++) desc[(size_t)i * 32 + k] = d.ptr()[k];
        }
        for (int i = 0; i < nC; i++)
            if (F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0) taken[i] = 1;
        orbm_map_points M = {n, px.data(), py.data(), pxr.data(), lvl.data(), vc.data(), flags.data(), desc.data()};
        orbm_cur_frame Cf = {nC, (const orbx_kp*)F.mvKeysUn.data(), F.mvuRight.empty() ? nullptr : F.mvuRight.data(),
                             F.mDescriptors.ptr(), taken.data()};
        std::vector<int32_t> curMatch(nC), mpMatch(n);
        int nmatches = 0;
        check(orbm_search_local_points(tls(), &P, mfNNratio, &M, &Cf, curMatch.data(), mpMatch.data(), &nmatches),
              "orbm_search_local_points");
        for (int i2 = 0; i2 < nC; i2++)
            if (curMatch[i2] >= 0) F.mvpMapPoints[i2] = vpMapPoints[curMatch[i2]];
        return nmatches;
    }

    // ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound,
    //                                const float th, const int ORBdist)
    // (src/ORBmatcher.cc:1468-1595; Tracking::Relocalization src/Tracking.cc:1552,1566) on the device. The per-point
    // gates that need MapPoint internals (scale-invariance range, PredictScale) stay on the host; projection, window
    // search, greedy claims and the rotation filter run in orbm_search_by_projection.
    template <class FrameT, class KeyFrameT, class SetT>
    int SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const SetT& sAlreadyFound, const float th, const int ORBdist) {
        orbm_proj_params P = orbm_proj_params();
        P.fx = CurrentFrame.fx; P.fy = CurrentFrame.fy; P.cx = CurrentFrame.cx; P.cy = CurrentFrame.cy; P.bf = CurrentFrame.mbf;
        P.minX = CurrentFrame.mnMinX; P.maxX = CurrentFrame.mnMaxX; P.minY = CurrentFrame.mnMinY; P.maxY = CurrentFrame.mnMaxY;
        P.grid_w_inv = CurrentFrame.mfGridElementWidthInv; P.grid_h_inv = CurrentFrame.mfGridElementHeightInv;
        P.nlevels = (int)CurrentFrame.mvScaleFactors.size();
        for (int i = 0; i < P.nlevels && i < ORBX_MAX_LEVELS; i++) P.scale[i] = CurrentFrame.mvScaleFactors[i];
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) P.Rcw[3 * r + c] = CurrentFrame.mTcw.template at<float>(r, c);
            P.tcw[r] = CurrentFrame.mTcw.template at<float>(r, 3);
        }
        P.th = th; P.th_high = ORBdist; P.check_orientation = mbCheckOrientation ? 1 : 0; P.skip_depth_test = 1;
        float Ow[3];  // -Rcw.t()*tcw: CV_32F product = double accumulation rounded once
        for (int r = 0; r < 3; r++) {
            double acc = 0;
            for (int c = 0; c < 3; c++) acc += (double)P.Rcw[3 * c + r] * (double)P.tcw[c];
            Ow[r] = -(float)acc;
        }
        const auto vpMPs = pKF->GetMapPointMatches();
        const int n = (int)vpMPs.size(), nC = CurrentFrame.N;
        std::vector<float> wp((size_t)n * 3, 0.f), ang(n, 0.f);
        std::vector<int32_t> lvl(n, 0);
        std::vector<uint8_t> flags(n, 0), desc((size_t)n * 32, 0), taken(nC, 0);
        for (int i = 0; i < n; i++) {
            auto* pMP = vpMPs[i];
            if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
            const cv::Mat x = pMP->GetWorldPos();
            double s = 0;
            for (int k = 0; k < 3; k++) {
                wp[3 * i + k] = x.template at<float>(k, 0);
                const float po = wp[3 * i + k] - Ow[k];
                s += (double)po * (double)po;
            }
            const float dist3D = (float)std::sqrt(s);  // cv::norm(PO)
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            lvl[i] = pMP->PredictScale(dist3D, &CurrentFrame);
            flags[i] = 3;
            ang[i] = pKF->mvKeysUn[i].angle;
            const cv::Mat d = pMP->GetDescriptor();
            for (int k = 0; k < 32; k++) desc[(size_t)i * 32 + k] = d.ptr()[k];
        }
        for (int i = 0; i < nC; i++) taken[i] = CurrentFrame.mvpMapPoints[i] != NULL;
        orbm_last_frame L = {n, wp.data(), flags.data(), lvl.data(), ang.data(), desc.data()};
        orbm_cur_frame Cf = {nC, (const orbx_kp*)CurrentFrame.mvKeysUn.data(), nullptr, CurrentFrame.mDescriptors.ptr(), taken.data()};
        std::vector<int32_t> curMatch(nC), kfMatch(n);
        std::vector<uint8_t> keep(n);
        int nmatches = 0;
        check(orbm_search_by_projection(tls(), &P, &L, &Cf, curMatch.data(), kfMatch.data(), keep.data(), &nmatches),
              "orbm_search_by_projection");
        for (int i2 = 0; i2 < nC; i2++)
            if (curMatch[i2] >= 0) CurrentFrame.mvpMapPoints[i2] = vpMPs[curMatch[i2]];
        return nmatches;
    }

    // ORBmatcher::Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th = 3.0)
    // (src/ORBmatcher.cc:821-971; LocalMapping::SearchInNeighbors src/LocalMapping.cc:487,512). Projection and the gates
    // that need MapPoint internals run on the host in the reference's arithmetic, the window search (grid,
    // GetFeaturesInArea, octave and chi-square gates, Hamming scan) in orbm_fuse_search, the map mutations on the host
    // in point order. (The Sim3 overload, :973-1096, uses the same entry point with check_chi2 = 0.)
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th = 3.0) {
        const cv::Mat Rcw = pKF->GetRotation(), tcw = pKF->GetTranslation(), Ow = pKF->GetCameraCenter();
        const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy, bf = pKF->mbf;
        const int nMPs = (int)vpMapPoints.size(), nK = pKF->N;
        std::vector<float> pu(nMPs, 0.f), pv(nMPs, 0.f), pur(nMPs, 0.f);
        std::vector<int32_t> lvl(nMPs, 0);
        std::vector<uint8_t> valid(nMPs, 0), desc((size_t)nMPs * 32, 0);
        for (int i = 0; i < nMPs; i++) {
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP || pMP->isBad()) continue;  // IsInKeyFrame is re-checked below: earlier fusions can change it
            const cv::Mat p3Dw = pMP->GetWorldPos();
            float pc[3], w[3];
            for (int k = 0; k < 3; k++) w[k] = p3Dw.template at<float>(k, 0);
            for (int r = 0; r < 3; r++) {  // Rcw*p3Dw + tcw (cv::gemm small-matrix path, see row3)
                const float Rr[3] = {Rcw.template at<float>(r, 0), Rcw.template at<float>(r, 1), Rcw.template at<float>(r, 2)};
                pc[r] = gemm_add(row3(Rr, w), tcw.template at<float>(r, 0));
            }
            if (pc[2] < 0.0f) continue;
            const float invz = 1 / pc[2];
            const float u = fx * (pc[0] * invz) + cx, v = fy * (pc[1] * invz) + cy;
            if (!pKF->IsInImage(u, v)) continue;
            double n2 = 0, dot = 0;
            const cv::Mat Pn = pMP->GetNormal();
            for (int k = 0; k < 3; k++) {
                const float po = w[k] - Ow.template at<float>(k, 0);
                n2 += (double)po * (double)po;
                dot += (double)po * (double)Pn.template at<float>(k, 0);
            }
            const float dist3D = (float)std::sqrt(n2);
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            if (dot < 0.5 * dist3D) continue;
            lvl[i] = pMP->PredictScale(dist3D, pKF);
            pu[i] = u; pv[i] = v; pur[i] = u - bf * invz;
            valid[i] = 1;
            const cv::Mat d = pMP->GetDescriptor();
            for (int k = 0; k < 32; k++) desc[(size_t)i * 32 + k] = d.ptr()[k];
        }
        orbm_proj_params P = orbm_proj_params();
        P.minX = pKF->mnMinX; P.maxX = pKF->mnMaxX; P.minY = pKF->mnMinY; P.maxY = pKF->mnMaxY;
        P.grid_w_inv = pKF->mfGridElementWidthInv; P.grid_h_inv = pKF->mfGridElementHeightInv;
        P.nlevels = (int)pKF->mvScaleFactors.size();
        for (int i = 0; i < P.nlevels && i < ORBX_MAX_LEVELS; i++) P.scale[i] = pKF->mvScaleFactors[i];
        P.th = th;
        orbm_fuse_points pts = {nMPs, pu.data(), pv.data(), pur.data(), lvl.data(), valid.data(), desc.data()};
        orbm_cur_frame kf = {nK, (const orbx_kp*)pKF->mvKeysUn.data(), pKF->mvuRight.empty() ? nullptr : pKF->mvuRight.data(),
                             pKF->mDescriptors.ptr(), nullptr};
        std::vector<int32_t> bestIdx(nMPs, -1);
        std::vector<uint16_t> bestDist(nMPs, 256);
        check(orbm_keyframe_hint(tls(), (uint64_t)pKF->mnId + 1), "orbm_keyframe_hint");  // keyframe arrays are immutable: cache
        check(orbm_fuse_search(tls(), &P, pKF->mvInvLevelSigma2.data(), 1, &pts, &kf, bestIdx.data(), bestDist.data()),
              "orbm_fuse_search");
        int nFused = 0;
        for (int i = 0; i < nMPs; i++) {
            MapPointT* pMP = vpMapPoints[i];
            if (!valid[i] || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
            if (bestIdx[i] < 0 || bestDist[i] > TH_LOW) continue;
            MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx[i]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) {
                    if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                    else pMPinKF->Replace(pMP);
                }
            } else {
                pMP->AddObservation(pKF, bestIdx[i]);
                pKF->AddMapPoint(pMP, bestIdx[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
    //                                     std::vector<int>& vnMatches12, int windowSize = 10)
    // (include/ORBmatcher.h:68, src/ORBmatcher.cc:400-515; Tracking::MonocularInitialization src/Tracking.cc:699)
    // entirely on the device: grid of F2, level-0 windows, the claim loop with stealing, rotation filter and the
    // vbPrevMatched update (orbm_search_for_initialization).
    template <class FrameT>
    int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                int windowSize = 10) {
        static_assert(sizeof(cv::Point2f) == 8 && sizeof(int) == 4, "vbPrevMatched / vnMatches12 cross the ABI in place");
        const int n1 = (int)F1.mvKeysUn.size(), n2 = (int)F2.mvKeysUn.size();
        vnMatches12 = std::vector<int>(n1, -1);
        orbm_proj_params P = orbm_proj_params();
        P.minX = F2.mnMinX; P.maxX = F2.mnMaxX; P.minY = F2.mnMinY; P.maxY = F2.mnMaxY;
        P.grid_w_inv = F2.mfGridElementWidthInv; P.grid_h_inv = F2.mfGridElementHeightInv;
        P.nlevels = (int)F2.mvScaleFactors.size();
        for (int i = 0; i < P.nlevels && i < ORBX_MAX_LEVELS; i++) P.scale[i] = F2.mvScaleFactors[i];
        P.check_orientation = mbCheckOrientation ? 1 : 0;
        orbm_cur_frame f2 = {n2, (const orbx_kp*)F2.mvKeysUn.data(), nullptr, F2.mDescriptors.ptr(), nullptr};
        int nmatches = 0;
        check(orbm_search_for_initialization(tls(), &P, mfNNratio, TH_LOW, windowSize, n1, (const orbx_kp*)F1.mvKeysUn.data(),
                                             F1.mDescriptors.ptr(), n1 ? &vbPrevMatched[0].x : nullptr, &f2,
                                             (int32_t*)vnMatches12.data(), &nmatches),
              "orbm_search_for_initialization");
        return nmatches;
    }

    // ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
    //                                std::vector<MapPoint*>& vpMatched, int th)
    // (include/ORBmatcher.h:64, src/ORBmatcher.cc:285-398; LoopClosing::ComputeSim3 src/LoopClosing.cc:411). The Sim3
    // projection and the MapPoint gates run on the host (project_sim3 below), the window search with the vpMatched
    // claims in orbm_search_kf_projection.
    template <class KeyFrameT, class MapPointT>
    int SearchByProjection(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints,
                           std::vector<MapPointT*>& vpMatched, int th) {
        const std::set<MapPointT*> spAlreadyFound(vpMatched.begin(), vpMatched.end());  // a NULL entry never matches a point
        Sim3Points S;
        project_sim3(pKF, Scw, vpPoints, spAlreadyFound, S);
        const int n = (int)vpPoints.size(), nK = (int)pKF->mvKeysUn.size();
        std::vector<uint8_t> taken(nK, 0);
        for (int i = 0; i < nK && i < (int)vpMatched.size(); i++) taken[i] = vpMatched[i] != NULL;
        orbm_proj_params P = kf_grid_params(pKF, (float)th);
        P.th_high = TH_LOW;
        orbm_fuse_points pts = {n, S.u.data(), S.v.data(), nullptr, S.lvl.data(), S.valid.data(), S.desc.data()};
        orbm_cur_frame kf = {nK, (const orbx_kp*)pKF->mvKeysUn.data(), nullptr, pKF->mDescriptors.ptr(), taken.data()};
        std::vector<int32_t> kfMatch(nK, -1), ptMatch(n, -1);
        int nmatches = 0;
        check(orbm_search_kf_projection(tls(), &P, &pts, &kf, kfMatch.data(), ptMatch.data(), &nmatches),
              "orbm_search_kf_projection");
        for (int idx = 0; idx < nK; idx++)
            if (kfMatch[idx] >= 0) vpMatched[idx] = vpPoints[kfMatch[idx]];
        return nmatches;
    }

    // ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
    //                  std::vector<MapPoint*>& vpReplacePoint)
    // (include/ORBmatcher.h:85, src/ORBmatcher.cc:973-1096; LoopClosing::SearchAndFuse src/LoopClosing.cc:633): no point
    // depends on another one's outcome here (spAlreadyFound is taken once, :989), so one orbm_fuse_search pass serves
    // all points; the map mutations follow on the host in point order.
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, float th,
             std::vector<MapPointT*>& vpReplacePoint) {
        const std::set<MapPointT*> spAlreadyFound = pKF->GetMapPoints();
        Sim3Points S;
        project_sim3(pKF, Scw, vpPoints, spAlreadyFound, S);
        const int n = (int)vpPoints.size(), nK = (int)pKF->mvKeysUn.size();
        orbm_proj_params P = kf_grid_params(pKF, th);
        std::vector<float> ur(n, 0.f), invSigma2(ORBX_MAX_LEVELS, 0.f);
        orbm_fuse_points pts = {n, S.u.data(), S.v.data(), ur.data(), S.lvl.data(), S.valid.data(), S.desc.data()};
        orbm_cur_frame kf = {nK, (const orbx_kp*)pKF->mvKeysUn.data(), nullptr, pKF->mDescriptors.ptr(), nullptr};
        std::vector<int32_t> bestIdx(n, -1);
        std::vector<uint16_t> bestDist(n, 256);
        check(orbm_keyframe_hint(tls(), (uint64_t)pKF->mnId + 1), "orbm_keyframe_hint");
        check(orbm_fuse_search(tls(), &P, invSigma2.data(), 0, &pts, &kf, bestIdx.data(), bestDist.data()), "orbm_fuse_search");
        int nFused = 0;
        for (int i = 0; i < n; i++) {
            if (!S.valid[i] || bestIdx[i] < 0 || bestDist[i] > TH_LOW) continue;
            MapPointT* pMP = vpPoints[i];
            MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx[i]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
            } else {
                pMP->AddObservation(pKF, bestIdx[i]);
                pKF->AddMapPoint(pMP, bestIdx[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12,
    //                          const cv::Mat& R12, const cv::Mat& t12, const float th)
    // (include/ORBmatcher.h:78, src/ORBmatcher.cc:1098-1322; LoopClosing::ComputeSim3 src/LoopClosing.cc:359): the two
    // directions are independent window searches (orbm_fuse_search without the chi-square gates, accept <= TH_HIGH),
    // then the mutual-agreement pass (:1303-1319).
    template <class KeyFrameT, class MapPointT>
    int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const float& s12,
                     const cv::Mat& R12, const cv::Mat& t12, const float th) {
        float R1w[9], t1w[3], R2w[9], t2w[3], sR12[9], sR21[9], t21[3], t12f[3];
        mat33(pKF1->GetRotation(), R1w); vec3(pKF1->GetTranslation(), t1w);
        mat33(pKF2->GetRotation(), R2w); vec3(pKF2->GetTranslation(), t2w);
        vec3(t12, t12f);
        const float inv = (float)(1.0 / s12);  // (1.0/s12)*R12.t(): the scale reaches the elements as a float factor
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) {
                sR12[3 * r + c] = s12 * R12.at<float>(r, c);
                sR21[3 * r + c] = inv * R12.at<float>(c, r);
            }
        for (int r = 0; r < 3; r++) t21[r] = -row3(sR21 + 3 * r, t12f);  // t21 = -sR21*t12: gemm with alpha = -1
        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
        std::vector<uint8_t> already1(N1, 0), already2(N2, 0);
        for (int i = 0; i < N1; i++) {
            MapPointT* pMP = vpMatches12[i];
            if (pMP) {
                already1[i] = 1;
                const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
                if (idx2 >= 0 && idx2 < N2) already2[idx2] = 1;
            }
        }
        std::vector<int> vnMatch1, vnMatch2;
        sim3_direction(pKF2, pKF1, vpMapPoints1, already1, R1w, t1w, sR21, t21, th, vnMatch1);  // KF1's points into KF2
        sim3_direction(pKF1, pKF1, vpMapPoints2, already2, R2w, t2w, sR12, t12f, th, vnMatch2);  // KF2's points into KF1
        int nFound = 0;
        for (int i1 = 0; i1 < N1; i1++) {
            const int idx2 = vnMatch1[i1];
            if (idx2 >= 0 && vnMatch2[idx2] == i1) {
                vpMatches12[i1] = vpMapPoints2[idx2];
                nFound++;
            }
        }
        return nFound;
    }

    // ORBmatcher::SearchByFboW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches)
    // (src/ORBmatcher.cc:157-283; TrackReferenceKeyFrame src/Tracking.cc:867, Relocalization :1476) on the device.
    // FeatVecT is fbow::fBow2 (std::map<uint32_t, std::vector<uint32_t>>).
    template <class KeyFrameT, class FrameT, class MapPointT>
    int SearchByFboW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
        const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
        vpMapPointMatches = std::vector<MapPointT*>(F.N, static_cast<MapPointT*>(NULL));
        const int nA = (int)vpMapPointsKF.size(), nB = F.N;
        std::vector<uint8_t> elig(nA, 0);
        std::vector<float> angA(nA), angB(nB);
        for (int i = 0; i < nA; i++) {
            elig[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();
            angA[i] = pKF->mvKeysUn[i].angle;
        }
        for (int i = 0; i < nB; i++) angB[i] = F.mvKeys[i].angle;
        BowCsr a(pKF->mFbowFeatVec), b(F.mFbowFeatVec);
        orbm_bow_side A = {nA, elig.data(), angA.data(), pKF->mDescriptors.ptr(), (int)a.id.size(), a.id.data(), a.off.data(), a.feat.data()};
        orbm_bow_side B = {nB, nullptr, angB.data(), F.mDescriptors.ptr(), (int)b.id.size(), b.id.data(), b.off.data(), b.feat.data()};
        std::vector<int32_t> match(nA, -1);
        int nmatches = 0;
        check(orbm_search_by_bow(tls(), &A, &B, a.off.back(), b.off.back(), mfNNratio, TH_LOW, 0, mbCheckOrientation ? 1 : 0,
                                 match.data(), &nmatches), "orbm_search_by_bow");
        for (int i = 0; i < nA; i++)
            if (match[i] >= 0) vpMapPointMatches[match[i]] = vpMapPointsKF[i];
        return nmatches;
    }

    // Batched form of SearchByFboW(KeyFrame*, Frame&, ...) for the candidate loop of Tracking::Relocalization
    // (src/Tracking.cc:1465-1490): every candidate keyframe against the SAME frame in one launch sequence
    // (orbm_search_by_bow_batch; the frame's arrays are staged once). vvpMapPointMatches[k] and the returned
    // nmatches[k] are what the k-th call of the single form would give; NULL / bad keyframes are the caller's to skip.
    // Not part of the reference interface: it replaces the loop around the call, not the call.
    template <class KeyFrameT, class FrameT, class MapPointT>
    std::vector<int> SearchByFboW(const std::vector<KeyFrameT*>& vpKFs, FrameT& F,
                                  std::vector<std::vector<MapPointT*> >& vvpMapPointMatches) {
        const int K = (int)vpKFs.size(), nB = F.N;
        vvpMapPointMatches.assign(K, std::vector<MapPointT*>(nB, static_cast<MapPointT*>(NULL)));
        std::vector<int> nm(K, 0);
        if (K == 0) return nm;
        std::vector<float> angB(nB);
        for (int i = 0; i < nB; i++) angB[i] = F.mvKeys[i].angle;
        BowCsr b(F.mFbowFeatVec);
        std::vector<std::vector<MapPointT*> > mps(K);
        std::vector<std::vector<uint8_t> > elig(K);
        std::vector<std::vector<float> > angA(K);
        std::vector<BowCsr> csr;
        csr.reserve(K);
        std::vector<orbm_bow_side> A(K), B(K);
        std::vector<int> npa(K), npb(K);
        std::vector<std::vector<int32_t> > match(K);
        std::vector<int32_t*> ptrs(K);
        for (int k = 0; k < K; k++) {
            KeyFrameT* pKF = vpKFs[k];
            mps[k] = pKF->GetMapPointMatches();
            const int nA = (int)mps[k].size();
            elig[k].assign(nA, 0);
            angA[k].resize(nA);
            for (int i = 0; i < nA; i++) {
                elig[k][i] = mps[k][i] && !mps[k][i]->isBad();
                angA[k][i] = pKF->mvKeysUn[i].angle;
            }
            csr.push_back(BowCsr(pKF->mFbowFeatVec));
            const BowCsr& a = csr.back();
            orbm_bow_side sa = {nA, elig[k].data(), angA[k].data(), pKF->mDescriptors.ptr(), (int)a.id.size(), a.id.data(), a.off.data(), a.feat.data()};
            orbm_bow_side sb = {nB, nullptr, angB.data(), F.mDescriptors.ptr(), (int)b.id.size(), b.id.data(), b.off.data(), b.feat.data()};
            A[k] = sa; B[k] = sb;
            npa[k] = a.off.back(); npb[k] = b.off.back();
            match[k].assign(nA, -1);
            ptrs[k] = match[k].data();
        }
        check(orbm_search_by_bow_batch(tls(), K, A.data(), B.data(), npa.data(), npb.data(), mfNNratio, TH_LOW, 0,
                                       mbCheckOrientation ? 1 : 0, ptrs.data(), nm.data()), "orbm_search_by_bow_batch");
        for (int k = 0; k < K; k++)
            for (size_t i = 0; i < match[k].size(); i++)
                if (match[k][i] >= 0) vvpMapPointMatches[k][match[k][i]] = mps[k][i];
        return nm;
    }

    // ORBmatcher::SearchByFboW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12)
    // (src/ORBmatcher.cc:517-650; LoopClosing::ComputeSim3 src/LoopClosing.cc:301) on the device.
    template <class KeyFrameT, class MapPointT>
    int SearchByFboW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12) {
        const std::vector<MapPointT*> mp1 = pKF1->GetMapPointMatches(), mp2 = pKF2->GetMapPointMatches();
        vpMatches12 = std::vector<MapPointT*>(mp1.size(), static_cast<MapPointT*>(NULL));
        const int nA = (int)mp1.size(), nB = (int)mp2.size();
        std::vector<uint8_t> e1(nA, 0), e2(nB, 0);
        std::vector<float> angA(nA), angB(nB);
        for (int i = 0; i < nA; i++) { e1[i] = mp1[i] && !mp1[i]->isBad(); angA[i] = pKF1->mvKeysUn[i].angle; }
        for (int i = 0; i < nB; i++) { e2[i] = mp2[i] && !mp2[i]->isBad(); angB[i] = pKF2->mvKeysUn[i].angle; }
        BowCsr a(pKF1->mFbowFeatVec), b(pKF2->mFbowFeatVec);
        orbm_bow_side A = {nA, e1.data(), angA.data(), pKF1->mDescriptors.ptr(), (int)a.id.size(), a.id.data(), a.off.data(), a.feat.data()};
        orbm_bow_side B = {nB, e2.data(), angB.data(), pKF2->mDescriptors.ptr(), (int)b.id.size(), b.id.data(), b.off.data(), b.feat.data()};
        std::vector<int32_t> match(nA, -1);
        int nmatches = 0;
        check(orbm_search_by_bow(tls(), &A, &B, a.off.back(), b.off.back(), mfNNratio, TH_LOW, 1, mbCheckOrientation ? 1 : 0,
                                 match.data(), &nmatches), "orbm_search_by_bow");
        for (int i = 0; i < nA; i++)
            if (match[i] >= 0) vpMatches12[i] = mp2[match[i]];
        return nmatches;
    }

    // ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
    //                                    std::vector<std::pair<size_t, size_t>>& vMatchedPairs, const bool bOnlyStereo)
    // (src/ORBmatcher.cc:652-819; LocalMapping::CreateNewMapPoints src/LocalMapping.cc:268) on the device.
    template <class KeyFrameT>
    int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, cv::Mat F12,
                               std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo) {
        // epipole of KF1's centre in KF2 (:657-664), CV_32F algebra as cv::Mat evaluates it
        // (R2w*Cw + t2w: cv::gemm small-matrix path, see row3 / gemm_add)
        const cv::Mat Cw = pKF1->GetCameraCenter(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
        float C2[3];
        float Cwf[3];
        vec3(Cw, Cwf);
        for (int r = 0; r < 3; r++) {
            const float Rr[3] = {R2w.template at<float>(r, 0), R2w.template at<float>(r, 1), R2w.template at<float>(r, 2)};
            C2[r] = gemm_add(row3(Rr, Cwf), t2w.template at<float>(r, 0));
        }
        const float invz = 1.0f / C2[2];
        orbm_triangulation T = orbm_triangulation();
        T.ex = pKF2->fx * C2[0] * invz + pKF2->cx;
        T.ey = pKF2->fy * C2[1] * invz + pKF2->cy;
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) T.F12[3 * r + c] = F12.at<float>(r, c);
        T.nlevels = (int)pKF2->mvScaleFactors.size();
        for (int i = 0; i < T.nlevels && i < ORBX_MAX_LEVELS; i++) {
            T.scale_factors[i] = pKF2->mvScaleFactors[i];
            T.level_sigma2[i] = pKF2->mvLevelSigma2[i];
        }
        const int nA = pKF1->N, nB = pKF2->N;
        std::vector<uint8_t> eA(nA), eB(nB), sA(nA), sB(nB);
        std::vector<float> angA(nA), angB(nB), ptA(2 * (size_t)nA), ptB(2 * (size_t)nB);
        std::vector<int32_t> octB(nB);
        for (int i = 0; i < nA; i++) {
            sA[i] = pKF1->mvuRight[i] >= 0;
            eA[i] = !pKF1->GetMapPoint(i) && (!bOnlyStereo || sA[i]);
            angA[i] = pKF1->mvKeysUn[i].angle;
            ptA[2 * i] = pKF1->mvKeysUn[i].pt.x; ptA[2 * i + 1] = pKF1->mvKeysUn[i].pt.y;
        }
        for (int i = 0; i < nB; i++) {
            sB[i] = pKF2->mvuRight[i] >= 0;
            eB[i] = !pKF2->GetMapPoint(i) && (!bOnlyStereo || sB[i]);
            angB[i] = pKF2->mvKeysUn[i].angle;
            ptB[2 * i] = pKF2->mvKeysUn[i].pt.x; ptB[2 * i + 1] = pKF2->mvKeysUn[i].pt.y;
            octB[i] = pKF2->mvKeysUn[i].octave;
        }
        T.pt_a = ptA.data(); T.pt_b = ptB.data(); T.octave_b = octB.data(); T.stereo_a = sA.data(); T.stereo_b = sB.data();
        BowCsr a(pKF1->mFbowFeatVec), b(pKF2->mFbowFeatVec);
        orbm_bow_side A = {nA, eA.data(), angA.data(), pKF1->mDescriptors.ptr(), (int)a.id.size(), a.id.data(), a.off.data(), a.feat.data()};
        orbm_bow_side B = {nB, eB.data(), angB.data(), pKF2->mDescriptors.ptr(), (int)b.id.size(), b.id.data(), b.off.data(), b.feat.data()};
        std::vector<int32_t> match(nA, -1);
        int nmatches = 0;
        check(orbm_search_for_triangulation(tls(), &A, &B, a.off.back(), b.off.back(), &T, TH_LOW, mbCheckOrientation ? 1 : 0,
                                            match.data(), &nmatches), "orbm_search_for_triangulation");
        vMatchedPairs.clear();
        vMatchedPairs.reserve(nmatches);
        for (int i = 0; i < nA; i++)
            if (match[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)match[i]));
        return nmatches;
    }

protected:
    // ---- host half of the Sim3 searches (src/ORBmatcher.cc:293-298, 313-351 and 1003-1041) ---------------------------
    struct Sim3Points {
        std::vector<float> u, v;
        std::vector<int32_t> lvl;
        std::vector<uint8_t> valid, desc;
    };
    static void mat33(const cv::Mat& M, float* R) {
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) R[3 * r + c] = M.at<float>(r, c);
    }
    static void vec3(const cv::Mat& M, float* t) {
        for (int r = 0; r < 3; r++) t[r] = M.at<float>(r, 0);
    }
    // How OpenCV evaluates the reference's CV_32F matrix expressions (cv::gemm, core/src/matmul.dispatch.cpp; pinned
    // against the real library by tests/golden/cv2_matops.npz and against the unmodified ORBmatcher.cc by the scene tests):
    //   * A*x (+ b), A 3x3, x 3x1, no transpose flag: the small-matrix path — the row's three products are formed and
    //     summed left to right IN FLOAT (no fused multiply-add), then d = (float)((double)t*alpha + (double)b*beta);
    //   * A.t()*x: the generic path — double accumulator, d = (float)(s*alpha);
    //   * Mat::dot, cv::norm: double accumulators.
    static float mul_rn(float a, float b) {  // a float product that the compiler cannot fuse into the following add
        volatile float p = a * b;
        return p;
    }
    static float row3(const float* a, const float* x) {
        float t = mul_rn(a[0], x[0]);
        t = t + mul_rn(a[1], x[1]);
        t = t + mul_rn(a[2], x[2]);
        return t;
    }
    static float gemm_add(float t, float c) { return (float)((double)t + (double)c); }
    // y = A*x + b
    static void mul_add(const float* A, const float* x, const float* b, float* y) {
        for (int r = 0; r < 3; r++) y[r] = gemm_add(row3(A + 3 * r, x), b[r]);
    }
    template <class KeyFrameT>
    static orbm_proj_params kf_grid_params(KeyFrameT* pKF, float th) {
        orbm_proj_params P = orbm_proj_params();
        P.minX = pKF->mnMinX; P.maxX = pKF->mnMaxX; P.minY = pKF->mnMinY; P.maxY = pKF->mnMaxY;
        P.grid_w_inv = pKF->mfGridElementWidthInv; P.grid_h_inv = pKF->mfGridElementHeightInv;
        P.nlevels = (int)pKF->mvScaleFactors.size();
        for (int i = 0; i < P.nlevels && i < ORBX_MAX_LEVELS; i++) P.scale[i] = pKF->mvScaleFactors[i];
        P.th = th;
        return P;
    }
    // Decompose Scw (:293-298) and run the per-point gates of :313-351 / :1003-1041; a point that passes gets
    // valid = 1, its projection, predicted level and descriptor.
    template <class KeyFrameT, class MapPointT, class SetT>
    static void project_sim3(KeyFrameT* pKF, const cv::Mat& Scw, const std::vector<MapPointT*>& vpPoints, const SetT& spAlreadyFound,
                             Sim3Points& S) {
        const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
        double dot = 0;  // sRcw.row(0).dot(sRcw.row(0)): cv::Mat::dot returns the double sum
        for (int c = 0; c < 3; c++) dot += (double)Scw.at<float>(0, c) * (double)Scw.at<float>(0, c);
        const float scw = (float)std::sqrt(dot);
        const float is = (float)(1.0 / scw);  // Mat / scalar: the elements are scaled by 1/scw as a float factor
        float Rcw[9], tcw[3], Ow[3];
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) Rcw[3 * r + c] = Scw.at<float>(r, c) * is;
            tcw[r] = Scw.at<float>(r, 3) * is;
        }
        for (int r = 0; r < 3; r++) {  // Ow = -Rcw.t()*tcw
            double acc = 0;
            for (int c = 0; c < 3; c++) acc += (double)Rcw[3 * c + r] * (double)tcw[c];
            Ow[r] = -(float)acc;
        }
        const int n = (int)vpPoints.size();
        S.u.assign(n, 0.f); S.v.assign(n, 0.f); S.lvl.assign(n, 0); S.valid.assign(n, 0); S.desc.assign((size_t)n * 32, 0);
        for (int i = 0; i < n; i++) {
            MapPointT* pMP = vpPoints[i];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
            float w[3], pc[3];
            vec3(pMP->GetWorldPos(), w);
            mul_add(Rcw, w, tcw, pc);
            if (pc[2] < 0.0f) continue;
            const float invz = (float)(1.0 / pc[2]);
            const float x = pc[0] * invz, y = pc[1] * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pKF->IsInImage(u, v)) continue;
            const cv::Mat Pn = pMP->GetNormal();
            double n2 = 0, d = 0;
            for (int k = 0; k < 3; k++) {
                const float po = w[k] - Ow[k];
                n2 += (double)po * (double)po;
                d += (double)po * (double)Pn.at<float>(k, 0);
            }
            const float dist = (float)std::sqrt(n2);  // cv::norm(PO)
            if (dist < pMP->GetMinDistanceInvariance() || dist > pMP->GetMaxDistanceInvariance()) continue;
            if (d < 0.5 * dist) continue;
            S.lvl[i] = pMP->PredictScale(dist, pKF);
            S.u[i] = u; S.v[i] = v;
            S.valid[i] = 1;
            const cv::Mat dMP = pMP->GetDescriptor();
            for (int k = 0; k < 32; k++) S.desc[(size_t)i * 32 + k] = dMP.ptr()[k];
        }
    }
    // One direction of SearchBySim3 (:1140-1218 / :1221-1299): the points of the source keyframe go through their own
    // camera pose (Rsw, tsw) and the relative Sim3 (sR, t) into the destination keyframe, whose grid is searched.
    template <class KeyFrameT, class MapPointT>
    int sim3_direction(KeyFrameT* pDst, KeyFrameT* pIntr, const std::vector<MapPointT*>& pts, const std::vector<uint8_t>& already, const float* Rsw,
                       const float* tsw, const float* sR, const float* t, float th, std::vector<int>& vnMatch) {
        const int n = (int)pts.size(), nK = (int)pDst->mvKeysUn.size();
        // both directions project with pKF1's intrinsics (src/ORBmatcher.cc:1101-1104)
        const float fx = pIntr->fx, fy = pIntr->fy, cx = pIntr->cx, cy = pIntr->cy;
        std::vector<float> u(n, 0.f), v(n, 0.f), ur(n, 0.f), invSigma2(ORBX_MAX_LEVELS, 0.f);
        std::vector<int32_t> lvl(n, 0);
        std::vector<uint8_t> valid(n, 0), desc((size_t)n * 32, 0);
        for (int i = 0; i < n; i++) {
            MapPointT* pMP = pts[i];
            if (!pMP || already[i] || pMP->isBad()) continue;
            float w[3], c1[3], c2[3];
            vec3(pMP->GetWorldPos(), w);
            mul_add(Rsw, w, tsw, c1);
            mul_add(sR, c1, t, c2);
            if (c2[2] < 0.0f) continue;
            const float invz = (float)(1.0 / c2[2]);
            const float x = c2[0] * invz, y = c2[1] * invz;
            const float pu = fx * x + cx, pv = fy * y + cy;
            if (!pDst->IsInImage(pu, pv)) continue;
            const float dist3D = (float)std::sqrt((double)c2[0] * c2[0] + (double)c2[1] * c2[1] + (double)c2[2] * c2[2]);
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            lvl[i] = pMP->PredictScale(dist3D, pDst);
            u[i] = pu; v[i] = pv;
            valid[i] = 1;
            const cv::Mat dMP = pMP->GetDescriptor();
            for (int k = 0; k < 32; k++) desc[(size_t)i * 32 + k] = dMP.ptr()[k];
        }
        orbm_proj_params P = kf_grid_params(pDst, th);
        orbm_fuse_points fp = {n, u.data(), v.data(), ur.data(), lvl.data(), valid.data(), desc.data()};
        orbm_cur_frame kf = {nK, (const orbx_kp*)pDst->mvKeysUn.data(), nullptr, pDst->mDescriptors.ptr(), nullptr};
        std::vector<int32_t> bestIdx(n, -1);
        std::vector<uint16_t> bestDist(n, 256);
        if (n) {
            check(orbm_keyframe_hint(tls(), (uint64_t)pDst->mnId + 1), "orbm_keyframe_hint");
            check(orbm_fuse_search(tls(), &P, invSigma2.data(), 0, &fp, &kf, bestIdx.data(), bestDist.data()), "orbm_fuse_search");
        }
        vnMatch.assign(n, -1);
        int found = 0;
        for (int i = 0; i < n; i++)
            if (valid[i] && bestIdx[i] >= 0 && bestDist[i] <= TH_HIGH) { vnMatch[i] = bestIdx[i]; found++; }
        return found;
    }

    // std::map<node, vector<feature>> (fbow::fBow2) flattened to the CSR form of the C ABI
    struct BowCsr {
        std::vector<uint32_t> id, feat;
        std::vector<int> off;
        template <class FeatVecT>
        explicit BowCsr(const FeatVecT& fv) {
            off.push_back(0);
            for (const auto& e : fv) {
                id.push_back(e.first);
                feat.insert(feat.end(), e.second.begin(), e.second.end());
                off.push_back((int)feat.size());
            }
        }
    };
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
