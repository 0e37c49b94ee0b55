// Compiles the reference-signature C++ wrappers (include/dropin) against the cv shim and drives them the way
// Frame::ExtractORB (/root/reference/src/Frame.cc:248-254) and Frame's stereo constructor (:61-117) do.
// Usage: dropin_run <w> <h> <left.raw> <right.raw> <out.bin>   (raw 8-bit images); built and run by tests/test_gpu_dropin.py
#include <cstdio>
#include <cstdlib>
#include <set>
#include <string>
#include <vector>

#include "dropin/ORBextractor.h"
#include "dropin/ORBmatcher.h"
#include "dropin/fbow.h"

static std::vector<unsigned char> slurp(const char* path, size_t n) {
    std::vector<unsigned char> b(n);
    FILE* f = fopen(path, "rb");
    if (!f || fread(b.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
    fclose(f);
    return b;
}

static void run_initialization(int w, int h, const std::vector<cv::KeyPoint>& kL, const cv::Mat& dL,
                               const std::vector<cv::KeyPoint>& kR, const cv::Mat& dR, const std::vector<float>& sf,
                               const char* out);
static int run_tracking_search(const char* in, const char* out);
static int run_sim3_projection(const char* in, const char* out);
static int run_search_by_sim3(const char* in, const char* out);
static int run_fuse(const char* in, const char* out);
static void run_bow_search(const std::vector<cv::KeyPoint>& kL, const cv::Mat& dL, const fbow::fBow2& fL,
                           const std::vector<cv::KeyPoint>& kR, const cv::Mat& dR, const fbow::fBow2& fR, const char* out);

int main(int argc, char** argv) {
    if (argc == 4 && std::string(argv[1]) == "track") return run_tracking_search(argv[2], argv[3]);
    if (argc == 4 && std::string(argv[1]) == "sim3") return run_sim3_projection(argv[2], argv[3]);
    if (argc == 4 && std::string(argv[1]) == "sim3b") return run_search_by_sim3(argv[2], argv[3]);
    if (argc == 4 && std::string(argv[1]) == "fuse") return run_fuse(argv[2], argv[3]);
    if (argc < 6) return 1;
    const int w = atoi(argv[1]), h = atoi(argv[2]);
    std::vector<unsigned char> L = slurp(argv[3], (size_t)w * h), R = slurp(argv[4], (size_t)w * h);
    cv::Mat imL(h, w, CV_8UC1, L.data(), (size_t)w), imR(h, w, CV_8UC1, R.data(), (size_t)w);
    ORB_SLAM2::ORBextractor exL(1000, 1.2f, 8, 20, 7, 31, 15, 19), exR(1000, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> kL, kR;
    cv::Mat dL, dR;
    exL(imL, cv::Mat(), kL, dL);
    exR(imR, cv::Mat(), kR, dR);
    std::vector<float> uR, depth;
    ORB_SLAM2::ComputeStereoMatchesDevice(exL, exR, (int)kL.size(), 40.0f, 0.0f, 400.0f, uR, depth);
    ORB_SLAM2::ORBmatcher matcher(0.7f);
    std::vector<int32_t> i1; std::vector<uint16_t> d1, d2;
    ORB_SLAM2::ORBmatcher::Knn2(dL, dR, i1, d1, d2);
    int accepted = 0;
    for (size_t i = 0; i < i1.size(); i++) accepted += matcher.AcceptRatio(d1[i], d2[i], ORB_SLAM2::ORBmatcher::TH_LOW);
    const int d01 = kL.size() > 1 ? ORB_SLAM2::ORBmatcher::DescriptorDistance(dL.rowRange(0, 1), dL.rowRange(1, 2)) : -1;
    FILE* f = fopen(argv[5], "wb");
    int hdr[6] = {(int)kL.size(), (int)kR.size(), accepted, d01, exL.GetLevels(), exL.mvImagePyramid[3].cols};
    fwrite(hdr, sizeof(int), 6, f);
    fwrite(kL.data(), sizeof(cv::KeyPoint), kL.size(), f);
    for (int i = 0; i < dL.rows; i++) fwrite(dL.ptr(i), 1, 32, f);
    fwrite(uR.data(), 4, uR.size(), f);
    fwrite(depth.data(), 4, depth.size(), f);
    fwrite(i1.data(), 4, i1.size(), f);
    // level-3 pyramid interior incl. what the ROI sees of its reflect border (first row above the ROI)
    const cv::Mat& p3 = exL.mvImagePyramid[3];
    fwrite(p3.data - p3.step, 1, p3.cols, f);
    for (int y = 0; y < p3.rows; y++) fwrite(p3.ptr(y), 1, p3.cols, f);
    fclose(f);
    printf("dropin ok: %zu/%zu keypoints, %d ratio-accepted\n", kL.size(), kR.size(), accepted);
    if (argc >= 8) {  // Frame::ComputeFboW (src/Frame.cc:395-400) + a KeyFrameDatabase-style score: <vocab.fbow> <out2.bin>
        fbow::Vocabulary voc;
        voc.readFromFile(argv[6]);
        fbow::fBow bL, bR;
        fbow::fBow2 fL, fR;
        voc.transform(dL, 4, bL, fL);
        voc.transform(dR, 4, bR, fR);
        fbow::fBow nL = voc.transform(dL), nR = voc.transform(dR);
        const double sc[2] = {fbow::fBow::score(nL, nR), fbow::fBow::score(nL, nL)};
        FILE* g = fopen(argv[7], "wb");
        int h2[4] = {(int)bL.size(), (int)fL.size(), (int)voc.getK(), (int)voc.getTotalWords()};
        fwrite(h2, sizeof(int), 4, g);
        for (const auto& e : bL) { float w = e.second; fwrite(&e.first, 4, 1, g); fwrite(&w, 4, 1, g); }
        for (const auto& e : fL) {
            int c = (int)e.second.size();
            fwrite(&e.first, 4, 1, g); fwrite(&c, 4, 1, g); fwrite(e.second.data(), 4, e.second.size(), g);
        }
        fwrite(sc, 8, 2, g);
        fclose(g);
        printf("fbow ok: %zu words, %zu nodes, score %.6f\n", bL.size(), fL.size(), sc[0]);
        if (argc >= 10) run_bow_search(kL, dL, fL, kR, dR, fR, argv[9]);  // <out4.bin>
    }
    if (argc >= 9) run_initialization(w, h, kL, dL, kR, dR, exL.GetScaleFactors(), argv[8]);  // <out3.bin>
    return 0;
}

// Compile-only instantiation of the device-resident SearchByProjection(Frame, Frame) wrapper against a stand-in with
// the members of ORB_SLAM2::Frame / MapPoint it reads (include/Frame.h:110-190, include/MapPoint.h:43-80).
namespace dropin_check {
struct MapPoint {
    cv::Mat pos, desc;
    int nObs = 0;
    bool mbTrackInView = false;
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackViewCos = 0;
    int mnTrackScaleLevel = 0;
    bool isBad() { return false; }
    cv::Mat GetWorldPos() { return pos; }
    cv::Mat GetDescriptor() { return desc; }
    int Observations() { return nObs; }
};
struct Frame {
    float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0, mb = 0;
    float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0, mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    int N = 0;
    cv::Mat mTcw, mDescriptors;
    std::vector<float> mvScaleFactors, mvuRight;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
};
int instantiate(ORB_SLAM2::ORBmatcher& m, Frame& cur, const Frame& last) { return m.SearchByProjection(cur, last, 15.f, false); }
void instantiate_prep(ORB_SLAM2::ORBextractor& e, const cv::Mat& M1, const cv::Mat& M2, const cv::Mat& K, const cv::Mat& D,
                      std::vector<cv::KeyPoint>& un) {
    e.SetInputFormat(ORBX_FMT_BGR8);
    e.SetRectifyMaps(M1, M2);
    e.UndistortKeyPoints(K, D, un);
}
struct KeyFrame {
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    fbow::fBow2 mFbowFeatVec;
    std::vector<MapPoint*> mps;
    std::vector<MapPoint*> GetMapPointMatches() { return mps; }
};
struct TriKeyFrame : KeyFrame {
    int N = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2;
    cv::Mat Ow, R, t;
    cv::Mat GetCameraCenter() { return Ow; }
    cv::Mat GetRotation() { return R; }
    cv::Mat GetTranslation() { return t; }
    MapPoint* GetMapPoint(int i) { return mps[i]; }
};
int instantiate5(ORB_SLAM2::ORBmatcher& m, TriKeyFrame* a, TriKeyFrame* b, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& out) {
    return m.SearchForTriangulation(a, b, F12, out, false);
}
struct RelocMapPoint : MapPoint {
    float GetMinDistanceInvariance() { return 0.f; }
    float GetMaxDistanceInvariance() { return 1e9f; }
    template <class F> int PredictScale(float, F*) { return 0; }
};
struct RelocFrame : Frame { std::vector<RelocMapPoint*> mvpMapPoints; };
struct RelocKeyFrame {
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<RelocMapPoint*> mps;
    std::vector<RelocMapPoint*> GetMapPointMatches() { return mps; }
};
int instantiate6(ORB_SLAM2::ORBmatcher& m, RelocFrame& F, RelocKeyFrame* kf, const std::set<RelocMapPoint*>& found) {
    return m.SearchByProjection(F, kf, found, 10.f, 100);
}
struct FuseKeyFrame;
struct FuseMapPoint : RelocMapPoint {
    cv::Mat normal;
    cv::Mat GetNormal() { return normal; }
    bool IsInKeyFrame(FuseKeyFrame*) { return false; }
    void Replace(FuseMapPoint*) {}
    void AddObservation(FuseKeyFrame*, int) {}
};
struct FuseKeyFrame {
    long unsigned int mnId = 0;
    int N = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0, mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::vector<float> mvScaleFactors, mvInvLevelSigma2, mvuRight;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors, R, t, Ow;
    std::vector<FuseMapPoint*> mps;
    cv::Mat GetRotation() { return R; }
    cv::Mat GetTranslation() { return t; }
    cv::Mat GetCameraCenter() { return Ow; }
    bool IsInImage(float x, float y) { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
    FuseMapPoint* GetMapPoint(int i) { return mps[i]; }
    void AddMapPoint(FuseMapPoint* p, int i) { mps[i] = p; }
};
int instantiate7(ORB_SLAM2::ORBmatcher& m, FuseKeyFrame* kf, const std::vector<FuseMapPoint*>& v) { return m.Fuse(kf, v, 3.0f); }
struct BowFrame : Frame { fbow::fBow2 mFbowFeatVec; };
int instantiate3(ORB_SLAM2::ORBmatcher& m, KeyFrame* kf, BowFrame& F, std::vector<MapPoint*>& out) { return m.SearchByFboW(kf, F, out); }
int instantiate4(ORB_SLAM2::ORBmatcher& m, KeyFrame* a, KeyFrame* b, std::vector<MapPoint*>& out) { return m.SearchByFboW(a, b, out); }
int instantiate2(ORB_SLAM2::ORBmatcher& m, Frame& F, const std::vector<MapPoint*>& v) { return m.SearchByProjection(F, v, 3.f); }
struct Sim3KeyFrame;
struct Sim3MapPoint : FuseMapPoint {
    int GetIndexInKeyFrame(Sim3KeyFrame*) { return -1; }
    void AddObservation(Sim3KeyFrame*, int) {}
    template <class K> int PredictScale(float, K*) { return 0; }
};
struct Sim3KeyFrame {
    long unsigned int mnId = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0, mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::vector<float> mvScaleFactors;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors, R, t;
    std::vector<Sim3MapPoint*> mps;
    cv::Mat GetRotation() { return R; }
    cv::Mat GetTranslation() { return t; }
    bool IsInImage(float x, float y) { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
    std::vector<Sim3MapPoint*> GetMapPointMatches() { return mps; }
    std::set<Sim3MapPoint*> GetMapPoints() { return std::set<Sim3MapPoint*>(mps.begin(), mps.end()); }
    Sim3MapPoint* GetMapPoint(int i) { return mps[i]; }
    void AddMapPoint(Sim3MapPoint* p, int i) { mps[i] = p; }
};
int instantiate8(ORB_SLAM2::ORBmatcher& m, Frame& F1, Frame& F2, std::vector<cv::Point2f>& prev, std::vector<int>& m12) {
    return m.SearchForInitialization(F1, F2, prev, m12, 100);
}
int instantiate9(ORB_SLAM2::ORBmatcher& m, Sim3KeyFrame* kf, cv::Mat Scw, const std::vector<Sim3MapPoint*>& pts,
                 std::vector<Sim3MapPoint*>& matched) {
    return m.SearchByProjection(kf, Scw, pts, matched, 10);
}
int instantiate10(ORB_SLAM2::ORBmatcher& m, Sim3KeyFrame* kf, cv::Mat Scw, const std::vector<Sim3MapPoint*>& pts,
                  std::vector<Sim3MapPoint*>& repl) {
    return m.Fuse(kf, Scw, pts, 4.0f, repl);
}
int instantiate11(ORB_SLAM2::ORBmatcher& m, Sim3KeyFrame* a, Sim3KeyFrame* b, std::vector<Sim3MapPoint*>& m12, const cv::Mat& R12,
                  const cv::Mat& t12) {
    const float s12 = 1.0f;
    return m.SearchBySim3(a, b, m12, s12, R12, t12, 7.5f);
}
std::vector<int> instantiate12(ORB_SLAM2::ORBmatcher& m, const std::vector<KeyFrame*>& kfs, BowFrame& F,
                               std::vector<std::vector<MapPoint*> >& out) {
    return m.SearchByFboW(kfs, F, out);
}
}  // namespace dropin_check

// Tracking::MonocularInitialization (/root/reference/src/Tracking.cc:690-699): ORBmatcher matcher(0.9, true);
// matcher.SearchForInitialization(mInitialFrame, mCurrentFrame, mvbPrevMatched, mvIniMatches, 100) through the drop-in
// template, on the two extracted frames. Output: n1, n2, nmatches, vnMatches12, vbPrevMatched, kR, dR.
static void run_initialization(int w, int h, const std::vector<cv::KeyPoint>& kL, const cv::Mat& dL,
                               const std::vector<cv::KeyPoint>& kR, const cv::Mat& dR, const std::vector<float>& sf,
                               const char* out) {
    dropin_check::Frame F1, F2;
    dropin_check::Frame* F[2] = {&F1, &F2};
    for (int k = 0; k < 2; k++) {
        F[k]->mnMinX = 0; F[k]->mnMaxX = (float)w; F[k]->mnMinY = 0; F[k]->mnMaxY = (float)h;
        F[k]->mfGridElementWidthInv = 64.0f / (F[k]->mnMaxX - F[k]->mnMinX);    // src/Frame.cc:101-102
        F[k]->mfGridElementHeightInv = 48.0f / (F[k]->mnMaxY - F[k]->mnMinY);
        F[k]->mvScaleFactors = sf;
    }
    F1.mvKeysUn = kL; F1.mDescriptors = dL; F1.N = (int)kL.size();
    F2.mvKeysUn = kR; F2.mDescriptors = dR; F2.N = (int)kR.size();
    std::vector<cv::Point2f> prev(kL.size());
    for (size_t i = 0; i < kL.size(); i++) prev[i] = kL[i].pt;
    std::vector<int> m12;
    ORB_SLAM2::ORBmatcher matcher(0.9f, true);
    const int nm = matcher.SearchForInitialization(F1, F2, prev, m12, 100);
    FILE* f = fopen(out, "wb");
    int hdr[3] = {(int)kL.size(), (int)kR.size(), nm};
    fwrite(hdr, sizeof(int), 3, f);
    fwrite(m12.data(), 4, m12.size(), f);
    fwrite(prev.data(), 8, prev.size(), f);
    fwrite(kR.data(), sizeof(cv::KeyPoint), kR.size(), f);
    for (int i = 0; i < dR.rows; i++) fwrite(dR.ptr(i), 1, 32, f);
    fclose(f);
    printf("initialization ok: %d matches\n", nm);
}

// Tracking::TrackWithMotionModel (/root/reference/src/Tracking.cc:985): matcher.SearchByProjection(mCurrentFrame,
// mLastFrame, th, bMono) through the drop-in template on a frame pair dumped by tests/test_gpu_dropin.py.
// Input: int32 nL, nC, stereo, bMono; float32 th, fx, fy, cx, cy, bf, w, h; Tcw_cur[16], Tcw_last[16], scale[8];
// last: world[nL*3] f32, valid[nL], outlier[nL], observed[nL] u8, keys[nL] (28 B), desc[nL*32];
// cur: keys[nC], uRight[nC] f32 (if stereo), desc[nC*32]. Output: nmatches, then per current keypoint the index of the
// last-frame map point left in mvpMapPoints (-1 = none).
static int run_tracking_search(const char* in, const char* out) {
    FILE* f = fopen(in, "rb");
    if (!f) return 2;
    int hdr[4];
    float cam[8], T[2][16], sc[8];
    if (fread(hdr, 4, 4, f) != 4 || fread(cam, 4, 8, f) != 8 || fread(T, 4, 32, f) != 32 || fread(sc, 4, 8, f) != 8) return 2;
    const int nL = hdr[0], nC = hdr[1], stereo = hdr[2], bMono = hdr[3];
    std::vector<float> world((size_t)nL * 3), uR(nC);
    std::vector<unsigned char> valid(nL), outlier(nL), observed(nL), dl((size_t)nL * 32), dc((size_t)nC * 32);
    std::vector<cv::KeyPoint> keysL(nL), keysC(nC);
    bool ok = fread(world.data(), 4, world.size(), f) == world.size() && fread(valid.data(), 1, nL, f) == (size_t)nL &&
              fread(outlier.data(), 1, nL, f) == (size_t)nL && fread(observed.data(), 1, nL, f) == (size_t)nL &&
              fread(keysL.data(), sizeof(cv::KeyPoint), nL, f) == (size_t)nL && fread(dl.data(), 1, dl.size(), f) == dl.size() &&
              fread(keysC.data(), sizeof(cv::KeyPoint), nC, f) == (size_t)nC;
    if (ok && stereo) ok = fread(uR.data(), 4, nC, f) == (size_t)nC;
    ok = ok && fread(dc.data(), 1, dc.size(), f) == dc.size();
    fclose(f);
    if (!ok) return 2;
    dropin_check::Frame cur, last;
    dropin_check::Frame* F[2] = {&cur, &last};
    for (int k = 0; k < 2; k++) {
        dropin_check::Frame& X = *F[k];
        X.fx = cam[1]; X.fy = cam[2]; X.cx = cam[3]; X.cy = cam[4]; X.mbf = cam[5]; X.mb = X.mbf / X.fx;
        X.mnMinX = 0; X.mnMaxX = cam[6]; X.mnMinY = 0; X.mnMaxY = cam[7];
        X.mfGridElementWidthInv = 64.0f / (X.mnMaxX - X.mnMinX);
        X.mfGridElementHeightInv = 48.0f / (X.mnMaxY - X.mnMinY);
        X.mvScaleFactors.assign(sc, sc + 8);
        X.mTcw = cv::Mat(4, 4, CV_32FC1, T[k], 4 * sizeof(float));  // header over the floats read above
    }
    std::vector<dropin_check::MapPoint> mps(nL);
    last.N = nL; last.mvKeys = keysL; last.mvKeysUn = keysL;
    last.mvpMapPoints.assign(nL, nullptr); last.mvbOutlier.assign(nL, false);
    for (int i = 0; i < nL; i++) {
        mps[i].pos = cv::Mat(3, 1, CV_32FC1, world.data() + 3 * (size_t)i, sizeof(float));
        mps[i].desc = cv::Mat(1, 32, CV_8UC1, dl.data() + (size_t)i * 32, 32);
        mps[i].nObs = observed[i] ? 1 : 0;
        if (valid[i]) last.mvpMapPoints[i] = &mps[i];
        last.mvbOutlier[i] = outlier[i] != 0;
    }
    cur.N = nC; cur.mvKeys = keysC; cur.mvKeysUn = keysC;
    cur.mDescriptors = cv::Mat(nC, 32, CV_8UC1, dc.data(), 32);
    if (stereo) cur.mvuRight = uR;
    cur.mvpMapPoints.assign(nC, nullptr);
    ORB_SLAM2::ORBmatcher matcher(0.9f, true);
    const int nm = matcher.SearchByProjection(cur, last, cam[0], bMono != 0);
    FILE* g = fopen(out, "wb");
    fwrite(&nm, 4, 1, g);
    for (int i = 0; i < nC; i++) {
        const int idx = cur.mvpMapPoints[i] ? (int)(cur.mvpMapPoints[i] - mps.data()) : -1;
        fwrite(&idx, 4, 1, g);
    }
    fclose(g);
    printf("tracking search ok: %d matches\n", nm);
    return 0;
}

// Tracking::TrackReferenceKeyFrame (/root/reference/src/Tracking.cc:862-867): ORBmatcher matcher(0.7, true);
// matcher.SearchByFboW(mpReferenceKF, mCurrentFrame, vpMapPointMatches) through the drop-in template: the left frame
// plays the keyframe (every 5th keypoint has no map point), the right one the current frame; the feature vectors are
// the ones fbow::Vocabulary::transform just produced. Output: nmatches, then per frame keypoint the keyframe index of
// the map point in vpMapPointMatches (-1 = none).
static void run_bow_search(const std::vector<cv::KeyPoint>& kL, const cv::Mat& dL, const fbow::fBow2& fL,
                           const std::vector<cv::KeyPoint>& kR, const cv::Mat& dR, const fbow::fBow2& fR, const char* out) {
    dropin_check::KeyFrame kf;
    kf.mvKeysUn = kL; kf.mDescriptors = dL; kf.mFbowFeatVec = fL;
    std::vector<dropin_check::MapPoint> mp(kL.size());
    kf.mps.assign(kL.size(), nullptr);
    for (size_t i = 0; i < kL.size(); i++)
        if (i % 5 != 4) kf.mps[i] = &mp[i];
    dropin_check::BowFrame F;
    F.N = (int)kR.size(); F.mvKeys = kR; F.mvKeysUn = kR; F.mDescriptors = dR; F.mFbowFeatVec = fR;
    std::vector<dropin_check::MapPoint*> matches;
    ORB_SLAM2::ORBmatcher matcher(0.7f, true);
    const int nm = matcher.SearchByFboW(&kf, F, matches);
    FILE* g = fopen(out, "wb");
    fwrite(&nm, 4, 1, g);
    for (size_t i = 0; i < matches.size(); i++) {
        const int idx = matches[i] ? (int)(matches[i] - mp.data()) : -1;
        fwrite(&idx, 4, 1, g);
    }
    fclose(g);
    printf("bow search ok: %d matches\n", nm);
}

// LoopClosing::ComputeSim3 (/root/reference/src/LoopClosing.cc:411): matcher.SearchByProjection(mpCurrentKF, mScw,
// mvpLoopMapPoints, mvpCurrentMatchedPoints, 10) through the drop-in template on dumped data.
// Input: int32 n, nK, th; float32 fx, fy, cx, cy, w, h; Scw[16]; scale[8]; points: world[n*3], normal[n*3], minDist[n],
// maxDist[n] f32, level[n] i32, bad[n] u8, desc[n*32]; keyframe: keys[nK], desc[nK*32], matched[nK] u8.
// Output: nmatches, then per keypoint: index in vpPoints of vpMatched[idx], -2 = the entry present on entry, -1 = NULL.
namespace run_sim3 {
struct KeyFrame;
struct MapPoint {
    const float *w, *nrm;
    float dmin, dmax;
    int level;
    bool bad;
    unsigned char* d;
    bool isBad() { return bad; }
    cv::Mat GetWorldPos() { return cv::Mat(3, 1, CV_32FC1, (void*)w, sizeof(float)); }
    cv::Mat GetNormal() { return cv::Mat(3, 1, CV_32FC1, (void*)nrm, sizeof(float)); }
    float GetMinDistanceInvariance() { return dmin; }
    float GetMaxDistanceInvariance() { return dmax; }
    int PredictScale(float, KeyFrame*) { return level; }
    cv::Mat GetDescriptor() { return cv::Mat(1, 32, CV_8UC1, d, 32); }
};
struct KeyFrame {
    long unsigned int mnId = 0;
    float fx, fy, cx, cy, mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<float> mvScaleFactors;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    bool IsInImage(const float& x, const float& y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
};
}  // namespace run_sim3
static int run_sim3_projection(const char* in, const char* out) {
    FILE* f = fopen(in, "rb");
    if (!f) return 2;
    int hdr[3];
    float cam[6], Scw[16], sc[8];
    if (fread(hdr, 4, 3, f) != 3 || fread(cam, 4, 6, f) != 6 || fread(Scw, 4, 16, f) != 16 || fread(sc, 4, 8, f) != 8) return 2;
    const int n = hdr[0], nK = hdr[1], th = hdr[2];
    std::vector<float> world((size_t)n * 3), normal((size_t)n * 3), dmin(n), dmax(n);
    std::vector<int> level(n);
    std::vector<unsigned char> bad(n), dp((size_t)n * 32), dk((size_t)nK * 32), matched(nK);
    std::vector<cv::KeyPoint> keys(nK);
    const bool ok = fread(world.data(), 4, world.size(), f) == world.size() && fread(normal.data(), 4, normal.size(), f) == normal.size() &&
                    fread(dmin.data(), 4, n, f) == (size_t)n && fread(dmax.data(), 4, n, f) == (size_t)n &&
                    fread(level.data(), 4, n, f) == (size_t)n && fread(bad.data(), 1, n, f) == (size_t)n &&
                    fread(dp.data(), 1, dp.size(), f) == dp.size() && fread(keys.data(), sizeof(cv::KeyPoint), nK, f) == (size_t)nK &&
                    fread(dk.data(), 1, dk.size(), f) == dk.size() && fread(matched.data(), 1, nK, f) == (size_t)nK;
    fclose(f);
    if (!ok) return 2;
    run_sim3::KeyFrame kf;
    kf.fx = cam[0]; kf.fy = cam[1]; kf.cx = cam[2]; kf.cy = cam[3];
    kf.mnMinX = 0; kf.mnMaxX = cam[4]; kf.mnMinY = 0; kf.mnMaxY = cam[5];
    kf.mfGridElementWidthInv = 64.0f / (kf.mnMaxX - kf.mnMinX);
    kf.mfGridElementHeightInv = 48.0f / (kf.mnMaxY - kf.mnMinY);
    kf.mvScaleFactors.assign(sc, sc + 8);
    kf.mvKeysUn = keys;
    kf.mDescriptors = cv::Mat(nK, 32, CV_8UC1, dk.data(), 32);
    std::vector<run_sim3::MapPoint> pts(n);
    std::vector<run_sim3::MapPoint*> vpPoints(n);
    for (int i = 0; i < n; i++) {
        run_sim3::MapPoint p = {world.data() + 3 * (size_t)i, normal.data() + 3 * (size_t)i, dmin[i], dmax[i], level[i], bad[i] != 0,
                                dp.data() + (size_t)i * 32};
        pts[i] = p;
        vpPoints[i] = &pts[i];
    }
    run_sim3::MapPoint present = pts.empty() ? run_sim3::MapPoint() : pts[0];  // stands for matches found earlier (not in vpPoints)
    std::vector<run_sim3::MapPoint*> vpMatched(nK, nullptr);
    for (int i = 0; i < nK; i++)
        if (matched[i]) vpMatched[i] = &present;
    cv::Mat S(4, 4, CV_32FC1, Scw, 4 * sizeof(float));
    ORB_SLAM2::ORBmatcher matcher(0.75f, true);
    const int nm = matcher.SearchByProjection(&kf, S, vpPoints, vpMatched, th);
    FILE* g = fopen(out, "wb");
    fwrite(&nm, 4, 1, g);
    for (int i = 0; i < nK; i++) {
        const int idx = !vpMatched[i] ? -1 : (vpMatched[i] == &present ? -2 : (int)(vpMatched[i] - pts.data()));
        fwrite(&idx, 4, 1, g);
    }
    fclose(g);
    printf("sim3 projection search ok: %d matches\n", nm);
    return 0;
}

// LoopClosing::ComputeSim3 (/root/reference/src/LoopClosing.cc:359): matcher.SearchBySim3(mpCurrentKF, pKF, vpMapPointMatches,
// s, R, t, 7.5) through the drop-in template on dumped data.
// Input: int32 n1, n2; float32 th, fx, fy, cx, cy, w, h, s12; R12[9], t12[3], R1w[9], t1w[3], R2w[9], t2w[3]; scale[8];
// then for KF1 and KF2 each: keys[n], kfdesc[n*32], has[n] u8, bad[n] u8, world[n*3], dmin[n], dmax[n] f32, level[n] i32,
// mpdesc[n*32]; finally pre[n1] i32: index in KF2 of the map point already in vpMatches12[i] (-1 = entry is NULL).
// Output: nFound, then per KF1 keypoint the KF2 index of vpMatches12[i] (-1 = NULL).
namespace run_sim3b {
struct KeyFrame;
struct MapPoint {
    const float* w;
    float dmin, dmax;
    int level, idxIn2;
    bool bad;
    unsigned char* d;
    bool isBad() { return bad; }
    cv::Mat GetWorldPos() { return cv::Mat(3, 1, CV_32FC1, (void*)w, sizeof(float)); }
    float GetMinDistanceInvariance() { return dmin; }
    float GetMaxDistanceInvariance() { return dmax; }
    int PredictScale(float, KeyFrame*) { return level; }
    int GetIndexInKeyFrame(KeyFrame*) { return idxIn2; }
    cv::Mat GetDescriptor() { return cv::Mat(1, 32, CV_8UC1, d, 32); }
};
struct KeyFrame {
    long unsigned int mnId = 0;
    float fx, fy, cx, cy, mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<float> mvScaleFactors;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    float R[9], t[3];
    std::vector<MapPoint*> mps;
    cv::Mat GetRotation() { return cv::Mat(3, 3, CV_32FC1, R, 3 * sizeof(float)); }
    cv::Mat GetTranslation() { return cv::Mat(3, 1, CV_32FC1, t, sizeof(float)); }
    std::vector<MapPoint*> GetMapPointMatches() { return mps; }
    bool IsInImage(const float& x, const float& y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
};
struct Side {
    std::vector<cv::KeyPoint> keys;
    std::vector<unsigned char> kd, has, bad, md;
    std::vector<float> world, dmin, dmax;
    std::vector<int> level;
    std::vector<MapPoint> pts;
    bool read(FILE* f, int n) {
        keys.resize(n); kd.resize((size_t)n * 32); has.resize(n); bad.resize(n); world.resize((size_t)n * 3);
        dmin.resize(n); dmax.resize(n); level.resize(n); md.resize((size_t)n * 32);
        return fread(keys.data(), sizeof(cv::KeyPoint), n, f) == (size_t)n && fread(kd.data(), 1, kd.size(), f) == kd.size() &&
               fread(has.data(), 1, n, f) == (size_t)n && fread(bad.data(), 1, n, f) == (size_t)n &&
               fread(world.data(), 4, world.size(), f) == world.size() && fread(dmin.data(), 4, n, f) == (size_t)n &&
               fread(dmax.data(), 4, n, f) == (size_t)n && fread(level.data(), 4, n, f) == (size_t)n &&
               fread(md.data(), 1, md.size(), f) == md.size();
    }
    void fill(KeyFrame& K, const float* cam, const float* sc, const float* R, const float* t) {
        const int n = (int)keys.size();
        K.fx = cam[1]; K.fy = cam[2]; K.cx = cam[3]; K.cy = cam[4];
        K.mnMinX = 0; K.mnMaxX = cam[5]; K.mnMinY = 0; K.mnMaxY = cam[6];
        K.mfGridElementWidthInv = 64.0f / (K.mnMaxX - K.mnMinX);
        K.mfGridElementHeightInv = 48.0f / (K.mnMaxY - K.mnMinY);
        K.mvScaleFactors.assign(sc, sc + 8);
        K.mvKeysUn = keys;
        K.mDescriptors = cv::Mat(n, 32, CV_8UC1, kd.data(), 32);
        for (int i = 0; i < 9; i++) K.R[i] = R[i];
        for (int i = 0; i < 3; i++) K.t[i] = t[i];
        pts.resize(n);
        K.mps.assign(n, nullptr);
        for (int i = 0; i < n; i++) {
            MapPoint p = {world.data() + 3 * (size_t)i, dmin[i], dmax[i], level[i], -1, bad[i] != 0, md.data() + (size_t)i * 32};
            pts[i] = p;
            if (has[i]) K.mps[i] = &pts[i];
        }
    }
};
}  // namespace run_sim3b
static int run_search_by_sim3(const char* in, const char* out) {
    FILE* f = fopen(in, "rb");
    if (!f) return 2;
    int hdr[2];
    float cam[8], R12[9], t12[3], R1w[9], t1w[3], R2w[9], t2w[3], sc[8];
    if (fread(hdr, 4, 2, f) != 2 || fread(cam, 4, 8, f) != 8 || fread(R12, 4, 9, f) != 9 || fread(t12, 4, 3, f) != 3 ||
        fread(R1w, 4, 9, f) != 9 || fread(t1w, 4, 3, f) != 3 || fread(R2w, 4, 9, f) != 9 || fread(t2w, 4, 3, f) != 3 ||
        fread(sc, 4, 8, f) != 8)
        return 2;
    const int n1 = hdr[0], n2 = hdr[1];
    run_sim3b::Side S1, S2;
    std::vector<int> pre(n1);
    const bool ok = S1.read(f, n1) && S2.read(f, n2) && fread(pre.data(), 4, n1, f) == (size_t)n1;
    fclose(f);
    if (!ok) return 2;
    run_sim3b::KeyFrame K1, K2;
    K1.mnId = 1; K2.mnId = 2;  // keyframe ids are unique (they key the device-side keyframe cache)
    S1.fill(K1, cam, sc, R1w, t1w);
    S2.fill(K2, cam, sc, R2w, t2w);
    for (int i = 0; i < n2; i++) S2.pts[i].idxIn2 = i;  // a map point of KF2 is observed in KF2 at its own index
    std::vector<run_sim3b::MapPoint*> vpMatches12(n1, nullptr);
    for (int i = 0; i < n1; i++)
        if (pre[i] >= 0) vpMatches12[i] = &S2.pts[pre[i]];
    const float s12 = cam[7];
    cv::Mat R(3, 3, CV_32FC1, R12, 3 * sizeof(float)), t(3, 1, CV_32FC1, t12, sizeof(float));
    ORB_SLAM2::ORBmatcher matcher(0.75f, true);
    const int nFound = matcher.SearchBySim3(&K1, &K2, vpMatches12, s12, R, t, cam[0]);
    FILE* g = fopen(out, "wb");
    fwrite(&nFound, 4, 1, g);
    for (int i = 0; i < n1; i++) {
        const int idx = vpMatches12[i] ? (int)(vpMatches12[i] - S2.pts.data()) : -1;
        fwrite(&idx, 4, 1, g);
    }
    fclose(g);
    printf("SearchBySim3 ok: %d found\n", nFound);
    return 0;
}

// LocalMapping::SearchInNeighbors (/root/reference/src/LocalMapping.cc:487): matcher.Fuse(pKFi, vpMapPointMatches) through
// the drop-in template on dumped data.
// Input: int32 n, nK; float32 th, fx, fy, cx, cy, bf, w, h; Rcw[9], tcw[3], Ow[3]; scale[8], invSigma2[8];
// points: world[n*3], normal[n*3], dmin[n], dmax[n] f32, level[n], nobs[n] i32, bad[n] u8, desc[n*32];
// keyframe: keys[nK], uRight[nK] f32, desc[nK*32], kfobs[nK] i32 (observations of the map point at that keypoint, -1 = none).
// Output: nFused, then per point: action (0 none, 1 pMP replaced by the keyframe's point, 2 keyframe's point replaced by
// pMP, 3 added as a new observation, 4 counted but the keyframe's point is bad) and the keypoint index.
namespace run_fuse_ns {
struct KeyFrame;
struct MapPoint {
    const float *w, *nrm;
    float dmin, dmax;
    int level, nobs;
    bool bad;
    unsigned char* d;
    int action, where;
    bool isBad() { return bad; }
    cv::Mat GetWorldPos() { return cv::Mat(3, 1, CV_32FC1, (void*)w, sizeof(float)); }
    cv::Mat GetNormal() { return cv::Mat(3, 1, CV_32FC1, (void*)nrm, sizeof(float)); }
    float GetMinDistanceInvariance() { return dmin; }
    float GetMaxDistanceInvariance() { return dmax; }
    int PredictScale(float, KeyFrame*) { return level; }
    cv::Mat GetDescriptor() { return cv::Mat(1, 32, CV_8UC1, d, 32); }
    int Observations() { return nobs; }
    bool IsInKeyFrame(KeyFrame*) { return false; }
    void Replace(MapPoint*) { action = -1; }   // marks "this point was replaced"; the caller records who replaced whom
    void AddObservation(KeyFrame*, int idx) { action = 3; where = idx; }
};
struct KeyFrame {
    long unsigned int mnId = 0;
    int N;
    float fx, fy, cx, cy, mbf, mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<float> mvScaleFactors, mvInvLevelSigma2, mvuRight;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    float R[9], t[3], O[3];
    std::vector<MapPoint*> mps;
    cv::Mat GetRotation() { return cv::Mat(3, 3, CV_32FC1, R, 3 * sizeof(float)); }
    cv::Mat GetTranslation() { return cv::Mat(3, 1, CV_32FC1, t, sizeof(float)); }
    cv::Mat GetCameraCenter() { return cv::Mat(3, 1, CV_32FC1, O, sizeof(float)); }
    bool IsInImage(const float& x, const float& y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
    MapPoint* GetMapPoint(int i) { return mps[i]; }
    void AddMapPoint(MapPoint* p, int i) { mps[i] = p; }
};
}  // namespace run_fuse_ns
static int run_fuse(const char* in, const char* out) {
    using namespace run_fuse_ns;
    FILE* f = fopen(in, "rb");
    if (!f) return 2;
    int hdr[2];
    float cam[8], pose[15], sc[8], is2[8];
    if (fread(hdr, 4, 2, f) != 2 || fread(cam, 4, 8, f) != 8 || fread(pose, 4, 15, f) != 15 || fread(sc, 4, 8, f) != 8 ||
        fread(is2, 4, 8, f) != 8)
        return 2;
    const int n = hdr[0], nK = hdr[1];
    std::vector<float> world((size_t)n * 3), normal((size_t)n * 3), dmin(n), dmax(n), uR(nK);
    std::vector<int> level(n), nobs(n), kfobs(nK);
    std::vector<unsigned char> bad(n), dp((size_t)n * 32), dk((size_t)nK * 32);
    std::vector<cv::KeyPoint> keys(nK);
    const bool ok = fread(world.data(), 4, world.size(), f) == world.size() && fread(normal.data(), 4, normal.size(), f) == normal.size() &&
                    fread(dmin.data(), 4, n, f) == (size_t)n && fread(dmax.data(), 4, n, f) == (size_t)n &&
                    fread(level.data(), 4, n, f) == (size_t)n && fread(nobs.data(), 4, n, f) == (size_t)n &&
                    fread(bad.data(), 1, n, f) == (size_t)n && fread(dp.data(), 1, dp.size(), f) == dp.size() &&
                    fread(keys.data(), sizeof(cv::KeyPoint), nK, f) == (size_t)nK && fread(uR.data(), 4, nK, f) == (size_t)nK &&
                    fread(dk.data(), 1, dk.size(), f) == dk.size() && fread(kfobs.data(), 4, nK, f) == (size_t)nK;
    fclose(f);
    if (!ok) return 2;
    KeyFrame kf;
    kf.N = nK;
    kf.fx = cam[1]; kf.fy = cam[2]; kf.cx = cam[3]; kf.cy = cam[4]; kf.mbf = cam[5];
    kf.mnMinX = 0; kf.mnMaxX = cam[6]; kf.mnMinY = 0; kf.mnMaxY = cam[7];
    kf.mfGridElementWidthInv = 64.0f / (kf.mnMaxX - kf.mnMinX);
    kf.mfGridElementHeightInv = 48.0f / (kf.mnMaxY - kf.mnMinY);
    kf.mvScaleFactors.assign(sc, sc + 8); kf.mvInvLevelSigma2.assign(is2, is2 + 8); kf.mvuRight = uR;
    kf.mvKeysUn = keys;
    kf.mDescriptors = cv::Mat(nK, 32, CV_8UC1, dk.data(), 32);
    for (int i = 0; i < 9; i++) kf.R[i] = pose[i];
    for (int i = 0; i < 3; i++) { kf.t[i] = pose[9 + i]; kf.O[i] = pose[12 + i]; }
    std::vector<MapPoint> own(nK), pts(n);
    kf.mps.assign(nK, nullptr);
    static float zero3[3] = {0, 0, 0};
    for (int i = 0; i < nK; i++)
        if (kfobs[i] >= 0) {
            MapPoint p = {zero3, zero3, 0, 0, 0, kfobs[i] % 1000, kfobs[i] >= 1000, nullptr, 0, -1};
            own[i] = p;
            kf.mps[i] = &own[i];
        }
    std::vector<MapPoint*> vp(n);
    for (int i = 0; i < n; i++) {
        MapPoint p = {world.data() + 3 * (size_t)i, normal.data() + 3 * (size_t)i, dmin[i], dmax[i], level[i], nobs[i], bad[i] != 0,
                      dp.data() + (size_t)i * 32, 0, -1};
        pts[i] = p;
        vp[i] = &pts[i];
    }
    ORB_SLAM2::ORBmatcher matcher(0.6f, true);
    const int nFused = matcher.Fuse(&kf, vp, cam[0]);
    FILE* g = fopen(out, "wb");
    fwrite(&nFused, 4, 1, g);
    for (int i = 0; i < n; i++) {
        int rec[2] = {pts[i].action, pts[i].where};
        fwrite(rec, 4, 2, g);
    }
    int nrep = 0;
    for (int i = 0; i < nK; i++) nrep += (kf.mps[i] == &own[i] && own[i].action == -1);
    fwrite(&nrep, 4, 1, g);
    fclose(g);
    printf("Fuse ok: %d fused\n", nFused);
    return 0;
}
