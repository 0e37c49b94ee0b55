// tests/cpp/scene_run.cpp — TEST INFRASTRUCTURE: one scene, one ORBmatcher method, a dump of everything it changed.
//
// The same program is compiled three ways (tests/scene_lib.py, oracle/Makefile):
//   -DSCENE_REF   `#include "ORBmatcher.h"` is the reference's include/ORBmatcher.h and the method that runs is the
//                 UNMODIFIED /root/reference/src/ORBmatcher.cc (oracle/_ref/libref_orbmatcher.so)  -> the pin
//   (default)     `dropin/ORBmatcher.h` over liborbx.so                                            -> the product on the GPU
//   (default) + tests/cpp/orbm_oracle_backend.cpp in place of liborbx.so                           -> host halves + oracle on CPU
// In all three the world state — ORB_SLAM2::Frame, KeyFrame, MapPoint — is the declarations of
// oracle/shim_matcher/ref_slam_types.h with the reference's own member bodies (libref_orbmatcher.so), so a call such as
// matcher.Fuse(pKF, vpMapPoints, th) is written once below, exactly as src/LocalMapping.cc:487 writes it, and what is
// compared between the variants is the return value, every output vector and the complete map state afterwards.
//
// usage: scene_run <scene.bin> <out.bin>     (format: tests/scene_lib.py)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <set>
#include <string>
#include <vector>

#ifdef SCENE_REF
#include "ORBmatcher.h"
#else
#include "ref_slam_types.h"
#include "dropin/ORBmatcher.h"
#endif

using namespace ORB_SLAM2;

namespace {

struct Reader {
    std::vector<unsigned char> b;
    size_t at = 0;
    explicit Reader(const char* path) {
        FILE* f = fopen(path, "rb");
        if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
        fseek(f, 0, SEEK_END);
        b.resize((size_t)ftell(f));
        fseek(f, 0, SEEK_SET);
        if (fread(b.data(), 1, b.size(), f) != b.size()) exit(2);
        fclose(f);
    }
    template <class T> T get() {
        T v;
        if (at + sizeof(T) > b.size()) { fprintf(stderr, "scene truncated\n"); exit(2); }
        memcpy(&v, b.data() + at, sizeof(T));
        at += sizeof(T);
        return v;
    }
    template <class T> std::vector<T> arr(size_t n) {
        std::vector<T> v(n);
        if (at + n * sizeof(T) > b.size()) { fprintf(stderr, "scene truncated\n"); exit(2); }
        if (n) memcpy(v.data(), b.data() + at, n * sizeof(T));
        at += n * sizeof(T);
        return v;
    }
};

struct RawKp { float x, y, size, angle, response; int octave, class_id; };

std::vector<cv::KeyPoint> keypoints(const std::vector<RawKp>& r) {
    std::vector<cv::KeyPoint> k(r.size());
    for (size_t i = 0; i < r.size(); i++) {
        k[i].pt.x = r[i].x; k[i].pt.y = r[i].y; k[i].size = r[i].size; k[i].angle = r[i].angle;
        k[i].response = r[i].response; k[i].octave = r[i].octave; k[i].class_id = r[i].class_id;
    }
    return k;
}
cv::Mat floatMat(int r, int c, const float* v) {
    cv::Mat m(r, c, CV_32F);
    for (int i = 0; i < r; i++)
        for (int j = 0; j < c; j++) m.at<float>(i, j) = v[i * c + j];
    return m;
}
cv::Mat byteMat(int r, int c, const uint8_t* v) {
    cv::Mat m(r, c, CV_8U);
    if (r * c) memcpy(m.data, v, (size_t)r * c);
    return m;
}

struct View {  // what the scene file holds per frame / keyframe
    int kind = 0, N = 0;
    std::vector<RawKp> keys, keysUn;
    std::vector<float> uRight, depth;
    std::vector<uint8_t> desc, outlier;
    std::vector<float> Tcw;
    std::vector<uint32_t> nodeId, nodeFeat;
    std::vector<int32_t> nodeOff, mp;
};

struct World {
    int nViews = 0, nMPs = 0;
    std::unique_ptr<Frame[]> frames;       // view v is frames[v] or kfs[v] depending on its kind; both arrays are
    std::unique_ptr<KeyFrame[]> kfs;       // contiguous so that std::map<KeyFrame*, size_t> iterates in view order
    std::unique_ptr<MapPoint[]> mps;
    std::vector<int> kind;
    Map map;
    MapPoint* mp(int id) { return id < 0 ? nullptr : &mps[id]; }
    int id(const MapPoint* p) const { return p ? (int)(p - &mps[0]) : -1; }
    std::vector<MapPoint*> mpList(const std::vector<int32_t>& ids) {
        std::vector<MapPoint*> v(ids.size());
        for (size_t i = 0; i < ids.size(); i++) v[i] = mp(ids[i]);
        return v;
    }
};

void put(std::vector<int32_t>& o, int32_t v) { o.push_back(v); }
void putList(std::vector<int32_t>& o, const std::vector<int32_t>& v) {
    o.push_back((int32_t)v.size());
    o.insert(o.end(), v.begin(), v.end());
}
int32_t bits(float f) { int32_t i; memcpy(&i, &f, 4); return i; }

}  // namespace

int main(int argc, char** argv) {
    if (argc != 3) { fprintf(stderr, "usage: scene_run <scene.bin> <out.bin>\n"); return 1; }
    Reader R(argv[1]);
    if (R.get<int32_t>() != 0x53434E31) { fprintf(stderr, "bad magic\n"); return 2; }
    const int method = R.get<int32_t>();
    World W;
    W.nViews = R.get<int32_t>(); W.nMPs = R.get<int32_t>();
    const std::vector<float> fpar = R.arr<float>(8);
    const std::vector<int32_t> ipar = R.arr<int32_t>(8);
    // calibration and scale tables, shared by every view (src/Frame.cc:76-115, src/KeyFrame.cc:31-57)
    const std::vector<float> cal = R.arr<float>(12);  // fx fy cx cy mbf mb thDepth minX maxX minY maxY logScaleFactor
    const int nlevels = R.get<int32_t>();
    const float scaleFactor = R.get<float>();
    const std::vector<float> sf = R.arr<float>(nlevels), isf = R.arr<float>(nlevels), s2 = R.arr<float>(nlevels),
                             is2 = R.arr<float>(nlevels);
    Frame::fx = cal[0]; Frame::fy = cal[1]; Frame::cx = cal[2]; Frame::cy = cal[3];
    Frame::invfx = 1.0f / cal[0]; Frame::invfy = 1.0f / cal[1];
    Frame::mnMinX = cal[7]; Frame::mnMaxX = cal[8]; Frame::mnMinY = cal[9]; Frame::mnMaxY = cal[10];
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);

    W.frames.reset(new Frame[W.nViews]);
    W.kfs.reset(new KeyFrame[W.nViews]);
    W.mps.reset(new MapPoint[W.nMPs > 0 ? W.nMPs : 1]);
    W.kind.resize(W.nViews);
    std::vector<View> views(W.nViews);
    for (int v = 0; v < W.nViews; v++) {
        View& V = views[v];
        V.kind = W.kind[v] = R.get<int32_t>();
        const int N = V.N = R.get<int32_t>();
        V.keys = R.arr<RawKp>(N); V.keysUn = R.arr<RawKp>(N);
        V.uRight = R.arr<float>(N); V.depth = R.arr<float>(N);
        V.desc = R.arr<uint8_t>((size_t)N * 32);
        V.Tcw = R.arr<float>(16);
        const int nNodes = R.get<int32_t>();
        V.nodeId = R.arr<uint32_t>(nNodes);
        V.nodeOff = R.arr<int32_t>(nNodes + 1);
        V.nodeFeat = R.arr<uint32_t>(nNodes ? V.nodeOff[nNodes] : 0);
        V.mp = R.arr<int32_t>(N);
        V.outlier = R.arr<uint8_t>(N);

        // a Frame is always filled: a KeyFrame is constructed from one (src/KeyFrame.cc:31-57), grid included
        Frame& F = W.frames[v];
        F.N = N;
        F.mbf = cal[4]; F.mb = cal[5]; F.mThDepth = cal[6];
        F.mvKeys = keypoints(V.keys); F.mvKeysUn = keypoints(V.keysUn);
        F.mvuRight = V.uRight; F.mvDepth = V.depth;
        F.mDescriptors = byteMat(N, 32, V.desc.data());
        F.mTcw = floatMat(4, 4, V.Tcw.data());
        F.mnScaleLevels = nlevels; F.mfScaleFactor = scaleFactor; F.mfLogScaleFactor = cal[11];
        F.mvScaleFactors = sf; F.mvInvScaleFactors = isf; F.mvLevelSigma2 = s2; F.mvInvLevelSigma2 = is2;
        for (int j = 0; j < nNodes; j++) {
            std::vector<uint32_t>& dst = F.mFbowFeatVec[V.nodeId[j]];
            dst.assign(V.nodeFeat.begin() + V.nodeOff[j], V.nodeFeat.begin() + V.nodeOff[j + 1]);
        }
        F.mvpMapPoints.assign(N, nullptr);
        F.mvbOutlier.assign(N, false);
        for (int i = 0; i < N; i++) { F.mvpMapPoints[i] = W.mp(V.mp[i]); F.mvbOutlier[i] = V.outlier[i] != 0; }
        F.AssignFeaturesToGrid();
        if (V.kind == 1) {
            KeyFrame& K = W.kfs[v];
            K.mnId = (long unsigned)v;
            K.mnGridCols = FRAME_GRID_COLS; K.mnGridRows = FRAME_GRID_ROWS;
            K.mfGridElementWidthInv = Frame::mfGridElementWidthInv; K.mfGridElementHeightInv = Frame::mfGridElementHeightInv;
            K.fx = Frame::fx; K.fy = Frame::fy; K.cx = Frame::cx; K.cy = Frame::cy; K.invfx = Frame::invfx; K.invfy = Frame::invfy;
            if (v == 1 && fpar[4] > 0.f) {  // a second keyframe from another camera (fpar[4..7] = its fx, fy, cx, cy)
                K.fx = fpar[4]; K.fy = fpar[5]; K.cx = fpar[6]; K.cy = fpar[7];
                K.invfx = 1.0f / K.fx; K.invfy = 1.0f / K.fy;
            }
            K.mbf = F.mbf; K.mb = F.mb; K.mThDepth = F.mThDepth; K.N = N;
            K.mvKeys = F.mvKeys; K.mvKeysUn = F.mvKeysUn; K.mvuRight = F.mvuRight; K.mvDepth = F.mvDepth;
            K.mDescriptors = F.mDescriptors.clone();
            K.mFbowFeatVec = F.mFbowFeatVec;
            K.mnScaleLevels = nlevels; K.mfScaleFactor = scaleFactor; K.mfLogScaleFactor = cal[11];
            K.mvScaleFactors = sf; K.mvLevelSigma2 = s2; K.mvInvLevelSigma2 = is2;
            K.mnMinX = (int)Frame::mnMinX; K.mnMinY = (int)Frame::mnMinY; K.mnMaxX = (int)Frame::mnMaxX; K.mnMaxY = (int)Frame::mnMaxY;
            K.mvpMapPoints = F.mvpMapPoints;
            K.mpMap = &W.map;
            K.mGrid.resize(K.mnGridCols);
            for (int i = 0; i < K.mnGridCols; i++) {
                K.mGrid[i].resize(K.mnGridRows);
                for (int j = 0; j < K.mnGridRows; j++) K.mGrid[i][j] = F.mGrid[i][j];
            }
            // KeyFrame::SetPose (src/KeyFrame.cc:65-79) without the stereo centre: Tcw, Ow = -Rcw.t()*tcw
            F.mTcw.copyTo(K.Tcw);
            cv::Mat Rcw = K.Tcw.rowRange(0, 3).colRange(0, 3);
            cv::Mat tcw = K.Tcw.rowRange(0, 3).col(3);
            cv::Mat Rwc = Rcw.t();
            K.Ow = -Rwc * tcw;
        }
    }
    for (int p = 0; p < W.nMPs; p++) {
        MapPoint& M = W.mps[p];
        M.mnId = (long unsigned)p;
        const std::vector<float> g = R.arr<float>(8);  // world[3] normal[3] minDist maxDist
        M.mWorldPos = floatMat(3, 1, g.data());
        M.mNormalVector = floatMat(3, 1, g.data() + 3);
        M.mfMinDistance = g[6]; M.mfMaxDistance = g[7];
        const std::vector<uint8_t> d = R.arr<uint8_t>(32);
        M.mDescriptor = byteMat(1, 32, d.data());
        const std::vector<int32_t> s = R.arr<int32_t>(6);  // bad nObsOverride visible found inView trackLevel
        M.mbBad = s[0] != 0; M.mnVisible = s[2]; M.mnFound = s[3];
        M.mbTrackInView = s[4] != 0; M.mnTrackScaleLevel = s[5];
        const std::vector<float> t = R.arr<float>(4);  // projX projY projXR viewCos
        M.mTrackProjX = t[0]; M.mTrackProjY = t[1]; M.mTrackProjXR = t[2]; M.mTrackViewCos = t[3];
        M.mpMap = &W.map;
        const int nobs = R.get<int32_t>();
        const std::vector<int32_t> ob = R.arr<int32_t>((size_t)nobs * 2);
        for (int k = 0; k < nobs; k++) {
            M.AddObservation(&W.kfs[ob[2 * k]], (size_t)ob[2 * k + 1]);
            if (!M.mpRefKF) M.mpRefKF = &W.kfs[ob[2 * k]];
        }
        if (s[1] >= 0) M.nObs = s[1];
    }
    const int nLists = R.get<int32_t>();
    std::vector<std::vector<int32_t>> lists(nLists);
    for (auto& l : lists) { const int n = R.get<int32_t>(); l = R.arr<int32_t>(n); }
    const int nFarr = R.get<int32_t>();
    std::vector<std::vector<float>> farr(nFarr);
    for (auto& l : farr) { const int n = R.get<int32_t>(); l = R.arr<float>(n); }

    ORBmatcher matcher(fpar[1], ipar[0] != 0);
    const float th = fpar[0];
    std::vector<int32_t> out;
    int ret = -1;
    switch (method) {
    case 1: {  // Tracking::SearchLocalPoints, src/Tracking.cc:1290
        const std::vector<MapPoint*> vpMapPoints = W.mpList(lists[0]);
        ret = matcher.SearchByProjection(W.frames[0], vpMapPoints, th);
        break;
    }
    case 2:  // Tracking::TrackWithMotionModel, src/Tracking.cc:985
        ret = matcher.SearchByProjection(W.frames[0], W.frames[1], th, ipar[1] != 0);
        break;
    case 3: {  // Tracking::Relocalization, src/Tracking.cc:1552
        std::set<MapPoint*> sFound;
        for (int32_t id : lists[0]) sFound.insert(W.mp(id));
        ret = matcher.SearchByProjection(W.frames[0], &W.kfs[1], sFound, th, ipar[1]);
        break;
    }
    case 4: {  // LoopClosing::ComputeSim3, src/LoopClosing.cc:411
        const std::vector<MapPoint*> vpPoints = W.mpList(lists[0]);
        std::vector<MapPoint*> vpMatched = W.mpList(lists[1]);
        ret = matcher.SearchByProjection(&W.kfs[0], floatMat(4, 4, farr[0].data()), vpPoints, vpMatched, ipar[1]);
        std::vector<int32_t> o;
        for (MapPoint* p : vpMatched) o.push_back(W.id(p));
        putList(out, o);
        break;
    }
    case 5: {  // Tracking::TrackReferenceKeyFrame, src/Tracking.cc:870
        std::vector<MapPoint*> vpMapPointMatches;
        ret = matcher.SearchByFboW(&W.kfs[0], W.frames[1], vpMapPointMatches);
        std::vector<int32_t> o;
        for (MapPoint* p : vpMapPointMatches) o.push_back(W.id(p));
        putList(out, o);
        break;
    }
    case 6: {  // LoopClosing::ComputeSim3, src/LoopClosing.cc:301
        std::vector<MapPoint*> vpMatches12;
        ret = matcher.SearchByFboW(&W.kfs[0], &W.kfs[1], vpMatches12);
        std::vector<int32_t> o;
        for (MapPoint* p : vpMatches12) o.push_back(W.id(p));
        putList(out, o);
        break;
    }
    case 7: {  // Tracking::MonocularInitialization, src/Tracking.cc:699
        std::vector<cv::Point2f> vbPrevMatched(farr[0].size() / 2);
        for (size_t i = 0; i < vbPrevMatched.size(); i++) vbPrevMatched[i] = cv::Point2f(farr[0][2 * i], farr[0][2 * i + 1]);
        std::vector<int> vnMatches12;
        ret = matcher.SearchForInitialization(W.frames[0], W.frames[1], vbPrevMatched, vnMatches12, ipar[1]);
        putList(out, std::vector<int32_t>(vnMatches12.begin(), vnMatches12.end()));
        std::vector<int32_t> o;
        for (const cv::Point2f& p : vbPrevMatched) { o.push_back(bits(p.x)); o.push_back(bits(p.y)); }
        putList(out, o);
        break;
    }
    case 8: {  // LocalMapping::CreateNewMapPoints, src/LocalMapping.cc:268
        std::vector<std::pair<size_t, size_t> > vMatchedPairs;
        ret = matcher.SearchForTriangulation(&W.kfs[0], &W.kfs[1], floatMat(3, 3, farr[0].data()), vMatchedPairs, ipar[1] != 0);
        std::vector<int32_t> o;
        for (const auto& p : vMatchedPairs) { o.push_back((int32_t)p.first); o.push_back((int32_t)p.second); }
        putList(out, o);
        break;
    }
    case 9: {  // LoopClosing::ComputeSim3, src/LoopClosing.cc:388
        std::vector<MapPoint*> vpMatches12 = W.mpList(lists[0]);
        const float s12 = fpar[2];
        ret = matcher.SearchBySim3(&W.kfs[0], &W.kfs[1], vpMatches12, s12, floatMat(3, 3, farr[0].data()),
                                   floatMat(3, 1, farr[1].data()), th);
        std::vector<int32_t> o;
        for (MapPoint* p : vpMatches12) o.push_back(W.id(p));
        putList(out, o);
        break;
    }
    case 10: {  // LocalMapping::SearchInNeighbors, src/LocalMapping.cc:487
        const std::vector<MapPoint*> vpMapPoints = W.mpList(lists[0]);
        ret = matcher.Fuse(&W.kfs[0], vpMapPoints, th);
        break;
    }
    case 11: {  // LoopClosing::SearchAndFuse, src/LoopClosing.cc:633
        const std::vector<MapPoint*> vpPoints = W.mpList(lists[0]);
        std::vector<MapPoint*> vpReplacePoints(vpPoints.size(), static_cast<MapPoint*>(NULL));
        ret = matcher.Fuse(&W.kfs[0], floatMat(4, 4, farr[0].data()), vpPoints, th, vpReplacePoints);
        std::vector<int32_t> o;
        for (MapPoint* p : vpReplacePoints) o.push_back(W.id(p));
        putList(out, o);
        break;
    }
    case 12: {  // Tracking::Relocalization, src/Tracking.cc:1465-1490: every candidate keyframe against the current frame.
                // The reference loops over the single call; the drop-in's batched overload replaces the loop.
        const int K = W.nViews - 1;
        std::vector<std::vector<MapPoint*> > vvp(K);
        std::vector<int> nm(K, 0);
#ifdef SCENE_REF
        for (int k = 0; k < K; k++) nm[k] = matcher.SearchByFboW(&W.kfs[k + 1], W.frames[0], vvp[k]);
#else
        std::vector<KeyFrame*> vpKFs(K);
        for (int k = 0; k < K; k++) vpKFs[k] = &W.kfs[k + 1];
        nm = matcher.SearchByFboW(vpKFs, W.frames[0], vvp);
#endif
        ret = 0;
        for (int k = 0; k < K; k++) {
            ret += nm[k];
            std::vector<int32_t> o(1, nm[k]);
            for (MapPoint* p : vvp[k]) o.push_back(W.id(p));
            putList(out, o);
        }
        break;
    }
    default:
        fprintf(stderr, "unknown method %d\n", method);
        return 2;
    }

    std::vector<int32_t> dump;
    put(dump, ret);
    put(dump, (int32_t)out.size());
    dump.insert(dump.end(), out.begin(), out.end());
    // the map state afterwards
    for (int v = 0; v < W.nViews; v++) {
        const std::vector<MapPoint*>& mp = W.kind[v] == 1 ? W.kfs[v].mvpMapPoints : W.frames[v].mvpMapPoints;
        put(dump, (int32_t)mp.size());
        for (MapPoint* p : mp) put(dump, W.id(p));
    }
    for (int p = 0; p < W.nMPs; p++) {
        MapPoint& M = W.mps[p];
        put(dump, M.mbBad); put(dump, M.nObs); put(dump, W.id(M.mpReplaced)); put(dump, M.mnVisible); put(dump, M.mnFound);
        put(dump, (int32_t)M.mObservations.size());
        for (const auto& o : M.mObservations) { put(dump, (int32_t)(o.first - &W.kfs[0])); put(dump, (int32_t)o.second); }
        int32_t d[8];
        memcpy(d, M.mDescriptor.data, 32);
        dump.insert(dump.end(), d, d + 8);
    }
    FILE* f = fopen(argv[2], "wb");
    if (!f) return 2;
    fwrite(dump.data(), 4, dump.size(), f);
    fclose(f);
    printf("method %d -> %d\n", method, ret);
    return 0;
}
