// oracle/shim_matcher/ref_slam_types.h — TEST INFRASTRUCTURE ONLY.
//
// Declarations of ORB_SLAM2::Frame, KeyFrame, MapPoint (and the few types they drag in) reduced to the members that
// /root/reference/src/ORBmatcher.cc and the extracted reference member functions touch. The real headers
// (/root/reference/include/{Frame,KeyFrame,MapPoint}.h) pull in boost, Eigen, g2o and fbow, none of which is installed
// here; this header is force-included (-include) before the unmodified ORBmatcher.cc, and because it defines the real
// headers' include guards, `#include "Frame.h"` etc. inside the reference's ORBmatcher.h become no-ops.
// Member names, types and static-ness follow the reference headers (include/Frame.h:43-212, include/KeyFrame.h:42-259,
// include/MapPoint.h:39-165). The BODIES of the member functions are not restated: oracle/Makefile extracts them by
// line range from the reference sources at build time (see ref_matcher_bridge.cpp).
#ifndef ORBX_REF_SLAM_TYPES_H
#define ORBX_REF_SLAM_TYPES_H

#define FRAME_H
#define KEYFRAME_H
#define MAPPOINT_H
#define ORBEXTRACTOR_H
#define MAP_H

#include <climits>
#include <cmath>
#include <cstdint>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <vector>

#include "cvmat_shim.hpp"

namespace fbow {
// Thirdparty/fbow/src/fbow.h: struct fBow2 : std::map<uint32_t, std::vector<uint32_t>>
struct fBow2 : std::map<uint32_t, std::vector<uint32_t>> {};
}  // namespace fbow

namespace ORB_SLAM2 {

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class MapPoint;
class KeyFrame;
class Frame;

class Map {
public:
    void EraseMapPoint(MapPoint*) {}
};

class ORBextractor {
public:
    std::vector<cv::Mat> mvImagePyramid;  // include/ORBextractor.h:86
};

class Frame {
public:
    Frame() {}
    bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                          const int maxLevel = -1) const;
    void ComputeStereoMatches();
    void AssignFeaturesToGrid();  // private in the reference; the bridge calls it like the constructors do

    ORBextractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;
    static float fx, fy, cx, cy, invfx, invfy;
    float mbf = 0, mb = 0, mThDepth = 0;
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    std::vector<float> mvuRight, mvDepth;
    fbow::fBow2 mFbowFeatVec;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    static float mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    cv::Mat mTcw;
    int mnScaleLevels = 0;
    float mfScaleFactor = 0, mfLogScaleFactor = 0;
    std::vector<float> mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
};

class KeyFrame {
public:
    KeyFrame() {}
    cv::Mat GetCameraCenter();
    cv::Mat GetRotation();
    cv::Mat GetTranslation();
    void AddMapPoint(MapPoint* pMP, const size_t& idx);
    void EraseMapPointMatch(const size_t& idx);
    void EraseMapPointMatch(MapPoint* pMP);
    void ReplaceMapPointMatch(const size_t& idx, MapPoint* pMP);
    std::set<MapPoint*> GetMapPoints();
    std::vector<MapPoint*> GetMapPointMatches();
    MapPoint* GetMapPoint(const size_t& idx);
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const;
    bool IsInImage(const float& x, const float& y) const;
    bool isBad() { return mbBad; }

    long unsigned int mnId = 0;
    int mnGridCols = 0, mnGridRows = 0;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0, invfx = 0, invfy = 0, mbf = 0, mb = 0, mThDepth = 0;
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight, mvDepth;
    cv::Mat mDescriptors;
    fbow::fBow2 mFbowFeatVec;
    int mnScaleLevels = 0;
    float mfScaleFactor = 0, mfLogScaleFactor = 0;
    std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    int mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;  // const int in the reference (include/KeyFrame.h:206-209)

    // protected in the reference; the bridge fills them
    cv::Mat Tcw, Ow;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<std::vector<std::vector<size_t>>> mGrid;
    bool mbBad = false;
    Map* mpMap = nullptr;
    std::mutex mMutexPose, mMutexConnections, mMutexFeatures;
};

class MapPoint {
public:
    MapPoint() {}
    cv::Mat GetWorldPos();
    cv::Mat GetNormal();
    std::map<KeyFrame*, size_t> GetObservations();
    int Observations();
    void AddObservation(KeyFrame* pKF, size_t idx);
    void EraseObservation(KeyFrame* pKF);
    int GetIndexInKeyFrame(KeyFrame* pKF);
    bool IsInKeyFrame(KeyFrame* pKF);
    void SetBadFlag();
    bool isBad();
    void Replace(MapPoint* pMP);
    void IncreaseVisible(int n = 1);
    void IncreaseFound(int n = 1);
    void ComputeDistinctiveDescriptors();
    cv::Mat GetDescriptor();
    float GetMinDistanceInvariance();
    float GetMaxDistanceInvariance();
    int PredictScale(const float& currentDist, KeyFrame* pKF);
    int PredictScale(const float& currentDist, Frame* pF);

    long unsigned int mnId = 0;
    int nObs = 0;
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
    bool mbTrackInView = false;
    int mnTrackScaleLevel = 0;
    float mTrackViewCos = 0;
    static std::mutex mGlobalMutex;

    // protected in the reference; the bridge fills them
    cv::Mat mWorldPos;
    std::map<KeyFrame*, size_t> mObservations;
    cv::Mat mNormalVector, mDescriptor;
    KeyFrame* mpRefKF = nullptr;
    int mnVisible = 1, mnFound = 1;
    bool mbBad = false;
    MapPoint* mpReplaced = nullptr;
    float mfMinDistance = 0, mfMaxDistance = 0;
    Map* mpMap = nullptr;
    std::mutex mMutexPos, mMutexFeatures;
};

}  // namespace ORB_SLAM2
#endif
