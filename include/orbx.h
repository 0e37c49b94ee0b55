/* orbx.h — C ABI of the B200-native ORB front end (liborbx.so).
 *
 * This is the drop-in boundary for ONE hot path of fabrizioromanelli/ORBSLAM2: ORB extraction and 256-bit Hamming
 * matching. The reference has no FFI layer for this path; the boundary it exposes is the C++ class API
 *   ORB_SLAM2::ORBextractor   /root/reference/include/ORBextractor.h:45-112, src/ORBextractor.cc:405-947
 *   ORB_SLAM2::ORBmatcher     /root/reference/include/ORBmatcher.h:38-102,   src/ORBmatcher.cc:1643-1659 + search loops
 *   Frame::ComputeStereoMatches /root/reference/include/Frame.h:92,          src/Frame.cc:464-642
 * Each entry point below names the reference interface it replaces. include/ORBextractor.h and include/ORBmatcher.h
 * (this repo) are the C++ wrappers with the reference's own signatures on top of this ABI; orbslam2_b200/ is the
 * Python mirror (ctypes). All functions return an ORBX_* status (0 = OK), never throw, never abort.
 * All kernels are hand-written CUDA for sm_100a; there is NO CPU fallback: without a CUDA device orbx_create and
 * orbm_create fail with ORBX_ECUDA.
 */
#ifndef ORBX_H
#define ORBX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_MAX_LEVELS 16

enum {
    ORBX_OK = 0,
    ORBX_EINVAL = 1, /* bad argument */
    ORBX_ESHAPE = 2, /* image shape unsupported (a pyramid level too small for one 30-px cell, w < h/2, ...) */
    ORBX_ECAP = 3,   /* caller's capacity too small; *n holds the required value */
    ORBX_ECUDA = 4,  /* CUDA runtime error or no device; see orbx_last_error_string() */
    ORBX_ESTATE = 5  /* call order (e.g. pyramid requested before any extract) */
};

enum { ORBX_MEM_HOST = 0, ORBX_MEM_DEVICE = 1 };

/* The eight constructor arguments of ORB_SLAM2::ORBextractor (include/ORBextractor.h:51, src/ORBextractor.cc:405-407). */
typedef struct {
    int nfeatures;
    float scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    int patchSize;     /* 31 */
    int halfPatchSize; /* 15 */
    int edgeThreshold; /* 19 */
} orbx_params;

/* Layout-identical to cv::KeyPoint (28 bytes) so a wrapper can memcpy into std::vector<cv::KeyPoint>. */
typedef struct {
    float x, y;     /* pt, level-0 coordinates */
    float size;     /* (int)(patchSize * scale[octave]) */
    float angle;    /* degrees [0,360) */
    float response; /* FAST score */
    int octave;
    int class_id;   /* -1 */
} orbx_kp;

typedef struct orbx_handle orbx_t;
typedef struct orbm_handle orbm_t;

/* ---- extractor: replaces ORBextractor::ORBextractor / ~ORBextractor (src/ORBextractor.cc:405-464) ------------------ */
int orbx_create(const orbx_params* p, int device, orbx_t** out);
int orbx_destroy(orbx_t* h);
/* Where the caller's image / result pointers live (default: host for both). Device pointers must be on `device`. */
int orbx_set_memory_space(orbx_t* h, int input_space, int output_space);

/* Replaces ORBextractor::operator() (src/ORBextractor.cc:858-919) for one 8-bit grey image. Keypoints come out
 * level-major and, inside a level, in canonical (y, x) order; desc is n x 32 bytes. */
int orbx_extract(orbx_t* h, const uint8_t* img, int w, int hgt, size_t stride, orbx_kp* kps, uint8_t* desc, int cap,
                 int* n);
/* Same for `batch` equally-shaped frames: frame f starts at imgs + f*img_stride; its results go to
 * kps + f*cap_per_frame, desc + f*cap_per_frame*32, n[f]. */
int orbx_extract_batch(orbx_t* h, const uint8_t* imgs, int batch, int w, int hgt, size_t stride, size_t img_stride,
                       orbx_kp* kps, uint8_t* desc, int cap_per_frame, int* n);
/* The same call in two halves, for callers that keep more than one batch in flight (one handle per batch in flight):
 * submit enqueues the copies and kernels and returns at once; wait blocks until that batch is complete, fills n[] and
 * reports device-side errors / ORBX_ECAP. The buffers passed to submit must stay valid until wait returns. One pending
 * batch per handle; orbx_extract_batch == submit + wait. With two handles the copy-in of batch k+1 and the copy-out of
 * batch k-1 overlap the kernels of batch k (the monocular Frame constructor's call, src/Frame.cc:247-253, issued for
 * a queue of camera frames instead of one). */
int orbx_extract_batch_submit(orbx_t* h, const uint8_t* imgs, int batch, int w, int hgt, size_t stride,
                              size_t img_stride, orbx_kp* kps, uint8_t* desc, int cap_per_frame);
int orbx_extract_batch_wait(orbx_t* h, int* n);
/* Single-process multi-GPU batch: `ndev` handles, one per device, each driven by its own host thread. Frames are split
 * into contiguous blocks (handle k owns the k-th block; the first batch % ndev blocks hold one frame more) and every
 * block's keypoints / descriptors / counts land in its own slice of the caller's HOST arrays; no collective, frames are
 * independent (the reference is a single process with std::thread workers, src/System.cc:104-112, and
 * has no multi-device counterpart). Same results, byte for byte, as one orbx_extract_batch on one device. */
int orbx_extract_batch_multi(orbx_t** handles, int ndev, const uint8_t* imgs, int batch, int w, int hgt, size_t stride,
                             size_t img_stride, orbx_kp* kps, uint8_t* desc, int cap, int* n);
/* ---- upstream image prep fused into the ingest (the calls the reference makes before the extractor) -----------------
 * Pixel format of the images handed to orbx_extract*: colour input is converted exactly as
 * cv::cvtColor(.., COLOR_{RGB,BGR,RGBA,BGRA}2GRAY) (src/Tracking.cc:269-294, 308-321, 337-350) while it is written into
 * pyramid level 0. `stride` of the extract calls is then the row stride in bytes of the interleaved image. */
enum { ORBX_FMT_GRAY8 = 0, ORBX_FMT_RGB8 = 1, ORBX_FMT_BGR8 = 2, ORBX_FMT_RGBA8 = 3, ORBX_FMT_BGRA8 = 4 };
int orbx_set_input_format(orbx_t* h, int fmt);
/* Rectification maps as cv::initUndistortRectifyMap(.., CV_32F, M1, M2) returns them (host pointers, out_w*out_h floats
 * each). Every following extract call remaps its source images as cv::remap(src, dst, M1, M2, cv::INTER_LINEAR) does
 * (Test/Replay/Stereo/stereo_euroc.cc:98-99, 136-137; BORDER_CONSTANT 0) into pyramid level 0: (w, hgt) of the extract
 * call describe the source, the pyramid has the maps' size. NULL, NULL switches rectification off. */
int orbx_set_rectify_maps(orbx_t* h, const float* map_x, const float* map_y, int out_w, int out_h);
/* Frame::UndistortKeyPoints (src/Frame.cc:402-432) for every frame of the last batch: kps_un[f*cap + i] = keypoint i of
 * frame f with pt replaced by cv::undistortPoints(pt, K, D, Mat(), K). K4 = fx, fy, cx, cy; dist = k1 k2 p1 p2 [k3 [k4
 * k5 k6]] (host pointers); dist[0] == 0 copies the keypoints (src/Frame.cc:404-408). kps_un lives in the output space. */
int orbx_undistort_keypoints(orbx_t* h, const float* K4, const float* dist, int ndist, orbx_kp* kps_un, int cap);
/* Smallest cap_per_frame that can never return ORBX_ECAP for this shape (sum over levels of quota + slack). */
int orbx_max_keypoints(orbx_t* h, int w, int hgt, int* cap);
/* With host buffers a batch is split into chunks of `frames` frames that are software-pipelined: the PCIe copy-in of
 * chunk c+1 and the copy-out of chunk c-1 overlap the kernels of chunk c. 0 = library default (48). */
int orbx_set_chunk(orbx_t* h, int frames);

/* Replaces the getters GetLevels / GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares /
 * GetInverseScaleSigmaSquares (include/ORBextractor.h:62-82) and exposes mnFeaturesPerLevel / umax
 * (src/ORBextractor.cc:411-463). Any output pointer may be NULL. */
int orbx_level_count(orbx_t* h);
int orbx_tables(orbx_t* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* features_per_level,
                int* umax /* halfPatchSize+1 */);

/* Replaces reads of the public member mvImagePyramid (include/ORBextractor.h:84; src/ORBextractor.cc:921-946) after the
 * last extract call: level size and pixels of frame `frame`; with_border adds the edgeThreshold-px BORDER_REFLECT_101
 * frame, i.e. the full padded buffer the reference's ROI lives in. dst is always a HOST pointer. */
int orbx_level_size(orbx_t* h, int level, int* w, int* hgt);
int orbx_pyramid_level(orbx_t* h, int frame, int level, uint8_t* dst, size_t dst_stride, int with_border);

/* Parity/debug taps (host outputs): the 7x7 sigma-2 blurred level (src/ORBextractor.cc:899-900) and the per-cell FAST
 * candidates handed to DistributeOctTree (src/ORBextractor.cc:783-823) as (x, y, score) int triples in level pixel
 * coordinates, unordered. */
/* The whole padded pyramid of one frame in ONE kernel launch and ONE device-to-host copy (what ORBextractor::mvImagePyramid
 * holds after operator(), src/ORBextractor.cc:921-946): level l lands at dst + offsets[l] as a dense (w+2e) x (h+2e)
 * buffer, e = edgeThreshold when with_border. offsets (nlevels entries) is an output. */
int orbx_pyramid_bytes(orbx_t* h, int with_border, size_t* bytes);
int orbx_pyramid_all(orbx_t* h, int frame, uint8_t* dst, size_t dst_bytes, int with_border, size_t* offsets);
int orbx_blurred_level(orbx_t* h, int frame, int level, uint8_t* dst, size_t dst_stride);
int orbx_candidates(orbx_t* h, int frame, int level, int32_t* xys, int cap, int* n);

/* ---- stereo: replaces Frame::ComputeStereoMatches (src/Frame.cc:464-642) ------------------------------------------
 * `left` and `right` must have just extracted the left and right images (same shape, same batch); works on frame
 * `frame` of each, entirely on the device (keypoints, descriptors and un-blurred pyramids are still resident).
 * minD/maxD are explicit (the reference derives maxD = mbf/mb; src/Frame.cc:498-500). uRight/depth: N_left floats
 * (-1 = no match), in the memory space selected for `left`'s outputs. */
int orbx_stereo_match(orbx_t* left, orbx_t* right, int frame, float mbf, float minD, float maxD, float* uRight,
                      float* depth);
/* All frames of the last batch; uRight/depth are batch x cap_per_frame. */
int orbx_stereo_match_batch(orbx_t* left, orbx_t* right, float mbf, float minD, float maxD, float* uRight,
                            float* depth, int cap_per_frame);

/* ---- matcher: replaces ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1643-1659) and the candidate-scan inner
 * loop shared by every ORBmatcher search (best / second-best with strict '<') -------------------- */
int orbm_create(int device, orbm_t** out);
int orbm_destroy(orbm_t* m);
int orbm_set_memory_space(orbm_t* m, int input_space, int output_space);
/* Upload (or adopt, for device memory) a train descriptor matrix nt x 32 that later calls reuse. */
int orbm_set_train(orbm_t* m, const uint8_t* t, int nt);
/* Brute-force 2-NN of every query against the train set: i1 = index of the best (lowest index wins ties),
 * d1 = its distance, d2 = second-smallest distance with multiplicity (256 if nt < 2). t == NULL reuses set_train. */
int orbm_knn2(orbm_t* m, const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* i1, uint16_t* d1, uint16_t* d2);
/* Which kernel orbm_knn2 runs. 0 (default): the tensor-core kernel (tcgen05.mma kind::i8: Hamming distance as a
 * binary GEMM, |q| + |t| - 2<q,t>; csrc/orbx_knn_tc.cuh) when nq * nt >= 2^24 and the train set holds a full 256-row
 * tile, else the POPC kernel; 1: POPC kernel only; 2: tensor cores whenever a full tile exists. Results are
 * bit-identical on every path. orbm_last_knn_path: 1 = POPC, 2 = tensor cores (+ POPC for the < 256-row remainder). */
int orbm_set_knn_mode(orbm_t* m, int mode);
int orbm_last_knn_path(orbm_t* m);
/* Allocation-free brute-force 2-NN. NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same results as orbm_knn2 (POPC path: k_knn2 over train splits + the ordered merge), but no handle and no allocation of
 * any kind: the per-split partial results live in a file-scope `static __device__` array inside liborbx (16 MiB, held by
 * every GPU that loads the library until exit). q (nq x 32), t (nt x 32), i1 / d1 / d2 (nq entries) are DEVICE pointers on
 * `device`, 16-byte aligned; the two kernels are enqueued on `stream` (a cudaStream_t, NULL = default stream) and the call
 * returns without synchronising: the outputs are valid once the stream reaches that point.
 * Limit: nq <= 65536 per call (ORBX_EINVAL beyond it; split the batch or use orbm_knn2); nt is unlimited.
 * Every call on a device shares the one scratch array: calls that overlap on the GPU (different streams, or two threads)
 * RACE — serialise them. Nothing in the array has to start at zero (each call writes every entry it later reads).
 * The first call on a device looks the array's address up (cudaGetSymbolAddress): make it outside graph capture. */
int orbm_knn2_static(int device, void* stream, const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* i1, uint16_t* d1,
                     uint16_t* d2);
/* Windowed search: query a scans candidates cand_idx[cand_off[a] .. cand_off[a+1]) in list order, skipping entries
 * with cand_skip[c] != 0 (cand_skip may be NULL). Outputs best/second index (-1 = none) and distance (256 = none). */
int orbm_window_search(orbm_t* m, const uint8_t* q, int nq, const int32_t* cand_off, const int32_t* cand_idx,
                       const uint8_t* cand_skip, const uint8_t* t, int nt, int32_t* best_idx, uint16_t* best_d,
                       uint16_t* second_d, int32_t* second_idx);
/* Batched DescriptorDistance: d[i] = hamming(a[i], b[i]). */
int orbm_distance(orbm_t* m, const uint8_t* a, const uint8_t* b, int n, uint16_t* d);

/* ---- device-resident windowed search: replaces ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame&
 * LastFrame, float th, bool bMono) (src/ORBmatcher.cc:1324-1466) together with what it calls on the current frame:
 * Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:231-246, 383-393), Frame::GetFeaturesInArea (src/Frame.cc:328-381),
 * DescriptorDistance and ComputeThreeMaxima (src/ORBmatcher.cc:1597-1638). Queries (last-frame map points) claim
 * current-frame keypoints greedily in index order exactly as the sequential loop does; the device reaches the same
 * result by a fixed-point iteration. All pointers live in the matcher's input / output memory space. ------ */
typedef struct {
    float fx, fy, cx, cy, bf;      /* Frame::fx, fy, cx, cy, mbf */
    float minX, maxX, minY, maxY;  /* Frame::mnMinX .. mnMaxY (src/Frame.cc:98-113) */
    float grid_w_inv, grid_h_inv;  /* Frame::mfGridElementWidthInv / HeightInv (src/Frame.cc:101-102); 64 x 48 cells */
    int nlevels;
    float scale[ORBX_MAX_LEVELS];  /* CurrentFrame.mvScaleFactors */
    float Rcw[9], tcw[3];          /* CurrentFrame.mTcw, row major */
    int forward, backward;         /* bForward / bBackward as computed at src/ORBmatcher.cc:1343-1344 */
    float th;                      /* window factor */
    int th_high;                   /* ORBmatcher::TH_HIGH (100) */
    int check_orientation;         /* mbCheckOrientation */
    int skip_depth_test;           /* 0: `if(invzc<0) continue` (src/ORBmatcher.cc:1360-1361); 1: the KeyFrame overload
                                      used by Relocalization has no such test (src/ORBmatcher.cc:1494-1505) */
} orbm_proj_params;

typedef struct {              /* LastFrame: n keypoints */
    int n;
    const float* world_pos;   /* n x 3, pMP->GetWorldPos() (ignored where flags bit0 is clear) */
    const uint8_t* flags;     /* bit0: mvpMapPoints[i] != NULL && !mvbOutlier[i]; bit1: pMP->Observations() > 0 */
    const int32_t* octave;    /* mvKeys[i].octave */
    const float* angle;       /* mvKeysUn[i].angle */
    const uint8_t* desc;      /* n x 32, pMP->GetDescriptor() */
} orbm_last_frame;

typedef struct {              /* CurrentFrame: n keypoints */
    int n;
    const orbx_kp* keys_un;   /* mvKeysUn (pt, octave, angle are read) */
    const float* u_right;     /* mvuRight, or NULL for monocular frames (treated as -1) */
    const uint8_t* desc;      /* mDescriptors, n x 32 */
    const uint8_t* taken;     /* 1 where mvpMapPoints[i] is set with Observations() > 0 on entry; may be NULL */
} orbm_cur_frame;

/* cur_match[i2]  (cur->n)  index into LastFrame of the map point left in CurrentFrame.mvpMapPoints[i2], -1 = none
 * last_match[i]  (last->n) keypoint i2 claimed by last-frame point i before the rotation filter, -1 = none
 * last_keep[i]   (last->n) 1 if that claim survived the rotation-consistency filter
 * nmatches       the function's return value (host pointer) */
/* The same entry point serves ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>&
 * sAlreadyFound, float th, int ORBdist) (src/ORBmatcher.cc:1468-1595; Tracking::Relocalization src/Tracking.cc:1552,
 * 1566): `last` holds pKF's map points with flags bit 0 = usable (not bad, not in sAlreadyFound, dist3D inside the
 * point's scale-invariance range) and bit 1 set, octave[i] = pMP->PredictScale(dist3D, &CurrentFrame) (host: it is a
 * logarithm per point), angle[i] = pKF->mvKeysUn[i].angle; forward = backward = 0, th_high = ORBdist,
 * skip_depth_test = 1, u_right = NULL, taken[i2] = CurrentFrame.mvpMapPoints[i2] != NULL. */
int orbm_search_by_projection(orbm_t* m, const orbm_proj_params* p, const orbm_last_frame* last,
                              const orbm_cur_frame* cur, int32_t* cur_match, int32_t* last_match, uint8_t* last_keep,
                              int* nmatches);
/* Replaces ORBmatcher::SearchByProjection(Frame &F, const std::vector<MapPoint*> &vpMapPoints, const float th)
 * (src/ORBmatcher.cc:43-127, called by Tracking::SearchLocalPoints, src/Tracking.cc:1290) on the device, with the same
 * grid / GetFeaturesInArea kernels. Of `p` the grid constants, scale[], th and th_high are read (no pose: the map
 * points carry the projection Frame::isInFrustum stored in them, src/Frame.cc:256-326). */
typedef struct {              /* vpMapPoints: n entries */
    int n;
    const float* proj_x;      /* pMP->mTrackProjX */
    const float* proj_y;      /* pMP->mTrackProjY */
    const float* proj_xr;     /* pMP->mTrackProjXR */
    const int32_t* level;     /* pMP->mnTrackScaleLevel */
    const float* view_cos;    /* pMP->mTrackViewCos */
    const uint8_t* flags;     /* bit0: pMP->mbTrackInView && !pMP->isBad(); bit1: pMP->Observations() > 0 */
    const uint8_t* desc;      /* n x 32, pMP->GetDescriptor() */
} orbm_map_points;
/* cur_match[idx] (cur->n): index into vpMapPoints of the point now in F.mvpMapPoints[idx], -1 = none;
 * mp_match[i] (mps->n): keypoint claimed by map point i, -1 = none; nn_ratio = ORBmatcher::mfNNratio. */
int orbm_search_local_points(orbm_t* m, const orbm_proj_params* p, float nn_ratio, const orbm_map_points* mps,
                             const orbm_cur_frame* cur, int32_t* cur_match, int32_t* mp_match, int* nmatches);
/* Batched MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307; called for every new or updated map point by
 * LocalMapping / Fuse): map point p owns the descriptor rows off[p] .. off[p+1]-1 of `desc` (its observations, in the
 * reference's iteration order). best_idx[p] = row (relative to off[p]) with the least median distance to the others —
 * the median is sorted_row[0.5*(N-1)], self distance included; the first such row wins — or -1 for a point without
 * observations. best_median may be NULL. At most 1024 observations per point. */
int orbm_distinctive_descriptors(orbm_t* m, const uint8_t* desc, const int32_t* off, int npoints, int32_t* best_idx,
                                 int32_t* best_median);
/* Diagnostics of the last orbm_search_by_projection call: fixed-point rounds and candidate pairs evaluated. */
int orbm_search_stats(orbm_t* m, int* rounds, int64_t* candidates);

/* ---- instrumentation ---------------------------------------------------------------------------------------------- */
/* CUDA stream (cudaStream_t) the handle launches on — for timing with events on the launching stream. */
int orbx_stream(orbx_t* h, void** stream);
int orbm_stream(orbm_t* m, void** stream);
/* Kernels launched by this handle since creation. */
int64_t orbx_launch_count(orbx_t* h);
int64_t orbm_launch_count(orbm_t* m);
/* Device milliseconds spent per stage during the last extract call (CUDA events on the handle's stream):
 * [0] H2D/ingest, [1] pyramid, [2] FAST, [3] quadtree, [4] blur, [5] orient+descriptor, [6] D2H. */
int orbx_stage_ms(orbx_t* h, float* ms7);
int orbx_set_profiling(orbx_t* h, int on);
const char* orbx_last_error_string(void);
const char* orbx_version(void);

/* The search half of ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th) (include/ORBmatcher.h:82,
 * src/ORBmatcher.cc:821-971; LocalMapping::SearchInNeighbors src/LocalMapping.cc:487,512) and of
 * Fuse(KeyFrame*, cv::Mat Scw, const vector<MapPoint*>&, th, vector<MapPoint*>&) (include/ORBmatcher.h:85,
 * src/ORBmatcher.cc:973-1096; LoopClosing::SearchAndFuse src/LoopClosing.cc:633). The caller projects the points
 * (positive depth, IsInImage, distance range, viewing angle, PredictScale need MapPoint internals) and passes
 * valid / u / v / ur / level; the device builds the KeyFrame's grid, walks KeyFrame::GetFeaturesInArea(u, v,
 * th*scale[level]) (src/KeyFrame.cc:563-602) in the reference's order, applies the octave gate [level-1, level], with
 * check_chi2 the reprojection gates (stereo 7.8 / mono 5.99 on e2*inv_level_sigma2[octave]) and returns per point the
 * best keypoint (first one on ties; -1 / 256 = none). The caller then applies `bestDist <= TH_LOW` and the map
 * mutations in point order. p supplies minX..grid_h_inv, nlevels, scale[], th. */
typedef struct {
    int n;
    const float* proj_x;  /* u */
    const float* proj_y;  /* v */
    const float* proj_xr; /* ur = u - bf*invz (first overload; ignored without check_chi2) */
    const int32_t* level; /* nPredictedLevel */
    const uint8_t* valid; /* every gate before GetFeaturesInArea passed */
    const uint8_t* desc;  /* [n][32] pMP->GetDescriptor() */
} orbm_fuse_points;
int orbm_fuse_search(orbm_t* m, const orbm_proj_params* p, const float* inv_level_sigma2, int check_chi2,
                     const orbm_fuse_points* pts, const orbm_cur_frame* kf, int32_t* best_idx, uint16_t* best_dist);
/* Keyframe cache. A KeyFrame's mvKeysUn / mvuRight / mDescriptors never change after construction (src/KeyFrame.cc:31-57),
 * yet Fuse, SearchBySim3 and SearchInNeighbors search the same keyframes again and again. orbm_keyframe_hint declares
 * that the `kf` argument of the NEXT orbm_fuse_search call is keyframe `key` (the drop-in passes pKF->mnId + 1; 0 = no
 * hint): if the handle still holds that keyframe (same n, same grid constants; 8 entries, least recently used replaced)
 * its upload and the four grid kernels are skipped, otherwise it is remembered. The hint is consumed by that call. */
int orbm_keyframe_hint(orbm_t* m, uint64_t key);
int64_t orbm_keyframe_cache_hits(orbm_t* m);

/* Replaces the search of ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints,
 * vector<MapPoint*>& vpMatched, int th) (include/ORBmatcher.h:64, src/ORBmatcher.cc:285-398; LoopClosing::ComputeSim3
 * src/LoopClosing.cc:411, th = 10). As for Fuse the caller projects the points with Scw (pts->valid / proj_x / proj_y /
 * level = nPredictedLevel; proj_xr is not read); the device builds pKF's grid, walks KeyFrame::GetFeaturesInArea(u, v,
 * th*scale[level]), applies the octave gate [level-1, level] and the rule "skip keypoints with vpMatched[idx] set":
 * kf->taken[idx] = vpMatched[idx] != NULL on entry, and every accepted point (bestDist <= p->th_high, pass TH_LOW)
 * blocks its keypoint for the later points, exactly as the sequential loop does (fixed point, see
 * orbm_search_by_projection). kf_match[idx] (kf->n) = index of the point newly written to vpMatched[idx], -1 = none;
 * pt_match[i] (pts->n) = keypoint taken by point i, -1 = none; *nmatches = the function's return value. */
int orbm_search_kf_projection(orbm_t* m, const orbm_proj_params* p, const orbm_fuse_points* pts, const orbm_cur_frame* kf,
                              int32_t* kf_match, int32_t* pt_match, int* nmatches);

/* Replaces ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched,
 * vector<int>& vnMatches12, int windowSize) (include/ORBmatcher.h:68, src/ORBmatcher.cc:400-515;
 * Tracking::MonocularInitialization src/Tracking.cc:699, windowSize = 100, ORBmatcher(0.9, true)). keys1 / desc1 =
 * F1.mvKeysUn / F1.mDescriptors (n1 rows; only octave-0 keypoints query, :417); prev_matched = vbPrevMatched as n1 x 2
 * floats, read as the window centres and UPDATED in place with F2.mvKeysUn[vnMatches12[i1]].pt for the final matches
 * (:508-510); f2 = F2 (keys_un, desc; u_right / taken are not read). The device builds F2's grid, walks
 * Frame::GetFeaturesInArea(x, y, windowSize, 0, 0) in the reference's order, and replays the claim loop in query order
 * including `vMatchedDistance[i2] <= dist` (:439), the best / second-best ratio rule (:454-456, th_low = TH_LOW,
 * nn_ratio = mfNNratio), the stealing of a keypoint from an earlier worse match (:458-466) and the rotation filter whose
 * histogram also counts the stolen accepts (:468-505). p supplies minX..grid_h_inv and check_orientation.
 * matches12[i1] (n1) = vnMatches12; *nmatches = the return value. */
int orbm_search_for_initialization(orbm_t* m, const orbm_proj_params* p, float nn_ratio, int th_low, int window_size,
                                   int n1, const orbx_kp* keys1, const uint8_t* desc1, float* prev_matched,
                                   const orbm_cur_frame* f2, int32_t* matches12, int* nmatches);

/* ORBmatcher::SearchByFboW, both overloads (include/ORBmatcher.h:54-55, 71): KeyFrame vs Frame
 * (src/ORBmatcher.cc:157-283; Tracking::TrackReferenceKeyFrame / Relocalization) with strict = 0, th = TH_LOW, and
 * KeyFrame vs KeyFrame (src/ORBmatcher.cc:517-650; LoopClosing) with strict = 1. A side is a frame's descriptors,
 * keypoint angles and its feature vector mFbowFeatVec in the CSR form orbv_transform_batch returns (node ids ascending).
 *   a (queries, KF / KF1): eligible[i] = vpMapPoints[i] && !isBad();   b: eligible = NULL (Frame: every feature is a
 *   candidate) or vpMapPoints2[i] && !isBad() (KF2).
 * match_a[i] = index in b matched to feature i of a after the rotation-histogram filter, else -1; the wrapper turns it
 * into vpMapPointMatches[match_a[i]] = vpMapPointsKF[i] (KF,F) or vpMatches12[i] = vpMapPoints2[match_a[i]] (KF,KF).
 * The "skip features already matched by an earlier query" rule is reproduced exactly (fixed point, see
 * orbm_search_by_projection). n_pos_a / n_pos_b = node_off[n_nodes] of each side. */
typedef struct {
    int n;
    const uint8_t* eligible; /* [n] */
    const float* angle;      /* [n] mvKeysUn[i].angle (KF) / mvKeys[i].angle (Frame) */
    const uint8_t* desc;     /* [n][32] */
    int n_nodes;
    const uint32_t* node_id; /* [n_nodes] ascending */
    const int* node_off;     /* [n_nodes + 1] */
    const uint32_t* node_feat;
} orbm_bow_side;
int orbm_search_by_bow(orbm_t* m, const orbm_bow_side* a, const orbm_bow_side* b, int n_pos_a, int n_pos_b, float nn_ratio,
                       int th, int strict, int check_orientation, int32_t* match_a, int* nmatches);
/* The same search for npairs independent (a[k], b[k]) pairs in ONE launch sequence: Relocalization matches the current
 * frame against every candidate keyframe (src/Tracking.cc:1465-1490: a[k] = candidate k, b[k] = the frame, the same
 * pointers for every k — a host array passed several times is staged once), LoopClosing::ComputeSim3 the current keyframe
 * against every loop candidate (src/LoopClosing.cc:290-312, strict = 1). A single pair is latency bound on the GPU; the
 * batch shares the launches and the fixed-point rounds. match_a[k] (a[k].n entries) and nmatches[k] as above. */
int orbm_search_by_bow_batch(orbm_t* m, int npairs, const orbm_bow_side* a, const orbm_bow_side* b, const int* n_pos_a,
                             const int* n_pos_b, float nn_ratio, int th, int strict, int check_orientation,
                             int32_t* const* match_a, int* nmatches);

/* ORBmatcher::SearchForTriangulation (include/ORBmatcher.h:76-77, src/ORBmatcher.cc:652-819; LocalMapping::
 * CreateNewMapPoints src/LocalMapping.cc:268) on the same machinery: sides as above with
 *   a.eligible[i] = !pKF1->GetMapPoint(i) && (!bOnlyStereo || mvuRight1[i] >= 0), b.eligible likewise for pKF2;
 * a candidate replaces the best when dist <= TH_LOW, dist <= best, it is not within 100*scale of the epipole (mono-mono
 * pairs only) and CheckDistEpipolarLine (src/ORBmatcher.cc:138-155) passes. match_a[i] = vMatches12[i]. */
typedef struct {
    const float* pt_a;        /* [a.n][2] mvKeysUn[i].pt of KF1 */
    const float* pt_b;        /* [b.n][2] mvKeysUn[i].pt of KF2 */
    const int32_t* octave_b;  /* [b.n]    mvKeysUn[i].octave of KF2 */
    const uint8_t* stereo_a;  /* [a.n]    mvuRight[i] >= 0 */
    const uint8_t* stereo_b;  /* [b.n] */
    float F12[9];             /* row major */
    float ex, ey;             /* epipole of KF1's centre in KF2 (src/ORBmatcher.cc:657-664) */
    int nlevels;
    float scale_factors[ORBX_MAX_LEVELS]; /* pKF2->mvScaleFactors */
    float level_sigma2[ORBX_MAX_LEVELS];  /* pKF2->mvLevelSigma2 */
} orbm_triangulation;
int orbm_search_for_triangulation(orbm_t* m, const orbm_bow_side* a, const orbm_bow_side* b, int n_pos_a, int n_pos_b,
                                  const orbm_triangulation* t, int th, int check_orientation, int32_t* match_a,
                                  int* nmatches);
/* The same for npairs (a[k], b[k], T[k]) pairs in one launch sequence: LocalMapping::CreateNewMapPoints matches the new
 * keyframe against each of its ~20 covisible neighbours (src/LocalMapping.cc:225-268: a[k] = the new keyframe for every
 * k, staged once; b[k], T[k] = neighbour k with its F12 and epipole). See orbm_search_by_bow_batch. */
int orbm_search_for_triangulation_batch(orbm_t* m, int npairs, const orbm_bow_side* a, const orbm_bow_side* b,
                                        const int* n_pos_a, const int* n_pos_b, const orbm_triangulation* T, int th,
                                        int check_orientation, int32_t* const* match_a, int* nmatches);

/* ---- vocabulary tree (bag of words): replaces fbow::Vocabulary as the reference uses it -------------------------------
 * (Thirdparty/fbow/src/fbow.h:51-125, fbow.cpp; called from Frame::ComputeFboW src/Frame.cc:395-400 and
 * KeyFrame::ComputeFboW src/KeyFrame.cc:62). */
typedef struct orbv_handle orbv_t;
/* Vocabulary::readFromFile / fromStream (fbow.cpp:156-183): blob = the bytes of a .fbow file. Only ORB vocabularies
 * (CV_8UC1, 32 bytes) are accepted; anything else is ORBX_EINVAL. */
int orbv_create(const void* blob, size_t size, int device, orbv_t** out);
int orbv_destroy(orbv_t* v);
/* getK / size / getDescSize / getTotalWords (fbow.h:100-108). Any output pointer may be NULL. */
int orbv_info(orbv_t* v, uint32_t* k, uint32_t* nblocks, uint32_t* desc_size, uint64_t* total_words);
int orbv_set_memory_space(orbv_t* v, int input_space, int output_space);
/* Vocabulary::transform(features, level, fBow&, fBow2&) (fbow.cpp:52-93; level >= 0) and Vocabulary::transform(features)
 * (fbow.cpp:95-153; level < 0 with normalize = 1) for `batch` descriptor matrices laid out as orbx_extract_batch
 * writes them: frame f = desc + f*stride*32 with n[f] rows. Per frame f the results start at index f*stride
 * (node_off: f*(stride+1)):
 *   fBow   word_id[0..n_words[f])  ascending ids, word_w = the summed leaf weights (float sum in feature order)
 *   fBow2  node_id[0..n_nodes[f])  ascending node ids; the features of node j are
 *          node_feat[node_off[j] .. node_off[j+1]) in ascending feature order
 * n, n_words, n_nodes are host arrays; the other pointers live in the handle's memory spaces. An empty frame gives
 * empty results (the reference throws on an empty matrix and its callers never pass one, src/Frame.cc:398). */
int orbv_transform_batch(orbv_t* v, const uint8_t* desc, const int* n, int batch, int stride, int level, int normalize,
                         uint32_t* word_id, float* word_w, int* n_words, uint32_t* node_id, int* node_off,
                         uint32_t* node_feat, int* n_nodes);
/* fBow::score(q, db[i]) (fbow.cpp:208-254) for m id-sorted vectors in CSR form (db_off: m+1 entries), as
 * KeyFrameDatabase evaluates it per candidate keyframe (src/KeyFrameDatabase.cc:130,247). Host pointers. v may be NULL
 * (fBow::score is static in the reference): the current device is used. */
int orbv_score_batch(orbv_t* v, const uint32_t* q_id, const float* q_w, int nq, const uint32_t* db_id, const float* db_w,
                     const int* db_off, int m, double* score);
int orbv_stream(orbv_t* v, void** stream);
int64_t orbv_launch_count(orbv_t* v);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H */
