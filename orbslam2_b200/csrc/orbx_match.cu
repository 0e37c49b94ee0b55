// orbx_match.cu — matcher half of the C ABI (include/orbx.h) and the stereo matcher.
//
// K7  k_knn2           brute-force 2-NN over a train matrix   (ref: inner scan of ORBmatcher::SearchByFboW,
//                                                              /root/reference/src/ORBmatcher.cc:195-229)
//                      small problems and the < 256-row train remainder; the bulk runs on the tensor cores:
// K7t k_knn_tc         orbx_knn_tc.cuh (tcgen05.mma kind::i8 + TMEM + TMA), bit-identical results
// K7m k_knn2_merge     ordered merge of train-split partials (lowest index wins ties, as a sequential scan would)
// K8  k_window_search  CSR candidate lists, one warp per query (ref: every ORBmatcher search loop)
// K8d k_distance       batched DescriptorDistance              (ref: src/ORBmatcher.cc:1643-1659)
// K9  k_stereo_match   row-band Hamming search + 11x11 SAD sweep + parabola (ref: src/Frame.cc:506-625)
// K10 k_stereo_cut     median outlier cut                      (ref: src/Frame.cc:628-641)
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <mutex>
#include <vector>

#include "orbx_internal.cuh"
#include "orbx_knn_tc.cuh"

struct orbx_handle;
struct OrbxView {
    const DevPlan* hp; DevPlan* dPlan; uint8_t* dPyr; orbx_kp* dKps; uint8_t* dDesc; int* dNOut; int lastBatch;
    cudaStream_t stream; int device; int outSpace; int* hCounts; const int* dSelCount;
};
OrbxView orbx_view(orbx_handle* h);
void orbx_count_launch(orbx_handle* h, int k);
int orbx_stereo_scratch(orbx_handle* h, size_t need, int** sad, float** u, float** z);

// ------------------------------------------------------------------------------------------------ K7 knn2
#define KNN_THREADS 128
#define KNN_QPT 2        // queries per thread
#define KNN_TILE 512     // train descriptors per shared-memory tile (16 KB)

struct Best2 {
    int d1, i1, d2;
};
__device__ __forceinline__ void best2_push(Best2& b, int d, int j) {
    // strict '<' in ascending index order: first (lowest index) wins ties; d2 = second smallest with multiplicity
    if (d < b.d1) {
        b.d2 = b.d1;
        b.d1 = d;
        b.i1 = j;
    } else if (d < b.d2)
        b.d2 = d;
}

__global__ void __launch_bounds__(KNN_THREADS) k_knn2(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t,
                                                      int nt, int chunk, int idxBase, int4* __restrict__ partial,
                                                      int32_t* __restrict__ i1, uint16_t* __restrict__ d1,
                                                      uint16_t* __restrict__ d2) {
    __shared__ uint4 tile[KNN_TILE * 2];
    const int tid = threadIdx.x;
    const int qbase = blockIdx.x * (KNN_THREADS * KNN_QPT);
    const int split = blockIdx.y;
    const int t0 = split * chunk, t1 = min(nt, t0 + chunk);
    uint4 qa[KNN_QPT][2];
    Best2 b[KNN_QPT];
#pragma unroll
    for (int k = 0; k < KNN_QPT; k++) {
        const int qi = min(qbase + k * KNN_THREADS + tid, nq - 1);
        qa[k][0] = q[2 * (size_t)qi];
        qa[k][1] = q[2 * (size_t)qi + 1];
        b[k].d1 = 256;
        b[k].d2 = 256;
        b[k].i1 = -1;
    }
    for (int base = t0; base < t1; base += KNN_TILE) {
        const int cnt = min(KNN_TILE, t1 - base);
        __syncthreads();
        for (int i = tid; i < cnt * 2; i += KNN_THREADS) tile[i] = t[2 * (size_t)base + i];
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < cnt; j++) {
            const uint4 x0 = tile[2 * j], x1 = tile[2 * j + 1];
#pragma unroll
            for (int k = 0; k < KNN_QPT; k++) best2_push(b[k], hamming256(qa[k][0], qa[k][1], x0, x1), base + j);
        }
    }
#pragma unroll
    for (int k = 0; k < KNN_QPT; k++) {
        const int qi = qbase + k * KNN_THREADS + tid;
        if (qi >= nq) continue;
        if (b[k].i1 >= 0) b[k].i1 += idxBase;
        if (partial) partial[(size_t)split * nq + qi] = make_int4(b[k].d1, b[k].i1, b[k].d2, 0);
        else {
            i1[qi] = b[k].i1;
            d1[qi] = (uint16_t)b[k].d1;
            d2[qi] = (uint16_t)b[k].d2;
        }
    }
}

__global__ void k_knn2_merge(const int4* __restrict__ partial, int nq, int splits, int32_t* __restrict__ i1,
                             uint16_t* __restrict__ d1, uint16_t* __restrict__ d2) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    Best2 b = {256, -1, 256};
    for (int s = 0; s < splits; s++) {  // ascending train order == the order a sequential scan visits them
        const int4 p = partial[(size_t)s * nq + qi];
        if (p.x < b.d1) {
            b.d2 = min(b.d1, p.z);
            b.d1 = p.x;
            b.i1 = p.y;
        } else
            b.d2 = min(b.d2, p.x);
    }
    i1[qi] = b.i1;
    d1[qi] = (uint16_t)b.d1;
    d2[qi] = (uint16_t)b.d2;
}

// ------------------------------------------------------------------------------------------------ K8 window search
__device__ __forceinline__ u64 shfl_min_u64(u64 v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const u64 w = __shfl_xor_sync(0xffffffffu, v, o);
        v = w < v ? w : v;
    }
    return v;
}

__global__ void __launch_bounds__(256) k_window_search(const uint4* __restrict__ q, int nq, const int* __restrict__ off,
                                                       const int* __restrict__ idx, const uint8_t* __restrict__ skip,
                                                       const uint4* __restrict__ t, int32_t* __restrict__ bi,
                                                       uint16_t* __restrict__ bd, uint16_t* __restrict__ sd,
                                                       int32_t* __restrict__ si) {
    const int lane = threadIdx.x & 31;
    const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (a >= nq) return;
    const uint4 q0 = q[2 * (size_t)a], q1 = q[2 * (size_t)a + 1];
    const int c0 = off[a], c1 = off[a + 1];
    const u64 none = ((u64)256 << 32) | 0xFFFFFFFFull;
    u64 k1 = none, k2 = none;  // two smallest (distance << 32 | list position) seen by this lane
    for (int c = c0 + lane; c < c1; c += 32) {
        if (skip && skip[c]) continue;
        const int j = idx[c];
        const int d = hamming256(q0, q1, t[2 * (size_t)j], t[2 * (size_t)j + 1]);
        const u64 k = ((u64)d << 32) | (unsigned)(c - c0);
        if (k < k1) {
            k2 = k1;
            k1 = k;
        } else if (k < k2)
            k2 = k;
    }
    const u64 best = shfl_min_u64(k1);
    const u64 second = shfl_min_u64(k1 == best && best != none ? k2 : k1);
    if (lane == 0) {
        const int db = (int)(best >> 32), ds = (int)(second >> 32);
        bd[a] = (uint16_t)db;
        sd[a] = (uint16_t)ds;
        bi[a] = (best == none || db >= 256) ? -1 : idx[c0 + (int)(best & 0xFFFFFFFFull)];
        si[a] = (second == none || ds >= 256) ? -1 : idx[c0 + (int)(second & 0xFFFFFFFFull)];
    }
}

__global__ void k_distance(const uint4* __restrict__ a, const uint4* __restrict__ b, int n, uint16_t* __restrict__ d) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    d[i] = (uint16_t)hamming256(a[2 * (size_t)i], a[2 * (size_t)i + 1], b[2 * (size_t)i], b[2 * (size_t)i + 1]);
}

// ------------------------------------------------------------------------------------------------ K9/K10 stereo
__device__ __forceinline__ int reflect101_i(int p, int len) {
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
    return p;
}

// One warp per left keypoint. Candidate scan: every right keypoint whose row band [floor(y-r), ceil(y+r)],
// r = 2*scale[octave], contains (int)vL, octave within +-1, uR in [uL-maxD, uL-minD]; best = strict '<' from TH_HIGH in
// ascending right index (packed (dist, iR) minimum). Then the 11-shift SAD sweep on the un-blurred pyramids.
// The right keypoints are stored level-major and, inside a level, by ascending y (the extractor's canonical order), and
// the band radius is constant inside a level: so per octave the candidates are ONE contiguous run found by two binary
// searches on y (a superset, (row - r - 1, row + r + 1); the exact band test is still applied) instead of a scan of all
// right keypoints — the role of vRowIndices in the reference (src/Frame.cc:476-492).
// Two kernels (round 2): the candidate search is a chain of dependent loads per warp (ncu of the fused kernel: 8 warps per
// issue slot waiting on the long scoreboard, issue 40 %), so it runs as its own kernel at 40 registers and 1.5x the resident
// warps, leaving the right keypoint's index (or -1) in sad[]; k_stereo_sad, which needs the 22 accumulators, then runs the
// 11-shift sweep for the keypoints that have a candidate.
__global__ void __launch_bounds__(256, 6) k_stereo_match(const DevPlan* __restrict__ P, const orbx_kp* __restrict__ kL,
                                                         const uint8_t* __restrict__ dL, const int* __restrict__ nL,
                                                         const orbx_kp* __restrict__ kR, const uint8_t* __restrict__ dR,
                                                         const int* __restrict__ nR, const int* __restrict__ selCountR,
                                                         float minD, float maxD, float* __restrict__ uRight,
                                                         float* __restrict__ depth, int* __restrict__ sad, int outStride) {
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int f = blockIdx.y;
    const int NL = nL[f], NR = nR[f];
    if (iL >= NL) return;
    const long long kb = (long long)f * P->selStride;
    const long long ob = (long long)f * outStride;
    if (lane == 0) {
        uRight[ob + iL] = -1.0f;
        depth[ob + iL] = -1.0f;
        sad[kb + iL] = -1;
    }
    const orbx_kp kp = kL[kb + iL];
    const int levelL = kp.octave;
    const float vL = kp.y, uL = kp.x;
    const int row = (int)vL;
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    if (maxU < 0) return;
    const uint4* qd = (const uint4*)(dL + (kb + iL) * 32);
    const uint4 q0 = qd[0], q1 = qd[1];
    u64 best = ((u64)100 << 32) | 0xFFFFFFFFull;  // ORBmatcher::TH_HIGH
    const int nl = P->nlevels;
    int segStart = 0;
    for (int l = 0; l < nl; l++) {
        const int segEnd = min(NR, segStart + selCountR[f * nl + l]);
        if (l >= levelL - 1 && l <= levelL + 1 && segEnd > segStart) {
            const float rad = __fmul_rn(2.0f, P->lv[l].scale);
            const float yLo = (float)row - rad - 1.001f, yHi = (float)row + rad + 1.001f;
            // first index of the segment with y > yLo: 32-ary search, the lanes probe 32 evenly spaced keypoints per
            // round (two dependent loads for a segment of <= 1024 keypoints instead of ten for a binary search)
            int lo = segStart, len = segEnd - segStart;
            while (len > 0) {
                const int step = (len + 31) >> 5;
                const int idx = lo + lane * step;
                const bool pred = idx < lo + len ? (kR[kb + idx].y > yLo) : true;
                const unsigned m = __ballot_sync(0xffffffffu, pred);
                const int fl = m ? __ffs(m) - 1 : 32;
                if (fl == 0) break;                       // the first probe already qualifies: answer = lo
                const int end = lo + len;
                lo = lo + (fl - 1) * step + 1;            // just after the last probe that failed
                len = min(lo + step - 1, end) - lo;       // up to (not including) the first probe that passed
            }
            // the run is short (band of a few rows): walk it 32 keypoints at a time until y leaves the band
            for (int base = lo; base < segEnd; base += 32) {
                const int iR = base + lane;
                const bool in = iR < segEnd;
                orbx_kp r;
                if (in) r = kR[kb + iR];
                const bool beyond = in && r.y >= yHi;
                if (in && !beyond && r.octave >= levelL - 1 && r.octave <= levelL + 1) {
                    const float radr = __fmul_rn(2.0f, P->lv[r.octave].scale);
                    const int maxr = (int)ceilf(__fadd_rn(r.y, radr)), minr = (int)floorf(__fsub_rn(r.y, radr));
                    if (row >= minr && row <= maxr && r.x >= minU && r.x <= maxU) {
                        const uint4* td = (const uint4*)(dR + (kb + iR) * 32);
                        const int d = hamming256(q0, q1, td[0], td[1]);
                        const u64 k = ((u64)d << 32) | (unsigned)iR;
                        best = k < best ? k : best;
                    }
                }
                if (__any_sync(0xffffffffu, beyond)) break;
            }
        }
        segStart = segEnd;
    }
    best = shfl_min_u64(best);
    const int bestDist = (int)(best >> 32);
    if (bestDist >= 75 || (best & 0xFFFFFFFFull) == 0xFFFFFFFFull) return;  // thOrbDist = (TH_HIGH+TH_LOW)/2
    if (lane == 0) sad[kb + iL] = (int)(best & 0xFFFFFFFFull);  // for k_stereo_sad
}

__global__ void __launch_bounds__(256) k_stereo_sad(const DevPlan* __restrict__ P, const orbx_kp* __restrict__ kL,
                                                    const int* __restrict__ nL, const uint8_t* __restrict__ pyrL,
                                                    const orbx_kp* __restrict__ kR, const uint8_t* __restrict__ pyrR, float mbf,
                                                    float minD, float maxD, float* __restrict__ uRight,
                                                    float* __restrict__ depth, int* __restrict__ sad, int outStride) {
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int f = blockIdx.y;
    if (iL >= nL[f]) return;
    const long long kb = (long long)f * P->selStride;
    const long long ob = (long long)f * outStride;
    const int bestIdxR = sad[kb + iL];
    if (bestIdxR < 0) return;
    __syncwarp();  // every lane has read the index before lane 0 overwrites it
    if (lane == 0) sad[kb + iL] = -1;
    const orbx_kp kp = kL[kb + iL];
    const int levelL = kp.octave;
    const float uL = kp.x;

    const LevelGeo& G = P->lv[levelL];
    const float uR0 = kR[kb + bestIdxR].x;
    const float sfac = __fdiv_rn(1.0f, G.scale);  // mvInvScaleFactors[octave] = 1.0f / mvScaleFactor[octave]
    const float scaleduL = roundf(__fmul_rn(kp.x, sfac));
    const float scaledvL = roundf(__fmul_rn(kp.y, sfac));
    const float scaleduR0 = roundf(__fmul_rn(uR0, sfac));
    const int w = 5, L = 5;
    const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
    if (iniu < 0 || endu >= (float)G.w) return;
    const int cy = (int)scaledvL, cxL = (int)scaleduL, cxR = (int)scaleduR0;
    const uint8_t* IL = pyrL + G.pyrOff + (long long)f * G.frameStride;
    const uint8_t* IR = pyrR + G.pyrOff + (long long)f * G.frameStride;
    const int cL = IL[(long long)cy * G.pitch + cxL];
    int acc[11], cR[11];
    // the 21 right-image columns cxR-10 .. cxR+10 lie inside the level for every keypoint the extractor can produce
    // (keypoints keep 16 px from the border at their own level); the reflected path covers anything else
    const bool inner = cxR - 2 * w >= 0 && cxR + 2 * w < G.w;
    const uint8_t* crow = IR + (long long)cy * G.pitch;
#pragma unroll
    for (int s = 0; s < 11; s++) {
        acc[s] = 0;
        cR[s] = inner ? crow[cxR + s - L] : crow[reflect101_i(cxR + s - L, G.w)];  // centre of the window shifted by s-L
    }
    for (int p = lane; p < 121; p += 32) {
        const int dy = p / 11 - w, dx = p % 11 - w;
        const int yy = min(max(cy + dy, 0), G.h - 1);
        const int a = IL[(long long)yy * G.pitch + cxL + dx] - cL;
        const uint8_t* rrow = IR + (long long)yy * G.pitch;
        if (inner) {
            const uint8_t* rr = rrow + cxR + dx - L;  // 11 consecutive bytes: one per shift
#pragma unroll
            for (int s = 0; s < 11; s++) acc[s] += abs(a - ((int)rr[s] - cR[s]));
        } else {
#pragma unroll
            for (int s = 0; s < 11; s++) acc[s] += abs(a - ((int)rrow[reflect101_i(cxR + s - L + dx, G.w)] - cR[s]));
        }
    }
#pragma unroll
    for (int s = 0; s < 11; s++)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[s] += __shfl_xor_sync(0xffffffffu, acc[s], o);
    if (lane != 0) return;
    int bestSad = 0x7fffffff, bestinc = 0;
#pragma unroll
    for (int s = 0; s < 11; s++)
        if (acc[s] < bestSad) {
            bestSad = acc[s];
            bestinc = s - L;
        }
    if (bestinc == -L || bestinc == L) return;
    float dist1 = 0, dist2 = 0, dist3 = 0;
#pragma unroll
    for (int s = 1; s < 10; s++)
        if (s - L == bestinc) {
            dist1 = (float)acc[s - 1];
            dist2 = (float)acc[s];
            dist3 = (float)acc[s + 1];
        }
    const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3),
                                   __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2))));
    if (deltaR < -1 || deltaR > 1) return;  // NaN (0/0) passes, as in the reference; rejected below by the disparity test
    float bestuR = __fmul_rn(G.scale, __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) {
            disparity = 0.01f;
            bestuR = (float)((double)uL - 0.01);
        }
        depth[ob + iL] = __fdiv_rn(mbf, disparity);
        uRight[ob + iL] = bestuR;
        sad[kb + iL] = bestSad;
    }
}

// One CTA per frame: sort accepted (SAD, iL), median = element size/2, threshold 1.5f*1.4f*median, drop SAD >= it.
__global__ void __launch_bounds__(1024) k_stereo_cut(const DevPlan* __restrict__ P, const int* __restrict__ nL,
                                                     const int* __restrict__ sad, float* __restrict__ uRight,
                                                     float* __restrict__ depth, int outStride, int P2) {
    extern __shared__ __align__(16) uint8_t cut_smem[];
    u64* buf = (u64*)cut_smem;
    __shared__ int cnt;
    const int f = blockIdx.x, tid = threadIdx.x;
    const int N = nL[f];
    const long long kb = (long long)f * P->selStride, ob = (long long)f * outStride;
    if (tid == 0) cnt = 0;
    for (int i = tid; i < P2; i += blockDim.x) buf[i] = ~0ull;
    __syncthreads();
    for (int i = tid; i < N; i += blockDim.x) {
        const int s = sad[kb + i];
        if (s >= 0) buf[atomicAdd(&cnt, 1)] = ((u64)(unsigned)s << 32) | (unsigned)i;
    }
    __syncthreads();
    const int m = cnt;
    if (m == 0) return;  // the reference indexes an empty vector here (UB); nothing to cut
    int p2 = 1;
    while (p2 < m) p2 <<= 1;
    for (int k = 2; k <= p2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < p2; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const bool up = ((i & k) == 0);
                    const u64 x = buf[i], y = buf[ixj];
                    if ((x > y) == up) {
                        buf[i] = y;
                        buf[ixj] = x;
                    }
                }
            }
            __syncthreads();
        }
    const float median = (float)(int)(buf[m / 2] >> 32);
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), median);
    for (int i = tid; i < m; i += blockDim.x) {
        const int s = (int)(buf[i] >> 32), il = (int)(buf[i] & 0xFFFFFFFFull);
        if (!((float)s < thDist)) {
            uRight[ob + il] = -1.0f;
            depth[ob + il] = -1.0f;
        }
    }
}

// ------------------------------------------------------------------------------------------------ matcher handle
struct orbm_handle {
    int device = 0;
    cudaStream_t stream = nullptr;
    int inSpace = ORBX_MEM_HOST, outSpace = ORBX_MEM_HOST;
    uint8_t* dTrain = nullptr; size_t trainCap = 0; int nTrain = 0; bool trainAdopted = false;
    uint8_t* dQ = nullptr; size_t qCap = 0;
    uint8_t* dB = nullptr; size_t bCap = 0;
    int4* dPartial = nullptr; size_t partialCap = 0;
    // tensor-core 2-NN (orbx_knn_tc.cuh): int8 expansions of the queries (per call) and of the resident train set
    uint8_t* dAm = nullptr; size_t amCap = 0; uint8_t* dAe = nullptr; size_t aeCap = 0;
    uint8_t* dBm = nullptr; size_t bmCap = 0; uint8_t* dBe = nullptr; size_t beCap = 0;
    bool trainExpanded = false;
    int knnMode = 0;     // 0 auto, 1 POPC kernel only, 2 tensor-core kernel whenever a full tile exists
    int lastKnnPath = 0; // 1 POPC, 2 tensor cores (+ POPC remainder)
    int32_t* dI1 = nullptr; int32_t* dI2 = nullptr; uint16_t* dD1 = nullptr; uint16_t* dD2 = nullptr; size_t outCap = 0;
    int* dOff = nullptr; size_t offCap = 0;
    int* dIdx = nullptr; size_t idxCap = 0;
    uint8_t* dSkip = nullptr; size_t skipCap = 0;
    uint8_t* dWs = nullptr; size_t wsCap = 0;      // orbm_search_by_projection: staged inputs + grid + outputs
    uint8_t* dCand = nullptr; size_t candCap = 0;  // candidate (keypoint, distance) lists
    int* hRes = nullptr;                           // pinned: [0] nmatches, [2] total candidates
    int* hFlags = nullptr; size_t hFlagsCap = 0;   // pinned copy of the per-round change flags
    uint8_t* hIn = nullptr; size_t hInCap = 0;     // pinned mirror of the staged inputs: many small arrays, one copy
    int smCount = 148;
    int64_t launches = 0;
    int lastRounds = 0; int64_t lastCandidates = 0;
    int gridFusedMax = 32768;  // keypoints up to which the 64 x 48 grid is built by one CTA (ORBX_GRID_FUSED_MAX)
    int bowFusedMax = 4096;  // feature-vector entries up to which a bag-of-words pair runs in one CTA (ORBX_BOW_FUSED_MAX)
    // keyframe cache (orbm_keyframe_hint): staged keypoints / uRight / descriptors and the built 64 x 48 grid of immutable
    // keyframes, so that repeated Fuse / SearchBySim3 searches against one keyframe skip the upload and four grid kernels
    struct KfEntry {
        uint64_t key = 0; int n = 0; bool hasUR = false; float geo[4] = {0, 0, 0, 0}; uint64_t stamp = 0;
        uint8_t* mem = nullptr; size_t cap = 0;
        orbx_kp* keys = nullptr; float* uRight = nullptr; uint4* desc = nullptr; int* cellStart = nullptr; int* items = nullptr;
    };
    KfEntry kfc[8];
    uint64_t kfHint = 0, kfClock = 0;
    int64_t kfHits = 0;
};

template <typename T>
static int ensure(T*& p, size_t& cap, size_t need) {
    if (need <= cap) return ORBX_OK;
    cudaFree(p);
    p = nullptr;
    cap = 0;
    CK(cudaMalloc(&p, need * sizeof(T)));
    cap = need;
    return ORBX_OK;
}

extern "C" int orbm_create(int device, orbm_t** out) {
    if (!out) FAIL(ORBX_EINVAL, "null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        FAIL(ORBX_ECUDA, "no CUDA device: liborbx has no CPU path");
    }
    if (device < 0 || device >= ndev) FAIL(ORBX_EINVAL, "device %d out of range", device);
    CK(cudaSetDevice(device));
    orbm_handle* m = new orbm_handle();
    m->device = device;
    cudaError_t e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete m;
        FAIL(ORBX_ECUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
    }
    cudaDeviceGetAttribute(&m->smCount, cudaDevAttrMultiProcessorCount, device);
    if (const char* v = getenv("ORBX_BOW_FUSED_MAX")) m->bowFusedMax = atoi(v);
    if (const char* v = getenv("ORBX_GRID_FUSED_MAX")) m->gridFusedMax = atoi(v);
    *out = m;
    return ORBX_OK;
}

extern "C" int orbm_destroy(orbm_t* m) {
    if (!m) return ORBX_OK;
    cudaSetDevice(m->device);
    cudaStreamSynchronize(m->stream);
    if (!m->trainAdopted) cudaFree(m->dTrain);
    cudaFree(m->dAm); cudaFree(m->dAe); cudaFree(m->dBm); cudaFree(m->dBe);
    cudaFree(m->dQ); cudaFree(m->dB); cudaFree(m->dPartial); cudaFree(m->dI1); cudaFree(m->dI2); cudaFree(m->dD1);
    cudaFree(m->dD2); cudaFree(m->dOff); cudaFree(m->dIdx); cudaFree(m->dSkip); cudaFree(m->dWs); cudaFree(m->dCand); if (m->hRes) cudaFreeHost(m->hRes); if (m->hFlags) cudaFreeHost(m->hFlags); if (m->hIn) cudaFreeHost(m->hIn);
    for (auto& e : m->kfc) cudaFree(e.mem);
    cudaStreamDestroy(m->stream);
    delete m;
    return ORBX_OK;
}

extern "C" int orbm_set_memory_space(orbm_t* m, int in, int out) {
    if (!m || in < 0 || in > 1 || out < 0 || out > 1) FAIL(ORBX_EINVAL, "bad memory space");
    m->inSpace = in;
    m->outSpace = out;
    return ORBX_OK;
}
extern "C" int orbm_stream(orbm_t* m, void** s) {
    if (!m || !s) FAIL(ORBX_EINVAL, "null argument");
    *s = (void*)m->stream;
    return ORBX_OK;
}
extern "C" int64_t orbm_launch_count(orbm_t* m) { return m ? m->launches : 0; }

extern "C" int orbm_set_train(orbm_t* m, const uint8_t* t, int nt) {
    if (!m || !t || nt < 1) FAIL(ORBX_EINVAL, "bad argument");
    CK(cudaSetDevice(m->device));
    if (m->inSpace == ORBX_MEM_DEVICE) {
        if (((uintptr_t)t & 15) != 0) FAIL(ORBX_EINVAL, "device descriptor matrix must be 16-byte aligned");
        if (!m->trainAdopted) cudaFree(m->dTrain);
        m->dTrain = (uint8_t*)t;
        m->trainCap = 0;
        m->trainAdopted = true;
    } else {
        if (m->trainAdopted) {
            m->dTrain = nullptr;
            m->trainCap = 0;
            m->trainAdopted = false;
        }
        int rc = ensure(m->dTrain, m->trainCap, (size_t)nt * 32);
        if (rc) return rc;
        CK(cudaMemcpyAsync(m->dTrain, t, (size_t)nt * 32, cudaMemcpyHostToDevice, m->stream));
    }
    m->nTrain = nt;
    m->trainExpanded = false;
    return ORBX_OK;
}

extern "C" int orbm_set_knn_mode(orbm_t* m, int mode) {
    if (!m || mode < 0 || mode > 2) FAIL(ORBX_EINVAL, "knn mode must be 0 (auto), 1 (POPC) or 2 (tensor cores)");
    m->knnMode = mode;
    return ORBX_OK;
}
extern "C" int orbm_last_knn_path(orbm_t* m) { return m ? m->lastKnnPath : 0; }

static ktc_encode_fn knn_encode() {
    static ktc_encode_fn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (ktc_encode_fn)p;
    });
    return fn;
}

// stage a caller array on the device if it is host memory
template <typename T>
static int stage_in(orbm_handle* m, const T* src, size_t count, T*& buf, size_t& cap, const T** dev) {
    if (m->inSpace == ORBX_MEM_DEVICE) {
        *dev = src;
        return ORBX_OK;
    }
    int rc = ensure(buf, cap, count);
    if (rc) return rc;
    CK(cudaMemcpyAsync(buf, src, count * sizeof(T), cudaMemcpyHostToDevice, m->stream));
    *dev = buf;
    return ORBX_OK;
}

static int ensure_out(orbm_handle* m, size_t nq) {
    if (nq <= m->outCap) return ORBX_OK;
    cudaFree(m->dI1); cudaFree(m->dI2); cudaFree(m->dD1); cudaFree(m->dD2);
    m->dI1 = m->dI2 = nullptr; m->dD1 = m->dD2 = nullptr; m->outCap = 0;
    CK(cudaMalloc(&m->dI1, nq * 4));
    CK(cudaMalloc(&m->dI2, nq * 4));
    CK(cudaMalloc(&m->dD1, nq * 2));
    CK(cudaMalloc(&m->dD2, nq * 2));
    m->outCap = nq;
    return ORBX_OK;
}

extern "C" int orbm_knn2(orbm_t* m, const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* i1, uint16_t* d1,
                         uint16_t* d2) {
    if (!m || !q || nq < 0 || !i1 || !d1 || !d2) FAIL(ORBX_EINVAL, "bad argument");
    if (nq == 0) return ORBX_OK;
    CK(cudaSetDevice(m->device));
    int rc;
    if (t) {
        if ((rc = orbm_set_train(m, t, nt))) return rc;
    } else if (!m->dTrain)
        FAIL(ORBX_ESTATE, "no train set: pass t or call orbm_set_train first");
    nt = m->nTrain;
    const uint8_t* dq = nullptr;
    if ((rc = stage_in(m, q, (size_t)nq * 32, m->dQ, m->qCap, &dq))) return rc;
    if (((uintptr_t)dq & 15) != 0) FAIL(ORBX_EINVAL, "device descriptor matrix must be 16-byte aligned");
    int32_t* o1 = i1; uint16_t* od1 = d1; uint16_t* od2 = d2;
    if (m->outSpace == ORBX_MEM_HOST) {
        if ((rc = ensure_out(m, nq))) return rc;
        o1 = m->dI1; od1 = m->dD1; od2 = m->dD2;
    }
    const int qblocks = (nq + KNN_THREADS * KNN_QPT - 1) / (KNN_THREADS * KNN_QPT);
    const int nTiles = nt / KTC_N;
    const bool useTc = m->knnMode != 1 && nTiles >= 1 && (m->knnMode == 2 || (size_t)nq * (size_t)nt >= ((size_t)1 << 24));
    if (useTc) {
        // ---- tensor-core path: full 256-row train tiles through k_knn_tc, the remainder through k_knn2, ordered merge
        ktc_encode_fn enc = knn_encode();
        if (!enc) FAIL(ORBX_ECUDA, "cuTensorMapEncodeTiled entry point not available from the driver");
        const int qRows = (nq + KTC_M - 1) / KTC_M * KTC_M, qTiles = qRows / KTC_M;
        const int bRows = nTiles * KTC_N, tail = nt - bRows;
        if ((rc = ensure(m->dAm, m->amCap, (size_t)qRows * KTC_KMAIN))) return rc;
        if ((rc = ensure(m->dAe, m->aeCap, (size_t)qRows * KTC_KEXT))) return rc;
        if (!m->trainExpanded) {
            if ((rc = ensure(m->dBm, m->bmCap, (size_t)bRows * KTC_KMAIN))) return rc;
            if ((rc = ensure(m->dBe, m->beCap, (size_t)bRows * KTC_KEXT))) return rc;
            k_knn_expand<<<(bRows * 8 + 255) / 256, 256, 0, m->stream>>>((const uint32_t*)m->dTrain, bRows, bRows, 0, m->dBm,
                                                                         m->dBe);
            m->launches++;
            m->trainExpanded = true;
        }
        k_knn_expand<<<(qRows * 8 + 255) / 256, 256, 0, m->stream>>>((const uint32_t*)dq, nq, qRows, 1, m->dAm, m->dAe);
        m->launches++;
        // train splits: rows inside a split must fit the 22-bit index of the running keys; beyond that, pick the count
        // that fills whole waves of one CTA per SM
        const int minSplits = (nTiles + 16383) / 16384;
        int splits = minSplits;
        double bestScore = -1;
        for (int sp = minSplits; sp <= 64 && (sp == minSplits || nTiles / sp >= 32); sp++) {
            const int tps = (nTiles + sp - 1) / sp, real = (nTiles + tps - 1) / tps;
            const double ctas = (double)qTiles * real, waves = ctas / m->smCount;
            const double eff = waves / ceil(waves);
            const double score = eff - 0.25 / (waves + 1) - 0.001 * real;   // prefer many short waves, few partials
            if (score > bestScore) {
                bestScore = score;
                splits = sp;
            }
        }
        const int tilesPerSplit = (nTiles + splits - 1) / splits;
        splits = (nTiles + tilesPerSplit - 1) / tilesPerSplit;
        const int parts = splits + (tail > 0 ? 1 : 0);
        if ((rc = ensure(m->dPartial, m->partialCap, (size_t)parts * nq))) return rc;
        const int lrc = ktc_launch(enc, m->dAm, m->dAe, qRows, m->dBm, m->dBe, nTiles, nq, tilesPerSplit, splits,
                                   m->dPartial, m->stream);
        if (lrc) FAIL(ORBX_ECUDA, "k_knn_tc launch failed at step %d: %s", lrc, cudaGetErrorString(cudaGetLastError()));
        m->launches++;
        if (tail > 0) {
            k_knn2<<<dim3(qblocks, 1), KNN_THREADS, 0, m->stream>>>((const uint4*)dq, nq, (const uint4*)m->dTrain + 2 * (size_t)bRows,
                                                                    tail, (tail + KNN_TILE - 1) / KNN_TILE * KNN_TILE, bRows,
                                                                    m->dPartial + (size_t)splits * nq, nullptr, nullptr, nullptr);
            m->launches++;
        }
        k_knn2_merge<<<(nq + 255) / 256, 256, 0, m->stream>>>(m->dPartial, nq, parts, o1, od1, od2);
        m->launches++;
        m->lastKnnPath = 2;
    } else {
    // split the train set so that the grid fills the machine for small query counts
    int splits = 1;
    const int target = m->smCount * 8;
    while (qblocks * splits < target && (nt / (splits * 2)) >= 4 * KNN_TILE) splits *= 2;
    int chunk = (nt + splits - 1) / splits;
    chunk = (chunk + KNN_TILE - 1) / KNN_TILE * KNN_TILE;
    splits = (nt + chunk - 1) / chunk;
    int4* partial = nullptr;
    if (splits > 1) {
        if ((rc = ensure(m->dPartial, m->partialCap, (size_t)splits * nq))) return rc;
        partial = m->dPartial;
    }
    k_knn2<<<dim3(qblocks, splits), KNN_THREADS, 0, m->stream>>>((const uint4*)dq, nq, (const uint4*)m->dTrain, nt, chunk, 0,
                                                                 partial, o1, od1, od2);
    m->launches++;
    if (splits > 1) {
        k_knn2_merge<<<(nq + 255) / 256, 256, 0, m->stream>>>(partial, nq, splits, o1, od1, od2);
        m->launches++;
    }
    m->lastKnnPath = 1;
    }
    CK(cudaGetLastError());
    if (m->outSpace == ORBX_MEM_HOST) {
        CK(cudaMemcpyAsync(i1, o1, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
        CK(cudaMemcpyAsync(d1, od1, (size_t)nq * 2, cudaMemcpyDeviceToHost, m->stream));
        CK(cudaMemcpyAsync(d2, od2, (size_t)nq * 2, cudaMemcpyDeviceToHost, m->stream));
    }
    CK(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------ allocation-free 2-NN
// orbm_knn2_static: k_knn2 + k_knn2_merge (the POPC path of orbm_knn2) for callers that cannot call cudaMalloc (CUDA graph
// capture, tight serving loops, harnesses that forbid it). The per-split partial results — the scratch orbm_knn2 keeps in
// m->dPartial — live in a file-scope device array of fixed size; every GPU that loads liborbx holds it until exit.
// No handle, no staging: inputs and outputs are DEVICE pointers; the kernels are enqueued on the caller's stream.
#define KNN_STATIC_MAX_QUERIES 65536
#define KNN_STATIC_MAX_SPLITS 16
static __device__ __align__(256) int4 g_knn2_static_partial[(size_t)KNN_STATIC_MAX_QUERIES * KNN_STATIC_MAX_SPLITS];  // 16 MiB

// In host code `g_knn2_static_partial` is only a stand-in for the symbol, NOT a device address: the address is fetched once
// per device, with that device current, by cudaGetSymbolAddress.
static int knn2_static_scratch(int device, int4** out, int* smCount) {
    static std::mutex mu;
    static int4* addr[64] = {nullptr};
    static int sms[64] = {0};
    if (device < 0 || device >= 64) FAIL(ORBX_EINVAL, "device %d out of range", device);
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[device]) {
        int4* p = nullptr;
        CK(cudaGetSymbolAddress((void**)&p, g_knn2_static_partial));
        if (!p) FAIL(ORBX_ECUDA, "cudaGetSymbolAddress returned a null address for the static 2-NN scratch");
        CK(cudaDeviceGetAttribute(&sms[device], cudaDevAttrMultiProcessorCount, device));
        addr[device] = p;
    }
    *out = addr[device];
    *smCount = sms[device];
    return ORBX_OK;
}

extern "C" int orbm_knn2_static(int device, void* stream, const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* i1,
                                uint16_t* d1, uint16_t* d2) {
    if (!q || !t || !i1 || !d1 || !d2 || nq < 0 || nt < 1) FAIL(ORBX_EINVAL, "bad argument");
    if (nq > KNN_STATIC_MAX_QUERIES)
        FAIL(ORBX_EINVAL, "orbm_knn2_static: %d queries exceed the static scratch (max %d per call; split the batch or use orbm_knn2)",
             nq, KNN_STATIC_MAX_QUERIES);
    if ((((uintptr_t)q) | ((uintptr_t)t)) & 15) FAIL(ORBX_EINVAL, "device descriptor matrices must be 16-byte aligned");
    if (nq == 0) return ORBX_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        FAIL(ORBX_ECUDA, "no CUDA device: liborbx has no CPU path");
    }
    CK(cudaSetDevice(device));
    int4* partial = nullptr;
    int smCount = 0;
    const int src = knn2_static_scratch(device, &partial, &smCount);
    if (src != ORBX_OK) return src;
    cudaStream_t st = (cudaStream_t)stream;
    // the same split rule as orbm_knn2's POPC path, capped by the scratch: splits * nq <= the array
    const int qblocks = (nq + KNN_THREADS * KNN_QPT - 1) / (KNN_THREADS * KNN_QPT);
    int splits = 1;
    const int target = smCount * 8;
    while (qblocks * splits < target && (nt / (splits * 2)) >= 4 * KNN_TILE && splits * 2 <= KNN_STATIC_MAX_SPLITS) splits *= 2;
    int chunk = (nt + splits - 1) / splits;
    chunk = (chunk + KNN_TILE - 1) / KNN_TILE * KNN_TILE;
    splits = (nt + chunk - 1) / chunk;
    // every (split, query) entry of the scratch that the merge reads is written by k_knn2 in this call: nothing has to start
    // at zero, and what an earlier call left behind is never read
    k_knn2<<<dim3(qblocks, splits), KNN_THREADS, 0, st>>>((const uint4*)q, nq, (const uint4*)t, nt, chunk, 0, partial, nullptr,
                                                          nullptr, nullptr);
    k_knn2_merge<<<(nq + 255) / 256, 256, 0, st>>>(partial, nq, splits, i1, d1, d2);
    CK(cudaGetLastError());
    return ORBX_OK;
}

extern "C" int orbm_window_search(orbm_t* m, const uint8_t* q, int nq, const int32_t* cand_off, const int32_t* cand_idx,
                                  const uint8_t* cand_skip, const uint8_t* t, int nt, int32_t* best_idx, uint16_t* best_d,
                                  uint16_t* second_d, int32_t* second_idx) {
    if (!m || !q || nq < 0 || !cand_off || !best_idx || !best_d || !second_d || !second_idx) FAIL(ORBX_EINVAL, "bad argument");
    if (nq == 0) return ORBX_OK;
    CK(cudaSetDevice(m->device));
    int rc;
    if (t) {
        if ((rc = orbm_set_train(m, t, nt))) return rc;
    } else if (!m->dTrain)
        FAIL(ORBX_ESTATE, "no train set");
    const uint8_t* dq = nullptr;
    if ((rc = stage_in(m, q, (size_t)nq * 32, m->dQ, m->qCap, &dq))) return rc;
    const int* doff = nullptr;
    const int* didx = nullptr;
    const uint8_t* dskip = nullptr;
    int ncand = 0;
    if (m->inSpace == ORBX_MEM_HOST) {
        ncand = cand_off[nq];
        for (int a = 0; a < nq; a++)
            if (cand_off[a] > cand_off[a + 1] || cand_off[a] < 0) FAIL(ORBX_EINVAL, "cand_off must be non-decreasing");
        if (ncand > 0 && !cand_idx) FAIL(ORBX_EINVAL, "null cand_idx");
        for (int c = 0; c < ncand; c++)
            if (cand_idx[c] < 0 || cand_idx[c] >= m->nTrain) FAIL(ORBX_EINVAL, "candidate index %d out of range", cand_idx[c]);
    } else {
        CK(cudaMemcpyAsync(&ncand, cand_off + nq, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
        CK(cudaStreamSynchronize(m->stream));
    }
    if ((rc = stage_in(m, cand_off, (size_t)nq + 1, m->dOff, m->offCap, &doff))) return rc;
    if (ncand > 0) {
        if ((rc = stage_in(m, cand_idx, (size_t)ncand, m->dIdx, m->idxCap, &didx))) return rc;
        if (cand_skip && (rc = stage_in(m, cand_skip, (size_t)ncand, m->dSkip, m->skipCap, &dskip))) return rc;
    }
    int32_t* o1 = best_idx; int32_t* o2 = second_idx; uint16_t* od1 = best_d; uint16_t* od2 = second_d;
    if (m->outSpace == ORBX_MEM_HOST) {
        if ((rc = ensure_out(m, nq))) return rc;
        o1 = m->dI1; o2 = m->dI2; od1 = m->dD1; od2 = m->dD2;
    }
    k_window_search<<<(nq + 7) / 8, 256, 0, m->stream>>>((const uint4*)dq, nq, doff, didx, dskip, (const uint4*)m->dTrain, o1,
                                                         od1, od2, o2);
    m->launches++;
    CK(cudaGetLastError());
    if (m->outSpace == ORBX_MEM_HOST) {
        CK(cudaMemcpyAsync(best_idx, o1, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
        CK(cudaMemcpyAsync(second_idx, o2, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
        CK(cudaMemcpyAsync(best_d, od1, (size_t)nq * 2, cudaMemcpyDeviceToHost, m->stream));
        CK(cudaMemcpyAsync(second_d, od2, (size_t)nq * 2, cudaMemcpyDeviceToHost, m->stream));
    }
    CK(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

extern "C" int orbm_distance(orbm_t* m, const uint8_t* a, const uint8_t* b, int n, uint16_t* d) {
    if (!m || !a || !b || n < 0 || !d) FAIL(ORBX_EINVAL, "bad argument");
    if (n == 0) return ORBX_OK;
    CK(cudaSetDevice(m->device));
    int rc;
    const uint8_t* da = nullptr;
    const uint8_t* db = nullptr;
    if ((rc = stage_in(m, a, (size_t)n * 32, m->dQ, m->qCap, &da))) return rc;
    if ((rc = stage_in(m, b, (size_t)n * 32, m->dB, m->bCap, &db))) return rc;
    uint16_t* od = d;
    if (m->outSpace == ORBX_MEM_HOST) {
        if ((rc = ensure_out(m, n))) return rc;
        od = m->dD1;
    }
    k_distance<<<(n + 255) / 256, 256, 0, m->stream>>>((const uint4*)da, (const uint4*)db, n, od);
    m->launches++;
    CK(cudaGetLastError());
    if (m->outSpace == ORBX_MEM_HOST) CK(cudaMemcpyAsync(d, od, (size_t)n * 2, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------ stereo ABI
// (the SAD / uRight / depth scratch lives on the left extractor's handle: orbx_stereo_scratch, freed by orbx_destroy)

extern "C" int orbx_stereo_match_batch(orbx_t* left, orbx_t* right, float mbf, float minD, float maxD, float* uRight,
                                       float* depth, int cap) {
    if (!left || !right || !uRight || !depth || cap < 1) FAIL(ORBX_EINVAL, "bad argument");
    OrbxView A = orbx_view((orbx_handle*)left), B = orbx_view((orbx_handle*)right);
    if (!A.lastBatch || A.lastBatch != B.lastBatch) FAIL(ORBX_ESTATE, "left and right must hold the same extracted batch");
    if (A.device != B.device) FAIL(ORBX_EINVAL, "left and right extractors must be on the same device");
    if (A.hp->nlevels != B.hp->nlevels || A.hp->selStride != B.hp->selStride || A.hp->lv[0].w != B.hp->lv[0].w ||
        A.hp->lv[0].h != B.hp->lv[0].h)
        FAIL(ORBX_EINVAL, "left and right extractors must share parameters and image shape");
    CK(cudaSetDevice(A.device));
    const int batch = A.lastBatch, stride = A.hp->selStride;
    const size_t need = (size_t)batch * stride;
    int* dSad = nullptr; float *dU = nullptr, *dZ = nullptr;
    { int rcS = orbx_stereo_scratch((orbx_handle*)left, need, &dSad, &dU, &dZ); if (rcS != ORBX_OK) return rcS; }
    cudaStream_t st = A.stream;
    CK(cudaStreamSynchronize(B.stream));  // right extractor's results must be complete (its extract call already synced)
    float* oU = A.outSpace == ORBX_MEM_DEVICE ? uRight : dU;
    float* oZ = A.outSpace == ORBX_MEM_DEVICE ? depth : dZ;
    const int oStride = A.outSpace == ORBX_MEM_DEVICE ? cap : stride;
    int maxN = 0;
    for (int f = 0; f < batch; f++) maxN = std::max(maxN, A.hCounts[f]);
    if (maxN > cap) FAIL(ORBX_ECAP, "cap_per_frame %d < %d left keypoints", cap, maxN);
    if (maxN == 0) maxN = 1;
    k_stereo_match<<<dim3((maxN + 7) / 8, batch), 256, 0, st>>>(A.dPlan, A.dKps, A.dDesc, A.dNOut, B.dKps, B.dDesc, B.dNOut,
                                                               B.dSelCount, minD, maxD, oU, oZ, dSad, oStride);
    k_stereo_sad<<<dim3((maxN + 7) / 8, batch), 256, 0, st>>>(A.dPlan, A.dKps, A.dNOut, A.dPyr, B.dKps, B.dPyr, mbf, minD, maxD, oU,
                                                             oZ, dSad, oStride);
    int P2 = 1;
    while (P2 < maxN) P2 <<= 1;
    // more than 6144 left keypoints need over 48 KB of dynamic shared memory: opt in (up to 227 KB = 16384 keypoints,
    // the 4K / 8000-keypoint config has ~8000), and say so clearly beyond that instead of failing the launch
    if ((size_t)P2 * 8 > 227 * 1024) FAIL(ORBX_ESHAPE, "%d left keypoints exceed the stereo median cut's 28k-entry sort buffer", maxN);
    if ((size_t)P2 * 8 > 48 * 1024) {
        static std::mutex mu;
        static int optedIn[64] = {0};
        std::lock_guard<std::mutex> lk(mu);
        if (A.device < 64 && optedIn[A.device] < P2 * 8) {
            CK(cudaFuncSetAttribute(k_stereo_cut, cudaFuncAttributeMaxDynamicSharedMemorySize, P2 * 8));
            optedIn[A.device] = P2 * 8;
        }
    }
    k_stereo_cut<<<batch, 1024, (size_t)P2 * 8, st>>>(A.dPlan, A.dNOut, dSad, oU, oZ, oStride, P2);
    orbx_count_launch((orbx_handle*)left, 3);
    CK(cudaGetLastError());
    if (A.outSpace == ORBX_MEM_HOST) {
        const int cols = std::min(cap, stride);
        CK(cudaMemcpy2DAsync(uRight, (size_t)cap * 4, oU, (size_t)stride * 4, (size_t)cols * 4, batch, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpy2DAsync(depth, (size_t)cap * 4, oZ, (size_t)stride * 4, (size_t)cols * 4, batch, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    return ORBX_OK;
}

extern "C" int orbx_stereo_match(orbx_t* left, orbx_t* right, int frame, float mbf, float minD, float maxD, float* uRight,
                                 float* depth) {
    if (!left || !right) FAIL(ORBX_EINVAL, "null handle");
    OrbxView A = orbx_view((orbx_handle*)left);
    if (frame < 0 || frame >= A.lastBatch) FAIL(ORBX_ESTATE, "frame %d not in the last extracted batch", frame);
    if (A.lastBatch == 1) return orbx_stereo_match_batch(left, right, mbf, minD, maxD, uRight, depth, std::max(A.hCounts[0], 1));
    // general case: run the batch and hand back one row (host outputs only)
    if (A.outSpace != ORBX_MEM_HOST) FAIL(ORBX_EINVAL, "single-frame stereo on a multi-frame batch needs host outputs");
    const int stride = A.hp->selStride;
    std::vector<float> u((size_t)A.lastBatch * stride), z((size_t)A.lastBatch * stride);
    int rc = orbx_stereo_match_batch(left, right, mbf, minD, maxD, u.data(), z.data(), stride);
    if (rc) return rc;
    const int n = A.hCounts[frame];
    std::copy(u.begin() + (size_t)frame * stride, u.begin() + (size_t)frame * stride + n, uRight);
    std::copy(z.begin() + (size_t)frame * stride, z.begin() + (size_t)frame * stride + n, depth);
    return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------ windowed search ABI
#include "orbx_search.cuh"

namespace {
struct WsCursor {  // bump allocator over the handle's workspace (256-byte aligned pieces)
    uint8_t* base; size_t off;
    template <typename T> T* take(size_t count) {
        T* p = (T*)(base + off);
        off += (count * sizeof(T) + 255) & ~(size_t)255;
        return p;
    }
};
}  // namespace

// Host-input staging: the caller's arrays (a dozen small ones per search) are gathered into a pinned mirror of the
// workspace prefix and cross PCIe as ONE copy; with device input the pointers are used in place.
struct Stager {
    orbm_handle* m; WsCursor* W; bool host; size_t begin;
    const void* in(const void* src, size_t bytes) {
        if (!src) return nullptr;
        if (!host) return src;
        uint8_t* d = W->take<uint8_t>(bytes);
        const size_t off = (size_t)(d - W->base);
        if (begin == (size_t)-1) begin = off;
        memcpy(m->hIn + (off - begin), src, bytes);
        return d;
    }
    int flush(cudaStream_t st) {
        if (!host || begin == (size_t)-1) return ORBX_OK;
        CK(cudaMemcpyAsync(W->base + begin, m->hIn, W->off - begin, cudaMemcpyHostToDevice, st));
        return ORBX_OK;
    }
};
static int ensure_host_in(orbm_handle* m, size_t bytes) {
    if (bytes <= m->hInCap) return ORBX_OK;
    if (m->hIn) cudaFreeHost(m->hIn);
    m->hIn = nullptr;
    m->hInCap = 0;
    CK(cudaMallocHost(&m->hIn, bytes));
    m->hInCap = bytes;
    return ORBX_OK;
}

// Queries of either search (mode 0: last-frame map points, mode 1: local map points with their tracking projection).
struct SpHostQueries {
    int n;
    const float* world_pos; const uint8_t* flags; const int32_t* octave; const float* angle; const uint8_t* desc;
    const float* proj_x; const float* proj_y; const float* proj_xr; const float* view_cos;
};

static int sp_run(orbm_handle* m, int mode, const orbm_proj_params* p, float nnRatio, const SpHostQueries* last,
                  const orbm_cur_frame* cur, int32_t* cur_match, int32_t* last_match, uint8_t* last_keep, int* nmatches) {
    const int nL = last->n, nC = cur->n;
    if (nL < 0 || nC < 0) FAIL(ORBX_EINVAL, "negative count");
    if ((nL && (!last->flags || !last->octave || !last->desc || !last_match)) || (nC && (!cur->keys_un || !cur->desc || !cur_match)))
        FAIL(ORBX_EINVAL, "null array");
    if (nL && mode == 0 && (!last->world_pos || !last->angle || !last_keep)) FAIL(ORBX_EINVAL, "null array");
    if (nL && mode == 1 && (!last->proj_x || !last->proj_y || !last->proj_xr || !last->view_cos)) FAIL(ORBX_EINVAL, "null array");
    if (nL && mode == 2 && (!last->proj_x || !last->proj_y)) FAIL(ORBX_EINVAL, "null array");
    if (p->nlevels < 1 || p->nlevels > ORBX_MAX_LEVELS) FAIL(ORBX_EINVAL, "nlevels out of range");
    if (nL > (1 << 20) || nC > (1 << 20)) FAIL(ORBX_EINVAL, "frame too large");
    *nmatches = 0;
    CK(cudaSetDevice(m->device));
    cudaStream_t st = m->stream;
    const bool hin = m->inSpace == ORBX_MEM_HOST, hout = m->outSpace == ORBX_MEM_HOST;
    if (nL == 0 || nC == 0) {  // nothing can match: outputs are all "none"
        if (hout) {
            for (int i = 0; i < nC; i++) cur_match[i] = -1;
            for (int i = 0; i < nL; i++) { last_match[i] = -1; if (last_keep) last_keep[i] = 0; }
        } else {
            if (nC) CK(cudaMemsetAsync(cur_match, 0xFF, (size_t)nC * 4, st));
            if (nL) { CK(cudaMemsetAsync(last_match, 0xFF, (size_t)nL * 4, st)); if (last_keep) CK(cudaMemsetAsync(last_keep, 0, nL, st)); }
            CK(cudaStreamSynchronize(st));
        }
        return ORBX_OK;
    }
    // workspace: staged inputs (host mode), grid, per-query state, outputs (host mode)
    const size_t need = ((size_t)nL * (12 + 1 + 4 + 4 + 32 + 4 + 4 + 4 + 4 + 1 + 4 + 16) + (size_t)nC * (28 + 4 + 32 + 1 + 4 + 4 + 4) +
                         (size_t)SP_CELLS * 12 + 64) + 256 * 40;
    int rc;
    if ((rc = ensure(m->dWs, m->wsCap, need))) return rc;
    if (!m->hRes) CK(cudaMallocHost(&m->hRes, 4 * sizeof(int)));
    if (m->hFlagsCap < (size_t)nL + 8) {
        if (m->hFlags) cudaFreeHost(m->hFlags);
        m->hFlags = nullptr;
        CK(cudaMallocHost(&m->hFlags, ((size_t)nL + 8) * sizeof(int)));
        m->hFlagsCap = (size_t)nL + 8;
    }
    WsCursor W{m->dWs, 0};
    SpDev S;
    S.p = *p;
    S.nLast = nL;
    S.nCur = nC;
    S.mode = mode;
    S.nnRatio = nnRatio;
    if ((rc = ensure_host_in(m, need))) return rc;  // staging of host inputs and the pinned mirror of the read-back span
    Stager stg{m, &W, hin, (size_t)-1};
    auto in = [&](const void* src, size_t bytes) -> const void* { return stg.in(src, bytes); };
    S.worldPos = (const float*)in(last->world_pos, (size_t)nL * 12);
    S.lflags = (const uint8_t*)in(last->flags, nL);
    S.loct = (const int32_t*)in(last->octave, (size_t)nL * 4);
    S.langle = (const float*)in(last->angle, (size_t)nL * 4);
    S.ldesc = (const uint4*)in(last->desc, (size_t)nL * 32);
    S.projX = (const float*)in(last->proj_x, (size_t)nL * 4);
    S.projY = (const float*)in(last->proj_y, (size_t)nL * 4);
    S.projXR = (const float*)in(last->proj_xr, (size_t)nL * 4);
    S.viewCos = (const float*)in(last->view_cos, (size_t)nL * 4);
    S.keys = (const orbx_kp*)in(cur->keys_un, (size_t)nC * sizeof(orbx_kp));
    S.uRight = (const float*)in(cur->u_right, (size_t)nC * 4);
    S.cdesc = (const uint4*)in(cur->desc, (size_t)nC * 32);
    S.taken = (const uint8_t*)in(cur->taken, nC);
    if ((((uintptr_t)S.ldesc) | ((uintptr_t)S.cdesc)) & 15) FAIL(ORBX_EINVAL, "device descriptor matrices must be 16-byte aligned");
    if ((rc = stg.flush(st))) return rc;
    int* cellCount = W.take<int>(SP_CELLS);
    int* cellStart = W.take<int>(SP_CELLS + 1);
    int* cursor = W.take<int>(SP_CELLS);
    int* items = W.take<int>(nC);
    int* candCount = W.take<int>(nL);
    int* candOff = W.take<int>(nL + 1);
    int* choice = W.take<int>(nL);
    int* owner = W.take<int>(nC);
    const int maxRounds = nL + 8;
    // everything the host reads back sits in one span: [outputs (host mode)] [dRes] [go flag, round flags]
    int32_t* oCur = hout ? W.take<int32_t>(nC) : cur_match;
    int32_t* oLast = hout ? W.take<int32_t>(nL) : last_match;
    uint8_t* oKeep = ((hout && last_keep) || (mode == 2 && !last_keep)) ? W.take<uint8_t>(nL) : last_keep;
    int* dRes = W.take<int>(4);
    int* flagsGo = W.take<int>(maxRounds + 1);
    int* flags = flagsGo + 1;  // flags[-1] = go
    if (W.off > m->wsCap) FAIL(ORBX_ECUDA, "internal: workspace estimate too small (%zu > %zu)", W.off, m->wsCap);
    const uint8_t* spanBegin = hout ? (const uint8_t*)oCur : (const uint8_t*)dRes;
    const int SPEC = 4;  // rounds enqueued before the first look at the flags
    const size_t spanBytes = (size_t)((const uint8_t*)(flags + SPEC) - spanBegin);
    const uint8_t* hs = m->hIn;  // pinned mirror of the span (the staged inputs have left it when the copy back starts)
    auto hostAt = [&](const void* dev) { return hs + ((const uint8_t*)dev - spanBegin); };

    // S1 grid
    if (nC <= m->gridFusedMax) {  // camera-sized frame: the whole grid build in one CTA, one launch
        k_sp_grid_build<<<1, 1024, 0, st>>>(S, cellStart, items);
        m->launches -= 3;  // (the call sites below count the four kernels of the sequence)
    } else {
    CK(cudaMemsetAsync(cellCount, 0, SP_CELLS * sizeof(int), st));
    k_sp_grid_count<<<(nC + 255) / 256, 256, 0, st>>>(S, cellCount);
    k_sp_grid_scan<<<1, 1024, 0, st>>>(cellCount, cellStart, cursor);
    k_sp_grid_fill<<<(nC + 255) / 256, 256, 0, st>>>(S, cursor, items);
    k_sp_grid_sort<<<(SP_CELLS + 255) / 256, 256, 0, st>>>(cellStart, items);
    }
    // S2 candidates: count, scan, fill — without a host round trip for the total (k_sp_guard)
    const int qBlocks = (nL + 7) / 8;
    k_sp_candidates<false><<<qBlocks, 256, 0, st>>>(S, cellStart, items, candCount, nullptr, nullptr, nullptr);
    k_sp_scan<<<1, 1024, 0, st>>>(candCount, nL, candOff);
    m->launches += 6;
    if ((rc = ensure(m->dCand, m->candCap, (std::max<size_t>((size_t)nL * 64, 1 << 16) + 64) * 6))) return rc;
    size_t total = 0;
    int round = 0, rounds = 0;
    int* candIdx = nullptr;
    unsigned short* candDist = nullptr;
    auto enqueueRounds = [&](int R) {
        for (int r = 0; r < R; r++, round++) {
            cudaMemsetAsync(owner, 0x7F, (size_t)nC * sizeof(int), st);
            k_sp_claim<<<(nL + 255) / 256, 256, 0, st>>>(S, choice, owner, flags, round);
            if (mode != 1) k_sp_choose<<<qBlocks, 256, 0, st>>>(S, candOff, candIdx, candDist, choice, owner, flags, round);
            else k_sp_choose_mp<<<(nL + 255) / 256, 256, 0, st>>>(S, candOff, candIdx, candDist, choice, owner, flags, round);
            m->launches += 2;
        }
    };
    auto finishAndFetch = [&]() -> int {
        if (mode != 1) k_sp_finish<<<1, 1024, 0, st>>>(S, choice, oCur, oLast, oKeep, dRes);
        else k_sp_finish_mp<<<1, 1024, 0, st>>>(S, choice, oCur, oLast, dRes);
        m->launches++;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(m->hIn, spanBegin, spanBytes, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return ORBX_OK;
    };
    for (int attempt = 0;; attempt++) {
        const size_t capEntries = ((m->candCap / 6) & ~(size_t)63) - 64;
        candIdx = (int*)m->dCand;
        candDist = (unsigned short*)(m->dCand + capEntries * 4);
        // S3 greedy claims as a fixed point: SPEC rounds and the finishing kernel are enqueued speculatively (converged
        // rounds return at once) and ONE copy brings back results, flags and the candidate total; more rounds follow
        // only if the last speculative one still changed something
        CK(cudaMemsetAsync(choice, 0xFF, (size_t)nL * sizeof(int), st));
        CK(cudaMemsetAsync(flagsGo, 0, (size_t)(maxRounds + 1) * sizeof(int), st));
        k_sp_guard<<<1, 1, 0, st>>>(candOff, nL, (int)std::min<size_t>(capEntries, 0x7FFFFFFF), flagsGo, dRes);
        k_sp_candidates<true><<<qBlocks, 256, 0, st>>>(S, cellStart, items, nullptr, candOff, candIdx, candDist, flagsGo);
        m->launches += 2;
        round = 0;
        enqueueRounds(SPEC);
        if ((rc = finishAndFetch())) return rc;
        total = (size_t)((const int*)hostAt(dRes))[2];
        if (((const int*)hostAt(flagsGo))[0]) break;
        if (attempt) FAIL(ORBX_ECUDA, "internal: candidate buffer still too small (%zu entries)", total);
        if ((rc = ensure(m->dCand, m->candCap, (total + total / 2 + 128) * 6))) return rc;
    }
    const int* hf = (const int*)hostAt(flags);
    rounds = 0;
    while (rounds < SPEC && hf[rounds]) rounds++;
    if (rounds == SPEC) {  // not converged yet: further rounds with the flags read in between, then finish again
        for (;;) {
            // rounds per host synchronisation double (4, 8, 16, ...): a long claim chain (dense repeated texture, every
            // query displaced by the previous one) costs O(log) synchronisations and at most twice the rounds it needs
            const int R = std::min(std::max(4, round), maxRounds - round);
            if (R <= 0) FAIL(ORBX_ECUDA, "internal: greedy resolution did not converge in %d rounds", maxRounds);
            enqueueRounds(R);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(m->hFlags, flags, (size_t)round * sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            rounds = 0;
            while (rounds < round && m->hFlags[rounds]) rounds++;
            if (rounds < round) break;  // round `rounds` changed nothing: fixed point reached
        }
        if ((rc = finishAndFetch())) return rc;
    }
    rounds++;  // the confirming round
    if (hout) {
        memcpy(cur_match, hostAt(oCur), (size_t)nC * 4);
        memcpy(last_match, hostAt(oLast), (size_t)nL * 4);
        if (last_keep) memcpy(last_keep, hostAt(oKeep), nL);
    }
    *nmatches = ((const int*)hostAt(dRes))[0];
    m->lastRounds = rounds;
    m->lastCandidates = (int64_t)total;
    return ORBX_OK;
}

extern "C" int orbm_search_by_projection(orbm_t* m, const orbm_proj_params* p, const orbm_last_frame* last,
                                         const orbm_cur_frame* cur, int32_t* cur_match, int32_t* last_match,
                                         uint8_t* last_keep, int* nmatches) {
    if (!m || !p || !last || !cur || !nmatches) FAIL(ORBX_EINVAL, "null argument");
    SpHostQueries Q = {last->n, last->world_pos, last->flags, last->octave, last->angle, last->desc, nullptr, nullptr, nullptr, nullptr};
    return sp_run(m, 0, p, 0.f, &Q, cur, cur_match, last_match, last_keep, nmatches);
}

extern "C" int orbm_search_local_points(orbm_t* m, const orbm_proj_params* p, float nn_ratio, const orbm_map_points* mps,
                                        const orbm_cur_frame* cur, int32_t* cur_match, int32_t* mp_match, int* nmatches) {
    if (!m || !p || !mps || !cur || !nmatches) FAIL(ORBX_EINVAL, "null argument");
    SpHostQueries Q = {mps->n, nullptr, mps->flags, mps->level, nullptr, mps->desc, mps->proj_x, mps->proj_y, mps->proj_xr, mps->view_cos};
    return sp_run(m, 1, p, nn_ratio, &Q, cur, cur_match, mp_match, nullptr, nmatches);
}

extern "C" int orbm_search_kf_projection(orbm_t* m, const orbm_proj_params* p, const orbm_fuse_points* pts,
                                        const orbm_cur_frame* kf, int32_t* kf_match, int32_t* pt_match, int* nmatches) {
    if (!m || !p || !pts || !kf || !nmatches) FAIL(ORBX_EINVAL, "null argument");
    orbm_proj_params P = *p;
    P.check_orientation = 0;  // the reference applies no rotation filter here
    SpHostQueries Q = {pts->n, nullptr, pts->valid, pts->level, nullptr, pts->desc, pts->proj_x, pts->proj_y, nullptr, nullptr};
    orbm_cur_frame K = *kf;
    K.u_right = nullptr;  // no stereo gate in this search
    return sp_run(m, 2, &P, 0.f, &Q, &K, kf_match, pt_match, nullptr, nmatches);
}

extern "C" int orbm_search_for_initialization(orbm_t* m, const orbm_proj_params* p, float nn_ratio, int th_low,
                                              int window_size, int n1, const orbx_kp* keys1, const uint8_t* desc1,
                                              float* prev_matched, const orbm_cur_frame* f2, int32_t* matches12,
                                              int* nmatches) {
    if (!m || !p || !f2 || !nmatches) FAIL(ORBX_EINVAL, "null argument");
    const int nL = n1, nC = f2->n;
    if (nL < 0 || nC < 0) FAIL(ORBX_EINVAL, "negative count");
    if ((nL && (!keys1 || !desc1 || !prev_matched || !matches12)) || (nC && (!f2->keys_un || !f2->desc)))
        FAIL(ORBX_EINVAL, "null array");
    if (nL > (1 << 20) || nC > (1 << 20)) FAIL(ORBX_EINVAL, "frame too large");
    *nmatches = 0;
    CK(cudaSetDevice(m->device));
    cudaStream_t st = m->stream;
    const bool hin = m->inSpace == ORBX_MEM_HOST, hout = m->outSpace == ORBX_MEM_HOST;
    if (nL == 0) return ORBX_OK;
    if (nC == 0) {  // vnMatches12 = all -1, vbPrevMatched untouched
        if (hout) for (int i = 0; i < nL; i++) matches12[i] = -1;
        else { CK(cudaMemsetAsync(matches12, 0xFF, (size_t)nL * 4, st)); CK(cudaStreamSynchronize(st)); }
        return ORBX_OK;
    }
    const size_t need = (size_t)nL * (28 + 32 + 8 + 4 + 4 + 4 + 4 + 4 + 8) + (size_t)nC * (28 + 32 + 4 + 4 + 4) +
                        (size_t)SP_CELLS * 12 + 256 * 32;
    int rc;
    if ((rc = ensure(m->dWs, m->wsCap, need))) return rc;
    if (!m->hRes) CK(cudaMallocHost(&m->hRes, 4 * sizeof(int)));
    if (hin && (rc = ensure_host_in(m, need))) return rc;
    WsCursor W{m->dWs, 0};
    Stager stg{m, &W, hin, (size_t)-1};
    SpDev S;
    memset(&S, 0, sizeof(S));
    S.p = *p;
    S.p.th = (float)window_size;  // GetFeaturesInArea(x, y, windowSize, 0, 0): the int becomes the float radius
    S.nLast = nL;
    S.nCur = nC;
    S.mode = 3;
    S.nnRatio = nn_ratio;
    S.qkeys = (const orbx_kp*)stg.in(keys1, (size_t)nL * sizeof(orbx_kp));
    S.ldesc = (const uint4*)stg.in(desc1, (size_t)nL * 32);
    S.projX = (const float*)stg.in(prev_matched, (size_t)nL * 8);
    S.keys = (const orbx_kp*)stg.in(f2->keys_un, (size_t)nC * sizeof(orbx_kp));
    S.cdesc = (const uint4*)stg.in(f2->desc, (size_t)nC * 32);
    if ((((uintptr_t)S.ldesc) | ((uintptr_t)S.cdesc)) & 15) FAIL(ORBX_EINVAL, "device descriptor matrices must be 16-byte aligned");
    if ((rc = stg.flush(st))) return rc;
    int* cellCount = W.take<int>(SP_CELLS);
    int* cellStart = W.take<int>(SP_CELLS + 1);
    int* cursor = W.take<int>(SP_CELLS);
    int* items = W.take<int>(nC);
    int* candCount = W.take<int>(nL);
    int* candOff = W.take<int>(nL + 1);
    int* matchedDist = W.take<int>(nC);
    int* match21 = W.take<int>(nC);
    int* bin12 = W.take<int>(nL);
    int* dRes = W.take<int>(4);
    int32_t* o12 = hout ? W.take<int32_t>(nL) : matches12;
    float* oPrev = hout ? W.take<float>((size_t)nL * 2) : prev_matched;
    if (W.off > m->wsCap) FAIL(ORBX_ECUDA, "internal: workspace estimate too small (%zu > %zu)", W.off, m->wsCap);

    if (nC <= m->gridFusedMax) {  // camera-sized frame: the whole grid build in one CTA, one launch
        k_sp_grid_build<<<1, 1024, 0, st>>>(S, cellStart, items);
        m->launches -= 3;  // (the call sites below count the four kernels of the sequence)
    } else {
    CK(cudaMemsetAsync(cellCount, 0, SP_CELLS * sizeof(int), st));
    k_sp_grid_count<<<(nC + 255) / 256, 256, 0, st>>>(S, cellCount);
    k_sp_grid_scan<<<1, 1024, 0, st>>>(cellCount, cellStart, cursor);
    k_sp_grid_fill<<<(nC + 255) / 256, 256, 0, st>>>(S, cursor, items);
    k_sp_grid_sort<<<(SP_CELLS + 255) / 256, 256, 0, st>>>(cellStart, items);
    }
    const int qBlocks = (nL + 7) / 8;
    k_sp_candidates<false><<<qBlocks, 256, 0, st>>>(S, cellStart, items, candCount, nullptr, nullptr, nullptr);
    k_sp_scan<<<1, 1024, 0, st>>>(candCount, nL, candOff);
    m->launches += 6;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(m->hRes + 2, candOff + nL, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    const size_t total = (size_t)m->hRes[2];
    if ((rc = ensure(m->dCand, m->candCap, (total + 64) * 6))) return rc;
    int* candIdx = (int*)m->dCand;
    unsigned short* candDist = (unsigned short*)(m->dCand + ((total + 63) & ~(size_t)63) * 4);
    k_sp_candidates<true><<<qBlocks, 256, 0, st>>>(S, cellStart, items, nullptr, candOff, candIdx, candDist);
    CK(cudaMemsetAsync(o12, 0xFF, (size_t)nL * sizeof(int), st));
    CK(cudaMemsetAsync(bin12, 0xFF, (size_t)nL * sizeof(int), st));
    const size_t stageBytes = 2 * ((size_t)INIT_STAGE * 4 + ((size_t)INIT_STAGE + 2) * 2) + 16;
    const size_t smem = (size_t)nC * 8 + stageBytes;
    if (smem <= 220 * 1024) {  // claim state in shared memory
        CK(cudaFuncSetAttribute(k_init_greedy<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_init_greedy<true><<<1, 32, smem, st>>>(S, candOff, candIdx, candDist, th_low, nullptr, nullptr, o12, bin12, oPrev, dRes);
    } else {
        CK(cudaMemsetAsync(matchedDist, 0x7F, (size_t)nC * sizeof(int), st));  // 0x7F7F7F7F: above every distance
        CK(cudaMemsetAsync(match21, 0xFF, (size_t)nC * sizeof(int), st));
        CK(cudaFuncSetAttribute(k_init_greedy<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stageBytes));
        k_init_greedy<false><<<1, 32, stageBytes, st>>>(S, candOff, candIdx, candDist, th_low, matchedDist, match21, o12, bin12, oPrev, dRes);
    }
    m->launches += 2;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(m->hRes, dRes, sizeof(int), cudaMemcpyDeviceToHost, st));
    if (hout) {
        CK(cudaMemcpyAsync(matches12, o12, (size_t)nL * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(prev_matched, oPrev, (size_t)nL * 8, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    *nmatches = m->hRes[0];
    m->lastRounds = 1;
    m->lastCandidates = (int64_t)total;
    return ORBX_OK;
}

extern "C" int orbm_distinctive_descriptors(orbm_t* m, const uint8_t* desc, const int32_t* off, int npoints,
                                            int32_t* best_idx, int32_t* best_median) {
    if (!m || !off || npoints < 0 || !best_idx) FAIL(ORBX_EINVAL, "bad argument");
    if (npoints == 0) return ORBX_OK;
    CK(cudaSetDevice(m->device));
    cudaStream_t st = m->stream;
    const bool hin = m->inSpace == ORBX_MEM_HOST, hout = m->outSpace == ORBX_MEM_HOST;
    // the offsets are needed on the host for the size checks (max observations -> shared memory)
    std::vector<int32_t> hoff;
    const int32_t* ho = off;
    if (!hin) {
        hoff.resize((size_t)npoints + 1);
        CK(cudaMemcpyAsync(hoff.data(), off, ((size_t)npoints + 1) * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        ho = hoff.data();
    }
    int maxN = 0;
    for (int p = 0; p < npoints; p++) {
        if (ho[p + 1] < ho[p]) FAIL(ORBX_EINVAL, "offsets must be non-decreasing");
        maxN = std::max(maxN, ho[p + 1] - ho[p]);
    }
    if (maxN > DD_MAX_OBS) FAIL(ORBX_ESHAPE, "a map point has %d observations (limit %d)", maxN, DD_MAX_OBS);
    const size_t total = (size_t)ho[npoints] - (size_t)ho[0];
    if (total && !desc) FAIL(ORBX_EINVAL, "null descriptors");
    const size_t need = total * 32 + ((size_t)npoints + 1) * 4 + (size_t)npoints * 8 + 1024;
    int rc;
    if ((rc = ensure(m->dWs, m->wsCap, need))) return rc;
    WsCursor W{m->dWs, 0};
    const uint4* dDesc = (const uint4*)desc;
    const int* dOff = off;
    if (hin) {
        uint8_t* d = W.take<uint8_t>(std::max<size_t>(total * 32, 16));
        if (total) CK(cudaMemcpyAsync(d, desc + (size_t)ho[0] * 32, total * 32, cudaMemcpyHostToDevice, st));
        dDesc = (const uint4*)d - 2 * (size_t)ho[0];  // so that row off[p] is addressed as given
        int* o = W.take<int>((size_t)npoints + 1);
        CK(cudaMemcpyAsync(o, off, ((size_t)npoints + 1) * 4, cudaMemcpyHostToDevice, st));
        dOff = o;
    } else if ((uintptr_t)desc & 15)
        FAIL(ORBX_EINVAL, "device descriptor matrix must be 16-byte aligned");
    int32_t* oIdx = hout ? W.take<int32_t>(npoints) : best_idx;
    int32_t* oMed = (hout || !best_median) ? W.take<int32_t>(npoints) : best_median;
    const size_t smem = (size_t)std::max(maxN, 1) * 32;
    if (smem > 48 * 1024) CK(cudaFuncSetAttribute(k_distinctive, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_distinctive<<<npoints, DD_THREADS, smem, st>>>(dDesc, dOff, npoints, oIdx, oMed);
    m->launches++;
    CK(cudaGetLastError());
    if (hout) {
        CK(cudaMemcpyAsync(best_idx, oIdx, (size_t)npoints * 4, cudaMemcpyDeviceToHost, st));
        if (best_median) CK(cudaMemcpyAsync(best_median, oMed, (size_t)npoints * 4, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    return ORBX_OK;
}

extern "C" int orbm_keyframe_hint(orbm_t* m, uint64_t key) {
    if (!m) FAIL(ORBX_EINVAL, "null handle");
    m->kfHint = key;
    return ORBX_OK;
}
extern "C" int64_t orbm_keyframe_cache_hits(orbm_t* m) { return m ? m->kfHits : 0; }

static orbm_handle::KfEntry* kf_lookup(orbm_handle* m, uint64_t key, int n, bool hasUR, const orbm_proj_params* p) {
    for (auto& e : m->kfc)
        if (e.key == key && e.n == n && e.hasUR == hasUR && e.geo[0] == p->minX && e.geo[1] == p->minY &&
            e.geo[2] == p->grid_w_inv && e.geo[3] == p->grid_h_inv) {
            e.stamp = ++m->kfClock;
            return &e;
        }
    return nullptr;
}
// (re)allocates the least recently used entry for a keyframe of n keypoints; device pointers are carved out of one block
static int kf_claim(orbm_handle* m, uint64_t key, int n, bool hasUR, const orbm_proj_params* p, orbm_handle::KfEntry** out) {
    orbm_handle::KfEntry* e = &m->kfc[0];
    for (auto& c : m->kfc)
        if (c.stamp < e->stamp) e = &c;
    const size_t a = 256;
    auto up = [&](size_t x) { return (x + a - 1) / a * a; };
    const size_t need = up((size_t)n * sizeof(orbx_kp)) + up((size_t)n * 4) + up((size_t)n * 32) + up((SP_CELLS + 1) * 4) +
                        up((size_t)std::max(n, 1) * 4);
    if (need > e->cap) {
        cudaFree(e->mem);
        e->mem = nullptr; e->cap = 0; e->key = 0;
        CK(cudaMalloc(&e->mem, need));
        e->cap = need;
    }
    uint8_t* q = e->mem;
    e->keys = (orbx_kp*)q; q += up((size_t)n * sizeof(orbx_kp));
    e->uRight = (float*)q; q += up((size_t)n * 4);
    e->desc = (uint4*)q; q += up((size_t)n * 32);
    e->cellStart = (int*)q; q += up((SP_CELLS + 1) * 4);
    e->items = (int*)q;
    e->key = key; e->n = n; e->hasUR = hasUR; e->stamp = ++m->kfClock;
    e->geo[0] = p->minX; e->geo[1] = p->minY; e->geo[2] = p->grid_w_inv; e->geo[3] = p->grid_h_inv;
    *out = e;
    return ORBX_OK;
}

extern "C" int orbm_fuse_search(orbm_t* m, const orbm_proj_params* p, const float* inv_level_sigma2, int check_chi2,
                                const orbm_fuse_points* pts, const orbm_cur_frame* kf, int32_t* best_idx,
                                uint16_t* best_dist) {
    if (!m || !p || !pts || !kf || !inv_level_sigma2) FAIL(ORBX_EINVAL, "null argument");
    const int n = pts->n, nC = kf->n;
    if (n < 0 || nC < 0) FAIL(ORBX_EINVAL, "negative count");
    if (n == 0) return ORBX_OK;
    if (!pts->proj_x || !pts->proj_y || !pts->proj_xr || !pts->level || !pts->valid || !pts->desc || !best_idx || !best_dist ||
        (nC && (!kf->keys_un || !kf->desc)))
        FAIL(ORBX_EINVAL, "null array");
    if (p->nlevels < 1 || p->nlevels > ORBX_MAX_LEVELS) FAIL(ORBX_EINVAL, "nlevels out of range");
    CK(cudaSetDevice(m->device));
    cudaStream_t st = m->stream;
    const bool hin = m->inSpace == ORBX_MEM_HOST, hout = m->outSpace == ORBX_MEM_HOST;
    const size_t need = (size_t)n * (4 * 4 + 1 + 32 + 4 + 2) + (size_t)nC * (28 + 4 + 32 + 4) + (size_t)SP_CELLS * 12 + 256 * 24;
    int rc;
    if ((rc = ensure(m->dWs, m->wsCap, need))) return rc;
    if (hin && (rc = ensure_host_in(m, need))) return rc;
    WsCursor W{m->dWs, 0};
    Stager stg{m, &W, hin, (size_t)-1};
    SpDev S;
    memset(&S, 0, sizeof(S));
    S.p = *p;
    S.nCur = nC;
    FuseDev F;
    F.n = n;
    F.u = (const float*)stg.in(pts->proj_x, (size_t)n * 4);
    F.v = (const float*)stg.in(pts->proj_y, (size_t)n * 4);
    F.ur = (const float*)stg.in(pts->proj_xr, (size_t)n * 4);
    F.level = (const int32_t*)stg.in(pts->level, (size_t)n * 4);
    F.valid = (const uint8_t*)stg.in(pts->valid, n);
    F.desc = (const uint4*)stg.in(pts->desc, (size_t)n * 32);
    // keyframe cache: a hinted keyframe (orbm_keyframe_hint) that is resident skips its upload and the grid build
    const uint64_t hint = m->kfHint;
    m->kfHint = 0;
    orbm_handle::KfEntry* hit = (hint && nC) ? kf_lookup(m, hint, nC, kf->u_right != nullptr, p) : nullptr;
    if (hit) {
        m->kfHits++;
        S.keys = hit->keys; S.uRight = hit->hasUR ? hit->uRight : nullptr; S.cdesc = hit->desc;
    } else {
        S.keys = (const orbx_kp*)stg.in(kf->keys_un, (size_t)nC * sizeof(orbx_kp));
        S.uRight = (const float*)stg.in(kf->u_right, (size_t)nC * 4);
        S.cdesc = (const uint4*)stg.in(kf->desc, (size_t)nC * 32);
    }
    for (int i = 0; i < ORBX_MAX_LEVELS; i++) F.invSigma2[i] = i < p->nlevels ? inv_level_sigma2[i] : 0.f;
    F.chi2 = check_chi2;
    if ((((uintptr_t)F.desc) | ((uintptr_t)S.cdesc)) & 15) FAIL(ORBX_EINVAL, "device descriptor matrices must be 16-byte aligned");
    if ((rc = stg.flush(st))) return rc;
    int* cellCount = W.take<int>(SP_CELLS);
    int* cellStart = W.take<int>(SP_CELLS + 1);
    int* cursor = W.take<int>(SP_CELLS);
    int* items = W.take<int>(std::max(nC, 1));
    int32_t* oIdx = hout ? W.take<int32_t>(n) : best_idx;
    unsigned short* oDist = hout ? W.take<unsigned short>(n) : best_dist;
    if (W.off > m->wsCap) FAIL(ORBX_ECUDA, "internal: workspace estimate too small (%zu > %zu)", W.off, m->wsCap);
    if (hit) {
        cellStart = hit->cellStart;
        items = hit->items;
    } else {
        if (nC <= m->gridFusedMax) {
            k_sp_grid_build<<<1, 1024, 0, st>>>(S, cellStart, items);
            m->launches += 1;
        } else {
            CK(cudaMemsetAsync(cellCount, 0, SP_CELLS * sizeof(int), st));
            if (nC) k_sp_grid_count<<<(nC + 255) / 256, 256, 0, st>>>(S, cellCount);
            k_sp_grid_scan<<<1, 1024, 0, st>>>(cellCount, cellStart, cursor);
            if (nC) k_sp_grid_fill<<<(nC + 255) / 256, 256, 0, st>>>(S, cursor, items);
            k_sp_grid_sort<<<(SP_CELLS + 255) / 256, 256, 0, st>>>(cellStart, items);
            m->launches += 4;
        }
        if (hint && nC) {  // remember this keyframe (device-to-device copies of what was just staged and built)
            orbm_handle::KfEntry* e = nullptr;
            if ((rc = kf_claim(m, hint, nC, kf->u_right != nullptr, p, &e))) return rc;
            CK(cudaMemcpyAsync(e->keys, S.keys, (size_t)nC * sizeof(orbx_kp), cudaMemcpyDeviceToDevice, st));
            if (S.uRight) CK(cudaMemcpyAsync(e->uRight, S.uRight, (size_t)nC * 4, cudaMemcpyDeviceToDevice, st));
            CK(cudaMemcpyAsync(e->desc, S.cdesc, (size_t)nC * 32, cudaMemcpyDeviceToDevice, st));
            CK(cudaMemcpyAsync(e->cellStart, cellStart, (SP_CELLS + 1) * sizeof(int), cudaMemcpyDeviceToDevice, st));
            CK(cudaMemcpyAsync(e->items, items, (size_t)nC * sizeof(int), cudaMemcpyDeviceToDevice, st));
        }
    }
    k_fuse_search<<<(n + 7) / 8, 256, 0, st>>>(S, F, cellStart, items, oIdx, oDist);
    m->launches += 1;
    CK(cudaGetLastError());
    if (hout) {
        CK(cudaMemcpyAsync(best_idx, oIdx, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(best_dist, oDist, (size_t)n * 2, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    return ORBX_OK;
}

#include "orbx_bowsearch.cuh"

extern "C" int orbm_search_stats(orbm_t* m, int* rounds, int64_t* candidates) {
    if (!m) FAIL(ORBX_EINVAL, "null handle");
    if (rounds) *rounds = m->lastRounds;
    if (candidates) *candidates = m->lastCandidates;
    return ORBX_OK;
}
