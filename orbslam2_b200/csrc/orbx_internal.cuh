// orbx_internal.cuh — shared host/device declarations of liborbx (not part of the public ABI).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>

#include "../../include/orbx.h"

typedef unsigned long long u64;

void orbx_set_err(const char* fmt, ...);

#define CK(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) {                                                                        \
            orbx_set_err("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));          \
            return ORBX_ECUDA;                                                                          \
        }                                                                                               \
    } while (0)

#define FAIL(code, ...)            \
    do {                           \
        orbx_set_err(__VA_ARGS__); \
        return (code);             \
    } while (0)

// Geometry of one pyramid level and of everything derived from it. Lives in device memory (DevPlan).
struct LevelGeo {
    int w, h, pitch;        // interior size in pixels; row pitch in bytes (multiple of 64)
    long long frameStride;  // pitch * h: frames of one level are stored back to back (level-major layout)
    long long pyrOff;       // byte offset of this level's block inside the pyramid (and the blurred) allocation
    // FAST cell grid (src/ORBextractor.cc:763-781 of the reference)
    int nCols, nRows, wCell, hCell;
    int stCX, stCY;         // FAST supertile = stCX x stCY cells (one CTA)
    int nStX, nStY, stBase; // supertile grid of this level; first block index in the fused FAST launch
    int boxW, boxH;         // TMA box of a supertile: roundup16(stCX*wCell + 28) x (stCY*hCell + 6)
    int maxSurv;            // upper bound of NMS survivors in one cell
    int fastUnits;          // upper bound of 4-pixel words in the detection area of one supertile (k_fast2 list capacity)
    // quadtree
    int quota, nIni, treeW, treeH;
    float hX;
    int candCap;            // candidate capacity per frame
    long long candOff;      // record offset of this level's candidate block (level-major: + frame*candCap)
    int selCap, selOff;     // selected-keypoint capacity, offset inside a frame's selection row
    float scale;            // mvScaleFactor[level]
    float sizeField;        // (int)(patchSize*scale)
    int xtabOff, ytabOff;   // offsets (in int2) of the bilinear tables (x table padded to a multiple of 4 entries)
    int resizeFast;         // 1: k_resize_fast applies (4 output pixels read <= 8 source bytes)
    int blurTileBase, blurTilesX, blurTilesY;
};

struct DevPlan {
    int nlevels, batch;
    int iniTh, minTh, halfPatch, edge;
    int totalSt, totalBlurTiles;
    int blurRows;   // output rows per k_blur strip (a multiple of 6)
    int selStride;  // selected-keypoint slots per frame (sum of selCap)
    int umax[32];
    LevelGeo lv[ORBX_MAX_LEVELS];
};

// One record per FAST supertile (blockIdx.x of k_fast_t), built by the host.
struct __align__(16) FastSt {
    unsigned short level, stCX, xFirst, xEnd, yFirst, yEnd, cj0, ci0;  // detection area [xFirst, xEnd) x [yFirst, yEnd), first cell
    short boxW, boxH, wCell, hCell, nlevels, minTh, iniTh, nCells;
    unsigned divCw, divCh;                                    // ceil(65536 / wCell), ceil(65536 / hCell)
    int candCap, pad;
    long long candOff, pad2;
};

struct TmapPack {
    CUtensorMap m[ORBX_MAX_LEVELS];
};

// 64-bit candidate record: (score << 32) | ~orderkey, orderkey = cellRow<<22 | cellCol<<12 | yin<<6 | xin.
// max() over records == reference's "largest response, first in original order" (src/ORBextractor.cc:744-751).
__host__ __device__ inline u64 cand_pack(int score, int ci, int cj, int yin, int xin) {
    unsigned ok = ((unsigned)ci << 22) | ((unsigned)cj << 12) | ((unsigned)yin << 6) | (unsigned)xin;
    return ((u64)(unsigned)score << 32) | (u64)(0xFFFFFFFFu - ok);
}
__host__ __device__ inline void cand_unpack(u64 r, const LevelGeo& G, int& xrel, int& yrel, int& score) {
    unsigned ok = 0xFFFFFFFFu - (unsigned)(r & 0xFFFFFFFFull);
    score = (int)(r >> 32);
    int ci = ok >> 22, cj = (ok >> 12) & 1023, yin = (ok >> 6) & 63, xin = ok & 63;
    xrel = cj * G.wCell + xin + 3;  // relative to (minBorderX, minBorderY) like the reference's vToDistributeKeys
    yrel = ci * G.hCell + yin + 3;
}

// 256-bit Hamming distance on 8 x u32 (ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1643-1659)
__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}
