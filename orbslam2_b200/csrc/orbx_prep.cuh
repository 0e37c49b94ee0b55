// orbx_prep.cuh — upstream image prep fused into the extractor's ingest.
//
// P1 k_cvt_gray   cv::cvtColor(.., COLOR_{RGB,BGR,RGBA,BGRA}2GRAY) on 8U (ref: src/Tracking.cc:269-294, 308-321, 337-350):
//                 (R*9798 + G*19235 + B*3735 + 2^14) >> 15, written straight into pyramid level 0
// P2 k_remap      cv::remap(.., CV_32F maps, INTER_LINEAR, BORDER_CONSTANT 0) on 8U (ref: Test/Replay/Stereo/
//                 stereo_euroc.cc:136-137): 1/32-px fixed point, 15-bit weights, written into pyramid level 0
// P3 k_undistort  cv::undistortPoints(.., K, D, Mat(), K) for the keypoints of the last batch (ref: Frame::
//                 UndistortKeyPoints, src/Frame.cc:402-432): double arithmetic, 5 iterations, no FMA contraction
#pragma once

// 4 pixels per thread; src rows are `ch` interleaved bytes per pixel. wide = 1: rows, frames and the base pointer are
// 4-byte (3 channels) / 16-byte (4 channels) aligned, so a thread's 12 / 16 source bytes are three / one vector loads.
__device__ __forceinline__ unsigned gray15(unsigned r, unsigned g, unsigned b) {
    return (r * 9798u + g * 19235u + b * 3735u + (1u << 14)) >> 15;
}

__global__ void __launch_bounds__(256) k_cvt_gray(const uint8_t* __restrict__ src, size_t sStride, size_t sImg, int ch,
                                                  int blueFirst, int wide, int w, int hgt, uint8_t* __restrict__ dst,
                                                  int dPitch, long long dFrame) {
    const int x0 = (blockIdx.x * 64 + threadIdx.x) * 4;
    const int y = blockIdx.y * 4 + threadIdx.y, f = blockIdx.z;
    if (x0 >= w || y >= hgt) return;
    const uint8_t* S = src + (size_t)f * sImg + (size_t)y * sStride + (size_t)x0 * ch;
    uint8_t* D = dst + (long long)f * dFrame + (long long)y * dPitch + x0;
    const int n = min(4, w - x0);
    unsigned packed = 0;
    if (wide && n == 4) {
        unsigned c0[4], c1[4], c2[4];  // channel 0, 1, 2 of the four pixels
        if (ch == 3) {
            const unsigned a = ((const unsigned*)S)[0], b = ((const unsigned*)S)[1], c = ((const unsigned*)S)[2];
            c0[0] = a & 255; c1[0] = (a >> 8) & 255; c2[0] = (a >> 16) & 255;
            c0[1] = a >> 24; c1[1] = b & 255; c2[1] = (b >> 8) & 255;
            c0[2] = (b >> 16) & 255; c1[2] = b >> 24; c2[2] = c & 255;
            c0[3] = (c >> 8) & 255; c1[3] = (c >> 16) & 255; c2[3] = c >> 24;
        } else {
            const uint4 v = *(const uint4*)S;
            const unsigned q[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; i++) { c0[i] = q[i] & 255; c1[i] = (q[i] >> 8) & 255; c2[i] = (q[i] >> 16) & 255; }
        }
#pragma unroll
        for (int i = 0; i < 4; i++) packed |= gray15(blueFirst ? c2[i] : c0[i], c1[i], blueFirst ? c0[i] : c2[i]) << (8 * i);
    } else {
        const int ri = blueFirst ? 2 : 0, bi = blueFirst ? 0 : 2;
#pragma unroll
        for (int i = 0; i < 4; i++)
            if (i < n) packed |= gray15(S[i * ch + ri], S[i * ch + 1], S[i * ch + bi]) << (8 * i);
    }
    if (n == 4 && ((dPitch | (int)(dFrame & 3)) & 3) == 0 && ((size_t)dst & 3) == 0) {
        *(unsigned*)D = packed;
    } else {
        for (int i = 0; i < n; i++) D[i] = (uint8_t)(packed >> (8 * i));
    }
}

// One thread owns one output position for REMAP_FRAMES frames of the batch: the map lookup, the fixed-point split, the
// weights and the four tap offsets are computed once and reused for every frame (they depend on the maps only).
#define REMAP_FRAMES 8
__global__ void __launch_bounds__(256) k_remap(const uint8_t* __restrict__ src, size_t sStride, size_t sImg, int sw, int sh,
                                               const float* __restrict__ mapx, const float* __restrict__ mapy, int dw,
                                               int dh, int nb, uint8_t* __restrict__ dst, int dPitch, long long dFrame) {
    const int x = blockIdx.x * 128 + threadIdx.x;
    const int y = blockIdx.y * 2 + threadIdx.y, fBase = blockIdx.z * REMAP_FRAMES;
    if (x >= dw || y >= dh) return;
    const size_t m = (size_t)y * dw + x;
    // cvRound(map * INTER_TAB_SIZE): float product, round half to even
    const int sx = __float2int_rn(mapx[m] * 32.f), sy = __float2int_rn(mapy[m] * 32.f);
    const int ix = min(32767, max(-32768, sx >> 5)), iy = min(32767, max(-32768, sy >> 5));
    const int fx = sx & 31, fy = sy & 31;
    int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32, w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
    if (w00 == 32768) { w00 = 32767; w11 = 1; }  // the saturated short and initInterTab2D's sum fix-up
    const bool x0 = ix >= 0 && ix < sw, x1 = ix + 1 >= 0 && ix + 1 < sw;
    const bool y0 = iy >= 0 && iy < sh, y1 = iy + 1 >= 0 && iy + 1 < sh;
    // taps outside the source count as 0 (BORDER_CONSTANT): zero their weight and clamp their address
    if (!(x0 && y0)) w00 = 0;
    if (!(x1 && y0)) w01 = 0;
    if (!(x0 && y1)) w10 = 0;
    if (!(x1 && y1)) w11 = 0;
    const int cx0 = min(max(ix, 0), sw - 1), cx1 = min(max(ix + 1, 0), sw - 1);
    const int cy0 = min(max(iy, 0), sh - 1), cy1 = min(max(iy + 1, 0), sh - 1);
    const size_t o00 = (size_t)cy0 * sStride + cx0, o01 = (size_t)cy0 * sStride + cx1;
    const size_t o10 = (size_t)cy1 * sStride + cx0, o11 = (size_t)cy1 * sStride + cx1;
    const uint8_t* S = src + (size_t)fBase * sImg;
    uint8_t* D = dst + (long long)fBase * dFrame + (long long)y * dPitch + x;
    const int nf = min(REMAP_FRAMES, nb - fBase);
    if (nf == REMAP_FRAMES) {
        int a[REMAP_FRAMES], b[REMAP_FRAMES], c[REMAP_FRAMES], d[REMAP_FRAMES];
#pragma unroll
        for (int f = 0; f < REMAP_FRAMES; f++) {
            const uint8_t* Sf = S + (size_t)f * sImg;
            a[f] = Sf[o00]; b[f] = Sf[o01]; c[f] = Sf[o10]; d[f] = Sf[o11];
        }
#pragma unroll
        for (int f = 0; f < REMAP_FRAMES; f++)
            D[(long long)f * dFrame] = (uint8_t)((a[f] * w00 + b[f] * w01 + c[f] * w10 + d[f] * w11 + (1 << 14)) >> 15);
    } else {
        for (int f = 0; f < nf; f++) {
            const uint8_t* Sf = S + (size_t)f * sImg;
            D[(long long)f * dFrame] = (uint8_t)((Sf[o00] * w00 + Sf[o01] * w01 + Sf[o10] * w10 + Sf[o11] * w11 + (1 << 14)) >> 15);
        }
    }
}

struct UndistortParams {
    double fx, fy, cx, cy, k[8];
    int identity;  // mDistCoef.at<float>(0) == 0: mvKeysUn = mvKeys (src/Frame.cc:404-408)
};

__global__ void __launch_bounds__(128) k_undistort(const orbx_kp* __restrict__ kps, const int* __restrict__ nOut, int stride,
                                                   UndistortParams P, orbx_kp* __restrict__ out) {
    const int f = blockIdx.y, i = blockIdx.x * 128 + threadIdx.x;
    if (i >= nOut[f]) return;
    orbx_kp kp = kps[(size_t)f * stride + i];
    if (!P.identity) {
        const double ifx = 1. / P.fx, ify = 1. / P.fy;
        double x = ((double)kp.x - P.cx) * ifx, y = ((double)kp.y - P.cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((P.k[7] * r2 + P.k[6]) * r2 + P.k[5]) * r2) / (1 + ((P.k[4] * r2 + P.k[1]) * r2 + P.k[0]) * r2);
            if (icdist < 0) { x = x0; y = y0; break; }
            const double dX = 2 * P.k[2] * x * y + P.k[3] * (r2 + 2 * x * x);
            const double dY = P.k[2] * (r2 + 2 * y * y) + 2 * P.k[3] * x * y;
            x = (x0 - dX) * icdist;
            y = (y0 - dY) * icdist;
        }
        const double xx = P.fx * x + 0. * y + P.cx, yy = 0. * x + P.fy * y + P.cy, ww = 1. / (0. * x + 0. * y + 1.);
        kp.x = (float)(xx * ww);
        kp.y = (float)(yy * ww);
    }
    out[(size_t)f * stride + i] = kp;
}

// Dense w x hgt frames (rows back to back, any alignment) -> pitched level-0 rows; one thread writes one 32-bit word.
__global__ void __launch_bounds__(256) k_repitch(const uint8_t* __restrict__ src, int w, int hgt, uint8_t* __restrict__ dst,
                                                 int pitch, long long frameStride) {
    const int x = (blockIdx.x * 64 + threadIdx.x) * 4, y = blockIdx.y * 4 + threadIdx.y;
    if (x >= w || y >= hgt) return;
    const uint8_t* s = src + ((size_t)blockIdx.z * hgt + y) * w + x;
    uint8_t* d = dst + (long long)blockIdx.z * frameStride + (long long)y * pitch + x;
    if (x + 4 <= w) {
        *(unsigned*)d = (unsigned)s[0] | ((unsigned)s[1] << 8) | ((unsigned)s[2] << 16) | ((unsigned)s[3] << 24);
    } else {
        for (int b = 0; b < w - x; b++) d[b] = s[b];
    }
}
