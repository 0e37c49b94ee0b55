// orbx_prep.cuh — upstream image prep fused into the extractor's ingest.
//
// P1 k_cvt_gray   cv::cvtColor(.., COLOR_{RGB,BGR,RGBA,BGRA}2GRAY) on 8U (ref: src/Tracking.cc:269-294, 308-321, 337-350):
//                 (R*9798 + G*19235 + B*3735 + 2^14) >> 15, written straight into pyramid level 0
// P2 k_remap      cv::remap(.., CV_32F maps, INTER_LINEAR, BORDER_CONSTANT 0) on 8U (ref: Test/Replay/Stereo/
//                 stereo_euroc.cc:136-137): 1/32-px fixed point, 15-bit weights, written into pyramid level 0
// P3 k_undistort  cv::undistortPoints(.., K, D, Mat(), K) for the keypoints of the last batch (ref: Frame::
//                 UndistortKeyPoints, src/Frame.cc:402-432): double arithmetic, 5 iterations, no FMA contraction
#pragma once

// 4 pixels per thread; src rows are `ch` interleaved bytes per pixel
__global__ void __launch_bounds__(64) k_cvt_gray(const uint8_t* __restrict__ src, size_t sStride, size_t sImg, int ch,
                                                 int blueFirst, int w, uint8_t* __restrict__ dst, int dPitch,
                                                 long long dFrame) {
    const int x0 = (blockIdx.x * 64 + threadIdx.x) * 4;
    if (x0 >= w) return;
    const int y = blockIdx.y, f = blockIdx.z;
    const uint8_t* S = src + (size_t)f * sImg + (size_t)y * sStride + (size_t)x0 * ch;
    uint8_t* D = dst + (long long)f * dFrame + (long long)y * dPitch + x0;
    const int ri = blueFirst ? 2 : 0, bi = blueFirst ? 0 : 2;
    const int n = min(4, w - x0);
    unsigned packed = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i < n) {
            const unsigned r = S[i * ch + ri], g = S[i * ch + 1], b = S[i * ch + bi];
            packed |= ((r * 9798u + g * 19235u + b * 3735u + (1u << 14)) >> 15) << (8 * i);
        }
    }
    if (n == 4 && ((dPitch | (int)(dFrame & 3)) & 3) == 0 && ((size_t)dst & 3) == 0) {
        *(unsigned*)D = packed;
    } else {
        for (int i = 0; i < n; i++) D[i] = (uint8_t)(packed >> (8 * i));
    }
}

__global__ void __launch_bounds__(128) k_remap(const uint8_t* __restrict__ src, size_t sStride, size_t sImg, int sw, int sh,
                                               const float* __restrict__ mapx, const float* __restrict__ mapy, int dw,
                                               uint8_t* __restrict__ dst, int dPitch, long long dFrame) {
    const int x = blockIdx.x * 128 + threadIdx.x;
    if (x >= dw) return;
    const int y = blockIdx.y, f = blockIdx.z;
    const size_t m = (size_t)y * dw + x;
    // cvRound(map * INTER_TAB_SIZE): float product, round half to even
    const int sx = __float2int_rn(mapx[m] * 32.f), sy = __float2int_rn(mapy[m] * 32.f);
    const int ix = min(32767, max(-32768, sx >> 5)), iy = min(32767, max(-32768, sy >> 5));
    const int fx = sx & 31, fy = sy & 31;
    int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32, w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
    if (w00 == 32768) { w00 = 32767; w11 = 1; }  // the saturated short and initInterTab2D's sum fix-up
    const uint8_t* S = src + (size_t)f * sImg;
    const bool x0 = ix >= 0 && ix < sw, x1 = ix + 1 >= 0 && ix + 1 < sw;
    const bool y0 = iy >= 0 && iy < sh, y1 = iy + 1 >= 0 && iy + 1 < sh;
    const int a = (x0 && y0) ? S[(size_t)iy * sStride + ix] : 0;
    const int b = (x1 && y0) ? S[(size_t)iy * sStride + ix + 1] : 0;
    const int c = (x0 && y1) ? S[(size_t)(iy + 1) * sStride + ix] : 0;
    const int d = (x1 && y1) ? S[(size_t)(iy + 1) * sStride + ix + 1] : 0;
    dst[(long long)f * dFrame + (long long)y * dPitch + x] = (uint8_t)((a * w00 + b * w01 + c * w10 + d * w11 + (1 << 14)) >> 15);
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
