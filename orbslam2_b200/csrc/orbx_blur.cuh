// orbx_blur.cuh — K4 k_blur: cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on 8-bit levels, bit exact.
// Reference call site: /root/reference/src/ORBextractor.cc:898-900. OpenCV's 8U path is fixed point: separable taps
// {18,34,48,56,48,34,18}/256 per axis, exact accumulation, ONE rounding at the end: (sum + 32768) >> 16.
//
// One warp = 128 px x BLUR_ROWS output rows of one level of one frame; a thread owns 4 adjacent pixels (one aligned
// 32-bit word) and slides down the rows, so every input row is loaded and filtered horizontally once per strip:
//   horizontal: two pixels per register in 16-bit lanes (a row sum is <= 255*256 = 65280, it fits), taps by IMAD;
//   vertical:   32-bit per pixel over a 7-row register window (loop unrolled by the window period 7, so the window is renamed, not moved);
//   reflect-101: row index reflection for y; left edge by byte permutation, right edge by an 8-byte gather (last lanes only).
#pragma once
#include "orbx_internal.cuh"

#define BLUR_ROWS 36  // + 6 halo rows = 42 = 6 x 7: the 7-row window unrolls exactly (small code, no i-cache misses)
#define BLUR_WARPS 4

__device__ __forceinline__ int blur_reflect(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
    return p;
}

__global__ void __launch_bounds__(32 * BLUR_WARPS) k_blur(const DevPlan* __restrict__ P, const uint8_t* __restrict__ pyr,
                                                          uint8_t* __restrict__ blur, int f0) {
    const int lane = threadIdx.x & 31;
    const int f = f0 + blockIdx.y;
    int unit = blockIdx.x * BLUR_WARPS + (threadIdx.x >> 5);
    if (unit >= P->totalBlurTiles) return;
    int level = 0;
    while (level + 1 < P->nlevels && unit >= P->lv[level + 1].blurTileBase) level++;
    const LevelGeo& G = P->lv[level];
    unit -= G.blurTileBase;
    const int ty = unit / G.blurTilesX, tx = unit - ty * G.blurTilesX;
    const int x0 = (tx * 32 + lane) * 4, y0 = ty * BLUR_ROWS;
    const int w = G.w, h = G.h;
    if (x0 >= w) return;
    const long long lvlOff = G.pyrOff + (long long)f * G.frameStride;
    const uint8_t* src = pyr + lvlOff;
    uint8_t* dst = blur + lvlOff;
    // x reflect-101: the left edge is a byte permutation of the first two words; at the right edge the 8 bytes at
    // x0..x0+7 are gathered through offsets computed once (only the last one or two lanes of a row take that path)
    const bool redge = x0 + 8 > w;
    int ridx[8];
#pragma unroll
    for (int j = 0; j < 8; j++) ridx[j] = redge ? blur_reflect(x0 + j, w) - x0 : j;

    unsigned hw[7][4];  // horizontal sums of the last 7 input rows, one 32-bit value per pixel
    static_assert((BLUR_ROWS + 6) % 7 == 0, "row loop is unrolled by the window period");
    for (int ib = 0; ib < BLUR_ROWS + 6; ib += 7)
#pragma unroll
    for (int i = 0; i < 7; i++) {   // i = row phase inside the 7-row window; absolute input row = ib + i
        const int yin = y0 + ib + i - 3;
        if (yin - 3 < h) {  // rows whose outputs all lie below the image are not needed
            const uint8_t* row = src + (long long)blur_reflect(yin, h) * G.pitch;
            const uint8_t* rp = row + x0;
            unsigned wm, w0, wp;
            if (redge) {
                w0 = rp[ridx[0]] | (rp[ridx[1]] << 8) | (rp[ridx[2]] << 16) | (rp[ridx[3]] << 24);
                wp = rp[ridx[4]] | (rp[ridx[5]] << 8) | (rp[ridx[6]] << 16) | (rp[ridx[7]] << 24);
            } else {
                w0 = *(const unsigned*)rp;
                wp = *(const unsigned*)(rp + 4);
            }
            wm = (x0 > 0) ? *(const unsigned*)(rp - 4) : __byte_perm(w0, wp, 0x1234);  // x = -4..-1 -> 4, 3, 2, 1
            // E[d+3] = pixels (x0+d, x0+d+2) in 16-bit lanes, d = -3..4
            unsigned E[8];
            E[0] = __byte_perm(wm, w0, 0x4321) & 0x00FF00FFu;
            E[1] = __byte_perm(wm, w0, 0x5432) & 0x00FF00FFu;
            E[2] = __byte_perm(wm, w0, 0x6543) & 0x00FF00FFu;
            E[3] = w0 & 0x00FF00FFu;
            E[4] = __byte_perm(w0, wp, 0x4321) & 0x00FF00FFu;
            E[5] = __byte_perm(w0, wp, 0x5432) & 0x00FF00FFu;
            E[6] = __byte_perm(w0, wp, 0x6543) & 0x00FF00FFu;
            E[7] = wp & 0x00FF00FFu;
            const unsigned he = 18u * (E[0] + E[6]) + 34u * (E[1] + E[5]) + 48u * (E[2] + E[4]) + 56u * E[3];  // px 0,2
            const unsigned ho = 18u * (E[1] + E[7]) + 34u * (E[2] + E[6]) + 48u * (E[3] + E[5]) + 56u * E[4];  // px 1,3
            hw[i % 7][0] = he & 0xFFFFu;
            hw[i % 7][1] = ho & 0xFFFFu;
            hw[i % 7][2] = he >> 16;
            hw[i % 7][3] = ho >> 16;
        }
        {
            const int yout = y0 + ib + i - 6;
            if (ib + i >= 6 && yout < h) {
                // window rows: oldest = (i+1)%7 ... newest = i%7
                unsigned out = 0;
#pragma unroll
                for (int p = 0; p < 4; p++) {
                    const unsigned s = 18u * (hw[(i + 1) % 7][p] + hw[i % 7][p]) + 34u * (hw[(i + 2) % 7][p] + hw[(i + 6) % 7][p]) +
                                       48u * (hw[(i + 3) % 7][p] + hw[(i + 5) % 7][p]) + 56u * hw[(i + 4) % 7][p];
                    out |= ((s + 32768u) >> 16) << (8 * p);
                }
                if (x0 + 4 <= w) {
                    *(unsigned*)(dst + (long long)yout * G.pitch + x0) = out;
                } else {
                    for (int p = 0; p < w - x0; p++) dst[(long long)yout * G.pitch + x0 + p] = (uint8_t)(out >> (8 * p));
                }
            }
        }
    }
}
