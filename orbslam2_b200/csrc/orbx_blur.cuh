// orbx_blur.cuh — K4 k_blur: cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on 8-bit levels, bit exact.
// Reference call site: /root/reference/src/ORBextractor.cc:898-900. OpenCV's 8U path is fixed point: separable taps
// {18,34,48,56,48,34,18}/256 per axis, exact accumulation, ONE rounding at the end: (sum + 32768) >> 16.
//
// One warp = 128 px x BLUR_ROWS output rows of one level of one frame; a thread owns 4 adjacent pixels (one aligned
// 32-bit word) and slides down the rows, so every input row is loaded and filtered horizontally once per strip:
//   horizontal: byte windows x-3..x and x+1..x+4 cut by PRMT, two DP4A per pixel (a row sum is <= 255*256 = 65280);
//   vertical:   row sums of adjacent rows packed in 16-bit pairs, three DP2A + one IMAD per pixel over a 6-pair register
//               window (loop unrolled by the window period 6, so the window is renamed, not moved);
//   pipes (B200, scripts/pipe_probe.cu): DP4A/DP2A/IMAD issue on the FMA pipe, PRMT on the ALU pipe, and they overlap;
//   reflect-101: row index reflection for y; left edge by byte permutation, right edge by an 8-byte gather (last lanes only).
#pragma once
#include "orbx_internal.cuh"

#define BLUR_ROWS 48  // + 6 halo rows = 54 = 9 x 6: the 6-pair window unrolls exactly (small code, no i-cache misses)
#define BLUR_WARPS 4

__device__ __forceinline__ unsigned blur_dp2a(unsigned a16x2, unsigned b8x4, unsigned c) {
    unsigned d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a16x2), "r"(b8x4), "r"(c));
    return d;
}

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

    // Horizontal pass: one DP4A pair per pixel on byte windows cut out of (wm, w0, wp) by PRMT.
    // Vertical pass: the 16-bit row sums of vertically adjacent rows are packed in pairs P_j = (H_j, H_j+1), so an
    // output is three DP2A and one IMAD: 18*H_n-6 + 34*H_n-5 | 48*H_n-4 + 56*H_n-3 | 48*H_n-2 + 34*H_n-1 | 18*H_n.
    const unsigned K1 = 18u | (34u << 8) | (48u << 16) | (56u << 24);  // taps of x-3 .. x
    const unsigned K2 = 48u | (34u << 8) | (18u << 16);                // taps of x+1 .. x+3 (x+4 weighs 0)
    const unsigned V01 = 18u | (34u << 8), V23 = 48u | (56u << 8), V45 = 48u | (34u << 8);
    unsigned PW[6][4];  // PW[j % 6][p] = H_j | H_j+1 << 16 for pixel p; the loop is unrolled by the window period 6
    unsigned hp[4];     // H of the previous input row
    static_assert((BLUR_ROWS + 6) % 6 == 0, "row loop is unrolled by the window period");
    for (int ib = 0; ib < BLUR_ROWS + 6; ib += 6) {
    if (y0 + ib - 6 >= h) break;  // every output of this and the following groups lies below the image
    // all loads of the 6-row group are issued before any is used: the kernel is bound by memory latency, not by
    // instruction issue, so what matters is how many rows each warp has in flight
    unsigned Wm[6], W0[6], Wp[6];
#pragma unroll
    for (int i = 0; i < 6; i++) {
        const int yin = y0 + ib + i - 3;
        const uint8_t* rp = src + (long long)blur_reflect(min(yin, h + 2), h) * G.pitch + x0;
        if (redge) {
            W0[i] = rp[ridx[0]] | (rp[ridx[1]] << 8) | (rp[ridx[2]] << 16) | (rp[ridx[3]] << 24);
            Wp[i] = rp[ridx[4]] | (rp[ridx[5]] << 8) | (rp[ridx[6]] << 16) | (rp[ridx[7]] << 24);
        } else {
            W0[i] = *(const unsigned*)rp;
            Wp[i] = *(const unsigned*)(rp + 4);
        }
        // (no use of loaded data in this loop — a predicated-off instruction still waits for its source registers, so
        // the left-edge reflection is applied in the compute loop; here the first column just re-reads its own word)
        Wm[i] = *(const unsigned*)(rp - (x0 > 0 ? 4 : 0));
    }
#pragma unroll
    for (int i = 0; i < 6; i++) {   // input row n = ib + i of the strip (absolute row y0 + n - 3)
        unsigned H[4];
        {
            const unsigned w0 = W0[i], wp = Wp[i];
            const unsigned wm = (x0 > 0) ? Wm[i] : __byte_perm(w0, wp, 0x1234);  // x = -4..-1 -> pixels 4, 3, 2, 1
            H[0] = __dp4a(__byte_perm(w0, wp, 0x4321), K2, __dp4a(__byte_perm(wm, w0, 0x4321), K1, 0u));
            H[1] = __dp4a(__byte_perm(w0, wp, 0x5432), K2, __dp4a(__byte_perm(wm, w0, 0x5432), K1, 0u));
            H[2] = __dp4a(__byte_perm(w0, wp, 0x6543), K2, __dp4a(__byte_perm(wm, w0, 0x6543), K1, 0u));
            H[3] = __dp4a(wp, K2, __dp4a(w0, K1, 0u));
        }
        if (ib + i >= 1) {
#pragma unroll
            for (int p = 0; p < 4; p++) PW[(i + 5) % 6][p] = __byte_perm(hp[p], H[p], 0x5410);  // P_n-1
        }
        if (ib + i >= 6) {
            const int yout = y0 + ib + i - 6;
            if (yout < h) {
                unsigned sum[4];
#pragma unroll
                for (int p = 0; p < 4; p++) {
                    unsigned acc = 18u * H[p] + 32768u;
                    acc = blur_dp2a(PW[i % 6][p], V01, acc);        // P_n-6
                    acc = blur_dp2a(PW[(i + 2) % 6][p], V23, acc);  // P_n-4
                    acc = blur_dp2a(PW[(i + 4) % 6][p], V45, acc);  // P_n-2
                    sum[p] = acc;
                }
                // result byte p = bits 16..23 of sum[p]
                const unsigned out = __byte_perm(__byte_perm(sum[0], sum[1], 0x0062), __byte_perm(sum[2], sum[3], 0x0062), 0x5410);
                if (x0 + 4 <= w) {
                    *(unsigned*)(dst + (long long)yout * G.pitch + x0) = out;
                } else {
                    for (int p = 0; p < w - x0; p++) dst[(long long)yout * G.pitch + x0 + p] = (uint8_t)(out >> (8 * p));
                }
            }
        }
#pragma unroll
        for (int p = 0; p < 4; p++) hp[p] = H[p];
    }
    }
}

