// orbx_blur.cuh — K4 k_blur: cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on 8-bit levels, bit exact.
// Reference call site: /root/reference/src/ORBextractor.cc:898-900. OpenCV's 8U path is fixed point: separable taps
// {18,34,48,56,48,34,18}/256 per axis, exact accumulation, ONE rounding at the end: (sum + 32768) >> 16.
//
// One half-warp = 64 px x BLUR_ROWS output rows of one level of one frame; a thread owns 4 adjacent pixels (one aligned
// 32-bit word) and slides down the rows, so every input row is loaded and filtered horizontally once per strip:
//   horizontal: the three aligned words around the thread's pixels against per-pixel tap vectors (the taps are shifted,
//               not the bytes): 2 + 3 + 3 + 2 DP4A per 4 pixels, no PRMT (a row sum is <= 255*256 = 65280);
//   vertical:   row sums of adjacent rows packed in 16-bit pairs, three DP2A + one IMAD per pixel over a 6-pair register
//               window (loop unrolled by the window period 6, so the window is renamed, not moved);
//   pipes (B200, scripts/pipe_probe.cu): DP4A/DP2A/IMAD issue on the FMA pipe, PRMT on the ALU pipe, and they overlap;
//   loads:      the 18 word loads of a 6-row group are issued together and nothing on the load path reads loaded data
//               (a predicated-off instruction still waits for its source registers — that serialised every warp);
//   reflect-101: row index reflection for y; the x reflection only exists in the EDGE instantiation, taken by the
//               first and last warp of a row (left edge: byte permutation; right edge: 8-byte gather of the last lanes).
#pragma once
#include "orbx_internal.cuh"

// output rows per strip (+ 6 halo rows; a multiple of the window period 6): 72 for batches (7.7 % halo instead of 11 %;
// measured 24 / 48 / 72 / 96 / 144 rows: 0.265 / 0.244 / 0.238 / 0.239 / 0.245 ms per 384 frames), 48 for plans of a few frames
// (more strips to spread over the SMs)
#define BLUR_ROWS 72
#define BLUR_ROWS_SMALL 48
#define BLUR_WARPS 4

__device__ __forceinline__ unsigned blur_dp2a(unsigned a16x2, unsigned b8x4, unsigned c) {
    unsigned d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a16x2), "r"(b8x4), "r"(c));
    return d;
}

// BORDER_REFLECT_101 of an index at most one period outside [0, len) (the blur reaches 3 past either end; len >= 30)
__device__ __forceinline__ int blur_reflect(int p, int len) {
    p = abs(p);
    return min(p, 2 * len - 2 - p);
}

struct BlurRows {
    unsigned m[6], z[6], q[6];  // words at x0-4, x0, x0+4 of six consecutive input rows
};

template <bool EDGE, int ROWS>
__device__ __forceinline__ void blur_strip(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int pitch, int w,
                                           int h, int x0, int y0) {
    const unsigned K1 = 18u | (34u << 8) | (48u << 16) | (56u << 24);  // taps of x-3 .. x
    const unsigned K2 = 48u | (34u << 8) | (18u << 16);                // taps of x+1 .. x+3 (x+4 weighs 0)
    const unsigned V01 = 18u | (34u << 8), V23 = 48u | (56u << 8), V45 = 48u | (34u << 8);
    // tap vectors of pixel x0+i against the aligned words (wm: x0-4..x0-1, w0: x0..x0+3, wp: x0+4..x0+7)
    const unsigned T0m = (18u << 8) | (34u << 16) | (48u << 24), T0z = 56u | (48u << 8) | (34u << 16) | (18u << 24);
    const unsigned T1m = (18u << 16) | (34u << 24), T1z = 48u | (56u << 8) | (48u << 16) | (34u << 24), T1q = 18u;
    const unsigned T2m = 18u << 24, T2z = 34u | (48u << 8) | (56u << 16) | (48u << 24), T2q = 34u | (18u << 8);
    const bool redge = EDGE && x0 + 8 > w;
    const bool ledge = EDGE && x0 == 0;
    // Right edge: bytes x0 .. x0+7 with BORDER_REFLECT_101 past w - 1. A reflected byte lies at most 4 left of x0, so every
    // byte comes from the 12-byte window (wm, w0, wp) the lane loads anyway: two PRMT per word with per-lane selectors
    // (identity for the lanes that are not on the edge) — no byte gathers, no divergence in the row loop.
    unsigned s1a = 0x7654u, s2a = 0x3210u, s1b = 0x7654u, s2b = 0x7654u;
    if (redge) {
        s1a = s2a = s1b = s2b = 0u;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            // index into (wm, w0, wp): 0 .. 11 for every byte a real pixel needs (x + 3 <= w + 2 reflects to >= w - 4 >= x0 - 3);
            // bytes that only feed the padding pixels past w - 1 may fall outside: clamped, their value does not matter
            const unsigned t = (unsigned)max(0, min(11, blur_reflect(x0 + j, w) - x0 + 4));
            const unsigned n1 = t < 8u ? t : 0u, n2 = t < 8u ? (unsigned)(j & 3) : t - 4u;
            if (j < 4) { s1a |= n1 << (4 * j); s2a |= n2 << (4 * j); }
            else { s1b |= n1 << (4 * (j - 4)); s2b |= n2 << (4 * (j - 4)); }
        }
    }
    const int lback = ledge ? 0 : 4;
    const uint8_t* col = src + x0;

    auto load_group = [&](int n0, BlurRows& R) {  // input rows n0 .. n0+5 of the strip (absolute row y0 + n - 3)
#pragma unroll
        for (int i = 0; i < 6; i++) {
            const int off = blur_reflect(min(y0 + n0 + i - 3, h + 2), h) * pitch;
            R.z[i] = *(const unsigned*)(col + off);
            R.q[i] = *(const unsigned*)(col + off + 4);  // may run into the row padding / next row: addressable
            R.m[i] = *(const unsigned*)(col + off - lback);
        }
    };
    // row sums H[0..3] of input row i of a loaded group
    auto hsum = [&](const BlurRows& R, int n, int i, unsigned (&H)[4]) {
        unsigned w0 = R.z[i], wp = R.q[i], wm = R.m[i];
        if (EDGE) {
            const unsigned n0 = __byte_perm(__byte_perm(wm, w0, s1a), wp, s2a);
            const unsigned n1 = __byte_perm(__byte_perm(wm, w0, s1b), wp, s2b);
            w0 = n0;
            wp = n1;
            if (ledge) wm = __byte_perm(w0, wp, 0x1234);  // x = -4..-1 -> pixels 4, 3, 2, 1
        }
        // the taps are shifted instead of the bytes: every DP4A reads an ALIGNED word against a constant tap vector
        // (10 DP4A on the FMA pipe and no PRMT, instead of 8 DP4A + 6 PRMT on the busier ALU pipe)
        H[0] = __dp4a(w0, T0z, __dp4a(wm, T0m, 0u));
        H[1] = __dp4a(wp, T1q, __dp4a(w0, T1z, __dp4a(wm, T1m, 0u)));
        H[2] = __dp4a(wp, T2q, __dp4a(w0, T2z, __dp4a(wm, T2m, 0u)));
        H[3] = __dp4a(wp, K2, __dp4a(w0, K1, 0u));
    };

    // Vertical pass state: PW[j % 6][p] = H_j | H_j+1 << 16 for pixel p, hp = H of the previous row.
    // Output row n-6 = 18*H_n-6 + 34*H_n-5 | 48*H_n-4 + 56*H_n-3 | 48*H_n-2 + 34*H_n-1 | 18*H_n: three DP2A + one IMAD.
    unsigned PW[6][4], hp[4];
    BlurRows R;
    load_group(0, R);
    {   // prologue: rows 0..5 only fill the window
        unsigned H[4];
        hsum(R, 0, 0, hp);
#pragma unroll
        for (int i = 1; i < 6; i++) {
            hsum(R, i, i, H);
#pragma unroll
            for (int p = 0; p < 4; p++) {
                PW[i - 1][p] = __byte_perm(hp[p], H[p], 0x5410);
                hp[p] = H[p];
            }
        }
    }
    static_assert(ROWS % 6 == 0, "row loop is unrolled by the window period");
    for (int n0 = 6; n0 < ROWS + 6; n0 += 6) {
        if (y0 + n0 - 6 >= h) break;  // every remaining output lies below the image
        load_group(n0, R);
        uint8_t* orow = dst + (y0 + n0 - 6) * pitch + x0;
#pragma unroll
        for (int i = 0; i < 6; i++) {  // input row n = n0 + i, output row n - 6
            unsigned H[4];
            hsum(R, n0 + i, i, H);
            unsigned sum[4];
#pragma unroll
            for (int p = 0; p < 4; p++) {
                PW[(i + 5) % 6][p] = __byte_perm(hp[p], H[p], 0x5410);  // P_n-1
                hp[p] = H[p];
                unsigned acc = 18u * H[p] + 32768u;
                acc = blur_dp2a(PW[i][p], V01, acc);            // P_n-6
                acc = blur_dp2a(PW[(i + 2) % 6][p], V23, acc);  // P_n-4
                acc = blur_dp2a(PW[(i + 4) % 6][p], V45, acc);  // P_n-2
                sum[p] = acc;
            }
            // result byte p = bits 16..23 of sum[p]
            const unsigned out = __byte_perm(__byte_perm(sum[0], sum[1], 0x0062), __byte_perm(sum[2], sum[3], 0x0062), 0x5410);
            if (y0 + n0 + i - 6 < h) {
                // (the last word of a row may reach into the row padding, pitch = roundup64(w): nothing reads it)
                *(unsigned*)(orow + i * pitch) = out;
            }
        }
    }
}

template <int ROWS>
__global__ void __launch_bounds__(32 * BLUR_WARPS, 6) k_blur(const DevPlan* __restrict__ P, const uint8_t* __restrict__ pyr,
                                                          uint8_t* __restrict__ blur, int f0, int tile0, int tileEnd) {
    const int lane = threadIdx.x & 31;
    const int f = f0 + blockIdx.y;
    int unit = tile0 + blockIdx.x * BLUR_WARPS + (threadIdx.x >> 5);  // units [tile0, tileEnd): all levels, or one level's
    if (unit >= tileEnd) return;
    int level = 0;
    while (level + 1 < P->nlevels && unit >= P->lv[level + 1].blurTileBase) level++;
    const LevelGeo& G = P->lv[level];
    unit -= G.blurTileBase;
    // A unit (one warp) is TWO strips of 64 px x BLUR_ROWS rows, consecutive in the level's row-major strip order: lanes 0-15
    // take strip 2u, lanes 16-31 strip 2u + 1. (With one 128-px strip per warp 17 % of the lanes of a 640x480 pyramid lay
    // past the right edge: level widths are 533, 444, 370, ...)
    const int nX = G.blurTilesX, nS = nX * G.blurTilesY;
    const int w = G.w, h = G.h;
    const int sA = 2 * unit, sB = sA + 1;
    const int tyA = sA / nX, txA = sA - tyA * nX;
    const int tyB = sB / nX, txB = sB - tyB * nX;
    // warp-uniform: does either strip hold the first column or a lane whose 8-byte window crosses the right border?
    const bool edge = txA == 0 || txA * 64 + 68 > w || (sB < nS && (txB == 0 || txB * 64 + 68 > w));
    const bool hi = lane >= 16;
    if (hi && sB >= nS) return;
    const int x0 = ((hi ? txB : txA) * 16 + (lane & 15)) * 4, y0 = (hi ? tyB : tyA) * ROWS;
    if (x0 >= w) return;
    const long long lvlOff = G.pyrOff + (long long)f * G.frameStride;
    if (edge) blur_strip<true, ROWS>(pyr + lvlOff, blur + lvlOff, G.pitch, w, h, x0, y0);
    else blur_strip<false, ROWS>(pyr + lvlOff, blur + lvlOff, G.pitch, w, h, x0, y0);
}
