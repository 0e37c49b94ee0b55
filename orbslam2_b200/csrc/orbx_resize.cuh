// orbx_resize.cuh — K1 k_resize_fast: one pyramid level from the previous one, cv::resize INTER_LINEAR on 8-bit data.
// Reference: ORBextractor::ComputePyramid, /root/reference/src/ORBextractor.cc:921-946 (resize at :934).
// OpenCV's 8U bilinear is fixed point (11-bit coefficients):
//     H(row, d) = S[row][s_d] * a0_d + S[row][s_d + 1] * a1_d
//     dst = (((b0 * (H(y0, d) >> 4)) >> 16) + ((b1 * (H(y1, d) >> 4)) >> 16) + 2) >> 2
// with {s_d, a0_d, a1_d} / {y0, b0, b1} from tables built on the host with OpenCV's own float expressions.
//
// One thread = 4 adjacent output pixels of RESIZE_ROWS consecutive rows (the x table entries and their PRMT selectors
// are shared by the rows, and all source words are in flight together). For down-scaling factors <= 2 their source pixels span at most 8
// bytes, so each of the two source rows is read as 3 aligned words, funnel-shifted to the first source pixel, and a
// pixel's (S[s], S[s+1]) pair is one PRMT; the horizontal blend is one DP2A against the table's packed (a0, a1).
// Down-scaling factors > 2 and OpenCV's exact-2x INTER_AREA shortcut use the generic k_resize.
#pragma once
#include "orbx_internal.cuh"

__device__ __forceinline__ unsigned dp2a_lo_u32(unsigned a16x2, unsigned b8x4, unsigned c) {
    unsigned d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a16x2), "r"(b8x4), "r"(c));
    return d;
}

#define RESIZE_ROWS 4  // output rows per thread

__global__ void __launch_bounds__(256) k_resize_fast(const __grid_constant__ DevPlan Pc, int level, uint8_t* __restrict__ pyr,
                                                     const int2* __restrict__ tabs, int f0) {
    const DevPlan* P = &Pc;  // the plan rides in the kernel parameters (constant bank): no global loads in the prologue
    const LevelGeo& D = P->lv[level];
    const LevelGeo& S = P->lv[level - 1];
    const int x4 = (blockIdx.x * 64 + threadIdx.x) * 4;
    const int yb = (blockIdx.y * 4 + threadIdx.y) * RESIZE_ROWS;
    const int f = f0 + blockIdx.z;
    if (x4 >= D.w || yb >= D.h) return;
    const uint8_t* src = pyr + S.pyrOff + (long long)f * S.frameStride;
    const int spitch = S.pitch, dpitch = D.pitch, sh1 = S.h - 1;
    // x table: 4 entries {s, a0 | a1 << 16} (the table is padded to a multiple of 4 entries); shared by the thread's rows
    const uint4* xt = (const uint4*)(tabs + D.xtabOff + x4);
    const uint4 t01 = xt[0], t23 = xt[1];
    int2 ty[RESIZE_ROWS];
#pragma unroll
    for (int r = 0; r < RESIZE_ROWS; r++) ty[r] = tabs[D.ytabOff + min(yb + r, D.h - 1)];
    const int sx0 = (int)t01.x;
    const int a0 = sx0 & ~3, sh = (sx0 & 3) * 8;
    const unsigned sx[4] = {t01.x, t01.z, t23.x, t23.z};
    const unsigned cf[4] = {t01.y, t01.w, t23.y, t23.w};
    unsigned sel[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const unsigned rel = min(sx[i] - (unsigned)sx0, 6u);  // <= 6 by construction; the clamp only guards the padding entries
        sel[i] = rel * 0x11u + 0x10u;                          // PRMT picks bytes (rel, rel+1) into byte 0, 1
    }
    // all source words of the thread's rows are requested before any is used (the kernel is latency bound)
    unsigned u[RESIZE_ROWS][3], v[RESIZE_ROWS][3];
#pragma unroll
    for (int r = 0; r < RESIZE_ROWS; r++) {
        const int sy0 = min(max(ty[r].x, 0), sh1), sy1 = min(max(ty[r].x + 1, 0), sh1);
        const unsigned* r0 = (const unsigned*)(src + sy0 * spitch + a0);
        const unsigned* r1 = (const unsigned*)(src + sy1 * spitch + a0);
        u[r][0] = r0[0]; u[r][1] = r0[1]; u[r][2] = r0[2];
        v[r][0] = r1[0]; v[r][1] = r1[1]; v[r][2] = r1[2];
    }
    uint8_t* dst = pyr + D.pyrOff + (long long)f * D.frameStride + yb * dpitch + x4;
    const unsigned keep = (x4 + 4 > D.w) ? 0xffffffffu >> (8 * (x4 + 4 - D.w)) : 0xffffffffu;  // keep the row padding zero
#pragma unroll
    for (int r = 0; r < RESIZE_ROWS; r++) {
        const unsigned b0s = (unsigned)(ty[r].y & 0xffff) << 16, b1s = (unsigned)ty[r].y & 0xffff0000u;  // b << 16: (b*x)>>16 == umulhi
        const unsigned ulo = __funnelshift_r(u[r][0], u[r][1], sh), uhi = __funnelshift_r(u[r][1], u[r][2], sh);  // source bytes sx0 .. sx0+7
        const unsigned vlo = __funnelshift_r(v[r][0], v[r][1], sh), vhi = __funnelshift_r(v[r][1], v[r][2], sh);
        unsigned out = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const unsigned H0 = dp2a_lo_u32(cf[i], __byte_perm(ulo, uhi, sel[i]), 0u);
            const unsigned H1 = dp2a_lo_u32(cf[i], __byte_perm(vlo, vhi, sel[i]), 0u);
            const unsigned val = (__umulhi(b0s, H0 >> 4) + __umulhi(b1s, H1 >> 4) + 2u) >> 2;
            out |= (val & 0xffu) << (8 * i);
        }
        if (yb + r < D.h) *(unsigned*)(dst + r * dpitch) = out & keep;
    }
}
