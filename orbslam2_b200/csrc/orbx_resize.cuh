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
#define RESIZE_BX 16   // CTA shape of k_resize_fast (threads)
#define RESIZE_BY 8    // measured per 384 frames: 16x16 0.278 ms, 16x8 0.269, 32x8 0.293, 8x16 0.290, 8x32 0.307, 4x64 0.444

// multipliers of the result packing, read from constant memory so that ptxas keeps the IMADs (FMA pipe) instead of
// strength-reducing them to LEA / shifts on the ALU pipe
__constant__ unsigned g_resize_mul[2] = {1u << 6, 1u << 22};
__device__ __forceinline__ unsigned resize_imad(unsigned a, unsigned b, unsigned c) {
    unsigned d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__global__ void __launch_bounds__(256) k_resize_fast(const __grid_constant__ DevPlan Pc, int level, uint8_t* __restrict__ pyr,
                                                     const int2* __restrict__ tabs, int f0) {
    const DevPlan* P = &Pc;  // the plan rides in the kernel parameters (constant bank): no global loads in the prologue
    const LevelGeo& D = P->lv[level];
    const LevelGeo& S = P->lv[level - 1];
    // CTA = 16 x 8 threads = 64 pixels x 32 rows, so a warp is 64 pixels x 8 rows: level widths are not multiples of anything
    // useful, and with 128-pixel warps (64 x 4 threads) 17 % of the lanes of a 640x480 pyramid lay past the right edge
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int yb = (blockIdx.y * blockDim.y + threadIdx.y) * RESIZE_ROWS;
    const int f = f0 + blockIdx.z;
    if (x4 >= D.w || yb >= D.h) return;
    const uint8_t* src = pyr + S.pyrOff + (long long)f * S.frameStride;
    const int spitch = S.pitch, dpitch = D.pitch;
    // x table: 4 entries {selector | s << 16, a0 | a1 << 16} (padded to a multiple of 4 entries); shared by the thread's rows
    const uint4* xt = (const uint4*)(tabs + D.xtabOff + x4);
    const uint4 t01 = xt[0], t23 = xt[1];
    // y table: {row0 | row1 << 16 (clamped on the host), b0 | b1 << 16}; padded to a multiple of 4 entries and 16-byte aligned,
    // so the four rows' entries are two vector loads issued next to the x table's — every table word is in flight at once
    // and all 24 source loads depend on one round trip only
    static_assert(RESIZE_ROWS == 4, "the y table is read as two uint4");
    const uint4* yt = (const uint4*)(tabs + D.ytabOff + yb);
    const uint4 y01 = yt[0], y23 = yt[1];
    const int2 ty[RESIZE_ROWS] = {make_int2((int)y01.x, (int)y01.y), make_int2((int)y01.z, (int)y01.w),
                                  make_int2((int)y23.x, (int)y23.y), make_int2((int)y23.z, (int)y23.w)};
    // an entry's .x is {PRMT selector of source bytes (rel, rel + 1) | source index << 16}, rel = offset from the group's first
    // source pixel (host-built, see build_plan); PRMT reads only the low 16 bits of its selector, so .x is used as it is
    const int sx0 = (int)(t01.x >> 16);
    const int a0 = sx0 & ~3, sh = (sx0 & 3) * 8;
    const unsigned sel[4] = {t01.x, t01.z, t23.x, t23.z};
    const unsigned cf[4] = {t01.y, t01.w, t23.y, t23.w};
    // all source words of the thread's rows are requested before any is used (the kernel is latency bound)
    unsigned u[RESIZE_ROWS][3], v[RESIZE_ROWS][3];
    const unsigned* sbase = (const unsigned*)(src + a0);
    const int spw = spitch >> 2;  // the pitch is a multiple of 64 bytes
#pragma unroll
    for (int r = 0; r < RESIZE_ROWS; r++) {
        const int sy0 = ty[r].x & 0xffff, sy1 = (int)((unsigned)ty[r].x >> 16);
        const unsigned* r0 = sbase + sy0 * spw;  // (pointer + int index: one IMAD.WIDE instead of a 64-bit add pair)
        const unsigned* r1 = sbase + sy1 * spw;
        u[r][0] = r0[0]; u[r][1] = r0[1]; u[r][2] = r0[2];
        v[r][0] = r1[0]; v[r][1] = r1[1]; v[r][2] = r1[2];
    }
    // Scheduling fence. ptxas sinks the loads of rows 1..3 below the arithmetic of row 0 (to save registers), and a warp then
    // waits for one memory round trip per row instead of one in all. An empty asm does not survive into PTX, so the fence is
    // a real data dependence: the funnel-shift amount of every row takes (all words of rows 1..3) & 0, with a zero the
    // compiler cannot see through (level is >= 1, so its sign bit is 0): 10 instructions per thread.
    unsigned dep = 0;
#pragma unroll
    for (int r = 1; r < RESIZE_ROWS; r++) dep |= u[r][0] | u[r][1] | u[r][2] | v[r][0] | v[r][1] | v[r][2];
    const int shd = sh | (int)(dep & ((unsigned)level >> 31));
    uint8_t* dst = pyr + D.pyrOff + (long long)f * D.frameStride + yb * dpitch + x4;
    const unsigned mulLo = g_resize_mul[0], mulHi = g_resize_mul[1];
    const unsigned keep = (x4 + 4 > D.w) ? 0xffffffffu >> (8 * (x4 + 4 - D.w)) : 0xffffffffu;  // keep the row padding zero
#pragma unroll
    for (int r = 0; r < RESIZE_ROWS; r++) {
        const unsigned b0s = (unsigned)(ty[r].y & 0xffff) << 16, b1s = (unsigned)ty[r].y & 0xffff0000u;  // b << 16: (b*x)>>16 == umulhi
        const unsigned ulo = __funnelshift_r(u[r][0], u[r][1], shd), uhi = __funnelshift_r(u[r][1], u[r][2], shd);  // source bytes sx0 .. sx0+7
        const unsigned vlo = __funnelshift_r(v[r][0], v[r][1], shd), vhi = __funnelshift_r(v[r][1], v[r][2], shd);
        // s = b0-part + b1-part <= 1020; the result byte is (s + 2) >> 2. Two sums are packed as (s + 2) << 6 into the 16-bit
        // lanes of one word by IMADs (FMA pipe; the + 2 of both lanes rides in the first addend), so each result byte is
        // byte 1 of its lane and one PRMT gathers the four: no add, shift, mask or OR per pixel on the ALU pipe.
        unsigned q[2];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const unsigned H0 = dp2a_lo_u32(cf[i], __byte_perm(ulo, uhi, sel[i]), 0u);
            const unsigned H1 = dp2a_lo_u32(cf[i], __byte_perm(vlo, vhi, sel[i]), 0u);
            const unsigned s = __umulhi(b0s, H0 >> 4) + __umulhi(b1s, H1 >> 4);
            q[i >> 1] = (i & 1) ? resize_imad(s, mulHi, q[i >> 1]) : resize_imad(s, mulLo, (2u << 6) | (2u << 22));
        }
        const unsigned out = __byte_perm(q[0], q[1], 0x7531);
        if (yb + r < D.h) *(unsigned*)(dst + r * dpitch) = out & keep;
    }
}
