// orbx_desc.cuh — K5 k_orient_desc: IC_Angle orientation + steered BRIEF-256 + keypoint assembly.
// Reference: IC_Angle /root/reference/src/ORBextractor.cc:72-99 (+ cv::fastAtan2), computeOrbDescriptor :103-142,
// border offset / octave / size :831-841, level scaling :909-915.
//
// One warp per keypoint, each warp walks KPW consecutive keypoints of one frame so that everything that depends only
// on the lane is computed once: the circular-patch byte masks (lane = patch row v = lane-15, |u| <= umax[|v|]) and the
// lane's 8 BRIEF pairs as floats (lane i produces descriptor byte i = pairs 8i..8i+7, LSB first).
//   staging:     the 31x31 un-blurred patch and the 37x37 blurred patch (pattern reach +-18) are copied to shared memory
//                with flat, row-crossing warp loads (~4 rows per load instead of one row per lane: 8x fewer L1 wavefronts)
//   orientation: 9 words per row from shared memory, funnel-shift to the patch start, mask, two DP4A per word
//                (m10 += sum u*I with constant s8 weights, row sum for m01 = v * sum I), warp-shuffle reduction,
//                OpenCV's fastAtan2 polynomial op for op in fp32 without FMA contraction.
//   descriptor:  a = cos(angle), b = sin(angle); sample (round(x*b + y*a), round(x*a - y*b)) from the blurred level;
//                rounding is round-half-even via the 1.5*2^23 magic add (what cvRound / rint do).
#pragma once
#include "orbx_internal.cuh"

#define DESC_KPW 4  // keypoints per warp
#define DESC_RAW_WORDS 288  // 31 rows x 9 words, rounded up to a multiple of 32
#define DESC_BLR_WORDS 384  // 37 rows x 10 words, rounded up to a multiple of 32

__device__ signed char d_pattern[1024];  // 256 x (x1, y1, x2, y2), the reference's bit_pattern_31_

__device__ __forceinline__ int dp4a_u8s8(unsigned a, int b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int round_half_even(float v) {  // exact for |v| < 2^22
    return __float_as_int(__fadd_rn(v, 12582912.0f)) - 0x4B400000;
}

__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.141592653589793238462643383279502884);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    const float eps = 2.2204460492503131e-16f;  // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f,
                      __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// s8x4 weights u = 4j-15 .. 4j-12 for word j of the 32-byte patch row (byte t <-> u = t - 15)
__host__ __device__ constexpr int desc_uweights(int j) {
    return (int)(((unsigned)((4 * j - 15) & 0xFF)) | ((unsigned)((4 * j - 14) & 0xFF) << 8) |
                 ((unsigned)((4 * j - 13) & 0xFF) << 16) | ((unsigned)((4 * j - 12) & 0xFF) << 24));
}

__global__ void __launch_bounds__(256, 3) k_orient_desc(const DevPlan* __restrict__ P, const uint8_t* __restrict__ pyr,
                                                     const uint8_t* __restrict__ blur, const uint2* __restrict__ sel,
                                                     const int* __restrict__ selCount, orbx_kp* __restrict__ kps,
                                                     uint8_t* __restrict__ desc, int* __restrict__ nOut, int f0) {
    const int lane = threadIdx.x & 31;
    const int f = f0 + blockIdx.y;
    const int nl = P->nlevels;
    // per-frame level prefix (lane l holds the end offset of level l in the compact output)
    int cnt = (lane < nl) ? selCount[f * nl + lane] : 0;
    int endOff = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, endOff, o);
        if (lane >= o) endOff += y;
    }
    const int total = __shfl_sync(0xffffffffu, endOff, nl - 1);
    const int warpGlobal = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (warpGlobal == 0 && lane == 0) nOut[f] = total;
    const int o0 = warpGlobal * DESC_KPW;
    if (o0 >= total) return;

    // lane constants: patch-row byte masks and the lane's 8 pattern pairs
    const int hp = P->halfPatch;  // 15 for the 31x31 patch (the ABI guarantees hp <= 15 for this kernel)
    const int v = lane - hp;
    const bool rowOn = lane <= 2 * hp;
    unsigned mask[8];
    {
        const int d = rowOn ? P->umax[v < 0 ? -v : v] : -1;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            unsigned m = 0;
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int u = 4 * j + b - 15;
                if (u >= -d && u <= d) m |= 0xFFu << (8 * b);
            }
            mask[j] = m;
        }
    }
    float px[16], py[16];
    {
        const uint4* pp = (const uint4*)(d_pattern + lane * 32);
        const uint4 p0 = pp[0], p1 = pp[1];
        const unsigned w[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            px[2 * k] = (float)(signed char)(w[k] & 0xFF);
            py[2 * k] = (float)(signed char)((w[k] >> 8) & 0xFF);
            px[2 * k + 1] = (float)(signed char)((w[k] >> 16) & 0xFF);
            py[2 * k + 1] = (float)(signed char)((w[k] >> 24) & 0xFF);
        }
    }

    // per-warp staging buffers: the un-blurred 31-row patch (9 words per row) and the blurred 37-row patch (10 words per
    // row). Both are fetched with flat lane -> (row, word) indexing, so one warp-load touches ~4 rows instead of 32.
    __shared__ unsigned sh_raw[256 / 32][DESC_RAW_WORDS];
    __shared__ unsigned sh_blr[256 / 32][DESC_BLR_WORDS];
    unsigned* raw = sh_raw[threadIdx.x >> 5];
    unsigned* blr = sh_blr[threadIdx.x >> 5];

    for (int o = o0; o < min(o0 + DESC_KPW, total); o++) {
        // level of output index o
        const unsigned ge = __ballot_sync(0xffffffffu, lane < nl && o >= endOff);
        const int level = __popc(ge);
        const int before = level ? __shfl_sync(0xffffffffu, endOff, level - 1) : 0;
        const LevelGeo& G = P->lv[level];
        const uint2 s = sel[(long long)f * P->selStride + G.selOff + (o - before)];
        const int x = s.x & 0xFFFF, y = s.x >> 16;
        const long long lvlOff = G.pyrOff + (long long)f * G.frameStride;
        const int pitch = G.pitch;

        // ---- stage both patches (all loads of the keypoint are in flight together)
        const int xs = x - 15, a0 = xs & ~3, sh = (xs & 3) * 8;
        const int xb = x - 18, b0 = xb & ~3, shB = xb & 3;
        {
            const uint8_t* rbase = pyr + lvlOff + (long long)(y - 15) * pitch + a0;
            const uint8_t* bbase = blur + lvlOff + (long long)(y - 18) * pitch + b0;
            unsigned rv[DESC_RAW_WORDS / 32], bv[DESC_BLR_WORDS / 32];
#pragma unroll
            for (int i = 0; i < DESC_RAW_WORDS / 32; i++) {
                const int idx = lane + 32 * i;
                const int row = (idx * 7282) >> 16;  // idx / 9 for idx < 288
                rv[i] = (idx < 31 * 9) ? *(const unsigned*)(rbase + row * pitch + 4 * (idx - row * 9)) : 0u;
            }
#pragma unroll
            for (int i = 0; i < DESC_BLR_WORDS / 32; i++) {
                const int idx = lane + 32 * i;
                const int row = (idx * 6554) >> 16;  // idx / 10 for idx < 384
                bv[i] = (idx < 37 * 10) ? *(const unsigned*)(bbase + row * pitch + 4 * (idx - row * 10)) : 0u;
            }
            __syncwarp();  // the previous keypoint's readers are done with the buffers
#pragma unroll
            for (int i = 0; i < DESC_RAW_WORDS / 32; i++) raw[lane + 32 * i] = rv[i];
#pragma unroll
            for (int i = 0; i < DESC_BLR_WORDS / 32; i++) blr[lane + 32 * i] = bv[i];
            __syncwarp();
        }

        // ---- IC_Angle
        int m10 = 0, rs = 0;
        if (rowOn) {
            const unsigned* rowp = raw + lane * 9;  // bank = (9 * lane + j) % 32: conflict free
            unsigned w[9];
#pragma unroll
            for (int j = 0; j < 9; j++) w[j] = rowp[j];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const unsigned a = __funnelshift_r(w[j], w[j + 1], sh) & mask[j];
                m10 = dp4a_u8s8(a, desc_uweights(j), m10);
                rs = dp4a_u8s8(a, 0x01010101, rs);
            }
        }
        int m01 = v * rs;
#pragma unroll
        for (int q = 16; q > 0; q >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, q);
            m01 += __shfl_xor_sync(0xffffffffu, m01, q);
        }
        const float angle = fast_atan2_deg((float)m01, (float)m10);

        // ---- steered BRIEF on the blurred level
        const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
        const float rad = __fmul_rn(angle, factorPI);
        // cos / sin of the float angle as glibc's cosf / sinf give them (double evaluation, one rounding); sincos shares the
        // argument reduction between the two (orient_desc 0.410 -> 0.383 ms per 384 frames). Measured and dropped: evaluating
        // the warp's four keypoints in one sincos (lane k = keypoint k) with four blurred patches staged per warp — 0.42 ms,
        // the two-phase loop costs more than the three saved evaluations.
        double sd, cd;
        sincos((double)rad, &sd, &cd);
        const float a = (float)cd, b = (float)sd;
        const uint8_t* ctr = (const uint8_t*)blr + 18 * 40 + 18 + shB;  // (row, col) of the keypoint inside the staged patch
        int t[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const int r = round_half_even(__fadd_rn(__fmul_rn(px[k], b), __fmul_rn(py[k], a)));
            const int c = round_half_even(__fsub_rn(__fmul_rn(px[k], a), __fmul_rn(py[k], b)));
            t[k] = ctr[r * 40 + c];
        }
        int val = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) val |= (t[2 * k] < t[2 * k + 1]) << k;
        const long long oo = (long long)f * P->selStride + o;
        desc[oo * 32 + lane] = (uint8_t)val;
        if (lane == 0) {
            orbx_kp kp;
            kp.x = (level == 0) ? (float)x : __fmul_rn((float)x, G.scale);
            kp.y = (level == 0) ? (float)y : __fmul_rn((float)y, G.scale);
            kp.size = G.sizeField;
            kp.angle = angle;
            kp.response = (float)s.y;
            kp.octave = level;
            kp.class_id = -1;
            kps[oo] = kp;
        }
    }
}
