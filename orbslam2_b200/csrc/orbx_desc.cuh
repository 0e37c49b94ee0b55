// orbx_desc.cuh — K5 k_orient_desc: IC_Angle orientation + steered BRIEF-256 + keypoint assembly.
// Reference: IC_Angle /root/reference/src/ORBextractor.cc:72-99 (+ cv::fastAtan2), computeOrbDescriptor :103-142,
// border offset / octave / size :831-841, level scaling :909-915.
//
// One warp per keypoint, each warp walks kpw consecutive keypoints of one frame so that everything that depends only
// on the lane is computed once: the circular-patch byte masks (lane = patch row v = lane-15, |u| <= umax[|v|]) and the
// lane's 8 BRIEF pairs as floats (lane i produces descriptor byte i = pairs 8i..8i+7, LSB first).
//   staging:     the 31x31 un-blurred patch and the 37x37 blurred patch (pattern reach +-18) arrive in shared memory as two
//                TMA box loads per keypoint (48 x 31 and 64 x 37 bytes, x start rounded down to 16), double buffered, with
//                the selection records fetched two keypoints ahead (round 1 / early round 2: 21 flat warp loads + 21 shared
//                stores per keypoint, which held the kernel at 74 % of the L1 data pipe's wavefront rate)
//   orientation: 9 words of the lane's box row from shared memory, funnel-shift to the patch start, mask, two DP4A per word
//                (m10 += sum u*I with constant s8 weights, row sum for m01 = v * sum I), warp-shuffle reduction,
//                OpenCV's fastAtan2 polynomial op for op in fp32 without FMA contraction.
//   descriptor:  a = cos(angle), b = sin(angle); sample (round(x*b + y*a), round(x*a - y*b)) from the blurred level;
//                rounding is round-half-even via the 1.5*2^23 magic add (what cvRound / rint do).
#pragma once
#include "orbx_internal.cuh"

#define DESC_KPW 1         // keypoints per warp on the single-frame path (measured 1 / 2 / 4 / 8: 0.094 / 0.097 / 0.111 / 0.121 ms per call)
#define DESC_KPW_BATCH 16  // ... and for large batches (measured 2 / 4 / 8 / 16 / 32: 0.330 / 0.300 / 0.292 / 0.290 / 0.305 ms)
#define DESC_RAW_BW 48      // TMA box width of the un-blurred patch: 15 bytes of alignment slack + 31, rounded up to 16
#define DESC_BLR_BW 64      // TMA box width of the blurred patch: 15 + 37, rounded up to 16
#define DESC_RAW_BYTES 1536 // 48 x 31 = 1488, rounded up to 128 (TMA destination alignment)
#define DESC_BUF_BYTES (DESC_RAW_BYTES + 2432)  // + 64 x 37 = 2368, rounded up to 128: one keypoint's two boxes
#define DESC_SMEM_BYTES (8 * 2 * DESC_BUF_BYTES + 8 * 2 * 8)  // 8 warps x 2 buffers + their mbarriers

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

// tmaps[l] / tmaps[ORBX_MAX_LEVELS + l]: 3-D (x, y, frame) u8 tensor maps of pyramid / blurred level l with boxes
// DESC_RAW_BW x 31 and DESC_BLR_BW x 37 (device memory, written by the host before the launch).
__global__ void __launch_bounds__(256, 3) k_orient_desc(const DevPlan* __restrict__ P, const CUtensorMap* __restrict__ tmaps,
                                                     const uint2* __restrict__ sel, const int* __restrict__ selCount,
                                                     orbx_kp* __restrict__ kps, uint8_t* __restrict__ desc,
                                                     int* __restrict__ nOut, int* __restrict__ err, int f0, int kpw) {
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
    const int o0 = warpGlobal * kpw;  // kpw consecutive keypoints per warp (host: 4 for a few frames, more for large batches)
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

    // Per-warp staging buffers, double buffered: the un-blurred 31-row patch and the blurred 37-row patch (pattern reach
    // +-18) of a keypoint arrive as TWO TMA box loads (cp.async.bulk.tensor, one elected lane, completion on an mbarrier) —
    // the patches of keypoint k+1 are in flight while keypoint k is evaluated. The box starts at the patch's x rounded
    // down to 16 bytes (TMA's innermost start coordinate must be 16-byte aligned on sm_100a), hence the 48- and 64-byte
    // rows. Before (21 flat warp loads + 21 shared stores per keypoint) the kernel sat at 74 % of the L1 data pipe's
    // wavefront rate: ~105 cache-line touches per keypoint for the loads alone. TMA traffic does not pass through the LSU.
    extern __shared__ __align__(128) uint8_t desc_smem[];
    const int wib = threadIdx.x >> 5;
    uint8_t* wbuf = desc_smem + wib * (2 * DESC_BUF_BYTES);
    unsigned long long* bars = (unsigned long long*)(desc_smem + 8 * 2 * DESC_BUF_BYTES) + 2 * wib;
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(bars);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    // selection records run two keypoints ahead of the evaluation (record k+2 is requested while keypoint k is evaluated,
    // its boxes one iteration later), so neither the record nor the patch latency sits on the warp's critical path
    const int nk = min(kpw, total - o0);
    auto fetch = [&](int k, int& level, uint2& s) {
        const int o = o0 + k;
        const unsigned ge = __ballot_sync(0xffffffffu, lane < nl && o >= endOff);
        level = __popc(ge);
        const int before = level ? __shfl_sync(0xffffffffu, endOff, level - 1) : 0;
        s = sel[(long long)f * P->selStride + P->lv[level].selOff + (o - before)];
    };
    int lvA = 0, lvB = 0, lvC = 0;
    uint2 svA = make_uint2(0u, 0u), svB = svA, svC = svA;
    fetch(0, lvA, svA);
    if (nk > 1) fetch(1, lvB, svB);
    auto stage = [&](int level, uint2 s, int buf) {  // lane 0: request both boxes of one keypoint
        if (lane == 0) {
            const int x = s.x & 0xFFFF, y = s.x >> 16;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(wbuf + buf * DESC_BUF_BYTES);
            const uint32_t bar = bar0 + 8u * buf;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
                         "r"((uint32_t)(DESC_RAW_BW * 31 + DESC_BLR_BW * 37))
                         : "memory");
            asm volatile(
                "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                ::"r"(dst), "l"(tmaps + level), "r"((x - 15) & ~15), "r"(y - 15), "r"(f), "r"(bar)
                : "memory");
            asm volatile(
                "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                ::"r"(dst + DESC_RAW_BYTES), "l"(tmaps + ORBX_MAX_LEVELS + level), "r"((x - 18) & ~15), "r"(y - 18), "r"(f), "r"(bar)
                : "memory");
        }
    };
    stage(lvA, svA, 0);

    for (int k = 0; k < nk; k++) {
        const int o = o0 + k;
        const int level = lvA;
        const uint2 s = svA;
        const LevelGeo& G = P->lv[level];
        const int x = s.x & 0xFFFF, y = s.x >> 16;
        const int xs = x - 15, sh = (xs & 3) * 8;
        if (k + 1 < nk) stage(lvB, svB, (k + 1) & 1);
        if (k + 2 < nk) fetch(k + 2, lvC, svC);
        {   // bounded wait for this keypoint's boxes (a lost TMA must not hang the GPU: flag the error and go on)
            const uint32_t bar = bar0 + 8u * (k & 1), parity = (uint32_t)((k >> 1) & 1);
            uint32_t done = 0;
            int spins = 0;
            while (!done) {
                asm volatile(
                    "{\n\t.reg .pred p;\n\t"
                    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                    "selp.u32 %0, 1, 0, p;\n\t}"
                    : "=r"(done)
                    : "r"(bar), "r"(parity)
                    : "memory");
                if (!done && ++spins > (1 << 22)) {
                    atomicExch(err, 6);
                    break;
                }
            }
        }
        const unsigned* raw = (const unsigned*)(wbuf + (k & 1) * DESC_BUF_BYTES);
        const uint8_t* blr = wbuf + (k & 1) * DESC_BUF_BYTES + DESC_RAW_BYTES;

        // ---- IC_Angle
        int m10 = 0, rs = 0;
        if (rowOn) {
            // row = lane; the patch starts (xs & 15) bytes into the 48-byte box row (12-word pitch: rows l, l+8, l+16, l+24
            // share banks, a 4-way conflict on 9 loads — still far fewer wavefronts than staging the patch by hand)
            const unsigned* rowp = raw + lane * (DESC_RAW_BW / 4) + ((xs & 15) >> 2);
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
        const uint8_t* ctr = blr + 18 * DESC_BLR_BW + 18 + ((x - 18) & 15);  // (row, col) of the keypoint inside the staged box
        int t[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const int r = round_half_even(__fadd_rn(__fmul_rn(px[k], b), __fmul_rn(py[k], a)));
            const int c = round_half_even(__fsub_rn(__fmul_rn(px[k], a), __fmul_rn(py[k], b)));
            t[k] = ctr[r * DESC_BLR_BW + c];
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
        __syncwarp();  // every lane is done with this buffer before lane 0 requests keypoint k + 2 into it
        lvA = lvB; svA = svB;
        lvB = lvC; svB = svC;
    }
}
