// orbx_fast.cuh — K2 k_fast: per-cell FAST-9/16 + cell-local NMS + ini/min threshold rule.
// Reference: ORBextractor::ComputeKeyPointsOctTree first half, /root/reference/src/ORBextractor.cc:759-823, and the
// external cv::FAST(img, kps, th, true) (TYPE_9_16) it calls per 30-px cell.
//
// Formulation (threshold independent, so the reference's "FAST(ini) else FAST(min)" retry needs no second pass):
//   for a centre v and its 16-pixel Bresenham circle c_0..c_15
//     M = max over the 16 arcs of 9 of min(c over the arc)      m = min over the arcs of max(c over the arc)
//     score = max(M - v, v - m) - 1   = the largest t at which the pixel is still a FAST corner (OpenCV cornerScore)
//   the pixel is a corner at threshold t  <=>  score >= t.
// The arc minima / maxima are computed for TWO pixels per instruction with the packed 16-bit DPX min/max
// (VIMNMX.U16x2 / VIMNMX3.U16x2) in the paired-arc formulation of fast_score2: 36 instructions per polarity and pixel
// pair. Measured on B200 (scripts/pipe_probe.cu): every ALU-pipe instruction — VIMNMX 2- or 3-input, HMNMX2, PRMT, LOP3,
// IADD3 — issues at 2 warp-inst/clk/SM, so this kernel is bound by its ALU-pipe instruction count; IMAD/DP4A go to the
// FMA pipe (also 2/clk/SM) and overlap. That is why 16 of the 72 min/max come from max(a, b) = a + b - min(a, b) as two
// IMADs (fast_score2<NI1, NI2>). One thread scores 4 horizontally adjacent pixels (one 32-bit word of the tile).
//
// One CTA = a "supertile" of stCX x stCY cells of one level of one frame (its geometry arrives as one host-built FastSt record):
//   TMA 3-D tile load (x start rounded down to 16 bytes — an unaligned innermost coordinate faults on sm_100a)
//   phase 1  score map  (u8, value = score - (minTh-1), 0 = not a corner at minTh), same coordinates as the tile
//   phase 2  SWAR non-maximum suppression, strictly greater than the 8 neighbours; a neighbour in another cell
//            counts 0 (the reference runs cv::FAST per cell ROI); survivors compacted into a shared-memory list
//   phase 3  per cell: keep survivors >= iniTh if any exist, else all; one global atomic per CTA; emit records.
#pragma once
#include <cuda_fp16.h>

#include "orbx_internal.cuh"

#define FAST_THREADS 256
#define FAST_MAX_ST_CELLS 16  // cells per supertile (stCX * stCY)

// bytes (x+dx .. x+dx+3) of a row given the three words around word x
__device__ __forceinline__ unsigned fast_align(unsigned wm, unsigned w0, unsigned wp, int dx) {
    switch (dx) {
        case -3: return __byte_perm(wm, w0, 0x4321);
        case -2: return __byte_perm(wm, w0, 0x5432);
        case -1: return __byte_perm(wm, w0, 0x6543);
        case 1: return __byte_perm(w0, wp, 0x4321);
        case 2: return __byte_perm(w0, wp, 0x5432);
        case 3: return __byte_perm(w0, wp, 0x6543);
        default: return w0;
    }
}

// Packed score of two pixels (16-bit lanes, pixel value in the HIGH byte, don't-care low byte): c[k] = circle values,
// vh = centres; returns max(score - (minTh-1), 0) as plain 16-bit lanes.
// The 16 arcs of 9 pair up around the 8 even-aligned runs of 8: with A8[k] = min(c[k..k+7]) (k even, by doubling),
//   arc9(k) = min(A8[k], c[k+8])   arc9(k-1) = min(c[k-1], A8[k])   =>   max of the two = min(A8[k], max(c[k+8], c[k-1]))
// so M = max over the 8 even k of min(A8[k], max(c[k+8], c[k-1])): 36 instructions per polarity instead of 40.
// NI1 of the 8 first-stage pairs and NI2 of the 8 (c[k+8], c[k-1]) pairs take their max from the identity
//   max(a, b) = a + b - min(a, b)
// which holds for the PACKED words modulo 2^32 (the per-lane results fit their 16 bits, so the carries between the lanes
// cancel): two IMADs on the otherwise idle FMA pipe (multipliers 1 and -1 read from constant memory so ptxas keeps them
// IMADs) replace one VIMNMX on the ALU pipe that bounds the kernel.
__constant__ unsigned g_fast_mul[2] = {1u, 0xFFFFFFFFu};
__device__ __forceinline__ unsigned fast_imad(unsigned a, unsigned b, unsigned c) {
    unsigned d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
template <int NI1, int NI2>
__device__ __forceinline__ unsigned fast_score2(const unsigned (&c)[16], unsigned vh, unsigned negThm1) {
    unsigned a2[8], b2[8], a4[8], b4[8];
    const unsigned one = g_fast_mul[0], mone = g_fast_mul[1];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        a2[j] = __vminu2(c[2 * j], c[2 * j + 1]);
        // odd j first so that the FMA-pipe pairs are spread over the tree when NI1 < 8
        if (((j >> 1) + ((j & 1) ? 0 : 4)) < NI1) b2[j] = fast_imad(a2[j], mone, fast_imad(c[2 * j], one, c[2 * j + 1]));
        else b2[j] = __vmaxu2(c[2 * j], c[2 * j + 1]);
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
        a4[j] = __vminu2(a2[j], a2[(j + 1) & 7]);
        b4[j] = __vmaxu2(b2[j], b2[(j + 1) & 7]);
    }
    unsigned r[8], q[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int k = 2 * j;
        const unsigned dmax = __vmaxu2(c[(k + 8) & 15], c[(k + 15) & 15]);
        unsigned dmin;
        if (((j >> 1) + ((j & 1) ? 0 : 4)) < NI2) dmin = fast_imad(dmax, mone, fast_imad(c[(k + 8) & 15], one, c[(k + 15) & 15]));
        else dmin = __vminu2(c[(k + 8) & 15], c[(k + 15) & 15]);
        r[j] = __vimin3_u16x2(a4[j], a4[(j + 2) & 7], dmax);
        q[j] = __vimax3_u16x2(b4[j], b4[(j + 2) & 7], dmin);
    }
    const unsigned Mh = __vimax3_u16x2(__vimax3_u16x2(r[0], r[1], r[2]), __vimax3_u16x2(r[3], r[4], r[5]), __vmaxu2(r[6], r[7]));
    const unsigned mh = __vimin3_u16x2(__vimin3_u16x2(q[0], q[1], q[2]), __vimin3_u16x2(q[3], q[4], q[5]), __vminu2(q[6], q[7]));
    // the pixel values are the high bytes of the lanes (see k_fast): bring them down before the signed 16-bit arithmetic
    const unsigned M = __byte_perm(Mh, 0u, 0x4341), m = __byte_perm(mh, 0u, 0x4341), v = __byte_perm(vh, 0u, 0x4341);
    // per 16-bit lane ~x == -x-1, so M + ~v = (M - v) - 1 and v + ~m = (v - m) - 1
    const unsigned sb = __viaddmax_s16x2(M, ~v, 0u);
    const unsigned sd = __viaddmax_s16x2(v, ~m, 0u);
    return __viaddmax_s16x2(__vmaxu2(sb, sd), negThm1, 0u);
}

// ---- k_fast variant 3 (round 2): the same score on a pre-expanded tile, min/max split over the ALU and the FMA pipe.
// Every tile pixel is expanded ONCE to a 16-bit lane 0x6400 | p — the fp16 number 1024 + p — and the lanes of one 32-bit word
// are the pixels (x, y) and (x, y + Q2) of the two halves ("bands") of the supertile's rows. A circle operand at ANY offset
// (dx, dy) is then one aligned LDS.32 (the 18 PRMT + 6 shifts per 4 pixels of variant 1 are gone), and because positive fp16
// numbers order like their bit patterns the lanes are valid operands for both VIMNMX.U16x2 (ALU pipe) and half2 arithmetic
// (FMA pipe). Where a min and a max of the SAME two operands are needed,
//     t = relu(a - b)   (one HFMA2.RELU)      max = b + t      min = a - t
// is exact (integers 1024..1279 and their differences are fp16 numbers) and moves two ALU-pipe instructions to three on the
// otherwise idle FMA pipe. NF1 of the 8 first-stage pairs and NF2 of the 8 (c[k+8], c[k-1]) pairs take that route.
__device__ __forceinline__ unsigned h2_relu_sub(unsigned a, unsigned b) {  // max(a - b, 0)
    unsigned d;
    asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(b), "r"(0xBC00BC00u), "r"(a));
    return d;
}
__device__ __forceinline__ unsigned h2_add(unsigned a, unsigned b) {
    unsigned d;
    asm("add.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ unsigned h2_sub(unsigned a, unsigned b) {
    unsigned d;
    asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ void h2_minmax(unsigned a, unsigned b, unsigned& mn, unsigned& mx) {
    const unsigned t = h2_relu_sub(a, b);
    mx = h2_add(b, t);
    mn = h2_sub(a, t);
}

// fast_score2 on fp16-biased lanes (0x6400 | pixel): returns 0x6400 | max(score - (minTh-1), 0) per lane.
// Cth = the fp16 pair (1024 - minTh).
template <int NF1, int NF2>
__device__ __forceinline__ unsigned fast_score2h(const unsigned (&c)[16], unsigned v, unsigned Cth) {
    unsigned a2[8], b2[8], a4[8], b4[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        // odd j first so that the FMA-pipe pairs are spread over the tree when NF1 < 8
        if (((j >> 1) + ((j & 1) ? 0 : 4)) < NF1) {
            h2_minmax(c[2 * j], c[2 * j + 1], a2[j], b2[j]);
        } else {
            a2[j] = __vminu2(c[2 * j], c[2 * j + 1]);
            b2[j] = __vmaxu2(c[2 * j], c[2 * j + 1]);
        }
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
        a4[j] = __vminu2(a2[j], a2[(j + 1) & 7]);
        b4[j] = __vmaxu2(b2[j], b2[(j + 1) & 7]);
    }
    unsigned r[8], q[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int k = 2 * j;
        unsigned dmax, dmin;
        if (((j >> 1) + ((j & 1) ? 0 : 4)) < NF2) {
            h2_minmax(c[(k + 8) & 15], c[(k + 15) & 15], dmin, dmax);
        } else {
            dmax = __vmaxu2(c[(k + 8) & 15], c[(k + 15) & 15]);
            dmin = __vminu2(c[(k + 8) & 15], c[(k + 15) & 15]);
        }
        r[j] = __vimin3_u16x2(a4[j], a4[(j + 2) & 7], dmax);
        q[j] = __vimax3_u16x2(b4[j], b4[(j + 2) & 7], dmin);
    }
    const unsigned M = __vimax3_u16x2(__vimax3_u16x2(r[0], r[1], r[2]), __vimax3_u16x2(r[3], r[4], r[5]), __vmaxu2(r[6], r[7]));
    const unsigned m = __vimin3_u16x2(__vimin3_u16x2(q[0], q[1], q[2]), __vimin3_u16x2(q[3], q[4], q[5]), __vminu2(q[6], q[7]));
    // 1024 + (M - v - minTh) and 1024 + (v - m - minTh): each partial sum is an integer of magnitude < 2048, i.e. exact
    const unsigned sb = h2_add(M, h2_sub(Cth, v));
    const unsigned sd = h2_add(v, h2_sub(Cth, m));
    return __vimax3_u16x2(sb, sd, 0x64006400u);
}

// flag (bit 7 of each byte) = (a > b) for unsigned bytes; a7 = a & 0x7f7f7f7f
__device__ __forceinline__ unsigned bytes_gt(unsigned a, unsigned a7, unsigned b) {
    const unsigned nb = ~b;
    const unsigned t = a7 + (nb & 0x7f7f7f7fu);      // carries out of the low 7 bits land in bit 7
    return (a & nb) | (a & t) | (nb & t);            // majority = carry out of bit 7 of a + ~b  <=>  a > b
}

struct FastSmemHdr {
    unsigned long long mbar;
    int nAll[FAST_MAX_ST_CELLS];
    int nIni[FAST_MAX_ST_CELLS];
    int emitBase[FAST_MAX_ST_CELLS];
    int total, gbase, slot, nList, nAct;
    unsigned firstCol[64], lastCol[64];  // per word column: bytes that are the first / last column of a cell
    unsigned valid[64];                  // per word column: bytes inside the detection columns
    unsigned char rowFlag[128];          // per detection row: bit0 first row of a cell, bit1 last row
};

// Variant 3 score loop. e0 = expanded word of (column 0, row 0) of the detection area, scb = its score-map byte.
// A warp takes (row, 32-column group) items; with EPC > 0 the expanded pitch is the compile-time constant EPC, so all
// 17 loads of an item are LDS [base + immediate] off one register.
#define FAST3_EP 144
template <int NF1, int NF2, int EPC>
__device__ __forceinline__ void fast3_scores(const unsigned* __restrict__ e0, int EPr, unsigned char* __restrict__ scb, int boxW,
                                             int dW, int Q2, int nR, unsigned Cth) {
    const int EP = EPC > 0 ? EPC : EPr;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int band1 = Q2 * boxW;
    // a warp owns rows warp, warp + 8, ... of band 0 and walks their 32-column groups (hCell is 30..32 on every level
    // of the usual configurations, i.e. Q2 = 15..16 or 30..32 rows: the 8 warps stay balanced)
    for (int row = warp; row < Q2; row += FAST_THREADS / 32) {
        const unsigned* p0 = e0 + row * EP + lane;
        unsigned char* o = scb + row * boxW + lane;
        const bool two = row + Q2 < nR;
        for (int xi = lane; xi < dW; xi += 32, p0 += 32, o += 32) {
            unsigned c[16];
            // Bresenham order (dx,dy): (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
            c[0] = p0[3 * EP]; c[1] = p0[3 * EP + 1]; c[2] = p0[2 * EP + 2]; c[3] = p0[EP + 3]; c[4] = p0[3];
            c[5] = p0[-EP + 3]; c[6] = p0[-2 * EP + 2]; c[7] = p0[-3 * EP + 1]; c[8] = p0[-3 * EP]; c[9] = p0[-3 * EP - 1];
            c[10] = p0[-2 * EP - 2]; c[11] = p0[-EP - 3]; c[12] = p0[-3]; c[13] = p0[EP - 3]; c[14] = p0[2 * EP - 2];
            c[15] = p0[3 * EP - 1];
            const unsigned s = fast_score2h<NF1, NF2>(c, p0[0], Cth);
            o[0] = (unsigned char)s;
            if (two) o[band1] = (unsigned char)(s >> 16);
        }
    }
}

// V = 1: variant 1 (byte tile, PRMT alignment, all min/max on the ALU pipe); V = 3: expanded fp16-lane tile (see above).
// Bytes of the third shared-memory region: variant 1 keeps the list of words that hold a corner there (u16 per word),
// variant 3 the expanded tile (boxW x (Q2 + 6) words), which the word list overlays once the scores are written.
__host__ __device__ inline size_t fast_region3_bytes(int V, int boxW, int boxH) {
    return V == 3 ? (size_t)boxW * (size_t)(((boxH - 6 + 1) >> 1) + 6) * 4 : (size_t)((((boxW * boxH) >> 1) + 15) & ~15);
}

template <int V, int NF1, int NF2>
__global__ void __launch_bounds__(FAST_THREADS, 4) k_fast_t(const __grid_constant__ TmapPack maps,
                                                         const FastSt* __restrict__ stTab, u64* __restrict__ cand,
                                                         int* __restrict__ candCount, int* __restrict__ err, int f0, int st0) {
    extern __shared__ __align__(128) uint8_t fast_smem[];
    const int tid = threadIdx.x;
    const int f = f0 + blockIdx.y;
    // everything a supertile needs arrives as ONE 64-byte record built by the host (build_plan): no level search, no
    // division, no chain of dependent global loads in front of the TMA request
    const FastSt S = stTab[st0 + blockIdx.x];  // (st0: the single-frame path launches the levels separately)
    const int level = S.level;
    const int cj0 = S.cj0, ci0 = S.ci0;
    const int xFirst = S.xFirst, xEnd = S.xEnd, yFirst = S.yFirst, yEnd = S.yEnd;
    if (xEnd <= xFirst || yEnd <= yFirst) return;

    const int boxW = S.boxW, boxH = S.boxH;
    const int pitchW = boxW >> 2;  // tile pitch in words
    const int tileBytes = boxW * boxH;
    unsigned* tile = (unsigned*)fast_smem;
    unsigned* sc = (unsigned*)(fast_smem + tileBytes);
    unsigned short* act = (unsigned short*)(fast_smem + 2 * tileBytes);  // units (words) with a corner, <= tileBytes/4 entries
    FastSmemHdr* H = (FastSmemHdr*)(fast_smem + 2 * tileBytes + fast_region3_bytes(V, boxW, boxH));
    const int oxa = (xFirst - 6) & ~15, oy = yFirst - 3;
    const int wx0 = (xFirst - oxa) >> 2, wx1 = (xEnd - 1 - oxa) >> 2;
    const int nW = wx1 - wx0 + 1, nR = yEnd - yFirst;

    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&H->mbar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        // the tile is requested before anything else so its latency overlaps the table set-up below
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)tileBytes) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
            ::"r"((uint32_t)__cvta_generic_to_shared(tile)), "l"(&maps.m[level]), "r"(oxa), "r"(oy), "r"(f), "r"(bar)
            : "memory");
        H->total = 0;
        H->slot = 0;
        H->nList = 0;
        H->nAct = 0;
    }
    if (tid < FAST_MAX_ST_CELLS) {
        H->nAll[tid] = 0;
        H->nIni[tid] = 0;
    }
    // column / row lookup tables (which bytes sit on a cell edge)
    const unsigned divCw = S.divCw, divCh = S.divCh;  // v / wCell == (v * divCw) >> 16 for v < 1024
    const int wCell = S.wCell, hCell = S.hCell, stCX = S.stCX;
    for (int i = tid; i < nW; i += FAST_THREADS) {
        unsigned fc = 0, lc = 0, vm = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int x = oxa + ((wx0 + i) << 2) + b;
            if (x >= xFirst && x < xEnd) {
                vm |= 0xFFu << (8 * b);
                const int v = x - xFirst;
                const int r = v - (int)(((unsigned)v * divCw) >> 16) * wCell;
                if (r == 0) fc |= 0xFFu << (8 * b);
                if (r == wCell - 1 || x == xEnd - 1) lc |= 0xFFu << (8 * b);
            }
        }
        H->firstCol[i] = fc;
        H->lastCol[i] = lc;
        H->valid[i] = vm;
    }
    for (int i = tid; i < nR; i += FAST_THREADS) {
        const int r = i - (int)(((unsigned)i * divCh) >> 16) * hCell;
        H->rowFlag[i] = (unsigned char)((r == 0 ? 1 : 0) | ((r == hCell - 1 || i == nR - 1) ? 2 : 0));
    }
    {   // zero the score map (its unused rim is what NMS reads as "no corner")
        uint4* z = (uint4*)sc;
        for (int i = tid; i < (tileBytes >> 4); i += FAST_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    {   // bounded wait on phase 0 (a lost TMA must not hang the GPU: flag the error and bail out)
        uint32_t done = 0;
        int spins = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(bar), "r"(0u)
                : "memory");
            if (!done && ++spins > (1 << 22)) {
                atomicExch(err, 2);
                break;
            }
        }
    }

    const int minTh = S.minTh, iniTh = S.iniTh;
    const unsigned negThm1 = (unsigned)((-(minTh - 1)) & 0xFFFF) * 0x10001u;
    const int nUnits = nW * nR;
    const unsigned divW = (1048576u + nW - 1) / nW;  // u / nW == (u * divW) >> 20 for u < 16384

    // ---- phase 1: scores
    if constexpr (V == 1)
    for (int u = tid; u < nUnits; u += FAST_THREADS) {
        const int wy = (int)(((unsigned)u * divW) >> 20), wi = u - wy * nW;
        const int wx = wx0 + wi;
        const unsigned* r0 = tile + (wy + 3) * pitchW + wx;  // centre row
        // Operands stay 16-bit lanes with the PIXEL IN THE HIGH BYTE and a neighbouring pixel as junk in the low byte:
        // the high byte of any min/max tree over such lanes is the same tree over the high bytes (lexicographic
        // order), so no masking is needed. With A(d) = bytes x+d .. x+d+3 of a row, the operand for pixels (1,3) at
        // circle offset dx is A(dx) itself and the one for pixels (0,2) is A(dx-1). Where only the high bytes matter a
        // plain shift of one word does (A(-1) ~ b << 8, A(3) ~ c << 8): IMAD.SHL runs on the idle FMA pipe.
        // Bresenham order (dx,dy): (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
        unsigned ce[16], co[16];
        {
            const unsigned* r = r0 + 3 * pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            const unsigned m2 = fast_align(a, b, c, -2), m1 = b << 8, p1 = fast_align(a, b, c, 1);
            ce[15] = m2; co[15] = m1; ce[0] = m1; co[0] = b; ce[1] = b; co[1] = p1;
        }
        {
            const unsigned* r = r0 + 2 * pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            ce[14] = fast_align(a, b, c, -3); co[14] = fast_align(a, b, c, -2);
            ce[2] = fast_align(a, b, c, 1); co[2] = fast_align(a, b, c, 2);
        }
        {
            const unsigned* r = r0 + pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            ce[13] = a; co[13] = fast_align(a, b, c, -3);
            ce[3] = fast_align(a, b, c, 2); co[3] = c << 8;
        }
        unsigned ve, vo;
        {
            const unsigned a = r0[-1], b = r0[0], c = r0[1];
            ce[12] = a; co[12] = fast_align(a, b, c, -3);
            ce[4] = fast_align(a, b, c, 2); co[4] = c << 8;
            ve = b << 8; vo = b;
        }
        {
            const unsigned* r = r0 - pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            ce[11] = a; co[11] = fast_align(a, b, c, -3);
            ce[5] = fast_align(a, b, c, 2); co[5] = c << 8;
        }
        {
            const unsigned* r = r0 - 2 * pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            ce[10] = fast_align(a, b, c, -3); co[10] = fast_align(a, b, c, -2);
            ce[6] = fast_align(a, b, c, 1); co[6] = fast_align(a, b, c, 2);
        }
        {
            const unsigned* r = r0 - 3 * pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            const unsigned m2 = fast_align(a, b, c, -2), m1 = b << 8, p1 = fast_align(a, b, c, 1);
            ce[9] = m2; co[9] = m1; ce[8] = m1; co[8] = b; ce[7] = b; co[7] = p1;
        }
        const unsigned se = fast_score2<NF1, NF2>(ce, ve, negThm1);
        const unsigned so = fast_score2<NF1, NF2>(co, vo, negThm1);
        const unsigned s = (se | (so << 8)) & H->valid[wi];  // drop bytes outside the detection columns
        sc[(wy + 3) * pitchW + wx] = s;
        if (s) {  // remember the words that hold a corner: phase 2 visits only those (warp-aggregated append)
            const unsigned m = __activemask();
            const int lane = tid & 31, leader = __ffs(m) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(&H->nAct, __popc(m));
            base = __shfl_sync(m, base, leader);
            act[base + __popc(m & ((1u << lane) - 1u))] = (unsigned short)u;
        }
    }
    if constexpr (V == 3) {
        const int Q2 = (nR + 1) >> 1;  // rows per band; lane 0 of a word = row y of band 0, lane 1 = row y + Q2
        const int wxE0 = (xFirst - 3 - oxa) >> 2, wxE1 = (xEnd + 2 - oxa) >> 2;  // tile words that hold columns xFirst-3 .. xEnd+2
        const int nWE = wxE1 - wxE0 + 1;
        const int EP = (nWE << 2) <= FAST3_EP && FAST3_EP <= boxW ? FAST3_EP : (nWE << 2);  // expanded pitch in words (= pixels)
        unsigned* ex = (unsigned*)(fast_smem + 2 * tileBytes);
        {   // expansion: 4 columns x 2 bands per item, one 16-byte store
            const int nItems = nWE * (Q2 + 6);
            int rr = tid / nWE, wi = tid - rr * nWE;
            const int rS = FAST_THREADS / nWE, wS = FAST_THREADS - rS * nWE;
            for (int i = tid; i < nItems; i += FAST_THREADS) {
                const unsigned w0 = tile[rr * pitchW + wxE0 + wi];
                const unsigned w1 = tile[min(rr + Q2, boxH - 1) * pitchW + wxE0 + wi];
                const unsigned lo = __byte_perm(w0, w1, 0x5140), hi = __byte_perm(w0, w1, 0x7362);
                *(uint4*)(ex + rr * EP + (wi << 2)) =
                    make_uint4(__byte_perm(lo, 0x64646464u, 0x4140), __byte_perm(lo, 0x64646464u, 0x4342),
                               __byte_perm(hi, 0x64646464u, 0x4140), __byte_perm(hi, 0x64646464u, 0x4342));
                rr += rS;
                wi += wS;
                if (wi >= nWE) { wi -= nWE; rr++; }
            }
        }
        __syncthreads();
        {
            const int colOff = xFirst - (oxa + (wxE0 << 2));  // >= 3
            const unsigned Cth = (unsigned)__half_as_ushort(__int2half_rn(1024 - minTh)) * 0x10001u;
            unsigned char* scb = (unsigned char*)sc + 3 * boxW + (xFirst - oxa);
            if (EP == FAST3_EP) fast3_scores<NF1, NF2, FAST3_EP>(ex + 3 * EP + colOff, EP, scb, boxW, xEnd - xFirst, Q2, nR, Cth);
            else fast3_scores<NF1, NF2, 0>(ex + 3 * EP + colOff, EP, scb, boxW, xEnd - xFirst, Q2, nR, Cth);
        }
        __syncthreads();  // the word list overlays the expanded tile
        {   // words that hold a corner: a lane looks at four words (one LDS.128); warp-uniform trip count, plain ballots
            const int pitchQ = pitchW >> 2, nQ = nR * pitchQ, lane = tid & 31;
            const unsigned divQ = (1048576u + pitchQ - 1) / pitchQ, lt = (1u << lane) - 1u;
            const uint4* sc4 = (const uint4*)(sc + 3 * pitchW);
            for (int qb = tid & ~31; qb < nQ; qb += FAST_THREADS) {
                const int q = qb + lane;
                const uint4 w = q < nQ ? sc4[q] : make_uint4(0u, 0u, 0u, 0u);
                const unsigned m0 = __ballot_sync(0xFFFFFFFFu, w.x != 0), m1 = __ballot_sync(0xFFFFFFFFu, w.y != 0);
                const unsigned m2 = __ballot_sync(0xFFFFFFFFu, w.z != 0), m3 = __ballot_sync(0xFFFFFFFFu, w.w != 0);
                if (m0 | m1 | m2 | m3) {
                    const int n0 = __popc(m0), n1 = n0 + __popc(m1), n2 = n1 + __popc(m2);
                    int base = 0;
                    if (lane == 0) base = atomicAdd(&H->nAct, n2 + __popc(m3));
                    base = __shfl_sync(0xFFFFFFFFu, base, 0);
                    const int wy = (int)(((unsigned)q * divQ) >> 20);
                    const int u0 = wy * nW + ((q - wy * pitchQ) << 2) - wx0;  // (only words wx0..wx1 can be non-zero)
                    if (w.x) act[base + __popc(m0 & lt)] = (unsigned short)u0;
                    if (w.y) act[base + n0 + __popc(m1 & lt)] = (unsigned short)(u0 + 1);
                    if (w.z) act[base + n1 + __popc(m2 & lt)] = (unsigned short)(u0 + 2);
                    if (w.w) act[base + n2 + __popc(m3 & lt)] = (unsigned short)(u0 + 3);
                }
            }
        }
    }
    __syncthreads();

    // ---- phase 2: NMS (strictly greater than all 8 neighbours of the same cell); survivors are compacted into a
    // list (packed wy<<16 | x<<8 | value) that overlays the no-longer-needed pixel tile
    unsigned* list = tile;
    const int listCap = tileBytes >> 2;
    const int nAct = H->nAct;
    for (int ia = tid; ia < nAct; ia += FAST_THREADS) {
        const int u = act[ia];
        const int wy = (int)(((unsigned)u * divW) >> 20), wi = u - wy * nW;
        const int wx = wx0 + wi;
        const unsigned* q = sc + (wy + 3) * pitchW + wx;
        const unsigned s = q[0];
        const unsigned s7 = s & 0x7f7f7f7fu;
        const unsigned nfc = ~H->firstCol[wi], nlc = ~H->lastCol[wi];
        const unsigned rf = H->rowFlag[wy];
        unsigned ok = bytes_gt(s, s7, __byte_perm(q[-1], s, 0x6543) & nfc) & bytes_gt(s, s7, __byte_perm(s, q[1], 0x4321) & nlc);
        if (!(rf & 1)) {
            const unsigned* r = q - pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            ok &= bytes_gt(s, s7, __byte_perm(a, b, 0x6543) & nfc) & bytes_gt(s, s7, b) &
                  bytes_gt(s, s7, __byte_perm(b, c, 0x4321) & nlc);
        }
        if (!(rf & 2)) {
            const unsigned* r = q + pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            ok &= bytes_gt(s, s7, __byte_perm(a, b, 0x6543) & nfc) & bytes_gt(s, s7, b) &
                  bytes_gt(s, s7, __byte_perm(b, c, 0x4321) & nlc);
        }
        ok &= 0x80808080u;  // (a zero byte is never greater than anything, so non-corners drop out here)
        if (ok) {
            const int n = __popc(ok);
            int pos = atomicAdd(&H->nList, n);
            const int base = (wy << 16) + (((wx << 2) + oxa - xFirst) << 8);  // x of byte 0 may be < 0 (masked bytes)
#pragma unroll
            for (int b = 0; b < 4; b++)
                if ((ok >> (8 * b + 7)) & 1u) {
                    if (pos < listCap) list[pos] = (unsigned)(base + (b << 8) + (int)((s >> (8 * b)) & 0xFFu));
                    pos++;
                }
        }
    }
    __syncthreads();
    // per-cell counts (all survivors / survivors >= iniTh)
    const int nList = min(H->nList, listCap);
    for (int i = tid; i < nList; i += FAST_THREADS) {
        const unsigned r = list[i];
        const unsigned wy = r >> 16, x = (r >> 8) & 0xFF, v = r & 0xFF;
        const int cell = (int)((wy * divCh) >> 16) * stCX + (int)((x * divCw) >> 16);
        atomicAdd(&H->nAll[cell], 1);
        if ((int)v + minTh - 1 >= iniTh) atomicAdd(&H->nIni[cell], 1);
    }
    __syncthreads();

    // ---- phase 3: per-cell rule, one global reservation per CTA, emit
    if (tid == 0) {
        int tot = 0;
        for (int c = 0; c < S.nCells; c++) tot += H->nIni[c] > 0 ? H->nIni[c] : H->nAll[c];
        H->total = tot;
        H->gbase = tot ? atomicAdd(&candCount[f * S.nlevels + level], tot) : 0;
    }
    __syncthreads();
    if (H->total == 0) return;
    const int gbase = H->gbase;
    u64* out = cand + S.candOff + (long long)f * S.candCap;
    for (int i = tid; i < nList; i += FAST_THREADS) {
        const unsigned r = list[i];
        const unsigned wy = r >> 16, x = (r >> 8) & 0xFF, v = r & 0xFF;
        const int cyl = (int)((wy * divCh) >> 16), yin = (int)wy - cyl * hCell;
        const int cxl = (int)((x * divCw) >> 16), xin = (int)x - cxl * wCell;
        const int score = (int)v + minTh - 1;
        if (H->nIni[cyl * stCX + cxl] > 0 && score < iniTh) continue;
        const int slot = gbase + atomicAdd(&H->slot, 1);
        if (slot < S.candCap) out[slot] = cand_pack(score, ci0 + cyl, cj0 + cxl, yin, xin);
        else atomicExch(err, 3);
    }
}
