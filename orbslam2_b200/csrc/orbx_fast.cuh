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
// FMA pipe (also 2/clk/SM) and overlap. One thread scores 4 horizontally adjacent pixels (one 32-bit word of the tile).
//
// One CTA = a "supertile" of stCX x stCY cells of one level of one frame:
//   TMA 3-D tile load (x start rounded down to 16 bytes — an unaligned innermost coordinate faults on sm_100a)
//   phase 1  score map  (u8, value = score - (minTh-1), 0 = not a corner at minTh), same coordinates as the tile
//   phase 2  SWAR non-maximum suppression, strictly greater than the 8 neighbours; a neighbour in another cell
//            counts 0 (the reference runs cv::FAST per cell ROI); survivors compacted into a shared-memory list
//   phase 3  per cell: keep survivors >= iniTh if any exist, else all; one global atomic per CTA; emit records.
#pragma once
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
__device__ __forceinline__ unsigned fast_score2(const unsigned (&c)[16], unsigned vh, unsigned negThm1) {
    unsigned a2[8], b2[8], a4[8], b4[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        a2[j] = __vminu2(c[2 * j], c[2 * j + 1]);
        b2[j] = __vmaxu2(c[2 * j], c[2 * j + 1]);
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
        const unsigned dmin = __vminu2(c[(k + 8) & 15], c[(k + 15) & 15]);
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

__global__ void __launch_bounds__(FAST_THREADS, 4) k_fast(const __grid_constant__ TmapPack maps,
                                                       const DevPlan* __restrict__ P, u64* __restrict__ cand,
                                                       int* __restrict__ candCount, int* __restrict__ err, int f0) {
    extern __shared__ __align__(128) uint8_t fast_smem[];
    const int tid = threadIdx.x;
    const int f = f0 + blockIdx.y;
    int level = 0;
    while (level + 1 < P->nlevels && (int)blockIdx.x >= P->lv[level + 1].stBase) level++;
    const LevelGeo& G = P->lv[level];
    const int st = blockIdx.x - G.stBase;
    const int sy = st / G.nStX, sx = st - sy * G.nStX;
    const int e = P->edge;
    const int cj0 = sx * G.stCX, ci0 = sy * G.stCY;
    const int ncx = min(G.stCX, G.nCols - cj0), ncy = min(G.stCY, G.nRows - ci0);
    const int xFirst = e + cj0 * G.wCell, xEnd = min(e + (cj0 + ncx) * G.wCell, G.w - e);
    const int yFirst = e + ci0 * G.hCell, yEnd = min(e + (ci0 + ncy) * G.hCell, G.h - e);
    if (xEnd <= xFirst || yEnd <= yFirst) return;

    const int boxW = G.boxW, boxH = G.boxH;
    const int pitchW = boxW >> 2;  // tile pitch in words
    const int tileBytes = boxW * boxH;
    unsigned* tile = (unsigned*)fast_smem;
    unsigned* sc = (unsigned*)(fast_smem + tileBytes);
    unsigned short* act = (unsigned short*)(fast_smem + 2 * tileBytes);  // units (words) with a corner, <= tileBytes/4 entries
    FastSmemHdr* H = (FastSmemHdr*)(fast_smem + 2 * tileBytes + (((tileBytes >> 1) + 15) & ~15));
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
    const unsigned divCw = (65536u + G.wCell - 1) / G.wCell;  // v / wCell == (v * divCw) >> 16 for v < 1024
    const unsigned divCh = (65536u + G.hCell - 1) / G.hCell;
    for (int i = tid; i < nW; i += FAST_THREADS) {
        unsigned fc = 0, lc = 0, vm = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int x = oxa + ((wx0 + i) << 2) + b;
            if (x >= xFirst && x < xEnd) {
                vm |= 0xFFu << (8 * b);
                const int v = x - xFirst;
                const int r = v - (int)(((unsigned)v * divCw) >> 16) * G.wCell;
                if (r == 0) fc |= 0xFFu << (8 * b);
                if (r == G.wCell - 1 || x == xEnd - 1) lc |= 0xFFu << (8 * b);
            }
        }
        H->firstCol[i] = fc;
        H->lastCol[i] = lc;
        H->valid[i] = vm;
    }
    for (int i = tid; i < nR; i += FAST_THREADS) {
        const int r = i - (int)(((unsigned)i * divCh) >> 16) * G.hCell;
        H->rowFlag[i] = (unsigned char)((r == 0 ? 1 : 0) | ((r == G.hCell - 1 || i == nR - 1) ? 2 : 0));
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

    const int minTh = P->minTh, iniTh = P->iniTh;
    const unsigned negThm1 = (unsigned)((-(minTh - 1)) & 0xFFFF) * 0x10001u;
    const int nUnits = nW * nR;
    const unsigned divW = (1048576u + nW - 1) / nW;  // u / nW == (u * divW) >> 20 for u < 16384

    // ---- phase 1: scores
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
        const unsigned se = fast_score2(ce, ve, negThm1);
        const unsigned so = fast_score2(co, vo, negThm1);
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
        const int cell = (int)((wy * divCh) >> 16) * G.stCX + (int)((x * divCw) >> 16);
        atomicAdd(&H->nAll[cell], 1);
        if ((int)v + minTh - 1 >= iniTh) atomicAdd(&H->nIni[cell], 1);
    }
    __syncthreads();

    // ---- phase 3: per-cell rule, one global reservation per CTA, emit
    if (tid == 0) {
        int tot = 0;
        for (int c = 0; c < G.stCX * G.stCY; c++) tot += H->nIni[c] > 0 ? H->nIni[c] : H->nAll[c];
        H->total = tot;
        H->gbase = tot ? atomicAdd(&candCount[f * P->nlevels + level], tot) : 0;
    }
    __syncthreads();
    if (H->total == 0) return;
    const int gbase = H->gbase;
    u64* out = cand + G.candOff + (long long)f * G.candCap;
    for (int i = tid; i < nList; i += FAST_THREADS) {
        const unsigned r = list[i];
        const unsigned wy = r >> 16, x = (r >> 8) & 0xFF, v = r & 0xFF;
        const int cyl = (int)((wy * divCh) >> 16), yin = (int)wy - cyl * G.hCell;
        const int cxl = (int)((x * divCw) >> 16), xin = (int)x - cxl * G.wCell;
        const int score = (int)v + minTh - 1;
        if (H->nIni[cyl * G.stCX + cxl] > 0 && score < iniTh) continue;
        const int slot = gbase + atomicAdd(&H->slot, 1);
        if (slot < G.candCap) out[slot] = cand_pack(score, ci0 + cyl, cj0 + cxl, yin, xin);
        else atomicExch(err, 3);
    }
}
