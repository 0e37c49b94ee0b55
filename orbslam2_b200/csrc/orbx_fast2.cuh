// orbx_fast2.cuh — K2 k_fast2: per-cell FAST-9/16 + cell-local NMS with the reference's two-threshold structure.
// Reference: ORBextractor::ComputeKeyPointsOctTree first half, /root/reference/src/ORBextractor.cc:759-823:
//     FAST(cell, vKeysCell, iniThFAST, true);  if (vKeysCell.empty()) FAST(cell, vKeysCell, minThFAST, true);
//
// k_fast (orbx_fast.cuh) computes the threshold-free score of EVERY pixel (2 x 79 packed min/max instructions per 4 pixels)
// and is bound by the ALU pipe. Only ~15 % of the pixels can be corners at iniThFAST and ~16 % of the cells ever need the
// minThFAST pass, so this kernel does what the reference does, in two passes over a supertile held in shared memory:
//   pass 1 (threshold iniTh, all words)          pass 2 (threshold minTh, only words touching a cell that pass 1 left empty)
//     A  pre-test: a run of 9 on the circle contains one pixel of every opposite pair, so a bright corner needs
//        min(max(c0,c8), max(c4,c12)) > v + t and a dark one max(min(c0,c8), min(c4,c12)) < v - t. Evaluated for the two
//        pixel pairs of a word (even / odd bytes, packed 16-bit lanes) and both polarities: ~30 ALU instructions per
//        word. A (word, parity, polarity) that passes is appended to one of four work lists.
//     B  the work lists are processed densely (no divergence: one loop per parity x polarity): the 36-instruction arc
//        tree of ONE polarity for ONE pixel pair, giving the exact threshold-free score (value = score - (minTh-1), as
//        k_fast stores it); OR-ed into the score map (a pixel cannot be a bright and a dark corner at once).
//     NMS pass 1: only pixels >= iniTh (their >= iniTh neighbours are all present, lower ones cannot suppress them);
//         survivors are emitted at once; cells without a survivor form the retry mask.
//     NMS pass 2: every scored pixel of a retry cell (all of its same-cell neighbours >= minTh are present by then).
// Results (candidate set and scores) are identical to k_fast's and to the oracle's; tests/test_gpu_extract.py.
#pragma once
#include "orbx_fast.cuh"

struct Fast2Hdr {
    unsigned long long mbar;
    int nIni[FAST_MAX_ST_CELLS];
    int nItem[4];  // [parity * 2 + polarity]
    int nAct, nList, slot, gbase;
    unsigned retryMask;
    unsigned firstCol[64], lastCol[64], valid[64];
    unsigned split[64];            // bytes of the word that belong to the NEXT cell column (0 if the word sits in one cell)
    unsigned char colCell[64];     // cell column (inside the supertile) of the word's first detection byte
    unsigned char rowFlag[128];
    unsigned char rowCell[128];    // cell row (inside the supertile) of a detection row
};

// circle operands of one pixel pair of word r0[0]: PAR 0 = pixels (0,2), PAR 1 = pixels (1,3); 16-bit lanes, pixel in the
// HIGH byte (see k_fast for the byte algebra)
template <int PAR>
__device__ __forceinline__ void fast2_circle(const unsigned* r0, int pitchW, unsigned (&c)[16], unsigned& vh) {
    {
        const unsigned* r = r0 + 3 * pitchW;
        const unsigned a = r[-1], b = r[0], cc = r[1];
        if (PAR == 0) { c[15] = fast_align(a, b, cc, -2); c[0] = b << 8; c[1] = b; }
        else { c[15] = b << 8; c[0] = b; c[1] = fast_align(a, b, cc, 1); }
    }
    {
        const unsigned* r = r0 + 2 * pitchW;
        const unsigned a = r[-1], b = r[0], cc = r[1];
        if (PAR == 0) { c[14] = fast_align(a, b, cc, -3); c[2] = fast_align(a, b, cc, 1); }
        else { c[14] = fast_align(a, b, cc, -2); c[2] = fast_align(a, b, cc, 2); }
    }
    {
        const unsigned* r = r0 + pitchW;
        const unsigned a = r[-1], b = r[0], cc = r[1];
        if (PAR == 0) { c[13] = a; c[3] = fast_align(a, b, cc, 2); }
        else { c[13] = fast_align(a, b, cc, -3); c[3] = cc << 8; }
    }
    {
        const unsigned a = r0[-1], b = r0[0], cc = r0[1];
        if (PAR == 0) { c[12] = a; c[4] = fast_align(a, b, cc, 2); vh = b << 8; }
        else { c[12] = fast_align(a, b, cc, -3); c[4] = cc << 8; vh = b; }
    }
    {
        const unsigned* r = r0 - pitchW;
        const unsigned a = r[-1], b = r[0], cc = r[1];
        if (PAR == 0) { c[11] = a; c[5] = fast_align(a, b, cc, 2); }
        else { c[11] = fast_align(a, b, cc, -3); c[5] = cc << 8; }
    }
    {
        const unsigned* r = r0 - 2 * pitchW;
        const unsigned a = r[-1], b = r[0], cc = r[1];
        if (PAR == 0) { c[10] = fast_align(a, b, cc, -3); c[6] = fast_align(a, b, cc, 1); }
        else { c[10] = fast_align(a, b, cc, -2); c[6] = fast_align(a, b, cc, 2); }
    }
    {
        const unsigned* r = r0 - 3 * pitchW;
        const unsigned a = r[-1], b = r[0], cc = r[1];
        if (PAR == 0) { c[9] = fast_align(a, b, cc, -2); c[8] = b << 8; c[7] = b; }
        else { c[9] = b << 8; c[8] = b; c[7] = fast_align(a, b, cc, 1); }
    }
}

// one polarity of fast_score2: POL 0 = bright (max over the 16 arcs of the arc minimum), POL 1 = dark.
// Returns max(score - (minTh-1), 0) per 16-bit lane.
template <int POL>
__device__ __forceinline__ unsigned fast2_score1(const unsigned (&c)[16], unsigned vh, unsigned negThm1) {
    unsigned a2[8], a4[8], r[8];
#pragma unroll
    for (int j = 0; j < 8; j++) a2[j] = POL == 0 ? __vminu2(c[2 * j], c[2 * j + 1]) : __vmaxu2(c[2 * j], c[2 * j + 1]);
#pragma unroll
    for (int j = 0; j < 8; j++) a4[j] = POL == 0 ? __vminu2(a2[j], a2[(j + 1) & 7]) : __vmaxu2(a2[j], a2[(j + 1) & 7]);
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int k = 2 * j;
        if (POL == 0) r[j] = __vimin3_u16x2(a4[j], a4[(j + 2) & 7], __vmaxu2(c[(k + 8) & 15], c[(k + 15) & 15]));
        else r[j] = __vimax3_u16x2(a4[j], a4[(j + 2) & 7], __vminu2(c[(k + 8) & 15], c[(k + 15) & 15]));
    }
    unsigned s;
    if (POL == 0) {
        const unsigned Mh = __vimax3_u16x2(__vimax3_u16x2(r[0], r[1], r[2]), __vimax3_u16x2(r[3], r[4], r[5]), __vmaxu2(r[6], r[7]));
        s = __viaddmax_s16x2(__byte_perm(Mh, 0u, 0x4341), ~__byte_perm(vh, 0u, 0x4341), 0u);   // (M - v) - 1
    } else {
        const unsigned mh = __vimin3_u16x2(__vimin3_u16x2(r[0], r[1], r[2]), __vimin3_u16x2(r[3], r[4], r[5]), __vminu2(r[6], r[7]));
        s = __viaddmax_s16x2(__byte_perm(vh, 0u, 0x4341), ~__byte_perm(mh, 0u, 0x4341), 0u);   // (v - m) - 1
    }
    return __viaddmax_s16x2(s, negThm1, 0u);
}

// warp-aggregated append of `val` to a shared-memory u16 list (called under divergence: uses the active mask)
__device__ __forceinline__ void fast2_append(unsigned short* list, int* counter, unsigned short val) {
    const unsigned m = __activemask();
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(m, base, leader);
    list[base + __popc(m & ((1u << lane) - 1u))] = val;
}

struct Fast2Ctx {
    const unsigned* tile;
    unsigned* sc;
    unsigned short* act;
    unsigned short* items;  // four lists of `cap` entries
    Fast2Hdr* H;
    int pitchW, wx0, nW, nUnits, cap, stCX;
    unsigned divW;
};

template <int PAR, int POL>
__device__ __forceinline__ void fast2_score_items(const Fast2Ctx& C, unsigned negThm1) {
    const unsigned short* list = C.items + (PAR * 2 + POL) * C.cap;
    const int n = C.H->nItem[PAR * 2 + POL];
    for (int it = threadIdx.x; it < n; it += FAST_THREADS) {
        const int u = list[it];
        const int wy = (int)(((unsigned)u * C.divW) >> 20), wi = u - wy * C.nW;
        const int idx = (wy + 3) * C.pitchW + C.wx0 + wi;
        unsigned c[16], vh;
        fast2_circle<PAR>(C.tile + idx, C.pitchW, c, vh);
        const unsigned v = fast2_score1<POL>(c, vh, negThm1);
        const unsigned word = (PAR ? (v << 8) : v) & C.H->valid[wi];
        if (word) {
            const unsigned old = atomicOr(&C.sc[idx], word);
            if (old == 0) fast2_append(C.act, &C.H->nAct, (unsigned short)u);
        }
    }
}

// pre-test of all words (PASS 1: at iniTh) or of the words touching a retry cell (PASS 2: at minTh, skipping what pass 1 scored)
template <int PASS>
__device__ __forceinline__ void fast2_pretest(const Fast2Ctx& C, unsigned negIniM1, unsigned negMinM1, unsigned retryMask) {
    Fast2Hdr* H = C.H;
    for (int u = threadIdx.x; u < C.nUnits; u += FAST_THREADS) {
        const int wy = (int)(((unsigned)u * C.divW) >> 20), wi = u - wy * C.nW;
        if (PASS == 2) {
            const int cell = H->rowCell[wy] * C.stCX + H->colCell[wi];
            unsigned touch = (retryMask >> cell) & 1u;
            if (H->split[wi]) touch |= (retryMask >> (cell + 1)) & 1u;
            if (!touch) continue;
        }
        const unsigned* r0 = C.tile + (wy + 3) * C.pitchW + C.wx0 + wi;
        const unsigned up = r0[3 * C.pitchW], dn = r0[-3 * C.pitchW];
        const unsigned a = r0[-1], b = r0[0], c = r0[1];
#pragma unroll
        for (int par = 0; par < 2; par++) {
            const unsigned c0 = par ? up : (up << 8), c8 = par ? dn : (dn << 8);
            const unsigned c12 = par ? fast_align(a, b, c, -3) : a, c4 = par ? (c << 8) : fast_align(a, b, c, 2);
            const unsigned vh = par ? b : (b << 8);
            const unsigned hb = __vminu2(__vmaxu2(c0, c8), __vmaxu2(c4, c12));
            const unsigned hd = __vmaxu2(__vminu2(c0, c8), __vminu2(c4, c12));
            const unsigned v = __byte_perm(vh, 0u, 0x4341);
            const unsigned sb = __viaddmax_s16x2(__byte_perm(hb, 0u, 0x4341), ~v, 0u);
            const unsigned sd = __viaddmax_s16x2(v, ~__byte_perm(hd, 0u, 0x4341), 0u);
            bool fb, fd;
            if (PASS == 1) {
                fb = __viaddmax_s16x2(sb, negIniM1, 0u) != 0u;
                fd = __viaddmax_s16x2(sd, negIniM1, 0u) != 0u;
            } else {
                fb = __viaddmax_s16x2(sb, negMinM1, 0u) != 0u && __viaddmax_s16x2(sb, negIniM1, 0u) == 0u;
                fd = __viaddmax_s16x2(sd, negMinM1, 0u) != 0u && __viaddmax_s16x2(sd, negIniM1, 0u) == 0u;
            }
            if (fb) fast2_append(C.items + (par * 2 + 0) * C.cap, &H->nItem[par * 2 + 0], (unsigned short)u);
            if (fd) fast2_append(C.items + (par * 2 + 1) * C.cap, &H->nItem[par * 2 + 1], (unsigned short)u);
        }
    }
}

// NMS over the words in C.act (see k_fast phase 2). PASS 1 keeps pixels >= iniTh only; PASS 2 keeps the pixels that lie in
// a retry cell. Survivors go to `list` as (wy << 16 | x << 8 | value), x relative to the supertile's first column.
template <int PASS>
__device__ __forceinline__ void fast2_nms(const Fast2Ctx& C, unsigned* list, int listCap, unsigned iniByteM1, unsigned retryMask,
                                          int oxa, int xFirst) {
    Fast2Hdr* H = C.H;
    const int nAct = H->nAct;
    for (int ia = threadIdx.x; ia < nAct; ia += FAST_THREADS) {
        const int u = C.act[ia];
        const int wy = (int)(((unsigned)u * C.divW) >> 20), wi = u - wy * C.nW;
        const int wx = C.wx0 + wi;
        const unsigned* q = C.sc + (wy + 3) * C.pitchW + wx;
        const unsigned s = q[0];
        const unsigned s7 = s & 0x7f7f7f7fu;
        const unsigned nfc = ~H->firstCol[wi], nlc = ~H->lastCol[wi];
        const unsigned rf = H->rowFlag[wy];
        unsigned ok = bytes_gt(s, s7, __byte_perm(q[-1], s, 0x6543) & nfc) & bytes_gt(s, s7, __byte_perm(s, q[1], 0x4321) & nlc);
        if (!(rf & 1)) {
            const unsigned* r = q - C.pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            ok &= bytes_gt(s, s7, __byte_perm(a, b, 0x6543) & nfc) & bytes_gt(s, s7, b) &
                  bytes_gt(s, s7, __byte_perm(b, c, 0x4321) & nlc);
        }
        if (!(rf & 2)) {
            const unsigned* r = q + C.pitchW;
            const unsigned a = r[-1], b = r[0], c = r[1];
            ok &= bytes_gt(s, s7, __byte_perm(a, b, 0x6543) & nfc) & bytes_gt(s, s7, b) &
                  bytes_gt(s, s7, __byte_perm(b, c, 0x4321) & nlc);
        }
        if (PASS == 1) ok &= bytes_gt(s, s7, iniByteM1);  // value > iniTh - minTh  <=>  score >= iniTh
        else {
            const int cell = H->rowCell[wy] * C.stCX + H->colCell[wi];
            const unsigned sp = H->split[wi];
            ok &= (((retryMask >> cell) & 1u) ? ~sp : 0u) | ((sp && ((retryMask >> (cell + 1)) & 1u)) ? sp : 0u);
        }
        ok &= 0x80808080u;
        if (ok) {
            const int n = __popc(ok);
            int pos = atomicAdd(&H->nList, n);
            const int base = (wy << 16) + (((wx << 2) + oxa - xFirst) << 8);
#pragma unroll
            for (int b = 0; b < 4; b++)
                if ((ok >> (8 * b + 7)) & 1u) {
                    if (pos < listCap) list[pos] = (unsigned)(base + (b << 8) + (int)((s >> (8 * b)) & 0xFFu));
                    pos++;
                }
        }
    }
}

__global__ void __launch_bounds__(FAST_THREADS, 4) k_fast2(const __grid_constant__ TmapPack maps,
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
    const int pitchW = boxW >> 2;
    const int tileBytes = boxW * boxH;
    const int cap = G.fastUnits;  // >= nW * nR of any supertile of this level
    unsigned* tile = (unsigned*)fast_smem;
    unsigned* sc = (unsigned*)(fast_smem + tileBytes);
    unsigned short* act = (unsigned short*)(fast_smem + 2 * tileBytes);
    unsigned short* items = (unsigned short*)(fast_smem + 2 * tileBytes + ((2 * cap + 15) & ~15));
    Fast2Hdr* H = (Fast2Hdr*)(fast_smem + 2 * tileBytes + ((2 * cap + 15) & ~15) + 8 * cap);
    unsigned* list = (unsigned*)items;   // survivor records overlay the work lists (free once a pass has been scored)
    const int listCap = 2 * cap;         // 8*cap bytes / 4
    const int oxa = (xFirst - 6) & ~15, oy = yFirst - 3;
    const int wx0 = (xFirst - oxa) >> 2, wx1 = (xEnd - 1 - oxa) >> 2;
    const int nW = wx1 - wx0 + 1, nR = yEnd - yFirst;

    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&H->mbar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)tileBytes) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
            ::"r"((uint32_t)__cvta_generic_to_shared(tile)), "l"(&maps.m[level]), "r"(oxa), "r"(oy), "r"(f), "r"(bar)
            : "memory");
        H->slot = 0; H->nList = 0; H->nAct = 0; H->retryMask = 0;
        H->nItem[0] = H->nItem[1] = H->nItem[2] = H->nItem[3] = 0;
    }
    if (tid < FAST_MAX_ST_CELLS) H->nIni[tid] = 0;
    const unsigned divCw = (65536u + G.wCell - 1) / G.wCell;
    const unsigned divCh = (65536u + G.hCell - 1) / G.hCell;
    for (int i = tid; i < nW; i += FAST_THREADS) {
        unsigned fc = 0, lc = 0, vm = 0, sp = 0;
        int c0 = -1;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int x = oxa + ((wx0 + i) << 2) + b;
            if (x >= xFirst && x < xEnd) {
                vm |= 0xFFu << (8 * b);
                const int v = x - xFirst;
                const int cx = (int)(((unsigned)v * divCw) >> 16);
                const int r = v - cx * G.wCell;
                if (c0 < 0) c0 = cx;
                else if (cx != c0) sp |= 0xFFu << (8 * b);
                if (r == 0) fc |= 0xFFu << (8 * b);
                if (r == G.wCell - 1 || x == xEnd - 1) lc |= 0xFFu << (8 * b);
            }
        }
        H->firstCol[i] = fc; H->lastCol[i] = lc; H->valid[i] = vm; H->split[i] = sp;
        H->colCell[i] = (unsigned char)max(c0, 0);
    }
    for (int i = tid; i < nR; i += FAST_THREADS) {
        const int cy = (int)(((unsigned)i * divCh) >> 16);
        const int r = i - cy * G.hCell;
        H->rowFlag[i] = (unsigned char)((r == 0 ? 1 : 0) | ((r == G.hCell - 1 || i == nR - 1) ? 2 : 0));
        H->rowCell[i] = (unsigned char)cy;
    }
    {
        uint4* z = (uint4*)sc;
        for (int i = tid; i < (tileBytes >> 4); i += FAST_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    {
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
    const unsigned negMinM1 = (unsigned)((-(minTh - 1)) & 0xFFFF) * 0x10001u;
    const unsigned negIniM1 = (unsigned)((-(iniTh - 1)) & 0xFFFF) * 0x10001u;
    const unsigned iniByteM1 = (unsigned)min(max(iniTh - minTh, 0), 255) * 0x01010101u;  // value > this  <=>  score >= iniTh
    Fast2Ctx C;
    C.tile = tile; C.sc = sc; C.act = act; C.items = items; C.H = H;
    C.pitchW = pitchW; C.wx0 = wx0; C.nW = nW; C.nUnits = nW * nR; C.cap = cap; C.stCX = G.stCX;
    C.divW = (1048576u + nW - 1) / nW;
    u64* out = cand + G.candOff + (long long)f * G.candCap;
    int* gcount = &candCount[f * P->nlevels + level];

    // ================= pass 1: iniTh
    fast2_pretest<1>(C, negIniM1, negMinM1, 0u);
    __syncthreads();
    fast2_score_items<0, 0>(C, negMinM1);
    fast2_score_items<0, 1>(C, negMinM1);
    fast2_score_items<1, 0>(C, negMinM1);
    fast2_score_items<1, 1>(C, negMinM1);
    __syncthreads();
    fast2_nms<1>(C, list, listCap, iniByteM1, 0u, oxa, xFirst);
    __syncthreads();
    int nList = min(H->nList, listCap);
    if (tid == 0) H->gbase = nList ? atomicAdd(gcount, nList) : 0;
    __syncthreads();
    for (int i = tid; i < nList; i += FAST_THREADS) {
        const unsigned r = list[i];
        const unsigned wy = r >> 16, x = (r >> 8) & 0xFF, v = r & 0xFF;
        const int cyl = (int)((wy * divCh) >> 16), yin = (int)wy - cyl * G.hCell;
        const int cxl = (int)((x * divCw) >> 16), xin = (int)x - cxl * G.wCell;
        atomicAdd(&H->nIni[cyl * G.stCX + cxl], 1);
        const int slot = H->gbase + i;
        if (slot < G.candCap) out[slot] = cand_pack((int)v + minTh - 1, ci0 + cyl, cj0 + cxl, yin, xin);
        else atomicExch(err, 3);
    }
    __syncthreads();
    if (tid == 0) {
        unsigned m = 0;
        for (int cy = 0; cy < ncy; cy++)
            for (int cx = 0; cx < ncx; cx++)
                if (H->nIni[cy * G.stCX + cx] == 0) m |= 1u << (cy * G.stCX + cx);
        H->retryMask = m;
        H->nItem[0] = H->nItem[1] = H->nItem[2] = H->nItem[3] = 0;
        H->nList = 0;
    }
    __syncthreads();
    const unsigned retryMask = H->retryMask;
    if (retryMask == 0u) return;

    // ================= pass 2: minTh, only around the cells pass 1 left empty
    fast2_pretest<2>(C, negIniM1, negMinM1, retryMask);
    __syncthreads();
    fast2_score_items<0, 0>(C, negMinM1);
    fast2_score_items<0, 1>(C, negMinM1);
    fast2_score_items<1, 0>(C, negMinM1);
    fast2_score_items<1, 1>(C, negMinM1);
    __syncthreads();
    if (tid == 0) H->nAct = 0;
    __syncthreads();
    for (int u = tid; u < C.nUnits; u += FAST_THREADS) {   // every scored word that touches a retry cell
        const int wy = (int)(((unsigned)u * C.divW) >> 20), wi = u - wy * nW;
        const int cell = H->rowCell[wy] * G.stCX + H->colCell[wi];
        unsigned touch = (retryMask >> cell) & 1u;
        if (H->split[wi]) touch |= (retryMask >> (cell + 1)) & 1u;
        if (touch && sc[(wy + 3) * pitchW + wx0 + wi]) fast2_append(act, &H->nAct, (unsigned short)u);
    }
    __syncthreads();
    fast2_nms<2>(C, list, listCap, iniByteM1, retryMask, oxa, xFirst);
    __syncthreads();
    nList = min(H->nList, listCap);
    if (nList == 0) return;
    if (tid == 0) H->gbase = atomicAdd(gcount, nList);
    __syncthreads();
    for (int i = tid; i < nList; i += FAST_THREADS) {
        const unsigned r = list[i];
        const unsigned wy = r >> 16, x = (r >> 8) & 0xFF, v = r & 0xFF;
        const int cyl = (int)((wy * divCh) >> 16), yin = (int)wy - cyl * G.hCell;
        const int cxl = (int)((x * divCw) >> 16), xin = (int)x - cxl * G.wCell;
        const int slot = H->gbase + i;
        if (slot < G.candCap) out[slot] = cand_pack((int)v + minTh - 1, ci0 + cyl, cj0 + cxl, yin, xin);
        else atomicExch(err, 3);
    }
}
