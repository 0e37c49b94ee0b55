// orbx_kernels.cuh — extraction kernels (sm_100a). Included once by orbx_extract.cu.
//
// K1 k_resize      bilinear 8U pyramid level from the previous level       (ref: src/ORBextractor.cc:921-946 + cv::resize)
// K2 k_fast        per-cell FAST-9/16 score + cell-local NMS + ini/min retry (ref: :759-823 + cv::FAST); orbx_fast.cuh
// K3 k_quadtree    DistributeOctTree as a level-synchronous CTA algorithm    (ref: :475-757)
// K4 k_blur        7x7 sigma-2 fixed-point Gaussian                          (ref: :899-900 + cv::GaussianBlur)
// K5 k_orient_desc IC_Angle + steered BRIEF + keypoint assembly              (ref: :72-142, :835-841, :909-917)
// K6 k_export      padded (reflect-101) copy of one level                    (ref: :936-942 + cv::copyMakeBorder)
#pragma once
#include "orbx_internal.cuh"
#include "../../include/orbx_pattern.inc"


// ------------------------------------------------------------------------------------------------ helpers
__device__ __forceinline__ int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
    return p;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Exclusive scan in place over a[0..m) in shared memory; every thread of the block must call; returns the total.
// ws: 33 ints of shared scratch.
__device__ int block_excl_scan(int* a, int m, int* ws) {
    const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = T >> 5;
    int carry = 0;
    for (int base = 0; base < m; base += T) {
        const int i = base + tid;
        const int v = (i < m) ? a[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) ws[wid] = x;
        __syncthreads();
        if (wid == 0) {
            const int s = (lane < nw) ? ws[lane] : 0;
            int t = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t += y;
            }
            ws[lane] = t - s;
            if (lane == 31) ws[32] = t;
        }
        __syncthreads();
        if (i < m) a[i] = carry + ws[wid] + x - v;
        carry += ws[32];
        __syncthreads();
    }
    return carry;
}

// Exclusive scan of a 0/1 predicate over i in [0, m): a[i] = number of true predicates before i; returns the total. Every
// thread of the block must call. Ballot + popc inside a warp, one barrier per blockDim-sized chunk for the warp totals
// (ws: 2 x 32 ints, alternated between chunks and calls by *phase, so no barrier is needed to recycle them): the quadtree
// sweeps are bound by their barriers, and three of their four scans are predicate scans (block_excl_scan costs three
// barriers a chunk plus one for the flags it reads).
template <typename F>
__device__ __forceinline__ int block_flag_scan(int* a, int m, int* ws, int* phase, F pred) {
    const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = T >> 5;
    int carry = 0;
    for (int base = 0; base < m; base += T) {
        const int i = base + tid;
        const bool p = i < m && pred(i);
        const unsigned bal = __ballot_sync(0xffffffffu, p);
        int* w = ws + ((*phase) & 1) * 32;
        (*phase)++;
        if (lane == 0) w[wid] = __popc(bal);
        __syncthreads();
        const int wt = lane < nw ? w[lane] : 0;
        const int pre = __reduce_add_sync(0xffffffffu, lane < wid ? wt : 0);
        const int tot = __reduce_add_sync(0xffffffffu, wt);
        if (i < m) a[i] = carry + pre + __popc(bal & ((1u << lane) - 1u));
        carry += tot;
    }
    return carry;
}

// The same for values in [0, 2^BITS): one ballot per bit.
template <int BITS, typename F>
__device__ __forceinline__ int block_small_scan(int* a, int m, int* ws, int* phase, F val) {
    const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = T >> 5;
    const unsigned lt = (1u << lane) - 1u;
    int carry = 0;
    for (int base = 0; base < m; base += T) {
        const int i = base + tid;
        const int v = i < m ? val(i) : 0;
        int inw = 0, wsum = 0;
#pragma unroll
        for (int b = 0; b < BITS; b++) {
            const unsigned bal = __ballot_sync(0xffffffffu, (v >> b) & 1);
            inw += __popc(bal & lt) << b;
            wsum += __popc(bal) << b;
        }
        int* w = ws + ((*phase) & 1) * 32;
        (*phase)++;
        if (lane == 0) w[wid] = wsum;
        __syncthreads();
        const int wt = lane < nw ? w[lane] : 0;
        const int pre = __reduce_add_sync(0xffffffffu, lane < wid ? wt : 0);
        const int tot = __reduce_add_sync(0xffffffffu, wt);
        if (i < m) a[i] = carry + pre + inw;
        carry += tot;
    }
    return carry;
}

// Ascending bitonic sort of a[0..P2) (P2 a power of two) in shared memory.
__device__ void block_bitonic_sort(u64* a, int P2) {
    for (int k = 2; k <= P2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < P2; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const bool up = ((i & k) == 0);
                    const u64 x = a[i], y = a[ixj];
                    if ((x > y) == up) {
                        a[i] = y;
                        a[ixj] = x;
                    }
                }
            }
            __syncthreads();
        }
}

__device__ __forceinline__ int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

// ------------------------------------------------------------------------------------------------ K1 resize
// OpenCV 8U INTER_LINEAR fixed point: H = S[s]*a0 + S[s+1]*a1 (scale 2^11); out = ((b0*(H0>>4))>>16 + (b1*(H1>>4))>>16 + 2)>>2.
// Tables (host-computed with OpenCV's float/double expressions): int2 {PRMT selector | source index << 16, a0 | a1<<16}; the y table holds
// {row0 | row1 << 16, b0 | b1 << 16} with both source rows already clamped to the image.
// One thread = 4 consecutive output pixels (one aligned 32-bit store).
__global__ void __launch_bounds__(256) k_resize(const __grid_constant__ DevPlan Pc, int level, uint8_t* __restrict__ pyr,
                                                const int2* __restrict__ tabs, int f0) {
    const DevPlan* P = &Pc;  // the plan rides in the kernel parameters (constant bank): no global loads in the prologue
    const LevelGeo& D = P->lv[level];
    const LevelGeo& S = P->lv[level - 1];
    const int x4 = (blockIdx.x * 64 + threadIdx.x) * 4;
    const int y = blockIdx.y * 4 + threadIdx.y;
    const int f = f0 + blockIdx.z;
    if (x4 >= D.w || y >= D.h) return;
    const uint8_t* src = pyr + S.pyrOff + (long long)f * S.frameStride;
    uint8_t* dst = pyr + D.pyrOff + (long long)f * D.frameStride + (long long)y * D.pitch + x4;
    uint32_t out = 0;
    if (S.w == 2 * D.w && S.h == 2 * D.h) {  // OpenCV's exact-2x INTER_AREA shortcut
        const uint8_t* r0 = src + (long long)(2 * y) * S.pitch;
        const uint8_t* r1 = r0 + S.pitch;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int x = x4 + i;
            if (x < D.w) {
                int v = (r0[2 * x] + r0[2 * x + 1] + r1[2 * x] + r1[2 * x + 1] + 2) >> 2;
                out |= (uint32_t)v << (8 * i);
            }
        }
        *(uint32_t*)dst = out;
        return;
    }
    const int2* xt = tabs + D.xtabOff;
    const int2 ty = tabs[D.ytabOff + y];
    const int sy0 = ty.x & 0xffff, sy1 = (int)((unsigned)ty.x >> 16);  // clamped on the host (linear_table)
    const int b0 = (short)(ty.y & 0xffff), b1 = ty.y >> 16;
    const uint8_t* r0 = src + (long long)sy0 * S.pitch;
    const uint8_t* r1 = src + (long long)sy1 * S.pitch;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int x = x4 + i;
        if (x < D.w) {
            const int2 tx = xt[x];
            const int sx = tx.x >> 16, sx1 = min(sx + 1, S.w - 1);  // (the low half is k_resize_fast's PRMT selector)
            const int a0 = (short)(tx.y & 0xffff), a1 = tx.y >> 16;
            const int H0 = r0[sx] * a0 + r0[sx1] * a1;
            const int H1 = r1[sx] * a0 + r1[sx1] * a1;
            const int v = (((b0 * (H0 >> 4)) >> 16) + ((b1 * (H1 >> 4)) >> 16) + 2) >> 2;
            out |= (uint32_t)(v & 0xff) << (8 * i);
        }
    }
    *(uint32_t*)dst = out;
}

#include "orbx_resize.cuh"

// ------------------------------------------------------------------------------------------------ K2 FAST
#include "orbx_fast.cuh"
#include "orbx_fast2.cuh"

// ------------------------------------------------------------------------------------------------ K3 quadtree
// DistributeOctTree without lists or pointers. Nodes live in shared memory as rectangles + key counts; every key
// carries the id of the node it is in. A sweep divides every node with > 1 key ("full" sweep) or, once
// size + 3*nToExpand > N, the largest nodes first until size >= N ("sorted" sweep: bitonic sort of
// (count desc, UL.y, UL.x), prefix sum of per-node gains, cut at the first prefix reaching N). The reference's
// termination rules (size >= N, or a sweep that did not grow the node set) are kept verbatim.
struct QtSmem {
    short4* geoA; short4* geoB;
    int* cntA; int* cntB;
    int* slotOf; int* remap; int* rankOfSlot; int* orderSlot; int* nodeOfSlot;
    int* childCnt; int* childNew; int* gain; int* flags;
    u64* sortbuf;
    int* ws;
};

#define QT_U 4  // candidates per thread and trip in the passes over the candidate list
#define QT_RANK_MAX 256  // up to this many keys a sort is a rank-by-counting pass

__device__ __forceinline__ int quadrant_of(const short4 g, int x, int y) {
    const int mx = g.x + ((g.y - g.x + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2)
    const int my = g.z + ((g.w - g.z + 1) >> 1);
    return (x < mx ? 0 : 1) + (y < my ? 0 : 2);
}

__global__ void k_quadtree(const __grid_constant__ DevPlan Pc, const u64* __restrict__ cand,
                           const int* __restrict__ candCount, int* __restrict__ nodeOf, uint2* __restrict__ sel,
                           int* __restrict__ selCount, int C, int C2, int* __restrict__ err, int f0, int level0) {
    const DevPlan* P = &Pc;  // the plan rides in the kernel parameters (constant bank): no global loads in the prologue
    extern __shared__ __align__(16) uint8_t qt_smem[];
    uint8_t* smem = qt_smem;
    // (frame-major launch order: the level-0 CTAs of all frames — the longest — are dispatched first)
    const int level = level0 + blockIdx.y, f = f0 + blockIdx.x;
    const LevelGeo& G = P->lv[level];
    const int nl = P->nlevels;
    const int tid = threadIdx.x, T = blockDim.x;
    int n = candCount[f * nl + level];
    if (n > G.candCap) n = G.candCap;
    if (n == 0) {
        if (tid == 0) selCount[f * nl + level] = 0;
        return;
    }
    const u64* keys = cand + G.candOff + (long long)f * G.candCap;
    int* nod = nodeOf + G.candOff + (long long)f * G.candCap;

    QtSmem S;
    {
        uint8_t* p = smem;
        S.sortbuf = (u64*)p; p += sizeof(u64) * C2;
        S.geoA = (short4*)p; p += sizeof(short4) * C;
        S.geoB = (short4*)p; p += sizeof(short4) * C;
        S.cntA = (int*)p; p += 4 * C;
        S.cntB = (int*)p; p += 4 * C;
        S.slotOf = (int*)p; p += 4 * C;
        S.remap = (int*)p; p += 4 * C;
        S.rankOfSlot = (int*)p; p += 4 * C;
        S.orderSlot = (int*)p; p += 4 * C;
        S.nodeOfSlot = (int*)p; p += 4 * C;
        S.gain = (int*)p; p += 4 * C;
        S.flags = (int*)p; p += 4 * C;
        S.childCnt = (int*)p; p += 16 * C;
        S.childNew = (int*)p; p += 16 * C;
        S.ws = (int*)p;  // 40 ints
    }
    const int N = G.quota;
    const float hX = G.hX;
    int nNodes = G.nIni;
    for (int i = tid; i < C; i += T) S.cntA[i] = 0;
    for (int i = tid; i < nNodes; i += T)
        S.geoA[i] = make_short4((short)(int)__fmul_rn(hX, (float)i), (short)(int)__fmul_rn(hX, (float)(i + 1)), 0,
                                (short)G.treeH);
    __syncthreads();
    // (every pass over the candidates takes QT_U of them per trip with all their loads issued first: a trip is one global
    // round trip, and the level-0 CTA of a large frame makes dozens of trips per pass — 4K frames: 558 us per launch before)
    for (int k0 = tid; k0 < n; k0 += QT_U * T) {
        u64 kk[QT_U];
#pragma unroll
        for (int j = 0; j < QT_U; j++) kk[j] = (k0 + j * T < n) ? keys[k0 + j * T] : 0ull;
#pragma unroll
        for (int j = 0; j < QT_U; j++) {
            const int k = k0 + j * T;
            if (k < n) {
                int x, y, s;
                cand_unpack(kk[j], G, x, y, s);
                int idx = (int)__fdiv_rn((float)x, hX);
                idx = min(idx, nNodes - 1);
                nod[k] = idx;
                atomicAdd(&S.cntA[idx], 1);
            }
        }
    }
    __syncthreads();

    bool sortedMode = false, first = true;
    int size = 0, prevSize = 0;
    __shared__ int qt_fs[64];  // block_flag_scan's warp totals
    int fsPhase = 0;
    size = block_flag_scan(S.flags, nNodes, qt_fs, &fsPhase, [&](int i) { return S.cntA[i] > 0; });  // alive count
    for (int sweep = 0; sweep < 64; sweep++) {
        // expandable nodes -> slots (a thread reads back only the flags it wrote itself)
        const int nE = block_flag_scan(S.flags, nNodes, qt_fs, &fsPhase, [&](int i) { return S.cntA[i] > 1; });
        if (!first) {
            if (size >= N || size == prevSize) break;
            if (!sortedMode && size + 3 * nE > N) sortedMode = true;
        }
        first = false;
        if (nE == 0) break;
        for (int i = tid; i < nNodes; i += T) {
            if (S.cntA[i] > 1) {
                const int s = S.flags[i];
                S.slotOf[i] = s;
                S.nodeOfSlot[s] = i;
            } else
                S.slotOf[i] = -1;
        }
        for (int i = tid; i < 4 * nE; i += T) S.childCnt[i] = 0;
        __syncthreads();
        if (sortedMode) {
            const int P2 = next_pow2(nE);
            for (int s = tid; s < P2; s += T) {
                u64 key = ~0ull;
                if (s < nE) {
                    const int nd = S.nodeOfSlot[s];
                    const short4 g = S.geoA[nd];
                    key = ((u64)(0x3FFFFFu - (unsigned)min(S.cntA[nd], 0x3FFFFF)) << 40) | ((u64)(unsigned short)g.z << 26) |
                          ((u64)(unsigned short)g.x << 12) | (u64)s;
                }
                S.sortbuf[s] = key;
            }
            __syncthreads();
            if (nE <= QT_RANK_MAX) {
                // few keys: rank by counting (the keys are distinct: the slot is their low bits) — one pass of broadcast
                // reads and no barrier, against log^2 barrier steps of the bitonic network
                for (int sl = tid; sl < nE; sl += T) {
                    const u64 key = S.sortbuf[sl];
                    int r = 0;
                    for (int j = 0; j < nE; j++) r += S.sortbuf[j] < key;
                    S.orderSlot[r] = sl;
                    S.rankOfSlot[sl] = r;
                }
            } else {
                block_bitonic_sort(S.sortbuf, P2);
                for (int r = tid; r < nE; r += T) {
                    const int s = (int)(S.sortbuf[r] & 0xFFFull);
                    S.orderSlot[r] = s;
                    S.rankOfSlot[s] = r;
                }
            }
        } else {
            for (int r = tid; r < nE; r += T) {
                S.orderSlot[r] = r;
                S.rankOfSlot[r] = r;
            }
        }
        __syncthreads();
        // child key counts
        for (int k0 = tid; k0 < n; k0 += QT_U * T) {
            u64 kk[QT_U];
            int nn[QT_U];
#pragma unroll
            for (int j = 0; j < QT_U; j++) {
                const bool in = k0 + j * T < n;
                kk[j] = in ? keys[k0 + j * T] : 0ull;
                nn[j] = in ? nod[k0 + j * T] : -1;
            }
#pragma unroll
            for (int j = 0; j < QT_U; j++) {
                if (nn[j] < 0) continue;
                const int nd = nn[j];
                const int s = S.slotOf[nd];
                if (s >= 0) {
                    int x, y, sc;
                    cand_unpack(kk[j], G, x, y, sc);
                    atomicAdd(&S.childCnt[4 * s + quadrant_of(S.geoA[nd], x, y)], 1);
                }
            }
        }
        __syncthreads();
        // gains in rank order, prefix sum, cut: gain + 1 = the number of non-empty children (0..4) is scanned bit by bit with
        // ballots (one barrier), and the prefix of the gains is that prefix minus the rank
        if (tid == 0) S.ws[36] = nE;  // expandCount
        block_small_scan<3>(S.gain, nE, qt_fs, &fsPhase, [&](int r) {
            const int* cc = S.childCnt + 4 * S.orderSlot[r];
            return (cc[0] > 0) + (cc[1] > 0) + (cc[2] > 0) + (cc[3] > 0);
        });
        for (int r = tid; r < nE; r += T) S.gain[r] -= r;  // (own elements)
        if (sortedMode) {
            for (int r = tid; r < nE; r += T) {
                const int* cc = S.childCnt + 4 * S.orderSlot[r];
                const int own = (cc[0] > 0) + (cc[1] > 0) + (cc[2] > 0) + (cc[3] > 0) - 1;
                if (size + S.gain[r] + own >= N) atomicMin(&S.ws[36], r + 1);
            }
        }
        __syncthreads();
        const int expandCount = S.ws[36];
        // survivors (alive, not divided in this sweep)
        const int nSurv = block_flag_scan(S.flags, nNodes, qt_fs, &fsPhase, [&](int i) {
            const int s = S.slotOf[i];
            return (S.cntA[i] > 0) && (s < 0 || S.rankOfSlot[s] >= expandCount);
        });
        const int nCh = block_flag_scan(S.childNew, 4 * nE, qt_fs, &fsPhase, [&](int j) {
            const int r = j >> 2, q = j & 3;
            return (r < expandCount) && (S.childCnt[4 * S.orderSlot[r] + q] > 0);
        });
        if (nSurv + nCh > C) {  // cannot happen (node bound); guard against smem overrun
            if (tid == 0) atomicExch(err, 4);
            break;
        }
        for (int i = tid; i < nNodes; i += T) {
            const int s = S.slotOf[i];
            if ((S.cntA[i] > 0) && (s < 0 || S.rankOfSlot[s] >= expandCount)) {
                const int id = S.flags[i];
                S.remap[i] = id;
                S.geoB[id] = S.geoA[i];
                S.cntB[id] = S.cntA[i];
            }
        }
        for (int j = tid; j < 4 * nE; j += T) {
            const int r = j >> 2, q = j & 3;
            if (r >= expandCount) continue;
            const int s = S.orderSlot[r];
            const int cc = S.childCnt[4 * s + q];
            if (cc == 0) continue;
            const short4 g = S.geoA[S.nodeOfSlot[s]];
            const short mx = g.x + ((g.y - g.x + 1) >> 1), my = g.z + ((g.w - g.z + 1) >> 1);
            short4 cg;
            cg.x = (q & 1) ? mx : g.x;
            cg.y = (q & 1) ? g.y : mx;
            cg.z = (q & 2) ? my : g.z;
            cg.w = (q & 2) ? g.w : my;
            const int id = nSurv + S.childNew[j];
            S.geoB[id] = cg;
            S.cntB[id] = cc;
        }
        __syncthreads();
        for (int k0 = tid; k0 < n; k0 += QT_U * T) {
            u64 kk[QT_U];
            int nn[QT_U];
#pragma unroll
            for (int j = 0; j < QT_U; j++) {
                const bool in = k0 + j * T < n;
                kk[j] = in ? keys[k0 + j * T] : 0ull;
                nn[j] = in ? nod[k0 + j * T] : -1;
            }
#pragma unroll
            for (int j = 0; j < QT_U; j++) {
                if (nn[j] < 0) continue;
                const int nd = nn[j];
                const int s = S.slotOf[nd];
                if (s >= 0 && S.rankOfSlot[s] < expandCount) {
                    int x, y, sc;
                    cand_unpack(kk[j], G, x, y, sc);
                    nod[k0 + j * T] = nSurv + S.childNew[4 * S.rankOfSlot[s] + quadrant_of(S.geoA[nd], x, y)];
                } else
                    nod[k0 + j * T] = S.remap[nd];
            }
        }
        __syncthreads();
        {
            short4* tg = S.geoA; S.geoA = S.geoB; S.geoB = tg;
            int* tc = S.cntA; S.cntA = S.cntB; S.cntB = tc;
        }
        prevSize = size;
        size = nSurv + nCh;
        nNodes = size;
    }

    // one keypoint per node: max (score, first in reference order)
    u64* best = (u64*)S.childCnt;  // 2*C u64 available
    for (int i = tid; i < nNodes; i += T) best[i] = 0;
    __syncthreads();
    for (int k0 = tid; k0 < n; k0 += QT_U * T) {
        u64 kk[QT_U];
        int nn[QT_U];
#pragma unroll
        for (int j = 0; j < QT_U; j++) {
            const bool in = k0 + j * T < n;
            kk[j] = in ? keys[k0 + j * T] : 0ull;
            nn[j] = in ? nod[k0 + j * T] : -1;
        }
#pragma unroll
        for (int j = 0; j < QT_U; j++)
            if (nn[j] >= 0) atomicMax(&best[nn[j]], kk[j]);
    }
    __syncthreads();
    const int P2 = next_pow2(nNodes);
    for (int i = tid; i < P2; i += T) {
        u64 key = ~0ull;
        if (i < nNodes && S.cntA[i] > 0) {
            int x, y, sc;
            cand_unpack(best[i], G, x, y, sc);
            x += P->edge - 3;
            y += P->edge - 3;
            key = ((u64)(unsigned)((y << 16) | x) << 32) | (u64)(unsigned)sc;
        }
        S.sortbuf[i] = key;
    }
    __syncthreads();
    const int nOut = min(size, G.selCap);
    if (size > G.selCap && tid == 0) atomicExch(err, 5);
    uint2* out = sel + (long long)f * P->selStride + G.selOff;
    if (nNodes <= QT_RANK_MAX) {
        // rank by counting (empty nodes carry the key ~0 and rank past every keypoint; equal keys ordered by index)
        for (int i = tid; i < nNodes; i += T) {
            const u64 key = S.sortbuf[i];
            if (key == ~0ull) continue;
            int r = 0;
            for (int j = 0; j < nNodes; j++) {
                const u64 o = S.sortbuf[j];
                r += (o < key) || (o == key && j < i);
            }
            if (r < nOut) out[r] = make_uint2((unsigned)(key >> 32), (unsigned)(key & 0xFFFFFFFFull));
        }
    } else {
        block_bitonic_sort(S.sortbuf, P2);
        for (int i = tid; i < nOut; i += T) {
            const u64 k = S.sortbuf[i];
            out[i] = make_uint2((unsigned)(k >> 32), (unsigned)(k & 0xFFFFFFFFull));
        }
    }
    if (tid == 0) selCount[f * nl + level] = nOut;
}

// ------------------------------------------------------------------------------------------------ K4 blur
#include "orbx_blur.cuh"

// ------------------------------------------------------------------------------------------------ K5 orientation + descriptor
#include "orbx_desc.cuh"
#include "orbx_prep.cuh"

// ------------------------------------------------------------------------------------------------ K6 export
// Copy one level of one frame into a dense (w+2e) x (h+2e) buffer with the BORDER_REFLECT_101 frame (e may be 0).
__global__ void k_export(const DevPlan* __restrict__ P, const uint8_t* __restrict__ src, int level, int f, int e,
                         uint8_t* __restrict__ dst) {
    const LevelGeo& G = P->lv[level];
    const int W = G.w + 2 * e, H = G.h + 2 * e;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    const uint8_t* s = src + G.pyrOff + (long long)f * G.frameStride;
    dst[(long long)y * W + x] = s[(long long)reflect101(y - e, G.h) * G.pitch + reflect101(x - e, G.w)];
}

// All levels of one frame in ONE launch: level l goes to dst + off[l] as a dense (w+2e) x (h+2e) reflect-101 buffer.
// blockIdx.z = level; the grid is sized for level 0 and the smaller levels exit early.
struct ExportOffsets { long long off[ORBX_MAX_LEVELS]; };
__global__ void k_export_all(const DevPlan* __restrict__ P, const uint8_t* __restrict__ src, int f, int e,
                             const __grid_constant__ ExportOffsets offs, uint8_t* __restrict__ dst) {
    const int level = blockIdx.z;
    const LevelGeo& G = P->lv[level];
    const int W = G.w + 2 * e, H = G.h + 2 * e;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    const uint8_t* s = src + G.pyrOff + (long long)f * G.frameStride;
    dst[offs.off[level] + (long long)y * W + x] = s[(long long)reflect101(y - e, G.h) * G.pitch + reflect101(x - e, G.w)];
}
