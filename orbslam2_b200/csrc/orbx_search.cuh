// orbx_search.cuh — device-resident windowed search (included by orbx_match.cu).
// Replaces ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono), /root/reference/src/ORBmatcher.cc:1324-1466,
// and the Frame functions it calls on the current frame: AssignFeaturesToGrid / PosInGrid (src/Frame.cc:231-246,
// 383-393) and GetFeaturesInArea (src/Frame.cc:328-381); rotation histogram + ComputeThreeMaxima (:1597-1638).
//
//   S1 k_sp_grid_count / k_sp_grid_fill   keypoints -> 64 x 48 cell lists in keypoint-index order (counting sort:
//                                          atomics place, a per-cell insertion sort restores the push_back order)
//   S2 k_sp_candidates<FILL>              one warp per last-frame map point: project, window, walk the cells in the
//                                          reference's order (ix outer, iy inner, items in insertion order), apply the
//                                          level / window / stereo gates, Hamming distance per surviving candidate.
//                                          Run twice: count, (scan), fill — the CSR lists keep the reference's order,
//                                          which is what breaks distance ties.
//   S3 k_sp_claim / k_sp_choose / k_sp_finish  the sequential loop's greedy claims ("skip a keypoint that an
//                                          earlier query already took") as a fixed point: every query picks its best
//                                          candidate among those not claimed by a lower-index query of the previous
//                                          round; rounds repeat until nothing changes. By induction on the query index
//                                          the fixed point is the sequential result (query 0 is right after round 1,
//                                          query k once 0..k-1 are). Then the rotation histogram, ComputeThreeMaxima,
//                                          the filter, and the final mvpMapPoints image.
#pragma once

#define SP_COLS 64  // FRAME_GRID_COLS, include/Frame.h:37
#define SP_ROWS 48  // FRAME_GRID_ROWS
#define SP_CELLS (SP_COLS * SP_ROWS)
#define SP_HISTO 30

struct SpDev {  // device copies of the call's arrays
    orbm_proj_params p;
    int nLast, nCur;
    const float* worldPos; const uint8_t* lflags; const int32_t* loct; const float* langle; const uint4* ldesc;
    const orbx_kp* keys; const float* uRight; const uint4* cdesc; const uint8_t* taken;
    // mode 2 = SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th): host-projected points (projX/projY, loct =
    // nPredictedLevel), radius th*scale[level], octave gate [level-1, level], every claim blocks later points
    // mode 1 = SearchByProjection(Frame&, vector<MapPoint*>, th): the queries come with their tracking projection
    // (MapPoint::mTrackProjX/Y/XR, mnTrackScaleLevel, mTrackViewCos set by Frame::isInFrustum, src/Frame.cc:256-326)
    // mode 3 = SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize): the queries are F1's level-0
    // keypoints (qkeys), window centre vbPrevMatched[i] (projX = the interleaved n x 2 array), radius th = windowSize,
    // level gate [0, 0]
    int mode;
    const float* projX; const float* projY; const float* projXR; const float* viewCos;
    float nnRatio;
    const orbx_kp* qkeys;
};

// Frame::PosInGrid (src/Frame.cc:383-393): round() = half away from zero, on the float product
__device__ __forceinline__ int sp_cell_of(const SpDev& S, int i) {
    const orbx_kp k = S.keys[i];
    const int posX = (int)round((double)__fmul_rn(__fsub_rn(k.x, S.p.minX), S.p.grid_w_inv));
    const int posY = (int)round((double)__fmul_rn(__fsub_rn(k.y, S.p.minY), S.p.grid_h_inv));
    if (posX < 0 || posX >= SP_COLS || posY < 0 || posY >= SP_ROWS) return -1;
    return posX * SP_ROWS + posY;
}

__global__ void k_sp_grid_count(SpDev S, int* __restrict__ cellCount) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S.nCur) return;
    const int c = sp_cell_of(S, i);
    if (c >= 0) atomicAdd(&cellCount[c], 1);
}

// one CTA: exclusive scan of the 3072 cell counts -> cellStart[0..3072]; fill cursors reset
__global__ void k_sp_grid_scan(const int* __restrict__ cellCount, int* __restrict__ cellStart, int* __restrict__ cursor) {
    __shared__ int part[1024];
    const int tid = threadIdx.x;  // 1024 threads, 3 cells each
    int c[3], s = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        c[k] = cellCount[tid * 3 + k];
        s += c[k];
    }
    part[tid] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = tid >= o ? part[tid - o] : 0;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    int base = part[tid] - s;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        cellStart[tid * 3 + k] = base;
        cursor[tid * 3 + k] = base;
        base += c[k];
    }
    if (tid == 1023) cellStart[SP_CELLS] = base;
}

__global__ void k_sp_grid_fill(SpDev S, int* __restrict__ cursor, int* __restrict__ items) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S.nCur) return;
    const int c = sp_cell_of(S, i);
    if (c >= 0) items[atomicAdd(&cursor[c], 1)] = i;
}

// thread per cell: restore keypoint-index order inside the cell (cells hold a handful of items)
__global__ void k_sp_grid_sort(const int* __restrict__ cellStart, int* __restrict__ items) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= SP_CELLS) return;
    const int s = cellStart[c], e = cellStart[c + 1];
    for (int a = s + 1; a < e; a++) {
        const int v = items[a];
        int b = a - 1;
        while (b >= s && items[b] > v) {
            items[b + 1] = items[b];
            b--;
        }
        items[b + 1] = v;
    }
}

// The whole grid build (clear, count, scan, fill, per-cell sort) in ONE 1024-thread CTA with the counters in shared memory:
// a camera frame has a few thousand keypoints, far too few to need a grid of CTAs, and the four launches + memset cost more
// than the work. Same cellStart / items as the kernel sequence above.
__global__ void __launch_bounds__(1024) k_sp_grid_build(SpDev S, int* __restrict__ cellStart, int* __restrict__ items) {
    __shared__ int cnt[SP_CELLS];
    __shared__ int part[1024];
    const int tid = threadIdx.x;
    for (int c = tid; c < SP_CELLS; c += 1024) cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < S.nCur; i += 1024) {
        const int c = sp_cell_of(S, i);
        if (c >= 0) atomicAdd(&cnt[c], 1);
    }
    __syncthreads();
    int c3[3], s = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        c3[k] = cnt[tid * 3 + k];
        s += c3[k];
    }
    part[tid] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = tid >= o ? part[tid - o] : 0;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    int base = part[tid] - s;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        cellStart[tid * 3 + k] = base;
        cnt[tid * 3 + k] = base;  // fill cursor
        base += c3[k];
    }
    if (tid == 1023) cellStart[SP_CELLS] = base;
    __syncthreads();
    for (int i = tid; i < S.nCur; i += 1024) {
        const int c = sp_cell_of(S, i);
        if (c >= 0) items[atomicAdd(&cnt[c], 1)] = i;
    }
    __syncthreads();
    for (int c = tid; c < SP_CELLS; c += 1024) {  // restore keypoint-index order inside the cell
        const int b0 = cellStart[c], e = cnt[c];  // the cursor ended at the cell's end
        for (int a = b0 + 1; a < e; a++) {
            const int v = items[a];
            int b = a - 1;
            while (b >= b0 && items[b] > v) {
                items[b + 1] = items[b];
                b--;
            }
            items[b + 1] = v;
        }
    }
}

// Projection of last-frame point i into the current frame and its search window (src/ORBmatcher.cc:1353-1386).
struct SpQuery {
    float u, v, invzc, radius, ur;  // ur = expected right-image coordinate for the stereo gate
    int minLevel, maxLevel;  // GetFeaturesInArea level arguments
    int x0, x1, y0, y1;      // cell range
    bool valid;
};

__device__ __forceinline__ SpQuery sp_query(const SpDev& S, int i) {
    SpQuery q;
    q.valid = false;
    if (S.mode == 3) {  // src/ORBmatcher.cc:415-420
        const int level1 = S.qkeys[i].octave;
        if (level1 > 0) return q;
        q.radius = S.p.th;
        q.u = S.projX[2 * i];
        q.v = S.projX[2 * i + 1];
        q.ur = 0.f;
        q.invzc = 0.f;
        q.minLevel = level1;
        q.maxLevel = level1;
    } else if (!(S.lflags[i] & 1)) {
        return q;
    } else if (S.mode == 2) {  // SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th): src/ORBmatcher.cc:349-352, 375
        const int lvl = S.loct[i];
        q.radius = __fmul_rn(S.p.th, S.p.scale[lvl]);
        q.u = S.projX[i];
        q.v = S.projY[i];
        q.ur = 0.f;
        q.invzc = 0.f;
        q.minLevel = lvl - 1;
        q.maxLevel = lvl;
    } else if (S.mode == 1) {  // src/ORBmatcher.cc:58-68
        const int lvl = S.loct[i];
        float r = (double)S.viewCos[i] > 0.998 ? 2.5f : 4.0f;  // RadiusByViewingCos, :129-135 (compares in double)
        if (S.p.th != 1.0f) r = __fmul_rn(r, S.p.th);
        q.radius = __fmul_rn(r, S.p.scale[lvl]);
        q.u = S.projX[i];
        q.v = S.projY[i];
        q.ur = S.projXR[i];
        q.invzc = 0.f;
        q.minLevel = lvl - 1;
        q.maxLevel = lvl;
    } else {
    const float X = S.worldPos[3 * i], Y = S.worldPos[3 * i + 1], Z = S.worldPos[3 * i + 2];
    float c3[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {  // Rcw*x3Dw + tcw as cv::gemm's small-matrix path evaluates it: float products summed
                                   // left to right in float (no FMA), then d = (float)((double)t + (double)tcw)
        const float t = __fadd_rn(__fadd_rn(__fmul_rn(S.p.Rcw[3 * r], X), __fmul_rn(S.p.Rcw[3 * r + 1], Y)),
                                  __fmul_rn(S.p.Rcw[3 * r + 2], Z));
        c3[r] = (float)__dadd_rn((double)t, (double)S.p.tcw[r]);
    }
    q.invzc = (float)(1.0 / (double)c3[2]);
    if (q.invzc < 0 && !S.p.skip_depth_test) return q;
    q.u = __fadd_rn(__fmul_rn(__fmul_rn(S.p.fx, c3[0]), q.invzc), S.p.cx);
    q.v = __fadd_rn(__fmul_rn(__fmul_rn(S.p.fy, c3[1]), q.invzc), S.p.cy);
    if (q.u < S.p.minX || q.u > S.p.maxX) return q;
    if (q.v < S.p.minY || q.v > S.p.maxY) return q;
    const int oct = S.loct[i];
    q.radius = __fmul_rn(S.p.th, S.p.scale[oct]);
    if (S.p.forward) { q.minLevel = oct; q.maxLevel = -1; }
    else if (S.p.backward) { q.minLevel = 0; q.maxLevel = oct; }
    else { q.minLevel = oct - 1; q.maxLevel = oct + 1; }
    q.ur = __fsub_rn(q.u, __fmul_rn(S.p.bf, q.invzc));
    }
    // Frame::GetFeaturesInArea cell range (src/Frame.cc:333-347)
    q.x0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(q.u, S.p.minX), q.radius), S.p.grid_w_inv)));
    if (q.x0 >= SP_COLS) return q;
    q.x1 = min(SP_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(q.u, S.p.minX), q.radius), S.p.grid_w_inv)));
    if (q.x1 < 0) return q;
    q.y0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(q.v, S.p.minY), q.radius), S.p.grid_h_inv)));
    if (q.y0 >= SP_ROWS) return q;
    q.y1 = min(SP_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(q.v, S.p.minY), q.radius), S.p.grid_h_inv)));
    if (q.y1 < 0) return q;
    q.valid = true;
    return q;
}

// One warp per query. FILL = false: candCount[i] = number of candidates. FILL = true: the (keypoint, distance) pairs
// are written at candOff[i] in the reference's visiting order.
template <bool FILL>
__global__ void __launch_bounds__(256) k_sp_candidates(SpDev S, const int* __restrict__ cellStart,
                                                       const int* __restrict__ items, int* __restrict__ candCount,
                                                       const int* __restrict__ candOff, int* __restrict__ candIdx,
                                                       unsigned short* __restrict__ candDist, const int* __restrict__ go = nullptr) {
    if (FILL && go && *go == 0) return;  // the lists would not fit: the host grows the buffer and runs the call again
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= S.nLast) return;
    const SpQuery q = sp_query(S, i);  // uniform across the warp
    int n = 0;
    if (q.valid) {
        const bool bCheckLevels = (q.minLevel > 0) || (q.maxLevel >= 0);
        uint4 d0, d1;
        int base = 0;
        if (FILL) {
            d0 = S.ldesc[2 * i];
            d1 = S.ldesc[2 * i + 1];
            base = candOff[i];
        }
        const float ur = q.ur;
        for (int ix = q.x0; ix <= q.x1; ix++) {
            // the cells (ix, y0..y1) are consecutive in the cell index, so their items are one contiguous run that is
            // already in the reference's order (iy ascending, then insertion order)
            const int s = cellStart[ix * SP_ROWS + q.y0], e = cellStart[ix * SP_ROWS + q.y1 + 1];
            for (int b = s; b < e; b += 32) {
                const int k = b + lane;
                bool pass = false;
                int i2 = -1;
                if (k < e) {
                    i2 = items[k];
                    const orbx_kp kp = S.keys[i2];
                    pass = true;
                    if (bCheckLevels) {
                        if (kp.octave < q.minLevel) pass = false;
                        if (q.maxLevel >= 0 && kp.octave > q.maxLevel) pass = false;
                    }
                    const float distx = __fsub_rn(kp.x, q.u), disty = __fsub_rn(kp.y, q.v);
                    if (!(fabsf(distx) < q.radius && fabsf(disty) < q.radius)) pass = false;
                    // static part of the skip test at src/ORBmatcher.cc:1399-1401 (points present on entry)
                    if (S.taken && S.taken[i2]) pass = false;
                    if (pass && S.uRight) {  // src/ORBmatcher.cc:1403-1409
                        const float uR = S.uRight[i2];
                        if (uR > 0) {
                            const float er = fabsf(__fsub_rn(ur, uR));
                            if (er > q.radius) pass = false;
                        }
                    }
                }
                const unsigned m = __ballot_sync(0xffffffffu, pass);
                if (FILL && pass) {
                    const int pos = base + n + __popc(m & ((1u << lane) - 1u));
                    const uint4 c0 = S.cdesc[2 * i2], c1 = S.cdesc[2 * i2 + 1];
                    candIdx[pos] = i2;
                    candDist[pos] = (unsigned short)hamming256(d0, d1, c0, c1);
                }
                n += __popc(m);
            }
        }
    }
    if (!FILL && lane == 0) candCount[i] = n;
}

// single CTA exclusive scan of n counts (n <= 1 << 20): off[0..n], returns total in off[n]
__global__ void k_sp_scan(const int* __restrict__ cnt, int n, int* __restrict__ off) {
    __shared__ int part[1024];
    __shared__ int carry;
    const int tid = threadIdx.x;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const int v = i < n ? cnt[i] : 0;
        part[tid] = v;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {
            const int a = tid >= o ? part[tid - o] : 0;
            __syncthreads();
            part[tid] += a;
            __syncthreads();
        }
        if (i < n) off[i] = carry + part[tid] - v;
        __syncthreads();
        if (tid == 1023) carry += part[1023];
        __syncthreads();
    }
    if (tid == 0) off[n] = carry;
}

// The candidate total is only known on the device. Instead of a host round trip to size the buffer, the call runs
// against the buffer it has: go = (total fits); everything downstream returns at once when go == 0 and the host, which
// reads go and the total together with the results, grows the buffer and repeats the call (first call of a handle only).
__global__ void k_sp_guard(const int* __restrict__ candOff, int n, int capEntries, int* __restrict__ go, int* __restrict__ res) {
    const int total = candOff[n];
    *go = total <= capEntries;
    res[2] = total;
}

// One round of the fixed point is two grid-wide kernels (the candidate lists of a dense frame are ~100 entries per
// query, too much for one CTA): k_sp_claim rebuilds owner[] from the previous round's choices, k_sp_choose re-picks.
// flag[r] = "round r changed something"; the kernels of round r + 1 return at once when flag[r] == 0 (converged), so
// the host can enqueue a few rounds speculatively and look at the flags once.
__global__ void k_sp_claim(SpDev S, const int* __restrict__ choice, int* __restrict__ owner, const int* __restrict__ flag,
                           int round) {
    if (flag[round - 1] == 0) return;  // flag[-1]: the call's go flag (0 = candidate buffer too small, see k_sp_guard)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S.nLast) return;
    // owner[i2] = lowest-index query that claimed i2 in the previous round with a point that blocks later queries
    // (CurrentFrame.mvpMapPoints[i2]->Observations() > 0, src/ORBmatcher.cc:1399-1401)
    const int c = choice[i];
    // (mode 2: vpMatched[bestIdx] = pMP blocks every later point, src/ORBmatcher.cc:370, 391)
    if (c >= 0 && ((S.lflags[i] & 2) || S.mode == 2)) atomicMin(&owner[c], i);
}

__global__ void __launch_bounds__(256) k_sp_choose(SpDev S, const int* __restrict__ candOff, const int* __restrict__ candIdx,
                                                   const unsigned short* __restrict__ candDist, int* __restrict__ choice,
                                                   const int* __restrict__ owner, int* __restrict__ flag, int round) {
    if (flag[round - 1] == 0) return;  // flag[-1]: the call's go flag (0 = candidate buffer too small, see k_sp_guard)
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= S.nLast) return;
    const int s = candOff[i], e = candOff[i + 1];
    unsigned best = 0xFFFFFFFFu;  // (distance << 22) | position: the minimum is the smallest distance, first in list order
    for (int c = s + lane; c < e; c += 32) {
        const int i2 = candIdx[c];
        if (owner[i2] < i) continue;  // taken by an earlier query
        const unsigned key = ((unsigned)candDist[c] << 22) | (unsigned)(c - s);
        best = min(best, key);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) {
        int nc = -1;
        if (best != 0xFFFFFFFFu && (int)(best >> 22) <= S.p.th_high) nc = candIdx[s + (int)(best & 0x3FFFFFu)];
        if (nc != choice[i]) {
            choice[i] = nc;
            flag[round] = 1;
        }
    }
}

__global__ void __launch_bounds__(1024) k_sp_finish(SpDev S, const int* __restrict__ choice, int32_t* __restrict__ curMatch,
                                                    int32_t* __restrict__ lastMatch, uint8_t* __restrict__ lastKeep,
                                                    int* __restrict__ result /* [0] nmatches */) {
    __shared__ int hist[SP_HISTO];
    __shared__ int keepBin[SP_HISTO];
    __shared__ int nKeep;
    const int tid = threadIdx.x, T = blockDim.x;
    const int nL = S.nLast, nC = S.nCur;
    // rotation histogram (src/ORBmatcher.cc:1427-1437) and ComputeThreeMaxima (:1597-1638)
    if (tid < SP_HISTO) {
        hist[tid] = 0;
        keepBin[tid] = 1;
    }
    if (tid == 0) nKeep = 0;
    for (int k = tid; k < nC; k += T) curMatch[k] = -1;
    __syncthreads();
    const float factor = 1.0f / SP_HISTO;
    for (int i = tid; i < nL; i += T) {
        const int c = choice[i];
        lastMatch[i] = c;
        if (c >= 0) {
            atomicMax(&curMatch[c], i);  // a later claim overwrites an earlier one
            if (S.p.check_orientation) {
                float rot = __fsub_rn(S.langle[i], S.keys[c].angle);
                if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                int bin = (int)roundf(__fmul_rn(rot, factor));
                if (bin == SP_HISTO) bin = 0;
                atomicAdd(&hist[bin], 1);
            }
        }
    }
    __syncthreads();
    if (tid == 0 && S.p.check_orientation) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < SP_HISTO; i++) {
            const int s = hist[i];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < 0.1f * (float)max1) { ind3 = -1; }
        for (int i = 0; i < SP_HISTO; i++) keepBin[i] = (i == ind1 || i == ind2 || i == ind3);
    }
    __syncthreads();
    for (int i = tid; i < nL; i += T) {
        const int c = choice[i];
        int keep = 0;
        if (c >= 0) {
            keep = 1;
            if (S.p.check_orientation) {
                float rot = __fsub_rn(S.langle[i], S.keys[c].angle);
                if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                int bin = (int)roundf(__fmul_rn(rot, factor));
                if (bin == SP_HISTO) bin = 0;
                keep = keepBin[bin];
            }
            if (keep) atomicAdd(&nKeep, 1);
            else curMatch[c] = -1;  // mvpMapPoints[i2] = NULL for every entry of a rejected bin
        }
        lastKeep[i] = (uint8_t)keep;
    }
    __syncthreads();
    if (tid == 0) result[0] = nKeep;
}

// ---- mode 1: ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*>&, th), src/ORBmatcher.cc:43-127 --------
// The accept rule needs the best and the second-best candidate and the pyramid level of each, where "second" follows
// the reference's update order (a demoted best replaces the current second even on equal distance), so each query
// replays the reference's scan over its short candidate list sequentially — one thread per query.
__global__ void __launch_bounds__(256) k_sp_choose_mp(SpDev S, const int* __restrict__ candOff,
                                                      const int* __restrict__ candIdx,
                                                      const unsigned short* __restrict__ candDist, int* __restrict__ choice,
                                                      const int* __restrict__ owner, int* __restrict__ flag, int round) {
    if (flag[round - 1] == 0) return;  // flag[-1]: the call's go flag (0 = candidate buffer too small, see k_sp_guard)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S.nLast) return;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int c = candOff[i], e = candOff[i + 1]; c < e; c++) {
        const int idx = candIdx[c];
        if (owner[idx] < i) continue;  // F.mvpMapPoints[idx] was set by an earlier map point with observations
        const int dist = candDist[c];
        if (dist < bestDist) {
            bestDist2 = bestDist;
            bestDist = dist;
            bestLevel2 = bestLevel;
            bestLevel = S.keys[idx].octave;
            bestIdx = idx;
        } else if (dist < bestDist2) {
            bestLevel2 = S.keys[idx].octave;
            bestDist2 = dist;
        }
    }
    int nc = -1;
    if (bestDist <= S.p.th_high) {
        if (!(bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(S.nnRatio, (float)bestDist2))) nc = bestIdx;
    }
    if (nc != choice[i]) {
        choice[i] = nc;
        flag[round] = 1;
    }
}

__global__ void __launch_bounds__(1024) k_sp_finish_mp(SpDev S, const int* __restrict__ choice, int32_t* __restrict__ curMatch,
                                                       int32_t* __restrict__ mpMatch, int* __restrict__ result) {
    __shared__ int nAcc;
    const int tid = threadIdx.x, T = blockDim.x;
    if (tid == 0) nAcc = 0;
    for (int k = tid; k < S.nCur; k += T) curMatch[k] = -1;
    __syncthreads();
    int mine = 0;
    for (int i = tid; i < S.nLast; i += T) {
        const int c = choice[i];
        mpMatch[i] = c;
        if (c >= 0) {
            atomicMax(&curMatch[c], i);  // a later claim overwrites an earlier one
            mine++;
        }
    }
    if (mine) atomicAdd(&nAcc, mine);
    __syncthreads();
    if (tid == 0) result[0] = nAcc;
}

// ---- mode 3: ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:400-515 ---------------------------------------------
// Here a later query can STEAL a keypoint from an earlier one (vMatchedDistance / vnMatches21, :439, :458-466), so what a
// query sees is the running minimum over the earlier accepted claims and not a first-claimer mark. The candidate lists
// and distances come from the parallel kernels above; the claim loop itself is replayed in query order by ONE warp whose
// lanes share each query's candidate list (best / second best with multiplicity, first position on ties, merged by
// shuffles). Only level-0 keypoints query (a few hundred), so the replay is short. The rotation histogram counts every
// accept, stolen or not, as the reference's rotHist does (:468-478).
// The replay is latency-critical (one warp, query after query), so nothing state-independent is loaded on the critical
// path: the offsets of 32 queries arrive with one coalesced load, the first 64 candidates (distance, keypoint) of the next
// non-empty query are fetched while the current one is resolved, and with SMEM the claim state (vMatchedDistance,
// vnMatches21) lives in shared memory.
#define INIT_STAGE 4096  // candidates per staging buffer (two buffers: 2 x 24 KB)
template <bool SMEM>
__global__ void __launch_bounds__(32) k_init_greedy(SpDev S, const int* __restrict__ candOff, const int* __restrict__ candIdx,
                                                    const unsigned short* __restrict__ candDist, int thLow,
                                                    int* gMatchedDist /* [nCur] preset >= INT_MAX/2 (SMEM = false) */,
                                                    int* gMatch21 /* [nCur] preset -1 (SMEM = false) */,
                                                    int32_t* match12 /* [nLast] preset -1 */,
                                                    int* bin12 /* [nLast] preset -1 */, float* prevOut /* [nLast][2] */,
                                                    int* __restrict__ result) {
    extern __shared__ int initSm[];
    __shared__ int hist[SP_HISTO];
    __shared__ int keepBin[SP_HISTO];
    const int lane = threadIdx.x;
    if (lane < SP_HISTO) { hist[lane] = 0; keepBin[lane] = 1; }
    volatile int* vMatchedDistance = SMEM ? initSm : gMatchedDist;
    volatile int* vnMatches21 = SMEM ? initSm + S.nCur : gMatch21;
    if (SMEM)
        for (int k = lane; k < S.nCur; k += 32) { vMatchedDistance[k] = INT_MAX; vnMatches21[k] = -1; }
    __syncwarp();
    int nm = 0;  // lane 0's count
    const float factor = 1.0f / SP_HISTO;
    const int nL = S.nLast;
    // Offsets of the queries of a block (lane l <-> query base + l) are loaded two blocks ahead. The candidates of 32
    // consecutive queries are one contiguous CSR range: the range of the NEXT block is copied into shared memory with
    // cp.async while the current block is replayed, so the replay itself reads no global memory.
    const int total = candOff[nL];
    int* stIdx[2] = {initSm + (SMEM ? 2 * S.nCur : 0), initSm + (SMEM ? 2 * S.nCur : 0) + INIT_STAGE};
    unsigned short* stDist[2] = {(unsigned short*)(stIdx[1] + INIT_STAGE), (unsigned short*)(stIdx[1] + INIT_STAGE) + INIT_STAGE + 2};
    auto loadOff = [&](int base, int& a, int& b) {
        const int q = base + lane;
        a = b = total;
        if (q < nL) { a = candOff[q]; b = candOff[q + 1]; }
    };
    auto stage = [&](int buf, int bs, int be) {  // candidates [bs, min(be, bs + INIT_STAGE)) -> buffer `buf`
        be = min(be, bs + INIT_STAGE);
        for (int c = bs + lane; c < be; c += 32) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(stIdx[buf] + (c - bs));
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(candIdx + c) : "memory");
        }
        const int w0 = bs >> 1, w1 = (be + 1) >> 1;  // distances travel as 4-byte words (two candidates)
        for (int w = w0 + lane; w < w1; w += 32) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared((unsigned*)stDist[buf] + (w - w0));
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"((const unsigned*)candDist + w) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int qs, qe, nqs, nqe;
    loadOff(0, qs, qe);
    loadOff(32, nqs, nqe);
    stage(0, __shfl_sync(0xffffffffu, qs, 0), __shfl_sync(0xffffffffu, qe, 31));
    for (int base = 0, buf = 0; base < nL; base += 32, buf ^= 1) {
        int n2s, n2e;
        loadOff(base + 64, n2s, n2e);
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncwarp();
        stage(buf ^ 1, __shfl_sync(0xffffffffu, nqs, 0), __shfl_sync(0xffffffffu, nqe, 31));
        const int bs = __shfl_sync(0xffffffffu, qs, 0);
        const int stEnd = min(__shfl_sync(0xffffffffu, qe, 31), bs + INIT_STAGE);
        const int* sIdx = stIdx[buf];
        const unsigned short* sDist = stDist[buf] - (bs & ~1);
        unsigned pending = __ballot_sync(0xffffffffu, qs < qe);
        while (pending) {
            const int q = __ffs(pending) - 1;
            pending &= pending - 1;
            const int i = base + q;
            const int s = __shfl_sync(0xffffffffu, qs, q), e = __shfl_sync(0xffffffffu, qe, q);
            int d[4], ix[4], md[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int c = s + lane + 32 * k;
                d[k] = -1;
                ix[k] = 0;
                if (c < e) {
                    if (c < stEnd) { d[k] = sDist[c]; ix[k] = sIdx[c - bs]; }
                    else { d[k] = candDist[c]; ix[k] = candIdx[c]; }
                }
            }
            // lane-local best / second best over its candidates in list order (positions s + lane + 32k)
            int b = INT_MAX, b2 = INT_MAX, bp = 0, bi = -1;
#pragma unroll
            for (int k = 0; k < 4; k++) md[k] = vMatchedDistance[ix[k]];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (d[k] < 0 || md[k] <= d[k]) continue;
                if (d[k] < b) { b2 = b; b = d[k]; bp = lane + 32 * k; bi = ix[k]; }
                else if (d[k] < b2) b2 = d[k];
            }
            for (int c0 = s + 128; c0 < e; c0 += 128) {  // long lists: further groups of 128, loads issued together
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int c = c0 + lane + 32 * k;
                    d[k] = -1;
                    ix[k] = 0;
                    if (c < e) {
                        if (c < stEnd) { d[k] = sDist[c]; ix[k] = sIdx[c - bs]; }
                        else { d[k] = candDist[c]; ix[k] = candIdx[c]; }
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; k++) md[k] = vMatchedDistance[ix[k]];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    if (d[k] < 0 || md[k] <= d[k]) continue;
                    if (d[k] < b) { b2 = b; b = d[k]; bp = c0 - s + lane + 32 * k; bi = ix[k]; }
                    else if (d[k] < b2) b2 = d[k];
                }
            }
            // warp merge: the winner is the smallest (distance, position); the second best is the smallest of every
            // other lane's best and the winner lane's own second (two REDUX + one shuffle)
            const unsigned key = b == INT_MAX ? 0xFFFFFFFFu : ((unsigned)b << 22) | (unsigned)bp;
            const unsigned wkey = __reduce_min_sync(0xffffffffu, key);
            const int wlane = (int)(wkey & 31u);  // position = lane + 32k
            b2 = (int)__reduce_min_sync(0xffffffffu, (unsigned)(lane == wlane ? b2 : b));
            bi = __shfl_sync(0xffffffffu, bi, wlane);
            b = wkey == 0xFFFFFFFFu ? INT_MAX : (int)(wkey >> 22);
            if (b <= thLow && (float)b < __fmul_rn((float)b2, S.nnRatio)) {
                if (lane == 0) {
                    const int bestIdx2 = bi;
                    const int prev = vnMatches21[bestIdx2];
                    int inc = 1;
                    if (prev >= 0) { match12[prev] = -1; inc = 0; }
                    match12[i] = bestIdx2;
                    vnMatches21[bestIdx2] = i;
                    vMatchedDistance[bestIdx2] = b;
                    nm += inc;
                    bin12[i] = bestIdx2;  // the accept is recorded; its histogram bin is computed after the replay
                }
                __syncwarp();
            }
        }
        qs = nqs; qe = nqe; nqs = n2s; nqe = n2e;
    }
    __syncwarp();
    if (S.p.check_orientation)  // rotHist of every accept (a query is accepted at most once), off the critical path
        for (int i = lane; i < nL; i += 32) {
            const int a = bin12[i];
            if (a < 0) continue;
            float rot = __fsub_rn(S.qkeys[i].angle, S.keys[a].angle);
            if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
            int bin = (int)roundf(__fmul_rn(rot, factor));
            if (bin == SP_HISTO) bin = 0;
            atomicAdd(&hist[bin], 1);
            bin12[i] = bin;
        }
    __syncwarp();
    if (lane == 0 && S.p.check_orientation) {  // ComputeThreeMaxima, :1597-1638
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < SP_HISTO; i++) {
            const int s = hist[i];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < 0.1f * (float)max1) { ind3 = -1; }
        for (int i = 0; i < SP_HISTO; i++) keepBin[i] = (i == ind1 || i == ind2 || i == ind3);
    }
    __syncwarp();
    int dropped = 0;
    for (int i = lane; i < nL; i += 32) {
        int c = match12[i];
        if (c >= 0 && S.p.check_orientation && !keepBin[bin12[i]]) {
            match12[i] = c = -1;
            dropped++;
        }
        float x = S.projX[2 * i], y = S.projX[2 * i + 1];
        if (c >= 0) {  // vbPrevMatched[i1] = F2.mvKeysUn[vnMatches12[i1]].pt, :508-510
            x = S.keys[c].x;
            y = S.keys[c].y;
        }
        prevOut[2 * i] = x;
        prevOut[2 * i + 1] = y;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dropped += __shfl_xor_sync(0xffffffffu, dropped, o);
    if (lane == 0) result[0] = nm - dropped;
}


// ---- ORBmatcher::Fuse: the search half (src/ORBmatcher.cc:821-971 and :973-1096) --------------------------------------
// Per map point (projected on the host, where the MapPoint-internal gates live): KeyFrame::GetFeaturesInArea(u, v,
// th*scale[level]) (src/KeyFrame.cc:563-602), octave in [level-1, level], the chi-square reprojection gates of the
// first overload (stereo 7.8 / mono 5.99 on e2 * invSigma2[octave]; chi2 = 0 for the Sim3 overload), best Hamming
// distance with the first candidate winning ties. No query depends on another one: the map mutations (Replace /
// AddObservation) are applied by the caller in point order from (best_idx, best_dist).
struct FuseDev {
    int n;
    const float *u, *v, *ur;
    const int32_t* level;
    const uint8_t* valid;
    const uint4* desc;
    float invSigma2[ORBX_MAX_LEVELS];
    int chi2;
};

__global__ void __launch_bounds__(256) k_fuse_search(SpDev S, FuseDev F, const int* __restrict__ cellStart,
                                                     const int* __restrict__ items, int32_t* __restrict__ bestIdx,
                                                     unsigned short* __restrict__ bestDist) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= F.n) return;
    unsigned best = 0xFFFFFFFFu;  // (distance << 22) | visiting position
    int bestI = -1;
    if (F.valid[i]) {
        const float u = F.u[i], v = F.v[i], ur = F.ur[i];
        const int lvl = F.level[i];
        const float r = __fmul_rn(S.p.th, S.p.scale[lvl]);
        const int x0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, S.p.minX), r), S.p.grid_w_inv)));
        const int x1 = min(SP_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, S.p.minX), r), S.p.grid_w_inv)));
        const int y0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, S.p.minY), r), S.p.grid_h_inv)));
        const int y1 = min(SP_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, S.p.minY), r), S.p.grid_h_inv)));
        if (x0 < SP_COLS && x1 >= 0 && y0 < SP_ROWS && y1 >= 0) {
            const uint4 d0 = F.desc[2 * (size_t)i], d1 = F.desc[2 * (size_t)i + 1];
            int pos = 0;
            for (int ix = x0; ix <= x1; ix++) {
                const int s = cellStart[ix * SP_ROWS + y0], e = cellStart[ix * SP_ROWS + y1 + 1];
                for (int b = s; b < e; b += 32, pos += 32) {
                    const int k = b + lane;
                    if (k >= e) continue;
                    const int i2 = items[k];
                    const orbx_kp kp = S.keys[i2];
                    if (!(fabsf(__fsub_rn(kp.x, u)) < r && fabsf(__fsub_rn(kp.y, v)) < r)) continue;
                    if (kp.octave < lvl - 1 || kp.octave > lvl) continue;
                    if (F.chi2) {
                        const float ex = __fsub_rn(u, kp.x), ey = __fsub_rn(v, kp.y);
                        const float kr = S.uRight ? S.uRight[i2] : -1.f;
                        if (kr >= 0) {
                            const float er = __fsub_rn(ur, kr);
                            const float e2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(er, er));
                            if ((double)__fmul_rn(e2, F.invSigma2[kp.octave]) > 7.8) continue;
                        } else {
                            const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                            if ((double)__fmul_rn(e2, F.invSigma2[kp.octave]) > 5.99) continue;
                        }
                    }
                    const unsigned d = (unsigned)hamming256(d0, d1, S.cdesc[2 * (size_t)i2], S.cdesc[2 * (size_t)i2 + 1]);
                    const unsigned key = (d << 22) | (unsigned)(pos + lane);
                    if (key < best) { best = key; bestI = i2; }
                }
            }
        }
    }
    // warp arg-min: smallest distance, first in visiting order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bestI, o);
        if (ob < best) { best = ob; bestI = oi; }
    }
    if (lane == 0) {
        bestIdx[i] = bestI;
        bestDist[i] = bestI >= 0 ? (unsigned short)(best >> 22) : (unsigned short)256;
    }
}

// ---- batched MapPoint::ComputeDistinctiveDescriptors, /root/reference/src/MapPoint.cc:242-307 -------------------------
// Per map point: all-pairs Hamming distances of its observations' descriptors, per row the median
// vDists[0.5*(N-1)] of the sorted row (self distance 0 included), result = first row with the least median.
// One CTA per map point; its descriptors sit in shared memory; a thread owns a row and finds the k-th smallest
// distance by bisection on the value (distances are 0..256: 9 counting passes over broadcast reads, no sort, no N x N
// matrix).
#define DD_THREADS 64
#define DD_MAX_OBS 1024

__global__ void __launch_bounds__(DD_THREADS) k_distinctive(const uint4* __restrict__ desc, const int* __restrict__ off,
                                                            int npoints, int32_t* __restrict__ bestIdx,
                                                            int32_t* __restrict__ bestMedian) {
    extern __shared__ uint4 dd_rows[];  // 2 x N
    __shared__ unsigned red[DD_THREADS / 32];
    const int p = blockIdx.x, tid = threadIdx.x;
    const int s = off[p], N = off[p + 1] - s;
    if (N <= 0) {
        if (tid == 0) {
            bestIdx[p] = -1;
            bestMedian[p] = 0x7fffffff;
        }
        return;
    }
    for (int i = tid; i < 2 * N; i += DD_THREADS) dd_rows[i] = desc[2 * (size_t)s + i];
    __syncthreads();
    const int k = (int)(0.5 * (N - 1));  // index into the sorted row, as the reference computes it
    unsigned best = 0xFFFFFFFFu;         // (median << 16) | row: the minimum is the least median, first row on ties
    for (int i = tid; i < N; i += DD_THREADS) {
        const uint4 a0 = dd_rows[2 * i], a1 = dd_rows[2 * i + 1];
        int lo = 0, hi = 256;  // smallest v with #{j : d(i,j) <= v} >= k + 1
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = 0; j < N; j++) cnt += hamming256(a0, a1, dd_rows[2 * j], dd_rows[2 * j + 1]) <= mid;
            if (cnt >= k + 1) hi = mid;
            else lo = mid + 1;
        }
        best = min(best, ((unsigned)lo << 16) | (unsigned)i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((tid & 31) == 0) red[tid >> 5] = best;
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < DD_THREADS / 32; w++) best = min(best, red[w]);
        bestIdx[p] = (int)(best & 0xFFFFu);
        bestMedian[p] = (int)(best >> 16);
    }
}
