// orbx_bowsearch.cuh — device-resident ORBmatcher::SearchByFboW (included at the end of orbx_match.cu).
//
// Replaces /root/reference/src/ORBmatcher.cc:157-283 (KeyFrame vs Frame: TrackReferenceKeyFrame, Relocalization) and
// :517-650 (KeyFrame vs KeyFrame: LoopClosing). Both walk the two feature vectors (fBow2: node -> feature indices) in
// node order; for every eligible feature of side A in a common node they scan side B's features of the same node with
// the best / second-best rule, skip B features already matched by an EARLIER A feature, accept on
// `best <= th` (`<` for KF-KF) and `(float)best < ratio * (float)second`, then filter by the rotation histogram.
//
// The sequential "already matched" dependence is resolved as in orbx_search.cuh: a fixed point over rounds in which
// every query replays its scan while skipping candidates claimed in the previous round by a query of lower rank
// (rank = position in A's feature vector = the reference's visiting order). By induction on the rank the fixed point is
// the sequential result. Node lists are short, so a thread per query computes its Hamming distances on the fly.
#pragma once

struct BowDev {
    int nA, nB, nPosA;                // features; entries of A's feature vector
    const uint8_t *eligA, *eligB;     // eligB may be NULL (every B feature is a candidate)
    const float *angA, *angB;
    const uint4 *descA, *descB;
    int nNodesA, nNodesB;
    const uint32_t *nodeIdA, *nodeIdB, *featA, *featB;
    const int *offA, *offB;
    float nnRatio;
    int th, strict, checkOri;
    // SearchForTriangulation only (tri = 1): undistorted keypoint positions, octaves of B, stereo flags, F12, epipole
    int tri;
    const float2 *ptA, *ptB;
    const int32_t* octB;
    const uint8_t *stereoA, *stereoB;
    float F[9], ex, ey, scale[ORBX_MAX_LEVELS], sigma2[ORBX_MAX_LEVELS];
};

// per entry p of A's feature vector: the matching node of B (or -1), found once
__device__ __forceinline__ void d_bow_pair_nodes(const BowDev& S, int* __restrict__ posNodeB, int j /* node of A */) {
    if (j >= S.nNodesA) return;
    const uint32_t id = S.nodeIdA[j];
    int lo = 0, hi = S.nNodesB - 1, hit = -1;
    while (lo <= hi) {
        const int mid = (lo + hi) >> 1;
        const uint32_t v = S.nodeIdB[mid];
        if (v == id) { hit = mid; break; }
        if (v < id) lo = mid + 1; else hi = mid - 1;
    }
    for (int p = S.offA[j]; p < S.offA[j + 1]; p++) posNodeB[p] = hit;
}
__global__ void k_bow_pair_nodes(BowDev S, int* __restrict__ posNodeB) {
    d_bow_pair_nodes(S, posNodeB, blockIdx.x * blockDim.x + threadIdx.x);
}

__device__ __forceinline__ void d_bow_claim(const BowDev& S, const int* __restrict__ choice, int* __restrict__ owner, int p) {
    if (p >= S.nPosA) return;
    const int c = choice[p];
    if (c >= 0) atomicMin(&owner[c], p);
}
__global__ void k_bow_claim(BowDev S, const int* __restrict__ choice, int* __restrict__ owner, const int* __restrict__ flag,
                            int round) {
    if (round > 0 && flag[round - 1] == 0) return;
    d_bow_claim(S, choice, owner, blockIdx.x * blockDim.x + threadIdx.x);
}

__device__ __forceinline__ void d_bow_choose(const BowDev& S, const int* __restrict__ posNodeB, int* __restrict__ choice,
                                             const int* __restrict__ owner, int* __restrict__ flag, int round, int p) {
    if (p >= S.nPosA) return;
    int nc = -1;
    const int jb = posNodeB[p];
    const uint32_t ia = S.featA[p];
    if (jb >= 0 && S.eligA[ia]) {
        const uint4 a0 = S.descA[2 * (size_t)ia], a1 = S.descA[2 * (size_t)ia + 1];
        int best1 = 256, best2 = 256, bestIdx = -1;
        for (int c = S.offB[jb], e = S.offB[jb + 1]; c < e; c++) {
            const uint32_t ib = S.featB[c];
            if (owner[ib] < p) continue;                 // vpMapPointMatches[realIdxF] / vbMatched2[idx2] set earlier
            if (S.eligB && !S.eligB[ib]) continue;       // !pMP2 || pMP2->isBad()
            const int d = hamming256(a0, a1, S.descB[2 * (size_t)ib], S.descB[2 * (size_t)ib + 1]);
            if (d < best1) { best2 = best1; best1 = d; bestIdx = (int)ib; }
            else if (d < best2) best2 = d;
        }
        const bool in = S.strict ? best1 < S.th : best1 <= S.th;
        if (in && (float)best1 < __fmul_rn(S.nnRatio, (float)best2)) nc = bestIdx;
    }
    if (nc != choice[p]) {
        choice[p] = nc;
        flag[round] = 1;
    }
}
__global__ void __launch_bounds__(128) k_bow_choose(BowDev S, const int* __restrict__ posNodeB, int* __restrict__ choice,
                                                    const int* __restrict__ owner, int* __restrict__ flag, int round) {
    if (round > 0 && flag[round - 1] == 0) return;
    d_bow_choose(S, posNodeB, choice, owner, flag, round, blockIdx.x * blockDim.x + threadIdx.x);
}

// ORBmatcher::SearchForTriangulation's scan (src/ORBmatcher.cc:697-749): best = TH_LOW; a candidate replaces the best
// when dist <= best (equal distance: the LAST one wins) and it passes the epipole and epipolar-line gates
// (CheckDistEpipolarLine, :138-155; float arithmetic in the reference's order, no FMA contraction).
__device__ __forceinline__ void d_tri_choose(const BowDev& S, const int* __restrict__ posNodeB, int* __restrict__ choice,
                                             const int* __restrict__ owner, int* __restrict__ flag, int round, int p) {
    if (p >= S.nPosA) return;
    int nc = -1;
    const int jb = posNodeB[p];
    const uint32_t ia = S.featA[p];
    if (jb >= 0 && S.eligA[ia]) {
        const uint4 a0 = S.descA[2 * (size_t)ia], a1 = S.descA[2 * (size_t)ia + 1];
        const float2 k1 = S.ptA[ia];
        const bool st1 = S.stereoA[ia] != 0;
        // epipolar line in the second image l = x1' F12 = [a b c]
        const float la = __fadd_rn(__fadd_rn(__fmul_rn(k1.x, S.F[0]), __fmul_rn(k1.y, S.F[3])), S.F[6]);
        const float lb = __fadd_rn(__fadd_rn(__fmul_rn(k1.x, S.F[1]), __fmul_rn(k1.y, S.F[4])), S.F[7]);
        const float lc = __fadd_rn(__fadd_rn(__fmul_rn(k1.x, S.F[2]), __fmul_rn(k1.y, S.F[5])), S.F[8]);
        const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
        int best = S.th;
        for (int c = S.offB[jb], e = S.offB[jb + 1]; c < e; c++) {
            const uint32_t ib = S.featB[c];
            if (owner[ib] < p) continue;     // vbMatched2[idx2]
            if (!S.eligB[ib]) continue;      // pMP2 (and the bOnlyStereo gate, folded in by the caller)
            const int d = hamming256(a0, a1, S.descB[2 * (size_t)ib], S.descB[2 * (size_t)ib + 1]);
            if (d > S.th || d > best) continue;
            const float2 k2 = S.ptB[ib];
            const int oct = S.octB[ib];
            if (!st1 && !S.stereoB[ib]) {
                const float dx = __fsub_rn(S.ex, k2.x), dy = __fsub_rn(S.ey, k2.y);
                if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, S.scale[oct])) continue;
            }
            const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, k2.x), __fmul_rn(lb, k2.y)), lc);
            if (den == 0.f) continue;
            const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
            if ((double)dsqr < 3.84 * (double)S.sigma2[oct]) { nc = (int)ib; best = d; }
        }
    }
    if (nc != choice[p]) {
        choice[p] = nc;
        flag[round] = 1;
    }
}
__global__ void __launch_bounds__(128) k_tri_choose(BowDev S, const int* __restrict__ posNodeB, int* __restrict__ choice,
                                                    const int* __restrict__ owner, int* __restrict__ flag, int round) {
    if (round > 0 && flag[round - 1] == 0) return;
    d_tri_choose(S, posNodeB, choice, owner, flag, round, blockIdx.x * blockDim.x + threadIdx.x);
}

__device__ __forceinline__ void d_bow_finish(const BowDev& S, const int* __restrict__ choice, int32_t* __restrict__ matchA,
                                             int* __restrict__ result) {
    __shared__ int hist[SP_HISTO];
    __shared__ int keepBin[SP_HISTO];
    __shared__ int nKeep;
    const int tid = threadIdx.x, T = blockDim.x;
    if (tid < SP_HISTO) { hist[tid] = 0; keepBin[tid] = 1; }
    if (tid == 0) nKeep = 0;
    for (int i = tid; i < S.nA; i += T) matchA[i] = -1;
    __syncthreads();
    const float factor = 1.0f / SP_HISTO;
    if (S.checkOri) {
        for (int p = tid; p < S.nPosA; p += T) {
            const int c = choice[p];
            if (c < 0) continue;
            float rot = __fsub_rn(S.angA[S.featA[p]], S.angB[c]);
            if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
            int bin = (int)roundf(__fmul_rn(rot, factor));
            if (bin == SP_HISTO) bin = 0;
            atomicAdd(&hist[bin], 1);
        }
        __syncthreads();
        if (tid == 0) {  // ComputeThreeMaxima, src/ORBmatcher.cc:1597-1638
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
    }
    int mine = 0;
    for (int p = tid; p < S.nPosA; p += T) {
        const int c = choice[p];
        if (c < 0) continue;
        int keep = 1;
        if (S.checkOri) {
            float rot = __fsub_rn(S.angA[S.featA[p]], S.angB[c]);
            if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
            int bin = (int)roundf(__fmul_rn(rot, factor));
            if (bin == SP_HISTO) bin = 0;
            keep = keepBin[bin];
        }
        if (keep) { matchA[S.featA[p]] = c; mine++; }
    }
    if (mine) atomicAdd(&nKeep, mine);
    __syncthreads();
    if (tid == 0) result[0] = nKeep;
}
__global__ void __launch_bounds__(1024) k_bow_finish(BowDev S, const int* __restrict__ choice, int32_t* __restrict__ matchA,
                                                     int* __restrict__ result) {
    d_bow_finish(S, choice, matchA, result);
}

// ---- one pair, one CTA: a single SearchByFboW / SearchForTriangulation call on camera-sized frames (~1000 features) is
// launch-latency bound as a kernel sequence (node pairing, 3-4 rounds of clear / claim / choose, finish: ~16 launches).
// Here one 1024-thread CTA runs the same device functions back to back with __syncthreads between the phases and loops
// until a round changes nothing — the same fixed point, one launch. result[1] = rounds executed.
__global__ void __launch_bounds__(1024) k_bow_fused(BowDev S, int* __restrict__ posNodeB, int* __restrict__ choice,
                                                    int* __restrict__ owner, int32_t* __restrict__ matchA,
                                                    int* __restrict__ result, int maxRounds) {
    __shared__ int changed[1];
    const int tid = threadIdx.x, T = blockDim.x;
    for (int j = tid; j < S.nNodesA; j += T) d_bow_pair_nodes(S, posNodeB, j);
    for (int p = tid; p < S.nPosA; p += T) choice[p] = -1;
    int round = 0;
    for (;; round++) {
        for (int i = tid; i < S.nB; i += T) owner[i] = 0x7F7F7F7F;
        if (tid == 0) changed[0] = 0;
        __syncthreads();
        for (int p = tid; p < S.nPosA; p += T) d_bow_claim(S, choice, owner, p);
        __syncthreads();
        for (int p = tid; p < S.nPosA; p += T) {
            if (S.tri) d_tri_choose(S, posNodeB, choice, owner, changed, 0, p);
            else d_bow_choose(S, posNodeB, choice, owner, changed, 0, p);
        }
        __syncthreads();
        const int ch = changed[0];
        __syncthreads();
        if (!ch || round + 1 >= maxRounds) break;
    }
    d_bow_finish(S, choice, matchA, result);
    if (tid == 0) result[1] = round + 1;
}

// ---- batched form: one launch sequence serves npairs independent (A, B) pairs (blockIdx.y = pair) -----------------------
// Relocalization matches the current frame against every candidate keyframe (src/Tracking.cc:1465-1490), LoopClosing a
// keyframe against every loop candidate (src/LoopClosing.cc:290-312): a single pair is latency bound on the GPU, the
// batch amortises the launches and the fixed-point rounds (a round ends for all pairs when no pair changed).
struct BowPair {
    BowDev S;
    int *posNodeB, *choice, *owner;
    int32_t* oMatch;
    int* result;
};
__global__ void k_bowb_pair_nodes(const BowPair* __restrict__ P) {
    const BowPair& Q = P[blockIdx.y];
    d_bow_pair_nodes(Q.S, Q.posNodeB, blockIdx.x * blockDim.x + threadIdx.x);
}
__global__ void k_bowb_claim(const BowPair* __restrict__ P, const int* __restrict__ flag, int round) {
    if (round > 0 && flag[round - 1] == 0) return;
    const BowPair& Q = P[blockIdx.y];
    d_bow_claim(Q.S, Q.choice, Q.owner, blockIdx.x * blockDim.x + threadIdx.x);
}
__global__ void __launch_bounds__(128) k_bowb_choose(const BowPair* __restrict__ P, int* __restrict__ flag, int round) {
    if (round > 0 && flag[round - 1] == 0) return;
    const BowPair& Q = P[blockIdx.y];
    if (Q.S.tri) d_tri_choose(Q.S, Q.posNodeB, Q.choice, Q.owner, flag, round, blockIdx.x * blockDim.x + threadIdx.x);
    else d_bow_choose(Q.S, Q.posNodeB, Q.choice, Q.owner, flag, round, blockIdx.x * blockDim.x + threadIdx.x);
}
__global__ void __launch_bounds__(1024) k_bowb_finish(const BowPair* __restrict__ P) {
    const BowPair& Q = P[blockIdx.y];
    d_bow_finish(Q.S, Q.choice, Q.oMatch, Q.result);
}

static int bow_side_ok(const orbm_bow_side* s) {
    if (s->n < 0 || s->n_nodes < 0) return 0;
    if (s->n && (!s->angle || !s->desc)) return 0;
    if (s->n_nodes && (!s->node_id || !s->node_off || !s->node_feat)) return 0;
    return 1;
}

static int bow_run(orbm_t* m, const orbm_bow_side* a, const orbm_bow_side* b, int n_pos_a, int n_pos_b, float nn_ratio,
                   int th, int strict, int check_orientation, const orbm_triangulation* T, int32_t* match_a,
                   int* nmatches) {
    if (!m || !a || !b || !nmatches || (a->n && !match_a)) FAIL(ORBX_EINVAL, "null argument");
    if (!bow_side_ok(a) || !bow_side_ok(b) || (a->n && !a->eligible) || n_pos_a < 0 || n_pos_b < 0) FAIL(ORBX_EINVAL, "bad feature-vector arrays");
    *nmatches = 0;
    CK(cudaSetDevice(m->device));
    cudaStream_t st = m->stream;
    const bool hin = m->inSpace == ORBX_MEM_HOST, hout = m->outSpace == ORBX_MEM_HOST;
    const int nA = a->n, nB = b->n;
    if (nA == 0) return ORBX_OK;
    if (nB == 0 || a->n_nodes == 0 || b->n_nodes == 0 || n_pos_a == 0) {
        if (hout) for (int i = 0; i < nA; i++) match_a[i] = -1;
        else { CK(cudaMemsetAsync(match_a, 0xFF, (size_t)nA * 4, st)); CK(cudaStreamSynchronize(st)); }
        return ORBX_OK;
    }
    const int maxRounds = n_pos_a + 8;
    if (T && (!T->pt_a || !T->pt_b || !T->octave_b || !T->stereo_a || !T->stereo_b || !b->eligible || T->nlevels < 1 ||
              T->nlevels > ORBX_MAX_LEVELS))
        FAIL(ORBX_EINVAL, "bad triangulation arguments");
    const size_t need = (size_t)nA * (1 + 4 + 32 + 4 + 8 + 1) + (size_t)nB * (1 + 4 + 32 + 4 + 8 + 4 + 1) + (size_t)(a->n_nodes + b->n_nodes) * 8 +
                        (size_t)(n_pos_a + n_pos_b) * 4 + (size_t)n_pos_a * 8 + (size_t)maxRounds * 4 + 256 * 24;
    int rc;
    if ((rc = ensure(m->dWs, m->wsCap, need))) return rc;
    if (!m->hRes) CK(cudaMallocHost(&m->hRes, 4 * sizeof(int)));
    if (m->hFlagsCap < (size_t)maxRounds) {
        if (m->hFlags) cudaFreeHost(m->hFlags);
        m->hFlags = nullptr;
        CK(cudaMallocHost(&m->hFlags, (size_t)maxRounds * sizeof(int)));
        m->hFlagsCap = (size_t)maxRounds;
    }
    WsCursor W{m->dWs, 0};
    if ((rc = ensure_host_in(m, need))) return rc;
    Stager stg{m, &W, hin, (size_t)-1};
    auto in = [&](const void* src, size_t bytes) -> const void* { return stg.in(src, bytes); };
    BowDev S;
    S.nA = nA; S.nB = nB; S.nPosA = n_pos_a;
    S.eligA = (const uint8_t*)in(a->eligible, nA);
    S.eligB = (const uint8_t*)in(b->eligible, nB);
    S.angA = (const float*)in(a->angle, (size_t)nA * 4);
    S.angB = (const float*)in(b->angle, (size_t)nB * 4);
    S.descA = (const uint4*)in(a->desc, (size_t)nA * 32);
    S.descB = (const uint4*)in(b->desc, (size_t)nB * 32);
    S.nNodesA = a->n_nodes; S.nNodesB = b->n_nodes;
    S.nodeIdA = (const uint32_t*)in(a->node_id, (size_t)a->n_nodes * 4);
    S.nodeIdB = (const uint32_t*)in(b->node_id, (size_t)b->n_nodes * 4);
    S.offA = (const int*)in(a->node_off, (size_t)(a->n_nodes + 1) * 4);
    S.offB = (const int*)in(b->node_off, (size_t)(b->n_nodes + 1) * 4);
    S.featA = (const uint32_t*)in(a->node_feat, (size_t)n_pos_a * 4);
    S.featB = (const uint32_t*)in(b->node_feat, (size_t)n_pos_b * 4);
    S.nnRatio = nn_ratio; S.th = th; S.strict = strict; S.checkOri = check_orientation;
    S.tri = T != nullptr;
    if (T) {
        S.ptA = (const float2*)in(T->pt_a, (size_t)nA * 8);
        S.ptB = (const float2*)in(T->pt_b, (size_t)nB * 8);
        S.octB = (const int32_t*)in(T->octave_b, (size_t)nB * 4);
        S.stereoA = (const uint8_t*)in(T->stereo_a, nA);
        S.stereoB = (const uint8_t*)in(T->stereo_b, nB);
        for (int i = 0; i < 9; i++) S.F[i] = T->F12[i];
        S.ex = T->ex; S.ey = T->ey;
        for (int i = 0; i < T->nlevels; i++) { S.scale[i] = T->scale_factors[i]; S.sigma2[i] = T->level_sigma2[i]; }
    }
    if ((((uintptr_t)S.descA) | ((uintptr_t)S.descB)) & 15) FAIL(ORBX_EINVAL, "device descriptor matrices must be 16-byte aligned");
    if ((rc = stg.flush(st))) return rc;
    int* posNodeB = W.take<int>(n_pos_a);
    int* choice = W.take<int>(n_pos_a);
    int* owner = W.take<int>(nB);
    // what the host reads back is one span: [dRes] [matches (host mode)] [round flags]
    int* dRes = W.take<int>(4);
    int32_t* oMatch = hout ? W.take<int32_t>(nA) : match_a;
    int* flags = W.take<int>(maxRounds);
    if (W.off > m->wsCap) FAIL(ORBX_ECUDA, "internal: workspace estimate too small (%zu > %zu)", W.off, m->wsCap);
    const int SPEC = 4;
    const uint8_t* spanBegin = (const uint8_t*)dRes;
    const size_t spanBytes = (size_t)((const uint8_t*)(flags + SPEC) - spanBegin);
    auto hostAt = [&](const void* dev) { return m->hIn + ((const uint8_t*)dev - spanBegin); };

    if (n_pos_a <= m->bowFusedMax) {  // camera-sized pair: everything in one CTA, one launch, one copy back
        k_bow_fused<<<1, 1024, 0, st>>>(S, posNodeB, choice, owner, oMatch, dRes, maxRounds);
        m->launches++;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(m->hIn, spanBegin, (size_t)((const uint8_t*)flags - spanBegin), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (hout) memcpy(match_a, hostAt(oMatch), (size_t)nA * 4);
        *nmatches = ((const int*)hostAt(dRes))[0];
        m->lastRounds = ((const int*)hostAt(dRes))[1];
        m->lastCandidates = 0;
        return ORBX_OK;
    }
    k_bow_pair_nodes<<<(a->n_nodes + 127) / 128, 128, 0, st>>>(S, posNodeB);
    m->launches++;
    CK(cudaMemsetAsync(choice, 0xFF, (size_t)n_pos_a * sizeof(int), st));
    CK(cudaMemsetAsync(flags, 0, (size_t)maxRounds * sizeof(int), st));
    int round = 0, rounds = 0;
    auto enqueueRounds = [&](int R) {
        for (int r = 0; r < R; r++, round++) {
            cudaMemsetAsync(owner, 0x7F, (size_t)nB * sizeof(int), st);
            k_bow_claim<<<(n_pos_a + 255) / 256, 256, 0, st>>>(S, choice, owner, flags, round);
            if (T) k_tri_choose<<<(n_pos_a + 127) / 128, 128, 0, st>>>(S, posNodeB, choice, owner, flags, round);
            else k_bow_choose<<<(n_pos_a + 127) / 128, 128, 0, st>>>(S, posNodeB, choice, owner, flags, round);
            m->launches += 2;
        }
    };
    auto finishAndFetch = [&]() -> int {
        k_bow_finish<<<1, 1024, 0, st>>>(S, choice, oMatch, dRes);
        m->launches++;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(m->hIn, spanBegin, spanBytes, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return ORBX_OK;
    };
    // SPEC rounds and the finishing kernel go out together; one copy returns matches, count and flags
    if (SPEC > maxRounds) FAIL(ORBX_ECUDA, "internal: round budget");
    enqueueRounds(SPEC);
    if ((rc = finishAndFetch())) return rc;
    const int* hf = (const int*)hostAt(flags);
    while (rounds < SPEC && hf[rounds]) rounds++;
    if (rounds == SPEC) {
        for (;;) {
            // rounds per host synchronisation double (4, 8, 16, ...): a long claim chain (dense repeated texture, every
            // query displaced by the previous one) costs O(log) synchronisations and at most twice the rounds it needs
            const int R = std::min(std::max(4, round), maxRounds - round);
            if (R <= 0) FAIL(ORBX_ECUDA, "internal: greedy resolution did not converge in %d rounds", maxRounds);
            enqueueRounds(R);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(m->hFlags, flags, (size_t)round * sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            rounds = 0;
            while (rounds < round && m->hFlags[rounds]) rounds++;
            if (rounds < round) break;
        }
        if ((rc = finishAndFetch())) return rc;
    }
    if (hout) memcpy(match_a, hostAt(oMatch), (size_t)nA * 4);
    *nmatches = ((const int*)hostAt(dRes))[0];
    m->lastRounds = rounds + 1;
    m->lastCandidates = 0;
    return ORBX_OK;
}

extern "C" int orbm_search_by_bow(orbm_t* m, const orbm_bow_side* a, const orbm_bow_side* b, int n_pos_a, int n_pos_b,
                                  float nn_ratio, int th, int strict, int check_orientation, int32_t* match_a,
                                  int* nmatches) {
    return bow_run(m, a, b, n_pos_a, n_pos_b, nn_ratio, th, strict, check_orientation, nullptr, match_a, nmatches);
}

extern "C" int orbm_search_for_triangulation(orbm_t* m, const orbm_bow_side* a, const orbm_bow_side* b, int n_pos_a,
                                             int n_pos_b, const orbm_triangulation* T, int th, int check_orientation,
                                             int32_t* match_a, int* nmatches) {
    if (!T) FAIL(ORBX_EINVAL, "null argument");
    return bow_run(m, a, b, n_pos_a, n_pos_b, 0.f, th, 0, check_orientation, T, match_a, nmatches);
}


static int bow_batch_run(orbm_t* m, int npairs, const orbm_bow_side* a, const orbm_bow_side* b, const int* n_pos_a,
                         const int* n_pos_b, float nn_ratio, int th, int strict, int check_orientation,
                         const orbm_triangulation* T /* [npairs] or NULL */, int32_t* const* match_a, int* nmatches) {
    if (!m || npairs < 0 || (npairs && (!a || !b || !n_pos_a || !n_pos_b || !match_a || !nmatches)))
        FAIL(ORBX_EINVAL, "null argument");
    if (npairs == 0) return ORBX_OK;
    if (npairs > 65535) FAIL(ORBX_EINVAL, "too many pairs");
    CK(cudaSetDevice(m->device));
    cudaStream_t st = m->stream;
    const bool hin = m->inSpace == ORBX_MEM_HOST, hout = m->outSpace == ORBX_MEM_HOST;
    // sizes; a host array passed for several pairs (the current frame of a relocalisation) is staged once
    std::vector<std::pair<const void*, const void*> > seen;  // (source, device copy)
    size_t need = (size_t)npairs * (sizeof(BowPair) + 256) + 256 * 8;
    int maxPosA = 0, maxNodesA = 0, sumPosA = 0;
    size_t sumA = 0;
    for (int k = 0; k < npairs; k++) {
        const orbm_bow_side *A = a + k, *B = b + k;
        if (!bow_side_ok(A) || !bow_side_ok(B) || (A->n && !A->eligible) || n_pos_a[k] < 0 || n_pos_b[k] < 0 || (A->n && !match_a[k]))
            FAIL(ORBX_EINVAL, "bad feature-vector arrays (pair %d)", k);
        if (T && (!T[k].pt_a || !T[k].pt_b || !T[k].octave_b || !T[k].stereo_a || !T[k].stereo_b || !B->eligible ||
                  T[k].nlevels < 1 || T[k].nlevels > ORBX_MAX_LEVELS))
            FAIL(ORBX_EINVAL, "bad triangulation arguments (pair %d)", k);
        need += (size_t)A->n * (1 + 4 + 32 + 4 + 8 + 1) + (size_t)B->n * (1 + 4 + 32 + 4 + 8 + 4 + 1) + (size_t)(A->n_nodes + B->n_nodes + 2) * 8 +
                (size_t)(n_pos_a[k] + n_pos_b[k]) * 4 + (size_t)n_pos_a[k] * 8 + 256 * 28;
        maxPosA = std::max(maxPosA, n_pos_a[k]);
        maxNodesA = std::max(maxNodesA, A->n_nodes);
        sumPosA += n_pos_a[k];
        sumA += (size_t)A->n;
        nmatches[k] = 0;
    }
    const int maxRounds = maxPosA + 8;
    need += (size_t)maxRounds * 4;
    int rc;
    if ((rc = ensure(m->dWs, m->wsCap, need))) return rc;
    if (m->hFlagsCap < (size_t)std::max(maxRounds, npairs)) {
        if (m->hFlags) cudaFreeHost(m->hFlags);
        m->hFlags = nullptr;
        CK(cudaMallocHost(&m->hFlags, (size_t)std::max(maxRounds, npairs) * sizeof(int)));
        m->hFlagsCap = (size_t)std::max(maxRounds, npairs);
    }
    if ((rc = ensure_host_in(m, need))) return rc;
    WsCursor W{m->dWs, 0};
    BowPair* dPairs = W.take<BowPair>(npairs);
    // what the host reads back is one span: [per-pair counts] [matches of all pairs (host mode)] [round flags]
    int* dRes = W.take<int>(npairs);
    int32_t* matchAll = hout ? W.take<int32_t>(std::max<size_t>(sumA, 1)) : nullptr;
    int* flags = W.take<int>(maxRounds);
    const int SPEC = 4;
    const uint8_t* spanBegin = (const uint8_t*)dRes;
    const size_t spanBytes = (size_t)((const uint8_t*)(flags + SPEC) - spanBegin);
    auto hostAt = [&](const void* dev) { return m->hIn + ((const uint8_t*)dev - spanBegin); };
    Stager stg{m, &W, hin, (size_t)-1};
    auto in = [&](const void* src, size_t bytes) -> const void* {
        if (!src || !hin) return src;
        for (size_t i = 0; i < seen.size(); i++)
            if (seen[i].first == src) return seen[i].second;
        const void* d = stg.in(src, bytes);
        seen.push_back(std::make_pair(src, d));
        return d;
    };
    std::vector<BowPair> hp((size_t)npairs);
    for (int k = 0; k < npairs; k++) {
        const orbm_bow_side *A = a + k, *B = b + k;
        BowDev& S = hp[k].S;
        memset(&S, 0, sizeof(S));
        S.nA = A->n; S.nB = B->n; S.nPosA = n_pos_a[k];
        S.eligA = (const uint8_t*)in(A->eligible, A->n);
        S.eligB = (const uint8_t*)in(B->eligible, B->n);
        S.angA = (const float*)in(A->angle, (size_t)A->n * 4);
        S.angB = (const float*)in(B->angle, (size_t)B->n * 4);
        S.descA = (const uint4*)in(A->desc, (size_t)A->n * 32);
        S.descB = (const uint4*)in(B->desc, (size_t)B->n * 32);
        S.nNodesA = A->n_nodes; S.nNodesB = B->n_nodes;
        S.nodeIdA = (const uint32_t*)in(A->node_id, (size_t)A->n_nodes * 4);
        S.nodeIdB = (const uint32_t*)in(B->node_id, (size_t)B->n_nodes * 4);
        S.offA = (const int*)in(A->node_off, (size_t)(A->n_nodes + 1) * 4);
        S.offB = (const int*)in(B->node_off, (size_t)(B->n_nodes + 1) * 4);
        S.featA = (const uint32_t*)in(A->node_feat, (size_t)n_pos_a[k] * 4);
        S.featB = (const uint32_t*)in(B->node_feat, (size_t)n_pos_b[k] * 4);
        S.nnRatio = nn_ratio; S.th = th; S.strict = strict; S.checkOri = check_orientation;
        S.tri = T != nullptr;
        if (T) {
            const orbm_triangulation& G = T[k];
            S.ptA = (const float2*)in(G.pt_a, (size_t)A->n * 8);
            S.ptB = (const float2*)in(G.pt_b, (size_t)B->n * 8);
            S.octB = (const int32_t*)in(G.octave_b, (size_t)B->n * 4);
            S.stereoA = (const uint8_t*)in(G.stereo_a, A->n);
            S.stereoB = (const uint8_t*)in(G.stereo_b, B->n);
            for (int i = 0; i < 9; i++) S.F[i] = G.F12[i];
            S.ex = G.ex; S.ey = G.ey;
            for (int i = 0; i < G.nlevels; i++) { S.scale[i] = G.scale_factors[i]; S.sigma2[i] = G.level_sigma2[i]; }
        }
        if ((((uintptr_t)S.descA) | ((uintptr_t)S.descB)) & 15) FAIL(ORBX_EINVAL, "device descriptor matrices must be 16-byte aligned");
        // a pair with an empty side or no nodes has no candidates: nPosA = 0 leaves match_a all -1
        if (B->n == 0 || A->n_nodes == 0 || B->n_nodes == 0) { S.nPosA = 0; S.nNodesA = 0; }
    }
    if ((rc = stg.flush(st))) return rc;
    // scratch: choice and owner of all pairs are contiguous so that one memset serves a round
    int* choiceAll = W.take<int>((size_t)std::max(sumPosA, 1));
    size_t sumB = 0;
    for (int k = 0; k < npairs; k++) sumB += (size_t)b[k].n;
    int* ownerAll = W.take<int>(std::max<size_t>(sumB, 1));
    size_t cOff = 0, oOff = 0, mOff = 0;
    for (int k = 0; k < npairs; k++) {
        hp[k].choice = choiceAll + cOff;
        hp[k].owner = ownerAll + oOff;
        cOff += (size_t)n_pos_a[k];
        oOff += (size_t)b[k].n;
        hp[k].posNodeB = W.take<int>((size_t)std::max(n_pos_a[k], 1));
        hp[k].oMatch = hout ? matchAll + mOff : match_a[k];
        mOff += (size_t)a[k].n;
        hp[k].result = dRes + k;
    }
    if (W.off > m->wsCap) FAIL(ORBX_ECUDA, "internal: workspace estimate too small (%zu > %zu)", W.off, m->wsCap);
    CK(cudaMemcpyAsync(dPairs, hp.data(), (size_t)npairs * sizeof(BowPair), cudaMemcpyHostToDevice, st));
    if (maxNodesA) k_bowb_pair_nodes<<<dim3((maxNodesA + 127) / 128, npairs), 128, 0, st>>>(dPairs);
    m->launches++;
    CK(cudaMemsetAsync(choiceAll, 0xFF, (size_t)std::max(sumPosA, 1) * sizeof(int), st));
    CK(cudaMemsetAsync(flags, 0, (size_t)maxRounds * sizeof(int), st));
    int round = 0, rounds = 0;
    auto enqueueRounds = [&](int R) {
        for (int r = 0; r < R; r++, round++) {
            cudaMemsetAsync(ownerAll, 0x7F, std::max<size_t>(sumB, 1) * sizeof(int), st);
            k_bowb_claim<<<dim3((maxPosA + 255) / 256, npairs), 256, 0, st>>>(dPairs, flags, round);
            k_bowb_choose<<<dim3((maxPosA + 127) / 128, npairs), 128, 0, st>>>(dPairs, flags, round);
            m->launches += 2;
        }
    };
    auto finishAndFetch = [&]() -> int {
        k_bowb_finish<<<dim3(1, npairs), 1024, 0, st>>>(dPairs);
        m->launches++;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(m->hIn, spanBegin, spanBytes, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return ORBX_OK;
    };
    if (maxPosA > 0) enqueueRounds(SPEC);  // SPEC rounds and the finishing kernel go out together
    if ((rc = finishAndFetch())) return rc;
    const int* hf = (const int*)hostAt(flags);
    while (rounds < SPEC && hf[rounds]) rounds++;
    if (maxPosA > 0 && rounds == SPEC) {
        for (;;) {
            // rounds per host synchronisation double (4, 8, 16, ...): a long claim chain (dense repeated texture, every
            // query displaced by the previous one) costs O(log) synchronisations and at most twice the rounds it needs
            const int R = std::min(std::max(4, round), maxRounds - round);
            if (R <= 0) FAIL(ORBX_ECUDA, "internal: greedy resolution did not converge in %d rounds", maxRounds);
            enqueueRounds(R);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(m->hFlags, flags, (size_t)round * sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            rounds = 0;
            while (rounds < round && m->hFlags[rounds]) rounds++;
            if (rounds < round) break;
        }
        if ((rc = finishAndFetch())) return rc;
    }
    const int* hres = (const int*)hostAt(dRes);
    const uint8_t* hmatch = hout ? hostAt(matchAll) : nullptr;
    mOff = 0;
    for (int k = 0; k < npairs; k++) {
        nmatches[k] = hres[k];
        if (hout && a[k].n) memcpy(match_a[k], hmatch + mOff * 4, (size_t)a[k].n * 4);
        mOff += (size_t)a[k].n;
    }
    m->lastRounds = rounds + 1;
    m->lastCandidates = 0;
    return ORBX_OK;
}

extern "C" int orbm_search_by_bow_batch(orbm_t* m, int npairs, const orbm_bow_side* a, const orbm_bow_side* b,
                                        const int* n_pos_a, const int* n_pos_b, float nn_ratio, int th, int strict,
                                        int check_orientation, int32_t* const* match_a, int* nmatches) {
    return bow_batch_run(m, npairs, a, b, n_pos_a, n_pos_b, nn_ratio, th, strict, check_orientation, nullptr, match_a, nmatches);
}

extern "C" int orbm_search_for_triangulation_batch(orbm_t* m, int npairs, const orbm_bow_side* a, const orbm_bow_side* b,
                                                   const int* n_pos_a, const int* n_pos_b, const orbm_triangulation* T,
                                                   int th, int check_orientation, int32_t* const* match_a, int* nmatches) {
    if (npairs > 0 && !T) FAIL(ORBX_EINVAL, "null argument");
    return bow_batch_run(m, npairs, a, b, n_pos_a, n_pos_b, 0.f, th, 0, check_orientation, T, match_a, nmatches);
}
