// orbx_fbow.cu — vocabulary-tree transform on the device: the orbv_* part of include/orbx.h.
//
// Replaces fbow::Vocabulary::fromStream / transform (/root/reference/Thirdparty/fbow/src/fbow.cpp:52-153,173-183;
// descent loops fbow.h:371-444) and fbow::fBow::score (fbow.cpp:208-254) as the reference calls them from
// Frame::ComputeFboW / KeyFrame::ComputeFboW (src/Frame.cc:395-400, src/KeyFrame.cc:62) and KeyFrameDatabase.
//
// V1 k_fbow_descend   thread per descriptor: walks the tree, 8 POPC per child, first-lowest child wins ties;
//                     emits (word, weight, node at the store level)
// V2 k_fbow_collect   CTA per frame: bitonic sort of (word | feature) and (node | feature) keys in shared memory,
//                     run heads by block scan, weights summed per word IN FEATURE ORDER (the float sum of the
//                     reference's `result[id] += weight` loop), node lists in push_back order; optional L2 normalisation
// V3 k_fbow_score     thread per database vector: ordered merge of two id-sorted sparse vectors, double accumulation
//
// The wire blob is re-laid out at load time: descriptors of a block's children contiguous (k x 32 B, no padding) so
// that a child is two 16-byte loads, node infos as uint2, child counts as one byte.
#include <algorithm>
#include <cstring>
#include <vector>

#include "orbx_internal.cuh"

#define FBOW_SIG 55824124ull
#define FBOW_NONE 0xFFFFFFFFu
#define FBOW_MAX_FEATURES 16384   // per frame: 2 x 128 KB of sort keys would not fit; one key array of 16384 x 8 B does
#define COLLECT_THREADS 512

struct orbv_handle {
    int device = 0;
    cudaStream_t stream = nullptr;
    uint32_t k = 0, nblocks = 0, descSize = 0;
    int nbits = 0;
    uint64_t totalWords = 0;
    uint4* dDesc = nullptr;     // [nblocks][k][2]
    uint2* dInfo = nullptr;     // [nblocks][k] {id_or_childblock, weight bits}
    uint8_t* dCount = nullptr;  // [nblocks]
    int inSpace = ORBX_MEM_HOST, outSpace = ORBX_MEM_HOST;
    // per-call scratch (grown on demand)
    size_t cap = 0;             // features
    uint8_t* dIn = nullptr;     // staged descriptors
    uint32_t *dWord = nullptr, *dNode = nullptr;
    float* dWeight = nullptr;
    int* dN = nullptr;
    size_t capN = 0;
    // staged outputs (host-output mode)
    uint32_t *oWid = nullptr, *oNid = nullptr, *oFeat = nullptr;
    float* oWw = nullptr;
    int *oNoff = nullptr, *oCounts = nullptr;
    size_t oCap = 0, oFrames = 0;
    int* dErr = nullptr;
    long long launches = 0;
};

// ------------------------------------------------------------------------------------------------ V1 descent
__global__ void __launch_bounds__(128) k_fbow_descend(const uint4* __restrict__ vdesc, const uint2* __restrict__ vinfo,
                                                      const uint8_t* __restrict__ vcount, int k, int nbits,
                                                      const uint4* __restrict__ feats, const int* __restrict__ nPer,
                                                      int stride, int storeLevel, uint32_t* __restrict__ word,
                                                      float* __restrict__ weight, uint32_t* __restrict__ node) {
    const int frame = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPer[frame]) return;
    const size_t at = (size_t)frame * stride + i;
    const uint4 f0 = feats[2 * at], f1 = feats[2 * at + 1];
    uint32_t block = 0, level = 0, curNode = 0, best = 0;
    uint32_t outWord = FBOW_NONE, outNode = FBOW_NONE;
    float outW = 0.f;
    for (;;) {
        const uint4* B = vdesc + (size_t)block * k * 2;
        const int n = vcount[block];
        int bestD = 0x7fffffff;
        for (int c = 0; c < n; c++) {  // strict '<' in ascending child order (fbow.h:421)
            const int d = hamming256(f0, f1, B[2 * c], B[2 * c + 1]);
            if (d < bestD) { bestD = d; best = c; }
        }
        if (storeLevel >= 0 && level == (uint32_t)storeLevel) outNode = curNode;  // fbow.h:424-425
        const uint2 info = vinfo[(size_t)block * k + best];
        if (info.x & 0x80000000u) {  // leaf: word id and weight (fbow.h:429-434)
            outWord = info.x & 0x7FFFFFFFu;
            outW = __uint_as_float(info.y);
            if (storeLevel >= 0 && level < (uint32_t)storeLevel) outNode = curNode;
            break;
        }
        if (info.x == 0) break;  // the loop condition `getId()!=0` (fbow.h:441)
        block = info.x;
        curNode = (curNode << nbits) | best;
        level++;
    }
    word[at] = outWord;
    weight[at] = outW;
    node[at] = outNode;
}

// ------------------------------------------------------------------------------------------------ V2 collect
__device__ __forceinline__ void bitonic_sort_u64(unsigned long long* key, int P) {
    for (int size = 2; size <= P; size <<= 1) {
        for (int strd = size >> 1; strd > 0; strd >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (P >> 1); t += blockDim.x) {
                const int lo = 2 * t - (t & (strd - 1));  // index with bit `strd` clear
                const int hi = lo + strd;
                const bool up = (lo & size) == 0;
                const unsigned long long a = key[lo], b = key[hi];
                if ((a > b) == up) { key[lo] = b; key[hi] = a; }
            }
        }
    }
    __syncthreads();
}

// exclusive scan of one int per thread over the CTA (COLLECT_THREADS threads); returns the prefix, total in *total
__device__ __forceinline__ int block_excl_scan(int v, int* warpSums, int* total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();
    if (lane == 31) warpSums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int s = lane < (COLLECT_THREADS / 32) ? warpSums[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += t;
        }
        if (lane < (COLLECT_THREADS / 32)) warpSums[lane] = s;  // inclusive
    }
    __syncthreads();
    *total = warpSums[COLLECT_THREADS / 32 - 1];
    return inc - v + (wid ? warpSums[wid - 1] : 0);
}

// Sorts (id << 32 | feature) keys of the frame's valid hits and writes the run heads. Returns the number of runs.
// headOut[u] = id of run u; offOut[u] = first sorted position of run u (offOut[runs] = number of valid keys).
__device__ int collect_runs(unsigned long long* key, int P, const uint32_t* __restrict__ id, int n, uint32_t* headOut,
                            int* offOut, int* warpSums) {
    for (int i = threadIdx.x; i < P; i += blockDim.x)
        key[i] = (i < n && id[i] != FBOW_NONE) ? (((unsigned long long)id[i] << 32) | (unsigned)i) : ~0ull;
    bitonic_sort_u64(key, P);
    // each thread owns a contiguous slice of the sorted keys
    const int per = (P + COLLECT_THREADS - 1) / COLLECT_THREADS;
    const int a = threadIdx.x * per, b = min(P, a + per);
    int heads = 0, valid = 0;
    for (int i = a; i < b; i++) {
        const unsigned long long kx = key[i];
        if (kx == ~0ull) break;
        valid++;
        if (i == 0 || (uint32_t)(key[i - 1] >> 32) != (uint32_t)(kx >> 32)) heads++;
    }
    int runs, nValid;
    int u = block_excl_scan(heads, warpSums, &runs);
    block_excl_scan(valid, warpSums, &nValid);
    for (int i = a; i < b; i++) {
        const unsigned long long kx = key[i];
        if (kx == ~0ull) break;
        if (i == 0 || (uint32_t)(key[i - 1] >> 32) != (uint32_t)(kx >> 32)) {
            headOut[u] = (uint32_t)(kx >> 32);
            offOut[u] = i;
            u++;
        }
    }
    if (threadIdx.x == 0) offOut[runs] = nValid;
    __syncthreads();
    return runs;
}

__global__ void __launch_bounds__(COLLECT_THREADS) k_fbow_collect(
    const uint32_t* __restrict__ word, const float* __restrict__ weight, const uint32_t* __restrict__ node,
    const int* __restrict__ nPer, int stride, int storeLevel, int normalize, uint32_t* __restrict__ wid,
    float* __restrict__ ww, int* __restrict__ nWords, uint32_t* __restrict__ nid, int* __restrict__ noff,
    uint32_t* __restrict__ nfeat, int* __restrict__ nNodes, int* __restrict__ runOff /* [frames][stride+1] scratch */) {
    extern __shared__ unsigned long long key[];
    __shared__ int warpSums[COLLECT_THREADS / 32];
    const int frame = blockIdx.x;
    const int n = nPer[frame];
    const size_t base = (size_t)frame * stride;
    int P = 2;
    while (P < n) P <<= 1;
    // ---- words: fBow
    int* wOff = runOff + (size_t)frame * (stride + 1);
    const int runs = collect_runs(key, P, word + base, n, wid + base, wOff, warpSums);
    for (int u = threadIdx.x; u < runs; u += blockDim.x) {
        float s = 0.f;  // map value starts at 0 (fbow.h:19), then += weight once per feature in ascending feature order
        for (int j = wOff[u]; j < wOff[u + 1]; j++) s = s + weight[base + (uint32_t)key[j]];
        ww[base + u] = s;
    }
    if (threadIdx.x == 0) nWords[frame] = runs;
    if (normalize) {  // fbow.cpp:141-152: double sum of float squares in id order, then v *= 1/sqrt(norm) in double
        __syncthreads();
        if (threadIdx.x == 0) {
            double norm = 0;
            for (int u = 0; u < runs; u++) norm += (double)(ww[base + u] * ww[base + u]);
            if (norm > 0.0) {
                const double inv = 1. / sqrt(norm);
                for (int u = 0; u < runs; u++) ww[base + u] = (float)((double)ww[base + u] * inv);
            }
        }
    }
    if (storeLevel < 0) {
        if (threadIdx.x == 0) nNodes[frame] = 0;
        return;
    }
    __syncthreads();
    // ---- nodes: fBow2
    int* nOff = noff + (size_t)frame * (stride + 1);
    const int nruns = collect_runs(key, P, node + base, n, nid + base, nOff, warpSums);
    const int nValid = nOff[nruns];
    for (int j = threadIdx.x; j < nValid; j += blockDim.x) nfeat[base + j] = (uint32_t)key[j];
    if (threadIdx.x == 0) nNodes[frame] = nruns;
}

// ------------------------------------------------------------------------------------------------ V3 score
__global__ void k_fbow_score(const uint32_t* __restrict__ qid, const float* __restrict__ qw, int nq,
                             const uint32_t* __restrict__ did, const float* __restrict__ dw,
                             const int* __restrict__ doff, int m, double* __restrict__ score) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= m) return;
    int a = 0, b = doff[v];
    const int bend = doff[v + 1];
    double s = 0;
    while (a < nq && b < bend) {  // fbow.cpp:218-246
        const uint32_t x = qid[a], y = did[b];
        if (x == y) { s += (double)(qw[a] * dw[b]); a++; b++; }
        else if (x < y) a++;
        else b++;
    }
    score[v] = s >= 1 ? 1.0 : 1.0 - sqrt(1.0 - s);
}

#include "orbx_fbow_host.cuh"
