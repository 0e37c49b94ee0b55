// orbx_extract.cu — host side of the extractor half of the C ABI (include/orbx.h): plan construction (level
// geometry, bilinear tables, TMA tensor maps, device arenas), the launch sequence and the result copies.
// Mirrors ORB_SLAM2::ORBextractor (/root/reference/src/ORBextractor.cc:405-464 ctor tables, :858-919 operator()).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "orbx_kernels.cuh"

// ------------------------------------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";
void orbx_set_err(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char* orbx_last_error_string(void) { return g_err; }
extern "C" const char* orbx_version(void) { return "orbx 0.1 (sm_100a)"; }

typedef void (*fast_kernel_t)(const TmapPack, const FastSt*, u64*, int*, int*, int, int);
// the FAST kernel instantiations: variant 1 (default), and variant 3 with NF1 / NF2 min-max pairs on the FMA pipe
// (8 / 4 balances the ALU pipe against the issue slots; 0 / 0 keeps every min/max on the ALU pipe)
typedef void (*blur_kernel_t)(const DevPlan*, const uint8_t*, uint8_t*, int, int, int);
static blur_kernel_t blur_kernel(int rows) { return rows == BLUR_ROWS_SMALL ? k_blur<BLUR_ROWS_SMALL> : k_blur<BLUR_ROWS>; }
static fast_kernel_t fast_kernel(int ver, int nf) {
    if (ver != 3) {  // variant 1: nf = which (NI1, NI2) max-by-IMAD instantiation (ORBX_FAST1_NI, A/B measurement only)
        switch (nf) {    // measured per 384 frames: (0,0) 1.178 ms, (8,0) 1.124, (8,4) 1.123, (8,8) 1.119
            case 0: return k_fast_t<1, 0, 0>;
            case 2: return k_fast_t<1, 8, 0>;
            case 5: return k_fast_t<1, 8, 4>;
            default: return k_fast_t<1, 8, 8>;
        }
    }
    return nf == 1 ? k_fast_t<3, 0, 0> : k_fast_t<3, 8, 4>;
}

// ------------------------------------------------------------------------------------------------ handle
struct orbx_handle {
    orbx_params prm;
    int device = 0;
    int fastVer = 1;      // FAST kernel: 1 = k_fast_t<1> (byte tile, every pixel scored in one pass; 16 of the 72 min/max per
                          // pixel pair as IMAD pairs on the FMA pipe, ORBX_FAST1_NI=0 keeps all on the ALU pipe);
                          // 3 = k_fast_t<3> (expanded fp16-lane tile, min/max split over the ALU and FMA pipes: identical
                          // candidates, measured the SAME speed — issue-slot bound instead of ALU bound,
                          // profiles/r02x_fast3_analysis.md);
                          // 2 = k_fast2 (two-threshold pre-test + work lists: identical candidates, measured 1.47x SLOWER,
                          // profiles/r02d_fast2_analysis.md). ORBX_FAST=1|2|3 selects; ORBX_FAST_V2=1 is the old spelling of 2.
    int fastNF = 0;       // variant 3: which (NF1, NF2) instantiation (ORBX_FAST3_NF); variant 1: which (NI1, NI2) (ORBX_FAST1_NI)
    int blurRows = BLUR_ROWS;  // k_blur strip height of the current plan (BLUR_ROWS, or BLUR_ROWS_SMALL for a few frames)
    int resizeBX = RESIZE_BX, resizeBY = RESIZE_BY;  // k_resize_fast CTA shape (ORBX_RESIZE_SHAPE=bx,by: A/B measurement only)
    int descKpw = 0;      // k_orient_desc keypoints per warp override (ORBX_DESC_KPW, tuning only)
    cudaStream_t stream = nullptr;
    int inSpace = ORBX_MEM_HOST, outSpace = ORBX_MEM_HOST;
    // reference ctor tables
    std::vector<float> sf, isf, sig2, isig2;
    std::vector<int> perLevel, umax;
    // plan
    int planW = 0, planH = 0, planBatch = 0;
    DevPlan hp;             // host copy
    DevPlan* dPlan = nullptr;
    FastSt* dFastSt = nullptr;  // one record per FAST supertile
    TmapPack maps;
    uint8_t* dPyr = nullptr;
    uint8_t* dBlur = nullptr;
    size_t pyrBytes = 0;
    int2* dTabs = nullptr;
    CUtensorMap* dDescMaps = nullptr;  // k_orient_desc: [level] pyramid boxes, [ORBX_MAX_LEVELS + level] blurred boxes
    u64* dCand = nullptr;
    int* dNodeOf = nullptr;
    int* dCandCount = nullptr;
    uint2* dSel = nullptr;
    int* dSelCount = nullptr;
    orbx_kp* dKps = nullptr;
    uint8_t* dDesc = nullptr;
    int* dNOut = nullptr;
    int* dErr = nullptr;
    uint8_t* dExport = nullptr;
    size_t exportBytes = 0;
    int* hCounts = nullptr;  // pinned: [batch] nOut + [1] err
    uint8_t* dOutBlock = nullptr;  // [dKps | dDesc | dNOut | dErr]
    size_t outBlockBytes = 0;
    uint8_t* hStage = nullptr;     // pinned mirror of the block (plans with batch == 1)
    // single-frame path as a CUDA graph: [clear candidate counts, the kernel chain with its side-stream blur, result block to
    // the pinned mirror] captured once per plan and replayed per call (the copy-in stays a stream operation: its source is
    // the caller's pointer). ORBX_NO_GRAPH=1 disables it.
    cudaGraphExec_t gExec = nullptr; int64_t gLaunches = 0; bool gDisabled = false;
    int qtC = 0, qtC2 = 0, qtThreads = 256;
    size_t qtSmem = 0, fastSmem = 0;
    int lastBatch = 0;
    cudaStream_t sSide = nullptr;  // small batches: blur beside FAST + quadtree
    cudaStream_t sLvl[ORBX_MAX_LEVELS] = {nullptr};  // small batches: FAST + quadtree of each level on its own stream
    cudaEvent_t evLvlFork[ORBX_MAX_LEVELS] = {nullptr}, evLvlJoin[ORBX_MAX_LEVELS] = {nullptr};
    bool levelStreams = true;  // ORBX_LEVEL_STREAMS=0: the single-frame path as one FAST and one quadtree launch
    cudaEvent_t evFork = nullptr, evJoin = nullptr;
    bool pendingOneCopy = false;
    orbx_kp* pendingKps = nullptr;
    uint8_t* pendingDesc = nullptr;
    int pendingBatch = 0, pendingCap = 0;  // orbx_extract_batch_submit without its wait (negative: empty-image batch)
    bool blurValid = false;
    int64_t launches = 0;
    bool profiling = false;
    cudaEvent_t ev[8] = {};
    // host-IO pipeline (orbx_extract_batch): copy-in, two compute and copy-out streams, one event pair per chunk
    int chunk = 0;
    cudaStream_t sIn = nullptr, sOut = nullptr, sComp[2] = {nullptr, nullptr};
    cudaEvent_t evStart = nullptr, evDone = nullptr;
    std::vector<cudaEvent_t> evIn, evComp;
    float stageMs[7] = {};
    // upstream image prep fused into the ingest (orbx_prep.cuh)
    int inFmt = ORBX_FMT_GRAY8;
    float *dMapX = nullptr, *dMapY = nullptr;  // rectification maps (output size mapW x mapH)
    int mapW = 0, mapH = 0;
    uint8_t *dRaw = nullptr, *dGray = nullptr;  // staged source images / grey source of the remap
    int* dStSad = nullptr; float *dStU = nullptr, *dStZ = nullptr; size_t stCap = 0;  // stereo scratch (orbx_stereo_match*)
    size_t rawCap = 0, grayCap = 0;
    orbx_kp* dKpsUn = nullptr;
    size_t kpsUnCap = 0;
};

static inline int round_half_even_f(float v) { return (int)lrintf(v); }

// ORBextractor ctor tables, src/ORBextractor.cc:409-463 (same expressions, same types).
static void init_tables(orbx_handle* h) {
    const orbx_params& p = h->prm;
    const int n = p.nlevels;
    const double scaleFactor = p.scaleFactor;  // the class member is a double initialised from the float argument
    h->sf.assign(n, 1.0f);
    h->sig2.assign(n, 1.0f);
    for (int i = 1; i < n; i++) {
        h->sf[i] = (float)(h->sf[i - 1] * scaleFactor);
        h->sig2[i] = h->sf[i] * h->sf[i];
    }
    h->isf.resize(n);
    h->isig2.resize(n);
    for (int i = 0; i < n; i++) {
        h->isf[i] = 1.0f / h->sf[i];
        h->isig2[i] = 1.0f / h->sig2[i];
    }
    h->perLevel.resize(n);
    const float factor = (float)(1.0f / scaleFactor);
    float nDesired = p.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)n));
    int sum = 0;
    for (int l = 0; l < n - 1; l++) {
        h->perLevel[l] = round_half_even_f(nDesired);
        sum += h->perLevel[l];
        nDesired *= factor;
    }
    h->perLevel[n - 1] = std::max(p.nfeatures - sum, 0);
    const int hpz = p.halfPatchSize;
    h->umax.assign(hpz + 2, 0);
    const int vmax = (int)floor(hpz * sqrtf(2.f) / 2 + 1);
    const int vmin = (int)ceil(hpz * sqrtf(2.f) / 2);
    const double hp2 = (double)hpz * hpz;
    for (int v = 0; v <= vmax; ++v) h->umax[v] = (int)lrint(sqrt(hp2 - v * v));
    for (int v = hpz, v0 = 0; v >= vmin; --v) {
        while (h->umax[v0] == h->umax[v0 + 1]) ++v0;
        h->umax[v] = v0;
        ++v0;
    }
    h->umax.resize(hpz + 1);
}

// OpenCV's INTER_LINEAR coefficient tables for one axis (resize.cpp: fx = (float)((d+0.5)*scale-0.5), 11-bit fixed point).
static void linear_table(int src, int dst, bool clampFrac, std::vector<int2>& out) {
    const double inv_scale = (double)dst / src;
    const double scale = 1.0 / inv_scale;
    for (int d = 0; d < dst; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= s;
        if (clampFrac) {
            if (s < 0) { f = 0; s = 0; }
            if (s >= src - 1) { f = 0; s = src - 1; }
        }
        const int a0 = (short)lrintf((1.f - f) * 2048.f), a1 = (short)lrintf(f * 2048.f);
        // x table: the source index; y table: the two source rows already clamped to the image, packed row0 | row1 << 16
        // (two instructions to unpack in the kernels instead of four min/max per output row)
        const int idx = clampFrac ? s : (std::min(std::max(s, 0), src - 1) | (std::min(std::max(s + 1, 0), src - 1) << 16));
        out.push_back(make_int2(idx, (a0 & 0xffff) | (a1 << 16)));
    }
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    });
    return fn;
}

static void free_plan(orbx_handle* h) {
    if (h->gExec) { cudaGraphExecDestroy(h->gExec); h->gExec = nullptr; }
    cudaFree(h->dPlan); cudaFree(h->dFastSt); cudaFree(h->dPyr); cudaFree(h->dBlur); cudaFree(h->dTabs); cudaFree(h->dDescMaps); cudaFree(h->dCand);
    cudaFree(h->dNodeOf); cudaFree(h->dCandCount); cudaFree(h->dSel); cudaFree(h->dSelCount); cudaFree(h->dOutBlock);
    cudaFree(h->dExport);
    if (h->hCounts) cudaFreeHost(h->hCounts);
    if (h->hStage) cudaFreeHost(h->hStage);
    h->hStage = nullptr; h->dOutBlock = nullptr; h->dErr = nullptr;
    h->dPlan = nullptr; h->dFastSt = nullptr; h->dPyr = h->dBlur = nullptr; h->dTabs = nullptr; h->dDescMaps = nullptr; h->dCand = nullptr; h->dNodeOf = nullptr;
    h->dCandCount = nullptr; h->dSel = nullptr; h->dSelCount = nullptr; h->dKps = nullptr; h->dDesc = nullptr;
    h->dNOut = nullptr; h->dExport = nullptr; h->exportBytes = 0; h->hCounts = nullptr;
    h->planW = h->planH = h->planBatch = 0;
}

static int build_plan(orbx_handle* h, int w, int hgt, int batch) {
    if (h->planW == w && h->planH == hgt && h->planBatch >= batch) return ORBX_OK;
    // grow the batch capacity geometrically so a caller alternating batch sizes does not rebuild every call
    if (h->planW == w && h->planH == hgt) batch = std::max(batch, h->planBatch * 2);
    free_plan(h);
    const orbx_params& p = h->prm;
    const int e = p.edgeThreshold;
    DevPlan& D = h->hp;
    memset(&D, 0, sizeof(D));
    D.nlevels = p.nlevels;
    D.batch = batch;
    D.iniTh = p.iniThFAST;
    D.minTh = p.minThFAST;
    D.halfPatch = p.halfPatchSize;
    h->blurRows = batch <= 4 ? BLUR_ROWS_SMALL : BLUR_ROWS;
    D.blurRows = h->blurRows;
    D.edge = e;
    for (int v = 0; v <= p.halfPatchSize; v++) D.umax[v] = h->umax[v];

    std::vector<int2> tabs;
    long long pyrOff = 0, candOff = 0;
    int cellBase = 0, selOff = 0, blurBase = 0, maxSel = 0;
    size_t fastSmem = 0;
    for (int l = 0; l < p.nlevels; l++) {
        LevelGeo& G = D.lv[l];
        G.w = round_half_even_f((float)w * h->isf[l]);  // src/ORBextractor.cc:925-926
        G.h = round_half_even_f((float)hgt * h->isf[l]);
        const int minB = e - 3, maxBX = G.w - e + 3, maxBY = G.h - e + 3;
        const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
        G.nCols = (int)(width / 30.f);
        G.nRows = (int)(height / 30.f);
        if (G.nCols < 1 || G.nRows < 1)
            FAIL(ORBX_ESHAPE, "level %d is %dx%d: smaller than one 30-px FAST cell inside the %d-px border", l, G.w, G.h, e);
        G.wCell = (int)ceilf(width / G.nCols);
        G.hCell = (int)ceilf(height / G.nRows);
        if (G.wCell > 63 || G.hCell > 63 || G.nCols > 1023 || G.nRows > 1023)
            FAIL(ORBX_ESHAPE, "level %d cell grid %dx%d of %dx%d px exceeds the candidate key layout", l, G.nCols, G.nRows,
                 G.wCell, G.hCell);
        G.pitch = (G.w + 63) & ~63;
        G.frameStride = (long long)G.pitch * G.h;
        G.pyrOff = pyrOff;
        pyrOff += ((G.frameStride * batch + 255) / 256) * 256;
        G.stCX = std::max(1, std::min(4, (256 - 28) / G.wCell));
        // two cell rows per supertile for throughput; a plan for a few frames takes one (twice the CTAs, each half as long:
        // single 752x480 frame 0.107 -> 0.094 ms per call, 640x480 0.097 -> 0.094)
        G.stCY = batch <= 4 ? 1 : 2;
        G.nStX = (G.nCols + G.stCX - 1) / G.stCX;
        G.nStY = (G.nRows + G.stCY - 1) / G.stCY;
        G.stBase = cellBase;
        cellBase += G.nStX * G.nStY;
        G.boxW = (G.stCX * G.wCell + 28 + 15) & ~15;  // 3-px circle rim + word/16-byte alignment slack (see k_fast)
        G.boxH = G.stCY * G.hCell + 6;
        G.maxSurv = ((G.wCell + 1) / 2) * ((G.hCell + 1) / 2);
        G.quota = h->perLevel[l];
        G.treeW = maxBX - minB;
        G.treeH = maxBY - minB;
        G.nIni = (int)roundf((float)G.treeW / (float)G.treeH);  // src/ORBextractor.cc:537
        if (G.nIni < 1) FAIL(ORBX_ESHAPE, "level %d: width/height rounds to 0 initial quadtree nodes (portrait image)", l);
        G.hX = (float)G.treeW / (float)G.nIni;
        G.candCap = G.nCols * G.nRows * G.maxSurv;
        G.candOff = candOff;
        candOff += (long long)G.candCap * batch;
        G.selCap = std::max(G.quota + 3, 4 * G.nIni) + 1;
        G.selOff = selOff;
        selOff += G.selCap;
        maxSel = std::max(maxSel, G.selCap);
        G.scale = h->sf[l];
        G.sizeField = (float)(int)(p.patchSize * h->sf[l]);  // src/ORBextractor.cc:831
        G.blurTilesX = (G.w + 63) / 64;  // strips of 64 px x BLUR_ROWS rows; a k_blur unit (warp) takes two consecutive ones
        G.blurTilesY = (G.h + h->blurRows - 1) / h->blurRows;
        G.blurTileBase = blurBase;
        blurBase += (G.blurTilesX * G.blurTilesY + 1) / 2;
        if (l > 0) {
            if (tabs.size() & 1) tabs.push_back(make_int2(0, 0));  // 16-byte alignment of the x table
            G.xtabOff = (int)tabs.size();
            linear_table(D.lv[l - 1].w, G.w, true, tabs);
            while ((tabs.size() - G.xtabOff) & 3) tabs.push_back(tabs.back());
            G.resizeFast = !(D.lv[l - 1].w == 2 * G.w && D.lv[l - 1].h == 2 * G.h);
            for (size_t i = G.xtabOff; i + 3 < tabs.size(); i += 4)
                if (tabs[i + 3].x - tabs[i].x > 6 || tabs[i + 3].x < tabs[i].x) G.resizeFast = 0;
            // x entries become {PRMT selector | source index << 16}: the selector picks source bytes (rel, rel + 1) of the
            // 8-byte window that starts at the first source pixel of the entry's group of four (k_resize_fast uses it as is)
            for (size_t g = G.xtabOff; g < tabs.size(); g += 4) {
                const int s0 = tabs[g].x;
                for (size_t i = g; i < g + 4; i++) {
                    const int rel = std::min(std::max(tabs[i].x - s0, 0), 6);  // <= 6 where k_resize_fast applies
                    tabs[i].x = (rel * 0x11 + 0x10) | (tabs[i].x << 16);
                }
            }
            G.ytabOff = (int)tabs.size();  // even (the x table is 16-byte aligned and padded to 4 entries)
            linear_table(D.lv[l - 1].h, G.h, false, tabs);
            while ((tabs.size() - G.ytabOff) & 3) tabs.push_back(tabs.back());  // k_resize_fast reads 4 rows' entries as two uint4
        }
        G.fastUnits = ((G.stCX * G.wCell + 3) / 4 + 2) * (G.stCY * G.hCell);
        if (G.fastUnits >= 16384 || (G.stCX * G.wCell + 3) / 4 + 2 > 64 || G.stCY * G.hCell > 128)
            FAIL(ORBX_ESHAPE, "level %d supertile of %d words exceeds k_fast2's tables", l, G.fastUnits);
        if (h->fastVer != 2)
            fastSmem = std::max(fastSmem, (size_t)2 * G.boxW * G.boxH + fast_region3_bytes(h->fastVer, G.boxW, G.boxH) + sizeof(FastSmemHdr));
        else
            fastSmem = std::max(fastSmem, (size_t)2 * G.boxW * G.boxH + ((2 * G.fastUnits + 15) & ~15) + (size_t)8 * G.fastUnits +
                                              sizeof(Fast2Hdr));
    }
    D.totalSt = cellBase;
    D.totalBlurTiles = blurBase;
    D.selStride = selOff;
    h->fastSmem = fastSmem;
    h->qtC = (maxSel + 7) & ~7;
    if (h->qtC > 4096) FAIL(ORBX_ESHAPE, "per-level keypoint quota %d exceeds the quadtree kernel's node capacity", maxSel);
    h->qtC2 = 1;
    while (h->qtC2 < h->qtC) h->qtC2 <<= 1;
    h->qtSmem = (size_t)8 * h->qtC2 + (size_t)(16 + 8 + 7 * 4 + 32) * h->qtC + 40 * 4 + 64;
    h->qtThreads = h->qtC > 1024 ? 1024 : (h->qtC > 512 ? 512 : 256);
    h->pyrBytes = (size_t)pyrOff + 256;

    CK(cudaMalloc(&h->dPlan, sizeof(DevPlan)));
    std::vector<FastSt> sts((size_t)D.totalSt);
    for (int l = 0; l < p.nlevels; l++) {
        const LevelGeo& G = D.lv[l];
        for (int st = 0; st < G.nStX * G.nStY; st++) {
            FastSt& S = sts[(size_t)G.stBase + st];
            memset(&S, 0, sizeof(S));
            const int sy = st / G.nStX, sx = st - sy * G.nStX;
            const int cj0 = sx * G.stCX, ci0 = sy * G.stCY;
            const int ncx = std::min(G.stCX, G.nCols - cj0), ncy = std::min(G.stCY, G.nRows - ci0);
            S.level = (unsigned short)l; S.stCX = (unsigned short)G.stCX; S.cj0 = (unsigned short)cj0; S.ci0 = (unsigned short)ci0;
            S.xFirst = (unsigned short)(e + cj0 * G.wCell); S.xEnd = (unsigned short)std::min(e + (cj0 + ncx) * G.wCell, G.w - e);
            S.yFirst = (unsigned short)(e + ci0 * G.hCell); S.yEnd = (unsigned short)std::min(e + (ci0 + ncy) * G.hCell, G.h - e);
            S.boxW = (short)G.boxW; S.boxH = (short)G.boxH; S.wCell = (short)G.wCell; S.hCell = (short)G.hCell;
            S.nlevels = (short)p.nlevels; S.minTh = (short)p.minThFAST; S.iniTh = (short)p.iniThFAST;
            S.nCells = (short)(G.stCX * G.stCY);
            S.divCw = (65536u + G.wCell - 1) / G.wCell; S.divCh = (65536u + G.hCell - 1) / G.hCell;
            S.candCap = G.candCap; S.candOff = G.candOff;
        }
    }
    CK(cudaMalloc(&h->dFastSt, std::max<size_t>(sts.size(), 1) * sizeof(FastSt)));
    CK(cudaMemcpyAsync(h->dFastSt, sts.data(), sts.size() * sizeof(FastSt), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMalloc(&h->dPyr, h->pyrBytes));
    CK(cudaMalloc(&h->dBlur, h->pyrBytes));
    CK(cudaMalloc(&h->dTabs, std::max<size_t>(tabs.size(), 1) * sizeof(int2)));
    CK(cudaMalloc(&h->dCand, (size_t)candOff * sizeof(u64)));
    CK(cudaMalloc(&h->dNodeOf, (size_t)candOff * sizeof(int)));
    CK(cudaMalloc(&h->dCandCount, (size_t)batch * p.nlevels * sizeof(int)));
    CK(cudaMalloc(&h->dSelCount, (size_t)batch * p.nlevels * sizeof(int)));
    CK(cudaMalloc(&h->dSel, (size_t)batch * D.selStride * sizeof(uint2)));
    // results live in ONE block [keypoints | descriptors | nOut[batch] | error flag] so that counts + flag are one
    // device->host copy and a single frame's whole result is one copy (single-frame latency)
    const size_t kpBytes = (size_t)batch * D.selStride * sizeof(orbx_kp), dsBytes = (size_t)batch * D.selStride * 32;
    h->outBlockBytes = kpBytes + dsBytes + (size_t)(batch + 1) * sizeof(int);
    CK(cudaMalloc(&h->dOutBlock, h->outBlockBytes));
    h->dKps = (orbx_kp*)h->dOutBlock;
    h->dDesc = h->dOutBlock + kpBytes;
    h->dNOut = (int*)(h->dOutBlock + kpBytes + dsBytes);
    h->dErr = h->dNOut + batch;
    CK(cudaMemsetAsync(h->dErr, 0, sizeof(int), h->stream));
    CK(cudaMallocHost(&h->hCounts, (size_t)(batch + 1) * sizeof(int)));
    if (batch == 1) CK(cudaMallocHost(&h->hStage, h->outBlockBytes));
    CK(cudaMemcpyAsync(h->dPlan, &D, sizeof(DevPlan), cudaMemcpyHostToDevice, h->stream));
    if (!tabs.empty())
        CK(cudaMemcpyAsync(h->dTabs, tabs.data(), tabs.size() * sizeof(int2), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemsetAsync(h->dPyr, 0, h->pyrBytes, h->stream));
    CK(cudaMemsetAsync(h->dBlur, 0, h->pyrBytes, h->stream));
    CK(cudaStreamSynchronize(h->stream));  // tabs is a local

    // TMA descriptors: one 3-D (x, y, frame) uint8 tensor per level, box = one FAST cell ROI
    PFN_encodeTiled enc = get_encode();
    if (!enc) FAIL(ORBX_ECUDA, "cuTensorMapEncodeTiled entry point not available from the driver");
    for (int l = 0; l < p.nlevels; l++) {
        const LevelGeo& G = D.lv[l];
        cuuint64_t gdim[3] = {(cuuint64_t)G.w, (cuuint64_t)G.h, (cuuint64_t)batch};
        cuuint64_t gstr[2] = {(cuuint64_t)G.pitch, (cuuint64_t)G.frameStride};
        cuuint32_t box[3] = {(cuuint32_t)G.boxW, (cuuint32_t)G.boxH, 1};
        cuuint32_t estr[3] = {1, 1, 1};
        CUresult r = enc(&h->maps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, h->dPyr + G.pyrOff, gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) FAIL(ORBX_ECUDA, "cuTensorMapEncodeTiled failed for level %d (CUresult %d)", l, (int)r);
    }
    {   // k_orient_desc fetches a keypoint's two patches as TMA boxes: same tensors, boxes 48 x 31 and 64 x 37
        std::vector<CUtensorMap> dm(2 * ORBX_MAX_LEVELS);
        memset(dm.data(), 0, dm.size() * sizeof(CUtensorMap));
        for (int l = 0; l < p.nlevels; l++) {
            const LevelGeo& G = D.lv[l];
            cuuint64_t gdim[3] = {(cuuint64_t)G.w, (cuuint64_t)G.h, (cuuint64_t)batch};
            cuuint64_t gstr[2] = {(cuuint64_t)G.pitch, (cuuint64_t)G.frameStride};
            cuuint32_t estr[3] = {1, 1, 1};
            for (int which = 0; which < 2; which++) {
                cuuint32_t box[3] = {(cuuint32_t)(which ? DESC_BLR_BW : DESC_RAW_BW), (cuuint32_t)(which ? 37 : 31), 1};
                CUresult r = enc(&dm[which * ORBX_MAX_LEVELS + l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3,
                                 (which ? h->dBlur : h->dPyr) + G.pyrOff, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r != CUDA_SUCCESS) FAIL(ORBX_ECUDA, "cuTensorMapEncodeTiled failed for the patch boxes of level %d (CUresult %d)", l, (int)r);
            }
        }
        CK(cudaMalloc(&h->dDescMaps, dm.size() * sizeof(CUtensorMap)));
        CK(cudaMemcpy(h->dDescMaps, dm.data(), dm.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
        CK(cudaFuncSetAttribute(k_orient_desc, cudaFuncAttributeMaxDynamicSharedMemorySize, DESC_SMEM_BYTES));
    }
    if (h->fastSmem > 227 * 1024) FAIL(ORBX_ESHAPE, "FAST supertile needs %zu bytes of shared memory", h->fastSmem);
    if (h->fastVer == 2) CK(cudaFuncSetAttribute(k_fast2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fastSmem));
    else CK(cudaFuncSetAttribute(fast_kernel(h->fastVer, h->fastNF), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fastSmem));
    CK(cudaFuncSetAttribute(k_quadtree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->qtSmem));
    h->planW = w;
    h->planH = hgt;
    h->planBatch = batch;
    h->lastBatch = 0;
    return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------ ABI
extern "C" int orbx_create(const orbx_params* p, int device, orbx_t** out) {
    if (!p || !out) FAIL(ORBX_EINVAL, "null argument");
    if (p->nlevels < 1 || p->nlevels > ORBX_MAX_LEVELS) FAIL(ORBX_EINVAL, "nlevels must be in [1,%d]", ORBX_MAX_LEVELS);
    if (p->nfeatures < 1 || !(p->scaleFactor > 1.0f)) FAIL(ORBX_EINVAL, "nfeatures >= 1 and scaleFactor > 1 required");
    if (p->minThFAST < 1 || p->iniThFAST < p->minThFAST || p->iniThFAST > 254)
        FAIL(ORBX_EINVAL, "need 1 <= minThFAST <= iniThFAST <= 254");
    if (p->halfPatchSize < 1 || p->halfPatchSize > 15 || p->patchSize < 1) FAIL(ORBX_EINVAL, "halfPatchSize in [1,15]");
    if (p->edgeThreshold < 19 || p->edgeThreshold < p->halfPatchSize + 4)
        FAIL(ORBX_EINVAL, "edgeThreshold must cover the descriptor reach (>= 19) and the orientation patch");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        FAIL(ORBX_ECUDA, "no CUDA device: liborbx has no CPU path");
    }
    if (device < 0 || device >= ndev) FAIL(ORBX_EINVAL, "device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    orbx_handle* h = new orbx_handle();
    h->prm = *p;
    h->device = device;
    { const char* v = getenv("ORBX_FAST_V2"); if (v && v[0] == '1') h->fastVer = 2; }
    { const char* v = getenv("ORBX_FAST"); if (v && v[0] >= '1' && v[0] <= '3') h->fastVer = v[0] - '0'; }
    { const char* v = getenv("ORBX_LEVEL_STREAMS"); if (v && v[0] == '0') h->levelStreams = false; }
    { const char* v = getenv("ORBX_DESC_KPW"); if (v && atoi(v) >= 1 && atoi(v) <= 64) h->descKpw = atoi(v); }
    { const char* v = getenv("ORBX_FAST3_NF"); if (v && v[0] >= '0' && v[0] <= '9') h->fastNF = v[0] - '0'; }
    if (h->fastVer == 1) { const char* v = getenv("ORBX_FAST1_NI"); h->fastNF = (v && v[0] >= '0' && v[0] <= '9') ? v[0] - '0' : 1; }
    { const char* v = getenv("ORBX_RESIZE_SHAPE"); int bx = 0, by = 0;
      if (v && sscanf(v, "%d,%d", &bx, &by) == 2 && bx >= 1 && by >= 1 && bx * by <= 256 && (bx * by) % 32 == 0) { h->resizeBX = bx; h->resizeBY = by; } }
    { const char* v = getenv("ORBX_NO_GRAPH"); h->gDisabled = v && v[0] == '1'; }
    init_tables(h);
    cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete h;
        FAIL(ORBX_ECUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
    }
    for (int i = 0; i < 8; i++) cudaEventCreate(&h->ev[i]);
    cudaMemcpyToSymbol(d_pattern, kOrbPattern, 1024);
    *out = h;
    return ORBX_OK;
}

extern "C" int orbx_destroy(orbx_t* h) {
    if (!h) return ORBX_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    free_plan(h);
    for (int i = 0; i < 8; i++) cudaEventDestroy(h->ev[i]);
    for (cudaEvent_t e : h->evIn) cudaEventDestroy(e);
    for (cudaEvent_t e : h->evComp) cudaEventDestroy(e);
    if (h->sIn) {
        cudaEventDestroy(h->evStart);
        cudaEventDestroy(h->evDone);
        cudaStreamDestroy(h->sIn);
        cudaStreamDestroy(h->sOut);
        cudaStreamDestroy(h->sComp[0]);
        cudaStreamDestroy(h->sComp[1]);
    }
    cudaFree(h->dMapX); cudaFree(h->dMapY); cudaFree(h->dRaw); cudaFree(h->dGray); cudaFree(h->dKpsUn);
    cudaFree(h->dStSad); cudaFree(h->dStU); cudaFree(h->dStZ);
    if (h->sSide) {
        cudaEventDestroy(h->evFork);
        cudaEventDestroy(h->evJoin);
        cudaStreamDestroy(h->sSide);
    }
    for (int l = 0; l < ORBX_MAX_LEVELS; l++)
        if (h->sLvl[l]) {
            cudaEventDestroy(h->evLvlFork[l]);
            cudaEventDestroy(h->evLvlJoin[l]);
            cudaStreamDestroy(h->sLvl[l]);
        }
    cudaStreamDestroy(h->stream);
    delete h;
    return ORBX_OK;
}

extern "C" int orbx_set_memory_space(orbx_t* h, int in, int out) {
    if (!h || in < 0 || in > 1 || out < 0 || out > 1) FAIL(ORBX_EINVAL, "bad memory space");
    h->inSpace = in;
    h->outSpace = out;
    return ORBX_OK;
}

extern "C" int orbx_level_count(orbx_t* h) { return h ? h->prm.nlevels : 0; }

extern "C" int orbx_tables(orbx_t* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* fpl,
                           int* umax) {
    if (!h) FAIL(ORBX_EINVAL, "null handle");
    const int n = h->prm.nlevels;
    for (int i = 0; i < n; i++) {
        if (scale) scale[i] = h->sf[i];
        if (inv_scale) inv_scale[i] = h->isf[i];
        if (sigma2) sigma2[i] = h->sig2[i];
        if (inv_sigma2) inv_sigma2[i] = h->isig2[i];
        if (fpl) fpl[i] = h->perLevel[i];
    }
    if (umax)
        for (int v = 0; v <= h->prm.halfPatchSize; v++) umax[v] = h->umax[v];
    return ORBX_OK;
}

extern "C" int orbx_max_keypoints(orbx_t* h, int w, int hgt, int* cap) {
    if (!h || !cap) FAIL(ORBX_EINVAL, "null argument");
    int total = 0;
    for (int l = 0; l < h->prm.nlevels; l++) {
        const int lw = round_half_even_f((float)w * h->isf[l]), lh = round_half_even_f((float)hgt * h->isf[l]);
        const int e = h->prm.edgeThreshold;
        const int tw = lw - 2 * e + 6, th = lh - 2 * e + 6;
        if (tw < 30 || th < 30) FAIL(ORBX_ESHAPE, "level %d too small", l);
        const int nIni = (int)roundf((float)tw / (float)th);
        total += std::max(h->perLevel[l] + 3, 4 * nIni) + 1;
    }
    *cap = total;
    return ORBX_OK;
}

static inline void stamp(orbx_handle* h, int i) {
    if (h->profiling) cudaEventRecord(h->ev[i], h->stream);
}

// ORBX_DEBUG_SYNC=1: synchronise after every launch and name the kernel that faulted (development aid).
static bool debug_sync() {
    static const bool on = getenv("ORBX_DEBUG_SYNC") != nullptr;
    return on;
}
#define KCHECK(h, name)                                                                              \
    do {                                                                                             \
        (h)->launches++;                                                                             \
        if (debug_sync()) {                                                                          \
            cudaError_t e_ = cudaStreamSynchronize((h)->stream);                                     \
            if (e_ == cudaSuccess) e_ = cudaGetLastError();                                          \
            if (e_ != cudaSuccess) FAIL(ORBX_ECUDA, "kernel %s failed: %s", name, cudaGetErrorString(e_)); \
        }                                                                                            \
    } while (0)

// side streams of the single-frame path (created outside any stream capture)
static int ensure_side_streams(orbx_handle* h, int nl) {
    if (!h->sSide) {
        CK(cudaStreamCreateWithFlags(&h->sSide, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&h->evFork, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&h->evJoin, cudaEventDisableTiming));
    }
    for (int l = 0; l < nl && l < ORBX_MAX_LEVELS; l++)
        if (!h->sLvl[l]) {
            CK(cudaStreamCreateWithFlags(&h->sLvl[l], cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&h->evLvlFork[l], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->evLvlJoin[l], cudaEventDisableTiming));
        }
    return ORBX_OK;
}

// One chunk of frames [f0, f0 + nb) through the whole kernel chain on stream `cs`. Every arena is indexed by the
// absolute frame number, so chunks of one batch never touch each other's memory.
static int launch_chain(orbx_handle* h, cudaStream_t cs, int f0, int nb, bool stamps, bool captured = false) {
    const DevPlan& D = h->hp;
    const int nl = D.nlevels;
    const int blurCtas = (D.totalBlurTiles + BLUR_WARPS - 1) / BLUR_WARPS;
    const int qtT = (nb <= 4 || (long long)nb * nl * 512 <= 148LL * 2048) ? std::max(512, h->qtThreads) : h->qtThreads;
    auto resize = [&](int l) {
        const LevelGeo& G = D.lv[l];
        dim3 grid((G.w + 255) / 256, (G.h + 3) / 4, nb), block(64, 4);
        const int bx = h->resizeBX, by = h->resizeBY;
        dim3 gridFast((G.w + 4 * bx - 1) / (4 * bx), (G.h + by * RESIZE_ROWS - 1) / (by * RESIZE_ROWS), nb);
        if (G.resizeFast) k_resize_fast<<<gridFast, dim3(bx, by), 0, cs>>>(h->hp, l, h->dPyr, h->dTabs, f0);
        else k_resize<<<grid, block, 0, cs>>>(h->hp, l, h->dPyr, h->dTabs, f0);
    };
    // A few frames do not fill the GPU, and the chain pyramid -> FAST -> quadtree is a sequence of latency-bound launches
    // (single 640x480 frame: 7 x 4.8 + 14 + 37 us). But level l's FAST and quadtree only need level l: on this path every
    // level gets its own stream, forked as soon as its resize is enqueued, so the critical path is the longest level chain
    // instead of the sum; the blur of level 0 runs beside them from the start, that of the other levels after the last resize.
    // (Large batches: every kernel fills the GPU, one launch per stage.)
    // Only inside the captured single-frame graph: as plain stream launches the 26 launches and 20 event operations cost the
    // host more than the overlap gives (measured 0.206 ms per call against 0.140); there the blur alone forks.
    const bool fork = nb <= 4 && !stamps && cs == h->stream;
    const bool levelPar = fork && captured && h->fastVer != 2 && h->levelStreams;
    if (levelPar) {
        { const int rcS = ensure_side_streams(h, nl); if (rcS != ORBX_OK) return rcS; }
        for (int l = 0; l < nl; l++) {
            const LevelGeo& G = D.lv[l];
            if (l > 0) {
                resize(l);
                KCHECK(h, "k_resize");
            }
            CK(cudaEventRecord(h->evLvlFork[l], cs));
            CK(cudaStreamWaitEvent(h->sLvl[l], h->evLvlFork[l], 0));
            if (l == 0) {
                CK(cudaStreamWaitEvent(h->sSide, h->evLvlFork[0], 0));
                const int t1 = nl > 1 ? D.lv[1].blurTileBase : D.totalBlurTiles;
                blur_kernel(h->blurRows)<<<dim3((t1 + BLUR_WARPS - 1) / BLUR_WARPS, nb), 32 * BLUR_WARPS, 0, h->sSide>>>(h->dPlan, h->dPyr, h->dBlur, f0, 0, t1);
                KCHECK(h, "k_blur");
                CK(cudaEventRecord(h->evJoin, h->sSide));
            }
            fast_kernel(h->fastVer, h->fastNF)<<<dim3(G.nStX * G.nStY, nb), FAST_THREADS, h->fastSmem, h->sLvl[l]>>>(
                h->maps, h->dFastSt, h->dCand, h->dCandCount, h->dErr, f0, G.stBase);
            KCHECK(h, "k_fast");
            k_quadtree<<<dim3(nb, 1), qtT, h->qtSmem, h->sLvl[l]>>>(h->hp, h->dCand, h->dCandCount, h->dNodeOf, h->dSel, h->dSelCount,
                                                                  h->qtC, h->qtC2, h->dErr, f0, l);
            KCHECK(h, "k_quadtree");
            CK(cudaEventRecord(h->evLvlJoin[l], h->sLvl[l]));
        }
        if (nl > 1) {
            const int t1 = D.lv[1].blurTileBase;
            blur_kernel(h->blurRows)<<<dim3((D.totalBlurTiles - t1 + BLUR_WARPS - 1) / BLUR_WARPS, nb), 32 * BLUR_WARPS, 0, cs>>>(h->dPlan, h->dPyr, h->dBlur, f0, t1, D.totalBlurTiles);
            KCHECK(h, "k_blur");
        }
        CK(cudaStreamWaitEvent(cs, h->evJoin, 0));
        for (int l = 0; l < nl; l++) CK(cudaStreamWaitEvent(cs, h->evLvlJoin[l], 0));
    } else {
        for (int l = 1; l < nl; l++) {
            resize(l);
            KCHECK(h, "k_resize");
        }
        if (stamps) stamp(h, 2);
        if (fork) {  // the blur (it reads only the pyramid) beside FAST and the quadtree
            const int rcS = ensure_side_streams(h, 0);
            if (rcS != ORBX_OK) return rcS;
            CK(cudaEventRecord(h->evFork, cs));
            CK(cudaStreamWaitEvent(h->sSide, h->evFork, 0));
            blur_kernel(h->blurRows)<<<dim3(blurCtas, nb), 32 * BLUR_WARPS, 0, h->sSide>>>(h->dPlan, h->dPyr, h->dBlur, f0, 0, D.totalBlurTiles);
            KCHECK(h, "k_blur");
            CK(cudaEventRecord(h->evJoin, h->sSide));
        }
        if (h->fastVer != 2)
            fast_kernel(h->fastVer, h->fastNF)<<<dim3(D.totalSt, nb), FAST_THREADS, h->fastSmem, cs>>>(h->maps, h->dFastSt, h->dCand, h->dCandCount, h->dErr, f0, 0);
        else
            k_fast2<<<dim3(D.totalSt, nb), FAST_THREADS, h->fastSmem, cs>>>(h->maps, h->dPlan, h->dCand, h->dCandCount, h->dErr, f0);
        KCHECK(h, "k_fast");
        if (stamps) stamp(h, 3);
        // a few frames: one CTA per level is all the parallelism there is, so give it the widest CTA
        // (measured, single 640x480 / 1241x376 frame: 256 thr 160 / 278 us per call, 512 thr 155 / 220, 1024 thr 160 / 215)
        // ... and the same holds while every CTA of the launch is resident at once (the kernel then lasts as long as its
        // level-0 CTA): 512 threads as long as nb x levels such CTAs fit on the 148 SMs (config 2, 32-64 frames of 2000
        // keypoints: 117 us at 256 threads, one wave, 30 % of the warp slots busy)
        k_quadtree<<<dim3(nb, nl), qtT, h->qtSmem, cs>>>(h->hp, h->dCand, h->dCandCount, h->dNodeOf, h->dSel, h->dSelCount, h->qtC,
                                                         h->qtC2, h->dErr, f0, 0);
        KCHECK(h, "k_quadtree");
        if (stamps) stamp(h, 4);
        if (fork) {
            CK(cudaStreamWaitEvent(cs, h->evJoin, 0));
        } else {
            blur_kernel(h->blurRows)<<<dim3(blurCtas, nb), 32 * BLUR_WARPS, 0, cs>>>(h->dPlan, h->dPyr, h->dBlur, f0, 0, D.totalBlurTiles);
            KCHECK(h, "k_blur");
        }
    }
    if (stamps) stamp(h, 5);
    {
        // keypoints per warp: a warp's prologue (level prefix, pattern, first record and first boxes: ~36 % of its time at 4
        // keypoints) amortises over more keypoints when there are enough frames to keep every SM busy anyway
        // (by the number of keypoint slots of the launch: 16 needs a few waves of warps to pay, 4K frames in small batches
        // measured slower with it: config 5, 16 frames x 8000 keypoints, 0.153 ms at 4 against 0.201 at 16)
        const long long slots = (long long)nb * D.selStride;
        int kpw = nb <= 4 ? DESC_KPW : slots >= 400000 ? DESC_KPW_BATCH : slots >= 150000 ? 8 : 4;
        if (h->descKpw > 0) kpw = h->descKpw;
        k_orient_desc<<<dim3((D.selStride + 8 * kpw - 1) / (8 * kpw), nb), 256, DESC_SMEM_BYTES, cs>>>(
            h->dPlan, h->dDescMaps, h->dSel, h->dSelCount, h->dKps, h->dDesc, h->dNOut, h->dErr, f0, kpw);
    }
    KCHECK(h, "k_orient_desc");
    if (stamps) stamp(h, 6);
    return ORBX_OK;
}

static inline int fmt_channels(int fmt) { return fmt == ORBX_FMT_GRAY8 ? 1 : (fmt == ORBX_FMT_RGB8 || fmt == ORBX_FMT_BGR8 ? 3 : 4); }

// staging for colour / rectified input: the raw source images (host input only) and the grey source of the remap
static int ensure_ingest(orbx_handle* h, int batch, int w, int hgt) {
    const int ch = fmt_channels(h->inFmt);
    if (ch == 1 && !h->dMapX) {
        // grey host input whose width is not the level-0 pitch (a multiple of 64): a pitched host-to-device copy moves such rows one DMA
        // descriptor at a time (measured 4.0 ms for 64 x 1241x376 against 0.6 ms flat), so the frames cross PCIe as one
        // flat copy into this staging block and k_repitch lays them out in level 0
        if (h->inSpace == ORBX_MEM_HOST && (w & 63) && (size_t)batch * hgt * w > h->rawCap) {
            cudaFree(h->dRaw);
            h->dRaw = nullptr;
            h->rawCap = 0;
            CK(cudaMalloc(&h->dRaw, (size_t)batch * hgt * w));
            h->rawCap = (size_t)batch * hgt * w;
        }
        return ORBX_OK;
    }
    const size_t raw = (size_t)batch * hgt * w * ch, gray = (size_t)batch * hgt * w;
    if (h->inSpace == ORBX_MEM_HOST && raw > h->rawCap) {
        cudaFree(h->dRaw);
        h->dRaw = nullptr;
        CK(cudaMalloc(&h->dRaw, raw));
        h->rawCap = raw;
    }
    if (ch > 1 && h->dMapX && gray > h->grayCap) {
        cudaFree(h->dGray);
        h->dGray = nullptr;
        CK(cudaMalloc(&h->dGray, gray));
        h->grayCap = gray;
    }
    return ORBX_OK;
}

static int copy_in(orbx_handle* h, cudaStream_t cs, const uint8_t* imgs, int f0, int nb, int w, int hgt, size_t stride,
                   size_t img_stride) {
    // ingest: level 0 of the pyramid IS the input image (src/ORBextractor.cc:939-943); one strided copy
    const cudaMemcpyKind kin = h->inSpace == ORBX_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    const LevelGeo& L0 = h->hp.lv[0];
    uint8_t* dst = h->dPyr + L0.pyrOff + (long long)f0 * L0.frameStride;
    const uint8_t* src = imgs + (size_t)f0 * img_stride;
    const int ch = fmt_channels(h->inFmt);
    if (ch > 1 || h->dMapX) {
        // colour and/or unrectified input: the source is staged (host input) or read in place (device input), then
        // converted / remapped straight into level 0
        const size_t rowBytes = (size_t)w * ch;
        size_t sStride = stride, sImg = img_stride;
        if (h->inSpace == ORBX_MEM_HOST) {
            uint8_t* raw = h->dRaw + (size_t)f0 * hgt * rowBytes;
            if (nb == 1 || img_stride == stride * (size_t)hgt) {
                CK(cudaMemcpy2DAsync(raw, rowBytes, src, stride, rowBytes, (size_t)hgt * nb, kin, cs));
            } else {
                for (int f = 0; f < nb; f++)
                    CK(cudaMemcpy2DAsync(raw + (size_t)f * hgt * rowBytes, rowBytes, src + f * img_stride, stride, rowBytes, hgt, kin, cs));
            }
            src = raw;
            sStride = rowBytes;
            sImg = rowBytes * hgt;
        }
        if (ch > 1) {
            uint8_t* g = h->dMapX ? h->dGray + (size_t)f0 * hgt * w : dst;
            const int gPitch = h->dMapX ? w : L0.pitch;
            const long long gFrame = h->dMapX ? (long long)hgt * w : L0.frameStride;
            const size_t al = ch == 3 ? 3 : 15;
            const int wide = (((size_t)src | sStride | sImg) & al) == 0;
            k_cvt_gray<<<dim3((w + 255) / 256, (hgt + 3) / 4, nb), dim3(64, 4), 0, cs>>>(
                src, sStride, sImg, ch, h->inFmt == ORBX_FMT_BGR8 || h->inFmt == ORBX_FMT_BGRA8, wide, w, hgt, g, gPitch, gFrame);
            KCHECK(h, "k_cvt_gray");
            src = g;
            sStride = (size_t)w;
            sImg = (size_t)hgt * w;
        }
        if (h->dMapX) {
            k_remap<<<dim3((L0.w + 127) / 128, (L0.h + 1) / 2, (nb + REMAP_FRAMES - 1) / REMAP_FRAMES), dim3(128, 2), 0, cs>>>(
                src, sStride, sImg, w, hgt, h->dMapX, h->dMapY, L0.w, L0.h, nb, dst, L0.pitch, L0.frameStride);
            KCHECK(h, "k_remap");
        }
        return ORBX_OK;
    }
    if (h->inSpace == ORBX_MEM_HOST && (w & 63) && stride == (size_t)w && h->dRaw &&
        (size_t)(f0 + nb) * hgt * w <= h->rawCap) {
        uint8_t* raw = h->dRaw + (size_t)f0 * hgt * w;
        if (nb == 1 || img_stride == stride * (size_t)hgt) {
            CK(cudaMemcpyAsync(raw, src, (size_t)nb * hgt * w, kin, cs));
        } else {
            for (int f = 0; f < nb; f++) CK(cudaMemcpyAsync(raw + (size_t)f * hgt * w, src + f * img_stride, (size_t)hgt * w, kin, cs));
        }
        k_repitch<<<dim3((w + 255) / 256, (hgt + 3) / 4, nb), dim3(64, 4), 0, cs>>>(raw, w, hgt, dst, L0.pitch, L0.frameStride);
        KCHECK(h, "k_repitch");
    } else if (nb == 1 || img_stride == stride * (size_t)hgt) {
        CK(cudaMemcpy2DAsync(dst, L0.pitch, src, stride, w, (size_t)hgt * nb, kin, cs));
    } else {
        for (int f = 0; f < nb; f++)
            CK(cudaMemcpy2DAsync(dst + f * L0.frameStride, L0.pitch, src + f * img_stride, stride, w, hgt, kin, cs));
    }
    return ORBX_OK;
}

static int copy_out(orbx_handle* h, cudaStream_t cs, orbx_kp* kps, uint8_t* desc, int cap, int f0, int nb) {
    const DevPlan& D = h->hp;
    const cudaMemcpyKind kout = h->outSpace == ORBX_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    const int rows = std::min(cap, D.selStride);
    if (rows > 0) {
        CK(cudaMemcpy2DAsync(kps + (size_t)f0 * cap, (size_t)cap * sizeof(orbx_kp), h->dKps + (size_t)f0 * D.selStride,
                             (size_t)D.selStride * sizeof(orbx_kp), (size_t)rows * sizeof(orbx_kp), nb, kout, cs));
        CK(cudaMemcpy2DAsync(desc + (size_t)f0 * cap * 32, (size_t)cap * 32, h->dDesc + (size_t)f0 * D.selStride * 32,
                             (size_t)D.selStride * 32, (size_t)rows * 32, nb, kout, cs));
    }
    return ORBX_OK;
}

static int ensure_pipeline(orbx_handle* h, int nChunks) {
    if (!h->sIn) {
        CK(cudaStreamCreateWithFlags(&h->sIn, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&h->sOut, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&h->sComp[0], cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&h->sComp[1], cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&h->evStart, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&h->evDone, cudaEventDisableTiming));
    }
    while ((int)h->evIn.size() < nChunks) {
        cudaEvent_t a, b;
        CK(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        h->evIn.push_back(a);
        h->evComp.push_back(b);
    }
    return ORBX_OK;
}

// Enqueue one batch (copies, kernels, count read-back) without waiting for it; orbx_extract_batch_wait() completes it.
// A caller with two handles keeps one batch on the PCIe link while the other computes.
extern "C" int orbx_extract_batch_submit(orbx_t* h, const uint8_t* imgs, int batch, int w, int hgt, size_t stride,
                                         size_t img_stride, orbx_kp* kps, uint8_t* desc, int cap) {
    if (!h) FAIL(ORBX_EINVAL, "null argument");
    if (h->pendingBatch != 0) FAIL(ORBX_EINVAL, "a submitted batch is still pending: call orbx_extract_batch_wait first");
    if (batch < 1) FAIL(ORBX_EINVAL, "batch must be >= 1");
    if (!imgs || w <= 0 || hgt <= 0) {  // empty image: the reference returns silently with outputs untouched (:861-862)
        h->pendingBatch = -batch;
        return ORBX_OK;
    }
    if (stride < (size_t)w * fmt_channels(h->inFmt) || (batch > 1 && img_stride < stride * (size_t)hgt))
        FAIL(ORBX_EINVAL, "bad strides");
    if (cap < 0 || (cap > 0 && (!kps || !desc))) FAIL(ORBX_EINVAL, "null output");
    CK(cudaSetDevice(h->device));
    // with rectification maps the pyramid has the maps' size; (w, hgt) describe the source images
    int rc = build_plan(h, h->dMapX ? h->mapW : w, h->dMapX ? h->mapH : hgt, batch);
    if (rc != ORBX_OK) return rc;
    rc = ensure_ingest(h, batch, w, hgt);
    if (rc != ORBX_OK) return rc;
    const DevPlan& D = h->hp;
    cudaStream_t st = h->stream;
    const int nl = D.nlevels;

    // Host buffers on either side: split the batch into chunks and software-pipeline them — chunk c+1 crosses PCIe
    // on the copy-in stream while chunk c computes (two compute streams, so one chunk's tail overlaps the next one's
    // head) and chunk c-1 leaves on the copy-out stream. Device buffers: one launch sequence for the whole batch.
    const bool hostIO = h->inSpace == ORBX_MEM_HOST || h->outSpace == ORBX_MEM_HOST;
    int chunk = batch;
    if (!h->profiling) {
        if (hostIO || h->chunk > 0) {
            // default chunk: about 14.7 MB of input (48 frames of 640x480, the measured optimum there), so that larger
            // frames still pipeline (KITTI-shape frames 31 at a time), but never fewer than 8 frames: the quadtree and the
            // launch sequence are latency bound, so 2-frame chunks of 4K images were measured slower than one 16-frame chunk
            const int byBytes = (int)std::max<size_t>(8, std::min<size_t>(48, (size_t)48 * 640 * 480 / ((size_t)w * hgt)));
            chunk = h->chunk > 0 ? h->chunk : byBytes;
            if (batch < 2 * chunk) chunk = h->chunk > 0 || batch < 4 ? batch : std::max(1, batch / 2);
        } else if (batch >= 96 || (batch >= 2 && (long long)batch * w * hgt >= 96LL * 640 * 480)) {  // (or as many pixels in fewer frames)
            chunk = (batch + 1) / 2;  // device buffers: two halves on the two compute streams, so one half's kernel
                                      // tails and its latency-bound quadtree overlap the other's kernels (measured +2 %)
        }
    }
    // chunk boundaries: the first two chunks are short (1/3 and 2/3 of a chunk) so the pipeline fills quickly — the
    // first copy-in is the only one nothing overlaps
    std::vector<int> cuts;
    if (chunk >= batch) {
        cuts = {0, batch};
    } else {
        cuts.push_back(0);
        int at = 0;
        if (chunk >= 12 && batch >= 3 * chunk) {
            at += chunk / 3; cuts.push_back(at);
            at += (2 * chunk) / 3; cuts.push_back(at);
        }
        while (at < batch) {  // (growing the later chunks was measured: no gain)
            at = std::min(batch, at + chunk);
            cuts.push_back(at);
        }
    }
    const int nChunks = (int)cuts.size() - 1;

    stamp(h, 0);
    CK(cudaMemsetAsync(h->dCandCount, 0, (size_t)batch * nl * sizeof(int), st));
    // single frame to host memory: the whole result block (keypoints, descriptors, count, error flag) leaves in one
    // copy into a pinned mirror; orbx_extract_batch_wait hands the rows to the caller
    const bool oneCopy = batch == 1 && h->planBatch == 1 && h->hStage && h->outSpace == ORBX_MEM_HOST && !h->profiling;
    h->pendingOneCopy = oneCopy;
    h->pendingKps = kps;
    h->pendingDesc = desc;
    if (nChunks == 1) {
        rc = copy_in(h, st, imgs, 0, batch, w, hgt, stride, img_stride);
        if (rc != ORBX_OK) return rc;
        stamp(h, 1);
        bool replayed = false;
        if (oneCopy && !h->gDisabled && !debug_sync()) {
            if (!h->gExec) {  // capture once per plan; any failure falls back to plain stream launches for good
                { const int rcS = ensure_side_streams(h, nl); if (rcS != ORBX_OK) return rcS; }
                const int64_t l0 = h->launches;
                cudaGraph_t g = nullptr;
                bool ok = cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
                if (ok) {
                    ok = cudaMemsetAsync(h->dCandCount, 0, (size_t)nl * sizeof(int), st) == cudaSuccess;
                    ok = ok && launch_chain(h, st, 0, 1, false, true) == ORBX_OK;
                    ok = ok && cudaMemcpyAsync(h->hStage, h->dOutBlock, h->outBlockBytes, cudaMemcpyDeviceToHost, st) == cudaSuccess;
                    ok = (cudaStreamEndCapture(st, &g) == cudaSuccess) && ok && g;
                }
                if (ok) ok = cudaGraphInstantiate(&h->gExec, g, 0) == cudaSuccess;
                if (g) cudaGraphDestroy(g);
                h->gLaunches = h->launches - l0;
                h->launches = l0;
                if (!ok) {
                    cudaGetLastError();
                    h->gExec = nullptr;
                    h->gDisabled = true;
                }
            }
            if (h->gExec) {
                CK(cudaGraphLaunch(h->gExec, st));
                h->launches += h->gLaunches;
                replayed = true;
            }
        }
        if (replayed) {
            // (the memset enqueued above ran before the copy-in; the graph clears the counts again: harmless)
        } else {
        rc = launch_chain(h, st, 0, batch, h->profiling);  // (stamps off => a small batch forks the blur onto the side stream)
        if (rc != ORBX_OK) return rc;
        CK(cudaGetLastError());
        }
        if (replayed) {
        } else if (oneCopy) {
            CK(cudaMemcpyAsync(h->hStage, h->dOutBlock, h->outBlockBytes, cudaMemcpyDeviceToHost, st));
        } else {
            rc = copy_out(h, st, kps, desc, cap, 0, batch);
            if (rc != ORBX_OK) return rc;
        }
    } else {
        rc = ensure_pipeline(h, nChunks);
        if (rc != ORBX_OK) return rc;
        CK(cudaEventRecord(h->evStart, st));
        CK(cudaStreamWaitEvent(h->sIn, h->evStart, 0));
        for (int c = 0; c < nChunks; c++) {
            const int f0 = cuts[c], nb = cuts[c + 1] - cuts[c];
            cudaStream_t cs = h->sComp[c & 1];
            rc = copy_in(h, h->sIn, imgs, f0, nb, w, hgt, stride, img_stride);
            if (rc != ORBX_OK) return rc;
            CK(cudaEventRecord(h->evIn[c], h->sIn));
            CK(cudaStreamWaitEvent(cs, h->evIn[c], 0));
            rc = launch_chain(h, cs, f0, nb, false);
            if (rc != ORBX_OK) return rc;
            CK(cudaGetLastError());
            CK(cudaEventRecord(h->evComp[c], cs));
            CK(cudaStreamWaitEvent(h->sOut, h->evComp[c], 0));
            rc = copy_out(h, h->sOut, kps, desc, cap, f0, nb);
            if (rc != ORBX_OK) return rc;
        }
        CK(cudaEventRecord(h->evDone, h->sOut));  // sOut is ordered after every compute chunk, hence after every copy-in
        CK(cudaStreamWaitEvent(st, h->evDone, 0));
    }
    h->lastBatch = batch;
    h->blurValid = true;
    if (oneCopy) {
        // counts and flag are the tail of the block already on its way
    } else if (batch == h->planBatch) {
        CK(cudaMemcpyAsync(h->hCounts, h->dNOut, (size_t)(batch + 1) * sizeof(int), cudaMemcpyDeviceToHost, st));  // nOut + flag
    } else {
        CK(cudaMemcpyAsync(h->hCounts, h->dNOut, (size_t)batch * sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h->hCounts + batch, h->dErr, sizeof(int), cudaMemcpyDeviceToHost, st));
    }
    stamp(h, 7);
    h->pendingBatch = batch;
    h->pendingCap = cap;
    return ORBX_OK;
}

extern "C" int orbx_extract_batch_wait(orbx_t* h, int* n) {
    if (!h || !n) FAIL(ORBX_EINVAL, "null argument");
    if (h->pendingBatch == 0) FAIL(ORBX_EINVAL, "no submitted batch to wait for");
    const int batch = std::abs(h->pendingBatch), cap = h->pendingCap;
    const bool empty = h->pendingBatch < 0;
    h->pendingBatch = 0;
    if (empty) {
        for (int f = 0; f < batch; f++) n[f] = 0;
        return ORBX_OK;
    }
    CK(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    CK(cudaStreamSynchronize(st));
    if (h->profiling)
        for (int i = 0; i < 7; i++) cudaEventElapsedTime(&h->stageMs[i], h->ev[i], h->ev[i + 1]);
    if (h->pendingOneCopy) {
        const DevPlan& D = h->hp;
        const size_t kpBytes = (size_t)D.selStride * sizeof(orbx_kp), dsBytes = (size_t)D.selStride * 32;
        memcpy(h->hCounts, h->hStage + kpBytes + dsBytes, 2 * sizeof(int));
        const int rows = std::min(std::min(cap, D.selStride), std::max(h->hCounts[0], 0));
        if (rows > 0 && h->hCounts[1] == 0) {
            memcpy(h->pendingKps, h->hStage, (size_t)rows * sizeof(orbx_kp));
            memcpy(h->pendingDesc, h->hStage + kpBytes, (size_t)rows * 32);
        }
    }
    if (h->hCounts[batch] != 0) {
        const int code = h->hCounts[batch];
        cudaMemsetAsync(h->dErr, 0, sizeof(int), st);
        FAIL(ORBX_ECUDA, "device-side error flag %d (2 = FAST TMA wait timed out, 3 = candidate overflow, 4/5 = quadtree bound, 6 = descriptor TMA wait timed out)", code);
    }
    int worst = 0;
    for (int f = 0; f < batch; f++) {
        n[f] = h->hCounts[f];
        worst = std::max(worst, h->hCounts[f]);
    }
    if (worst > cap) FAIL(ORBX_ECAP, "cap_per_frame %d < %d keypoints found", cap, worst);
    return ORBX_OK;
}

extern "C" int orbx_extract_batch(orbx_t* h, const uint8_t* imgs, int batch, int w, int hgt, size_t stride,
                                  size_t img_stride, orbx_kp* kps, uint8_t* desc, int cap, int* n) {
    if (!h || !n) FAIL(ORBX_EINVAL, "null argument");
    int rc = orbx_extract_batch_submit(h, imgs, batch, w, hgt, stride, img_stride, kps, desc, cap);
    if (rc != ORBX_OK) {
        h->pendingBatch = 0;
        return rc;
    }
    return orbx_extract_batch_wait(h, n);
}

/* Single-process multi-GPU form of orbx_extract_batch: `ndev` handles, one per device (created with
 * orbx_create(.., device, ..)), each driven by its own host thread; the batch is split into contiguous blocks — handle k
 * owns frames [k*base + min(k, extra), ...) with base = batch / ndev, extra = batch % ndev, the same rule as
 * orbslam2_b200/sharding.py:shard_range — and every block's results land in its own slice of the caller's arrays. No
 * collective: frames are independent. Host buffers only (each device copies its own slice). Returns the first error. */
extern "C" int orbx_extract_batch_multi(orbx_t** hs, int ndev, const uint8_t* imgs, int batch, int w, int hgt, size_t stride,
                                        size_t img_stride, orbx_kp* kps, uint8_t* desc, int cap, int* n) {
    if (!hs || ndev < 1 || !n || batch < 0) FAIL(ORBX_EINVAL, "bad argument");
    for (int k = 0; k < ndev; k++)
        if (!hs[k] || hs[k]->inSpace != ORBX_MEM_HOST || hs[k]->outSpace != ORBX_MEM_HOST)
            FAIL(ORBX_EINVAL, "orbx_extract_batch_multi takes host buffers (handle %d)", k);
    std::vector<int> rc(ndev, ORBX_OK);
    std::vector<std::string> msg(ndev);
    std::vector<std::thread> th;
    const int base = batch / ndev, extra = batch % ndev;
    for (int k = 0; k < ndev; k++) {
        const int lo = k * base + std::min(k, extra), cnt = base + (k < extra ? 1 : 0);
        if (cnt == 0) continue;
        th.emplace_back([=, &rc, &msg]() {
            rc[k] = orbx_extract_batch(hs[k], imgs + (size_t)lo * img_stride, cnt, w, hgt, stride, img_stride,
                                       kps + (size_t)lo * cap, desc + (size_t)lo * cap * 32, cap, n + lo);
            if (rc[k] != ORBX_OK) msg[k] = orbx_last_error_string();  // the error text is thread-local
        });
    }
    for (auto& t : th) t.join();
    for (int k = 0; k < ndev; k++)
        if (rc[k] != ORBX_OK) FAIL(rc[k], "device block %d: %s", k, msg[k].c_str());
    return ORBX_OK;
}

/* Pixel format of the images handed to orbx_extract*: colour input is converted as cv::cvtColor(.., COLOR_x2GRAY) does
 * (src/Tracking.cc:269-294) while it is written into pyramid level 0. Strides are in bytes. */
extern "C" int orbx_set_input_format(orbx_t* h, int fmt) {
    if (!h || fmt < ORBX_FMT_GRAY8 || fmt > ORBX_FMT_BGRA8) FAIL(ORBX_EINVAL, "bad input format");
    h->inFmt = fmt;
    return ORBX_OK;
}

/* Rectification maps as cv::initUndistortRectifyMap(.., CV_32F, M1, M2) returns them (host pointers, out_w x out_h
 * floats each): every following extract call remaps its source images like cv::remap(src, dst, M1, M2, INTER_LINEAR)
 * (Test/Replay/Stereo/stereo_euroc.cc:98-99,136-137) into pyramid level 0. NULL maps switch it off. */
extern "C" int orbx_set_rectify_maps(orbx_t* h, const float* map_x, const float* map_y, int out_w, int out_h) {
    if (!h) FAIL(ORBX_EINVAL, "null handle");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    cudaFree(h->dMapX); cudaFree(h->dMapY);
    h->dMapX = h->dMapY = nullptr;
    h->mapW = h->mapH = 0;
    if (!map_x && !map_y) return ORBX_OK;
    if (!map_x || !map_y || out_w <= 0 || out_h <= 0) FAIL(ORBX_EINVAL, "bad maps");
    const size_t bytes = (size_t)out_w * out_h * sizeof(float);
    CK(cudaMalloc(&h->dMapX, bytes));
    CK(cudaMalloc(&h->dMapY, bytes));
    CK(cudaMemcpy(h->dMapX, map_x, bytes, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(h->dMapY, map_y, bytes, cudaMemcpyHostToDevice));
    h->mapW = out_w;
    h->mapH = out_h;
    return ORBX_OK;
}

/* Frame::UndistortKeyPoints (src/Frame.cc:402-432) for every frame of the last batch: kps_un[f*cap + i] is keypoint i
 * of frame f with pt replaced by cv::undistortPoints(pt, K, D, Mat(), K); K4 = fx, fy, cx, cy; dist = k1 k2 p1 p2 [k3
 * [k4 k5 k6]] (host pointers). dist[0] == 0 (or ndist == 0) copies the keypoints unchanged, as the reference does. */
extern "C" int orbx_undistort_keypoints(orbx_t* h, const float* K4, const float* dist, int ndist, orbx_kp* kps_un, int cap) {
    if (!h || !K4 || !kps_un || ndist < 0 || ndist > 8 || (ndist > 0 && !dist)) FAIL(ORBX_EINVAL, "bad argument");
    if (!h->lastBatch) FAIL(ORBX_ESTATE, "no extracted batch to undistort");
    CK(cudaSetDevice(h->device));
    const DevPlan& D = h->hp;
    const int batch = h->lastBatch;
    const size_t need = (size_t)batch * D.selStride;
    if (need > h->kpsUnCap) {
        cudaFree(h->dKpsUn);
        h->dKpsUn = nullptr;
        CK(cudaMalloc(&h->dKpsUn, need * sizeof(orbx_kp)));
        h->kpsUnCap = need;
    }
    UndistortParams P;
    memset(&P, 0, sizeof(P));
    P.fx = K4[0]; P.fy = K4[1]; P.cx = K4[2]; P.cy = K4[3];
    for (int i = 0; i < ndist; i++) P.k[i] = dist[i];
    P.identity = ndist == 0 || dist[0] == 0.0f;
    cudaStream_t st = h->stream;
    k_undistort<<<dim3((D.selStride + 127) / 128, batch), 128, 0, st>>>(h->dKps, h->dNOut, D.selStride, P, h->dKpsUn);
    KCHECK(h, "k_undistort");
    CK(cudaGetLastError());
    const cudaMemcpyKind kout = h->outSpace == ORBX_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    const int rows = std::min(cap, D.selStride);
    if (rows > 0)
        CK(cudaMemcpy2DAsync(kps_un, (size_t)cap * sizeof(orbx_kp), h->dKpsUn, (size_t)D.selStride * sizeof(orbx_kp),
                             (size_t)rows * sizeof(orbx_kp), batch, kout, st));
    CK(cudaStreamSynchronize(st));
    return ORBX_OK;
}

/* Frames per pipeline chunk when a buffer lives in host memory (0 = default). */
extern "C" int orbx_set_chunk(orbx_t* h, int frames) {
    if (!h || frames < 0) FAIL(ORBX_EINVAL, "bad chunk");
    h->chunk = frames;
    return ORBX_OK;
}

extern "C" int orbx_extract(orbx_t* h, const uint8_t* img, int w, int hgt, size_t stride, orbx_kp* kps, uint8_t* desc,
                            int cap, int* n) {
    return orbx_extract_batch(h, img, 1, w, hgt, stride, stride * (size_t)(hgt > 0 ? hgt : 0), kps, desc, cap, n);
}

extern "C" int orbx_level_size(orbx_t* h, int level, int* w, int* hgt) {
    if (!h || level < 0 || level >= h->prm.nlevels) FAIL(ORBX_EINVAL, "bad level");
    if (!h->planW) FAIL(ORBX_ESTATE, "no image has been extracted yet");
    if (w) *w = h->hp.lv[level].w;
    if (hgt) *hgt = h->hp.lv[level].h;
    return ORBX_OK;
}

static int export_level(orbx_handle* h, const uint8_t* base, int frame, int level, uint8_t* dst, size_t dst_stride, int e) {
    if (!h || !dst || level < 0 || level >= h->prm.nlevels) FAIL(ORBX_EINVAL, "bad argument");
    if (!h->lastBatch || frame < 0 || frame >= h->lastBatch) FAIL(ORBX_ESTATE, "frame %d not in the last extracted batch", frame);
    CK(cudaSetDevice(h->device));
    const LevelGeo& G = h->hp.lv[level];
    const int W = G.w + 2 * e, H = G.h + 2 * e;
    if (dst_stride < (size_t)W) FAIL(ORBX_EINVAL, "dst_stride too small");
    const size_t need = (size_t)W * H;
    if (need > h->exportBytes) {
        cudaFree(h->dExport);
        h->dExport = nullptr;
        h->exportBytes = 0;
        CK(cudaMalloc(&h->dExport, need));
        h->exportBytes = need;
    }
    dim3 block(32, 8), grid((W + 31) / 32, (H + 7) / 8);
    k_export<<<grid, block, 0, h->stream>>>(h->dPlan, base, level, frame, e, h->dExport);
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpy2DAsync(dst, dst_stride, h->dExport, W, W, H, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

extern "C" int orbx_pyramid_level(orbx_t* h, int frame, int level, uint8_t* dst, size_t dst_stride, int with_border) {
    if (!h) FAIL(ORBX_EINVAL, "null handle");
    return export_level(h, h->dPyr, frame, level, dst, dst_stride, with_border ? h->prm.edgeThreshold : 0);
}

/* The whole pyramid of one frame in one launch and one copy: level l is written to dst + offsets[l] (offsets are an
 * OUTPUT: cumulative dense sizes) as a dense (w + 2e) x (h + 2e) buffer, e = edgeThreshold if with_border else 0.
 * dst_bytes must hold the sum (orbx_pyramid_bytes gives it). Replaces nlevels orbx_pyramid_level calls (each its own
 * kernel, copy and synchronisation) in the drop-in's mvImagePyramid refresh. dst is a host pointer. */
extern "C" int orbx_pyramid_bytes(orbx_t* h, int with_border, size_t* bytes) {
    if (!h || !bytes) FAIL(ORBX_EINVAL, "null argument");
    if (!h->lastBatch) FAIL(ORBX_ESTATE, "no frame extracted yet");
    const int e = with_border ? h->prm.edgeThreshold : 0;
    size_t tot = 0;
    for (int l = 0; l < h->prm.nlevels; l++) tot += (size_t)(h->hp.lv[l].w + 2 * e) * (h->hp.lv[l].h + 2 * e);
    *bytes = tot;
    return ORBX_OK;
}
extern "C" int orbx_pyramid_all(orbx_t* h, int frame, uint8_t* dst, size_t dst_bytes, int with_border, size_t* offsets) {
    if (!h || !dst || !offsets) FAIL(ORBX_EINVAL, "null argument");
    if (!h->lastBatch || frame < 0 || frame >= h->lastBatch) FAIL(ORBX_ESTATE, "frame %d not in the last extracted batch", frame);
    CK(cudaSetDevice(h->device));
    const int e = with_border ? h->prm.edgeThreshold : 0, nl = h->prm.nlevels;
    ExportOffsets offs;
    size_t tot = 0;
    for (int l = 0; l < nl; l++) {
        offs.off[l] = (long long)tot;
        offsets[l] = tot;
        tot += (size_t)(h->hp.lv[l].w + 2 * e) * (h->hp.lv[l].h + 2 * e);
    }
    if (dst_bytes < tot) FAIL(ORBX_ECAP, "dst holds %zu bytes, the pyramid needs %zu", dst_bytes, tot);
    if (tot > h->exportBytes) {
        cudaFree(h->dExport);
        h->dExport = nullptr;
        h->exportBytes = 0;
        CK(cudaMalloc(&h->dExport, tot));
        h->exportBytes = tot;
    }
    const int W0 = h->hp.lv[0].w + 2 * e, H0 = h->hp.lv[0].h + 2 * e;
    dim3 block(64, 4), grid((W0 + 63) / 64, (H0 + 3) / 4, nl);
    k_export_all<<<grid, block, 0, h->stream>>>(h->dPlan, h->dPyr, frame, e, offs, h->dExport);
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dst, h->dExport, tot, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

extern "C" int orbx_blurred_level(orbx_t* h, int frame, int level, uint8_t* dst, size_t dst_stride) {
    if (!h) FAIL(ORBX_EINVAL, "null handle");
    if (!h->blurValid) FAIL(ORBX_ESTATE, "no blurred pyramid resident");
    return export_level(h, h->dBlur, frame, level, dst, dst_stride, 0);
}

extern "C" int orbx_candidates(orbx_t* h, int frame, int level, int32_t* xys, int cap, int* n) {
    if (!h || !n || level < 0 || level >= h->prm.nlevels) FAIL(ORBX_EINVAL, "bad argument");
    if (!h->lastBatch || frame < 0 || frame >= h->lastBatch) FAIL(ORBX_ESTATE, "frame %d not in the last extracted batch", frame);
    CK(cudaSetDevice(h->device));
    const LevelGeo& G = h->hp.lv[level];
    int cnt = 0;
    CK(cudaMemcpyAsync(&cnt, h->dCandCount + frame * h->prm.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *n = cnt;
    if (cnt > cap) FAIL(ORBX_ECAP, "cap %d < %d candidates", cap, cnt);
    if (cnt == 0) return ORBX_OK;
    std::vector<u64> rec(cnt);
    CK(cudaMemcpyAsync(rec.data(), h->dCand + G.candOff + (long long)frame * G.candCap, (size_t)cnt * sizeof(u64),
                       cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (int i = 0; i < cnt; i++) {  // pure unpacking of the device records (layout documented in orbx_internal.cuh)
        int x, y, s;
        cand_unpack(rec[i], G, x, y, s);
        xys[3 * i] = x + h->prm.edgeThreshold - 3;
        xys[3 * i + 1] = y + h->prm.edgeThreshold - 3;
        xys[3 * i + 2] = s;
    }
    return ORBX_OK;
}

extern "C" int orbx_stream(orbx_t* h, void** stream) {
    if (!h || !stream) FAIL(ORBX_EINVAL, "null argument");
    *stream = (void*)h->stream;
    return ORBX_OK;
}
extern "C" int64_t orbx_launch_count(orbx_t* h) { return h ? h->launches : 0; }
extern "C" int orbx_set_profiling(orbx_t* h, int on) {
    if (!h) FAIL(ORBX_EINVAL, "null handle");
    h->profiling = on != 0;
    return ORBX_OK;
}
extern "C" int orbx_stage_ms(orbx_t* h, float* ms7) {
    if (!h || !ms7) FAIL(ORBX_EINVAL, "null argument");
    for (int i = 0; i < 7; i++) ms7[i] = h->stageMs[i];
    return ORBX_OK;
}

// accessors used by orbx_match.cu (stereo works on the extractor's resident state)
struct OrbxView {
    const DevPlan* hp; DevPlan* dPlan; uint8_t* dPyr; orbx_kp* dKps; uint8_t* dDesc; int* dNOut; int lastBatch;
    cudaStream_t stream; int device; int outSpace; int* hCounts; const int* dSelCount;
};
OrbxView orbx_view(orbx_handle* h) {
    return OrbxView{&h->hp, h->dPlan, h->dPyr, h->dKps, h->dDesc, h->dNOut, h->lastBatch, h->stream, h->device, h->outSpace,
                    h->hCounts, h->dSelCount};
}
void orbx_count_launch(orbx_handle* h, int k) { h->launches += k; }
// Scratch of Frame::ComputeStereoMatches on the LEFT extractor's handle (freed by orbx_destroy): SAD per left keypoint and,
// for host outputs, the device copies of mvuRight / mvDepth.
int orbx_stereo_scratch(orbx_handle* h, size_t need, int** sad, float** u, float** z) {
    if (need > h->stCap) {
        cudaFree(h->dStSad); cudaFree(h->dStU); cudaFree(h->dStZ);
        h->dStSad = nullptr; h->dStU = h->dStZ = nullptr; h->stCap = 0;
        CK(cudaMalloc(&h->dStSad, need * 4));
        CK(cudaMalloc(&h->dStU, need * 4));
        CK(cudaMalloc(&h->dStZ, need * 4));
        h->stCap = need;
    }
    *sad = h->dStSad; *u = h->dStU; *z = h->dStZ;
    return ORBX_OK;
}
