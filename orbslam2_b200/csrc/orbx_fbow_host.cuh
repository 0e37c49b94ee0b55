// orbx_fbow_host.cuh — host side of the orbv_* entry points (included at the end of orbx_fbow.cu).
#pragma once

static void orbv_free_scratch(orbv_handle* v) {
    cudaFree(v->dIn); cudaFree(v->dWord); cudaFree(v->dNode); cudaFree(v->dWeight); cudaFree(v->dN);
    cudaFree(v->oWid); cudaFree(v->oNid); cudaFree(v->oFeat); cudaFree(v->oWw); cudaFree(v->oNoff); cudaFree(v->oCounts);
    v->dIn = nullptr; v->dWord = v->dNode = nullptr; v->dWeight = nullptr; v->dN = nullptr;
    v->oWid = v->oNid = v->oFeat = nullptr; v->oWw = nullptr; v->oNoff = v->oCounts = nullptr;
    v->cap = v->capN = v->oCap = v->oFrames = 0;
}

// Parses the bytes of a .fbow file (Vocabulary::toStream, fbow.cpp:164-171; params struct fbow.h:115-125) and uploads
// the tree in the device layout. Replaces Vocabulary::readFromFile / fromStream.
extern "C" int orbv_create(const void* blob_, size_t size, int device, orbv_t** out) {
    if (!blob_ || !out) FAIL(ORBX_EINVAL, "null argument");
    const uint8_t* blob = (const uint8_t*)blob_;
    if (size < 128) FAIL(ORBX_EINVAL, "vocabulary blob shorter than its header");
    uint64_t sig;
    memcpy(&sig, blob, 8);
    if (sig != FBOW_SIG) FAIL(ORBX_EINVAL, "invalid vocabulary signature (fbow.cpp:177)");
    const uint8_t* P = blob + 8;
    uint32_t nblocks, k;
    uint64_t descWp, blockSize, featOff, childOff, total;
    int32_t descType, descSize;
    memcpy(&nblocks, P + 56, 4);
    memcpy(&descWp, P + 64, 8);
    memcpy(&blockSize, P + 72, 8);
    memcpy(&featOff, P + 80, 8);
    memcpy(&childOff, P + 88, 8);
    memcpy(&total, P + 96, 8);
    memcpy(&descType, P + 104, 4);
    memcpy(&descSize, P + 108, 4);
    memcpy(&k, P + 112, 4);
    if (descType != 0 || descSize != 32)
        FAIL(ORBX_EINVAL, "only CV_8UC1 32-byte (ORB) vocabularies are supported (type %d, %d bytes)", descType, descSize);
    if (k < 1 || k > 255 || nblocks < 1) FAIL(ORBX_EINVAL, "bad vocabulary shape k=%u blocks=%u", k, nblocks);
    if (total != blockSize * (uint64_t)nblocks || size < 128 + total || descWp < 32 || childOff < featOff + k * descWp ||
        blockSize < childOff + 8ull * k)
        FAIL(ORBX_EINVAL, "inconsistent vocabulary header");
    const uint8_t* data = blob + 128;

    std::vector<uint8_t> desc((size_t)nblocks * k * 32), count(nblocks);
    std::vector<uint2> info((size_t)nblocks * k);
    uint64_t words = 0;
    for (uint32_t b = 0; b < nblocks; b++) {
        const uint8_t* B = data + (size_t)b * blockSize;
        uint16_t n, leaf;
        memcpy(&n, B, 2);
        memcpy(&leaf, B + 2, 2);
        if (n > k) FAIL(ORBX_EINVAL, "block %u holds %u nodes, more than k=%u", b, n, k);
        count[b] = (uint8_t)n;
        if (leaf) words += n;  // getTotalWords (fbow.cpp:185-198)
        for (uint32_t c = 0; c < k; c++) {
            memcpy(&desc[((size_t)b * k + c) * 32], B + featOff + c * descWp, 32);
            uint32_t id, w;
            memcpy(&id, B + childOff + 8 * c, 4);
            memcpy(&w, B + childOff + 8 * c + 4, 4);
            if (c < n && !(id & 0x80000000u) && id >= nblocks) FAIL(ORBX_EINVAL, "block %u child %u points outside the tree", b, c);
            info[(size_t)b * k + c] = make_uint2(id, w);
        }
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) FAIL(ORBX_ECUDA, "no CUDA device %d", device);
    CK(cudaSetDevice(device));
    orbv_handle* v = new orbv_handle();
    v->device = device;
    v->k = k; v->nblocks = nblocks; v->descSize = 32; v->totalWords = words;
    while ((1u << v->nbits) < k) v->nbits++;  // ceil(log2(k)) (fbow.h:411)
    CK(cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking));
    CK(cudaMalloc(&v->dDesc, desc.size()));
    CK(cudaMalloc(&v->dInfo, info.size() * sizeof(uint2)));
    CK(cudaMalloc(&v->dCount, count.size()));
    CK(cudaMalloc(&v->dErr, sizeof(int)));
    CK(cudaMemcpyAsync(v->dDesc, desc.data(), desc.size(), cudaMemcpyHostToDevice, v->stream));
    CK(cudaMemcpyAsync(v->dInfo, info.data(), info.size() * sizeof(uint2), cudaMemcpyHostToDevice, v->stream));
    CK(cudaMemcpyAsync(v->dCount, count.data(), count.size(), cudaMemcpyHostToDevice, v->stream));
    CK(cudaFuncSetAttribute(k_fbow_collect, cudaFuncAttributeMaxDynamicSharedMemorySize, FBOW_MAX_FEATURES * 8));
    CK(cudaStreamSynchronize(v->stream));
    *out = v;
    return ORBX_OK;
}

extern "C" int orbv_destroy(orbv_t* v) {
    if (!v) return ORBX_OK;
    cudaSetDevice(v->device);
    cudaStreamSynchronize(v->stream);
    orbv_free_scratch(v);
    cudaFree(v->dDesc); cudaFree(v->dInfo); cudaFree(v->dCount); cudaFree(v->dErr);
    cudaStreamDestroy(v->stream);
    delete v;
    return ORBX_OK;
}

extern "C" int orbv_info(orbv_t* v, uint32_t* k, uint32_t* nblocks, uint32_t* desc_size, uint64_t* total_words) {
    if (!v) FAIL(ORBX_EINVAL, "null handle");
    if (k) *k = v->k;
    if (nblocks) *nblocks = v->nblocks;
    if (desc_size) *desc_size = v->descSize;
    if (total_words) *total_words = v->totalWords;
    return ORBX_OK;
}

extern "C" int orbv_set_memory_space(orbv_t* v, int in, int out) {
    if (!v || (in != ORBX_MEM_HOST && in != ORBX_MEM_DEVICE) || (out != ORBX_MEM_HOST && out != ORBX_MEM_DEVICE))
        FAIL(ORBX_EINVAL, "bad memory space");
    v->inSpace = in;
    v->outSpace = out;
    return ORBX_OK;
}

extern "C" int orbv_stream(orbv_t* v, void** stream) {
    if (!v || !stream) FAIL(ORBX_EINVAL, "null argument");
    *stream = (void*)v->stream;
    return ORBX_OK;
}
extern "C" int64_t orbv_launch_count(orbv_t* v) { return v ? v->launches : 0; }

static int orbv_reserve(orbv_handle* v, size_t frames, size_t stride) {
    const size_t feats = frames * stride;
    if (feats > v->cap || frames > v->capN) {
        cudaFree(v->dWord); cudaFree(v->dNode); cudaFree(v->dWeight); cudaFree(v->dN); cudaFree(v->dIn); cudaFree(v->oCounts);
        v->dIn = nullptr;
        CK(cudaMalloc(&v->dWord, feats * 4));
        CK(cudaMalloc(&v->dNode, feats * 4));
        CK(cudaMalloc(&v->dWeight, feats * 4));
        CK(cudaMalloc(&v->dN, (frames * (stride + 2)) * sizeof(int)));  // counts + run-offset scratch
        CK(cudaMalloc(&v->oCounts, 2 * frames * sizeof(int)));
        v->cap = feats;
        v->capN = frames;
    }
    if (v->inSpace == ORBX_MEM_HOST && !v->dIn) CK(cudaMalloc(&v->dIn, v->cap * 32));
    if (v->outSpace == ORBX_MEM_HOST && (feats > v->oCap || frames > v->oFrames)) {
        cudaFree(v->oWid); cudaFree(v->oNid); cudaFree(v->oFeat); cudaFree(v->oWw); cudaFree(v->oNoff);
        CK(cudaMalloc(&v->oWid, feats * 4));
        CK(cudaMalloc(&v->oNid, feats * 4));
        CK(cudaMalloc(&v->oFeat, feats * 4));
        CK(cudaMalloc(&v->oWw, feats * 4));
        CK(cudaMalloc(&v->oNoff, (feats + frames) * sizeof(int)));
        v->oCap = feats;
        v->oFrames = frames;
    }
    return ORBX_OK;
}

// Vocabulary::transform(features, level, fBow&, fBow2&) (level >= 0) and Vocabulary::transform(features) (level < 0,
// normalize = 1) for `batch` descriptor matrices: frame f = desc + f*stride*32, n[f] rows. n, n_words, n_nodes are
// host arrays; the other pointers live in the handle's memory spaces.
extern "C" int orbv_transform_batch(orbv_t* v, const uint8_t* desc, const int* n, int batch, int stride, int level,
                                    int normalize, uint32_t* word_id, float* word_w, int* n_words, uint32_t* node_id,
                                    int* node_off, uint32_t* node_feat, int* n_nodes) {
    if (!v || !desc || !n || !word_id || !word_w || !n_words) FAIL(ORBX_EINVAL, "null argument");
    if (level >= 0 && (!node_id || !node_off || !node_feat || !n_nodes)) FAIL(ORBX_EINVAL, "null node outputs");
    if (batch < 1 || stride < 1) FAIL(ORBX_EINVAL, "bad batch/stride");
    int nmax = 0;
    for (int f = 0; f < batch; f++) {
        // the reference throws "No input data" on an empty matrix (fbow.cpp:53); Frame::ComputeFboW never calls it
        // then (src/Frame.cc:398) — here an empty frame simply yields empty results
        if (n[f] < 0 || n[f] > stride) FAIL(ORBX_EINVAL, "frame %d: %d features with stride %d", f, n[f], stride);
        nmax = std::max(nmax, n[f]);
    }
    if (nmax > FBOW_MAX_FEATURES) FAIL(ORBX_ECAP, "%d features in one frame (limit %d)", nmax, FBOW_MAX_FEATURES);
    CK(cudaSetDevice(v->device));
    int rc = orbv_reserve(v, batch, stride);
    if (rc != ORBX_OK) return rc;
    cudaStream_t st = v->stream;
    const size_t feats = (size_t)batch * stride;
    const uint8_t* dDesc = desc;
    if (v->inSpace == ORBX_MEM_HOST) {
        const size_t bytes = batch == 1 ? (size_t)n[0] * 32 : feats * 32;
        if (bytes) CK(cudaMemcpyAsync(v->dIn, desc, bytes, cudaMemcpyHostToDevice, st));
        dDesc = v->dIn;
    }
    int* dN = v->dN;
    int* dRunOff = v->dN + batch;
    CK(cudaMemcpyAsync(dN, n, batch * sizeof(int), cudaMemcpyHostToDevice, st));
    const bool hostOut = v->outSpace == ORBX_MEM_HOST;
    uint32_t* oWid = hostOut ? v->oWid : word_id;
    float* oWw = hostOut ? v->oWw : word_w;
    uint32_t* oNid = hostOut ? v->oNid : node_id;
    int* oNoff = hostOut ? v->oNoff : node_off;
    uint32_t* oFeat = hostOut ? v->oFeat : node_feat;
    if (nmax > 0) {
        k_fbow_descend<<<dim3((nmax + 127) / 128, batch), 128, 0, st>>>(v->dDesc, v->dInfo, v->dCount, (int)v->k, v->nbits,
                                                                       (const uint4*)dDesc, dN, stride, level, v->dWord,
                                                                       v->dWeight, v->dNode);
        v->launches++;
    }
    int P = 2;
    while (P < nmax) P <<= 1;
    k_fbow_collect<<<batch, COLLECT_THREADS, (size_t)P * 8, st>>>(v->dWord, v->dWeight, v->dNode, dN, stride, level,
                                                                  normalize, oWid, oWw, v->oCounts, oNid, oNoff, oFeat,
                                                                  v->oCounts + batch, dRunOff);
    v->launches++;
    CK(cudaGetLastError());
    if (hostOut) {
        CK(cudaMemcpyAsync(word_id, oWid, feats * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(word_w, oWw, feats * 4, cudaMemcpyDeviceToHost, st));
        if (level >= 0) {
            CK(cudaMemcpyAsync(node_id, oNid, feats * 4, cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpyAsync(node_off, oNoff, (feats + batch) * sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpyAsync(node_feat, oFeat, feats * 4, cudaMemcpyDeviceToHost, st));
        }
    }
    CK(cudaMemcpyAsync(n_words, v->oCounts, batch * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (n_nodes) CK(cudaMemcpyAsync(n_nodes, v->oCounts + batch, batch * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return ORBX_OK;
}

// fBow::score(q, db[i]) for m id-sorted database vectors in CSR form (db_off has m+1 entries): what
// KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates evaluate per candidate keyframe
// (src/KeyFrameDatabase.cc:130,247). All pointers are host pointers.
extern "C" int orbv_score_batch(orbv_t* v, const uint32_t* q_id, const float* q_w, int nq, const uint32_t* db_id,
                                const float* db_w, const int* db_off, int m, double* score) {
    if (!db_off || !score || nq < 0 || m < 0) FAIL(ORBX_EINVAL, "bad argument");
    if (m == 0) return ORBX_OK;
    if (v) CK(cudaSetDevice(v->device));
    const int total = db_off[m];
    cudaStream_t st = v ? v->stream : cudaStreamPerThread;  // fBow::score is static: no vocabulary needed
    uint8_t* buf = nullptr;
    const size_t bq = (size_t)std::max(nq, 1) * 4, bd = (size_t)std::max(total, 1) * 4, bo = (size_t)(m + 1) * 4;
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    CK(cudaMallocAsync((void**)&buf, 2 * up(bq) + 2 * up(bd) + up(bo) + up((size_t)m * 8), st));
    uint8_t* p = buf;
    uint32_t* dq = (uint32_t*)p; p += up(bq);
    float* dqw = (float*)p; p += up(bq);
    uint32_t* dd = (uint32_t*)p; p += up(bd);
    float* ddw = (float*)p; p += up(bd);
    int* doff = (int*)p; p += up(bo);
    double* ds = (double*)p;
    if (nq) {
        CK(cudaMemcpyAsync(dq, q_id, (size_t)nq * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(dqw, q_w, (size_t)nq * 4, cudaMemcpyHostToDevice, st));
    }
    if (total) {
        CK(cudaMemcpyAsync(dd, db_id, (size_t)total * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(ddw, db_w, (size_t)total * 4, cudaMemcpyHostToDevice, st));
    }
    CK(cudaMemcpyAsync(doff, db_off, bo, cudaMemcpyHostToDevice, st));
    k_fbow_score<<<(m + 127) / 128, 128, 0, st>>>(dq, dqw, nq, dd, ddw, doff, m, ds);
    if (v) v->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(score, ds, (size_t)m * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaFreeAsync(buf, st));
    CK(cudaStreamSynchronize(st));
    return ORBX_OK;
}
