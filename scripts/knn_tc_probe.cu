// Developer probe for csrc/orbx_knn_tc.cuh (run on the GPU box through gpurun):
//   knn_tc_probe pack            prints what tcgen05.ld .pack::16b returns for a known TMEM pattern
//   knn_tc_probe run NQ NT [S]   expands, runs k_knn_tc, compares with a CPU brute force, prints the rate
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o scripts/knn_tc_probe scripts/knn_tc_probe.cu -lcuda
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "../orbslam2_b200/csrc/orbx_knn_tc.cuh"

#define CHECK(x)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (x);                                                                 \
        if (e_ != cudaSuccess) {                                                              \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);   \
            exit(2);                                                                          \
        }                                                                                     \
    } while (0)

__global__ void k_pack_probe(uint32_t* out) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(ktc::s32(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    ktc::tc_fence_before();
    __syncthreads();
    ktc::tc_fence_after();
    const uint32_t tmem = slot;
    const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16);
    // column c of row r holds 0x7000_0000 | (c+1) << 16 ... low half = 0x100 * r + c
    for (int c = 0; c < 128; c += 8) {
        uint32_t v[8];
        for (int j = 0; j < 8; j++) v[j] = 0x70000000u | ((uint32_t)(c + j + 1) << 16) | (uint32_t)((warp * 32 + lane) * 256 + c + j);
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(addr + c), "r"(v[0]),
                     "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                     : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    uint32_t r[32], p[32];
    KTC_LD32("", r, addr);
    KTC_WAIT32(r);
    KTC_LD32(".pack::16b", p, addr);
    KTC_WAIT32(p);
    if (threadIdx.x == 37) {
        for (int i = 0; i < 32; i++) out[i] = r[i];
        for (int i = 0; i < 32; i++) out[32 + i] = p[i];
    }
    ktc::tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem) : "memory");
}

static int popc256(const uint32_t* a, const uint32_t* b) {
    int d = 0;
    for (int i = 0; i < 8; i++) d += __builtin_popcount(a[i] ^ b[i]);
    return d;
}

int main(int argc, char** argv) {
    const char* mode = argc > 1 ? argv[1] : "pack";
    if (!strcmp(mode, "pack")) {
        uint32_t* d;
        CHECK(cudaMalloc(&d, 64 * 4));
        k_pack_probe<<<1, 128>>>(d);
        CHECK(cudaDeviceSynchronize());
        uint32_t h[64];
        CHECK(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
        printf("thread 37 (row 37), plain .x32:\n");
        for (int i = 0; i < 32; i++) printf("%08x%c", h[i], i % 8 == 7 ? '\n' : ' ');
        printf("pack::16b .x32:\n");
        for (int i = 0; i < 32; i++) printf("%08x%c", h[32 + i], i % 8 == 7 ? '\n' : ' ');
        return 0;
    }
    const int nq = argc > 2 ? atoi(argv[2]) : 1000, nt = argc > 3 ? atoi(argv[3]) : 4096;
    int splits = argc > 4 ? atoi(argv[4]) : 1;
    const int reps = argc > 5 ? atoi(argv[5]) : 1;
    const int nTiles = nt / KTC_N;
    if (nTiles < 1) { printf("nt too small\n"); return 1; }
    int tilesPerSplit = (nTiles + splits - 1) / splits;
    splits = (nTiles + tilesPerSplit - 1) / tilesPerSplit;
    const int qRows = (nq + KTC_M - 1) / KTC_M * KTC_M;
    std::mt19937 rng(5);
    std::vector<uint32_t> hq((size_t)nq * 8), ht((size_t)nt * 8);
    for (auto& x : ht) x = rng();
    for (int i = 0; i < nq; i++) {
        if (i % 3 == 0) for (int k = 0; k < 8; k++) hq[(size_t)i * 8 + k] = rng();
        else {   // a train row with a few flipped bits (ties and near matches)
            const int src = rng() % nt;
            for (int k = 0; k < 8; k++) hq[(size_t)i * 8 + k] = ht[(size_t)src * 8 + k];
            const int flips = rng() % 40;
            for (int f = 0; f < flips; f++) { const int b = rng() % 256; hq[(size_t)i * 8 + b / 32] ^= 1u << (b % 32); }
        }
    }
    for (int i = 0; i + 1 < nt && i < 2000; i += 97)   // exact duplicates in the train set: lowest index must win
        for (int k = 0; k < 8; k++) ht[(size_t)(nt - 1 - i) * 8 + k] = ht[(size_t)i * 8 + k];
    if (nq > 2) { for (int k = 0; k < 8; k++) { hq[k] = 0; hq[8 + k] = 0xFFFFFFFFu; } }
    uint32_t *dq, *dt;
    uint8_t *dAm, *dAe, *dBm, *dBe;
    int4* dPart;
    CHECK(cudaMalloc(&dq, hq.size() * 4));
    CHECK(cudaMalloc(&dt, ht.size() * 4));
    CHECK(cudaMalloc(&dAm, (size_t)qRows * KTC_KMAIN));
    CHECK(cudaMalloc(&dAe, (size_t)qRows * KTC_KEXT));
    CHECK(cudaMalloc(&dBm, (size_t)nTiles * KTC_N * KTC_KMAIN));
    CHECK(cudaMalloc(&dBe, (size_t)nTiles * KTC_N * KTC_KEXT));
    CHECK(cudaMalloc(&dPart, (size_t)splits * nq * sizeof(int4)));
    CHECK(cudaMemcpy(dq, hq.data(), hq.size() * 4, cudaMemcpyHostToDevice));
    CHECK(cudaMemcpy(dt, ht.data(), ht.size() * 4, cudaMemcpyHostToDevice));
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr));
    ktc_encode_fn enc = (ktc_encode_fn)p;
    cudaEvent_t e0, e1, e2;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
    float msExpand = 0, msMain = 0;
    for (int rep = 0; rep < reps + 1; rep++) {
        cudaEventRecord(e0);
        k_knn_expand<<<(qRows * 8 + 255) / 256, 256>>>(dq, nq, qRows, 1, dAm, dAe);
        k_knn_expand<<<(nTiles * KTC_N * 8 + 255) / 256, 256>>>(dt, nTiles * KTC_N, nTiles * KTC_N, 0, dBm, dBe);
        cudaEventRecord(e1);
        const int rc = ktc_launch(enc, dAm, dAe, qRows, dBm, dBe, nTiles, nq, tilesPerSplit, splits, dPart, 0);
        if (rc) { printf("ktc_launch rc=%d (%s)\n", rc, cudaGetErrorString(cudaGetLastError())); return 2; }
        cudaEventRecord(e2);
        CHECK(cudaDeviceSynchronize());
        cudaEventElapsedTime(&msExpand, e0, e1);
        cudaEventElapsedTime(&msMain, e1, e2);
    }
    const double pairs = (double)nq * nTiles * KTC_N;
    printf("nq %d nt %d (tiles %d, splits %d x %d): expand %.3f ms, k_knn_tc %.3f ms = %.1f Gpairs/s\n", nq, nt, nTiles,
           splits, tilesPerSplit, msExpand, msMain, pairs / msMain / 1e6);
    std::vector<int4> part((size_t)splits * nq);
    CHECK(cudaMemcpy(part.data(), dPart, part.size() * sizeof(int4), cudaMemcpyDeviceToHost));
    // CPU check on a sample of queries
    const int step = nq > 600 ? nq / 600 : 1;
    int bad = 0, checked = 0;
    for (int q = 0; q < nq; q += step) {
        for (int s = 0; s < splits; s++) {
            const int r0 = s * tilesPerSplit * KTC_N, r1 = std::min(nTiles, (s + 1) * tilesPerSplit) * KTC_N;
            int d1 = 256, i1 = -1, d2 = 256;
            for (int j = r0; j < r1; j++) {
                const int d = popc256(&hq[(size_t)q * 8], &ht[(size_t)j * 8]);
                if (d < d1) { d2 = d1; d1 = d; i1 = j; } else if (d < d2) d2 = d;
            }
            const int4 g = part[(size_t)s * nq + q];
            checked++;
            if (g.x != d1 || g.y != i1 || g.z != d2) {
                if (bad < 10) printf("MISMATCH q %d split %d: gpu (%d,%d,%d) cpu (%d,%d,%d)\n", q, s, g.x, g.y, g.z, d1, i1, d2);
                bad++;
            }
        }
    }
    printf("checked %d (query, split) results: %d mismatches\n", checked, bad);
    return bad ? 1 : 0;
}
