// orbx_knn_tc.cuh — brute-force Hamming 2-NN on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// Reference: the distance is ORBmatcher::DescriptorDistance (/root/reference/src/ORBmatcher.cc:1643-1659), the scan
// rule (strict '<' in ascending index order, second best with multiplicity) is the loop at src/ORBmatcher.cc:195-229.
//
// Hamming distance over 256-bit descriptors is a binary GEMM:  d(q,t) = |q| + |t| - 2 <q,t>.  The descriptors are
// expanded once to int8 rows of K' = 288 entries:
//     queries  (A): bit -> -16,  then 32 extra entries  [ a0..a4 (sum = 2|q|), 64 x5, 1, 0... ]
//     train    (B): bit -> +16,  then 32 extra entries  [ 64 x5, b0..b4 (sum = 2|t|), (col & 255) >> 1, 0... ]
// so that ONE tcgen05.mma.kind::i8 chain (9 instructions of K = 32) leaves in TMEM, for every (query, train) pair,
//     key = 128 * d(q,t) + ((col & 255) >> 1)           0 <= key <= 32895, exact in the int32 accumulator.
// The epilogue reads the accumulators with tcgen05.ld ... .pack::16b (two adjacent columns -> one 32-bit register) and
// keeps the two smallest keys of the even and of the odd columns with packed 16-bit min/max (VIMNMX.U16x2): 2.5
// instructions per register = 1.25 per descriptor pair. The column in the low bits makes the lowest index win ties; the
// second-smallest KEY is the second-smallest distance with multiplicity.
//
// CTA = 128 queries x one contiguous range of 256-row train tiles. Warp 0: TMA producer (cp.async.bulk.tensor, 128B
// swizzle), warp 1: TMEM allocator + MMA issuer, warps 2..9: epilogue (two sets of four warps, one per TMEM lane
// quadrant; the sets alternate tiles = accumulator buffers, so the MMA of tile i+1 runs under the epilogue of tile i).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define KTC_M 128            // queries per CTA  (UMMA M)
#define KTC_N 256            // train rows per tile (UMMA N)
#define KTC_KMAIN 256        // int8 entries carrying the descriptor bits
#define KTC_KEXT 32          // extra entries (norms + column index)
#define KTC_STAGES 2
#define KTC_THREADS 320
#define KTC_A_MAIN_BYTES (KTC_M * 128)          // one 128-byte K block of the A tile
#define KTC_A_EXT_BYTES (KTC_M * KTC_KEXT)
#define KTC_B_MAIN_BYTES (KTC_N * 128)
#define KTC_B_EXT_BYTES (KTC_N * KTC_KEXT)
#define KTC_A_BYTES (2 * KTC_A_MAIN_BYTES + KTC_A_EXT_BYTES)   // 36864
#define KTC_B_BYTES (2 * KTC_B_MAIN_BYTES + KTC_B_EXT_BYTES)   // 73728
#define KTC_SMEM (1024 + KTC_A_BYTES + KTC_STAGES * KTC_B_BYTES + 256)

namespace ktc {

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Bounded wait: a protocol error traps (the launch fails) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (int spins = 0; !done; ++spins) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (!done && (spins & 1023) == 1023) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 4000000000ll) __trap();   // ~2 s at 1.9 GHz
        }
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
}
// K-major shared-memory matrix descriptors (cute::UMMA::SmemDescriptor): start address >> 4 in [0,14), leading byte
// offset >> 4 in [16,30) (unused for swizzled K-major: 1), stride byte offset >> 4 in [32,46) = distance between
// 8-row groups, version 1 in [46,48), layout type in [61,64): 2 = SWIZZLE_128B, 6 = SWIZZLE_32B.
__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr) {
    return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ uint64_t desc_sw32(uint32_t addr) {
    return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) | (6ull << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D = S32, A = B = signed int8, both K-major, N >> 3, M >> 4.
__host__ __device__ constexpr uint32_t idesc_i8(int M, int N) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

#define KTC_LD32(PACKQ, r, addr)                                                                                       \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32" PACKQ ".b32 "                                                    \
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"                                             \
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"                            \
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),     \
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),            \
                   "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]),          \
                   "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]),          \
                   "=r"(r[29]), "=r"(r[30]), "=r"(r[31])                                                               \
                 : "r"(addr))
// The wait names the destination registers as in/out operands so that no use of them can be scheduled above it.
#define KTC_WAIT32(r)                                                                                                  \
    asm volatile("tcgen05.wait::ld.sync.aligned;"                                                                      \
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),     \
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]),            \
                   "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]),          \
                   "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]),          \
                   "+r"(r[29]), "+r"(r[30]), "+r"(r[31])                                                               \
                 :                                                                                                     \
                 : "memory")

// two smallest of {m1 <= m2, k0, k1} per 16-bit lane: 5 packed instructions per two registers
__device__ __forceinline__ void push2(uint32_t& m1, uint32_t& m2, uint32_t k0, uint32_t k1) {
    const uint32_t lo = __vminu2(k0, k1), hi = __vmaxu2(k0, k1);
    m2 = __vimin3_u16x2(m2, __vmaxu2(m1, lo), hi);
    m1 = __vminu2(m1, lo);
}
__device__ __forceinline__ void push_full(uint32_t& g1, uint32_t& g2, uint32_t k) {
    const uint32_t t = max(g1, k);
    g1 = min(g1, k);
    g2 = min(g2, t);
}

}  // namespace ktc

// ------------------------------------------------------------------------------------------------ expansion
// One thread per 32-bit word of a descriptor: 32 int8 entries (two 16-byte stores). The eight threads of a row reduce
// the popcount; the first writes the row's 32 extra entries. Rows >= n (padding of the last query tile) are zero.
__global__ void __launch_bounds__(256) k_knn_expand(const uint32_t* __restrict__ bits, int n, int rowsPadded, int isA,
                                                    uint8_t* __restrict__ mainOut, uint8_t* __restrict__ extOut) {
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const int row = gid >> 3, w = gid & 7;
    const bool live = row < rowsPadded;
    const uint32_t x = (live && row < n) ? bits[(size_t)row * 8 + w] : 0u;
    const uint32_t val = isA ? 0xF0u : 0x10u;   // -16 / +16
    uint32_t o[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t nib = (x >> (4 * i)) & 15u;
        o[i] = ((nib * 0x00204081u) & 0x01010101u) * val;   // bit j of the nibble -> byte j
    }
    int pc = __popc(x);
    pc += __shfl_xor_sync(0xffffffffu, pc, 1);
    pc += __shfl_xor_sync(0xffffffffu, pc, 2);
    pc += __shfl_xor_sync(0xffffffffu, pc, 4);
    if (!live) return;
    uint4* mo = (uint4*)(mainOut + (size_t)row * KTC_KMAIN + w * 32);
    mo[0] = make_uint4(o[0], o[1], o[2], o[3]);
    mo[1] = make_uint4(o[4], o[5], o[6], o[7]);
    if (w == 0) {
        uint32_t p[5];
        int rem = 2 * pc;   // <= 512, split over five entries of at most 127
#pragma unroll
        for (int i = 0; i < 5; i++) {
            p[i] = rem < 127 ? rem : 127;
            rem -= p[i];
        }
        // entries 0..4 | 5..9 | 10 : A = norm parts, 64 x5, 1 ; B = 64 x5, norm parts, (column in tile) >> 1
        const uint32_t n0 = p[0] | (p[1] << 8) | (p[2] << 16) | (p[3] << 24), c0 = 0x40404040u;
        const uint32_t last = isA ? 1u : (uint32_t)((row & (KTC_N - 1)) >> 1);
        uint4 e0;
        if (isA) e0 = make_uint4(n0, p[4] | (c0 << 8), 0x4040u | (last << 16), 0u);
        else e0 = make_uint4(c0, 0x40u | (n0 << 8), (n0 >> 24) | (p[4] << 8) | (last << 16), 0u);
        if (row >= n) e0 = make_uint4(0u, 0u, 0u, 0u);   // padding row: all zero (its results are never written)
        uint4* eo = (uint4*)(extOut + (size_t)row * KTC_KEXT);
        eo[0] = e0;
        eo[1] = make_uint4(0u, 0u, 0u, 0u);
    }
}

// ------------------------------------------------------------------------------------------------ main kernel
// grid (query tiles, splits). Split s covers train tiles [s*tilesPerSplit, min(nTiles, (s+1)*tilesPerSplit)).
// partial[(size_t)s * nq + q] = (d1, i1, d2, 0), i1 a global train index.
__global__ void __launch_bounds__(KTC_THREADS, 1)
k_knn_tc(const __grid_constant__ CUtensorMap mapAm, const __grid_constant__ CUtensorMap mapAe,
         const __grid_constant__ CUtensorMap mapBm, const __grid_constant__ CUtensorMap mapBe, int nq, int nTiles,
         int tilesPerSplit, int4* __restrict__ partial) {
    using namespace ktc;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (s32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sA = base;                                   // [Amain k0 | Amain k1 | Aext]
    const uint32_t sB = base + KTC_A_BYTES;                     // stages of [Bmain k0 | Bmain k1 | Bext]
    const uint32_t sBar = sB + KTC_STAGES * KTC_B_BYTES;        // barriers (8 B each)
    const uint32_t barA = sBar, barBFull = sBar + 8, barBEmpty = barBFull + 8 * KTC_STAGES,
                   barAccFull = barBEmpty + 8 * KTC_STAGES, barAccEmpty = barAccFull + 16;
    uint8_t* gen = smem_raw + (base - s32(smem_raw));
    uint32_t* tmemSlot = (uint32_t*)(gen + KTC_A_BYTES + KTC_STAGES * KTC_B_BYTES + 128);
    uint2* xchg = (uint2*)gen;   // reused after the main loop (A tile is dead by then): 128 x (g1, g2)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qTile = blockIdx.x, split = blockIdx.y;
    const int tile0 = split * tilesPerSplit;
    const int nLocal = min(nTiles, tile0 + tilesPerSplit) - tile0;

    if (threadIdx.x == 0) {
        mbar_init(barA, 1);
        for (int s = 0; s < KTC_STAGES; s++) {
            mbar_init(barBFull + 8 * s, 1);
            mbar_init(barBEmpty + 8 * s, 1);
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(barAccFull + 8 * b, 1);
            mbar_init(barAccEmpty + 8 * b, 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(s32(tmemSlot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmemSlot;

    uint32_t g1 = 0xFFFFFFFFu, g2 = 0xFFFFFFFFu;   // (dist << 22 | train row inside this split), two smallest

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(barA, KTC_A_BYTES);
            tma_load_2d(sA, &mapAm, 0, qTile * KTC_M, barA);
            tma_load_2d(sA + KTC_A_MAIN_BYTES, &mapAm, 128, qTile * KTC_M, barA);
            tma_load_2d(sA + 2 * KTC_A_MAIN_BYTES, &mapAe, 0, qTile * KTC_M, barA);
            for (int t = 0; t < nLocal; t++) {
                const int s = t % KTC_STAGES, use = t / KTC_STAGES;
                mbar_wait(barBEmpty + 8 * s, (use & 1) ^ 1);
                const uint32_t dst = sB + s * KTC_B_BYTES, bar = barBFull + 8 * s;
                const int row = (tile0 + t) * KTC_N;
                mbar_expect_tx(bar, KTC_B_BYTES);
                tma_load_2d(dst, &mapBm, 0, row, bar);
                tma_load_2d(dst + KTC_B_MAIN_BYTES, &mapBm, 128, row, bar);
                tma_load_2d(dst + 2 * KTC_B_MAIN_BYTES, &mapBe, 0, row, bar);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = idesc_i8(KTC_M, KTC_N);
            mbar_wait(barA, 0);
            for (int t = 0; t < nLocal; t++) {
                const int s = t % KTC_STAGES, use = t / KTC_STAGES;
                const int buf = t & 1, bufUse = t >> 1;
                mbar_wait(barAccEmpty + 8 * buf, (bufUse & 1) ^ 1);
                mbar_wait(barBFull + 8 * s, use & 1);
                tc_fence_after();
                const uint32_t d = tmem + buf * KTC_N;
                const uint32_t bS = sB + s * KTC_B_BYTES;
#pragma unroll
                for (int kb = 0; kb < 2; kb++)
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        mma_i8(d, desc_sw128(sA + kb * KTC_A_MAIN_BYTES + k * 32),
                               desc_sw128(bS + kb * KTC_B_MAIN_BYTES + k * 32), idesc, (kb | k) != 0);
                mma_i8(d, desc_sw32(sA + 2 * KTC_A_MAIN_BYTES), desc_sw32(bS + 2 * KTC_B_MAIN_BYTES), idesc, 1);
                mma_commit(barBEmpty + 8 * s);      // the stage's shared memory may be refilled
                mma_commit(barAccFull + 8 * buf);   // the accumulator is complete
            }
        }
    } else {
        const int quad = warp & 3, set = (warp - 2) >> 2;
        const uint32_t laneBase = tmem + ((uint32_t)(quad * 32) << 16);
        for (int t = set; t < nLocal; t += 2) {
            const int bufUse = t >> 1;
            mbar_wait(barAccFull + 8 * set, bufUse & 1);
            tc_fence_after();
            const uint32_t acc = laneBase + set * KTC_N;
            uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
            uint32_t ra[32], rb[32];
            KTC_LD32(".pack::16b", ra, acc);
            KTC_LD32(".pack::16b", rb, acc + 64);
            KTC_WAIT32(ra);
            KTC_WAIT32(rb);
#pragma unroll
            for (int i = 0; i < 32; i += 2) push2(m1, m2, ra[i], ra[i + 1]);
            KTC_LD32(".pack::16b", ra, acc + 128);
#pragma unroll
            for (int i = 0; i < 32; i += 2) push2(m1, m2, rb[i], rb[i + 1]);
            KTC_LD32(".pack::16b", rb, acc + 192);
            KTC_WAIT32(ra);
            KTC_WAIT32(rb);
            tc_fence_before();
            mbar_arrive(barAccEmpty + 8 * set);
#pragma unroll
            for (int i = 0; i < 32; i += 2) push2(m1, m2, ra[i], ra[i + 1]);
#pragma unroll
            for (int i = 0; i < 32; i += 2) push2(m1, m2, rb[i], rb[i + 1]);
            // tile result -> running best two of this thread's row
            const uint32_t e1 = m1 & 0xFFFFu, o1 = m1 >> 16;
            if ((min(e1, o1) >> 7) <= (g2 >> 22)) {
                const uint32_t rowBase = (uint32_t)t * KTC_N;
                const uint32_t e2 = m2 & 0xFFFFu, o2 = m2 >> 16;
                push_full(g1, g2, ((e1 >> 7) << 22) | (rowBase + 2 * (e1 & 127)));
                push_full(g1, g2, ((o1 >> 7) << 22) | (rowBase + 2 * (o1 & 127) + 1));
                push_full(g1, g2, ((e2 >> 7) << 22) | (rowBase + 2 * (e2 & 127)));
                push_full(g1, g2, ((o2 >> 7) << 22) | (rowBase + 2 * (o2 & 127) + 1));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    // the odd-tile set hands its pair to the even-tile set through shared memory (the A tile is no longer read)
    const int quad = warp & 3, row = quad * 32 + lane;
    if (warp >= 6) xchg[row] = make_uint2(g1, g2);
    __syncthreads();
    if (warp >= 2 && warp < 6) {
        const uint2 o = xchg[row];
        push_full(g1, g2, o.x);
        push_full(g1, g2, o.y);
        const int q = qTile * KTC_M + row;
        if (q < nq) {
            const int rowOff = tile0 * KTC_N;
            const int d1 = (int)(g1 >> 22), d2 = g2 == 0xFFFFFFFFu ? 256 : (int)(g2 >> 22);
            partial[(size_t)split * nq + q] =
                g1 == 0xFFFFFFFFu ? make_int4(256, -1, 256, 0) : make_int4(d1, rowOff + (int)(g1 & 0x3FFFFFu), d2, 0);
        }
    }
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*ktc_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 2-D row-major int8 matrix [rows][rowBytes], box = boxInner bytes x boxRows rows
static inline CUresult ktc_make_map(ktc_encode_fn enc, CUtensorMap* m, const void* basePtr, uint64_t rows,
                                    uint32_t rowBytes, uint32_t boxInner, uint32_t boxRows, CUtensorMapSwizzle sw) {
    cuuint64_t gdim[2] = {rowBytes, rows};
    cuuint64_t gstr[1] = {rowBytes};
    cuuint32_t box[2] = {boxInner, boxRows};
    cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(basePtr), gdim, gstr, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

// Expanded operands in, per-split partial results out. qRows = padded query rows (multiple of KTC_M).
static inline int ktc_launch(ktc_encode_fn enc, const uint8_t* dAm, const uint8_t* dAe, int qRows, const uint8_t* dBm,
                             const uint8_t* dBe, int nTiles, int nq, int tilesPerSplit, int splits, int4* partial,
                             cudaStream_t stream) {
    CUtensorMap mAm, mAe, mBm, mBe;
    if (ktc_make_map(enc, &mAm, dAm, qRows, KTC_KMAIN, 128, KTC_M, CU_TENSOR_MAP_SWIZZLE_128B) != CUDA_SUCCESS) return 1;
    if (ktc_make_map(enc, &mAe, dAe, qRows, KTC_KEXT, KTC_KEXT, KTC_M, CU_TENSOR_MAP_SWIZZLE_32B) != CUDA_SUCCESS) return 2;
    if (ktc_make_map(enc, &mBm, dBm, (uint64_t)nTiles * KTC_N, KTC_KMAIN, 128, KTC_N, CU_TENSOR_MAP_SWIZZLE_128B) !=
        CUDA_SUCCESS)
        return 3;
    if (ktc_make_map(enc, &mBe, dBe, (uint64_t)nTiles * KTC_N, KTC_KEXT, KTC_KEXT, KTC_N, CU_TENSOR_MAP_SWIZZLE_32B) !=
        CUDA_SUCCESS)
        return 4;
    if (cudaFuncSetAttribute(k_knn_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, KTC_SMEM) != cudaSuccess) return 5;
    k_knn_tc<<<dim3(qRows / KTC_M, splits), KTC_THREADS, KTC_SMEM, stream>>>(mAm, mAe, mBm, mBe, nq, nTiles,
                                                                             tilesPerSplit, partial);
    return cudaGetLastError() == cudaSuccess ? 0 : 6;
}
