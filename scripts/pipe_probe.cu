// pipe_probe.cu — issue-rate probe for the instructions the FAST / blur kernels are built from (sm_100a).
// Each kernel runs a long dependent-free stream of ONE instruction kind (8 independent chains per thread) so the
// measured rate is the pipe's, not the latency's. Output: warp-instructions per clock per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/pipe_probe scripts/pipe_probe.cu && scripts/pipe_probe
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdio>

#define ITERS 4096
#define CHAINS 8

template <int OP>
__device__ __forceinline__ unsigned op(unsigned a, unsigned b, unsigned c) {
    if (OP == 0) return __vimin3_u16x2(a, b, c);
    if (OP == 1) return __vimax3_u16x2(a, b, c);
    if (OP == 2) {  // HMNMX2 on integer-valued 16-bit lanes (fp16 subnormals order like integers)
        __half2 x = *(__half2*)&a, y = *(__half2*)&b;
        __half2 r = __hmin2(x, y);
        return *(unsigned*)&r;
    }
    if (OP == 3) return __byte_perm(a, b, c);
    if (OP == 4) return (a & b) | c;           // LOP3
    if (OP == 5) return a * b + c;             // IMAD
    if (OP == 6) return __funnelshift_r(a, b, c);
    if (OP == 7) return __vminu2(a, b);        // VIMNMX.U16x2 two-input
    if (OP == 8) return __viaddmax_s16x2(a, b, c);
    if (OP == 9) return __umulhi(a, b) + c;    // IMAD.HI
    if (OP == 10) return __popc(a) + b;
    if (OP == 11) return (unsigned)__dp4a(a, b, c);
    if (OP == 12) { unsigned d; asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
    if (OP == 13) return a + b + c;            // IADD3
    if (OP == 14) return (a > b) ? c : a;      // ISETP + SEL
    if (OP == 15) return (a << 8);             // shift (IMAD.SHL or SHF)
    return a;
}

template <int OP>
__global__ void probe(unsigned* out, unsigned seed) {
    unsigned r[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; i++) r[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
    unsigned b = seed ^ 0x5bd1e995u, c = (OP == 3) ? 0x5410u : ((OP == 6) ? 8u : (seed | 1u));
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < CHAINS; i++) r[i] = op<OP>(r[i], b, c);
    }
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) s ^= r[i];
    if (s == 0x12345678u) out[0] = s;
}

// cross-chain variant: operand b is the neighbouring chain, so nothing is loop invariant or idempotent
template <int OP>
__global__ void probe_x(unsigned* out, unsigned seed) {
    unsigned r[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; i++) r[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
    unsigned c = seed | 1u;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < CHAINS; i++) {
            if (OP == 0) asm("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(r[(i + 1) % CHAINS]), "r"(c));
            if (OP == 1) asm("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(r[i]) : "r"(r[(i + 1) % CHAINS]), "r"(c));
            if (OP == 2) r[i] = __vminu2(r[i], r[(i + 1) % CHAINS]) + 0u;
            if (OP == 3) r[i] = r[i] + r[(i + 1) % CHAINS] + c;
            if (OP == 4) r[i] = __popc(r[i] ^ r[(i + 1) % CHAINS]);
            if (OP == 5) {
                __half2 x = *(__half2*)&r[i], y = *(__half2*)&r[(i + 1) % CHAINS];
                __half2 z = __hmin2(x, y);
                r[i] = *(unsigned*)&z + 0u;
            }
            if (OP == 6) {  // half the chains VIMNMX3, half HMNMX2
                if (i & 1) {
                    __half2 x = *(__half2*)&r[i], y = *(__half2*)&r[(i + 2) % CHAINS];
                    __half2 z = __hmax2(x, y);
                    r[i] = *(unsigned*)&z;
                } else
                    r[i] = __vimin3_u16x2(r[i], r[(i + 2) % CHAINS], c);
            }
            if (OP == 7) {  // half VIMNMX3, half IMAD
                if (i & 1) r[i] = r[i] * r[(i + 2) % CHAINS] + c;
                else r[i] = __vimin3_u16x2(r[i], r[(i + 2) % CHAINS], c);
            }
        }
    }
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) s ^= r[i];
    if (s == 0x12345678u) out[0] = s;
}

// mix: A of op X then B of op Y per iteration, independent chains
template <int X, int Y, int NX, int NY>
__global__ void probe_mix(unsigned* out, unsigned seed) {
    unsigned r[NX + NY];
#pragma unroll
    for (int i = 0; i < NX + NY; i++) r[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
    unsigned b = seed ^ 0x5bd1e995u, c = seed | 1u;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NX; i++) r[i] = op<X>(r[i], b, c);
#pragma unroll
        for (int i = 0; i < NY; i++) r[NX + i] = op<Y>(r[NX + i], b, c);
    }
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < NX + NY; i++) s ^= r[i];
    if (s == 0x12345678u) out[0] = s;
}

template <typename K>
static void run(const char* name, K kern, double instPerThread) {
    unsigned* d;
    cudaMalloc(&d, 4);
    int dev = 0, sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const int blocks = sms * 4, threads = 512;
    kern<<<blocks, threads>>>(d, 12345u);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < 5; i++) kern<<<blocks, threads>>>(d, 12345u + i);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double warpInst = 5.0 * blocks * (threads / 32) * instPerThread;
    const double clk = ms * 1e-3 * khz * 1e3;
    printf("%-28s %7.3f warp-inst/clk/SM  (%.3f ms, %s)\n", name, warpInst / clk / sms, ms, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d);
}

int main() {
    const double n = (double)ITERS * CHAINS;
    run("VIMNMX3.U16x2 min", probe<0>, n);
    run("VIMNMX3.U16x2 max", probe<1>, n);
    run("HMNMX2", probe<2>, n);
    run("PRMT", probe<3>, n);
    run("LOP3", probe<4>, n);
    run("IMAD", probe<5>, n);
    run("SHF", probe<6>, n);
    run("VIMNMX.U16x2 (2-in)", probe<7>, n);
    run("VIADDMNMX.S16x2", probe<8>, n);
    run("IMAD.HI", probe<9>, n);
    run("POPC+IADD", probe<10>, n);
    run("DP4A", probe<11>, n);
    run("DP2A", probe<12>, n);
    run("IADD3", probe<13>, n);
    run("ISETP+SEL (2 inst)", probe<14>, 2 * n);
    run("SHL", probe<15>, n);
    run("x LOP3 xor3", probe_x<0>, n);
    run("x LOP3 maj", probe_x<1>, n);
    run("x VIMNMX.U16x2 2-in", probe_x<2>, n);
    run("x IADD3", probe_x<3>, n);
    run("x XOR+POPC (2 inst)", probe_x<4>, 2 * n);
    run("x HMNMX2", probe_x<5>, n);
    run("x VIMNMX3 : HMNMX2 1:1", probe_x<6>, n);
    run("x VIMNMX3 : IMAD 1:1", probe_x<7>, n);
    run("mix DP4A 4 : PRMT 4", probe_mix<11, 3, 4, 4>, (double)ITERS * 8);
    run("mix DP4A 4 : IMAD 4", probe_mix<11, 5, 4, 4>, (double)ITERS * 8);
    run("mix VIMNMX3 4 : HMNMX2 4", probe_mix<0, 2, 4, 4>, (double)ITERS * 8);
    run("mix VIMNMX3 4 : IMAD 4", probe_mix<0, 5, 4, 4>, (double)ITERS * 8);
    run("mix PRMT 4 : IMAD 4", probe_mix<3, 5, 4, 4>, (double)ITERS * 8);
    run("mix VIMNMX3 4 : PRMT 4", probe_mix<0, 3, 4, 4>, (double)ITERS * 8);
    run("mix LOP3 4 : HMNMX2 4", probe_mix<4, 2, 4, 4>, (double)ITERS * 8);
    return 0;
}
