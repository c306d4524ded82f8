// int_probe.cu — integer-pipe throughput micro-benchmarks for the NTT roofline (SURVEY §8(d) asks the
// builder to measure the integer peak on the box).  Each kernel runs ILP independent dependency chains
// of one instruction flavour per thread; the host prints warp-level G-instructions/s per flavour.
//   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/int_probe tools/int_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

typedef unsigned long long u64;
typedef unsigned int u32;
constexpr int ILP = 8;

enum { K_MADLO, K_MADWIDE, K_MADHI, K_MULHI64, K_MULLO64, K_SHOUP, K_SHOUP_APPROX, K_IADD3, K_MULWIDE, K_MADWIDE_IND, K_MULHI, K_BFLY, K_MULWIDE_BOTH, K_MADWIDE_BOTH, K_BFLY_V2, K_BFLY_V3, K_COUNT };
static const char* names[] = {"mad.lo.u32", "mad.wide.u32", "mad.hi.u32", "__umul64hi", "mul.lo.u64 (+add)",
                              "shoup_lazy (mulhi64 + 2 mullo64)", "shoup_approx (3-product hi)", "add.u32 x3 (alu)",
                              "mul.wide.u32 (no addend)", "mad.wide.u32 (addend not from the multiplier chain)", "mul.hi.u32 (no addend)",
                              "NTT butterfly multiplier as the kernels write it (2 mul.hi + 3 wide + 4 mad.lo)",
                              "mul.wide.u32, both halves of the product consumed", "mad.wide.u32 (64-bit addend), both halves consumed",
                              "butterfly multiplier v2: 5 mul.wide (no addend) + 64-bit adds on the ALU + 4 mad.lo",
                              "butterfly multiplier v3: 2 mul.hi + 3 mul.wide (no addend) + ALU adds + 4 mad.lo"};

template <int K>
__global__ void __launch_bounds__(256) probe(u64* out, int iters, u32 seed) {
    u32 a[ILP], b = seed | 1u, c = seed * 3u + threadIdx.x;
    u64 x[ILP], w = 0x9E3779B97F4A7C15ull * (seed | 1u), wp = w ^ 0x5555555555555555ull, nq = 0ull - (w >> 4);
#pragma unroll
    for (int k = 0; k < ILP; k++) { a[k] = threadIdx.x * 2654435761u + k + blockIdx.x; x[k] = ((u64)a[k] << 32) | (a[k] * 77u + k); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < ILP; k++) {
            if (K == K_MADLO) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(b), "r"(c));
            if (K == K_MADWIDE) x[k] = (u64)(u32)(x[k] >> 32) * (u64)b + x[k];
            if (K == K_MADHI) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(b), "r"(c));
            if (K == K_MULHI64) x[k] = __umul64hi(x[k], w) + 1;
            if (K == K_MULLO64) x[k] = x[k] * w + wp;
            if (K == K_SHOUP) { u64 qh = __umul64hi(x[k], wp); x[k] = x[k] * w + qh * nq; }
            if (K == K_SHOUP_APPROX) {
                u32 y0 = (u32)x[k], y1 = (u32)(x[k] >> 32), p0 = (u32)wp, p1 = (u32)(wp >> 32);
                u64 qh = (u64)y1 * p1 + __umulhi(y1, p0) + __umulhi(y0, p1);
                x[k] = x[k] * w + qh * nq;
            }
            if (K == K_MULWIDE) x[k] = (u64)(u32)x[k] * (u64)b;
            if (K == K_MADWIDE_IND) { u64 t = (u64)a[k] * (u64)b + x[k]; x[k] = t; a[k] = (u32)t ^ c; }
            if (K == K_MULHI) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[k]) : "r"(b));
            if (K == K_BFLY) {
                u32 x0 = (u32)x[k], x1 = (u32)(x[k] >> 32), w0 = (u32)w, w1 = (u32)(w >> 32);
                u32 p0 = (u32)wp, p1 = (u32)(wp >> 32), n0 = (u32)nq, n1 = (u32)(nq >> 32), t1, t2, lo, hi;
                asm("mul.hi.u32 %0, %1, %2;" : "=r"(t1) : "r"(x1), "r"(p0));
                asm("mul.hi.u32 %0, %1, %2;" : "=r"(t2) : "r"(x0), "r"(p1));
                u64 qh = (u64)t1 + (u64)t2, acc;
                asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(qh) : "r"(x1), "r"(p1));
                u32 h0 = (u32)qh, h1 = (u32)(qh >> 32);
                asm("mul.wide.u32 %0, %1, %2;" : "=l"(acc) : "r"(x0), "r"(w0));
                asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(h0), "r"(n0));
                asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(acc));
                asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(x0), "r"(w1));
                asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(x1), "r"(w0));
                asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(h0), "r"(n1));
                asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(h1), "r"(n0));
                asm("mov.b64 %0, {%1, %2};" : "=l"(x[k]) : "r"(lo), "r"(hi));
            }
            if (K == K_MULWIDE_BOTH) { u64 t; asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(t) : "r"((u32)x[k]), "r"(b)); x[k] = t ^ (t >> 32); }
            if (K == K_MADWIDE_BOTH) { u64 t = x[k]; asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(t) : "r"((u32)x[k]), "r"(b)); x[k] = t ^ (t >> 32); }
            if (K == K_BFLY_V2 || K == K_BFLY_V3) {
                u32 x0 = (u32)x[k], x1 = (u32)(x[k] >> 32), w0 = (u32)w, w1 = (u32)(w >> 32);
                u32 p0 = (u32)wp, p1 = (u32)(wp >> 32), n0 = (u32)nq, n1 = (u32)(nq >> 32), t1, t2, lo, hi;
                u64 m0, m1, m2, qh, a0, a1, acc;
                if (K == K_BFLY_V2) {
                    asm("mul.wide.u32 %0, %1, %2;" : "=l"(m0) : "r"(x1), "r"(p0));
                    asm("mul.wide.u32 %0, %1, %2;" : "=l"(m1) : "r"(x0), "r"(p1));
                    t1 = (u32)(m0 >> 32); t2 = (u32)(m1 >> 32);
                } else {
                    asm("mul.hi.u32 %0, %1, %2;" : "=r"(t1) : "r"(x1), "r"(p0));
                    asm("mul.hi.u32 %0, %1, %2;" : "=r"(t2) : "r"(x0), "r"(p1));
                }
                asm("mul.wide.u32 %0, %1, %2;" : "=l"(m2) : "r"(x1), "r"(p1));
                qh = m2 + (u64)t1 + (u64)t2;
                u32 h0 = (u32)qh, h1 = (u32)(qh >> 32);
                asm("mul.wide.u32 %0, %1, %2;" : "=l"(a0) : "r"(x0), "r"(w0));
                asm("mul.wide.u32 %0, %1, %2;" : "=l"(a1) : "r"(h0), "r"(n0));
                acc = a0 + a1;
                asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(acc));
                asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(x0), "r"(w1));
                asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(x1), "r"(w0));
                asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(h0), "r"(n1));
                asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(h1), "r"(n0));
                asm("mov.b64 %0, {%1, %2};" : "=l"(x[k]) : "r"(lo), "r"(hi));
            }
            if (K == K_IADD3) asm volatile("add.u32 %0, %0, %1; add.u32 %0, %0, %2; add.u32 %0, %0, %1;" : "+r"(a[k]) : "r"(b), "r"(c));
        }
    }
    u64 s = 0;
#pragma unroll
    for (int k = 0; k < ILP; k++) s ^= x[k] ^ a[k];
    if (s == 0x1234567ull) out[0] = s;
}

template <int K>
static void run(u64* d, int blocks, int iters) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<K><<<blocks, 256>>>(d, 64, 12345u);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        probe<K><<<blocks, 256>>>(d, iters, 12345u + r);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    double ops = (double)blocks * 256 * iters * ILP;
    printf("{\"probe\": \"%s\", \"ms\": %.4f, \"thread_ops_per_s\": %.4e, \"per_sm_per_clk_at_1965MHz\": %.2f}\n", names[K], best,
           ops / (best * 1e-3), ops / (best * 1e-3) / 148 / 1.965e9);
}

int main(int argc, char** argv) {
    int blocks = 148 * 8, iters = argc > 1 ? atoi(argv[1]) : 8192;
    u64* d; cudaMalloc(&d, 256);
    run<K_MADLO>(d, blocks, iters);
    run<K_MADWIDE>(d, blocks, iters);
    run<K_MADHI>(d, blocks, iters);
    run<K_MULHI64>(d, blocks, iters);
    run<K_MULLO64>(d, blocks, iters);
    run<K_SHOUP>(d, blocks, iters);
    run<K_SHOUP_APPROX>(d, blocks, iters);
    run<K_IADD3>(d, blocks, iters);
    run<K_MULWIDE>(d, blocks, iters);
    run<K_MADWIDE_IND>(d, blocks, iters);
    run<K_MULHI>(d, blocks, iters);
    run<K_BFLY>(d, blocks, iters);
    run<K_MULWIDE_BOTH>(d, blocks, iters);
    run<K_MADWIDE_BOTH>(d, blocks, iters);
    run<K_BFLY_V2>(d, blocks, iters);
    run<K_BFLY_V3>(d, blocks, iters);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
