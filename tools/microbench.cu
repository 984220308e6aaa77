// Pipe-rate microbenchmarks for the roofline denominators the driver does not measure
// (MEASURED_PEAKS.json has HBM and bf16 only): FFMA, packed FFMA2, MUFU ex2/rcp/lg2,
// an FFMA+MUFU mix, SHFL and LDS, in warp-instructions per clock per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define UNROLL 8

template <int OP>
__global__ void __launch_bounds__(1024) k(float* out, long long* cyc, float seed) {
    float a[UNROLL];
    float2 p[UNROLL];
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) { a[i] = seed + threadIdx.x * 1e-3f + i; p[i] = make_float2(a[i], a[i] + 1.f); }
    const float m = 0.999f + seed * 1e-6f, c = 1e-3f * seed;
    unsigned long long mm, cc;
    asm volatile("mov.b64 %0, {%1, %2};" : "=l"(mm) : "f"(m), "f"(m * 0.9999f));
    asm volatile("mov.b64 %0, {%1, %2};" : "=l"(cc) : "f"(c), "f"(c * 1.0001f));
    __shared__ float sh[1024];
    sh[threadIdx.x] = seed;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < UNROLL; ++i) {
            if (OP == 0) a[i] = fmaf(a[i], m, c);
            if (OP == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(*(const unsigned long long*)&p[(i + 1) % UNROLL]), "l"(*(const unsigned long long*)&p[(i + 2) % UNROLL]));
            if (OP == 2) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            if (OP == 3) { asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i])); a[i] += c; }   // FADD keeps ptxas from folding rcp(rcp(x))
            if (OP == 4) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            if (OP == 5) { a[i] = fmaf(a[i], m, c); a[i] = fmaf(a[i], m, c); a[i] = fmaf(a[i], m, c); a[i] = fmaf(a[i], m, c);
                           a[i] = fmaf(a[i], m, c); a[i] = fmaf(a[i], m, c); a[i] = fmaf(a[i], m, c);
                           asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])); }     // 7 FFMA : 1 MUFU
            if (OP == 6) a[i] = __shfl_xor_sync(0xffffffffu, a[i], 1);
            if (OP == 7) a[i] = sh[(threadIdx.x + (int)a[i]) & 1023];
            if (OP == 8) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(a[i]));
            if (OP == 9) { a[i] = fmaf(a[i], m, c); a[i] = fmaf(a[i], m, c); a[i] = fmaf(a[i], m, c);
                           asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])); }     // 3 FFMA : 1 MUFU
            if (OP == 10) a[i] = (a[i] > 0.5f) ? a[(i + 1) % UNROLL] : m;                 // FSETP+SEL (alu pipe)
            // packed-pair variants: which operand patterns reach one instruction per two clocks per scheduler
            if (OP == 11) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc));   // invariant multiplier and addend
            if (OP == 12) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc));   // accumulate form
            if (OP == 13) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm));
            if (OP == 14) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(*(unsigned long long*)&p[i]) : "l"(cc));
            if (OP == 15) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc)); a[i] = fmaf(a[i], m, c); }  // 1 FFMA2 + 1 FFMA
            if (OP == 16) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc));
                            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc));
                            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc));
                            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])); }     // 3 FFMA2 : 1 MUFU
            if (OP == 17) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(*(const unsigned long long*)&p[(i + 1) % UNROLL]), "l"(cc));  // two varying sources
            if (OP == 18) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc));
                            a[i] = (a[i] > 0.5f) ? a[(i + 1) % UNROLL] : m; }   // 1 FFMA2 + FSETP + SEL (alu pipe)
            // does a packed instruction hold the issue port for its two pipe cycles?  (2 FFMA2 + 1 LOP3: 4 clocks if not, 5 if so)
            if (OP == 19) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc));
                            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[(i + 4) % UNROLL]) : "l"(mm), "l"(cc));
                            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(*(unsigned*)&a[i]) : "r"(*(const unsigned*)&m), "r"(*(const unsigned*)&c)); }
            if (OP == 20) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(*(unsigned long long*)&p[i]) : "l"(mm), "l"(cc));
                            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(*(unsigned*)&a[i]) : "r"(*(const unsigned*)&m), "r"(*(const unsigned*)&c)); }
            if (OP == 21) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(*(unsigned*)&a[i]) : "r"(*(const unsigned*)&m), "r"(*(const unsigned*)&c));
            if (OP == 22) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(*(unsigned long long*)&p[i]) : "l"(*(const unsigned long long*)&p[(i + 1) % UNROLL]), "l"(*(const unsigned long long*)&p[(i + 3) % UNROLL]));  // accumulate, three varying
            if (OP == 23) { a[i] = fmaf(a[i], m, c); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(*(unsigned*)&p[i].x) : "r"(*(const unsigned*)&m), "r"(*(const unsigned*)&c)); }  // FFMA + LOP3
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) s += a[i] + p[i].x + p[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int per_iter, int threads) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    float* out; long long* cyc;
    cudaMalloc(&out, sizeof(float) * sms * 1024);
    cudaMalloc(&cyc, sizeof(long long) * sms);
    k<OP><<<sms, threads>>>(out, cyc, 1.0f);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<OP><<<sms, threads>>>(out, cyc, 1.0f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[256];
    cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < sms; ++i) mean += h[i]; mean /= sms;
    double winst = (double)ITERS * UNROLL * per_iter * (threads / 32);
    printf("%-28s threads/SM=%4d  warp-inst/clk/SM=%7.3f  lane-ops/clk/SM=%8.2f  cycles=%.0f  ms=%.3f  implied_MHz=%.0f\n",
           name, threads, winst / mean, winst * 32 / mean, mean, ms, mean / (ms * 1e3));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("device %s sm_%d%d SMs=%d clock=%d kHz\n", p.name, p.major, p.minor, p.multiProcessorCount, p.clockRate);
    for (int threads : {1024, 512, 256}) {
        run<0>("FFMA", 1, threads);
        run<1>("FFMA2 (fma.rn.f32x2)", 1, threads);
        run<2>("MUFU.EX2", 1, threads);
        run<3>("MUFU.RCP", 1, threads);
        run<4>("MUFU.LG2", 1, threads);
        run<8>("MUFU.TANH", 1, threads);
        run<5>("7 FFMA + 1 MUFU.EX2", 8, threads);
        run<9>("3 FFMA + 1 MUFU.EX2", 4, threads);
        run<6>("SHFL.BFLY", 1, threads);
        run<7>("LDS", 1, threads);
        run<10>("FSETP+SEL", 2, threads);
        run<11>("FFMA2 d=d*M+C (invariant M,C)", 1, threads);
        run<12>("FFMA2 d=M*C+d (accumulate)", 1, threads);
        run<13>("FMUL2 d=d*M", 1, threads);
        run<14>("FADD2 d=d+C", 1, threads);
        run<15>("1 FFMA2 + 1 FFMA", 2, threads);
        run<16>("3 FFMA2 + 1 MUFU.EX2", 4, threads);
        run<17>("FFMA2 d=d*p+C (two varying)", 1, threads);
        run<18>("1 FFMA2 + FSETP + SEL", 3, threads);
        run<19>("2 FFMA2 + 1 LOP3", 3, threads);
        run<20>("1 FFMA2 + 1 LOP3", 2, threads);
        run<21>("LOP3", 1, threads);
        run<22>("FFMA2 d=a*b+d (three varying)", 1, threads);
        run<23>("1 FFMA + 1 LOP3", 2, threads);
    }
    return 0;
}
