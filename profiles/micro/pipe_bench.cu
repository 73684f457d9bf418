// Issue-rate micro-benchmark for the integer instructions the MFCC kernel is made of.
// For each instruction (or mix) a kernel runs long dependency chains, 8 independent per
// thread; the result is warp-instructions per cycle per SM sub-partition (SMSP).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_bench pipe_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
#define CH 8

template <int OP>
__global__ void k(int *out, int a0, int b0)
{
    int v[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) v[i] = a0 + threadIdx.x + i;
    int b = b0;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
      for (int rep = 0; rep < 4; ++rep) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (OP == 0) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(b), "r"(it));
            if (OP == 1) asm volatile("mul.hi.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b));
            if (OP == 2) asm volatile("shr.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b + rep));
            if (OP == 3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[i]) : "r"(b), "r"(it));
            if (OP == 4) asm volatile("prmt.b32 %0, %0, %1, 0x7632;" : "+r"(v[i]) : "r"(b));
            if (OP == 5) asm volatile("add.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(v[(i + 3) % CH]));
            if (OP == 6) {   // 1:1 mix of IMAD and SHF, independent chains
                if (i & 1) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(b), "r"(it));
                else       asm volatile("shr.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b + rep));
            }
            if (OP == 7) {   // 64-bit multiply-add (IMAD.WIDE)
                unsigned long long t = (unsigned long long)(unsigned)v[i] * (unsigned)b + (unsigned)it;
                v[i] = (int)(t >> 7);
            }
            if (OP == 8) asm volatile("shf.l.wrap.b32 %0, %0, %1, 15;" : "+r"(v[i]) : "r"(b));
            if (OP == 9) {   // lea-like: (v << 15) + b
                asm volatile("{ .reg .s32 t; shl.b32 t, %0, 15; add.s32 %0, t, %1; }" : "+r"(v[i]) : "r"(b));
            }
            if (OP == 10) {  // 2:1:1 IMAD : SHF : IADD, the butterfly's mix
                if ((i & 3) < 2) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(b), "r"(it));
                else if ((i & 3) == 2) asm volatile("shr.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b + rep));
                else asm volatile("add.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(v[(i + 3) % CH]));
            }
            if (OP == 11) asm volatile("mad.lo.s32 %0, %0, 11585, %1;" : "+r"(v[i]) : "r"(it));       // IMAD R, R, imm, R
            if (OP == 12) asm volatile("mad.lo.s32 %0, %0, %1, 16382;" : "+r"(v[i]) : "r"(b));        // IMAD R, R, R, imm
            if (OP == 13) {  // 1:1 IMAD-imm : SHF
                if (i & 1) asm volatile("mad.lo.s32 %0, %0, 11585, %1;" : "+r"(v[i]) : "r"(it));
                else       asm volatile("shr.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b + rep));
            }
            if (OP == 14) {  // 2:1 IMAD-imm : SHF
                if ((i % 3) < 2) asm volatile("mad.lo.s32 %0, %0, 11585, %1;" : "+r"(v[i]) : "r"(it));
                else             asm volatile("shr.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b + rep));
            }
            if (OP == 15) asm volatile("shf.r.clamp.b32 %0, %0, %1, 1;" : "+r"(v[i]) : "r"(b));        // SHF with an immediate count
            if (OP == 16) {  // float pipe for comparison: FFMA 3-register form
                float f = __int_as_float(v[i]);
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(__int_as_float(b)), "f"(__int_as_float(it)));
                v[i] = __float_as_int(f);
            }
            if (OP == 17) {  // FFMA with an immediate multiplier
                float f = __int_as_float(v[i]);
                asm volatile("fma.rn.f32 %0, %0, 0f3F800100, %1;" : "+f"(f) : "f"(__int_as_float(it)));
                v[i] = __float_as_int(f);
            }
            if (OP == 18) {  // 1:1:1 IMAD : SHF : FFMA (three pipes?)
                if (i % 3 == 0) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(b), "r"(it));
                else if (i % 3 == 1) asm volatile("shr.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(b + rep));
                else {
                    float f = __int_as_float(v[i]);
                    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(__int_as_float(b)), "f"(__int_as_float(it)));
                    v[i] = __float_as_int(f);
                }
            }
        }
      }
    }
    int s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s ^= v[i];
    if (s == 0x12345678) out[threadIdx.x] = s;
}

template <int OP>
double run(const char *name, int sms, double clk_hz, double per_iter)
{
    int *d; cudaMalloc(&d, 4096);
    const int blocks = sms * 4, threads = 256;        // 8 warps per SMSP
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP><<<blocks, threads>>>(d, 1, 3);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<OP><<<blocks, threads>>>(d, 1, 3);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double winst = (double)blocks * (threads / 32) * ITERS * 4.0 * CH * per_iter;
    const double per_smsp_clk = winst / (ms * 1e-3 * clk_hz) / (sms * 4);
    printf("{\"op\": \"%s\", \"ms\": %.4f, \"warp_inst_per_clk_per_smsp\": %.3f, \"lane_inst_per_s\": %.4g}\n",
           name, ms, per_smsp_clk, winst * 32 / (ms * 1e-3));
    cudaFree(d);
    return per_smsp_clk;
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double clk = clk_khz * 1e3;
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_mhz\": %.0f}\n", p.name, p.multiProcessorCount, clk / 1e6);
    const int s = p.multiProcessorCount;
    run<0>("IMAD (mad.lo.s32)", s, clk, 1);
    run<1>("IMAD.HI (mul.hi.s32)", s, clk, 1);
    run<2>("SHF (shr.s32)", s, clk, 1);
    run<3>("LOP3", s, clk, 1);
    run<4>("PRMT", s, clk, 1);
    run<5>("IADD", s, clk, 1);
    run<6>("IMAD+SHF 1:1", s, clk, 1);
    run<7>("IMAD.WIDE.U32 (+shift)", s, clk, 2);
    run<8>("SHF.L funnel", s, clk, 1);
    run<9>("shl+add (LEA?)", s, clk, 1);
    run<10>("IMAD:SHF:IADD 2:1:1", s, clk, 1);
    run<11>("IMAD R,R,imm,R", s, clk, 1);
    run<12>("IMAD R,R,R,imm", s, clk, 1);
    run<13>("IMAD-imm+SHF 1:1", s, clk, 1);
    run<14>("IMAD-imm:SHF 2:1", s, clk, 1);
    run<15>("SHF imm count", s, clk, 1);
    run<16>("FFMA 3-reg", s, clk, 1);
    run<17>("FFMA imm", s, clk, 1);
    run<18>("IMAD:SHF:FFMA 1:1:1", s, clk, 1);
    return 0;
}
