// pipe_probe.cu -- integer-pipe throughput probe for sm_100a (development aid, not product code).
// Measures warp-instructions per cycle per SM sub-partition for the instruction forms a 256-bit Montgomery
// product can be built from, so that the IMAD roofline of DESIGN.md uses the right denominator.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../zklaim_b200/csrc/fp.cuh"

#define CHK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// form 0: mad.wide.u32 (IMAD.WIDE.U32): acc_j += a_j * lo(acc_j), 8 independent chains per thread
__global__ void k_wide(uint64_t *out, int iters, uint32_t seed)
{
    uint32_t a[8];
    uint64_t acc[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { a[j] = mix(seed + threadIdx.x * 8 + j) | 1u; acc[j] = mix(a[j] + blockIdx.x); }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int j = 0; j < 8; j++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[j]) : "r"(a[j]), "r"((uint32_t)acc[j]));
    }
    uint64_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s ^= acc[j];
    if (s == 0x1234567ull) out[0] = s;
}

// form 1: carry chains mad.lo.cc / madc.hi.cc (-> IMAD.WIDE.U32.X), the multiplier is the other chain's low word
__global__ void k_chain(uint64_t *out, int iters, uint32_t seed)
{
    uint32_t a[4], x[8], y[8];
#pragma unroll
    for (int j = 0; j < 4; j++) a[j] = mix(seed + threadIdx.x * 4 + j) | 1u;
#pragma unroll
    for (int j = 0; j < 8; j++) { x[j] = mix(a[0] + j + blockIdx.x); y[j] = mix(a[1] + j); }
    for (int i = 0; i < iters; i++) {
        asm volatile("mad.lo.cc.u32  %0, %8,  %12, %0;\n\t"
            "madc.hi.cc.u32 %1, %8,  %12, %1;\n\t"
            "madc.lo.cc.u32 %2, %9,  %12, %2;\n\t"
            "madc.hi.cc.u32 %3, %9,  %12, %3;\n\t"
            "madc.lo.cc.u32 %4, %10, %12, %4;\n\t"
            "madc.hi.cc.u32 %5, %10, %12, %5;\n\t"
            "madc.lo.cc.u32 %6, %11, %12, %6;\n\t"
            "madc.hi.u32    %7, %11, %12, %7;"
            : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(y[0]));
        asm volatile("mad.lo.cc.u32  %0, %8,  %12, %0;\n\t"
            "madc.hi.cc.u32 %1, %8,  %12, %1;\n\t"
            "madc.lo.cc.u32 %2, %9,  %12, %2;\n\t"
            "madc.hi.cc.u32 %3, %9,  %12, %3;\n\t"
            "madc.lo.cc.u32 %4, %10, %12, %4;\n\t"
            "madc.hi.cc.u32 %5, %10, %12, %5;\n\t"
            "madc.lo.cc.u32 %6, %11, %12, %6;\n\t"
            "madc.hi.u32    %7, %11, %12, %7;"
            : "+r"(y[0]), "+r"(y[1]), "+r"(y[2]), "+r"(y[3]), "+r"(y[4]), "+r"(y[5]), "+r"(y[6]), "+r"(y[7])
            : "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[0]), "r"(x[0]));
    }
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s ^= x[j] ^ y[j];
    if (s == 0x1234567u) out[0] = s;
}

// form 2: 32-bit mad.lo (IMAD): acc_j = a_j * acc_j + acc_j, 16 chains
__global__ void k_lo(uint64_t *out, int iters, uint32_t seed)
{
    uint32_t a[16], acc[16];
#pragma unroll
    for (int j = 0; j < 16; j++) { a[j] = mix(seed + threadIdx.x * 16 + j) | 1u; acc[j] = mix(a[j] + blockIdx.x); }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) asm volatile("mad.lo.u32 %0, %1, %0, %2;" : "+r"(acc[j]) : "r"(a[j]), "r"(a[(j + 1) & 15]));
    }
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) s ^= acc[j];
    if (s == 0x1234567u) out[0] = s;
}

// form 3: mad.hi.u32 (IMAD.HI.U32), 16 chains
__global__ void k_hi(uint64_t *out, int iters, uint32_t seed)
{
    uint32_t a[16], acc[16];
#pragma unroll
    for (int j = 0; j < 16; j++) { a[j] = mix(seed + threadIdx.x * 16 + j) | 0x80000001u; acc[j] = mix(a[j] + blockIdx.x); }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) asm volatile("mad.hi.u32 %0, %1, %0, %2;" : "+r"(acc[j]) : "r"(a[j]), "r"(a[(j + 1) & 15]));
    }
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) s ^= acc[j];
    if (s == 0x1234567u) out[0] = s;
}

// form 4: 8 x mul.wide.u32 into fresh registers (fma pipe) + one 16-word add.cc chain (alu pipe) per iteration
__global__ void k_wide_plus_add(uint64_t *out, int iters, uint32_t seed)
{
    uint32_t x[16], a[8];
#pragma unroll
    for (int j = 0; j < 16; j++) x[j] = mix(seed + threadIdx.x * 16 + j + blockIdx.x);
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = mix(x[j]) | 1u;
    for (int i = 0; i < iters; i++) {
        uint64_t p[8];
#pragma unroll
        for (int j = 0; j < 8; j++) asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p[j]) : "r"(a[j]), "r"(x[2 * j]));
        uint32_t lo[8], hi[8];
#pragma unroll
        for (int j = 0; j < 8; j++) { lo[j] = (uint32_t)p[j]; hi[j] = (uint32_t)(p[j] >> 32); }
        asm volatile("add.cc.u32 %0, %0, %16;\n\taddc.cc.u32 %1, %1, %17;\n\taddc.cc.u32 %2, %2, %18;\n\taddc.cc.u32 %3, %3, %19;\n\t"
                     "addc.cc.u32 %4, %4, %20;\n\taddc.cc.u32 %5, %5, %21;\n\taddc.cc.u32 %6, %6, %22;\n\taddc.cc.u32 %7, %7, %23;\n\t"
                     "addc.cc.u32 %8, %8, %24;\n\taddc.cc.u32 %9, %9, %25;\n\taddc.cc.u32 %10, %10, %26;\n\taddc.cc.u32 %11, %11, %27;\n\t"
                     "addc.cc.u32 %12, %12, %28;\n\taddc.cc.u32 %13, %13, %29;\n\taddc.cc.u32 %14, %14, %30;\n\taddc.u32 %15, %15, %31;"
                     : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]), "+r"(x[8]),
                       "+r"(x[9]), "+r"(x[10]), "+r"(x[11]), "+r"(x[12]), "+r"(x[13]), "+r"(x[14]), "+r"(x[15])
                     : "r"(lo[0]), "r"(hi[0]), "r"(lo[1]), "r"(hi[1]), "r"(lo[2]), "r"(hi[2]), "r"(lo[3]), "r"(hi[3]), "r"(lo[4]),
                       "r"(hi[4]), "r"(lo[5]), "r"(hi[5]), "r"(lo[6]), "r"(hi[6]), "r"(lo[7]), "r"(hi[7]));
    }
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) s ^= x[j];
    if (s == 0x1234567u) out[0] = s;
}

// form 5: the product itself: 2 independent chains of Fq Montgomery products per thread (fp.cuh), 136 IMAD-class each
__global__ void __launch_bounds__(128) k_modmul(uint64_t *out, int iters, uint32_t seed)
{
    zkb::Fq a[2], b[2];
#pragma unroll
    for (int c = 0; c < 2; c++)
#pragma unroll
        for (int j = 0; j < 8; j++) {
            a[c].v[j] = mix(seed + threadIdx.x * 16 + j + 97 * c + blockIdx.x) >> (j == 7 ? 3 : 0);
            b[c].v[j] = mix(a[c].v[j] + 1) >> (j == 7 ? 3 : 0);
        }
    for (int i = 0; i < iters / 8; i++) {
#pragma unroll
        for (int c = 0; c < 2; c++) {
            a[c] = a[c] * b[c];
            b[c] = b[c] * a[c];
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int c = 0; c < 2; c++)
#pragma unroll
        for (int j = 0; j < 8; j++) s ^= a[c].v[j] ^ b[c].v[j];
    if (s == 0x1234567u) out[0] = s;
}

typedef void (*kern_t)(uint64_t *, int, uint32_t);

int main(int argc, char **argv)
{
    int dev = 0;
    CHK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CHK(cudaGetDeviceProperties(&prop, dev));
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
    printf("device %s SMs %d clockRate %d kHz\n", prop.name, prop.multiProcessorCount, clk_khz);
    uint64_t *d;
    CHK(cudaMalloc(&d, 64));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    struct { const char *name; kern_t k; double mul_instr_per_iter; } forms[] = {
        {"mad.wide.u32 (IMAD.WIDE.U32), 8 accumulators", k_wide, 16},
        {"mad.lo.cc/madc.hi.cc chains (IMAD.WIDE.U32.X pairs), 2 chains -> 8 wide per iter", k_chain, 8},
        {"mad.lo.u32 (IMAD)", k_lo, 16},
        {"mad.hi.u32 (IMAD.HI.U32)", k_hi, 16},
        {"mul.wide.u32 x8 + 16-word add.cc chain", k_wide_plus_add, 8},
        {"Fq Montgomery product (fp.cuh), 136 IMAD-class instr each: counted in IMAD", k_modmul, 4.0 * 136 / 8},
    };
    const int iters = 4096;
    for (int wps = 1; wps <= 16; wps *= 2) { // warps per SM sub-partition
        const int threads = 128, blocks = prop.multiProcessorCount * wps;
        for (auto &f : forms) {
            double best = 1e30;
            for (int rep = 0; rep < 4; rep++) {
                cudaEventRecord(e0);
                f.k<<<blocks, threads>>>(d, iters, 12345u + rep);
                cudaEventRecord(e1);
                CHK(cudaEventSynchronize(e1));
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                if (rep && ms < best) best = ms;
            }
            const double warp_instr = (double)blocks * (threads / 32) * iters * f.mul_instr_per_iter;
            const double per_s = warp_instr * 32 / (best * 1e-3);
            const double cyc = best * 1e-3 * 1.965e9; // at max clock
            const double per_smsp_cycle = warp_instr / (prop.multiProcessorCount * 4.0) / cyc;
            printf("wps %2d  %-80s %8.3f ms  %7.2f T mul-instr(thread)/s  %.3f warp-instr/clk/SMSP (@1965 MHz) -> %.2f clk each\n",
                   wps, f.name, best, per_s / 1e12, per_smsp_cycle, 1.0 / per_smsp_cycle);
        }
    }
    return 0;
}
