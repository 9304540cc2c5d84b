// pipe_probe.cu -- integer-pipe throughput probe for sm_100a (development aid, not product code).
// Measures warp-instructions per cycle per SM sub-partition for the instruction forms a 256-bit Montgomery
// product can be built from, so that the IMAD roofline of DESIGN.md uses the right denominator.
#include <cstdio>
#include <cstdint>
#include <cstring>
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

// ---- the other roof: FP64.  B200 issues DFMA at twice the IMAD.WIDE rate (64 vs 32 per clock per SM), and a pair of
// fma.rz.f64 yields the exact 104-bit product of two 52-bit integers held in doubles (Emmart / Luitjens / Weems,
// "Faster modular exponentiation using double precision floating point arithmetic on the GPU"):
//     hi = fma_rz(a, b, 2^104)                -> 2^104 + 2^52 * floor(a b / 2^52)           (ulp of [2^104, 2^105) is 2^52)
//     lo = fma_rz(a, b, (2^104 + 2^52) - hi)  -> 2^52 + (a b mod 2^52)                      (exact)
// The mantissa bits of hi / lo ARE the two 52-bit halves, so column sums are int64 additions of the raw bit patterns
// (the exponent patterns are subtracted per column at the end).
// form 6: fma.rz.f64 alone, 16 independent chains
__global__ void k_dfma(uint64_t *out, int iters, uint32_t seed)
{
    double a[16], acc[16];
#pragma unroll
    for (int j = 0; j < 16; j++) { a[j] = 1.0 + 1e-9 * (mix(seed + threadIdx.x * 16 + j) & 1023); acc[j] = 0.5 + 1e-3 * (mix(j + blockIdx.x) & 255); }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) asm volatile("fma.rz.f64 %0, %1, %0, %2;" : "+d"(acc[j]) : "d"(a[j]), "d"(a[(j + 1) & 15]));
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) s += acc[j];
    if (s == 0.123456789) out[0] = 1;
}

constexpr double D_2_52 = 4503599627370496.0;             // 2^52
constexpr double D_2_104 = 20282409603651670423947251286016.0; // 2^104
constexpr unsigned long long B_2_52 = 0x4330000000000000ull, B_2_104 = 0x4670000000000000ull, MASK52 = 0x000fffffffffffffull;

__device__ __forceinline__ void mul_hilo(double a, double b, long long &hi, long long &lo)
{
    double h, l;
    asm("fma.rz.f64 %0, %1, %2, %3;" : "=d"(h) : "d"(a), "d"(b), "d"(D_2_104));
    const double sub = (D_2_104 + D_2_52) - h;
    asm("fma.rz.f64 %0, %1, %2, %3;" : "=d"(l) : "d"(a), "d"(b), "d"(sub));
    hi = __double_as_longlong(h);
    lo = __double_as_longlong(l);
}
__device__ __forceinline__ double int52_to_double(unsigned long long x) { return __longlong_as_double((long long)(x | B_2_52)) - D_2_52; }

// Montgomery product on 5 x 52-bit limbs, R = 2^260: r = a b R^-1 mod p, inputs and output < 2 p (no final subtraction is
// needed: 4 p < 2^256 << R).  p, pinv = -p^-1 mod 2^52 as doubles.  55 limb products = 110 DFMA.
struct F52 {
    double v[5];
};
__device__ __forceinline__ F52 mont52(const F52 &a, const F52 &b, const double *p, double pinv)
{
    unsigned long long t[11];
#pragma unroll
    for (int k = 0; k < 11; k++) t[k] = 0;
#pragma unroll
    for (int i = 0; i < 5; i++)
#pragma unroll
        for (int j = 0; j < 5; j++) {
            long long hi, lo;
            mul_hilo(a.v[i], b.v[j], hi, lo);
            t[i + j] += (unsigned long long)lo;
            t[i + j + 1] += (unsigned long long)hi;
        }
    // remove the exponent patterns: column k received n lo-terms and m hi-terms
#pragma unroll
    for (int k = 0; k < 10; k++) {
        const int nlo = k <= 4 ? k + 1 : (k <= 8 ? 9 - k : 0), nhi = k == 0 ? 0 : (k <= 5 ? k : 10 - k);
        t[k] -= (unsigned long long)nlo * B_2_52 + (unsigned long long)nhi * B_2_104;
    }
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const double xd = int52_to_double(t[i] & MASK52);
        long long qh, ql;
        mul_hilo(xd, pinv, qh, ql);
        const double qd = int52_to_double((unsigned long long)ql & MASK52);
#pragma unroll
        for (int j = 0; j < 5; j++) {
            long long hi, lo;
            mul_hilo(qd, p[j], hi, lo);
            t[i + j] += (unsigned long long)lo - B_2_52;
            t[i + j + 1] += (unsigned long long)hi - B_2_104;
        }
        t[i + 1] += t[i] >> 52; // t[i] is now a multiple of 2^52
    }
    F52 r;
#pragma unroll
    for (int j = 0; j < 5; j++) {
        if (j < 4) t[5 + j + 1] += t[5 + j] >> 52;
        r.v[j] = int52_to_double(j < 4 ? (t[5 + j] & MASK52) : t[5 + j]);
    }
    return r;
}

__constant__ double c_p52[5];
__constant__ double c_pinv52;

// form 7: the FP64 Montgomery product, 2 independent chains per thread; counted in products
__global__ void __launch_bounds__(128) k_modmul_dfma(uint64_t *out, int iters, uint32_t seed)
{
    F52 a[2], b[2];
    double p[5];
#pragma unroll
    for (int j = 0; j < 5; j++) p[j] = c_p52[j];
    const double pinv = c_pinv52;
#pragma unroll
    for (int c = 0; c < 2; c++)
#pragma unroll
        for (int j = 0; j < 5; j++) {
            const unsigned long long x = ((unsigned long long)mix(seed + threadIdx.x * 16 + j + 97 * c + blockIdx.x) << 20) ^ mix(j + c);
            a[c].v[j] = int52_to_double(x & (j == 4 ? 0x3fffffffffffull : MASK52));
            b[c].v[j] = int52_to_double((x * 0x9e3779b97f4a7c15ull >> 12) & (j == 4 ? 0x3fffffffffffull : MASK52));
        }
    for (int i = 0; i < iters / 8; i++) {
#pragma unroll
        for (int c = 0; c < 2; c++) {
            a[c] = mont52(a[c], b[c], p, pinv);
            b[c] = mont52(b[c], a[c], p, pinv);
        }
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < 2; c++)
#pragma unroll
        for (int j = 0; j < 5; j++) s += a[c].v[j] + b[c].v[j];
    if (s == 0.123456789) out[0] = 1;
}

// correctness of mont52: out[0..4] = limbs of mont52(a, b) for the limbs in[0..4], in[5..9] (checked on the host with
// 128-bit integer arithmetic against a b 2^-260 mod p)
__global__ void k_mont52_check(const double *in, double *outv)
{
    F52 a, b;
    double p[5];
    for (int j = 0; j < 5; j++) { a.v[j] = in[j]; b.v[j] = in[5 + j]; p[j] = c_p52[j]; }
    const F52 r = mont52(a, b, p, c_pinv52);
    for (int j = 0; j < 5; j++) outv[j] = r.v[j];
}

typedef void (*kern_t)(uint64_t *, int, uint32_t);

// q in 52-bit limbs, -q^-1 mod 2^52, and a check of mont52 against big-integer arithmetic
typedef unsigned __int128 u128;
static const uint64_t Q64[4] = {0x3c208c16d87cfd47ull, 0x97816a916871ca8dull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
struct Big { uint64_t w[10]; }; // little-endian, up to 640 bits
static void big_mul(const uint64_t *a, int na, const uint64_t *b, int nb, uint64_t *r)
{
    for (int i = 0; i < na + nb; i++) r[i] = 0;
    for (int i = 0; i < na; i++) {
        u128 c = 0;
        for (int j = 0; j < nb; j++) {
            c += (u128)a[i] * b[j] + r[i + j];
            r[i + j] = (uint64_t)c;
            c >>= 64;
        }
        r[i + nb] = (uint64_t)c;
    }
}
static bool geq(const uint64_t *a, const uint64_t *b, int n) { for (int i = n - 1; i >= 0; i--) if (a[i] != b[i]) return a[i] > b[i]; return true; }
static void sub_n(uint64_t *a, const uint64_t *b, int n) { uint64_t br = 0; for (int i = 0; i < n; i++) { u128 d = (u128)a[i] - b[i] - br; a[i] = (uint64_t)d; br = (uint64_t)(d >> 64) & 1; } }
// x mod q for x of n words (schoolbook shift-subtract; n <= 9)
static void mod_q(uint64_t *x, int n, uint64_t *r4)
{
    for (int bit = n * 64 - 255; bit >= 0; bit--) { // subtract q << bit while possible
        uint64_t sh[10] = {0};
        const int ws = bit / 64, bs = bit % 64;
        for (int i = 0; i < 4; i++) { sh[i + ws] |= Q64[i] << bs; if (bs && i + ws + 1 < 10) sh[i + ws + 1] |= Q64[i] >> (64 - bs); }
        if (geq(x, sh, n)) sub_n(x, sh, n);
    }
    for (int i = 0; i < 4; i++) r4[i] = x[i];
}
static void limbs52(const uint64_t *x4, double *l)
{
    for (int j = 0; j < 5; j++) {
        const int bit = 52 * j, w = bit / 64, o = bit % 64;
        uint64_t v = x4[w] >> o;
        if (o > 12 && w + 1 < 4) v |= x4[w + 1] << (64 - o);
        l[j] = (double)(v & 0xfffffffffffffull);
    }
}
static int setup_and_check_mont52()
{
    double p52[5];
    limbs52(Q64, p52);
    uint64_t inv = 1; // q^-1 mod 2^64 by Newton, then negate, mod 2^52
    for (int i = 0; i < 6; i++) inv *= 2 - Q64[0] * inv;
    const double pinv = (double)((0 - inv) & 0xfffffffffffffull);
    CHK(cudaMemcpyToSymbol(c_p52, p52, sizeof(p52)));
    CHK(cudaMemcpyToSymbol(c_pinv52, &pinv, sizeof(pinv)));
    double *d_in, *d_out;
    CHK(cudaMalloc(&d_in, 80));
    CHK(cudaMalloc(&d_out, 40));
    int bad = 0;
    uint64_t st = 88172645463325252ull;
    for (int trial = 0; trial < 64; trial++) {
        uint64_t a[4], b[4];
        for (int i = 0; i < 4; i++) { st ^= st << 13; st ^= st >> 7; st ^= st << 17; a[i] = st; st ^= st << 13; st ^= st >> 7; st ^= st << 17; b[i] = st; }
        a[3] &= 0x3fffffffffffffffull; b[3] &= 0x3fffffffffffffffull; // < 2^254 < 2 q
        double in[10], outv[5];
        limbs52(a, in);
        limbs52(b, in + 5);
        CHK(cudaMemcpy(d_in, in, 80, cudaMemcpyHostToDevice));
        k_mont52_check<<<1, 1>>>(d_in, d_out);
        CHK(cudaMemcpy(outv, d_out, 40, cudaMemcpyDeviceToHost));
        // r * 2^260 == a * b (mod q) ?
        uint64_t r5[5] = {0, 0, 0, 0, 0};
        for (int j = 0; j < 5; j++) { // r = sum limb_j 2^(52 j)
            const uint64_t v = (uint64_t)outv[j];
            const int bit = 52 * j, w = bit / 64, o = bit % 64;
            u128 add = (u128)v << o;
            u128 c = (u128)r5[w] + (uint64_t)add;
            r5[w] = (uint64_t)c;
            c = (c >> 64) + (uint64_t)(add >> 64);
            for (int k = w + 1; k < 5 && c; k++) { c += r5[k]; r5[k] = (uint64_t)c; c >>= 64; }
        }
        uint64_t sh[10] = {0}; // r << 260
        for (int i = 0; i < 5; i++) { sh[i + 4] |= r5[i] << 4; if (i + 5 < 10) sh[i + 5] |= r5[i] >> 60; }
        uint64_t lhs[4], rhs[4], prod[10] = {0};
        mod_q(sh, 10, lhs);
        big_mul(a, 4, b, 4, prod);
        mod_q(prod, 10, rhs);
        if (memcmp(lhs, rhs, 32)) bad++;
    }
    printf("mont52 (5 x 52-bit limbs on fma.rz.f64) against big-integer arithmetic: %d of 64 products wrong\n", bad);
    return bad ? 1 : 0;
}

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
        {"fma.rz.f64 (DFMA), 16 accumulators", k_dfma, 16},
        {"Fq Montgomery product on 5 x 52-bit limbs, fma.rz.f64 hi/lo pairs: counted in PRODUCTS", k_modmul_dfma, 4.0 / 8},
    };
    if (setup_and_check_mont52()) return 1;
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
