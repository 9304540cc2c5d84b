// CPU build of zklaim_b200/csrc/fp.cuh's portable path: lets the (GPU-less) test-suite check
// the carry-chain Montgomery algorithm itself against big-integer arithmetic.
#include "../../zklaim_b200/csrc/fp.cuh"
#include <cstring>
using namespace zkb;
template <class F> static void run(int op, const uint32_t *a, const uint32_t *b, uint32_t *o, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        F x, y, z;
        memcpy(x.v, a + 8 * i, 32);
        if (b) memcpy(y.v, b + 8 * i, 32); else y = F::zero();
        switch (op) {
        case 0: z = x + y; break;
        case 1: z = x - y; break;
        case 2: z = x * y; break;
        case 3: z = x.sqr(); break;
        case 4: z = x.inverse(); break;
        case 5: z = x.neg(); break;
        case 6: z = x.from_mont(); break;
        case 7: z = x.to_mont(); break;
        default: z = F::zero();
        }
        memcpy(o + 8 * i, z.v, 32);
    }
}
extern "C" void hostfp_op(int field, int op, const uint32_t *a, const uint32_t *b, uint32_t *o, size_t n)
{
    if (field == 0) run<Fr>(op, a, b, o, n); else run<Fq>(op, a, b, o, n);
}
extern "C" void hostfq2_op(int op, const uint32_t *a, const uint32_t *b, uint32_t *o, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        Fq2 x, y, z;
        memcpy(&x, a + 16 * i, 64);
        if (b) memcpy(&y, b + 16 * i, 64); else y = Fq2::zero();
        switch (op) {
        case 0: z = x + y; break;
        case 1: z = x - y; break;
        case 2: z = x * y; break;
        case 3: z = x.sqr(); break;
        case 4: z = x.inverse(); break;
        case 5: z = x.neg(); break;
        default: z = Fq2::zero();
        }
        memcpy(o + 16 * i, &z, 64);
    }
}
