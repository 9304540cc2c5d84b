// glv_host.cpp -- zklaim_b200/csrc/glv.cuh compiled by the host compiler (tests/test_glv_host.py)
#include "../../zklaim_b200/csrc/glv.cuh"

extern "C" void hostglv_decompose(const uint32_t *k, unsigned long n, uint32_t *k1, uint32_t *k2abs)
{
    for (unsigned long i = 0; i < n; i++) glv::decompose(k + 8 * i, k1 + 5 * i, k2abs + 5 * i);
}

extern "C" void hostglv_beta_mont(uint32_t *out)
{
    const uint32_t b[8] = ZKB_GLV_BETA_MONT;
    for (int i = 0; i < 8; i++) out[i] = b[i];
}
