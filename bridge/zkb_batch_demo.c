/*
 * zkb_batch_demo.c -- a C caller shaped like zklaim/main_benchmark.c:174-198 (credential after credential through the
 * prover), with no Python and no torch: `count` credentials go through zkb_libsnark_prove_batch, which splits them over
 * every device of the bridge's session (ZKB_DEVICES="0,1,..." / ZKB_DEVICE=n / all visible GPUs: one host thread and one
 * proof stream per GPU, zkb_pool_prove_batch), and every proof is checked by the reference's own libsnark_verify
 * (zklaim/libsnark_wrapper.cpp:252-276).  Exit code 0 = all accepted.
 *
 *   usage: zkb_batch_demo [count = 64] [payloads = 1]
 * Links against bridge/_build/libzkb_zklaim.so only (plain C ABI: the zkbr_* helpers build the credentials).
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

void *zkbr_new(size_t payloads, uint64_t seed);
void zkbr_free(void *h);
int zkbr_roundtrip(void *h, size_t count, uint64_t seed0, double *seconds);
int zkb_libsnark_devices(void);

int main(int argc, char **argv)
{
    const size_t count = argc > 1 ? (size_t)atol(argv[1]) : 64;
    const size_t payloads = argc > 2 ? (size_t)atol(argv[2]) : 1;
    void *h = zkbr_new(payloads, 5);
    if (!h) {
        fprintf(stderr, "setup failed (no GPU?)\n");
        return 2;
    }
    const int devices = zkb_libsnark_devices();
    double sec[2] = {0, 0};
    int ok = zkbr_roundtrip(h, count, 31000, sec); /* warm: key tuning, scratch growth */
    if (ok == (int)count) ok = zkbr_roundtrip(h, count, 32000, sec);
    printf("{\"devices\": %d, \"credentials\": %zu, \"accepted_by_libsnark_verify\": %d, \"prove_batch_s\": %.4f, \"proofs_per_s\": %.1f}\n",
           devices, count, ok, sec[0], ok > 0 ? count / sec[0] : 0.0);
    zkbr_free(h);
    return ok == (int)count ? 0 : 1;
}
