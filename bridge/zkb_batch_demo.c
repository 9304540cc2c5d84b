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
#include <execinfo.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>

void *zkbr_new(size_t payloads, uint64_t seed);
void zkbr_free(void *h);
int zkbr_roundtrip(void *h, size_t count, uint64_t seed0, double *seconds);
int zkb_libsnark_devices(void);
void zkb_bridge_reset(void);

/* a crash in a caller with no debugger at hand: leave the frames on stderr */
static void on_fault(int sig)
{
    void *frames[48];
    const int n = backtrace(frames, 48);
    dprintf(2, "zkb_batch_demo: signal %d\n", sig);
    backtrace_symbols_fd(frames, n, 2);
    _exit(128 + sig);
}

int main(int argc, char **argv)
{
    const size_t count = argc > 1 ? (size_t)atol(argv[1]) : 64;
    const size_t payloads = argc > 2 ? (size_t)atol(argv[2]) : 1;
    signal(SIGSEGV, on_fault);
    signal(SIGBUS, on_fault);
    signal(SIGABRT, on_fault);
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
    fflush(stdout);
    zkbr_free(h);
    zkb_bridge_reset(); /* device keys, contexts: released here, not by static destructors after the CUDA runtime is gone */
    return ok == (int)count ? 0 : 1;
}
