// stubs.cu -- entry points of include/zkb.h whose kernels are not built yet.
#include "common.cuh"
using namespace zkb;
extern "C" {
int zkb_key_upload(zkb_ctx *ctx, const zkb_key_desc *, zkb_key **) { return fail(ctx, ZKB_ERR_ARG, "not implemented"); }
void zkb_key_free(zkb_key *) {}
size_t zkb_key_device_bytes(const zkb_key *) { return 0; }
int zkb_key_domain(const zkb_key *, size_t *, int *) { return ZKB_ERR_ARG; }
int zkb_witness_map(zkb_ctx *ctx, const zkb_key *, const uint64_t *, uint64_t *) { return fail(ctx, ZKB_ERR_ARG, "not implemented"); }
int zkb_prove(zkb_ctx *ctx, const zkb_key *, const uint64_t *, const uint64_t *, const uint64_t *, zkb_proof *, uint64_t *) { return fail(ctx, ZKB_ERR_ARG, "not implemented"); }
int zkb_prove_batch(zkb_ctx *ctx, const zkb_key *, size_t, const uint64_t *, const uint64_t *, const uint64_t *, zkb_proof *) { return fail(ctx, ZKB_ERR_ARG, "not implemented"); }
}
