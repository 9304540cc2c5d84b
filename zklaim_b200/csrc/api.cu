// api.cu -- context lifetime, error reporting and measurement helpers of libzkb.so.
#include "common.cuh"

namespace zkb {
thread_local std::string g_init_err;

int ensure_pinned(zkb_ctx *ctx, size_t bytes)
{
    if (bytes <= ctx->pinned_cap) return 0;
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    ctx->pinned = nullptr;
    ctx->pinned_cap = 0;
    size_t want = bytes < 4096 ? 4096 : bytes;
    cudaError_t e = cudaMallocHost(&ctx->pinned, want);
    if (e != cudaSuccess) {
        ctx->err = std::string("cudaMallocHost: ") + cudaGetErrorString(e);
        return ZKB_ERR_CUDA;
    }
    ctx->pinned_cap = want;
    return 0;
}
void prof_collect(zkb_ctx *ctx)
{
    if (!ctx->prof) return;
    for (int g = 0; g < 2; g++) {
        if (!ctx->prof_pending[g]) continue;
        float ms = 0;
        if (cudaEventElapsedTime(&ms, ctx->pe[g][0], ctx->pe[g][1]) == cudaSuccess) {
            ctx->prof_ms[g] += ms;
            ctx->prof_launches[g]++;
            ctx->prof_pairs[g] += ctx->prof_pairs_host[g];
        }
        ctx->prof_pending[g] = false;
    }
}
} // namespace zkb

using namespace zkb;

namespace {
// independent 64-bit multiply-add chains: 8 accumulators per thread, all lanes busy
__global__ void k_imad_peak(uint64_t *out, int iters, uint32_t seed)
{
    uint32_t a = seed + threadIdx.x, b = seed * 2654435761u + blockIdx.x;
    uint64_t acc[8];
#pragma unroll
    for (int j = 0; j < 8; j++) acc[j] = (uint64_t)(j + 1) * a;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++)
            asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[j]) : "r"(a + j), "r"(b));
#pragma unroll
        for (int j = 0; j < 8; j++)
            asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[j]) : "r"(b + j), "r"(a));
    }
    uint64_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s ^= acc[j];
    if (s == 0x1234567ull) out[0] = s; // never true in practice; keeps the chains alive
}
} // namespace

extern "C" {

const char *zkb_version(void) { return "zkb 0.1 (sm_100a)"; }

int zkb_init(int device_id, zkb_ctx **out)
{
    if (!out) return fail(nullptr, ZKB_ERR_ARG, "zkb_init: out is NULL");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, ZKB_ERR_CUDA,
                    std::string("zkb_init: no CUDA device (") + cudaGetErrorString(e) +
                        "); this library has no CPU fallback");
    if (device_id < 0 || device_id >= count) return fail(nullptr, ZKB_ERR_ARG, "zkb_init: bad device id");
    ZKB_CUDA(nullptr, cudaSetDevice(device_id));
    zkb_ctx *ctx = new zkb_ctx();
    ctx->dev = device_id;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device_id) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess) {
        delete ctx;
        return fail(nullptr, ZKB_ERR_CUDA, "zkb_init: stream/event creation failed");
    }
    *out = ctx;
    return ZKB_OK;
}

void zkb_shutdown(zkb_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->dev);
    cudaStreamSynchronize(ctx->stream);
    for (auto &s : ctx->scratch) s.release();
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    if (ctx->prof_pairs_host) {
        cudaFreeHost(ctx->prof_pairs_host);
        for (int g = 0; g < 2; g++)
            for (int k = 0; k < 2; k++) cudaEventDestroy(ctx->pe[g][k]);
    }
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *zkb_last_error(const zkb_ctx *ctx) { return ctx ? ctx->err.c_str() : g_init_err.c_str(); }

int zkb_profile(zkb_ctx *ctx, int enable)
{
    if (!ctx) return ZKB_ERR_ARG;
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    if (enable && !ctx->prof_pairs_host) {
        for (int g = 0; g < 2; g++)
            for (int k = 0; k < 2; k++) ZKB_CUDA(ctx, cudaEventCreate(&ctx->pe[g][k]));
        ZKB_CUDA(ctx, cudaMallocHost((void **)&ctx->prof_pairs_host, 16));
    }
    ctx->prof = enable != 0;
    for (int g = 0; g < 2; g++) {
        ctx->prof_ms[g] = ctx->prof_pairs[g] = 0;
        ctx->prof_launches[g] = 0;
        ctx->prof_pending[g] = false;
    }
    return ZKB_OK;
}

int zkb_profile_read(const zkb_ctx *ctx, double out[6])
{
    if (!ctx || !out) return ZKB_ERR_ARG;
    for (int g = 0; g < 2; g++) {
        out[3 * g] = ctx->prof_ms[g];
        out[3 * g + 1] = (double)ctx->prof_launches[g];
        out[3 * g + 2] = ctx->prof_pairs[g];
    }
    return ZKB_OK;
}

double zkb_last_device_ms(const zkb_ctx *ctx) { return ctx ? ctx->last_ms : -1.0; }
long zkb_last_kernel_launches(const zkb_ctx *ctx) { return ctx ? ctx->launches : -1; }

double zkb_measure_imad_peak(zkb_ctx *ctx, int iters)
{
    if (!ctx) return -1.0;
    if (iters <= 0) iters = 4096;
    if (cudaSetDevice(ctx->dev) != cudaSuccess) return -1.0;
    if (ctx->scratch[SB_MISC].ensure(64) != cudaSuccess) return -1.0;
    const int blocks = ctx->sm_count * 8, threads = 256;
    double best = 0;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(ctx->ev0, ctx->stream);
        k_imad_peak<<<blocks, threads, 0, ctx->stream>>>(ctx->scratch[SB_MISC].as<uint64_t>(), iters, 12345u + rep);
        cudaEventRecord(ctx->ev1, ctx->stream);
        if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) return -1.0;
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
        const double macs = (double)blocks * threads * (double)iters * 16.0;
        const double rate = macs / (ms * 1e-3);
        if (rep > 0 && rate > best) best = rate;
    }
    return best;
}

} // extern "C"
