// api.cu -- context lifetime, error reporting and measurement helpers of libzkb.so.
#include "common.cuh"
#include <algorithm>
#include <atomic>
#include <string.h>
#include <thread>

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
int upload_rows(zkb_ctx *ctx, void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t rows, cudaStream_t st)
{
    if (!width || !rows) return 0;
    const size_t total = width * rows;
    // Pageable memory reaches the GPU through the driver's own bounce buffer with ONE host thread copying (~11 GB/s
    // measured here: 11 ms for the 124 MB witness of the 2^22 circuit).  From 16 MB on the rows are cut into 2 MB pieces
    // that several host threads copy into a pinned area, each piece followed by its own asynchronous transfer on `st`.
    bool staged = total >= ((size_t)16 << 20) && total <= ((size_t)1 << 30);
    if (staged) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, src) != cudaSuccess || at.type != cudaMemoryTypeUnregistered) staged = false; // pinned / device
        cudaGetLastError();
    }
    const char *knob = getenv("ZKB_STAGED_UPLOAD");
    if (knob && atoi(knob) == 0) staged = false;
    if (staged && ctx->stage_cap < total) {
        if (ctx->stage) cudaFreeHost(ctx->stage);
        ctx->stage = nullptr;
        ctx->stage_cap = 0;
        if (cudaMallocHost(&ctx->stage, total) == cudaSuccess) {
            ctx->stage_cap = total;
        } else {
            cudaGetLastError();
            staged = false; // no pinned memory to be had: the plain copy still works
        }
    }
    if (!staged) {
        cudaError_t e = cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyDefault, st);
        if (e != cudaSuccess) {
            ctx->err = std::string("cudaMemcpy2DAsync: ") + cudaGetErrorString(e);
            return ZKB_ERR_CUDA;
        }
        return 0;
    }
    const size_t piece = (size_t)2 << 20, per_row = (width + piece - 1) / piece, npieces = per_row * rows;
    std::atomic<size_t> next{0};
    std::atomic<int> bad{0};
    auto work = [&]() {
        cudaSetDevice(ctx->dev);
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= npieces || bad.load()) return;
            const size_t row = i / per_row, off = (i % per_row) * piece, len = std::min(piece, width - off);
            char *mid = (char *)ctx->stage + row * width + off;
            memcpy(mid, (const char *)src + row * spitch + off, len);
            if (cudaMemcpyAsync((char *)dst + row * dpitch + off, mid, len, cudaMemcpyHostToDevice, st) != cudaSuccess) bad.store(1);
        }
    };
    unsigned nt = std::thread::hardware_concurrency() / 2;
    nt = nt < 1 ? 1 : nt > 8 ? 8 : nt;
    if ((size_t)nt > npieces) nt = (unsigned)npieces;
    std::vector<std::thread> helpers;
    try {
        for (unsigned t = 1; t < nt; t++) helpers.emplace_back(work);
    } catch (const std::exception &) { // no more threads to be had: the ones that started and this one share the pieces
    }
    work();
    for (auto &h : helpers) h.join();
    if (bad.load()) {
        ctx->err = std::string("upload_rows: ") + cudaGetErrorString(cudaGetLastError());
        return ZKB_ERR_CUDA;
    }
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
// Integer multiply-add peak, measured with the instruction mix the whole library is made of: chains of Fq
// Montgomery products (fp.cuh), each 128 IMAD.WIDE.U32(.X) + 8 IMAD = 136 multiply-adds of 32 x 32 (+ 64) bits.
// Two independent chains per thread, 16 warps per SM sub-partition; every product depends on the previous one
// of its chain, so ptxas can neither hoist nor strength-reduce anything (an earlier probe made of
// mad.wide.u32 with loop-invariant operands was rewritten by ptxas into IADD3 chains and measured the ALU
// pipe instead).  On B200 a warp-wide IMAD.WIDE occupies the fma-heavy pipe of its sub-partition for 4 cycles
// (scripts/probe/pipe_probe.cu; profiles/r01_pipe_probe.txt), i.e. 32 wide multiply-adds per clock and SM:
// 148 x 32 x 1.965 GHz = 9.31e12/s nominal; this kernel sustains ~97 % of that.
__device__ __forceinline__ uint32_t mix32(uint32_t x)
{
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
constexpr int IMAD_PER_MODMUL = 136;
__global__ void __launch_bounds__(128) k_imad_peak(uint64_t *out, int iters, uint32_t seed)
{
    Fq a[2], b[2];
#pragma unroll
    for (int c = 0; c < 2; c++)
#pragma unroll
        for (int j = 0; j < 8; j++) {
            a[c].v[j] = mix32(seed + threadIdx.x * 16 + j + 97 * c + blockIdx.x) >> (j == 7 ? 3 : 0);
            b[c].v[j] = mix32(a[c].v[j] + 1) >> (j == 7 ? 3 : 0);
        }
    for (int i = 0; i < iters; i++) {
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
    if (s == 0x1234567u) out[0] = s; // never true in practice; keeps the chains alive
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
    // the main lane (witness map -> H + L sum) is the critical path of a proof; the witness-only sums of the side lanes
    // fill the SMs it leaves free.  ZKB_LANE_PRIO=0 gives all lanes one priority (measurement knob).
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi); // lo = numerically greatest = least urgent
    const char *pe = getenv("ZKB_LANE_PRIO");
    if (pe && atoi(pe) == 0) prio_hi = prio_lo;
    if (cudaStreamCreateWithPriority(&ctx->main_stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
        cudaStreamCreateWithPriority(&ctx->side_stream[0], cudaStreamNonBlocking, prio_lo) != cudaSuccess ||
        cudaStreamCreateWithPriority(&ctx->side_stream[1], cudaStreamNonBlocking, prio_lo) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_in[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_in[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_join[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_join[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess ||
        cudaEventCreate(&ctx->ev2) != cudaSuccess || cudaEventCreate(&ctx->ev3) != cudaSuccess) {
        delete ctx;
        return fail(nullptr, ZKB_ERR_CUDA, "zkb_init: stream/event creation failed");
    }
    ctx->stream = ctx->main_stream;
    *out = ctx;
    return ZKB_OK;
}

void zkb_shutdown(zkb_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->dev);
    cudaDeviceSynchronize();
    for (auto &s : ctx->scratch) s.release();
    for (int k = 0; k < 2; k++) {
        ctx->in_w[k].release();
        ctx->in_rs[k].release();
        cudaEventDestroy(ctx->ev_in[k]);
    }
    cudaStreamDestroy(ctx->copy_stream);
    if (ctx->stage) cudaFreeHost(ctx->stage);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    if (ctx->prof_pairs_host) {
        cudaFreeHost(ctx->prof_pairs_host);
        for (int g = 0; g < 2; g++)
            for (int k = 0; k < 2; k++) cudaEventDestroy(ctx->pe[g][k]);
    }
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev2);
    cudaEventDestroy(ctx->ev3);
    cudaEventDestroy(ctx->ev1);
    cudaEventDestroy(ctx->ev_fork);
    cudaEventDestroy(ctx->ev_join[0]);
    cudaEventDestroy(ctx->ev_join[1]);
    cudaStreamDestroy(ctx->main_stream);
    cudaStreamDestroy(ctx->side_stream[0]);
    cudaStreamDestroy(ctx->side_stream[1]);
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
    if (iters <= 0) iters = 512;
    if (cudaSetDevice(ctx->dev) != cudaSuccess) return -1.0;
    if (ctx->scratch[SB_MISC].ensure(64) != cudaSuccess) return -1.0;
    const int blocks = ctx->sm_count * 16, threads = 128;
    double best = 0;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(ctx->ev0, ctx->stream);
        k_imad_peak<<<blocks, threads, 0, ctx->stream>>>(ctx->scratch[SB_MISC].as<uint64_t>(), iters, 12345u + rep);
        cudaEventRecord(ctx->ev1, ctx->stream);
        if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) return -1.0;
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
        const double macs = (double)blocks * threads * (double)iters * 4.0 * IMAD_PER_MODMUL;
        const double rate = macs / (ms * 1e-3);
        if (rep > 0 && rate > best) best = rate;
    }
    return best;
}

} // extern "C"
