// common.cuh -- context, error handling, scratch memory and vectorised point I/O shared by
// the translation units of libzkb.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/zkb.h"
#include "ec.cuh"

namespace zkb {

// grow-only device buffer (scratch memory owned by a context)
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T *as() const { return (T *)p; }
};

} // namespace zkb

struct zkb_ctx {
    int dev = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;          // the stream work is enqueued on right now (main unless a LaneScope is active)
    cudaStream_t main_stream = nullptr;     // owns the timing events and every synchronisation with the host
    cudaStream_t side_stream[2] = {nullptr, nullptr}; // prover: witness-only MSMs overlap the witness map
    cudaEvent_t ev_fork = nullptr, ev_join[2] = {nullptr, nullptr};
    uint32_t func_attr_done = 0;            // cudaFuncSetAttribute calls already made on this device (bit per kernel)
    int lane_bias = 0;                      // scratch-slot offset of the active MSM lane
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev2 = nullptr, ev3 = nullptr; // second pair: the later phases of a split proof
    // pipelined batches (zkb_prove_batch_submit / _collect): inputs of the next batch travel on copy_stream into one of
    // two staging slots while the current batch computes
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_in[2] = {nullptr, nullptr};
    zkb::DevBuf in_w[2], in_rs[2];
    const zkb_key *in_key[2] = {nullptr, nullptr};
    size_t in_count[2] = {0, 0};
    int in_head = 0, in_pending = 0;
    std::vector<uint64_t> vf_abc_host; // verify.cu: the gamma_ABC_g1 array whose window table is resident (SV_TAB)
    std::vector<int32_t> batch_status; // per proof of the most recent zkb_prove* call: 0 = ok, ZKB_ERR_UNSAT
    std::string err;
    double last_ms = 0.0;
    long launches = 0;
    // scratch (grow-only); slot numbers: enum below
    zkb::DevBuf scratch[64];
    void *pinned = nullptr; // small pinned staging area for results
    size_t pinned_cap = 0;
    void *stage = nullptr; // pinned staging area for large inputs that arrive in pageable memory (upload_rows)
    size_t stage_cap = 0;
    // optional profiling of the dominant kernel (bucket accumulation), per group: [0] = G1, [1] = G2
    bool prof = false;
    cudaEvent_t pe[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
    uint32_t *prof_pairs_host = nullptr; // pinned, 2 words
    bool prof_pending[2] = {false, false};
    double prof_ms[2] = {0, 0}, prof_pairs[2] = {0, 0};
    long prof_launches[2] = {0, 0};
};

namespace zkb {

// scratch slots of a context.  An MSM run owns one "lane" of slots; the prover runs three lanes concurrently
// (G1 job C on the main stream, G1 jobs A + B1 and the G2 job on the side streams)
enum {
    SB_HIST = 0, SB_START, SB_CURSOR, SB_BLOCKSUM, SB_ENT, SB_EBKT, SB_PARTS, SB_BUCKETS, SB_L1, SB_L2, SB_RES, SB_WIDE, SB_WIDESUM,
    SB_PER_GROUP,                      // slots above are per lane: index + msm_slot(ctx, group)
    SB_SCAL = 3 * SB_PER_GROUP, SB_PLAIN, SB_STAGE_A, SB_STAGE_B, SB_STAGE_O, SB_MISC,
    SB_PROVER0 // first slot of prover.cu
};
inline int group_slot(int group) { return group == ZKB_G2 ? SB_PER_GROUP : 0; }
inline int msm_slot(const zkb_ctx *ctx, int group) { return group_slot(group) + ctx->lane_bias; }
constexpr int LANE_G1_SIDE = 2 * SB_PER_GROUP; // lane_bias of the second G1 lane

// enqueue on a side stream / another scratch lane for the lifetime of the scope (host-side switch only)
struct LaneScope {
    zkb_ctx *ctx;
    cudaStream_t saved_stream;
    int saved_bias;
    LaneScope(zkb_ctx *c, cudaStream_t st, int bias) : ctx(c), saved_stream(c->stream), saved_bias(c->lane_bias)
    {
        c->stream = st;
        c->lane_bias = bias;
    }
    ~LaneScope()
    {
        ctx->stream = saved_stream;
        ctx->lane_bias = saved_bias;
    }
};

extern thread_local std::string g_init_err;

inline int fail(zkb_ctx *ctx, int code, const std::string &msg)
{
    if (ctx)
        ctx->err = msg;
    else
        g_init_err = msg;
    return code;
}

#define ZKB_CUDA(ctx, call)                                                                                          \
    do {                                                                                                             \
        cudaError_t e__ = (call);                                                                                    \
        if (e__ != cudaSuccess)                                                                                      \
            return zkb::fail(ctx, ZKB_ERR_CUDA,                                                                      \
                             std::string(#call) + ": " + cudaGetErrorString(e__) + " (" + __FILE__ + ":" +           \
                                 std::to_string(__LINE__) + ")");                                                    \
    } while (0)

#define ZKB_LAUNCH_CHECK(ctx)                                                                                        \
    do {                                                                                                             \
        (ctx)->launches++;                                                                                           \
        cudaError_t e__ = cudaGetLastError();                                                                        \
        if (e__ != cudaSuccess)                                                                                      \
            return zkb::fail(ctx, ZKB_ERR_CUDA,                                                                      \
                             std::string("kernel launch: ") + cudaGetErrorString(e__) + " (" + __FILE__ + ":" +      \
                                 std::to_string(__LINE__) + ")");                                                    \
    } while (0)

// ---- vectorised 16-byte global access for field elements / points -----------------------
template <class T> __device__ __forceinline__ T ld_vec(const T *p)
{
    static_assert(sizeof(T) % 16 == 0, "16-byte multiples only");
    T r;
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) d[i] = s[i];
    return r;
}
template <class T> __device__ __forceinline__ T ld_vec_ro(const T *p)
{
    static_assert(sizeof(T) % 16 == 0, "16-byte multiples only");
    T r;
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) d[i] = __ldg(s + i);
    return r;
}
template <class T> __device__ __forceinline__ void st_vec(T *p, const T &v)
{
    static_assert(sizeof(T) % 16 == 0, "16-byte multiples only");
    uint4 *d = reinterpret_cast<uint4 *>(p);
    const uint4 *s = reinterpret_cast<const uint4 *>(&v);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) d[i] = s[i];
}

inline unsigned ceil_div(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

// host-side helpers implemented in api.cu
int ensure_pinned(zkb_ctx *ctx, size_t bytes);
// `rows` rows of `width` bytes from host (pageable or pinned) or device memory at src (pitch spitch) to device rows at
// dst (pitch dpitch), ordered on `st`.  Large pageable inputs go through ctx->stage with several host threads copying.
int upload_rows(zkb_ctx *ctx, void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t rows, cudaStream_t st);
// folds the event pairs recorded since the last call into the profile counters (stream must be idle)
void prof_collect(zkb_ctx *ctx);

} // namespace zkb
