// elem.cu -- element-wise field / group entry points (K1-K3 parity surface).
//
// One thread per element; used by the parity tests to pin the device field and group
// arithmetic against libff's Fp_model / Fp2_model (fp.tcc, fp2.tcc) and alt_bn128_G1/G2
// (alt_bn128_g1.cpp:139-358, alt_bn128_g2.cpp:149-365), and by the prover's final assembly.
#include "common.cuh"

using namespace zkb;

namespace {
enum { SB_A = SB_STAGE_A, SB_B = SB_STAGE_B, SB_O = SB_STAGE_O };

template <class F> __global__ void k_field_op(int op, const F *a, const F *b, F *o, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const F x = ld_vec(a + i);
    F y = F::zero();
    if (b) y = ld_vec(b + i);
    F z;
    switch (op) {
    case ZKB_OP_ADD: z = x + y; break;
    case ZKB_OP_SUB: z = x - y; break;
    case ZKB_OP_MUL: z = x * y; break;
    case ZKB_OP_SQR: z = x.sqr(); break;
    case ZKB_OP_INV: z = x.inverse(); break;
    default: z = x.neg(); break;
    }
    st_vec(o + i, z);
}

template <class F> __global__ void k_group_op(int op, const Jacobian<F> *a, const Jacobian<F> *b, Jacobian<F> *o, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    XYZZ<F> x = XYZZ<F>::from_jacobian(ld_vec(a + i));
    switch (op) {
    case ZKB_GOP_ADD: x.add(XYZZ<F>::from_jacobian(ld_vec(b + i))); break;
    case ZKB_GOP_DBL: x = x.dbl(); break;
    case ZKB_GOP_MIXED_ADD: {
        const Jacobian<F> q = ld_vec(b + i);
        Affine<F> p;
        if (q.Z.is_zero()) {
            p.x = F::zero();
            p.y = F::zero();
        } else {
            p.x = q.X;
            p.y = q.Y;
        }
        x.madd(p);
        break;
    }
    case ZKB_GOP_NORMALIZE: {
        // the reference's to_affine_coordinates (alt_bn128_g1.cpp:60-77): (X/Z^2, Y/Z^3, 1) or zero (0, 1, 0)
        Jacobian<F> j;
        if (x.is_inf()) {
            j.X = F::zero();
            j.Y = F::one();
            j.Z = F::zero();
        } else {
            const F inv = (x.ZZ * x.ZZZ).inverse();
            j.X = x.X * (inv * x.ZZZ);
            j.Y = x.Y * (inv * x.ZZ);
            j.Z = F::one();
        }
        st_vec(o + i, j);
        return;
    }
    default: x = x.neg(); break;
    }
    st_vec(o + i, x.to_jacobian());
}

// o[0] = sum of n Jacobian points (one block; the partial sums of a range-sharded MSM, multiexp.tcc:433-438)
template <class F> __global__ void __launch_bounds__(32) k_group_sum(const Jacobian<F> *a, Jacobian<F> *o, size_t n)
{
    __shared__ XYZZ<F> sh[32];
    const int t = threadIdx.x;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (size_t i = t; i < n; i += 32) acc.add(XYZZ<F>::from_jacobian(ld_vec(a + i)));
    sh[t] = acc;
    __syncwarp();
    for (int d = 16; d > 0; d >>= 1) {
        if (t < d) {
            acc.add(sh[t + d]);
            sh[t] = acc;
        }
        __syncwarp();
    }
    if (t == 0) st_vec(o, acc.to_jacobian());
}

template <class F> __global__ void k_scalar_mul(const Fr *s, const Jacobian<F> *p, Jacobian<F> *o, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Fr k = ld_vec(s + i).from_mont();
    const XYZZ<F> r = scalar_mul<F>(k.v, XYZZ<F>::from_jacobian(ld_vec(p + i)));
    st_vec(o + i, r.to_jacobian());
}

int stage(zkb_ctx *ctx, int slot, const void *host, size_t bytes, void **dev)
{
    *dev = nullptr;
    if (!host) return ZKB_OK;
    ZKB_CUDA(ctx, ctx->scratch[slot].ensure(bytes));
    ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->scratch[slot].p, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    *dev = ctx->scratch[slot].p;
    return ZKB_OK;
}
} // namespace

extern "C" {

int zkb_field_op(zkb_ctx *ctx, int field, int op, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t n)
{
    if (!ctx || !a || !out || field < ZKB_FR || field > ZKB_FQ2 || op < 0 || op > ZKB_OP_NEG)
        return fail(ctx, ZKB_ERR_ARG, "zkb_field_op: bad argument");
    if ((op == ZKB_OP_ADD || op == ZKB_OP_SUB || op == ZKB_OP_MUL) && !b)
        return fail(ctx, ZKB_ERR_ARG, "zkb_field_op: binary op needs b");
    if (n == 0) return ZKB_OK;
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const size_t es = field == ZKB_FQ2 ? 64 : 32;
    void *da, *db;
    int rc;
    if ((rc = stage(ctx, SB_A, a, n * es, &da))) return rc;
    if ((rc = stage(ctx, SB_B, b, n * es, &db))) return rc;
    ZKB_CUDA(ctx, ctx->scratch[SB_O].ensure(n * es));
    void *d_o = ctx->scratch[SB_O].p;
    const unsigned g = ceil_div(n, 128);
    if (field == ZKB_FR)
        k_field_op<Fr><<<g, 128, 0, ctx->stream>>>(op, (const Fr *)da, (const Fr *)db, (Fr *)d_o, n);
    else if (field == ZKB_FQ)
        k_field_op<Fq><<<g, 128, 0, ctx->stream>>>(op, (const Fq *)da, (const Fq *)db, (Fq *)d_o, n);
    else
        k_field_op<Fq2><<<g, 128, 0, ctx->stream>>>(op, (const Fq2 *)da, (const Fq2 *)db, (Fq2 *)d_o, n);
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaMemcpyAsync(out, d_o, n * es, cudaMemcpyDeviceToHost, ctx->stream));
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZKB_OK;
}

int zkb_group_op(zkb_ctx *ctx, int group, int op, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t n)
{
    if (!ctx || !a || !out || (group != ZKB_G1 && group != ZKB_G2) || op < 0 || op > ZKB_GOP_NORMALIZE)
        return fail(ctx, ZKB_ERR_ARG, "zkb_group_op: bad argument");
    if ((op == ZKB_GOP_ADD || op == ZKB_GOP_MIXED_ADD) && !b)
        return fail(ctx, ZKB_ERR_ARG, "zkb_group_op: binary op needs b");
    if (n == 0) return ZKB_OK;
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const size_t es = group == ZKB_G1 ? 96 : 192;
    void *da, *db;
    int rc;
    if ((rc = stage(ctx, SB_A, a, n * es, &da))) return rc;
    if ((rc = stage(ctx, SB_B, b, n * es, &db))) return rc;
    ZKB_CUDA(ctx, ctx->scratch[SB_O].ensure(n * es));
    void *d_o = ctx->scratch[SB_O].p;
    const unsigned g = ceil_div(n, 64);
    if (group == ZKB_G1)
        k_group_op<Fq><<<g, 64, 0, ctx->stream>>>(op, (const G1J *)da, (const G1J *)db, (G1J *)d_o, n);
    else
        k_group_op<Fq2><<<g, 64, 0, ctx->stream>>>(op, (const G2J *)da, (const G2J *)db, (G2J *)d_o, n);
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaMemcpyAsync(out, d_o, n * es, cudaMemcpyDeviceToHost, ctx->stream));
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZKB_OK;
}

int zkb_group_sum(zkb_ctx *ctx, int group, const uint64_t *points, size_t n, uint64_t *out)
{
    if (!ctx || !out || (!points && n) || (group != ZKB_G1 && group != ZKB_G2))
        return fail(ctx, ZKB_ERR_ARG, "zkb_group_sum: bad argument");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const size_t es = group == ZKB_G1 ? 96 : 192;
    void *da = nullptr;
    int rc;
    if (n && (rc = stage(ctx, SB_A, points, n * es, &da))) return rc;
    ZKB_CUDA(ctx, ctx->scratch[SB_O].ensure(es));
    void *d_o = ctx->scratch[SB_O].p;
    if (group == ZKB_G1)
        k_group_sum<Fq><<<1, 32, 0, ctx->stream>>>((const G1J *)da, (G1J *)d_o, n);
    else
        k_group_sum<Fq2><<<1, 32, 0, ctx->stream>>>((const G2J *)da, (G2J *)d_o, n);
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaMemcpyAsync(out, d_o, es, cudaMemcpyDeviceToHost, ctx->stream));
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZKB_OK;
}

int zkb_group_scalar_mul(zkb_ctx *ctx, int group, const uint64_t *scalars, const uint64_t *points, uint64_t *out, size_t n)
{
    if (!ctx || !scalars || !points || !out || (group != ZKB_G1 && group != ZKB_G2))
        return fail(ctx, ZKB_ERR_ARG, "zkb_group_scalar_mul: bad argument");
    if (n == 0) return ZKB_OK;
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const size_t es = group == ZKB_G1 ? 96 : 192;
    void *ds, *dp;
    int rc;
    if ((rc = stage(ctx, SB_A, scalars, n * 32, &ds))) return rc;
    if ((rc = stage(ctx, SB_B, points, n * es, &dp))) return rc;
    ZKB_CUDA(ctx, ctx->scratch[SB_O].ensure(n * es));
    void *d_o = ctx->scratch[SB_O].p;
    const unsigned g = ceil_div(n, 32);
    if (group == ZKB_G1)
        k_scalar_mul<Fq><<<g, 32, 0, ctx->stream>>>((const Fr *)ds, (const G1J *)dp, (G1J *)d_o, n);
    else
        k_scalar_mul<Fq2><<<g, 32, 0, ctx->stream>>>((const Fr *)ds, (const G2J *)dp, (G2J *)d_o, n);
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaMemcpyAsync(out, d_o, n * es, cudaMemcpyDeviceToHost, ctx->stream));
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZKB_OK;
}

} // extern "C"
