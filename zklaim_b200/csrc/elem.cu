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

// ---- f1: point decompression (alt_bn128_g1.cpp:436-450, alt_bn128_g2.cpp:449-463) ---------------------------------
// The reference stores a key element as (is_zero, X, lsb of Y) and recovers Y = sqrt(X^3 + b) with Tonelli-Shanks
// (fp.tcc:724-776, fp2.tcc:146-198) on import, then picks the root whose plain-integer least significant bit (G2: of
// Y.c0) equals the stored one.  q = 3 (mod 4), so a square root in Fq is the single fixed-exponent power
// a^((q + 1) / 4); which root the algorithm returns does not matter, the parity rule selects the same Y.
__device__ __forceinline__ Fq fq_sqrt_candidate(const Fq &a)
{
    uint32_t e[8]; // (q + 1) / 4: q + 1 does not carry out of the lowest word
#pragma unroll
    for (int i = 0; i < 8; i++) e[i] = FqP::P(i);
    e[0] += 1u;
#pragma unroll
    for (int i = 0; i < 8; i++) e[i] = (e[i] >> 2) | (i < 7 ? (e[i + 1] << 30) : 0u);
    Fq r = Fq::one();
    for (int i = 253; i >= 0; i--) { // (q + 1) / 4 < 2^252
        r = r.sqr();
        if ((e[i >> 5] >> (i & 31)) & 1) r = r * a;
    }
    return r;
}
__device__ __forceinline__ bool fq_sqrt(const Fq &a, Fq &r)
{
    r = fq_sqrt_candidate(a);
    return r.sqr() == a;
}
// square root in Fq2 = Fq[u] / (u^2 + 1) by the norm: a0 + a1 u = (x0 + x1 u)^2 with x0^2 = (a0 +- sqrt(a0^2 + a1^2)) / 2
__device__ __forceinline__ bool fq2_sqrt(const Fq2 &a, const Fq &inv2, Fq2 &r)
{
    if (a.c1.is_zero()) {
        Fq t;
        if (fq_sqrt(a.c0, t)) {
            r = Fq2{t, Fq::zero()};
            return true;
        }
        if (!fq_sqrt(a.c0.neg(), t)) return false;
        r = Fq2{Fq::zero(), t};
        return true;
    }
    Fq s;
    if (!fq_sqrt(a.c0.sqr() + a.c1.sqr(), s)) return false;
    Fq x0;
    if (!fq_sqrt((a.c0 + s) * inv2, x0) && !fq_sqrt((a.c0 - s) * inv2, x0)) return false;
    r = Fq2{x0, a.c1 * x0.dbl().inverse()};
    return r.sqr() == a;
}
__device__ __forceinline__ uint32_t plain_lsb(const Fq &y) { return y.from_mont().v[0] & 1u; }
__device__ __forceinline__ uint32_t plain_lsb(const Fq2 &y) { return y.c0.from_mont().v[0] & 1u; }

// flags[i]: bit 0 = the stored lsb of Y, bit 1 = is_zero.  status[i] = 1 ok, 0 = X is not the abscissa of a curve point
__global__ void k_decompress_g1(const Fq *__restrict__ x, const uint8_t *__restrict__ flags, size_t n, Fq b,
                                Affine<Fq> *__restrict__ out, int32_t *__restrict__ status)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Affine<Fq> p;
    p.x = Fq::zero();
    p.y = Fq::zero();
    int ok = 1;
    if (!(flags[i] & 2)) {
        const Fq X = ld_vec(x + i);
        Fq Y;
        ok = fq_sqrt(X.sqr() * X + b, Y) ? 1 : 0;
        if (ok) {
            if (plain_lsb(Y) != (uint32_t)(flags[i] & 1)) Y = Y.neg();
            p.x = X;
            p.y = Y;
        }
    }
    st_vec(out + i, p);
    status[i] = ok;
}
__global__ void k_decompress_g2(const Fq2 *__restrict__ x, const uint8_t *__restrict__ flags, size_t n, Fq2 b, Fq inv2,
                                Affine<Fq2> *__restrict__ out, int32_t *__restrict__ status)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Affine<Fq2> p;
    p.x = Fq2::zero();
    p.y = Fq2::zero();
    int ok = 1;
    if (!(flags[i] & 2)) {
        const Fq2 X = ld_vec(x + i);
        Fq2 Y;
        ok = fq2_sqrt(X.sqr() * X + b, inv2, Y) ? 1 : 0;
        if (ok) {
            if (plain_lsb(Y) != (uint32_t)(flags[i] & 1)) Y = Y.neg();
            p.x = X;
            p.y = Y;
        }
    }
    st_vec(out + i, p);
    status[i] = ok;
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


int zkb_points_decompress(zkb_ctx *ctx, int group, const uint64_t *x, const uint8_t *flags, size_t n, uint64_t *out_affine,
                          int32_t *status)
{
    if (!ctx || !x || !flags || !out_affine || (group != ZKB_G1 && group != ZKB_G2))
        return fail(ctx, ZKB_ERR_ARG, "zkb_points_decompress: bad argument");
    if (n == 0) return ZKB_OK;
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const size_t xs = group == ZKB_G1 ? 32 : 64, ps = 2 * xs;
    void *dx, *df;
    int rc;
    if ((rc = stage(ctx, SB_A, x, n * xs, &dx))) return rc;
    if ((rc = stage(ctx, SB_B, flags, n, &df))) return rc;
    ZKB_CUDA(ctx, ctx->scratch[SB_O].ensure(n * ps + n * 4));
    char *d_o = ctx->scratch[SB_O].as<char>();
    int32_t *d_st = (int32_t *)(d_o + n * ps);
    // curve coefficients (alt_bn128_init.cpp:124-128): b = 3, twist b = 3 / (9 + u); host side of fp.cuh
    Fq three = Fq::zero();
    three.v[0] = 3;
    three = three.to_mont();
    const unsigned g = ceil_div(n, 64);
    if (group == ZKB_G1) {
        k_decompress_g1<<<g, 64, 0, ctx->stream>>>((const Fq *)dx, (const uint8_t *)df, n, three, (Affine<Fq> *)d_o, d_st);
    } else {
        Fq nine = Fq::zero();
        nine.v[0] = 9;
        nine = nine.to_mont();
        const Fq2 xi_inv = Fq2{nine, Fq::one()}.inverse();
        const Fq2 twist_b = Fq2{three * xi_inv.c0, three * xi_inv.c1};
        const Fq inv2 = Fq::one().dbl().inverse();
        k_decompress_g2<<<g, 64, 0, ctx->stream>>>((const Fq2 *)dx, (const uint8_t *)df, n, twist_b, inv2, (Affine<Fq2> *)d_o, d_st);
    }
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaMemcpyAsync(out_affine, d_o, n * ps, cudaMemcpyDeviceToHost, ctx->stream));
    std::vector<int32_t> st_host;
    int32_t *st = status;
    if (!st) {
        st_host.resize(n);
        st = st_host.data();
    }
    ZKB_CUDA(ctx, cudaMemcpyAsync(st, d_st, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < n; i++)
        if (!st[i]) return fail(ctx, ZKB_ERR_ARG, "zkb_points_decompress: an abscissa is not on the curve");
    return ZKB_OK;
}

} // extern "C"
