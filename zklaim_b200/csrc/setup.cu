// setup.cu -- K11: the Groth16 generator for fixed secret randomness.
//
// Replaces r1cs_gg_ppzksnark_generator (r1cs_gg_ppzksnark.tcc:211-388) from the point where the reference has drawn
// t, alpha, beta, gamma, delta and the two group generators.  Every value below is a field or group ELEMENT, so any
// order of evaluation gives the reference's bits; the data-parallel shape is chosen for the GPU:
//
//   k_lagrange     u_i = L_i(t) for the m points of the domain, one thread per point: a closed form with one field
//                  inversion each (the reference walks the domain sequentially, basic_radix2_domain_aux.tcc:225-233)
//   k_col_eval     A_j(t) = sum_i u_i * A[i][j]: a product with the TRANSPOSE of the constraint matrices, one warp
//                  per variable over a CSC index built on the host (r1cs_to_qap.tcc:124-154 scatters row by row)
//   k_combine      (beta A_j + alpha B_j + C_j) / gamma | delta           (r1cs_gg_ppzksnark.tcc:254-275)
//   k_pow_scaled   Zt / delta * t^i                                        (r1cs_to_qap.tcc:156-161, tcc:320)
//   k_fb_table     d * 2^(8 j) * G for the 32 windows of 8 bits           (multiexp.tcc:547-596 get_window_table)
//   k_fixed_base   k * G = sum_j table[j][byte j of k]: 32 mixed additions and one inversion per scalar, affine out
//                  (multiexp.tcc:598-681 windowed_exp / batch_exp, batch_to_special)
//
// All kernels are bound by the integer multiply pipe like the rest of the library; the window table (0.5 MB in G1,
// 1 MB in G2) lives in L2.  One generator run for 2^22 constraints is ~2 * 10^10 modular products, well under a second.
#include <algorithm>

#include "common.cuh"
#include "ntt.cuh"

using namespace zkb;

namespace {

constexpr int FB_W = 8;                  // window bits
constexpr int FB_WINDOWS = 256 / FB_W;   // 32
constexpr int FB_ROW = (1 << FB_W) - 1;  // non-zero digits per window

__device__ __forceinline__ Fr dev_pow(Fr b, size_t e)
{
    Fr acc = Fr::one();
    for (; e; e >>= 1) {
        if (e & 1) acc = acc * b;
        b = b.sqr();
    }
    return acc;
}

// out[i] = l0 * w^i / ((tt - w^i) * (use2 ? e^i - sub : 1))
__global__ void k_lagrange(Fr *__restrict__ out, size_t n, Fr w, Fr l0, Fr tt, Fr e, Fr sub, int use2)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Fr wi = dev_pow(w, i);
    Fr den = tt - wi;
    if (use2) den = den * (dev_pow(e, i) - sub);
    st_vec(out + i, l0 * wi * den.inverse());
}

// out[i] = scale * base^i
__global__ void k_pow_scaled(Fr *__restrict__ out, size_t n, Fr base, Fr scale)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    st_vec(out + i, scale * dev_pow(base, i));
}

__device__ __forceinline__ Fr warp_sum(Fr a)
{
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        Fr o;
#pragma unroll
        for (int k = 0; k < 8; k++) o.v[k] = __shfl_xor_sync(0xffffffffu, a.v[k], d);
        a = a + o;
    }
    return a;
}

// One warp per column j of a constraint matrix in CSC form: out[j] = sum over the column of u[row] * coeff[pos]
// (+ u[nc + j] for j <= ni when `inputs` is set: the input-consistency rows of r1cs_to_qap.tcc:119-122).
__global__ void __launch_bounds__(128) k_col_eval(const uint32_t *__restrict__ col_ptr, const uint2 *__restrict__ ent,
                                                   const Fr *__restrict__ coeff, const Fr *__restrict__ u, size_t ncols, size_t nc,
                                                   size_t ni, int inputs, Fr *__restrict__ out)
{
    const size_t j = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (j >= ncols) return;
    Fr acc = Fr::zero();
    const uint32_t hi = col_ptr[j + 1];
    for (uint32_t e = col_ptr[j] + lane; e < hi; e += 32) {
        const uint2 rp = ent[e];
        acc = acc + ld_vec_ro(u + rp.x) * ld_vec_ro(coeff + rp.y);
    }
    acc = warp_sum(acc);
    if (lane == 0) {
        if (inputs && j <= ni) acc = acc + ld_vec_ro(u + nc + j);
        st_vec(out + j, acc);
    }
}

// x_j = (beta A_j + alpha B_j + C_j) * (j <= ni ? 1 / gamma : 1 / delta)
__global__ void k_combine(const Fr *__restrict__ A, const Fr *__restrict__ B, const Fr *__restrict__ C, size_t n, size_t ni, Fr alpha,
                          Fr beta, Fr ginv, Fr dinv, Fr *__restrict__ out)
{
    const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const Fr x = beta * ld_vec(A + j) + alpha * ld_vec(B + j) + ld_vec(C + j);
    st_vec(out + j, x * (j <= ni ? ginv : dinv));
}

template <class F> __device__ __forceinline__ Affine<F> to_affine(const XYZZ<F> &p)
{
    Affine<F> r;
    if (p.is_inf()) {
        r.x = F::zero();
        r.y = F::zero();
        return r;
    }
    const F a = p.ZZZ.inverse();         // ZZ^3 = ZZZ^2  =>  1 / ZZ = (ZZ / ZZZ)^2
    r.x = p.X * (p.ZZ * a).sqr();
    r.y = p.Y * a;
    return r;
}

// table[j * FB_ROW + d - 1] = d * 2^(8 j) * G, affine
template <class F> __global__ void k_fb_table(XYZZ<F> gen, Affine<F> *__restrict__ table)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= FB_WINDOWS * FB_ROW) return;
    const int j = idx / FB_ROW, d = idx % FB_ROW + 1;
    uint32_t k[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    k[j >> 2] = (uint32_t)d << ((j & 3) * 8);
    st_vec(table + idx, to_affine(scalar_mul<F>(k, gen)));
}

template <class F> __global__ void __launch_bounds__(128) k_fixed_base(const Affine<F> *__restrict__ table, const Fr *__restrict__ s,
                                                                        size_t n, Affine<F> *__restrict__ out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Fr k = ld_vec(s + i).from_mont();
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int j = 0; j < FB_WINDOWS; j++) {
        const uint32_t d = (k.v[j >> 2] >> ((j & 3) * 8)) & 255u;
        if (d) acc.madd(ld_vec_ro(table + j * FB_ROW + d - 1));
    }
    st_vec(out + i, to_affine(acc));
}

} // namespace

namespace {

__global__ void k_put3(Fr *dst, Fr a, Fr b, Fr c)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        dst[0] = a;
        dst[1] = b;
        dst[2] = c;
    }
}

// device allocations of one generator run, released on every exit path
struct Tmp {
    std::vector<void *> ptrs;
    ~Tmp()
    {
        for (void *p : ptrs) cudaFree(p);
    }
    template <class T> cudaError_t get(T **p, size_t count)
    {
        void *q = nullptr;
        const cudaError_t e = cudaMalloc(&q, (count ? count : 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(q);
        *p = (T *)q;
        return e;
    }
};

inline Fr ld_fr(const uint64_t *p)
{
    Fr r;
    memcpy(r.v, p, 32);
    return r;
}
inline bool fr_canonical(const Fr &a)
{
    for (int i = 7; i >= 0; i--) {
        if (a.v[i] < FrP::P(i)) return true;
        if (a.v[i] > FrP::P(i)) return false;
    }
    return false;
}

struct Csc {
    uint32_t *col_ptr = nullptr;
    uint2 *ent = nullptr; // (row, position of the coefficient in the CSR arrays)
    Fr *coeff = nullptr;
};

// Transposed index of one constraint matrix (index work only; the arithmetic is k_col_eval's).
int upload_transposed(zkb_ctx *ctx, Tmp &tmp, const zkb_csr &M, size_t nc, size_t ncols, Csc *out)
{
    const size_t nnz = M.row_ptr[nc];
    if (nnz >= 0xffffffffull || nc >= 0xffffffffull) return fail(ctx, ZKB_ERR_ARG, "zkb_setup: matrix too large for 32-bit indices");
    std::vector<uint32_t> ptr(ncols + 1, 0);
    for (size_t e = 0; e < nnz; e++) {
        if (M.col[e] >= ncols) return fail(ctx, ZKB_ERR_ARG, "zkb_setup: column index out of range");
        ptr[M.col[e] + 1]++;
    }
    for (size_t j = 0; j < ncols; j++) ptr[j + 1] += ptr[j];
    std::vector<uint2> ent(nnz);
    {
        std::vector<uint32_t> cur(ptr.begin(), ptr.end() - 1);
        for (size_t i = 0; i < nc; i++) {
            if (M.row_ptr[i + 1] < M.row_ptr[i] || M.row_ptr[i + 1] > nnz) return fail(ctx, ZKB_ERR_ARG, "zkb_setup: bad row_ptr");
            for (size_t e = M.row_ptr[i]; e < M.row_ptr[i + 1]; e++) ent[cur[M.col[e]]++] = make_uint2((uint32_t)i, (uint32_t)e);
        }
    }
    ZKB_CUDA(ctx, tmp.get(&out->col_ptr, ncols + 1));
    ZKB_CUDA(ctx, tmp.get(&out->ent, nnz));
    ZKB_CUDA(ctx, tmp.get(&out->coeff, nnz));
    // stream-ordered copies: a synchronous cudaMemcpy from pageable memory may return before its DMA has landed, and the
    // context's streams are non-blocking, so the kernel that follows on ctx->stream would not wait for it
    cudaStream_t st = ctx->stream;
    ZKB_CUDA(ctx, cudaMemcpyAsync(out->col_ptr, ptr.data(), (ncols + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    if (nnz) {
        ZKB_CUDA(ctx, cudaMemcpyAsync(out->ent, ent.data(), nnz * sizeof(uint2), cudaMemcpyHostToDevice, st));
        ZKB_CUDA(ctx, cudaMemcpyAsync(out->coeff, M.coeff, nnz * sizeof(Fr), cudaMemcpyHostToDevice, st));
    }
    ZKB_CUDA(ctx, cudaStreamSynchronize(st)); // the host vectors go out of scope here
    return ZKB_OK;
}

// r1cs_constraint_system::swap_AB_if_beneficial (r1cs.tcc:182-231): B is swapped with A when it touches more variables
bool swap_is_beneficial(const zkb_setup_desc *d)
{
    const size_t ncols = d->num_variables + 1;
    std::vector<uint8_t> ta(ncols, 0), tb(ncols, 0);
    const size_t na = d->A.row_ptr[d->num_constraints], nb = d->B.row_ptr[d->num_constraints];
    for (size_t e = 0; e < na; e++)
        if (d->A.col[e] < ncols) ta[d->A.col[e]] = 1;
    for (size_t e = 0; e < nb; e++)
        if (d->B.col[e] < ncols) tb[d->B.col[e]] = 1;
    size_t ca = 0, cb = 0;
    for (size_t j = 0; j < ncols; j++) {
        ca += ta[j];
        cb += tb[j];
    }
    return cb > ca;
}

template <class F>
int fixed_base(zkb_ctx *ctx, Tmp &tmp, const uint64_t *gen_jac, const Fr *d_scalars, size_t n, Affine<F> **d_out)
{
    Jacobian<F> g;
    memcpy(&g, gen_jac, sizeof(g));
    Affine<F> *table;
    ZKB_CUDA(ctx, tmp.get(&table, (size_t)FB_WINDOWS * FB_ROW));
    ZKB_CUDA(ctx, tmp.get(d_out, n));
    k_fb_table<F><<<ceil_div(FB_WINDOWS * FB_ROW, 64), 64, 0, ctx->stream>>>(XYZZ<F>::from_jacobian(g), table);
    ZKB_LAUNCH_CHECK(ctx);
    k_fixed_base<F><<<ceil_div(n, 128), 128, 0, ctx->stream>>>(table, d_scalars, n, *d_out);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

} // namespace

extern "C" int zkb_setup(zkb_ctx *ctx, const zkb_setup_desc *d, zkb_setup_out *o, int *swapped)
{
    if (!ctx || !d || !o) return fail(ctx, ZKB_ERR_ARG, "zkb_setup: NULL argument");
    if (!d->t || !d->alpha || !d->beta || !d->gamma || !d->delta || !d->g1_generator || !d->g2_generator || !d->A.row_ptr ||
        !d->B.row_ptr || !d->C.row_ptr)
        return fail(ctx, ZKB_ERR_ARG, "zkb_setup: incomplete description");
    if (!o->alpha_g1 || !o->beta_g1 || !o->delta_g1 || !o->beta_g2 || !o->delta_g2 || !o->gamma_g2 || !o->A_query || !o->B_g1 ||
        !o->B_g2 || !o->H_query || !o->L_query || !o->gamma_ABC_g1)
        return fail(ctx, ZKB_ERR_ARG, "zkb_setup: every output array must be allocated by the caller");
    const size_t nc = d->num_constraints, nv = d->num_variables, ni = d->num_inputs, ncols = nv + 1;
    if (ni > nv) return fail(ctx, ZKB_ERR_ARG, "zkb_setup: num_inputs > num_variables");
    size_t m = 0;
    int kind = 0;
    if (domain_select_host(nc + ni + 1, &m, &kind)) return fail(ctx, ZKB_ERR_DOMAIN, "zkb_setup: no supported evaluation domain for this size");
    const Fr t = ld_fr(d->t), alpha = ld_fr(d->alpha), beta = ld_fr(d->beta), gamma = ld_fr(d->gamma), delta = ld_fr(d->delta);
    if (!fr_canonical(t) || !fr_canonical(alpha) || !fr_canonical(beta) || !fr_canonical(gamma) || !fr_canonical(delta))
        return fail(ctx, ZKB_ERR_ARG, "zkb_setup: scalar not reduced");
    if (gamma.is_zero() || delta.is_zero()) return fail(ctx, ZKB_ERR_ARG, "zkb_setup: gamma and delta must be invertible");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    cudaStream_t st = ctx->stream;
    ctx->launches = 0;

    const bool swap = swap_is_beneficial(d);
    if (swapped) *swapped = swap ? 1 : 0;
    const zkb_csr &MA = swap ? d->B : d->A, &MB = swap ? d->A : d->B;

    // ---- constants of the domain at t (a few dozen field operations on the host, like domain_build) ----
    const Fr one = Fr::one();
    Fr Zt;
    struct Part {
        size_t off, n;
        Fr w, l0, tt, e, sub;
        int use2;
    } part[2];
    int nparts = 0;
    if (kind == ZKB_DOMAIN_BASIC_RADIX2) {
        // basic_radix2_domain_aux.tcc:183-236: u_i = (t^m - 1) / m * w^i / (t - w^i)
        Zt = host_fr_pow(t, m) - one;
        part[nparts++] = Part{0, m, host_fr_root_of_unity(host_ceil_log2(m)), Zt * host_fr_from_u64(m).inverse(), t, one, one, 0};
    } else {
        // step_radix2_domain.tcc:161-186 (and :202-205 for Z)
        const int lg = host_ceil_log2(m);
        const size_t big = (size_t)1 << (lg - 1), small = m - big;
        const Fr omega = host_fr_root_of_unity(lg), big_omega = omega.sqr(), small_omega = host_fr_root_of_unity(host_ceil_log2(small));
        const Fr om_s = host_fr_pow(omega, small);
        const Fr Zb = host_fr_pow(t, big) - one, L0 = host_fr_pow(t, small) - om_s;
        Zt = Zb * L0;
        const Fr tt = t * omega.inverse();
        const Fr Zs = host_fr_pow(tt, small) - one;
        if (!Zt.is_zero()) {
            const Fr L1 = Zb * (host_fr_pow(omega, big) - one).inverse();
            part[nparts++] = Part{0, big, big_omega, Zb * host_fr_from_u64(big).inverse() * L0, t, host_fr_pow(big_omega, small), om_s, 1};
            part[nparts++] = Part{big, small, small_omega, Zs * host_fr_from_u64(small).inverse() * L1, tt, one, one, 0};
        }
    }
    if (Zt.is_zero()) return fail(ctx, ZKB_ERR_ARG, "zkb_setup: t is an element of the evaluation domain");
    const Fr ginv = gamma.inverse(), dinv = delta.inverse();

    Tmp tmp;
    int rc;
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
    Fr *u;
    ZKB_CUDA(ctx, tmp.get(&u, m));
    for (int p = 0; p < nparts; p++) {
        const Part &q = part[p];
        k_lagrange<<<ceil_div(q.n, 128), 128, 0, st>>>(u + q.off, q.n, q.w, q.l0, q.tt, q.e, q.sub, q.use2);
        ZKB_LAUNCH_CHECK(ctx);
    }

    // ---- G1 scalars: [A_j(t) | B_j(t) | Zt / delta * t^i | (beta A + alpha B + C) / (gamma | delta) | alpha beta delta] ----
    const size_t nh = m - 1;
    const size_t oA = 0, oB = ncols, oH = 2 * ncols, oX = oH + nh, oS = oX + ncols, n1 = oS + 3;
    Fr *s1, *Ct, *s2;
    ZKB_CUDA(ctx, tmp.get(&s1, n1));
    ZKB_CUDA(ctx, tmp.get(&Ct, ncols));
    ZKB_CUDA(ctx, tmp.get(&s2, ncols + 3));
    const zkb_csr *mats[3] = {&MA, &MB, &d->C};
    Fr *dst[3] = {s1 + oA, s1 + oB, Ct};
    for (int k = 0; k < 3; k++) {
        Csc csc;
        if ((rc = upload_transposed(ctx, tmp, *mats[k], nc, ncols, &csc))) return rc;
        k_col_eval<<<ceil_div(ncols * 32, 128), 128, 0, st>>>(csc.col_ptr, csc.ent, csc.coeff, u, ncols, nc, ni, k == 0, dst[k]);
        ZKB_LAUNCH_CHECK(ctx);
    }
    if (nh) {
        k_pow_scaled<<<ceil_div(nh, 128), 128, 0, st>>>(s1 + oH, nh, t, Zt * dinv);
        ZKB_LAUNCH_CHECK(ctx);
    }
    k_combine<<<ceil_div(ncols, 128), 128, 0, st>>>(s1 + oA, s1 + oB, Ct, ncols, ni, alpha, beta, ginv, dinv, s1 + oX);
    ZKB_LAUNCH_CHECK(ctx);
    k_put3<<<1, 32, 0, st>>>(s1 + oS, alpha, beta, delta);
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaMemcpyAsync(s2, s1 + oB, ncols * sizeof(Fr), cudaMemcpyDeviceToDevice, st));
    k_put3<<<1, 32, 0, st>>>(s2 + ncols, beta, delta, gamma);
    ZKB_LAUNCH_CHECK(ctx);

    // ---- every exponentiation of the key ----
    G1Affine *o1;
    G2Affine *o2;
    if ((rc = fixed_base<Fq>(ctx, tmp, d->g1_generator, s1, n1, &o1))) return rc;
    if ((rc = fixed_base<Fq2>(ctx, tmp, d->g2_generator, s2, ncols + 3, &o2))) return rc;
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev1, st));

    const size_t a1 = sizeof(G1Affine), a2 = sizeof(G2Affine);
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->A_query, o1 + oA, ncols * a1, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->B_g1, o1 + oB, ncols * a1, cudaMemcpyDeviceToHost, st));
    if (nh) ZKB_CUDA(ctx, cudaMemcpyAsync(o->H_query, o1 + oH, nh * a1, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->gamma_ABC_g1, o1 + oX, (ni + 1) * a1, cudaMemcpyDeviceToHost, st));
    if (nv > ni) ZKB_CUDA(ctx, cudaMemcpyAsync(o->L_query, o1 + oX + ni + 1, (nv - ni) * a1, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->alpha_g1, o1 + oS, a1, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->beta_g1, o1 + oS + 1, a1, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->delta_g1, o1 + oS + 2, a1, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->B_g2, o2, ncols * a2, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->beta_g2, o2 + ncols, a2, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->delta_g2, o2 + ncols + 1, a2, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(o->gamma_g2, o2 + ncols + 2, a2, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_ms = ms;
    return ZKB_OK;
}
