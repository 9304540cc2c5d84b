// pool.cu -- several B200s behind the C ABI ("zkb_init over a device list", "one proof
// stream per GPU", "MSM point ranges split per GPU with a tiny host reduction of partial sums").
//
// A pool owns one zkb_ctx per device and drives them with one host thread per device; nothing here touches a
// kernel: every device runs the single-device entry points of zkb.h on its share of the work.
//   * zkb_pool_prove_batch   independent proofs (zklaim/main_benchmark.c:174-198 proves credential after credential):
//                            the batch is cut into one contiguous slice per device, the key is replicated;
//   * zkb_pool_msm           libff::multi_exp's own chunking (multiexp.tcc:417-431): device i holds the bases of
//                            chunk i, computes its partial sum, the 96- / 192-byte partials are added on device 0
//                            (multiexp.tcc:433-438 -> zkb_group_sum);
//   * zkb_pool_prove_sharded ONE proof over all devices: every device holds one range of the four queries
//                            (zkb_key_upload_shard), computes its shares (zkb_prove_partial), device 0 combines.
#include "common.cuh"
#include <string.h>
#include <thread>

struct zkb_pool {
    std::vector<zkb_ctx *> ctx;
    std::string err;
};
struct zkb_key_set {
    zkb_pool *pool = nullptr;
    bool sharded = false;
    std::vector<zkb_key *> keys;
    std::vector<char> owns; // contexts of one device share ONE replica (a device key is immutable): the first one owns it
};
struct zkb_bases_set {
    zkb_pool *pool = nullptr;
    int group = 0;
    size_t n = 0;
    std::vector<zkb_bases *> parts; // parts[i] holds the bases [lo[i], hi[i])
    std::vector<size_t> lo, hi;
};

namespace {

int pool_fail(zkb_pool *p, int code, const std::string &msg)
{
    if (p) p->err = msg;
    return code;
}

// runs fn(i) for every device of the pool on its own host thread; returns the first non-zero result
template <class Fn> int on_all(zkb_pool *p, Fn fn)
{
    const size_t n = p->ctx.size();
    std::vector<int> rc(n, 0);
    if (n == 1) {
        rc[0] = fn(0);
    } else {
        std::vector<std::thread> th;
        for (size_t i = 0; i < n; i++) th.emplace_back([&, i] { rc[i] = fn((int)i); });
        for (auto &t : th) t.join();
    }
    int first = 0;
    for (size_t i = 0; i < n; i++)
        if (rc[i] && !first) {
            first = rc[i];
            p->err = std::string("device ") + std::to_string(p->ctx[i]->dev) + ": " + zkb_last_error(p->ctx[i]);
        }
    return first;
}

// chunk i of `total` items over `world` devices, as libff::multi_exp cuts them (multiexp.tcc:417-431)
void chunk(size_t total, int i, int world, size_t *lo, size_t *hi)
{
    if (total < (size_t)world || world == 1) {
        *lo = 0;
        *hi = i == 0 ? total : 0;
        return;
    }
    const size_t one = total / world;
    *lo = (size_t)i * one;
    *hi = i == world - 1 ? total : *lo + one;
}

} // namespace

extern "C" {

int zkb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int zkb_pool_init(const int *device_ids, int n, zkb_pool **out)
{
    if (!out || (n > 0 && !device_ids)) return zkb::fail(nullptr, ZKB_ERR_ARG, "zkb_pool_init: bad argument");
    std::vector<int> ids;
    if (n <= 0) {
        const int all = zkb_device_count();
        for (int i = 0; i < all; i++) ids.push_back(i);
    } else {
        ids.assign(device_ids, device_ids + n);
    }
    if (ids.empty()) return zkb::fail(nullptr, ZKB_ERR_CUDA, "zkb_pool_init: no CUDA device (there is no CPU fallback)");
    zkb_pool *p = new zkb_pool();
    for (int id : ids) {
        zkb_ctx *c = nullptr;
        const int rc = zkb_init(id, &c);
        if (rc) {
            for (zkb_ctx *x : p->ctx) zkb_shutdown(x);
            delete p;
            return rc;
        }
        p->ctx.push_back(c);
    }
    *out = p;
    return ZKB_OK;
}

void zkb_pool_shutdown(zkb_pool *p)
{
    if (!p) return;
    for (zkb_ctx *c : p->ctx) zkb_shutdown(c);
    delete p;
}

int zkb_pool_size(const zkb_pool *p) { return p ? (int)p->ctx.size() : 0; }
zkb_ctx *zkb_pool_ctx(zkb_pool *p, int i) { return p && i >= 0 && (size_t)i < p->ctx.size() ? p->ctx[i] : nullptr; }
const char *zkb_pool_last_error(const zkb_pool *p) { return p ? p->err.c_str() : zkb_last_error(nullptr); }

static int key_set_build(zkb_pool *p, const zkb_key_desc *desc, bool sharded, zkb_key_set **out)
{
    if (!p || !desc || !out) return pool_fail(p, ZKB_ERR_ARG, "zkb_pool_key_upload: bad argument");
    zkb_key_set *ks = new zkb_key_set();
    ks->pool = p;
    ks->sharded = sharded;
    ks->keys.assign(p->ctx.size(), nullptr);
    ks->owns.assign(p->ctx.size(), 1);
    const int world = (int)p->ctx.size();
    std::vector<int> first(p->ctx.size()); // the first context on the same device
    for (size_t i = 0; i < p->ctx.size(); i++) {
        first[i] = (int)i;
        for (size_t j = 0; j < i && !sharded; j++)
            if (p->ctx[j]->dev == p->ctx[i]->dev) {
                first[i] = (int)j;
                ks->owns[i] = 0;
                break;
            }
    }
    const int rc = on_all(p, [&](int i) {
        if (!ks->owns[i]) return (int)ZKB_OK;
        return sharded ? zkb_key_upload_shard(p->ctx[i], desc, i, world, &ks->keys[i]) : zkb_key_upload(p->ctx[i], desc, &ks->keys[i]);
    });
    if (rc) {
        for (size_t i = 0; i < ks->keys.size(); i++)
            if (ks->owns[i]) zkb_key_free(ks->keys[i]);
        delete ks;
        return rc;
    }
    for (size_t i = 0; i < ks->keys.size(); i++)
        if (!ks->owns[i]) ks->keys[i] = ks->keys[first[i]];
    *out = ks;
    return ZKB_OK;
}

int zkb_pool_key_upload(zkb_pool *p, const zkb_key_desc *desc, zkb_key_set **out) { return key_set_build(p, desc, false, out); }
int zkb_pool_key_upload_sharded(zkb_pool *p, const zkb_key_desc *desc, zkb_key_set **out) { return key_set_build(p, desc, true, out); }

int zkb_pool_key_tune(zkb_pool *p, zkb_key_set *ks, const uint64_t *w)
{
    if (!p || !ks || !w || ks->pool != p) return pool_fail(p, ZKB_ERR_ARG, "zkb_pool_key_tune: bad argument");
    return on_all(p, [&](int i) { return ks->owns[i] ? zkb_key_tune(p->ctx[i], ks->keys[i], w) : (int)ZKB_OK; });
}

const zkb_key *zkb_pool_key(const zkb_key_set *ks, int i) { return ks && i >= 0 && (size_t)i < ks->keys.size() ? ks->keys[i] : nullptr; }

void zkb_pool_key_free(zkb_key_set *ks)
{
    if (!ks) return;
    for (size_t i = 0; i < ks->keys.size(); i++)
        if (ks->owns[i]) zkb_key_free(ks->keys[i]);
    delete ks;
}

int zkb_pool_prove_batch(zkb_pool *p, const zkb_key_set *ks, size_t count, const uint64_t *w, const uint64_t *r, const uint64_t *s,
                         zkb_proof *out, int32_t *status)
{
    if (!p || !ks || ks->pool != p || ks->sharded || (count && (!w || !r || !s || !out)))
        return pool_fail(p, ZKB_ERR_ARG, "zkb_pool_prove_batch: bad argument (a sharded key set serves zkb_pool_prove_sharded)");
    if (!count) return ZKB_OK;
    const int world = (int)p->ctx.size();
    // witnesses per proof: the key knows; ask through the domain-independent accessor
    size_t nv = 0;
    if (zkb_key_num_variables(ks->keys[0], &nv)) return pool_fail(p, ZKB_ERR_ARG, "zkb_pool_prove_batch: bad key");
    if (status) memset(status, 0, count * sizeof(int32_t));
    return on_all(p, [&](int i) {
        // even slices (the first count % world devices take one more)
        const size_t base = count / world, extra = count % world;
        const size_t lo = (size_t)i * base + ((size_t)i < extra ? (size_t)i : extra);
        const size_t n = base + ((size_t)i < extra ? 1 : 0);
        if (!n) return (int)ZKB_OK;
        const int rc = zkb_prove_batch(p->ctx[i], ks->keys[i], n, w + lo * nv * 4, r + lo * 4, s + lo * 4, out + lo);
        if (status) zkb_prove_batch_status(p->ctx[i], status + lo, n);
        return rc;
    });
}

int zkb_pool_prove_sharded(zkb_pool *p, const zkb_key_set *ks, const uint64_t *w, const uint64_t r[4], const uint64_t s[4],
                           zkb_proof *out)
{
    if (!p || !ks || ks->pool != p || !ks->sharded || !w || !r || !s || !out)
        return pool_fail(p, ZKB_ERR_ARG, "zkb_pool_prove_sharded: bad argument (needs a key set made by zkb_pool_key_upload_sharded)");
    std::vector<zkb_partial> parts(p->ctx.size());
    int rc = on_all(p, [&](int i) { return zkb_prove_partial(p->ctx[i], ks->keys[i], w, r, s, &parts[i], nullptr); });
    if (rc) return rc;
    rc = zkb_prove_combine(p->ctx[0], parts.data(), parts.size(), r, s, out);
    if (rc) p->err = zkb_last_error(p->ctx[0]);
    return rc;
}

int zkb_pool_bases_upload(zkb_pool *p, int group, const uint64_t *affine, size_t n, int window_bits, int precompute_factor,
                          zkb_bases_set **out)
{
    if (!p || !affine || !out || !n || (group != ZKB_G1 && group != ZKB_G2))
        return pool_fail(p, ZKB_ERR_ARG, "zkb_pool_bases_upload: bad argument");
    zkb_bases_set *bs = new zkb_bases_set();
    bs->pool = p;
    bs->group = group;
    bs->n = n;
    const int world = (int)p->ctx.size();
    bs->parts.assign(world, nullptr);
    bs->lo.resize(world);
    bs->hi.resize(world);
    for (int i = 0; i < world; i++) chunk(n, i, world, &bs->lo[i], &bs->hi[i]);
    const size_t words = group == ZKB_G1 ? 8 : 16;
    const int rc = on_all(p, [&](int i) {
        if (bs->hi[i] == bs->lo[i]) return (int)ZKB_OK;
        return zkb_bases_upload(p->ctx[i], group, affine + bs->lo[i] * words, bs->hi[i] - bs->lo[i], window_bits, precompute_factor,
                                &bs->parts[i]);
    });
    if (rc) {
        for (zkb_bases *b : bs->parts) zkb_bases_free(b);
        delete bs;
        return rc;
    }
    *out = bs;
    return ZKB_OK;
}

void zkb_pool_bases_free(zkb_bases_set *bs)
{
    if (!bs) return;
    for (zkb_bases *b : bs->parts) zkb_bases_free(b);
    delete bs;
}

int zkb_pool_msm(zkb_pool *p, const zkb_bases_set *bs, const uint64_t *scalars, size_t n, uint64_t *out)
{
    if (!p || !bs || bs->pool != p || !out || (!scalars && n)) return pool_fail(p, ZKB_ERR_ARG, "zkb_pool_msm: bad argument");
    if (n != bs->n) return pool_fail(p, ZKB_ERR_ARG, "zkb_pool_msm: one scalar per uploaded base is required");
    const size_t words = bs->group == ZKB_G1 ? 12 : 24;
    const size_t world = p->ctx.size();
    std::vector<uint64_t> parts(world * words);
    std::vector<char> live(world, 0);
    int rc = on_all(p, [&](int i) {
        if (!bs->parts[i]) return (int)ZKB_OK;
        live[i] = 1;
        return zkb_msm(p->ctx[i], bs->parts[i], 0, scalars + bs->lo[i] * 4, bs->hi[i] - bs->lo[i], &parts[(size_t)i * words]);
    });
    if (rc) return rc;
    std::vector<uint64_t> have;
    for (size_t i = 0; i < world; i++)
        if (live[i]) have.insert(have.end(), parts.begin() + i * words, parts.begin() + (i + 1) * words);
    rc = zkb_group_sum(p->ctx[0], bs->group, have.data(), have.size() / words, out); // multiexp.tcc:433-438
    if (rc) p->err = zkb_last_error(p->ctx[0]);
    return rc;
}

} // extern "C"
