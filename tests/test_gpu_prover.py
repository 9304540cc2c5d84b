"""Full-prover parity: zkb_prove / zkb_prove_batch / zkb_witness_map against the reference prover
(r1cs_gg_ppzksnark.tcc:390-505, r1cs_to_qap.tcc:205-334) run with the same injected (r, s):
  * coefficients_for_H bit-identical,
  * proof elements (A, B, C) bit-identical after the reference's own affine normalisation, and the
    serialised proof (r1cs_gg_ppzksnark.tcc:168-189) byte-identical,
  * the reference verifier (r1cs_gg_ppzksnark.tcc:523-576) accepts the GPU proof.
Circuits: the zklaim credential circuit of zklaim/main.c (1 payload; 5 payloads = step domain),
libsnark's generate_r1cs_example_with_field_input (dense witnesses; basic and step domains).
"""
import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu


def _norm_proof(ref, pr):
    a = ref.to_affine_jac(zk.G1, pr[None, 0:12])[0]
    b = ref.to_affine_jac(zk.G2, pr[None, 12:36])[0]
    c = ref.to_affine_jac(zk.G1, pr[None, 36:48])[0]
    return np.concatenate([a, b, c])


def _check_against_reference(zctx, ref, circ, key, w, seed):
    d = circ.dims()
    r = ref.sha512_rng(seed, 1)[0]
    s = ref.sha512_rng(seed + 1, 1)[0]
    exp = circ.prove(w, r, s)
    got, H = zctx.prove(key, w, r, s, want_H=True)
    assert np.array_equal(H, exp["H"]), "coefficients_for_H differ"
    assert np.array_equal(_norm_proof(ref, got), _norm_proof(ref, exp["proof"])), "proof elements differ"
    assert ref.proof_bytes(got) == ref.proof_bytes(exp["proof"])
    assert circ.verify(w[: d["num_inputs"]], got), "reference verifier rejects the GPU proof"
    return got


@pytest.fixture(scope="module")
def zklaim1(zctx, ref):
    circ = ref.Circuit.zklaim(1, 1)
    assert circ.is_satisfied()
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    yield circ, key
    key.free()
    circ.free()


def test_zklaim_circuit_proof_bit_exact(zctx, ref, zklaim1):
    circ, key = zklaim1
    d = circ.dims()
    assert key.domain() == (d["m"], d["kind"]) == (32768, zk.DOMAIN_BASIC_RADIX2)
    w = circ.witness()
    _check_against_reference(zctx, ref, circ, key, w, 1000)
    # witness map alone
    Hexp, _ = circ.witness_map(w)
    assert np.array_equal(zctx.witness_map(key, w), Hexp)


def test_zklaim_batch_of_distinct_witnesses(zctx, ref, zklaim1):
    circ, key = zklaim1
    d = circ.dims()
    n = 5
    ws = np.stack([ref.zklaim_witness(1, 100 + i, d["num_variables"]) for i in range(n)])
    r = ref.sha512_rng(5000, n)
    s = ref.sha512_rng(6000, n)
    got = zctx.prove_batch(key, ws, r, s)
    for i in range(n):
        exp = circ.prove(ws[i], r[i], s[i], want_H=False)
        assert np.array_equal(_norm_proof(ref, got[i]), _norm_proof(ref, exp["proof"])), i
        assert circ.verify(ws[i][: d["num_inputs"]], got[i])


def test_contexts_of_one_device_share_a_key(zctx, ref, zklaim1):
    """include/zkb.h, zkb_init: several contexts on one device, one host thread each, all proving under ONE uploaded key
    at the same time (bench.py's proof streams).  Every stream's proofs must equal the reference prover's for the same
    (w, r, s), whatever the other streams are doing."""
    import threading
    circ, key = zklaim1
    d = circ.dims()
    streams, rounds, n = 3, 3, 4
    ctxs = [zk.Context(0) for _ in range(streams)]
    ws = [np.stack([ref.zklaim_witness(1, 900 + 10 * j + i, d["num_variables"]) for i in range(n)]) for j in range(streams)]
    rs = [ref.sha512_rng(9100 + j, n) for j in range(streams)]
    ss = [ref.sha512_rng(9200 + j, n) for j in range(streams)]
    out = [[None] * rounds for _ in range(streams)]
    err = []

    def body(j):
        try:
            for k in range(rounds):
                out[j][k] = ctxs[j].prove_batch(key, ws[j], rs[j], ss[j])
        except Exception as e:
            err.append(e)

    th = [threading.Thread(target=body, args=(j,)) for j in range(streams)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not err, err
    for j in range(streams):
        want = [_norm_proof(ref, circ.prove(ws[j][i], rs[j][i], ss[j][i], want_H=False)["proof"]) for i in range(n)]
        for k in range(rounds):
            for i in range(n):
                assert np.array_equal(_norm_proof(ref, out[j][k][i]), want[i]), (j, k, i)
    for c in ctxs:
        c.close()


def test_pipelined_batches_equal_plain_batches(zctx, ref, zklaim1):
    """zkb_prove_batch_submit / _collect (inputs of batch k + 1 travel while batch k computes, two staging slots):
    same proofs as zkb_prove_batch, in submission order; an unsatisfying batch is reported by ITS collect only; a
    third submit without a collect and a collect with nothing pending are argument errors."""
    circ, key = zklaim1
    d = circ.dims()
    sizes = [3, 1, 4, 2]
    ws = [np.stack([ref.zklaim_witness(1, 300 + 10 * b + i, d["num_variables"]) for i in range(n)]) for b, n in enumerate(sizes)]
    rs = [ref.sha512_rng(7000 + b, n) for b, n in enumerate(sizes)]
    ss = [ref.sha512_rng(8000 + b, n) for b, n in enumerate(sizes)]
    want = [zctx.prove_batch(key, ws[b], rs[b], ss[b]) for b in range(len(sizes))]
    got = []
    zctx.prove_batch_submit(key, ws[0], rs[0], ss[0])
    for b in range(len(sizes)):
        if b + 1 < len(sizes):
            zctx.prove_batch_submit(key, ws[b + 1], rs[b + 1], ss[b + 1])
        got.append(zctx.prove_batch_collect())
    def same(a, b):  # as group elements: the Jacobian representative depends on the order pairs reach a bucket
        return all(np.array_equal(_norm_proof(ref, x), _norm_proof(ref, y)) for x, y in zip(a, b))

    for b in range(len(sizes)):
        assert got[b].shape == want[b].shape and same(got[b], want[b]), b
    with pytest.raises(zk.ZkbError) as e:
        zctx.prove_batch_collect()
    assert e.value.code == zk.ERR_ARG
    bad = ws[1].copy()
    bad[0, d["num_inputs"] + 7] = ws[1][0, d["num_inputs"] + 8] ^ np.uint64(1)
    zctx.prove_batch_submit(key, ws[0], rs[0], ss[0])
    zctx.prove_batch_submit(key, bad, rs[1], ss[1])
    with pytest.raises(zk.ZkbError) as e:
        zctx.prove_batch_submit(key, ws[1], rs[1], ss[1])  # a third batch before any collect
    assert e.value.code == zk.ERR_ARG
    assert same(zctx.prove_batch_collect(), want[0])
    with pytest.raises(zk.ZkbError) as e:
        zctx.prove_batch_collect()
    assert e.value.code == zk.ERR_UNSAT
    zctx.prove_batch_submit(key, ws[2], rs[2], ss[2])  # the pipeline is usable afterwards
    assert same(zctx.prove_batch_collect(), want[2])


def test_key_tune_keeps_proofs_bit_exact(zctx, ref):
    """zkb_key_tune re-plans the MSM windows from a sample witness: narrower windows for the 0 / 1-dominated
    A / B queries, proofs unchanged (single proof, batch, and a dense witness under the tuned key)."""
    circ = ref.Circuit.zklaim(1, 1)
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    d = circ.dims()
    w = circ.witness()
    before = key.windows()
    key.tune(w)
    after = key.windows()
    assert after[0] < before[0] and after[1] < before[1] and after[3] < before[3], (before, after)
    assert after[2] >= before[2] - 1  # the H query stays dense
    _check_against_reference(zctx, ref, circ, key, w, 4242)
    n = 3
    ws = np.stack([ref.zklaim_witness(1, 900 + i, d["num_variables"]) for i in range(n)])
    r = ref.sha512_rng(5100, n)
    s = ref.sha512_rng(6100, n)
    got = zctx.prove_batch(key, ws, r, s)
    for i in range(n):
        exp = circ.prove(ws[i], r[i], s[i], want_H=False)
        assert ref.proof_bytes(got[i]) == ref.proof_bytes(exp["proof"]), i
    key.tune(w)  # idempotent
    assert key.windows() == after
    key.free()
    circ.free()


def test_key_tune_dense_witness(zctx, ref):
    circ = ref.Circuit.example(1013, 10, seed=5)
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    key.tune(circ.witness())
    _check_against_reference(zctx, ref, circ, key, circ.witness(), 99)
    key.free()
    circ.free()


def test_blinding_edge_values(zctx, ref):
    """(r, s) = (0, 1), (1, 0), (r - 1, 2), (17, 16): the windowed scalar multiplications s * g1_A and r * g1_B
    (r1cs_gg_ppzksnark.tcc:495) with empty / single / all-ones / colliding 4-bit windows."""
    circ = ref.Circuit.example(253, 5, seed=3)
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    w = circ.witness()
    zero = np.zeros(4, dtype=np.uint64)
    one = ref.field_const(zk.FR, 1)
    small = lambda v: ref.from_plain(zk.FR, np.array([[v, 0, 0, 0]], dtype=np.uint64))[0]
    minus1 = ref.field_op(zk.FR, 1, zero[None], one[None])[0]
    for r, s in ((zero, one), (one, zero), (minus1, small(2)), (small(17), small(16)), (zero, zero)):
        exp = circ.prove(w, r, s, want_H=False)
        got = zctx.prove(key, w, r, s)
        assert ref.proof_bytes(got) == ref.proof_bytes(exp["proof"])
    key.free()
    circ.free()


def test_batch_larger_than_one_chunk(zctx, ref):
    """zkb_prove_batch cuts a batch into chunks of at most 64 proofs: 70 proofs of a small circuit with distinct
    (r, s); proofs 0, 63, 64 and 69 (both sides of the chunk boundary) against the reference."""
    circ = ref.Circuit.example(253, 5, seed=3)
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    w = circ.witness()
    n = 70
    ws = np.stack([w] * n)
    r = ref.sha512_rng(7000, n)
    s = ref.sha512_rng(8000, n)
    got = zctx.prove_batch(key, ws, r, s)
    for i in (0, 63, 64, 69):
        exp = circ.prove(w, r[i], s[i], want_H=False)
        assert ref.proof_bytes(got[i]) == ref.proof_bytes(exp["proof"]), i
    key.free()
    circ.free()


def test_sha_chain_circuit(zctx, ref):
    """BASELINE configs[4] in miniature: two chained sha256_two_to_one_hash_gadget blocks (54 560 constraints, 256 public
    inputs, domain 2^16), key tuned; the full-size run (153 blocks, 2^22) is scripts/s22_prove.py, profiles/r01_s22_full_prove.txt"""
    circ = ref.Circuit.sha_chain(2, 11)
    assert circ.is_satisfied()
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    key.tune(circ.witness())
    _check_against_reference(zctx, ref, circ, key, circ.witness(), 515)
    key.free()
    circ.free()


@pytest.mark.parametrize("world", [2, 3, 8])
def test_one_proof_sharded_over_ranks(zctx, ref, world):
    """one proof over `world` shards (one GPU plays every rank): every shard holds its range of the four
    queries, computes its shares, zkb_prove_combine sums and assembles -> bytes identical to the reference proof.
    A whole key refuses zkb_prove_partial-less use of a shard key."""
    circ = ref.Circuit.example(1013, 10, seed=21)
    circ.setup()
    kd = circ.export_key()
    w = circ.witness()
    r = ref.sha512_rng(31, 1)[0]
    s = ref.sha512_rng(32, 1)[0]
    exp = circ.prove(w, r, s, want_H=False)
    parts = []
    for rank in range(world):
        key = zctx.upload_key(kd, rank, world)
        if rank == 1:
            key.tune(w)  # tuning a shard must not change its share
            with pytest.raises(zk.ZkbError):
                zctx.prove(key, w, r, s)
        parts.append(zctx.prove_partial(key, w, r, s))
        key.free()
    got = zctx.prove_combine(np.stack(parts), r, s)
    assert ref.proof_bytes(got) == ref.proof_bytes(exp["proof"])
    assert circ.verify(w[:10], got)
    circ.free()


def test_sharded_zklaim_proof_more_ranks_than_some_ranges(zctx, ref, zklaim1):
    """the zklaim circuit over 4 shards through sharding.sharded_prove's building blocks; plus world = 1 == zkb_prove"""
    from zklaim_b200 import sharding
    circ, key1 = zklaim1
    kd = circ.export_key()
    w = circ.witness()
    r = ref.sha512_rng(41, 1)[0]
    s = ref.sha512_rng(42, 1)[0]
    whole = zctx.prove(key1, w, r, s)
    parts = []
    for rank in range(4):
        key = zctx.upload_key(kd, rank, 4)
        parts.append(zctx.prove_partial(key, w, r, s))
        key.free()
    got = zctx.prove_combine(np.stack(parts), r, s)
    assert ref.proof_bytes(got) == ref.proof_bytes(whole)
    solo = sharding.sharded_prove(zctx, key1, w, r, s)  # no process group: a world of one
    assert ref.proof_bytes(solo) == ref.proof_bytes(whole)


def test_unsatisfied_witness_is_reported(zctx, ref, zklaim1):
    circ, key = zklaim1
    w = circ.witness().copy()
    w[100] = ref.sha512_rng(1, 1)[0]
    r = ref.sha512_rng(1, 1)[0]
    with pytest.raises(zk.ZkbError) as e:
        zctx.prove(key, w, r, r)
    assert e.value.code == zk.ERR_UNSAT


@pytest.mark.parametrize("nc,ni", [(20, 3), (1013, 10), (2048 - 11, 10), (2048 + 256 - 11, 10), (5000, 100), (20000, 10)])
def test_example_circuits_dense_witness(zctx, ref, nc, ni):
    circ = ref.Circuit.example(nc, ni, seed=nc)
    assert circ.is_satisfied()
    circ.setup()
    d = circ.dims()
    key = zctx.upload_key(circ.export_key())
    assert key.domain() == (d["m"], d["kind"])
    _check_against_reference(zctx, ref, circ, key, circ.witness(), 77 + nc)
    key.free()
    circ.free()


def test_zklaim_five_payloads_step_domain(zctx, ref):
    circ = ref.Circuit.zklaim(5, 3)
    circ.setup()
    d = circ.dims()
    assert d["kind"] == zk.DOMAIN_STEP_RADIX2 and d["m"] == 147456
    key = zctx.upload_key(circ.export_key())
    _check_against_reference(zctx, ref, circ, key, circ.witness(), 31)
    key.free()
    circ.free()


@pytest.mark.parametrize("nc,kind", [((1 << 18) - 11, zk.DOMAIN_BASIC_RADIX2), ((1 << 17) + (1 << 14) - 11, zk.DOMAIN_STEP_RADIX2)])
def test_witness_map_at_scale_bit_exact(zctx, ref, nc, kind):
    """BASELINE.json configs[3] (r1cs_to_qap_witness_map, r1cs_to_qap.tcc:205-334) on
    generate_r1cs_example_with_field_input(m - 11, 10): all m + 1 H coefficients == the reference, at 2^18 and at a
    step-radix-2 size.  The last constraint of these instances has ~m terms (the giant-row SpMV path).
    The witness map needs the constraint system only, so the queries of the key are one generator each."""
    mt = ref.available(mt=True)
    circ = ref.Circuit.example(nc, 10, seed=nc, mt=mt)
    d = circ.dims()
    assert d["kind"] == kind
    w = circ.witness()
    g1 = ref.bases(zk.G1, 1, 1)[0]
    g2 = ref.bases(zk.G2, 1, 1)[0]
    nv, ni, m = d["num_variables"], d["num_inputs"], d["m"]
    kd = dict(num_constraints=d["num_constraints"], num_variables=nv, num_inputs=ni,
              csr_A=circ.csr(0), csr_B=circ.csr(1), csr_C=circ.csr(2),
              alpha_g1=g1, beta_g1=g1, delta_g1=g1, beta_g2=g2, delta_g2=g2, A_query=np.tile(g1, (nv + 1, 1)),
              B_index=np.zeros(1, dtype=np.uint32), B_g1=g1[None], B_g2=g2[None], H_query=np.tile(g1, (m - 1, 1)),
              L_query=np.tile(g1, (nv - ni, 1)))
    key = zctx.upload_key(kd)
    H = zctx.witness_map(key, w)
    Href, _ = circ.witness_map(w)
    assert H.shape == Href.shape and np.array_equal(H, Href)
    w2 = w.copy()
    w2[ni + 5] = w[ni + 6]  # an unsatisfying assignment is reported, not mapped
    if not np.array_equal(w2, w):
        with pytest.raises(zk.ZkbError) as e:
            zctx.witness_map(key, w2)
        assert e.value.code == zk.ERR_UNSAT
    key.free()
    circ.free()


@pytest.mark.parametrize("world", [2, 3, 5])
def test_one_proof_with_the_witness_map_split_by_polynomial(zctx, ref, world):
    """SURVEY 8(e) optional row: ranks 0 / 1 / 2 compute the coset evaluations of A / B / C (zkb_prove_split_begin), one
    rank divides and transforms back (zkb_prove_split_combine), every rank finishes its share from the H coefficients
    (zkb_prove_split_end).  One GPU plays every rank (a context per rank, device buffers through torch): the proof bytes
    equal the reference's, on the zklaim circuit and on a step-domain example; the checking rank reports an
    unsatisfying witness."""
    import torch
    for circ in (ref.Circuit.zklaim(1, 1), ref.Circuit.example(2048 + 256 - 11, 10, seed=9)):
        circ.setup()
        kd = circ.export_key()
        w = circ.witness()
        r, s = ref.sha512_rng(51, 1)[0], ref.sha512_rng(52, 1)[0]
        exp = circ.prove(w, r, s, want_H=False)
        m = circ.dims()["m"]
        for witness, unsat in ((w, False), (None, True)):
            if unsat:
                witness = w.copy()
                witness[circ.dims()["num_inputs"] + 3] = ref.sha512_rng(7, 1)[0]
            ctxs = [zk.Context(0) for _ in range(world)]
            keys = [ctxs[k].upload_key(kd, k, world) for k in range(world)]
            abc = torch.empty((3, m, 4), dtype=torch.int64, device="cuda")
            H = torch.empty((m, 4), dtype=torch.int64, device="cuda")
            for k in range(world):
                ctxs[k].prove_split_begin(keys[k], witness, r, s, k if k < 3 else -1, check=(k == world - 1),
                                          d_part=abc[k].data_ptr() if k < 3 else 0)
            if world == 2:  # rank 0 takes A and C (zkb_prove_split_chain)
                ctxs[0].prove_split_chain(keys[0], 2, abc[2].data_ptr())
            ctxs[0].prove_split_combine(keys[0], abc.data_ptr(), H.data_ptr())
            parts, failed = [], 0
            for k in range(world):
                try:
                    parts.append(ctxs[k].prove_split_end(keys[k], H.data_ptr()))
                except zk.ZkbError as e:
                    assert e.code == zk.ERR_UNSAT and k == world - 1
                    failed += 1
            if unsat:
                assert failed == 1
            else:
                got = ctxs[0].prove_combine(np.stack(parts), r, s)
                assert ref.proof_bytes(got) == ref.proof_bytes(exp["proof"])
            for k in range(world):
                keys[k].free()
                ctxs[k].close()
        circ.free()
