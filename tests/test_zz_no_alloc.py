"""zkb_fft_noalloc (include/zkb.h): the entry point that allocates nothing -- all its scratch is a static device array
inside libzkb.so -- against the existing entry point zkb_domain_op on the same inputs, bit for bit (integer work: the
tolerance of every test here is exact equality), on two calls in a row (the calls share the array and the second must
not see the first one's twiddles or data), then an in-place call, and the size limit.  No counterpart in the reference.
Collected last on purpose."""
import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu


def _rand_fr(rng, n):
    a = rng.integers(0, 1 << 64, size=(n, 4), dtype=np.uint64)
    a[:, 3] &= np.uint64((1 << 60) - 1)  # < 2^252 < r: canonical residues
    return a


def test_fft_noalloc_equals_domain_op_on_two_calls_in_a_row(zctx):
    import torch
    lib = zk.load_library()
    kmax = int(lib.zkb_fft_noalloc_max_log2())
    assert kmax == 14
    rng = np.random.default_rng(2024)
    # buffers exist before the calls under test: the entry point itself allocates nothing
    bufs = {k: (torch.empty((1 << k, 4), dtype=torch.int64, device="cuda"), torch.empty((1 << k, 4), dtype=torch.int64, device="cuda"))
            for k in (kmax, 9, 3)}
    torch.cuda.synchronize()
    # two calls in a row: a forward transform at the maximum size, then an inverse one at another size
    for k, inverse in ((kmax, False), (9, True), (kmax, True), (3, False)):
        n = 1 << k
        a = _rand_fr(rng, n)
        want = zctx.domain_op(n, zk.NTT_IFFT if inverse else zk.NTT_FFT, a)
        d_in, d_out = bufs[k]
        d_in.copy_(torch.from_numpy(a.view(np.int64)))
        d_out.zero_()
        torch.cuda.synchronize()
        zctx.fft_noalloc(k, inverse, d_in.data_ptr(), d_out.data_ptr())
        got = d_out.cpu().numpy().view(np.uint64)
        assert np.array_equal(got, want), (k, inverse)
        assert np.array_equal(d_in.cpu().numpy().view(np.uint64), a)  # the input is left alone
    # in place, and the round trip iFFT(FFT(a)) == a through the no-alloc path alone
    k = 9
    a = _rand_fr(rng, 1 << k)
    d_in, _ = bufs[k]
    d_in.copy_(torch.from_numpy(a.view(np.int64)))
    torch.cuda.synchronize()
    zctx.fft_noalloc(k, False, d_in.data_ptr(), d_in.data_ptr())
    assert np.array_equal(d_in.cpu().numpy().view(np.uint64), zctx.domain_op(1 << k, zk.NTT_FFT, a))
    zctx.fft_noalloc(k, True, d_in.data_ptr(), d_in.data_ptr())
    assert np.array_equal(d_in.cpu().numpy().view(np.uint64), a)


def test_fft_noalloc_rejects_sizes_beyond_the_static_array(zctx):
    import torch
    d = torch.zeros((8, 4), dtype=torch.int64, device="cuda")
    for bad in (0, 15, 28):
        with pytest.raises(zk.ZkbError) as e:
            zctx.fft_noalloc(bad, False, d.data_ptr(), d.data_ptr())
        assert e.value.code == zk.ERR_ARG
    with pytest.raises(zk.ZkbError):
        zctx.fft_noalloc(3, False, 0, d.data_ptr())
