"""zkb_domain_select == libfqfft::get_evaluation_domain<Fr> (get_evaluation_domain.tcc:33-52) for every
size it is asked about here (host-only logic: runs without a GPU)."""
import pytest

import zklaim_b200 as zk


def test_domain_select_matches_reference(ref):
    sizes = list(range(2, 600)) + [29363, 58722, 146799, (1 << 20) - 3, (1 << 20) + 1, (1 << 22) - 11, 4177921,
                                   (1 << 27) + 5, 1 << 28]
    for s in sizes:
        assert zk.domain_select(s) == ref.domain_select(s), s


def test_domain_select_rejects_unsupported():
    for s in (0, 1, (1 << 28) + 1, 1 << 29):
        with pytest.raises(zk.ZkbError):
            zk.domain_select(s)
