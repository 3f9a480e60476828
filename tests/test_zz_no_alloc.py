"""aha_b200_scan_device_noalloc (include/aha_b200.h): the allocation-free entry point against aha_b200_scan_device on
the same device buffers, through the C ABI.  Bit-exact (integer work, no tolerance).  Collected last on purpose: the
entry point has no counterpart in the reference and a failure here must not hide the parity tests under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

import aha_b200
from aha_b200 import _lib, synth

pytestmark = pytest.mark.gpu

LIMIT = 64 << 20  # bytes of text per call the static scratch is sized for


def _batch(corpus, seed, n_docs, doc_bytes, ragged_seed):
    """documents of ragged lengths (some empty, some ending mid chunk) over the synthetic stream"""
    rng = np.random.default_rng(ragged_seed)
    lens = rng.integers(0, 2 * doc_bytes, n_docs)
    lens[rng.integers(0, n_docs, max(1, n_docs // 8))] = 0
    lens[-1] = 0  # trailing empty document
    off = np.zeros(n_docs + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    text = synth.gen_host(corpus, seed, 0, int(off[-1]))
    return text, off


def _compare(ac, text_h, off_h, cap_slack=0):
    import torch
    text = torch.from_numpy(text_h.copy()).cuda() if text_h.size else torch.zeros(0, dtype=torch.uint8, device="cuda")
    off = torch.from_numpy(off_h).cuda()
    want = ac.scan_device(text, off)  # the existing entry point (default device path)
    want_hits = want.hits_tensor().cpu().numpy().reshape(-1, 3)
    want_off = want.doc_hit_off()
    cap = want.n_hits + cap_slack
    hits = torch.full((max(cap, 1), 3), -7, dtype=torch.int32, device="cuda")
    doc_hit_off = torch.full((off_h.size,), -7, dtype=torch.int64, device="cuda")
    n_hits = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    ac.scan_device_noalloc(text, off, hits[:cap], doc_hit_off, n_hits)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the call changed the device's free memory"
    assert int(n_hits.item()) == want.n_hits
    assert np.array_equal(doc_hit_off.cpu().numpy(), want_off)
    assert np.array_equal(hits[:want.n_hits].cpu().numpy(), want_hits)
    return want.n_hits


@pytest.mark.parametrize("force_byte", [0, 1])
def test_no_alloc_matches_scan_device_twice_in_a_row(force_byte):
    L = _lib.lib()
    L.aha_b200_debug_set(0, force_byte)  # what the EXISTING entry point runs: character path / byte path
    try:
        ac = aha_b200.AC.compile(synth.patterns_zh_ngrams(2000) + synth.patterns_en_words(500))
        # first call: a large batch that fills most of the scratch; second call: a smaller, different one, so that
        # anything left over from the first call would show
        t1, o1 = _batch(synth.MIX, 11, 96, 256 << 10, 1)
        t2, o2 = _batch(synth.ZH, 12, 37, 40 << 10, 2)
        assert t1.size <= LIMIT
        n1 = _compare(ac, t1, o1)
        n2 = _compare(ac, t2, o2, cap_slack=5)
        assert n1 > 0 and n2 > 0 and n1 != n2
        _compare(ac, t1, o1)  # and back again
    finally:
        L.aha_b200_debug_set(0, 0)


def test_no_alloc_edge_cases():
    import torch
    ac = aha_b200.AC.compile([b"a", b"ab", b"bab", b"\xe4\xb8\xad"])
    # only empty documents, no documents at all, one byte, hits at the very first and last byte
    for docs in ([b"", b"", b""], [], [b"a"], [b"ab", b"", b"babab", b"", b""], [b"x" * 511 + b"ab", b"ab" * 700]):
        text = np.frombuffer(b"".join(docs), np.uint8)
        off = np.zeros(len(docs) + 1, np.int64)
        np.cumsum([len(d) for d in docs], out=off[1:])
        _compare(ac, text, off)
    # a hit buffer that is too small: the total is still reported, the records that fit are the first ones
    docs = [b"abab" * 100]
    text = torch.from_numpy(np.frombuffer(b"".join(docs), np.uint8).copy()).cuda()
    off = torch.tensor([0, text.numel()], dtype=torch.int64, device="cuda")
    want = ac.scan_device(text, off)
    full = want.hits_tensor().cpu().numpy().reshape(-1, 3)
    cap = 10
    hits = torch.full((cap + 4, 3), -7, dtype=torch.int32, device="cuda")
    dho = torch.zeros(2, dtype=torch.int64, device="cuda")
    n = torch.zeros(1, dtype=torch.int64, device="cuda")
    ac.scan_device_noalloc(text, off, hits[:cap], dho, n)
    torch.cuda.synchronize()
    assert int(n.item()) == want.n_hits > cap
    assert np.array_equal(hits[:cap].cpu().numpy(), full[:cap])
    assert (hits[cap:].cpu().numpy() == -7).all(), "wrote beyond the caller's capacity"


def test_no_alloc_rejects_batches_beyond_the_static_scratch():
    import torch
    ac = aha_b200.AC.compile([b"abc"])
    off = torch.tensor([0, LIMIT + 1], dtype=torch.int64, device="cuda")
    one = torch.zeros(8, dtype=torch.int64, device="cuda")
    L = _lib.lib()
    # rejected before anything is read: the text pointer only has to be non-null
    rc = L.aha_b200_scan_device_noalloc(ac._dfa, one.data_ptr(), off.data_ptr(), 1, LIMIT + 1, None, 0, one.data_ptr(),
                                        one.data_ptr(), None)
    assert rc == -1
    msg = L.aha_b200_last_error().decode()
    assert "64 MiB" in msg and str(LIMIT + 1) in msg
    # exactly the limit is accepted
    text = torch.zeros(LIMIT, dtype=torch.uint8, device="cuda")
    off = torch.tensor([0, LIMIT], dtype=torch.int64, device="cuda")
    dho = torch.full((2,), -1, dtype=torch.int64, device="cuda")
    n = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    ac.scan_device_noalloc(text, off, None, dho, n)
    torch.cuda.synchronize()
    assert int(n.item()) == 0 and dho.cpu().tolist() == [0, 0]
