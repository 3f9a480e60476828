"""GPU parity tests: the CUDA path, called through the C ABI (aha_b200.AC -> libaha_b200.so), against the
CPU oracle on the same inputs.  Bit-exact: integer/byte work, no tolerance."""
import os
import random

import numpy as np
import pytest

import aha_b200
from aha_b200 import synth
from oracle import oracle as O
from tests.test_oracle_kat import _rand_keys

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["default", "byte", "two-pass", "span"])
def scan_impl(request):
    """every parity test runs against the default device path (character automaton in shared memory where the
    dictionary allows: walk / scan / expand kernels, else the fused byte-automaton kernel), the fused kernel forced to
    the byte automaton (once with staged tiles / the default span, once with 64-byte spans read from global memory), and
    the two-pass checker kernels (the path of keys > 1024 bytes); the heavy full-size tests run on the default path only.  The knobs are test hooks of the C ABI (aha_b200_debug_set), not product switches."""
    from aha_b200 import _lib
    heavy = request.node.get_closest_marker("heavy") is not None
    if heavy and request.param != "default":
        pytest.skip("full-size test: default path only")
    L = _lib.lib()
    L.aha_b200_debug_set(0, 1 if request.param in ("byte", "span") else 0)
    L.aha_b200_debug_set(1, 1 if request.param == "two-pass" else 0)
    # "span": the byte kernel's span mode with 64-byte spans, so that batches of small documents (average >= 256 bytes)
    # exercise it too; the other runs keep the default span, which only engages for documents of kilobytes
    L.aha_b200_debug_set(4, 64 if request.param == "span" else -1)
    yield request.param
    L.aha_b200_debug_set(0, 0)
    L.aha_b200_debug_set(1, 0)
    L.aha_b200_debug_set(4, -1)


def as_tuples(h):
    return [(int(x["start"]), int(x["end"]), int(x["value"])) for x in h]


def oracle_batch(oac, docs, mode=0, sep=None, sep_size=256):
    hits, off = [], [0]
    for d in docs:
        h = oac.scan(d, mode & ~aha_b200.SEP, sep=sep, sep_size=sep_size)
        hits.append(h)
        off.append(off[-1] + len(h))
    return (np.concatenate(hits) if hits else np.zeros(0, O.HIT_DTYPE)), np.array(off, np.int64)


def check_batch(keys, docs, mode=0, sep=None, sep_size=256):
    g = aha_b200.AC.compile(keys)
    o = O.AC.compile(keys)
    res = g.scan_batch(docs, mode, sep, sep_size)
    want, want_off = oracle_batch(o, docs, mode, sep, sep_size)
    got = res.hits()
    assert res.n_hits == len(want)
    assert np.array_equal(res.doc_hit_off(), want_off)
    assert got.tobytes() == want.tobytes()
    return res


def test_spec_vectors():
    m = aha_b200.AC.compile(["我", "我是", "是中"])  # spec/ac_spec.cr:5-12
    assert [(h[1], h[2]) for h in as_tuples(m.match("我是中国人"))] == [(1, 0), (2, 1), (3, 2)]
    assert as_tuples(m.match("我是中国人".encode())) == [(0, 3, 0), (0, 6, 1), (3, 9, 2)]
    m = aha_b200.AC.compile(["a", "aa"])  # spec/ac_spec.cr:25-34
    assert [(h[1], h[2]) for h in as_tuples(m.match("a aaa", sep=[" "]))] == [(1, 0)]
    t = aha_b200.Cedar()  # spec/ac_longest_match_spec.cr:5-18
    for s in ["Ruby", "ruby", "rub"]:
        t.insert(s)
    ac = aha_b200.AC.compile(t)
    assert [(s, e, ac[v]) for s, e, v in as_tuples(ac.match_longest("Ruby on rub"))] == [(0, 4, "Ruby"), (8, 11, "rub")]
    t = aha_b200.Cedar()  # spec/ac_longest_match_spec.cr:20-33
    for s in ["Ruby", "ruby", "uby "]:
        t.insert(s)
    ac = aha_b200.AC.compile(t)
    assert [(s, e, ac[v]) for s, e, v in as_tuples(ac.match_longest("ruby ", True))] == [(0, 4, "ruby"), (1, 5, "uby ")]


def test_quirks_q1_q5_q7():
    assert as_tuples(aha_b200.AC.compile(["abcd", "bc"]).match(b"abcd")) == [(0, 4, 0)]
    assert as_tuples(aha_b200.AC.compile(["xab", "abz", "b"]).match(b"xab")) == [(0, 3, 0)]
    assert as_tuples(aha_b200.AC.compile(["ab", "b"]).match_longest(b"abb")) == [(0, 2, 0)]
    m = aha_b200.AC.compile(["a"])
    assert m.scan_batch([b"\xe4a\xe4"], 0, sep=[" "], sep_size=128).n_hits == 1
    assert m.scan_batch([b"\xe4a\xe4"], 0, sep=[" "], sep_size=256).n_hits == 0


def test_q4_stale_end_flag_vector():
    """the hand-derived Q4 vector of tests/test_oracle_kat.py (stale END flag on the interior node "bc"), on the GPU"""
    g = aha_b200.AC.compile(["bba", "bca"])
    assert as_tuples(g.match_longest(b"bcbba")) == []
    assert as_tuples(g.match_longest(b"bcbba", True)) == [(2, 5, 0)]
    assert as_tuples(g.match(b"bcbba")) == [(2, 5, 0)]
    assert as_tuples(aha_b200.AC.compile(["bca", "bba"]).match_longest(b"bcbba")) == [(2, 5, 1)]


def test_errors():
    with pytest.raises(aha_b200.AhaError, match="appear twice"):
        aha_b200.AC.compile(["a", "a"])
    with pytest.raises(aha_b200.AhaError, match="sep BitArray size > 256"):
        aha_b200.AC.compile(["a"]).scan_batch([b"a"], 0, sep=[" "], sep_size=300)


def _docs(rng, alphabet, n_docs, max_len):
    docs = []
    for _ in range(n_docs):
        L = rng.choice([0, 0, 1, 2, rng.randint(0, max_len), max_len])
        docs.append(bytes(rng.choice(alphabet) for _ in range(L)))
    return docs


@pytest.mark.parametrize("seed", range(8))
def test_random_ragged_batches(seed):
    rng = random.Random(seed)
    alphabet = [b"ab", b"abc", b"abcdefgh", bytes(range(1, 256))][seed % 4]
    keys = _rand_keys(rng, rng.choice([5, 50, 500, 3000]), alphabet, rng.choice([2, 4, 9]))
    docs = _docs(rng, alphabet, rng.choice([1, 7, 40]), rng.choice([10, 700, 5000]))
    check_batch(keys, docs)
    check_batch(keys, docs, sep=[alphabet[0]])
    check_batch(keys, docs, aha_b200.LONGEST)
    check_batch(keys, docs, aha_b200.LONGEST | aha_b200.INTERSECTABLE)


def test_empty_inputs():
    keys = ["ab", "b"]
    check_batch(keys, [])
    check_batch(keys, [b""])
    check_batch(keys, [b"", b"", b"ab", b"", b""])
    check_batch(keys, [b"", b""], aha_b200.LONGEST)


def test_char_units_utf8():
    rng = random.Random(5)
    zh = "的一是在不了有和人这中大为上个国我以要他时来用们生到作地于出就分对成会可主发年动同工也能下过子说产种面而方后多定行学法所民得经"
    keys = list(dict.fromkeys("".join(rng.choice(zh) for _ in range(rng.randint(1, 3))) for _ in range(300)))
    keys += ["ab", "b c", "é", "日本"]
    docs = ["".join(rng.choice(zh + "ab cé日本") for _ in range(rng.randint(0, 2000))) for _ in range(12)]
    check_batch(keys, docs, aha_b200.UNITS_CHAR)
    check_batch(keys, docs, aha_b200.UNITS_CHAR, sep=[" "])
    check_batch(keys, docs, aha_b200.UNITS_CHAR | aha_b200.LONGEST)
    check_batch(keys, docs, aha_b200.UNITS_CHAR | aha_b200.LONGEST | aha_b200.INTERSECTABLE)


def _split_docs(text, size):
    return [text[i:i + size].tobytes() for i in range(0, len(text), size)]


@pytest.mark.parametrize("cfg", ["c1_en_1k", "c2_zh_10k", "c3_mix_100k"])
def test_scaled_baseline_configs(cfg):
    if cfg == "c1_en_1k":
        keys, corpus, seed = synth.patterns_en_words(1000), synth.EN, synth.SEED_TEXT
    elif cfg == "c2_zh_10k":
        keys, corpus, seed = synth.patterns_zh_ngrams(10000), synth.ZH, synth.SEED_TEXT + 1
    else:
        keys, corpus, seed = synth.patterns_mix(100000), synth.MIX, synth.SEED_TEXT + 2
    D = synth.doc_bytes(1 << 20)
    text = synth.gen_host(corpus, seed, 0, 8 * D)
    docs = _split_docs(text, D)
    res = check_batch(keys, docs)
    assert res.n_hits > 0
    check_batch(keys, docs[:2], aha_b200.UNITS_CHAR)


def test_device_api_and_generator_parity():
    import torch
    keys = synth.patterns_zh_ngrams(2000)
    g = aha_b200.AC.compile(keys)
    o = O.AC.compile(keys)
    D = synth.doc_bytes(1 << 18)
    n_docs = 16
    dev = synth.gen_device(synth.ZH, synth.SEED_TEXT + 1, 0, n_docs * D)
    host = synth.gen_host(synth.ZH, synth.SEED_TEXT + 1, 0, n_docs * D)
    assert np.array_equal(dev.cpu().numpy(), host)
    doc_off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
    res = g.scan_device(dev, doc_off)
    got = res.hits_tensor().cpu().numpy()
    want, want_off = oracle_batch(o, _split_docs(host, D))
    assert np.array_equal(res.doc_hit_off(), want_off)
    assert got.astype("<i4").tobytes() == want.tobytes()
    assert res.scan_ms > 0


def test_blob_roundtrip_and_save_load(tmp_path):
    keys = ["我", "我是", "是中", "abc"]
    g = aha_b200.AC.compile(keys)
    g2 = aha_b200.AC.from_blob(g.blob())
    assert as_tuples(g2.match("我是中国人abc")) == as_tuples(g.match("我是中国人abc"))
    p = str(tmp_path / "aha.bin")
    g.save(p)  # spec/ac_spec.cr:14-23
    g3 = aha_b200.AC.load(p)
    assert [(h[1], h[2]) for h in as_tuples(g3.match("我是中国人"))] == [(1, 0), (2, 1), (3, 2)]
    assert g3[1] == "我是"


def _trie3():
    t = aha_b200.Cedar()
    for s in ["Ruby", "ruby", "rub"]:
        t.insert(s)
    return t


def test_cedar_match_longest_spec_vectors():  # spec/cedar_match_spec.cr:4-30, 51-63
    t = _trie3()
    got = [(s, e, t[v]) for s, e, v in as_tuples(t.match_longest("Ruby on rub", False))]
    assert got == [(0, 4, "Ruby"), (8, 11, "rub")]
    got = sorted((s, e, t[v]) for s, e, v in as_tuples(t.match_longest("Ruby on ruby", True)))
    assert got == [(0, 4, "Ruby"), (8, 12, "ruby")]
    got = sorted((s, e, t[v]) for s, e, v in as_tuples(t.match_longest("Ruby rRuby on rubyx ruby", True, sep=[" "])))
    assert got == [(0, 4, "Ruby"), (20, 24, "ruby")]


def test_cedar_wide_separator_is_refused():
    t = aha_b200.Cedar()
    t.insert("ruby")
    with pytest.raises(aha_b200.AhaError):
        t.match_longest("ruby，ruby", sep=["，"])  # cedar.cr:1071-1073 would accept it; a 256-bit set cannot hold it


def test_cedar_gsub_spec_vectors():  # spec/cedar_match_spec.cr:32-49, 65-81
    t = _trie3()
    names = {"ruby": "crystal", "Ruby": "Crystal"}
    block = lambda m: names.get(t[m[2]], "Unknown")  # noqa: E731
    assert t.gsub("Ruby on ruby rub", True, block=block) == "Crystal on crystal Unknown"
    assert t.gsub("Ruby rRuby on rubyx ruby", True, sep=[" "], block=block) == "Crystal rRuby on rubyx crystal"


@pytest.mark.parametrize("seed", range(4))
def test_cedar_gsub_random(seed):
    rng = random.Random(100 + seed)
    letters = "abAB cCdé日本Я" if seed % 2 else "abAB "
    words = list(dict.fromkeys("".join(rng.choice(letters.replace(" ", "")) for _ in range(rng.randint(1, 4))) for _ in range(60)))
    gt, ot = aha_b200.Cedar(), O.Cedar()
    for w in words:
        assert gt.insert(w) == ot.insert(w)
    docs = ["".join(rng.choice(letters) for _ in range(rng.randint(0, 120))) for _ in range(12)] + ["", " ", "a", "  a  "]
    block = lambda m: "<%d:%d>" % (m[2], m[1] - m[0])  # noqa: E731
    for icase in (False, True):
        for sep in (None, [" "]):
            got = gt.gsub_batch(docs, icase, sep, block)
            want = [ot.gsub(d, icase, sep, block) for d in docs]
            assert got == want


@pytest.mark.parametrize("seed", range(6))
def test_cedar_match_longest_random(seed):
    rng = random.Random(seed)
    letters = "abAB cCdé日本Я" if seed % 2 else "abAB "
    words = list(dict.fromkeys("".join(rng.choice(letters.replace(" ", "")) for _ in range(rng.randint(1, 5))) for _ in range(200)))
    words += ["a b", "ab a"]
    words = list(dict.fromkeys(words))
    gt, ot = aha_b200.Cedar(), O.Cedar()
    for w in words:
        assert gt.insert(w) == ot.insert(w)
    docs = ["".join(rng.choice(letters) for _ in range(rng.randint(0, 300))) for _ in range(30)] + ["", " ", "a"]
    for icase in (False, True):
        for sep in (None, [" "]):
            res = gt.match_longest_batch(docs, icase, sep)
            want = [ot.match_longest(d, icase, sep) for d in docs]
            off = np.cumsum([0] + [len(w) for w in want])
            assert np.array_equal(res.doc_hit_off(), off)
            assert res.hits().tobytes() == np.concatenate(want).tobytes()


@pytest.mark.parametrize("seed", range(4))
def test_cedar_ignore_case_variant_order(seed):
    """ignore_case runs over the case-folded trie: many keys that differ only in case (several original nodes behind one
    folded node), so the reference's choice -- the first candidate of the greatest length in frontier order, i.e. "as the
    text stands" before "case swapped", character by character -- decides nearly every hit.  ASCII, Latin-1, Greek and
    Cyrillic pairs, uncased characters in between, keys up to 64 characters."""
    rng = random.Random(400 + seed)
    base = ["ab", "aB", "Ab", "AB", "abc", "ABC", "Abc", "aBc", "éa", "Éa", "éA", "ÉA", "жд", "Жд", "жД", "αβ", "Αβ", "αΒ", "a日b", "A日B", "a日B",
            "x" * 40 + "Y" * 24, "X" * 64, "q" * 63 + "Q"]
    alpha = "abABcC éÉжЖдДαΑβΒ日"
    words = base + ["".join(rng.choice(alpha.replace(" ", "")) for _ in range(rng.randint(1, 4))) for _ in range(300)]
    words = list(dict.fromkeys(words))
    rng.shuffle(words)
    gt, ot = aha_b200.Cedar(), O.Cedar()
    for w in words:
        assert gt.insert(w) == ot.insert(w)
    docs = ["".join(rng.choice(alpha) for _ in range(rng.randint(0, 400))) for _ in range(40)]
    docs += ["AB ab Ab aB ABC abc aBC", "ÉA éa Éa", "ЖД жд", "ΑΒ αβ", "A日b a日b", "x" * 64 + " " + "X" * 64 + " " + "x" * 40 + "y" * 24, "Q" * 64 + "q" * 64, ""]
    for sep in (None, [" "]):
        res = gt.match_longest_batch(docs, True, sep)
        want = [ot.match_longest(d, True, sep) for d in docs]
        off = np.cumsum([0] + [len(w) for w in want])
        assert np.array_equal(res.doc_hit_off(), off)
        assert res.hits().tobytes() == np.concatenate(want).tobytes()


def test_cedar_match_longest_long_documents():
    """documents of several windows of the warp-per-document kernel: multi-byte characters and matches across window
    boundaries, keys that contain the separator, with and without sep / ignore_case, one very long document"""
    rng = random.Random(77)
    letters = "abAB cCdé日本Я "
    words = list(dict.fromkeys("".join(rng.choice(letters.replace(" ", "")) for _ in range(rng.randint(1, 6))) for _ in range(300)))
    words += ["a b", "ab a", "日本 é", "b  b"]
    words = list(dict.fromkeys(words))
    gt, ot = aha_b200.Cedar(), O.Cedar()
    for w in words:
        assert gt.insert(w) == ot.insert(w)
    docs = ["".join(rng.choice(letters) for _ in range(n)) for n in (1023, 1024, 1025, 2047, 3000, 5000, 40000)] + ["", "a"]
    for icase in (False, True):
        for sep in (None, [" "]):
            res = gt.match_longest_batch(docs, icase, sep)
            want = [ot.match_longest(d, icase, sep) for d in docs]
            off = np.cumsum([0] + [len(w) for w in want])
            assert np.array_equal(res.doc_hit_off(), off)
            assert res.hits().tobytes() == np.concatenate(want).tobytes()


ZH = "的一是在不了有和人这中大为上个国我以要他时来用们生到作地于出就分对成会可主发年动同工也能下过子说产种面而方后多定行学法所民得经"


def _utf8_keys(rng, alpha, n, maxlen):
    return list(dict.fromkeys("".join(rng.choice(alpha) for _ in range(rng.randint(1, maxlen))) for _ in range(n)))


@pytest.mark.parametrize("seed", range(6))
def test_char_automaton_paths(seed):
    """UTF-8 dictionaries (eligible for the shared-memory character automaton): ragged / tiny / empty documents,
    dense hits (event-slot overflow -> direct pass), malformed UTF-8 documents (byte-automaton fallback)."""
    rng = random.Random(1000 + seed)
    alpha = [ZH, "abcé日本語Я𝄞 ", "ab", ZH[:6] + "ab "][seed % 4]
    keys = _utf8_keys(rng, alpha, rng.choice([30, 600, 4000]), rng.choice([2, 4, 6]))
    docs = []
    for _ in range(rng.choice([3, 40, 400])):
        L = rng.choice([0, 1, 3, rng.randint(0, 50), rng.randint(0, 3000)])
        docs.append("".join(rng.choice(alpha) for _ in range(L)).encode())
    check_batch(keys, docs)
    check_batch(keys, docs, aha_b200.UNITS_CHAR)
    check_batch(keys, docs, sep=[" "])
    # malformed UTF-8 sprinkled in (truncated units, stray continuation bytes, 0xFF): byte offsets only
    bad = []
    for d in docs[:50]:
        b = bytearray(d)
        for _ in range(rng.randint(0, 3)):
            if b:
                b[rng.randrange(len(b))] = rng.choice([0x80, 0xBF, 0xE4, 0xFF, 0xC3, 0x41])
        bad.append(bytes(b))
    check_batch(keys, bad)


def test_char_automaton_dense_hits_and_many_tiny_docs():
    keys = ["的", "的的", "的的的", "一", "的一", "一的"]
    check_batch(keys, ["的" * 5000, "的一" * 3000, ("一的的" * 700)])
    check_batch(keys, ["的一"] * 3000 + [""] * 50 + ["的"] * 500, aha_b200.UNITS_CHAR)


def _stream_hash(hits_i32, base=0):
    """Order-sensitive 64-bit checksum of a hit stream on the device: sum_i mix(i + base, start, end, value) mod 2^64."""
    import torch
    h = hits_i32.view(-1, 3).to(torch.int64)
    idx = torch.arange(h.shape[0], dtype=torch.int64, device=h.device) + base
    x = h[:, 0] * -7046029254386353131 + h[:, 1] * -4417276706812531889 + h[:, 2] * 1609587929392839161 + idx * -8796714831421723037
    x = (x ^ (x >> 29)) * -4658895280553007687
    return int(x.sum().item())


@pytest.mark.heavy
def test_full_size_c2_properties():
    """BASELINE config C2 at full size (10 k CJK patterns, 2 048 documents of 1 MiB = 2.1 GB resident in HBM).  The oracle
    cannot replay 80 M hits record by record in test time, so the full-size run is checked through size-independent
    properties: total hit count against the oracle's multi-threaded counting walk over ALL the text, every 97th
    document hit for hit, offsets monotone and consistent, and the stream of the whole batch equal (order-sensitive
    checksum) to the streams of its two halves scanned separately."""
    import torch
    keys = synth.patterns_zh_ngrams(10000)
    g = aha_b200.AC.compile(keys)
    o = O.AC.compile(keys)
    D = synth.doc_bytes(1 << 20)
    n_docs = 2048
    dev = synth.gen_device(synth.ZH, synth.SEED_TEXT + 1, 0, n_docs * D)
    doc_off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
    res = g.scan_device(dev, doc_off)
    off = res.doc_hit_off_tensor().cpu().numpy()
    assert off[0] == 0 and off[-1] == res.n_hits and np.all(np.diff(off) >= 0)
    host = dev.cpu().numpy()
    total = o.scan_count_mt(host, np.arange(n_docs + 1, dtype=np.int64) * D, os.cpu_count() or 4)
    assert total == res.n_hits
    hits = res.hits_tensor()
    for d in range(0, n_docs, 97):
        want = o.scan(host[d * D:(d + 1) * D])
        got = hits[int(off[d]):int(off[d + 1])].cpu().numpy().astype("<i4")
        assert got.tobytes() == want.tobytes(), "document %d" % d
    whole = _stream_hash(hits)
    half = n_docs // 2
    a = g.scan_device(dev[:half * D], doc_off[:half + 1])
    ha = _stream_hash(a.hits_tensor())
    na = a.n_hits
    assert na == off[half]
    a = None
    b = g.scan_device(dev[half * D:], doc_off[:n_docs - half + 1])
    assert na + b.n_hits == res.n_hits
    hb = _stream_hash(b.hits_tensor(), na)
    assert (ha + hb) % (1 << 64) == whole % (1 << 64)


def _doc_stream_hash(res):
    """(hits, checksum) of a device result in the oracle's batch formula (bench.gpu_stream_hash: mix of document,
    index in the document, start, end, value)"""
    import bench
    return bench.gpu_stream_hash(res.hits_tensor(), res.doc_hit_off_tensor(), res.n_docs)


@pytest.mark.heavy
@pytest.mark.parametrize("cfg", ["c2", "c3"])
def test_modes_at_config_scale(cfg):
    """ACX#match(sep), match_longest (+ intersectable) and the String forms on inputs of the configurations' shape
    (64 documents of 1 MiB, the configuration's dictionary): hit count and order-sensitive checksum of every
    document's stream against the oracle's document-parallel batch scan."""
    import torch
    if cfg == "c2":
        keys, corpus, seed = synth.patterns_zh_ngrams(10000), synth.ZH, synth.SEED_TEXT + 1
    else:
        keys, corpus, seed = synth.patterns_mix(100000), synth.MIX, synth.SEED_TEXT + 2
    g = aha_b200.AC.compile(keys)
    o = O.AC.compile(keys)
    D = synth.doc_bytes(1 << 20)
    n_docs = 64
    dev = synth.gen_device(corpus, seed, 0, n_docs * D)
    doc_off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
    host = dev.cpu().numpy()
    host_off = np.arange(n_docs + 1, dtype=np.int64) * D
    threads = os.cpu_count() or 4
    L, I, U = aha_b200.LONGEST, aha_b200.INTERSECTABLE, aha_b200.UNITS_CHAR
    for mode, sep in [(0, [" "]), (0, ["\n", " "]), (U, [" "]), (L, None), (L | I, None), (L | U, None), (U, None)]:
        res = g.scan_device(dev, doc_off, mode, sep)
        want_n, want_h = o.scan_mode_mt(host, host_off, threads, mode, sep)
        got_n, got_h = _doc_stream_hash(res)
        assert (res.n_hits, got_n, got_h) == (want_n, want_n, want_h), "mode %d sep %r" % (mode, sep)


def test_char_array_overloads_and_code_point_sep():
    """Array(Char) / Slice(Char) overloads (ac.cr:288-295, 312-319, 342-364): character offsets, and with `sep` the
    neighbour test on the neighbouring character's code point (AHA_SEP_CODEPOINT)."""
    g = aha_b200.AC.compile(["我", "我是", "是中"])  # spec/ac_spec.cr:36-43
    assert [(int(h["end"]), int(h["value"])) for h in g.match(list("我是中国人"))] == [(1, 0), (2, 1), (3, 2)]
    g = aha_b200.AC.compile(["a", "aa"])  # spec/ac_spec.cr:45-54
    assert [(int(h["end"]), int(h["value"])) for h in g.match(list("a aaa"), sep=[" "])] == [(1, 0)]
    keys = ["我", "a", "é", "éa", "的的", "ab", "b "]
    g, o = aha_b200.AC.compile(keys), O.AC.compile(keys)
    for text in ["我a是", "éa", "a é", "我 a", "aé的的é ab", "é", "", "的的a的的 éa b ", "ÿaÿ þa", "a\u0100a \u00ffa"]:
        for sep, size in [([" "], 256), ([" ", "é"], 256), ([" "], 0xE9), ([" "], 0x61), ([], 256)]:
            assert as_tuples(g.scan_batch([text], aha_b200.UNITS_CHAR | aha_b200.SEP_CODEPOINT, sep, size).hits()) == \
                as_tuples(o.scan(text, aha_b200.UNITS_CHAR | aha_b200.SEP_CODEPOINT, sep, size)), (text, sep, size)
            assert as_tuples(g.match(list(text), sep=sep)) == as_tuples(o.match(list(text), sep=sep))
    rng = random.Random(9)
    alpha = "ab é的我ÿ"
    big = _utf8_keys(rng, alpha, 300, 4)
    docs = ["".join(rng.choice(alpha) for _ in range(rng.choice([0, 3, 40, 900]))) for _ in range(200)]
    check_batch(big, docs, aha_b200.UNITS_CHAR | aha_b200.SEP_CODEPOINT, sep=[" "])
    check_batch(big, docs, aha_b200.SEP_CODEPOINT, sep=[" ", "é"], sep_size=240)


def test_deeply_nested_keys_on_constant_text():
    """700 nested keys a, aa, ..., a^700 over a^n: every position ends up to 700 keys, a lane of a tile that goes through
    the direct pass counts far more than 65 535 hits (tile_meta holds a full 32-bit count per lane for such tiles)."""
    keys = ["a" * k for k in range(1, 701)]
    res = check_batch(keys, ["a" * 9000, "", "a" * 3, "的" + "a" * 2500])
    assert res.n_hits > 6_000_000
    check_batch(keys, ["a" * 5000], aha_b200.UNITS_CHAR)


def test_sep_on_the_character_path():
    """ACX#match(seq, sep) with byte offsets runs on the shared-memory character automaton: the vetoes are applied by
    the expansion.  Neighbours across document boundaries never veto; sep sizes below 256 (Q7); dense nested hits."""
    keys = ["的", "的的", "一的", "ab", "b", "abc", "c d"]
    docs = ["的的 的一的 的", "的", "", "的的", " 的 ", "ab abc b", "abc", "c d c d", "xab", "ab的", "的ab 的"]
    for sep, size in [([" "], 256), ([" "], 33), ([" "], 32), ([" ", "a"], 128), ([], 256), ([" "], 0)]:
        check_batch(keys, docs, sep=sep, sep_size=size)
    # a hit that ends a document followed by a document that starts with a non-separator, many tiny documents
    check_batch(keys, ["的", "的的", "ab", "b", ""] * 700, sep=[" "])
    # dense hits: event slots overflow -> direct pass with the same vetoes
    check_batch(keys, ["的" * 6000, "的 " * 3000, "ab " * 2000 + "的一的" * 800], sep=[" "])
    rng = random.Random(77)
    alpha = ZH[:5] + "ab  "
    big = _utf8_keys(rng, alpha, 3000, 5)
    docs = ["".join(rng.choice(alpha) for _ in range(rng.choice([0, 5, 700, 9000]))) for _ in range(120)]
    check_batch(big, docs, sep=[" "])
    check_batch(big, docs, sep=[" ", "a"], sep_size=200)
    # the String overloads (character offsets) on the same path: the expansion counts the characters
    U = aha_b200.UNITS_CHAR
    check_batch(big, docs, U, sep=[" "])
    check_batch(big, docs, U, sep=[" ", "a"], sep_size=200)
    check_batch(keys, ["的的 的一的 的", "的", "", "的的", " 的 ", "ab abc b", "abc", "c d c d", "xab", "ab的", "的ab 的"], U, sep=[" "])
    check_batch(keys, ["的", "的的", "ab", "b", ""] * 700, U, sep=[" "])
    check_batch(keys, ["的" * 6000, "的 " * 3000, "ab " * 2000 + "的一的" * 800, "é的 ab é" * 900], U, sep=[" "])


def test_match_longest_inside_long_documents():
    """ACX#match_longest runs as speculative 512-byte chunks inside a document: long documents (many chunks), document
    ends inside and exactly at chunk borders, pending ends carried across chunks, and texts on which the machine never
    re-synchronises with a guessed start (periodic text: the settle rounds walk the document chunk by chunk)."""
    L, I, U = aha_b200.LONGEST, aha_b200.INTERSECTABLE, aha_b200.UNITS_CHAR
    rng = random.Random(5)
    keys = ["a", "aa", "aaa", "aaaaa", "ab", "abab", "b", "ba" * 20, "的", "的的的", "的一", "一" * 7, "x" * 300]
    docs = ["a" * 5000, "ab" * 3000, "a" * 512, "b" * 511, "a", "", "a" * 1024, "的" * 2000, "的一" * 1500 + "的的的的",
            "x" * 299 + "ab" + "x" * 700, "ba" * 400 + "a" * 300, "a" * 513]
    docs += ["".join(rng.choice("ab的一x ") for _ in range(rng.choice([0, 7, 511, 512, 513, 4000]))) for _ in range(60)]
    for mode in (L, L | I, L | U, L | I | U):
        check_batch(keys, docs, mode)
    big = _utf8_keys(rng, ZH[:12] + "ab", 3000, 6)
    docs = ["".join(rng.choice(ZH[:12] + "ab") for _ in range(rng.choice([100, 3000, 20000]))) for _ in range(40)]
    for mode in (L, L | I, L | U):
        check_batch(big, docs, mode)


def test_hit_density_swings_between_calls():
    """One dictionary, batches whose hit density swings by orders of magnitude between calls: the hit buffer is sized
    per call after the walk has counted, whatever the previous call looked like."""
    keys = ["ab", "b", "日本", "本"]
    g = aha_b200.AC.compile(keys)
    o = O.AC.compile(keys)
    batches = [["x" * 100000], ["ab日本" * 30000, "b" * 50000], ["ab" * 10, "", "x" * 5000], ["本" * 200000]]
    for docs in batches:
        for mode in (0, aha_b200.UNITS_CHAR):
            res = g.scan_batch(docs, mode)
            want, want_off = oracle_batch(o, docs, mode)
            assert res.n_hits == len(want)
            assert np.array_equal(res.doc_hit_off(), want_off)
            assert res.hits().tobytes() == want.tobytes()


@pytest.mark.heavy
def test_large_c3_properties():
    """BASELINE config C3's dictionary (100 k mixed patterns: the byte automaton in global memory) over 1 GiB of MIX
    text, checked like the full-size C2 run: oracle total count over all text, every 61st document hit for hit,
    both offset units, and halves against the whole."""
    import torch
    keys = synth.patterns_mix(100000)
    g = aha_b200.AC.compile(keys)
    o = O.AC.compile(keys)
    D = synth.doc_bytes(1 << 20)
    n_docs = 1024
    dev = synth.gen_device(synth.MIX, synth.SEED_TEXT + 2, 0, n_docs * D)
    doc_off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
    res = g.scan_device(dev, doc_off)
    off = res.doc_hit_off_tensor().cpu().numpy()
    assert off[0] == 0 and off[-1] == res.n_hits and np.all(np.diff(off) >= 0)
    host = dev.cpu().numpy()
    total = o.scan_count_mt(host, np.arange(n_docs + 1, dtype=np.int64) * D, os.cpu_count() or 4)
    assert total == res.n_hits
    hits = res.hits_tensor()
    for d in range(0, n_docs, 61):
        want = o.scan(host[d * D:(d + 1) * D])
        got = hits[int(off[d]):int(off[d + 1])].cpu().numpy().astype("<i4")
        assert got.tobytes() == want.tobytes(), "document %d" % d
    whole = _stream_hash(hits)
    n_all = res.n_hits
    hits = None
    res = None
    half = n_docs // 2
    a = g.scan_device(dev[:half * D], doc_off[:half + 1])
    ha, na = _stream_hash(a.hits_tensor()), a.n_hits
    assert na == off[half]
    a = None
    b = g.scan_device(dev[half * D:], doc_off[:n_docs - half + 1])
    assert na + b.n_hits == n_all
    assert (ha + _stream_hash(b.hits_tensor(), na)) % (1 << 64) == whole % (1 << 64)
    b = None
    # character offsets (String overloads) on a slice: every hit, against the oracle
    docs = [host[d * D:(d + 1) * D].tobytes() for d in (0, 511)]
    r = g.scan_batch(docs, aha_b200.UNITS_CHAR)
    want, want_off = oracle_batch(o, docs, aha_b200.UNITS_CHAR)
    assert np.array_equal(r.doc_hit_off(), want_off) and r.hits().tobytes() == want.tobytes()


def test_large_dictionary_without_dense_table():
    """BASELINE config C3's dictionary has no dense table (722 k states x 121 classes > 128 MB): every mode walks the
    sparse rows -- match (all three device paths of the fixture), sep, both longest modes, character units, and the trie
    path -- on a few documents, every hit against the oracle"""
    keys = synth.patterns_mix(100000)
    D = synth.doc_bytes(1 << 15)
    text = synth.gen_host(synth.MIX, synth.SEED_TEXT + 2, 0, 6 * D)
    docs = _split_docs(text, D) + [b"", b"ab"]
    g = aha_b200.AC.compile(keys)
    o = O.AC.compile(keys)
    assert g.info()["n_states"] * g.info()["n_classes"] * 4 > (128 << 20)
    for mode, sep in ((0, None), (0, [" "]), (aha_b200.LONGEST, None), (aha_b200.LONGEST | aha_b200.INTERSECTABLE, None),
                      (aha_b200.UNITS_CHAR, None)):
        res = g.scan_batch(docs, mode, sep)
        want, want_off = oracle_batch(o, docs, mode, sep)
        assert np.array_equal(res.doc_hit_off(), want_off), mode
        assert res.hits().tobytes() == want.tobytes(), mode
    gt, ot = aha_b200.Cedar(), O.Cedar()
    for k in keys[:30000]:
        assert gt.insert(k) == ot.insert(k)
    sdocs = [d.decode("utf-8") for d in docs[:3]]
    for icase, sep in ((True, [" "]), (False, None)):
        res = gt.match_longest_batch(sdocs, icase, sep)
        want = [ot.match_longest(d, icase, sep) for d in sdocs]
        assert np.array_equal(res.doc_hit_off(), np.cumsum([0] + [len(w) for w in want]))
        assert res.hits().tobytes() == np.concatenate(want).tobytes()


@pytest.mark.heavy
def test_scaled_c5_million_patterns():
    """BASELINE config C5's dictionary (1 M mixed patterns: a 3.2 GB byte automaton, far beyond L2) over 256 MiB of
    MIX text: oracle total count over all text, every 37th document hit for hit, halves against the whole, and the
    same automaton rebuilt from its device blob (what a rank does after the ncclBroadcast) giving the same stream."""
    import torch
    keys = synth.patterns_mix(1000000)
    g = aha_b200.AC.compile(keys)
    o = O.AC.compile(keys)
    info = g.info()
    assert info["n_keys"] == 1000000 and info["n_states"] > 3000000
    D = synth.doc_bytes(1 << 20)
    n_docs = 256
    dev = synth.gen_device(synth.MIX, synth.SEED_TEXT + 2, 0, n_docs * D)
    doc_off = torch.arange(0, n_docs + 1, dtype=torch.int64, device="cuda") * D
    res = g.scan_device(dev, doc_off)
    off = res.doc_hit_off_tensor().cpu().numpy()
    assert off[0] == 0 and off[-1] == res.n_hits and np.all(np.diff(off) >= 0)
    host = dev.cpu().numpy()
    total = o.scan_count_mt(host, np.arange(n_docs + 1, dtype=np.int64) * D, os.cpu_count() or 4)
    assert total == res.n_hits
    hits = res.hits_tensor()
    for d in range(0, n_docs, 37):
        want = o.scan(host[d * D:(d + 1) * D])
        got = hits[int(off[d]):int(off[d + 1])].cpu().numpy().astype("<i4")
        assert got.tobytes() == want.tobytes(), "document %d" % d
    whole, n_all = _stream_hash(hits), res.n_hits
    hits = res = None
    # a second rank's view: the blob arrives in device memory, the automaton is rebuilt from it
    blob = torch.from_numpy(g.blob_view().copy()).cuda()
    g2 = aha_b200.AC.from_device_blob(blob)
    half = n_docs // 2
    a = g2.scan_device(dev[:half * D], doc_off[:half + 1])
    ha, na = _stream_hash(a.hits_tensor()), a.n_hits
    assert na == off[half]
    a = None
    b = g2.scan_device(dev[half * D:], doc_off[:n_docs - half + 1])
    assert na + b.n_hits == n_all
    assert (ha + _stream_hash(b.hits_tensor(), na)) % (1 << 64) == whole % (1 << 64)


def test_scaled_c4_trie_longest():
    """BASELINE config C4 scaled down: 20 k EN words and phrases in mixed case, 3 MB of EN text as 4 KiB documents,
    CedarX#match_longest with ignore_case and sep = [' '] (character offsets), every hit against the oracle."""
    keys = synth.patterns_en_phrases(20000)
    gt, ot = aha_b200.Cedar(), O.Cedar()
    for k in keys:
        assert gt.insert(k) == ot.insert(k)
    D = 4096 // synth.BLOCK * synth.BLOCK
    text = synth.gen_host(synth.EN, synth.SEED_TEXT + 3, 0, 768 * D)
    docs = [text[i:i + D].tobytes().decode("ascii") for i in range(0, len(text), D)]
    for icase, sep in ((True, [" "]), (False, None)):
        res = gt.match_longest_batch(docs, icase, sep)
        want = [ot.match_longest(d, icase, sep) for d in docs]
        off = np.cumsum([0] + [len(w) for w in want])
        assert res.n_hits > 1000
        assert np.array_equal(res.doc_hit_off(), off)
        assert res.hits().tobytes() == np.concatenate(want).tobytes()
