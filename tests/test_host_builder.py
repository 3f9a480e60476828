"""CPU tests of the product's host side: the C ABI library loads and exports every declared symbol,
the host cedar builder reproduces the oracle's (= reference-layout) arrays bit for bit, and the
flattened DFA, walked in pure Python, yields the oracle's hit stream."""
import ctypes as C
import os
import random
import re

import numpy as np
import pytest

import aha_b200
from aha_b200 import _lib
from oracle import oracle as O
from tests.test_oracle_kat import _rand_keys, sev

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "aha_b200.h")).read()
    declared = set(re.findall(r"\b(aha_b200_[a-z0-9_]+)\s*\(", hdr))
    L = _lib.lib()
    assert declared, "no declarations found"
    for name in sorted(declared):
        assert hasattr(L, name), name
    assert declared == set(_lib.EXPORTS)


def _key_sets():
    rng = random.Random(7)
    yield ["我", "我是", "是中"]
    yield ["Ruby", "ruby", "rub"]
    yield open(os.path.join(ROOT, "tests", "golden", "cedar_words.txt")).read().split()
    for seed in range(10):
        r = random.Random(seed)
        alphabet = [b"ab", b"abc", b"abcdefgh", bytes(range(1, 256))][seed % 4]
        yield _rand_keys(r, r.choice([30, 300, 3000]), alphabet, r.choice([3, 5, 9]))
    words = ["".join(rng.choice("abcdefghijklmnopqrstuvwxyz") for _ in range(rng.randint(2, 10))) for _ in range(20000)]
    yield list(dict.fromkeys(words))


@pytest.mark.parametrize("idx", range(14))
def test_builder_matches_oracle_layout(idx):
    keys = list(_key_sets())[idx]
    o = O.AC.compile(keys)
    da = aha_b200.Cedar()
    for i, k in enumerate(keys):
        assert da.insert(k) == i
    p_nodes, p_fails, p_out, p_kl = da.compiled_arrays()
    o_nodes, o_fails, o_out, o_kl = o.arrays()
    assert np.array_equal(p_nodes, o_nodes)
    assert np.array_equal(p_fails, o_fails)
    assert np.array_equal(p_out, o_out)
    assert np.array_equal(p_kl, o_kl)
    for i in (0, len(keys) // 2, len(keys) - 1):
        k = keys[i]
        assert da.key_bytes(i) == (k.encode() if isinstance(k, str) else bytes(k))
        assert da.get(k) == i


def test_builder_errors_match_reference():
    da = aha_b200.Cedar()
    with pytest.raises(aha_b200.AhaError, match="Cannot insert empty key"):  # cedar.cr:756
        da.insert("")
    with pytest.raises(aha_b200.AhaError, match=r"key\[1\] is zero"):  # cedar.cr:235
        da.insert(b"a\x00b")
    da = aha_b200.Cedar()  # (a raised insert leaves a half-built path behind, in the reference too)
    assert da.insert("a") == 0
    assert da.insert("a") == 0  # existing key returns its id (cedar.cr:757-758)


class BlobDfa:
    """pure-Python reader + walker of the flattened DFA blob (test-only)."""

    def __init__(self, arrays):
        nodes, fails, output, kl = arrays
        blob, n = C.c_void_p(), C.c_int64()
        L = _lib.lib()
        _lib.check(L.aha_b200_flatten_host(nodes.ctypes.data, len(fails), fails.ctypes.data, output.ctypes.data,
                                           kl.ctypes.data if len(kl) else None, len(kl), C.byref(blob), C.byref(n)))
        raw = C.string_at(blob.value, n.value)
        L.aha_b200_blob_free(blob)
        hdr = np.frombuffer(raw[:120], dtype=np.uint8)
        u64 = lambda o: int(np.frombuffer(raw[o:o + 8], "<u8")[0])
        u32 = lambda o: int(np.frombuffer(raw[o:o + 4], "<u4")[0])
        self.C, self.S, self.K, self.n_out, self.maxlen = u32(12), u64(16), u64(24), u64(32), u32(40)
        offs = [u64(48 + 8 * i) for i in range(9)]
        self.classmap = np.frombuffer(raw, np.uint8, 256, offs[0])
        self.delta = np.frombuffer(raw, "<u4", self.S * self.C, offs[1]).reshape(self.S, self.C)
        self.out_off = np.frombuffer(raw, "<u4", self.S + 1, offs[2])
        self.out_val = np.frombuffer(raw, "<i4", self.n_out, offs[3])
        self.key_len = np.frombuffer(raw, "<u4", self.K, offs[4])
        self.key_tail = np.frombuffer(raw, "<u4", self.K, offs[5])
        assert offs[8] == len(raw)
        self.fail = np.frombuffer(raw, '<u4', self.S, offs[7])
        self.depth = np.frombuffer(raw, '<u2', self.S, offs[6])

    def match(self, text):
        hits, s = [], 0
        for i, b in enumerate(text):
            e = int(self.delta[s, self.classmap[b]])
            s = e & 0x1FFFFFFF
            if e & 0x80000000:
                for j in range(self.out_off[s], self.out_off[s + 1]):
                    v = int(self.out_val[j])
                    hits.append((i + 1 - int(self.key_len[v]), i + 1, v))
        return hits

    def match_longest(self, text, inter):
        hits, s, prev_i, prev_s = [], 0, -1, 0

        def flush():
            j = self.out_off[prev_s]
            if j < self.out_off[prev_s + 1]:
                v = int(self.out_val[j])
                hits.append((prev_i + 1 - int(self.key_len[v]), prev_i + 1, v))

        for i, b in enumerate(text):
            e = int(self.delta[s, self.classmap[b]])
            if e & 0x20000000:
                s = e & 0x1FFFFFFF
                if e & 0x40000000:
                    prev_i, prev_s = i, s
            else:
                if prev_i >= 0:
                    flush()
                    prev_i = -1
                    if not inter:
                        s = 0
                        continue
                s = e & 0x1FFFFFFF
                if s != 0 and (e & 0x40000000):
                    prev_i, prev_s = i, s
        if prev_i >= 0:
            flush()
        return hits


@pytest.mark.parametrize("seed", range(8))
def test_flattened_dfa_reproduces_oracle_stream(seed):
    rng = random.Random(100 + seed)
    alphabet = [b"ab", b"abc", b"abcdefgh", bytes(range(1, 256))][seed % 4]
    keys = _rand_keys(rng, rng.choice([5, 50, 500]), alphabet, rng.choice([2, 4, 7]))
    text = bytes(rng.choice(alphabet) for _ in range(3000))
    o = O.AC.compile(keys)
    da = aha_b200.Cedar()
    for k in keys:
        da.insert(k)
    dfa = BlobDfa(da.compiled_arrays())
    assert dfa.maxlen == max(len(k) for k in keys)
    assert dfa.match(text) == sev(o.match(text))
    for inter in (False, True):  # includes the stale-END cases (Q4): the DFA carries raw is_end bits
        assert dfa.match_longest(text, inter) == sev(o.match_longest(text, inter))


def test_generator_is_deterministic_and_nul_free():
    from aha_b200 import synth
    for corpus in (synth.EN, synth.ZH, synth.MIX):
        a = synth.gen_host(corpus, 42, 0, 192 * 50)
        b = synth.gen_host(corpus, 42, 192 * 10 + 7, 1000)
        assert np.array_equal(a[192 * 10 + 7:192 * 10 + 1007], b)
        assert (a != 0).all()
        a.tobytes().decode("utf-8")  # every block is valid UTF-8


def _char_scan(blob, text, chunk, budget=200 * 1024):
    L = _lib.lib()
    a = np.frombuffer(text, dtype=np.uint8)
    cap = 1 << 20
    hits = np.zeros(cap, dtype=aha_b200.HIT_DTYPE)
    n, info = C.c_int64(), (C.c_int64 * 8)()
    bl = np.frombuffer(blob, dtype=np.uint8)
    _lib.check(L.aha_b200_debug_char_scan(bl.ctypes.data, bl.size, a.ctypes.data if a.size else None, a.size, chunk, budget,
                                          hits.ctypes.data, cap, C.byref(n), info))
    return hits[: n.value], list(info)


def _blob_of(keys):
    da = aha_b200.Cedar()
    for k in keys:
        da.insert(k)
    nodes, fails, output, kl = da.compiled_arrays()
    blob, n = C.c_void_p(), C.c_int64()
    L = _lib.lib()
    _lib.check(L.aha_b200_flatten_host(nodes.ctypes.data, len(fails), fails.ctypes.data, output.ctypes.data,
                                       kl.ctypes.data if len(kl) else None, len(kl), C.byref(blob), C.byref(n)))
    raw = C.string_at(blob.value, n.value)
    L.aha_b200_blob_free(blob)
    return raw


@pytest.mark.parametrize("seed", range(6))
def test_char_automaton_replay_matches_oracle(seed):
    """the shared-memory character automaton (tables + per-lane walk, replayed on the host) == oracle"""
    rng = random.Random(seed)
    alpha = ["ab", "abcé日本語Я𝄞 ", "的一是在不了有和人这中大为上个国我以要他时来用们生到作地于出就分对成会可主发年动同工也能下过子说产种面而方后多定行学法所民得经", "abc"][seed % 4]
    keys = list(dict.fromkeys("".join(rng.choice(alpha) for _ in range(rng.randint(1, 5))) for _ in range(rng.choice([20, 400, 3000]))))
    text = "".join(rng.choice(alpha) for _ in range(6000)).encode()
    o = O.AC.compile(keys)
    want = o.scan(text)
    blob = _blob_of(keys)
    for chunk in (1, 7, 132, 100000):
        got, info = _char_scan(blob, text, chunk)
        assert info[0] == 1 and info[1] == 1, info
        assert got.tobytes() == want.tobytes(), (chunk, len(got), len(want))


def test_char_automaton_eligibility_and_bad_text():
    blob = _blob_of([b"\xe4\xb8", b"ab"])  # a key that stops inside a unit
    _, info = _char_scan(blob, b"abc", 16)
    assert info[0] == 0
    blob = _blob_of([b"\x80x"])  # a key that starts with a continuation byte
    _, info = _char_scan(blob, b"abc", 16)
    assert info[0] == 0
    blob = _blob_of(["日本", "ab"])
    for bad in (b"ab\xe6\x97", b"ab\x80\x80ab", b"\xff", b"\xe6ab"):
        _, info = _char_scan(blob, bad, 16)
        assert info[0] == 1 and info[1] == 0, bad


def test_char_automaton_c2_dictionary_fits_shared_memory():
    from aha_b200 import synth
    keys = synth.patterns_zh_ngrams(10000)
    blob = _blob_of(keys)
    text = synth.gen_host(synth.ZH, synth.SEED_TEXT + 1, 0, 192 * 600).tobytes()
    got, info = _char_scan(blob, text, 132, 227 * 1024 - 16 * (4352 + 128) - 512)  # what build_hot_table leaves for the tables
    assert info[0] == 1 and info[1] == 1, info
    want = O.AC.compile(keys).scan(text)
    assert got.tobytes() == want.tobytes()


def test_threaded_flatten_is_byte_identical(monkeypatch):
    """the level-synchronous / two-pass flatten on several host threads (large dictionaries) writes the same blob as the
    sequential passes; both with a dense table (small class count) and without one (sparse rows only)"""
    rng = random.Random(31)
    for alpha, n_keys, lens in (("abcdefghijklmnopqrstuvwxyz", 70000, (3, 12)), ("的一是在不了有和人这中大为上个国我以要他时来用们生到作地于出就分对成会可主发年动同工也能下过子说产种面而方后多定行学法所民得经abcdefgh ", 120000, (2, 9))):
        keys = list(dict.fromkeys("".join(rng.choice(alpha) for _ in range(rng.randint(*lens))) for _ in range(n_keys)))
        monkeypatch.setenv("AHA_B200_FLATTEN_THREADS", "1")
        one = _blob_of(keys)
        monkeypatch.setenv("AHA_B200_FLATTEN_THREADS", "7")
        many = _blob_of(keys)
        assert len(one) > (1 << 22) and one == many


def _folded_longest(blob, text, sep=None):
    L = _lib.lib()
    a = np.frombuffer(text, dtype=np.uint8)
    cap = 1 << 18
    hits = np.zeros(cap, dtype=aha_b200.HIT_DTYPE)
    n, info = C.c_int64(), (C.c_int64 * 8)()
    bl = np.frombuffer(blob, dtype=np.uint8)
    mode = aha_b200.TRIE_LONGEST | aha_b200.IGNORE_CASE | (aha_b200.SEP if sep is not None else 0)
    bits, size = O.sep_bits(sep) if sep is not None else (bytes(32), 0)
    _lib.check(L.aha_b200_debug_folded_longest(bl.ctypes.data, bl.size, a.ctypes.data if a.size else None, a.size, mode, bits, size,
                                               hits.ctypes.data, cap, C.byref(n), info))
    return hits[: n.value], list(info)


@pytest.mark.parametrize("seed", range(8))
def test_folded_trie_replay_matches_oracle(seed):
    """CedarX#match_longest(ignore_case: true) over the case-folded trie (table builder + walk + candidate pick, replayed
    on the host) == the oracle's frontier walk (cedar.cr:984-1027, 1075-1121), keys that differ only in case included"""
    rng = random.Random(100 + seed)
    alpha = ["abAB", "abcABC ,", "aAbBéÉñÑ -", "αβγΑΒΓдДжЖ ab", "abAB日本 ", "xyXY"][seed % 6]
    n_keys = rng.choice([6, 40, 300])
    keys = list(dict.fromkeys("".join(rng.choice(alpha.replace(" ", "").replace(",", "").replace("-", "")) for _ in range(rng.randint(1, 6)))
                              for _ in range(n_keys)))
    if seed % 6 == 5:  # four and more case variants of one key under one folded node: the overflow records
        keys += ["xyx", "Xyx", "xYx", "xyX", "XYx", "XyX", "xYX", "XYX"]
        keys = list(dict.fromkeys(keys))
    da, o = aha_b200.Cedar(), O.Cedar()
    for k in keys:
        da.insert(k)
        o.insert(k)
    nodes, fails, output, kl = da.compiled_arrays()
    blob, n = C.c_void_p(), C.c_int64()
    L = _lib.lib()
    _lib.check(L.aha_b200_flatten_host(nodes.ctypes.data, len(fails), fails.ctypes.data, output.ctypes.data,
                                       kl.ctypes.data if len(kl) else None, len(kl), C.byref(blob), C.byref(n)))
    raw = C.string_at(blob.value, n.value)
    L.aha_b200_blob_free(blob)
    for trial in range(4):
        text = "".join(rng.choice(alpha) for _ in range(rng.choice([0, 1, 50, 3000]))).encode()
        for sep in (None, [" ", ",", "-"]):
            got, info = _folded_longest(raw, text, sep)
            assert info[0] == 1, info
            want = o.match_longest(text, ignore_case=True, sep=sep)
            assert got.tobytes() == want.tobytes(), (seed, trial, sep, len(got), len(want))
    if seed % 6 == 5:
        assert info[5] >= 1  # the eight variants share one node: its record is the overflow form


def test_folded_trie_refuses_what_it_cannot_hold():
    blob = _blob_of(["a" * 65, "b"])  # a key of more than 64 characters: the depth-first kernel keeps the dictionary
    _, info = _folded_longest(blob, b"aaab")
    assert info[0] == 0
    blob = _blob_of([b"\xe4\xb8", b"ab"])  # a key that is not valid UTF-8
    _, info = _folded_longest(blob, b"ab")
    assert info[0] == 0
