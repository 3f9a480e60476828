"""Pins the CPU oracle against every known-answer vector the reference's own specs hold for the
Aha::AC path, then cross-checks it against an independent model."""
import random

import numpy as np
import pytest

from oracle import oracle as O
from tests.pymodel import PyAC


def ev(h):
    return [(int(x["end"]), int(x["value"])) for x in h]


def sev(h):
    return [(int(x["start"]), int(x["end"]), int(x["value"])) for x in h]


def test_ac_spec_basic():  # spec/ac_spec.cr:5-12, README.md:29-38
    m = O.AC.compile(["我", "我是", "是中"])
    assert ev(m.match("我是中国人")) == [(1, 0), (2, 1), (3, 2)]
    # Bytes overload yields byte offsets (ac.cr:269-273)
    assert sev(m.match("我是中国人".encode())) == [(0, 3, 0), (0, 6, 1), (3, 9, 2)]


def test_ac_spec_sep():  # spec/ac_spec.cr:25-34, 45-54
    m = O.AC.compile(["a", "aa"])
    assert ev(m.match("a aaa", sep=[" "])) == [(1, 0)]


def test_ac_spec_char_array():  # spec/ac_spec.cr:36-43, 45-54: the Array(Char) overloads
    m = O.AC.compile(["我", "我是", "是中"])
    assert ev(m.match(list("我是中国人"))) == [(1, 0), (2, 1), (3, 2)]
    m = O.AC.compile(["a", "aa"])
    assert ev(m.match(list("a aaa"), sep=[" "])) == [(1, 0)]


def test_ac_char_array_sep_tests_code_points():
    """ac.cr:342-364 against ac.cr:321-340, derived from the source: the Array(Char) form asks `chr.ord < sep.size`, so a
    neighbour at or above U+0100 can never veto, while the Bytes / String form sees its UTF-8 bytes (0x80..0xFF < 256,
    none of them a separator) and vetoes.  A Latin-1 neighbour (U+00E9 < 256) vetoes in both forms."""
    m = O.AC.compile(["我", "a"])
    assert sev(m.match(list("我a是"), sep=[" "])) == [(1, 2, 1)]   # 'a' between two CJK characters passes; "我" is vetoed by 'a'
    assert sev(m.match("我a是", sep=[" "])) == []                   # bytes: every neighbour vetoes
    assert sev(m.match(list("éa"), sep=[" "])) == []                # U+00E9 in front of 'a' vetoes
    assert sev(m.match(list("a é"), sep=[" "])) == [(0, 1, 1)]
    assert sev(m.match(list("我 a"), sep=[" "])) == [(0, 1, 0), (2, 3, 1)]


def test_ac_longest_not_intersectable():  # spec/ac_longest_match_spec.cr:5-18, 35-48
    t = O.Cedar()
    assert [t.insert(s) for s in ["Ruby", "ruby", "rub"]] == [0, 1, 2]
    ac = O.AC.compile(t)
    got = [(s, e, ac[v]) for s, e, v in sev(ac.match_longest("Ruby on rub"))]
    assert got == [(0, 4, "Ruby"), (8, 11, "rub")]


def test_ac_longest_intersectable():  # spec/ac_longest_match_spec.cr:20-33, 50-63
    t = O.Cedar()
    for s in ["Ruby", "ruby", "uby "]:
        t.insert(s)
    ac = O.AC.compile(t)
    got = [(s, e, ac[v]) for s, e, v in sev(ac.match_longest("ruby ", True))]
    assert got == [(0, 4, "ruby"), (1, 5, "uby ")]


def _trie():
    t = O.Cedar()
    for s in ["Ruby", "ruby", "rub"]:
        t.insert(s)
    return t


def test_cedar_match():  # spec/cedar_match_spec.cr:4-16
    t = _trie()
    got = [(s, e, t[v]) for s, e, v in sev(t.match_longest("Ruby on rub", False))]
    assert got == [(0, 4, "Ruby"), (8, 11, "rub")]


def test_cedar_match_ignore_case():  # spec/cedar_match_spec.cr:18-30
    t = _trie()
    got = sorted((s, e, t[v]) for s, e, v in sev(t.match_longest("Ruby on ruby", True)))
    assert got == [(0, 4, "Ruby"), (8, 12, "ruby")]


def test_cedar_match_sep():  # spec/cedar_match_spec.cr:51-63
    t = _trie()
    for early in (True, False):
        got = sorted((s, e, t[v]) for s, e, v in sev(t.match_longest("Ruby rRuby on rubyx ruby", True, sep=[" "], early_exit=early)))
        assert got == [(0, 4, "Ruby"), (20, 24, "ruby")]


def _crystal_block(t):
    names = {"ruby": "crystal", "Ruby": "Crystal"}
    return lambda m: names.get(t[m[2]], "Unknown")


def test_cedar_gsub():  # spec/cedar_match_spec.cr:32-49
    t = _trie()
    assert t.gsub("Ruby on ruby rub", True, block=_crystal_block(t)) == "Crystal on crystal Unknown"


def test_cedar_gsub_sep():  # spec/cedar_match_spec.cr:65-81
    t = _trie()
    assert t.gsub("Ruby rRuby on rubyx ruby", True, sep=[" "], block=_crystal_block(t)) == "Crystal rRuby on rubyx crystal"


def test_cedar_insert_ids():  # spec/cedar_spec.cr:4-13 (insert part)
    t = O.Cedar()
    assert [t.insert(s) for s in ["Ruby", "ruby", "rb"]] == [0, 1, 2]
    assert t.insert("ruby") == 1


WORDS = open(__file__.replace("test_oracle_kat.py", "golden/cedar_words.txt")).read().split()


def test_cedar_words_roundtrip():  # spec/cedar_spec.cr:28-293
    t = O.Cedar()
    for w in WORDS:
        t.insert(w)
    assert t.size == len(WORDS)
    for i, w in enumerate(WORDS):
        assert t[i] == w
        assert t[w] == i


def test_compile_errors():  # ac.cr:66, cedar.cr:756, cedar.cr:235
    with pytest.raises(RuntimeError):
        O.AC.compile(["a", "a"])
    with pytest.raises(RuntimeError):
        O.AC.compile(["a", ""])
    with pytest.raises(RuntimeError):
        O.AC.compile([b"a\x00b"])


def test_quirks():
    # Q1: outputs only at end states; chain stops at first non-end fail
    assert sev(O.AC.compile(["abcd", "bc"]).match(b"abcd")) == [(0, 4, 0)]
    assert sev(O.AC.compile(["xab", "abz", "b"]).match(b"xab")) == [(0, 3, 0)]
    # Q5: longest mode drops the failing byte on reset
    assert sev(O.AC.compile(["ab", "b"]).match_longest(b"abb")) == [(0, 2, 0)]
    # Q7: neighbour bytes >= sep.size pass
    m = O.AC.compile(["a"])
    assert len(m.scan(b"\xe4a\xe4", 0, sep=[" "], sep_size=128)) == 1
    assert len(m.scan(b"\xe4a\xe4", 0, sep=[" "], sep_size=256)) == 0


def _rand_keys(rng, n, alphabet, maxlen):
    keys = set()
    space = sum(len(alphabet) ** l for l in range(1, maxlen + 1))
    n = min(n, space // 2)
    while len(keys) < n:
        L = rng.randint(1, maxlen)
        keys.add(bytes(rng.choice(alphabet) for _ in range(L)))
    keys = sorted(keys)  # set order of bytes depends on PYTHONHASHSEED: sort, then shuffle with the seeded generator
    rng.shuffle(keys)
    return keys


def test_q4_stale_end_flag_hand_derived():
    """SURVEY Q4, pinned by reasoning about the Crystal source, not by a second port.

    Keys "bba" (id 0), "bca" (id 1); a = 97, b = 98, c = 99.  Trace of cedar.cr / ac.cr, slot numbers in decimal:

    insert "bba" (cedar.cr:755-778 -> get 224-241 -> follow 244-259):
      * root 0 has base 0 (value -1): 'b' goes to slot 0 ^ 98 = 98, free, popped (pop_enode 327-357); root.child = 98.
      * node 98 is fresh (value = value_limit, base = -(MAX + 1) wraps negative): find_place (527-531) sees no block
        on the Closed / Open rings (block 0 is never on one), so add_block (298-313) creates block 1 (slots 256-511,
        ehead 256): 'b' lands in 256, node 98 gets base 256 ^ 98 = 354, child 98.
      * node 256 is fresh: find_place -> ehead of block 1 = 257: 'a' lands in 257, node 256 gets base 257 ^ 97 = 352.
      * node 257 keeps value_limit >= 0, so get returns 257 itself (line 240); insert sets value = 0 and end!
        (flags = 0x200, line 773-774).
    insert "bca":
      * 'b': slot 98 exists.  'c' from node 98: to = 354 ^ 99 = 257 -- occupied by the child of node 256 -> resolve
        (585-655) with from_n = 98, base_n = 354, label_n = 99, to_pn = 257, from_p = 256, base_p = 352.
        consult (428-430): child_num(98) = 1 < child_num(256) = 1 is false, so the OCCUPANT's family moves:
        children = ['a'], find_place -> ehead of block 1 = 258, base = 258 ^ 97 = 355, node 256 gets base 355.
        The child is copied to slot 258 (value 0, flags 0x200 copied at line 620, leafs[0] = 258).  Since
        to_ == to_pn (line 635) slot 257 is taken over for the new child: push_sibling (636), child = 0 (638),
        value = value_limit (639), check = 98 (640) -- and flags is NOT touched: slot 257, now the interior node
        "bc", still carries END = 0x200.  (push_enode, the only place that clears flags, 394-396, is the else branch.)
      * 'a' from node 257: fresh node -> find_place = 259; node 257 gets base 259 ^ 97 = 354, child 97, and
        child_num= (69-71) keeps the END bit: flags = 0x201.  Node 259 gets value = 1, end!.
    Result: is_end?(257 = "bc") is true (657-660) although "bc" is no key.

    ACX.compile (ac.cr:71-112): fails[98] = 0; children of 98 in sibling order: 'c' (257) then 'b' (256);
    fails[257] = 0, fails[256] = 98; node 257 is_end? -> vk = value(257) = -1 (726-734: its base slot 354 is free), so
    key_lens[-1] = 2 (unchecked write) and output[257].value = -1; fails[259] = 0, fails[258] = 0;
    output[259].value = 1, output[258].value = 0, key_lens = [3, 3].

    match_longest_("bcbba", false) (ac.cr:118-143): b -> 98; c -> 257, is_end? -> prev = (1, 257); b: child(257, 98)
    = 354 ^ 98 = 256 whose check is 98 -> miss, prev_i != -1 -> yield (1, 257): fetch_one (249-263) finds
    output[257].value = -1 and yields nothing; prev_i = -1, nid = 0, `break if nid == 0` drops this 'b' (Q5);
    b -> 98; a: child(98, 97) = 354 ^ 97 = 259 whose check is 257 -> miss, nid = fails[98] = 0, child(0, 97) misses,
    break.  End of text with prev_i == -1: NO hit at all -- a textbook longest-match reports "bba" at [2, 5).
    With intersectable = true nid stays 257 after the empty yield, fails to 0, then b -> 98, b -> 256, a -> 258 (end):
    [(2, 5, 0)].  Plain match (176-192, 265-278): [(2, 5, 0)] as well (the spurious end state yields an empty list).
    """
    keys = ["bba", "bca"]
    ac = O.AC.compile(keys)
    nodes, fails, output, kl = ac.arrays()
    nd = np.asarray(nodes).view(np.uint8).reshape(-1, 12)
    flags = nd[:, 10].astype(np.int32) | (nd[:, 11].astype(np.int32) << 8)
    check = nd[:, 4:8].copy().view("<i4").reshape(-1)
    value = nd[:, 0:4].copy().view("<i4").reshape(-1)
    assert (int(check[257]), int(flags[257]), int(value[257])) == (98, 0x201, -355)   # "bc": stale END, base 354
    assert (int(check[258]), int(flags[258]), int(value[258])) == (256, 0x200, 0)     # "bba" after the move
    assert (int(check[259]), int(flags[259]), int(value[259])) == (257, 0x200, 1)     # "bca"
    assert int(value[256]) == -356 and int(value[98]) == -355
    assert [int(fails[i]) for i in (98, 257, 256, 259, 258)] == [0, 0, 98, 0, 0]
    assert sev(ac.match_longest(b"bcbba", False)) == []
    assert sev(ac.match_longest(b"bcbba", True)) == [(2, 5, 0)]
    assert sev(ac.match(b"bcbba")) == [(2, 5, 0)]
    # the same dictionary without the collision: insertion order "bca", "bba" leaves no stale flag
    ac2 = O.AC.compile(["bca", "bba"])
    assert sev(ac2.match_longest(b"bcbba", False)) == [(2, 5, 1)]


def _has_stale_end_flag(nodes_u8, keys):
    """Q4: an interior node that carries the END flag (bit 9) although no key ends there."""
    nd = nodes_u8.view(np.dtype([("value", "<i4"), ("check", "<i4"), ("sibling", "u1"), ("child", "u1"), ("flags", "<u2")]))
    n = len(nd)

    def child(i, b):  # cedar.cr:441-447
        base = -(int(nd["value"][i]) + 1)
        if base < 0:
            return -1
        c = base ^ b
        return c if 0 <= c < n and int(nd["check"][c]) == i else -1

    ends, visited = set(), set()
    for k in keys:
        i = 0
        for b in k:
            i = child(i, b)
            assert i >= 0
            visited.add(i)
        ends.add(i)
    return any((int(nd["flags"][i]) & 0x200) and i not in ends for i in visited)


@pytest.mark.parametrize("seed", range(12))
def test_oracle_vs_independent_model(seed):
    rng = random.Random(seed)
    alphabet = [b"ab", b"abc", b"abcdefgh", bytes(range(1, 256))][seed % 4]
    keys = _rand_keys(rng, rng.choice([3, 30, 300, 2000]), alphabet, rng.choice([2, 4, 8]))
    text = bytes(rng.choice(alphabet) for _ in range(5000))
    ac = O.AC.compile(keys)
    py = PyAC(keys)
    assert sev(ac.match(text)) == py.match(text)
    sepset = {alphabet[0]}
    assert sev(ac.scan(text, 0, sep=list(sepset))) == py.match(text, sep=sepset)
    # longest modes agree with the model whenever no stale END flag exists (Q4): the model
    # assumes is_end == has key; detect disagreement on is_end and skip in that case.
    nodes, fails, output, kl = ac.arrays()
    for inter in (False, True):
        got = sev(ac.match_longest(text, inter))
        want = py.match_longest(text, inter)
        if got != want:
            assert _has_stale_end_flag(nodes, keys), "longest mode differs from the model without a stale END flag"
            pytest.xfail("stale END flag (Q4) changes longest-mode stream for seed %d" % seed)
    for i, k in enumerate(keys):
        assert ac.da.key(i) == k
        assert kl[i] == len(k)
