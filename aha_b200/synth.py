"""Synthetic corpora and dictionaries of SURVEY.md 8(d) (generators live in csrc/gen.cuh)."""
import ctypes as C

import numpy as np

from ._lib import check, lib

EN, ZH, MIX = 0, 1, 2
BLOCK = 192
SEED_TEXT = 0xA11A0002
SEED_PAT = 0xA11A0003


def doc_bytes(target):
    """document size: the multiple of the 192-byte generator block closest below `target`."""
    return max(BLOCK, (target // BLOCK) * BLOCK)


_host_generator = None


def set_host_generator(fn):
    """fn(corpus, seed, first, nbytes) -> uint8 array.  bench.py's reference arm installs the oracle's generator
    (same include/aha_synth.h stream) so that it builds its corpus and patterns without loading libaha_b200.so."""
    global _host_generator
    _host_generator = fn


def gen_host(corpus, seed, first, nbytes):
    if _host_generator is not None:
        return _host_generator(corpus, seed, first, nbytes)
    out = np.empty(nbytes, np.uint8)
    check(lib().aha_b200_gen_text_host(corpus, seed, first, nbytes, out.ctypes.data))
    return out


def gen_device(corpus, seed, first, nbytes, device="cuda"):
    import torch
    out = torch.empty(nbytes, dtype=torch.uint8, device=device)
    check(lib().aha_b200_gen_text_device(corpus, seed, first, nbytes, out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out


def _sample_text(corpus, seed, nbytes=8 << 20):
    return gen_host(corpus, seed, 0, doc_bytes(nbytes)).tobytes()


def patterns_en_words(n, seed_text=SEED_TEXT, seed=SEED_PAT, min_len=3):
    """n distinct words in order of first appearance in the EN stream, thinned deterministically."""
    words = _sample_text(EN, seed_text).split()
    seen, out = set(), []
    for w in words:
        if len(w) >= min_len and w not in seen:
            seen.add(w)
            out.append(w)
    rng = np.random.default_rng(seed)
    if len(out) > n:
        idx = np.sort(rng.choice(len(out), n, replace=False))
        out = [out[i] for i in idx]
    return out


def patterns_zh_ngrams(n, seed_text=SEED_TEXT + 1, seed=SEED_PAT + 1, corpus=ZH, lens=(2, 3, 4)):
    """n distinct 2-4 character n-grams sampled at random character positions of the ZH stream."""
    text = _sample_text(corpus, seed_text, 16 << 20).decode("utf-8", "ignore")
    rng = np.random.default_rng(seed)
    seen, out = set(), []
    while len(out) < n:
        pos = rng.integers(0, len(text) - 8, size=4 * n)
        ls = rng.choice(lens, size=4 * n)
        for p, l in zip(pos, ls):
            g = text[p:p + l]
            if any(ord(c) < 0x3400 or ord(c) >= 0xFF00 for c in g) or g in seen:
                continue
            seen.add(g)
            out.append(g.encode())
            if len(out) == n:
                break
    return out


def patterns_mix(n, seed_text=SEED_TEXT + 2, seed=SEED_PAT + 2, en_frac=0.6):
    """60 % EN words / bigrams / trigrams, 40 % ZH n-grams, all sampled from the MIX stream."""
    raw = _sample_text(MIX, seed_text, 64 << 20 if n > 200000 else 24 << 20)
    rng = np.random.default_rng(seed)
    n_en = int(n * en_frac)
    toks = raw.split()
    toks = [t for t in toks if t[0] < 0x80]
    seen, out_en = set(), []
    order = rng.permutation(len(toks) - 3)
    for i in order:
        k = int(rng.integers(1, 4))
        g = b" ".join(toks[i:i + k])
        if len(g) < 3 or g in seen:
            continue
        seen.add(g)
        out_en.append(g)
        if len(out_en) == n_en:
            break
    out_zh = patterns_zh_ngrams(n - len(out_en), seed_text, seed + 7, corpus=MIX)
    out = out_en + [z for z in out_zh if z not in seen]
    perm = rng.permutation(len(out))
    return [out[i] for i in perm]


def patterns_en_phrases(n, seed_text=SEED_TEXT + 3, seed=SEED_PAT + 3):
    """Config C4: n distinct EN words and two-word phrases sampled from the EN stream, every third one
    with its case changed (upper / capitalised), for the ignore_case + sep scan of CedarX#match_longest."""
    toks = _sample_text(EN, seed_text, 16 << 20).split()
    rng = np.random.default_rng(seed)
    seen, out = set(), []
    while len(out) < n:
        pos = rng.integers(0, len(toks) - 2, size=4 * n)
        two = rng.random(4 * n) < 0.4
        style = rng.integers(0, 6, size=4 * n)
        for p, t, st in zip(pos, two, style):
            w = toks[p] + (b" " + toks[p + 1] if t else b"")
            if st == 0:
                w = w.upper()
            elif st == 1:
                w = w.capitalize()
            if len(w) < 3 or len(w) > 60 or w in seen:
                continue
            seen.add(w)
            out.append(w)
            if len(out) == n:
                break
    return out
