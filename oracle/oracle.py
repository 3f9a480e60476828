"""ctypes loader for the CPU oracle (oracle/aha_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(aha_b200/) must never import this module.

The classes mirror the reference API names (Aha::Cedar, Aha::AC; src/aha/cedar.cr,
src/aha/ac.cr) so the known-answer tests read like the reference specs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libaha_oracle.so")

UNITS_CHAR, SEP, LONGEST, INTERSECTABLE, SEP_CODEPOINT = 1, 2, 4, 8, 128

HIT_DTYPE = np.dtype([("start", "<i4"), ("end", "<i4"), ("value", "<i4")])


def build(force=False):
    src = os.path.join(_HERE, "aha_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B" if force else "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        vp, i32, i64, u8p = C.c_void_p, C.c_int32, C.c_int64, C.c_char_p
        L.orc_cedar_new.restype = vp
        L.orc_cedar_free.argtypes = [vp]
        L.orc_cedar_insert.argtypes = [vp, u8p, i32]
        L.orc_cedar_insert.restype = i32
        L.orc_cedar_lookup.argtypes = [vp, u8p, i32]
        L.orc_cedar_lookup.restype = i32
        L.orc_cedar_error.argtypes = [vp]
        L.orc_cedar_error.restype = C.c_char_p
        for f in ("orc_cedar_array_size", "orc_cedar_leaf_size", "orc_cedar_key_num"):
            getattr(L, f).argtypes = [vp]
            getattr(L, f).restype = i32
        L.orc_cedar_nodes.argtypes = [vp]
        L.orc_cedar_nodes.restype = vp
        L.orc_cedar_key.argtypes = [vp, i32, C.c_char_p, i32]
        L.orc_cedar_key.restype = i32
        L.orc_ac_compile.argtypes = [vp]
        L.orc_ac_compile.restype = vp
        L.orc_ac_free.argtypes = [vp]
        for f in ("orc_ac_fails", "orc_ac_output", "orc_ac_key_lens", "orc_ac_hits"):
            getattr(L, f).argtypes = [vp]
            getattr(L, f).restype = vp
        L.orc_ac_scan.argtypes = [vp, vp, i64, C.c_uint32, C.c_char_p, i32]
        L.orc_ac_scan.restype = i64
        L.orc_ac_scan_count.argtypes = [vp, vp, i64, C.POINTER(C.c_uint64)]
        L.orc_ac_scan_count.restype = i64
        L.orc_ac_scan_count_mt.argtypes = [vp, vp, vp, i64, i32]
        L.orc_ac_scan_count_mt.restype = i64
        L.orc_ac_scan_hash_mt.argtypes = [vp, vp, vp, i64, i32, C.POINTER(C.c_uint64)]
        L.orc_ac_scan_hash_mt.restype = i64
        L.orc_ac_scan_mode_mt.argtypes = [vp, vp, vp, i64, C.c_uint32, C.c_char_p, i32, i32, C.POINTER(C.c_uint64)]
        L.orc_ac_scan_mode_mt.restype = i64
        L.orc_gen_text.argtypes = [i32, C.c_uint64, i64, i64, vp, i32]
        L.orc_gen_text.restype = i32
        L.orc_cedar_match_longest.argtypes = [vp, vp, i64, i32, i32, C.c_char_p, i32, i32]
        L.orc_cedar_match_longest.restype = i64
        L.orc_trie_hits.restype = vp
        L.orc_cedar_match_longest_mt.argtypes = [vp, vp, vp, i64, i32, i32, C.c_char_p, i32, i32, C.POINTER(C.c_uint64)]
        L.orc_cedar_match_longest_mt.restype = i64
        _lib = L
    return _lib


def gen_text(corpus, seed, first, nbytes, threads=None):
    """include/aha_synth.h stream on the host cores (the reference arm of bench.py never loads the product)."""
    out = np.empty(nbytes, np.uint8)
    rc = lib().orc_gen_text(corpus, seed, first, nbytes, out.ctypes.data, threads or os.cpu_count() or 1)
    if rc != 0:
        raise ValueError("orc_gen_text: bad argument")
    return out


def _bytes(x):
    if isinstance(x, str):
        return x.encode("utf-8")
    return bytes(x)


def sep_bits(sep, size=256):
    """sep: iterable of byte values / 1-char strings -> (32-byte bitmap, BitArray size)."""
    bits = bytearray(32)
    for s in sep:
        o = ord(s) if isinstance(s, str) else int(s)
        if o < size:
            bits[o >> 3] |= 1 << (o & 7)
    return bytes(bits), size


def _as_u8(text):
    if isinstance(text, np.ndarray):
        a = np.ascontiguousarray(text, dtype=np.uint8)
    else:
        a = np.frombuffer(_bytes(text), dtype=np.uint8)
    return a


class Cedar:
    """Aha::Cedar (cedar.cr) -- insert / [] / []? / match_longest."""

    def __init__(self):
        self._L = lib()
        self._h = self._L.orc_cedar_new()

    def __del__(self):
        try:
            self._L.orc_cedar_free(self._h)
        except Exception:
            pass

    def insert(self, key):
        k = _bytes(key)
        r = self._L.orc_cedar_insert(self._h, k, len(k))
        if r == -2:
            raise RuntimeError(self._L.orc_cedar_error(self._h).decode())
        return r

    def get(self, key):  # []?
        k = _bytes(key)
        r = self._L.orc_cedar_lookup(self._h, k, len(k))
        return None if r < 0 else r

    def key(self, kid):  # trie[i]
        buf = C.create_string_buffer(1 << 16)
        n = self._L.orc_cedar_key(self._h, kid, buf, len(buf))
        if n < 0:
            raise IndexError(kid)
        return buf.raw[:n]

    def __getitem__(self, k):
        if isinstance(k, int):
            return self.key(k).decode("utf-8", "surrogateescape")
        r = self.get(k)
        if r is None:
            raise IndexError(k)
        return r

    @property
    def size(self):
        return self._L.orc_cedar_key_num(self._h)

    @property
    def array_size(self):
        return self._L.orc_cedar_array_size(self._h)

    @property
    def leaf_size(self):
        return self._L.orc_cedar_leaf_size(self._h)

    def nodes(self):
        """raw Node(Int32)[array_size] as a uint8 copy (12 bytes per node)."""
        n = self.array_size * 12
        return np.frombuffer(C.string_at(self._L.orc_cedar_nodes(self._h), n), dtype=np.uint8).copy()

    def match_longest(self, text, ignore_case=False, sep=None, early_exit=True):
        a = _as_u8(text)
        bits, size = (sep_bits(sep) if sep is not None else (bytes(32), 0))
        n = self._L.orc_cedar_match_longest(self._h, a.ctypes.data, a.size, int(ignore_case), int(sep is not None),
                                            bits, size, int(early_exit))
        buf = C.string_at(self._L.orc_trie_hits(), n * 12) if n else b""
        return np.frombuffer(buf, dtype=HIT_DTYPE).copy()

    def _mt(self, text, doc_off, ignore_case, sep, threads, want_hash):
        a = _as_u8(text)
        off = np.ascontiguousarray(doc_off, dtype=np.int64)
        bits, size = (sep_bits(sep) if sep is not None else (bytes(32), 0))
        h = C.c_uint64()
        n = self._L.orc_cedar_match_longest_mt(self._h, a.ctypes.data, off.ctypes.data, len(off) - 1, int(ignore_case),
                                               int(sep is not None), bits, size, threads, C.byref(h) if want_hash else None)
        return n, h.value

    def match_longest_count_mt(self, text, doc_off, ignore_case, sep, threads):
        """hits of CedarX#match_longest over a batch of documents, document-parallel (early-exit form, Q9)"""
        return self._mt(text, doc_off, ignore_case, sep, threads, False)[0]

    def match_longest_hash_mt(self, text, doc_off, ignore_case, sep, threads):
        """(hit count, order-sensitive checksum of the per-document streams)"""
        return self._mt(text, doc_off, ignore_case, sep, threads, True)

    def gsub(self, text, ignore_case=False, sep=None, block=None):
        """CedarX#gsub(String, ignore_case, sep) { |hit| ... } (cedar.cr:1123-1177), restated loop by loop.

        `prefix(s + offset, ...)` with the longest-wins rule of cedar.cr:1153-1160 is played by the C restatement of
        match_longest (cedar.cr:1098-1121) run on the suffix: at a non-separator offset its first iteration is that
        very computation, so the longest prefix match is its first hit iff that hit starts at 0.  O(n^2): tests only."""
        s = text if isinstance(text, str) else bytes(text).decode("utf-8")
        seps = None if sep is None else set(c if isinstance(c, str) else chr(c) for c in sep)
        sb, offset = [], 0
        while offset < len(s):
            if seps is not None:
                while offset < len(s) and s[offset] in seps:  # cedar.cr:1148-1151
                    sb.append(s[offset])
                    offset += 1
                if offset >= len(s):
                    break
            h = self.match_longest(s[offset:], ignore_case, sep)
            last_vk, last_char_num = (int(h[0]["value"]), int(h[0]["end"])) if len(h) and h[0]["start"] == 0 else (0, 0)
            if last_char_num == 0:  # cedar.cr:1161-1171
                sb.append(s[offset])
                offset += 1
                if seps is not None:
                    while offset < len(s) and s[offset] not in seps:
                        sb.append(s[offset])
                        offset += 1
            else:  # cedar.cr:1172-1176
                sb.append(str(block((offset, offset + last_char_num, last_vk))))
                offset += last_char_num
        return "".join(sb)


class AC:
    """Aha::AC (ac.cr)."""

    def __init__(self, cedar):
        self._L = lib()
        self.da = cedar
        self._h = self._L.orc_ac_compile(cedar._h)

    @classmethod
    def compile(cls, keys):
        if isinstance(keys, Cedar):
            return cls(keys)
        da = Cedar()
        for idx, k in enumerate(keys):
            kid = da.insert(k)
            if kid != idx:
                raise RuntimeError("key:%r appear twice." % (k,))
        return cls(da)

    def __del__(self):
        try:
            self._L.orc_ac_free(self._h)
        except Exception:
            pass

    def __getitem__(self, k):
        return self.da[k]

    def arrays(self):
        """(nodes u8[array_size*12], fails i32[array_size], output i32[array_size,2], key_lens u32[leaf_size])"""
        n, k = self.da.array_size, max(self.da.leaf_size, 1)
        fails = np.frombuffer(C.string_at(self._L.orc_ac_fails(self._h), n * 4), dtype=np.int32).copy()
        output = np.frombuffer(C.string_at(self._L.orc_ac_output(self._h), n * 8), dtype=np.int32).reshape(n, 2).copy()
        kl = np.frombuffer(C.string_at(self._L.orc_ac_key_lens(self._h), k * 4), dtype=np.uint32).copy()
        return self.da.nodes(), fails, output, kl[: self.da.leaf_size]

    def scan(self, text, mode=0, sep=None, sep_size=256):
        a = _as_u8(text)
        bits, size = (sep_bits(sep, sep_size) if sep is not None else (bytes(32), 0))
        if sep is not None:
            mode |= SEP
        n = self._L.orc_ac_scan(self._h, a.ctypes.data, a.size, mode, bits, size)
        buf = C.string_at(self._L.orc_ac_hits(self._h), n * 12) if n else b""
        return np.frombuffer(buf, dtype=HIT_DTYPE).copy()

    # reference-shaped entry points; str -> character offsets, bytes -> byte offsets
    def match(self, seq, sep=None):
        if isinstance(seq, (list, tuple)):  # Array(Char) / Slice(Char) (ac.cr:288-295, 342-364)
            return self.scan("".join(seq), UNITS_CHAR | (SEP_CODEPOINT if sep is not None else 0), sep)
        return self.scan(seq, UNITS_CHAR if isinstance(seq, str) else 0, sep)

    def match_longest(self, seq, intersectable=False):
        if isinstance(seq, (list, tuple)):  # ac.cr:312-319
            seq = "".join(seq)
        m = LONGEST | (INTERSECTABLE if intersectable else 0) | (UNITS_CHAR if isinstance(seq, str) else 0)
        return self.scan(seq, m)

    def scan_count(self, text):
        a = _as_u8(text)
        cs = C.c_uint64(0)
        n = self._L.orc_ac_scan_count(self._h, a.ctypes.data, a.size, C.byref(cs))
        return n, cs.value

    def scan_count_mt(self, text, doc_off, threads):
        a = _as_u8(text)
        off = np.ascontiguousarray(doc_off, dtype=np.int64)
        return self._L.orc_ac_scan_count_mt(self._h, a.ctypes.data, off.ctypes.data, len(off) - 1, threads)

    def scan_mode_mt(self, text, doc_off, threads, mode=0, sep=None, sep_size=256):
        """any mode of scan() over a batch: (hit count, order-sensitive checksum of the per-document hit streams)"""
        a = _as_u8(text)
        off = np.ascontiguousarray(doc_off, dtype=np.int64)
        bits, size = (sep_bits(sep, sep_size) if sep is not None else (bytes(32), 0))
        if sep is not None:
            mode |= SEP
        h = C.c_uint64()
        n = self._L.orc_ac_scan_mode_mt(self._h, a.ctypes.data, off.ctypes.data, len(off) - 1, mode, bits, size, threads, C.byref(h))
        return n, h.value

    def scan_hash_mt(self, text, doc_off, threads):
        """(hit count, order-sensitive 64-bit checksum of the per-document hit streams)"""
        a = _as_u8(text)
        off = np.ascontiguousarray(doc_off, dtype=np.int64)
        h = C.c_uint64()
        n = self._L.orc_ac_scan_hash_mt(self._h, a.ctypes.data, off.ctypes.data, len(off) - 1, threads, C.byref(h))
        return n, h.value
