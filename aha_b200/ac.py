"""Host-side mirror of the reference's public API for the scan path: `Aha::Cedar` (insert / [] / []?),
`Aha::AC` (compile / match / match_longest / save / load) and `Aha::Hit`, on top of the C ABI.

Crystal is not available in this image, so this module plays the role of src/aha_gpu.cr: same
method names, argument meaning and error behaviour as src/aha/ac.cr:62-69, 280-340 and
src/aha/matcher.cr:34-46.  `str` arguments take the String overloads (character offsets),
`bytes`/uint8 arrays the Bytes overloads (byte offsets).  All matching runs on the GPU.
"""
import ctypes as C
import struct

import numpy as np

from . import _lib
from ._lib import AhaError, Hits, check, lib

UNITS_CHAR, SEP, LONGEST, INTERSECTABLE, TRIE_LONGEST, IGNORE_CASE, COUNT_ONLY, SEP_CODEPOINT = 1, 2, 4, 8, 16, 32, 64, 128

HIT_DTYPE = np.dtype([("start", "<i4"), ("end", "<i4"), ("value", "<i4")])
NODE_DTYPE = np.dtype([("value", "<i4"), ("check", "<i4"), ("sibling", "u1"), ("child", "u1"), ("flags", "<u2")])


def _b(x):
    return x.encode("utf-8") if isinstance(x, str) else bytes(x)


def sep_bitmap(sep, size=256):
    """BitArray(size) with the given bytes / 1-char strings set -> (32-byte bitmap, size)."""
    if size > 256:
        raise AhaError(-1, "sep BitArray size > 256 is not supported")  # ac.cr:322
    bits = bytearray(32)
    for s in sep:
        o = ord(s) if isinstance(s, str) else int(s)
        if o < size:
            bits[o >> 3] |= 1 << (o & 7)
    return bytes(bits), size


class Cedar:
    """Aha::Cedar: insertion-ordered key ids over a double-array trie (cedar.cr:751-778, 817-834)."""

    def __init__(self):
        h = C.c_void_p()
        check(lib().aha_b200_builder_new(C.byref(h)))
        self._h = h
        self._ac = None

    def __del__(self):
        try:
            lib().aha_b200_builder_free(self._h)
        except Exception:
            pass

    def insert(self, key):
        k = _b(key)
        out = C.c_int32()
        check(lib().aha_b200_builder_insert(self._h, k, len(k), C.byref(out)))
        self._ac = None  # the cached automaton is stale
        return out.value

    def get(self, key):  # []?
        k = _b(key)
        out = C.c_int32()
        check(lib().aha_b200_builder_lookup(self._h, k, len(k), C.byref(out)))
        return None if out.value < 0 else out.value

    def key_bytes(self, kid):
        buf = C.create_string_buffer(1 << 16)
        n = C.c_int32()
        rc = lib().aha_b200_builder_key(self._h, int(kid), buf, len(buf), C.byref(n))
        if rc != 0:
            raise IndexError(kid)
        return buf.raw[: n.value]

    def __getitem__(self, k):
        if isinstance(k, (int, np.integer)):
            return self.key_bytes(k).decode("utf-8", "surrogateescape")
        r = self.get(k)
        if r is None:
            raise IndexError(k)
        return r

    def match_longest(self, text, ignore_case=False, sep=None):
        """CedarX#match_longest(String, ignore_case, sep) (cedar.cr:1075-1121): leftmost-longest,
        character offsets.  Runs on the GPU over the trie part of the flattened automaton."""
        return self.match_longest_batch([text], ignore_case, sep).hits()

    def to_gpu(self):
        """CedarX#to_gpu (src/aha_gpu.cr): the trie alone on the GPU, through ACX.compile(da) (ac.cr:71-112)."""
        if self._ac is None:
            self._ac = AC.compile(self)
        return self._ac

    def match_longest_batch(self, docs, ignore_case=False, sep=None):
        if sep is not None and any((ord(c) if isinstance(c, str) else int(c)) >= 256 for c in sep):
            # cedar.cr:1071-1073 takes any Array(Char); the kernel's separator set is 256 bits: refuse instead of
            # silently ignoring the character
            raise AhaError(-1, "CedarX#match_longest on the GPU: separators must be code points below U+0100")
        self.to_gpu()
        return self._ac.scan_batch(docs, TRIE_LONGEST | (IGNORE_CASE if ignore_case else 0), sep)

    def gsub(self, text, ignore_case=False, sep=None, block=None):
        """CedarX#gsub(String, ignore_case, sep) { |hit| replacement } (cedar.cr:1123-1177).

        The reference's gsub loop is its match_longest loop (cedar.cr:1098-1121) with `sb <<` on every branch: each
        character is copied once, except inside a hit, where the block's return value goes instead.  So the string
        is rebuilt on the host from the GPU's TRIE_LONGEST hit stream.  `block` receives (start, end, value)."""
        return self.gsub_batch([text], ignore_case, sep, block)[0]

    def gsub_batch(self, docs, ignore_case=False, sep=None, block=None):
        if block is None:
            raise TypeError("gsub needs a block: hit -> replacement")
        res = self.match_longest_batch(docs, ignore_case, sep)
        hits, off = res.hits(), res.doc_hit_off()
        out = []
        for d, doc in enumerate(docs):
            s = doc if isinstance(doc, str) else bytes(doc).decode("utf-8")
            sb, pos = [], 0
            for h in hits[int(off[d]):int(off[d + 1])]:
                start, end = int(h["start"]), int(h["end"])
                sb.append(s[pos:start])
                sb.append(str(block((start, end, int(h["value"])))))
                pos = end
            sb.append(s[pos:])
            out.append("".join(sb))
        return out

    def compiled_arrays(self):
        """Runs ACX.compile(da) (ac.cr:71-112); returns copies of (nodes, fails, output, key_lens)."""
        nodes, fails, output, kl = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        n, k = C.c_int64(), C.c_int64()
        check(lib().aha_b200_builder_compile(self._h, C.byref(nodes), C.byref(n), C.byref(fails), C.byref(output),
                                             C.byref(kl), C.byref(k)))
        n, k = n.value, k.value
        a_nodes = np.frombuffer(C.string_at(nodes.value, n * 12), dtype=np.uint8).copy()
        a_fails = np.frombuffer(C.string_at(fails.value, n * 4), dtype=np.int32).copy()
        a_out = np.frombuffer(C.string_at(output.value, n * 8), dtype=np.int32).reshape(n, 2).copy()
        a_kl = np.frombuffer(C.string_at(kl.value, k * 4), dtype=np.uint32).copy() if k else np.zeros(0, np.uint32)
        return a_nodes, a_fails, a_out, a_kl


class _DevArr:
    def __init__(self, ptr, n, typestr, owner):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr or 0, False), "version": 2}
        self._owner = owner


class ScanResult:
    """Hit stream of one batch scan.  Host result: numpy views; device result: zero-copy torch tensors."""

    def __init__(self, ph):
        self._ph = ph
        h = ph.contents
        self.n_hits, self.n_docs, self.on_device, self.scan_ms = h.n_hits, h.n_docs, bool(h.on_device), h.scan_ms
        self.walk_ms, self.expand_ms = h.walk_ms, h.expand_ms

    def __del__(self):
        try:
            lib().aha_b200_hits_free(self._ph)
        except Exception:
            pass

    def hits(self):
        h = self._ph.contents
        if self.on_device:
            raise AhaError(-1, "device result: use hits_tensor()")
        if not h.hits or h.n_hits == 0:
            return np.zeros(0, HIT_DTYPE)
        return np.frombuffer(C.string_at(h.hits, h.n_hits * 12), dtype=HIT_DTYPE)

    def doc_hit_off(self):
        h = self._ph.contents
        if self.on_device:
            return self.doc_hit_off_tensor().cpu().numpy()
        return np.frombuffer(C.string_at(h.doc_hit_off, (h.n_docs + 1) * 8), dtype=np.int64)

    def hits_tensor(self):
        """int32 [n_hits, 3] CUDA tensor aliasing the library buffer (kept alive by this object)."""
        import torch
        h = self._ph.contents
        if h.n_hits == 0 or not h.hits:
            return torch.zeros((0, 3), dtype=torch.int32, device="cuda")
        return torch.as_tensor(_DevArr(h.hits, h.n_hits * 3, "<i4", self), device="cuda").view(-1, 3)

    def doc_hit_off_tensor(self):
        import torch
        h = self._ph.contents
        return torch.as_tensor(_DevArr(h.doc_hit_off, h.n_docs + 1, "<i8", self), device="cuda")


class AC:
    """Aha::AC (ac.cr).  compile -> flattened DFA resident on the GPU; match* run the CUDA scan."""

    def __init__(self, dfa_handle, arrays=None, cedar=None):
        self._dfa = dfa_handle
        self._arrays = arrays
        self._cedar = cedar
        self._leaf_of = None

    def __del__(self):
        try:
            lib().aha_b200_dfa_free(self._dfa)
        except Exception:
            pass

    # -- construction ------------------------------------------------------------------------
    @classmethod
    def compile(cls, keys):
        """AC.compile(keys) (ac.cr:62-69) or AC.compile(trie) (ac.cr:71-112)."""
        if isinstance(keys, Cedar):
            da = keys
        else:
            da = Cedar()
            for idx, k in enumerate(keys):
                kid = da.insert(k)
                if kid != idx:
                    raise AhaError(-3, "key:%s appear twice." % (k if isinstance(k, str) else _b(k).decode("utf-8", "replace")))
        arrays = da.compiled_arrays()
        return cls(cls._dfa_from_arrays(arrays), arrays, da)

    @staticmethod
    def _dfa_from_arrays(arrays):
        nodes, fails, output, kl = arrays
        h = C.c_void_p()
        check(lib().aha_b200_dfa_from_cedar(nodes.ctypes.data, len(fails), fails.ctypes.data, output.ctypes.data,
                                            kl.ctypes.data if len(kl) else None, len(kl), C.byref(h)))
        return h

    @classmethod
    def from_blob(cls, blob):
        a = np.frombuffer(blob, dtype=np.uint8)
        h = C.c_void_p()
        check(lib().aha_b200_dfa_from_blob(a.ctypes.data, a.size, C.byref(h)))
        return cls(h)

    @classmethod
    def from_device_blob(cls, tensor):
        h = C.c_void_p()
        check(lib().aha_b200_dfa_from_device_blob(tensor.data_ptr(), tensor.numel(), C.byref(h)))
        return cls(h)

    def blob_view(self):
        """The serialised DFA as a uint8 array over library-owned memory (valid while this object lives); no copy,
        so a multi-GB automaton (1 M patterns: 3.2 GB) can be handed to the broadcast as it is."""
        p, n = C.c_void_p(), C.c_int64()
        check(lib().aha_b200_dfa_blob(self._dfa, C.byref(p), C.byref(n)))
        a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,))
        a.flags.writeable = False
        return a

    def blob(self):
        return self.blob_view().tobytes()

    def info(self):
        names = ["n_states", "n_classes", "n_keys", "max_key_len", "n_out", "blob_bytes", "hot_states", "entry_bytes"]
        return {k: lib().aha_b200_dfa_info(self._dfa, i) for i, k in enumerate(names)}

    # -- save / load (ac.cr:13,45-60).  Container is our own: SuperIO's byte layout is unpinned. --
    MAGIC = b"AHAB200\x01"

    def save(self, path):
        if self._arrays is None:
            raise AhaError(-5, "automaton was built from a blob; nothing to save")
        with open(path, "wb") as f:
            f.write(self.MAGIC)
            for a in self._arrays:
                raw = np.ascontiguousarray(a).tobytes()
                f.write(struct.pack("<q", len(raw)))
                f.write(raw)

    @classmethod
    def load(cls, path):
        with open(path, "rb") as f:
            if f.read(8) != cls.MAGIC:
                raise AhaError(-1, "not an aha_b200 automaton file")
            parts = []
            for _ in range(4):
                head = f.read(8)
                if len(head) != 8:
                    raise AhaError(-1, "corrupt automaton file: truncated")
                (n,) = struct.unpack("<q", head)
                if n < 0:
                    raise AhaError(-1, "corrupt automaton file")
                parts.append(f.read(n))
                if len(parts[-1]) != n:
                    raise AhaError(-1, "corrupt automaton file: truncated")
        if len(parts[1]) % 4 or len(parts[3]) % 4:
            raise AhaError(-1, "corrupt automaton file")
        n = len(parts[1]) // 4
        if n < 256 or len(parts[0]) != 12 * n or len(parts[2]) != 8 * n:  # the file is an untrusted input
            raise AhaError(-1, "corrupt automaton file: array sizes disagree")
        nodes = np.frombuffer(parts[0], dtype=np.uint8).copy()
        fails = np.frombuffer(parts[1], dtype=np.int32).copy()
        output = np.frombuffer(parts[2], dtype=np.int32).reshape(-1, 2).copy()
        kl = np.frombuffer(parts[3], dtype=np.uint32).copy()
        arrays = (nodes, fails, output, kl)
        return cls(cls._dfa_from_arrays(arrays), arrays, None)

    # -- delegated dictionary lookups (ac.cr:41-43) -------------------------------------------
    def key_bytes(self, kid):
        if self._cedar is not None:
            return self._cedar.key_bytes(kid)
        if self._arrays is None:
            raise IndexError(kid)
        nodes = self._arrays[0].view(NODE_DTYPE)
        if self._leaf_of is None:
            out = self._arrays[2]
            own = np.nonzero(out[:, 1] >= 0)[0]
            self._leaf_of = {int(out[s, 1]): int(s) for s in own}
        at = self._leaf_of[int(kid)]
        rev = bytearray()
        while at > 0:  # cedar.cr:707-722
            up = int(nodes["check"][at])
            ch = (-(int(nodes["value"][up]) + 1)) ^ at
            if ch != 0:
                rev.append(ch)
            at = up
        return bytes(reversed(rev))

    def __getitem__(self, k):
        if isinstance(k, (int, np.integer)):
            return self.key_bytes(k).decode("utf-8", "surrogateescape")
        if self._cedar is None:
            raise IndexError(k)
        return self._cedar[k]

    # -- scanning ----------------------------------------------------------------------------
    def scan_batch(self, docs, mode=0, sep=None, sep_size=256):
        """Batch form over host documents: list of bytes/str, or (uint8 array, int64 doc_off)."""
        if isinstance(docs, tuple):
            text, doc_off = docs
            text = np.ascontiguousarray(text, dtype=np.uint8)
            doc_off = np.ascontiguousarray(doc_off, dtype=np.int64)
        else:
            raw = [_b(d) for d in docs]
            doc_off = np.zeros(len(raw) + 1, np.int64)
            np.cumsum([len(r) for r in raw], out=doc_off[1:])
            text = np.frombuffer(b"".join(raw), dtype=np.uint8)
        bits, size = (sep_bitmap(sep, sep_size) if sep is not None else (None, 0))
        if sep is not None:
            mode |= SEP
        ph = C.POINTER(Hits)()
        check(lib().aha_b200_scan(self._dfa, text.ctypes.data if text.size else None, doc_off.ctypes.data,
                                  len(doc_off) - 1, mode, bits, size, C.byref(ph)))
        return ScanResult(ph)

    def scan_device(self, text, doc_off, mode=0, sep=None, sep_size=256, stream=None):
        """text: uint8 CUDA tensor, doc_off: int64 CUDA tensor [n_docs+1].  Results stay on the device."""
        bits, size = (sep_bitmap(sep, sep_size) if sep is not None else (None, 0))
        if sep is not None:
            mode |= SEP
        if stream is None:
            import torch
            stream = torch.cuda.current_stream().cuda_stream
        ph = C.POINTER(Hits)()
        check(lib().aha_b200_scan_device(self._dfa, text.data_ptr(), doc_off.data_ptr(), doc_off.numel() - 1,
                                         text.numel(), mode, bits, size, stream, C.byref(ph)))
        return ScanResult(ph)

    def scan_device_noalloc(self, text, doc_off, hits_out, doc_hit_off_out, n_hits_out, stream=None):
        """aha_b200_scan_device_noalloc: ACX#match (byte offsets) into caller-provided CUDA tensors, no allocation,
        no host wait.  hits_out: int32 [cap, 3]; doc_hit_off_out: int64 [n_docs + 1]; n_hits_out: int64 [1].  At most
        64 MiB of text per call (static scratch); calls on one device must not overlap on the GPU."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream().cuda_stream
        check(lib().aha_b200_scan_device_noalloc(self._dfa, text.data_ptr(), doc_off.data_ptr(), doc_off.numel() - 1,
                                                 text.numel(), hits_out.data_ptr() if hits_out is not None else None,
                                                 hits_out.shape[0] if hits_out is not None else 0,
                                                 doc_hit_off_out.data_ptr(), n_hits_out.data_ptr(), stream))

    def _one(self, seq, mode, sep=None):
        """seq: bytes (byte offsets), str (String overloads: character offsets), or a list / tuple of 1-character strings
        (the Array(Char) / Slice(Char) overloads, ac.cr:288-295, 312-319, 342-364: character offsets, and with `sep` the
        neighbour test on the neighbouring character's code point)."""
        if isinstance(seq, str):
            mode |= UNITS_CHAR
        elif isinstance(seq, (list, tuple)) and all(isinstance(c, str) for c in seq):
            seq = "".join(seq)
            mode |= UNITS_CHAR | (SEP_CODEPOINT if sep is not None else 0)
        return self.scan_batch([seq], mode, sep).hits()

    def match(self, seq, sep=None):
        """ACX#match(seq) / match(seq, sep) (ac.cr:280-295, 321-364; matcher.cr:34-46)."""
        return self._one(seq, 0, sep)

    def match_longest(self, seq, intersectable=False):
        """ACX#match_longest (ac.cr:297-319)."""
        return self._one(seq, LONGEST | (INTERSECTABLE if intersectable else 0))
