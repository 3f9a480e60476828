"""Independent plain-trie model of the reference's Aho-Corasick rules (dict-of-dicts trie, no
double array).  Used only to cross-check the cedar-layout oracle: different data structure,
same rules (ac.cr:71-112, 118-143, 176-192, 249-278).  Valid when no stale END
flags exist (Q4), i.e. is_end(node) == "node holds a key"."""
from collections import deque


class PyAC:
    def __init__(self, keys):
        self.keys = [k.encode() if isinstance(k, str) else bytes(k) for k in keys]
        self.goto = [{}]
        self.val = [-1]
        for i, k in enumerate(self.keys):
            s = 0
            for b in k:
                if b not in self.goto[s]:
                    self.goto.append({})
                    self.val.append(-1)
                    self.goto[s][b] = len(self.goto) - 1
                s = self.goto[s][b]
            assert self.val[s] == -1, "duplicate key"
            self.val[s] = i
        n = len(self.goto)
        self.fail = [0] * n
        self.onext = [-1] * n
        q = deque()
        for b, c in self.goto[0].items():
            q.append(c)
        while q:
            s = q.popleft()
            for b, c in self.goto[s].items():
                q.append(c)
                fid = s
                while fid != 0:
                    fs = self.fail[fid]
                    if b in self.goto[fs]:
                        fid = self.goto[fs][b]
                        break
                    fid = self.fail[fid]
                self.fail[c] = fid
                if self.val[fid] >= 0:  # "is_end?(fid)" (ac.cr:106-108)
                    self.onext[c] = fid

    def _fetch(self, i, s, one=False):
        out = []
        while self.val[s] >= 0:
            v = self.val[s]
            out.append((i - len(self.keys[v]) + 1, i + 1, v))
            if one or self.onext[s] < 0:
                break
            s = self.onext[s]
        return out

    def match(self, text, sep=None, sep_size=256):
        hits = []
        s = 0
        n = len(text)

        def rejects(ch):
            return ch < sep_size and ch not in sep

        for i, b in enumerate(text):
            while True:
                if b in self.goto[s]:
                    s = self.goto[s][b]
                    if self.val[s] >= 0:
                        if sep is not None and i + 1 < n and rejects(text[i + 1]):
                            break
                        for h in self._fetch(i, s):
                            if sep is not None and h[0] > 0 and rejects(text[h[0] - 1]):
                                continue
                            hits.append(h)
                    break
                if s == 0:
                    break
                s = self.fail[s]
        return hits

    def match_longest(self, text, intersectable=False):
        hits = []
        s = 0
        prev_i, prev_s = -1, -1
        for i, b in enumerate(text):
            while True:
                if b in self.goto[s]:
                    s = self.goto[s][b]
                    if self.val[s] >= 0:
                        prev_i, prev_s = i, s
                    break
                if prev_i != -1:
                    hits += self._fetch(prev_i, prev_s, one=True)
                    prev_i = -1
                    if not intersectable:
                        s = 0
                if s == 0:
                    break
                s = self.fail[s]
        if prev_i != -1:
            hits += self._fetch(prev_i, prev_s, one=True)
        return hits
