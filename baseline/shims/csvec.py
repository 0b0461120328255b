"""Stand-in for the external ``csvec`` package (github.com/nikitaivkin/csh) that the
reference imports but does not ship or pin (reference: fed_aggregator.py:5,
fed_worker.py:10; setup.py:11-13 has the dependency commented out) and that is not
installed on this image.

Pure PyTorch, same algorithmic structure as the public package -- per-row and per-block
Python loops over ATen ``bincount`` / gather / ``median`` / ``topk`` -- so the reference
baseline pays the cost profile it would pay with the real dependency.  None of the
commefficient_b200 kernels are used here.
"""
import numpy as np
import torch

import _ref_compat  # noqa: F401

LARGEPRIME = 2 ** 61 - 1
cache = {}


class CSVec(object):
    def __init__(self, d, c, r, doInitialize=True, device=None, numBlocks=1):
        self.r, self.c, self.d = r, c, int(d)
        self.numBlocks = numBlocks
        if device is None:
            device = "cuda" if torch.cuda.is_available() else "cpu"
        self.device = device
        if not doInitialize:
            return
        self.table = torch.zeros((r, c), device=self.device)
        key = (d, c, r, numBlocks, str(device))
        if key in cache:
            hit = cache[key]
            self.signs, self.buckets = hit["signs"], hit["buckets"]
            if numBlocks > 1:
                self.blockSigns, self.blockOffsets = hit["blockSigns"], hit["blockOffsets"]
            return
        rand_state = torch.random.get_rng_state()
        torch.random.manual_seed(42)
        hashes = torch.randint(0, LARGEPRIME, (r, 6), dtype=torch.int64, device="cpu")
        if numBlocks > 1:
            nTokens = self.d // numBlocks
            if self.d % numBlocks != 0:
                nTokens += 1
            self.blockSigns = (torch.randint(0, 2, size=(numBlocks,), device="cpu") * 2 - 1).to(self.device)
            self.blockOffsets = torch.randint(0, self.c, size=(numBlocks,), device="cpu").to(self.device)
        else:
            nTokens = self.d
        torch.random.set_rng_state(rand_state)
        tokens = torch.arange(nTokens, dtype=torch.int64, device="cpu").reshape((1, nTokens))
        h1, h2, h3, h4 = hashes[:, 2:3], hashes[:, 3:4], hashes[:, 4:5], hashes[:, 5:6]
        self.signs = (((h1 * tokens + h2) * tokens + h3) * tokens + h4)
        self.signs = ((self.signs % LARGEPRIME % 2) * 2 - 1).float().to(self.device)
        h1, h2 = hashes[:, 0:1], hashes[:, 1:2]
        self.buckets = (((h1 * tokens) + h2) % LARGEPRIME % self.c).to(self.device)
        cache[key] = {"signs": self.signs, "buckets": self.buckets}
        if numBlocks > 1:
            cache[key].update({"blockSigns": self.blockSigns, "blockOffsets": self.blockOffsets})

    def zero(self):
        self.table.zero_()

    def __deepcopy__(self, memodict={}):
        new = CSVec(d=self.d, c=self.c, r=self.r, doInitialize=False, device=self.device,
                    numBlocks=self.numBlocks)
        new.table = self.table.clone()
        new.signs, new.buckets = self.signs, self.buckets
        if self.numBlocks > 1:
            new.blockSigns, new.blockOffsets = self.blockSigns, self.blockOffsets
        return new

    def __imul__(self, other):
        if isinstance(other, (int, float)):
            self.table = self.table.mul_(other)
        else:
            raise ValueError("Can't multiply a CSVec by {}".format(other))
        return self

    def __truediv__(self, other):
        if isinstance(other, (int, float)):
            self.table = self.table.div_(other)
        else:
            raise ValueError("Can't divide a CSVec by {}".format(other))
        return self

    def __add__(self, other):
        import copy
        out = copy.deepcopy(self)
        out += other
        return out

    def __iadd__(self, other):
        if isinstance(other, CSVec):
            self.accumulateCSVec(other)
        else:
            raise ValueError("Can't add this to a CSVec: {}".format(other))
        return self

    def accumulateVec(self, vec):
        assert len(vec.size()) == 1 and vec.size()[0] == self.d
        for r in range(self.r):
            buckets = self.buckets[r, :].to(self.device)
            signs = self.signs[r, :].to(self.device)
            for blockId in range(self.numBlocks):
                start = blockId * buckets.size()[0]
                end = min((blockId + 1) * buckets.size()[0], self.d)
                offsetBuckets = buckets[:end - start].clone()
                offsetSigns = signs[:end - start].clone()
                if self.numBlocks > 1:
                    offsetBuckets += self.blockOffsets[blockId]
                    offsetBuckets %= self.c
                    offsetSigns *= self.blockSigns[blockId]
                self.table[r, :] += torch.bincount(input=offsetBuckets,
                                                   weights=offsetSigns * vec[start:end],
                                                   minlength=self.c)

    def accumulateCSVec(self, csvec):
        assert self.d == csvec.d and self.c == csvec.c and self.r == csvec.r
        assert self.numBlocks == csvec.numBlocks
        self.table += csvec.table

    def accumulateTable(self, table):
        if table.size() != self.table.size():
            raise ValueError("Passed in table has size {}, expecting {}".format(table.size(), self.table.size()))
        self.table += table

    def _findAllValues(self):
        if self.numBlocks == 1:
            vals = torch.zeros(self.r, self.d, device=self.device)
            for r in range(self.r):
                vals[r] = self.table[r, self.buckets[r, :]] * self.signs[r, :]
            return vals.median(dim=0)[0]
        medians = torch.zeros(self.d, device=self.device)
        n = self.buckets.size()[1]
        for blockId in range(self.numBlocks):
            start = blockId * n
            end = min((blockId + 1) * n, self.d)
            vals = torch.zeros(self.r, end - start, device=self.device)
            for r in range(self.r):
                buckets = self.buckets[r, :end - start]
                signs = self.signs[r, :end - start]
                offsetBuckets = (buckets + self.blockOffsets[blockId]) % self.c
                offsetSigns = signs * self.blockSigns[blockId]
                vals[r] = self.table[r, offsetBuckets] * offsetSigns
            medians[start:end] = vals.median(dim=0)[0]
        return medians

    def _findHHK(self, k):
        vals = self._findAllValues()
        outVals = torch.zeros(k, device=vals.device)
        HHs = torch.zeros(k, device=vals.device).long()
        torch.topk(vals ** 2, k, sorted=False, out=(outVals, HHs))
        return HHs, vals[HHs]

    def _findHHThr(self, thr):
        vals = self._findAllValues()
        HHs = vals.abs() >= thr
        return HHs, vals[HHs]

    def unSketch(self, k=None, epsilon=None):
        if epsilon is None:
            thr = None
        else:
            thr = epsilon * self.l2estimate()
        hhs = self._findHHK(k) if k is not None else self._findHHThr(thr)
        if k is not None:
            assert len(hhs[1]) == k
        unSketched = torch.zeros(self.d, device=self.device)
        unSketched[hhs[0]] = hhs[1]
        return unSketched

    def l2estimate(self):
        return np.sqrt(torch.median(torch.sum(self.table ** 2, 1)).item())

    @classmethod
    def median(cls, csvecs):
        d, c, r = csvecs[0].d, csvecs[0].c, csvecs[0].r
        tables = [csvec.table for csvec in csvecs]
        med = torch.median(torch.stack(tables), dim=0)[0]
        out = CSVec(d, c, r, doInitialize=False, device=csvecs[0].device, numBlocks=csvecs[0].numBlocks)
        out.table = med
        return out
