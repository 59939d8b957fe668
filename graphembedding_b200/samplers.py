"""Pair samplers with the reference's exact draw order (model/Siamese/samplers.py:19-65).

``RandomSampler``: consecutive neighbours of a list that is reshuffled IN PLACE with a fresh
``random.Random(123)`` on every wrap (:24-31); the list object is shared with the owning
``ModelGraphList`` (F10), which this mirror keeps.  ``DistributionSampler``: density-sorted
bins (:37-65)."""
import random

import networkx as nx


class Sampler(object):
    def __init__(self, gs, sample_num, sampler_duplicate_removal):
        assert len(gs) >= 2
        self.gs = gs
        self.sample_num = sample_num
        self.sampler_duplicate_removal = sampler_duplicate_removal

    def get_pair(self):
        raise NotImplementedError()

    def get_triple_for_hinge_loss(self):
        raise NotImplementedError()


class RandomSampler(Sampler):
    def __init__(self, gs, sample_num, sampler_duplicate_removal):
        super().__init__(gs, sample_num, sampler_duplicate_removal)
        self.idx = 0

    def get_pair(self):
        first = self.gs[self.idx]
        nxt = self.idx + 1
        if nxt >= len(self.gs):
            random.Random(123).shuffle(self.gs)  # shared list, fixed seed (samplers.py:28)
            nxt = 0
        self.idx = nxt
        return first, self.gs[nxt]


class DistributionSampler(Sampler):
    def __init__(self, gs, sample_num, sampler_duplicate_removal, bin_size=5):
        super().__init__(gs, sample_num, sampler_duplicate_removal)
        dens = [nx.density(g.get_nxgraph()) for g in self.gs]
        self.dens_list = sorted((d, i) for i, d in enumerate(dens))
        self.bin_size = bin_size
        self.bin_number = int(len(self.gs) / bin_size)
        self.bin_idx = self.shuffle_idx()
        self.item_idx = random.Random(123).randint(0, bin_size - 1)
        self.cur = 0

    def shuffle_idx(self):
        bins = list(range(self.bin_number))
        random.Random(123).shuffle(bins)
        if self.sample_num > 0:
            bins = bins[:2 * self.sample_num]
        return bins

    def _pick(self, slot):
        return self.gs[self.dens_list[self.bin_idx[slot] * self.bin_size + self.item_idx][1]]

    def get_pair(self):
        pair = (self._pick(self.cur), self._pick(self.cur + 1))
        self.cur += 2
        if self.cur >= len(self.bin_idx) - 1:
            self.cur = 0
        self.item_idx = random.Random(123 + self.cur).randint(0, self.bin_size - 1)
        return pair
