"""Drop-in for the reference's utils/input_partition.py (hot-path part: the partition grid).

`partition` and `partitioned_ranges` keep the reference signatures and return types (dict / list of dicts),
but the list is a `PartitionList`: it remembers the grid it came from, so `utils.prune.sound_prune*` can
hand whole index ranges to the GPU (device-side box generation, `fb_grid_boxes`) instead of rebuilding each
box from its dict.  (For grids too large to hold as dicts -- stress/AC has 3.3 M -- use
fairify_b200.engine.Grid directly; it addresses partitions by index and never builds them on the host.)
"""
import copy
import itertools
import random  # noqa: F401  (the reference's partitioned_ranges_df uses `random` without importing it)

import numpy as np   # noqa: F401  (leaked to the drivers through `import *`, like the reference)
import pandas as pd  # noqa: F401

from utils.verif_utils import *  # noqa: F401,F403  (the reference re-exports verif_utils from here)


def partition(r_dict, partition_size):
    """Chunks of `partition_size` for every attribute whose range is larger than that
    (reference utils/input_partition.py:17-46)."""
    partition_dict = {}
    for col in r_dict.keys():
        low, high = r_dict[col][0], r_dict[col][1]
        if high - low + 1 <= partition_size:
            continue
        parts = []
        cur = low
        while cur <= high:
            parts.append([cur, min(cur + partition_size - 1, high)])
            cur += partition_size
        partition_dict[col] = parts
    return partition_dict


class PartitionBox(dict):
    """One partition: a plain dict {attr: [lo, hi]} that also knows its canonical index in its grid."""
    __slots__ = ("grid_ref", "index")


class PartitionList(list):
    """list of PartitionBox in itertools.product order (shuffle-able like any list)."""

    def __init__(self, items, A, p_dict, range_dict):
        super().__init__(items)
        self.attributes = list(A)
        self.p_dict = p_dict
        self.range_dict = {k: list(v) for k, v in range_dict.items()}

    def position(self, box):
        """Index of `box` (by identity) in the list's CURRENT order, or None; survives `shuffle`."""
        pos = self.__dict__.get("_pos")
        k = pos.get(id(box)) if pos is not None else None
        if k is None or k >= len(self) or self[k] is not box:
            pos = self._pos = {id(b): i for i, b in enumerate(self)}
            k = pos.get(id(box))
            if k is not None and self[k] is not box:
                k = None
        return k

    def uniform_threshold(self):
        """The partition size if every split attribute was cut with the same size (what `partition` does)."""
        sizes = {parts[0][1] - parts[0][0] + 1 for parts in self.p_dict.values() if len(parts) > 1}
        return sizes.pop() if len(sizes) == 1 else None


def partitioned_ranges(A, PA, p_dict, range_dict):
    """Cartesian product of the per-attribute chunks, last partitioned attribute fastest; un-partitioned
    attributes first in each dict (reference utils/input_partition.py:48-76)."""
    base = [(attr, range_dict[attr]) for attr in A if attr not in p_dict]
    keys = list(p_dict.keys())
    out = []
    for idx, comb in enumerate(itertools.product(*[p_dict[k] for k in keys])):
        # every box owns its [lo, hi] lists, like the reference's deepcopy, at a tenth of its cost
        box = PartitionBox([(attr, list(r)) for attr, r in base])
        for k, part in zip(keys, comb):
            box[k] = list(part)
        box.index = idx
        out.append(box)
    plist = PartitionList(out, A, p_dict, range_dict)
    for box in plist:
        box.grid_ref = plist
    return plist


def partition_df(r_dict, partition_size):
    """Default-credit variant (reference :78-109): split only when hi - lo > partition_size."""
    partition_dict = {}
    for col, (low, high) in r_dict.items():
        if high - low <= partition_size:
            continue
        parts = []
        cur = low
        while cur < high:
            top = min(cur + partition_size - 1, high)
            parts.append([cur, top])
            cur = top + 1
        if parts:
            partition_dict[col] = parts
    return partition_dict


def partitioned_ranges_df(A, PA, p_dict, range_dict, max_partitions=100):
    """Capped variant (reference :111-182): protected attributes are partitioned first, others only while the
    product stays <= max_partitions; beyond that a random sample of the combinations is kept."""
    new_ranges = {attr: range_dict[attr] for attr in A if attr not in p_dict}
    if not p_dict:
        return [new_ranges]
    chosen = [a for a in PA if a in p_dict]
    est = 1
    for a in chosen:
        est *= len(p_dict[a])
    for a in [a for a in p_dict if a not in chosen]:
        if est * len(p_dict[a]) <= max_partitions:
            est *= len(p_dict[a]); chosen.append(a)
        else:
            new_ranges[a] = range_dict[a]
    if not chosen:
        return [new_ranges]
    combs = list(itertools.product(*[p_dict[a] for a in chosen]))
    if len(combs) > max_partitions:
        combs = random.sample(combs, max_partitions)
    out = []
    for comb in combs:
        box = copy.deepcopy(new_ranges)
        for a, part in zip(chosen, comb):
            box[a] = part
        out.append(box)
    return out


def p_list_density(range_dict, p_list, df):
    """Fraction of dataset rows falling in each partition (reference :184-218).  For a list that came from
    `partitioned_ranges` the rows are located on the device by index arithmetic (`fb_grid_locate`); any other list
    of boxes is compared on the host."""
    data = df.drop(labels=['income-per-year'], axis=1, inplace=False)
    total = df.shape[0]
    if isinstance(p_list, PartitionList) and p_list.uniform_threshold() is not None and len(p_list) > 0 \
            and all(isinstance(p, PartitionBox) for p in p_list):
        from fairify_b200 import engine
        cols = list(data.columns)
        grid = engine.Grid(cols, p_list.range_dict, p_list.uniform_threshold())
        if len(grid) == len(p_list):
            _, dens = grid.locate(data[cols].to_numpy().astype(np.float64), want_index=False, want_density=True)
            dens = dens.cpu().numpy()
            return [int(dens[p.index]) / total for p in p_list]
    cols = list(range_dict.keys())
    X = data[cols].to_numpy()
    out = []
    for p in p_list:
        lo = np.array([p[c][0] for c in cols]); hi = np.array([p[c][1] for c in cols])
        out.append(int(np.all((X >= lo) & (X <= hi), axis=1).sum()) / total)
    return out
