"""Sharding of the analysed individuals over ranks (one engine per GPU) -- SURVEY.md section 8(e).

Each (analysed individual, chromosome) unit is independent (cnF2freq.cpp:5283-5295); the only coupling between
units is the additive accumulation into the statistics rows of the individuals of their focus trees
(:5893-5902).  Analysed individuals are dealt to ranks in sibship blocks (same two parents together, so a
rank touches as few parameter rows as possible); the rows that more than one rank contributes to are the
payload of the one sum all-reduce per chromosome (the stale MPI reduce at :6241-6256).
"""
import numpy as np


def focus_tree(parents, ind):
    """Ids of the 7-slot focus tree of `ind` (fixtrees, cnF2freq.cpp:3099-3187); 0 = nobody."""
    ids = [int(ind)]
    for p in parents[ind]:
        ids.append(int(p))
        ids.extend(int(x) for x in (parents[p] if p else (0, 0)))
    return [i for i in ids if i]


def shard_dous(parents, dous, world):
    """Deals the analysed individuals to `world` ranks: sibships (same parent pair) stay together, the largest
    first, each to the least loaded rank.  Returns a list of int32 arrays, each in the original dous order."""
    parents = np.asarray(parents)
    dous = np.asarray(dous, np.int32)
    groups = {}
    for pos, d in enumerate(dous):
        groups.setdefault((int(parents[d, 0]), int(parents[d, 1])), []).append(pos)
    load = [0] * world
    picks = [[] for _ in range(world)]
    for key in sorted(groups, key=lambda k: (-len(groups[k]), k)):
        members = groups[key]
        # a sibship larger than a fair share is split: units are independent, only the row traffic grows
        fair = -(-len(dous) // world)
        for i in range(0, len(members), fair):
            r = min(range(world), key=lambda q: (load[q], q))
            chunk = members[i:i + fair]
            picks[r].extend(chunk)
            load[r] += len(chunk)
    return [dous[np.sort(np.asarray(p, np.int64))] if p else np.zeros(0, np.int32) for p in picks]


def shared_rows(parents, shards):
    """Individuals whose statistics rows receive contributions from more than one rank, sorted by id
    (the same list, in the same order, on every rank: cnf2_set_shared_rows)."""
    parents = np.asarray(parents)
    seen = {}
    for r, sh in enumerate(shards):
        for d in sh:
            for i in focus_tree(parents, d):
                seen.setdefault(i, set()).add(r)
    return np.array(sorted(i for i, rs in seen.items() if len(rs) > 1), np.int32)
