"""Synthetic ORB vocabularies in the text format DBoW2's TemplatedVocabulary::loadFromTextFile reads
(/root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1421; written by saveToTextFile :1423-1450):

    line 1:       k L scoringType weightingType
    node lines:   parentId isLeaf d0 d1 ... d31 weight        (node ids 1, 2, ... in file order; 0 is the root)

The reference's vocabulary blob (Vocabulary/ORBvoc.txt.tar.gz, k = 10, L = 6, L1 scoring, TF-IDF weighting) is not in
the reference tree, so only the FORMAT and the ALGORITHM are exercised: a hierarchical random tree -- the children of
a node are its descriptor with bits flipped at a rate that falls with depth, as k-means children of a cluster would
be --, every leaf at depth L, weights random with a few stop words (weight 0, transform() skips them :1154, :1185).

The file is written WITHOUT a newline after the last node: the reference's reader loops `while(!f.eof())`, so a final
newline makes it parse one more, empty, line into a node with an indeterminate descriptor hung under the root.

Deterministic in (k, L, seed, ...): tests generate what they need into a temporary directory; nothing large is committed.
Run as a script to write one file:  python tests/golden/make_vocabulary.py out.txt 10 6 [seed]
"""
from __future__ import annotations

import sys

import numpy as np


def make_tree(k: int, L: int, seed: int = 0, ragged: bool = False, stop_rate: float = 0.03):
    """-> parent [N] int32, is_leaf [N] uint8, desc [N, 32] uint8, weight [N] float64 for nodes 1..N (index i = node
    i + 1), in the order saveToTextFile writes them: a parent's children are consecutive, parents before children."""
    rng = np.random.default_rng(1000 * seed + 17 * k + L)
    flip = [0.5, 0.22, 0.16, 0.12, 0.09, 0.07, 0.05, 0.04, 0.03, 0.02]
    parent, leaf, desc, weight = [], [], [], []
    # breadth first: level by level, (node id, descriptor bits)
    frontier = [(0, None)]
    next_id = 1
    for level in range(1, L + 1):
        new_frontier = []
        for pid, pbits in frontier:
            nch = k if not ragged else int(rng.integers(max(2, k - 3), k + 1))
            if pbits is None:
                bits = rng.integers(0, 2, (nch, 256), dtype=np.uint8)
            else:
                bits = pbits[None, :] ^ (rng.random((nch, 256)) < flip[min(level - 1, len(flip) - 1)]).astype(np.uint8)
            for c in range(nch):
                parent.append(pid)
                leaf.append(1 if level == L else 0)
                desc.append(np.packbits(bits[c], bitorder="little"))
                if level == L:
                    w = 0.0 if rng.random() < stop_rate else float(rng.uniform(0.05, 12.0))
                else:
                    w = 0.0
                weight.append(w)
                new_frontier.append((next_id, bits[c]))
                next_id += 1
        frontier = new_frontier
    return (np.asarray(parent, np.int32), np.asarray(leaf, np.uint8), np.stack(desc).astype(np.uint8),
            np.asarray(weight, np.float64))


def write_vocabulary(path: str, k: int, L: int, seed: int = 0, scoring: int = 0, weighting: int = 0,
                     ragged: bool = False, stop_rate: float = 0.03) -> int:
    """Writes the text file; returns the number of nodes (without the root)."""
    parent, leaf, desc, weight = make_tree(k, L, seed, ragged, stop_rate)
    lines = [f"{k} {L} {scoring} {weighting}"]
    for i in range(len(parent)):
        lines.append(f"{parent[i]} {leaf[i]} " + " ".join(map(str, desc[i].tolist())) + f" {float(weight[i])!r}")
    with open(path, "w") as f:
        f.write("\n".join(lines))  # no final newline, see the module docstring
    return len(parent)


def words_for(desc_tree: np.ndarray, leaf: np.ndarray, n: int, seed: int, noise_bits: int = 24) -> np.ndarray:
    """n query descriptors: random leaves of the tree with up to noise_bits bits flipped (so that queries spread
    over the words the way real descriptors spread over a trained vocabulary)."""
    rng = np.random.default_rng(seed)
    leaves = np.nonzero(leaf)[0]
    pick = leaves[rng.integers(0, len(leaves), n)]
    d = desc_tree[pick].copy()
    for i in range(n):
        for b in rng.integers(0, 256, rng.integers(0, noise_bits + 1)):
            d[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return d


if __name__ == "__main__":
    out, k, L = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    seed = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    print(write_vocabulary(out, k, L, seed), "nodes ->", out)
