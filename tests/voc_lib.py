"""Synthetic DBoW2 vocabularies (text format of TemplatedVocabulary::loadFromTextFile) and access to
tests/cpp/_build/libdropin_voc.so (product ORBVocabulary.h + the vendored DBoW2 it derives from)."""
import ctypes as C
import os

import numpy as np

from orb_slam2_e_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
VOC_PATH = os.path.join(HERE, "cpp", "_build", "libdropin_voc.so")


def make_vocabulary(seed, k=10, L=3, stop_frac=0.05, ragged=True):
    """Tree in file order (a node's children follow it depth-first, as DBoW2 saves them): returns
    (parent, is_leaf, desc, weight) with node 0 = root.  Children are noisy copies of their parent
    so descents are decided by small margins (and exact ties)."""
    rng = np.random.default_rng(seed)
    parent, leaf, desc, weight = [0], [0], [np.zeros(32, np.uint8)], [0.0]

    def grow(pid, level, centre):
        n_child = k if not ragged or level == 1 else int(rng.integers(max(2, k - 3), k + 1))
        for c in range(n_child):
            d = synth.flip_bits(centre, int(rng.integers(20, 90) if level == 1 else rng.integers(0, 40)), rng)
            if c and rng.random() < 0.1:
                d = desc[-1].copy() if parent[-1] == pid else d  # duplicate sibling: exact tie, first child must win
            nid = len(parent)
            parent.append(pid)
            is_leaf = level == L or (ragged and level >= 2 and rng.random() < 0.05)
            leaf.append(int(is_leaf))
            desc.append(d)
            weight.append(0.0 if not is_leaf else (0.0 if rng.random() < stop_frac else float(np.round(rng.uniform(0.1, 9.0), 6))))
            if not is_leaf:
                grow(nid, level + 1, d)

    grow(0, 1, synth.random_descriptors(1, seed)[0])
    return (np.array(parent, np.int32), np.array(leaf, np.uint8), np.stack(desc), np.array(weight, np.float64))


def write_text(path, k, L, voc, scoring=0, weighting=0):
    parent, leaf, desc, weight = voc
    lines = [f"{k} {L} {scoring} {weighting}"]
    for i in range(1, len(parent)):
        lines.append(f"{parent[i]} {leaf[i]} " + " ".join(str(int(b)) for b in desc[i]) + f" {weight[i]:.6f}")
    with open(path, "w") as f:
        f.write("\n".join(lines))  # no trailing newline: the reference's loader turns an empty last line into a node


def descend_model(voc, feats, L, levelsup):
    """numpy model of TemplatedVocabulary.h:1217-1259."""
    parent, leaf, desc, weight = voc
    children = [[] for _ in parent]
    for i in range(1, len(parent)):
        children[parent[i]].append(i)
    word_of = np.cumsum(leaf) - 1
    pop = np.array([bin(i).count("1") for i in range(256)], np.int32)
    words, nodes, ws = [], [], []
    nid_level = L - levelsup
    for f in feats:
        node, level, at = 0, 0, 0
        while children[node]:
            level += 1
            ch = children[node]
            d = pop[desc[ch] ^ f].sum(1)
            node = ch[int(np.argmin(d))]  # first minimum
            if level == nid_level:
                at = node
        words.append(word_of[node] if leaf[node] else 0)
        nodes.append(at)
        ws.append(weight[node])
    return np.array(words, np.int32), np.array(nodes, np.int32), np.array(ws, np.float64)


def features(voc, n, seed):
    parent, leaf, desc, weight = voc
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(leaf)
    f = synth.random_descriptors(n, seed + 1)
    for i in range(n):
        if rng.random() < 0.8:
            f[i] = synth.flip_bits(desc[rng.choice(leaves)], int(rng.integers(0, 60)), rng)
    return f


_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(VOC_PATH):
            return None
        L = C.CDLL(VOC_PATH)
        L.voc_error.restype = C.c_char_p
        L.voc_load.argtypes = [C.c_char_p]
        L.voc_load.restype = C.c_void_p
        L.voc_free.argtypes = [C.c_void_p]
        L.voc_words.argtypes = [C.c_void_p]
        L.voc_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 2 + [C.POINTER(C.c_int)] + \
                                   [C.c_void_p] * 3 + [C.POINTER(C.c_int)]
        _lib = L
    return _lib


def transform(L, v, feats, levelsup, which):
    n = len(feats)
    feats = np.ascontiguousarray(feats, np.uint8)
    bid, bval = np.zeros(n + 1, np.uint32), np.zeros(n + 1, np.float64)
    fnode, foff, fidx = np.zeros(n + 1, np.uint32), np.zeros(n + 2, np.int32), np.zeros(n + 1, np.uint32)
    nb, nf = C.c_int(0), C.c_int(0)
    rc = L.voc_transform(v, feats.ctypes.data, n, levelsup, which, bid.ctypes.data, bval.ctypes.data, C.byref(nb), fnode.ctypes.data,
                         foff.ctypes.data, fidx.ctypes.data, C.byref(nf))
    assert rc == 0, L.voc_error()
    return (bid[:nb.value].copy(), bval[:nb.value].copy(), fnode[:nf.value].copy(), foff[:nf.value + 1].copy(),
            fidx[:foff[nf.value]].copy())
