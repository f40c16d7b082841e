"""Vocabulary-tree transform on the GPU (orbm_vocab_*, include/ORBVocabulary.h) against
(a) a numpy model of the descent rule and (b) the reference's own vendored DBoW2 transform, run on
the very same object, on synthetic vocabularies with near-ties, exact ties, ragged depth and stopped
words.  Word / node ids and FeatureVector must be equal; BowVector values must be the same doubles."""
import ctypes as C

import numpy as np
import pytest

import voc_lib
from orb_slam2_e_b200 import _lib
from orb_slam2_e_b200._lib import check, ptr

pytestmark = pytest.mark.gpu


def _create(voc, k, L):
    parent, leaf, desc, weight = voc
    h = C.c_void_p()
    check(_lib.load().orbm_vocab_create(0, k, L, len(parent), ptr(parent), ptr(leaf), ptr(np.ascontiguousarray(desc)), ptr(weight),
                                        C.byref(h)))
    return h


@pytest.mark.parametrize("k,L,levelsup,ragged", [(10, 3, 1, False), (10, 4, 2, True), (7, 5, 4, True), (10, 3, 5, False), (3, 6, 4, True)])
def test_descent_matches_model(k, L, levelsup, ragged):
    voc = voc_lib.make_vocabulary(5 + k + L, k, L, ragged=ragged)
    feats = voc_lib.features(voc, 3000, 9)
    h = _create(voc, k, L)
    word, node, w = np.zeros(len(feats), np.int32), np.zeros(len(feats), np.int32), np.zeros(len(feats), np.float64)
    check(_lib.load().orbm_vocab_transform(h, ptr(feats), len(feats), levelsup, ptr(word), ptr(node), ptr(w)))
    mw, mn, mws = voc_lib.descend_model(voc, feats, L, levelsup)
    assert np.array_equal(word, mw) and np.array_equal(node, mn) and np.array_equal(w, mws)
    nw = C.c_int(0)
    check(_lib.load().orbm_vocab_info(h, None, None, None, C.byref(nw)))
    assert nw.value == int(voc[1].sum())
    check(_lib.load().orbm_vocab_destroy(h))


def test_text_loader_equals_flat_arrays(tmp_path):
    voc = voc_lib.make_vocabulary(3, 10, 3)
    path = str(tmp_path / "voc.txt")
    voc_lib.write_text(path, 10, 3, voc)
    feats = voc_lib.features(voc, 500, 4)
    outs = []
    for make in (lambda: _create(voc, 10, 3), None):
        h = make() if make else C.c_void_p()
        if not make:
            check(_lib.load().orbm_vocab_load_text(0, path.encode(), C.byref(h)))
        word, node, w = np.zeros(len(feats), np.int32), np.zeros(len(feats), np.int32), np.zeros(len(feats), np.float64)
        check(_lib.load().orbm_vocab_transform(h, ptr(feats), len(feats), 4, ptr(word), ptr(node), ptr(w)))
        outs.append((word, node, w))
        check(_lib.load().orbm_vocab_destroy(h))
    for a, b in zip(*outs):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("k,L,levelsup,weighting", [(10, 3, 1, 0), (10, 4, 4, 0), (6, 4, 2, 1), (10, 3, 2, 3)])
def test_orbvocabulary_transform_equals_reference_dbow2(tmp_path, k, L, levelsup, weighting):
    """Frame::ComputeBoW's call (Frame.cc:415): the product's ORBVocabulary::transform vs the base class
    DBoW2 implementation compiled unmodified from the reference tree."""
    lib = voc_lib.load()
    if lib is None:
        pytest.skip("tests/cpp/_build/libdropin_voc.so not built (needs /root/reference at build time)")
    voc = voc_lib.make_vocabulary(11 + k, k, L)
    path = str(tmp_path / "voc.txt")
    voc_lib.write_text(path, k, L, voc, scoring=0, weighting=weighting)
    v = lib.voc_load(path.encode())
    assert v, lib.voc_error()
    assert lib.voc_words(v) == int(voc[1].sum())
    for n in (0, 1, 1200):
        feats = voc_lib.features(voc, n, 21 + n) if n else np.zeros((0, 32), np.uint8)
        got = voc_lib.transform(lib, v, feats, levelsup, 0)
        ref = voc_lib.transform(lib, v, feats, levelsup, 1)
        for a, b in zip(got, ref):
            assert a.tobytes() == b.tobytes()
        if n > 100:
            assert len(ref[0]) > 50 and len(ref[2]) >= (1 if levelsup >= L else 5)
    lib.voc_free(v)
