"""CPU tests of the bag-of-words oracle (DBoW2 half): the C restatement (oracle/visnav_oracle_bow.c)
against the committed fixture tests/golden/bow.npz -- produced by the REFERENCE's own DBoW2
(tools/make_golden.py bow) -- and, where oracle/_ref was built, against that library live.  Bit-exact: the word
ids are integer work and the doubles are accumulated in the reference's order."""
import os

import numpy as np
import pytest

from oracle import oracle as oo
from visnav_b200 import workloads as wl

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCORINGS = (0, 1, 2, 4, 5)
WEIGHTINGS = (0, 1, 2, 3)


def retyped_vocab(tmp_path, scoring, weighting):
    """The golden vocabulary text with another scoring / weighting type in its header line."""
    lines = open(os.path.join(GOLD, "bow_vocab.txt")).read().split("\n")
    k, L = lines[0].split()[:2]
    p = os.path.join(str(tmp_path), "voc_%d_%d.txt" % (scoring, weighting))
    open(p, "w").write("\n".join(["%s %s %d %d" % (k, L, scoring, weighting)] + lines[1:]))
    return p


def same_bits(a, b):
    return np.array_equal(np.asarray(a, np.float64).view(np.uint64), np.asarray(b, np.float64).view(np.uint64))


@pytest.mark.parametrize("scoring", SCORINGS)
def test_port_matches_golden(tmp_path, scoring):
    g = np.load(os.path.join(GOLD, "bow.npz"))
    for weighting in WEIGHTINGS:
        v = oo.BowVocab.from_text(retyped_vocab(tmp_path, scoring, weighting))
        assert (v.k, v.L, v.n_words) == (5, 3, int(g["n_words"]))
        vecs = []
        for i in range(4):
            for lu in (2, 4):
                ids, vals, fw, fn, fwt = v.transform(g["set%d" % i], lu)
                key = "s%d_w%d_set%d_l%d_" % (scoring, weighting, i, lu)
                assert np.array_equal(ids, g[key + "ids"]), key
                assert same_bits(vals, g[key + "vals"]), key
                assert np.array_equal(np.where(fwt > 0, fn, 0xFFFFFFFF), g[key + "fv"]), key
                assert np.array_equal(fw, g["words_set%d" % i])
            vecs.append((ids, vals))
        got = np.array([[v.score(a, b) for b in vecs] for a in vecs])
        assert same_bits(got, g["s%d_w%d_scores" % (scoring, weighting)]), (scoring, weighting)
        v.close()


def test_port_matches_reference_library(tmp_path):
    """Unbalanced tree (early leaves), stopped words, ties: restatement == the reference's DBoW2, every type."""
    if oo.BowVocabRef.lib() is None:
        pytest.skip("oracle/_ref not built")
    voc = wl.synth_vocabulary(7, 4, 11, prune=0.2, stop=0.1)
    voc["desc"][3] = voc["desc"][2]          # two identical children of the root: the first wins every tie
    path = os.path.join(str(tmp_path), "v.txt")
    for scoring in SCORINGS:
        for weighting in WEIGHTINGS:
            wl.write_vocabulary_text(path, voc, scoring, weighting)
            p, r = oo.BowVocab.from_text(path), oo.BowVocabRef.from_text(path)
            assert (p.k, p.L, p.scoring, p.weighting, p.n_words) == (r.k, r.L, r.scoring, r.weighting, r.n_words)
            vecs = []
            for s in range(3):
                f = wl.vocabulary_features(voc, 700 - 250 * s, s, 10 + 40 * s)
                # nid level 1..3 (levelsup 3..1): above the shallowest leaves the reference's node id is undefined
                ids, vals, fw, fn, fwt = p.transform(f, 3)
                rids, rvals, rfv = r.transform(f, 3)
                assert np.array_equal(ids, rids) and same_bits(vals, rvals)
                assert np.array_equal(fw, r.words(f))
                assert np.array_equal(np.where(fwt > 0, fn, 0xFFFFFFFF), rfv)
                vecs.append((ids, vals))
            for a in vecs:
                for b in vecs:
                    assert same_bits(p.score(a, b), r.score(a, b)), (scoring, weighting)
            p.close()
            r.close()


def test_text_loader_edge_cases(tmp_path):
    voc = wl.synth_vocabulary(4, 2, 5)
    a, b = os.path.join(str(tmp_path), "a.txt"), os.path.join(str(tmp_path), "b.txt")
    wl.write_vocabulary_text(a, voc)
    wl.write_vocabulary_text(b, voc, trailing_newline=True)   # what saveToTextFile writes; the blank line is skipped
    va, vb = oo.BowVocab.from_text(a), oo.BowVocab.from_text(b)
    assert (va.n_nodes, va.n_words) == (vb.n_nodes, vb.n_words) == (21, 16)
    f = wl.vocabulary_features(voc, 100, 1)
    assert all(np.array_equal(x, y) for x, y in zip(va.transform(f), vb.transform(f)))
    k, L, sc, wt, arrays = wl.read_vocabulary_text(a)
    vc = oo.BowVocab.from_arrays(k, L, sc, wt, arrays["parent"], arrays["is_leaf"], arrays["desc"], arrays["weight"])
    assert all(np.array_equal(x, y) for x, y in zip(va.transform(f), vc.transform(f)))
    bad = os.path.join(str(tmp_path), "bad.txt")
    open(bad, "w").write("25 3 0 0\n")                          # k > 20: "not a correct text file"
    with pytest.raises(ValueError):
        oo.BowVocab.from_text(bad)
    # no features / nothing in common
    ids, vals, *_ = va.transform(np.zeros((0, 8), np.uint32))
    assert len(ids) == 0
    e = (np.zeros(0, np.uint32), np.zeros(0))
    assert va.score(e, e) == 0.0
