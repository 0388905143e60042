"""GPU parity of the bag-of-words kernels (DBoW2 half: vslam_bow_transform_batch, vslam_bow_score,
vslam_bow_db_query) through the C ABI, against the committed fixture the REFERENCE's DBoW2 produced
(tests/golden/bow.npz) and against the CPU oracle on seeded inputs.  Bit-exact, doubles included: the kernels
accumulate in the reference's order."""
import os

import numpy as np
import pytest

from oracle import oracle as oo
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import BowDatabase, BowVocabulary, VslamError

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCORINGS = (0, 1, 2, 4, 5)
WEIGHTINGS = (0, 1, 2, 3)


def retyped_vocab(tmp_path, scoring, weighting):
    lines = open(os.path.join(GOLD, "bow_vocab.txt")).read().split("\n")
    k, L = lines[0].split()[:2]
    p = os.path.join(str(tmp_path), "voc_%d_%d.txt" % (scoring, weighting))
    open(p, "w").write("\n".join(["%s %s %d %d" % (k, L, scoring, weighting)] + lines[1:]))
    return p


def same_bits(a, b):
    return np.array_equal(np.asarray(a, np.float64).view(np.uint64), np.asarray(b, np.float64).view(np.uint64))


@pytest.mark.parametrize("scoring", SCORINGS)
def test_golden_reference_dbow2(ctx, tmp_path, scoring):
    g = np.load(os.path.join(GOLD, "bow.npz"))
    sets = [g["set%d" % i] for i in range(4)]
    for weighting in WEIGHTINGS:
        v = BowVocabulary.load_text(ctx, retyped_vocab(tmp_path, scoring, weighting))
        assert (v.k, v.L, v.scoring, v.weighting, v.n_words) == (5, 3, scoring, weighting, int(g["n_words"]))
        for lu in (2, 4):
            vecs, fw, fn, fwt, begin = v.transform_batch(sets, lu)
            for i in range(4):
                key = "s%d_w%d_set%d_l%d_" % (scoring, weighting, i, lu)
                assert np.array_equal(vecs[i][0], g[key + "ids"]), key
                assert same_bits(vecs[i][1], g[key + "vals"]), key
                sl = slice(begin[i], begin[i + 1])
                assert np.array_equal(np.where(fwt[sl] > 0, fn[sl], 0xFFFFFFFF), g[key + "fv"]), key
                assert np.array_equal(fw[sl], g["words_set%d" % i])
        want = g["s%d_w%d_scores" % (scoring, weighting)]
        got = np.array([[v.score(a, b) for b in vecs] for a in vecs])
        assert same_bits(got, want), (scoring, weighting)
        db = BowDatabase(ctx, v)
        for a in vecs:
            db.add(a)
        for i, a in enumerate(vecs):
            scores, common = db.query(a)
            assert same_bits(scores, want[i])
            assert np.array_equal(common, [len(np.intersect1d(a[0], b[0])) for b in vecs])
        db.close()
        v.close()


@pytest.mark.parametrize("scoring,weighting", [(0, 0), (1, 0), (5, 1), (2, 2), (4, 3)])
def test_against_oracle_large_unbalanced(ctx, scoring, weighting):
    """k = 10, L = 4, early leaves, stopped words, identical siblings; 1500-feature images in one batch, ragged
    and empty sets; every levelsup."""
    voc = wl.synth_vocabulary(10, 4, 21, prune=0.1, stop=0.08)
    voc["desc"][5] = voc["desc"][4]
    o = oo.BowVocab.from_arrays(10, 4, scoring, weighting, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    v = BowVocabulary.from_arrays(ctx, 10, 4, scoring, weighting, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    assert (v.n_nodes, v.n_words) == (o.n_nodes, o.n_words)
    sizes = [1500, 1500, 0, 1, 31, 32, 33, 1024, 1025, 700]
    sets = [wl.vocabulary_features(voc, n, 100 + i, 8 + 6 * (i % 5)) for i, n in enumerate(sizes)]
    sets[1] = np.concatenate([sets[0][:700], sets[1][:800]])      # shares words with image 0
    for lu in range(0, 6):
        vecs, fw, fn, fwt, begin = v.transform_batch(sets, lu)
        for i, s in enumerate(sets):
            ids, vals, ofw, ofn, ofwt = o.transform(s, lu)
            sl = slice(begin[i], begin[i + 1])
            assert np.array_equal(vecs[i][0], ids) and same_bits(vecs[i][1], vals), (lu, i)
            assert np.array_equal(fw[sl], ofw) and np.array_equal(fn[sl], ofn) and same_bits(fwt[sl], ofwt), (lu, i)
    db = BowDatabase(ctx, v)
    for a in vecs:
        db.add(a)
    assert len(db) == len(vecs)
    for a in vecs:
        scores, common = db.query(a)
        assert same_bits(scores, [o.score(a, b) for b in vecs])
        assert np.array_equal(common, [len(np.intersect1d(a[0], b[0])) for b in vecs])
    assert same_bits(v.score(vecs[0], vecs[1]), o.score(vecs[0], vecs[1]))
    assert v.score(vecs[0], vecs[1]) != v.score(vecs[0], vecs[0])
    one, fv = v.transform(sets[0], 2)
    assert np.array_equal(one[0], vecs[0][0]) and same_bits(one[1], vecs[0][1])   # levelsup does not change the BowVector
    assert sorted(f for fs in fv.values() for f in fs) == list(np.nonzero(o.transform(sets[0], 2)[4] > 0)[0])
    db.close()
    v.close()
    o.close()


def test_first_minimum_wins_and_wide_nodes(ctx):
    """A root with 70 children (more than one warp pass), all at the same distance from the feature but one: the
    strict '<' scan keeps the first minimum."""
    rng = np.random.default_rng(3)
    n = 70
    base = wl.random_descriptors(1, 9)[0]
    desc = np.repeat(base[None], n, 0).copy()
    for c in range(n):                      # child c differs from the base in bit c (distance 1) ...
        desc[c, c // 32] ^= np.uint32(1) << np.uint32(c % 32)
    parent = np.zeros(n, np.int32)
    leaf = np.ones(n, np.uint8)
    weight = rng.random(n) + 0.5
    v = BowVocabulary.from_arrays(ctx, 20, 1, 0, 0, parent, leaf, desc, weight)
    o = oo.BowVocab.from_arrays(20, 1, 0, 0, parent, leaf, desc, weight)
    feats = np.stack([base, desc[41], desc[69], desc[0] ^ desc[1] ^ base])   # tie of all; exact hits; tie of children 0, 1
    vecs, fw, fn, fwt, _ = v.transform_batch([feats], 0)
    assert list(fw) == [0, 41, 69, 0] == list(o.transform(feats, 0)[2])
    assert np.array_equal(fn, o.transform(feats, 0)[3])
    v.close()
    o.close()


def test_edge_cases_and_errors(ctx, tmp_path):
    voc = wl.synth_vocabulary(4, 2, 5)
    a, b = os.path.join(str(tmp_path), "a.txt"), os.path.join(str(tmp_path), "b.txt")
    wl.write_vocabulary_text(a, voc)
    wl.write_vocabulary_text(b, voc, trailing_newline=True)
    va, vb = BowVocabulary.load_text(ctx, a), BowVocabulary.load_text(ctx, b)
    assert (va.n_nodes, va.n_words) == (vb.n_nodes, vb.n_words) == (21, 16)
    f = wl.vocabulary_features(voc, 200, 1)
    (ia, xa), _ = va.transform(f)
    (ib, xb), _ = vb.transform(f)
    assert np.array_equal(ia, ib) and same_bits(xa, xb)
    assert abs(xa.sum() - 1.0) < 1e-12                     # L1-normalised
    # no images, an image without features, an empty database
    assert va.transform_batch([])[0] == []
    (ie, xe), fv = va.transform(np.zeros((0, 8), np.uint32))
    assert len(ie) == 0 and fv == {}
    db = BowDatabase(ctx, va)
    assert len(db.query((ia, xa))[0]) == 0
    db.add((ie, xe))
    db.add((ia, xa))
    scores, common = db.query((ia, xa))
    assert list(common) == [0, len(ia)] and scores[1] == va.score((ia, xa), (ia, xa))
    assert scores[0] == 0.0 and np.signbit(scores[0])      # -0.0: "score = -score/2.0" of an empty sum
    # errors: capacity, unsorted vector, too many features, KL scoring, malformed file
    with pytest.raises(VslamError):
        va.transform_batch([f], cap=2)
    with pytest.raises(VslamError):
        va.score((ia[::-1].copy(), xa), (ia, xa))
    with pytest.raises(VslamError):
        va.transform_batch([np.zeros((8193, 8), np.uint32)])
    with pytest.raises(VslamError):
        BowVocabulary.from_arrays(ctx, 4, 2, 3, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    bad = os.path.join(str(tmp_path), "bad.txt")
    open(bad, "w").write("25 3 0 0\n")
    with pytest.raises(VslamError):
        BowVocabulary.load_text(ctx, bad)
    with pytest.raises(VslamError):
        BowVocabulary.load_text(ctx, os.path.join(str(tmp_path), "missing.txt"))
    # vectors longer than the shared-memory query window (hand-made: a transform yields at most 8192 words)
    rng = np.random.default_rng(8)
    big = []
    for _ in range(2):
        ids = np.sort(rng.choice(40000, 9000, replace=False)).astype(np.uint32)
        x = rng.random(9000)
        big.append((ids, x / x.sum()))
    o = oo.BowVocab.from_arrays(4, 2, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    assert same_bits(va.score(big[0], big[1]), o.score(big[0], big[1]))
    assert same_bits(va.score(big[0], (ia, xa)), o.score(big[0], (ia, xa)))
    o.close()
    # a vocabulary without words: transform clears both vectors (TemplatedVocabulary.h:1135-1138)
    v0 = BowVocabulary.from_arrays(ctx, 10, 5, 0, 0, np.zeros(0, np.int32), np.zeros(0, np.uint8),
                                   np.zeros((0, 8), np.uint32), np.zeros(0))
    assert len(v0.transform(f)[0][0]) == 0
    for x in (db, va, vb, v0):
        x.close()


def test_orbvoc_shape_full_size(ctx):
    """A tree of ORBvoc's shape (k = 10, L = 6: 1 111 110 nodes, 10^6 words) with 2000 images' worth of features:
    size-independent properties -- every feature built from a word descends to that word when unperturbed, the
    BowVector is L1-normalised, scores lie in [0, 1], an image scores 1 against itself."""
    voc = wl.synth_vocabulary(10, 6, 31, flips=(0, 30, 24, 18, 14, 10))   # decisive at every level
    v = BowVocabulary.from_arrays(ctx, 10, 6, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    assert (v.n_nodes, v.n_words) == (1111111, 1000000)
    rng = np.random.default_rng(5)
    words = np.nonzero(voc["is_leaf"])[0]
    pick = words[rng.integers(0, len(words), 1500)]
    exact = np.ascontiguousarray(voc["desc"][pick])
    vecs, fw, fn, fwt, _ = v.transform_batch([exact, wl.flip_bits(exact, 6, rng)], 4)
    word_of_node = np.cumsum(voc["is_leaf"]) - 1
    want = word_of_node[pick]
    assert np.array_equal(fw[:1500], want)
    assert (fw[1500:] == want).mean() > 0.95
    par = voc["parent"]
    lvl2 = par[par[par[par[pick] - 1] - 1] - 1]              # the ancestor at level 2 = L - 4 (node ids are 1-based)
    assert np.array_equal(fn[:1500], lvl2)
    for ids, vals in vecs:
        assert abs(vals.sum() - 1.0) < 1e-12 and np.all(np.diff(ids.astype(np.int64)) > 0)
    assert abs(v.score(vecs[0], vecs[0]) - 1.0) < 1e-12
    assert 0.5 < v.score(vecs[0], vecs[1]) < 1.0
    v.close()
