"""Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:395-402, src/KeyFrame.cc:59-68) =
ORBVocabulary::transform(descriptors, BowVector, FeatureVector, 4) (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1194),
and Vocabulary::score (L1Scoring::score, Thirdparty/DBoW2/DBoW2/ScoringObject.cpp:25-72) for batches of keyframes.
CPU: the restatement (oracle/vocab_port.cc) against the reference's own DBoW2 + Frame code (oracle/_ref/libframeref.so), the
vocabulary going through the reference's loadFromTextFile; GPU: orbgpu_vocabulary_* bit-for-bit against the restatement
(word ids, the doubles of the BowVector by their bit patterns, node ids, feature lists, scores)."""
import numpy as np
import pytest

import oracle_lib

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")


def make_tree(seed, kmax, depth, min_leaf_depth=None, pool=None, zero_weights=0.05, fixed_k=False):
    """A vocabulary tree in DBoW2's node order (HKmeansStep: the children of a node get consecutive ids, then each child is
    expanded in turn).  Node descriptors come from a small pool when `pool` is given (equal distances: the first child
    must win).  Leaves may appear from min_leaf_depth on; every node at `depth` is a leaf."""
    rng = np.random.default_rng(seed)
    min_leaf_depth = depth if min_leaf_depth is None else min_leaf_depth
    parent, leaf, level = [0], [0], [0]

    def expand(node, lv):
        nc = kmax if fixed_k else int(rng.integers(1, kmax + 1))
        kids = []
        for _ in range(nc):
            parent.append(node); level.append(lv + 1); leaf.append(0); kids.append(len(parent) - 1)
        for c in kids:
            if lv + 1 >= depth or (lv + 1 >= min_leaf_depth and rng.random() < 0.3):
                leaf[c] = 1
            else:
                expand(c, lv + 1)
    expand(0, 0)
    n = len(parent)
    if pool is None:
        desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    else:
        base = rng.integers(0, 256, (pool, 32), dtype=np.uint8)
        desc = base[rng.integers(0, pool, n)].copy()
    weight = np.where(np.array(leaf) > 0, rng.uniform(0.1, 9.0, n), 0.0)
    stop = (rng.random(n) < zero_weights) & (np.array(leaf) > 0)
    weight[stop] = 0.0                                               # stopped words (TemplatedVocabulary.h:1157)
    return dict(k=min(kmax, 20), L=depth, parent=np.array(parent, np.int32), is_leaf=np.array(leaf, np.uint8), desc=desc,
                weight=weight, level=np.array(level))


def features_for(tree, seed, n, noise_bits=20):
    """Descriptors near leaf descriptors (so that the walk is not a coin toss), some pure noise, some exact copies."""
    rng = np.random.default_rng(seed)
    leaves = np.nonzero(tree["is_leaf"])[0]
    d = tree["desc"][leaves[rng.integers(0, len(leaves), n)]].copy()
    for _ in range(noise_bits):
        hit = rng.random(n) < 0.7
        d[hit, rng.integers(0, 32, hit.sum())] ^= (np.uint8(1) << rng.integers(0, 8, hit.sum()).astype(np.uint8))
    noise = rng.random(n) < 0.1
    d[noise] = rng.integers(0, 256, (int(noise.sum()), 32), dtype=np.uint8)
    return d


def vocab_args(t, scoring=0, weighting=0):
    return (t["k"], t["L"], t["parent"], t["is_leaf"], t["desc"], t["weight"], scoring, weighting)


def same_bow(a, b):
    names = ("words", "values", "node_ids", "start", "idx")
    for x, y, nm in zip(a, b, names):
        assert len(x) == len(y), (nm, len(x), len(y))
        if x.dtype == np.float64:
            bad = np.nonzero(x.view(np.uint64) != y.view(np.uint64))[0]
            assert len(bad) == 0, (nm, bad[:5], x[bad[:5]], y[bad[:5]])
        else:
            assert np.array_equal(x, y), (nm, np.nonzero(x != y)[0][:5])
    return True


def random_bow_vectors(rng, n_words, nk, lo=50, hi=900, share=None):
    """L1-normalised BowVectors over a word universe (as transform leaves them)"""
    out = []
    for _ in range(nk):
        m = int(rng.integers(lo, hi))
        w = np.sort(rng.choice(n_words, size=min(m, n_words), replace=False)).astype(np.uint32)
        if share is not None and rng.random() < 0.7:
            w = np.unique(np.concatenate([w, rng.choice(share, size=len(share) // 2, replace=False)])).astype(np.uint32)
        v = rng.uniform(0.01, 5.0, len(w))
        out.append((w, v / v.sum()))
    return out


def csr(vectors):
    ks = np.zeros(len(vectors) + 1, np.int32)
    ks[1:] = np.cumsum([len(v[0]) for v in vectors])
    kw = np.concatenate([v[0] for v in vectors]) if vectors else np.zeros(0, np.uint32)
    kv = np.concatenate([v[1] for v in vectors]) if vectors else np.zeros(0, np.float64)
    return kw.astype(np.uint32), kv.astype(np.float64), ks


TREES = {
    "k10-depth4": dict(seed=1, kmax=10, depth=4, fixed_k=True),
    "ragged": dict(seed=2, kmax=9, depth=5, min_leaf_depth=3),
    "ties": dict(seed=3, kmax=8, depth=4, pool=6, fixed_k=True),
    "wide": dict(seed=4, kmax=20, depth=2, fixed_k=True),
}


# ------------------------------------------------------------------------------------------------------------- CPU side
def test_port_bow_vector_properties():
    t = make_tree(1, 10, 4, fixed_k=True)
    v = oracle_lib.PortVocabulary(*vocab_args(t))
    d = features_for(t, 5, 1200)
    words, values, nodes, start, idx = v.transform(d, 2)
    assert len(words) > 300 and np.all(np.diff(words.astype(np.int64)) > 0) and np.all(np.diff(nodes.astype(np.int64)) > 0)
    assert abs(values.sum() - 1.0) < 1e-12 and (values > 0).all()
    assert start[0] == 0 and start[-1] == len(idx) and len(np.unique(idx)) == len(idx)
    assert (t["level"][nodes] == t["L"] - 2).all()                   # FeatureVector nodes: `levelsup` above the leaves
    for j in range(len(nodes)):
        assert np.all(np.diff(idx[start[j]:start[j + 1]].astype(np.int64)) > 0)
    # every word of the BowVector hangs below one of the FeatureVector's nodes, and every such node holds a word
    leaves = np.nonzero(t["is_leaf"])[0]
    anc = leaves[words]
    while (t["level"][anc] > t["L"] - 2).any():
        up = t["level"][anc] > t["L"] - 2
        anc[up] = t["parent"][anc[up]]
    assert set(anc.tolist()) == set(nodes.tolist())


@needs_ref
@pytest.mark.parametrize("name", list(TREES))
@pytest.mark.parametrize("mode", [(0, 0), (0, 1), (0, 2), (0, 3), (1, 0), (5, 0), (5, 1), (2, 0)],
                         ids=["L1-tfidf", "L1-tf", "L1-idf", "L1-binary", "L2-tfidf", "dot-tfidf", "dot-tf", "chi2-tfidf"])
def test_vocab_restatement_equals_reference(pkg, tmp_path, name, mode):
    scoring, weighting = mode
    t = make_tree(**TREES[name])
    path = tmp_path / "voc.txt"
    pkg.write_vocabulary_text(path, t["k"], t["L"], t["parent"], t["is_leaf"], t["desc"], t["weight"], scoring, weighting)
    ref = oracle_lib.RefVocabulary(path)
    port = oracle_lib.PortVocabulary(*vocab_args(t, scoring, weighting))
    assert ref.info() == (t["k"], t["L"], scoring, weighting, len(t["parent"]), int(t["is_leaf"].sum()))
    min_leaf = int(t["level"][t["is_leaf"] > 0].min())
    for seed, n in [(11, 1000), (12, 37), (13, 1)]:
        d = features_for(t, seed, n)
        for levelsup in (0, 1, t["L"] - min_leaf, t["L"], t["L"] + 2):
            if t["L"] - levelsup > min_leaf:
                continue                                             # the reference reads an uninitialised nid there
            assert same_bow(port.transform(d, levelsup), ref.transform(d, levelsup))
    if t["L"] - 4 <= min_leaf:
        d = features_for(t, 14, 800)
        assert same_bow(port.transform(d, 4), ref.compute_bow(d))    # Frame::ComputeBoW: levelsup = 4


@needs_ref
def test_score_restatement_equals_reference(pkg, tmp_path):
    t = make_tree(1, 10, 4, fixed_k=True)
    path = tmp_path / "voc.txt"
    pkg.write_vocabulary_text(path, t["k"], t["L"], t["parent"], t["is_leaf"], t["desc"], t["weight"])
    ref = oracle_lib.RefVocabulary(path)
    port = oracle_lib.PortVocabulary(*vocab_args(t))
    rng = np.random.default_rng(7)
    nw = int(t["is_leaf"].sum())
    q = random_bow_vectors(rng, nw, 1)[0]
    kfs = random_bow_vectors(rng, nw, 60, share=q[0]) + [q, (np.zeros(0, np.uint32), np.zeros(0))]
    kw, kv, ks = csr(kfs)
    want = ref.score_csr(q[0], q[1], kw, kv, ks)
    got = port.score_csr(q[0], q[1], kw, kv, ks)
    assert np.array_equal(want.view(np.uint64), got.view(np.uint64))
    assert abs(want[-2] - 1.0) < 1e-12 and want[-1] == 0.0 and (want[:-2] > 0).any()
    # BowVectors made by transform itself
    frames = [port.transform(features_for(t, 100 + i, 700), 2)[:2] for i in range(8)]
    kw, kv, ks = csr(frames)
    assert np.array_equal(ref.score_csr(frames[0][0], frames[0][1], kw, kv, ks).view(np.uint64),
                          port.score_csr(frames[0][0], frames[0][1], kw, kv, ks).view(np.uint64))


# ------------------------------------------------------------------------------------------------------------- GPU side
@pytest.mark.gpu
@pytest.mark.parametrize("name", list(TREES))
@pytest.mark.parametrize("mode", [(0, 0), (0, 1), (0, 2), (0, 3), (1, 0), (5, 0), (5, 1)],
                         ids=["L1-tfidf", "L1-tf", "L1-idf", "L1-binary", "L2-tfidf", "dot-tfidf", "dot-tf"])
def test_device_transform_equals_oracle(pkg, name, mode):
    scoring, weighting = mode
    t = make_tree(**TREES[name])
    port = oracle_lib.PortVocabulary(*vocab_args(t, scoring, weighting))
    dev = pkg.ORBVocabulary(*vocab_args(t, scoring, weighting))
    assert dev.info() == (t["k"], t["L"], scoring, weighting, len(t["parent"]), int(t["is_leaf"].sum()))
    for seed, n in [(11, 1000), (12, 37), (13, 1), (14, 2048), (15, 2049)]:
        d = features_for(t, seed, n)
        for levelsup in (0, 1, 2, t["L"], t["L"] + 2, 4):
            assert same_bow(dev.transform(d, levelsup), port.transform(d, levelsup))
    dev.close()


@pytest.mark.gpu
def test_device_transform_edge_cases(pkg, tmp_path):
    t = make_tree(5, 40, 3, min_leaf_depth=1)                         # up to 40 children per node, leaves at every depth
    port = oracle_lib.PortVocabulary(*vocab_args(t))
    dev = pkg.ORBVocabulary(*vocab_args(t))
    for n in (0, 1, 5000, 16384):
        d = features_for(t, n + 1, n) if n else np.zeros((0, 32), np.uint8)
        assert same_bow(dev.transform(d, 1), port.transform(d, 1))
    with pytest.raises(pkg.OrbGpuError):
        dev.transform(np.zeros((16385, 32), np.uint8))
    dev.close()
    # every word stopped -> empty vectors; a vocabulary without words -> empty vectors
    t0 = make_tree(6, 4, 2, zero_weights=2.0, fixed_k=True)
    dev = pkg.ORBVocabulary(*vocab_args(t0))
    out = dev.transform(features_for(t0, 1, 100), 1)
    assert all(len(x) == 0 for x in (out[0], out[1], out[2], out[4])) and out[3].tolist() == [0]
    dev.close()
    dev = pkg.ORBVocabulary(10, 6, [0], [0], np.zeros((1, 32), np.uint8), [0.0])
    assert all(len(x) == 0 for x in dev.transform(np.zeros((10, 32), np.uint8))[:3])
    dev.close()
    # the text loader builds the same vocabulary as the arrays (with and without the trailing newline saveToTextFile leaves)
    t = make_tree(2, 9, 5, min_leaf_depth=3)
    port = oracle_lib.PortVocabulary(*vocab_args(t, 1, 0))
    d = features_for(t, 3, 900)
    for trailing in (False, True):
        path = tmp_path / ("voc%d.txt" % trailing)
        pkg.write_vocabulary_text(path, t["k"], t["L"], t["parent"], t["is_leaf"], t["desc"], t["weight"], 1, 0, trailing_newline=trailing)
        dev = pkg.ORBVocabulary.load_from_text_file(path)
        assert dev.info() == (t["k"], t["L"], 1, 0, len(t["parent"]), int(t["is_leaf"].sum()))
        assert same_bow(dev.transform(d, 2), port.transform(d, 2))
        dev.close()
    bad = tmp_path / "bad.txt"
    bad.write_text("30 6 0 0\n0 1 " + " ".join(["0"] * 32) + " 1.0")
    with pytest.raises(pkg.OrbGpuError):
        pkg.ORBVocabulary.load_from_text_file(bad)                   # k > 20: "not a correct text file" (:1358-1362)


@pytest.mark.gpu
def test_device_transform_batch_equals_per_frame_oracle(pkg):
    t = make_tree(2, 9, 5, min_leaf_depth=3)
    port = oracle_lib.PortVocabulary(*vocab_args(t))
    dev = pkg.ORBVocabulary(*vocab_args(t))
    sizes = [1000, 0, 1, 2500, 64, 65, 0, 1013] + [1000] * 40
    descs = [features_for(t, 300 + i, n) if n else np.zeros((0, 32), np.uint8) for i, n in enumerate(sizes)]
    descs[3][:] = descs[3][0]                                        # 2500 copies of one descriptor: one word, added 2500 times
    got = dev.transform_batch(descs, 2)
    assert len(got) == len(sizes)
    for d, g in zip(descs, got):
        assert same_bow(g, port.transform(d, 2))
    assert dev.transform_batch([], 2) == []
    assert same_bow(dev.transform_batch([descs[0]], 2)[0], got[0])
    dev.close()


@pytest.mark.gpu
def test_device_score_equals_oracle(pkg):
    t = make_tree(1, 10, 4, fixed_k=True)
    port = oracle_lib.PortVocabulary(*vocab_args(t))
    dev = pkg.ORBVocabulary(*vocab_args(t))
    rng = np.random.default_rng(9)
    nw = int(t["is_leaf"].sum())
    for nk in (1, 33, 700):
        q = random_bow_vectors(rng, nw, 1)[0]
        kfs = random_bow_vectors(rng, nw, nk, share=q[0]) + [q, (np.zeros(0, np.uint32), np.zeros(0))]
        kw, kv, ks = csr(kfs)
        want = port.score_csr(q[0], q[1], kw, kv, ks)
        got = dev.score_csr(q[0], q[1], kw, kv, ks)
        assert np.array_equal(want.view(np.uint64), got.view(np.uint64))
        assert np.array_equal(dev.score(q, kfs).view(np.uint64), want.view(np.uint64))
    # an empty query, no keyframes
    kw, kv, ks = csr(random_bow_vectors(rng, nw, 5))
    assert np.array_equal(dev.score_csr(np.zeros(0, np.uint32), np.zeros(0), kw, kv, ks), np.zeros(5))
    assert len(dev.score_csr(q[0], q[1], np.zeros(0, np.uint32), np.zeros(0), np.zeros(1, np.int32))) == 0
    # BowVectors made by the device transform, scored against each other
    frames = [dev.transform(features_for(t, 200 + i, 900), 2)[:2] for i in range(12)]
    kw, kv, ks = csr(frames)
    want = port.score_csr(frames[3][0], frames[3][1], kw, kv, ks)
    got = dev.score_csr(frames[3][0], frames[3][1], kw, kv, ks)
    assert np.array_equal(want.view(np.uint64), got.view(np.uint64)) and abs(got[3] - 1.0) < 1e-12
    dev.close()
    other = pkg.ORBVocabulary(*vocab_args(t, 1, 0))
    with pytest.raises(pkg.OrbGpuError):
        other.score_csr(q[0], q[1], kw, kv, ks)                      # only L1_NORM vocabularies are scored on the device
    other.close()


# ------------------------------------------------------------------------------------------- KeyFrameDatabase (src/KeyFrameDatabase.cc)
def db_scene(seed, n_words, n_kf=90, reloc=False):
    """Keyframes with overlapping BowVectors (a few near-copies of the query), some erased from the database, some connected
    to the query keyframe, random ordered covisibility lists (more than 10 entries for some), scores left by earlier queries."""
    rng = np.random.default_rng(seed)
    q = random_bow_vectors(rng, n_words, 1, 300, 600)[0]
    kfs = random_bow_vectors(rng, n_words, n_kf, 100, 600, share=q[0])
    for i in rng.choice(n_kf, 6, replace=False):                      # near-copies: high scores, many common words
        keep = rng.random(len(q[0])) < 0.9
        v = q[1][keep] * rng.uniform(0.8, 1.2, int(keep.sum()))
        kfs[i] = (q[0][keep], v / v.sum())
    kfs[int(rng.integers(n_kf))] = (np.zeros(0, np.uint32), np.zeros(0))
    erased = rng.random(n_kf) < 0.1
    connected = (rng.random(n_kf) < 0.15) & (not reloc)
    neighbours = [rng.choice(n_kf, int(rng.integers(0, 16)), replace=False).tolist() for _ in range(n_kf)]
    prev = rng.uniform(0, 0.05, n_kf).astype(np.float32)
    min_score = float(rng.choice([0.0, 0.01, 0.03]))
    return kfs, erased, q, connected, neighbours, prev, min_score


@needs_ref
@pytest.mark.parametrize("reloc", [False, True], ids=["loop", "reloc"])
@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_keyframe_database_restatement_equals_reference(pkg, tmp_path, seed, reloc):
    t = make_tree(1, 10, 4, fixed_k=True)
    path = tmp_path / "voc.txt"
    pkg.write_vocabulary_text(path, t["k"], t["L"], t["parent"], t["is_leaf"], t["desc"], t["weight"])
    ref = oracle_lib.RefVocabulary(path)
    kfs, erased, q, connected, neighbours, prev, min_score = db_scene(seed, int(t["is_leaf"].sum()), reloc=reloc)
    args, keep = oracle_lib.db_args(kfs, erased, q, connected, neighbours, prev, min_score, reloc)
    wm, ww, ws, wc = oracle_lib.ref_db_detect(ref, args)
    gm, gw, gs, gc = oracle_lib.port_db_detect(args)
    assert wm.sum() > 20 and len(wc) >= 1
    assert np.array_equal(gm, wm) and np.array_equal(gw[wm], ww[wm])
    assert np.array_equal(gs.view(np.uint32), ws.view(np.uint32))
    assert np.array_equal(gc, wc)
    assert not wm[erased].any() and not wm[np.asarray(connected, bool)].any()


@pytest.mark.gpu
@pytest.mark.parametrize("reloc", [False, True], ids=["loop", "reloc"])
@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6, 7, 8])
def test_device_keyframe_database_equals_oracle(pkg, seed, reloc):
    kfs, erased, q, connected, neighbours, prev, min_score = db_scene(seed, 10000, n_kf=90 if seed < 7 else 1500, reloc=reloc)
    args, keep = oracle_lib.db_args(kfs, erased, q, connected, neighbours, prev, min_score, reloc)
    wm, ww, ws, wc = oracle_lib.port_db_detect(args)
    db = pkg.KeyFrameDatabase()
    for k in kfs:
        db.add(*k)
    for i in np.nonzero(erased)[0]:
        db.erase(int(i))
    assert len(db) == len(kfs)
    # the kernel's three arrays against plain set arithmetic
    common, first, scores = db.query(*q)
    for i in (0, 5, len(kfs) - 1):
        both = np.intersect1d(q[0], kfs[i][0])
        assert common[i] == (0 if erased[i] else len(both))
        assert first[i] == (len(q[0]) if erased[i] or len(both) == 0 else int(np.searchsorted(q[0], both[0])))
    state = db.reloc_score if reloc else db.loop_score
    state[:] = list(prev)
    gm, gw, gc = db._detect(q, reloc, connected, lambda k: neighbours[k], min_score)
    assert np.array_equal(gm, wm) and np.array_equal(gw[wm], ww[wm]) and list(gc) == wc.tolist() and len(wc) >= 1
    assert np.array_equal(np.array(state, np.float32).view(np.uint32), ws.view(np.uint32))
    if reloc:
        assert db.detect_relocalization_candidates(q, lambda k: neighbours[k]) == list(gc)   # scores now left by the first query
    db.clear()
    assert len(db) == 0 and db.detect_relocalization_candidates(q, lambda k: []) == []
    s = db.add(*kfs[0])
    assert s == 0 and db.query(*kfs[0])[0][0] == len(kfs[0][0])
    db.close()
