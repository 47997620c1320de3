"""Python mirror of the DBoW2 side of Frame::ComputeBoW (reference src/Frame.cc:572-583 ->
ORBVocabulary::transform, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1259): the per-descriptor tree descent runs
on the GPU (orbm_bow_transform); the folding of (word, weight, node) triples into BowVector / FeatureVector is the
caller-side container logic of the reference (TF-IDF weighting + L1 normalisation, as ORB-SLAM2 configures DBoW2)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check


def synthetic_vocabulary(k=10, L=3, seed=0, stop_p=0.02):
    """Random k-ary tree of depth L (ORBvoc.txt is not available offline): dict in the flat form orbm_vocab_create takes."""
    rng = np.random.Generator(np.random.PCG64(seed))
    child_start, child_ids, level_of = [0], [], [0]
    frontier, n = [0], 1
    children_of = {}
    for lvl in range(1, L + 1):
        nxt = []
        for p in frontier:
            ids = list(range(n, n + k))
            n += k
            children_of[p] = ids
            nxt.extend(ids)
            level_of.extend([lvl] * k)
        frontier = nxt
    for i in range(n):
        child_ids.extend(children_of.get(i, []))
        child_start.append(len(child_ids))
    word_id = np.full(n, -1, np.int32)
    leaves = [i for i in range(n) if i not in children_of]
    word_id[leaves] = np.arange(len(leaves), dtype=np.int32)
    weight = np.zeros(n, np.float64)
    weight[leaves] = np.where(rng.random(len(leaves)) < stop_p, 0.0, rng.random(len(leaves)) * 8 + 0.1)
    return dict(L=L, child_start=np.array(child_start, np.int32), child_ids=np.array(child_ids, np.int32),
                node_desc=rng.integers(0, 256, (n, 32), dtype=np.uint8), word_id=word_id, weight=weight)


def fold(word, node, weight):
    """TemplatedVocabulary::transform(features, v, fv, levelsup) container logic (:1127-1192) for TF_IDF + L1:
    BowVector as (ascending word ids, values), FeatureVector as CSR (ascending node ids)."""
    bow = {}
    fv = {}
    for i in range(len(word)):
        w = float(weight[i])
        if w > 0:  # not stopped
            wid = int(word[i])
            bow[wid] = bow[wid] + w if wid in bow else w          # BowVector::addWeight
            fv.setdefault(int(node[i]), []).append(i)             # FeatureVector::addFeature
    ids = sorted(bow)
    vals = np.array([bow[i] for i in ids], np.float64)
    norm = 0.0
    for v in vals:                                                 # BowVector::normalize(L1): sequential, map order
        norm += abs(float(v))
    if norm > 0.0:
        vals = vals / norm
    nids = sorted(fv)
    start, feat = [0], []
    for nid in nids:
        feat.extend(fv[nid])
        start.append(len(feat))
    return (np.array(ids, np.int32), vals), dict(node_id=np.array(nids, np.int32), node_start=np.array(start, np.int32),
                                                 node_feat=np.array(feat, np.int32))


class ORBVocabulary:
    def __init__(self, vocab, matcher):
        self._L = _lib.load()
        self._matcher = matcher
        self._keep = [np.ascontiguousarray(vocab[k]) for k in ("child_start", "child_ids", "node_desc", "word_id", "weight")]
        v = C.c_void_p()
        check(self._L.orbm_vocab_create(matcher._m, len(vocab["word_id"]), int(vocab["L"]), *[a.ctypes.data for a in self._keep],
                                        C.byref(v)), "orbm_vocab_create")
        self._v = v

    def close(self):
        if getattr(self, "_v", None):
            self._L.orbm_vocab_destroy(self._v)
            self._v = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform_features(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8)
        n = len(desc)
        word, node, weight = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        check(self._L.orbm_bow_transform(self._matcher._m, self._v, desc.ctypes.data, n, int(levelsup), word.ctypes.data,
                                         node.ctypes.data, weight.ctypes.data), "orbm_bow_transform")
        return word[:n], node[:n], weight[:n]

    def transform(self, desc, levelsup=4):
        """(BowVector, FeatureVector) like ORBVocabulary::transform(vCurrentDesc, mBowVec, mFeatVec, 4)."""
        return fold(*self.transform_features(desc, levelsup))
