"""Float64 CPU restatement of the CIDEr-D scorer.  TEST INFRASTRUCTURE ONLY.

PARITY STATUS: **unpinned for the scorer arithmetic.**  The reference imports
its scorer from two un-vendored, unpinned checkouts
(``/root/reference/README.md:9-12`` -> github.com/xiadingZ/cider, package
``pyciderevalcap``; github.com/xiadingZ/coco-caption, package
``pycocoevalcap``; import sites ``/root/reference/utils.py:8-17``,
``train.py:23-30``, ``compute_scores.py:21-28``).  Neither is present under
``/root/reference`` nor installed, and the reference holds no tests or golden
vectors.  What follows restates the published algorithm of that lineage
(ruotianluo/cider ``ciderD_scorer.py`` and tylin/coco-caption
``cider_scorer.py``) keeping its float64 operation order, and is anchored on
the pieces the reference *does* contain:

* n-gram extraction      -> ``/root/reference/compute_ciderdf.py:19-34`` (precook)
* per-video ref cooking  -> ``/root/reference/compute_ciderdf.py:37-53``
* document frequency     -> ``/root/reference/compute_ciderdf.py:56-70``
* df pickle schema       -> ``/root/reference/compute_ciderdf.py:123-129``
* call protocol (df mode)-> ``/root/reference/utils.py:194-208, 260-274``
* call protocol (corpus) -> ``/root/reference/compute_scores.py:81-103``

``oracle/make_golden.py`` runs the reference's own ``compute_ciderdf.py`` and
``utils.py`` (unmodified, with this scorer injected for the absent import) to
pin everything except the scorer arithmetic itself.
"""
from __future__ import annotations

import math
import pickle
from collections import defaultdict

import numpy as np

N_ORDERS = 4
SIGMA = 6.0


# ---------------------------------------------------------------------------
# n-grams and document frequency (restates compute_ciderdf.py:19-70)
# ---------------------------------------------------------------------------
def ngram_counts(sentence, n=N_ORDERS):
    """Multiset of 1..n-grams of a whitespace-tokenised sentence.

    Follows ``compute_ciderdf.py:19-34``: keys are tuples of the split words,
    orders outer, positions inner (so dict insertion order is (order, first
    position) -- the order the CUDA parity probe reports n-grams in).
    """
    toks = sentence.split()
    table = defaultdict(int)
    for order in range(1, n + 1):
        last = len(toks) - order
        pos = 0
        while pos <= last:
            table[tuple(toks[pos:pos + order])] += 1
            pos += 1
    return table


def cook_video(refs, n=N_ORDERS):
    """One count table per reference caption (``compute_ciderdf.py:37-45``)."""
    return [ngram_counts(r, n) for r in refs]


def document_frequency(cooked_videos):
    """df[g] = number of videos whose reference set holds g at least once.

    Follows ``compute_ciderdf.py:56-70``: per video the *set union* of its
    refs' n-grams is taken, each member gets += 1 (float valued).
    """
    df = defaultdict(float)
    for cooked in cooked_videos:
        present = set()
        for table in cooked:
            present.update(table.keys())
        for g in present:
            df[g] += 1
    return df


def build_df_from_strings(videos_refs):
    """``compute_ciderdf.build_dict`` for already id-joined strings
    (``compute_ciderdf.py:73-93``): returns (df dict, number of videos)."""
    cooked = [cook_video(refs) for refs in videos_refs]
    return document_frequency(cooked), len(videos_refs)


# ---------------------------------------------------------------------------
# scorer core (restates upstream CiderScorer.counts2vec / sim / compute_cider)
# ---------------------------------------------------------------------------
class _ScorerCore:
    """tf-idf vectors, clipped cosine and Gaussian length penalty in float64."""

    def __init__(self, df, log_ref_len, n=N_ORDERS, sigma=SIGMA):
        self.df = df                    # mapping ngram tuple -> float count
        self.log_ref_len = log_ref_len  # np.log(float(number of documents))
        self.n = n
        self.sigma = sigma

    def vectorise(self, counts):
        """upstream ``counts2vec``: vec[n][g] = tf * (ref_len - log(max(1, df)))."""
        vec = [dict() for _ in range(self.n)]
        norm = [0.0] * self.n
        length = 0
        for g, tf in counts.items():
            # unseen n-grams count as df = 1 (upstream clamps with max(1.0, .))
            logdf = np.log(max(1.0, self.df.get(g, 0.0)))
            o = len(g) - 1
            w = float(tf) * (self.log_ref_len - logdf)
            vec[o][g] = w
            norm[o] += pow(w, 2)
            if o == 1:                  # the "length" is the bigram count
                length += tf
        return vec, [np.sqrt(x) for x in norm], length

    def similarity(self, vh, vr, nh, nr, lh, lr):
        """upstream ``sim``: clipped product, cosine normalisation, penalty."""
        delta = float(lh - lr)
        out = np.array([0.0 for _ in range(self.n)])
        for o in range(self.n):
            ref_o = vr[o]
            for g, wh in vh[o].items():
                wr = ref_o.get(g, 0.0)
                out[o] += min(wh, wr) * wr
            if nh[o] != 0 and nr[o] != 0:
                out[o] /= (nh[o] * nr[o])
            assert not math.isnan(out[o])
            out[o] *= np.e ** (-(delta ** 2) / (2 * self.sigma ** 2))
        return out

    def score_one(self, hyp_counts, refs_counts):
        """upstream ``compute_cider`` body for one (hypothesis, refs) pair."""
        vh, nh, lh = self.vectorise(hyp_counts)
        acc = np.array([0.0 for _ in range(self.n)])
        for rc in refs_counts:
            vr, nr, lr = self.vectorise(rc)
            acc += self.similarity(vh, vr, nh, nr, lh, lr)
        s = np.mean(acc)
        s /= len(refs_counts)
        s *= 10.0
        return s


def load_df_pickle(path):
    """Read the ``{'document_frequency', 'ref_len'}`` pickle written at
    ``compute_ciderdf.py:123-129`` -> (df mapping, document count)."""
    with open(path, 'rb') as fh:
        blob = pickle.load(fh, encoding='latin1')
    return blob['document_frequency'], blob['ref_len']


class CiderD:
    """Stand-in for ``pyciderevalcap.ciderD.ciderD.CiderD`` in cached-df mode.

    Call sites: ``train.py:121`` (construction with a pickle path),
    ``utils.py:208,274`` (``compute_score(gts, res)`` with ``res`` a list of
    ``{'image_id', 'caption': [str]}``).  ``df`` may be a pickle path or a
    ``(mapping, n_documents)`` pair.
    """

    def __init__(self, n=N_ORDERS, sigma=SIGMA, df=None):
        if isinstance(df, (str, bytes)):
            table, ndoc = load_df_pickle(df)
        else:
            table, ndoc = df
        self._n = n
        self._sigma = sigma
        self._core = _ScorerCore(table, np.log(float(ndoc)), n, sigma)
        self._cook_cache = None

    def compute_score(self, gts, res):
        core = self._core
        scores = []
        for entry in res:
            hyp = entry['caption']
            refs = gts[entry['image_id']]
            assert type(hyp) is list and len(hyp) == 1
            assert type(refs) is list and len(refs) > 0
            # upstream re-cooks the refs for every hypothesis (cook_append)
            hyp_counts = ngram_counts(hyp[0], self._n)
            refs_counts = [ngram_counts(r, self._n) for r in refs]
            scores.append(core.score_one(hyp_counts, refs_counts))
        arr = np.array(scores)
        return np.mean(arr), arr


class Cider:
    """Stand-in for coco-caption's ``pycocoevalcap.cider.cider.Cider``
    (corpus mode): df is recomputed from the refs of *this* call and
    ``ref_len = log(number of images in the call)``.  Call site
    ``compute_scores.py:67-68, 92``: ``compute_score(gts, res)`` with both
    dicts keyed by image id, ``res[id]`` a one-element list.
    """

    def __init__(self, n=N_ORDERS, sigma=SIGMA):
        self._n = n
        self._sigma = sigma

    def compute_score(self, gts, res):
        assert gts.keys() == res.keys()
        ids = list(gts.keys())
        cooked_h, cooked_r = [], []
        for i in ids:
            hyp, refs = res[i], gts[i]
            assert type(hyp) is list and len(hyp) == 1
            assert type(refs) is list and len(refs) > 0
            cooked_h.append(ngram_counts(hyp[0], self._n))
            cooked_r.append([ngram_counts(r, self._n) for r in refs])
        df = document_frequency(cooked_r)
        assert len(cooked_h) >= max(df.values())
        core = _ScorerCore(df, np.log(float(len(cooked_r))), self._n, self._sigma)
        scores = [core.score_one(h, r) for h, r in zip(cooked_h, cooked_r)]
        arr = np.array(scores)
        return np.mean(arr), arr
