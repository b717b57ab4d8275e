"""Float64 CPU restatement of the BLEU and ROUGE-L sentence scorers.  TEST INFRASTRUCTURE ONLY.

PARITY STATUS: **unpinned.**  The reference builds its reward scorer from
``{'Bleu_4': Bleu(), 'CIDEr': CiderD(...), 'METEOR': Meteor(), 'ROUGE_L': Rouge()}[opt.eval_metric]``
(``/root/reference/train.py:119-124``) and imports ``Bleu`` / ``Rouge`` from the un-vendored, unpinned
``coco-caption`` checkout (``/root/reference/README.md:9-12`` -> github.com/xiadingZ/coco-caption,
package ``pycocoevalcap``; import sites ``utils.py:10-17``, ``train.py:23-30``).  Neither file is under
``/root/reference`` nor installed, and the reference holds no tests or golden vectors for them.  What
follows restates the published algorithm of that lineage (tylin/coco-caption ``bleu/bleu.py``,
``bleu/bleu_scorer.py``, ``rouge/rouge.py``), keeping its float64 operation order, anchored on the
call protocol the reference does contain:

* how the scorer is called        -> ``utils.py:176-208`` (self-critical), ``utils.py:246-274`` (CST):
  ``gts``: dict id -> list of reference strings, ``res``: dict id -> [candidate string] (the list-of-dict
  form of ``utils.py:194-198`` is for ``CiderD`` only)
* what is taken from the result   -> ``utils.py:210-218, 276-282``: for ``Bleu`` the last of the four
  (score, scores) entries, i.e. BLEU-4 per sentence; lists become arrays
* the strings                     -> ``utils.py:135-152`` (``array_to_str``)

``tests/test_metrics.py`` holds hand-derived values for both scorers.  METEOR is a Java
program in that checkout and has no counterpart here.
"""
from __future__ import annotations

import math
from collections import defaultdict

import numpy as np

SMALL = 1e-9
TINY = 1e-15      # "so that if guess is 0 still return 0"


# ---------------------------------------------------------------------------
# BLEU (bleu_scorer.py: precook / cook_refs / cook_test / compute_score(option='closest'))
# ---------------------------------------------------------------------------
def _precook(sentence, n=4):
    """(number of words, {n-gram tuple: count}) for orders 1..n; words = ``sentence.split()``."""
    words = sentence.split()
    counts = defaultdict(int)
    for k in range(1, n + 1):
        for i in range(len(words) - k + 1):
            counts[tuple(words[i:i + k])] += 1
    return len(words), counts


def _cook_refs(refs, n=4):
    """(list of reference lengths, {n-gram: largest count in any one reference})."""
    lengths, maxcounts = [], {}
    for ref in refs:
        length, counts = _precook(ref, n)
        lengths.append(length)
        for gram, c in counts.items():
            maxcounts[gram] = max(maxcounts.get(gram, 0), c)
    return lengths, maxcounts


def _cook_test(test, cooked_refs, n=4):
    """Sufficient statistics of one candidate: length, guessed and clipped-correct n-gram counts."""
    ref_lengths, ref_max = cooked_refs
    testlen, counts = _precook(test, n)
    comps = {'reflen': ref_lengths, 'testlen': testlen,
             'guess': [max(0, testlen - k + 1) for k in range(1, n + 1)],
             'correct': [0] * n}
    for gram, c in counts.items():
        comps['correct'][len(gram) - 1] += min(ref_max.get(gram, 0), c)
    return comps


def _closest_reflen(ref_lengths, testlen):
    """``min((abs(l - testlen), l) for l in reflens)[1]``: nearest length, the shorter one on a tie."""
    return min((abs(l - testlen), l) for l in ref_lengths)[1]


class Bleu:
    """``pycocoevalcap.bleu.bleu.Bleu(n=4)``: ``compute_score(gts, res)`` -> (4 corpus scores,
    4 lists of sentence scores).  The reference's reward reads entry ``[-1]`` of both."""

    def __init__(self, n=4):
        self._n = n

    def compute_score(self, gts, res):
        assert list(gts.keys()) == list(res.keys())
        n = self._n
        ctest = []
        for key in gts.keys():
            hypo, ref = res[key], gts[key]
            assert type(hypo) is list and len(hypo) == 1
            assert type(ref) is list and len(ref) >= 1
            ctest.append(_cook_test(hypo[0], _cook_refs(ref, n), n))
        # BleuScorer.compute_score(option='closest')
        bleu_list = [[] for _ in range(n)]
        tot_test = tot_ref = 0
        total = {'guess': [0] * n, 'correct': [0] * n}
        for comps in ctest:
            testlen = comps['testlen']
            tot_test += testlen
            reflen = _closest_reflen(comps['reflen'], testlen)
            tot_ref += reflen
            for key in ('guess', 'correct'):
                for k in range(n):
                    total[key][k] += comps[key][k]
            bleu = 1.
            for k in range(n):
                bleu *= (float(comps['correct'][k]) + TINY) / (float(comps['guess'][k]) + SMALL)
                bleu_list[k].append(bleu ** (1. / (k + 1)))
            ratio = (testlen + TINY) / (reflen + SMALL)      # N.B.: avoids a zero division
            if ratio < 1:
                for k in range(n):
                    bleu_list[k][-1] *= math.exp(1 - 1 / ratio)
        bleus = []
        bleu = 1.
        for k in range(n):
            bleu *= float(total['correct'][k] + TINY) / (total['guess'][k] + SMALL)
            bleus.append(bleu ** (1. / (k + 1)))
        ratio = (tot_test + TINY) / (tot_ref + SMALL)
        if ratio < 1:
            for k in range(n):
                bleus[k] *= math.exp(1 - 1 / ratio)
        return bleus, bleu_list


# ---------------------------------------------------------------------------
# ROUGE-L (rouge.py: my_lcs / Rouge.calc_score / compute_score)
# ---------------------------------------------------------------------------
def lcs_length(a, b):
    """Length of the longest common subsequence of two token lists (the textbook table)."""
    if len(a) < len(b):
        a, b = b, a
    table = [[0] * (len(b) + 1) for _ in range(len(a) + 1)]
    for j in range(1, len(b) + 1):
        for i in range(1, len(a) + 1):
            if a[i - 1] == b[j - 1]:
                table[i][j] = table[i - 1][j - 1] + 1
            else:
                table[i][j] = max(table[i - 1][j], table[i][j - 1])
    return table[len(a)][len(b)]


class Rouge:
    """``pycocoevalcap.rouge.rouge.Rouge``: F-measure (beta = 1.2) of the best LCS precision and the
    best LCS recall over the references.  Sentences are cut with ``split(" ")``: the empty string is
    ONE empty token (so an empty candidate has length 1 and matches only an empty reference)."""

    def __init__(self):
        self.beta = 1.2

    def calc_score(self, candidate, refs):
        assert len(candidate) == 1
        assert len(refs) > 0
        prec, rec = [], []
        token_c = candidate[0].split(" ")
        for reference in refs:
            token_r = reference.split(" ")
            lcs = lcs_length(token_r, token_c)
            prec.append(lcs / float(len(token_c)))
            rec.append(lcs / float(len(token_r)))
        prec_max, rec_max = max(prec), max(rec)
        if prec_max != 0 and rec_max != 0:
            return ((1 + self.beta ** 2) * prec_max * rec_max) / float(rec_max + self.beta ** 2 * prec_max)
        return 0.0

    def compute_score(self, gts, res):
        assert list(gts.keys()) == list(res.keys())
        score = []
        for key in gts.keys():
            hypo, ref = res[key], gts[key]
            score.append(self.calc_score(hypo, ref))
            assert type(hypo) is list and len(hypo) == 1
            assert type(ref) is list and len(ref) > 0
        return np.mean(np.array(score)), np.array(score)
