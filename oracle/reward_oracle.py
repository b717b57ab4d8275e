"""CPU restatement of the reference's reward functions.  TEST INFRASTRUCTURE ONLY.

PARITY STATUS: pinned against the reference itself -- ``oracle/make_golden.py``
imports the unmodified ``/root/reference/utils.py`` in the build container and
the fixtures under ``tests/golden/`` hold its outputs; ``tests/test_oracle.py``
checks this file against them.  (The scorer those functions call is the
restated one in ``ciderd_oracle.py``; its arithmetic is unpinned, see there.)

This file exists because ``/root/reference`` is not present on the GPU box:
the CPU baseline timed by ``bench.py`` and the checker used by the ``-m gpu``
tests need the same behaviour from code that travels with the repo.
"""
from __future__ import annotations

import numpy as np

from .ciderd_oracle import CiderD


def row_to_words(row, use_eos=0):
    """Canonical token list of one id row (``utils.py:135-152``).

    * without ``use_eos`` the first 0 ends the caption and is dropped;
    * id 1 (``<start>``) is skipped wherever it occurs;
    * with ``use_eos`` the first 0 is emitted, then the caption ends.
    """
    words = []
    for tok in row:
        tok = int(tok)
        if tok == 0 and not use_eos:
            break
        if tok == 1:
            continue
        words.append(tok)
        if tok == 0:
            break
    return words


def row_to_str(row, use_eos=0):
    return ' '.join(str(t) for t in row_to_words(row, use_eos))


def _strings_for(model_rows, data_gts, use_eos):
    hyp = [[row_to_str(r, use_eos)] for r in model_rows]
    refs = [[row_to_str(r, use_eos) for r in video] for video in data_gts]
    return hyp, refs


def _video_of(i, n_cand, expand_feat, seq_per_img):
    """candidate index -> index into data_gts (``utils.py:200-206, 266-272``)."""
    j = i % n_cand
    return j // seq_per_img if expand_feat == 1 else j


def score_rows(model_rows, data_gts, scorer, expand_feat, seq_per_img, use_eos,
               n_cand=None):
    """Scores of every row of ``model_rows`` against its video's refs through
    ``scorer.compute_score``: the list-of-dict protocol for ``CiderD``
    (``utils.py:194-198, 260-264``), the dict protocol for any other scorer;
    of ``Bleu``'s four score lists the last one, BLEU-4, counts
    (``utils.py:210-218, 276-282``)."""
    from .bleu_rouge_oracle import Bleu
    hyp, refs = _strings_for(model_rows, data_gts, use_eos)
    total = len(hyp)
    n_cand = total if n_cand is None else n_cand
    if isinstance(scorer, CiderD):
        res = [{'image_id': i, 'caption': hyp[i]} for i in range(total)]
    else:
        res = {i: hyp[i] for i in range(total)}
    gts = {i: refs[_video_of(i, n_cand, expand_feat, seq_per_img)]
           for i in range(total)}
    _, scores = scorer.compute_score(gts, res)
    if isinstance(scorer, Bleu):
        scores = scores[-1]
    if type(scores) == list:
        scores = np.array(scores)
    return scores


def get_self_critical_reward(model_res, greedy_res, data_gts, bcmr_scorer,
                             expand_feat=0, seq_per_img=20, use_eos=0):
    """``utils.py:169-226``: sampled and greedy rows scored in one call,
    reward = sampled - greedy, broadcast over the T columns."""
    n = model_res.shape[0]
    both = list(model_res) + list(greedy_res)
    scores = score_rows(both, data_gts, bcmr_scorer, expand_feat, seq_per_img,
                        use_eos, n_cand=n)
    m_score = np.mean(scores[:n])
    g_score = np.mean(scores[n:])
    diff = scores[:n] - scores[n:]
    rewards = np.repeat(diff[:, np.newaxis], model_res.shape[1], 1)
    return rewards, m_score, g_score


def get_cst_reward(model_res, data_gts, bcmr_scorer, bcmrscores=None,
                   expand_feat=0, seq_per_img=20, scb_captions=20,
                   scb_baseline=1, use_eos=0, use_mixer=0):
    """``utils.py:229-324``: CST reward with the GT-consensus (1) or
    model-sample (2) SCB baseline = mean of the ``scb_captions`` LOWEST scores
    of the row (the caption itself included)."""
    if bcmrscores is None or use_mixer == 1:
        scores = score_rows(list(model_res), data_gts, bcmr_scorer,
                            expand_feat, seq_per_img, use_eos)
        scores = scores.reshape(-1, seq_per_img)
    else:
        scores = bcmrscores.copy()

    if scb_captions > 0:
        by_row = np.sort(scores, axis=1)
        if scb_baseline == 1:
            base_rows = np.sort(bcmrscores, axis=1)
            m_score = np.mean(scores)
            b_score = np.mean(bcmrscores)
        elif scb_baseline == 2:
            base_rows = by_row
            m_score = np.mean(by_row)
            b_score = np.mean(by_row[:, :scb_captions])
        else:
            raise ValueError('unknown scb_baseline!')
        for v in range(scores.shape[0]):
            scores[v] = scores[v] - np.mean(base_rows[v, :scb_captions])
    else:
        m_score = np.mean(scores)
        b_score = 0

    flat = scores.reshape(-1)
    rewards = np.repeat(flat[:, np.newaxis], model_res.shape[1], 1)
    return rewards, m_score, b_score


def consensus_scores(videos_refs, seq_per_img=20, remove_in_ref=True):
    """CIDEr column of ``compute_scores.py:75-103``: caption i of every video
    scored against the video's other captions, corpus-mode df per pass."""
    from .ciderd_oracle import Cider
    vids = list(range(len(videos_refs)))
    out = np.zeros((len(vids), seq_per_img))
    scorer = Cider()
    for i in range(seq_per_img):
        preds = {v: [videos_refs[v][i]] for v in vids}
        if remove_in_ref:
            gts = {v: videos_refs[v][:i] + videos_refs[v][i + 1:] for v in vids}
        else:
            gts = {v: list(videos_refs[v]) for v in vids}
        _, s = scorer.compute_score(gts, preds)
        out[:, i] = s
    return out
