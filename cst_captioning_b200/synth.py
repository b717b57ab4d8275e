"""Synthetic MSR-VTT-shaped token corpus.

There is no network for the real dataset, so tests and ``bench.py`` draw token
rows with the shapes and conventions the reference's pipeline produces:

* ids 0 = ``<end>``, 1 = ``<start>``, 2 = ``<unk>`` (``build_vocab.py:41``),
  words 3..vocab-1 drawn Zipf ``p(rank) ~ rank**-1.05``;
* words per caption ``clip(round(N(9.3, 3.5)), 2, 28)``;
* reference rows ``[1, w_1..w_L, 0, 0-pad]`` of width 30
  (``create_sequencelabel.py:42-52, 82-84``);
* sampled rows ``[w.., 0, 0-pad]`` of width 28 without ``<start>``
  (``model.py:255-270``): a reference of the same video with each word
  replaced by a fresh Zipf draw with probability ``p``.

Pure numpy and vectorised; everything is a function of the seed
(reference default seed 123, ``opts.py:231-235``).
"""
from __future__ import annotations

import numpy as np

VOCAB = 10000
REF_WIDTH = 30
CAND_WIDTH = 28
FIRST_WORD = 3
SEED = 123


def _zipf_cdf(vocab=VOCAB, s=1.05):
    ranks = np.arange(1, vocab - FIRST_WORD + 1, dtype=np.float64)
    p = ranks ** (-s)
    return np.cumsum(p / p.sum())


def _draw_words(rng, shape, cdf):
    u = rng.random(shape)
    idx = np.searchsorted(cdf, u, side='right')
    np.minimum(idx, len(cdf) - 1, out=idx)
    return (idx + FIRST_WORD).astype(np.int32)


def make_refs(rng, n_videos, n_refs=20, width=REF_WIDTH, vocab=VOCAB):
    """-> (rows int32 (n_videos, n_refs, width), n_words int32 (n_videos, n_refs))."""
    cdf = _zipf_cdf(vocab)
    max_words = min(28, width - 2)
    n_words = np.clip(np.rint(rng.normal(9.3, 3.5, (n_videos, n_refs))), 2,
                      max_words).astype(np.int32)
    words = _draw_words(rng, (n_videos, n_refs, max_words), cdf)
    rows = np.zeros((n_videos, n_refs, width), dtype=np.int32)
    col = np.arange(max_words, dtype=np.int32)
    keep = col[None, None, :] < n_words[:, :, None]
    rows[:, :, 1:1 + max_words] = np.where(keep, words, 0)
    rows[:, :, 0] = 1
    return rows, n_words


def corrupt(rng, refs, n_words, per_video, p, width=CAND_WIDTH, vocab=VOCAB):
    """``per_video`` corrupted copies of a randomly chosen reference per video.

    -> rows int32 (n_videos * per_video, width), video-major (the layout
    ``expand_feat=1`` assumes: candidate i belongs to video i // per_video).
    """
    cdf = _zipf_cdf(vocab)
    n_videos, n_refs, ref_width = refs.shape
    src = rng.integers(0, n_refs, (n_videos, per_video))
    vid = np.arange(n_videos)[:, None]
    base = refs[vid, src][:, :, 1:1 + width]            # drop <start>
    if base.shape[2] < width:
        base = np.pad(base, ((0, 0), (0, 0), (0, width - base.shape[2])))
    length = np.minimum(n_words[vid, src], width)
    fresh = _draw_words(rng, base.shape, cdf)
    swap = rng.random(base.shape) < p
    col = np.arange(width)[None, None, :]
    is_word = col < length[:, :, None]
    out = np.where(is_word & swap, fresh, base)
    out = np.where(is_word, out, 0).astype(np.int32)
    return out.reshape(n_videos * per_video, width)


def make_batch(n_videos, seq_per_img=20, n_refs=20, seed=SEED, greedy=False,
               vocab=VOCAB):
    """One synthetic reward-path batch.

    -> dict with ``refs`` (n_videos, n_refs, 30) int32, ``n_words``,
    ``sampled`` (n_videos*seq_per_img, 28) int32 and, if asked, ``greedy``
    of the same shape (one corruption per video replicated seq_per_img times,
    as ``model.sample`` with ``expand_feat=1`` yields identical greedy rows).
    """
    rng = np.random.default_rng(seed)
    refs, n_words = make_refs(rng, n_videos, n_refs, vocab=vocab)
    out = {'refs': refs, 'n_words': n_words,
           'sampled': corrupt(rng, refs, n_words, seq_per_img, 0.3, vocab=vocab)}
    if greedy:
        one = corrupt(rng, refs, n_words, 1, 0.15, vocab=vocab)
        out['greedy'] = np.repeat(one, seq_per_img, axis=0)
    return out


def df_corpus_strings(refs, n_words):
    """The id-joined strings ``compute_ciderdf.main`` builds
    (``compute_ciderdf.py:112-118, 84``): words then ``<end>`` (id 0), no
    ``<start>``.  -> list (videos) of lists (captions) of str."""
    out = []
    for v in range(refs.shape[0]):
        caps = []
        for r in range(refs.shape[1]):
            w = refs[v, r, 1:1 + int(n_words[v, r])]
            caps.append(' '.join(map(str, w.tolist())) + ' 0')
        out.append(caps)
    return out
