"""``CiderD`` with the signature of ``pyciderevalcap.ciderD.ciderD.CiderD``.

Construction site ``train.py:121`` (``CiderD(df=opt.train_cached_tokens)``),
call sites ``utils.py:72, 208, 274`` (``compute_score(gts, res)``).  The object
owns a device-resident :class:`DfTable`; the reward functions in
``cst_captioning_b200.reward`` use it directly on token rows, while
``compute_score`` keeps the string protocol so the reference's unmodified
``utils.py`` can drive the GPU path.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from ._lib import lib, check
from .df_table import DfTable, _row_dtype


def _parse(sentence, width):
    """An array_to_str string (utils.py:135-152) -> its token ids."""
    toks = sentence.split()
    if len(toks) > width:
        raise ValueError('caption of %d tokens exceeds the kernel row width %d' % (len(toks), width))
    try:
        ids = [int(t) for t in toks]
    except ValueError:
        raise ValueError('the GPU scorer works on token-id strings (utils.array_to_str output); '
                         'got %r' % sentence)
    if 1 in ids or 0 in ids[:-1]:
        raise ValueError('not an array_to_str string (<start> or an inner <end>): %r' % sentence)
    return ids


class WordIds:
    """Interns the words of one ``compute_score`` call: word -> token id.

    The reference hands its scorers raw caption WORDS wherever it scores text rather than
    model output (``compute_scores.py:81-92`` from the cocofmt json; ``utils.language_eval``,
    ``utils.py:114-132``).  CIDEr-D only ever compares n-grams for equality, so any injective
    word -> id map yields the scores of the strings themselves; the kernels take ids up to
    65534, which bounds the distinct words of one call.

    ``first_id``: lowest id handed out (ids below it stay free for words that ARE ids, see
    ``CiderD.compute_score``).  Ids 0 (<end>) and 1 (<start>) are never handed out."""

    def __init__(self, first_id=2, max_tokens=_lib.MAX_ROW_WIDTH, on_long='error'):
        if on_long not in ('error', 'truncate'):
            raise ValueError("on_long must be 'error' or 'truncate'")
        self.first_id = max(2, int(first_id))
        self.next_id = self.first_id
        self.map = {}
        self.max_tokens = int(max_tokens)
        self.on_long = on_long
        self.truncated = 0

    def ids(self, sentence, keep_numbers=False, sep=None):
        """``sep``: what ``str.split`` is given -- None for the scorers that cut on any white space
        (``precook``), ``' '`` for ROUGE's ``split(" ")``, whose empty tokens are words like any other."""
        toks = sentence.split(sep)
        if len(toks) > self.max_tokens:
            if self.on_long == 'error':
                raise ValueError('caption of %d words exceeds the %d tokens a kernel row holds (pass '
                                 "on_long='truncate' to score its first %d words instead): %r"
                                 % (len(toks), self.max_tokens, self.max_tokens, sentence[:80]))
            toks = toks[:self.max_tokens]
            self.truncated += 1
        out = []
        for t in toks:
            if keep_numbers and t.isdigit() and len(t) <= 5 and int(t) < self.first_id:
                out.append(int(t))
                continue
            i = self.map.get(t)
            if i is None:
                i = self.next_id
                if i > _lib.MAX_TOKEN_ID:
                    raise ValueError('more than %d distinct words in one compute_score call'
                                     % (_lib.MAX_TOKEN_ID - self.first_id + 1))
                self.map[t] = i
                self.next_id += 1
            out.append(i)
        return out


def _row(ids, width):
    row = np.full(width, -1, dtype=np.int32)          # -1 = padding: ends the caption, never emitted
    row[:len(ids)] = ids
    return row


class CiderD:
    """CIDEr-D scorer in cached-df mode (n = 4, sigma = 6.0, the defaults the
    reference constructs it with)."""

    def __init__(self, n=4, sigma=6.0, df=None, device=None):
        if n != 4 or float(sigma) != 6.0:
            raise ValueError('the sm_100a kernels implement n=4, sigma=6.0 (the reference defaults)')
        if isinstance(df, DfTable):
            self.table = df
        elif isinstance(df, (str, bytes)):
            self.table = DfTable.from_pickle(df, device)
        elif isinstance(df, tuple) and len(df) == 2:
            self.table = DfTable.from_dict(df[0], df[1], device)
        else:
            raise ValueError("df must be a *_ciderdf.pkl path, a (mapping, ref_len) pair or a DfTable; "
                             "the 'corpus' mode lives in cst_captioning_b200.consensus")
        self._n = n
        self._sigma = sigma

    # -- tensor level ----------------------------------------------------------
    def score_rows(self, cand, refs, cands_per_video, use_eos=0, greedy=None):
        """CIDEr-D of token rows.  ``cand`` (N, T) CUDA int tensor, ``refs`` a
        :class:`~cst_captioning_b200.reward.RefSet`; candidate i is scored against
        video ``i // cands_per_video``.  -> float64 CUDA tensor (N,), or (2N,) with
        ``greedy`` rows appended (``utils.py:180-183`` order)."""
        dev = self.table.device
        n = cand.shape[0]
        out = torch.empty(n * (2 if greedy is not None else 1), dtype=torch.float64, device=dev)
        check(lib.cst_cider_scores(
            self.table.handle, cand.data_ptr(), greedy.data_ptr() if greedy is not None else None,
            _row_dtype(cand), n, cand.shape[1], cand.stride(0),
            refs.rows.data_ptr(), _row_dtype(refs.rows), refs.rows.shape[1], refs.rows.stride(0),
            refs.offsets.data_ptr() if refs.offsets is not None else None, refs.max_refs,
            n // cands_per_video, cands_per_video, int(use_eos), out.data_ptr(),
            torch.cuda.current_stream(dev).cuda_stream))
        return out

    # -- string level (the reference protocol) -----------------------------------
    def compute_score(self, gts, res, on_long='error'):
        """``gts``: {image_id: [ref strings]}, ``res``: [{'image_id', 'caption': [str]}]
        -> (mean, float64 ndarray).  Strings are space-joined token ids as
        ``utils.array_to_str`` emits them (``utils.py:135-152``): those are looked up in the df
        table as they are.  Any other word (the generic call site ``utils.py:58-72`` feeds
        prediction text) cannot be in a table keyed by ids; it is interned above the table's
        largest id, i.e. scored as the unseen word it is."""
        from .reward import RefSet
        width = _lib.MAX_ROW_WIDTH
        n = len(res)
        if n == 0:
            return float('nan'), np.zeros(0)
        words = None

        def _ids(sentence):
            nonlocal words
            try:
                return _parse(sentence, width)
            except ValueError:
                if all(t.isdigit() for t in sentence.split()):
                    raise                                  # an id string that breaks the protocol, or too long
                if words is None:
                    words = WordIds(first_id=self.table.max_token_id() + 1, max_tokens=width, on_long=on_long)
                return words.ids(sentence, keep_numbers=True)

        cand = np.full((n, width), -1, dtype=np.int32)
        groups, ref_rows, counts = [], [], []
        last = None
        for i, entry in enumerate(res):
            hyp = entry['caption']
            refs = gts[entry['image_id']]
            assert type(hyp) is list and len(hyp) == 1
            assert type(refs) is list and len(refs) > 0
            ids = _ids(hyp[0])
            cand[i, :len(ids)] = ids
            if refs is last:
                groups[-1] += 1
                continue
            last = refs
            groups.append(1)
            counts.append(len(refs))
            for r in refs:
                ref_rows.append(_row(_ids(r), width))
        # The strings are already canonical: rows are the ids followed by the padding
        # id -1 (ends a caption without being emitted) and are scored with use_eos=1,
        # which keeps a trailing <end> = 0 exactly when the string has one.
        dev = self.table.device
        per = groups[0] if len(set(groups)) == 1 else None
        if per is None:        # uneven sharing of ref lists: one "video" per hypothesis
            expanded, counts2 = [], []
            at = 0
            for g, c in zip(groups, counts):
                for _ in range(g):
                    expanded.extend(ref_rows[at:at + c])
                    counts2.append(c)
                at += c
            ref_rows, counts, per = expanded, counts2, 1
        rows = torch.from_numpy(np.stack(ref_rows)).to(dev)
        if len(set(counts)) == 1:
            refset = RefSet(rows, per_video=counts[0])
        else:
            refset = RefSet(rows, offsets=np.concatenate([[0], np.cumsum(counts)]))
        scores = self.score_rows(torch.from_numpy(cand).to(dev), refset, per, use_eos=1)
        arr = scores.cpu().numpy()
        return np.mean(arr), arr


class Cider:
    """``pycocoevalcap.cider.cider.Cider`` (corpus mode) with the string protocol of
    ``compute_scores.py:67-68, 92``: ``compute_score(gts, res)`` with both dicts keyed by
    image id and ``res[id]`` a one-element list.  The document frequency is rebuilt on the
    device from the references of THIS call (``ref_len = log(number of images)``), then
    every hypothesis is scored with the same kernel as the RL path.

    The strings are whatever the caller scores: raw caption words (``compute_scores.py:81-92``,
    ``utils.language_eval``) or token-id strings; every distinct word of the call is interned to
    an id (:class:`WordIds`), so at most 65533 distinct words per call.  A kernel row holds 32
    tokens: a longer caption raises ``ValueError`` (``on_long='error'``, the default) or is cut
    to its first 32 words (``on_long='truncate'``; ``truncated`` counts them) -- the reference's
    own label rows are cut at 30 (``create_sequencelabel.py:48-52``).
    """

    def __init__(self, n=4, sigma=6.0, device=None, on_long='error'):
        if n != 4 or float(sigma) != 6.0:
            raise ValueError('the sm_100a kernels implement n=4, sigma=6.0 (the reference defaults)')
        self._device = torch.device(device if device is not None else 'cuda')
        self._on_long = on_long
        self.truncated = 0

    def compute_score(self, gts, res):
        from .reward import RefSet
        assert gts.keys() == res.keys()
        ids = list(gts.keys())
        width = _lib.MAX_ROW_WIDTH
        words = WordIds(max_tokens=width, on_long=self._on_long)
        counts, ref_rows = [], []
        cand = np.full((len(ids), width), -1, dtype=np.int32)
        for i, key in enumerate(ids):
            hyp, refs = res[key], gts[key]
            assert type(hyp) is list and len(hyp) == 1
            assert type(refs) is list and len(refs) > 0
            toks = words.ids(hyp[0])
            cand[i, :len(toks)] = toks
            counts.append(len(refs))
            for r in refs:
                ref_rows.append(_row(words.ids(r), width))
        self.truncated += words.truncated
        dev = self._device
        rows = torch.from_numpy(np.stack(ref_rows)).to(dev)
        offsets = np.concatenate([[0], np.cumsum(counts)])
        table = DfTable.from_rows(rows, video_offsets=offsets, use_eos=0)
        scorer = CiderD(df=table)
        refset = RefSet(rows, per_video=counts[0]) if len(set(counts)) == 1 else RefSet(rows, offsets=offsets)
        arr = scorer.score_rows(torch.from_numpy(cand).to(dev), refset, 1, use_eos=0).cpu().numpy()
        table.close()
        return np.mean(arr), arr


def load_gt_refs(cocofmt_file):
    """``utils.load_gt_refs`` (``utils.py:50-55``): {image_id: [caption, ...]} of a cocofmt json."""
    import json
    with open(cocofmt_file) as fh:
        d = json.load(fh)
    out = {}
    for a in d['annotations']:
        out.setdefault(a['image_id'], []).append(a['caption'])
    return out


_PTB_PUNCT = ("''", "'", "``", "`", "-LRB-", "-RRB-", "-LCB-", "-RCB-", ".", "?", "!", ",", ":", "-", "--", "...", ";")


def _tokenize(caption):
    """What coco-caption's PTBTokenizer leaves of a caption, without the Java tool: lower case,
    the punctuation it drops removed.  (MSR-VTT captions are lower-case words already.)"""
    text = caption.lower().replace('\n', ' ')
    for ch in '.,?!:;"()':
        text = text.replace(ch, ' ')
    return ' '.join(t for t in text.split() if t not in _PTB_PUNCT)


def language_eval(gold_file, pred_file, device=None, on_long='truncate'):
    """``utils.language_eval`` (``utils.py:114-132``), which ``train.py:35-42, 353-358`` calls on the
    validation predictions and ``check_model`` (``train.py:386-391``) reads through ``opt.eval_metric``:
    the corpus scores COCOEvalCap reports for ``pred_file``'s captions against the references of
    ``gold_file`` (both cocofmt json), on the device -> {'Bleu_1' .. 'Bleu_4', 'ROUGE_L', 'CIDEr'}, rounded
    to 5 places like the reference.  METEOR is a Java program and is not computed; the PTB tokenizer (Java
    as well) is replaced by lower-casing and its punctuation list."""
    import json
    from .metrics import Bleu, Rouge
    gts = load_gt_refs(gold_file)
    with open(pred_file) as fh:
        preds = json.load(fh)
    if isinstance(preds, dict):
        preds = preds.get('predictions', preds)
    res = {p['image_id']: [_tokenize(p['caption'])] for p in preds}
    gts = {k: [_tokenize(c) for c in gts[k]] for k in res}
    out = {}
    bleus, _ = Bleu(device=device, on_long=on_long).compute_score(gts, res)
    for k, b in enumerate(bleus):
        out['Bleu_%d' % (k + 1)] = round(float(b), 5)
    rouge, _ = Rouge(device=device, on_long=on_long).compute_score(gts, res)
    out['ROUGE_L'] = round(float(rouge), 5)
    score, _ = Cider(device=device, on_long=on_long).compute_score(gts, res)
    out['CIDEr'] = round(float(score), 5)
    return out
