"""Raw caption WORDS through the scorers: the reference scores text, not
token ids, in ``compute_scores.py:81-92`` (GT consensus scores from the cocofmt json) and in
``utils.language_eval`` (``utils.py:114-132``, read by ``check_model``, ``train.py:386-391``).
The loops below are the reference's own, with the scorer swapped."""
import json

import numpy as np
import pytest

from oracle import ciderd_oracle as co

WORDS = ['a', 'man', 'woman', 'is', 'are', 'playing', 'guitar', 'the', 'dog', 'runs', 'on', 'grass', 'cooking',
         'kitchen', 'in', 'car', 'drives', 'fast', 'people', 'dance', 'stage', 'and', 'sings', 'song', "it's", '42']


def _captions(n_videos, per_video, seed):
    rng = np.random.default_rng(seed)
    out = {}
    for v in range(n_videos):
        topic = rng.choice(len(WORDS), 8, replace=False)
        caps = []
        for _ in range(per_video):
            n = int(rng.integers(1, 12))
            caps.append(' '.join(WORDS[topic[int(rng.integers(0, 8))]] if rng.random() < 0.8
                                 else WORDS[int(rng.integers(0, len(WORDS)))] for _ in range(n)))
        out['video%d' % v] = caps
    return out


def test_word_ids_are_injective_and_bounded():
    import cst_captioning_b200 as cst
    w = cst.WordIds()
    a = w.ids('the dog runs the dog')
    assert a[0] == a[3] and a[1] == a[4] and len(set(a)) == 3 and min(a) >= 2      # never <end> / <start>
    assert w.ids('dog') == [a[1]]
    with pytest.raises(ValueError):
        cst.WordIds(max_tokens=4).ids('one two three four five')
    t = cst.WordIds(max_tokens=4, on_long='truncate')
    assert len(t.ids('one two three four five')) == 4 and t.truncated == 1
    small = cst.WordIds(first_id=65533)
    small.ids('x y')
    with pytest.raises(ValueError):
        small.ids('z')


@pytest.mark.gpu
@pytest.mark.parametrize('remove_in_ref', [True, False])
def test_compute_scores_loop_on_raw_words(remove_in_ref):
    """compute_scores.py:75-103 with ``Cider()`` swapped for this package's: leave-one-out
    consensus scores of every caption, corpus-mode df per pass, raw words in."""
    import cst_captioning_b200 as cst
    seq_per_img = 6
    gt_refs = _captions(30, seq_per_img, seed=1)

    def run(scorer):
        videos = sorted(gt_refs.keys())
        gt_scores = np.zeros((len(gt_refs), seq_per_img))
        for i in range(seq_per_img):
            preds_i = {v: [gt_refs[v][i]] for v in videos}
            if remove_in_ref:
                gt_refs_i = {v: gt_refs[v][:i] + gt_refs[v][i + 1:] for v in videos}
            else:
                gt_refs_i = gt_refs
            score_i, scores_i = scorer.compute_score(gt_refs_i, preds_i)
            gt_scores[:, i] = scores_i
            assert score_i == pytest.approx(np.mean(scores_i))
        return gt_scores

    got, want = run(cst.Cider()), run(co.Cider())
    assert want.max() > 0
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-12)


@pytest.mark.gpu
def test_cached_df_scorer_takes_words_it_has_no_ids_for():
    """utils.compute_score (utils.py:58-72) hands CiderD whatever text the caller has.  Id
    strings are looked up as ids; a word the table cannot know scores as an unseen n-gram."""
    import cst_captioning_b200 as cst
    df, n = co.build_df_from_strings([['3 4 5', '3 6 0'], ['3 4', '7 8 0'], ['9 4 0']])
    gts = {0: ['3 4 5 0', 'zebra 4 5'], 1: ['7 8 0', '3 unknown word 0']}
    res = [{'image_id': 0, 'caption': ['3 4 zebra']}, {'image_id': 1, 'caption': ['7 unknown word']},
           {'image_id': 1, 'caption': ['3 4 0']}]
    _, want = co.CiderD(df=(df, n)).compute_score(gts, res)
    scorer = cst.CiderD(df=(df, n))
    assert scorer.table.max_token_id() == 9
    _, got = scorer.compute_score(gts, res)
    assert want.max() > 0
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-12)


@pytest.mark.gpu
def test_long_captions_and_language_eval(tmp_path):
    import cst_captioning_b200 as cst
    long_cap = ' '.join('w%d' % i for i in range(40))
    gts, res = {0: [long_cap, 'w1 w2 w3']}, {0: ['w1 w2 w3 w4']}
    with pytest.raises(ValueError, match='32 tokens'):
        cst.Cider().compute_score(gts, res)
    cut = cst.Cider(on_long='truncate')
    _, got = cut.compute_score(gts, res)
    _, want = co.Cider().compute_score({0: [' '.join(long_cap.split()[:32]), 'w1 w2 w3']}, res)
    np.testing.assert_allclose(got, want, rtol=1e-9)
    assert cut.truncated == 1

    # utils.language_eval (utils.py:114-132) on cocofmt files -> the CIDEr check_model reads
    refs = _captions(12, 5, seed=2)
    ids = sorted(refs)
    gold = {'images': [{'id': k} for k in ids],
            'annotations': [{'image_id': k, 'id': j, 'caption': c.capitalize() + '.'}
                            for j, (k, c) in enumerate((k, c) for k in ids for c in refs[k])]}
    preds = [{'image_id': k, 'caption': refs[k][0] + ' , the'} for k in ids]
    (tmp_path / 'gold.json').write_text(json.dumps(gold))
    (tmp_path / 'pred.json').write_text(json.dumps(preds))
    out = cst.language_eval(str(tmp_path / 'gold.json'), str(tmp_path / 'pred.json'))
    g2, r2 = {k: refs[k] for k in ids}, {k: [refs[k][0] + ' the'] for k in ids}
    score, _ = co.Cider().compute_score(g2, r2)
    assert out['CIDEr'] == round(float(score), 5)
    # ... and the other corpus scores COCOEvalCap reports (Bleu_1..4, ROUGE_L; METEOR is a Java tool)
    from oracle import bleu_rouge_oracle as bo
    bleus, _ = bo.Bleu().compute_score(g2, r2)
    rouge, _ = bo.Rouge().compute_score(g2, r2)
    want = {'Bleu_%d' % (k + 1): round(float(b), 5) for k, b in enumerate(bleus)}
    want.update({'ROUGE_L': round(float(rouge), 5), 'CIDEr': round(float(score), 5)})
    assert out == want
