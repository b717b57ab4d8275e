"""B200-native CIDEr-D reward path of CST captioning (one hot path, drop-in).

Public surface (mirrors the reference's interfaces for this path):

* :class:`CiderD`                      <- ``pyciderevalcap.ciderD.ciderD.CiderD``
* :class:`Cider`                       <- ``pycocoevalcap.cider.cider.Cider`` (corpus mode, ``compute_scores.py:67-68``;
                                       raw caption words or id strings)
* :class:`Bleu`, :class:`Rouge`        <- ``pycocoevalcap`` ``Bleu`` / ``Rouge``: the other reward scorers of
                                       ``train.py:119-124`` (``--eval_metric Bleu_4`` / ``ROUGE_L``)
* :func:`language_eval`                <- ``utils.language_eval`` (``utils.py:114-132``): Bleu_1..4, ROUGE_L, CIDEr
* :func:`get_cst_reward`               <- ``utils.get_cst_reward`` (``utils.py:229``)
* :func:`get_self_critical_reward`     <- ``utils.get_self_critical_reward`` (``utils.py:169``)
* :class:`DfTable`                     <- the ``*_ciderdf.pkl`` of ``compute_ciderdf.py``
* :class:`RefSet`                      device-resident reference rows
* :class:`CookedRefs`                  references staged once (the kernel's reference half kept in HBM)
* :class:`RewardCriterion`             <- ``model.RewardCriterion`` (``model.py:7-23``), fused fwd/bwd
* :class:`HostRewardPipeline`          the same reward for pinned host batches (upload overlapped with scoring)

Importing this package loads ``libcst_reward.so`` (hand-written CUDA for
sm_100a behind the C ABI of ``include/cst_reward.h``) and fails if it has not
been built: there is no CPU fallback.
"""
from ._lib import CstError, LIB_PATH, lib          # noqa: F401  (loads the .so)
from .df_table import DfTable, pack_ngrams, unpack_keys  # noqa: F401
from .ciderd import Cider, CiderD, WordIds, language_eval, load_gt_refs  # noqa: F401
from .metrics import Bleu, Rouge                    # noqa: F401
from .reward import CookedRefs, RefSet, RefStore, get_cst_reward, get_self_critical_reward  # noqa: F401
from .pipeline import HostRewardPipeline            # noqa: F401
from .criterion import RewardCriterion              # noqa: F401
