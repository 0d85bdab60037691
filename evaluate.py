"""Top-level shim so that the reference scripts' ``from evaluate import R1_mAP_eval``
(zero_shot_learning.py:10, prompt_learning.py:19, ...) resolves to the B200 implementation when
this repository precedes the reference on ``sys.path``."""
import importlib as _importlib

_impl = _importlib.import_module("multimodal-reid_b200.evaluate")

euclidean_distance = _impl.euclidean_distance
cosine_similarity = _impl.cosine_similarity
eval_func = _impl.eval_func
R1_mAP_eval = _impl.R1_mAP_eval
re_ranking = _impl.re_ranking
