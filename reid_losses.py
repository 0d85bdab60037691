"""Top-level shim for ``from losses import euclidean_dist, hard_example_mining, WeightedRegularizedTriplet…``
(the distance / hard-mining part of the reference's losses.py)."""
import importlib as _importlib

_impl = _importlib.import_module("multimodal-reid_b200.losses")

euclidean_dist = _impl.euclidean_dist
hard_example_mining = _impl.hard_example_mining
WeightedRegularizedTriplet = _impl.WeightedRegularizedTriplet
WeightedRegularizedTripletXBM = _impl.WeightedRegularizedTripletXBM
