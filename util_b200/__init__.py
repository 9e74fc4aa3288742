"""util_b200 -- B200-native drop-in for util.approximate's distance-based predictors.

    from util_b200 import NearestNeighbor, Shepard, ShepMod, LSHEP, Voronoi

Same fit(x, y) / predict(x) / __call__ contract as the reference
(util/approximate/base.py); every model computes through libb200approx.so
(hand-written CUDA for sm_100a) and fails loudly when it is unavailable.
"""
from .base import (Approximator, WeightedApproximator, ImproperUsage, UnexpectedType,
                   UnexpectedShape, MissingOperator, UnderdefinedModel, class_name)
from .nearest_neighbor import NearestNeighbor, Average, auto
from .shepard import Shepard
from .voronoi import Voronoi, DuplicateInterpolationPoints
from .modified_shepard import ShepMod, TooFewDataPoints
from .linear_shepard import LSHEP
from .wrappers import Split, unique, condition

__all__ = ["Approximator", "WeightedApproximator", "NearestNeighbor", "Average", "auto", "Shepard",
           "Voronoi", "ShepMod", "LSHEP", "Split", "unique", "condition", "TooFewDataPoints", "DuplicateInterpolationPoints",
           "ImproperUsage", "UnexpectedType", "UnexpectedShape", "MissingOperator", "UnderdefinedModel"]
