"""Approximator / WeightedApproximator: the fit / predict / __call__ contract of
util/approximate/base.py:25-108,122-173, kept name for name so the B200 models
are a drop-in.  Validation, reshaping and the exception classes are host logic;
every distance, weight and (numeric) weighted sum is computed by libb200approx.
"""
import numpy as np


# util/approximate/base.py:7-11
class ImproperUsage(Exception): pass
class UnexpectedType(Exception): pass
class UnexpectedShape(Exception): pass
class MissingOperator(Exception): pass
class UnderdefinedModel(Exception): pass


def is_numeric(obj):
    # util.math.is_numeric as used by base.py:35,130
    try:
        abs((1. * obj + 1 * obj) ** 1 - (obj / 1. + obj / 1) ** 1.0)
        return True
    except Exception:
        return False


def abs_diff(v1, v2):
    # util/math/__init__.py:47-52 (default metric of nearest_neighbor.auto)
    if hasattr(v1, "__iter__"):
        try: return sum(abs(v1 - v2))
        except Exception: return int(v1 == v2)
    try: return abs(v1 - v2)
    except Exception: return int(v1 == v2)


def class_name(obj):
    # util/approximate/__init__.py:6-8
    if type(obj) == type: return (repr(obj)[1:-2].split(".")[-1])
    else: return (repr(obj)[1:-2].split(".")[-1]).split(" ")[0]


class Approximator:
    """base.py:25-108.  Subclasses implement _fit(x2d, y2d) and _predict(x2d)."""
    _response_dim = 2
    classifier = False

    def __str__(self):
        return class_name(self)

    def fit(self, x, y, classifier=None, *args, **kwargs):
        if classifier is not None: self.classifier = classifier
        elif not is_numeric(y[0]): self.classifier = True
        if (not issubclass(type(x), np.ndarray)) or (len(x.shape) != 2):
            raise UnexpectedType("Provided 'x' should be a 2D numpy array.")
        if self.classifier:
            # base.py:40-46 imports util.data.regular_simplex, which the reference
            # snapshot does not export: regression only here too.
            raise ImproperUsage("classifier mode of Approximator.fit is not available")
        if type(y) == list: y = np.array(y)
        if not issubclass(type(y), np.ndarray):
            raise UnexpectedType("Provided 'y' should be a 1D or 2D numpy array.")
        elif len(y.shape) == 1:
            y = np.reshape(y, (y.shape[0], 1))
            self._response_dim = 1
        elif len(y.shape) == 2:
            pass
        else:
            raise UnexpectedShape("Provided 'y' should be a 1D or 2D numpy array.")
        return self._fit(x.copy(), y.copy(), *args, **kwargs)

    def predict(self, x, *args, **kwargs):
        if (not issubclass(type(x), np.ndarray)) or (0 >= len(x.shape) < 2):
            raise UnexpectedType("Provided 'x' should be a 1D or 2D numpy array.")
        single_response = len(x.shape) == 1
        if single_response:
            x = np.reshape(x, (1, len(x)))
        response = np.asarray(self._predict(x.copy(), *args, **kwargs), dtype=float)
        if (self._response_dim == 1) and (len(response.shape) > 1) and (response.shape[1] == 1):
            response = response[:, 0]
        if single_response:
            response = response[0]
        return response

    def __call__(self, *args, **kwargs):
        return self.predict(*args, **kwargs)


class WeightedApproximator(Approximator):
    """base.py:122-173.  _predict returns (indices, weights) per query.

    When y is a numeric 1-D/2-D array the sums sum(y[i]*w) are evaluated on the
    GPU by the subclass's `_predict_values` (same left-to-right order, separate
    multiply and add roundings as base.py:169); any other y (objects with + and
    *, categories) goes through the reference's own Python loop on the
    (indices, weights) the GPU returned.
    """
    WeightedApproximator = True
    y = None
    _y_numeric = None

    def fit(self, x, y=None, classifier=None, **kwargs):
        if classifier is not None: self.classifier = classifier
        elif (y is not None) and (not is_numeric(y[0])): self.classifier = True
        if (not issubclass(type(x), np.ndarray)) or (len(x.shape) != 2):
            raise UnexpectedType("Provided 'x' should be a 2D numpy array.")
        self._y_numeric = None
        if y is not None:
            if not hasattr(y, "__len__"):
                raise MissingOperator("Provided 'y' to fit must have defined '__len__' operator.")
            elif not hasattr(y, "__getitem__"):
                raise MissingOperator("Provided 'y' to fit must have defined '__getitem__' operator.")
            elif not hasattr(y[0], "__add__"):
                raise MissingOperator("Elements of provided 'y' must have defined '__add__' operator.")
            elif not hasattr(y[0], "__mul__"):
                raise MissingOperator("Elements of provided 'y' must have defined '__mul__' operator.")
            self.y = y
            if (not self.classifier) and isinstance(y, np.ndarray) and y.ndim == 1 \
               and y.dtype == np.float64 and len(y) == len(x):
                self._y_numeric = y
        result = self._fit(x.copy(), **kwargs)
        # numeric responses go to the device once, here, not with every predict call
        model = getattr(self, "model", None)
        if self._y_numeric is not None and model is not None and hasattr(model, "set_values"):
            model.set_values(self._y_numeric)
        return result

    def _predict_values(self, x, *args, **kwargs):
        """GPU weighted sums for numeric y; None when the subclass has no fused path."""
        return None

    def predict(self, x, *args, **kwargs):
        if (not issubclass(type(x), np.ndarray)) or (0 >= len(x.shape) < 2):
            raise UnexpectedType("Provided 'x' should be a 1D or 2D numpy array.")
        single_response = len(x.shape) == 1
        if single_response:
            x = np.reshape(x, (1, len(x)))
        response = None
        if self._y_numeric is not None:
            vals = self._predict_values(x.copy(), *args, **kwargs)
            if vals is not None:
                response = list(vals)  # the reference returns a list of np.float64 (base.py:157-169)
        if response is None:
            indices, weights = self._predict(x.copy(), *args, **kwargs)
            if self.y is None:
                response = [(ids, wts) for (ids, wts) in zip(indices, weights)]
            else:
                response = []
                for ids, wts in zip(indices, weights):
                    if self.classifier:
                        val_weights = {}
                        for i, w in zip(ids, wts):
                            val_weights[self.y[i]] = val_weights.get(self.y[i], 0.) + w
                        response.append(max(val_weights.items(), key=lambda i: i[-1])[0])
                    else:
                        response.append(sum(self.y[i] * w for (i, w) in zip(ids, wts)))
        if single_response: response = response[0]
        return response

    def __call__(self, *args, **kwargs):
        return self.predict(*args, **kwargs)
