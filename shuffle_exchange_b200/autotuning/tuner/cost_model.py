"""Surrogate model of the tuner: predicts the (normalised) metric of a configuration from its feature vector.

The reference uses XGBoost rank regression (``tuner/cost_model.py``); XGBoost is not part of this image, so the model is
a gradient-boosted regression tree ensemble from scikit-learn when available, else ridge regression on the features and
their pairwise products (closed form, numpy only)."""
import numpy as np

try:
    from sklearn.ensemble import GradientBoostingRegressor
except Exception:  # pragma: no cover
    GradientBoostingRegressor = None


class XGBoostCostModel:
    def __init__(self, loss_type="reg", num_threads=None, log_interval=25, upper_model=None):
        self.loss_type = loss_type
        self.model = None
        self._w = None

    @staticmethod
    def _expand(x):
        x = np.asarray(x, dtype=np.float64)
        cross = np.einsum("ni,nj->nij", x, x).reshape(len(x), -1)
        return np.concatenate([np.ones((len(x), 1)), x, cross], axis=1)

    def fit(self, xs, ys):
        x, y = np.asarray(xs, dtype=np.float64), np.asarray(ys, dtype=np.float64)
        y = y / max(np.max(np.abs(y)), 1e-9)
        if GradientBoostingRegressor is not None and len(x) >= 4:
            self.model = GradientBoostingRegressor(n_estimators=60, max_depth=3, learning_rate=0.1, subsample=1.0)
            self.model.fit(x, y)
        else:
            phi = self._expand(x)
            self._w = np.linalg.solve(phi.T @ phi + 1e-3 * np.eye(phi.shape[1]), phi.T @ y)

    def predict(self, xs):
        x = np.asarray(xs, dtype=np.float64)
        if self.model is not None:
            return self.model.predict(x)
        if self._w is None:
            return np.zeros(len(x))
        return self._expand(x) @ self._w
