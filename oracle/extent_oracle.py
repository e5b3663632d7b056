"""TEST INFRASTRUCTURE ONLY -- CPU restatement (oracle) of the extent parser and scaler.

Only ``tests/`` may import this file.  Pin: ``tests/test_oracle_golden.py`` checks it against
``tests/golden/extent_text.npz`` -- outputs of the reference's own ``state_to_records``
(scripts/pre_experiments/model_myself/utils/extent.py:94-136, imported unmodified by oracle/ref_import.py) and of
sklearn's ``StandardScaler`` (the third-party routine the reference calls at vae_extent_search.py:394-397;
unpinned there, scikit-learn 1.9.0 in this image), plus the printed state of the reference README (:86-108) as
a known-answer vector.
"""
import re

import numpy as np


def state_to_extents(state_strs):
    """extent.py:101-131: per state the `(0,n)` extents followed by the auto_unroll values ([0.0] if none)."""
    rows = []
    for state in state_strs:
        schedule = state.strip().split("Placeholder")[-1][2:]
        ext = [float(v) for v in re.findall(r"\(0,\s*(\d+)\)", schedule)]
        unr = [float(v) for v in re.findall(r"auto_unroll:\s*(\d+)", schedule)] or [0.0]
        rows.append(np.array(ext + unr, dtype=np.float32))
    return rows


def scaler_fit(raw):
    """vae_extent_search.py:394-397: log1p then StandardScaler.fit.  Two-pass fp64 statistics as in sklearn's
    _incremental_mean_and_var (first batch) with _is_constant_feature / _handle_zeros_in_scale."""
    x = np.log1p(np.asarray(raw, dtype=np.float32))
    n = x.shape[0]
    s = x.sum(axis=0, dtype=np.float64)
    mean = s / n
    t = x.astype(np.float64) - mean
    corr = t.sum(axis=0)
    var = ((t * t).sum(axis=0) - corr * corr / n) / n
    eps = np.finfo(np.float64).eps
    const = var <= n * eps * var + (n * mean * eps) ** 2
    scale = np.sqrt(var)
    scale[const | (scale == 0.0)] = 1.0
    return mean, var, scale


def scaler_transform(raw, mean, scale):
    """StandardScaler.transform on the fp32 array: `X -= mean_.astype(X.dtype); X /= scale_.astype(X.dtype)`
    (sklearn/preprocessing/_data.py, scikit-learn 1.9.0: the statistics are cast to X's dtype first)."""
    x = np.log1p(np.asarray(raw, dtype=np.float32))
    x -= np.asarray(mean).astype(np.float32)
    x /= np.asarray(scale).astype(np.float32)
    return x


def scaler_inverse(x_scaled, mean, scale):
    """expm1(x*scale_+mean_), rounded, clamped at 1 (the inverse of the cell above; SURVEY §8c iii)."""
    y = np.expm1(np.asarray(x_scaled, dtype=np.float64) * scale + mean)
    return np.maximum(np.rint(y), 1).astype(np.int64)
