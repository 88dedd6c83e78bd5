"""CPU oracle for the approximator edge of the sampling path.  TEST INFRASTRUCTURE ONLY — see
oracle/__init__.py; nothing under bayesflow_b200/ imports this.

* `OracleStandardize` restates bayesflow/networks/helpers/standardization/standardize.py:11-170 in torch FP32 (the
  reference's own two-pass batch moments + pairwise merge).  Pinned by the reference's tests
  tests/test_networks/test_standardization.py:11-170, restated in tests/test_oracle_edge.py.
* `repeat_and_flatten_conditions` / `unflatten_samples` restate approximators/helpers/samplers.py:51-69.
* `philox4x32_10` / `normal_stream` restate the noise generator of `bfk_normal_fill`.  The reference draws z with
  `keras.random.normal` (distributions/diagonal_normal.py:100-105), whose bits depend on the backend generator, so there
  is nothing to match bit-for-bit on the reference side: parity for the noise is (i) the published Philox4x32-10
  known-answer vectors (Random123 kat_vectors: Salmon et al., SC'11), (ii) this restatement of the Box-Muller map, and
  (iii) the moment checks the reference applies to its own samplers (tests/test_utils/test_integrate.py:96-222 style).
"""
from __future__ import annotations

import numpy as np
import torch

# ---------------------------------------------------------------------------------------------------------------------
_M0, _M1, _W0, _W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85


def philox4x32_10(counter: np.ndarray, key: np.ndarray) -> np.ndarray:
    """counter (n, 4) uint32, key (2,) or (n, 2) uint32 -> (n, 4) uint32.  Ten rounds of the Philox-4x32 bijection."""
    c = np.asarray(counter, dtype=np.uint64).reshape(-1, 4).copy()
    k = np.broadcast_to(np.asarray(key, dtype=np.uint64).reshape(-1, 2), (c.shape[0], 2)).copy()
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(_M0) * c[:, 0]
        p1 = np.uint64(_M1) * c[:, 2]
        n0 = ((p1 >> np.uint64(32)) ^ c[:, 1] ^ k[:, 0]) & mask
        n2 = ((p0 >> np.uint64(32)) ^ c[:, 3] ^ k[:, 1]) & mask
        c = np.stack([n0, p1 & mask, n2, p0 & mask], axis=1)
        k[:, 0] = (k[:, 0] + np.uint64(_W0)) & mask
        k[:, 1] = (k[:, 1] + np.uint64(_W1)) & mask
    return c.astype(np.uint32)


def philox_words(n: int, seed: int, offset: int = 0) -> np.ndarray:
    """raw uint32 words of global elements [offset, offset + n): element e is word e % 4 of counter {e // 4, 0, 0}."""
    q0, q1 = offset // 4, (offset + n - 1) // 4
    q = np.arange(q0, q1 + 1, dtype=np.uint64)
    ctr = np.stack([q & np.uint64(0xFFFFFFFF), q >> np.uint64(32), np.zeros_like(q), np.zeros_like(q)], axis=1)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint64)
    words = philox4x32_10(ctr, key).reshape(-1)
    lo = offset - 4 * q0
    return words[lo:lo + n]


def _u01(r: np.ndarray) -> np.ndarray:
    return ((r >> np.uint32(8)).astype(np.float64) + 0.5) * 2.0 ** -24


def normal_stream(rows: int, D: int, seed: int, row_offset: int = 0, mean=None, std=None) -> np.ndarray:
    """(rows, D) float64 restatement of bfk_normal_fill: Box-Muller on word pairs (0,1) and (2,3) of each counter."""
    n, e0 = rows * D, row_offset * D
    q0, q1 = e0 // 4, (e0 + n - 1) // 4
    words = philox_words(4 * (q1 - q0 + 1), seed, 4 * q0).reshape(-1, 2, 2)
    rad = np.sqrt(-2.0 * np.log(_u01(words[..., 0])))
    ang = 2.0 * np.pi * (_u01(words[..., 1]) - 0.5)  # centred: |angle| <= pi
    z = np.stack([rad * np.cos(ang), rad * np.sin(ang)], axis=-1).reshape(-1)
    z = z[e0 - 4 * q0: e0 - 4 * q0 + n].reshape(rows, D)
    if std is not None:
        z = z * np.asarray(std, dtype=np.float64)
    if mean is not None:
        z = z + np.asarray(mean, dtype=np.float64)
    return z


# ---------------------------------------------------------------------------------------------------------------------
class OracleStandardize:
    """standardize.py:11-170 for a single tensor input (index 0)."""

    def __init__(self, F: int, dtype=torch.float32):
        self.moving_mean = torch.zeros(F, dtype=dtype)   # :43-45
        self.moving_m2 = torch.zeros(F, dtype=dtype)     # :46-48
        self.count = torch.zeros((), dtype=dtype)        # :49

    def moving_std(self):
        # :27-37
        return torch.where(self.moving_m2 > 0, torch.sqrt(self.moving_m2 / self.count), torch.ones_like(self.moving_m2))

    def update_moments(self, x: torch.Tensor):
        # :135-170
        x = x.to(self.moving_mean.dtype)
        axes = tuple(range(x.ndim - 1))
        batch_count = torch.tensor(float(np.prod(x.shape[:-1])), dtype=self.count.dtype)
        batch_mean = x.mean(dim=axes)
        batch_m2 = ((x - batch_mean) ** 2).sum(dim=axes)
        total = self.count + batch_count
        delta = batch_mean - self.moving_mean
        self.moving_mean = self.moving_mean + delta * (batch_count / total)
        self.moving_m2 = self.moving_m2 + batch_m2 + delta ** 2 * (self.count * batch_count / total)
        self.count = total

    def __call__(self, x, stage="inference", forward=True, log_det_jac=False, transformation_type="location_scale"):
        # :50-133
        if (forward or log_det_jac) and transformation_type != "location_scale":
            raise ValueError('Non-default transformation (i.e. transformation_type != "location_scale") '
                             "is not supported for forward or log_det_jac.")
        if stage == "training":
            self.update_moments(x)
        mean, std = self.moving_mean, self.moving_std()
        x = x.to(mean.dtype)
        if forward:
            out = (x - mean) / std
        elif transformation_type == "location_scale":
            out = x * std + mean
        elif transformation_type == "both_sides_scale":
            s = std.reshape((1,) * (x.ndim - 1) + (-1,))
            out = x * s * s.movedim(-1, -2)
        elif transformation_type == "left_side_scale":
            out = x * std.reshape((1,) * (x.ndim - 1) + (-1,)).movedim(-1, -2)
        elif transformation_type == "right_side_scale_inverse":
            out = x / std
        elif transformation_type == "identity":
            out = x
        else:
            raise NotImplementedError(f"No implementation for transformation type {transformation_type!r}")
        if log_det_jac:
            ldj = torch.log(torch.abs(std)).sum(-1).expand(x.shape[:-1])
            return out, (-ldj if forward else ldj)
        return out


def repeat_and_flatten_conditions(conditions, num_samples: int):
    # samplers.py:51-63
    if conditions is None:
        return None
    n, rest = conditions.shape[0], tuple(conditions.shape[1:])
    return conditions.unsqueeze(1).expand(n, num_samples, *rest).reshape(n * num_samples, *rest)


def unflatten_samples(samples, num_samples: int):
    # samplers.py:65-69
    return samples.reshape(-1, num_samples, *samples.shape[1:])
