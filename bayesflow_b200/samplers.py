"""`Sampler` — batched, repeated sampling from an inference network (approximators/helpers/samplers.py:13-155) without
materialising the repeated conditions.

The reference broadcasts the (N, ...) conditions to (N * num_samples, ...) before `inference_network.sample`
(samplers.py:51-63): for BASELINE configs[1] that is a 128 MB tensor whose rows are 1000-fold copies.  Here the (N, C)
conditions go to the kernels as they are, with `conditions_repeats=num_samples` (row r reads dataset r // num_samples),
z is generated in HBM (`bfk_normal_fill`) and, if a `Standardize` layer is given, the draws are mapped back to the data
scale in place (`bfk_standardize`, forward=False — continuous_approximator.py's inverse standardization)."""
from __future__ import annotations

from typing import Sequence

import torch


class Sampler:
    def infer_sample_shape(self, conditions, sample_shape):
        # samplers.py:30-49
        if sample_shape == "infer":
            if conditions is None:
                return ()
            return tuple(conditions.shape[1:-1])
        if isinstance(sample_shape, int):
            return (sample_shape,)
        if isinstance(sample_shape, (tuple, list)):
            return tuple(sample_shape)
        raise ValueError(f"sample_shape must be 'infer', an int, or a tuple/list of ints, but got {type(sample_shape)}.")

    def repeat_and_flatten_conditions(self, conditions, num_samples: int):
        """samplers.py:51-63 — kept for callers that want the materialised tensor; `sample` does not use it."""
        if conditions is None:
            return None
        n = conditions.shape[0]
        rest = tuple(conditions.shape[1:])
        return conditions.unsqueeze(1).expand(n, num_samples, *rest).reshape(n * num_samples, *rest)

    def unflatten_samples(self, samples, num_samples: int):
        # samplers.py:65-69
        return samples.reshape(-1, num_samples, *samples.shape[1:])

    def sample(self, inference_network, num_samples: int, conditions=None, batch_size: int | None = None,
               sample_shape="infer", seed=None, standardize=None, **kwargs):
        """-> (N, num_samples, D) draws (or (num_samples, D) without conditions); samplers.py:71-148."""
        if conditions is None:
            out = inference_network.sample((num_samples,) + self.infer_sample_shape(None, sample_shape), conditions=None,
                                           seed=seed, **kwargs)
            return self._unstandardize(out, standardize)
        if conditions.ndim != 2:
            raise NotImplementedError("structured (set- or sequence-valued) conditions are outside the hot path "
                                      "(SURVEY.md §8): expected conditions of shape (num_datasets, C)")
        if self.infer_sample_shape(conditions, sample_shape) != ():
            raise NotImplementedError("structural sample dimensions are outside the hot path (SURVEY.md §8)")
        n = conditions.shape[0]
        if batch_size is None:
            batch_size = n
        batches = []
        for lo in range(0, n, batch_size):
            hi = min(n, lo + batch_size)
            # tensor-valued kwargs with a batch axis follow the conditions slice (samplers.py:100-103, 126-133)
            kw = {k: (v[lo:hi] if torch.is_tensor(v) and v.ndim >= 2 else v) for k, v in kwargs.items()}
            for k, v in kw.items():
                if torch.is_tensor(v) and v.ndim >= 2:
                    kw[k] = self.repeat_and_flatten_conditions(v, num_samples)
            flat = inference_network.sample(((hi - lo) * num_samples,), conditions=conditions[lo:hi].contiguous(),
                                            conditions_repeats=num_samples, seed=seed, row_offset=lo * num_samples, **kw)
            batches.append(self.unflatten_samples(flat, num_samples))
        out = batches[0] if len(batches) == 1 else torch.cat(batches, dim=0)
        return self._unstandardize(out, standardize)

    @staticmethod
    def _unstandardize(samples, standardize):
        if standardize is None:
            return samples
        return standardize._location_scale(samples, 0, forward=False, out=samples if samples.is_contiguous() else None)

    def get_config(self) -> dict:
        return {}
