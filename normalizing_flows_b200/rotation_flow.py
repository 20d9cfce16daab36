"""`OrthogonalProjection2D` — the tutorial bijector of `src/rotation_flow.py:13-37` / `README.md:32-57`.

Same constructor and methods as the reference class; the rotation runs in `rotate2d_kernel`
(csrc/core.cu).  Reference quirk preserved and flagged: `forward_and_log_det` and
`inverse_and_log_det` return the Python constant `1` as the log-determinant, although a rotation
has log|det| = 0.  `logdet="exact"` returns zeros with the batch shape instead, which is what a
`Transformed.log_prob` through this bijector needs to be a density.
"""
from __future__ import annotations

import numpy as np
import torch

from . import distrax_api as distrax
from . import ops


class OrthogonalProjection2D(distrax.Bijector):
    def __init__(self, theta, logdet: str = "reference"):
        super().__init__(event_ndims_in=1, event_ndims_out=1)
        if logdet not in ("reference", "exact"):
            raise ValueError("logdet must be 'reference' (constant 1, rotation_flow.py:31,36) or 'exact' (0)")
        self.thetas = theta
        t = np.float32(float(theta))
        self.sin_theta = float(np.sin(t))
        self.cos_theta = float(np.cos(t))
        # R = [[cos, -sin], [sin, cos]].T  (rotation_flow.py:19-21)
        self.R = torch.tensor([[self.cos_theta, -self.sin_theta], [self.sin_theta, self.cos_theta]]).T.contiguous()
        self._logdet = logdet

    def _ld(self, x):
        if self._logdet == "reference":
            return 1
        return torch.zeros(x.shape[:-1], device=x.device, dtype=torch.float32)

    def forward(self, x):
        return ops.rotate2d(x, self.cos_theta, self.sin_theta, inverse=False)

    def inverse(self, x):
        return ops.rotate2d(x, self.cos_theta, self.sin_theta, inverse=True)

    def forward_and_log_det(self, x):
        return self.forward(x), self._ld(x)

    def inverse_and_log_det(self, x):
        return self.inverse(x), self._ld(x)


__all__ = ["OrthogonalProjection2D"]
