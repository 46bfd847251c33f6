"""The reference's filter table (mpi/mpi_convolution.c:89-102).  Only ``gaussian`` is live
there (:99); ``box`` and ``edge`` are the commented-out alternatives (:98,100).  ``edge`` is a
blur despite its name (all weights positive, /28)."""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class Filter:
    name: str
    weights: tuple
    divisor: int

    @property
    def exact_integer(self) -> bool:
        """True when out = (sum w*p) >> log2(divisor) bit-for-bit (power-of-two divisor)."""
        return self.divisor & (self.divisor - 1) == 0

    def as_float32(self) -> np.ndarray:
        """The fp32 weights the reference builds: float(w / divisor) (:97-100)."""
        return (np.array(self.weights, dtype=np.float64) / float(self.divisor)).astype(np.float32)


FILTERS = {
    "box": Filter("box", ((1, 1, 1), (1, 1, 1), (1, 1, 1)), 9),
    "gaussian": Filter("gaussian", ((1, 2, 1), (2, 4, 2), (1, 2, 1)), 16),
    "edge": Filter("edge", ((1, 4, 1), (4, 8, 4), (1, 4, 1)), 28),
}
