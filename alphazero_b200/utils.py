"""Small helpers shared by the drop-in layer (alphazero/utils.py:22-34)."""
import numpy as np

from .base import dotdict  # noqa: F401  (re-exported like the reference's utils.dotdict)


def fair_max(elements, key=lambda x: x):
    """arg-max with a uniform random tie-break drawn from NumPy's global RNG (utils.py:28-34); a unique maximum
    consumes no random number."""
    elements = list(elements)
    best = key(max(elements, key=key))
    ties = [x for x in elements if key(x) == best]
    return ties[np.random.choice(len(ties))]
