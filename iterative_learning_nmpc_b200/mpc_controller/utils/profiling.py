"""Wall-clock timing decorator, API of the reference's `utils/profiling.py:5-31`:
`@time_fn(name)` appends milliseconds to `self.timings[name]` when `self.compute_timings`;
`print_timings` reports mean/std/max over all but the first sample, and the first one."""
import time
from functools import wraps
from typing import Dict, List

import numpy as np


def time_fn(method_name):
    def decorator(func):
        @wraps(func)
        def wrapper(self, *args, **kwargs):
            if not getattr(self, "compute_timings", False):
                return func(self, *args, **kwargs)
            t0 = time.time()
            out = func(self, *args, **kwargs)
            self.timings[method_name].append((time.time() - t0) * 1000.0)
            return out
        return wrapper
    return decorator


def print_timings(timings: Dict[str, List[float]]):
    for name, samples in timings.items():
        rest = np.asarray(samples[1:], dtype=float)
        print("---", name, "---")
        if rest.size:
            print(f"mean: {rest.mean():.2f} ms")
            print(f"std: {rest.std():.2f} ms")
            print(f"max: {rest.max():.2f} ms")
        print(f"first: {samples[0]:.2f} ms")
