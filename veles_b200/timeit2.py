"""timeit helper returning (result, seconds) (reference: veles/timeit2.py)."""
import time


def timeit(fn, *args, **kwargs):
    t0 = time.perf_counter()
    res = fn(*args, **kwargs)
    return res, time.perf_counter() - t0
