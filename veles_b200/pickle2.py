"""Pickle helpers (reference: veles/pickle2.py:46, 66-111): ``best_protocol`` and a debug
mode that locates the exact attribute path that fails to pickle."""
import pickle

best_protocol = pickle.HIGHEST_PROTOCOL
pickle_debug = [False]


def setup_pickle_debug(enabled=True):
    pickle_debug[0] = bool(enabled)


def find_unpicklable(obj, path="obj", seen=None, depth=0, max_depth=12):
    """Return a list of (path, error) for the deepest members that cannot be pickled."""
    seen = seen if seen is not None else set()
    if id(obj) in seen or depth > max_depth:
        return []
    seen.add(id(obj))
    try:
        pickle.dumps(obj, protocol=best_protocol)
        return []
    except Exception as e:  # noqa: BLE001 - we report whatever pickle raised
        err = e
    children = []
    state = None
    if hasattr(obj, "__getstate__"):
        try:
            state = obj.__getstate__()
        except Exception:  # noqa: BLE001
            state = None
    if state is None and hasattr(obj, "__dict__"):
        state = obj.__dict__
    if isinstance(state, dict):
        for k, v in state.items():
            children += find_unpicklable(v, "%s.%s" % (path, k), seen, depth + 1, max_depth)
    elif isinstance(obj, (list, tuple, set, frozenset)):
        for i, v in enumerate(obj):
            children += find_unpicklable(v, "%s[%d]" % (path, i), seen, depth + 1, max_depth)
    elif isinstance(obj, dict):
        for k, v in obj.items():
            children += find_unpicklable(v, "%s[%r]" % (path, k), seen, depth + 1, max_depth)
    return children or [(path, err)]


def dumps(obj, protocol=None):
    try:
        return pickle.dumps(obj, protocol=best_protocol if protocol is None else protocol)
    except Exception:
        if pickle_debug[0]:
            for path, err in find_unpicklable(obj):
                print("unpicklable: %s -> %r" % (path, err))
        raise


loads = pickle.loads
dump = pickle.dump
load = pickle.load
