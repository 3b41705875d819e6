"""Stand-in for h5py used ONLY to run the unmodified reference (baseline/_ref) on this image, which has no
h5py / libhdf5.  Implements the tiny slice of the API the reference's BertH5pyData touches
(``h5py.File(path, 'r', libver=..., swmr=...)`` as a context manager, ``f[key][i]``, ``len(f[key])``) on top of a
NumPy ``.npz`` file with the same six datasets.  It is a dependency stub, not part of either framework."""
import numpy as np

_CACHE = {}


class File:
    def __init__(self, path, mode="r", **kwargs):
        if path not in _CACHE:
            with np.load(path) as z:
                _CACHE[path] = {k: z[k] for k in z.files}
        self._d = _CACHE[path]

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def __getitem__(self, key):
        return self._d[key]

    def keys(self):
        return self._d.keys()

    def flush(self):
        pass

    def close(self):
        pass
