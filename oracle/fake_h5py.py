"""ORACLE-side test double (test infrastructure only): an in-memory stand-in for the tiny part of the h5py API the reference's
lazy-loading writer / reader use (h5py is not installed in this image): File(path, mode) as a context manager, nested groups,
create_group / create_dataset(data=...), keys() / items(), item access returning numpy arrays. Files live in a process-wide
registry keyed by path, so a file written in "a" mode can be re-opened for reading."""
import numpy as np

_REGISTRY = {}


class Group(dict):
    def create_group(self, name):
        g = Group()
        self[name] = g
        return g

    def create_dataset(self, name, data=None, **kw):
        self[name] = np.array(data)
        return self[name]


class File(Group):
    def __new__(cls, path, mode="r", **kw):
        path = str(path)
        if mode in ("r", "r+") and path not in _REGISTRY:
            raise FileNotFoundError(path)
        if mode == "w" or path not in _REGISTRY:
            _REGISTRY[path] = super().__new__(cls)
            dict.__init__(_REGISTRY[path])
        return _REGISTRY[path]

    def __init__(self, path, mode="r", **kw):
        self.filename = str(path)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def close(self):
        pass


def reset():
    _REGISTRY.clear()


def paths():
    return sorted(_REGISTRY)
