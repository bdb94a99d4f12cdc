"""LazyArray — a host-array stand-in whose data lives on the GPU until the host asks for it.

The reference's drivers shuttle every render through numpy (train_model.py:240-243:
render_batch -> numpy -> prepare_model_input -> torch -> .to(device)).  To keep those drivers
unchanged without paying a device->host->device round trip per refinement iteration, the
shadow `render_batch` returns LazyArrays: `prepare_model_input` consumes the device tensor
directly; anything else (np.max, indexing, .astype, plotting) sees an ordinary ndarray,
materialised on first touch."""
import numpy as np


class LazyArray:
    __array_priority__ = 100.0

    def __init__(self, device_tensor):
        self._dev = device_tensor
        self._host = None

    # -- what consumers inside this package use
    @property
    def device_tensor(self):
        return self._dev

    # -- ndarray protocol
    def _materialise(self):
        if self._host is None:
            self._host = self._dev.detach().cpu().numpy()
        return self._host

    def __array__(self, dtype=None, copy=None):
        a = self._materialise()
        return a if dtype is None else a.astype(dtype, copy=False)

    @property
    def shape(self):
        return tuple(self._dev.shape)

    @property
    def ndim(self):
        return self._dev.dim()

    @property
    def dtype(self):
        return np.dtype(str(self._dev.dtype).replace("torch.", ""))

    def __len__(self):
        return self._dev.shape[0]

    def __getitem__(self, idx):
        return self._materialise()[idx]

    def __iter__(self):
        return iter(self._materialise())

    def __getattr__(self, name):   # anything else ndarray offers (astype, max, copy, ...)
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self._materialise(), name)

    def __repr__(self):
        return "LazyArray(shape=%s, dtype=%s, materialised=%s)" % (self.shape, self.dtype, self._host is not None)


def _binop(name):
    def f(self, other):
        return getattr(self._materialise(), name)(np.asarray(other))
    return f


for _n in ("__add__", "__radd__", "__sub__", "__rsub__", "__mul__", "__rmul__", "__truediv__", "__rtruediv__",
           "__gt__", "__lt__", "__ge__", "__le__", "__eq__", "__ne__"):
    setattr(LazyArray, _n, _binop(_n))
LazyArray.__hash__ = None
