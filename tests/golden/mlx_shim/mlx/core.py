"""A numpy-backed stand-in for the slice of ``mlx.core`` that the reference's host code uses.

TEST INFRASTRUCTURE ONLY.  It exists so that tests/golden/gen_golden_mlx.py can import and execute the
UNMODIFIED reference modules that need MLX (matching.py, image.py, geometry.py, global_opt.py,
tracker.py, liegroups/, optimizer.py, frame.py) in the build container, where no MLX / Metal exists,
and record their outputs as golden vectors.  Nothing in the product imports it.

Semantics reproduced (what the reference relies on):
  * ``array`` is an ndarray subclass; default dtypes are MLX's: float32 / int32 / bool.  There is no
    float64 or int64 anywhere: python scalars are weakly typed, an integer array meeting a float
    becomes float32 BEFORE the operation (so the arithmetic is float32, as on the device), and any
    64-bit result is narrowed.
  * functional updates ``a.at[idx].add(v)`` (copy, then ``+=`` on the indexed view).
  * ``eval`` is a no-op (numpy is eager).
Not reproduced: MLX's internal reduction order for ``sum`` (numpy's is used) -- for the 3-term sums on
the path this is the one documented unpinned detail.
"""
from __future__ import annotations

import builtins as _b

import numpy as _np

float32 = _np.float32
float16 = _np.float16
int32 = _np.int32
int64 = _np.int64
uint8 = _np.uint8
uint32 = _np.uint32
bool_ = _np.bool_
Dtype = type


def _narrow_dtype(dt):
    if dt == _np.float64:
        return _np.dtype(_np.float32)
    if dt == _np.int64:
        return _np.dtype(_np.int32)
    if dt == _np.uint64:
        return _np.dtype(_np.uint32)
    return None


def _wrap(x):
    """ndarray / scalar -> array with MLX's dtypes."""
    if isinstance(x, tuple):
        return tuple(_wrap(v) for v in x)
    if isinstance(x, list):
        return [_wrap(v) for v in x]
    if not isinstance(x, (_np.ndarray, _np.generic)):
        return x
    a = _np.asarray(x)
    nd = _narrow_dtype(a.dtype)
    if nd is not None:
        a = a.astype(nd)
    return a.view(array)


class _At:
    def __init__(self, arr, idx=None):
        self._arr, self._idx = arr, idx

    def __getitem__(self, idx):
        return _At(self._arr, idx)

    def _apply(self, fn, v):
        out = _np.array(self._arr, copy=True)
        v = _np.asarray(v)
        fn(out, self._idx, v.astype(out.dtype) if v.dtype != out.dtype else v)
        return out.view(array)

    def add(self, v):
        def f(o, i, v):
            _np.add.at(o, i, v) if _needs_at(i) else o.__setitem__(i, o[i] + v)
        return self._apply(f, v)

    def subtract(self, v):
        return self.add(-_np.asarray(v))

    def multiply(self, v):
        def f(o, i, v):
            o[i] = o[i] * v
        return self._apply(f, v)


def _needs_at(idx):
    """integer-array indices may repeat -> accumulate like MLX's scatter-add"""
    idx = idx if isinstance(idx, tuple) else (idx,)
    return _b.any(isinstance(i, _np.ndarray) and i.dtype != _np.bool_ for i in idx)


def _is_float_like(x):
    if isinstance(x, (float, _np.floating)):
        return True
    return isinstance(x, _np.ndarray) and x.dtype.kind == "f"


class array(_np.ndarray):
    __array_priority__ = 100

    def __new__(cls, data=None, dtype=None):
        if isinstance(data, _np.ndarray):
            a = _np.array(data, copy=True)
        else:
            a = _np.array(data)
            if a.dtype == _np.object_:
                a = _np.array([_np.asarray(d) for d in data])
        if dtype is not None:
            a = a.astype(dtype)
        else:
            nd = _narrow_dtype(a.dtype)
            if nd is not None:
                a = a.astype(nd)
        return a.view(cls)

    @property
    def at(self):
        return _At(self)

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kwargs):
        ins = [i.view(_np.ndarray) if isinstance(i, array) else i for i in inputs]
        any_float = _b.any(_is_float_like(i) for i in ins) or ufunc in (_np.true_divide,)
        if any_float:
            # MLX: int (*) float -> float32, computed in float32
            ins = [i.astype(_np.float32) if isinstance(i, _np.ndarray) and i.dtype.kind in "iub"
                   and ufunc not in (_np.logical_and, _np.logical_or, _np.logical_not) else i for i in ins]
        # 0-d / numpy scalars must not drag float32 arrays to float64
        ins = [float(i) if isinstance(i, _np.floating) and i.dtype == _np.float64 else i for i in ins]
        if out is not None:
            kwargs["out"] = tuple(o.view(_np.ndarray) if isinstance(o, array) else o for o in out)
        res = getattr(ufunc, method)(*ins, **kwargs)
        if res is NotImplemented:
            return NotImplemented
        return _wrap(res)

    def __getitem__(self, idx):
        if isinstance(idx, array):
            idx = idx.view(_np.ndarray)
        elif isinstance(idx, tuple):
            idx = tuple(i.view(_np.ndarray) if isinstance(i, array) else i for i in idx)
        r = _np.ndarray.__getitem__(self.view(_np.ndarray), idx)
        return _wrap(r) if isinstance(r, _np.ndarray) else _wrap(_np.asarray(r))

    def astype(self, dtype, *a, **k):
        return _np.ndarray.astype(self.view(_np.ndarray), dtype).view(array)

    def item(self):
        return _np.ndarray.item(self.view(_np.ndarray))

    def tolist(self):
        return _np.ndarray.tolist(self.view(_np.ndarray))

    def squeeze(self, axis=None):
        return _np.squeeze(self.view(_np.ndarray), axis=axis).view(array)

    def __matmul__(self, o):
        return _wrap(_np.matmul(self.view(_np.ndarray), _np.asarray(o)))

    def __rmatmul__(self, o):
        return _wrap(_np.matmul(_np.asarray(o), self.view(_np.ndarray)))

    def sum(self, axis=None, keepdims=False):
        return sum_(self, axis=axis, keepdims=keepdims)


def _n(x):
    return x.view(_np.ndarray) if isinstance(x, array) else x


def _fn(npf):
    def f(*args, **kw):
        return _wrap(npf(*[_n(a) for a in args], **{k: _n(v) for k, v in kw.items()}))
    f.__name__ = npf.__name__
    return f


def eval(*args, **kw):  # noqa: A001
    return None


def zeros(shape, dtype=float32):
    return _np.zeros(shape, dtype).view(array)


def ones(shape, dtype=float32):
    return _np.ones(shape, dtype).view(array)


def zeros_like(a):
    return _np.zeros_like(_n(a)).view(array)


def ones_like(a):
    return _np.ones_like(_n(a)).view(array)


def eye(n, dtype=float32):
    return _np.eye(n, dtype=dtype).view(array)


def arange(*a, dtype=None):
    r = _np.arange(*a)
    if dtype is not None:
        r = r.astype(dtype)
    return _wrap(r)


def linspace(a, b, n, dtype=float32):
    return _np.linspace(a, b, n).astype(dtype).view(array)


def concatenate(arrs, axis=0):
    arrs = [_np.asarray(_n(a)) for a in arrs]
    # empty float32 placeholders ([]) adopt the other operands' shape, as MLX does for 0-size inputs
    nz = [a for a in arrs if a.size]
    if nz and len(nz) != len(arrs):
        dt = _np.result_type(*nz)
        arrs = nz if _b.all(a.ndim == nz[0].ndim for a in nz) else arrs
        return _wrap(_np.concatenate(arrs, axis=axis).astype(dt))
    return _wrap(_np.concatenate(arrs, axis=axis))


def stack(arrs, axis=0):
    return _wrap(_np.stack([_n(a) for a in arrs], axis=axis))


def sum_(a, axis=None, keepdims=False):
    a = _np.asarray(_n(a))
    if a.dtype == _np.bool_:
        a = a.astype(_np.int32)
    return _wrap(_np.sum(a, axis=axis, keepdims=keepdims))


sum = sum_  # noqa: A001


def mean(a, axis=None, keepdims=False):
    return _wrap(_np.mean(_n(a), axis=axis, keepdims=keepdims, dtype=_np.float32))


def max(a, axis=None, keepdims=False):  # noqa: A001
    return _wrap(_np.max(_n(a), axis=axis, keepdims=keepdims))


def min(a, axis=None, keepdims=False):  # noqa: A001
    return _wrap(_np.min(_n(a), axis=axis, keepdims=keepdims))


def prod(a, axis=None, keepdims=False):
    return _wrap(_np.prod(_n(a), axis=axis, keepdims=keepdims))


def any(a, axis=None, keepdims=False):  # noqa: A001
    return _wrap(_np.any(_n(a), axis=axis, keepdims=keepdims))


def all(a, axis=None, keepdims=False):  # noqa: A001
    return _wrap(_np.all(_n(a), axis=axis, keepdims=keepdims))


def median(a, axis=None):
    return _wrap(_np.median(_n(a), axis=axis).astype(_np.float32))


def clip(a, lo, hi):
    a = _np.asarray(_n(a))
    if a.dtype.kind == "f":
        lo = None if lo is None else _np.float32(lo)
        hi = None if hi is None else _np.float32(hi)
    return _wrap(_np.clip(a, lo, hi))


def where(c, a, b):
    a, b = _n(a), _n(b)
    if isinstance(a, (int, float)) and isinstance(b, _np.ndarray):
        a = _np.asarray(a, b.dtype)
    if isinstance(b, (int, float)) and isinstance(a, _np.ndarray):
        b = _np.asarray(b, a.dtype)
    return _wrap(_np.where(_n(c), a, b))


def broadcast_to(a, shape):
    return _np.array(_np.broadcast_to(_n(a), shape)).view(array)


def transpose(a, axes=None):
    return _np.transpose(_n(a), axes).view(array)


def swapaxes(a, i, j):
    return _np.swapaxes(_n(a), i, j).view(array)


def meshgrid(*a, indexing="xy"):
    return [_wrap(g) for g in _np.meshgrid(*[_n(x) for x in a], indexing=indexing)]


def unique(a):
    return _wrap(_np.unique(_n(a)))


def einsum(spec, *ops):
    return _wrap(_np.einsum(spec, *[_n(o) for o in ops]))


def _unary(npf):
    def f(a):
        a = _np.asarray(_n(a))
        if a.dtype.kind in "iub":
            a = a.astype(_np.float32)
        return _wrap(npf(a))
    f.__name__ = npf.__name__
    return f


sqrt, sin, cos, exp, log, arccos, arctan, floor, abs = (  # noqa: A001
    _unary(_np.sqrt), _unary(_np.sin), _unary(_np.cos), _unary(_np.exp), _unary(_np.log),
    _unary(_np.arccos), _unary(_np.arctan), _unary(_np.floor), _fn(_np.abs))
minimum, maximum, arctan2 = _fn(_np.minimum), _fn(_np.maximum), _fn(_np.arctan2)


class _Linalg:
    @staticmethod
    def cholesky(a):
        return _wrap(_np.linalg.cholesky(_n(a)))

    @staticmethod
    def solve_triangular(a, b, upper=False):
        import scipy.linalg as sl
        return _wrap(sl.solve_triangular(_n(a), _n(b), lower=not upper).astype(_np.float32))

    @staticmethod
    def pinv(a):
        return _wrap(_np.linalg.pinv(_n(a)))


linalg = _Linalg()
