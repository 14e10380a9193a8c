"""Value-type small vectors / matrices of IEEE binary32 elements for the taichi shim."""
from __future__ import annotations

import numpy as np


def _flatten_to_array(values):
    """Nested lists / Matrix / scalars -> float32 ndarray of 1 or 2 dimensions."""
    if isinstance(values, Matrix):
        return values.a.copy()
    if isinstance(values, np.ndarray):
        return values.astype(np.float32)
    rows = []
    for v in values:
        if isinstance(v, Matrix):
            rows.append(v.a)
        elif isinstance(v, (list, tuple)):
            rows.append(np.array([np.float32(x) for x in v], np.float32))
        else:
            rows.append(np.float32(v))
    return np.array(rows, np.float32)


class Matrix:
    __slots__ = ("a",)
    __array_priority__ = 1000  # numpy scalars defer to our reflected operators

    def __init__(self, values):
        self.a = _flatten_to_array(values)

    # -- structure ------------------------------------------------------------------------------
    @property
    def n(self):
        return self.a.shape[0]

    @property
    def m(self):
        return self.a.shape[1] if self.a.ndim == 2 else 1

    def copy(self):
        out = Matrix.__new__(Matrix)
        out.a = self.a.copy()
        return out

    @staticmethod
    def _wrap(a):
        out = Matrix.__new__(Matrix)
        out.a = np.asarray(a, np.float32)
        return out

    def map(self, fn):
        flat = [fn(v) for v in self.a.reshape(-1)]
        return Matrix._wrap(np.array(flat, np.float32).reshape(self.a.shape))

    def __getitem__(self, idx):
        return self.a[idx]

    def __setitem__(self, idx, v):
        self.a[idx] = np.float32(v)

    def __iter__(self):
        return iter(self.a)

    def __len__(self):
        return self.a.shape[0]

    def __repr__(self):
        return f"Matrix({self.a.tolist()})"

    x = property(lambda s: s.a[0], lambda s, v: s.a.__setitem__(0, np.float32(v)))
    y = property(lambda s: s.a[1], lambda s, v: s.a.__setitem__(1, np.float32(v)))
    z = property(lambda s: s.a[2], lambda s, v: s.a.__setitem__(2, np.float32(v)))
    w = property(lambda s: s.a[3], lambda s, v: s.a.__setitem__(3, np.float32(v)))

    # -- element-wise arithmetic (one binary32 rounding per element per operator) ---------------
    @staticmethod
    def _operand(o):
        if isinstance(o, Matrix):
            return o.a
        return np.float32(o)

    def __add__(self, o): return Matrix._wrap(self.a + Matrix._operand(o))
    def __radd__(self, o): return Matrix._wrap(Matrix._operand(o) + self.a)
    def __sub__(self, o): return Matrix._wrap(self.a - Matrix._operand(o))
    def __rsub__(self, o): return Matrix._wrap(Matrix._operand(o) - self.a)
    def __mul__(self, o): return Matrix._wrap(self.a * Matrix._operand(o))
    def __rmul__(self, o): return Matrix._wrap(Matrix._operand(o) * self.a)
    def __truediv__(self, o): return Matrix._wrap(self.a / Matrix._operand(o))
    def __rtruediv__(self, o): return Matrix._wrap(Matrix._operand(o) / self.a)
    def __neg__(self): return Matrix._wrap(-self.a)
    def __pos__(self): return self.copy()

    # -- Taichi matrix_ops --------------------------------------------------------------------------
    def __matmul__(self, o):
        a = self.a
        b = o.a if isinstance(o, Matrix) else np.asarray(o, np.float32)
        vec = b.ndim == 1
        if vec:
            b = b.reshape(-1, 1)
        row = a.ndim == 1
        if row:
            a = a.reshape(1, -1)
        assert a.shape[1] == b.shape[0], (a.shape, b.shape)
        acc = a[:, 0:1] * b[0:1, :]
        for k in range(1, a.shape[1]):
            acc = acc + a[:, k:k + 1] * b[k:k + 1, :]
        if row and vec:
            return acc[0, 0]       # vector @ vector: the dot product
        if vec:
            return Matrix._wrap(acc[:, 0])
        if row:
            return Matrix._wrap(acc[0, :])
        return Matrix._wrap(acc)

    def outer_product(self, o):
        return Matrix._wrap(self.a.reshape(-1, 1) * o.a.reshape(1, -1))

    def transpose(self):
        return Matrix._wrap(self.a.T.copy())

    def sum(self):
        flat = self.a.reshape(-1)
        acc = flat[0]
        for v in flat[1:]:
            acc = acc + v
        return acc

    def dot(self, o):
        return (self * o).sum()

    def norm_sqr(self):
        return (self * self).sum()

    def norm(self, eps=0.0):
        return np.sqrt(np.float32(self.norm_sqr() + np.float32(eps)))

    def normalized(self, eps=0.0):
        invlen = np.float32(1.0) / (self.norm() + np.float32(eps))
        return invlen * self

    def trace(self):
        acc = self.a[0, 0]
        for i in range(1, self.a.shape[0]):
            acc = acc + self.a[i, i]
        return acc

    def determinant(self):
        a = self.a
        if a.shape == (2, 2):
            return a[0, 0] * a[1, 1] - a[0, 1] * a[1, 0]
        if a.shape == (3, 3):
            return (a[0, 0] * (a[1, 1] * a[2, 2] - a[2, 1] * a[1, 2])
                    - a[1, 0] * (a[0, 1] * a[2, 2] - a[2, 1] * a[0, 2])
                    + a[2, 0] * (a[0, 1] * a[1, 2] - a[1, 1] * a[0, 2]))
        raise NotImplementedError(a.shape)

    def inverse(self):
        a = self.a
        if a.shape == (2, 2):
            inv_det = np.float32(1.0) / self.determinant()
            return inv_det * Matrix([[a[1, 1], -a[0, 1]], [-a[1, 0], a[0, 0]]])
        raise NotImplementedError(a.shape)


def make_vector_type(n):
    def ctor(*args):
        if len(args) == 1 and isinstance(args[0], (list, tuple, Matrix, np.ndarray)):
            vals = args[0]
        else:
            vals = list(args)

        def flat(seq):  # vec4(vec3, 1) / vec4([a, b, c], d) style concatenation
            out = []
            for x in seq:
                if isinstance(x, Matrix):
                    out.extend(list(x.a.reshape(-1)))
                elif isinstance(x, (list, tuple, np.ndarray)):
                    out.extend(flat(x))
                else:
                    out.append(x)
            return out

        m = Matrix(flat(vals))
        if m.a.shape == () or m.a.size == 1 and n > 1:
            m = Matrix([np.float32(m.a.reshape(-1)[0])] * n)
        assert m.a.shape == (n,), (m.a.shape, n)
        return m
    ctor.n = n
    return ctor


def make_matrix_type(n, m):
    def ctor(*args):
        vals = args[0] if len(args) == 1 else list(args)
        out = Matrix(vals)
        if out.a.ndim == 1 and out.a.size == n * m:
            out = Matrix._wrap(out.a.reshape(n, m))
        assert out.a.shape == (n, m), (out.a.shape, (n, m))
        return out
    ctor.shape = (n, m)
    return ctor
