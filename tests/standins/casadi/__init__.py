"""A STAND-IN for the parts of casadi's Python API that cocofest_b200.bioptim_seam touches -- TEST INFRASTRUCTURE ONLY.

casadi is not installable in this image, so the `casadi.Callback` pair of `as_casadi_function` could never run.  This
module reproduces the CONTRACT casadi imposes on a Callback subclass (casadi/core/callback.hpp, function.hpp), and is
strict about it, so that the wrappers are executed and their declarations checked:

  * `construct(name, opts)` finalises the object: it calls `init()` and then reads `get_n_in / get_n_out`, the names and
    the sparsities ONCE; after that the Function is immutable;
  * a call converts every argument to a `DM` of the declared input shape (a mismatch raises, as casadi does), calls
    `eval(list_of_DM)` and requires as many results as declared outputs, each of the declared shape and -- casadi
    projects results onto the declared sparsity -- with no value outside a structurally-zero block;
  * `jacobian()` asks `has_jacobian()` and `get_jacobian(name, inames, onames, opts)` with casadi's naming:
    name = "jac_<f>", inames = inputs + "out_<o>", onames = "jac_<o>_<i>" with OUTPUTS outermost; the returned Function
    must have n_in + n_out inputs and n_in * n_out outputs whose shapes are (numel(out_o), numel(in_i));
  * `map(n)` evaluates n horizontally concatenated columns.

It does no symbolic algebra and no AD: nothing here can stand in for casadi at run time.
"""
import numpy as np

__version__ = "0.0-standin"


class Sparsity:
    """Shape + pattern; only 'dense' and 'empty' patterns exist here (all that the seam declares)."""

    def __init__(self, rows=0, cols=0, _dense=False):
        self._r, self._c, self._dense = int(rows), int(cols), bool(_dense)

    @staticmethod
    def dense(rows, cols=1):
        return Sparsity(rows, cols, _dense=True)

    def size1(self):
        return self._r

    def size2(self):
        return self._c

    def shape(self):
        return (self._r, self._c)

    def numel(self):
        return self._r * self._c

    def nnz(self):
        return self._r * self._c if self._dense else 0

    def is_dense(self):
        return self._dense or self.numel() == 0

    def __repr__(self):
        return f"Sparsity({self._r}x{self._c}, {'dense' if self._dense else 'empty'})"


class DM:
    """Dense numeric matrix (column vectors are [n, 1])."""

    def __init__(self, x=None, cols=None):
        if isinstance(x, Sparsity):
            self._a = np.zeros(x.shape())
            self._sp = x
            return
        if isinstance(x, DM):
            x = x._a
        if cols is not None:
            self._a = np.zeros((int(x), int(cols)))
        elif x is None:
            self._a = np.zeros((0, 0))
        else:
            a = np.asarray(x, dtype=np.float64)
            if a.ndim == 0:
                a = a.reshape(1, 1)
            elif a.ndim == 1:
                a = a.reshape(-1, 1)      # casadi: a 1-D sequence becomes a column
            elif a.ndim != 2:
                raise TypeError("DM holds matrices")
            self._a = a.copy()
        self._sp = Sparsity.dense(*self._a.shape)

    @property
    def shape(self):
        return self._a.shape

    def sparsity(self):
        return self._sp

    def nnz(self):
        return self._sp.nnz()

    def full(self):
        return self._a.copy()

    def __array__(self, dtype=None, copy=None):
        return self._a.astype(dtype) if dtype is not None else self._a.copy()

    def __repr__(self):
        return f"DM({self._a!r})"


def _as_dm(x, sp, what):
    d = x if isinstance(x, DM) else DM(np.zeros((0, 1)) if (x is None or (hasattr(x, "__len__") and len(x) == 0)) else x)
    if d.shape != sp.shape():
        if d._a.size == sp.numel() and 1 in (d.shape + sp.shape()):
            d = DM(d._a.reshape(sp.shape()))          # casadi accepts a row for a column of the same length
        else:
            raise RuntimeError(f"{what}: dimension mismatch, got {d.shape}, declared {sp.shape()}")
    return d


class Function:
    def name(self):
        return self._name

    def n_in(self):
        return self._n_in

    def n_out(self):
        return self._n_out

    def name_in(self, i=None):
        return list(self._names_in) if i is None else self._names_in[i]

    def name_out(self, i=None):
        return list(self._names_out) if i is None else self._names_out[i]

    def sparsity_in(self, i):
        return self._sp_in[i]

    def sparsity_out(self, i):
        return self._sp_out[i]

    def size_in(self, i):
        return self._sp_in[i].shape()

    def size_out(self, i):
        return self._sp_out[i].shape()


class Callback(Function):
    def __init__(self):
        self._constructed = False

    # defaults of casadi::Callback
    def init(self):
        pass

    def get_n_in(self):
        return 1

    def get_n_out(self):
        return 1

    def get_name_in(self, i):
        return f"i{i}"

    def get_name_out(self, i):
        return f"o{i}"

    def get_sparsity_in(self, i):
        return Sparsity.dense(1, 1)

    def get_sparsity_out(self, i):
        return Sparsity.dense(1, 1)

    def has_jacobian(self):
        return False

    def get_jacobian(self, name, inames, onames, opts):
        raise RuntimeError("get_jacobian not defined")

    def eval(self, arg):
        raise RuntimeError("eval not defined")

    def construct(self, name, opts=None):
        if self._constructed:
            raise RuntimeError("construct called twice")
        if not isinstance(name, str) or not name.replace("_", "a").isalnum() or name[0].isdigit():
            raise RuntimeError(f"invalid function name {name!r}")     # casadi rejects names that are not identifiers
        self._name, self._opts = name, dict(opts or {})
        self.init()
        self._n_in, self._n_out = int(self.get_n_in()), int(self.get_n_out())
        self._names_in = [self.get_name_in(i) for i in range(self._n_in)]
        self._names_out = [self.get_name_out(i) for i in range(self._n_out)]
        self._sp_in = [self.get_sparsity_in(i) for i in range(self._n_in)]
        self._sp_out = [self.get_sparsity_out(i) for i in range(self._n_out)]
        for sp in self._sp_in + self._sp_out:
            if not isinstance(sp, Sparsity):
                raise RuntimeError("get_sparsity_* must return a Sparsity")
        self._constructed = True
        self._jac_fun = None

    def __call__(self, *args, **kwargs):
        if not self._constructed:
            raise RuntimeError("Callback used before construct()")
        if kwargs:
            args = [kwargs.get(n) for n in self._names_in]
        if len(args) != self._n_in:
            raise RuntimeError(f"{self._name}: {len(args)} arguments for {self._n_in} inputs")
        arg = [_as_dm(a, sp, f"{self._name} input {i} ({self._names_in[i]})") for i, (a, sp) in enumerate(zip(args, self._sp_in))]
        res = self.eval(arg)
        if not isinstance(res, (list, tuple)) or len(res) != self._n_out:
            raise RuntimeError(f"{self._name}.eval returned {len(res) if hasattr(res, '__len__') else res} results for {self._n_out} outputs")
        out = []
        for i, (r, sp) in enumerate(zip(res, self._sp_out)):
            d = _as_dm(r, sp, f"{self._name} output {i} ({self._names_out[i]})")
            if sp.nnz() == 0 and sp.numel() and np.any(d._a != 0):
                raise RuntimeError(f"{self._name} output {i} is declared structurally zero but eval returned non-zeros")
            out.append(d)
        return out[0] if self._n_out == 1 else tuple(out)

    def jacobian(self):
        if not self.has_jacobian():
            raise RuntimeError(f"{self._name}: no Jacobian (casadi would fall back to finite differences)")
        if self._jac_fun is None:
            inames = list(self._names_in) + [f"out_{n}" for n in self._names_out]
            onames = [f"jac_{o}_{i}" for o in self._names_out for i in self._names_in]
            J = self.get_jacobian(f"jac_{self._name}", inames, onames, {})
            if J.n_in() != self._n_in + self._n_out or J.n_out() != self._n_in * self._n_out:
                raise RuntimeError(f"jac_{self._name}: expected {self._n_in + self._n_out} inputs / {self._n_in * self._n_out} outputs, "
                                   f"got {J.n_in()} / {J.n_out()}")
            for k in range(self._n_in + self._n_out):
                want = (self._sp_in + self._sp_out)[k].shape()
                if J.size_in(k) != want:
                    raise RuntimeError(f"jac_{self._name} input {k}: shape {J.size_in(k)}, expected {want}")
            for o in range(self._n_out):
                for i in range(self._n_in):
                    want = (self._sp_out[o].numel(), self._sp_in[i].numel())
                    if J.size_out(o * self._n_in + i) != want:
                        raise RuntimeError(f"jac_{self._name} output {onames[o * self._n_in + i]}: shape "
                                           f"{J.size_out(o * self._n_in + i)}, expected {want}")
            J._names_out_expected = onames
            self._jac_fun = J
        return self._jac_fun

    def map(self, n, *parallelization):
        """n evaluations over horizontally concatenated arguments (Function.map)."""
        f = self

        def mapped(*args):
            cols = []
            for k in range(n):
                sub = []
                for a, sp in zip(args, f._sp_in):
                    A = np.asarray(a, dtype=np.float64).reshape(sp.size1(), -1) if sp.numel() else np.zeros((0, 1))
                    w = sp.size2()
                    sub.append(A[:, k * w:(k + 1) * w] if (sp.numel() and A.shape[1] == n * w) else A)   # one column set = shared by all
                r = f(*sub)
                cols.append(r if isinstance(r, tuple) else (r,))
            return tuple(DM(np.concatenate([c[o]._a for c in cols], axis=1)) for o in range(f._n_out))

        return mapped
