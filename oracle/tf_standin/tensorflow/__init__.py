"""TEST INFRASTRUCTURE ONLY (oracle). Not imported by the product package.

A torch-CPU stand-in for the ~60 ``tensorflow`` symbols the reference's head path calls, so that
``/root/reference/yolov3/{yolov3_decoder,label_decoder,yolov3_loss}.py`` and the label half of
``/root/reference/dataset/file_util.py`` can be IMPORTED AND RUN UNMODIFIED in the build container
(TensorFlow itself is absent and cannot be installed).  ``tests/golden/make_loss_golden.py`` puts this directory and
``/root/reference`` on ``sys.path`` and writes the fixtures that pin ``oracle/ref_numpy.py`` / ``ref_torch.py``
and the CUDA kernels.  torch autograd plays the role of ``tf.gradients``.

Every mapping is one op onto one torch op; the TF semantics each one assumes (TF 1.13 python/ops sources:
``math_ops.py``, ``clip_ops.py``, ``array_ops.py``, ``math_grad.py``, ``nn_grad.py``; CPU functors) is written next to it.
What remains an ASSUMPTION (cannot be executed here) is listed once:

 A1 graph-mode reads happen before writes inside one ``session.run``: both ``tf.cond`` predicates of
    yolov3_loss.py:125,131 see ``current_num`` before the in-branch ``update_add`` (:152).  The stand-in queues
    ``update_add`` and applies it in ``end_step()``.
 A2 Python scalars / lists / NumPy arrays met by an op are materialised as constants of the *tensor's* dtype
    (``ops.convert_to_tensor(v, dtype=x.dtype)``): rounded once to fp32 in the fp32 graph.
 A3 ``tf.sigmoid`` = 1 / (1 + exp(-x)) (Eigen ``scalar_logistic_op`` generic form; gradient ``dy*y*(1-y)``),
    ``tf.nn.softmax`` = exp(x - max x) * (1 / sum) (``softmax_op_functor.h``; gradient
    ``(dy - sum(dy*y)) * y``).  exp/log themselves are the platform's (torch CPU here, libdevice on the GPU):
    <= 1-2 ulp apart, which is what the 1e-5 relative tolerance of the value tests absorbs.
 A4 ``clip_by_value`` is the composite ``maximum(minimum(t, hi), lo)``; Minimum/Maximum pass the gradient to ``t``
    at equality (``math_grad._MaximumMinimumGrad``: ``x >= y`` / ``x <= y``).
 A5 ``reduce_max`` over an empty axis is -inf; its gradient is split equally among ties.
 A6 ``gather_nd`` gradient is ``scatter_nd`` (duplicates add); ``sparse_to_dense(validate_indices=False)`` accepts
    duplicate / unsorted indices; ``where(cond)`` lists indices in row-major order; ``arg_max`` returns the first
    maximum; float -> int32 ``cast`` truncates.

Precision twin: ``standin.set_precision('float64')`` widens ``tf.float32`` to fp64 for the ground-truth run.  Constants
keep their fp32 values (rounded to fp32 first, then widened), so the twin is "the same fp32 graph constants and inputs,
exact arithmetic".
"""
import builtins
import contextlib
import types

import numpy as np
import torch

from . import keras  # noqa: F401  (``from tensorflow import keras``)

__version__ = "1.13.1-standin"


# ------------------------------------------------------------------------------------------------ dtypes
class DType:
    def __init__(self, name, floating):
        self.name, self.is_floating = name, floating

    @property
    def torch(self):
        if self.name == "float32":
            return _STATE["float"]
        return getattr(torch, self.name)

    def __repr__(self):
        return "tf." + self.name


float32 = DType("float32", True)
float64 = DType("float64", True)
int32 = DType("int32", False)
int64 = DType("int64", False)
bool_ = DType("bool", False)
string = DType("string", False)

_STATE = {"float": torch.float32, "pending": [], "taps": []}


def _as_dtype(dt):
    if dt is None:
        return None
    if isinstance(dt, DType):
        return dt.torch
    if isinstance(dt, torch.dtype):
        return dt
    raise TypeError(dt)


# ------------------------------------------------------------------------------------------------ tensor
class Tensor:
    """Wraps a torch tensor; operators convert the other operand the way TF's overloads do (A2)."""
    __array_priority__ = 1000          # ndarray.__mul__(Tensor) defers to Tensor.__rmul__

    def __init__(self, t):
        self.t = t

    # -- TF-like attributes
    @property
    def shape(self):
        return tuple(self.t.shape)

    @property
    def dtype(self):
        return self.t.dtype

    def numpy(self):
        return self.t.detach().cpu().numpy()

    def __repr__(self):
        return "standin.Tensor(%r)" % (self.t,)

    def __len__(self):
        return self.t.shape[0]

    def __iter__(self):
        for i in builtins.range(self.t.shape[0]):
            yield Tensor(self.t[i])

    def __bool__(self):
        return bool(self.t)

    def __int__(self):
        return int(self.t)

    def __float__(self):
        return float(self.t)

    def __index__(self):
        return int(self.t)

    # -- indexing: strided_slice; negative steps (file_util.py:48 ``[::-1]``) become flips
    def __getitem__(self, key):
        if not isinstance(key, tuple):
            key = (key,)
        out, flips, dim = [], [], 0
        n_real = sum(1 for k in key if k is not Ellipsis and k is not None)
        for k in key:
            if k is Ellipsis:
                dim += self.t.dim() - n_real
                out.append(k)
                continue
            if k is None:
                out.append(k)
                continue
            if isinstance(k, Tensor):
                k = k.t if k.t.dim() else int(k.t)
            if isinstance(k, slice) and k.step is not None and k.step < 0:
                if not (k.start is None and k.stop is None and k.step == -1):
                    raise NotImplementedError("only [::-1]")
                flips.append(dim)
                k = slice(None)
            out.append(k)
            dim += 1
        r = self.t[tuple(out)]
        if flips:
            if any(isinstance(k, int) for k in out):
                raise NotImplementedError
            r = torch.flip(r, flips)
        return Tensor(r)

    # -- arithmetic / comparison overloads (math_ops.py: ``_OverrideBinaryOperatorHelper``)
    def _other(self, o):
        return _convert(o, like=self.t)

    def __add__(self, o): return Tensor(self.t + self._other(o))
    def __radd__(self, o): return Tensor(self._other(o) + self.t)
    def __sub__(self, o): return Tensor(self.t - self._other(o))
    def __rsub__(self, o): return Tensor(self._other(o) - self.t)
    def __mul__(self, o): return Tensor(self.t * self._other(o))
    def __rmul__(self, o): return Tensor(self._other(o) * self.t)
    def __truediv__(self, o): return Tensor(self.t / self._other(o))
    def __rtruediv__(self, o): return Tensor(self._other(o) / self.t)
    def __neg__(self): return Tensor(-self.t)
    def __ge__(self, o): return Tensor(self.t >= self._other(o))
    def __gt__(self, o): return Tensor(self.t > self._other(o))
    def __le__(self, o): return Tensor(self.t <= self._other(o))
    def __lt__(self, o): return Tensor(self.t < self._other(o))


def _convert(v, like=None, dtype=None):
    """ops.convert_to_tensor(v, dtype): a constant of the requested (or the partner's) dtype (A2).
    Float constants are rounded to fp32 first, also in the widened twin."""
    if isinstance(v, Tensor):
        return v.t
    if isinstance(v, Variable):
        return v.read().t
    if isinstance(v, torch.Tensor):
        return v
    want = dtype if dtype is not None else (like.dtype if like is not None else None)
    if isinstance(v, (list, tuple)) and any(isinstance(e, (Tensor, Variable)) for e in v):
        return torch.stack([_convert(e, like=like, dtype=dtype) for e in v])
    a = np.asarray(v)
    if want is None:
        want = _STATE["float"] if a.dtype.kind == "f" else {"i": torch.int32, "u": torch.int32, "b": torch.bool}[a.dtype.kind]
    if want in (torch.float32, torch.float64):
        return torch.from_numpy(np.ascontiguousarray(a.astype(np.float32)).reshape(a.shape)).to(want)
    return torch.from_numpy(np.ascontiguousarray(a).reshape(a.shape)).to(want)


def _t(x, like=None):
    return _convert(x, like=like)


def _tap(name, value):
    for fn in _STATE["taps"]:
        fn(name, value)
    return value


# ------------------------------------------------------------------------------------------------ construction
def constant(value, dtype=None, shape=None, name=None):
    t = _convert(value, dtype=_as_dtype(dtype))
    if shape is not None:
        t = t.reshape(tuple(shape)) if t.numel() > 1 else t.expand(tuple(shape)).clone()
    return Tensor(t)


def convert_to_tensor(value, dtype=None, name=None):
    return Tensor(_convert(value, dtype=_as_dtype(dtype)))


def zeros(shape, dtype=float32, name=None):
    return Tensor(torch.zeros(tuple(int(s) for s in shape), dtype=_as_dtype(dtype)))


def range(start, limit=None, delta=1, dtype=None, name=None):     # noqa: A001  (tf.range; int32 for Python ints)
    if limit is None:
        start, limit = 0, start
    return Tensor(torch.arange(int(start), int(limit), int(delta), dtype=_as_dtype(dtype) or torch.int32))


def meshgrid(*args, indexing="xy"):
    """array_ops.meshgrid, default 'xy' indexing: outputs have shape (len(y), len(x))."""
    xs = [_t(a) for a in args]
    out = torch.meshgrid(*xs, indexing=indexing)
    return [Tensor(o) for o in out]


def shape(x, name=None, out_type=None):
    return Tensor(torch.tensor(list(_t(x).shape), dtype=torch.int32))


def cast(x, dtype, name=None):
    """float -> int truncates toward zero (A6); float32 -> float32 is the identity and carries the gradient."""
    return Tensor(_t(x).to(_as_dtype(dtype)))


# ------------------------------------------------------------------------------------------------ elementwise math
class _Sigmoid(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        one = torch.ones((), dtype=x.dtype)
        y = one / (one + torch.exp(-x))                          # A3
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return dy * (y * (1 - y))                                # math_grad._SigmoidGrad -> sigmoid_grad(y, dy)


class _Softmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        e = torch.exp(x - torch.amax(x, dim=-1, keepdim=True))
        one = torch.ones((), dtype=x.dtype)
        y = e * (one / torch.sum(e, dim=-1, keepdim=True))       # A3: multiply by the reciprocal of the sum
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return (dy - torch.sum(dy * y, dim=-1, keepdim=True)) * y   # nn_grad._SoftmaxGrad


def sigmoid(x, name=None):
    return Tensor(_Sigmoid.apply(_t(x)))


def exp(x, name=None):
    return Tensor(torch.exp(_t(x)))


def log(x, name=None):
    return Tensor(torch.log(_t(x)))


def square(x, name=None):
    x = _t(x)
    return Tensor(x * x)


def floor(x, name=None):
    return Tensor(torch.floor(_t(x)))


def pow(x, y, name=None):                                          # noqa: A001
    x = _t(x)
    return Tensor(torch.pow(x, _convert(y, like=x)))


def add(x, y, name=None):
    x = _t(x)
    return Tensor(x + _convert(y, like=x))


def multiply(x, y, name=None):
    if isinstance(x, (Tensor, Variable)):
        a = _t(x)
        b = _convert(y, like=a)
    else:                                                         # tf.multiply(python_float, tensor): yolov3_loss.py:157
        b = _t(y)
        a = _convert(x, like=b)
    return Tensor(_tap("multiply", a * b))


def maximum(x, y, name=None):
    """Gradient goes to x where x >= y, else to y (A4) -- torch.maximum splits 0.5/0.5 at equality, so route it."""
    x = _t(x)
    y = _convert(y, like=x)
    return Tensor(torch.where(x >= y, x, y))


def minimum(x, y, name=None):
    x = _t(x)
    y = _convert(y, like=x)
    return Tensor(torch.where(x <= y, x, y))


def clip_by_value(t, clip_value_min, clip_value_max, name=None):
    """clip_ops.clip_by_value: maximum(minimum(t, max), min) (A4)."""
    return maximum(minimum(t, clip_value_max), clip_value_min)


def logical_and(x, y, name=None):
    return Tensor(torch.logical_and(_t(x), _t(y)))


# ------------------------------------------------------------------------------------------------ reductions
def _axes(axis):
    if axis is None:
        return None
    if isinstance(axis, (list, tuple)):
        return [int(a) for a in axis]
    return [int(axis)]


def reduce_sum(x, axis=None, keepdims=False, name=None):
    x = _t(x)
    ax = _axes(axis)
    return Tensor(torch.sum(x) if ax is None else torch.sum(x, dim=ax, keepdim=keepdims))


def reduce_mean(x, axis=None, keepdims=False, name=None):
    """Eigen MeanReducer: sum, then one division by the count."""
    x = _t(x)
    ax = _axes(axis)
    if ax is None:
        r = torch.sum(x) / x.numel()
        return Tensor(r.reshape((1,) * x.dim()) if keepdims else r)
    cnt = 1
    for a in ax:
        cnt *= x.shape[a]
    return Tensor(torch.sum(x, dim=ax, keepdim=keepdims) / cnt)


def reduce_max(x, axis=None, keepdims=False, name=None):
    """-inf over an empty axis; gradient split equally among ties (A5; torch.amax does the same)."""
    x = _t(x)
    ax = _axes(axis)
    if ax is None:
        ax = list(np.arange(x.dim()))
    if any(x.shape[a] == 0 for a in ax):
        shp = [1 if (i in [a % x.dim() for a in ax]) else s for i, s in enumerate(x.shape)]
        r = torch.full(shp, float("-inf"), dtype=x.dtype)
        if not keepdims:
            r = r.reshape([s for i, s in enumerate(x.shape) if i not in [a % x.dim() for a in ax]])
        return Tensor(r)
    return Tensor(torch.amax(x, dim=ax, keepdim=keepdims))


def count_nonzero(x, axis=None, name=None):
    return Tensor(torch.count_nonzero(_t(x)).to(torch.int64))


def arg_max(x, dimension, output_type=int64, name=None):
    """first index among equal maxima (A6)."""
    x = _tap("arg_max_input", _t(x))
    return Tensor(torch.argmax(x, dim=int(dimension)).to(_as_dtype(output_type)))


argmax = arg_max


# ------------------------------------------------------------------------------------------------ shape ops
def _ints(seq):
    return [int(s) for s in seq]


def reshape(x, shape, name=None):            # noqa: A002
    return Tensor(_t(x).reshape(_ints(shape)))


def expand_dims(x, axis, name=None):
    return Tensor(_t(x).unsqueeze(int(axis)))


def concat(values, axis, name=None):
    ts = [_t(v) for v in values]
    want = ts[0].dtype
    return Tensor(torch.cat([t.to(want) for t in ts], dim=int(axis)))


def stack(values, axis=0, name=None):
    return Tensor(torch.stack([_t(v) for v in values], dim=int(axis)))


def tile(x, multiples, name=None):
    return Tensor(_t(x).repeat(*_ints(multiples)))


def reverse(x, axis, name=None):
    return Tensor(torch.flip(_t(x), _ints(axis)))


def where(condition, x=None, y=None, name=None):
    """one-argument form: (k, rank) int64 coordinates in row-major order (A6)."""
    if x is None and y is None:
        return Tensor(torch.nonzero(_t(condition)))
    c = _t(condition)
    a = _t(x)
    return Tensor(torch.where(c, a, _convert(y, like=a)))


def gather_nd(params, indices, name=None):
    """indices (..., K) address the first K axes of params; K = 0 columns never occurs here.
    A row of indices out of range raises (TF CPU kernel does too).  Gradient accumulates duplicates (A6)."""
    p = _t(params)
    idx = _t(indices).long()
    k = idx.shape[-1]
    for d in np.arange(k):
        if idx.numel() and (int(idx[..., d].min()) < 0 or int(idx[..., d].max()) >= p.shape[d]):
            raise IndexError("gather_nd: index out of range (InvalidArgumentError on the TF CPU kernel)")
    return Tensor(p[tuple(idx[..., d] for d in np.arange(k))])


def sparse_to_dense(sparse_indices, output_shape, sparse_values, default_value=0, validate_indices=True, name=None):
    idx = _t(sparse_indices).long()
    if validate_indices:
        raise NotImplementedError("the reference passes validate_indices=False (yolov3_loss.py:328)")
    out = torch.full(_ints(output_shape), float(default_value), dtype=_STATE["float"])
    if idx.numel():
        out[tuple(idx[:, d] for d in np.arange(idx.shape[1]))] = float(sparse_values)
    return Tensor(_tap("sparse_to_dense", out))


def one_hot(indices, depth, name=None):
    """rows for indices outside [0, depth) are all zero; on_value/off_value 1/0 in float32."""
    i = _t(indices).long()
    return Tensor((i.unsqueeze(-1) == torch.arange(int(depth))).to(_STATE["float"]))


# ------------------------------------------------------------------------------------------------ control flow
def map_fn(fn, elems, dtype=None, parallel_iterations=None, name=None):
    """serial over the leading axis, results stacked (functional_ops.map_fn)."""
    single = not isinstance(elems, (tuple, list))
    es = [elems] if single else list(elems)
    n = _t(es[0]).shape[0]
    outs = []
    for i in np.arange(n):
        item = [Tensor(_t(e)[int(i)]) for e in es]
        outs.append(_t(fn(item[0] if single else tuple(item))))
    return Tensor(torch.stack(outs, dim=0))


def cond(pred, true_fn=None, false_fn=None, name=None):
    """only the taken branch's ops (incl. its update_add) run (control_flow_ops.cond)."""
    p = pred
    if isinstance(p, Tensor):
        p = bool(p.t)
    return true_fn() if p else false_fn()


@contextlib.contextmanager
def control_dependencies(control_inputs):
    yield


@contextlib.contextmanager
def name_scope(name, default_name=None, values=None):
    yield


@contextlib.contextmanager
def variable_scope(name_or_scope, *a, **k):
    yield


# ------------------------------------------------------------------------------------------------ variables
class Variable:
    """A graph variable.  ``assign`` takes effect at once (the loss only writes the detail variables);
    ``update_add`` is queued until end_step() (A1)."""

    def __init__(self, value, dtype, name):
        self.name = name
        self.value = _convert(value, dtype=dtype).clone()

    def read(self):
        return Tensor(self.value.clone())

    def assign(self, value, name=None):
        self.value = _t(value).detach().clone().to(self.value.dtype)
        return self

    def numpy(self):
        return self.value.numpy()

    def __le__(self, o): return Tensor(self.value <= _convert(o, like=self.value))
    def __lt__(self, o): return Tensor(self.value < _convert(o, like=self.value))
    def __ge__(self, o): return Tensor(self.value >= _convert(o, like=self.value))
    def __gt__(self, o): return Tensor(self.value > _convert(o, like=self.value))


def constant_initializer(value=0, dtype=float32):
    return lambda shape, dt: torch.full(tuple(shape), float(value), dtype=dt)


def get_variable(name, shape=None, dtype=float32, initializer=None, **kw):   # noqa: A002
    dt = _as_dtype(dtype)
    return Variable(initializer(shape, dt), dt, name)


# ------------------------------------------------------------------------------------------------ gradients
def gradients(ys, xs, grad_ys=None, name=None):
    """tf.gradients(sum of ys, xs) -- what the optimizer calls (utils/radam.py:57 get_gradients)."""
    ys = ys if isinstance(ys, (list, tuple)) else [ys]
    xs_l = xs if isinstance(xs, (list, tuple)) else [xs]
    total = sum(_t(y).sum() for y in ys)
    g = torch.autograd.grad(total, [_t(x) for x in xs_l], allow_unused=True)
    return [None if v is None else Tensor(v) for v in g]


# ------------------------------------------------------------------------------------------------ strings / files
# (label half of dataset/file_util.py:22-52; strings are Python str, 1-D string tensors are lists)
class _Sparse:
    def __init__(self, values):
        self.values = values


def string_split(source, delimiter=" ", skip_empty=True):
    out = []
    for s in source:
        s = s.decode() if isinstance(s, bytes) else s
        out += [f for f in s.split(delimiter) if (f != "" or not skip_empty)]
    return _Sparse(out)


def string_join(inputs, separator="", name=None):
    return separator.join(inputs)


def string_to_number(string_tensor, out_type=float32, name=None):
    """strtof: decimal text -> nearest float32."""
    vals = [np.float32(s) for s in string_tensor]
    return Tensor(torch.tensor(np.asarray(vals, dtype=np.float32)).to(_as_dtype(out_type)))


def read_file(filename, name=None):
    with open(filename, "rb") as f:
        return f.read()


def _decode_jpeg(contents, channels=0, **kw):
    import cv2
    img = cv2.imdecode(np.frombuffer(contents, np.uint8), cv2.IMREAD_COLOR)
    return Tensor(torch.from_numpy(np.ascontiguousarray(img[:, :, ::-1])))      # RGB uint8 (H, W, 3)


def _resize_image_with_pad(image, target_height, target_width, method=None):
    """Pixels are outside the head path: returns a zero canvas of the target size (only the label maths is pinned)."""
    c = _t(image).shape[-1]
    return Tensor(torch.zeros((int(target_height), int(target_width), c), dtype=_t(image).dtype))


def _convert_image_dtype(image, dtype, saturate=False, name=None):
    return Tensor(_t(image).to(_as_dtype(dtype)) / 255.0)


image = types.SimpleNamespace(
    decode_jpeg=_decode_jpeg, resize_image_with_pad=_resize_image_with_pad, convert_image_dtype=_convert_image_dtype,
    ResizeMethod=types.SimpleNamespace(NEAREST_NEIGHBOR=1, BILINEAR=0))

nn = types.SimpleNamespace(softmax=lambda logits, axis=-1, name=None: Tensor(_Softmax.apply(_t(logits))))


def _softmax_axis_check(logits, axis=-1, name=None):
    if axis not in (-1, _t(logits).dim() - 1):
        raise NotImplementedError
    return Tensor(_Softmax.apply(_t(logits)))


nn.softmax = _softmax_axis_check


class TensorShape(list):
    pass


# ------------------------------------------------------------------------------------------------ stand-in controls
class _Standin:
    """Controls used by the golden generator only (not TF API)."""

    @staticmethod
    def set_precision(name):
        _STATE["float"] = {"float32": torch.float32, "float64": torch.float64}[name]

    @staticmethod
    def end_step():
        """apply the queued ``update_add``s: the end of one ``session.run`` (A1)."""
        for var, inc in _STATE["pending"]:
            var.value = var.value + int(inc)
        _STATE["pending"].clear()

    @staticmethod
    def input(array, requires_grad=False):
        """a placeholder fed with fp32 data (widened in the twin); a leaf for ``tf.gradients``."""
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(array, dtype=np.float32))).to(_STATE["float"])
        return Tensor(t.clone().requires_grad_(requires_grad))

    @staticmethod
    @contextlib.contextmanager
    def tap(fn):
        """fn(op_name, torch_tensor) is called with the result of sparse_to_dense / multiply and the input of arg_max
        while active."""
        _STATE["taps"].append(fn)
        try:
            yield
        finally:
            _STATE["taps"].remove(fn)


standin = _Standin
