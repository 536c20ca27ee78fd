"""Minimal stand-in for the `tensorflow` module, backed by torch-CPU.

TEST INFRASTRUCTURE ONLY.  TensorFlow cannot be installed in the build container (no
network), so the reference's Python files (model.py, loss.py, transformations.py,
resnet/models.py, old/image_warp.py, old/loss.py) are IMPORTED UNMODIFIED from
/root/reference with this package first on sys.path, and executed to produce the golden
vectors under tests/golden/ (see make_golden.py).  What is pinned that way is the
reference's own composition of ops - index arithmetic, the row/col swap of warp_features,
weight formulas, loss normalisation, layer wiring.  What is restated here (from TF's
documented semantics) is each primitive op: floor/cast/clip, gather / gather_nd, pad, tile,
reduce_*, image.resize (bilinear, half-pixel centres, no antialias), Conv2D 'same' padding,
inference-mode BatchNormalization (eps 1e-3), LeakyReLU alpha 0.3, MaxPool2D 2x2/2.

Tensors are a torch.Tensor subclass whose every op is appended to a tape, so that a Keras
functional `Model(inputs, outputs)` can be re-applied to new inputs by replaying the tape
segment that produced `outputs` from `inputs`.  Gradients come from torch autograd, whose
rules for these primitives coincide with TF's (floor / integer casts / clips of indices
carry none; gather <-> scatter-add).
"""
import builtins

import torch
import torch.nn.functional as _F

float32 = torch.float32
int32 = torch.int32
int64 = torch.int64

_DT = {"int32": torch.int32, "float32": torch.float32, "int64": torch.int64}

TAPE = []


class Tensor(torch.Tensor):
    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        out = super().__torch_function__(func, types, args, kwargs)
        TAPE.append((func, args, kwargs, out))
        return out


def _wrap(t):
    if isinstance(t, Tensor):
        return t
    return t.as_subclass(Tensor)


def _dtype(d):
    return _DT[d] if isinstance(d, str) else d


def _t(x, dtype=None):
    if isinstance(x, torch.Tensor):
        return _wrap(x)
    return _wrap(torch.as_tensor(x, dtype=dtype))


def convert_to_tensor(x, dtype=None):
    return _t(x, _dtype(dtype) if dtype is not None else None)


constant = convert_to_tensor


# ---- shapes / creation -----------------------------------------------------------------
def shape(x):
    return [int(s) for s in x.shape]


def unstack(x, axis=0):
    if isinstance(x, (list, tuple)):
        return list(x)
    return list(torch.unbind(x, dim=axis))


def zeros(shape_, dtype=float32):
    return _wrap(torch.zeros(tuple(shape_), dtype=_dtype(dtype)))


def ones(shape_, dtype=float32):
    return _wrap(torch.ones(tuple(shape_), dtype=_dtype(dtype)))


def ones_like(x):
    return torch.ones_like(x)


def range(*args):  # noqa: A001 - mirrors tf.range
    return _wrap(torch.arange(*[int(a) for a in args], dtype=torch.int32))


def meshgrid(a, b, indexing="xy"):
    return list(torch.meshgrid(a, b, indexing=indexing))


def cast(x, dtype):
    dtype = _dtype(dtype)
    if not isinstance(x, torch.Tensor):
        return _wrap(torch.as_tensor(x, dtype=dtype))
    return x.to(dtype)


# ---- elementwise -----------------------------------------------------------------------
def floor(x):
    return torch.floor(x)


def square(x):
    return x * x


def sqrt(x):
    return torch.sqrt(x)


def sin(x):
    return torch.sin(x)


def cos(x):
    return torch.cos(x)


def clip_by_value(x, lo, hi):
    return torch.clamp(x, min=lo, max=hi)


# ---- structure ---------------------------------------------------------------------------
def reshape(x, shape_):
    return torch.reshape(x, tuple(int(s) for s in shape_))


def expand_dims(x, axis):
    return torch.unsqueeze(x, axis)


def squeeze(x, axis=None):
    return torch.squeeze(x) if axis is None else torch.squeeze(x, axis)


def tile(x, multiples):
    return x.repeat(*[int(m) for m in multiples])


def stack(xs, axis=0):
    return torch.stack(list(xs), dim=axis)


def concat(xs, axis):
    return torch.cat([_t(x) for x in xs], dim=axis)


def transpose(x, perm=None):
    return x.permute(*perm) if perm is not None else x.t()


def pad(x, paddings):
    flat = []
    for lo, hi in reversed([tuple(p) for p in paddings]):
        flat += [int(lo), int(hi)]
    return _F.pad(x, flat)


def gather(params, indices, axis=0):
    assert axis == 0
    return params[indices.long()]


def gather_nd(params, indices):
    idx = indices.long()
    n = idx.shape[-1]
    return params[tuple(idx[..., k] for k in builtins.range(n))]


# ---- reductions --------------------------------------------------------------------------
def reduce_sum(x, axis=None):
    return torch.sum(x) if axis is None else torch.sum(x, dim=axis)


def reduce_mean(x, axis=None):
    return torch.mean(x) if axis is None else torch.mean(x, dim=axis)


def add_n(xs):
    out = xs[0]
    for x in xs[1:]:
        out = out + x
    return out


def norm(x, axis=None, keepdims=False):
    return torch.linalg.norm(x) if axis is None else torch.linalg.norm(x, dim=axis, keepdim=keepdims)


def argmax(x, axis=None):
    return torch.argmax(x, dim=axis)


class _Math:
    abs = staticmethod(torch.abs)
    accumulate_n = staticmethod(add_n)
    floordiv = staticmethod(lambda a, b: torch.div(a, b, rounding_mode="floor"))
    floormod = staticmethod(lambda a, b: torch.remainder(a, b))


math = _Math()


class _Linalg:
    matvec = staticmethod(lambda a, b: torch.matmul(_t(a), _t(b).unsqueeze(-1)).squeeze(-1))


linalg = _Linalg()


class _Image:
    @staticmethod
    def resize(images, size):
        """tf.image.resize default: bilinear, half_pixel_centers, antialias=False."""
        oh, ow = int(size[0]), int(size[1])
        y = _F.interpolate(images.permute(0, 3, 1, 2), size=(oh, ow), mode="bilinear", align_corners=False,
                           antialias=False)
        return y.permute(0, 2, 3, 1)


image = _Image()


# ---- autodiff / misc --------------------------------------------------------------------
def function(fn=None, **_kw):
    return fn if fn is not None else (lambda f: f)


class GradientTape:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def gradient(self, target, sources):
        gs = torch.autograd.grad(target, list(sources), allow_unused=True)
        return list(gs)


from . import keras  # noqa: E402,F401
