"""Keras subset used by the reference's model.py / resnet/models.py (see ../__init__.py)."""
import torch

from .. import TAPE, Tensor, _wrap
from . import layers, regularizers  # noqa: F401


def Input(shape):
    """Placeholder: zeros with the batch size set in layers.BATCH (static shapes, like a traced
    @tf.function).  The Model built from it is re-applied to real inputs by tape replay."""
    t = _wrap(torch.zeros((layers.BATCH,) + tuple(shape), dtype=torch.float32))
    t._tape_start = len(TAPE)
    return t


def _map(obj, env):
    if isinstance(obj, torch.Tensor):
        return env.get(id(obj), obj)
    if isinstance(obj, (list, tuple)):
        return type(obj)(_map(o, env) for o in obj)
    if isinstance(obj, dict):
        return {k: _map(v, env) for k, v in obj.items()}
    return obj


def _bind(old, new, env):
    if isinstance(old, torch.Tensor):
        env[id(old)] = new
    elif isinstance(old, (list, tuple)):
        for o, n in zip(old, new):
            _bind(o, n, env)


class Model:
    def __init__(self, inputs, outputs, name=None):
        self.inputs, self.outputs, self.name = inputs, outputs, name
        self.segment = list(TAPE[inputs._tape_start:])
        self._keep = self.segment            # keeps recorded tensors alive so ids stay unique

    def __call__(self, x, training=None):
        assert tuple(x.shape) == tuple(self.inputs.shape), (x.shape, self.inputs.shape)
        env = {id(self.inputs): x}
        for func, args, kwargs, out in self.segment:
            new = func(*_map(args, env), **_map(kwargs, env))
            _bind(out, new, env)
        return _map(self.outputs, env)

    @property
    def trainable_weights(self):
        return [w for _, w in layers.all_weights(trainable=True)]

    def summary(self):
        for n, w in layers.all_weights(trainable=True):
            print(n, tuple(w.shape))


class Sequential:      # imported by the reference, unused on the hot path
    pass


class utils:           # imported by the reference, unused on the hot path
    pass
