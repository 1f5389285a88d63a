"""``Equation`` - base of ``Valid`` / ``Infer``.

The reference ships only the SUBCLASS side of this contract
(examples/digit-recognizer/executors/{valid,infer}.py import
``mlcomp.worker.executors.valid.Valid`` / ``infer.Infer``, files that are
absent from the snapshot - SURVEY.md section 8 row A13).  Reconstructed from
the call sites:

* ctor takes ``cache_names``, ``layout`` and the executor's YAML mapping
  (examples/digit-recognizer/all.yml:22-30, pipe.yml:7-27, plus ``model_id`` /
  ``model_name`` injected by server/back/create_dags/model_start.py:40-42);
  every YAML key becomes an attribute (worker/executors/kaggle.py:87-101
  relies on ``hasattr(self, 'model_id')`` / ``self.suffix``);
* YAML values that are expressions over named variables and functions, e.g.
  ``y: torch(x, file='net.pth', batch_size=128)`` (all.yml:25), are
  *equations*; ``solve('y')`` evaluates one for the current part;
* ``[0, count())`` is processed in parts ``(lo, hi)`` handed to
  ``adjust_part``.
"""
import ast
import operator
import os

from .base import Executor

_BIN = {ast.Add: operator.add, ast.Sub: operator.sub,
        ast.Mult: operator.mul, ast.Div: operator.truediv,
        ast.Pow: operator.pow}


class Equation(Executor):
    # names that may appear as YAML keys with plain (non-equation) values
    _plain = {'type', 'gpu', 'cpu', 'memory', 'depends', 'grid', 'distr',
              'name', 'suffix', 'model_name', 'model_id', 'layout'}

    def __init__(self, cache_names=None, layout=None, name=None,
                 model_id=None, model_name=None, suffix='', max_count=None,
                 part_size=None, test=False, prepare_submit=False,
                 plot_count=0, **kwargs):
        super().__init__(**kwargs)
        self.cache_names = list(cache_names or [])
        self.layout = layout
        self.name = name
        self.model_id = model_id
        self.model_name = model_name
        self.suffix = suffix
        self.max_count = max_count
        self.part_size = part_size
        self.test = test
        self.prepare_submit = prepare_submit
        self.plot_count = plot_count
        self.equations = {}
        self.part = None
        self._cache = {}
        texts = {k: v for k, v in kwargs.items()
                 if isinstance(v, str) and k not in self._plain}
        # equations: anything with a function call, then (to a fixed point)
        # arithmetic over names that are themselves equations, e.g.
        # ``y: (y1 + y2) / 2``; 'data/train.csv' stays a plain string
        for k, v in texts.items():
            if self._has_call(v):
                self.equations[k] = v
        grew = True
        while grew:
            grew = False
            for k, v in texts.items():
                if k not in self.equations and \
                        self._is_arithmetic_over(v, self.equations):
                    self.equations[k] = v
                    grew = True
        for k, v in kwargs.items():
            if k not in self.equations and not hasattr(type(self), k):
                setattr(self, k, v)          # YAML key -> attribute

    # -- parsing ---------------------------------------------------------------
    @staticmethod
    def _parse(text):
        try:
            return ast.parse(text.strip(), mode='eval')
        except SyntaxError:
            return None

    @classmethod
    def _has_call(cls, text):
        tree = cls._parse(text)
        return tree is not None and any(
            isinstance(n, ast.Call) and isinstance(n.func, ast.Name)
            for n in ast.walk(tree))

    @classmethod
    def _is_arithmetic_over(cls, text, names):
        tree = cls._parse(text)
        if tree is None:
            return False
        seen = [n.id for n in ast.walk(tree) if isinstance(n, ast.Name)]
        ok = (ast.Expression, ast.BinOp, ast.UnaryOp, ast.Name, ast.Constant,
              ast.Load, ast.operator, ast.unaryop)
        return bool(seen) and all(n in names for n in seen) and \
            all(isinstance(n, ok) for n in ast.walk(tree))

    def functions(self):
        """Callables usable inside equations."""
        return {'torch': self.torch, 'load': self.load}

    def torch(self, x, file, **kwargs):
        """``torch(x, file=..., batch_size=...)`` of the YAML: forward pass
        of a traced model (mlcomp/utils/torch.py:51-70), logits kept on the
        device."""
        from mlcomp_b200.utils.torch import infer
        path = file if os.path.exists(file) else os.path.join('models', file)
        return infer(x, path, **kwargs)

    def load(self, file, **kwargs):
        """``load('valid.logits.npy')``: arrays dumped by InferBestCallback
        (contrib/catalyst/callbacks/inference.py:38-49), sliced to the part."""
        import numpy as np
        arr = np.load(file, mmap_mode='r')
        if self.part is not None:
            arr = arr[self.part[0]:self.part[1]]
        return arr

    def _eval(self, node):
        if isinstance(node, ast.Expression):
            return self._eval(node.body)
        if isinstance(node, ast.Constant):
            return node.value
        if isinstance(node, ast.Name):
            if node.id in self.equations:
                return self.solve(node.id)
            if hasattr(self, node.id):
                return getattr(self, node.id)
            raise NameError(f'equation variable {node.id!r} is not defined')
        if isinstance(node, ast.BinOp) and type(node.op) in _BIN:
            return _BIN[type(node.op)](self._eval(node.left),
                                       self._eval(node.right))
        if isinstance(node, ast.UnaryOp) and isinstance(node.op, ast.USub):
            return -self._eval(node.operand)
        if isinstance(node, (ast.List, ast.Tuple)):
            return [self._eval(e) for e in node.elts]
        if isinstance(node, ast.Call) and isinstance(node.func, ast.Name):
            fns = self.functions()
            if node.func.id not in fns:
                raise NameError(f'equation function {node.func.id!r} is not '
                                'defined')
            args = [self._eval(a) for a in node.args]
            kw = {k.arg: self._eval(k.value) for k in node.keywords}
            return fns[node.func.id](*args, **kw)
        raise SyntaxError(f'unsupported syntax in equation: {ast.dump(node)}')

    def solve(self, name):
        """Value of equation ``name`` for the current part (cached per part
        when listed in ``cache_names``)."""
        key = (name, self.part)
        if name in self.cache_names and key in self._cache:
            return self._cache[key]
        if name not in self.equations:
            raise KeyError(f'no equation named {name!r} in the executor config')
        val = self._eval(ast.parse(self.equations[name].strip(), mode='eval'))
        if name in self.cache_names:
            self._cache = {k: v for k, v in self._cache.items()
                           if k[1] == self.part}
            self._cache[key] = val
        return val

    # -- parts -----------------------------------------------------------------
    def generate_parts(self, count):
        if self.max_count:
            count = min(count, int(self.max_count))
        size = int(self.part_size) if self.part_size else max(count, 1)
        return [(lo, min(count, lo + size)) for lo in range(0, count, size)]

    def set_part(self, part):
        self.part = tuple(part)
        self.adjust_part(self.part)

    # hooks every subclass provides (valid.py:32-57, infer.py:43-60)
    def create_base(self):
        pass

    def count(self):
        raise NotImplementedError

    def adjust_part(self, part):
        pass


__all__ = ['Equation']
