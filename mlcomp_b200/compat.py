"""Make this package answer to the reference's module names.

MLComp finds executors by module path: ``Storage.import_executor`` scans for a
class whose name matches the YAML ``type`` and ``importlib.import_module``s its
file (mlcomp/worker/storage.py:249-306), and user code imports the bases as
``from mlcomp.worker.executors.valid import Valid``
(examples/digit-recognizer/executors/valid.py:6-8).  ``install_as_mlcomp()``
registers aliases in ``sys.modules`` so that those imports resolve to the
classes of this package - for the hot-path modules only:

    mlcomp.worker.executors           (Executor, StepWrap, ...)
    mlcomp.worker.executors.{valid,infer,equation}
    mlcomp.worker.reports.{classification,segmenation}
    mlcomp.contrib.metrics.dice       dice_numpy
    mlcomp.contrib.transform.{rle,tta}
    mlcomp.contrib.torch.tensors      flip
    mlcomp.contrib.catalyst.callbacks.inference   InferBestCallback
    mlcomp.utils.torch                apply_activation, infer

Nothing is imported or written as a side effect (the reference's own
``mlcomp/__init__.py:7-51`` creates folders under the home directory; this
shim does not).  A real ``mlcomp`` that is already imported is left alone
unless ``force=True``.
"""
import importlib
import sys
import types

ALIASES = {
    'mlcomp.worker.executors': 'mlcomp_b200.worker.executors',
    'mlcomp.worker.executors.valid': 'mlcomp_b200.worker.executors.valid',
    'mlcomp.worker.executors.infer': 'mlcomp_b200.worker.executors.infer',
    'mlcomp.worker.executors.equation':
        'mlcomp_b200.worker.executors.equation',
    'mlcomp.worker.executors.base': 'mlcomp_b200.worker.executors.base',
    'mlcomp.worker.reports': 'mlcomp_b200.worker.reports',
    'mlcomp.worker.reports.classification':
        'mlcomp_b200.worker.reports.classification',
    'mlcomp.worker.reports.segmenation':
        'mlcomp_b200.worker.reports.segmenation',
    'mlcomp.contrib.metrics.dice': 'mlcomp_b200.contrib.metrics.dice',
    'mlcomp.contrib.transform.rle': 'mlcomp_b200.contrib.transform.rle',
    'mlcomp.contrib.transform.tta': 'mlcomp_b200.contrib.transform.tta',
    'mlcomp.contrib.torch.tensors': 'mlcomp_b200.contrib.torch.tensors',
    'mlcomp.contrib.catalyst.callbacks.inference':
        'mlcomp_b200.contrib.catalyst.callbacks.inference',
    'mlcomp.utils.torch': 'mlcomp_b200.utils.torch',
}
_PACKAGES = ('mlcomp', 'mlcomp.worker', 'mlcomp.contrib',
             'mlcomp.contrib.metrics', 'mlcomp.contrib.transform',
             'mlcomp.contrib.torch', 'mlcomp.contrib.catalyst',
             'mlcomp.contrib.catalyst.callbacks', 'mlcomp.utils')


def install_as_mlcomp(force=False):
    """Register the aliases; returns the list of module names provided."""
    real = sys.modules.get('mlcomp')
    if real is not None and not getattr(real, '__mlcomp_b200_shim__', False) \
            and not force:
        raise RuntimeError('a real `mlcomp` package is already imported; '
                           'pass force=True to shadow its hot-path modules')
    for name in _PACKAGES:
        if name not in sys.modules or force:
            pkg = types.ModuleType(name)
            pkg.__path__ = []              # a namespace with no files
            pkg.__mlcomp_b200_shim__ = True
            sys.modules[name] = pkg
    for alias, target in ALIASES.items():
        mod = importlib.import_module(target)
        sys.modules[alias] = mod
        parent, _, leaf = alias.rpartition('.')
        setattr(sys.modules[parent], leaf, mod)
    for name in _PACKAGES[1:]:
        parent, _, leaf = name.rpartition('.')
        setattr(sys.modules[parent], leaf, sys.modules[name])
    return sorted(ALIASES)


def uninstall():
    for name in list(sys.modules):
        if name == 'mlcomp' or name.startswith('mlcomp.'):
            mod = sys.modules[name]
            if getattr(mod, '__mlcomp_b200_shim__', False) or \
                    getattr(mod, '__name__', '').startswith('mlcomp_b200'):
                del sys.modules[name]


__all__ = ['install_as_mlcomp', 'uninstall', 'ALIASES']
