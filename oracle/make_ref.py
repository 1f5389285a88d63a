"""Recipe for ``oracle/_ref/``: the reference's OWN leaf modules of the path,
taken unmodified from ``/root/reference`` so that the CPU baseline of
``bench.py`` can time the reference itself (``cpu_baseline.kind ==
"reference"``) on the GPU box, where ``/root/reference`` does not exist.

    python oracle/make_ref.py        (run by __graft_entry__.build() as well)

``oracle/_ref/`` is git-ignored (the files are the reference's, not ours, and
never enter the history) but not gpurun-ignored, so it travels with the
snapshot like the built .so.  The three modules import nothing but numpy /
torch; the rest of the reference (``mlcomp/__init__.py`` needs sqlalchemy and
writes to the home directory) is never imported.

TEST INFRASTRUCTURE ONLY - see oracle/__init__.py.
"""
import importlib.util
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = '/root/reference'
DEST = os.path.join(HERE, '_ref')
FILES = {
    'dice.py': 'mlcomp/contrib/metrics/dice.py',        # dice_numpy
    'rle.py': 'mlcomp/contrib/transform/rle.py',        # mask2rle, rle2mask
    'tensors.py': 'mlcomp/contrib/torch/tensors.py',    # flip
}
# the reference's only in-tree users of the Valid / Infer bases: run as they
# are by tests/test_reference_examples.py on the bases of this package
EXAMPLES = {
    'examples/valid.py': 'examples/digit-recognizer/executors/valid.py',
    'examples/infer.py': 'examples/digit-recognizer/executors/infer.py',
    'examples/dataset.py': 'examples/digit-recognizer/dataset.py',
}


def make():
    """Copy the files; returns True when oracle/_ref is complete."""
    if os.path.isdir(REF_ROOT):
        os.makedirs(DEST, exist_ok=True)
        os.makedirs(os.path.join(DEST, 'examples'), exist_ok=True)
        for name, rel in list(FILES.items()) + list(EXAMPLES.items()):
            shutil.copyfile(os.path.join(REF_ROOT, rel),
                            os.path.join(DEST, name))
    return all(os.path.exists(os.path.join(DEST, n)) for n in FILES)


def load():
    """The reference's functions, or None when oracle/_ref is absent:
    a namespace with dice_numpy, mask2rle, rle2mask, flip."""
    if not all(os.path.exists(os.path.join(DEST, n)) for n in FILES):
        return None
    mods = {}
    for name in FILES:
        spec = importlib.util.spec_from_file_location(
            'mlcomp_ref_' + name[:-3], os.path.join(DEST, name))
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        mods[name] = m

    class Ref:
        dice_numpy = staticmethod(mods['dice.py'].dice_numpy)
        mask2rle = staticmethod(mods['rle.py'].mask2rle)
        rle2mask = staticmethod(mods['rle.py'].rle2mask)
        flip = staticmethod(mods['tensors.py'].flip)
    return Ref


if __name__ == '__main__':
    print('oracle/_ref complete:', make())
