"""CPU oracle for the MLComp scoring / post-processing hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``mlcomp_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs use it, and only as the checker
(or as the CPU baseline that is timed *beside* the CUDA path).

Parity status: the reference (deepalcoholic/mlcomp 20.3.1.1) ships no golden
vectors, known-answer tests or fixtures for this path (its only test file is
``mlcomp/db/tests/test_project.py``), so this oracle is pinned instead against
the outputs of the reference's own functions, executed in the build container
by ``tests/golden/make_golden.py`` (which loads ``dice.py``, ``rle.py`` and
``tensors.py`` from ``/root/reference`` by file path) and committed as
``tests/golden/*.npz``; ``tests/golden/make_golden_composite.py`` does the same
for ``encode_pred`` / ``plot_mask`` (the two methods are cut out of
``segmenation.py`` with ``ast`` and run with this container's OpenCV 4.13).
``tests/test_oracle_golden.py`` replays them.
"""
from .scoring import *  # noqa: F401,F403
from .scoring import __all__  # noqa: F401
