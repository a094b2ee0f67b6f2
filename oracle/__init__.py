"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU checker for the mx-rcnn RPN proposal / box-processing hot path.  Nothing in the
product package (``mxrcnn_b200/``) imports this directory; only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may.

Contents
--------
* ``restate.py``   numpy restatement of the reference algorithm, every function citing
                   the reference ``file:line`` it follows.  Travels to the GPU box.
* ``synth.py``     seeded synthetic inputs (SURVEY.md section 8d recipes).
* ``ref_shim.py``  loader that imports the *real* reference from ``/root/reference``
                   (in memory, nothing copied).  Works only in the build container;
                   never used by ``-m gpu`` tests, ``smoke()`` or ``bench.py``.
* ``make_golden.py`` runs the real reference through the shim and writes the fixtures
                   under ``tests/golden/`` that pin ``restate.py`` and the CUDA path.

Parity status: PINNED -- the reference holds no golden vectors of its own for this path
(SURVEY.md section 4), so the restatement is pinned against outputs of the reference
code itself, executed in the build container by ``make_golden.py`` and committed as
``tests/golden/*.npz``.
"""
