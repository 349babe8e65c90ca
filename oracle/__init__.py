"""CPU oracle for the Pynama KLE hot path -- TEST INFRASTRUCTURE ONLY.

This package is a numpy/scipy restatement of the reference's algorithm for the
hot path named in BASELINE.json (spectral-element KLE matrices and operators,
their scatter-add assembly, the KLE solve, the vorticity-transport RHS and the
RK5(4)BS step).  Every function cites the reference file:line it follows.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it.  Nothing under ``pynama_b200/`` does:
the product path is CUDA behind ``libpynama_b200.so`` and fails loudly without
it.

Modules: sem (tables, element matrices), assembly (scatter-add, also restricted to a
sample of rows), timestep (KLE solve, RHS, RK5BS fixed and adaptive, preconditioners as
numpy), ibm (immersed boundary), boxmesh (standalone box-mesh tables), cport + ckle.c
(C + OpenMP assembly / SpMV / CG for the CPU arm of bench.py at the full size).

Pinning: the element tables and element matrices are pinned against golden
vectors generated from the reference's own ``Spectral`` class
(``tests/golden/make_golden.py`` -> ``tests/golden/*.npz``); the assembled
solve is pinned through the reference's known-answer tests KAT-1/2/3
(``src/tests/test_solver.py``); the assembled matrices (patterns bit for bit),
the no-slip dof classes, the KLE solves, evalRHS and fixed RK5BS steps are pinned
against outputs of the reference's own unmodified modules run over a numpy
stand-in for petsc4py (``tests/refshim/``, ``tests/golden/make_golden_refpath.py``
-> ``tests/golden/refpath_*.npz``, ``tests/test_reference_path.py``).  Global node numbering of PETSc DMPlex for
ngl >= 3, PETSc's TS adaptivity and KSP internals are NOT in the reference
tree: for those parts parity is unpinned.
"""
