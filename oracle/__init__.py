"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU restatement (cpu_ref.c / transition_ref.py) of the reference's algorithms for the hot path and
the build recipe for the reference's own CUDA kernels (ref_build/ → oracle/_ref/).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package;
nothing under geot_b200/ does, and the product fails loudly without its CUDA library.
"""
