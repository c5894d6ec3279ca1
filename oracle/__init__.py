"""oracle/ -- TEST INFRASTRUCTURE, never part of the product path.

* ``legged_oracle.py``  CPU (torch fp32) restatement of the reference's per-env step path, each
  function citing the reference file:line it follows.  Travels to the GPU box; the ``-m gpu``
  parity tests, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
  legs are the only importers.
* ``reference_loader.py`` + ``refshim/``  import the reference's own, unmodified ``legged_gym``
  from /root/reference over a fake Isaac Gym (this container only) -- used to validate the
  restatement and to generate ``tests/golden`` fixtures (``make_golden.py``).

Parity status: PINNED against outputs of the reference itself run in this container (the reference
ships no golden vectors or asserting tests for this path: legged_gym/tests/test_env.py:42-52).
The six ``isaacgym.torch_utils`` functions are a closed third-party dependency restated from the
published Preview-4 source: unpinned at that boundary.

Nothing under ``legged_gym_b200/`` imports this package.
"""
