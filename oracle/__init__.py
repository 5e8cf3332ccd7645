"""CPU oracle for the quantum-simulation path of sq-learn.

TEST INFRASTRUCTURE ONLY.  Nothing under ``sq-learn_b200/`` imports this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the CPU-baseline legs
of ``bench.py`` may.  It restates, in plain NumPy, the algorithms of the
reference files cited in each function, and is pinned against the reference's
own source run in this container (``oracle/refload.py`` +
``tests/golden/gen_golden.py``; fixtures committed under ``tests/golden/``).

Parity status: the reference ships no test, golden vector or known-answer
table for any quantum symbol, so every pin below comes from
outputs of the reference itself, generated here with the committed script.
"""
