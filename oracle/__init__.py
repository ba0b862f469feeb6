"""CPU oracle (test infrastructure only) — see oracle/step_tail.py for the contract."""
