"""Oracle = test infrastructure (CPU restatement of the reference path + golden generators).
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import it."""
