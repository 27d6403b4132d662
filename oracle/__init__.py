"""CPU oracle for the Krylov hot path -- TEST INFRASTRUCTURE (see krylov_oracle.py header).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import it."""
