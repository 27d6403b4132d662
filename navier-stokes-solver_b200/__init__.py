"""B200-native Krylov solvers (drop-in for the reference CPU Krylov loops)."""
