"""ngsolve.krylovspace stand-in (test infrastructure): templates/NavierStokesSIMPLE_iterative.py:2 imports MinRes from
here at module level; the methods run for the golden vectors (DoTimeStep, Project) never call it."""


def MinRes(*a, **k):
    raise NotImplementedError("ngsolve.krylovspace.MinRes is NGSolve's own solver; not part of the stand-in")
