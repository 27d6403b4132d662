"""B200-native Krylov solvers for saddle-point flow problems: a drop-in for the CPU Krylov loops of
matschiner/navier-stokes-solver (minres.py, bramble_pasciak_cg.py, solvers/bramblepasciak_new.py,
orthonormalization.py) behind the same call signatures.  Compute lives in libnsb200.so (hand-written
sm_100a CUDA behind the C ABI of include/nsb200.h); this package is the host-side mirror of the
reference's operator/vector protocol."""
from . import _lib, la, solvers, simple  # noqa: F401
from .simple import ExtendedInverse, PermutationMatrix, SimpleStepper  # noqa: F401
from .la import (AuxiliarySpacePreconditioner, BaseVector, BlockMatrix, BlockSmoother, BlockVector, CondensedOperator, Context, DiagonalPreconditioner, IdentityMatrix,  # noqa: F401
                 InnerProduct, Norm, Preconditioner, SparseMatrix, default_context, set_default_context)
from .solvers import (BramblePasciakCG, CGSolve, CGSolver, EigenValues_Preconditioner, MinRes,  # noqa: F401
                      ImplicitRungeKuttaMethodWeights, bramble_pasciak_cg, heat_large_step,
                      linear_implicit_runge_kutta_step,
                      orthonormalize)

__version__ = "0.1.0"
