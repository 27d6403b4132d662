import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["NSB_SELL_MIN_NNZ"]="0"; os.environ["NSB_SELL_DEBUG"]="1"
import numpy as np
import navier_stokes_solver_b200 as nsb
from oracle.cases import Case
c=Case("chan3d")
A=nsb.SparseMatrix.from_scipy(c.A); print("A",A.format())
B=nsb.SparseMatrix.from_scipy(c.B); print("B",B.format())
BT=B.T; print("BT",BT.format())
