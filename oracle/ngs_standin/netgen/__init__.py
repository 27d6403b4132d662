"""netgen stand-in (test infrastructure): heat.py:5,10 import ``netgen.geom2d.unit_square`` and ``netgen.gui``."""
