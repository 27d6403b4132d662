"""``unit_square.GenerateMesh(maxh=..)`` placeholder: the stand-in space of ngsolve/fem.py carries its own nodes."""


class _UnitSquare:
    def GenerateMesh(self, maxh=None, **k):
        return ("unit_square", maxh)


unit_square = _UnitSquare()
