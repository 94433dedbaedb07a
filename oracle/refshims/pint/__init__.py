"""Stand-in for the `pint` package (absent from this image, no network).

TEST INFRASTRUCTURE ONLY: lets the read-only reference at /root/reference be
imported in the build container so that golden vectors can be generated from
its Numba backend (see oracle/gen_golden.py).  Only what
PySDM/physics/constants.py:17-19 and physics/impl/fake_unit_registry.py:8-58
touch is provided: unit name -> SI base magnitude.
"""
_PREFIX_LONG = {"nano": 1e-9, "micro": 1e-6, "milli": 1e-3, "centi": 1e-2,
                "deci": 1e-1, "": 1.0, "hecto": 1e2, "kilo": 1e3}
_PREFIX_SHORT = {"n": 1e-9, "u": 1e-6, "m": 1e-3, "c": 1e-2, "d": 1e-1,
                 "": 1.0, "h": 1e2, "k": 1e3}
_UNIT_LONG = {"bar": 1e5, "metre": 1.0, "meter": 1.0, "gram": 1e-3,
              "hertz": 1.0, "mole": 1.0, "joule": 1.0, "kelvin": 1.0,
              "second": 1.0, "minute": 60.0, "pascal": 1.0, "litre": 1e-3,
              "liter": 1e-3, "hour": 3600.0, "newton": 1.0, "watt": 1.0}
_UNIT_SHORT = {"b": 1e-28, "m": 1.0, "g": 1e-3, "Hz": 1.0, "mol": 1.0,
               "J": 1.0, "K": 1.0, "s": 1.0, "min": 60.0, "day": 86400.0,
               "Pa": 1.0, "l": 1e-3, "h": 3600.0, "bar": 1e5, "N": 1.0,
               "W": 1.0}


class Quantity:
    def __init__(self, magnitude):
        self.magnitude = float(magnitude)

    def to_base_units(self):
        return self

    def __rmul__(self, other):
        return Quantity(other * self.magnitude)

    __mul__ = __rmul__


class Unit(Quantity):
    pass


def _resolve(name):
    for plural in (False, True):
        base = name[:-1] if plural and name.endswith("s") else name
        if plural and not name.endswith("s"):
            continue
        for pfx, pm in _PREFIX_LONG.items():
            if base.startswith(pfx) and base[len(pfx):] in _UNIT_LONG:
                return pm * _UNIT_LONG[base[len(pfx):]]
    # short symbols: try the bare symbol first (e.g. "min", "mol", "m")
    if name in _UNIT_SHORT:
        return _UNIT_SHORT[name]
    for pfx, pm in _PREFIX_SHORT.items():
        if pfx and name.startswith(pfx) and name[len(pfx):] in _UNIT_SHORT:
            return pm * _UNIT_SHORT[name[len(pfx):]]
    raise AttributeError(name)


class UnitRegistry:
    dimensionless = 1.0
    Quantity = Quantity
    Unit = Unit

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return Unit(_resolve(name))

    def parse_expression(self, _):
        raise NotImplementedError("pint stand-in: products are out of scope")
