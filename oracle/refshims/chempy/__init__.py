"""Stand-in for `chempy` (absent).  TEST INFRASTRUCTURE ONLY.  Molar masses are
needed only by the reference's aqueous-chemistry mix-in, executed at backend
construction (dynamics/impl/chemistry_utils.py:130-146); values are irrelevant
to the collision path."""


class Substance:
    def __init__(self, mass):
        self.mass = mass

    @staticmethod
    def from_formula(_formula):
        return Substance(1.0)
