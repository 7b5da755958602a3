"""Cell / collision type codes (reference: src/LBM/utils/types.py:7-15, 61-107).

Only FLUID (1) and OBSTACLE (2) are ever tested by a solver op.  Every KBC
predicate of the reference masks with the shared high bit 0b1000_0000, so all
of them are true for every KBC_* value and false for LBGK / None; that quirk is
part of the behaviour and is kept.
"""
from enum import Enum


class CellType(Enum):
    NOTHING = 0
    FLUID = 1
    OBSTACLE = 2
    EMPTY = 4
    INFLOW = 8
    OUTFLOW = 16
    OPEN = 32
    STICK = 64

    def __int__(self):
        return self.value


def _masked(mask):
    def pred(code) -> bool:
        return code is not None and (int(code) & mask) > 0
    return staticmethod(pred)


class KBCType(Enum):
    LBGK = 0
    KBC_A = 0b10000101
    KBC_B = 0b10000110
    KBC_C = 0b10001001
    KBC_D = 0b10001010

    def __int__(self):
        return self.value

    is_KBC = _masked(0b10000000)
    is_KBC_AC = _masked(0b10000001)
    is_KBC_BD = _masked(0b10000010)
    is_KBC_AB = _masked(0b10000100)
    is_KBC_CD = _masked(0b10001000)
