"""Drop-in for /root/reference/code/siamese3.py (sigmoid-1 head, code/siamese3.py:14-28, predict :102-103)."""
from .siamese import SiameseNetwork as _Base


class SiameseNetwork(_Base):
    _C = 1
    _ACT = "sigmoid"


class SmallRes(SiameseNetwork):
    """512/64/1 head of code/siamese3.py:134-138 (same widths as the base class)."""
