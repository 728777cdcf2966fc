"""The correctly rounded power behind (1 + bounces) ** (1 + falloff) (csrc/vt_pow_cr.h, reference lib.py:450, 465), on the
host: equal to the exact value rounded once (70-digit decimal arithmetic) on every argument tried; glibc's pow -- what the
reference's interpreter calls -- is the same number except on the few arguments it misrounds, which are counted."""
import math
import random
from decimal import Decimal, getcontext


def test_pow_is_correctly_rounded():
    from python_raytracer_b200 import abi
    lib = abi.load()
    getcontext().prec = 70
    rng = random.Random(7)
    absorb = [0.25, 0.5, 1.0, 1.5, 2.0, 0.1, 0.75]
    args = []
    for _ in range(1500):                                  # 1 + sums of material absorptions, like the kernel asks for
        b = 0.0
        for _ in range(rng.randint(1, 6)):
            b += rng.choice(absorb)
        args.append((1.0 + b, 1.0 + rng.choice([0.25, 0.0, 0.5, 1.0, 0.1, 2.0])))
    for _ in range(1500):
        args.append((1.0 + rng.random() * rng.choice([1, 8, 60]), 1.0 + rng.random() * 2.5))
    args += [(1.0, 1.25), (2.0, 2.0), (4.0, 1.5), (1.0 + 2.0 ** -52, 3.0), (1e3, 1.25), (0.5, 1.25), (0.001, 2.5)]
    wrong, glibc_differs = 0, 0
    for x, y in args:
        exact = float(Decimal(x) ** Decimal(y))
        wrong += lib.vt_pow_host(x, y) != exact
        glibc_differs += math.pow(x, y) != exact
    assert wrong == 0
    assert glibc_differs <= len(args) // 200               # glibc: correctly rounded on > 99.5 % of arguments
