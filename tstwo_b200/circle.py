"""Host-side index arithmetic of the circle group: mirrors CirclePointIndex / Coset (packages/core/src/circle.ts:152-291),
CanonicCoset (poly/circle/canonic.ts:24-130) and CircleDomain (poly/circle/domain.ts:12-148) of the reference.
Only indices live on the host; points and twiddles are computed on the device."""
from __future__ import annotations

from dataclasses import dataclass

M31_CIRCLE_LOG_ORDER = 31
_MASK = (1 << 31) - 1
P = (1 << 31) - 1
M31_CIRCLE_GEN = (2, 1268011823)  # circle.ts:137


@dataclass(frozen=True)
class CirclePointIndex:
    value: int

    @staticmethod
    def zero(): return CirclePointIndex(0)

    @staticmethod
    def generator(): return CirclePointIndex(1)

    @staticmethod
    def subgroup_gen(log_size: int):
        if log_size > M31_CIRCLE_LOG_ORDER:
            raise ValueError("log_size too large")
        return CirclePointIndex((1 << (M31_CIRCLE_LOG_ORDER - log_size)) & _MASK)

    def reduce(self): return CirclePointIndex(self.value & _MASK)
    def add(self, rhs): return CirclePointIndex((self.value + rhs.value) & _MASK)
    def sub(self, rhs): return CirclePointIndex((self.value - rhs.value) & _MASK)
    def mul(self, k: int): return CirclePointIndex((self.value * k) & _MASK)
    def neg(self): return CirclePointIndex((-self.value) & _MASK)

    def half(self):
        if self.value & 1:
            raise ValueError("not even")
        return CirclePointIndex(self.value >> 1)

    def to_point(self):
        """(x, y) of index * G by double-and-add (circle.ts:172-174)."""
        def add(p, q):
            return ((p[0] * q[0] - p[1] * q[1]) % P, (p[0] * q[1] + p[1] * q[0]) % P)
        r, cur, s = (1, 0), M31_CIRCLE_GEN, self.value & _MASK
        while s:
            if s & 1:
                r = add(r, cur)
            cur = add(cur, cur)
            s >>= 1
        return r


@dataclass(frozen=True)
class Coset:
    initial_index: CirclePointIndex
    log_size: int

    @staticmethod
    def new(initial_index: CirclePointIndex, log_size: int):
        if log_size > M31_CIRCLE_LOG_ORDER:
            raise ValueError("log_size too large")
        return Coset(initial_index.reduce(), log_size)

    @staticmethod
    def subgroup(log_size): return Coset.new(CirclePointIndex.zero(), log_size)

    @staticmethod
    def odds(log_size): return Coset.new(CirclePointIndex.subgroup_gen(log_size + 1), log_size)

    @staticmethod
    def half_odds(log_size): return Coset.new(CirclePointIndex.subgroup_gen(log_size + 2), log_size)

    @property
    def step_size(self): return CirclePointIndex.subgroup_gen(self.log_size)

    def size(self): return 1 << self.log_size
    def logSize(self): return self.log_size

    def double(self):
        if self.log_size <= 0:
            raise ValueError("log_size must be >0 to double")
        return Coset.new(self.initial_index.mul(2), self.log_size - 1)

    def repeated_double(self, n):
        c = self
        for _ in range(n):
            c = c.double()
        return c

    def is_doubling_of(self, other) -> bool:
        return self.log_size <= other.log_size and self == other.repeated_double(other.log_size - self.log_size)

    def index_at(self, i: int): return self.initial_index.add(self.step_size.mul(i))
    def at(self, i: int): return self.index_at(i).to_point()
    def shift(self, s: CirclePointIndex): return Coset.new(self.initial_index.add(s), self.log_size)
    def conjugate(self): return Coset.new(self.initial_index.neg(), self.log_size)


@dataclass(frozen=True)
class CircleDomain:
    halfCoset: Coset

    @staticmethod
    def new(half_coset: Coset): return CircleDomain(half_coset)

    def logSize(self): return self.halfCoset.log_size + 1
    log_size = logSize
    def size(self): return 1 << self.logSize()

    def indexAt(self, i: int):
        if i < 0:
            raise ValueError("i must be a non-negative integer")
        half = self.halfCoset.size()
        return self.halfCoset.index_at(i) if i < half else self.halfCoset.index_at(i - half).neg()
    index_at = indexAt

    def at(self, i: int): return self.indexAt(i).to_point()

    def isCanonic(self) -> bool:
        return self.halfCoset.initial_index.mul(4).value == self.halfCoset.step_size.value
    is_canonic = isCanonic


@dataclass(frozen=True)
class CanonicCoset:
    coset: Coset

    @staticmethod
    def new(log_size: int):
        if not isinstance(log_size, int) or log_size <= 0:
            raise ValueError("log_size must be a positive integer")
        return CanonicCoset(Coset.odds(log_size))

    def log_size(self): return self.coset.log_size
    logSize = log_size
    def size(self): return self.coset.size()
    def half_coset(self): return Coset.half_odds(self.log_size() - 1)
    halfCoset = half_coset
    def circle_domain(self): return CircleDomain.new(self.half_coset())
    circleDomain = circle_domain


@dataclass(frozen=True)
class LineDomain:
    """poly/line.ts:18-115 (index form)."""
    _coset: Coset

    @staticmethod
    def new(coset: Coset): return LineDomain(coset)
    def coset(self): return self._coset
    def size(self): return self._coset.size()
    def logSize(self): return self._coset.log_size
    log_size = logSize
    def double(self): return LineDomain(self._coset.double())
    def at(self, i: int): return self._coset.at(i)[0]
