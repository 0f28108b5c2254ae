"""Half-open key ranges with the reference's even division.

ref: src/util/range.h:10-131.  `even_divide` reproduces the long-double arithmetic of
Range<T>::evenDivide (range.h:99-107) so server shards and feature blocks land on the same
boundaries as the reference for any uint64 range.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

U64_MAX = (1 << 64) - 1
_LD = np.longdouble
_HAS_X87 = np.finfo(_LD).nmant >= 63  # x86-64: 80-bit extended, 64-bit significand


def _to_ld(x: int):
    hi, lo = divmod(int(x), 1 << 32)
    return _LD(hi) * _LD(1 << 32) + _LD(lo)


def _from_ld(v) -> int:
    v = np.floor(v)
    hi = int(np.floor(v / _LD(1 << 32)))
    lo = int(v - _LD(hi) * _LD(1 << 32))
    return hi * (1 << 32) + lo


@dataclass(frozen=True)
class Range:
    begin: int
    end: int

    @staticmethod
    def all() -> "Range":
        """[0, 2^64-1) — the reference's Range<Key>::all() (range.h:88-92)."""
        return Range(0, U64_MAX)

    def valid(self) -> bool:
        return self.end >= self.begin

    def empty(self) -> bool:
        return self.begin >= self.end

    def size(self) -> int:
        return max(0, self.end - self.begin)

    def contains(self, x: int) -> bool:
        return self.begin <= x < self.end

    def intersection(self, o: "Range") -> "Range":
        return Range(max(self.begin, o.begin), min(self.end, o.end))

    def even_divide(self, n: int, i: int) -> "Range":
        assert self.valid() and n > 0 and 0 <= i < n
        if _HAS_X87:
            itv = _to_ld(self.end - self.begin) / _LD(n)
            b = _from_ld(_to_ld(self.begin) + itv * _LD(i))
            e = _from_ld(_to_ld(self.begin) + itv * _LD(i + 1))
        else:  # exact rational fallback on platforms without 80-bit long double
            span = self.end - self.begin
            b = self.begin + span * i // n
            e = self.begin + span * (i + 1) // n
        if i == n - 1:
            # long-double truncation can land one short of `end` (e.g. [3,1003)/7 -> 1002), which
            # in the reference silently drops the largest key from the last block; we close the
            # cover instead (documented deviation)
            e = self.end
        return Range(b, e)

    def __add__(self, k: int) -> "Range":
        return Range(self.begin + k, self.end + k)

    def __sub__(self, k: int) -> "Range":
        return Range(self.begin - k, self.end - k)

    def __str__(self) -> str:
        return f"[{self.begin},{self.end})"
