"""Seeded random source with NumPy ``RandomState`` (MT19937) bit-compatibility.

Parity: reference ``master/pkg/nprand/nprand.go:1-168`` — the Go master re-implements NumPy's
legacy Mersenne Twister so hyperparameter samples, request ids and trial seeds are reproducible and
checkable against NumPy.  Here NumPy's ``RandomState`` *is* the generator; the bounded-integer
(masked rejection) and unit-interval constructions are spelled out on top of its raw 32-bit
stream so the sequence of draws matches the reference's for the same seed, and the state can be
snapshotted as JSON for master restarts.
"""
import base64
import pickle
from typing import Any, Dict, List

import numpy as np


class State:
    def __init__(self, seed: int) -> None:
        self._rs = np.random.RandomState(int(seed) & 0xFFFFFFFF)

    # -- raw stream ----------------------------------------------------------------------------
    def bits32(self) -> int:
        return int(self._rs.randint(0, 1 << 32, dtype=np.uint32))

    def bits64(self) -> int:
        hi = self.bits32()
        return (hi << 32) | self.bits32()

    def _bits_limit(self, limit: int) -> int:
        """Uniform integer in [0, limit] by masked rejection (numpy ``rk_interval``)."""
        if limit == 0:
            return 0
        mask = limit
        for s in (1, 2, 4, 8, 16, 32):
            mask |= mask >> s
        if limit <= 0xFFFFFFFF:
            while True:
                v = self.bits32() & mask
                if v <= limit:
                    return v
        while True:
            v = self.bits64() & mask
            if v <= limit:
                return v

    # -- derived draws ---------------------------------------------------------------------------
    def int64(self, low: int, high: int) -> int:
        if high <= low:
            raise ValueError(f"nprand int64: high {high} <= low {low}")
        return low + self._bits_limit(high - low - 1)

    def intn(self, n: int) -> int:
        if n <= 0:
            raise ValueError(f"nprand intn: n {n} <= 0")
        return self._bits_limit(n - 1)

    def unit_interval(self) -> float:
        a = self.bits32() >> 5
        b = self.bits32() >> 6
        return (a * 67108864.0 + b) / 9007199254740992.0

    def uniform(self, low: float, high: float) -> float:
        if high <= low:
            raise ValueError(f"nprand uniform: high {high} <= low {low}")
        return low + (high - low) * self.unit_interval()

    def read(self, n: int) -> bytes:
        out = bytearray()
        val, left = 0, 0
        for _ in range(n):
            if left == 0:
                val, left = self.bits32(), 4
            out.append(val & 0xFF)
            val >>= 8
            left -= 1
        return bytes(out)

    def uuid4(self) -> str:
        b = bytearray(self.read(16))
        b[6] = (b[6] & 0x0F) | 0x40
        b[8] = (b[8] & 0x3F) | 0x80
        h = b.hex()
        return f"{h[0:8]}-{h[8:12]}-{h[12:16]}-{h[16:20]}-{h[20:32]}"

    # -- snapshot ----------------------------------------------------------------------------------
    def snapshot(self) -> Dict[str, Any]:
        name, key, pos, has_gauss, gauss = self._rs.get_state()
        return {"key": [int(k) for k in key], "pos": int(pos)}

    @classmethod
    def restore(cls, snap: Dict[str, Any]) -> "State":
        s = cls(0)
        s._rs.set_state(("MT19937", np.asarray(snap["key"], dtype=np.uint32), int(snap["pos"]), 0, 0.0))
        return s
