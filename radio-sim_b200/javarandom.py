"""java.util.Random, bit for bit, for the host side of the drop-in.

The reference's stochastic reception draws from ONE `java.util.Random` owned by the simulator
(`Simulator.java:64,83-94`: `simulator.getRandom().nextDouble()` in `UDGMRadioMedium.java:89,106`,
`N2NRadioMedium.java:45,62`).  The device never draws: the tick gets the next stretch of that sequence as
host-supplied uniforms (`rmb_tick(..., uniforms, n_uniforms, ...)`) and reports how many it consumed
(`rmb_result.n_uniforms_used`); the adapter then advances the generator by exactly that many draws.  Hence the
two operations a plain generator lacks: `peek_doubles(n)` (look ahead without consuming) and `skip_doubles(k)`.

Algorithm (JDK specification of java.util.Random): 48-bit LCG, seed' = seed * 0x5DEECE66D + 0xB mod 2^48,
`next(bits) = seed' >>> (48 - bits)`, `nextDouble = ((next(26) << 27) + next(27)) * 2^-53`.  Runs of states are
produced by block doubling (the k-step map of an affine LCG is affine), so a million draws cost a handful of
numpy operations.  The synthetic deployments of bench.py / tests are drawn from this class too, so that a Java
harness regenerates them bit for bit (`java/tools/GenGolden.java`).
"""
import numpy as np

_MULT = 0x5DEECE66D
_ADD = 0xB
_MASK = (1 << 48) - 1


def _states(seed, count):
    """The `count` LCG states that follow `seed` (uint64 array), by block doubling."""
    out = np.empty(max(count, 1), dtype=np.uint64)
    out[0] = (seed * _MULT + _ADD) & _MASK
    a, c, m = _MULT, _ADD, 1                     # the m-step map s -> a*s + c (mod 2^48)
    mask = np.uint64(_MASK)
    while m < count:
        k = min(m, count - m)
        with np.errstate(over="ignore"):
            out[m:m + k] = (out[:k] * np.uint64(a) + np.uint64(c)) & mask
        a, c = (a * a) & _MASK, (a * c + c) & _MASK
        m *= 2
    return out[:count]


def _advance(seed, steps):
    """seed after `steps` LCG steps, O(log steps)."""
    a, c = _MULT, _ADD
    ra, rc = 1, 0
    while steps:
        if steps & 1:
            ra, rc = (a * ra) & _MASK, (a * rc + c) & _MASK
        a, c = (a * a) & _MASK, (a * c + c) & _MASK
        steps >>= 1
    return (ra * seed + rc) & _MASK


class JavaRandom:
    """`new java.util.Random(seed)`."""

    def __init__(self, seed):
        self.seed = (int(seed) ^ _MULT) & _MASK          # Random.initialScramble

    def _next(self, bits):
        self.seed = (self.seed * _MULT + _ADD) & _MASK
        v = self.seed >> (48 - bits)
        return v - (1 << bits) if bits == 32 and v >= (1 << 31) else v

    def next_double(self):
        return ((self._next(26) << 27) + self._next(27)) * (1.0 / (1 << 53))

    def next_int(self, bound=None):
        """Random.nextInt() / nextInt(bound), including the rejection loop."""
        if bound is None:
            return self._next(32)
        if bound <= 0:
            raise ValueError("bound must be positive")
        if bound & (bound - 1) == 0:
            return (bound * self._next(31)) >> 31
        while True:
            bits = self._next(31)
            val = bits % bound
            if bits - val + (bound - 1) < (1 << 31):     # Java: int overflow => retry
                return val

    def peek_doubles(self, n):
        """The next n nextDouble() values without consuming them."""
        if n <= 0:
            return np.zeros(0, dtype=np.float64)
        s = _states(self.seed, 2 * n)
        hi, lo = s[0::2] >> np.uint64(22), s[1::2] >> np.uint64(21)
        return ((hi << np.uint64(27)) + lo).astype(np.float64) * (1.0 / (1 << 53))

    def skip_doubles(self, k):
        self.seed = _advance(self.seed, 2 * int(k))

    def next_doubles(self, n):
        u = self.peek_doubles(n)
        self.skip_doubles(n)
        return u

    def sample_distinct(self, n, k):
        """First k entries of a Fisher-Yates shuffle of 0..n-1 (for i < k: swap(i, i + nextInt(n - i))): k distinct
        indices in k draws.  Java: the same loop over an int[] (GenGolden.java `sampleDistinct`)."""
        moved = {}
        out = np.empty(k, dtype=np.int32)
        for i in range(k):
            j = i + self.next_int(n - i)
            vi, vj = moved.get(i, i), moved.get(j, j)
            moved[i], moved[j] = vj, vi
            out[i] = vj
        return out
