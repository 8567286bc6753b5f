"""Host-side restatements of the JVM random streams that leak into Spark's results.

* XORShiftRandom  - org.apache.spark.util.random.XORShiftRandom (validated against
  result.txt:105-125): state = hashSeed(seed) over a 64-BYTE big-endian buffer, MurmurHash3.
* JavaRandom      - java.util.Random LCG (feature-subset seeds in RandomForest.run).
"""
import struct

_M32 = 0xFFFFFFFF
_M64 = 0xFFFFFFFFFFFFFFFF


def _rotl32(x, r):
    return ((x << r) | (x >> (32 - r))) & _M32


def murmur3_bytes_hash(data, seed):
    """scala.util.hashing.MurmurHash3.bytesHash == MurmurHash3_x86_32."""
    h = seed & _M32
    n = len(data)
    for i in range(0, n - n % 4, 4):
        k = struct.unpack_from("<I", data, i)[0]
        k = (k * 0xCC9E2D51) & _M32
        k = _rotl32(k, 15)
        k = (k * 0x1B873593) & _M32
        h ^= k
        h = _rotl32(h, 13)
        h = (h * 5 + 0xE6546B64) & _M32
    rem = n % 4
    if rem:
        k = 0
        tail = data[n - rem:]
        if rem == 3:
            k ^= tail[2] << 16
        if rem >= 2:
            k ^= tail[1] << 8
        k ^= tail[0]
        k = (k * 0xCC9E2D51) & _M32
        k = _rotl32(k, 15)
        k = (k * 0x1B873593) & _M32
        h ^= k
    h ^= n
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & _M32
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & _M32
    h ^= h >> 16
    return h


def xorshift_hash_seed(seed):
    buf = struct.pack(">q", _to_signed64(seed)) + b"\x00" * 56     # ByteBuffer.allocate(Long.SIZE = 64)
    lo = murmur3_bytes_hash(buf, 0x3C074A61)
    hi = murmur3_bytes_hash(buf, lo)
    return ((hi << 32) | lo) & _M64


def _to_signed64(x):
    x &= _M64
    return x - (1 << 64) if x >= (1 << 63) else x


def _to_signed32(x):
    x &= _M32
    return x - (1 << 32) if x >= (1 << 31) else x


class XORShiftRandom:
    def __init__(self, seed):
        self.s = xorshift_hash_seed(seed)

    def _next(self, bits):
        s = self.s
        s ^= (s << 21) & _M64
        s ^= s >> 35
        s ^= (s << 4) & _M64
        self.s = s
        return s & ((1 << bits) - 1)

    def nextDouble(self):
        return ((self._next(26) << 27) + self._next(27)) * (1.0 / (1 << 53))

    def nextInt(self):
        return _to_signed32(self._next(32))

    def nextLong(self):
        return _to_signed64((self._next(32) << 32) + _to_signed32(self._next(32)))


class JavaRandom:
    def __init__(self, seed):
        self.s = (seed ^ 0x5DEECE66D) & ((1 << 48) - 1)

    def _next(self, bits):
        self.s = (self.s * 0x5DEECE66D + 0xB) & ((1 << 48) - 1)
        return _to_signed32(self.s >> (48 - bits))

    def nextInt(self):
        return self._next(32)

    def nextLong(self):
        return _to_signed64((self._next(32) << 32) + self._next(32))

    def nextDouble(self):
        return ((self._next(26) & 0x3FFFFFF) * (1 << 27) + (self._next(27) & 0x7FFFFFF)) * (1.0 / (1 << 53))
