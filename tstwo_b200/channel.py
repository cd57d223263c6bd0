"""Fiat-Shamir channel of the commitment scheme: host-side mirror of TStwo's Blake2sChannel
(packages/core/src/channel/blake2.ts:25-224), ChannelTime (channel/index.ts:28-109) and Blake2sMerkleChannel
(vcs/blake2_merkle.ts:27-32).  A channel hashes a few dozen bytes per protocol step; like the reference it runs on
the host (hashlib's BLAKE2s), the GPU only ever sees the 32-byte roots it mixes and the field elements it draws.
Method names, argument checks and error types follow the reference so its channel tests read the same here."""
import hashlib
from typing import List, Sequence

P = 2147483647
BLAKE_BYTES_PER_HASH = 32
FELTS_PER_HASH = 8
SECURE_EXTENSION_DEGREE = 4


class ChannelTime:
    """Immutable (n_challenges, n_sent) pair, channel/index.ts:28-89."""

    def __init__(self, n_challenges: int = 0, n_sent: int = 0):
        self.n_challenges, self.n_sent = n_challenges, n_sent

    def __eq__(self, other):
        return (self.n_challenges, self.n_sent) == (other.n_challenges, other.n_sent)

    def __repr__(self):
        return f"ChannelTime(n_challenges={self.n_challenges}, n_sent={self.n_sent})"


def _blake2s(data: bytes) -> bytes:
    return hashlib.blake2s(data, digest_size=32).digest()


class Blake2sChannel:
    BYTES_PER_HASH = BLAKE_BYTES_PER_HASH

    def __init__(self, digest: bytes = bytes(32), n_challenges: int = 0, n_sent: int = 0, base_queue: Sequence[int] = ()):
        self._digest = bytes(digest)
        self._n_challenges, self._n_sent = n_challenges, n_sent
        self._base_queue: List[int] = list(base_queue)

    @staticmethod
    def create() -> "Blake2sChannel":
        return Blake2sChannel()

    @staticmethod
    def fromState(digest: bytes, channel_time: ChannelTime, base_queue: Sequence[int] = ()) -> "Blake2sChannel":
        return Blake2sChannel(digest, channel_time.n_challenges, channel_time.n_sent, base_queue)

    def clone(self) -> "Blake2sChannel":
        return Blake2sChannel(self._digest, self._n_challenges, self._n_sent, self._base_queue)

    def digest(self) -> bytes: return self._digest
    def digestBytes(self) -> bytes: return self._digest
    def getChannelTime(self) -> ChannelTime: return ChannelTime(self._n_challenges, self._n_sent)

    def updateDigest(self, new_digest: bytes) -> None:
        """blake2.ts:75-79: replaces the digest; a new challenge resets the sent counter."""
        if len(new_digest) != 32:
            raise ValueError("Blake2sHash must be 32 bytes")
        self._digest = bytes(new_digest)
        self._n_challenges += 1
        self._n_sent = 0

    def trailing_zeros(self) -> int:
        v = int.from_bytes(self._digest[:16], "little")
        return 128 if v == 0 else (v & -v).bit_length() - 1

    def mix_felts(self, felts: Sequence[Sequence[int]]) -> None:
        """felts: QM31 values as 4 M31 coordinates each (to_m31_array order, qm31.ts:139-141)."""
        buf = bytearray()
        for f in felts:
            if len(f) != SECURE_EXTENSION_DEGREE:
                raise TypeError("mix_felts takes secure-field elements (4 coordinates)")
            for w in f:
                buf += int(w).to_bytes(4, "little")
        self.updateDigest(_blake2s(self._digest + bytes(buf)))

    def mix_u32s(self, data: Sequence[int]) -> None:
        for w in data:
            if not isinstance(w, int) or isinstance(w, bool) or w < 0 or w > 0xFFFFFFFF:
                raise TypeError(f"Invalid u32 value: {w}")
        self.updateDigest(_blake2s(self._digest + b"".join(w.to_bytes(4, "little") for w in data)))

    def mix_u64(self, value: int) -> None:
        if not isinstance(value, int) or isinstance(value, bool) or value < 0 or value >> 64:
            raise TypeError(f"Invalid u64 value: {value}")
        self.mix_u32s([value & 0xFFFFFFFF, value >> 32])

    def draw_random_bytes(self) -> bytes:
        out = _blake2s(self._digest + self._n_sent.to_bytes(4, "little") + bytes(BLAKE_BYTES_PER_HASH - 4))
        self._n_sent += 1
        return out

    def _draw_base_felts(self) -> List[int]:
        while True:   # retry probability ~2^-28 per round
            b = self.draw_random_bytes()
            u = [int.from_bytes(b[4 * i:4 * i + 4], "little") for i in range(FELTS_PER_HASH)]
            if all(x < 2 * P for x in u):
                return [x - P if x >= P else x for x in u]

    def draw_felt(self) -> List[int]:
        while len(self._base_queue) < SECURE_EXTENSION_DEGREE:
            self._base_queue += self._draw_base_felts()
        out = self._base_queue[:SECURE_EXTENSION_DEGREE]
        del self._base_queue[:SECURE_EXTENSION_DEGREE]
        return out

    def draw_felts(self, n_felts: int) -> List[List[int]]:
        if not isinstance(n_felts, int) or isinstance(n_felts, bool) or n_felts < 0:
            raise TypeError(f"Invalid felt count: {n_felts}")
        out, queue = [], []
        for _ in range(n_felts):
            while len(queue) < SECURE_EXTENSION_DEGREE:
                queue += self._draw_base_felts()
            out.append(queue[:SECURE_EXTENSION_DEGREE])
            del queue[:SECURE_EXTENSION_DEGREE]
        return out


class Blake2sMerkleChannel:
    """vcs/blake2_merkle.ts:27-32: the root enters the transcript as H(digest || root)."""

    @staticmethod
    def mix_root(channel: Blake2sChannel, root: bytes) -> None:
        channel.updateDigest(_blake2s(channel.digest() + bytes(root)))
