"""Blake2sChannel: the product's host mirror (hashlib) and the oracle's restatement (own BLAKE2s) against the digests the
reference's tests carry over from Stwo, and against each other on random transcripts.  Mirrors
packages/core/test/channel/channel_exact_rust_tests.test.ts:14-118 and blake2_channel.test.ts."""
import random

import pytest

MIX_U64 = bytes([0xbc, 0x9e, 0x3f, 0xc1, 0xd2, 0x4e, 0x88, 0x97, 0x95, 0x6d, 0x33, 0x59, 0x32, 0x73, 0x97, 0x24, 0x9d, 0x6b, 0xca, 0xcd,
                 0x22, 0x4d, 0x92, 0x74, 0x04, 0xe7, 0xba, 0x4a, 0x77, 0xdc, 0x6e, 0xce])   # channel_exact_rust_tests.test.ts:97-101
MIX_U32S = bytes([0x70, 0x91, 0x76, 0x83, 0x57, 0xbb, 0x1b, 0xb3, 0x34, 0x6f, 0xda, 0xb6, 0xb3, 0x57, 0xd7, 0xfa, 0x46, 0xb8, 0xfb, 0xe3,
                  0x2c, 0x2e, 0x43, 0x24, 0xa0, 0xff, 0xc2, 0x94, 0xcb, 0xf9, 0xa1, 0xc7])  # :112-116
P = 2147483647


def channels(orc):
    from tstwo_b200.channel import Blake2sChannel
    return Blake2sChannel.create(), orc.Channel()


def test_rust_digests(orc):
    prod, ref = channels(orc)
    prod.mix_u64(0x1111222233334444); ref.mix_u64(0x1111222233334444)
    assert prod.digest() == MIX_U64 and ref.digest == MIX_U64
    prod2, ref2 = channels(orc)
    prod2.mix_u32s([0x33334444, 0x11112222]); ref2.mix_u32s([0x33334444, 0x11112222])
    assert prod2.digest() == MIX_U64 and ref2.digest == MIX_U64
    prod3, ref3 = channels(orc)
    prod3.mix_u32s(list(range(1, 10))); ref3.mix_u32s(list(range(1, 10)))
    assert prod3.digest() == MIX_U32S and ref3.digest == MIX_U32S


def test_channel_time(orc):
    prod, _ = channels(orc)
    t = prod.getChannelTime()
    assert (t.n_challenges, t.n_sent) == (0, 0)
    prod.draw_random_bytes()
    assert (prod.getChannelTime().n_challenges, prod.getChannelTime().n_sent) == (0, 1)
    prod.draw_felts(9)        # 9 secure felts = 36 base felts = 5 hashes
    assert (prod.getChannelTime().n_challenges, prod.getChannelTime().n_sent) == (0, 6)
    prod.mix_u64(7)
    assert (prod.getChannelTime().n_challenges, prod.getChannelTime().n_sent) == (1, 0)


def test_draws_differ_and_are_reduced(orc):
    prod, _ = channels(orc)
    a, b = prod.draw_random_bytes(), prod.draw_random_bytes()
    assert a != b
    felts = prod.draw_felts(5)
    assert len({tuple(f) for f in felts}) == 5 and all(0 <= w < P for f in felts for w in f)
    assert prod.draw_felt() != prod.draw_felt()


def test_argument_checks(orc):
    prod, _ = channels(orc)
    for bad in ([-1], [0x100000000], [1.5]):
        with pytest.raises(TypeError):
            prod.mix_u32s(bad)
    for bad in (-1, 1.5, 1 << 64):
        with pytest.raises(TypeError):
            prod.mix_u64(bad)
    with pytest.raises(TypeError):
        prod.draw_felts(-1)


def test_random_transcripts_agree(orc):
    from tstwo_b200.channel import Blake2sMerkleChannel
    rnd = random.Random(5)
    for _ in range(20):
        prod, ref = channels(orc)
        for _ in range(30):
            op = rnd.randrange(7)
            if op == 0:
                root = bytes(rnd.randrange(256) for _ in range(32))
                Blake2sMerkleChannel.mix_root(prod, root); ref.mix_root(root)
            elif op == 1:
                f = [[rnd.randrange(P) for _ in range(4)] for _ in range(rnd.randrange(4))]
                prod.mix_felts(f); ref.mix_felts(f)
            elif op == 2:
                d = [rnd.randrange(1 << 32) for _ in range(rnd.randrange(6))]
                prod.mix_u32s(d); ref.mix_u32s(d)
            elif op == 3:
                v = rnd.randrange(1 << 64)
                prod.mix_u64(v); ref.mix_u64(v)
            elif op == 4:
                assert prod.draw_felt() == ref.draw_felt()
            elif op == 5:
                n = rnd.randrange(4)
                assert prod.draw_felts(n) == ref.draw_felts(n)
            else:
                assert prod.draw_random_bytes() == ref.draw_random_bytes()
            assert prod.digest() == ref.digest and prod.trailing_zeros() == ref.trailing_zeros()
            t = prod.getChannelTime()
            assert (t.n_challenges, t.n_sent) == (ref.n_challenges, ref.n_sent)
        c = prod.clone()
        assert c.draw_felt() == prod.draw_felt() and c.digest() == prod.digest()
