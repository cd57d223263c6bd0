"""Reference property tests for the LDE path, re-run on the oracle (the reference holds no literal
outputs for these -- SURVEY F12).  Ports of packages/core/test/backend/cpu/circle.test.ts:52-141,236-256
plus an independent big-int model (tests/pyfield.py) of the circle-polynomial basis."""
import numpy as np
import pytest

import pyfield as F


def _canonic(orc, log):
    return orc.canonic_half_coset(log)           # (initial_index, half_log)


def _tw(orc, log):
    init, hl = _canonic(orc, log)
    tw, itw = orc.precompute_twiddles(init, hl)
    return (init, hl), tw, itw


def test_generator_and_points(orc):
    assert orc.index_to_point(1) == (2, 1268011823)          # circle.ts:137
    for idx in [0, 1, 5, 1 << 30, (1 << 31) - 1, 123456789]:
        assert orc.index_to_point(idx) == F.to_point(idx)
    # x^2 + y^2 = 1
    x, y = orc.index_to_point(987654321)
    assert (x * x + y * y) % F.P == 1


def test_twiddle_layout(orc):
    # backend/cpu/circle.ts:210-221: per layer, bit-reversed x of first half of the i-times doubled coset; pad 1
    for log in range(0, 8):
        init = orc.half_odds_index(log)
        tw, itw = orc.precompute_twiddles(init, log)
        exp = []
        ci, cl = init, log
        for _ in range(log):
            step = 1 << (31 - cl)
            xs = [F.to_point(ci + k * step)[0] for k in range((1 << cl) // 2)]
            m = len(xs).bit_length() - 1
            exp += [xs[F.bitrev(j, m)] for j in range(len(xs))]
            ci, cl = (2 * ci) % (1 << 31), cl - 1
        exp.append(1)
        assert tw.tolist() == exp
        assert all(int(a) * int(b) % F.P == 1 for a, b in zip(tw, itw))


@pytest.mark.parametrize("log", [1, 2, 3, 4, 5, 6, 9])
def test_evaluate_matches_eval_at_point(orc, log):
    # circle.test.ts:52-97: evaluation.bitReverse()[i] == eval_at_point(domain.iter()[i]).
    # With the TS conjugate() quirk (F9) iter order differs from at(i) in the second half; the TS
    # evaluate compensates only at log 3 (the swap).  We check (a) quirks=False against the correct
    # enumeration at(i) for every size and (b) the TS behaviour at the sizes the reference tests (1..3).
    root, tw, itw = _tw(orc, log)
    coeffs = np.arange(1, (1 << log) + 1, dtype=np.uint32)
    ev = orc.evaluate(coeffs, log, root[0], root, tw, quirks=False)
    nat = orc.bit_reverse(ev)
    for i in range(1 << log):
        x, y = orc.domain_point(root[0], root[1], i)
        got = orc.eval_at_point(coeffs, [x, 0, 0, 0], [y, 0, 0, 0])
        assert got == [int(nat[i]), 0, 0, 0]
        assert F.circle_basis_eval([int(c) for c in coeffs], x, y) == int(nat[i])
    if log <= 3:
        evq = orc.bit_reverse(orc.evaluate(coeffs, log, root[0], root, tw, quirks=True))
        for i in range(1 << log):
            x, y = orc.domain_point(root[0], root[1], i, iter_order=True)
            assert orc.eval_at_point(coeffs, [x, 0, 0, 0], [y, 0, 0, 0]) == [int(evq[i]), 0, 0, 0]


def test_log3_quirk_is_a_swap(orc):
    root, tw, itw = _tw(orc, 3)
    coeffs = np.arange(1, 9, dtype=np.uint32)
    a = orc.evaluate(coeffs, 3, root[0], root, tw, quirks=False)
    b = orc.evaluate(coeffs, 3, root[0], root, tw, quirks=True)
    a2 = a.copy(); a2[[5, 7]] = a2[[7, 5]]
    assert (a2 == b).all() and not (a == b).all()


@pytest.mark.parametrize("log", list(range(1, 13)))
@pytest.mark.parametrize("quirks", [False, True])
def test_roundtrip(orc, log, quirks):
    # circle.test.ts:99-141, 236-256
    root, tw, itw = _tw(orc, log)
    rng = np.random.default_rng(log)
    coeffs = rng.integers(0, F.P, 1 << log, dtype=np.uint32)
    ev = orc.evaluate(coeffs, log, root[0], root, tw, quirks=quirks)
    back = orc.interpolate(ev, root[0], root, itw, quirks=quirks)
    assert (back == coeffs).all()


def test_extend_blowup(orc):
    # evaluate on a larger canonic domain using the big twiddle tree (poly/circle/ops.ts:89-101)
    log, blow = 6, 2
    root, tw, itw = _tw(orc, log + blow)
    rng = np.random.default_rng(1)
    coeffs = rng.integers(0, F.P, 1 << log, dtype=np.uint32)
    ev = orc.evaluate(coeffs, log + blow, root[0], root, tw)
    nat = orc.bit_reverse(ev)
    for i in [0, 1, 17, 100, 255]:
        x, y = orc.domain_point(root[0], root[1], i)
        assert F.circle_basis_eval([int(c) for c in coeffs], x, y) == int(nat[i])
    # interpolate of a smaller domain with the bigger tree (is_doubling_of)
    sm = orc.canonic_half_coset(log)
    ev_s = orc.evaluate(coeffs, log, sm[0], root, tw)
    assert (orc.interpolate(ev_s, sm[0], root, itw) == coeffs).all()


def test_twiddle_mismatch(orc):
    root, tw, itw = _tw(orc, 5)
    with pytest.raises(orc.TwiddleMismatch):
        orc.evaluate(np.ones(32, dtype=np.uint32), 5, root[0] + 2, root, tw)
    with pytest.raises(orc.TwiddleMismatch):     # domain larger than the tree
        big = orc.canonic_half_coset(7)
        orc.evaluate(np.ones(128, dtype=np.uint32), 7, big[0], root, tw)


def test_eval_at_point_small(orc):
    # circle.test.ts:8-50
    x, y = [5, 0, 0, 0], [8, 0, 0, 0]
    assert orc.eval_at_point([1, 3, 2, 4], x, y) == [(1 + 3 * 8 + 2 * 5 + 4 * 40) % F.P, 0, 0, 0]
    assert orc.eval_at_point([1, 2], x, y) == [17, 0, 0, 0]
    assert orc.eval_at_point([1], x, y) == [1, 0, 0, 0]
