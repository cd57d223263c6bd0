"""Host-side FRI bookkeeping (queries.ts, poly/line.ts LinePoly, fri.ts:346-384 positions) and the oracle's FRI flow on
its own: a low-degree column passes the last-layer degree check, a random one does not (core/src/fri.ts:733-748)."""
import numpy as np
import pytest

P = 2147483647


def test_queries_and_line_poly_host_logic():
    from tstwo_b200.fri import LinePoly, Queries, decommitment_positions, get_query_positions_by_log_size
    q = Queries.from_positions([1, 2, 5, 9, 14], 4)
    assert q.fold(1).positions == [0, 1, 2, 4, 7] and q.fold(1).log_domain_size == 3
    assert get_query_positions_by_log_size(q, {4, 2}) == {4: [1, 2, 5, 9, 14], 2: [0, 1, 2, 3]}
    assert decommitment_positions([1, 2, 5], 1) == ([0, 1, 2, 3, 4, 5], [0, 3, 4])
    with pytest.raises(TypeError):
        Queries.from_positions([3, 1], 4)
    with pytest.raises(TypeError):
        Queries.from_positions([16], 4)
    p = LinePoly.from_ordered_coefficients([[1, 0, 0, 0], [2, 0, 0, 0], [3, 0, 0, 0], [4, 0, 0, 0]])
    assert p.coeffs == [[1, 0, 0, 0], [3, 0, 0, 0], [2, 0, 0, 0], [4, 0, 0, 0]] and p.into_ordered_coefficients()[1] == [2, 0, 0, 0]
    with pytest.raises(ValueError):
        LinePoly.new([[0, 0, 0, 0]] * 3)


def test_oracle_fri_flow_degree_check(orc):
    rng = np.random.default_rng(8)
    log_degree, blow = 5, 1
    log = log_degree + blow
    root_init, root_log = orc.canonic_half_coset(log)
    tw, _ = orc.precompute_twiddles(root_init, root_log)
    coords = [orc.evaluate(rng.integers(0, P, 1 << log_degree, dtype=np.uint32), log, root_init, (root_init, root_log), tw) for _ in range(4)]
    ch = orc.Channel()
    state = orc.fri_commit(ch, 1, blow, [(log, coords)])
    assert len(state["inner"]) == log - 1 - 2 and len(state["last_poly"]) == 2 and len(state["alphas"]) == len(state["inner"]) + 1
    assert ch.n_challenges == 1 + len(state["inner"]) + 1
    # the transcript binds the input: another column gives other alphas
    coords2 = [c.copy() for c in coords]; coords2[0][0] ^= 1
    with pytest.raises(ValueError, match="invalid degree"):
        orc.fri_commit(orc.Channel(), 1, blow, [(log, coords2)])
    out = orc.fri_decommit_on_queries(state, [3, 17, 40], log)
    assert len(out) == 1 + len(state["inner"])
    w, (queried, hw, cw), root = out[0]
    pos = orc.fri_decommitment_positions([3, 17, 40], 1)[0]
    orc.merkle_verify(root, [log] * 4, {log: pos}, queried, hw, cw)
    assert len(w) == 3 and len(queried) == 4 * len(pos)


def test_hash_rows_behaves_like_a_list_of_bytes():
    """HashRows (the witness-hash sequence of the one-call FRI decommit) against the list of bytes objects it stands for."""
    from tstwo_b200.backend import HashRows
    a = np.arange(5 * 32, dtype=np.uint8).reshape(5, 32)
    h, ref = HashRows(a), [bytes(r) for r in a]
    assert len(h) == 5 and h[0] == ref[0] and h[4] == ref[4] and list(h) == ref
    assert h == ref and h == HashRows(a.copy()) and not (h == ref[:4]) and not (h == [b"\x00" * 32] * 5)
    assert list(h[1:3]) == ref[1:3] and len(HashRows(a[:0])) == 0 and HashRows(a[:0]) == []
