"""CPU: the oracle's wide-Fibonacci proof flow (oracle/stark.py) is self-consistent: the composition polynomial recombined from
its sampled coordinates equals the constraints evaluated at the sampled mask values, every tree decommitment verifies against its
root at the FRI query positions, and an invalid trace fails the out-of-domain check."""
import numpy as np
import pytest

P = 2147483647


@pytest.mark.parametrize("log,n_cols,blow", [(5, 6, 1), (4, 4, 2)])
def test_oracle_prove_wide_fibonacci(orc, log, n_cols, blow):
    from oracle import stark
    pr = stark.prove_wide_fibonacci(log, n_cols, 4, 1, blow, 3)
    assert pr["composition_at_oods"] == pr["constraints_at_oods"]
    assert len(pr["commitments"]) == 3 and [len(t) for t in pr["sampled_values"]] == [0, n_cols, 4]
    for root, logs, (queried, hw, cw) in zip(pr["commitments"], pr["tree_logs"], pr["decommitments"]):
        orc.merkle_verify(root, logs, pr["queries_by_log"], list(queried), hw, list(cw))
    c = orc.Channel(); c.digest = pr["channel_digest"]
    assert len(pr["positions"]) == 3
    # composition polynomial: degree bound 2^log + 1 basis functions (tests/test_oracle_air.py)
    for co in pr["composition_coeffs"]:
        assert not co[(1 << log) + 1:].any()


def test_oracle_prove_rejects_invalid_trace(orc):
    from oracle import stark
    bad = stark.fib_trace(5, 6)
    bad[4][7] = (int(bad[4][7]) + 1) % P
    # the interpolated quotient is still a polynomial of the committed degree, so FRI passes; what catches it is prove()'s check of
    # the composition polynomial against the constraints at the out-of-domain point (ConstraintsNotSatisfied)
    pr = stark.prove_wide_fibonacci(5, 6, 4, 1, 1, 3, trace=bad)
    assert pr["composition_at_oods"] != pr["constraints_at_oods"]
    assert any(co[(1 << 5) + 1:].any() for co in pr["composition_coeffs"])
