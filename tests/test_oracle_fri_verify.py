"""The oracle's FRI proofs are accepted by a host FRI verifier (oracle/fri_verify.py: the bodies the reference leaves as
placeholders, core/src/fri.ts:544-617 and 899-919), including columns of several degrees, and tampered proofs are rejected.
This pins WHICH alpha folds a circle column that joins a later layer: the alpha that produced that layer (fri.ts:704-711)."""
import numpy as np
import pytest

P = 2147483647


def low_degree_column(orc, log_degree, log_blowup, seed):
    rng = np.random.default_rng(seed)
    log = log_degree + log_blowup
    root_init, root_log = orc.canonic_half_coset(log)
    tw, _ = orc.precompute_twiddles(root_init, root_log)
    return log, [orc.evaluate(rng.integers(0, P, 1 << log_degree, dtype=np.uint32), log, root_init, (root_init, root_log), tw)
                 for _ in range(4)]


def prove(orc, log_degrees, log_last, blow, positions, seed=0):
    cols = [low_degree_column(orc, d, blow, 17 * d + seed) for d in log_degrees]
    ch = orc.Channel(); ch.mix_u64(7)
    state = orc.fri_commit(ch, log_last, blow, cols)
    max_log = cols[0][0]
    dec = orc.fri_decommit_on_queries(state, positions, max_log)
    layers = [(w, (hw, cw), root) for w, (_, hw, cw), root in dec]
    evals = []
    for log, col in cols:
        q = sorted({p >> (max_log - log) for p in positions})
        evals.append([[int(c[p]) for c in col] for p in q])
    return cols, state, layers, evals


@pytest.mark.parametrize("log_degrees,log_last,blow,positions", [
    ((5,), 0, 1, [3, 17, 40]), ((7, 5, 4), 1, 1, [0, 9, 77, 78, 200, 255]), ((6, 3), 0, 2, [1, 100, 254]), ((7, 3), 2, 1, [5, 64, 127, 250])])
def test_oracle_proof_verifies(orc, log_degrees, log_last, blow, positions):
    from oracle.fri_verify import fri_verify
    cols, state, layers, evals = prove(orc, log_degrees, log_last, blow, positions)
    vch = orc.Channel(); vch.mix_u64(7)
    alphas = fri_verify(orc, vch, log_last, blow, [log for log, _ in cols], layers, state["last_poly"], positions, evals)
    assert [tuple(int(v) for v in a) for a in state["alphas"]] == alphas


def test_tampered_proofs_are_rejected(orc):
    from oracle.fri_verify import fri_verify
    positions = [2, 31, 90, 201]
    cols, state, layers, evals = prove(orc, (7, 5), 1, 1, positions, seed=3)
    logs = [log for log, _ in cols]
    def run(layers=layers, evals=evals, last=state["last_poly"], pos=positions):
        ch = orc.Channel(); ch.mix_u64(7)
        return fri_verify(orc, ch, 1, 1, logs, layers, last, pos, evals)
    run()
    bad = [list(map(list, e)) for e in evals]; bad[0][1][2] ^= 1                 # a query evaluation of the big column
    with pytest.raises(ValueError, match="RootMismatch"):
        run(evals=bad)
    bad = [list(map(list, e)) for e in evals]; bad[1][0][0] = (bad[1][0][0] + 1) % P   # of the small column (joins layer 2)
    with pytest.raises(ValueError, match="RootMismatch"):
        run(evals=bad)
    w, dec, root = layers[2]
    bad_layers = list(layers); bad_layers[2] = ([[(v + 1) % P for v in w[0]]] + w[1:], dec, root)   # an inner-layer FRI witness
    with pytest.raises(ValueError, match="RootMismatch"):
        run(layers=bad_layers)
    last = [list(c) for c in state["last_poly"]]; last[0][0] = (last[0][0] + 1) % P
    with pytest.raises(ValueError):                                                # other transcript -> other alphas -> mismatch
        run(last=last)
    with pytest.raises(ValueError, match="InvalidNumFriLayers"):
        run(layers=layers[:-1])


def test_verifier_pins_the_circle_alpha(orc, monkeypatch):
    """A prover that folded the later circle column with the FIRST alpha instead of the joining layer's produces a proof this
    verifier rejects (and vice versa nothing else changes): the choice documented in DESIGN 2 is the one being checked."""
    import oracle as o
    from oracle.fri_verify import fri_verify
    positions = [4, 55, 130]
    real = o.fold_circle_into_line
    seen = []
    def first_alpha_only(dst4, src4, half_init, alpha):
        seen.append(alpha)
        return real(dst4, src4, half_init, seen[0])
    monkeypatch.setattr(o, "fold_circle_into_line", first_alpha_only)
    cols, state, layers, evals = prove(o, (7, 4), 1, 1, positions, seed=5)
    assert len(seen) == 2 and tuple(seen[0]) != tuple(seen[1])
    ch = o.Channel(); ch.mix_u64(7)
    with pytest.raises(ValueError, match="RootMismatch|LastLayerEvaluationsInvalid"):
        fri_verify(o, ch, 1, 1, [log for log, _ in cols], layers, state["last_poly"], positions, evals)
