"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the wide-Fibonacci proof flow, built from the
oracle's own pieces, against which tests compare tstwo_b200.prover.prove field by field.

Follows prover/index.ts:582-666 (prove), with the components the TS code simulates replaced by what its comments name, as in
Stwo: composition polynomial = constraint quotients on the evaluation domain (oracle.widefib_constraints) interpolated; last
tree = its 4 coordinate polynomials; out-of-domain point from the channel (circle.ts:126-133); prove_values (pcs/prover.ts:85-156
Rust text): sample, mix, DEEP quotients with Stwo's denominators, FRI commit, proof of work, queries, decommit.
parity unpinned: the reference's prove() is a stub and no proof golden exists; this pins the product to one restated flow."""
import numpy as np

import oracle as orc

P = 2147483647
ONE = [1, 0, 0, 0]


def _qm(op, a, b=None): return orc.qm31_op(op, a, b)     # 0 add, 1 sub, 2 mul, 3 neg, 4 inv, 6 mul by M31


def fib_trace(log_n_rows: int, n_cols: int) -> np.ndarray:
    """generateTestTrace, examples/fibonacci.ts:245-258, rows in natural order."""
    n = 1 << log_n_rows
    cols = [np.ones(n, dtype=np.uint64), np.arange(n, dtype=np.uint64)]
    for k in range(2, n_cols):
        cols.append((cols[k - 2] ** 2 % P + cols[k - 1] ** 2 % P) % P)
    return np.stack(cols).astype(np.uint32)


def random_point(channel):
    """CirclePoint.get_random_point, circle.ts:126-133."""
    t = channel.draw_felt()
    t2 = _qm(2, t, t)
    inv = _qm(4, _qm(0, t2, ONE))
    return tuple(_qm(2, _qm(1, ONE, t2), inv)), tuple(_qm(2, _qm(0, t, t), inv))


def coset_vanishing_secure(coset_initial_index: int, coset_log_size: int, p):
    """constraints.ts:32-47 at a secure-field point."""
    ix, iy = orc.index_to_point(coset_initial_index)
    hx, hy = orc.index_to_point(orc.subgroup_gen(coset_log_size) >> 1)
    sx = _qm(0, _qm(6, p[0], [ix]), _qm(6, p[1], [iy]))
    sy = _qm(1, _qm(6, p[1], [ix]), _qm(6, p[0], [iy]))
    x = _qm(1, _qm(6, sx, [hx]), _qm(6, sy, [hy]))
    for _ in range(1, coset_log_size):
        x = _qm(1, _qm(6, _qm(2, x, x), [2]), ONE)
    return x


def generate_queries(channel, log_domain_size: int, n_queries: int):
    """Queries.generate, queries.ts:50-82."""
    found, mask = set(), (1 << log_domain_size) - 1
    while len(found) < n_queries:
        b = channel.draw_random_bytes()
        for i in range(0, 32, 4):
            found.add(int.from_bytes(b[i:i + 4], "little") & mask)
            if len(found) == n_queries:
                break
    return sorted(found)


def prove_wide_fibonacci(log_n_rows: int, n_cols: int, pow_bits: int, log_last_layer_degree_bound: int, log_blowup: int, n_queries: int,
                         trace=None):
    """Returns a dict of every value the proof carries plus the intermediate values a test may want to compare."""
    trace = fib_trace(log_n_rows, n_cols) if trace is None else trace
    comp_log = log_n_rows + 1                                   # maxConstraintLogDegreeBound
    top_log = comp_log + log_blowup
    hi, hl = orc.canonic_half_coset(top_log)
    tw, itw = orc.precompute_twiddles(hi, hl)
    root = (hi, hl)
    half = lambda log: orc.canonic_half_coset(log)[0]
    channel = orc.Channel()

    # tree 0: the (empty) preprocessed trace; tree 1: the trace
    trees = []          # per tree: (column log sizes, evaluations, layers, coefficient arrays)
    layers0 = orc.merkle_commit([])
    channel.mix_root(layers0[0][0].tobytes())
    trees.append(([], [], layers0, []))
    coeffs = [orc.interpolate(c, half(log_n_rows), root, itw, quirks=False) for c in trace]
    evals = [orc.evaluate(c, log_n_rows + log_blowup, half(log_n_rows + log_blowup), root, tw, quirks=False) for c in coeffs]
    layers1 = orc.merkle_commit(evals)
    channel.mix_root(layers1[0][0].tobytes())
    trees.append(([log_n_rows + log_blowup] * n_cols, evals, layers1, coeffs))

    # composition polynomial
    random_coeff = channel.draw_felt()
    n_con = n_cols - 2
    pows = orc.generate_secure_powers(random_coeff, n_con)
    cons_coeffs = [[int(x) for x in pows[n_con - 1 - i]] for i in range(n_con)]
    dom_evals = evals if log_blowup == 1 else [orc.evaluate(c, comp_log, half(comp_log), root, tw, quirks=False) for c in coeffs]
    zero = [np.zeros(1 << comp_log, dtype=np.uint32) for _ in range(4)]
    quot = orc.widefib_constraints(log_n_rows, comp_log, dom_evals, cons_coeffs, orc.widefib_denom_inv(log_n_rows, comp_log), zero)
    comp_coeffs = [orc.interpolate(c, half(comp_log), root, itw, quirks=False) for c in quot]
    comp_evals = [orc.evaluate(c, top_log, half(top_log), root, tw, quirks=False) for c in comp_coeffs]
    layers2 = orc.merkle_commit(comp_evals)
    channel.mix_root(layers2[0][0].tobytes())
    trees.append(([top_log] * 4, comp_evals, layers2, comp_coeffs))

    # out-of-domain sampling
    oods = random_point(channel)
    sampled = [[[orc.eval_at_point(c, oods[0], oods[1])] for c in t[3]] for t in trees]
    channel.mix_felts([v for tree in sampled for col in tree for v in col])

    # DEEP quotients: columns of all trees, largest first (stable); one batch per size since every column has the one point
    fri_random_coeff = channel.draw_felt()
    flat = [(log, ev, sampled[ti][ci][0]) for ti, t in enumerate(trees) for ci, (log, ev) in enumerate(zip(t[0], t[1]))]
    fri_columns = []
    for log in sorted({f[0] for f in flat}, reverse=True):
        group = [f for f in flat if f[0] == log]
        batches = [(oods, [(i, g[2]) for i, g in enumerate(group)])]
        q = orc.accumulate_quotients(log, half(log), [g[1] for g in group], fri_random_coeff, batches, ts_denominators=False)
        fri_columns.append((log, q))
    fri = orc.fri_commit(channel, log_last_layer_degree_bound, log_blowup, fri_columns)
    nonce = orc.grind(channel, pow_bits)
    channel.mix_u64(nonce)
    positions = generate_queries(channel, top_log, n_queries)
    fri_layers = orc.fri_decommit_on_queries(fri, positions, top_log)
    fold = lambda pos, k: sorted({p >> k for p in pos})
    by_log = {log: fold(positions, top_log - log) for log in {f[0] for f in flat}}
    decommitments = [orc.merkle_decommit(t[2], by_log, t[1]) for t in trees]

    # the sanity check of prove(): composition value from the samples == constraints at the sampled mask values
    comp_at_oods = sampled[2][0][0]
    for k, unit in ((1, [0, 1, 0, 0]), (2, [0, 0, 1, 0]), (3, [0, 0, 0, 1])):
        comp_at_oods = _qm(0, comp_at_oods, _qm(2, sampled[2][k][0], unit))
    denom_inv = _qm(4, coset_vanishing_secure(orc.subgroup_gen(log_n_rows + 1), log_n_rows, oods))
    acc = [0, 0, 0, 0]
    for i in range(n_con):
        a, b, c = sampled[1][i][0], sampled[1][i + 1][0], sampled[1][i + 2][0]
        con = _qm(1, c, _qm(0, _qm(2, a, a), _qm(2, b, b)))
        acc = _qm(0, _qm(2, acc, random_coeff), _qm(2, denom_inv, con))
    return {"commitments": [t[2][0][0].tobytes() for t in trees], "sampled_values": sampled, "random_coeff": random_coeff, "oods": oods,
            "composition_coeffs": comp_coeffs, "fri": fri, "fri_layers": fri_layers, "proof_of_work": nonce, "positions": positions,
            "decommitments": decommitments, "queries_by_log": by_log, "tree_logs": [t[0] for t in trees],
            "composition_at_oods": comp_at_oods, "constraints_at_oods": acc, "channel_digest": channel.digest}


def serialize_proof(exp: dict, pow_bits: int, log_last_layer_degree_bound: int, log_blowup: int, n_queries: int) -> bytes:
    """The oracle's proof (the dict prove_wide_fibonacci returns) in the byte format of tstwo_b200/proof_format.py, written
    independently of it (struct.pack, value by value): config, commitments, sampled values, decommitments, queried values,
    proof of work, FRI proof.  Tests compare these bytes with StarkProof.toBytes() of the device prover."""
    import struct
    out = [b"TSTWOPF1", struct.pack("<IIII", pow_bits, log_last_layer_degree_bound, log_blowup, n_queries)]
    u32 = lambda v: struct.pack("<I", int(v))
    def words(ws):
        ws = [int(w) for w in ws]
        return struct.pack(f"<{len(ws)}I", *ws)
    def decommitment(hash_witness, column_witness):
        return [u32(len(hash_witness)), b"".join(bytes(h) for h in hash_witness), u32(len(column_witness)), words(column_witness)]
    out.append(u32(len(exp["commitments"])))
    out += [bytes(c) for c in exp["commitments"]]
    out.append(u32(len(exp["sampled_values"])))
    for tree in exp["sampled_values"]:
        out.append(u32(len(tree)))
        for col in tree:
            out.append(u32(len(col)))
            for value in col:
                out.append(words(value))
    out.append(u32(len(exp["decommitments"])))
    for queried, hw, cw in exp["decommitments"]:
        out += decommitment(hw, cw)
    out.append(u32(len(exp["decommitments"])))
    for queried, hw, cw in exp["decommitments"]:
        out += [u32(len(queried)), words(queried)]
    out.append(struct.pack("<Q", exp["proof_of_work"]))
    layers = exp["fri_layers"]
    def fri_layer(layer):
        witness, (queried, hw, cw), root = layer
        r = [u32(len(witness))] + [words(w) for w in witness]
        return r + decommitment(hw, cw) + [bytes(root)]
    out += fri_layer(layers[0])
    out.append(u32(len(layers) - 1))
    for layer in layers[1:]:
        out += fri_layer(layer)
    last = exp["fri"]["last_poly"]
    out.append(u32(len(last)))
    out += [words(c) for c in last]
    return b"".join(out)
