"""TEST INFRASTRUCTURE (oracle): a host FRI verifier, so that the proofs the device FriProver and the oracle's fri_commit /
fri_decommit_on_queries produce are checked for what they are for -- a verifier accepts them -- and not only against each other.

The reference's verifier is a skeleton (core/src/fri.ts:544-617, 791-976: `verify`, `verifyAndFold` and `decommitInnerLayers`
are placeholders); this file fills those bodies in with the semantics its own prover text implies (fri.ts:676-716):

  * transcript: mix_root(first layer), alpha_0 = draw_felt; per inner layer mix_root, alpha_k = draw_felt; mix_felts(last poly)
    (FriVerifier.commit, fri.ts:809-886, with the channel calls in place of the "placeholder" comments);
  * a circle column whose folded size equals an inner layer's size joins that layer, folded with the alpha that produced the
    layer (the PREVIOUS layer's folding alpha; the first column uses alpha_0): fri.ts:704-711 passes the loop's `foldingAlpha`
    to both fold_line and fold_circle_into_line.  accumulate_line (fri.ts:453-462): eval = eval * alpha^2 + folded;
  * SparseEvaluation (fri.ts:283-335) rebuilt from the query evaluations and the FRI witness (fri.ts:386-450), every layer's
    Merkle decommitment checked with the MerkleVerifier restatement (oracle.merkle_verify);
  * last layer: the folded evaluations equal last_layer_poly at the query points (fri.ts:921-945).

Plain Python integers (independent of the C oracle's field code)."""
from __future__ import annotations

P = 2147483647


def _inv(a: int) -> int:
    return pow(a % P, P - 2, P)


def q_add(x, y): return tuple((a + b) % P for a, b in zip(x, y))
def q_sub(x, y): return tuple((a - b) % P for a, b in zip(x, y))
def q_scale(x, m): return tuple(a * m % P for a in x)


def _c_mul(x, y):
    return ((x[0] * y[0] - x[1] * y[1]) % P, (x[0] * y[1] + x[1] * y[0]) % P)


def q_mul(x, y):
    """QM31 = CM31[u] / (u^2 - 2 - i), fields/qm31.ts."""
    a, b, c, d = (x[0], x[1]), (x[2], x[3]), (y[0], y[1]), (y[2], y[3])
    bd = _c_mul(b, d)
    rbd = ((2 * bd[0] - bd[1]) % P, (bd[0] + 2 * bd[1]) % P)
    ac = _c_mul(a, c)
    ad, bc = _c_mul(a, d), _c_mul(b, c)
    return ((ac[0] + rbd[0]) % P, (ac[1] + rbd[1]) % P, (ad[0] + bc[0]) % P, (ad[1] + bc[1]) % P)


def _bitrev(i: int, log: int) -> int:
    return int(format(i, f"0{log}b")[::-1], 2) if log else 0


def _point(index: int):
    """CirclePointIndex.toPoint by square-and-multiply on the generator (circle.ts:172-174, M31_CIRCLE_GEN = (2, 1268011823))."""
    x, y = 1, 0
    gx, gy = 2, 1268011823
    index &= 0x7FFFFFFF
    while index:
        if index & 1:
            x, y = (x * gx - y * gy) % P, (x * gy + y * gx) % P
        gx, gy = (gx * gx - gy * gy) % P, (2 * gx * gy) % P
        index >>= 1
    return x, y


def _circle_domain_at(log: int, i: int):
    """CircleDomain.at(i) of CanonicCoset(log).circleDomain(): half coset G_{log+1}-odd start, step G_{log-1}; second half is
    the conjugate half (domain.ts: index i >= half -> -halfCoset.indexAt(i - half))."""
    half = 1 << (log - 1)
    init, step = 1 << (30 - log), 1 << (31 - (log - 1))          # CanonicCoset.new(log).halfCoset() = Coset.new(G_{4n}, log-1)
    if i < half:
        return _point(init + i * step)
    x, y = _point(init + (i - half) * step)
    return x, (P - y) % P


def _line_domain_at(log: int, i: int) -> int:
    """LineDomain.new(Coset.half_odds(log)).at(i) = x of half_odds(log).at(i) (line.ts, circle.ts half_odds = G_{4n} start, step G_n)."""
    init, step = 1 << (31 - (log + 2)), 1 << (31 - log)
    return _point(init + i * step)[0]


def _sparse(queries, query_evals, witness, fold_step=1):
    """SparseEvaluation of fri.ts:386-450: per coset of 2^fold_step positions the evaluations (query value or next witness
    value) and the coset's first position.  Returns (subsets, starts, all decommitment positions)."""
    subsets, starts, positions = [], [], []
    qi, i = 0, 0
    witness = iter(witness)
    while i < len(queries):
        coset = queries[i] >> fold_step
        here = {}
        while i < len(queries) and queries[i] >> fold_step == coset:
            here[queries[i]] = tuple(int(v) for v in query_evals[qi]); qi += 1; i += 1
        vals = []
        for pos in range(coset << fold_step, (coset + 1) << fold_step):
            positions.append(pos)
            if pos in here:
                vals.append(here[pos])
            else:
                try:
                    vals.append(tuple(int(v) for v in next(witness)))
                except StopIteration:
                    raise ValueError("InsufficientWitness")
        subsets.append(vals); starts.append(coset << fold_step)
    if next(witness, None) is not None:
        raise ValueError("WitnessTooLong")
    return subsets, starts, positions


def fri_verify(orc, channel, log_last_layer_degree_bound: int, log_blowup: int, column_logs, layers, last_poly, query_positions,
               first_layer_query_evals):
    """column_logs: commitment-domain log sizes, strictly decreasing; layers: [(fri_witness, (hash_witness, column_witness), root)]
    first layer first; last_poly: bit-reversed-order coefficients (LinePoly.coeffs); query_positions on the largest domain;
    first_layer_query_evals[c][k] = column c at its k-th (folded) query position, 4 coordinates.  Raises ValueError."""
    # ---- commit phase (transcript)
    if any(a <= b for a, b in zip(column_logs, column_logs[1:])):
        raise ValueError("InvalidNumFriLayers")
    first, inner = layers[0], layers[1:]
    channel.mix_root(first[2])
    alphas = [tuple(channel.draw_felt())]
    layer_log = column_logs[0] - 1
    last_log = log_last_layer_degree_bound + log_blowup
    if len(inner) != layer_log - last_log:
        raise ValueError("InvalidNumFriLayers")
    for _, _, root in inner:
        channel.mix_root(root)
        alphas.append(tuple(channel.draw_felt()))
    if len(last_poly) > 1 << log_last_layer_degree_bound:
        raise ValueError("LastLayerDegreeInvalid")
    channel.mix_felts(last_poly)
    # a whole-proof verifier draws the queries (and derives the evaluations from the trace decommitments) only now, from the
    # transcript as it stands after the commit phase: it passes a callable instead of the positions
    if callable(query_positions):
        query_positions, first_layer_query_evals = query_positions(channel)

    # ---- first layer: sparse evaluations per column + ONE Merkle decommitment over all coordinate columns
    max_log = column_logs[0]
    fold_q = lambda pos, k: sorted({p >> k for p in pos})
    sparse_by_col, pos_by_log, queried = [], {}, {}
    wit = iter(first[0])
    for c, log in enumerate(column_logs):
        col_q = fold_q(query_positions, max_log - log)
        # the witness of the first layer is the concatenation of the columns' witnesses: take exactly this column's share
        n_wit = len({p >> 1 for p in col_q}) * 2 - len(col_q)
        mine = []
        for _ in range(n_wit):
            try:
                mine.append(next(wit))
            except StopIteration:
                raise ValueError("InsufficientWitness")
        subsets, starts, positions = _sparse(col_q, first_layer_query_evals[c], mine)
        sparse_by_col.append((log, subsets, starts))
        pos_by_log[log] = positions
        queried[log] = [v for vals in subsets for ev in vals for v in ev]
    if next(wit, None) is not None:
        raise ValueError("WitnessTooLong")
    flat = [v for log in sorted(queried, reverse=True) for v in queried[log]]
    orc.merkle_verify(first[2], [log for log in column_logs for _ in range(4)], pos_by_log, flat, first[1][0], first[1][1])

    # ---- inner layers
    def fold_pair(f0, f1, inv_tw, alpha):        # ibutterfly then f0' + alpha f1' (fri.ts:120-192 on a 2-element evaluation)
        return q_add(q_add(f0, f1), q_mul(alpha, q_scale(q_sub(f0, f1), inv_tw)))

    def fold_in_circle(evals, queries, log_line, prev_alpha):
        while sparse_by_col and sparse_by_col[0][0] - 1 == log_line:
            log, subsets, starts = sparse_by_col.pop(0)
            if len(subsets) != len(evals):
                raise ValueError("FirstLayerEvaluationsInvalid")
            a2 = q_mul(prev_alpha, prev_alpha)
            for i, (vals, start) in enumerate(zip(subsets, starts)):
                if start >> 1 != queries[i]:
                    raise ValueError("FirstLayerEvaluationsInvalid")
                y = _circle_domain_at(log, _bitrev(start, log))[1]
                evals[i] = q_add(q_mul(evals[i], a2), fold_pair(vals[0], vals[1], _inv(y), prev_alpha))

    queries = fold_q(query_positions, 1)
    evals = [(0, 0, 0, 0)] * len(queries)
    prev_alpha = alphas[0]
    for k, (witness, (hash_w, col_w), root) in enumerate(inner):
        fold_in_circle(evals, queries, layer_log, prev_alpha)
        subsets, starts, positions = _sparse(queries, evals, witness)
        flat = [v for vals in subsets for ev in vals for v in ev]
        orc.merkle_verify(root, [layer_log] * 4, {layer_log: positions}, flat, hash_w, col_w)
        alpha = alphas[k + 1]
        evals = [fold_pair(vals[0], vals[1], _inv(_line_domain_at(layer_log, _bitrev(start, layer_log))), alpha)
                 for vals, start in zip(subsets, starts)]
        queries = fold_q(queries, 1)
        layer_log -= 1
        prev_alpha = alpha
    fold_in_circle(evals, queries, layer_log, prev_alpha)      # a column may join the last layer itself (fri.ts:707-711)
    if sparse_by_col:
        raise ValueError("FirstLayerEvaluationsInvalid")

    # ---- last layer: LinePoly.evalAtPoint (poly/line.ts:225-246) at the query points
    if layer_log != last_log:
        raise ValueError("InvalidNumFriLayers")
    coeffs = [tuple(int(v) for v in c) for c in last_poly]
    log_n = (len(coeffs)).bit_length() - 1
    for pos, ev in zip(queries, evals):
        x = _line_domain_at(layer_log, _bitrev(pos, layer_log))
        # bit-reversed coefficient order: fold with the doublings of x, highest first (line.ts evalAtPoint)
        doublings = []
        t = x
        for _ in range(log_n):
            doublings.append(t)
            t = (2 * t * t - 1) % P
        vals = list(coeffs)
        for d in reversed(doublings):            # poly/utils.ts fold(): pairs (lo, hi) -> lo + hi * factor, last factor first
            vals = [q_add(vals[2 * i], q_scale(vals[2 * i + 1], d)) for i in range(len(vals) // 2)]
        if vals[0] != tuple(ev):
            raise ValueError("LastLayerEvaluationsInvalid")
    return alphas
