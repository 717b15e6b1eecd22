"""ORACLE (test infrastructure only).  Independent big-int restatement of SNARK::prove -- written from the Rust sources, not
from oracle/snark.hpp -- for small instances: it pins the C++ restatement through tests/golden/snark_prove_py_golden.json
(two readings of the reference that agree byte for byte).  PARITY UNPINNED against a cargo build, like the rest.

  SNARK::prove                                lib.rs:335-403
  R1CSEvalProof::prove                        r1csinstance.rs:310-329
  SparseMatPolyEvalProof::prove               sparse_mlpoly.rs:1471-1538
  PolyEvalNetwork::new, Layers                sparse_mlpoly.rs:516-668
  ProductLayerProof::prove                    sparse_mlpoly.rs:1052-1230
  HashLayerProof::prove, DerefsEvalProof      sparse_mlpoly.rs:74-152, 723-837
  ProductCircuit, DotProductCircuit,
  ProductCircuitEvalProofBatched::prove       product_tree.rs:10-115, 257-382
  SumcheckInstanceProof::prove_cubic_batched  sumcheck.rs:252-422
  PolyEvalProof::prove                        dense_mlpoly.rs:309-362
"""
import struct

import snark_spec as E
import spartan_spec as S

Q = S.ELL


def bound_top(Z, r):  # DensePolynomial::bound_poly_var_top, dense_mlpoly.rs:214-221
    n = len(Z) // 2
    return [(Z[i] + r * (Z[i + n] - Z[i])) % Q for i in range(n)]


def bound_bot(Z, r):  # bound_poly_var_bot, dense_mlpoly.rs:223-230
    n = len(Z) // 2
    return [(Z[2 * i] + r * (Z[2 * i + 1] - Z[2 * i])) % Q for i in range(n)]


def evaluate(Z, r):  # DensePolynomial::evaluate, dense_mlpoly.rs:233-239
    chis = S.eq_evals(r)
    assert len(chis) == len(Z)
    return sum(a * b for a, b in zip(Z, chis)) % Q


def append_scalars(tr, label, v):  # impl AppendToTranscript for [Scalar], transcript.rs:46-54
    tr.append_message(label, b"begin_append_vector")
    for x in v:
        tr.append_scalar(label, x)
    tr.append_message(label, b"end_append_vector")


def append_poly_commitment(tr, label, C):  # dense_mlpoly.rs:289-297
    tr.append_message(label, b"poly_commitment_begin")
    for c in C:
        tr.append_point(b"poly_commitment_share", c)
    tr.append_message(label, b"poly_commitment_end")


class ProductCircuit:  # product_tree.rs:10-64
    def __init__(self, poly):
        num_layers = S.log2(len(poly))
        half = len(poly) // 2
        self.left, self.right = [poly[:half]], [poly[half:]]
        for i in range(num_layers - 1):
            l, r = self.left[i], self.right[i]
            n = len(l) + len(r)
            self.left.append([l[k] * r[k] % Q for k in range(n // 4)])
            self.right.append([l[k] * r[k] % Q for k in range(n // 4, n // 2)])

    def evaluate(self):
        assert len(self.left[-1]) == 1 and len(self.right[-1]) == 1
        return self.left[-1][0] * self.right[-1][0] % Q


class DotProductCircuit:  # product_tree.rs:66-115
    def __init__(self, left, right, weight):
        self.left, self.right, self.weight = left, right, weight

    def evaluate(self):
        return sum(a * b * c for a, b, c in zip(self.left, self.right, self.weight)) % Q

    def split(self):
        h = len(self.left) // 2
        return (DotProductCircuit(self.left[:h], self.right[:h], self.weight[:h]),
                DotProductCircuit(self.left[h:], self.right[h:], self.weight[h:]))


def prove_cubic_batched(claim, num_rounds, par, seq, coeffs, tr):
    """sumcheck.rs:252-422 with comb_func = A*B*C.  par = (list of A, list of B, C); seq = (list of A, list of B, list of C);
    the lists are replaced by their bound forms.  Returns (compressed polys, r, claims_prod, claims_dotp)."""
    A_par, B_par, C_par = par
    A_seq, B_seq, C_seq = seq
    e, r, polys = claim, [], []

    def evals3(A, B, C):
        n = len(A) // 2
        e0 = e2 = e3 = 0
        for i in range(n):
            e0 += A[i] * B[i] * C[i]
            a, b, c = 2 * A[n + i] - A[i], 2 * B[n + i] - B[i], 2 * C[n + i] - C[i]
            e2 += a * b * c
            a, b, c = a + A[n + i] - A[i], b + B[n + i] - B[i], c + C[n + i] - C[i]
            e3 += a * b * c
        return e0 % Q, e2 % Q, e3 % Q

    for _ in range(num_rounds):
        ev = [evals3(A, B, C_par) for A, B in zip(A_par, B_par)] + [evals3(A, B, C) for A, B, C in zip(A_seq, B_seq, C_seq)]
        c0 = sum(x[0] * k for x, k in zip(ev, coeffs)) % Q
        c2 = sum(x[1] * k for x, k in zip(ev, coeffs)) % Q
        c3 = sum(x[2] * k for x, k in zip(ev, coeffs)) % Q
        poly = S.unipoly_from_evals([c0, (e - c0) % Q, c2, c3])
        tr.append_message(b"poly", b"UniPoly_begin")  # unipoly.rs:112-120
        for x in poly:
            tr.append_scalar(b"coeff", x)
        tr.append_message(b"poly", b"UniPoly_end")
        r_j = tr.challenge_scalar(b"challenge_nextround")
        r.append(r_j)
        A_par[:] = [bound_top(A, r_j) for A in A_par]
        B_par[:] = [bound_top(B, r_j) for B in B_par]
        C_par = bound_top(C_par, r_j)
        A_seq[:] = [bound_top(A, r_j) for A in A_seq]
        B_seq[:] = [bound_top(B, r_j) for B in B_seq]
        C_seq[:] = [bound_top(C, r_j) for C in C_seq]
        e = S.unipoly_eval(poly, r_j)
        polys.append([poly[0]] + poly[2:])  # CompressedUniPoly: the linear term is dropped, unipoly.rs:82-88
    claims_prod = ([A[0] for A in A_par], [B[0] for B in B_par])
    claims_dotp = ([A[0] for A in A_seq], [B[0] for B in B_seq], [C[0] for C in C_seq])
    return polys, r, claims_prod, claims_dotp


def product_batched_prove(prods, dotps, tr):
    """ProductCircuitEvalProofBatched::prove, product_tree.rs:257-382.  Returns (layer proofs, claims_dotp_final, rand)."""
    layers, claims_dotp_final = [], ([], [], [])
    num_layers = len(prods[0].left)
    claims = [c.evaluate() for c in prods]
    rand = []
    for layer_id in reversed(range(num_layers)):
        n = len(prods[0].left[layer_id]) + len(prods[0].right[layer_id])
        C_par = S.eq_evals(rand)
        assert len(C_par) == n // 2
        num_rounds = S.log2(len(C_par))
        A_par = [c.left[layer_id] for c in prods]
        B_par = [c.right[layer_id] for c in prods]
        A_seq, B_seq, C_seq = [], [], []
        if layer_id == 0 and dotps:
            for d in dotps:
                claims.append(d.evaluate())
                assert len(d.left) == n // 2
            A_seq, B_seq, C_seq = [d.left for d in dotps], [d.right for d in dotps], [d.weight for d in dotps]
        coeffs = tr.challenge_vector(b"rand_coeffs_next_layer", len(claims))
        claim = sum(a * b for a, b in zip(claims, coeffs)) % Q
        polys, rand_prod, (cl, cr), cd = prove_cubic_batched(claim, num_rounds, (A_par, B_par, C_par), (A_seq, B_seq, C_seq), coeffs, tr)
        for i in range(len(prods)):
            tr.append_scalar(b"claim_prod_left", cl[i])
            tr.append_scalar(b"claim_prod_right", cr[i])
        if layer_id == 0 and dotps:
            for i in range(len(dotps)):
                tr.append_scalar(b"claim_dotp_left", cd[0][i])
                tr.append_scalar(b"claim_dotp_right", cd[1][i])
                tr.append_scalar(b"claim_dotp_weight", cd[2][i])
            claims_dotp_final = cd
        r_layer = tr.challenge_scalar(b"challenge_r_layer")
        claims = [(l + r_layer * (r - l)) % Q for l, r in zip(cl, cr)]
        rand = [r_layer] + rand_prod
        layers.append((polys, cl, cr))
    return layers, claims_dotp_final, rand


def polyeval_prove_plain(poly, r, Zr, gens_n, gens_1, tr, tape):
    """PolyEvalProof::prove(poly, None, r, Zr, None, ...), dense_mlpoly.rs:309-362."""
    tr.append_protocol_name(b"polynomial evaluation proof")
    assert len(poly) == 1 << len(r)
    lnv = len(r) // 2
    Ls, Rs = 1 << lnv, 1 << (len(r) - lnv)
    Lv, Rv = S.eq_evals(r[:lnv]), S.eq_evals(r[lnv:])
    LZ = [sum(Lv[j] * poly[j * Rs + i] for j in range(Ls)) % Q for i in range(Rs)]
    dpl, _, _ = S.dotproductlog_prove(gens_n, gens_1, tr, tape, LZ, 0, Rv, Zr, 0)
    return dpl


def poly_gens(num_vars, label):
    n = 1 << (num_vars - num_vars // 2)
    return S.MultiCommitGens.new(n + 1, label).split_at(n)


def combine(tr, label, evals, r):
    ch = tr.challenge_vector(label, S.log2(len(evals)))
    for c in reversed(ch):
        evals = bound_bot(evals, c)
    assert len(evals) == 1
    return ch + list(r), evals[0]


def sparse_eval_prove(inst, num_nz_entries, rx, ry, evals, tr, tape, label=b"gens_r1cs_eval"):
    """SparseMatPolyEvalProof::prove.  Returns the proof as a dict."""
    tr.append_protocol_name(b"Sparse polynomial evaluation proof")
    N, cells, comb_ops, comb_mem = E.sparse_to_dense(inst)
    mats = (inst.A, inst.B, inst.C)
    rows = [[e[0] for e in M] + [0] * (N - len(M)) for M in mats]
    cols = [[e[1] for e in M] + [0] * (N - len(M)) for M in mats]
    vals = [[e[2] for e in M] + [0] * (N - len(M)) for M in mats]
    row_ts, row_audit = E.addr_timestamps(cells, N, rows)
    col_ts, col_audit = E.addr_timestamps(cells, N, cols)
    rx, ry = list(rx), list(ry)
    if len(rx) < len(ry):
        rx = [0] * (len(ry) - len(rx)) + rx
    if len(ry) < len(rx):
        ry = [0] * (len(rx) - len(ry)) + ry
    mem_rx, mem_ry = S.eq_evals(rx), S.eq_evals(ry)
    row_val = [[mem_rx[a] for a in rows[m]] for m in range(3)]
    col_val = [[mem_ry[a] for a in cols[m]] for m in range(3)]
    derefs_comb = [x for v in row_val + col_val for x in v]
    derefs_comb += [0] * (E.next_pow2(len(derefs_comb)) - len(derefs_comb))
    lg = S.log2(E.next_pow2(num_nz_entries))
    nv_ops, nv_derefs = lg + S.log2(E.next_pow2(15)), lg + S.log2(E.next_pow2(6))
    nv_mem = max(S.log2(inst.num_cons), S.log2(2 * inst.num_vars)) + 1
    g_ops, g_mem, g_derefs = poly_gens(nv_ops, label), poly_gens(nv_mem, label), poly_gens(nv_derefs, label)
    comm_derefs = E.commit_unblinded(derefs_comb, g_derefs[0])
    tr.append_message(b"derefs_commitment", b"begin_derefs_commitment")
    append_poly_commitment(tr, b"comm_poly_row_col_ops_val", comm_derefs)
    tr.append_message(b"derefs_commitment", b"end_derefs_commitment")
    r_hash, r_ms = tr.challenge_vector(b"challenge_r_hash", 2)
    h2 = r_hash * r_hash % Q
    hsh = lambda a, v, ts: (ts * h2 + v * r_hash + a - r_ms) % Q

    def side(mem, addrs, derefs, read_ts, audit_ts):  # Layers::new, :608-649
        assert len(mem) == len(audit_ts)
        init = ProductCircuit([hsh(i, mem[i], 0) for i in range(len(mem))])
        audit = ProductCircuit([hsh(i, mem[i], audit_ts[i]) for i in range(len(mem))])
        read = [ProductCircuit([hsh(a, v, t) for a, v, t in zip(addrs[m], derefs[m], read_ts[m])]) for m in range(3)]
        write = [ProductCircuit([hsh(a, v, t + 1) for a, v, t in zip(addrs[m], derefs[m], read_ts[m])]) for m in range(3)]
        return init, read, write, audit

    row = side(mem_rx, rows, row_val, row_ts, row_audit)
    col = side(mem_ry, cols, col_val, col_ts, col_audit)
    tr.append_protocol_name(b"Sparse polynomial evaluation proof")  # PolyEvalNetworkProof::protocol_name
    # ---- ProductLayerProof::prove
    tr.append_protocol_name(b"Sparse polynomial product layer proof")
    pl = {}
    for name, (init, read, write, audit) in ((b"row", row), (b"col", col)):
        ei, ea = init.evaluate(), audit.evaluate()
        er, ew = [c.evaluate() for c in read], [c.evaluate() for c in write]
        ws = rs = 1
        for x in ew:
            ws = ws * x % Q
        for x in er:
            rs = rs * x % Q
        assert ei * ws % Q == rs * ea % Q
        tr.append_scalar(b"claim_" + name + b"_eval_init", ei)
        append_scalars(tr, b"claim_" + name + b"_eval_read", er)
        append_scalars(tr, b"claim_" + name + b"_eval_write", ew)
        tr.append_scalar(b"claim_" + name + b"_eval_audit", ea)
        pl[name] = (ei, er, ew, ea)
    dl, dr, el, er_ = [], [], [], []
    for m in range(3):
        left, right = DotProductCircuit(list(row_val[m]), list(col_val[m]), list(vals[m])).split()
        a, b = left.evaluate(), right.evaluate()
        tr.append_scalar(b"claim_eval_dotp_left", a)
        tr.append_scalar(b"claim_eval_dotp_right", b)
        assert (a + b) % Q == evals[m]
        el.append(a)
        er_.append(b)
        dl.append(left)
        dr.append(right)
    ops_circuits = row[1] + row[2] + col[1] + col[2]
    dotps = [dl[0], dr[0], dl[1], dr[1], dl[2], dr[2]]
    proof_ops, claims_dotp, rand_ops = product_batched_prove(ops_circuits, dotps, tr)
    proof_mem, claims_dotp_mem, rand_mem = product_batched_prove([row[0], row[3], col[0], col[3]], [], tr)
    # ---- HashLayerProof::prove
    tr.append_protocol_name(b"Sparse polynomial hash layer proof")
    ev_row = [evaluate(row_val[m], rand_ops) for m in range(3)]
    ev_col = [evaluate(col_val[m], rand_ops) for m in range(3)]
    tr.append_protocol_name(b"Derefs evaluation proof")
    ev = ev_row + ev_col
    ev += [0] * (E.next_pow2(len(ev)) - len(ev))
    append_scalars(tr, b"evals_ops_val", ev)
    r_joint, joint = combine(tr, b"challenge_combine_n_to_one", ev, rand_ops)
    tr.append_scalar(b"joint_claim_eval", joint)
    proof_derefs = polyeval_prove_plain(derefs_comb, r_joint, joint, g_derefs[0], g_derefs[1], tr, tape)
    e_row_addr = [evaluate(rows[m], rand_ops) for m in range(3)]
    e_row_ts = [evaluate(row_ts[m], rand_ops) for m in range(3)]
    e_row_audit = evaluate(row_audit, rand_mem)
    e_col_addr = [evaluate(cols[m], rand_ops) for m in range(3)]
    e_col_ts = [evaluate(col_ts[m], rand_ops) for m in range(3)]
    e_col_audit = evaluate(col_audit, rand_mem)
    e_val = [evaluate(vals[m], rand_ops) for m in range(3)]
    ev = e_row_addr + e_row_ts + e_col_addr + e_col_ts + e_val
    ev += [0] * (E.next_pow2(len(ev)) - len(ev))
    append_scalars(tr, b"claim_evals_ops", ev)
    r_joint, joint = combine(tr, b"challenge_combine_n_to_one", ev, rand_ops)
    tr.append_scalar(b"joint_claim_eval_ops", joint)
    proof_ops_eval = polyeval_prove_plain(comb_ops, r_joint, joint, g_ops[0], g_ops[1], tr, tape)
    ev = [e_row_audit, e_col_audit]
    append_scalars(tr, b"claim_evals_mem", ev)
    r_joint, joint = combine(tr, b"challenge_combine_two_to_one", ev, rand_mem)
    tr.append_scalar(b"joint_claim_eval_mem", joint)
    proof_mem_eval = polyeval_prove_plain(comb_mem, r_joint, joint, g_mem[0], g_mem[1], tr, tape)
    return {"comm_derefs": comm_derefs, "row": pl[b"row"], "col": pl[b"col"], "eval_dotp": (el, er_),
            "proof_mem": (proof_mem, claims_dotp_mem), "proof_ops": (proof_ops, claims_dotp),
            "hash": {"row": (e_row_addr, e_row_ts, e_row_audit), "col": (e_col_addr, e_col_ts, e_col_audit), "val": e_val,
                     "derefs": (ev_row, ev_col), "proof_ops": proof_ops_eval, "proof_mem": proof_mem_eval, "proof_derefs": proof_derefs}}


def inst_evaluate(inst, rx, ry):  # R1CSInstance::evaluate, r1csinstance.rs:264-270
    ex, ey = S.eq_evals(rx), S.eq_evals(ry)
    return [sum(ex[r] * ey[c] * v for r, c, v in M) % Q for M in (inst.A, inst.B, inst.C)]


def snark_prove(inst, vars_, inputs, num_nz_entries, tr, tape_seed):
    """SNARK::prove, lib.rs:335-403.  Returns bincode(SNARK)."""
    tape = S.RandomTape(b"proof", tape_seed)
    tr.append_protocol_name(b"Spartan SNARK proof")
    vars_ = list(vars_) + [0] * (inst.num_vars - len(vars_))
    gens_sat = S.NIZKGens(inst.num_cons, inst.num_vars, inst.num_inputs)  # R1CSGens::new(b"gens_r1cs_sat", ...)
    sat = S.r1cs_prove(inst, vars_, inputs, gens_sat, tr, tape)
    rx, ry = sat["rx"], sat["ry"]
    evals = inst_evaluate(inst, rx, ry)
    tr.append_scalar(b"Ar_claim", evals[0])
    tr.append_scalar(b"Br_claim", evals[1])
    tr.append_scalar(b"Cr_claim", evals[2])
    ev = sparse_eval_prove(inst, num_nz_entries, rx, ry, evals, tr, tape)
    return serialize_snark(sat, evals, ev)


def _sc(x):
    return S.fl_ser(x)


def _vsc(v):
    return struct.pack("<Q", len(v)) + b"".join(_sc(x) for x in v)


def _vpt(v):
    return struct.pack("<Q", len(v)) + b"".join(v)


def _batched(p):
    layers, cd = p
    out = struct.pack("<Q", len(layers))
    for polys, cl, cr in layers:
        out += struct.pack("<Q", len(polys)) + b"".join(_vsc(c) for c in polys) + _vsc(cl) + _vsc(cr)
    return out + _vsc(cd[0]) + _vsc(cd[1]) + _vsc(cd[2])


def _dlog(d):
    return _vpt(d["bullet"]["L_vec"]) + _vpt(d["bullet"]["R_vec"]) + d["delta"] + d["beta"] + _sc(d["z1"]) + _sc(d["z2"])


def serialize_snark(sat, evals, ev):
    out = S.serialize_r1cs_proof(sat) + b"".join(_sc(x) for x in evals)
    out += _vpt(ev["comm_derefs"])
    for side in ("row", "col"):
        ei, er, ew, ea = ev[side]
        out += _sc(ei) + _vsc(er) + _vsc(ew) + _sc(ea)
    out += _vsc(ev["eval_dotp"][0]) + _vsc(ev["eval_dotp"][1])
    out += _batched(ev["proof_mem"]) + _batched(ev["proof_ops"])
    h = ev["hash"]
    for side in ("row", "col"):
        a, t, au = h[side]
        out += _vsc(a) + _vsc(t) + _sc(au)
    out += _vsc(h["val"]) + _vsc(h["derefs"][0]) + _vsc(h["derefs"][1])
    out += _dlog(h["proof_ops"]) + _dlog(h["proof_mem"]) + _dlog(h["proof_derefs"])
    return out
