"""ORACLE (test infrastructure only).  Independent big-int restatement of SNARK::encode -- the public, deterministic
"computation commitment" of Spartan's SNARK mode: lib.rs:321-333 -> R1CSInstance::commit
r1csinstance.rs:285-297 -> SparseMatPolynomial::multi_commit sparse_mlpoly.rs:485-506.

PARITY UNPINNED against a cargo build (no Rust toolchain here); it pins oracle/spartan.hpp's C++ restatement of the same
functions through tests/golden/snark_encode_golden.json.
"""
import struct

import spartan_spec as S


def next_pow2(n):
    """usize::next_power_of_two (0 -> 1)."""
    return 1 if n <= 1 else 1 << (n - 1).bit_length()


def addr_timestamps(num_cells, num_ops, ops_addr):
    """AddrTimestamps::new, sparse_mlpoly.rs:233-266: the audit counters run on across the matrices of the batch."""
    audit = [0] * num_cells
    read = []
    for inst in ops_addr:
        assert len(inst) == num_ops
        rts = [0] * num_ops
        for i, addr in enumerate(inst):
            assert addr < num_cells
            rts[i] = audit[addr]
            audit[addr] += 1
        read.append(rts)
    return read, audit


def sparse_to_dense(inst):
    """multi_sparse_to_dense_rep, sparse_mlpoly.rs:371-422.  Returns (N, num_mem_cells, comb_ops, comb_mem) as integers."""
    mats = (inst.A, inst.B, inst.C)
    N = next_pow2(max(next_pow2(len(M)) for M in mats))
    nvx, nvy = S.log2(inst.num_cons), S.log2(2 * inst.num_vars)  # r1csinstance.rs:96-97
    cells = 1 << max(nvx, nvy)
    rows = [[e[0] for e in M] + [0] * (N - len(M)) for M in mats]  # sparse_to_dense_vecs :353-369
    cols = [[e[1] for e in M] + [0] * (N - len(M)) for M in mats]
    vals = [[e[2] for e in M] + [0] * (N - len(M)) for M in mats]
    r_read, r_audit = addr_timestamps(cells, N, rows)
    c_read, c_audit = addr_timestamps(cells, N, cols)
    comb_ops = [x for group in (rows, r_read, cols, c_read, vals) for v in group for x in v]  # DensePolynomial::merge
    comb_ops += [0] * (next_pow2(len(comb_ops)) - len(comb_ops))
    return N, cells, comb_ops, r_audit + c_audit


def poly_gens_n(num_vars, label):
    """PolyCommitmentGens::new(num_vars, label).gens.gens_n, dense_mlpoly.rs:28-35 + nizk/mod.rs:409-414."""
    n = 1 << (num_vars - num_vars // 2)
    return S.MultiCommitGens.new(n + 1, label).split_at(n)[0]


def commit_unblinded(Z, gn):
    """DensePolynomial::commit(gens, None), dense_mlpoly.rs:178-203 (zero blinds)."""
    ell = S.log2(len(Z))
    assert len(Z) == 1 << ell
    L = 1 << (ell // 2)
    R = len(Z) // L
    return [S.commit_vec(Z[R * i:R * (i + 1)], 0, gn).compress() for i in range(L)]


def snark_encode(inst, num_nz_entries, label=b"gens_r1cs_eval"):
    """Returns the bytes the oracle and the library emit: num_cons, num_vars, num_inputs (R1CSCommitment is not
    Serialize in the reference), then bincode of SparseMatPolyCommitment (sparse_mlpoly.rs:329-337)."""
    N, cells, comb_ops, comb_mem = sparse_to_dense(inst)
    nvx, nvy = S.log2(inst.num_cons), S.log2(2 * inst.num_vars)
    nv_ops = S.log2(next_pow2(num_nz_entries)) + S.log2(next_pow2(3 * 5))  # sparse_mlpoly.rs:308-309
    nv_mem = max(nvx, nvy) + 1
    c_ops = commit_unblinded(comb_ops, poly_gens_n(nv_ops, label))
    c_mem = commit_unblinded(comb_mem, poly_gens_n(nv_mem, label))
    out = struct.pack("<6Q", inst.num_cons, inst.num_vars, inst.num_inputs, 3, N, cells)
    out += struct.pack("<Q", len(c_ops)) + b"".join(c_ops)
    out += struct.pack("<Q", len(c_mem)) + b"".join(c_mem)
    return out, comb_ops, comb_mem


def commit_derefs(inst, num_nz_entries, rx, ry, label=b"gens_r1cs_eval"):
    """The commit_nondet_witness step of SparseMatPolyEvalProof::prove (sparse_mlpoly.rs:1485-1503): equalize (:1453-1469),
    eq tables, deref (:268-284, :508-514), Derefs::new (:47-62), Derefs::commit (:64-67).  Returns (points, comb)."""
    rx, ry = list(rx), list(ry)
    if len(rx) < len(ry):
        rx = [0] * (len(ry) - len(rx)) + rx
    if len(ry) < len(rx):
        ry = [0] * (len(rx) - len(ry)) + ry
    mem_rx, mem_ry = S.eq_evals(rx), S.eq_evals(ry)
    mats = (inst.A, inst.B, inst.C)
    N = next_pow2(max(next_pow2(len(M)) for M in mats))
    comb = []
    for k, mem in ((0, mem_rx), (1, mem_ry)):
        for M in mats:
            comb += [mem[e[k]] for e in M] + [mem[0]] * (N - len(M))
    comb += [0] * (next_pow2(len(comb)) - len(comb))
    nv_derefs = S.log2(next_pow2(num_nz_entries)) + S.log2(next_pow2(3 * 2))  # sparse_mlpoly.rs:315-316
    return commit_unblinded(comb, poly_gens_n(nv_derefs, label)), comb


def network_products(inst, rx, ry, r_hash, r_multiset):
    """PolyEvalNetwork::new (sparse_mlpoly.rs:652-668): build_hash_layer (:531-606) and the product circuits
    (product_tree.rs:16-64) of both sides, reduced to the 16 values ProductCircuit::evaluate returns.  A product circuit
    multiplies all its inputs, so the product is taken directly (field arithmetic is exact)."""
    rx, ry = list(rx), list(ry)
    if len(rx) < len(ry):
        rx = [0] * (len(ry) - len(rx)) + rx
    if len(ry) < len(rx):
        ry = [0] * (len(rx) - len(ry)) + ry
    mats = (inst.A, inst.B, inst.C)
    N = next_pow2(max(next_pow2(len(M)) for M in mats))
    cells = 1 << max(S.log2(inst.num_cons), S.log2(2 * inst.num_vars))
    q = S.ELL
    h2 = r_hash * r_hash % q

    def prod(xs):
        assert len(xs) >= 2
        p = 1
        for x in xs:
            p = p * x % q
        return p

    out = []
    for k, mem in ((0, S.eq_evals(rx)), (1, S.eq_evals(ry))):
        assert len(mem) == cells
        addrs = [[e[k] for e in M] + [0] * (N - len(M)) for M in mats]
        read, audit = addr_timestamps(cells, N, addrs)
        hsh = lambda a, v, ts: (ts * h2 + v * r_hash + a - r_multiset) % q
        side = [prod([hsh(i, mem[i], 0) for i in range(cells)])]
        side += [prod([hsh(a, mem[a], t) for a, t in zip(addrs[m], read[m])]) for m in range(3)]
        side += [prod([hsh(a, mem[a], t + 1) for a, t in zip(addrs[m], read[m])]) for m in range(3)]
        side.append(prod([hsh(i, mem[i], audit[i]) for i in range(cells)]))
        out += side
    return out
