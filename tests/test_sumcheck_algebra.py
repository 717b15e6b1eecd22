"""The two algebraic shortcuts of the device sum-checks, restated with Python integers and checked against the plain
round computation of the reference (sumcheck.rs:453-484, 618-674; dense_mlpoly.rs:214-221).  No GPU, no library: this pins
the identities `prover.inc` `EqFactor` / `HalfZero` and `kernels.cu` `sc_cubic_eq_kernel` / `sc_quad_half0_kernel` rely on;
the byte-level parity of the proofs they produce is in tests/test_gpu_parity.py and tests/test_sharded.py."""
import random

ELL = 2 ** 252 + 27742317777372353535851937790883648493


def eq_table(tau):
    """EqPolynomial::evals, dense_mlpoly.rs:67-83: index bit (len-1-j) selects tau[j]."""
    t = [1]
    for x in tau:
        t = [w for v in t for w in (v * (1 - x) % ELL, v * x % ELL)]
    return t


def bind_top(t, r):
    h = len(t) // 2
    return [(t[i] + r * (t[i + h] - t[i])) % ELL for i in range(h)]


def cubic_round(T, A, B, C):
    """e(0), e(2), e(3) of sum_x T*(A*B - C), sumcheck.rs:622-650."""
    h = len(T) // 2
    e = [0, 0, 0]
    for i in range(h):
        lo = [X[i] for X in (T, A, B, C)]
        hi = [X[i + h] for X in (T, A, B, C)]
        for k, pt in enumerate((0, 2, 3)):
            v = [(l + pt * (u - l)) % ELL for l, u in zip(lo, hi)]
            e[k] = (e[k] + v[0] * (v[1] * v[2] - v[3])) % ELL
    return e


def test_eq_factored_out_of_the_cubic_round():
    rng = random.Random(5)
    m = 6
    n = 1 << m
    tau = [rng.randrange(ELL) for _ in range(m)]
    A = [rng.randrange(ELL) for _ in range(n)]
    B = [rng.randrange(ELL) for _ in range(n)]
    # a satisfied instance: C = A*B, so the claim of round 0 is 0 as in r1csproof.rs:199
    C = [a * b % ELL for a, b in zip(A, B)]
    T = eq_table(tau)
    claim, S = 0, 1
    for j in range(m):
        want = cubic_round(T, A, B, C)
        assert (want[0] + (claim - want[0])) % ELL == claim
        # what the factored kernel returns: q(0) and q's leading coefficient over the suffix table E_j
        h = len(A) // 2
        E = eq_table(tau[j + 1:])
        q0 = sum(E[x] * (A[x] * B[x] - C[x]) for x in range(h)) % ELL
        qa = sum(E[x] * (A[x + h] - A[x]) * (B[x + h] - B[x]) for x in range(h)) % ELL
        # host side (EqFactor::evals): q(1) from the claim, then s(X) = S * eq(tau_j, X) * q(X)
        c = claim * pow(S, -1, ELL) % ELL  # the prover tracks c = claim / S without dividing
        q1 = (c - (1 - tau[j]) * q0) * pow(tau[j], -1, ELL) % ELL
        qb = (q1 - q0 - qa) % ELL
        got = []
        for pt in (0, 2, 3):
            lin = S * ((1 - tau[j]) * (1 - pt) + tau[j] * pt) % ELL
            got.append(lin * (q0 + qb * pt + qa * pt * pt) % ELL)
        assert got == want
        r = rng.randrange(ELL)
        # the round polynomial through e(0), e(1) = claim - e(0), e(2), e(3), evaluated at r: next claim
        e1 = (claim - want[0]) % ELL
        pts, vals = (0, 1, 2, 3), (want[0], e1, want[1], want[2])
        nxt = 0
        for i, (xi, yi) in enumerate(zip(pts, vals)):
            num, den = 1, 1
            for k, xk in enumerate(pts):
                if k != i:
                    num = num * (r - xk) % ELL
                    den = den * (xi - xk) % ELL
            nxt = (nxt + yi * num * pow(den, -1, ELL)) % ELL
        # EqFactor::bound: S picks up eq(tau_j, r), c becomes q(r) -- and S * c is the next claim
        S_next = S * ((1 - tau[j]) * (1 - r) + tau[j] * r) % ELL
        c_next = (q0 + r * (qb + r * qa)) % ELL
        assert S_next * c_next % ELL == nxt
        claim, S = nxt, S_next
        T, A, B, C = (bind_top(X, r) for X in (T, A, B, C))
    assert len(T) == 1


def quad_round(Z, Q):
    h = len(Z) // 2
    e0 = sum(Z[i] * Q[i] for i in range(h)) % ELL
    e2 = sum((2 * Z[i + h] - Z[i]) * (2 * Q[i + h] - Q[i]) for i in range(h)) % ELL
    return e0, e2


def test_half_empty_tables_in_the_quadratic_rounds():
    rng = random.Random(6)
    N, k0 = 32, 5  # z = (w, 1, io, 0...): the upper half is zero past k0 entries, and so is the combined matrix table
    Z = [rng.randrange(ELL) for _ in range(N)] + [rng.randrange(ELL) for _ in range(k0)] + [0] * (N - k0)
    Q = [rng.randrange(ELL) for _ in range(N)] + [rng.randrange(ELL) for _ in range(k0)] + [0] * (N - k0)
    # round 0: one product per pair past the leading entries
    e0 = sum(Z[i] * Q[i] for i in range(N)) % ELL
    e2 = (sum((2 * Z[i + N] - Z[i]) * (2 * Q[i + N] - Q[i]) for i in range(k0)) + sum(Z[i] * Q[i] for i in range(k0, N))) % ELL
    assert (e0, e2) == quad_round(Z, Q)
    r0 = rng.randrange(ELL)
    Zb, Qb = bind_top(Z, r0), bind_top(Q, r0)
    # instead of the bind: adjust the k0 leading entries, keep s = 1 - r0 outside
    s = (1 - r0) % ELL
    t = r0 * pow(s, -1, ELL) % ELL
    Zs = [(Z[i] + t * Z[i + N]) % ELL if i < k0 else Z[i] for i in range(N)]
    Qs = [(Q[i] + t * Q[i + N]) % ELL if i < k0 else Q[i] for i in range(N)]
    assert [s * v % ELL for v in Zs] == Zb and [s * v % ELL for v in Qs] == Qb
    # later rounds run on the unscaled tables; their sums carry s^2, values that reach the host carry s
    for _ in range(3):
        want = quad_round(Zb, Qb)
        got = quad_round(Zs, Qs)
        assert tuple(s * s * v % ELL for v in got) == want
        r = rng.randrange(ELL)
        Zb, Qb, Zs, Qs = (bind_top(X, r) for X in (Zb, Qb, Zs, Qs))
        assert [s * v % ELL for v in Zs] == Zb
