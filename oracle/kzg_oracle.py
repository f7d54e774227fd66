"""CPU oracle for the KZG hot path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module.  The product path (kzg-commitments_b200/) never does; it fails loudly without its CUDA library.

This is a plain-Python (big-int) restatement of what the reference computes on the path
SURVEY.md §8 names.  It follows the reference's *mathematical* definition — a naive sum of
independent scalar multiplications — not MIRACL's internal GLV / constant-time machinery, because the
parity target is the unique affine serialisation of the resulting group element:

  polyeval_G1 / polyeval_G2      src/trusted_setup.cpp:149-201   -> msm_g1 / msm_g2
  create_commit / create_proof   src/trusted_setup.cpp:137-142, 203-228
  evaluate_polynomial_points     src/util.cpp:186-211            -> poly_eval_range
  linear_roots_and_polyfit       src/util.cpp:172-178            -> vanishing_poly / interpolate
  (P - I) / Z                    src/trusted_setup.cpp:225       -> poly_quotient
  serialize_ECP                  src/util.cpp:78-94 (ECP_toOctet, miracl-core/cpp/ecp.cpp:445-491)
  export_setup / file ctor       src/trusted_setup.cpp:76-121, 256-287
  blob::from_string / from_bytes src/blob.cpp:3-48
  serialize_ZZ_pX                src/util.cpp:118-144

Parity pinning: the reference stores NO golden vectors, so this oracle is pinned against
outputs of the reference itself, built here by oracle/build_ref.sh (oracle/_ref/kzgref-<curve>):
tests/golden/*.json were produced by tools/gen_golden.py from that binary and tests/test_oracle.py
checks this module against every one of them.
"""
import struct

from .curve_constants import CURVES


class Curve:
    def __init__(self, name):
        c = CURVES[name]
        self.name = name
        self.id = c["id"]
        self.p = c["p"]
        self.r = c["r"]
        self.b = c["b"]
        self.b2 = c["b2"]
        self.g1 = (c["gx"], c["gy"])
        self.g2 = (c["qx"], c["qy"])
        self.M = c["modbytes"]
        self.SB = c["scalar_bytes"]

    # ------------------------------------------------------------------ Fp2 = Fp[i]/(i^2+1)
    def f2mul(self, a, b):
        p = self.p
        return ((a[0] * b[0] - a[1] * b[1]) % p, (a[0] * b[1] + a[1] * b[0]) % p)

    def f2sqr(self, a):
        p = self.p
        return ((a[0] + a[1]) * (a[0] - a[1]) % p, 2 * a[0] * a[1] % p)

    def f2add(self, a, b):
        return ((a[0] + b[0]) % self.p, (a[1] + b[1]) % self.p)

    def f2sub(self, a, b):
        return ((a[0] - b[0]) % self.p, (a[1] - b[1]) % self.p)

    def f2inv(self, a):
        p = self.p
        n = pow(a[0] * a[0] + a[1] * a[1], -1, p)
        return (a[0] * n % p, -a[1] * n % p)

    # ------------------------------------------------------------------ group law (Jacobian, a = 0)
    # Points: None = infinity, else affine (x, y).  `ops` = (mul, sqr, add, sub, inv, zero, one)
    def _ops(self, g2):
        p = self.p
        if not g2:
            return (lambda a, b: a * b % p, lambda a: a * a % p, lambda a, b: (a + b) % p,
                    lambda a, b: (a - b) % p, lambda a: pow(a, -1, p), 0, 1)
        return (self.f2mul, self.f2sqr, self.f2add, self.f2sub, self.f2inv, (0, 0), (1, 0))

    def _jdbl(self, P, o):
        mul, sqr, add, sub, inv, zero, one = o
        X, Y, Z = P
        if Z == zero or Y == zero:
            return (one, one, zero)
        A = sqr(X); B = sqr(Y); C = sqr(B)
        t = sub(sub(sqr(add(X, B)), A), C)
        D = add(t, t)
        E = add(add(A, A), A)
        F = sqr(E)
        X3 = sub(F, add(D, D))
        C8 = add(C, C); C8 = add(C8, C8); C8 = add(C8, C8)
        Y3 = sub(mul(E, sub(D, X3)), C8)
        Z3 = mul(add(Y, Y), Z)
        return (X3, Y3, Z3)

    def _jadd(self, P, Q, o):
        mul, sqr, add, sub, inv, zero, one = o
        if P[2] == zero:
            return Q
        if Q[2] == zero:
            return P
        Z1Z1 = sqr(P[2]); Z2Z2 = sqr(Q[2])
        U1 = mul(P[0], Z2Z2); U2 = mul(Q[0], Z1Z1)
        S1 = mul(mul(P[1], Q[2]), Z2Z2); S2 = mul(mul(Q[1], P[2]), Z1Z1)
        if U1 == U2:
            if S1 == S2:
                return self._jdbl(P, o)
            return (one, one, zero)
        H = sub(U2, U1); R = sub(S2, S1)
        HH = sqr(H); HHH = mul(H, HH); V = mul(U1, HH)
        X3 = sub(sub(sqr(R), HHH), add(V, V))
        Y3 = sub(mul(R, sub(V, X3)), mul(S1, HHH))
        Z3 = mul(mul(P[2], Q[2]), H)
        return (X3, Y3, Z3)

    def _to_affine(self, P, o):
        mul, sqr, add, sub, inv, zero, one = o
        if P[2] == zero:
            return None
        zi = inv(P[2]); zi2 = sqr(zi)
        return (mul(P[0], zi2), mul(P[1], mul(zi2, zi)))

    def _from_affine(self, P, o):
        return (o[6], o[6], o[5]) if P is None else (P[0], P[1], o[6])

    def add(self, P, Q, g2=False):
        o = self._ops(g2)
        return self._to_affine(self._jadd(self._from_affine(P, o), self._from_affine(Q, o), o), o)

    def neg(self, P, g2=False):
        if P is None:
            return None
        o = self._ops(g2)
        return (P[0], o[3](o[5], P[1]))

    def mul(self, k, P, g2=False):
        """k * P by left-to-right double-and-add (what PAIR_G1mul / PAIR_G2mul compute, pair.cpp:876-983)."""
        o = self._ops(g2)
        return self._to_affine(self._jmul(k, P, o), o)

    def _jmul(self, k, P, o):
        k %= self.r
        acc = (o[6], o[6], o[5])
        if P is None or k == 0:
            return acc
        J = self._from_affine(P, o)
        for bit in bin(k)[2:]:
            acc = self._jdbl(acc, o)
            if bit == "1":
                acc = self._jadd(acc, J, o)
        return acc

    def msm(self, scalars, points, g2=False):
        """sum_i scalars[i] * points[i]  — polyeval_G1 / polyeval_G2 (src/trusted_setup.cpp:149-201).
        Windowed evaluation of the same sum (the reference's per-term loop is a speed detail, the
        group element is what is compared)."""
        o = self._ops(g2)
        inf = (o[6], o[6], o[5])
        n = len(scalars)
        if n == 0:
            return None
        c = 4 if n < 32 else 8
        nb = self.r.bit_length()
        total = inf
        jp = [self._from_affine(P, o) for P in points[:n]]
        for w in reversed(range((nb + c - 1) // c)):
            for _ in range(c):
                total = self._jdbl(total, o)
            buckets = [inf] * (1 << c)
            for s, P in zip(scalars, jp):
                d = ((s % self.r) >> (w * c)) & ((1 << c) - 1)
                if d:
                    buckets[d] = self._jadd(buckets[d], P, o)
            run = inf; acc = inf
            for d in range((1 << c) - 1, 0, -1):
                run = self._jadd(run, buckets[d], o)
                acc = self._jadd(acc, run, o)
            total = self._jadd(total, acc, o)
        return self._to_affine(total, o)

    def on_curve(self, P, g2=False):
        if P is None:
            return True
        if not g2:
            return (P[1] * P[1] - P[0] ** 3 - self.b) % self.p == 0
        return self.f2sqr(P[1]) == self.f2add(self.f2mul(self.f2sqr(P[0]), P[0]), self.b2)

    # ------------------------------------------------------------------ byte formats
    def g1_to_octet(self, P):
        """ECP_toOctet(.., compress=false): 04 | x | y big-endian; infinity -> 04 | 0..0 (ecp.cpp:445-491)."""
        M = self.M
        if P is None:
            return b"\x04" + bytes(2 * M)
        return b"\x04" + P[0].to_bytes(M, "big") + P[1].to_bytes(M, "big")

    def g2_to_octet(self, Q):
        """ECP2_toOctet: 04 | x.b | x.a | y.b | y.a (imaginary part first; fp2.cpp:83-87, ecp2.cpp:209-215).
        The reference's FP2 (a + ib) is stored here as the tuple (a, b)."""
        M = self.M
        if Q is None:
            return b"\x04" + bytes(4 * M)
        (xa, xb), (ya, yb) = Q
        return b"\x04" + b"".join(v.to_bytes(M, "big") for v in (xb, xa, yb, ya))

    def serialize_g1(self, P):
        """serialize_ECP (src/util.cpp:78-94): uint32 len (native LE) | octet."""
        o = self.g1_to_octet(P)
        return struct.pack("<I", len(o)) + o

    def serialize_g2(self, Q):
        o = self.g2_to_octet(Q)
        return struct.pack("<I", len(o)) + o

    def deserialize_g1(self, raw):
        """deserialize_ECP (src/util.cpp:96-115): anything not on the curve decodes to infinity."""
        M = self.M
        (ln,) = struct.unpack_from("<I", raw, 0)
        body = raw[4:4 + ln]
        if ln != 2 * M + 1 or body[0] != 4:
            return None
        P = (int.from_bytes(body[1:1 + M], "big"), int.from_bytes(body[1 + M:], "big"))
        return P if self.on_curve(P) and P[0] < self.p and P[1] < self.p else None

    def scalars_to_bytes(self, scalars):
        return b"".join((s % self.r).to_bytes(self.SB, "little") for s in scalars)

    def scalars_from_bytes(self, raw):
        SB = self.SB
        return [int.from_bytes(raw[i:i + SB], "little") for i in range(0, len(raw), SB)]

    def srs_to_file_bytes(self, g1, g2):
        """export_setup (src/trusted_setup.cpp:256-287)."""
        out = [struct.pack("<Q", len(g1))]
        out += [self.serialize_g1(P) for P in g1]
        out += [self.serialize_g2(Q) for Q in g2]
        return b"".join(out)

    def srs_from_file_bytes(self, raw):
        """trusted_setup(const string&) (src/trusted_setup.cpp:76-121)."""
        M = self.M
        (n,) = struct.unpack_from("<Q", raw, 0)
        off = 8
        g1, g2 = [], []
        for _ in range(n):
            (ln,) = struct.unpack_from("<I", raw, off); off += 4
            assert ln == 2 * M + 1 and raw[off] == 4
            g1.append((int.from_bytes(raw[off + 1:off + 1 + M], "big"),
                       int.from_bytes(raw[off + 1 + M:off + 1 + 2 * M], "big")))
            off += ln
        for _ in range(n):
            (ln,) = struct.unpack_from("<I", raw, off); off += 4
            assert ln == 4 * M + 1 and raw[off] == 4
            v = [int.from_bytes(raw[off + 1 + j * M:off + 1 + (j + 1) * M], "big") for j in range(4)]
            g2.append(((v[1], v[0]), (v[3], v[2])))
            off += ln
        return g1, g2

    def setup(self, s, n):
        """trusted_setup(int) with a KNOWN secret (src/trusted_setup.cpp:21-74, 123-135)."""
        g1, g2, si = [], [], 1
        for _ in range(n):
            g1.append(self.mul(si, self.g1))
            g2.append(self.mul(si, self.g2, g2=True))
            si = si * s % self.r
        return g1, g2

    # ------------------------------------------------------------------ Fr polynomials (coeff lists, low first)
    def poly_trim(self, P):
        P = [c % self.r for c in P]
        while P and P[-1] == 0:
            P.pop()
        return P

    def poly_eval(self, P, x):
        y = 0
        for c in reversed(P):
            y = (y * x + c) % self.r
        return y

    def poly_eval_range(self, P, off, k):
        """evaluate_polynomial_points (src/util.cpp:186-211): y_i = P(off+i)."""
        return [self.poly_eval(P, (off + i) % self.r) for i in range(k)]

    def vanishing_poly(self, off, k):
        """Z = prod (x - (off+i)) — build_linear_roots_tree (src/util.cpp:269-284)."""
        Z = [1]
        for i in range(k):
            a = (off + i) % self.r
            nz = [0] * (len(Z) + 1)
            for j, c in enumerate(Z):
                nz[j + 1] = (nz[j + 1] + c) % self.r
                nz[j] = (nz[j] - a * c) % self.r
            Z = nz
        return Z

    def poly_divmod(self, P, Z):
        r = self.r
        P = self.poly_trim(P)
        Z = self.poly_trim(Z)
        if len(P) < len(Z):
            return [], P
        R = list(P)
        inv = pow(Z[-1], -1, r)
        q = [0] * (len(P) - len(Z) + 1)
        for i in range(len(q) - 1, -1, -1):
            c = R[i + len(Z) - 1] * inv % r
            q[i] = c
            if c:
                for j, z in enumerate(Z):
                    R[i + j] = (R[i + j] - c * z) % r
        return q, self.poly_trim(R[:len(Z) - 1])

    def poly_quotient(self, P, off, k):
        """q = (P - I) / Z (src/trusted_setup.cpp:218-225) with I = P mod Z: returns (q, I, Z)."""
        Z = self.vanishing_poly(off, k)
        q, I = self.poly_divmod(P, Z)
        return q, I, Z

    def interpolate(self, xs, ys):
        """polyfit (src/util.cpp:180-184): the unique poly of degree < k through the points (Lagrange)."""
        r = self.r
        k = len(xs)
        res = [0] * k
        for i in range(k):
            num = [1]
            den = 1
            for j in range(k):
                if j == i:
                    continue
                nz = [0] * (len(num) + 1)
                for t, c in enumerate(num):
                    nz[t + 1] = (nz[t + 1] + c) % r
                    nz[t] = (nz[t] - xs[j] * c) % r
                num = nz
                den = den * (xs[i] - xs[j]) % r
            s = ys[i] * pow(den, -1, r) % r
            for t, c in enumerate(num):
                res[t] = (res[t] + s * c) % r
        return self.poly_trim(res)

    def serialize_poly(self, P):
        """serialize_ZZ_pX (src/util.cpp:118-144): long degree | per coeff [u8 nbytes | LE magnitude]."""
        P = self.poly_trim(P)
        out = [struct.pack("<q", len(P) - 1)]
        for c in P:
            nb = (c.bit_length() + 7) // 8
            out.append(bytes([nb]) + c.to_bytes(nb, "little"))
        return b"".join(out)

    # ------------------------------------------------------------------ blobs (src/blob.cpp)
    def blob_from_string(self, s, offset=0):
        data = s.encode("latin-1") if isinstance(s, str) else s
        # `ZZ_y = s[i]` converts a (signed) char; the reference's tests only use ASCII
        return [((i + offset) % self.r, (c if c < 128 else c - 256) % self.r) for i, c in enumerate(data)]

    def blob_from_bytes(self, data, byte_offset, byte_length, chunk_size):
        if chunk_size > self.SB - 1:
            raise ValueError("chunk_size must be at most MAX_CHUNK_BYTES.")
        if byte_offset % chunk_size:
            raise ValueError("byte_offset is not a multiple of chunk_size.")
        if byte_length % chunk_size:
            raise ValueError("byte_length is not a multiple of chunk_size.")
        co, cl = byte_offset // chunk_size, byte_length // chunk_size
        return [((co + i) % self.r, int.from_bytes(data[i * chunk_size:(i + 1) * chunk_size], "little") % self.r)
                for i in range(cl)]

    # ------------------------------------------------------------------ the KZG operations
    def create_commit(self, g1, P):
        P = self.poly_trim(P)
        if len(P) >= len(g1):
            raise ValueError("polynomial degree be at most one less than the setup size (num_coeffs)")
        return self.msm(P, g1)

    def create_proof(self, g1, P, off, k):
        if k < 1:
            raise ValueError("chunk_length must be 1 or greater")
        q, _, _ = self.poly_quotient(P, off, k)
        return self.msm(q, g1)


def get_curve(name):
    return Curve({"bn158": "BN158", "bn254": "BN254", "bls12381": "BLS12381"}.get(name.lower(), name))
