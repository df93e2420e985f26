#!/usr/bin/env python3
"""Regenerate the 360 Poseidon-Goldilocks round constants offline and emit C headers.

The reference mount is empty, so the constants cannot be read
from plonky2's `hash/poseidon_goldilocks.rs`.  They are reproduced from the
published generation procedure: ChaCha8Rng seeded with
`seed_from_u64(0)`, 360 uniform samples in [0, p).  The result is pinned by the
published all-zeros known-answer vector of plonky2's `check_test_vectors`
(SURVEY.md section 4), asserted below before anything is written.

Usage: python tools/gen_poseidon_constants.py   (writes both headers in place)
"""
import os
import sys

P = 0xFFFFFFFF00000001
M32 = 0xFFFFFFFF
M64 = (1 << 64) - 1


def pcg32_seed_words(seed):
    """rand_core::SeedableRng::seed_from_u64 — PCG32 expansion to a 32-byte key."""
    words = []
    state = seed
    for _ in range(8):
        state = (state * 6364136223846793005 + 11634580027462260723) & M64
        x = (((state >> 18) ^ state) >> 27) & M32
        rot = state >> 59
        words.append(((x >> rot) | (x << ((32 - rot) & 31))) & M32 if rot else x)
    return words


def rotl(v, n):
    return ((v << n) | (v >> (32 - n))) & M32


def chacha_block(key, counter, rounds):
    st = [0x61707865, 0x3320646E, 0x79622D32, 0x6B206574] + list(key) + [
        counter & M32, (counter >> 32) & M32, 0, 0]
    w = list(st)

    def qr(a, b, c, d):
        w[a] = (w[a] + w[b]) & M32; w[d] = rotl(w[d] ^ w[a], 16)
        w[c] = (w[c] + w[d]) & M32; w[b] = rotl(w[b] ^ w[c], 12)
        w[a] = (w[a] + w[b]) & M32; w[d] = rotl(w[d] ^ w[a], 8)
        w[c] = (w[c] + w[d]) & M32; w[b] = rotl(w[b] ^ w[c], 7)

    for _ in range(rounds // 2):
        qr(0, 4, 8, 12); qr(1, 5, 9, 13); qr(2, 6, 10, 14); qr(3, 7, 11, 15)
        qr(0, 5, 10, 15); qr(1, 6, 11, 12); qr(2, 7, 8, 13); qr(3, 4, 9, 14)
    return [(w[i] + st[i]) & M32 for i in range(16)]


class ChaCha8Rng:
    def __init__(self, seed):
        self.key = pcg32_seed_words(seed)
        self.counter = 0
        self.buf = []

    def next_u32(self):
        if not self.buf:
            self.buf = chacha_block(self.key, self.counter, 8)
            self.counter += 1
        return self.buf.pop(0)

    def next_u64(self):
        lo = self.next_u32()
        hi = self.next_u32()
        return (hi << 32) | lo


def sample_field(rng):
    """rand 0.8 `gen_range(0..p)` for u64: widening-multiply rejection sampling."""
    while True:
        v = rng.next_u64()
        m = v * P
        if (m & M64) <= P - 1:
            return m >> 64


def round_constants():
    rng = ChaCha8Rng(0)
    return [sample_field(rng) for _ in range(360)]


MDS_CIRC = [17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20]
MDS_DIAG = [8] + [0] * 11


def mds(s):
    return [(sum(MDS_CIRC[i] * s[(i + r) % 12] for i in range(12)) + MDS_DIAG[r] * s[r]) % P
            for r in range(12)]


def poseidon_naive(state, rc):
    s = list(state)
    for r in range(30):
        s = [(s[i] + rc[12 * r + i]) % P for i in range(12)]
        if r < 4 or r >= 26:
            s = [pow(x, 7, P) for x in s]
        else:
            s[0] = pow(s[0], 7, P)
        s = mds(s)
    return s


KAT_ZERO = [0x3c18a9786cb0b359, 0xc4055e3364a246c3, 0x7953db0ab48808f4, 0xc71603f33a1144ca,
            0xd7709673896996dc, 0x46a84e87642f44ed, 0xd032648251ee0b3c, 0x1c687363b207df62,
            0xdf8565563e8045fe, 0x40f5b37ff4254dae, 0xd070f637b431067c, 0x1792b1c4342109d7]
FIRST12 = [0xb585f766f2144405, 0x7746a55f43921ad7, 0xb2fb0d31cee799b4, 0x0f6760a4803427d7,
           0xe10d666650f4e012, 0x8cae14cb07d09bf1, 0xd438539c95f63e9f, 0xef781c7ce35b4c3d,
           0xcdc4a239b0c44426, 0x277fa208bf337bff, 0xe17653a29da578a1, 0xc54302f225db2c76]


# ---------------------------------------------------------------------------------------------
# Fast partial rounds (same permutation, fewer multiplications).  The 22 partial rounds
#     x += c_p ; x_0 <- x_0^7 ; x <- M x
# are rewritten, as in the Poseidon paper's optimised form (plonky2 hash/poseidon.rs
# `partial_first_constant_layer`, `mds_partial_layer_init`, `mds_partial_layer_fast`), into
#     x += FIRST ; x[1:] <- INIT x[1:] ; for p: x_0 <- x_0^7 + K[p] ;
#                                              (x_0, x[1:]) <- (25 x_0 + V[p].x[1:], x[1:] + x_0 W[p])
# The factorisation is derived here from RC and the MDS matrix and checked against the naive
# permutation; it need not (and is not claimed to) reproduce plonky2's constant tables word for
# word -- only the permutation, which the KATs pin.
def mds_matrix():
    m = [[0] * 12 for _ in range(12)]
    for r in range(12):
        for i in range(12):
            m[r][(i + r) % 12] = (m[r][(i + r) % 12] + MDS_CIRC[i]) % P
        m[r][r] = (m[r][r] + MDS_DIAG[r]) % P
    return m


def mat_mul(a, b):
    n, k, m = len(a), len(b), len(b[0])
    return [[sum(a[i][t] * b[t][j] for t in range(k)) % P for j in range(m)] for i in range(n)]


def mat_vec(a, v):
    return [sum(a[i][j] * v[j] for j in range(len(v))) % P for i in range(len(a))]


def mat_inv(a):
    n = len(a)
    a = [list(row) + [1 if i == j else 0 for j in range(n)] for i, row in enumerate(a)]
    for c in range(n):
        piv = next(r for r in range(c, n) if a[r][c])
        a[c], a[piv] = a[piv], a[c]
        inv = pow(a[c][c], P - 2, P)
        a[c] = [x * inv % P for x in a[c]]
        for r in range(n):
            if r != c and a[r][c]:
                f = a[r][c]
                a[r] = [(x - f * y) % P for x, y in zip(a[r], a[c])]
    return [row[n:] for row in a]


def fast_partial_constants(rc):
    M = mds_matrix()
    Minv = mat_inv(M)
    # constants: push every partial-round constant vector back through the linear layers
    K = [0] * 22
    acc = list(rc[12 * 25:12 * 26])  # constants of the last partial round (global round 25)
    for p in range(21, 0, -1):
        d = mat_vec(Minv, acc)
        K[p - 1] = d[0]
        prev = rc[12 * (4 + p - 1):12 * (4 + p)]
        acc = [prev[0]] + [(prev[i] + d[i]) % P for i in range(1, 12)]
    first = acc
    # matrices: M_eff = SPARSE_p * DENSE'_p, DENSE' = diag(1, Mhat) moves back one round
    V, W = [None] * 22, [None] * 22
    Meff = M
    for p in range(21, -1, -1):
        Mhat = [row[1:] for row in Meff[1:]]
        w = [Meff[i][0] for i in range(1, 12)]
        v = Meff[0][1:]
        MhatInv = mat_inv(Mhat)
        vhat = [sum(v[t] * MhatInv[t][j] for t in range(11)) % P for j in range(11)]  # v^T Mhat^-1
        assert Meff[0][0] == 25
        V[p], W[p] = vhat, w
        Mprime = [[1] + [0] * 11] + [[0] + row for row in Mhat]
        Meff = mat_mul(Mprime, M)
    init = [row[1:] for row in Mprime[1:]]  # 11 x 11, applied to x[1:] before the first partial round
    return first, init, K, V, W


def poseidon_fast(state, rc, fc):
    first, init, K, V, W = fc
    s = list(state)
    for r in range(4):
        s = [pow((s[i] + rc[12 * r + i]) % P, 7, P) for i in range(12)]
        s = mds(s)
    s = [(s[i] + first[i]) % P for i in range(12)]
    s = [s[0]] + mat_vec(init, s[1:])
    for p in range(22):
        s0 = (pow(s[0], 7, P) + K[p]) % P
        d = (25 * s0 + sum(V[p][i] * s[1 + i] for i in range(11))) % P
        s = [d] + [(s[1 + i] + s0 * W[p][i]) % P for i in range(11)]
    for r in range(26, 30):
        s = [pow((s[i] + rc[12 * r + i]) % P, 7, P) for i in range(12)]
        s = mds(s)
    return s


def mds_plane_fft(x, k):
    """The multiplication-free MDS layer of csrc/poseidon_mds.cuh on one limb plane (wrapping u32):
    4-point DFT over the index 3a + b, three tiny 3-convolutions, inverse DFT."""
    A, Bm, R, I = [0] * 3, [0] * 3, [0] * 3, [0] * 3
    for b in range(3):
        x0, x1, x2, x3 = x[b], x[3 + b], x[6 + b], x[9 + b]
        A[b], Bm[b], R[b], I[b] = x0 + x2 + x1 + x3, x0 + x2 - x1 - x3, x0 - x2, x1 - x3
    T = sum(A)
    P1 = [16 * (T + A[2]), 16 * (T + A[0]), 16 * (T + A[1])]
    Pm = [8 * Bm[2] - Bm[0] - 2 * Bm[1], -8 * Bm[0] - Bm[1] - 2 * Bm[2], 2 * Bm[0] - 8 * Bm[1] - Bm[2]]
    Pr = [2 * R[0] - I[0] + R[1] - 16 * I[1] + R[2] + 4 * I[2],
          -4 * R[0] + I[0] + 2 * R[1] - I[1] + R[2] - 16 * I[2],
          16 * R[0] + I[0] - 4 * R[1] + I[1] + 2 * R[2] - I[2]]
    Pi = [R[0] + 2 * I[0] + 16 * R[1] + I[1] - 4 * R[2] + I[2],
          -R[0] - 4 * I[0] + R[1] + 2 * I[1] + 16 * R[2] + I[2],
          -R[0] + 16 * I[0] - R[1] - 4 * I[1] + R[2] + 2 * I[2]]
    o = [0] * 12
    for b in range(3):
        p, q = P1[b] + Pm[b], P1[b] - Pm[b]
        o[b], o[3 + b], o[6 + b], o[9 + b] = p + Pr[b] + k[b], q + Pi[b] + k[3 + b], p - Pr[b] + k[6 + b], q - Pi[b] + k[9 + b]
    o[0] += 8 * x[0]
    return [v & M32 for v in o]


def mds_fft_check():
    """Frequency-domain kernel of the circulant (must be the tiny integers the CUDA code hard-wires) and
    the limb-plane transform against the dense form."""
    K = [MDS_CIRC[(-m) % 12] for m in range(12)]
    Kb = [[K[3 * a + b] for a in range(4)] for b in range(3)]
    assert [sum(k) // 4 for k in Kb] == [16, 32, 16]
    assert [(k[0] - k[1] + k[2] - k[3]) // 4 for k in Kb] == [-1, -8, 2]
    assert [((k[0] - k[2]) // 2, (k[1] - k[3]) // 2) for k in Kb] == [(2, 1), (-4, -1), (16, -1)]
    import random
    rnd = random.Random(7)
    for t in range(200):
        x = [rnd.randrange(1 << 22) if t else (1 << 22) - 1 for _ in range(12)]
        k = [rnd.randrange(1 << 22) if t else (1 << 22) - 1 for _ in range(12)]
        want = [sum(MDS_CIRC[i] * x[(i + r) % 12] for i in range(12)) + MDS_DIAG[r] * x[r] + k[r] for r in range(12)]
        assert max(want) <= M32 and want == mds_plane_fft(x, k)



def split22(v):
    return [v & 0x3FFFFF, (v >> 22) & 0x3FFFFF, v >> 44]


def merge22_lazy(o):
    """poseidon_merge22 of csrc/poseidon.cuh: o0 + o1 2^22 + o2 2^44 (each < 2^32) -> some u64 representative"""
    assert all(0 <= v <= M32 for v in o)
    w = o[0] + (o[1] << 22) + (o[2] << 44)
    lo, w2 = w & ((1 << 64) - 1), w >> 64
    assert w2 < (1 << 13)
    lo += w2 * 0xFFFFFFFF
    if lo >> 64:
        lo = (lo & ((1 << 64) - 1)) + 0xFFFFFFFF
    assert lo < (1 << 64)
    return lo


def s32(v):
    v &= M32
    return v - (1 << 32) if v >> 31 else v


def norm22(y):
    """Limb normalisation between two plane-resident partial rounds (wrapping u32, arithmetic shifts):
    fold the bits above 2^64 back with 2^64 == 2^32 - 1, then propagate the limb carries."""
    y0, y1, y2 = y
    assert 0 <= y2 < (1 << 31)
    top = y2 >> 20
    y2 &= 0xFFFFF
    assert y1 + (top << 10) + (1 << 10) < (1 << 31)  # the carry out of limb 1 is read with an arithmetic shift
    y1 = (y1 + (top << 10)) & M32
    y0 = (y0 - top) & M32
    c0 = s32(y0) >> 22
    y0 &= 0x3FFFFF
    y1 = (y1 + c0) & M32
    c1 = s32(y1) >> 22
    y1 &= 0x3FFFFF
    y2 = (y2 + c1) & M32
    assert 0 <= y2 < (1 << 20) + (1 << 11)
    return [y0, y1, y2]


def poseidon_planes(state, rc, stats=None):
    """The shipped device schedule: full rounds on u64 lanes with the limb-plane MDS; the 22 partial rounds keep
    lanes 1..11 as 22-bit limb planes from round to round (dense MDS, no multiplications), only lane 0 is merged
    for its S-box.  post constants = the next round's RC as third addends of the planes."""
    post = [[split22(c) for c in rc[12 * (r + 1):12 * (r + 2)]] if r < 29 else [[0, 0, 0]] * 12 for r in range(30)]

    def planes(x, r):  # x: 12 x 3 limbs -> 12 x 3 limbs (unnormalised)
        ys = [mds_plane_fft([x[i][k] for i in range(12)], [post[r][i][k] for i in range(12)]) for k in range(3)]
        return [[ys[k][i] for k in range(3)] for i in range(12)]

    s = [(state[i] + rc[i]) % P for i in range(12)]
    for r in range(4):
        s = [pow(v, 7, P) for v in s]
        s = [merge22_lazy(y) % P for y in planes([split22(v) for v in s], r)]
    x = [split22(v) for v in s]
    y0 = x[0]
    for r in range(4, 26):
        x[0] = split22(pow(merge22_lazy(y0) % P, 7, P))
        y = planes(x, r)
        if stats is not None:
            stats[0] = max(stats[0], max(max(l) for l in y))
        y0 = y[0]
        x = [None] + [norm22(l) for l in y[1:]]
    s = [merge22_lazy(y0) % P] + [merge22_lazy(l) % P for l in x[1:]]
    for r in range(26, 30):
        s = [pow(v, 7, P) for v in s]
        s = [merge22_lazy(y) % P for y in planes([split22(v) for v in s], r)]
    return s


def c_list32(name, vals, per_line=8):
    out = "#define %s \\\n" % name
    for i in range(0, len(vals), per_line):
        out += "  " + ", ".join("0x%06xu" % c for c in vals[i:i + per_line])
        out += ", \\\n" if i + per_line < len(vals) else "\n"
    return out


def c_list(name, vals, per_line=4):
    out = "#define %s \\\n" % name
    for i in range(0, len(vals), per_line):
        out += "  " + ", ".join("0x%016xULL" % c for c in vals[i:i + per_line])
        out += ", \\\n" if i + per_line < len(vals) else "\n"
    return out


def emit(path, rc, guard, fc=None):
    with open(path, "w") as f:
        f.write("/* Generated by tools/gen_poseidon_constants.py -- do not edit.\n"
                " * Poseidon-Goldilocks width-12 round constants (plonky2 hash/poseidon_goldilocks.rs\n"
                " * ALL_ROUND_CONSTANTS), regenerated offline and pinned by the all-zeros KAT. */\n")
        f.write("#ifndef %s\n#define %s\n" % (guard, guard))
        f.write("#define POSEIDON_N_RC 360\n")
        f.write(c_list("POSEIDON_RC_LIST", rc))
        if fc is not None:
            first, init, K, V, W = fc
            f.write("/* fast partial rounds: derived from RC and the MDS matrix, checked against the naive form */\n")
            f.write(c_list("POSEIDON_FAST_FIRST_LIST", first))
            f.write(c_list("POSEIDON_FAST_INIT_LIST", [x for row in init for x in row]))  # [11][11] row-major
            f.write(c_list("POSEIDON_FAST_K_LIST", K))
            f.write(c_list("POSEIDON_FAST_V_LIST", [x for row in V for x in row]))  # [22][11]
            f.write(c_list("POSEIDON_FAST_W_LIST", [x for row in W for x in row]))  # [22][11]
            # constants added right after the MDS of full round j (j = 0..7): the next round's RC, the
            # fast-partial FIRST vector after round 3, nothing after the last round; 32-bit halves as u64
            post = [rc[12:24], rc[24:36], rc[36:48], first, rc[12 * 27:12 * 28], rc[12 * 28:12 * 29],
                    rc[12 * 29:12 * 30], [0] * 12]
            flat = [x for row in post for x in row]
            f.write(c_list("POSEIDON_POST_LO_LIST", [x & M32 for x in flat]))  # [8][12]
            f.write(c_list("POSEIDON_POST_HI_LIST", [x >> 32 for x in flat]))  # [8][12]
            # the same vectors cut into 22-bit limb planes for the multiplication-free MDS: [8][3][12]
            l22 = [(x >> (22 * k)) & 0x3FFFFF for row in post for k in range(3) for x in row]
            f.write(c_list32("POSEIDON_POST_L22_LIST", l22))
            # plane-resident schedule: after the MDS of round r (all 30) the next round's RC, [30][3][12]
            post30 = [rc[12 * (r + 1):12 * (r + 2)] if r < 29 else [0] * 12 for r in range(30)]
            f.write(c_list32("POSEIDON_POST30_L22_LIST",
                             [(x >> (22 * k)) & 0x3FFFFF for row in post30 for k in range(3) for x in row]))
        f.write("#endif\n")


def main():
    rc = round_constants()
    assert rc[:12] == FIRST12, "first 12 constants differ from the recalled upstream values"
    assert max(rc) < 0xfffeeac900011537
    assert poseidon_naive([0] * 12, rc) == KAT_ZERO, "all-zeros KAT mismatch"
    fc = fast_partial_constants(rc)
    mds_fft_check()
    import random
    rnd = random.Random(1)
    for t in range(6):
        st = [0] * 12 if t == 0 else [P - 1] * 12 if t == 1 else [rnd.randrange(P) for _ in range(12)]
        assert poseidon_fast(st, rc, fc) == poseidon_naive(st, rc), "fast partial rounds differ from naive"
    stats = [0]
    for t in range(300):
        st = [0] * 12 if t == 0 else [P - 1] * 12 if t == 1 else [rnd.randrange(P) for _ in range(12)]
        assert poseidon_planes(st, rc, stats) == poseidon_naive(st, rc), "plane-resident partial rounds differ from naive"
    assert stats[0] < (1 << 31)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    emit(os.path.join(root, "oracle", "poseidon_rc.h"), rc, "ORACLE_POSEIDON_RC_H")
    emit(os.path.join(root, "0-kno-blobstreamx_b200", "csrc", "poseidon_rc.h"), rc, "BSX_POSEIDON_RC_H", fc)
    print("ok: 360 constants, KAT(zeros) matches, fast partial rounds == naive")
    for name, inp in (("0..11", list(range(12))), ("p-1", [P - 1] * 12)):
        print(name, " ".join("%016x" % x for x in poseidon_naive(inp, rc)))
    print("max |init|", max(max(r) for r in fc[1]), "K[21]", fc[2][21])


if __name__ == "__main__":
    sys.exit(main())
