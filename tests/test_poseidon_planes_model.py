"""Host model of the device Poseidon schedule (tools/gen_poseidon_constants.py `poseidon_planes`): full rounds on
u64 lanes, partial rounds on resident 22-bit limb planes with the limb normalisation of csrc/poseidon.cuh
`poseidon_norm22`.  Checks the model against the naive permutation (plonky2 hash/poseidon.rs `poseidon`), the
32-bit bounds the wrapping arithmetic relies on, and that the generated header carries the constants the model
uses."""
import importlib.util
import os
import random
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gen():
    spec = importlib.util.spec_from_file_location("gen_poseidon_constants",
                                                  os.path.join(ROOT, "tools", "gen_poseidon_constants.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_planes_schedule_matches_naive_permutation():
    g = _gen()
    rc = g.round_constants()
    rnd = random.Random(20)
    stats = [0]
    states = [[0] * 12, [g.P - 1] * 12, list(range(12)), [1 << 63] * 12, [0xFFFFFFFF] * 12,
              [0xFFFFFFFF00000000] * 12]
    states += [[rnd.randrange(g.P) for _ in range(12)] for _ in range(40)]
    for st in states:
        assert g.poseidon_planes(st, rc, stats) == g.poseidon_naive(st, rc)
    assert stats[0] < (1 << 31)  # plane outputs stay inside the signed 32-bit range the carries assume


def test_norm22_bounds_and_value():
    g = _gen()
    rnd = random.Random(21)
    top = 265 << 22  # plane outputs: at most (256 + 8) * (2^22 - 1) + a 22-bit constant
    for t in range(2000):
        y = [rnd.randrange(top) for _ in range(3)] if t else [top - 1] * 3
        x = g.norm22(list(y))
        assert 0 <= x[0] < (1 << 22) and 0 <= x[1] < (1 << 22) and 0 <= x[2] < (1 << 20) + (1 << 11)
        val = lambda l: (l[0] + (l[1] << 22) + (l[2] << 44)) % g.P
        assert val(x) == val(y)


def test_header_carries_the_plane_constants():
    g = _gen()
    rc = g.round_constants()
    hdr = open(os.path.join(ROOT, "0-kno-blobstreamx_b200", "csrc", "poseidon_rc.h")).read()
    body = hdr.split("#define POSEIDON_POST30_L22_LIST", 1)[1].split("#endif", 1)[0]
    vals = [int(v, 16) for v in re.findall(r"0x([0-9a-f]+)u", body)]
    assert len(vals) == 30 * 36
    for r in (0, 3, 4, 25, 28):
        want = [(c >> (22 * k)) & 0x3FFFFF for k in range(3) for c in rc[12 * (r + 1):12 * (r + 2)]]
        assert vals[36 * r:36 * (r + 1)] == want
    assert vals[36 * 29:] == [0] * 36
