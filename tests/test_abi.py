"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/bsx_b200.h declares, and the product fails loudly without a GPU (no fallback)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
import __graft_entry__ as G


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "bsx_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bsx_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    if not os.path.exists(pkg.library_path()):
        G.build()
    L = ctypes.CDLL(pkg.library_path())
    declared = _header_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(L, name), "libbsx_b200.so does not export %s" % name
    assert sorted(pkg.EXPORTED_SYMBOLS) == declared, "binding.py and the header disagree"


def test_no_cpu_fallback(pkg):
    L = pkg.lib()
    if L.bsx_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.BsxError):
        pkg.Context(0)


def test_product_never_imports_oracle():
    """The shipped path must not import, link, or execute anything under oracle/ ("oracle" as plonky2's
    word for a committed PolynomialBatch -- `oracles`, `oracle_index` -- is fine)."""
    pkg_dir = os.path.join(ROOT, "0-kno-blobstreamx_b200")
    banned = re.compile(r"(from\s+oracle\b|import\s+oracle\b|oracle/|oracle\.py|liboracle|plonky2_oracle|\borc_\w+)")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or f == "Makefile":
                src = open(os.path.join(dirpath, f)).read().replace("fri/oracle.rs", "")
                m = banned.search(src)
                assert m is None, "%s references the test oracle: %r" % (f, m.group(0))


def test_fri_reduction_strategy(pkg):
    # ConstantArityBits(4, 5) at standard_recursion_config
    assert pkg.fri_reduction_arity_bits(20, 3, 4) == [4, 4, 4, 4]
    assert pkg.fri_reduction_arity_bits(14, 3, 4) == [4, 4, 4]
    assert pkg.fri_reduction_arity_bits(5, 3, 4) == []
    p = pkg.FriConfig().fri_params(20)
    assert p.lde_bits() == 23 and p.final_poly_len() == 16


def test_header_is_plain_c_and_links():
    """include/bsx_b200.h compiles as C11 (-Wall -Wextra -Werror) and a C program links against the library using
    nothing else.  Without a GPU the program must fail loudly at bsx_ctx_create (no CPU fallback)."""
    import subprocess
    import c_harness_build
    exe = c_harness_build.build()
    r = subprocess.run([exe, "5", "3", "1", "1", "7"], capture_output=True, text=True, timeout=120)
    if r.returncode != 0:
        assert r.returncode == 1 and "no CPU fallback" in r.stderr, r.stderr
    else:
        assert r.stdout.startswith("cap ")


def _gen_rust_sys():
    import importlib.util
    spec = importlib.util.spec_from_file_location("gen_rust_sys", os.path.join(ROOT, "tools", "gen_rust_sys.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_rust_sys_bindings_match_the_header():
    """rust/bsx-b200-sys/src/lib.rs is what tools/gen_rust_sys.py derives from the header (no rustc here to run
    bindgen): one `pub fn` per exported symbol, nothing stale."""
    g = _gen_rust_sys()
    text, names = g.render()
    assert open(g.OUT).read() == text, "run `python tools/gen_rust_sys.py`: the committed bindings are stale"
    assert sorted(names) == _header_symbols()
    for name in names:
        assert len(re.findall(r"pub fn %s\(" % name, text)) == 1


def test_rust_shim_calls_exist_with_the_header_arity():
    """Every `sys::bsx_*(..)` call in rust/plonky2-bsx-shim names a function of the header and passes as many arguments
    as the header declares (the crate cannot be compiled in this image)."""
    import glob
    g = _gen_rust_sys()
    consts, opaque, structs, funcs = g.parse(open(g.HEADER).read())
    arity = {n: len(p) for n, p, _ in funcs}
    types = {o[1] for o in opaque} | {s[0] for s in structs}
    n_calls = 0
    for path in glob.glob(os.path.join(ROOT, "rust", "plonky2-bsx-shim", "src", "*.rs")):
        src = open(path).read()
        for m in re.finditer(r"sys::(bsx_\w+)\s*\(", src):
            name = m.group(1)
            if name in types:
                continue
            assert name in arity, "%s calls %s, which the header does not declare" % (os.path.basename(path), name)
            i, depth, commas, body = m.end(), 1, 0, ""
            while depth:
                ch = src[i]
                depth += ch in "([{"
                depth -= ch in ")]}"
                if depth == 1 and ch == ",":
                    commas += 1
                elif depth >= 1:
                    body += ch
                i += 1
            n_args = commas + (1 if body.strip() else 0)
            assert n_args == arity[name], "%s: %s called with %d arguments, declared with %d" % (
                os.path.basename(path), name, n_args, arity[name])
            n_calls += 1
    assert n_calls >= 20
