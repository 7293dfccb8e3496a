"""Decimal -> double conversion of the GPU scene ingest (csrc/obj_number.cuh), compiled for the host with g++ and checked
against glibc strtod — the function behind the reference's atof calls (MTPC/sceneManagement.cpp:87-123).  Bit-exact."""
import ctypes
import os
import random
import struct
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_libc = ctypes.CDLL("libc.so.6")
_libc.strtod.restype = ctypes.c_double
_libc.strtod.argtypes = [ctypes.c_char_p, ctypes.c_void_p]
_libc.atol.restype = ctypes.c_long
_libc.atol.argtypes = [ctypes.c_char_p]


def _bits(x):
    return struct.unpack("<Q", struct.pack("<d", x))[0]


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("objnum") / "harness")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "cpu", "obj_number_harness.cpp")])

    def run(tokens):
        out = subprocess.run([exe], input="\n".join(tokens) + "\n", capture_output=True, text=True, check=True).stdout.split("\n")
        rows = [r.split() for r in out if r]
        assert len(rows) == len(tokens)
        return [(int(r[0]), int(r[1], 16), int(r[2]), int(r[3], 16), int(r[4])) for r in rows]
    return run


def _tokens(seed):
    rng = random.Random(seed)
    t = []
    for _ in range(60000):      # what OBJ writers print: fixed decimals, %g, repr
        x = rng.uniform(-100, 100) * 10 ** rng.randint(-6, 3)
        t += ["%.6f" % x, "%g" % x, "%.9g" % x, repr(x), "%.17g" % x, "%.4e" % x]
    for _ in range(20000):      # full double range incl. subnormals, through the 896-bit path
        x = struct.unpack("<d", struct.pack("<Q", rng.getrandbits(64) & 0x7FEFFFFFFFFFFFFF))[0]
        t += [repr(x), "%.20e" % x, "%.30g" % x]
    for _ in range(20000):      # random digit strings with random exponents
        nd = rng.randint(1, 38)
        digs = "".join(rng.choice("0123456789") for _ in range(nd))
        dot = rng.randint(0, nd)
        s = rng.choice(["", "-", "+"]) + digs[:dot] + rng.choice([".", ""]) * (dot < nd or rng.random() < 0.5) + digs[dot:]
        if "." not in s and dot < nd:
            s = s[:len(s) - (nd - dot)] + "." + digs[dot:]
        t.append(s + rng.choice(["", "e%d" % rng.randint(-350, 320), "E+%d" % rng.randint(0, 30), "e-%d" % rng.randint(0, 30)]))
    return t


def test_matches_strtod_bit_for_bit(harness):
    toks = _tokens(7)
    res = harness(toks)
    deferred = 0
    for tok, (sf, bf, sb, bb, _) in zip(toks, res):
        want = _bits(_libc.strtod(tok.encode(), None))
        assert sb == 0 and bb == want, (tok, hex(bb), hex(want))
        if sf == 0:
            assert bf == want, (tok, hex(bf), hex(want))
        else:
            assert sf == 1
            deferred += 1
    # the main kernel's own paths cover everything OBJ writers print with <= 19 digits and a small exponent
    assert all(r[0] == 0 for r in res[:360000])
    assert deferred < len(toks) // 4


def test_halfway_and_boundary_cases(harness):
    toks = ["9007199254740993", "9007199254740992", "9007199254740995", "9007199254740993.0000000000000001",
            "4.9406564584124654e-324", "2.4703282292062327e-324", "2.4703282292062328e-324", "2.2250738585072011e-308",
            "2.2250738585072014e-308", "1.7976931348623157e308", "1.7976931348623158e308", "1.7976931348623159e308",
            "1e309", "1e-400", "1e400", "-1e400", "0e999999", "0.0", "-0.0", "-0", "+0", "1e23", "8.5e-324", "1e22", "1e-22",
            "123456789012345678", "1234567890123456789", "12345678901234567890", "0.1", "0.3", "1000000000000000000000",
            "1000000000000000000001", "0.000000000000000000000000000000000000001", "5e-324", "3e-324",
            "179769313486231580793728971405303415079934132710037826936173778980444968292764750946649017977587207096330286416692887910946555547851940402630657488671505820681908902000708383676273854845817711531764475730270069855571366959622842914819860834936475292719074168444365510704342711559699508093042880177904174497791.9999999999999999999999999999999999999999999999999999999999999999999999"]
    res = harness(toks)
    for tok, (sf, bf, sb, bb, _) in zip(toks, res):
        want = _bits(_libc.strtod(tok.encode(), None))
        if sb == 2:          # more than 40 significant digits: reported, never guessed
            assert sum(c.isdigit() for c in tok) > 40, tok
            continue
        assert sb == 0 and bb == want, (tok, hex(bb), hex(want))
        assert sf in (0, 1) and (sf == 1 or bf == want), tok


def test_prefix_semantics_and_unsupported_tokens(harness):
    ok = ["1.5abc", "1e", "1e+", "1e-", ".5", "5.", ".", "+.e1", "-", "--1", "1/2", "abc", "", "1.2.3", "-.5e1x", "1e5.5", "e5", "+-1",
          "12//3", "7/8/9"]
    res = harness(ok)
    for tok, (sf, bf, sb, bb, lv) in zip(ok, res):
        want = _bits(_libc.strtod(tok.encode(), None))
        assert sf == 0 and sb == 0 and bf == want and bb == want, tok
        assert lv == _libc.atol(tok.encode()), tok
    bad = ["inf", "-Infinity", "nan", "NaN(1)", "0x1p3", "-0X10"]
    for tok, (sf, _, sb, _, _) in zip(bad, harness(bad)):
        assert sf == 2 and sb == 2, tok
