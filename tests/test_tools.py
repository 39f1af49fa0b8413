"""SURVEY 8(f).4 host utilities: stringCheck (C ABI edsbwt_string_check + the drop-in executable) against the unmodified
reference's stringCheck compiled by oracle/Makefile (when /root/reference is present) and against committed expectations;
the MSA pattern extractor against the reference script's contract.  Host-only code: runs without a GPU."""
import importlib
import os
import random
import subprocess
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))

CASES = [
    (b"ACGT{A,C}GT{T,}A", b"{ACGT}{A,C}{GT}{T,}{A}"),
    (b"{ACGT}{A,C}{GT}", b"{ACGT}{A,C}{GT}"),
    (b"A{C,G}", b"{A{C,G}"),                     # the FIRST character is never looked ahead from (stringCheck.cpp:29-37): reference quirk
    (b"{A,C}{G,T}", b"{A,C}{G,T}"),
    (b"AC<comment>GT{A,}T", b"{ACGT}{A,}{T}"),
    (b"AC<c>{A,}T<d>x", b"{AC}{A,}{Tx}"),
    (b"AC<c>{A,}T", b"{AC}{A,}{T}"),
    (b"ACGT\n", b"{ACGT\n}"),                    # the trailing newline is a letter for the reference
    (b"A", b"{A"),                               # single character: the loop that would close the bracket never runs
    (b"{A}", b"{A}"),
]


@pytest.fixture(scope="module")
def eb():
    m = importlib.import_module("eds-bwt_b200")
    if not os.path.exists(m.LIB_PATH):
        importlib.import_module("__graft_entry__").build()
    return m


@pytest.mark.parametrize("src,want", CASES)
def test_string_check_expected(eb, tmp_path, src, want):
    inp, out = tmp_path / "in.txt", tmp_path / "out.eds"
    inp.write_bytes(src)
    eb.string_check(str(inp), str(out))
    assert out.read_bytes() == want


def test_string_check_reserved_character(eb, tmp_path):
    inp = tmp_path / "in.txt"
    inp.write_bytes(b"ACGZT")
    with pytest.raises(eb.EdsbwtError) as ei:
        eb.string_check(str(inp), str(tmp_path / "o"))
    assert ei.value.code == -2
    with pytest.raises(eb.EdsbwtError) as ei:
        eb.string_check(str(tmp_path / "missing"), str(tmp_path / "o"))
    assert ei.value.code == -1


def test_string_check_vs_unmodified_reference(eb, oracle_built, tmp_path):
    ref = os.path.join(oracle_built.REF_DIR, "stringCheck")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref/stringCheck not built (no /root/reference here)")
    rng = random.Random(5)
    exe = os.path.join(os.path.dirname(eb.__file__), "stringCheck")
    for case in range(60):
        parts = []
        for _ in range(rng.randint(1, 12)):
            r = rng.random()
            if r < 0.45:
                parts.append("".join(rng.choice("ACGT") for _ in range(rng.randint(1, 9))))
            elif r < 0.85:
                parts.append("{" + ",".join("".join(rng.choice("ACGT") for _ in range(rng.randint(0, 4))) for _ in range(rng.randint(1, 4))) + "}")
            else:
                parts.append("<" + "".join(rng.choice("acgt ") for _ in range(rng.randint(0, 5))) + ">")
        text = "".join(parts)
        if text.endswith(">") or len(text) < 2:
            text += "A"                     # a comment at EOF makes the reference loop forever on some libstdc++ versions
        if rng.random() < 0.3:
            text += "\n"
        inp = tmp_path / ("c%d.txt" % case)
        inp.write_text(text)
        r1 = subprocess.run([ref, str(inp), str(tmp_path / ("ref%d" % case))], capture_output=True, text=True, timeout=20)
        r2 = subprocess.run([exe, str(inp), str(tmp_path / ("our%d" % case))], capture_output=True, text=True, timeout=20)
        assert (tmp_path / ("ref%d.eds" % case)).read_bytes() == (tmp_path / ("our%d.eds" % case)).read_bytes(), text
        assert r1.stdout.replace("ref%d" % case, "X") == r2.stdout.replace("our%d" % case, "X")
    # reserved character: both exit(1) with the same message
    inp = tmp_path / "bad.txt"
    inp.write_text("ACZGT")
    r1 = subprocess.run([ref, str(inp), str(tmp_path / "b1")], capture_output=True, text=True)
    r2 = subprocess.run([exe, str(inp), str(tmp_path / "b2")], capture_output=True, text=True)
    assert r1.returncode == r2.returncode == 1 and r1.stderr == r2.stderr
    # usage
    r1 = subprocess.run([ref], capture_output=True, text=True)
    r2 = subprocess.run([exe], capture_output=True, text=True)
    assert r1.returncode == r2.returncode == 1
    assert r1.stderr.replace(ref, "X") == r2.stderr.replace(exe, "X")


def test_extract_patterns_from_msa(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "eds-bwt_b200", "tools"))
    import extract_patterns_from_msa as X
    msa = tmp_path / "a.msa"
    msa.write_text(">s1\nAC-GT\nAC--G\n>s2\n--TTTTGG\n>s3\nAC\n")
    seqs = X.load_sequences(str(msa))
    assert seqs == ["ACGTACG", "TTTTGG", "AC"]
    out = tmp_path / "p.txt"
    X.main([str(msa), str(out), "-l", "4", "-n", "5", "--seed", "3"])
    text = out.read_text()
    pats = text.split("\n")
    assert len(pats) == 5 and not text.endswith("\n") and all(len(p) == 4 for p in pats)
    windows = [s[i:i + 4] for s in seqs for i in range(len(s) - 3)]
    assert all(p in windows for p in pats)
    X.main([str(msa), str(out), "-l", "4", "-n", "7", "--seed", "1"])          # all 7 windows: a permutation of them
    assert sorted(out.read_text().split("\n")) == sorted(windows)
    with pytest.raises(ValueError):
        X.main([str(msa), str(out), "-l", "4", "-n", "8"])


def test_cli_report_formatter(tmp_path):
    """The executables format "Found N starting positions ..." by hand (10 M lines per batch): same bytes as printf."""
    sys.path.insert(0, ROOT)
    import __graft_entry__ as G
    G.build()
    exe = str(tmp_path / "cli_format_check")
    libdir = os.path.join(ROOT, "eds-bwt_b200")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-o", exe,
                    os.path.join(ROOT, "tests", "native", "cli_format_check.cpp"), "-L", libdir, "-ledsbwt_cuda",
                    "-Wl,-rpath," + libdir], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "format ok", r.stdout
