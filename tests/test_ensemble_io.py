"""ensemble reader / writer (host code behind the C ABI), /root/reference/src/
strucreader.f90: rdensembleparam 268-309, rdensemble 433-481, coordline 2005-2022, e2i 2120-2147,
grepenergy 2225-2260, wrxyz 1679-1698.  No Fortran compiler exists here, so the checks are: the
reference's own example trajectory parsed by our reader == the golden arrays parsed independently by
tests/make_golden.py (numpy), known-answer lines for grepenergy / e2i derived by hand from the
Fortran list-directed rules, exact byte layout of the writer's formats, and write -> read round
trips."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from crest_b200 import synth

TRJ = "/root/reference/examples/expl-3/xtb.trj"


@pytest.fixture(scope="module")
def cb(cuda_lib):
    import crest_b200
    return crest_b200


@pytest.mark.skipif(not os.path.exists(TRJ), reason="reference example not present on this machine")
def test_reads_reference_example(cb):
    g = np.load(os.path.join(GOLDEN, "expl3.npz"))
    nat, nall = cb.rdensembleparam(TRJ)
    assert (nat, nall) == (g["raw_xyz"].shape[1], g["raw_xyz"].shape[0])
    ens = cb.rdensemble(TRJ)
    assert np.array_equal(ens["at"], g["at"])
    assert np.array_equal(ens["xyz"], g["raw_xyz"])          # same decimal strings -> same doubles
    assert np.array_equal(ens["er"], g["raw_energy"])
    assert ens["comments"][0].lstrip().startswith("energy:")


def test_grepenergy_known_answers(cb):
    kat = [
        (" energy: -1.629564241025 gnorm: 0.000113350403 xtb: 6.2.2 (89a525f)   !mtd1", -1.629564241025),
        ("  -12.5   0.3", -12.5),                   # first float of the line
        ("Etot= -3.25d0 x", -3.25),                 # conversion errors drop one character at a time
        ("energy=abc", 0.0),                        # keyword present, unreadable value -> 0
        ("blah energy= 7.5e-1, 3", 0.75),
        ("", 0.0), ("abc", 0.0),
        ("SCF done -5.0E+01 !x", -50.0),
        (" ,3.0", 0.0),                             # null first item: energy keeps its initial 0.0
        ("     -100.12345678901234       0.25", -100.12345678901234),   # what cregen writes (er, p)
        ("1.5+3", 1500.0),                          # Fortran exponent without a letter
    ]
    for line, want in kat:
        assert cb.grepenergy(line) == want, line


def test_e2i_known_answers(cb):
    kat = {"c": 6, "C": 6, "Cl": 17, "CL": 17, "cl1": 17, "6": 6, "h": 1, "D": 1, "T": 1, "13C": 6,
           "Xx": 0, "og": 118, "119": 0, "Fe2+": 26, "na_1": 11, "n_a": 0, "he": 2}
    for lab, z in kat.items():
        assert cb.e2i(lab) == z, lab


def test_writer_byte_layout(cb, tmp_path):
    at = np.array([6, 17, 1, 118], np.int32)
    xyz = np.array([[[0.0, -1.5, 2.25], [10.123456789012, -0.00000000004, 1e3], [1, 2, 3], [-1, -2, -3]]])
    f = tmp_path / "one.xyz"
    cb.wrensemble(f, at, xyz, comments=["   -5.25000000000000       1.00000000000000     "])
    lines = f.read_text().split("\n")
    assert lines[0] == "  4"                                             # '(2x,i0)'
    assert lines[1] == "   -5.25000000000000       1.00000000000000"     # trim(comment)
    # '(1x,a2,1x,3f20.10)': blank, symbol left-justified in 2, blank, three 20-wide fixed fields
    def fline(sym, v):
        return " " + sym.ljust(2) + " " + "".join(f"{x:20.10f}" for x in v)
    assert lines[2] == fline("C", xyz[0, 0]) == " C  " + " " * 8 + "0.0000000000" + " " * 7 + "-1.5000000000" + " " * 8 + "2.2500000000"
    assert lines[3] == fline("Cl", xyz[0, 1])
    assert lines[5] == fline("Og", xyz[0, 3])
    assert all(len(l) == 64 for l in lines[2:6])                         # 1 + 2 + 1 + 3*20


@pytest.mark.parametrize("nall,nat", [(1, 3), (37, 50), (400, 12)])
def test_round_trip(cb, tmp_path, nall, nat):
    ens = synth.make_ensemble(nall, nat, max(3, nat // 2), seed=nall)
    f = tmp_path / "ens.xyz"
    comments = [f"  {e:.12f}   0.5  !origin{i}" for i, e in enumerate(ens["er"])]
    cb.wrensemble(f, ens["at"], ens["xyz"], comments=comments)
    assert cb.rdensembleparam(f) == (nat, nall)
    back = cb.rdensemble(f)
    assert np.array_equal(back["at"], ens["at"])
    assert np.abs(back["xyz"] - ens["xyz"]).max() <= 0.5e-10 * 1.0001     # f20.10
    assert np.abs(back["er"] - ens["er"]).max() < 1e-12
    assert back["comments"] == [c.rstrip() for c in comments]
    # energy-only comment, then the reader feeds crest_cuda_cregen_esort's input
    cb.wrensemble(f, ens["at"], ens["xyz"], er=ens["er"])
    again = cb.rdensemble(f, want_comments=False)
    assert np.abs(again["er"] - ens["er"]).max() < 1e-13


def test_reader_irregular_file(cb, tmp_path):
    f = tmp_path / "messy.xyz"
    f.write_text(
        "\n"                       # blank records before a count are skipped (list-directed read)
        "3\n"
        " energy= -1.5 kcal\n"
        "  O   0.0 0.0 0.1\n"
        "  H   0.0 0.7 -0.4\n"
        "  H   0.0 -0.7 -0.4\n"
        "2\n"                      # a structure with another atom count: skipped by the reader
        "foreign\n"
        " H 0 0 0\n"
        " H 0 0 0.7\n"
        "not a number\n"           # conversion error on a count line: `cycle`
        "   3\n"
        "-2.5d0\n"
        " 0.0, 0.0, 0.2  o\n"       # "x y z sym" order, comma separators, lower-case label
        " 0.0  0.7 -0.3  1\n"       # numeric label
        " 0.0 -0.7 -0.3  h\n")
    nat, nall = cb.rdensembleparam(f)
    # rdensembleparam reads atom records as `str,x,y,z` only (strucreader.f90:297): it counts the
    # 2-atom structure but not the "x y z sym" one -> 2; the reader (coordline) then skips the
    # 2-atom structure (ensemble_strucskip) and reads the other two
    assert (nat, nall) == (3, 2)
    ens = cb.rdensemble(f)
    assert len(ens["er"]) == 2
    assert np.array_equal(ens["er"], [-1.5, -2.5])
    assert np.array_equal(ens["at"], [8, 1, 1])
    assert np.array_equal(ens["xyz"][1], [[0, 0, 0.2], [0, 0.7, -0.3], [0, -0.7, -0.3]])
    assert cb.rdensembleparam(tmp_path / "absent.xyz") == (0, 0)
    with pytest.raises(cb.CrestCudaError):
        cb.rdensemble(tmp_path / "absent.xyz")


def test_reader_truncated_file_is_an_error(cb, tmp_path):
    f = tmp_path / "cut.xyz"
    f.write_text("2\nc\n H 0 0 0\n H 0 0 1\n2\nc\n H 0 0 0\n")
    assert cb.rdensembleparam(f) == (2, 1)
    ens = cb.rdensemble(f)                 # asks for the one complete structure only
    assert ens["xyz"].shape == (1, 2, 3)


@pytest.mark.gpu
def test_device_parser_equals_record_reader(tmp_path):
    """Large files convert their atom records on the device (ensemble_fast.cu: exact fast path of the
    decimal -> binary64 conversion, one thread per record); records beyond the fast path are re-read on
    the host.  Bit-identical arrays to the record-by-record reader (CREST_CUDA_IO=host), for the
    reference's own '(1x,a2,1x,3f20.10)' layout and for awkward but legal records."""
    import crest_b200 as cb
    from crest_b200 import synth, _lib
    if _lib.lib().crest_cuda_device_count() == 0:
        pytest.skip("needs a CUDA device")
    ens = synth.make_config("C3", nall=4000, seed=8)
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    fn = str(tmp_path / "big.xyz")
    cb.wrensemble(fn, at, xyz, er)
    # rewrite some records in other legal forms: "x y z sym", exponents, d-exponents, 17+ digits, commas
    lines = open(fn).read().split("\n")
    rng = np.random.default_rng(1)
    nat = xyz.shape[1]
    for s in rng.integers(0, 4000, 300):
        j = int(rng.integers(0, nat))
        k = s * (nat + 2) + 2 + j
        sym, x, y, z = lines[k].split()
        form = int(rng.integers(0, 5))
        if form == 0:
            lines[k] = f"  {x}   {y}  {z}   {sym}"
        elif form == 1:
            lines[k] = f" {sym} {float(x):.12e} {float(y):.12E} {float(z):.9e}"
        elif form == 2:
            lines[k] = f" {sym} {float(x):.10e}".replace("e", "d") + f" {y} {z}"
        elif form == 3:
            lines[k] = f" {sym} {float(x):.25f} {y}, {z}"
        else:
            lines[k] = f"\t{sym}\t{x}\t{y}\t{z}\r"
    open(fn, "w").write("\n".join(lines))
    os.environ["CREST_CUDA_IO"] = "host"
    try:
        ph = cb.rdensembleparam(fn)
        h = cb.rdensemble(fn)
    finally:
        os.environ.pop("CREST_CUDA_IO")
    pd = cb.rdensembleparam(fn)
    d = cb.rdensemble(fn)
    # rdensembleparam reads `str x y z` (strucreader.f90:268-309): structures holding an "x y z sym" record
    # are not counted - by either path
    assert ph == pd and ph[0] == nat and 3900 < ph[1] < 4000
    assert np.array_equal(h["at"], d["at"]) and np.array_equal(h["er"], d["er"])
    assert np.array_equal(h["xyz"].view(np.int64), d["xyz"].view(np.int64))       # bit for bit
    assert h["comments"] == d["comments"]
    # a record that does not scan is reported like the record reader reports it
    lines[5 * (nat + 2) + 2 + 3] = " C 1.0 abc 2.0"
    open(fn, "w").write("\n".join(lines))
    with pytest.raises(cb.CrestCudaError, match="structure 6, atom 4"):
        cb.rdensemble(fn)
    assert cb.rdensembleparam(fn) == (nat, ph[1] - 1)
