"""CPU-side checks: the C-ABI library loads and exports every declared symbol, refuses to
run without a GPU (no CPU fallback), and the host mirror of the lab's interface (file
formats, loop control, generators) behaves like the reference."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import HAS_GPU, ROOT
from parallel_comp_labs_b200 import capi, generators, lab


def test_library_exports_every_declared_symbol():
    L = capi.lib()
    names = capi.declared_symbols()
    assert len(names) >= 20 and "nbody_step" in names and "nbody_rhs" in names
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert b"sm_100a" in L.nbody_version()


def test_library_has_sm100a_code_and_tma():
    so = capi.LIB_PATH
    elf = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout
    assert "sm_100a" in elf
    names = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    fn = [l.split()[-1] for l in names.splitlines() if "Function :" in l and "force_kernelId" in l][0]
    sass = subprocess.run(["cuobjdump", "-sass", "-fun", fn, so], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass          # cp.async.bulk (TMA engine) stages the j-tiles
    assert "MUFU.RSQ64H" in sass and "DFMA" in sass and "SYNCS" in sass


@pytest.mark.skipif(HAS_GPU, reason="this box has a GPU")
def test_no_cpu_fallback():
    with pytest.raises(capi.NBodyError) as e:
        capi.Context(16, 64)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)
    with pytest.raises(capi.NBodyError):
        capi.measure_fma_peak(64, 1.0)
    # the C++ host driver fails loudly too
    exe = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
    inp = os.path.join(ROOT, "tests", "golden", "_4body.txt")
    r = subprocess.run([exe, "--input", inp, "--system", "default_test", "--no-traj"],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr


def test_argument_errors_without_gpu():
    L = capi.lib()
    h = C.c_void_p()
    assert L.nbody_create(C.byref(h), 0, 64, 0, 0, 1, None) == 1          # n == 0
    assert L.nbody_create(C.byref(h), 8, 16, 0, 0, 1, None) == 1          # precision
    assert L.nbody_create(C.byref(h), 4, 64, 0, 0, 5, None) == 1          # world > n
    assert b"greater than the problem dimension" in L.nbody_last_error(None)
    assert h.value is None


def test_read_data_formats(tmp_path, golden):
    p = tmp_path / "b.txt"
    p.write_bytes(b"3 \r\n8.81032e+09  -61.3392 12.7171 -99.7497 -2.99417 -0.402539 1.70019\r\n"
                  b"1\t2\t3\t4\t\t5\t6\t7\n  1e3 0 0 0 0 0 0")
    b = lab.read_data(str(p))
    assert b.shape == (3, 7) and b[0, 0] == 8.81032e9 and b[1, 6] == 7 and b[2, 0] == 1000
    with pytest.raises(IOError):
        lab.read_data(str(tmp_path / "missing.txt"))


def test_row_format_matches_reference_bytes(golden):
    """write_bodies_coords output of the reference build, byte for byte."""
    want = bytes(golden["ref_row_bytes"])
    got = lab.format_row(0.30000000000000004, golden["body4"][:, 1:4]).encode()
    assert got == want
    assert bytes(golden["ref_header"]).decode() == lab.LAB4_HEADER
    assert lab.lab5_header(4) == ("time;r_1_1;r_1_2;r_1_3;r_2_1;r_2_2;r_2_3;r_3_1;r_3_2;r_3_3;"
                                  "r_4_1;r_4_2;r_4_3\n")


def test_step_count_matches_oracle(oracle):
    for tau, T in ((0.1, 1), (0.1, 20), (0.05, 20), (0.025, 20), (0.01, 1)):
        assert lab.step_count(0.0, T, tau) == oracle.step_count(0.0, T, tau)
        assert lab.step_count(0.0, T, tau, np.float32) == oracle.step_count_f32(0.0, T, tau)


def test_generators_are_deterministic_and_sane():
    a, b = generators.uniform_bodies(1000), generators.uniform_bodies(1000)
    assert np.array_equal(a, b)
    assert a[:, 0].min() >= 1e6 and a[:, 0].max() <= 1e8
    assert np.abs(a[:, 1:4]).max() <= 100 and np.abs(a[:, 4:7]).max() <= 10
    p = generators.plummer_bodies(20000, seed=0x5EED0003)
    assert np.array_equal(p, generators.plummer_bodies(20000, seed=0x5EED0003))
    G = 6.67e-11
    assert abs(p[:, 0].sum() * G - 1.0) < 1e-12                 # G*M = 1
    assert np.abs(p[:, 1:4].mean(axis=0)).max() < 1e-12          # centred
    r = np.linalg.norm(p[:, 1:4], axis=1)
    assert 1.2 < np.median(r) < 1.4 and r.max() < 24              # half-mass radius 1.305a
    v2 = (p[:, 4:7] ** 2).sum(axis=1).mean()
    assert 0.25 < v2 < 0.35                                      # <v^2> = 3 pi/32 (G M / a)


def test_host_driver_cli_input_errors(tmp_path):
    exe = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
    r = subprocess.run([exe, "--input", str(tmp_path / "nope.txt")], capture_output=True, text=True)
    assert r.returncode == 1 and "file is't open" in r.stdout   # the reference's throw 1


def test_cxx_generator_matches_python(tmp_path):
    exe = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
    u, p = str(tmp_path / "u.txt"), str(tmp_path / "p.txt")
    assert subprocess.run([exe, "--generate", "uniform", "--n", "777", "--generate-out", u]).returncode == 0
    assert subprocess.run([exe, "--generate", "plummer", "--n", "1500", "--seed", "0x5EED0003",
                           "--generate-out", p]).returncode == 0
    assert np.array_equal(lab.read_data(u), generators.uniform_bodies(777))
    a, b = lab.read_data(p), generators.plummer_bodies(1500, seed=0x5EED0003)
    assert np.abs(a - b).max() / np.abs(b).max() < 1e-14


def test_cxx_read_data_matches_python(tmp_path):
    """The C++ host's from_chars reader accepts what operator>> accepts (tabs, CRLF, '+', exponents,
    a short last record) and agrees with the Python mirror; --reformat needs no GPU."""
    exe = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
    src = tmp_path / "in.txt"
    src.write_bytes(b"4 \r\n8.81032e+09  -61.3392 12.7171 -99.7497 -2.99417 -0.402539 1.70019\r\n"
                    b"1\t2\t3\t4\t\t5\t6\t7\n  +1e3 0 0 0 0 0 0\n 2.5 1 1 1")
    out = tmp_path / "out.txt"
    r = subprocess.run([exe, "--reformat", str(src), "--reformat-out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    a, b = lab.read_data(str(out)), lab.read_data(str(src))
    assert a.shape == (4, 7) and np.array_equal(a[:3], b[:3])
    assert a[2, 0] == 1000.0
    # the truncated 4th record "2.5 1 1 1": operator>> assigns m and r and leaves v at the
    # previous body's values (reference :27-33)
    assert np.array_equal(a[3], [2.5, 1, 1, 1, 0, 0, 0]) and np.array_equal(a, b)
    assert subprocess.run([exe, "--reformat", str(tmp_path / "none.txt")], capture_output=True).returncode == 1


def test_short_file_semantics_match_reference(reflab4, tmp_path):
    exe = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
    src = tmp_path / "short.txt"
    src.write_text("5\n1 2 3 4 5 6 7\n8 9 10 11 12 13 14\n15 16 17\n")
    want = reflab4.read_data(str(src))
    out = tmp_path / "o.txt"
    assert subprocess.run([exe, "--reformat", str(src), "--reformat-out", str(out)]).returncode == 0
    assert np.array_equal(lab.read_data(str(out)), want)
    assert np.array_equal(lab.read_data(str(src)), want)
    assert np.array_equal(want[2], [15, 16, 17, 11, 12, 13, 14]) and np.array_equal(want[4], want[2])


def test_cxx_number_formatting_matches_printf(reflab4, tmp_path):
    """std::to_chars(scientific, 16) in the C++ host == the reference's iostream formatting
    (scientific, precision 16) == printf %.16e, including 3-digit exponents, denormals, -0."""
    exe = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
    vals = [1e300, -2.5e-300, 4.9e-324, -0.0, 123456789.12345679, 1.0 / 3.0, 8810324116.227]
    src = tmp_path / "v.txt"
    src.write_text("1\n" + " ".join(repr(x) for x in vals) + "\n")
    out = tmp_path / "o.txt"
    assert subprocess.run([exe, "--reformat", str(src), "--reformat-out", str(out)]).returncode == 0
    toks = open(out).read().split()[1:]
    assert toks == ["%.16e" % x for x in vals]
    # and the reference's own writer on the same numbers (positions only: fields 1..3)
    ref_csv = tmp_path / "r.csv"
    body = np.array([vals])
    reflab4.write_bodies_coords(str(ref_csv), vals[0], body)
    assert open(ref_csv).read() == lab.format_row(vals[0], body[:, 1:4])


def test_hot_loop_instruction_budget():
    """Regression guard on the SASS of the default FP64 force kernels (tools/sass_mix.py): the
    hot loop has 16 FP64 instructions per pair (15 with the mass factored out), no clamp
    compare/select, one MUFU per pair, and fewer than 4.2 other instructions per pair."""
    import re
    out = subprocess.run([os.sys.executable, os.path.join(ROOT, "tools", "sass_mix.py"), capi.LIB_PATH],
                         capture_output=True, text=True).stdout
    rows = {}
    for line in out.splitlines():
        m = re.match(r"double, \(int\)128, \(int\)2, \(int\)3, \(int\)8, \(int\)([01]), \(int\)8, \(int\)2 "
                     r"loop=\s*(\d+) pairs=\s*(\d+) fp64/pair=([\d.]+) other/pair=([\d.]+)", line)
        # the hot loop is the fully unrolled full-tile loop (16 pairs per trip); the ragged-tile
        # loop (4 pairs) and the exact-clamp redo loop (no VIMNMX3) are off the hot path
        if m and "VIMNMX3" in line and int(m.group(3)) == 16:
            rows[int(m.group(1))] = (float(m.group(4)), float(m.group(5)), line)
    assert set(rows) == {0, 1}, out[-2000:]
    assert rows[0][0] == 16.0 and rows[1][0] == 15.0
    for um in (0, 1):
        assert rows[um][1] < 4.2, rows[um][2]
        assert "'ISETP'" not in rows[um][2] and "'SEL'" not in rows[um][2], rows[um][2]


def _sym_emulator(tmp_path_factory=None):
    """tests/tools/sym_emulate.cpp built with g++ (walks the pair-symmetric kernel's
    decomposition through the schedule header the kernel itself includes)."""
    import ctypes as C
    import tempfile
    out = os.path.join(tempfile.mkdtemp(prefix="symemu"), "libsymemu.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", out,
                    os.path.join(ROOT, "tests", "tools", "sym_emulate.cpp")], check=True)
    L = C.CDLL(out)
    L.sym_emulate_rhs.argtypes = [C.c_uint, C.c_void_p, C.c_double, C.c_double, C.c_uint, C.c_uint,
                                  C.c_int, C.c_void_p, C.c_void_p]
    return L


def test_pair_symmetric_schedule_covers_every_pair_once(oracle):
    """The pair-symmetric kernel's walk (sym_schedule.h): every ordered pair of existing bodies
    is credited exactly once, rotating j-bodies come home, and slice rows + j-side sums add up
    to the reference operator -- odd and even block counts, ragged last block, phantom tiles,
    every i-bodies-per-thread setting the library instantiates, general and equal masses."""
    from parallel_comp_labs_b200 import generators
    L = _sym_emulator()
    for n in (1, 4, 255, 256, 257, 513, 1025, 1537, 2600, 3073):
        for ipt in (2, 4, 6, 8):
            for slices, um in ((1, 0), (3, 1), (7, 0)):
                b = generators.uniform_bodies(n, seed=n + ipt)
                if um:
                    b[:, 0] = 3.0e7
                acc = np.zeros((n, 3))
                cov = np.zeros((n, n), dtype=np.uint8)
                rc = L.sym_emulate_rhs(n, b.ctypes.data, 6.67e-11, 1e-6, ipt, slices, um,
                                       acc.ctypes.data, cov.ctypes.data)
                assert rc == 0, (n, ipt, slices, um, rc)
                assert (cov == 1).all(), (n, ipt, slices, um, int(cov.min()), int(cov.max()))
                want = oracle.rhs(b)[:, 4:7]
                assert np.abs(acc - want).max() <= 1e-12 * max(np.abs(want).max(), 1e-300)


def test_cxx_tableaux_match_python_table():
    """nbody_tableaux.hpp (C++ driver) and tableaux.py (Python mirror) hold the same schemes."""
    from parallel_comp_labs_b200 import tableaux
    exe = os.path.join(ROOT, "parallel_comp_labs_b200", "bin", "nbody_b200")
    out = subprocess.run([exe, "--list-tableaux"], capture_output=True, text=True, check=True).stdout
    seen = set()
    for line in out.splitlines():
        f = line.split()
        name, s, order, err_order = f[0], int(f[1]), int(f[2]), int(f[3])
        v = np.array([float(x) for x in f[4:]])
        assert v.size == s * s + 2 * s
        t = tableaux.TABLEAUX[name]
        assert t["order"] == order and (t["err_order"] or 0) == err_order
        assert np.array_equal(v[:s * s].reshape(s, s), t["a"])
        assert np.array_equal(v[s * s:s * s + s], t["b"])
        if t["b_err"] is not None:
            assert np.array_equal(v[s * s + s:], t["b_err"])
        assert tableaux.check_order_conditions(name) < 1e-14
        seen.add(name)
    assert seen == set(tableaux.TABLEAUX)


def test_automatic_pair_symmetric_build_divides_every_rank_block():
    """Several GPUs: the ranks own whole kernel blocks, and nbody_partition pads a rank's block to
    256 rows only -- so whatever n and world, the automatic choice must name a build whose block
    divides it (round 2 chose 512-/1024-row builds from preference lists that could all miss:
    n = 1 100 000 on 2 GPUs, n = 2 000 000 on 8, n = 1 000 960 on 2 failed with NBODY_ERR_STATE)."""
    named = [(1_100_000, 2), (2_000_000, 8), (1_000_960, 2)]
    rng = np.random.default_rng(5)
    sizes = [int(x) for x in rng.integers(2048, 3_000_000, 300)] + [2048, 20_000, 262_144, 1 << 20]
    cases = named + [(n, w) for n in sizes for w in (1, 2, 3, 4, 8)]
    fallback_seen = False
    for n, world in cases:
        for prec in (64, 32):
            for um in (True, False):
                for det in (False, True):
                    v, kb, rb = capi.sym_variant_for(n, prec, world, um, det)
                    assert rb % 256 == 0 and kb % 256 == 0
                    assert world == 1 or rb % kb == 0, (n, world, prec, um, det, v, kb, rb)
                    fallback_seen |= world > 1 and kb == 256 and (rb // 256) % 2 == 1 and n > 800_000
    assert fallback_seen            # the named cases do end on a 256-row build
    # one GPU: the measured winners are untouched (profiles/r02_sym_sweep2.md, r02_sym_crossover.md)
    assert capi.sym_variant_for(1 << 20, 64, 1, True)[0] == 8
    assert capi.sym_variant_for(1 << 20, 64, 1, False)[0] == 7
    assert capi.sym_variant_for(262_144, 64, 1, True)[0] == 9
    assert capi.sym_variant_for(20_000, 64, 1, True)[0] == 5
    assert capi.sym_variant_for(3000, 64, 1, True)[0] == 3
    assert capi.sym_variant_for(1 << 20, 32, 1, True)[0] == 1
    assert capi.sym_variant_for(1 << 20, 64, 1, True, deterministic=True)[0] == 11
    with pytest.raises(ValueError):
        capi.sym_variant_for(4, 64, 5)
