"""Host side of `analyze krd` (jplace ingest, matrix writer, CLI) against fixtures generated from the
compiled reference (tests/golden/make_golden_host.py) and — where /root/reference is present —
against the reference run live.  Bar: BIT-EXACT mass arrays and BYTE-IDENTICAL CSV output."""
import ctypes
import glob
import gzip
import os
import subprocess

import numpy as np
import pytest

from gappa_b200 import host

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
JP = os.path.join(GOLD, "jplace")
REF = "/root/reference"
OPTS = {"default": {}, "pm": dict(point_mass=True), "im": dict(ignore_multiplicities=True)}


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "host_ingest.npz"))


def test_host_library_exports():
    L = ctypes.CDLL(host.LIB_PATH)
    hdr = open(os.path.join(os.path.dirname(HERE), "include", "krd_host.h")).read()
    for name in host.EXPORTS:
        assert name + "(" in hdr, name
        assert hasattr(L, name), name


@pytest.mark.parametrize("name", list(OPTS))
def test_ingest_matches_reference_fixture(gold, name):
    tree, masses, ex = host.load_jplace([JP], **OPTS[name])
    for k in ("parent", "child", "bl"):
        assert np.array_equal(gold[f"{name}_{k}"], tree[k]), k
    assert tree["root"] == int(gold[f"{name}_root"])
    assert np.array_equal(gold[f"{name}_edge_num"], ex["edge_num"])
    for k in ("off", "edge", "pos", "mass"):
        assert np.array_equal(gold[f"{name}_{k}"], masses[k]), k      # bit-exact doubles
    assert np.array_equal(gold[f"{name}_pendant"], ex["pendant"])
    assert float(gold[f"{name}_diameter"]) == ex["diameter"]
    assert ex["labels"] == [str(x) for x in gold["labels"]]


def test_paths_files_and_dirs_are_pooled_and_sorted(gold):
    files = sorted(glob.glob(JP + "/*"))
    t1, m1, e1 = host.load_jplace(list(reversed(files)))
    t2, m2, e2 = host.load_jplace([JP + "/"])
    assert e1["labels"] == e2["labels"] == [str(x) for x in gold["labels"]]
    assert np.array_equal(m1["mass"], m2["mass"])


def test_threads_do_not_change_result():
    a = host.load_jplace([JP], threads=1)
    b = host.load_jplace([JP], threads=5)
    for k in a[1]:
        assert np.array_equal(a[1][k], b[1][k])


def test_number_accumulation_matches_reference_quirk():
    # genesis utils/formats/json/reader.cpp:266-296: dx += digit * pos with pos /= 10 — not correctly rounded.
    def model(txt):
        ip, fp = txt.split(".")
        dx, pos = float(int(ip)), 1.0
        for ch in fp:
            pos /= 10.0
            dx += int(ch) * pos
        return dx
    n_diff = 0
    rng = np.random.default_rng(5)
    for _ in range(2000):
        txt = "%d.%s" % (rng.integers(0, 3), "".join(str(d) for d in rng.integers(0, 10, rng.integers(1, 18))))
        got = host.parse_number(txt)
        assert got == model(txt), txt
        n_diff += got != float(txt)
    assert n_diff > 0          # the quirk is real: some values differ from strtod by an ulp
    assert host.parse_number("17") == 17.0
    assert host.parse_number("-3") == -3.0
    assert host.parse_number("1.5e3") == 1500.0
    assert host.parse_number("25E-2") == 25 * (0.1 * 0.1)
    assert host.parse_number("-0.125") == -0.125


def _bad(tmp_path, mutate):
    doc = __import__("json").load(open(os.path.join(JP, "s_a.jplace")))
    mutate(doc)
    p = tmp_path / "bad.jplace"
    p.write_text(__import__("json").dumps(doc))
    return [str(p)]


@pytest.mark.parametrize("mutate,msg", [
    (lambda d: d.pop("tree"), "does not contain a valid Newick tree at key 'tree'"),
    (lambda d: d.pop("fields"), "does not contain field names at key 'fields'"),
    (lambda d: d.pop("placements"), "does not contain pqueries at key 'placements'"),
    (lambda d: d["fields"].remove("pendant_length") or [r.pop() for q in d["placements"] for r in q["p"]],
     "does not contain necessary field 'pendant_length'"),
    (lambda d: d["placements"][0]["p"][0].append(1), "different size than the fields name array"),
    (lambda d: d["placements"][0]["p"][0].__setitem__(0, 9999), "does not correspond to any edge_num"),
    (lambda d: d["placements"][0].update(n=["a"], nm=[["a", 1]]), "both an 'n' and an 'nm' key"),
    (lambda d: [d["placements"][0].pop(k, None) for k in ("n", "nm", "m")], "neither an 'n' nor an 'nm' key"),
    (lambda d: d["placements"][0].__setitem__("p", []), "does not contain any placements at key 'p'"),
])
def test_malformed_documents_raise_reference_messages(tmp_path, mutate, msg):
    with pytest.raises(RuntimeError, match=msg.replace("'", ".")):
        host.load_jplace(_bad(tmp_path, mutate))


def test_different_topology_is_rejected(tmp_path):
    doc = __import__("json").load(open(os.path.join(JP, "s_a.jplace")))
    doc["tree"] = "((A:1{0},B:1{1}):1{2},C:1{3});"
    for q in doc["placements"]:
        for r in q["p"]:
            r[0] = 0
    p = tmp_path / "other.jplace"
    p.write_text(__import__("json").dumps(doc))
    with pytest.raises(RuntimeError, match="Sample reference trees do not have identical topology."):
        host.load_jplace([os.path.join(JP, "s_a.jplace"), str(p)])


@pytest.mark.parametrize("fmt", ["matrix", "list", "triangular"])
@pytest.mark.parametrize("labels", [True, False])
def test_matrix_writer_bytes(gold, tmp_path, fmt, labels):
    want = open(os.path.join(GOLD, "csv", fmt + ("" if labels else "_nolabels") + ".csv"), "rb").read()
    out = tmp_path / "m.csv"
    n = host.write_matrix(out, gold["default_m_p1.0"], fmt.upper() if fmt == "list" else fmt,
                          [str(x) for x in gold["labels"]] if labels else None, threads=3)
    assert out.read_bytes() == want
    assert n == len(want)
    gz = tmp_path / "m.csv.gz"
    host.write_matrix(gz, gold["default_m_p1.0"], fmt, [str(x) for x in gold["labels"]] if labels else None, gzip=True)
    assert gzip.open(gz, "rb").read() == want


def test_matrix_writer_value_text_equals_printf_g(tmp_path):
    """The writer formats values without printf where it can decide the six digits with certainty and falls back to printf
    otherwise: every cell must read exactly like "%g" (= the reference's default ostream formatting, writer.hpp:72,237) —
    checked on random magnitudes, values on and next to six-digit rounding ties, powers of ten, zeros, infinities, NaN."""
    rng = np.random.default_rng(11)
    vals = [0.0, -0.0, 1.0, 10.0, 100000.0, 999999.0, 999999.5, 1e6, 1e-4, 1e-5, 9.9999949999e-5, 1e-17, 1e22, 5e-324, 1e308,
            123456.5, 0.1, 0.5, 2.5, 1234565.0, float("inf"), float("-inf"), float("nan"), 0.30000000000000004, 4.35, 0.000123456789]
    vals += list(rng.random(20000) * 5.0) + list(np.exp(rng.uniform(-60, 60, 20000)) * rng.choice([-1.0, 1.0], 20000))
    for X in range(-19, 24):
        for n in rng.integers(100000, 1000000, 40):
            base = (int(n) + 0.5) * 10.0 ** (X - 5)
            vals += [base, np.nextafter(base, 0.0), np.nextafter(base, np.inf), int(n) * 10.0 ** (X - 5)]
    k = int(np.ceil(np.sqrt(len(vals))))
    m = np.zeros(k * k)
    m[: len(vals)] = vals
    out = tmp_path / "v.csv"
    host.write_matrix(str(out), m.reshape(k, k), "matrix", None, threads=3)
    got = out.read_text().replace("\n", "\t").split("\t")
    want = ["%g" % v for v in m]
    assert got[: len(want)] == want


def test_cfg1_csv_is_7847_bytes(tmp_path):
    # gappa test/scripts/analyze-krd-1.sh:7 expects krd_matrix.csv of exactly 7847 bytes
    g = np.load(os.path.join(GOLD, "cfg1.npz"))
    out = tmp_path / "krd_matrix.csv"
    host.write_matrix(out, g["m_p1.0"], "matrix", [str(x) for x in g["names"]])
    want = open(os.path.join(GOLD, "csv", "cfg1_krd_matrix.csv"), "rb").read()
    assert len(want) == 7847 and out.read_bytes() == want


def test_cli_errors_before_any_gpu_work(tmp_path):
    cli = host.CLI_PATH
    r = subprocess.run([cli, "analyze", "krd", "--jplace-path", os.path.join(JP, "s_a.jplace"), "--out-dir", str(tmp_path)],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "Cannot run krd with fewer than 2 samples." in r.stderr     # krd.cpp:131-133
    (tmp_path / "krd_matrix.csv").write_text("x")
    r = subprocess.run([cli, "analyze", "krd", "--jplace-path", JP, "--out-dir", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode != 0 and "Output file already exists" in r.stderr and "--allow-file-overwriting" in r.stderr
    r = subprocess.run([cli, "analyze", "krd", "--jplace-path", JP, "--matrix-format", "nope"], capture_output=True, text=True)
    assert r.returncode != 0
    r = subprocess.run([cli, "analyze", "krd", "--jplace-path", str(tmp_path / "missing")], capture_output=True, text=True)
    assert r.returncode != 0 and "does not exist" in r.stderr
    r = subprocess.run([cli, "analyze", "krd", "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--exponent" in r.stdout and "--point-mass" in r.stdout


# ---- live against the compiled reference (build container only) ----------------------------------

needs_ref = pytest.mark.skipif(not os.path.isdir(REF + "/test/data/jplace"), reason="/root/reference not present")


@needs_ref
@pytest.mark.parametrize("name", list(OPTS))
def test_cfg1_ingest_equals_reference_live(name):
    import oracle
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built")
    paths = sorted(glob.glob(REF + "/test/data/jplace/*.jplace.gz"))
    t1, m1, x1 = oracle.ref_jplace(paths, **OPTS[name])
    t2, m2, x2 = host.load_jplace([REF + "/test/data/jplace"], **OPTS[name])
    for k in ("parent", "child", "bl"):
        assert np.array_equal(t1[k], t2[k])
    for k in m1:
        assert np.array_equal(m1[k], m2[k]), k
    assert x1["diameter"] == x2["diameter"] and np.array_equal(x1["pendant"], x2["pendant"])


@needs_ref
def test_genesis_unit_test_documents_live():
    import oracle
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built")
    d = REF + "/libs/genesis/test/data/placement/"
    for f in ("test_a", "test_b", "test_c", "version_1", "version_2", "version_3", "rooted", "unrooted", "duplicates_a", "duplicates_b"):
        t1, m1, x1 = oracle.ref_jplace([d + f + ".jplace"])
        t2, m2, x2 = host.load_jplace([d + f + ".jplace"])
        for k in m1:
            assert np.array_equal(m1[k], m2[k]), (f, k)
        assert np.array_equal(t1["parent"], t2["parent"]) and np.array_equal(t1["bl"], t2["bl"]), f


def test_pendant_work_secondary_surface():
    # placement::earth_movers_distance(SampleSet, p, with_pendant_length) — genesis' own unit-test values
    # (libs/genesis/test/src/placement/measures.cpp:84-85, 92-97, 104-109): 3.4916666 off the diagonal,
    # 0.52499998 / 0.76666665 on it
    z = np.load(os.path.join(GOLD, "test_ab.npz"))
    m = host.add_pendant_work(z["m_p1.0"], z["pendant"])
    assert np.float32(m[0, 1]) == np.float32(3.4916666) and m[0, 1] == m[1, 0]
    assert np.float32(m[0, 0]) == np.float32(0.52499998) and np.float32(m[1, 1]) == np.float32(0.76666665)
    assert np.array_equal(z["m_p1.0"], z["m_p1.0"].T) and z["m_p1.0"][0, 0] == 0.0   # input untouched


@pytest.mark.gpu
def test_pendant_work_on_gpu_matrix():
    z = np.load(os.path.join(GOLD, "test_ab.npz"))
    import gappa_b200
    tree = dict(parent=z["parent"], child=z["child"], bl=z["bl"], root=int(z["root"]))
    mass = dict(off=z["off"], edge=z["edge"], pos=z["pos"], mass=z["mass"])
    m = host.add_pendant_work(gappa_b200.earth_movers_distance(tree, mass, p=1.0), z["pendant"])
    assert np.float32(m[0, 1]) == np.float32(3.4916666)
    assert np.float32(m[0, 0]) == np.float32(0.52499998) and np.float32(m[1, 1]) == np.float32(0.76666665)


# ---- the CLI end to end on a B200 ------------------------------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("args,csv,key", [
    ([], "matrix.csv", "default_m_p1.0"),
    (["--matrix-format", "triangular"], "triangular.csv", "default_m_p1.0"),
    (["--matrix-format", "list", "--omit-matrix-labels"], "list_nolabels.csv", "default_m_p1.0"),
    (["--normalize"], "matrix_normalized.csv", None),
    (["--point-mass", "--krd-exponent", "2"], "matrix_pm_p2.csv", "pm_m_p2.0"),
])
def test_cli_end_to_end(gold, tmp_path, args, csv, key):
    r = subprocess.run([host.CLI_PATH, "analyze", "krd", "--jplace-path", JP, "--out-dir", str(tmp_path / "o"),
                        "--write-binary", str(tmp_path / "m.bin"), "--verbose"] + args, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert (tmp_path / "o" / "krd_matrix.csv").read_bytes() == open(os.path.join(GOLD, "csv", csv), "rb").read()
    got = np.fromfile(tmp_path / "m.bin").reshape(5, 5)
    want = gold[key] if key else gold["default_m_p1.0"] / float(gold["default_diameter"])
    assert np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)) <= 1e-10   # north_star tolerance, FP64
    assert "Finished" in r.stdout


@pytest.mark.gpu
def test_cli_multi_gpu_matches_single(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    # --exchange nccl (default): NCCL all-gather of the result tiles inside the one-process CLI; host: per-GPU D2H of own tiles
    for g, d, ex in ((1, "o1", "nccl"), (2, "o2", "nccl"), (2, "o3", "host")):
        r = subprocess.run([host.CLI_PATH, "analyze", "krd", "--jplace-path", JP, "--out-dir", str(tmp_path / d), "--gpus", str(g),
                            "--exchange", ex, "--write-binary", str(tmp_path / f"{d}.bin")], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    assert (tmp_path / "o1.bin").read_bytes() == (tmp_path / "o2.bin").read_bytes() == (tmp_path / "o3.bin").read_bytes()
    assert (tmp_path / "o2" / "krd_matrix.csv").read_bytes() == open(os.path.join(GOLD, "csv", "matrix.csv"), "rb").read()


@pytest.mark.gpu
def test_cli_gpus_beyond_device_count_fails_loudly(tmp_path):
    r = subprocess.run([host.CLI_PATH, "analyze", "krd", "--jplace-path", JP, "--out-dir", str(tmp_path / "o"), "--gpus", "64"],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "Error:" in r.stderr


@pytest.mark.gpu
def test_cli_squash_end_to_end(tmp_path):
    r = subprocess.run([host.CLI_PATH, "analyze", "squash", "--jplace-path", JP, "--out-dir", str(tmp_path / "o")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert (tmp_path / "o" / "cluster.newick").read_text() == open(os.path.join(GOLD, "csv", "cluster.newick")).read()
    r = subprocess.run([host.CLI_PATH, "analyze", "squash", "--jplace-path", JP, "--out-dir", str(tmp_path / "o")],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "Output file already exists" in r.stderr


@pytest.mark.gpu
def test_cli_phylogenetic_kmeans_end_to_end(tmp_path):
    base = [host.CLI_PATH, "analyze", "phylogenetic-kmeans", "--jplace-path", JP, "--out-dir", str(tmp_path / "o"), "--bins", "5"]
    r = subprocess.run(base + ["--k", "2", "--init-centroids", "0,3", "--write-overview-file"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for name in ("k_2_assignments.csv", "overview.csv"):
        assert (tmp_path / "o" / name).read_text() == open(os.path.join(GOLD, "csv", name)).read()
    # k-means++ start (seeded): every k of the range list gets its table; clusters are non-empty
    r = subprocess.run(base + ["--k", "1-3", "--seed", "7", "--file-prefix", "pp_"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for k in (1, 2, 3):
        rows = (tmp_path / "o" / f"pp_k_{k}_assignments.csv").read_text().splitlines()
        assert rows[0] == "Sample\tCluster\tDistance" and len(rows) == 6
        assert sorted({int(x.split("\t")[1]) for x in rows[1:]}) == list(range(k))
    r = subprocess.run(base + ["--k", "2,x"], capture_output=True, text=True)
    assert r.returncode != 0 and "Invalid list of values for k" in r.stderr
    # gappa's default --bins 0 (branch-centred centroids): runs, every sample gets a cluster
    r = subprocess.run(base[:-2] + ["--k", "2", "--seed", "3", "--file-prefix", "b0_"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert len((tmp_path / "o" / "b0_k_2_assignments.csv").read_text().splitlines()) == 6


@pytest.mark.gpu
@pytest.mark.parametrize("bins", [1, 4, 37])
def test_cli_bins_ships_compact_form_and_matches_full_form(tmp_path, bins):
    """`analyze krd --bins B`: the command bins on the host and uploads (cell, mass) — krd_set_masses_binned16 here; the
    matrix must equal, bit for bit, what the library computes from the full (edge, pos, mass) form with bins = B."""
    import gappa_b200
    r = subprocess.run([host.CLI_PATH, "analyze", "krd", "--jplace-path", JP, "--out-dir", str(tmp_path / "o"), "--bins", str(bins),
                        "--write-binary", str(tmp_path / "m.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    tree, masses, _ = host.load_jplace([JP])
    want = gappa_b200.earth_movers_distance(tree, masses, p=1.0, bins=bins)
    got = np.fromfile(tmp_path / "m.bin").reshape(want.shape)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_cli_exponent_error_message(tmp_path):
    r = subprocess.run([host.CLI_PATH, "analyze", "krd", "--jplace-path", JP, "--out-dir", str(tmp_path), "--exponent", "-1"],
                       capture_output=True, text=True)
    assert r.returncode != 0
    assert "Invalid exponent value p for earth mover's distance calculation. Has to be > 0.0." in r.stderr   # emd.cpp:186-190
