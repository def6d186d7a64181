"""The in-memory formulation must reproduce the reference's Matrix_Formulation_Class bit for bit
(goldens: tools/gen_golden.py ran the reference class on ibmpg1 lines)."""
import numpy as np
import pytest

from proton_b200 import formulation as F
from conftest import split_blob


def test_formulation_matches_reference_goldens(formulation_golden, tmp_path):
    g = formulation_golden
    csvs = split_blob(g["csv_blob"])
    assert len(csvs) == len(g["names"]) >= 150
    for i, name in enumerate(g["names"]):
        p = tmp_path / f"{name}.csv"
        p.write_text(csvs[i])
        kw = {"sp_step": float(g["param"][i])} if g["mode"][i] == "step" else {"given_disc_point": int(g["param"][i])}
        li = F.formulate_csv(str(p), technology=str(g["tech"]), temperature=float(g["temperature"]),
                             givenWidth=float(g["width"]), **kw)
        t0, t1 = g["trip_off"][i], g["trip_off"][i + 1]
        u0, u1 = g["u_off"][i], g["u_off"][i + 1]
        assert li.num_nodes == int(g["num_nodes"][i]), name
        assert li.a_coeff == float(g["a_coeff"][i]), name
        assert np.array_equal(li.rows, g["rows"][t0:t1]) and np.array_equal(li.cols, g["cols"][t0:t1]), name
        assert np.array_equal(li.vals, g["vals"][t0:t1]), name
        assert np.array_equal(li.curden, g["curden"][u0:u1]), name


def test_error_messages_match_reference():
    segs = [(" n1_0_5", " n1_10_5", 10.0, 1e-3), (" n1_10_5", " n1_12_5", 2.0, 1e-3)]
    with pytest.raises(F.FormulationError, match="not dense enough"):
        F.formulate_segments(segs, sp_step=5.0)
    with pytest.raises(F.FormulationError, match="no discretization method"):
        F.formulate_segments(segs)
    with pytest.raises(F.FormulationError, match="CSV file not found"):
        F.formulate_csv("/nonexistent/line.csv", sp_step=1.0)


def test_single_segment_policy_gives_one_node_and_negative_b(ibmpg1_lines, tmp_path):
    """under `analyze`'s step policy a one-segment line has N = 1 and B = [-bcoef]."""
    names, texts = ibmpg1_lines
    done = 0
    for name, text in zip(names, texts):
        if text.count("\nR") != 1:
            continue
        p = tmp_path / f"{name}.csv"
        p.write_text(text)
        segs, _ = F.parse_line_csv(str(p))
        li = F.formulate_segments(segs, sp_step=F.analyze_policy_step(segs))
        assert li.num_nodes == 1 and len(li.vals) == 1 and li.vals[0] == -li.bcoef
        done += 1
        if done == 5:
            break
    assert done > 0


def test_read_line_inputs_round_trip(tmp_path, oracle):
    li = F.formulate_segments([(" n1_0_5", " n1_100_5", 100.0, 2e-3), (" n1_100_5", " n1_130_5", 30.0, -1e-3)],
                              sp_step=2.0)
    d = tmp_path / "input" / "x" / "cfg"
    oracle.write_line_inputs(d, li.num_nodes, li.a_coeff, li.rows, li.cols, li.vals, li.curden, 1e6)
    (d / "analytical.txt").write_text(f"nx_total = {li.num_nodes}\nAcoef = {li.a_coeff}\n")
    back = F.read_line_inputs(str(d))
    assert back.num_nodes == li.num_nodes and back.a_coeff == li.a_coeff
    assert np.array_equal(back.rows, li.rows) and np.array_equal(back.vals, li.vals)
    assert np.array_equal(back.curden, li.curden)


FIELDS = ("num_nodes", "a_coeff", "trip_off", "trip_row", "trip_col", "trip_val", "u_off", "curden")


def _same_batch(a, b):
    for f in FIELDS:
        assert np.array_equal(getattr(a, f), getattr(b, f)), f


@pytest.mark.parametrize("engine", ["native", "numpy"])
def test_vectorised_discretiser_equals_the_per_line_twin_on_ibmpg1(engine):
    """formulate_table (all lines at once) against formulate_segments line by line: the analyze
    policy step and several fixed steps, including coarse ones where segments get one point and vias collide with the
    first / last entry of B."""
    from proton_b200 import workloads as W
    from proton_b200.api import LineBatch
    lines = W.ibmpg1_segments()
    tab = F.SegmentTable.from_lines(lines)
    assert tab.num_lines == 1717
    got = F.formulate_table(tab, engine=engine)
    want = LineBatch.from_lines([F.formulate_segments(s, sp_step=F.analyze_policy_step(s), name=n) for n, s in lines])
    _same_batch(got, want)
    assert got.names == want.names
    for step in (0.5, 2.0, 5.0):
        ok = [(n, s) for n, s in lines if min(x[2] for x in s) >= step]
        got = F.formulate_table(F.SegmentTable.from_lines(ok), sp_step=step, engine=engine)
        want = LineBatch.from_lines([F.formulate_segments(s, sp_step=step, name=n) for n, s in ok])
        _same_batch(got, want)
    with pytest.raises(F.FormulationError, match="not dense enough"):
        F.formulate_table(tab, sp_step=5.0, engine=engine)


@pytest.mark.parametrize("engine", ["native", "numpy"])
def test_vectorised_discretiser_on_the_reference_goldens(formulation_golden, tmp_path, engine):
    """The step-mode goldens produced by the reference's own Matrix_Formulation_Class, through the vectorised path."""
    g = formulation_golden
    csvs = split_blob(g["csv_blob"])
    done = 0
    for i, name in enumerate(g["names"]):
        if g["mode"][i] != "step":
            continue
        p = tmp_path / f"{name}.csv"
        p.write_text(csvs[i])
        segs, _ = F.parse_line_csv(str(p))
        b = F.formulate_table(F.SegmentTable.from_lines([(str(name), segs)]), sp_step=float(g["param"][i]),
                              technology=str(g["tech"]), temperature=float(g["temperature"]), givenWidth=float(g["width"]), engine=engine)
        t0, t1 = g["trip_off"][i], g["trip_off"][i + 1]
        u0, u1 = g["u_off"][i], g["u_off"][i + 1]
        assert int(b.num_nodes[0]) == int(g["num_nodes"][i]) and float(b.a_coeff[0]) == float(g["a_coeff"][i]), name
        assert np.array_equal(b.trip_row, g["rows"][t0:t1]) and np.array_equal(b.trip_col, g["cols"][t0:t1]), name
        assert np.array_equal(b.trip_val, g["vals"][t0:t1]) and np.array_equal(b.curden, g["curden"][u0:u1]), name
        done += 1
    assert done >= 50


@pytest.mark.parametrize("engine", ["native", "numpy"])
def test_repeated_via_keys_and_one_point_segments(engine):
    """Hand-made lines that hit the reference's dict overwrite (two joins at the same node) and the overwrite of the
    first / last entry of B by a via."""
    from proton_b200.api import LineBatch
    lines = [
        ("a", [(" n1_0_5", " n1_10_5", 10.0, 1e-3), (" n1_10_5", " n1_20_5", 10.0, 2e-3), (" n1_20_5", " n1_30_5", 10.0, -1e-3)]),
        ("b", [(" n1_0_5", " n1_1_5", 1.0, 1e-3), (" n1_1_5", " n1_2_5", 1.0, 2e-3), (" n1_2_5", " n1_40_5", 38.0, 1e-3)]),
        ("c", [(" n1_0_5", " n1_0_9", 4.0, 1e-3), (" n1_0_9", " n1_7_9", 7.0, 1e-3)]),
        ("d", [(" n1_0_5", " n1_10_5", 10.0, 1e-3)]),
    ]
    for step in (10.0, 1.0, 0.99):
        ok = [(n, s) for n, s in lines if min(x[2] for x in s) >= step]
        got = F.formulate_table(F.SegmentTable.from_lines(ok), sp_step=step, engine=engine)
        want = LineBatch.from_lines([F.formulate_segments(s, sp_step=step, name=n) for n, s in ok])
        _same_batch(got, want)


def test_workload_batches_equal_their_per_line_construction():
    from proton_b200 import workloads as W
    from proton_b200.api import LineBatch
    _same_batch(W.ibmpg1_batch(), LineBatch.from_lines(W.ibmpg1_lines()))
    a, b = W.ibmpg2_shaped_batch(replicas=3, seed=2), LineBatch.from_lines(W.ibmpg2_shaped_lines(replicas=3, seed=2))
    _same_batch(a, b)
    assert a.names == b.names


def test_native_discretiser_equals_numpy_on_random_tables():
    """proton_em_formulate_segments against the NumPy formulation on random segment tables built to hit the corners:
    lengths at exact half-multiples of the step (round-half-to-even), one-point segments (vias on the first / last
    node), runs of joins at one node (the dict overwrite), unnamed nodes, per-line steps."""
    rng = np.random.default_rng(77)
    for trial in range(40):
        L = int(rng.integers(1, 60))
        counts = rng.integers(1, 12, L)
        K = int(counts.sum())
        seg_off = np.zeros(L + 1, np.int64)
        np.cumsum(counts, out=seg_off[1:])
        step = np.repeat(rng.choice([0.5, 1.0, 2.0, 0.75], L), 1)
        mult = rng.choice([1.0, 1.0, 1.5, 2.5, 3.0, 3.5, 1.0000001, 7.25, 40.0], K)
        length = np.repeat(step, counts) * mult
        tab = F.SegmentTable(names=[f"l{i}" for i in range(L)], seg_off=seg_off, length_um=length,
                             current=rng.normal(0, 1e-3, K), lcoord=rng.integers(0, 3, (K, 2)).astype(np.int64),
                             rcoord=rng.integers(0, 3, (K, 2)).astype(np.int64), named=rng.random(K) < 0.9)
        for sp in (step, None if trial % 2 else float(step.min())):
            try:
                want = F.formulate_table(tab, sp_step=sp, engine="numpy")
            except F.FormulationError as e:
                with pytest.raises(F.FormulationError, match=str(e)[:20]):
                    F.formulate_table(tab, sp_step=sp, engine="native")
                continue
            _same_batch(F.formulate_table(tab, sp_step=sp, engine="native"), want)
    empty = F.SegmentTable(names=["x"], seg_off=np.array([0, 0]), length_um=np.zeros(0), current=np.zeros(0),
                           lcoord=np.zeros((0, 2), np.int64), rcoord=np.zeros((0, 2), np.int64), named=np.zeros(0, bool))
    for eng in ("native", "numpy"):
        with pytest.raises(F.FormulationError, match="Corrupted"):
            F.formulate_table(empty, engine=eng)
