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
