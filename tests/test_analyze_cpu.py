"""Host-side pieces of the batched analyze driver that need no GPU."""
import numpy as np


def test_writer_follows_the_filename_collision_rule(tmp_path):
    """stress_<%f>.txt has six decimals: timestamps that print alike share a file and the last one wins, as with the
    reference's one-process-per-(line, t) loop; files hold one %.18g value per node."""
    from proton_b200.analyze import write_reference_outputs
    from proton_b200.api import EvalResult
    ts = [1.0, 2.0000001, 2.0000004, 6.38e8]                 # the middle two both print as 2.000000
    node_off = np.array([0, 3, 5])
    T = len(ts)
    stress = np.arange(5 * T, dtype=np.float64) * 1.25 + 0.1
    res = EvalResult(stress, np.zeros((2, T)), np.zeros((2, T), np.int32), np.zeros((2, T), np.uint8), node_off, T)
    n = write_reference_outputs(str(tmp_path), ["M1_a", "M2_b"], res, ts, "CuDD", 378.0, 0.1)
    assert n == 2 * 3
    d = tmp_path / "output" / "M1_a" / "CuDD_378.0_0.1"
    assert sorted(p.name for p in d.iterdir()) == ["stress_1.000000.txt", "stress_2.000000.txt", "stress_638000000.000000.txt"]
    got = [float(x) for x in (d / "stress_2.000000.txt").read_text().split()]
    assert got == list(res.line(0)[2])                       # the later of the two colliding timestamps
    assert (d / "stress_2.000000.txt").read_text().splitlines()[0] == "%.18g" % res.line(0)[2][0]
    got = [float(x) for x in (tmp_path / "output" / "M2_b" / "CuDD_378.0_0.1" / "stress_638000000.000000.txt").read_text().split()]
    assert got == list(res.line(1)[3])
