"""tools/dc_solve.py — the DC-current provider of the fixtures: a deck small enough to solve by hand,
Kirchhoff's current law at every free node, and the file layout the reference's SPICE parser reads
(spice_function_cli.py:94-115: lines with '|Current', name = words[3] + words[4][:-1], amps = float(words[5]))."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tools"))
import dc_solve  # noqa: E402

DECK = """* a 1 V source feeds a ladder: n1 -1ohm- n2 -2ohm- n3, n3 -1ohm- ground, and a 0.1 A load at n2
V1 n1_0_0 0 1.0
R1 n1_0_0 n1_10_0 1.0
R2 n1_10_0 n1_20_0 2.0
R3 n1_20_0 0 1.0
I1 n1_10_0 0 0.1
.op
.end
"""


def test_ladder_by_hand_and_kcl(tmp_path):
    deck = tmp_path / "ladder.spice"
    deck.write_text(DECK)
    cur, resid = dc_solve.solve_deck(str(deck))
    # unknowns v2, v3:  (1 - v2)/1 = (v2 - v3)/2 + 0.1,  (v2 - v3)/2 = v3/1  ->  v3 = v2/3, v2 = 0.9/(4/3) = 0.675
    v2 = 0.675
    v3 = v2 / 3
    assert np.isclose(cur["R1"], 1.0 - v2, rtol=1e-12)
    assert np.isclose(cur["R2"], (v2 - v3) / 2, rtol=1e-12)
    assert np.isclose(cur["R3"], v3, rtol=1e-12)
    assert np.isclose(cur["R1"], cur["R2"] + 0.1, rtol=1e-12)          # KCL at n2
    assert resid <= 1e-12
    out = tmp_path / "DC_analysis_ladder.txt"
    dc_solve.write_dc_file(cur, str(out))
    parsed = {}
    for line in out.read_text().splitlines():
        w = line.split()
        if "|Current" in w:
            parsed[w[3] + w[4][:-1]] = float(w[5])
    assert parsed == cur


def test_zero_volt_sources_merge_nodes(tmp_path):
    """IBM decks stitch layers with 0 V sources (vias): the two nodes are one node for the solve."""
    deck = tmp_path / "via.spice"
    deck.write_text("V1 a_0_0 0 2.0\nR1 a_0_0 b_0_0 1.0\nV2 b_0_0 c_0_0 0.0\nR2 c_0_0 0 3.0\n")
    cur, _ = dc_solve.solve_deck(str(deck))
    assert np.isclose(cur["R1"], 0.5) and np.isclose(cur["R2"], 0.5)
