"""The EMtool_analytical-compatible launcher: argument / config / file errors must print what the reference binary
prints (captured by running bin/EMtool_analytical under oracle/shim) and exit 1 — without a GPU."""
import subprocess
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def launcher():
    from proton_b200 import build
    exe = build.build_launcher()
    assert exe.exists()
    return exe


def run(exe, *args):
    r = subprocess.run([str(exe), *args], capture_output=True, text=True, timeout=120)
    return r.returncode, r.stdout


def reference(oracle, *args):
    import os
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = str(oracle.REF_DIR)
    r = subprocess.run([str(oracle.REF_BIN), *args], capture_output=True, text=True, env=env, timeout=120)
    return r.returncode, r.stdout


CASES = {
    "no_input_files": "num_nodes 5\nA_coeff 1e-9\nsimulation_time 1e6\n",
    "no_num_nodes": "input_files {d}/\nA_coeff 1e-9\nsimulation_time 1e6\n",
    "no_a": "input_files {d}/\nnum_nodes 5\nsimulation_time 1e6\n",
    "no_time": "input_files {d}/\nnum_nodes 5\nA_coeff 1e-9\n",
}
EXPECT = {
    "no_input_files": "{cfg} The configuration file does not include the input directory.\n",
    "no_num_nodes": "The configuration file does not include the number of nodes.\n",
    "no_a": "The configuration file does not include the A coefficient.\n",
    "no_time": "The configuration file does not include the target simulation time.\n",
}


def test_wrong_argc(launcher):
    assert run(launcher) == (1, "Wrong number of arguments. Configuration file expected.\n")
    assert run(launcher, "a", "b") == (1, "Wrong number of arguments. Configuration file expected.\n")


@pytest.mark.parametrize("case", sorted(CASES))
def test_missing_keys(launcher, oracle, tmp_path, case):
    d = tmp_path / "input"
    d.mkdir()
    cfg = d / "analytical.cfg"
    cfg.write_text(CASES[case].format(d=d))
    rc, out = run(launcher, str(cfg))
    assert rc == 1 and out == EXPECT[case].format(cfg=cfg)
    if oracle.ref_binary_available():
        assert (rc, out) == reference(oracle, str(cfg))


def test_missing_and_malformed_b_csv(launcher, oracle, tmp_path):
    d = tmp_path / "input"
    d.mkdir()
    cfg = d / "analytical.cfg"
    cfg.write_text(f"input_files {d}/\nnum_nodes 5\nA_coeff 1e-9\nsimulation_time 1e6\n")
    (d / "curden.csv").write_text("1e10\n")
    want = f"Terminating\nThe file {d}/B.csv was not found.\nTerminating.\n"
    assert run(launcher, str(cfg)) == (1, want)
    if oracle.ref_binary_available():
        assert reference(oracle, str(cfg)) == (1, want)
    for body, word in (("0,0,1e-13\n4\n", "second"), ("0,0,1e-13\n4,0\n", "third")):
        (d / "B.csv").write_text(body)
        want = f"Required {word} value at line 2 of the file {d}/B.csv.\nTerminating.\n"
        assert run(launcher, str(cfg)) == (1, want)
        if oracle.ref_binary_available():
            assert reference(oracle, str(cfg)) == (1, want)


def test_nonexistent_cfg(launcher, oracle):
    rc, out = run(launcher, "/nonexistent/analytical.cfg")
    assert rc == 1 and out == "/nonexistent/analytical.cfg The configuration file does not include the input directory.\n"
    if oracle.ref_binary_available():
        assert (rc, out) == reference(oracle, "/nonexistent/analytical.cfg")
