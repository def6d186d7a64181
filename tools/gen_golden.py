"""tools/gen_golden.py — regenerates tests/golden/*.npz.  Runs ONLY where /root/reference exists.

The reference ships no golden vectors, so the fixtures are manufactured from the
reference itself:
  * per-line CSVs: the reference's own SPICE parser (spice_function_cli.Spice_Parser_Class) on
    benchmarks/ibmpg1.spice, fed resistor currents from tools/dc_solve.py (the DC file is a missing blob);
  * kernel inputs: the reference's own Matrix_Formulation_Class (matrix_formulation_func_cli.py);
  * kernel outputs: the UNMODIFIED reference binary bin/EMtool_analytical under oracle/shim.

Outputs
  tests/golden/ibmpg1_lines.npz        every ibmpg1 line CSV (text), for the full-grid workloads
  tests/golden/formulation_golden.npz  reference formulation outputs for a subset of lines
  tests/golden/binary_golden.npz       reference binary stress vectors (ibmpg1 subset, config-1 line,
                                       synthetic edge cases)
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import tempfile
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[1]
REF = Path("/root/reference")
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tools"))

from oracle import restatement as R  # noqa: E402
import dc_solve  # noqa: E402

GOLD = REPO / "tests" / "golden"
TIMES = [1.0e6, 1.0e7, 6.38e8]
TECH, TEMP, WIDTH = "CuDD", 378.0, 0.1


def ragged(list_of_arrays, dtype):
    off = np.zeros(len(list_of_arrays) + 1, dtype=np.int64)
    for i, a in enumerate(list_of_arrays):
        off[i + 1] = off[i] + len(a)
    flat = np.concatenate([np.asarray(a, dtype=dtype) for a in list_of_arrays]) if list_of_arrays else np.zeros(0, dtype)
    return flat, off


def reference_formulation(csv_path, proj, sp_step=None, points=None):
    """Run the reference class, then read back the files it wrote."""
    sys.path.insert(0, str(REF))
    import matrix_formulation_func_cli as mf
    from proton_b200.formulation import read_line_inputs
    with contextlib.redirect_stdout(io.StringIO()):
        msg = mf.Matrix_Formulation_Class(csv_file=str(csv_path), project_location=str(proj), sp_step=sp_step,
                                          given_disc_point=points, technology=TECH, temperature=float(TEMP),
                                          givenWidth=float(WIDTH)).matrix_formulation()
    if "seconds" not in msg:
        return msg, None
    name = Path(csv_path).stem
    d = Path(proj) / "input" / name / f"{TECH}_{float(TEMP)}_{float(WIDTH)}"
    return msg, read_line_inputs(str(d), name)


def main():
    if not REF.exists():
        raise SystemExit("needs /root/reference")
    R.build()
    GOLD.mkdir(parents=True, exist_ok=True)
    rng = np.random.default_rng(20260101)
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        (tmp / "benchmarks" / "DC_analyses").mkdir(parents=True)
        cur, resid = dc_solve.solve_deck(str(REF / "benchmarks" / "ibmpg1.spice"))
        dc_solve.write_dc_file(cur, str(tmp / "benchmarks" / "DC_analyses" / "DC_analysis_ibmpg1.txt"))
        proj = tmp / "proj"
        proj.mkdir()
        sys.path.insert(0, str(REF))
        import spice_function_cli as sfc
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                msg = sfc.Spice_Parser_Class(str(REF / "benchmarks" / "ibmpg1.spice"), str(proj), False,
                                             str(REF)).spice_parser()
        finally:
            os.chdir(cwd)
        assert "seconds" in msg, msg
        csvs = sorted((proj / "ibmpg1").glob("*/*.csv"), key=lambda p: (p.parent.name, p.stem.split("_")[1],
                                                                       int(p.stem.split("_")[2])))
        print(f"{len(csvs)} lines parsed by the reference parser (DC residual {resid:.2e})")

        # ---- 1. every line CSV as text ----
        names = [p.stem for p in csvs]
        texts = [p.read_text() for p in csvs]
        blob = "\x00".join(texts).encode()
        np.savez_compressed(GOLD / "ibmpg1_lines.npz", names=np.array(names), csv_blob=np.frombuffer(blob, dtype=np.uint8),
                            dc_residual=np.float64(resid))

        # ---- 2. formulation goldens (reference class) + 3. binary outputs ----
        from proton_b200.formulation import parse_line_csv, analyze_policy_step
        pick = []
        by_S = {}
        for p in csvs:
            segs, _ = parse_line_csv(str(p))
            by_S.setdefault(len(segs), []).append(p)
        for S in sorted(by_S):
            lst = by_S[S]
            take = lst[:2] if S > 3 else lst[:4]
            pick.extend(take)
        extra = [p for p in csvs if p not in pick]
        pick.extend(list(rng.choice(np.array(extra, dtype=object), size=40, replace=False)))
        f_names, f_mode, f_param, f_N, f_a, f_rows, f_cols, f_vals, f_u, f_csv = [], [], [], [], [], [], [], [], [], []
        b_tag, b_N, b_a, b_rows, b_cols, b_vals, b_u, b_t, b_stress = [], [], [], [], [], [], [], [], []

        def add_binary(tag, li, times):
            for t in times:
                st = R.reference_binary_line(li.num_nodes, li.a_coeff, li.rows, li.cols, li.vals, li.curden, t)
                assert len(st) == li.num_nodes
                b_tag.append(tag), b_N.append(li.num_nodes), b_a.append(li.a_coeff)
                b_rows.append(li.rows), b_cols.append(li.cols), b_vals.append(li.vals), b_u.append(li.curden)
                b_t.append(t), b_stress.append(st)

        for p in pick:
            segs, _ = parse_line_csv(str(p))
            step = analyze_policy_step(segs)
            total = sum(s[2] for s in segs)
            for mode, param in (("step", step), ("points", 60)):
                if mode == "step" and total / step > 2500:
                    continue
                msg, li = reference_formulation(p, proj, sp_step=param if mode == "step" else None,
                                                points=param if mode == "points" else None)
                if li is None:
                    continue
                f_names.append(p.stem), f_mode.append(mode), f_param.append(float(param))
                f_N.append(li.num_nodes), f_a.append(li.a_coeff)
                f_rows.append(li.rows), f_cols.append(li.cols), f_vals.append(li.vals), f_u.append(li.curden)
                f_csv.append(p.read_text())
                if li.num_nodes <= 1500 and len(b_tag) < 260:
                    add_binary(f"ibmpg1:{p.stem}:{mode}", li, TIMES if li.num_nodes <= 400 else TIMES[-1:])
        print(f"formulation goldens: {len(f_names)}; binary goldens so far: {len(b_tag)}")

        # config 1: M5_n1_1, --points 1000 (examples/line_EM_analysis.tcl:20-28), analyze_line 6.38e8
        cfg1 = proj / "ibmpg1" / "M5" / "M5_n1_1.csv"
        msg, li = reference_formulation(cfg1, proj, points=1000)
        assert li is not None, msg
        f_names.append("M5_n1_1"), f_mode.append("points"), f_param.append(1000.0)
        f_N.append(li.num_nodes), f_a.append(li.a_coeff)
        f_rows.append(li.rows), f_cols.append(li.cols), f_vals.append(li.vals), f_u.append(li.curden)
        f_csv.append(cfg1.read_text())
        add_binary("config1:M5_n1_1:points1000", li, TIMES)
        print(f"config 1: N={li.num_nodes} a={li.a_coeff!r}")

        fr, fro = ragged(f_rows, np.int32)
        fc, _ = ragged(f_cols, np.int32)
        fv, _ = ragged(f_vals, np.float64)
        fu, fuo = ragged(f_u, np.float64)
        np.savez_compressed(GOLD / "formulation_golden.npz", names=np.array(f_names), mode=np.array(f_mode),
                            param=np.array(f_param), num_nodes=np.array(f_N, np.int32), a_coeff=np.array(f_a),
                            rows=fr, cols=fc, vals=fv, trip_off=fro, curden=fu, u_off=fuo,
                            csv_blob=np.frombuffer("\x00".join(f_csv).encode(), dtype=np.uint8),
                            tech=np.array(TECH), temperature=np.float64(TEMP), width=np.float64(WIDTH))

        # ---- synthetic edge cases through the binary ----
        from types import SimpleNamespace

        def synth(N, S, a=None):
            bcoef = 8.29e-14 * rng.uniform(0.5, 2.0)
            a = 3.3586e-21 / (rng.uniform(1, 12) * 1e-6) ** 2 if a is None else a
            if N == 1:
                rows, cols, vals = [0], [S - 1], [-bcoef]
            else:
                S = min(S, N - 1) if N > 2 else 1
                vias = np.sort(rng.choice(np.arange(1, N - 1), size=S - 1, replace=False)) if S > 1 else []
                rows, cols, vals = [0], [0], [bcoef]
                for s, v in enumerate(vias, start=1):
                    rows += [int(v), int(v)]
                    cols += [s - 1, s]
                    vals += [-bcoef, bcoef]
                rows.append(N - 1), cols.append(S - 1), vals.append(-bcoef)
            u = np.float32(rng.normal(0, 1e11, S)).astype(np.float64)
            return SimpleNamespace(num_nodes=N, a_coeff=float(a), rows=np.array(rows, np.int32),
                                   cols=np.array(cols, np.int32), vals=np.array(vals), curden=u)

        sizes = [1, 2, 3, 4, 5, 7, 8, 15, 16, 17, 31, 32, 33, 61, 64, 97, 100, 127, 128, 129, 255, 256, 257, 499, 509,
                 512, 513, 600, 1000, 1021, 1024, 1025, 2047, 2048, 2049]
        for N in sizes:
            for S in (1, 5):
                add_binary(f"synth:N{N}:S{S}", synth(N, S), [1.0e6, 6.38e8])
        # large a*t (saturating high modes, exp underflow) and tiny a*t (cancellation regime, F10)
        add_binary("synth:saturate", synth(300, 4, a=3.3586e-21 / (0.05e-6) ** 2), [1.0e6, 6.38e8])
        add_binary("synth:tiny_at", synth(300, 4, a=3.3586e-21 / (40e-6) ** 2), [1.0e3, 1.0e6])
        # duplicate triplets and an unsorted B.csv (Eigen setFromTriplets sums duplicates)
        d = synth(50, 3)
        d.rows = np.concatenate([d.rows[::-1], d.rows[:2]]).astype(np.int32)
        d.cols = np.concatenate([d.cols[::-1], d.cols[:2]]).astype(np.int32)
        d.vals = np.concatenate([d.vals[::-1], d.vals[:2] * 0.25])
        add_binary("synth:duplicates", d, [6.38e8])

        br, bro = ragged(b_rows, np.int32)
        bc, _ = ragged(b_cols, np.int32)
        bv, _ = ragged(b_vals, np.float64)
        bu, buo = ragged(b_u, np.float64)
        bs, bso = ragged(b_stress, np.float64)
        np.savez_compressed(GOLD / "binary_golden.npz", tag=np.array(b_tag), num_nodes=np.array(b_N, np.int32),
                            a_coeff=np.array(b_a), rows=br, cols=bc, vals=bv, trip_off=bro, curden=bu, u_off=buo,
                            t=np.array(b_t), stress=bs, stress_off=bso)
        print(f"binary goldens: {len(b_tag)} (line, t) pairs, {bs.size} stress values")


if __name__ == "__main__":
    main()
