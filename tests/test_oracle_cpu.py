"""The oracle (oracle/adk_oracle.c) pinned against the reference's golden vectors and live reference outputs.
CPU only; this is what makes the oracle trustworthy as the checker of the CUDA path."""
import os
import random

import pytest

from util import ROOT, xy, assert_same_solution, float32_equal, same_cycle


def test_oracle_reproduces_reference_known_answers_path_A(golden, oracle):
    """data/samples/tests/*/results.json, the assertions of src/test/main.cpp:93-101, all 100 instances."""
    for name, t in golden["ref_tests"].items():
        x, y = xy(golden["points"][name])
        r = oracle.solve_area(x, y, t["max_count"], t["max_area"], True)
        ca = t["convex_area"]
        assert bool(r["found"]) == (ca is not None), name
        if ca:
            assert r["count"] == ca["count"], name
            assert sorted(r["diam"]) == sorted(ca["diameter_indices"]), name
            assert float32_equal(r["area"], ca["area"]), (name, r["area"], ca["area"])
            assert same_cycle(ca["hull_indices"], r["hull"]), name


def test_oracle_matches_live_reference_path_A_with_counters(golden, oracle):
    for r in golden["runs"]["test.A"]:
        x, y = xy(golden["points"][r["name"]])
        g = oracle.solve_area(x, y, r["kmax"], r["amax"], True)
        assert_same_solution(g, r, r["name"])
        assert g["entries_alloc"] == r["entries_alloc"] and g["entries_required"] == r["entries_required"], r["name"]


@pytest.mark.parametrize("tag,nmax", [("test.ADopt", 70), ("bind.ADopt.50", 80), ("bind.ADopt.75", 70)])
def test_oracle_matches_live_reference_path_AD(golden, oracle, tag, nmax):
    """Non-binding diameter (src/test/main.cpp:113) and binding diameters (no upstream golden: live reference runs)."""
    for r in golden["runs"][tag]:
        n = int(r["name"].split("/")[1])
        if n > nmax:
            continue
        x, y = xy(golden["points"][r["name"]])
        g = oracle.solve_area_diameter(x, y, r["kmax"], r["amax"], r["dmax"], True)
        assert_same_solution(g, r, (tag, r["name"]))
        assert g["entries_required"] == r["entries_required"], (tag, r["name"])


def test_oracle_matches_reference_bench_call_shape(golden, oracle):
    """K = inf, A = 4, the diameter sweep of src/bench/main.cpp:57-59 on the shipped uniform / gaussian / real sets."""
    for r in golden["runs"]["bench.ADopt"] + golden["runs"]["real.ADopt.d1"]:
        if r["solve_ms"] > 400:
            continue
        x, y = xy(golden["points"][r["name"]])
        g = oracle.solve_area_diameter(x, y, r["kmax"], r["amax"], r["dmax"], True)
        assert_same_solution(g, r, (r["name"], r["dmax"]))
        assert g["entries_required"] == r["entries_required"]
    for r in golden["runs"]["bench.A"][:3]:
        x, y = xy(golden["points"][r["name"]])
        g = oracle.solve_area(x, y, r["kmax"], r["amax"], True)
        assert_same_solution(g, r, r["name"])
        assert g["entries_required"] == r["entries_required"]


def test_sweep_tables_equal_their_definitions(golden, oracle):
    """Utils.cpp:74-143 (sweep) against the definitions of the reference's own disabled self-check (Utils.cpp:145-223),
    including the real sets with equal-x / equal-y pairs and an eps-collinear triple."""
    for name in ["tests/10/0", "tests/40/2", "tests/100/1", "uniform/10/0", "real/9/0", "real/6/0", "real/8/0"]:
        x, y = xy(golden["points"][name])
        assert oracle.count_tables(x, y) == oracle.count_tables_bruteforce(x, y), name


def test_oracle_against_reference_library_on_random_inputs(oracle):
    """Direct differential test against oracle/_ref (only where the reference was built, i.e. the build container)."""
    from oracle.refpy import RefLib, ref_available
    if not ref_available("x87"):
        pytest.skip("oracle/_ref not built on this machine")
    ref = RefLib("x87")
    rng = random.Random(11)
    for trial in range(12):
        n = rng.choice([3, 4, 7, 15, 25, 40])
        x = [rng.uniform(0, 10) for _ in range(n)]
        y = [rng.uniform(0, 10) for _ in range(n)]
        k = rng.choice([None, 3, 5, n])
        a = rng.choice([2.0, 8.0, 30.0, float("inf")])
        d = rng.choice([1.5, 4.0, 20.0])
        assert_same_solution(oracle.solve_area(x, y, k, a, True), ref.solve(0, x, y, k, a, float("inf"), True), ("A", trial))
        assert_same_solution(oracle.solve_area_diameter(x, y, k, a, d, True), ref.solve(2, x, y, k, a, d, True), ("AD", trial))
        assert oracle.cw_order(x, y) == ref.cw_order(x, y)
        assert oracle.count_tables(x, y) == ref.count_tables(x, y)


def test_binary64_reference_variant_agrees_with_x87(golden):
    """The tie-sensitivity detector: the reference compiled with 64-bit long double gives the same answers."""
    from oracle.refpy import RefLib, ref_available
    if not (ref_available("x87") and ref_available("f64")):
        pytest.skip("oracle/_ref not built on this machine")
    f64 = RefLib("f64")
    for r in golden["runs"]["test.A"]:
        if int(r["name"].split("/")[1]) > 40:
            continue
        x, y = xy(golden["points"][r["name"]])
        assert_same_solution(f64.solve(0, x, y, r["kmax"], r["amax"], float("inf"), True), r, r["name"])


def test_edge_cases(oracle):
    assert oracle.solve_area([0, 1], [0, 1])["found"] == 0
    assert oracle.solve_area([0, 1, 0], [0, 0, 1], 2)["found"] == 0
    r = oracle.solve_area([0, 1, 0], [0, 0, 1], 3, 10.0)
    assert r["found"] and r["count"] == 3 and abs(r["area"] - 0.5) < 1e-15
    assert oracle.solve_area([0, 1, 0], [0, 0, 1], 3, 0.4)["found"] == 0
    r = oracle.solve_area_diameter([0, 1, 0], [0, 0, 1], 3, 10.0, 10.0)
    assert r["found"] and r["count"] == 3
    assert oracle.solve_area_diameter([0, 1, 0], [0, 0, 1], 3, 10.0, 1.2)["found"] == 0


def test_sharded_oracle_solves_merge_to_the_whole_solve(golden, oracle):
    """adko_solve_area_diameter_shard (every nshards-th candidate pair) + merge_shards == the whole solve: answer, diameter pair, and
    the cell / state / pair counters.  This is what lets tests/golden/make_oracle_sharded.py afford the dense bench sets."""
    from oracle.oraclepy import Oracle
    done = 0
    for r in golden["runs"]["bind.ADopt.75"][:12] + golden["runs"]["test.ADopt"][:6]:
        x, y = xy(golden["points"][r["name"]])
        whole = oracle.solve_area_diameter(x, y, r["kmax"], r["amax"], r["dmax"], False)
        for nshards in (1, 3, 7):
            parts = [oracle.solve_area_diameter_shard(x, y, r["kmax"], r["amax"], r["dmax"], s, nshards) for s in range(nshards)]
            m = Oracle.merge_shards(parts)
            assert (m["found"], m["count"], m["diam"]) == (whole["found"], whole["count"], whole["diam"]), (r["name"], nshards)
            assert not whole["found"] or m["area"] == whole["area"], (r["name"], nshards)
            assert (m["cells"], m["states"], m["pairs"]) == (whole["cells"], whole["states"], whole["pairs"]), (r["name"], nshards)
        done += 1
    assert done >= 12


def test_recorded_gpu_answers_of_the_dense_sweeps_equal_the_sharded_oracle():
    """The bench lines the B200 printed for the dense sets SURVEY 8d names (profiles/r2_bench_sigma1.0_n1.json, _sigma0.5_) against
    the binary64 oracle's answers for the same solves (tests/golden/oracle_sharded.json, computed in the build container with the
    candidate pairs sharded over processes): count, diameter pair, area to 1e-9; and, once all five diameters of a sweep are held,
    the sweep's cell and state totals.  A check of RECORDED device output — the live one is the -m gpu test of the same name."""
    import json
    with open(os.path.join(ROOT, "tests", "golden", "oracle_sharded.json")) as f:
        runs = json.load(f)["runs"]
    by_sigma = {}
    for r in runs:
        by_sigma.setdefault(float(r["sigma"]), {})[float(r["dmax"])] = r
    compared = 0
    for sigma, held in by_sigma.items():
        path = os.path.join(ROOT, "profiles", f"r2_bench_sigma{sigma}_n1.json")
        if not os.path.exists(path):
            continue
        with open(path) as f:
            line = [json.loads(l) for l in f if l.lstrip().startswith("{")][-1]
        assert line["config"]["sigma"] == sigma and line["config"]["amax"] == next(iter(held.values()))["amax"]
        for a in line["answers"]:
            o = held.get(float(a["D"]))
            if o is None:
                continue
            assert a["count"] == o["count"] and sorted(a["diam"]) == sorted(o["diam"]), (sigma, a, o)
            assert abs(a["area"] - o["area"]) <= 1e-9 * o["area"], (sigma, a["area"], o["area"])
            compared += 1
        if all(float(d) in held for d in line["config"]["diameters"]):
            assert sum(held[float(d)]["cells"] for d in line["config"]["diameters"]) == line["config"]["cells_per_step"], sigma
            assert sum(held[float(d)]["states"] for d in line["config"]["diameters"]) == line["config"]["states_per_step"], sigma
    assert compared >= 1


def test_reference_optimisation_flag_changes_counters_only(golden):
    """aShouldEnableOptimizations (Eppstein.cpp:186-189, AntipodalOptimized.cpp:505-508) prunes anchors / candidate pairs that cannot
    beat the best count so far: with the flag on the unmodified reference returns the same answer and never larger BenchmarkInfo
    counters.  This is why the GPU library may ignore the flag and why the frozen runs, generated with the
    flag off as the reference's bench does (src/CMakeLists.txt:88), pin the counters."""
    from oracle.refpy import RefLib, ref_available
    if not ref_available("x87"):
        pytest.skip("oracle/_ref is built only where /root/reference exists")
    ref = RefLib("x87")
    smaller = 0
    for algo, tag in ((0, "test.A"), (2, "test.ADopt"), (2, "bind.ADopt.75")):
        for r in golden["runs"][tag][:10]:
            x, y = xy(golden["points"][r["name"]])
            off = ref.solve(algo, x, y, r["kmax"], r["amax"], r["dmax"], True, False)
            on = ref.solve(algo, x, y, r["kmax"], r["amax"], r["dmax"], True, True)
            assert (on["found"], on["count"], on["area"], on["hull"], on["diam"]) == (off["found"], off["count"], off["area"], off["hull"], off["diam"]), r["name"]
            assert (off["entries_alloc"], off["entries_required"]) == (r["entries_alloc"], r["entries_required"]), r["name"]
            assert on["entries_required"] <= off["entries_required"] and on["entries_alloc"] <= off["entries_alloc"], r["name"]
            smaller += on["entries_required"] < off["entries_required"]
    assert smaller >= 1


def test_sharded_oracle_area_solves_merge_to_the_whole_solve(golden, oracle):
    """adko_solve_area_shard (every nshards-th anchor of the bottom-up order) + merge_area_shards == the whole solve: answer, hull
    cycle, diameter pair, and the cell / prefix-minimum counters.  This is what lets make_oracle_sharded.py --area afford config 3's
    n = 1000 set."""
    from oracle.oraclepy import Oracle
    done = 0
    for r in golden["runs"]["test.A"][:40:2] + golden["runs"]["bench.A"][:4]:
        x, y = xy(golden["points"][r["name"]])
        if len(x) > 120:
            continue
        whole = oracle.solve_area(x, y, r["kmax"], r["amax"], True)
        for nshards in (1, 3, 7):
            parts = [oracle.solve_area_shard(x, y, r["kmax"], r["amax"], True, s, nshards) for s in range(nshards)]
            m = Oracle.merge_area_shards(parts)
            assert (m["found"], m["count"], m["hull"], m["diam"]) == (whole["found"], whole["count"], whole["hull"], whole["diam"]), (r["name"], nshards)
            assert not whole["found"] or m["area"] == whole["area"], (r["name"], nshards)
            assert (m["cells"], m["entries_required"]) == (whole["cells"], whole["entries_required"]), (r["name"], nshards)
        done += 1
    assert done >= 12


def test_recorded_gpu_answer_of_config3_equals_the_sharded_oracle():
    """The answer the B200 printed for config 3 at full size (profiles/r2_bench_area_n1000_n1.json: area path, Gaussian n = 1000, K = inf)
    against the binary64 oracle's answer for the same solve with the anchors sharded over processes
    (tests/golden/oracle_sharded_area.json): count, hull cycle, area to 1e-9.  A check of RECORDED device output — the live one is the
    -m gpu test test_config3_n1000_equals_the_sharded_oracle."""
    import json
    with open(os.path.join(ROOT, "tests", "golden", "oracle_sharded_area.json")) as f:
        o = next(r for r in json.load(f)["runs"] if r["tag"].startswith("area.n1000.K"))
    with open(os.path.join(ROOT, "profiles", "r2_bench_area_n1000_n1.json")) as f:
        line = [json.loads(l) for l in f if l.lstrip().startswith("{")][-1]
    a = line["answers"][0]
    assert o["count"] < o["kmax"] - 1                      # the cap of the oracle run did not bind
    assert a["count"] == o["count"] and same_cycle(a["hull"], o["hull"]), (a, o["count"], o["hull"])
    assert abs(a["area"] - o["area"]) <= 1e-9 * o["area"], (a["area"], o["area"])


def test_oracle_reproduces_the_frozen_reference_on_the_wider_bench_sweeps(golden, oracle):
    """The wider sets of the bench fixture (Gaussian n = 200, sigma = 3.0 and 2.0): the oracle against the unmodified x87 reference on
    every frozen solve of the sigma = 3.0 diameter sweep and on D = 2, 3 of sigma = 2.0 (D = 4, 5, 6 of sigma = 2.0 take minutes of one
    core and were compared once when they were frozen: equal, entries included) — answer, hull cycle, diameter pair and hash entries."""
    fx = golden["bench_n200"]
    done = 0
    for sigma, dmaxs in (("3.0", (2, 3, 4, 5, 6)), ("2.0", (2, 3))):
        pts = fx["sets"][sigma]["points"]
        x = [float(p[0]) for p in pts]
        y = [float(p[1]) for p in pts]
        for d in dmaxs:
            smp = next((s_ for s_ in fx["sets"][sigma]["sample"] if float(s_["dmax"]) == d), None)
            if smp is not None:
                want = {"found": 1, "count": smp["count"], "area": float(smp["area"]), "hull": smp["hull"], "diam": smp["diam"],
                        "entries_required": smp["ref_entries"]}
            else:
                want = golden["heavy"][f"bench.sigma{sigma}.D{d}"]
            r = oracle.solve_area_diameter(x, y, None, fx["amax"], float(d), True)
            assert_same_solution(r, want, (sigma, d))
            assert r["entries_required"] == want["entries_required"], (sigma, d)
            done += 1
    assert done == 7
