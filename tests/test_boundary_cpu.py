"""The drop-in boundary, checked without a GPU: the C ABI library loads and exports what include/adk.h declares,
fails loudly without a device, and the C++ mirror defines the reference's public symbols."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    with open(os.path.join(ROOT, "include", "adk.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(adk_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.load_library()
    names = _declared_symbols()
    assert len(names) >= 18
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert sorted(pkg.EXPORTS) == names


def test_no_cpu_fallback_without_a_device(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the loud-failure path is covered on CPU-only machines")
    with pytest.raises(pkg.AdkError) as e:
        pkg.Solver(0)
    assert "adk_create failed" in str(e.value)


def test_merge_is_lexicographic(pkg):
    lib = pkg.load_library()
    from importlib import import_module
    capi = import_module("area_diameter_algorithm_b200.capi")
    parts = [
        dict(found=0, count=99, area=0.1, tie_key=0, w0=0, w1=0),
        dict(found=1, count=7, area=3.0, tie_key=5, w0=1, w1=2),
        dict(found=1, count=8, area=3.5, tie_key=9, w0=3, w1=4),
        dict(found=1, count=8, area=3.25, tie_key=7, w0=5, w1=6),
        dict(found=1, count=8, area=3.25, tie_key=6, w0=7, w1=8),
    ]
    arr = (capi.AdkPartial * len(parts))(*[capi.AdkPartial(p["found"], p["count"], p["area"], p["tie_key"], p["w0"], p["w1"]) for p in parts])
    out = capi.AdkPartial()
    assert lib.adk_merge(arr, ctypes.c_uint32(len(parts)), ctypes.byref(out)) == 0
    assert (out.found, out.count, out.area, out.tie_key, out.w0, out.w1) == (1, 8, 3.25, 6, 7, 8)
    dist = import_module("area_diameter_algorithm_b200.dist")
    assert dist.merge_partials(parts) == parts[4]
    assert dist.unpack_partial(dist.pack_partial(parts[3])) == parts[3]
    assert dist.RECORD_BYTES == ctypes.sizeof(capi.AdkPartial) == 32


def test_host_mirror_defines_the_reference_entry_points(built):
    so = os.path.join(ROOT, "area-diameter-algorithm_b200", "libadk_host.so")
    out = subprocess.run("nm -D --defined-only '%s' | c++filt" % so, shell=True, capture_output=True, text=True, check=True).stdout
    for sym in ["MT::EppsteinAlgorithm(", "MT::EppsteinAlgorithmWithBenchmarkInfo(", "MT::AntipodalAlgorithm(",
                "MT::AntipodalAlgorithmWithBenchmarkInfo(", "MT::AntipodalOptimizedAlgorithm(",
                "MT::AntipodalOptimizedAlgorithmWithBenchmarkInfo(", "MT::CM::Distance2(", "MT::CM::AreCollinear(",
                "MT::CM::Orientation(", "MT::CM::SignedArea(", "MT::Diameter::operator==(", "MT::CountPointsBelowAllSegments(",
                "MT::SortPointsClockWiseAroundPoint(", "MT::ComputeDiameter(", "MT::PointsInTriangle(", "MT::GetHullPoints(",
                "MT::B200::SetDevices("]:
        assert sym in out, sym


def test_reference_callers_build_unmodified_against_this_library(built):
    """src/test/main.cpp and src/py-binding/main.cpp of the reference, compiled as they are, link against libadk_host.so."""
    if not os.path.isdir("/root/reference/src/test"):
        pytest.skip("the reference tree is only present in the build container")
    subprocess.run([os.path.join(ROOT, "tests", "dropin", "build_dropin.sh")], check=True, capture_output=True)
    assert os.path.exists(os.path.join(ROOT, "tests", "dropin", "_build", "ref_test_dropin"))


def test_public_helpers_of_utils_h_match_the_reference(built):
    """A caller of MT::ForAllAntipodalPairs (Utils.h:57-109), MT::ComputeDiameter (Utils.cpp:318-338) and MT::PointsInTriangle
    (Utils.h:45-49), compiled against this build's headers (tests/dropin/helpers_caller.cpp), reports what the unmodified
    reference reports: the same antipodal pairs in the same order, the same diameter pair, the same counts.  Host code only."""
    from oracle.refpy import RefLib, ref_available
    if not ref_available("x87"):
        pytest.skip("oracle/_ref is built in the build container")
    exe = os.path.join(ROOT, "tests", "dropin", "_build", "helpers_caller")
    subprocess.run([os.path.join(ROOT, "tests", "dropin", "build_dropin.sh")], check=True, capture_output=True)
    assert os.path.exists(exe)
    import json
    import random
    with open(os.path.join(ROOT, "tests", "golden", "points.json")) as f:
        points = json.load(f)
    with open(os.path.join(ROOT, "tests", "golden", "ref_outputs.json")) as f:
        runs = [r for r in json.load(f)["runs"] if r["found"] and len(r["hull"]) >= 3]
    ref = RefLib("x87")
    rng = random.Random(5)
    checked = 0
    for r in rng.sample(runs, 40):
        pts = points[r["name"]]
        if len(pts) > 120:
            continue
        x = [p[0] for p in pts]
        y = [p[1] for p in pts]
        n = len(x)
        below, coll, rb = ref.count_tables(x, y)
        tri = [rng.sample(range(n), 3) for _ in range(50)]
        text = [str(n)] + [f"{a!r} {b!r}" for a, b in zip(x, y)] + [str(len(r["hull"])), " ".join(map(str, r["hull"])), str(len(tri))]
        text += [" ".join(map(str, t)) for t in tri] + [" ".join(map(str, below)), " ".join(map(str, coll)), " ".join(map(str, rb))]
        out = subprocess.run([exe], input="\n".join(text) + "\n", capture_output=True, text=True, check=True).stdout.split()
        k = int(out[1])
        pairs = [(int(out[2 + 2 * t]), int(out[3 + 2 * t])) for t in range(k)]
        at = 2 + 2 * k
        assert out[at] == "diameter"
        diam = [int(out[at + 1]), int(out[at + 2])]
        assert out[at + 3] == "pit"
        pit = [int(v) for v in out[at + 5:at + 5 + len(tri)]]
        assert pairs == ref.antipodal_pairs(x, y, r["hull"]), r["name"]
        assert diam == ref.compute_diameter(x, y, r["hull"]), r["name"]
        assert pit == ref.points_in_triangle(x, y, tri), r["name"]
        checked += 1
    assert checked >= 20


def _gloo_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import __graft_entry__ as ge
    ge.load_package()
    from importlib import import_module
    d = import_module("area_diameter_algorithm_b200.dist")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    mine = [dict(found=1, count=9, area=2.5, tie_key=40, w0=4, w1=5), dict(found=1, count=9, area=2.5, tie_key=17, w0=1, w1=7)][rank]
    parts = d.gather_partials(mine)
    best = d.merge_partials(parts)
    tot = d.gather_stats_sum([10.0 * (rank + 1), 1.0])
    owner = d.owner_of(best, parts)
    res = {"found": 1, "count": 9, "area": 2.5, "hull": [1, 3, 7, 2], "diam": [1, 7], "kernel_ms": 0.1} if rank == owner else None
    shared = d.share_result(res, owner, 12)
    dist.destroy_process_group()
    q.put((rank, best, len(parts), tot, owner, shared))


def test_two_rank_gather_and_merge_over_gloo(built):
    """The exchanges of the multi-GPU path (one 32-byte record per rank, then the witness from its owner), world_size 2 on CPU."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, best, nparts, tot, owner, shared in got:
        assert nparts == 2
        assert best == dict(found=1, count=9, area=2.5, tie_key=17, w0=1, w1=7)
        assert tot == [30.0, 2.0]
        assert owner == 1                                    # the rank whose partial is the merged best reconstructs ...
        assert (shared["found"], shared["count"], shared["area"], shared["hull"], shared["diam"]) == (1, 9, 2.5, [1, 3, 7, 2], [1, 7])   # ... and every rank gets the witness


def test_batch_assignment_is_lpt_and_deterministic(pkg):
    """adk_batch_assign (SURVEY 8e, C5 batching): heaviest instance first onto the least loaded rank; no device needed."""
    import numpy as np
    sizes = [699, 623, 594, 526, 492, 474, 412, 382, 377, 362]            # the real cell samples
    sets = [np.zeros((n, 2)) for n in sizes]
    own = pkg.batch_assign(sets, 8)
    assert own == pkg.batch_assign(sets, 8)
    assert own[:8] == list(range(8))                                      # the eight largest start one rank each
    load = [0.0] * 8
    for n, o in zip(sizes, own):
        load[o] += float(n) ** 4
    assert max(load) == 699.0 ** 4                                        # nothing is stacked on the densest slide
    assert pkg.batch_assign(sets, 1) == [0] * 10
    assert pkg.batch_assign([], 4) == []
    # order of the input does not matter for who gets the big ones
    own_r = pkg.batch_assign(sets[::-1], 2)
    tot = [sum(float(n) ** 4 for n, o in zip(sizes[::-1], own_r) if o == r) for r in range(2)]
    assert abs(tot[0] - tot[1]) <= 699.0 ** 4


def _gloo_batch_worker(rank, world, port, q):
    import torch.distributed as dist
    import __graft_entry__ as ge
    ge.load_package()
    from importlib import import_module
    d = import_module("area_diameter_algorithm_b200.dist")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    mine = [({"count": 10 + i, "owner": rank} if i % world == rank else None) for i in range(5)]
    allr = d.gather_batch_results(mine)
    dist.destroy_process_group()
    q.put((rank, allr))


def test_two_rank_batch_gather_over_gloo(built):
    """Instance-level sharding: each rank holds the results it owns, the host gather completes the list on every rank."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_batch_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, allr in got:
        assert allr == [{"count": 10 + i, "owner": i % 2} for i in range(5)]


def test_tiled_workflow_window_grid_equals_the_reference(built):
    """tiling.make_patches against the reference's own create_patch_dict (frozen by tests/golden/make_tiled_golden.py):
    same number of non-empty windows, visited by population, largest first (no device needed)."""
    import json
    from importlib import import_module
    built.load_package()
    tiling = import_module("area_diameter_algorithm_b200.tiling")
    with open(os.path.join(ROOT, "tests", "golden", "tiled_area.json")) as f:
        cases = json.load(f)["cases"]
    with open(os.path.join(ROOT, "tests", "golden", "points.json")) as f:
        pts = json.load(f)
    for c in cases:
        xy, keep, order, member = tiling.make_patches(pts[c["name"]], c["patch_size"], c["step_size"])
        assert len(order) == c["patches"], c["name"]
        pops = [len(member(w)) for w in order]
        assert pops == sorted(pops, reverse=True) and pops[-1] >= 1
        assert len(xy) == len(set(map(tuple, pts[c["name"]])))


def test_bench_witness_check_accepts_the_frozen_answers_and_rejects_tampered_ones():
    """bench.py checks every answer of a sweep that has no frozen reference output by its witness alone (sigma = 1.0 / 0.5 beyond D = 2):
    the check must accept the unmodified reference's own answers and notice a wrong count, a dropped vertex or a diameter bound
    that the polygon does not meet."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("adk_bench", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    checked = 0
    for sigma in (1.5, 1.0):
        w = bench.load_workload(sigma)
        frozen = {float(r["dmax"]): r for r in w["sample"]}
        frozen.update(bench.frozen_answers(sigma))
        assert frozen, sigma
        for d, r in frozen.items():
            sol = {"found": True, "count": r["count"], "area": float(r["area"]), "hull": list(r["hull"])}
            assert bench.witness_is_valid(w["x"], w["y"], sol, w["amax"], d), (sigma, d)
            assert not bench.witness_is_valid(w["x"], w["y"], dict(sol, count=sol["count"] + 1), w["amax"], d)
            assert not bench.witness_is_valid(w["x"], w["y"], dict(sol, hull=sol["hull"][:-1]), w["amax"], d)
            assert not bench.witness_is_valid(w["x"], w["y"], sol, w["amax"], 0.5 * d)
            assert not bench.witness_is_valid(w["x"], w["y"], sol, 0.5 * float(r["area"]), d)
            checked += 1
    assert checked >= 6          # sigma 1.5: D = 2..6, sigma 1.0: D = 2


def test_header_is_plain_c(tmp_path):
    """The drop-in boundary is a C ABI: include/adk.h must compile as C99 (no C++ in the signatures) and adk_partial, the record
    the ranks exchange, must stay 32 bytes."""
    src = tmp_path / "t.c"
    src.write_text('#include "adk.h"\n'
                   'int main(void) { adk_result r; adk_stats s; (void) r; (void) s; return sizeof(adk_partial) == 32 ? 0 : 1; }\n')
    exe = tmp_path / "t"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                   check=True, capture_output=True)
    assert subprocess.run([str(exe)]).returncode == 0


def test_area_slot_plan_invariants(tmp_path):
    """The table-slot planner of the area path (csrc/adk_area_plan.h, host code without CUDA): on 4 000 random anchor populations
    every slot holds its own anchor and anything the queue can hand out later, every ring lies inside its slot and is at least the
    shortest ring the sweep can run with, the total stays inside the budget, and a refusal only happens when one anchor alone
    does not fit."""
    exe = tmp_path / "slot_plan_check"
    subprocess.run(["g++", "-O1", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "area-diameter-algorithm_b200", "csrc"),
                    os.path.join(ROOT, "tests", "dropin", "slot_plan_check.cpp"), "-o", str(exe)], check=True, capture_output=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
