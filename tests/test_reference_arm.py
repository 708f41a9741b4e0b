"""bench.py's reference arm: the compiled reference of oracle/_ref (oracle/make_ref.py) reproduces BASELINE.json configs[0]'s
known answer, and both arms describe the workload with the same `config` object."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from oracle import ref_baseline as rb  # noqa: E402

needs_ref = pytest.mark.skipif(not rb.available(), reason="oracle/_ref not built (python -m oracle.make_ref in the build container)")


@needs_ref
def test_compiled_reference_reproduces_config1_known_answer():
    r = bench.config1_reference(2)
    assert r["frames"] == 37 and r["final_state"] == 8
    assert r["final_pos"] == [812.0036311681246, 542.0]            # BASELINE.json configs[0]
    assert r["known_answer_ok"]


@needs_ref
def test_reference_pool_counts_ticks_of_the_workload_policy():
    arr, lens = bench.load_maps("simplest_with_mines")
    slices, wall, procs, envs = rb.run_slices([arr[k, :lens[k]] for k in range(4)], 2, 0.3, 2, envs_per_process=2)
    assert procs == 2 and envs == 4 and len(slices) == 2
    assert all(rate > 100 and ticks > 30 for rate, ticks in slices)


def test_both_arms_describe_the_workload_identically():
    """The driver compares the two arms' `config` objects.  Both arms must build theirs from the command line through the same
    function (and through nothing else), and that function must not depend on --impl or on anything only one arm knows."""
    import inspect
    import json
    for fn in (bench.run_reference, bench.main):
        src = inspect.getsource(fn)
        assert src.count('"config": headline_config(args, ') == 1, fn.__name__
        assert "config_dict(" not in src.replace("headline_config(", ""), fn.__name__
    n_levels = len(bench.load_maps(bench.WORKLOADS["c3"][0])[0])
    for flags in ([], ["--envs", "131072", "--workload", "c4"], ["--no-flush", "--placement", "adjacent"]):
        ours = bench.build_parser().parse_args(flags + ["--gpus", "1", "--steps", "20", "--warmup", "5"])
        ref = bench.build_parser().parse_args(flags + ["--impl", "reference", "--gpus", "1", "--steps", "20", "--warmup", "5"])
        a, b = json.dumps(bench.headline_config(ours, n_levels)), json.dumps(bench.headline_config(ref, n_levels))
        assert a == b
        cfg = json.loads(a)
        assert set(cfg) >= {"workload", "envs_per_gpu", "frame_skip", "max_frames", "distinct_levels"} and "model" not in cfg
    assert json.loads(a)["l2"] == "not flushed" and json.loads(a)["caller_env_order"] == "adjacent"


def test_reference_modules_are_compiler_output_only():
    """oracle/_ref holds no reference source text (only marshalled code objects and a manifest)."""
    root = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(root):
        pytest.skip("oracle/_ref not built")
    for d, _, files in os.walk(root):
        for f in files:
            assert f.endswith(".refc") or f == "MANIFEST.txt", os.path.join(d, f)
