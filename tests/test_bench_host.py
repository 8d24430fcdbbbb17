"""CPU: the host side of bench.py - the workloads both arms are timed on, the `config` object they report, and the
committed ncu figure behind `roofline.traffic`.  No GPU, no compute."""
import json
import os

import numpy as np
import pytest

import bench


@pytest.mark.parametrize("wl", ["default", "config4"])
def test_workload_inputs_are_deterministic_and_well_formed(wl):
    a, b = bench.make_inputs(0, wl), bench.make_inputs(0, wl)
    for k in a:
        assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), k
    faces, frames = a["crops"].shape[0], a["frames"].shape[0]
    assert faces == bench.WORKLOADS[wl]["faces"] and frames == bench.WORKLOADS[wl]["frames"]
    assert a["crops"].shape[1:] == (256, 256, 3) and a["frames"].shape[1:] == (bench.H, bench.W, 3)
    assert a["masks"].shape == (faces, 224, 224) and a["tfms"].shape == (faces, 2, 3)
    ff, fs = np.asarray(a["face_frame"]), np.asarray(a["face_source"])
    assert ff.shape == (faces,) and fs.shape == (faces,)
    assert np.all(np.diff(ff) >= 0) and ff.min() == 0 and ff.max() == frames - 1     # faces come frame by frame
    assert fs.min() >= 0 and fs.max() < a["sources"].shape[0]
    assert a["crops"].dtype == np.uint8 and a["frames"].dtype == np.uint8
    if wl == "config4":   # four faces per frame, each swapped with its own source identity
        assert np.all(np.bincount(ff) == 4) and a["sources"].shape[0] == 4


def test_other_seeds_give_other_inputs():
    a, b = bench.make_inputs(0, "config4"), bench.make_inputs(1, "config4")
    assert not np.array_equal(a["crops"], b["crops"])


@pytest.mark.parametrize("wl", ["default", "config4"])
def test_both_arms_report_the_same_config(wl):
    """The driver compares the `config` objects of the two arms: same workload string, same keys, no host-specific
    entries (the NUMA node is a top-level key of our line)."""
    c1, c8 = bench.workload_config(1, wl), bench.workload_config(8, wl)
    assert c1 == bench.workload_config(1, wl)
    assert set(c1) == set(c8) and "host_numa_node" not in c1
    assert "workload" in c1 and "model" not in c1
    assert c1["faces_per_step_per_gpu"] == bench.WORKLOADS[wl]["faces"]
    json.dumps(c1)


def test_config5_config_names_the_sr_workload():
    c = bench.config5_config(1)
    assert "configs[4]" in c["workload"] and "model" not in c
    json.dumps(c)


def test_roofline_traffic_comes_from_the_committed_ncu_capture():
    t = bench.ncu_traffic()
    path = os.path.join(bench.ROOT, "profiles", "ncu_traffic.json")
    assert os.path.exists(path)
    with open(path) as f:
        j = json.load(f)
    assert t == j["igemm_kernel"] and 1e6 < t < 1e10 and "ncu" in j["note"]
    assert os.path.exists(os.path.join(bench.ROOT, "profiles", "r2_ncu_by_op.csv"))
