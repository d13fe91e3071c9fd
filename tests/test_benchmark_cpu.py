"""Reader of the reference's preset benchmark samples (Env(preset=True) / modify_preset,
DCML_BID_FIRST_MA_ENV_SingleProcess.py:25-32, 344-353) against what the reference itself holds after reading
``data/dcml_benchmark/Sample_1*`` (fixture ``preset_sample1``, tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from conftest import golden
from mat_dcml_b200.benchmark import PresetSamples

REF_DIR = "/root/reference/data/dcml_benchmark"


def _from_fixture():
    g = golden("preset_sample1")
    return g, PresetSamples(g["master"], g["worker_prs"], g["disable_rates"])


def test_preset_files_round_trip(tmp_path):
    g, s = _from_fixture()
    mp, wp = str(tmp_path / "Sample_1master_states.npy"), str(tmp_path / "Sample_1worker_states.npy")
    s.save(mp, wp)
    t = PresetSamples.load(mp, wp)
    assert len(t) == len(g["master"]) and t.n_workers == 100
    np.testing.assert_array_equal(t.master_status, g["master"])
    np.testing.assert_array_equal(t.worker_Prs, g["worker_prs"])
    np.testing.assert_array_equal(t.disable_rates, g["disable_rates"])
    assert t.disable_rates.dtype == np.int64


def test_modify_preset_matches_reference():
    g, s = _from_fixture()
    s.modify_preset(R=(2 ** 19), C=(9 * (2 ** 10) / 10))
    np.testing.assert_array_equal(s.master_status, g["master_RC"])
    s.modify_preset(disable_rate=24)
    np.testing.assert_array_equal(s.disable_rates, g["disable_rates_24"])
    s.modify_preset(Pr=0.3)
    np.testing.assert_array_equal(s.worker_Prs, g["worker_prs_03"])
    np.testing.assert_array_equal(s.master_status[:, 2], g["master"][:, 2])      # the master's Pr column is left alone


def test_bad_shapes_rejected():
    g, _ = _from_fixture()
    with pytest.raises(ValueError):
        PresetSamples(g["master"][:, :2], g["worker_prs"], g["disable_rates"])
    with pytest.raises(ValueError):
        PresetSamples(g["master"], g["worker_prs"][:-1], g["disable_rates"])


@pytest.mark.skipif(not os.path.isdir(REF_DIR), reason="reference data only exists in the build container")
def test_reads_the_reference_files():
    g, _ = _from_fixture()
    s = PresetSamples.load(os.path.join(REF_DIR, "Sample_1master_states.npy"), os.path.join(REF_DIR, "Sample_1worker_states.npy"))
    n = len(g["master"])
    assert len(s) == int(g["n_rows"])
    np.testing.assert_array_equal(s.master_status[:n], g["master"])
    np.testing.assert_array_equal(s.worker_Prs[:n], g["worker_prs"])
    np.testing.assert_array_equal(s.disable_rates[:n], g["disable_rates"])
