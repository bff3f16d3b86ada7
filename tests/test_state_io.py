"""State file bridge (wxfactory_b200/state_io.py) against a restart file written by the UNMODIFIED reference
(tests/golden/state_dcmip21/, made by oracle/refharness/make_state_golden.py): same name, same bytes, readable,
and the partitioned gather / scatter round-trips for 1, 2, 4 and 8 ranks."""

import glob
import os
import types

import numpy

from conftest import GOLDEN
from wxfactory_b200 import state_io
from wxfactory_b200.layout import local_state, scatter_back
from wxfactory_b200.partition import build_partition

FILE = glob.glob(os.path.join(GOLDEN, "state_dcmip21", "state_vector_*.npy"))[0]


def test_reads_the_reference_file_and_reproduces_its_name(golden):
    prob, ex = golden("dcmip21")
    state, version, content = state_io.load_state(FILE)
    assert version == state_io.STATE_VERSION
    assert numpy.array_equal(state, ex["Q0"])  # the reference saved its initial state
    assert "[Time_integration]" in content or "dt" in content
    want = state_io.state_file_name(os.path.dirname(FILE), 7, 25.0, prob.nh, prob.nz, prob.n)
    assert want == FILE
    assert state_io.config_hash(25, 3, 4, 3) == state_io.config_hash(25.0, 3, 4, 3)


def test_writes_the_same_bytes(golden, tmp_path):
    prob, ex = golden("dcmip21")
    _, version, _ = state_io.load_state(FILE)
    raw = open(FILE, "rb").read()
    # the INI text is stored verbatim after the version line: recover it from the file itself
    head = state_io.dumps(ex["Q0"], "", version)
    content = raw[len(head):].decode("utf-8")
    out = tmp_path / "s.npy"
    state_io.save_state(ex["Q0"], content, str(out), version)
    assert open(out, "rb").read() == raw


def test_partitioned_round_trip(golden):
    for name in ("dcmip31", "dcmip21", "case6"):
        prob, ex = golden(name)
        euler = prob.equations == "euler"
        for nranks in (1, 2, 4, 8):
            if (6 * prob.nh) % nranks:
                continue  # the partition cuts whole rows of element columns (partition.py)
            out = numpy.full(prob.state_shape, numpy.nan)
            for r in range(nranks):
                part = build_partition(prob.nh, prob.nz if euler else 1, prob.n, prob.bnd_sn, prob.bnd_we, nranks=nranks, rank=r,
                                       num_var=6 if euler else 3, nfaces=6 if euler else 4, trace_layout="element" if euler else "field")
                scatter_back(prob, part, local_state(prob, part, ex["Q0"]), out)
            assert numpy.array_equal(out, ex["Q0"].reshape(prob.state_shape))


def test_restart_helpers_single_rank(golden, tmp_path):
    prob, ex = golden("dcmip21")
    part = build_partition(prob.nh, prob.nz, prob.n, prob.bnd_sn, prob.bnd_we, nranks=1, rank=0, num_var=6, nfaces=6, trace_layout="element")
    rhs = types.SimpleNamespace(prob=prob, part=part, nranks=1, rank=0, num_trace=6, local_state=lambda Q: local_state(prob, part, Q))
    name = state_io.write_restart(rhs, local_state(prob, part, ex["Q0"]), "[General]\ndt = 25\n", str(tmp_path), 3, 25.0)
    assert os.path.basename(name) == os.path.basename(FILE).replace("00000007", "00000003")
    loc, content = state_io.read_restart(rhs, name)
    assert numpy.array_equal(loc, local_state(prob, part, ex["Q0"])) and content.startswith("[General]")


def test_save_load_works_like_the_reference_unit_test(tmp_path):
    """tests/unit/output/test_state.py::test_save_load_works of the reference, replayed: a seeded 5-vector
    (random.Random(5646459), uniform in [-10, 10]) and an INI text survive save_state / load_state unchanged."""
    import random

    rnd = random.Random(5646459)
    arr = numpy.array([rnd.uniform(-10, 10) for _ in range(5)])
    ini = "[General]\nequations = euler\n\n[Grid]\nnum_solpts = 4\n"
    out = tmp_path / "test_state_data"
    state_io.save_state(arr, ini, str(out))
    data, version, content = state_io.load_state(str(out))
    assert data.ndim == 1 and data.shape[0] == 5
    assert all(arr[i] == data[i] for i in range(5))
    assert version == state_io.STATE_VERSION
    import configparser

    a, b = configparser.ConfigParser(), configparser.ConfigParser()
    a.read_string(ini)
    b.read_string(content)
    assert {s: dict(a[s]) for s in a.sections()} == {s: dict(b[s]) for s in b.sections()}


def test_read_restart_rejects_a_state_of_the_wrong_shape(golden, tmp_path):
    """OutputManager.load_state_from_file raises ValueError when the stored state does not fit the run
    (wx_factory/output/output_manager.py:106-110): same here."""
    import pytest

    prob, ex = golden("dcmip21")
    part = build_partition(prob.nh, prob.nz, prob.n, prob.bnd_sn, prob.bnd_we, nranks=1, rank=0, num_var=6, nfaces=6, trace_layout="element")
    rhs = types.SimpleNamespace(prob=prob, part=part, nranks=1, rank=0, num_trace=6, local_state=lambda Q: local_state(prob, part, Q))
    bad = tmp_path / "bad.npy"
    state_io.save_state(numpy.zeros((6, 5, 2, 3, 3, 27)), "[General]\n", str(bad))
    with pytest.raises(ValueError):
        state_io.read_restart(rhs, str(bad))
