"""Golden restart file -- TEST INFRASTRUCTURE ONLY: the UNMODIFIED reference's ``save_state``
(wx_factory/output/state.py:9-17) and ``OutputManager.state_file_name`` hash (output_manager.py:76-95) for
config/dcmip21.ini, written to tests/golden/state_dcmip21/ (the file the B200 state bridge must reproduce byte for byte
and be able to read).

    python oracle/refharness/make_state_golden.py
"""

import os
import sys

import numpy

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
sys.path.insert(0, _HERE)
sys.path.insert(0, _ROOT)

import harness  # noqa: E402


def main():
    harness.bootstrap()
    st = harness.build(harness.read_ini("dcmip21.ini"))
    from output.state import load_state, save_state

    cfg = st[0].config
    total = numpy.stack([s.Q for s in st])  # gather_cube with one rank per panel (process_topology.py:444-470)
    state_params = (cfg.dt, cfg.num_elements_horizontal_total, cfg.num_elements_vertical, cfg.num_solpts)
    h = state_params.__hash__() & 0xFFFFFFFFFFFF
    out_dir = os.path.join(_ROOT, "tests", "golden", "state_dcmip21")
    os.makedirs(out_dir, exist_ok=True)
    name = os.path.join(out_dir, f"state_vector_{h:012x}.{7:08d}.npy")
    save_state(total, cfg, name, device=st[0].device)
    back, conf = load_state(name, schema=cfg.schema, device=st[0].device)
    assert numpy.array_equal(back, total) and conf.dt == cfg.dt
    print(name, os.path.getsize(name), "bytes; dt", cfg.dt, "nh", cfg.num_elements_horizontal_total, "nz", cfg.num_elements_vertical, "n", cfg.num_solpts)


if __name__ == "__main__":
    main()
