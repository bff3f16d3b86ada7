"""State file bridge: read and write the reference's restart files directly from the
partitioned state of a B200 run, so that runs can start from, and be compared with, reference runs at any rank count.

File format (wx_factory/output/state.py:9-33): ``numpy.save`` of the GLOBAL state
``(6, num_var, [nz,] nh, nh, P)`` (``ProcessTopology.gather_cube``, wx_factory/process_topology.py:444-470: panels
stacked on the first axis), followed by one text line with the configuration schema version and the raw INI text.
File name (wx_factory/output/output_manager.py:76-95): ``state_vector_<hash>.<step:08d>.npy`` with
``hash = hash((dt, nh_total, nz, n)) & 0xFFFFFFFFFFFF``.

A rank of a 1/2/4/8-way partition holds the rows of element columns its slabs cover (partition.py); ``gather_state``
/ ``scatter_state`` move between that flat local layout and the global array (through torch.distributed when there is
more than one rank; the file itself is written / read by rank 0 like the reference's ``SingleProcess`` block).
"""

import io
import os

import numpy

from .layout import local_state, scatter_back

STATE_VERSION = "0.1"  # config/config-format.json "version" of the reference this was written against


def config_hash(dt, num_elements_horizontal_total, num_elements_vertical, num_solpts):
    """output_manager.py:76-83 (Python's tuple hash of numbers is deterministic)."""
    return (dt, num_elements_horizontal_total, num_elements_vertical, num_solpts).__hash__() & 0xFFFFFFFFFFFF


def state_file_name(output_dir, step_id, dt, num_elements_horizontal_total, num_elements_vertical, num_solpts):
    """output_manager.py:91-95."""
    h = config_hash(dt, num_elements_horizontal_total, num_elements_vertical, num_solpts)
    return f"{output_dir}/state_vector_{h:012x}.{step_id:08d}.npy"


def save_state(state, config_content, output_file_name, state_version=STATE_VERSION):
    """state.py:9-17: the array, then ``<version>\\n`` and the INI text."""
    with open(output_file_name, "wb+") as f:
        numpy.save(f, numpy.asarray(state))
        f.write(bytes(f"{state_version}\n", "utf-8"))
        f.write(bytes(config_content, "utf-8"))


def load_state(input_file_name):
    """state.py:20-33 without building a Configuration: returns (state, state_version, config_content)."""
    with open(input_file_name, "rb") as f:
        state = numpy.load(f)
        version = str(f.readline(), "utf-8").strip()
        content = "\n".join(str(line, "utf-8") for line in f.readlines()).strip()
    return state, version, content


def _peer_partition(rhs, r):
    from .partition import build_partition

    prob = rhs.prob
    if r == rhs.rank:
        return rhs.part
    euler = prob.equations == "euler"
    return build_partition(prob.nh, prob.nz if euler else 1, prob.n, prob.bnd_sn, prob.bnd_we, nranks=rhs.nranks, rank=r,
                           num_var=rhs.num_trace, nfaces=6 if euler else 4, trace_layout="element" if euler else "field")


def _wire(group):
    """Where point-to-point tensors must live for the group's backend: NCCL moves device memory, gloo host memory."""
    import torch
    import torch.distributed as dist

    if dist.get_backend(group) == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def _global_rank(group, r):
    import torch.distributed as dist

    return r if group is None else dist.get_global_rank(group, r)


def gather_state(rhs, local, group=None):
    """Flat local state of every rank (numpy or CUDA tensor) -> global ``prob.state_shape`` numpy array on rank 0
    (None elsewhere).  ``rhs``: this rank's ``RhsB200`` (knows the problem and the partition).  The pieces travel one
    at a time as float64 tensors, point to point (ProcessTopology.gather_cube, process_topology.py:444-470, moves
    arrays, not pickles): rank 0 holds one piece besides the result, whatever the number of ranks."""
    prob = rhs.prob
    if rhs.nranks == 1:
        loc = local.detach().cpu().numpy() if hasattr(local, "detach") else numpy.asarray(local)
        return scatter_back(prob, rhs.part, loc.reshape(-1), numpy.empty(prob.state_shape))
    import torch
    import torch.distributed as dist

    wire = _wire(group)
    mine = local.detach() if hasattr(local, "detach") else torch.from_numpy(numpy.ascontiguousarray(local, dtype=numpy.float64))
    mine = mine.reshape(-1).to(wire)
    if rhs.rank != 0:
        dist.send(mine.contiguous(), dst=_global_rank(group, 0), group=group)
        return None
    out = numpy.empty(prob.state_shape)
    scatter_back(prob, rhs.part, mine.cpu().numpy(), out)
    buf = torch.empty_like(mine)  # 6 nh rows split evenly: every rank's piece has this rank's size
    for r in range(1, rhs.nranks):
        dist.recv(buf, src=_global_rank(group, r), group=group)
        scatter_back(prob, _peer_partition(rhs, r), buf.cpu().numpy(), out)
    return out


def scatter_state(rhs, global_state, group=None):
    """Global state on rank 0 (ignored elsewhere) -> this rank's flat local numpy state.  Rank 0 cuts every rank's piece
    out of the global array and sends only that (ProcessTopology.distribute_cube, process_topology.py:472-539)."""
    if rhs.nranks == 1:
        return rhs.local_state(global_state)
    import torch
    import torch.distributed as dist

    wire = _wire(group)
    if rhs.rank == 0:
        for r in range(1, rhs.nranks):
            tiled = getattr(rhs.prob, "local_state", None)  # synthetic tiled problems cut their own pieces (rhs.local_state)
            part = _peer_partition(rhs, r)
            piece = tiled(part, global_state) if tiled is not None else local_state(rhs.prob, part, global_state)
            piece = numpy.ascontiguousarray(piece, dtype=numpy.float64).reshape(-1)
            dist.send(torch.from_numpy(piece).to(wire), dst=_global_rank(group, r), group=group)
        return rhs.local_state(global_state)
    size = getattr(rhs, "size", None) or int(numpy.prod(rhs.prob.state_shape)) // rhs.nranks  # 6 nh rows split evenly
    buf = torch.empty(int(size), dtype=torch.float64, device=wire)
    dist.recv(buf, src=_global_rank(group, 0), group=group)
    return buf.cpu().numpy()


def write_restart(rhs, local, config_content, output_dir, step_id, dt, group=None):
    """What OutputManager.step does when save_state_freq hits (output_manager.py:136-143)."""
    total = gather_state(rhs, local, group)
    if rhs.rank != 0:
        return None
    prob = rhs.prob
    os.makedirs(os.path.abspath(output_dir), exist_ok=True)
    name = state_file_name(output_dir, step_id, dt, prob.nh, prob.nz if prob.equations == "euler" else 1, prob.n)
    save_state(total, config_content, name)
    return name


def read_restart(rhs, input_file_name, group=None):
    """OutputManager.load_state_from_file (output_manager.py:97-124): the local state of this rank + the INI text.
    Rank 0 reads the file; a missing / corrupt file or a state of the wrong shape raises on EVERY rank (the reference
    raises on every rank after distributing the state), never on rank 0 alone while the others wait in a collective."""
    state = content = error = None
    if rhs.rank == 0:
        try:
            state, _, content = load_state(input_file_name)
            if tuple(state.shape) != tuple(rhs.prob.state_shape):
                raise ValueError(f"ERROR reading state vector from file. The shape is wrong! ({state.shape}, should be {rhs.prob.state_shape})")
        except Exception as e:  # noqa: BLE001 -- shipped to the other ranks below
            state, error = None, e
    if rhs.nranks > 1:
        import torch.distributed as dist

        box = [(type(error).__name__, str(error)) if error is not None else None, content]
        dist.broadcast_object_list(box, src=0, group=group)
        if box[0] is not None and error is None:
            error = (ValueError if box[0][0] == "ValueError" else OSError)(box[0][1])
        content = box[1]
    if error is not None:
        raise error
    return scatter_state(rhs, state, group), content


def dumps(state, config_content, state_version=STATE_VERSION):
    """The file's bytes (for tests)."""
    b = io.BytesIO()
    numpy.save(b, numpy.asarray(state))
    b.write(bytes(f"{state_version}\n", "utf-8"))
    b.write(bytes(config_content, "utf-8"))
    return b.getvalue()
