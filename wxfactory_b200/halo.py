"""Face-trace halo exchange between ranks: the one collective on the path.

Replaces the reference's ``Ineighbor_alltoall`` on a dist-graph communicator
(wx_factory/process_topology.py:269-386, wx_factory/rhs/rhs_dfr.py:141-201) by point-to-point sends of the packed
trace records each peer needs (`Partition.send_peers` / `recv_peers`, built in partition.py so that the sender's
record order IS the receiver's ghost order).  Works on any torch.distributed backend: NCCL on the GPUs, gloo in the
CPU tests.
"""

import torch.distributed as dist


def exchange(part, send_t, recv_t, rec, group=None):
    """Send `part.send_peers` slices of `send_t` and receive `part.recv_peers` slices into `recv_t`
    (flat tensors; `rec` = doubles per record).  Blocks the host until the operations have been enqueued
    (NCCL) or completed (gloo)."""
    ops = []
    for peer, first, cnt in part.send_peers:
        ops.append(dist.P2POp(dist.isend, send_t[first * rec : (first + cnt) * rec], peer, group=group))
    for peer, first, cnt in part.recv_peers:
        ops.append(dist.P2POp(dist.irecv, recv_t[first * rec : (first + cnt) * rec], peer, group=group))
    if not ops:
        return
    for w in dist.batch_isend_irecv(ops):
        w.wait()
