"""`torchrun ... server.py` — same command line as the upstream script of this name.

Positional arguments are unchanged; any `FedRecConfig` field can follow
as `--name value`. The work happens in `fedrec_b200.engine.federated`.
"""
import sys

from fedrec_b200.engine.federated import run_entrypoint
from fedrec_b200.utils.transport import StateDictServer, average_state_dicts


def handle_client(conn, addr, no, out_dir="."):
    """Receive one client's state_dict upload into ``received_model_{no}.pt`` (upstream server.py:25-35)."""
    srv = StateDictServer.__new__(StateDictServer)
    srv.out_dir = out_dir
    paths = [None] * no
    srv._handle(conn, no, paths)
    return paths[no - 1]


def average_all_params_rcvd(world_size=None, out_dir=".", weights=None):
    """Uniform (or weighted) mean of ``received_model_{1..world_size-1}.pt`` (upstream server.py:37-55)."""
    import os
    import torch.distributed as dist
    if world_size is None:
        world_size = dist.get_world_size()
    paths = [os.path.join(out_dir, f"received_model_{i}.pt") for i in range(1, world_size)]
    return average_state_dicts(paths, weights)


def main(argv=None):
    return run_entrypoint("server", list(sys.argv[1:] if argv is None else argv))


if __name__ == "__main__":
    main()
