"""`torchrun ... server.py` — same command line as the upstream script of this name.

Positional arguments are unchanged; any `FedRecConfig` field can follow
as `--name value`. The work happens in `fedrec_b200.engine.federated`.
"""
import sys

from fedrec_b200.engine.federated import run_entrypoint


def main(argv=None):
    return run_entrypoint("server", list(sys.argv[1:] if argv is None else argv))


if __name__ == "__main__":
    main()
