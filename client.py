"""`torchrun ... client.py` — same command line as the upstream script of this name.

Positional arguments are unchanged; any `FedRecConfig` field can follow
as `--name value`. The work happens in `fedrec_b200.engine.federated`.
"""
import sys

from fedrec_b200.engine.federated import run_entrypoint
from fedrec_b200.engine.trainer import Trainer, process_news_grad, process_user_grad, train_on_step  # noqa: F401
from fedrec_b200.utils.transport import send_saved_model  # noqa: F401  (upstream client.py:191-210)


def main(argv=None):
    return run_entrypoint("client", list(sys.argv[1:] if argv is None else argv))


if __name__ == "__main__":
    main()
