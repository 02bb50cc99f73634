"""Configuration for every entrypoint.

The reference has no config system: positional ``sys.argv`` plus constants
scattered through the sources (reference main.py:178-184,
client.py:297-312, server.py:108-114, dataset.py:8-9, model.py:22-23,
encoder.py:38-43).  Here all of them are fields of one dataclass; the positional
argv of each reference entrypoint still parses (``from_argv``), and every field
can also be set with ``--name value`` or the ``FEDREC_<NAME>`` environment
variable.
"""
from __future__ import annotations

import dataclasses
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence


@dataclass
class FedRecConfig:
    # ---- reference argv -------------------------------------
    total_epochs: int = 1          # local steps per round in client mode (client.py:284)
    batch_size: int = 50           # users per round per client (README.md:65)
    save_every: int = 1
    dp_epsilon: float = 0.0        # 0 disables DP (client.py:304)
    run_name: str = "fedrec"
    global_epochs: int = 1         # server.py:113
    # ---- model constants ---------------------------------------------------
    news_dim: int = 400            # encoder.py:38
    num_heads: int = 20            # encoder.py:39
    head_dim: int = 20             # encoder.py:44
    query_dim: int = 200           # encoder.py:40
    dropout: float = 0.2           # encoder.py:43
    bert_dim: int = 768
    bert_layers: int = 6
    bert_heads: int = 12
    bert_ffn: int = 3072
    vocab_size: int = 30522
    max_positions: int = 512
    bert_path: str = ""            # local HF-format DistilBERT directory (config.json + model.safetensors / pytorch_model.bin);
                                   # empty: random-init trunk of the shape above (upstream downloads distilbert-base-uncased, encoder.py:19)
    # ---- data constants ------------------------------------------------------
    npratio: int = 4               # dataset.py:8
    max_his_len: int = 50          # dataset.py:9
    title_len: int = 30            # BASELINE.json config 4 (toy shard ships 50)
    data_dir: str = "UserData"     # main.py:146
    synthetic: bool = False        # generate MIND-shaped data instead of reading data_dir
    num_news: int = 65536
    num_train: int = 4000
    num_valid: int = 200
    # ---- optimiser -------------------------------------------------------------
    lr: float = 5e-5               # model.py:22-23
    beta1: float = 0.9
    beta2: float = 0.999
    adam_eps: float = 1e-8
    # ---- federated / privacy ----------------------------------------------------
    mode: str = "ddp"              # ddp | grad_avg | param_avg | client_server
    local_steps: int = 1           # steps between aggregations (param_avg / client_server)
    clip_norm: float = 0.0         # per-client L2 clip C (0 = off); client.py:220 says 2
    noise_multiplier: float = 0.0  # sigma; noise std on the clipped gradient is sigma*C
    dp_delta: float = 1e-5         # client.py:222
    dp_epochs: int = 50            # client.py:223
    news_grad_noise: float = 0.0   # reference behaviour: N(0, s) on news-vector grads (client.py:87-89)
    server_rank: int = 1           # client.py:257 hard-codes src=1
    client_weights: Optional[List[float]] = None  # FedAvg weights; None = uniform
    average_frozen: bool = False   # include the frozen DistilBERT in parameter averaging (PA:144)
    transport: str = "nvlink"      # nvlink | collective | tcp   (client/server upload path)
    server_host: str = "127.0.0.1"  # client.py:193 hard-codes an EC2 address
    server_port: int = 12345       # client.py:194
    # ---- runtime -----------------------------------------------------------------
    backend: str = "auto"          # auto → nccl when CUDA is present, else gloo
    device: str = "auto"
    dtype: str = "bf16"            # bf16 = the sm_100a kernels (bf16 tensor cores, fp32 accumulate); fp32 = the PyTorch step (debugging)
    fused: bool = True             # use the sm_100a kernels when available
    dedup: bool = True             # encode each distinct news id once per step
    mask_padding: bool = False     # the reference never masks
    trunk_dropout_in_update: bool = False  # model.py:73 runs DistilBERT in train() for the 2nd pass
    cuda_graph: bool = True
    trunk_cache: bool = False      # keep the frozen trunk's hidden states of every news item seen resident in HBM
    ln_fold: bool = True           # fold 11 of the trunk's 12 LayerNorms into the neighbouring GEMM epilogues (exact algebra)
    seed: int = 1234
    snapshot_path: str = "snapshot.pt"
    log_dir: str = "logs"
    precomputed_news: bool = False  # BASELINE.json config 1: user encoder over cached news vectors
    timeout_s: float = 1800.0      # host-side waits: process group, store barriers (upstream: two days, client.py:227)
    agg_timeout_s: float = 30.0    # in-kernel bound on waiting for a peer GPU; on expiry the run degrades to NCCL
    agg_blocks: int = 0            # grid of the aggregation kernels (0 = default)
    pipeline_optimizer: bool = True  # apply step i's aggregation + Adam at the start of step i+1, beside the frozen trunk
    steps_per_epoch: int = 0       # 0 = one pass over the client's shard per epoch (upstream main.py:55-91)
    participation: float = 1.0     # param_avg / client_server: fraction of the clients that train in a round (FedAvg's C); the rest enter the average with weight 0
    validate_every: int = 0        # ddp / grad_avg / param_avg: validate every N epochs (0 = never, as upstream; the client role always does)
    native_loader: bool = False    # batches from the C++ producer thread (csrc/host_loader.cu) through TrainStep.run_host
    heartbeat_s: float = 5.0       # straggler/failure detector period when world_size > 1 (0 = off)

    # ------------------------------------------------------------------------------
    @classmethod
    def field_names(cls) -> List[str]:
        return [f.name for f in dataclasses.fields(cls)]

    @classmethod
    def from_env(cls, **overrides) -> "FedRecConfig":
        cfg = cls()
        for f in dataclasses.fields(cls):
            key = "FEDREC_" + f.name.upper()
            if key in os.environ:
                setattr(cfg, f.name, _coerce(f, os.environ[key]))
        for k, v in overrides.items():
            setattr(cfg, k, v)
        return cfg

    @classmethod
    def from_argv(cls, argv: Sequence[str], positional: Sequence[str], **overrides) -> "FedRecConfig":
        """Parse ``argv`` (without the program name).

        Leading bare values fill ``positional`` in order, exactly like the reference
        entrypoints; anything of the form ``--name value`` / ``--name=value`` /
        ``--flag`` sets the matching field.
        """
        cfg = cls.from_env(**overrides)
        fields = {f.name: f for f in dataclasses.fields(cls)}
        pos = list(positional)
        i = 0
        argv = list(argv)
        while i < len(argv):
            tok = argv[i]
            if tok.startswith("--"):
                name, eq, val = tok[2:].partition("=")
                name = name.replace("-", "_")
                if name.startswith("no_") and name[3:] in fields and fields[name[3:]].type in ("bool", bool):
                    setattr(cfg, name[3:], False)
                    i += 1
                    continue
                if name not in fields:
                    raise SystemExit(f"unknown option --{name}; known: {sorted(fields)}")
                f = fields[name]
                if not eq:
                    if f.type in ("bool", bool) and (i + 1 >= len(argv) or argv[i + 1].startswith("--")):
                        val = "1"
                    else:
                        i += 1
                        if i >= len(argv):
                            raise SystemExit(f"option --{name} needs a value")
                        val = argv[i]
                setattr(cfg, name, _coerce(f, val))
            else:
                if not pos:
                    raise SystemExit(f"unexpected positional argument {tok!r}")
                name = pos.pop(0)
                setattr(cfg, name, _coerce(fields[name], tok))
            i += 1
        return cfg

    # derived ----------------------------------------------------------------------
    @property
    def num_candidates(self) -> int:
        return 1 + self.npratio

    def resolved_backend(self) -> str:
        if self.backend != "auto":
            return self.backend
        import torch
        return "nccl" if torch.cuda.is_available() else "gloo"

    def resolved_device(self):
        import torch
        if self.device != "auto":
            return torch.device(self.device)
        if torch.cuda.is_available():
            return torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
        return torch.device("cpu")

    def to_dict(self):
        return dataclasses.asdict(self)


def _coerce(f: dataclasses.Field, text: str):
    t = f.type if isinstance(f.type, str) else getattr(f.type, "__name__", str(f.type))
    if t == "bool":
        return str(text).lower() in ("1", "true", "yes", "on")
    if t == "int":
        return int(float(text))
    if t == "float":
        return float(text)
    if t.startswith("Optional[List[float]]"):
        if text in ("", "none", "None"):
            return None
        return [float(x) for x in str(text).split(",")]
    return text
