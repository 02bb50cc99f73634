"""Import shim that lets the *unmodified* upstream sources in ``baseline/_ref`` run on this box.

Nothing in ``baseline/_ref`` is edited. What the shim supplies are the things upstream assumes
from its environment:

* ``pickle5`` — does not exist for Python >= 3.8: aliased to the standard ``pickle``;
* ``DistilBertModel.from_pretrained("distilbert-base-uncased")`` needs a network download:
  replaced by a random-init model from the default ``DistilBertConfig`` (BASELINE.json fixes all
  runs to random-init weights);
* ``wandb`` / ``opacus`` stubs, only needed if ``client.py`` is imported.

The upstream modules import each other by bare name, so ``baseline/_ref`` is put first on
``sys.path`` while they load and the framework's own same-named root modules are shadowed.
"""
from __future__ import annotations

import importlib
import pickle
import sys
import types


def _install_stubs() -> None:
    sys.modules.setdefault("pickle5", pickle)
    if "wandb" not in sys.modules:
        w = types.ModuleType("wandb")
        w.login = lambda *a, **k: None
        w.init = lambda *a, **k: None
        w.log = lambda *a, **k: None
        sys.modules["wandb"] = w
    if "opacus" not in sys.modules:
        o = types.ModuleType("opacus")

        class PrivacyEngine:                                   # only ever used upstream to read a sigma
            def make_private_with_epsilon(self, **kw):
                from fedrec_b200.utils.privacy import get_noise_multiplier
                dl = kw["data_loader"]
                q = min(1.0, dl.batch_size / max(len(dl.dataset), 1))
                opt = types.SimpleNamespace(noise_multiplier=get_noise_multiplier(
                    kw["target_epsilon"], kw["target_delta"], q, epochs=kw["epochs"]))
                return kw["module"], opt, dl
        o.PrivacyEngine = PrivacyEngine
        sys.modules["opacus"] = o
    import transformers

    def _from_config(*_a, **_k):
        return transformers.DistilBertModel(transformers.DistilBertConfig())
    transformers.DistilBertModel.from_pretrained = staticmethod(_from_config)


def load_reference(ref_dir: str, module: str = "Gradient_Averaging_main"):
    _install_stubs()
    shadowed = ["model", "encoder", "attention", "dataset", "evaluation_functions", "main", "client", "server",
                "Gradient_Averaging_main", "Parameter_Averaging_main"]
    saved = {k: sys.modules.pop(k) for k in shadowed if k in sys.modules}
    sys.path.insert(0, ref_dir)
    try:
        mod = importlib.import_module(module)
        for k in shadowed:                                     # keep upstream modules under private names
            if k in sys.modules:
                sys.modules["_upstream_" + k] = sys.modules[k]
    finally:
        sys.path.remove(ref_dir)
        for k in shadowed:
            sys.modules.pop(k, None)
        sys.modules.update(saved)
    return mod
