"""Process-wide runtime behind the reference-facing classes: one ``TacoEngine`` per (process, device).

The reference builds a TF graph and owns variables through ``tf.variable_scope('model')`` +
``tf.train.Saver`` (Utils/Tacotron_synthesizer.py:24,49-50).  Here the variables are a plain
``name -> ndarray`` container (weights.py; names = reference variable names, TF layouts) that is
uploaded once through ``taco_set_weight``; the classes in ``TacotronModel/modules`` are
descriptors that drive the C ABI through the engine returned by :func:`engine`.
"""
from __future__ import annotations

import os

_state = {"engine": None, "key": None, "weights": None, "precision": os.environ.get("TACO_PRECISION", "bf16"),
          "trainer": None, "trainer_key": None}


def set_weights(weights):
    """Install the variable container used by the next :func:`engine` call (``saver.restore``)."""
    _state["weights"] = weights
    close()


def set_precision(precision):
    _state["precision"] = precision
    close()


def weights(hparams):
    if _state["weights"] is None:
        from . import weights as W
        # no checkpoint: TF-default random init with the reference's seed (Hyperparams.py:162)
        _state["weights"] = W.init_weights(hparams)
    return _state["weights"]


def device_index():
    import torch
    if not torch.cuda.is_available():
        from ._lib import TacoError
        raise TacoError("no CUDA device: the B200 path has no CPU fallback")
    return int(os.environ.get("LOCAL_RANK", torch.cuda.current_device()))


def engine(hparams):
    """The engine of this process (created on first use; one handle per GPU, include/taco_b200.h)."""
    from .engine import TacoEngine
    dev = device_index()
    key = (dev, _state["precision"], id(_state["weights"]))
    if _state["engine"] is None or _state["key"] != key:
        close()
        _state["engine"] = TacoEngine(hparams, weights(hparams), device=dev, precision=_state["precision"])
        _state["key"] = (dev, _state["precision"], id(_state["weights"]))
    return _state["engine"]


def trainer(hparams):
    """The data-parallel trainer of this process (train/trainer.py): flat parameter / gradient / Adam buffers created
    from the installed weight container on first use; ``set_weights`` drops it."""
    from .train.trainer import DataParallelTrainer
    dev = device_index()
    key = (dev, id(_state["weights"]))
    if _state["trainer"] is None or _state["trainer_key"] != key:
        _state["trainer"] = DataParallelTrainer(hparams, weights(hparams), device=dev)
        _state["trainer_key"] = (dev, id(_state["weights"]))
    return _state["trainer"]


def close():
    if _state["engine"] is not None:
        _state["engine"].close()
    _state["engine"] = None
    _state["key"] = None
    _state["trainer"] = None
    _state["trainer_key"] = None


def install_aliases():
    """Expose the mirror under the reference's top-level import names (``TacotronModel.*``,
    ``Utils.*``) so that reference call sites import unchanged."""
    import importlib
    import sys
    for short in ("TacotronModel", "TacotronModel.modules", "TacotronModel.modules.Tacotron",
                  "TacotronModel.modules.Encoder", "TacotronModel.modules.Decoder",
                  "TacotronModel.modules.Attention", "TacotronModel.modules.Postnet", "TacotronModel.synthesize",
                  "Utils", "Utils.Helpers", "Utils.Hyperparams", "Utils.Infolog", "Utils.Utils",
                  "Utils.Tacotron_feeder", "Utils.Tacotron_synthesizer"):
        sys.modules.setdefault(short, importlib.import_module(__package__ + "." + short))
