"""``from Utils.Hyperparams import hparams`` (Utils/Hyperparams.py:4) without TensorFlow."""
from ..hparams import HParams, default_hparams  # noqa: F401

hparams = default_hparams()


def hparams_debug_string():
    values = hparams.values()
    hp = ["  %s: %s" % (name, values[name]) for name in sorted(values) if name != "sentences"]
    return "Hyperparameters:\n" + "\n".join(hp)
