import pytest

from tacotron_prosody_multi_gpu_b200.hparams import default_hparams


def test_parse_typed_overrides():
    hp = default_hparams()
    hp.parse("max_iters=720,tacotron_synthesis_batch_size=64,stop_at_any=False,prenet_layers=[128,128],dropout_rate=0.25")
    assert hp.max_iters == 720 and hp.tacotron_synthesis_batch_size == 64
    assert hp.stop_at_any is False and hp.prenet_layers == [128, 128] and hp.dropout_rate == 0.25
    assert hp.values()["max_iters"] == 720
    with pytest.raises(ValueError):
        hp.parse("no_such_field=1")
    assert default_hparams().outputs_per_step == 2 and default_hparams().tacotron_random_seed == 45454
